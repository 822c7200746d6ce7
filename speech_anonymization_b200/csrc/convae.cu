// ConvAutoencoder forward (models/ConvAutoEncoder.py:136-200) lowered to tap GEMMs.
//
// The reference flattens the [B,T,F] feature grid to ONE 1-D signal of length L = T*F
// with a single channel (:181,188) and runs ten 1-D convolutions over it; the adversarial
// TDNN sex classifier (:30-69) reads the 128-channel bottleneck.  Here:
//   encoder.0 (1->32, k15)            first_layer_kernel   (CUDA cores; K = 15 is too thin for UMMA)
//   encoder.2..11, decoder.0..8, tdnn tap GEMMs            (tapgemm.cuh; tcgen05 in fp16 mode)
//   InstanceNorm1d                    statistics in the producer's epilogue, affine + SiLU in
//                                     the consumer's operand load (instnorm_finalize_kernel
//                                     turns the sums into per-(b,c) scale/shift)
//   eval-mode BatchNorm1d             folded into per-channel scale/shift (bn_fold_kernel)
//   reshape + StatisticsPooling       pool_kernel on the reference's channels-first memory image
//   classify MLP + log_softmax        clf_head_kernel
// Utterances are processed in chunks sized so that a layer's output is still in the 126 MB L2
// when the next layer reads it.
#include "tapgemm.cuh"
#include "pack.cuh"
#include <algorithm>
#include <math.h>
#include <stdlib.h>

namespace spanon {
int fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t st);

namespace {

// ------------------------------------------------------------------ weight packing (pack.cuh: PackMode, packed_src)
// fp32 [ntaps][cin_v][cout_v]
__global__ void pack_f32_kernel(const float* __restrict__ w, float* __restrict__ dst, int mode, int cin, int cout,
                                int k, int ntaps, int cin_v, int cout_v) {
  const int64_t n = (int64_t)ntaps * cin_v * cout_v;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int cov = (int)(i % cout_v);
    const int civ = (int)((i / cout_v) % cin_v);
    const int tap = (int)(i / ((int64_t)cout_v * cin_v));
    dst[i] = packed_src(w, mode, cin, cout, k, tap, civ, cov);
  }
}

// fp16 UMMA K-major core-matrix layout [ntaps][cin_v/8][cout_v][8]
__global__ void pack_umma_kernel(const float* __restrict__ w, __half* __restrict__ dst, int mode, int cin,
                                 int cout, int k, int ntaps, int cin_v, int cout_v) {
  const int64_t n = (int64_t)ntaps * cin_v * cout_v;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int e = (int)(i % 8);
    const int cov = (int)((i / 8) % cout_v);
    const int kc = (int)((i / (8 * (int64_t)cout_v)) % (cin_v / 8));
    const int tap = (int)(i / ((int64_t)cout_v * cin_v));
    dst[i] = __float2half_rn(packed_src(w, mode, cin, cout, k, tap, kc * 8 + e, cov));
  }
}

// ------------------------------------------------------------------ decoder.4 o decoder.5 as ONE transposed convolution
// models/ConvAutoEncoder.py:166-167: Conv1d(C0->C1,k5,p2) is followed directly (no norm, no activation) by
// ConvTranspose1d(C1->C2,k5,s2,p2,op1), so their composition is linear:
//   h[i]      = b4 + sum_{s=-2..2} x[i+s] W4[:, :, s+2]
//   y[2m+ph]  = b5 + sum_d h[m+d] W5[:, :, 2 - 2d + ph]          (d = -1,0,1 even phase; d = 0,1 odd phase)
//             = bc[ph] + sum_{u=-3..3} x[m+u] Wc[ph][u],   Wc[ph][u] = sum_{d, s=u-d} W4[s+2]^T W5[2-2d+ph]
// i.e. a 7-tap phase-view tap GEMM (tap u = -3 feeds only the even phase): one launch and one HBM round trip instead of
// two, 6.5 instead of 7.5 tap-slices of MMAs.  The composition assumes h exists at i = -1 and i = L2 (computed from the
// zero-padded x); the reference has no such rows, so dec45_edge_kernel subtracts their contribution from the three
// affected output rows of each utterance (y[0], y[2*L2-2], y[2*L2-1]).
// Composed weights in the fp16 UMMA layout [7][C0/8][2*C2][8]; w4 = Conv1d weight [C1][C0][5], w5 = ConvT weight [C1][C2][5].
__global__ void compose_dec45_kernel(const float* __restrict__ w4, const float* __restrict__ w5, __half* __restrict__ dst,
                                     int C0, int C1, int C2) {
  const int cout_v = 2 * C2;
  const int64_t n = (int64_t)7 * C0 * cout_v;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int e = (int)(i % 8);
    const int cov = (int)((i / 8) % cout_v);
    const int kc = (int)((i / (8 * (int64_t)cout_v)) % (C0 / 8));
    const int u = (int)(i / ((int64_t)cout_v * C0)) - 3;
    const int c0 = kc * 8 + e, ph = cov / C2, c2 = cov - ph * C2;
    float acc = 0.f;
    for (int d = -1; d <= 1; ++d) {
      const int sft = u - d, t = 2 - 2 * d + ph;
      if (sft < -2 || sft > 2 || t > 4) continue;
      for (int c1 = 0; c1 < C1; ++c1)
        acc = fmaf(__ldg(w4 + ((int64_t)c1 * C0 + c0) * 5 + sft + 2), __ldg(w5 + ((int64_t)c1 * C2 + c2) * 5 + t), acc);
    }
    dst[i] = __float2half_rn(acc);
  }
}

// The same composition on ROW PAIRS of the input (L_DEC45P): view row m = input rows (2m, 2m+1) -> 128 input channels
// (q, c0); the four output rows 4m .. 4m+3 = (input row s, phase ph) -> 128 columns (2s + ph, c2); pair tap p = -2 .. 2 holds
// Wc[ph][u = 2p + q - s] (zero for |u| > 3).  fp16 UMMA layout [5][128/8][128][8].
__global__ void compose_dec45_pair_kernel(const float* __restrict__ w4, const float* __restrict__ w5, __half* __restrict__ dst,
                                          int C0, int C1, int C2) {
  const int cin_v = 2 * C0, cout_v = 4 * C2;
  const int64_t n = (int64_t)5 * cin_v * cout_v;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int e = (int)(i % 8);
    const int cov = (int)((i / 8) % cout_v);
    const int kc = (int)((i / (8 * (int64_t)cout_v)) % (cin_v / 8));
    const int p = (int)(i / ((int64_t)cout_v * cin_v)) - 2;
    const int civ = kc * 8 + e, q = civ / C0, c0 = civ - q * C0;
    const int sp = cov / C2, c2 = cov - sp * C2, sr = sp >> 1, ph = sp & 1;
    const int u = 2 * p + q - sr;
    float acc = 0.f;
    if (u >= -3 && u <= 3) {
      for (int d = -1; d <= 1; ++d) {
        const int sft = u - d, t = 2 - 2 * d + ph;
        if (sft < -2 || sft > 2 || t > 4) continue;
        for (int c1 = 0; c1 < C1; ++c1)
          acc = fmaf(__ldg(w4 + ((int64_t)c1 * C0 + c0) * 5 + sft + 2), __ldg(w5 + ((int64_t)c1 * C2 + c2) * 5 + t), acc);
      }
    }
    dst[i] = __float2half_rn(acc);
  }
}

// bc[cov] = b5[c2] + sum_c1 b4[c1] * sum_{t = ph (mod 2)} W5[c1][c2][t],  ph = (cov / C2) & 1, c2 = cov % C2
// (n_cols = 2*C2 for the row view, 4*C2 for the pair view: the two input rows of a pair share the constants)
__global__ void compose_dec45_bias_kernel(const float* __restrict__ b4, const float* __restrict__ w5,
                                          const float* __restrict__ b5, float* __restrict__ bc, int C1, int C2, int n_cols) {
  const int cov = blockIdx.x * blockDim.x + threadIdx.x;
  if (cov >= n_cols) return;
  const int ph = (cov / C2) & 1, c2 = cov % C2;
  float acc = b5[c2];
  for (int c1 = 0; c1 < C1; ++c1) {
    float ws = 0.f;
    for (int t = ph; t < 5; t += 2) ws += w5[((int64_t)c1 * C2 + c2) * 5 + t];
    acc = fmaf(b4[c1], ws, acc);
  }
  bc[cov] = acc;
}

// The correction is linear in the four activated input rows x[0], x[1], x[L2-2], x[L2-1]; its matrices are composed
// once with the weights: E[j][r][c0][c2] (j = y[0] | y[2*L2-2] | y[2*L2-1]; r = nearer | farther input row) followed by
// the constants Ec[j][c2] = sum_c1 b4[c1] W5[c1][c2][tap_j].
constexpr int EDGE45_FLOATS = 3 * 2 * 64 * 32 + 3 * 32;
__global__ void compose_dec45_edge_kernel(const float* __restrict__ w4, const float* __restrict__ b4,
                                          const float* __restrict__ w5, float* __restrict__ E, int C0, int C1, int C2) {
  const int n_mat = 3 * 2 * C0 * C2;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_mat + 3 * C2) return;
  const int tap5[3] = {4, 0, 1};
  float acc = 0.f;
  if (i < n_mat) {
    const int c2 = i % C2, c0 = (i / C2) % C0, r = (i / (C2 * C0)) % 2, j = i / (2 * C2 * C0);
    // h[-1] = b4 + x[0] W4[k=3] + x[1] W4[k=4];   h[L2] = b4 + x[L2-1] W4[k=1] + x[L2-2] W4[k=0]
    const int k4 = j == 0 ? 3 + r : 1 - r;
    for (int c1 = 0; c1 < C1; ++c1)
      acc = fmaf(w4[((int64_t)c1 * C0 + c0) * 5 + k4], w5[((int64_t)c1 * C2 + c2) * 5 + tap5[j]], acc);
  } else {
    const int c2 = (i - n_mat) % C2, j = (i - n_mat) / C2;
    for (int c1 = 0; c1 < C1; ++c1) acc = fmaf(b4[c1], w5[((int64_t)c1 * C2 + c2) * 5 + tap5[j]], acc);
  }
  E[i] = acc;
}

// One block per utterance, 384 threads = 3 output rows x 4 slices of c0 x 32 channels.  x: raw planes [utt][8][L2][8]
// fp16 of the composed layer's input, scale/shift: its InstanceNorm affine [utt][64]; y: planes [utt][4][2*L2][8] fp16.
__global__ void __launch_bounds__(384)
dec45_edge_kernel(const __half* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ shift,
                  const float* __restrict__ E, __half* __restrict__ y, int L2, double* __restrict__ stats, int folded) {
  constexpr int C0 = 64, C2 = 32;
  __shared__ float xs[4][C0], part[3][4][C2];
  const int b = blockIdx.x, tid = threadIdx.x;
  if (tid < 4 * C0) {   // rows 0, 1, L2-2, L2-1 of the activated input
    const int r = tid / C0, c = tid % C0;
    const int row = r < 2 ? r : L2 - 4 + r;
    // (folded: the input planes are the pair view [utt][16][L2/2][8], plane (row & 1) * 8 + c / 8, row >> 1)
    const float v = folded ? __half2float(x[(((int64_t)b * (C0 / 4) + (row & 1) * (C0 / 8) + c / 8) * (L2 / 2) + (row >> 1)) * 8 + (c & 7)])
                           : __half2float(x[(((int64_t)b * (C0 / 8) + c / 8) * L2 + row) * 8 + (c & 7)]);
    xs[r][c] = silu_fast(fmaf(v, scale[(int64_t)b * C0 + c], shift[(int64_t)b * C0 + c]));
  }
  __syncthreads();
  {
    const int c2 = tid % C2, sl = (tid / C2) % 4, j = tid / (4 * C2);
    const float* near = xs[j == 0 ? 0 : 3];
    const float* far = xs[j == 0 ? 1 : 2];
    const float* e0 = E + ((int64_t)(j * 2 + 0) * C0) * C2 + c2;
    const float* e1 = E + ((int64_t)(j * 2 + 1) * C0) * C2 + c2;
    float acc = sl == 0 ? E[3 * 2 * C0 * C2 + j * C2 + c2] : 0.f;
#pragma unroll
    for (int c0 = sl * 16; c0 < sl * 16 + 16; ++c0) {
      acc = fmaf(near[c0], e0[c0 * C2], acc);
      acc = fmaf(far[c0], e1[c0 * C2], acc);
    }
    part[j][sl][c2] = acc;
  }
  __syncthreads();
  if (tid < 3 * C2) {
    const int c2 = tid % C2, j = tid / C2;
    const float corr = part[j][0][c2] + part[j][1][c2] + part[j][2][c2] + part[j][3][c2];
    const int64_t row = j == 0 ? 0 : 2 * (int64_t)L2 - 3 + j;
    __half* q = y + (((int64_t)b * (C2 / 8) + c2 / 8) * (2 * (int64_t)L2) + row) * 8 + (c2 & 7);
    const float was = __half2float(*q);
    const __half now = __float2half_rn(was - corr);
    *q = now;
    if (stats) {   // statistics already reduced by the producing kernel's epilogue: move them with the corrected sample
      const float nw = __half2float(now);
      atomicAdd(stats + ((int64_t)b * C2 + c2) * 2, (double)(nw - was));
      atomicAdd(stats + ((int64_t)b * C2 + c2) * 2 + 1, (double)nw * nw - (double)was * was);
    }
  }
}

// ------------------------------------------------------------------ small fusions
__global__ void bn_fold_kernel(SpanonBatchNorm bn, int C, float* __restrict__ scale, float* __restrict__ shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) {
    const float s = bn.weight[c] / sqrtf(bn.running_var[c] + 1e-5f);
    scale[c] = s;
    shift[c] = bn.bias[c] - bn.running_mean[c] * s;
  }
}

// D += A * B for one m16n8k16 fp16 tile with fp32 accumulation (legacy warp-level MMA; the tensor pipe is idle here)
__device__ __forceinline__ void mma16816(float* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// nn.InstanceNorm1d(affine=True): biased variance over the sequence, eps 1e-5
__global__ void instnorm_finalize_kernel(const double* stats, const float* __restrict__ gamma,
                                         const float* __restrict__ beta, int B, int C, double count,
                                         float* __restrict__ scale, float* __restrict__ shift) {
  pdl_launch_dependents();      // (launched as a programmatic dependent of the producer: common.cuh)
  pdl_wait();
  stats = pdl_acquire(stats);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B * C) {
    const int c = i % C;
    const double mean = stats[2 * i] / count;
    double var = stats[2 * i + 1] / count - mean * mean;
    if (var < 0.0) var = 0.0;
    const float rstd = (float)(1.0 / sqrt(var + 1e-5));
    const float s = gamma[c] * rstd;
    scale[i] = s;
    shift[i] = beta[c] - (float)mean * s;
  }
}

// InstanceNorm statistics of a chunk-planar fp16 plane set [B][P][R][8] (tcgen05 layer output) + finalize.
// The channel of (plane p, element j) is (p % (C/8))*8 + j (pair-folded outputs store each channel in two planes).
// grid = (row splits, P, B), block = 256: a thread accumulates 8 channels over its rows (one 16 B load per row),
// the block reduces through shuffles + shared memory and adds to stats[b][c][2] (fp64); instnorm_finalize_kernel then
// converts the sums to scale/shift.  HBM/L2-bound: one read of the plane.
constexpr int ST_NT = 256, ST_ROWS = 32;   // rows per thread

// The per-element work of the obvious loop (2 cvt + add + fma per value) made this pass issue-bound at ~50 % of HBM
// speed, so the sums are formed on the tensor pipe instead.  A lane holds two rows (16 B = 4 channel pairs each).
// For channel pair q the two 32-bit words (wA, wB) are used, unmoved, as the B fragment (k = this lane's 4 values,
// n = lane/4) AND as the A fragment (rows lane/4: low halves only; rows lane/4+8: high halves only).  Then
//   D[g][g]   = sum over the group's 8 values of lo^2  (even channel),  D[g+8][g] = sum of hi^2 (odd channel),
// which lane (g, t = g/2) holds in c[g&1] / c[2+(g&1)].  A second MMA with a constant 0/1 A fragment gives the plain
// sums: row 0 = even k (low halves), row 1 = odd k.  Products of fp16 values are exact in fp32; accumulation is fp32.
__global__ void __launch_bounds__(ST_NT)
instnorm_stats_kernel(const __half* __restrict__ x, int P, int R, int C, double* __restrict__ stats) {
  __shared__ float red[ST_NT / 32][16];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int p = blockIdx.y, b = blockIdx.z;
  const uint4* base = reinterpret_cast<const uint4*>(x) + ((int64_t)b * P + p) * R;
  const int r0 = blockIdx.x * (ST_NT * ST_ROWS);
  const int g = lane >> 2, t = lane & 3;
  const uint32_t one_a = g == 0 ? 0x00003C00u : g == 1 ? 0x3C000000u : 0u;   // (1,0) for row 0, (0,1) for row 1
  float sq[4][4], sm[4][4];
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int j = 0; j < 4; ++j) { sq[q][j] = 0.f; sm[q][j] = 0.f; }
  constexpr int ST_BATCH = 8;   // 16 B loads in flight per thread (issued back to back before any use)
#pragma unroll 1
  for (int i0 = 0; i0 < ST_ROWS; i0 += ST_BATCH) {
    uint4 v[ST_BATCH];
#pragma unroll
    for (int u = 0; u < ST_BATCH; ++u) {
      const int r = r0 + (i0 + u) * ST_NT + tid;
      // volatile asm keeps the whole batch of loads ahead of the MMAs (the compiler otherwise sinks them next to
      // their first use, leaving two loads in flight per thread)
      const uint4* src = base + (r < R ? r : 0);
      asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                   : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src));
      if (r >= R) v[u] = make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int u = 0; u < ST_BATCH; u += 2) {
      const uint32_t wa[4] = {v[u].x, v[u].y, v[u].z, v[u].w}, wb[4] = {v[u + 1].x, v[u + 1].y, v[u + 1].z, v[u + 1].w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        mma16816(sq[q], wa[q] & 0x0000FFFFu, wa[q] & 0xFFFF0000u, wb[q] & 0x0000FFFFu, wb[q] & 0xFFFF0000u, wa[q], wb[q]);
        mma16816(sm[q], one_a, 0u, one_a, 0u, wa[q], wb[q]);
      }
    }
  }
  float s[8], q[8];
  const bool diag = t == (g >> 1);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    q[2 * k] = diag ? ((g & 1) ? sq[k][1] : sq[k][0]) : 0.f;
    q[2 * k + 1] = diag ? ((g & 1) ? sq[k][3] : sq[k][2]) : 0.f;
    s[2 * k] = g == 0 ? sm[k][0] + sm[k][1] : 0.f;
    s[2 * k + 1] = g == 1 ? sm[k][0] + sm[k][1] : 0.f;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = warp_sum(s[j]); q[j] = warp_sum(q[j]); }
  if (lane == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { red[warp][j] = s[j]; red[warp][8 + j] = q[j]; }
  }
  __syncthreads();
  if (tid < 16) {
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < ST_NT / 32; ++w) a += red[w][tid];
    const int c = (p % (C / 8)) * 8 + (tid & 7);
    atomicAdd(stats + ((int64_t)b * C + c) * 2 + (tid >> 3), (double)a);   // fire and forget: the block retires at once
  }
}

// InstanceNorm statistics of an fp32 channels-last tensor [B][R][32] (ConvReconstruction's encoder.0 output on the fp32
// path): grid = (row splits, B), block 256 = 8 rows x 32 channels per step.
__global__ void __launch_bounds__(256)
instnorm_stats_f32c32_kernel(const float* __restrict__ x, int R, double* __restrict__ stats) {
  __shared__ float red[2][8][32];
  const int tid = threadIdx.x, c = tid & 31, rr = tid >> 5, b = blockIdx.y;
  const float* base = x + (int64_t)b * R * 32;
  const int rows_per_block = (R + (int)gridDim.x - 1) / (int)gridDim.x;
  const int r0 = blockIdx.x * rows_per_block, r1 = min(R, r0 + rows_per_block);
  float s = 0.f, q = 0.f;
  for (int r = r0 + rr; r < r1; r += 8) {
    const float v = __ldg(base + (int64_t)r * 32 + c);
    s += v; q = fmaf(v, v, q);
  }
  red[0][rr][c] = s; red[1][rr][c] = q;
  __syncthreads();
  if (tid < 64) {
    const int k = tid >> 5;
    double a = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += (double)red[k][w][c];
    atomicAdd(stats + ((int64_t)b * 32 + c) * 2 + k, a);
  }
}

// ------------------------------------------------------------------ encoder.0 (1 -> 32, k15, p7) + SiLU
constexpr int FL_NT = 256;
constexpr int FL_CHUNKS = 3;     // 256-row chunks per block of first_layer_mma_kernel
// ACT = false stores the raw convolution (ConvReconstruction puts an InstanceNorm between encoder.0 and its "GLU")
template <typename TAct, bool FAST_ACT, bool ACT = true>
__global__ void __launch_bounds__(FL_NT)
first_layer_kernel(const float* __restrict__ feats, int64_t feats_bstride, int64_t n_valid /*valid elems*/, int L,
                   int F, const float* __restrict__ norm_mean, const float* __restrict__ norm_std,
                   const int32_t* __restrict__ utt_max, float top_db, const float* __restrict__ w /*[32][1][15]*/,
                   const float* __restrict__ bias, TAct* __restrict__ out, int64_t out_bstride) {
  __shared__ float sig[FL_NT + 14];
  __shared__ __align__(16) float ws[15 * 32];
  __shared__ float bs[32];
  const int tid = threadIdx.x, b = blockIdx.y;
  const int l0 = blockIdx.x * FL_NT;
  const float* x = feats + (int64_t)b * feats_bstride;
  const float floor_db = utt_max ? key_to_float(utt_max[b]) - top_db : -INFINITY;
  for (int i = tid; i < FL_NT + 14; i += FL_NT) {
    const int l = l0 - 7 + i;
    float v = 0.f;
    if (l >= 0 && l < L && (int64_t)l < n_valid) {
      v = __ldg(x + l);
      if (norm_mean) {
        const int f = l % F;
        v = (fmaxf(v, floor_db) - __ldg(norm_mean + f)) / __ldg(norm_std + f);
      }
    }
    sig[i] = v;
  }
  for (int i = tid; i < 15 * 32; i += FL_NT) {
    const int t = i / 32, c = i - t * 32;
    ws[i] = __ldg(w + c * 15 + t);
  }
  if (tid < 32) bs[tid] = __ldg(bias + tid);
  __syncthreads();
  const int l = l0 + tid;
  float acc[32];
#pragma unroll
  for (int c = 0; c < 32; ++c) acc[c] = bs[c];
#pragma unroll
  for (int t = 0; t < 15; ++t) {
    const float xv = sig[tid + t];
    const float4* w4 = reinterpret_cast<const float4*>(ws + t * 32);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float4 wv = w4[q];
      acc[4 * q + 0] = fmaf(xv, wv.x, acc[4 * q + 0]);
      acc[4 * q + 1] = fmaf(xv, wv.y, acc[4 * q + 1]);
      acc[4 * q + 2] = fmaf(xv, wv.z, acc[4 * q + 2]);
      acc[4 * q + 3] = fmaf(xv, wv.w, acc[4 * q + 3]);
    }
  }
  if (l < L) {
    TAct* o = out + (int64_t)b * out_bstride + (int64_t)l * 32;
#pragma unroll
    for (int c = 0; c < 32; ++c) acc[c] = !ACT ? acc[c] : FAST_ACT ? silu_fast(acc[c]) : silu_precise(acc[c]);
    if (sizeof(TAct) == 4) {
      float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
      for (int q = 0; q < 8; ++q) o4[q] = make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
    } else {
      // fp16 planes are chunk-planar in the CONSUMER's view: encoder.2 (stride 2) sees row pairs as one row of 64
      // channels -> plane (l & 1) * 4 + q, row l >> 1  (tapgemm.cuh OUT_PAIR_FOLD)
      TAct* ob = out + (int64_t)b * out_bstride;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4* o4 = reinterpret_cast<uint4*>(ob + (((int64_t)((l & 1) * 4 + q) * (L >> 1) + (l >> 1)) << 3));
        __half2 h0 = __floats2half2_rn(acc[8 * q + 0], acc[8 * q + 1]);
        __half2 h1 = __floats2half2_rn(acc[8 * q + 2], acc[8 * q + 3]);
        __half2 h2 = __floats2half2_rn(acc[8 * q + 4], acc[8 * q + 5]);
        __half2 h3 = __floats2half2_rn(acc[8 * q + 6], acc[8 * q + 7]);
        uint4 r;
        r.x = *reinterpret_cast<uint32_t*>(&h0); r.y = *reinterpret_cast<uint32_t*>(&h1);
        r.z = *reinterpret_cast<uint32_t*>(&h2); r.w = *reinterpret_cast<uint32_t*>(&h3);
        *o4 = r;
      }
    }
  }
}

// Tensor-pipe form of encoder.0 for the fp16 path.  The CUDA-core kernel above issues 480 FMAs per output row and is
// issue-bound at ~1.5 TB/s of output; here a warp computes 32 rows x 32 channels with warp-level MMAs over a Toeplitz
// view of the signal:  A[i][t] = x[l + i + t - 7]  (16 x 16, tap 15 is a zero weight),  B[t][c] = w[c][t] / 2.
// Because consecutive taps are consecutive samples, an A fragment register is just two neighbouring samples: three
// 32-bit shared-memory loads per 16 rows (a second copy of the signal shifted by one sample keeps odd rows aligned).
// The accumulators start at bias/2, so the MMA result is h = y/2 and SiLU(y) = h + h*tanh(h) (one MUFU per value).
template <bool ACT>
__global__ void __launch_bounds__(FL_NT)
first_layer_mma_kernel(const float* __restrict__ feats, int64_t feats_bstride, int64_t n_valid, int L, int F,
                       const float* __restrict__ norm_mean, const float* __restrict__ norm_std,
                       const int32_t* __restrict__ utt_max, float top_db, const float* __restrict__ w /*[32][1][15]*/,
                       const float* __restrict__ bias, __half* __restrict__ out, int64_t out_bstride) {
  // A block walks FL_CHUNKS consecutive 256-row chunks: the weight fragments below (a third of the per-chunk
  // instructions of this issue-bound kernel) are then built once per 768 rows.  Two copies of the signal buffer
  // alternate, so one barrier per chunk is enough.
  __shared__ __align__(16) __half shbuf[2][2][FL_NT + 32];    // [buffer]: sh[0][i] = x[l0 - 7 + i], sh[1][i] = x[l0 - 6 + i]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, b = blockIdx.y;
  const float* x = feats + (int64_t)b * feats_bstride;
  const float floor_db = utt_max ? key_to_float(utt_max[b]) - top_db : -INFINITY;
  const int g = lane >> 2, t4 = lane & 3;
  uint32_t bw[4][2];       // B fragments: channel 8*nt + g, taps (2*t4, 2*t4+1) and (2*t4+8, 2*t4+9)
  float bz[4][2];          // bias/2 of channels 8*nt + 2*t4, +1 (accumulator columns)
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) {
    const float* wc = w + (8 * nt + g) * 15;
    const int ta = 2 * t4, tb = 2 * t4 + 8;
    constexpr float kh = ACT ? 0.5f : 1.f;     // SiLU runs on y/2; the raw store (ACT = false) keeps y
    const __half2 lo = __floats2half2_rn(kh * __ldg(wc + ta), kh * __ldg(wc + ta + 1));
    const __half2 hi = __floats2half2_rn(kh * __ldg(wc + tb), tb + 1 < 15 ? kh * __ldg(wc + tb + 1) : 0.f);
    bw[nt][0] = *reinterpret_cast<const uint32_t*>(&lo);
    bw[nt][1] = *reinterpret_cast<const uint32_t*>(&hi);
    bz[nt][0] = kh * __ldg(bias + 8 * nt + 2 * t4);
    bz[nt][1] = kh * __ldg(bias + 8 * nt + 2 * t4 + 1);
  }
  __half* ob = out + (int64_t)b * out_bstride;
  const int64_t plane_rows = L >> 1;
#pragma unroll 1
  for (int ck = 0; ck < FL_CHUNKS; ++ck) {
  const int l0 = (blockIdx.x * FL_CHUNKS + ck) * FL_NT;
  if (l0 >= L) break;
  __half (*sh)[FL_NT + 32] = shbuf[ck & 1];
  for (int i = tid; i < FL_NT + 32; i += FL_NT) {
    const int l = l0 - 7 + i;
    float v = 0.f;
    if (l >= 0 && l < L && (int64_t)l < n_valid) {
      v = __ldg(x + l);
      if (norm_mean) {
        const int f = l % F;
        v = (fmaxf(v, floor_db) - __ldg(norm_mean + f)) / __ldg(norm_std + f);
      }
    }
    const __half hv = __float2half_rn(v);
    sh[0][i] = hv;
    if (i > 0) sh[1][i - 1] = hv;
  }
  __syncthreads();
#pragma unroll
  for (int mt = 0; mt < 2; ++mt) {
    const int base = warp * 32 + mt * 16;              // first row of this 16-row tile inside the block
    const int e = base + g + 2 * t4;                   // sample index of (row g, tap 2*t4)
    const uint32_t* src = reinterpret_cast<const uint32_t*>((e & 1) ? &sh[1][e - 1] : &sh[0][e]);
    const uint32_t a0 = src[0], a12 = src[4], a3 = src[8];   // samples e.., e+8.., e+16..
    float acc[4][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      acc[nt][0] = bz[nt][0]; acc[nt][1] = bz[nt][1]; acc[nt][2] = bz[nt][0]; acc[nt][3] = bz[nt][1];
      mma16816(acc[nt], a0, a12, a12, a3, bw[nt][0], bw[nt][1]);
    }
#pragma unroll
    for (int hr = 0; hr < 2; ++hr) {                   // rows g and g + 8
      const int l = l0 + base + g + 8 * hr;
      if (l < L) {
        // fp16 planes are chunk-planar in the CONSUMER's view (encoder.2, stride 2): plane (l & 1) * 4 + q, row l >> 1
        __half* o = ob + ((((int64_t)(l & 1) * 4) * plane_rows + (l >> 1)) << 3) + 2 * t4;
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          __half2 hh = __floats2half2_rn(acc[nt][2 * hr], acc[nt][2 * hr + 1]);
          if (ACT) {
            uint32_t th, in = *reinterpret_cast<uint32_t*>(&hh);
            asm("tanh.approx.f16x2 %0, %1;" : "=r"(th) : "r"(in));
            hh = __hfma2(hh, *reinterpret_cast<__half2*>(&th), hh);
          }
          *reinterpret_cast<__half2*>(o + ((int64_t)nt * plane_rows << 3)) = hh;
        }
      }
    }
  }
  }   // chunk
}

// ------------------------------------------------------------------ reshape + StatisticsPooling
// ConvAutoEncoder.py:61 reinterprets the channels-first [128][L3] block as [L3][128]; pooled
// column j therefore collects flat[r*128 + j].  grid = (chunks, B), block = 256.
__global__ void __launch_bounds__(256)
pool_kernel(const float* __restrict__ x, int64_t bstride, int64_t n_elems /*128*L3*/, double* __restrict__ pool) {
  __shared__ double red[256][2];
  const int b = blockIdx.y, tid = threadIdx.x;
  const float* p = x + (int64_t)b * bstride;
  double s = 0.0, q = 0.0;
  // stride 256*gridDim.x keeps (index % 128) == tid % 128 for every element this thread reads
  for (int64_t i = blockIdx.x * 256 + tid; i < n_elems; i += (int64_t)gridDim.x * 256) {
    const float v = __ldg(p + i);
    s += (double)v;
    q += (double)v * (double)v;
  }
  red[tid][0] = s; red[tid][1] = q;
  __syncthreads();
  if (tid < 128) {
    s = red[tid][0] + red[tid + 128][0];
    q = red[tid][1] + red[tid + 128][1];
    atomicAdd(pool + ((int64_t)b * 128 + tid) * 2 + 0, s);
    atomicAdd(pool + ((int64_t)b * 128 + tid) * 2 + 1, q);
  }
}

__device__ __forceinline__ float bn_eval(float x, const SpanonBatchNorm& bn, int c) {
  return (x - bn.running_mean[c]) / sqrtf(bn.running_var[c] + 1e-5f) * bn.weight[c] + bn.bias[c];
}

// classify: Linear(256,128) ReLU BN | Linear(128,64) ReLU BN | Linear(64,2) | log_softmax.  grid = B, block = 256
__global__ void __launch_bounds__(256)
clf_head_kernel(const double* pool, const float* __restrict__ noise, int64_t n /*L3*/,
                SpanonConvAEWeights w, float* __restrict__ logp, const unsigned* err_flag) {
  // 256 threads: eight warps share the output rows (a warp per row, the lanes read 128 contiguous bytes of the weight row per
  // step, sixteen rows' loads in flight per warp).  The 160 KB of weights are cold in L2 every step (5.5 GB of planes pass
  // between two calls), so the launch is one DRAM-latency chain per batch of rows: a thread per row (round 1) took 36 us,
  // four warps 26 us.
  __shared__ float s0[256], s1[128], s2[64], s3[2];
  const int b = blockIdx.x, tid = threadIdx.x, t128 = tid & 127, t64 = tid & 63;
  pdl_launch_dependents();      // (common.cuh: launched as a programmatic dependent of tdnn.6 / the pooling pass)
  // (bias / BatchNorm parameters are fetched up front, before the wait for the producer)
  const float b0 = w.cls_b[0][t128], g0 = w.cls_bn[0].weight[t128], be0 = w.cls_bn[0].bias[t128],
              m0 = w.cls_bn[0].running_mean[t128], v0 = w.cls_bn[0].running_var[t128];
  const float b1 = w.cls_b[1][t64], g1 = w.cls_bn[1].weight[t64], be1 = w.cls_bn[1].bias[t64], m1 = w.cls_bn[1].running_mean[t64],
              v1 = w.cls_bn[1].running_var[t64];
  pdl_wait();
  pool = pdl_acquire(pool);     // (produced by the predecessor: see common.cuh)
  err_flag = pdl_acquire(err_flag);
  if (tid < 128) {
    const double s = pool[((int64_t)b * 128 + tid) * 2 + 0], q = pool[((int64_t)b * 128 + tid) * 2 + 1];
    const double mean = s / (double)n;
    double var = n > 1 ? (q - s * s / (double)n) / (double)(n - 1) : (double)NAN;
    if (var < 0.0) var = 0.0;
    s0[tid] = (float)mean + (noise ? noise[(int64_t)b * 128 + tid] : 0.f);
    s0[128 + tid] = (float)sqrt(var) + 1e-5f;
  }
  __syncthreads();
  const int lane = tid & 31, warp = tid >> 5;
#pragma unroll 16
  for (int j = warp; j < 128; j += 8) {
    const float* wr = w.cls_w[0] + j * 256 + lane;
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) a = fmaf(s0[lane + 32 * k], __ldg(wr + 32 * k), a);
    a = warp_sum(a);
    if (lane == 0) s1[j] = a;
  }
  __syncthreads();
  if (tid < 128) s1[tid] = (fmaxf(s1[tid] + b0, 0.f) - m0) / sqrtf(v0 + 1e-5f) * g0 + be0;
  __syncthreads();
#pragma unroll 8
  for (int j = warp; j < 64; j += 8) {
    const float* wr = w.cls_w[1] + j * 128 + lane;
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) a = fmaf(s1[lane + 32 * k], __ldg(wr + 32 * k), a);
    a = warp_sum(a);
    if (lane == 0) s2[j] = a;
  }
  __syncthreads();
  if (tid < 64) s2[tid] = (fmaxf(s2[tid] + b1, 0.f) - m1) / sqrtf(v1 + 1e-5f) * g1 + be1;
  __syncthreads();
  if (warp < 2) {
    const float a = warp_sum(fmaf(s2[lane], __ldg(w.cls_w[2] + warp * 64 + lane), s2[lane + 32] * __ldg(w.cls_w[2] + warp * 64 + lane + 32)));
    if (lane == 0) s3[warp] = a + w.cls_b[2][warp];
  }
  __syncthreads();
  if (tid < 2) {
    const float m = fmaxf(s3[0], s3[1]);
    const float lse = m + logf(expf(s3[0] - m) + expf(s3[1] - m));
    // a dataflow wait of the TDNN launch timed out: its planes are incomplete, say so instead of returning numbers
    logp[(int64_t)b * 2 + tid] = (err_flag && *err_flag) ? NAN : s3[tid] - lse;
  }
}

// Same pooling for the tcgen05 path, whose tdnn.6 output is the chunk-planar fp16 layout [B][P][R][8] (plane p =
// channels 8p..8p+7): element (c, l) of the channels-first image has flat index c*R + l and is pooled into column
// (c*R + l) % 128.  grid = (row splits, P, B); a block's first row is a multiple of 128, so row % 128 == tid % 128
// for every row a thread reads and its eight (channel -> column) bins are fixed: register accumulators, then an
// eight-round conflict-free shared-memory fold and one fp64 atomic pair per column.
constexpr int PL_NT = 256, PL_ROWS = 32;
__global__ void __launch_bounds__(PL_NT)
pool_planar_kernel(const __half* __restrict__ x, int P, int R, const float* __restrict__ bias,
                   const float* __restrict__ bn_scale, const float* __restrict__ bn_shift, double* __restrict__ pool) {
  __shared__ float part[2][8][128];
  __shared__ float bins[2][128];
  const int tid = threadIdx.x, p = blockIdx.y, b = blockIdx.z;
  const uint4* base = reinterpret_cast<const uint4*>(x) + ((int64_t)b * P + p) * R;
  const int r0 = blockIdx.x * (PL_NT * PL_ROWS);
  float s[8], q[8], cb[8], cs[8], ch[8];   // the plane holds the raw tdnn.6 conv: ReLU(x + bias) * bn_scale + bn_shift here
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    s[j] = 0.f; q[j] = 0.f;
    cb[j] = __ldg(bias + 8 * p + j); cs[j] = __ldg(bn_scale + 8 * p + j); ch[j] = __ldg(bn_shift + 8 * p + j);
  }
  constexpr int PL_BATCH = 8;   // 16 B loads in flight per thread (issued back to back before any use, as in instnorm_stats_kernel)
#pragma unroll 1
  for (int i0 = 0; i0 < PL_ROWS; i0 += PL_BATCH) {
    uint4 v[PL_BATCH];
#pragma unroll
    for (int u = 0; u < PL_BATCH; ++u) {
      const int r = r0 + (i0 + u) * PL_NT + tid;
      const uint4* src = base + (r < R ? r : 0);
      asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                   : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src));
    }
#pragma unroll
    for (int u = 0; u < PL_BATCH; ++u) {
      if (r0 + (i0 + u) * PL_NT + tid < R) {
        const __half2* h = reinterpret_cast<const __half2*>(&v[u]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float2 f = __half22float2(h[j]);
          f.x = fmaf(fmaxf(f.x + cb[2 * j], 0.f), cs[2 * j], ch[2 * j]);
          f.y = fmaf(fmaxf(f.y + cb[2 * j + 1], 0.f), cs[2 * j + 1], ch[2 * j + 1]);
          s[2 * j] += f.x; s[2 * j + 1] += f.y;
          q[2 * j] = fmaf(f.x, f.x, q[2 * j]); q[2 * j + 1] = fmaf(f.y, f.y, q[2 * j + 1]);
        }
      }
    }
  }
  if (tid >= 128) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { part[0][j][tid - 128] = s[j]; part[1][j][tid - 128] = q[j]; }
  } else {
    bins[0][tid] = 0.f; bins[1][tid] = 0.f;
  }
  __syncthreads();
  const int Rm = R & 127;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (tid < 128) {   // for a fixed j, tid -> bin is a bijection on [0,128)
      const int bin = ((8 * p + j) * Rm + tid) & 127;
      bins[0][bin] += s[j] + part[0][j][tid];
      bins[1][bin] += q[j] + part[1][j][tid];
    }
    __syncthreads();
  }
  if (tid < 128) {
    atomicAdd(pool + ((int64_t)b * 128 + tid) * 2 + 0, (double)bins[0][tid]);
    atomicAdd(pool + ((int64_t)b * 128 + tid) * 2 + 1, (double)bins[1][tid]);
  }
}

// ------------------------------------------------------------------ layer table
struct LayerSpec {
  int mode;                 // PackMode
  int cin, cout, k;         // real sizes
  int cin_v, cout_v;        // view widths
  int ntaps;
  int shift[MAX_TAPS];
  int rows_div;             // rows = L / rows_div (view rows of the input)
  int rows_shrink;          // rows_out = rows_in - rows_shrink (TDNN, no padding)
};

enum LayerId { L_ENC2, L_ENC5, L_ENC8, L_ENC11, L_TD0, L_TD3, L_TD6, L_DEC0, L_DEC1, L_DEC4, L_DEC5, L_DEC8, L_DEC45, L_ENC5P, L_DEC45P, N_LAYERS };
const char* const kLayerTag[N_LAYERS] = {"tapgemm:enc2", "tapgemm:enc5", "tapgemm:enc8", "tapgemm:enc11", "tapgemm:tdnn0", "tapgemm:tdnn3",
                                        "tapgemm:tdnn6", "tapgemm:dec0", "tapgemm:dec1", "tapgemm:dec4", "tapgemm:dec5", "tapgemm:dec8",
                                        "tapgemm:dec45", "tapgemm:enc5", "tapgemm:dec45"};

// tc = this layer runs on the tcgen05 kernel (fp16 UMMA-packed weights)
LayerSpec make_spec(int id, bool tc) {
  LayerSpec s{};
  auto taps = [&](std::initializer_list<int> sh) { s.ntaps = 0; for (int v : sh) s.shift[s.ntaps++] = v; };
  switch (id) {
    case L_ENC2:  s = {PACK_CONV_S2, 32, 64, 5, 64, 64, 0, {}, 2, 0};    taps({-1, 0, 1}); break;
    case L_ENC5:  s = {PACK_CONV_S1, 64, 64, 5, 64, 64, 0, {}, 2, 0};    taps({-2, -1, 0, 1, 2}); break;
    // encoder.5 on ROW PAIRS (tcgen05 path): 64 -> 64 channels, k5 becomes a 3-tap 128 -> 128 layer over L/4 pair rows (10 of
    // the 12 64x64 blocks non-zero: 20 % more MMA work, but 128 columns per tile instead of 64 -- the N=64 form is bound by
    // the A-operand fetch).  Its input is encoder.2's output pair-folded, its 128 output columns ARE encoder.8's pair view.
    case L_ENC5P: s = {PACK_CONV_S1_PAIR, 64, 64, 5, 128, 128, 0, {}, 4, 0}; taps({-1, 0, 1}); break;
    case L_ENC8:  s = {PACK_CONV_S2, 64, 128, 5, 128, 128, 0, {}, 4, 0}; taps({-1, 0, 1}); break;
    case L_ENC11: s = {PACK_CONV_S1, 128, 128, 5, 128, 128, 0, {}, 4, 0}; taps({-2, -1, 0, 1, 2}); break;
    case L_TD0:   s = {PACK_CONV_S1, 128, 128, 5, 128, 128, 0, {}, 4, 4}; taps({0, 1, 2, 3, 4}); break;
    case L_TD3:   s = {PACK_CONV_S1, 128, 128, 3, 128, 128, 0, {}, 4, 4}; taps({0, 2, 4}); break;
    case L_TD6:   s = {PACK_CONV_S1, 128, 128, 3, 128, 128, 0, {}, 4, 6}; taps({0, 3, 6}); break;
    case L_DEC0:  s = {PACK_CONV_S1, 128, 128, 5, 128, 128, 0, {}, 4, 0}; taps({-2, -1, 0, 1, 2}); break;
    case L_DEC1:  s = {PACK_CONVT_S2, 128, 64, 5, 128, 128, 0, {}, 4, 0}; taps({-1, 0, 1}); break;
    case L_DEC4:  s = {PACK_CONV_S1, 64, 64, 5, 64, 64, 0, {}, 2, 0};    taps({-2, -1, 0, 1, 2}); break;
    case L_DEC5:  s = {PACK_CONVT_S2, 64, 32, 5, 64, 64, 0, {}, 2, 0};   taps({-1, 0, 1}); break;
    case L_DEC45: s = {PACK_COMPOSED, 64, 32, 7, 64, 64, 0, {}, 2, 0};   taps({-3, -2, -1, 0, 1, 2, 3}); break;
    // the composed layer on ROW PAIRS (compose_dec45_pair_kernel): 5 pair taps, 128 -> 128 columns over L/4 pair rows
    case L_DEC45P: s = {PACK_COMPOSED, 64, 32, 7, 128, 128, 0, {}, 4, 0}; taps({-2, -1, 0, 1, 2}); break;
    case L_DEC8:
      if (tc) { s = {PACK_TOEPLITZ, 32, 1, 15, 32, 16, 0, {}, 1, 0}; taps({-7}); }   // Toeplitz GEMM, see conv_tc.cu
      else { s = {PACK_CONV_S1, 32, 1, 15, 32, 4, 0, {}, 1, 0}; taps({-7, -6, -5, -4, -3, -2, -1, 0, 1, 2, 3, 4, 5, 6, 7}); }
      break;
  }
  return s;
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Workspace {
  // carved in this order from the caller's buffer (all 256-byte aligned)
  void* packed[N_LAYERS];
  float* bn_scale[4]; float* bn_shift[4];      // clf norm, tdnn bn 0..2
  float* ones;                                 // [128] 1.0f (unit scale of a bias-only affine on load)
  float* bias45;                               // [64 | 128] per-column bias of the composed decoder.4 o decoder.5 layer
  float* edge45;                               // [EDGE45_FLOATS] its edge-correction matrices (dec45_edge_kernel)
  double* stats[6];                            // enc3, enc6, enc9, enc12, dec2, dec6   [chunk][C][2]
  float* in_scale[6]; float* in_shift[6];      // [chunk][C]
  double* pool;                                // [chunk][128][2]
  unsigned* flags;                             // [2][chunk * ceil(L/4 / 128)] per-tile counters of the TDNN dataflow launch
  size_t zero_bytes;                           // stats[0] .. end of flags: cleared per chunk
  unsigned* counter;                           // last-CTA ticket of the tcgen05 kernel's fused finalize
  void* act[4];                                // activation ring, each chunk*L*32 elements
  float* tdnn_out;                             // [chunk][128][L4-14] fp32
  size_t total;
  const int* inC;                              // channels of each InstanceNorm slot
};

const int kInC[6] = {64, 64, 128, 128, 64, 32};        // ConvAutoencoder: channels of the six InstanceNorms
const int kInC_recon[6] = {32, 64, 64, 32, 8, 8};      // ConvReconstruction: encoder.{1,4,7,10} (two unused slots)

int default_chunk(int64_t L, size_t esize) {
  // Utterances per pass over the layer list.  Fewer, longer launches (better tile balance of the persistent kernels
  // over 148 SMs) beat keeping a layer pair's planes in the 126 MB L2 (bench.py --chunk on 60 x 10 s: a chunk of 6 is
  // ~1.5x slower per step than the whole batch); the only cap is ~2 GB of activation planes per workspace.
  const double per = (double)L * 32.0 * (double)esize;      // one plane of one utterance
  int c = (int)(2.0e9 / (6.0 * per));
  return std::max(1, std::min(c, 64));
}

void carve(Workspace& ws, char* base, int64_t chunk, int64_t L, bool tc, size_t esize, const int* inC = kInC,
           bool with_tdnn = true) {
  ws.inC = inC;
  size_t off = 0;
  auto take = [&](size_t bytes) { void* p = base ? base + off : nullptr; off += align_up(bytes, 256); return p; };
  for (int i = 0; i < N_LAYERS; ++i) {
    LayerSpec s = make_spec(i, tc);
    ws.packed[i] = take((size_t)s.ntaps * s.cin_v * std::max(s.cout_v, 16) * 4);
  }
  for (int i = 0; i < 4; ++i) { ws.bn_scale[i] = (float*)take(128 * 4); ws.bn_shift[i] = (float*)take(128 * 4); }
  ws.ones = (float*)take(128 * 4);
  ws.bias45 = (float*)take(128 * 4);
  ws.edge45 = (float*)take(EDGE45_FLOATS * 4);
  // stats, pool and flags are contiguous so one memset clears them
  const size_t stats_off = off;
  for (int i = 0; i < 6; ++i) ws.stats[i] = (double*)take((size_t)chunk * inC[i] * 2 * sizeof(double));
  ws.pool = (double*)take((size_t)chunk * 128 * 2 * sizeof(double));
  ws.flags = (unsigned*)take(with_tdnn ? 2 * (size_t)chunk * (size_t)ceil_div(std::max<int64_t>(L / 4, 1), 128) * sizeof(unsigned) : 0);
  ws.zero_bytes = off - stats_off;
  ws.counter = (unsigned*)take(256);
  for (int i = 0; i < 6; ++i) {
    ws.in_scale[i] = (float*)take((size_t)chunk * inC[i] * 4);
    ws.in_shift[i] = (float*)take((size_t)chunk * inC[i] * 4);
  }
  for (int i = 0; i < 4; ++i) ws.act[i] = take((size_t)chunk * L * 32 * esize + 1024);
  ws.tdnn_out = (float*)take(with_tdnn ? (size_t)chunk * 128 * (size_t)std::max<int64_t>(L / 4 - 14, 1) * 4 : 256);
  ws.total = off;
}

struct Ctx {
  const SpanonConvAEWeights* w;
  Workspace ws;
  bool tc;       // fp16 activations (+ tcgen05 where supported)
  int n_sm;
  int64_t L;
  cudaStream_t st;
  bool dec45_pair = false; // decoder.4 o decoder.5 as a 5-tap 128 -> 128 layer on row pairs (L_DEC45P; opt-in with $SPANON_DEC45_PAIR=1)
  bool enc5_pair = false;  // encoder.5 as a 128 -> 128 layer on row pairs (L_ENC5P; tcgen05 path, $SPANON_ENC5_PAIR=0: N=64 form)
  bool prepared = false;   // weights_prepared call: no weight-preparation kernel runs ahead of the first layer
};

bool layer_on_tc(const Ctx& c, int id) { return c.tc; (void)id; }

// $SPANON_COMPOSE_DEC=0 runs decoder.4 and decoder.5 as two launches on the tcgen05 path too (A/B switch)
bool compose_dec45(const Ctx& c) {
  static const bool on = !(getenv("SPANON_COMPOSE_DEC") && getenv("SPANON_COMPOSE_DEC")[0] == '0');
  return c.tc && on;
}

int pack_layer(const Ctx& c, int id, const float* w) {
  const bool on_tc = layer_on_tc(c, id);
  LayerSpec s = make_spec(id, on_tc);
  const int64_t n = (int64_t)s.ntaps * s.cin_v * s.cout_v;
  const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(n, 256), 1024);
  if (on_tc)
    pack_umma_kernel<<<grid, 256, 0, c.st>>>(w, (__half*)c.ws.packed[id], s.mode, s.cin, s.cout, s.k, s.ntaps,
                                             s.cin_v, s.cout_v);
  else
    pack_f32_kernel<<<grid, 256, 0, c.st>>>(w, (float*)c.ws.packed[id], s.mode, s.cin, s.cout, s.k, s.ntaps, s.cin_v,
                                            s.cout_v);
  return check_launch("pack_kernel");
}

struct LayerIO {
  const void* in; void* out;
  int in_norm;        // index into stats/in_scale of the InstanceNorm applied on load (+SiLU), -1 none
  int t2_bn;          // bn_scale index applied after the activation on load, -1 none
  const float* bias;
  int relu;
  int e_bn;           // bn index folded into the epilogue, -1 none
  int out_stats;      // stats index accumulated in the epilogue, -1 none
  int out_mode;
  const float* fin_gamma = nullptr;  // affine of the InstanceNorm that consumes out_stats
  const float* fin_beta = nullptr;
  double fin_count = 0.0;
  int out_fold = OUT_DIRECT;         // fp16 mode: how the consumer views this layer's output planes
  const float* t1s = nullptr;        // tcgen05 path: per-channel affine on load (overrides in_norm), e.g. the
  const float* t1h = nullptr;        //   producer's bias as (ones, bias)
  int act = -1;                      // activation on load (TapGemmParams::silu); -1 = SiLU iff in_norm >= 0
  int out_half = 0;                  // tcgen05 path, OUT_CH0_F32: fp16 output (recon_fp16)
  const TapGemmParams* gen = nullptr; // tcgen05 path: gen_* fields of a launch that generates its input (encoder.0 fused)
};

TapGemmParams make_params(const Ctx& c, int id, const LayerIO& io, int nb) {
  const bool on_tc = layer_on_tc(c, id);
  LayerSpec s = make_spec(id, on_tc);
  TapGemmParams p{};
  p.in = io.in; p.out = io.out;
  p.rows_in = (int)(c.L / s.rows_div);
  if (id == L_TD3) p.rows_in -= 4;
  if (id == L_TD6) p.rows_in -= 8;
  p.rows_out = p.rows_in - s.rows_shrink;
  p.cin_v = s.cin_v; p.cout_v = s.cout_v;
  p.cin_mod = s.cin; p.cout_mod = s.cout;
  p.in_bstride = (int64_t)p.rows_in * s.cin_v;
  p.out_bstride = io.out_mode == OUT_CH0_F32 ? (int64_t)p.rows_out : (int64_t)p.rows_out * s.cout_v;
  p.ntaps = s.ntaps;
  for (int t = 0; t < s.ntaps; ++t) {
    p.shift[t] = s.shift[t];
    p.tap_kmax[t] = s.cin_v;
    p.tap_nmax[t] = s.cout_v;
  }
  p.tap_step = s.ntaps > 1 ? s.shift[1] - s.shift[0] : 0;
  if (s.mode == PACK_CONV_S2) p.tap_kmax[s.ntaps - 1] = s.cin;   // tap (2*tap+1) = 5 does not exist
  if ((s.mode == PACK_CONVT_S2 || s.mode == PACK_COMPOSED) && id != L_DEC45P) p.tap_nmax[0] = s.cout;   // the first shift only feeds the even phase
  p.w = c.ws.packed[id];
  p.t1_bstride = s.cin;
  if (io.in_norm >= 0) { p.t1_scale = c.ws.in_scale[io.in_norm]; p.t1_shift = c.ws.in_shift[io.in_norm]; p.silu = 1; }
  if (io.t1s) { p.t1_scale = io.t1s; p.t1_shift = io.t1h; p.t1_bstride = 0; }
  if (io.act >= 0) p.silu = io.act;
  if (io.t2_bn >= 0) { p.t2_scale = c.ws.bn_scale[io.t2_bn]; p.t2_shift = c.ws.bn_shift[io.t2_bn]; }
  p.bias = io.bias;
  p.bias_v = (id == L_DEC45 || id == L_DEC45P) ? 1 : 0;
  p.relu = io.relu;
  if (io.e_bn >= 0) { p.e_scale = c.ws.bn_scale[io.e_bn]; p.e_shift = c.ws.bn_shift[io.e_bn]; }
  p.stats = io.out_stats >= 0 ? c.ws.stats[io.out_stats] : nullptr;
  p.out_mode = io.out_mode;
  p.out_fold = io.out_fold;
  p.out_plane_rows = io.out_fold == OUT_PAIR_FOLD ? p.rows_out / 2 : io.out_fold == OUT_PHASE_UNFOLD ? p.rows_out * 2
                     : io.out_fold == OUT_PHASE_UNFOLD4 ? p.rows_out * 4 : p.rows_out;
  p.in_planar = c.tc ? 1 : 0;
  p.out_half = io.out_half;
  if (io.gen) {
    p.gen_feats = io.gen->gen_feats; p.gen_bstride = io.gen->gen_bstride; p.gen_nvalid = io.gen->gen_nvalid;
    p.gen_F = io.gen->gen_F; p.gen_mean = io.gen->gen_mean; p.gen_std = io.gen->gen_std; p.gen_umax = io.gen->gen_umax;
    p.gen_topdb = io.gen->gen_topdb; p.gen_w = io.gen->gen_w; p.gen_b = io.gen->gen_b;
  }
  p.B = nb;
  p.mma_group = (id == L_DEC45 || id == L_TD3 || id == L_TD6) ? 4 : 1;
  p.tag = io.gen ? "tapgemm:enc02" : kLayerTag[id];   // enc02 = encoder.0 generated inside encoder.2's launch
  return p;
}

int run_layer(const Ctx& c, int id, const LayerIO& io, int nb) {
  const bool on_tc = layer_on_tc(c, id);
  LayerSpec s = make_spec(id, on_tc);
  TapGemmParams p = make_params(c, id, io, nb);
  // InstanceNorm statistics of a tcgen05 layer's output: reduced by the producing kernel's epilogue (default), or by a
  // separate bandwidth-bound pass over the stored plane (instnorm_stats_kernel, $SPANON_FUSED_STATS=0).  Measured on
  // B200 (tools/ab_env.sh SPANON_FUSED_STATS 0 1, 60 x 10 s): +25..+48 us per producing launch against 66 + 6 us for
  // the separate pass and its 310 MB of re-read, 2.33 -> 2.20 ms per step.  (Two earlier forms, the transform warps
  // reading the stored tiles back through L2, and the butterfly on four epilogue warps, were slower than the pass.)
  static const bool fused_stats = !(getenv("SPANON_FUSED_STATS") && getenv("SPANON_FUSED_STATS")[0] == '0');
  const bool fuse = on_tc && io.out_stats >= 0 && fused_stats;
  const bool split_stats = on_tc && io.out_stats >= 0 && !fuse;
  if (split_stats) p.stats = nullptr;
  if (on_tc) {
    // programmatic dependent launch for every tcgen05 layer but the first of a forward (its predecessor may be a weight
    // preparation kernel, whose output the prologue reads)
    p.pdl = id != L_ENC2 || c.prepared;
    // $SPANON_TC_STAGES: cap on the one-SM kernel's operand ring, 8 .. 32 stages of 16 channels.  Default 16: a deeper ring
    // measured no gain there (encoder.5 178 vs 180 us, decoder.4o5 210 vs 211), unlike on the CTA-pair kernel, whose ring
    // holds 8 stages per tile (32 stages: tdnn.3 122 -> 114, decoder.0 153 -> 146, tdnn.6 157 -> 152 us).
    static const int ring_cap_env = getenv("SPANON_TC_STAGES") ? atoi(getenv("SPANON_TC_STAGES")) : 16;
    p.ring_cap = ring_cap_env;
    // the 128 -> 128 layers without output statistics on CTA pairs (conv_tc2.cu); $SPANON_TC2=0: one-SM kernel
    static const bool tc2_env = !(getenv("SPANON_TC2") && getenv("SPANON_TC2")[0] == '0');
    if (tc2_env && tapgemm_tc2_supported(p)) SPANON_TRY(launch_tapgemm_tc2(p, c.n_sm, c.st));
    else
    SPANON_TRY(launch_tapgemm_tc(p, c.n_sm, c.st));
    if (id == L_DEC45 || id == L_DEC45P) {   // take the two rows of decoder.4 the reference does not have back out (before the statistics)
      const int folded = id == L_DEC45P;
      dec45_edge_kernel<<<(unsigned)nb, 384, 0, c.st>>>((const __half*)io.in, p.t1_scale, p.t1_shift, c.ws.edge45,
                                                        (__half*)io.out, folded ? 2 * p.rows_in : p.rows_in,
                                                        fuse ? c.ws.stats[io.out_stats] : nullptr, folded);
      SPANON_TRY(check_launch("dec45_edge_kernel"));
    }
    if (split_stats) {   // one bandwidth-bound pass over the fp16 plane just written (cheaper than reducing in the epilogue)
      const int P = p.out_fold == OUT_PAIR_FOLD ? s.cout_v / 4 : p.out_fold == OUT_PHASE_UNFOLD ? s.cout_v / 16
                    : p.out_fold == OUT_PHASE_UNFOLD4 ? s.cout_v / 32 : s.cout_v / 8;
      dim3 grid((unsigned)ceil_div(p.out_plane_rows, ST_NT * ST_ROWS), (unsigned)P, (unsigned)nb);
      instnorm_stats_kernel<<<grid, ST_NT, 0, c.st>>>((const __half*)io.out, P, p.out_plane_rows, s.cout,
                                                      c.ws.stats[io.out_stats]);
      SPANON_TRY(check_launch("instnorm_stats_kernel"));
      // (a separate tiny launch instead of a last-block-done counter: blocks then end without waiting on an atomic)
      instnorm_finalize_kernel<<<(unsigned)ceil_div((int64_t)nb * s.cout, 128), 128, 0, c.st>>>(
          c.ws.stats[io.out_stats], io.fin_gamma, io.fin_beta, nb, s.cout, io.fin_count, c.ws.in_scale[io.out_stats],
          c.ws.in_shift[io.out_stats]);
      return check_launch("instnorm_finalize_kernel");
    }
    if (fuse) {
      SPANON_CUDA_TRY(launch_pdl(instnorm_finalize_kernel, dim3((unsigned)ceil_div((int64_t)nb * s.cout, 128)), dim3(128), 0, c.st,
                                 id != L_DEC45, (const double*)c.ws.stats[io.out_stats], io.fin_gamma, io.fin_beta, nb, s.cout,
                                 io.fin_count, c.ws.in_scale[io.out_stats], c.ws.in_shift[io.out_stats]));
      return check_launch("instnorm_finalize_kernel");
    }
    return SPANON_OK;
  }
  if (c.tc) return launch_tapgemm_simt<__half>(p, c.st);
  return launch_tapgemm_simt<float>(p, c.st);
}

int finalize_in(const Ctx& c, int producer, int idx, const float* gamma, const float* beta, int nb, double count) {
  if (layer_on_tc(c, producer)) return SPANON_OK;  // done by the producing tcgen05 kernel
  const int C = c.ws.inC[idx];
  instnorm_finalize_kernel<<<(unsigned)ceil_div((int64_t)nb * C, 128), 128, 0, c.st>>>(
      c.ws.stats[idx], gamma, beta, nb, C, count, c.ws.in_scale[idx], c.ws.in_shift[idx]);
  return check_launch("instnorm_finalize_kernel");
}

}  // namespace
}  // namespace spanon

using namespace spanon;

static int resolve_opts(const SpanonConvAEOpts* o, int64_t L, bool* tc, int* chunk) {
  const int prec = o ? o->precision : SPANON_PREC_FP16;
  if (prec != SPANON_PREC_FP32 && prec != SPANON_PREC_FP16) return SPANON_ERR_UNSUPPORTED;
  *tc = prec == SPANON_PREC_FP16;
  int ch = o ? o->chunk_utts : 0;
  if (ch <= 0) ch = default_chunk(L, *tc ? 2 : 4);
  *chunk = ch;
  return SPANON_OK;
}

extern "C" size_t spanon_convae_workspace_bytes(int64_t B, int64_t T, int64_t F, const SpanonConvAEOpts* opts) {
  if (B <= 0 || T <= 0 || F <= 0) return 0;
  bool tc; int chunk;
  if (resolve_opts(opts, T * F, &tc, &chunk) != SPANON_OK) return 0;
  chunk = (int)std::min<int64_t>(chunk, B);
  Workspace ws;
  carve(ws, nullptr, chunk, T * F, tc, tc ? 2 : 4);
  return ws.total + 256;
}

extern "C" int spanon_convae_fwd(const float* feats, const SpanonConvAEWeights* w, const float* pool_noise,
                                 const float* norm_mean, const float* norm_std, const int32_t* utt_max, float top_db,
                                 int64_t T_valid, float* recon, float* logp, void* workspace, size_t workspace_bytes,
                                 int64_t B, int64_t T, int64_t F, const SpanonConvAEOpts* opts, void* stream) {
  if (!w) return SPANON_ERR_NULL_POINTER;
  if (B < 0 || T <= 0 || F <= 0) return SPANON_ERR_BAD_SHAPE;
  const int64_t L = T * F;
  // the reference fails in its final reshape unless two stride-2 stages divide L (ConvAutoEncoder.py:197-198)
  if (L % 4 != 0 || L > (int64_t)1 << 30) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!feats || !recon || !workspace) return SPANON_ERR_NULL_POINTER;
  if ((norm_mean == nullptr) != (norm_std == nullptr)) return SPANON_ERR_NULL_POINTER;
  const bool fused_norm = norm_mean != nullptr;
  if (!fused_norm) T_valid = T;
  if (T_valid < 0 || T_valid > T) return SPANON_ERR_BAD_SHAPE;
  const bool run_clf = logp != nullptr && !(opts && opts->skip_classifier);
  const bool recon_half = opts && opts->recon_fp16;
  if (recon_half && (!opts || opts->precision != SPANON_PREC_FP16)) return SPANON_ERR_UNSUPPORTED;   // the 16-bit mode's output
  if (run_clf && L / 4 < 16) return SPANON_ERR_BAD_SHAPE;  // TDNN (k5, k3 d2, k3 d3) needs >= 15 rows
  Ctx c{};
  int chunk;
  SPANON_TRY(resolve_opts(opts, L, &c.tc, &chunk));
  chunk = (int)std::min<int64_t>(chunk, B);
  if (workspace_bytes < spanon_convae_workspace_bytes(B, T, F, opts)) return SPANON_ERR_WORKSPACE;
  char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<uintptr_t>(workspace), 256));
  const size_t esize = c.tc ? 2 : 4;
  carve(c.ws, base, chunk, L, c.tc, esize);
  profile_entry();
  c.w = w; c.L = L; c.st = as_stream(stream);

  {
    int dev = 0;
    SPANON_CUDA_TRY(cudaGetDevice(&dev));
    SPANON_CUDA_TRY(cudaDeviceGetAttribute(&c.n_sm, cudaDevAttrMultiProcessorCount, dev));
  }
  // ---- per-call weight preparation (weights may have changed since the last call) ----
  static const bool enc5_pair_env = !(getenv("SPANON_ENC5_PAIR") && getenv("SPANON_ENC5_PAIR")[0] == '0');
  c.enc5_pair = c.tc && enc5_pair_env;
  // (opt-in: measured 225 - 235 us against 213 us for the 7-tap N=64 form -- 43 % more MMA work outweighs the wider tile)
  static const bool dec45_pair_env = getenv("SPANON_DEC45_PAIR") && getenv("SPANON_DEC45_PAIR")[0] == '1';
  c.dec45_pair = c.tc && dec45_pair_env && compose_dec45(c);
  c.prepared = opts && opts->weights_prepared;
  if (!c.prepared) {
  SPANON_TRY(pack_layer(c, L_ENC2, w->enc_w[1]));
  SPANON_TRY(pack_layer(c, c.enc5_pair ? L_ENC5P : L_ENC5, w->enc_w[2]));
  SPANON_TRY(pack_layer(c, L_ENC8, w->enc_w[3]));
  SPANON_TRY(pack_layer(c, L_ENC11, w->enc_w[4]));
  SPANON_TRY(pack_layer(c, L_DEC0, w->dec_w[0]));
  SPANON_TRY(pack_layer(c, L_DEC1, w->dec_w[1]));
  if (compose_dec45(c)) {
    if (c.dec45_pair)
      compose_dec45_pair_kernel<<<160, 256, 0, c.st>>>(w->dec_w[2], w->dec_w[3], (__half*)c.ws.packed[L_DEC45P], 64, 64, 32);
    else
      compose_dec45_kernel<<<112, 256, 0, c.st>>>(w->dec_w[2], w->dec_w[3], (__half*)c.ws.packed[L_DEC45], 64, 64, 32);
    SPANON_TRY(check_launch("compose_dec45_kernel"));
    compose_dec45_bias_kernel<<<1, 128, 0, c.st>>>(w->dec_b[2], w->dec_w[3], w->dec_b[3], c.ws.bias45, 64, 32,
                                                   c.dec45_pair ? 128 : 64);
    SPANON_TRY(check_launch("compose_dec45_bias_kernel"));
    compose_dec45_edge_kernel<<<(unsigned)ceil_div(EDGE45_FLOATS, 256), 256, 0, c.st>>>(w->dec_w[2], w->dec_b[2], w->dec_w[3],
                                                                                        c.ws.edge45, 64, 64, 32);
    SPANON_TRY(check_launch("compose_dec45_edge_kernel"));
  } else {
    SPANON_TRY(pack_layer(c, L_DEC4, w->dec_w[2]));
    SPANON_TRY(pack_layer(c, L_DEC5, w->dec_w[3]));
  }
  SPANON_TRY(pack_layer(c, L_DEC8, w->dec_w[4]));
  // The classifier's constants are packed whenever its parameters are given, whether or not THIS call runs it: a later
  // call with weights_prepared = 1 may ask for the log-probabilities (a pipe slot first used with logp_host == NULL).
  if (run_clf || (w->tdnn_w[0] && w->tdnn_w[1] && w->tdnn_w[2] && w->clf_norm.weight && w->tdnn_bn[0].weight &&
                  w->tdnn_bn[1].weight && w->tdnn_bn[2].weight)) {
    SPANON_TRY(pack_layer(c, L_TD0, w->tdnn_w[0]));
    SPANON_TRY(pack_layer(c, L_TD3, w->tdnn_w[1]));
    SPANON_TRY(pack_layer(c, L_TD6, w->tdnn_w[2]));
    SPANON_TRY(fill_i32(reinterpret_cast<int32_t*>(c.ws.ones), 128, 0x3F800000, c.st));   // 1.0f
    bn_fold_kernel<<<1, 128, 0, c.st>>>(w->clf_norm, 128, c.ws.bn_scale[0], c.ws.bn_shift[0]);
    SPANON_TRY(check_launch("bn_fold_kernel"));
    for (int i = 0; i < 3; ++i) {
      bn_fold_kernel<<<1, 128, 0, c.st>>>(w->tdnn_bn[i], 128, c.ws.bn_scale[1 + i], c.ws.bn_shift[1 + i]);
      SPANON_TRY(check_launch("bn_fold_kernel"));
    }
  }
  }
  const int64_t L2 = L / 2, L4 = L / 4;
  const size_t stats_bytes = c.ws.zero_bytes;
  SPANON_CUDA_TRY(cudaMemsetAsync(c.ws.counter, 0, 256, c.st));
  for (int64_t b0 = 0; b0 < B; b0 += chunk) {
    const int nb = (int)std::min<int64_t>(chunk, B - b0);
    SPANON_CUDA_TRY(cudaMemsetAsync(c.ws.stats[0], 0, stats_bytes, c.st));
    void** X = c.ws.act;
    // encoder.0 + GLU.  16-bit path: generated inside encoder.2's operand production (conv_tc.cu, XF_GEN) -- the
    // 32-channel plane (5.16 MB per utterance written and read back) never exists; $SPANON_FUSE_ENC0=0 keeps the
    // separate first_layer_mma_kernel launch (A/B switch).
    static const bool fuse_enc0_env = !(getenv("SPANON_FUSE_ENC0") && getenv("SPANON_FUSE_ENC0")[0] == '0');
    const bool fuse_enc0 = c.tc && fuse_enc0_env && L >= 512;
    TapGemmParams gen{};
    if (fuse_enc0) {
      const int64_t fstride = fused_norm ? T_valid * F : L;
      gen.gen_feats = feats + b0 * fstride; gen.gen_bstride = fstride; gen.gen_nvalid = T_valid * F; gen.gen_F = (int)F;
      gen.gen_mean = norm_mean; gen.gen_std = norm_std; gen.gen_umax = utt_max ? utt_max + b0 : nullptr;
      gen.gen_topdb = top_db; gen.gen_w = w->enc_w[0]; gen.gen_b = w->enc_b[0];
    } else {
      const int64_t fstride = fused_norm ? T_valid * F : L;
      const float* src = feats + b0 * fstride;
      dim3 grid((unsigned)ceil_div(L, FL_NT), (unsigned)nb);
      const int32_t* um = utt_max ? utt_max + b0 : nullptr;
      if (c.tc)
        first_layer_mma_kernel<true><<<dim3((unsigned)ceil_div(L, FL_NT * FL_CHUNKS), (unsigned)nb), FL_NT, 0, c.st>>>(src, fstride, T_valid * F, (int)L, (int)F, norm_mean, norm_std, um,
                                                         top_db, w->enc_w[0], w->enc_b[0], (__half*)X[0], L * 32);
      else
        first_layer_kernel<float, false><<<grid, FL_NT, 0, c.st>>>(src, fstride, T_valid * F, (int)L, (int)F,
                                                                   norm_mean, norm_std, um, top_db, w->enc_w[0],
                                                                   w->enc_b[0], (float*)X[0], L * 32);
      SPANON_TRY(check_launch("first_layer_kernel"));
    }
    // A bias in front of an InstanceNorm only moves the mean the norm removes again: the tcgen05 path drops it (nb_).
    // Its TDNN layers store the raw convolution; bias + ReLU + BatchNorm run in the CONSUMER's operand transform
    // (eight otherwise idle warps) instead of the four epilogue warps.
    const bool tcp = c.tc;
    auto nb_ = [&](const float* bias) { return tcp ? nullptr : bias; };
    //                         in    out   in_norm t2  bias          relu e_bn stats out_mode
    {
      LayerIO e2{X[0], X[1], -1, -1, nb_(w->enc_b[1]), 0, -1, 0, OUT_CHANNELS_LAST, w->enc_in_w[0], w->enc_in_b[0], (double)L2};
      if (fuse_enc0) e2.gen = &gen;
      if (c.enc5_pair) e2.out_fold = OUT_PAIR_FOLD;      // encoder.5 reads row pairs
      SPANON_TRY(run_layer(c, L_ENC2, e2, nb));
    }
    SPANON_TRY(finalize_in(c, L_ENC2, 0, w->enc_in_w[0], w->enc_in_b[0], nb, (double)L2));
    if (c.enc5_pair)
      SPANON_TRY(run_layer(c, L_ENC5P, {X[1], X[2], 0, -1, nullptr, 0, -1, 1, OUT_CHANNELS_LAST, w->enc_in_w[1], w->enc_in_b[1], (double)L2,
                                        OUT_DIRECT}, nb));
    else
    SPANON_TRY(run_layer(c, L_ENC5, {X[1], X[2], 0, -1, nb_(w->enc_b[2]), 0, -1, 1, OUT_CHANNELS_LAST, w->enc_in_w[1], w->enc_in_b[1], (double)L2,
                                     OUT_PAIR_FOLD}, nb));
    SPANON_TRY(finalize_in(c, L_ENC5, 1, w->enc_in_w[1], w->enc_in_b[1], nb, (double)L2));
    SPANON_TRY(run_layer(c, L_ENC8, {X[2], X[3], 1, -1, nb_(w->enc_b[3]), 0, -1, 2, OUT_CHANNELS_LAST, w->enc_in_w[2], w->enc_in_b[2], (double)L4}, nb));
    SPANON_TRY(finalize_in(c, L_ENC8, 2, w->enc_in_w[2], w->enc_in_b[2], nb, (double)L4));
    SPANON_TRY(run_layer(c, L_ENC11, {X[3], X[0], 2, -1, nb_(w->enc_b[4]), 0, -1, 3, OUT_CHANNELS_LAST, w->enc_in_w[3], w->enc_in_b[3], (double)L4}, nb));
    SPANON_TRY(finalize_in(c, L_ENC11, 3, w->enc_in_w[3], w->enc_in_b[3], nb, (double)L4));
    // X[0] = raw encoder.11 output; the bottleneck is SiLU(IN(X[0])), applied by both consumers on load
    if (run_clf) {
      const int64_t L3 = L4 - 14;
      bool pooled = false;
      if (c.tc) {
        LayerIO t0{X[0], X[1], 3, 0, nullptr, 0, -1, -1, OUT_CHANNELS_LAST};
        LayerIO t3{X[1], X[2], -1, 1, nullptr, 0, -1, -1, OUT_CHANNELS_LAST};
        t3.t1s = c.ws.ones; t3.t1h = w->tdnn_b[0]; t3.act = 2;     // ReLU(x + b0) -> BatchNorm 1
        LayerIO t6{X[2], X[3], -1, 2, nullptr, 0, -1, -1, OUT_CHANNELS_LAST};   // fp16 planes (X[3] is free here)
        t6.t1s = c.ws.ones; t6.t1h = w->tdnn_b[1]; t6.act = 2;     // ReLU(x + b1) -> BatchNorm 2
        // tdnn.0 -> tdnn.3 -> tdnn.6 as ONE dataflow launch (conv_tc.cu, launch_tapgemm_tc_chain): three groups of CTAs,
        // each with one layer's weights resident, the intermediate planes handed over tile by tile through L2 with
        // per-tile counters, so the two planes are read back from L2 while they are hot instead of from HBM after a
        // launch boundary.  $SPANON_TDNN_CHAIN=0: three launches; $SPANON_TDNN_SPLIT=a,b,c: CTAs per layer;
        // $SPANON_TDNN_GROUP: tiles per round-robin group.
        static const bool chain_env = getenv("SPANON_TDNN_CHAIN") && getenv("SPANON_TDNN_CHAIN")[0] == '1';
        if (chain_env && c.n_sm >= 12) {
          TapGemmParams ps[3] = {make_params(c, L_TD0, t0, nb), make_params(c, L_TD3, t3, nb), make_params(c, L_TD6, t6, nb)};
          const int tpu0 = (int)ceil_div(ps[0].rows_out, 128), tpu1 = (int)ceil_div(ps[1].rows_out, 128);
          unsigned* f0 = c.ws.flags;
          unsigned* f1 = c.ws.flags + (size_t)chunk * (size_t)ceil_div(L4, 128);
          ps[0].flag_out = f0;
          ps[1].flag_in = f0; ps[1].flag_in_tpu = tpu0; ps[1].flag_target = 4; ps[1].flag_out = f1;
          ps[2].flag_in = f1; ps[2].flag_in_tpu = tpu1; ps[2].flag_target = 4;
          for (int i = 0; i < 3; ++i) ps[i].err_flag = c.ws.counter + 1;
          ps[0].tag = "tapgemm:tdnn036";
          // CTAs per layer in proportion to the measured per-tile cost (5 taps : 3 taps : 3 taps of 128 x 128 MMAs)
          int ctas[3] = {(c.n_sm * 46 + 50) / 100, 0, 0};
          ctas[1] = (c.n_sm - ctas[0]) / 2; ctas[2] = c.n_sm - ctas[0] - ctas[1];
          if (const char* e = getenv("SPANON_TDNN_SPLIT")) {
            int a = 0, b = 0, d = 0;
            if (sscanf(e, "%d,%d,%d", &a, &b, &d) == 3 && a > 0 && b > 0 && d > 0 && a + b + d <= c.n_sm) { ctas[0] = a; ctas[1] = b; ctas[2] = d; }
          }
          int group = 8;
          if (const char* e = getenv("SPANON_TDNN_GROUP")) { const int g = atoi(e); if (g >= 1 && g <= 64) group = g; }
          SPANON_TRY(launch_tapgemm_tc_chain(ps, 3, ctas, group, c.st));
        } else {
        SPANON_TRY(run_layer(c, L_TD0, t0, nb));
        SPANON_TRY(run_layer(c, L_TD3, t3, nb));
        // tdnn.6 on the CTA-pair kernel keeps ReLU -> BatchNorm 3 -> reshape -> pooling sums in its epilogue (running sums per
        // row residue and channel in tensor memory, conv_tc2.cu X2_POOL): its output plane is never written and
        // pool_planar_kernel's pass over it disappears.  $SPANON_POOL_FUSE=0: store the plane, pool in a second pass.
        static const bool pool_fuse = !(getenv("SPANON_POOL_FUSE") && getenv("SPANON_POOL_FUSE")[0] == '0') &&
                                      !(getenv("SPANON_TC2") && getenv("SPANON_TC2")[0] == '0');
        if (pool_fuse) {
          LayerIO t6p = t6;
          t6p.bias = w->tdnn_b[2]; t6p.relu = 1; t6p.e_bn = 3;
          TapGemmParams pp = make_params(c, L_TD6, t6p, nb);
          pp.pool = c.ws.pool;
          pp.pdl = 1;
          if (tapgemm_tc2_supported(pp)) {
            SPANON_TRY(launch_tapgemm_tc2(pp, c.n_sm, c.st));
            pooled = true;
          }
        }
        if (!pooled) SPANON_TRY(run_layer(c, L_TD6, t6, nb));
        }
        // (pooling in tdnn.6's epilogue instead -- a lane rotation per column hands each value to the lane owning its bin
        // -- was measured at +140 us on that launch against 85 us for this pass: ~20 dependent instructions per column on
        // two warps per TMEM lane quarter run at ~3 cycles each)
        if (!pooled) {
        dim3 pg((unsigned)ceil_div(L3, PL_NT * PL_ROWS), 16u, (unsigned)nb);   // ReLU(x + b2) -> BatchNorm 3 -> pooling
        pool_planar_kernel<<<pg, PL_NT, 0, c.st>>>((const __half*)X[3], 16, (int)L3, w->tdnn_b[2], c.ws.bn_scale[3],
                                                   c.ws.bn_shift[3], c.ws.pool);
        SPANON_TRY(check_launch("pool_kernel"));
        }
      } else {
        SPANON_TRY(run_layer(c, L_TD0, {X[0], X[1], 3, 0, w->tdnn_b[0], 1, 1, -1, OUT_CHANNELS_LAST}, nb));
        SPANON_TRY(run_layer(c, L_TD3, {X[1], X[2], -1, -1, w->tdnn_b[1], 1, 2, -1, OUT_CHANNELS_LAST}, nb));
        SPANON_TRY(run_layer(c, L_TD6, {X[2], c.ws.tdnn_out, -1, -1, w->tdnn_b[2], 1, 3, -1, OUT_CHANNELS_FIRST_F32}, nb));
        const int64_t n_el = 128 * L3;
        dim3 pg((unsigned)std::min<int64_t>(ceil_div(n_el, 256 * 16), 256), (unsigned)nb);
        pool_kernel<<<pg, 256, 0, c.st>>>(c.ws.tdnn_out, n_el, n_el, c.ws.pool);
        SPANON_TRY(check_launch("pool_kernel"));
      }
      SPANON_CUDA_TRY(launch_pdl(clf_head_kernel, dim3((unsigned)nb), dim3(256), 0, c.st, true, (const double*)c.ws.pool,
                                 pool_noise ? pool_noise + b0 * 128 : (const float*)nullptr, L3, *w, logp + b0 * 2,
                                 (const unsigned*)(c.ws.counter + 1)));
      SPANON_TRY(check_launch("clf_head_kernel"));
    }
    SPANON_TRY(run_layer(c, L_DEC0, {X[0], X[1], 3, -1, w->dec_b[0], 0, -1, -1, OUT_CHANNELS_LAST}, nb));
    // (pair form of the composed layer: decoder.1's 128 columns [even | odd phase] ARE its input's pair view -- no unfold)
    SPANON_TRY(run_layer(c, L_DEC1, {X[1], X[2], -1, -1, nb_(w->dec_b[1]), 0, -1, 4, OUT_CHANNELS_LAST, w->dec_in_w[0], w->dec_in_b[0], (double)L2,
                                     c.dec45_pair ? OUT_DIRECT : OUT_PHASE_UNFOLD}, nb));
    SPANON_TRY(finalize_in(c, L_DEC1, 4, w->dec_in_w[0], w->dec_in_b[0], nb, (double)L2));
    if (c.dec45_pair) {       // ... as one 5-tap 128 -> 128 layer on row pairs (compose_dec45_pair_kernel)
      SPANON_TRY(run_layer(c, L_DEC45P, {X[2], X[3], 4, -1, c.ws.bias45, 0, -1, 5, OUT_CHANNELS_LAST, w->dec_in_w[1], w->dec_in_b[1], (double)L,
                                         OUT_PHASE_UNFOLD4}, nb));
    } else if (compose_dec45(c)) {   // decoder.4 o decoder.5 as one 7-tap phase-view layer (compose_dec45_kernel)
      SPANON_TRY(run_layer(c, L_DEC45, {X[2], X[3], 4, -1, c.ws.bias45, 0, -1, 5, OUT_CHANNELS_LAST, w->dec_in_w[1], w->dec_in_b[1], (double)L,
                                        OUT_PHASE_UNFOLD}, nb));
    } else {
      SPANON_TRY(run_layer(c, L_DEC4, {X[2], X[1], 4, -1, w->dec_b[2], 0, -1, -1, OUT_CHANNELS_LAST}, nb));
      SPANON_TRY(run_layer(c, L_DEC5, {X[1], X[3], -1, -1, nb_(w->dec_b[3]), 0, -1, 5, OUT_CHANNELS_LAST, w->dec_in_w[1], w->dec_in_b[1], (double)L,
                                       OUT_PHASE_UNFOLD}, nb));
      SPANON_TRY(finalize_in(c, L_DEC5, 5, w->dec_in_w[1], w->dec_in_b[1], nb, (double)L));
    }
    {
      LayerIO d8{X[3], recon + b0 * L, 5, -1, w->dec_b[4], 0, -1, -1, OUT_CH0_F32};
      if (recon_half) { d8.out = reinterpret_cast<__half*>(recon) + b0 * L; d8.out_half = 1; }
      SPANON_TRY(run_layer(c, L_DEC8, d8, nb));
    }
  }
  return SPANON_OK;
}

// ====================================================================================================================
// ConvReconstruction ("endtoend" anonymizer, models/EndToEnd.py:36-87): the same tap-GEMM kernels over a shorter layer
// list with an InstanceNorm behind EVERY convolution but the last:
//   encoder.0 Conv1d(1->32,k15,p7) -> IN(32) -> GLU -> encoder.3 Conv1d(32->64,k5,s2,p2) -> IN(64) -> GLU
//   -> encoder.6 Conv1d(64->64,k5,p2) -> IN(64) -> GLU -> encoder.9 ConvTranspose1d(64->32,k5,s2,p2,op1) -> IN(32) -> GLU
//   -> encoder.12 Conv1d(32->1,k15,p7)
// (the external x-vector sex classifier the reference then runs on the output is scope row 8(f)-2, a separate call).
extern "C" size_t spanon_convrecon_workspace_bytes(int64_t B, int64_t T, int64_t F, const SpanonConvAEOpts* opts) {
  if (B <= 0 || T <= 0 || F <= 0) return 0;
  bool tc; int chunk;
  if (resolve_opts(opts, T * F, &tc, &chunk) != SPANON_OK) return 0;
  chunk = (int)std::min<int64_t>(chunk, B);
  Workspace ws;
  carve(ws, nullptr, chunk, T * F, tc, tc ? 2 : 4, kInC_recon, false);
  return ws.total + 256;
}

extern "C" int spanon_convrecon_fwd(const float* feats, const SpanonConvReconWeights* w, const float* norm_mean,
                                    const float* norm_std, const int32_t* utt_max, float top_db, int64_t T_valid,
                                    float* recon, void* workspace, size_t workspace_bytes, int64_t B, int64_t T,
                                    int64_t F, const SpanonConvAEOpts* opts, void* stream) {
  if (!w) return SPANON_ERR_NULL_POINTER;
  if (B < 0 || T <= 0 || F <= 0) return SPANON_ERR_BAD_SHAPE;
  const int64_t L = T * F;
  // one stride-2 stage: the reference's final reshape fails for odd L (EndToEnd.py:78)
  if (L % 2 != 0 || L > (int64_t)1 << 30) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!feats || !recon || !workspace) return SPANON_ERR_NULL_POINTER;
  if ((norm_mean == nullptr) != (norm_std == nullptr)) return SPANON_ERR_NULL_POINTER;
  const bool fused_norm = norm_mean != nullptr;
  if (!fused_norm) T_valid = T;
  if (T_valid < 0 || T_valid > T) return SPANON_ERR_BAD_SHAPE;
  if (opts && opts->recon_fp16) return SPANON_ERR_UNSUPPORTED;   // fp16 output: ConvAutoencoder entry point only
  Ctx c{};
  int chunk;
  SPANON_TRY(resolve_opts(opts, L, &c.tc, &chunk));
  chunk = (int)std::min<int64_t>(chunk, B);
  if (workspace_bytes < spanon_convrecon_workspace_bytes(B, T, F, opts)) return SPANON_ERR_WORKSPACE;
  char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<uintptr_t>(workspace), 256));
  const size_t esize = c.tc ? 2 : 4;
  carve(c.ws, base, chunk, L, c.tc, esize, kInC_recon, false);
  profile_entry();
  c.w = nullptr; c.L = L; c.st = as_stream(stream);
  {
    int dev = 0;
    SPANON_CUDA_TRY(cudaGetDevice(&dev));
    SPANON_CUDA_TRY(cudaDeviceGetAttribute(&c.n_sm, cudaDevAttrMultiProcessorCount, dev));
  }
  c.prepared = opts && opts->weights_prepared;
  if (!c.prepared) {
    SPANON_TRY(pack_layer(c, L_ENC2, w->conv_w[1]));
    SPANON_TRY(pack_layer(c, L_ENC5, w->conv_w[2]));
    SPANON_TRY(pack_layer(c, L_DEC5, w->conv_w[3]));
    SPANON_TRY(pack_layer(c, L_DEC8, w->conv_w[4]));
  }
  const int64_t L2 = L / 2;
  const size_t stats_bytes = reinterpret_cast<char*>(c.ws.pool) - reinterpret_cast<char*>(c.ws.stats[0]);
  for (int64_t b0 = 0; b0 < B; b0 += chunk) {
    const int nb = (int)std::min<int64_t>(chunk, B - b0);
    SPANON_CUDA_TRY(cudaMemsetAsync(c.ws.stats[0], 0, stats_bytes, c.st));
    void** X = c.ws.act;
    {   // encoder.0, raw (the InstanceNorm + GLU behind it run in encoder.3's operand load)
      const int64_t fstride = fused_norm ? T_valid * F : L;
      const float* src = feats + b0 * fstride;
      dim3 grid((unsigned)ceil_div(L, FL_NT), (unsigned)nb);
      const int32_t* um = utt_max ? utt_max + b0 : nullptr;
      if (c.tc) {
        first_layer_mma_kernel<false><<<dim3((unsigned)ceil_div(L, FL_NT * FL_CHUNKS), (unsigned)nb), FL_NT, 0, c.st>>>(src, fstride, T_valid * F, (int)L, (int)F, norm_mean, norm_std,
                                                                um, top_db, w->conv_w[0], w->conv_b[0], (__half*)X[0], L * 32);
        SPANON_TRY(check_launch("first_layer_kernel"));
        dim3 sg((unsigned)ceil_div(L2, ST_NT * ST_ROWS), 8u, (unsigned)nb);    // pair-folded: 8 planes of L/2 rows
        instnorm_stats_kernel<<<sg, ST_NT, 0, c.st>>>((const __half*)X[0], 8, (int)L2, 32, c.ws.stats[0]);
        SPANON_TRY(check_launch("instnorm_stats_kernel"));
      } else {
        first_layer_kernel<float, false, false><<<grid, FL_NT, 0, c.st>>>(src, fstride, T_valid * F, (int)L, (int)F,
                                                                          norm_mean, norm_std, um, top_db, w->conv_w[0],
                                                                          w->conv_b[0], (float*)X[0], L * 32);
        SPANON_TRY(check_launch("first_layer_kernel"));
        dim3 sg((unsigned)std::min<int64_t>(ceil_div(L, 2048), 1024), (unsigned)nb);
        instnorm_stats_f32c32_kernel<<<sg, 256, 0, c.st>>>((const float*)X[0], (int)L, c.ws.stats[0]);
        SPANON_TRY(check_launch("instnorm_stats_kernel"));
      }
      instnorm_finalize_kernel<<<(unsigned)ceil_div((int64_t)nb * 32, 128), 128, 0, c.st>>>(
          c.ws.stats[0], w->in_w[0], w->in_b[0], nb, 32, (double)L, c.ws.in_scale[0], c.ws.in_shift[0]);
      SPANON_TRY(check_launch("instnorm_finalize_kernel"));
    }
    const bool tcp = c.tc;
    auto nb_ = [&](const float* bias) { return tcp ? nullptr : bias; };   // a bias in front of an InstanceNorm is a no-op
    SPANON_TRY(run_layer(c, L_ENC2, {X[0], X[1], 0, -1, nb_(w->conv_b[1]), 0, -1, 1, OUT_CHANNELS_LAST, w->in_w[1], w->in_b[1], (double)L2}, nb));
    SPANON_TRY(finalize_in(c, L_ENC2, 1, w->in_w[1], w->in_b[1], nb, (double)L2));
    SPANON_TRY(run_layer(c, L_ENC5, {X[1], X[2], 1, -1, nb_(w->conv_b[2]), 0, -1, 2, OUT_CHANNELS_LAST, w->in_w[2], w->in_b[2], (double)L2}, nb));
    SPANON_TRY(finalize_in(c, L_ENC5, 2, w->in_w[2], w->in_b[2], nb, (double)L2));
    SPANON_TRY(run_layer(c, L_DEC5, {X[2], X[3], 2, -1, nb_(w->conv_b[3]), 0, -1, 3, OUT_CHANNELS_LAST, w->in_w[3], w->in_b[3], (double)L,
                                     OUT_PHASE_UNFOLD}, nb));
    SPANON_TRY(finalize_in(c, L_DEC5, 3, w->in_w[3], w->in_b[3], nb, (double)L));
    SPANON_TRY(run_layer(c, L_DEC8, {X[3], recon + b0 * L, 3, -1, w->conv_b[4], 0, -1, -1, OUT_CH0_F32}, nb));
  }
  return SPANON_OK;
}
