// x-vector gender classifier on anonymized features for sm_100a.
//
// Reference behaviour reproduced:
//   Xvector.forward            speechbrain/speechbrain/lobes/models/Xvector.py:45-108   (evaluator_inference.yaml:34-41:
//                              TDNN 80 -> 512 -> 512 -> 512 -> 512 -> 1500, kernels 5,3,3,1,1, dilations 1,2,3,1,1, each
//                              Conv1d("same", reflect padding; nnet/CNN.py:349-352,400-403) -> LeakyReLU -> BatchNorm1d,
//                              then StatisticsPooling over the valid frames and Linear(3000 -> 128))
//   Classifier                 Xvector.py:111-172 (LeakyReLU -> BatchNorm -> Linear -> LeakyReLU -> BatchNorm -> Linear ->
//                              log-softmax; evaluator_inference.yaml:43-48)
//   EncoderClassifier.classify_batch_feats   speechbrain/speechbrain/pretrained/interfaces.py:825-829 (+ :762-776)
//   StatisticsPooling.forward  speechbrain/speechbrain/nnet/pooling.py:263-312 (lengths: round(len * T) valid frames)
//
// Layout: activations are fp16 chunk-planar per utterance, [utt][C/8][Rp][8], with H = 4 halo rows in front of row 0 and
// behind row T-1 that hold the REFLECTED rows (the "same" padding of the next conv), written by the producing epilogue, so
// a conv tap is a row shift of one shared-memory slab exactly as in conv_tc.cu.  Every TDNN layer is one launch of
// xv_gemm_kernel: tile = 128 frames x 256 output channels, K loop over 16-channel stages x taps, A slab (2 bulk copies)
// and the stage's weight slice [tap][2][256][8] (1 bulk copy, weights are streamed from L2, not resident: 512 x 512 x 3
// does not fit) into a ring of shared-memory stages -> tcgen05.mma M=128 N=256 K=16, fp32 accumulators double-buffered in
// TMEM (2 x 256 columns) -> epilogue: + bias, LeakyReLU, folded eval-mode BatchNorm, fp16 store (+ reflected copies).
// Roles: warps 0-3 epilogue, warp 4 MMA issuer, warp 5 loader.  No operand transform: activation + BatchNorm are applied
// by the producer's epilogue, which is exact here because nothing (no InstanceNorm) sits between them and the next conv.
#include "common.cuh"
#include "tc_prims.cuh"
#include <algorithm>

namespace spanon {
namespace {

using namespace tc;

constexpr int XV_H = 4;               // halo rows on each side of an utterance's planes (max reflect padding is 3)
constexpr int XV_BN = 256;            // output channels per tile = UMMA N
constexpr int XV_KS = 16;             // input channels per stage = UMMA K
constexpr int XV_THREADS = 6 * 32;
constexpr int XV_MAX_STAGES = 8;
constexpr size_t XV_SMEM_LIMIT = 232448;
constexpr float XV_LEAKY = 0.01f;     // torch.nn.LeakyReLU default negative_slope
constexpr int XV_NL = 5;
constexpr int XV_CIN[XV_NL] = {80, 512, 512, 512, 512};
constexpr int XV_COUT[XV_NL] = {512, 512, 512, 512, 1500};
constexpr int XV_K[XV_NL] = {5, 3, 3, 1, 1};
constexpr int XV_DIL[XV_NL] = {1, 2, 3, 1, 1};
constexpr int XV_EMB = 128, XV_POOLC = 1500;

inline int pad_to(int v, int m) { return (v + m - 1) / m * m; }

struct XvGemmParams {
  const __half* in; __half* out; const __half* w; const float* ec;   // ec: [3][cout_pad] bias | bn scale | bn shift
  int B, T, Rp;
  int cin, cout_pad;          // cin is a multiple of 16, cout_pad of 256
  int ntaps, dil, pad;
  int n_rt, n_nt, n_stages, slab_rows;   // slab_rows = 128 + 2 * pad
  uint32_t lbo_a, stage_bytes;
};

enum { XB_FULL = 0, XB_EMPTY = 8, XB_TFULL = 16, XB_TEMPTY = 18, XB_TPTR = 20 };

__global__ void __launch_bounds__(XV_THREADS, 1) xv_gemm_kernel(XvGemmParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t s_base = smem_u32(smem);
  auto BAR = [&](int i) { return s_base + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(smem + 8 * XB_TPTR);
  const uint32_t a0 = 256u;                                  // stages start behind the barrier block
  const int n_ks = p.cin / XV_KS;
  const int total = p.B * p.n_rt * p.n_nt;
  const int t_begin = (int)((int64_t)blockIdx.x * total / (int64_t)gridDim.x);
  const int n_my = (int)((int64_t)(blockIdx.x + 1) * total / (int64_t)gridDim.x) - t_begin;

  if (warp == 4) {
    if (lane == 0) {
      for (int i = 0; i < p.n_stages; ++i) { mbar_init(BAR(XB_FULL + i), 1); mbar_init(BAR(XB_EMPTY + i), 1); }
      for (int i = 0; i < 2; ++i) { mbar_init(BAR(XB_TFULL + i), 1); mbar_init(BAR(XB_TEMPTY + i), 4); }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_ptr_smem)),
                 "r"(2 * XV_BN)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const int planes_in = p.cin / 8, planes_out = p.cout_pad / 8;
  const uint32_t b_bytes = (uint32_t)p.ntaps * (2u * XV_BN * 16u);

  if (warp == 5) {
    // =========================== LOADER ===========================
    uint32_t stage = 0, ph = 0;
    const uint32_t a_bytes = (uint32_t)p.slab_rows * 16u;
    for (int k = 0; k < n_my; ++k) {
      const int t = t_begin + k;
      const int nt = t % p.n_nt, rt = (t / p.n_nt) % p.n_rt, b = t / (p.n_nt * p.n_rt);
      const __half* srcA = p.in + (((int64_t)b * planes_in) * p.Rp + (XV_H + rt * TC_M - p.pad)) * 8;
      const __half* srcB = p.w + (int64_t)nt * n_ks * p.ntaps * (2 * XV_BN * 8);
      for (int s = 0; s < n_ks; ++s) {
        mbar_wait(BAR(XB_EMPTY + stage), ph ^ 1u);
        mbar_expect_tx_elect(BAR(XB_FULL + stage), 2u * a_bytes + b_bytes);
        const uint32_t dst = s_base + a0 + stage * p.stage_bytes;
        bulk_g2s_elect(dst, srcA, a_bytes, BAR(XB_FULL + stage));
        bulk_g2s_elect(dst + p.lbo_a, srcA + (int64_t)p.Rp * 8, a_bytes, BAR(XB_FULL + stage));
        bulk_g2s_elect(dst + 2u * p.lbo_a, srcB, b_bytes, BAR(XB_FULL + stage));
        srcA += (int64_t)2 * p.Rp * 8;
        srcB += (int64_t)p.ntaps * (2 * XV_BN * 8);
        if (++stage == (uint32_t)p.n_stages) { stage = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 4) {
    // =========================== MMA ISSUER ===========================
    const uint32_t desc_hi = (uint32_t)(umma_desc(0, 0) >> 32);
    const uint32_t a_lo_const = (uint32_t)umma_desc(0, p.lbo_a);
    const uint32_t b_lo_const = (uint32_t)umma_desc(0, (uint32_t)XV_BN * 16u);
    const uint32_t idesc = umma_idesc(XV_BN);
    const uint32_t b_tap = 2u * XV_BN;                 // 16 B units per tap slice
    const uint32_t dil = (uint32_t)p.dil;
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    uint32_t stage = 0, ph = 0;
    for (int k = 0; k < n_my; ++k) {
      const uint32_t buf = (uint32_t)k & 1u;
      mbar_wait(BAR(XB_TEMPTY + buf), (((uint32_t)k >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_u + buf * (uint32_t)XV_BN;
      for (int s = 0; s < n_ks; ++s) {
        const uint32_t sa = s_base + a0 + stage * p.stage_bytes;
        uint32_t a_lo = a_lo_const + (sa >> 4);
        uint32_t b_lo = b_lo_const + ((sa + 2u * p.lbo_a) >> 4);
        mbar_wait(BAR(XB_FULL + stage), ph);
        tc_fence_after();
        umma_f16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo, ((uint64_t)desc_hi << 32) | b_lo, idesc, s != 0);
#pragma unroll 1
        for (int tap = 1; tap < p.ntaps; ++tap) {
          a_lo += dil;
          b_lo += b_tap;
          umma_f16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo, ((uint64_t)desc_hi << 32) | b_lo, idesc, 1u);
        }
        tc_commit_elect(BAR(XB_EMPTY + stage));
        if (++stage == (uint32_t)p.n_stages) { stage = 0; ph ^= 1u; }
      }
      tc_commit_elect(BAR(XB_TFULL + buf));
    }
  } else {
    // =========================== EPILOGUE ===========================
    const int mloc = warp * 32 + lane;
    for (int k = 0; k < n_my; ++k) {
      const int t = t_begin + k;
      const int nt = t % p.n_nt, rt = (t / p.n_nt) % p.n_rt, b = t / (p.n_nt * p.n_rt);
      const int m = rt * TC_M + mloc;
      const bool valid = m < p.T;
      // reflected copies for the consumer's "same" padding: row -j = row j, row T-1+j = row T-1-j  (j = 1..H)
      const int mir_lo = (valid && m >= 1 && m <= XV_H) ? XV_H - m : -1;
      const int mir_hi = (valid && m <= p.T - 2 && m >= p.T - 1 - XV_H) ? XV_H + 2 * (p.T - 1) - m : -1;
      const uint32_t buf = (uint32_t)k & 1u;
      mbar_wait(BAR(XB_TFULL + buf), ((uint32_t)k >> 1) & 1u);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * (uint32_t)XV_BN + ((uint32_t)(warp * 32) << 16);
      __half* ob = p.out + ((int64_t)b * planes_out + nt * (XV_BN / 8)) * p.Rp * 8;
      const float* ec = p.ec + nt * XV_BN;
#pragma unroll 1
      for (int cb = 0; cb < XV_BN; cb += 32) {
        float v[32];
        tmem_ld32(taddr + (uint32_t)cb, v);
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 bz = __ldg(reinterpret_cast<const float4*>(ec + cb + j));
          const float4 es = __ldg(reinterpret_cast<const float4*>(ec + p.cout_pad + cb + j));
          const float4 eh = __ldg(reinterpret_cast<const float4*>(ec + 2 * p.cout_pad + cb + j));
          float x;
          x = v[j] + bz.x;     x = x > 0.f ? x : XV_LEAKY * x; v[j] = fmaf(x, es.x, eh.x);
          x = v[j + 1] + bz.y; x = x > 0.f ? x : XV_LEAKY * x; v[j + 1] = fmaf(x, es.y, eh.y);
          x = v[j + 2] + bz.z; x = x > 0.f ? x : XV_LEAKY * x; v[j + 2] = fmaf(x, es.z, eh.z);
          x = v[j + 3] + bz.w; x = x > 0.f ? x : XV_LEAKY * x; v[j + 3] = fmaf(x, es.w, eh.w);
        }
        if (valid) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 o;
            __half2* oh = reinterpret_cast<__half2*>(&o);
#pragma unroll
            for (int j = 0; j < 4; ++j) oh[j] = __floats2half2_rn(v[8 * q + 2 * j], v[8 * q + 2 * j + 1]);
            __half* pl = ob + (int64_t)((cb >> 3) + q) * p.Rp * 8;
            *reinterpret_cast<uint4*>(pl + (int64_t)(XV_H + m) * 8) = o;
            if (mir_lo >= 0) *reinterpret_cast<uint4*>(pl + (int64_t)mir_lo * 8) = o;
            if (mir_hi >= 0) *reinterpret_cast<uint4*>(pl + (int64_t)mir_hi * 8) = o;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(BAR(XB_TEMPTY + buf));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * XV_BN) : "memory");
  }
}

// feats [B][T][80] fp32 -> planes [B][10][Rp][8] fp16 with reflected halo rows
__global__ void xv_input_kernel(const float* __restrict__ feats, __half* __restrict__ out, int T, int Rp, int F) {
  // grid = (row blocks, planes, B): consecutive lanes = consecutive rows of ONE plane, so a warp writes 512 contiguous
  // bytes (the first version walked (row, plane) pairs: 16-byte pieces scattered over ten planes, 0.2 ms for 29 MB)
  const int b = blockIdx.z, pl = blockIdx.y, planes = F / 8;
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= T + 2 * XV_H) return;
  int m = r - XV_H;
  if (m < 0) m = -m;
  if (m >= T) m = 2 * (T - 1) - m;
  uint4 o = make_uint4(0u, 0u, 0u, 0u);
  if (m >= 0 && m < T) {
    const float4* src = reinterpret_cast<const float4*>(feats + ((int64_t)b * T + m) * F + pl * 8);
    const float4 x0 = __ldg(src), x1 = __ldg(src + 1);
    __half2* oh = reinterpret_cast<__half2*>(&o);
    oh[0] = __floats2half2_rn(x0.x, x0.y); oh[1] = __floats2half2_rn(x0.z, x0.w);
    oh[2] = __floats2half2_rn(x1.x, x1.y); oh[3] = __floats2half2_rn(x1.z, x1.w);
  }
  *reinterpret_cast<uint4*>(out + (((int64_t)b * planes + pl) * Rp + r) * 8) = o;
}

// Conv1d weight [Cout][Cin][k] fp32 -> [n_tile][k_stage][tap][2 chunks][256 cols][8] fp16 (UMMA K-major B operand, LBO = 4 KB)
__global__ void xv_pack_kernel(const float* __restrict__ w, __half* __restrict__ dst, int cin, int cout, int k, int cin_pad,
                               int cout_pad) {
  const int n_ks = cin_pad / XV_KS;
  const int64_t n = (int64_t)(cout_pad / XV_BN) * n_ks * k * 2 * XV_BN * 8;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i;
    const int e = (int)(r % 8); r /= 8;
    const int col = (int)(r % XV_BN); r /= XV_BN;
    const int ch = (int)(r % 2); r /= 2;
    const int tap = (int)(r % k); r /= k;
    const int ks = (int)(r % n_ks); r /= n_ks;
    const int nt = (int)r;
    const int ci = ks * XV_KS + ch * 8 + e, co = nt * XV_BN + col;
    dst[i] = __float2half_rn((ci < cin && co < cout) ? __ldg(w + ((int64_t)co * cin + ci) * k + tap) : 0.f);
  }
}

// bias | folded eval-mode BatchNorm scale | shift, zero padded to cout_pad
__global__ void xv_epi_const_kernel(const float* __restrict__ bias, SpanonBatchNorm bn, int cout, int cout_pad,
                                    float* __restrict__ ec) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cout_pad) return;
  float b = 0.f, s = 0.f, h = 0.f;
  if (c < cout) {
    b = bias[c];
    s = bn.weight[c] * rsqrtf(bn.running_var[c] + 1e-5f);
    h = bn.bias[c] - bn.running_mean[c] * s;
  }
  ec[c] = b; ec[cout_pad + c] = s; ec[2 * cout_pad + c] = h;
}

// StatisticsPooling over the valid frames of planes [B][P][Rp][8]: pooled[b][c] = mean + noise, pooled[b][C + c] = std + eps
__global__ void __launch_bounds__(256)
xv_pool_kernel(const __half* __restrict__ x, const float* __restrict__ rel_len, const float* __restrict__ noise, int T,
               int Rp, int P, int C, float* __restrict__ pooled) {
  __shared__ double red[2][8][8];
  const int p = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int n = T;
  if (rel_len) n = (int)rintf(rel_len[b] * (float)T);     // int(torch.round(lengths * T)), pooling.py:281
  n = max(0, min(n, T));
  const uint4* base = reinterpret_cast<const uint4*>(x) + ((int64_t)b * P + p) * Rp + XV_H;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = 0.f; q[j] = 0.f; }
  for (int r = tid; r < n; r += 256) {
    const uint4 v = __ldg(base + r);
    const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = __half22float2(h[j]);
      s[2 * j] += f.x; s[2 * j + 1] += f.y;
      q[2 * j] = fmaf(f.x, f.x, q[2 * j]); q[2 * j + 1] = fmaf(f.y, f.y, q[2 * j + 1]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float a = warp_sum(s[j]), c = warp_sum(q[j]);        // (<= 4 rows per lane: fp32 is exact enough here;
    if (lane == 0) { red[0][warp][j] = (double)a; red[1][warp][j] = (double)c; }   //  the cross-warp sums are fp64)
  }
  __syncthreads();
  if (tid < 8) {
    const int c = p * 8 + tid;
    if (c < C) {
      double a = 0.0, qq = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) { a += red[0][w][tid]; qq += red[1][w][tid]; }
      const double mean = a / (double)n;
      double var = (qq - (double)n * mean * mean) / (double)(n - 1);   // unbiased, like torch.std (NaN for n = 1)
      if (var < 0.0) var = 0.0;
      pooled[(int64_t)b * 2 * C + c] = (float)mean + (noise ? noise[(int64_t)b * C + c] : 0.f);
      pooled[(int64_t)b * 2 * C + C + c] = (float)sqrt(var) + 1e-5f;
    }
  }
}

__device__ __forceinline__ float xv_leaky(float x) { return x > 0.f ? x : XV_LEAKY * x; }
__device__ __forceinline__ float xv_bn(float x, const SpanonBatchNorm& bn, int c) {
  return (x - bn.running_mean[c]) * rsqrtf(bn.running_var[c] + 1e-5f) * bn.weight[c] + bn.bias[c];
}

// Linear(3000 -> 128) embedding + the Classifier MLP + log-softmax: one block per utterance.  XV_HEAD_NT threads: sixteen
// warps share the 128 embedding rows (a warp per row, coalesced weight reads) -- with four warps the 1.5 MB weight matrix was
// streamed by 240 warps on the whole GPU, latency-bound at 0.19 ms.
constexpr int XV_HEAD_NT = 512;
__global__ void __launch_bounds__(XV_HEAD_NT)
xv_head_kernel(const float* __restrict__ pooled, SpanonXvectorWeights w, int n_classes, float* __restrict__ emb_out,
               float* __restrict__ logp) {
  __shared__ float sp[2 * XV_POOLC];
  __shared__ float e[XV_EMB], h[XV_EMB], o[8];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 2 * XV_POOLC; i += XV_HEAD_NT) sp[i] = pooled[(int64_t)b * 2 * XV_POOLC + i];
  __syncthreads();
  for (int j = warp; j < XV_EMB; j += XV_HEAD_NT / 32) {        // a warp per output: coalesced reads of the weight row
    const float* wr = w.emb_w + (int64_t)j * 2 * XV_POOLC;
    float a0 = 0.f, a1 = 0.f;
    int i = lane;
    for (; i + 32 < 2 * XV_POOLC; i += 64) { a0 = fmaf(__ldg(wr + i), sp[i], a0); a1 = fmaf(__ldg(wr + i + 32), sp[i + 32], a1); }
    if (i < 2 * XV_POOLC) a0 = fmaf(__ldg(wr + i), sp[i], a0);
    const float a = warp_sum(a0 + a1);
    if (lane == 0) e[j] = a + w.emb_b[j];
  }
  __syncthreads();
  if (emb_out && tid < XV_EMB) emb_out[(int64_t)b * XV_EMB + tid] = e[tid];
  if (!logp) return;                                              // Xvector.forward alone: the embedding is the result
  float x0 = 0.f;
  if (tid < XV_EMB) x0 = xv_bn(xv_leaky(e[tid]), w.clf_norm, tid);  // Classifier: act -> norm
  __syncthreads();
  if (tid < XV_EMB) e[tid] = x0;
  __syncthreads();
  for (int j = warp; j < XV_EMB; j += XV_HEAD_NT / 32) {        // DNN.block_0: linear -> act -> norm
    const float* wr = w.dnn_w + (int64_t)j * XV_EMB;
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < XV_EMB / 32; ++k) a = fmaf(__ldg(wr + lane + 32 * k), e[lane + 32 * k], a);
    a = warp_sum(a);
    if (lane == 0) h[j] = xv_bn(xv_leaky(a + w.dnn_b[j]), w.dnn_bn, j);
  }
  __syncthreads();
  if (warp == 0) {
    for (int c = 0; c < n_classes; ++c) {
      float a = 0.f;
      for (int i = lane; i < XV_EMB; i += 32) a = fmaf(__ldg(w.out_w + c * XV_EMB + i), h[i], a);
      a = warp_sum(a);
      if (lane == 0) o[c] = a + w.out_b[c];
    }
    __syncwarp();
    if (lane == 0) {
      float mx = o[0];
      for (int c = 1; c < n_classes; ++c) mx = fmaxf(mx, o[c]);
      float se = 0.f;
      for (int c = 0; c < n_classes; ++c) se += expf(o[c] - mx);
      const float lse = mx + logf(se);
      for (int c = 0; c < n_classes; ++c) logp[(int64_t)b * n_classes + c] = o[c] - lse;
    }
  }
}

inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }

struct XvWs {
  __half* packed[XV_NL]; float* ec[XV_NL];
  __half* in; __half* act[2]; __half* last; float* pooled;
  size_t total;
};

void xv_carve(XvWs& ws, char* base, int64_t B, int Rp) {
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += up256(bytes); return p; };
  for (int l = 0; l < XV_NL; ++l) {
    const int cin_pad = pad_to(XV_CIN[l], XV_KS), cout_pad = pad_to(XV_COUT[l], XV_BN);
    ws.packed[l] = (__half*)take((size_t)cin_pad * cout_pad * XV_K[l] * 2);
    ws.ec[l] = (float*)take((size_t)3 * cout_pad * 4);
  }
  ws.in = (__half*)take((size_t)B * (XV_CIN[0] / 8) * Rp * 16 + 4096);
  for (int i = 0; i < 2; ++i) ws.act[i] = (__half*)take((size_t)B * (512 / 8) * Rp * 16 + 4096);
  ws.last = (__half*)take((size_t)B * (pad_to(XV_POOLC, XV_BN) / 8) * Rp * 16 + 4096);
  ws.pooled = (float*)take((size_t)B * 2 * XV_POOLC * 4);
  ws.total = off;
}

int xv_rows(int64_t T) { return (int)(ceil_div(T, TC_M) * TC_M + 2 * XV_H); }

}  // namespace
}  // namespace spanon

using namespace spanon;

extern "C" size_t spanon_xvector_workspace_bytes(int64_t B, int64_t T) {
  if (B <= 0 || T <= 0 || T > (1 << 24)) return 0;
  XvWs ws;
  xv_carve(ws, nullptr, B, xv_rows(T));
  return ws.total + 256;
}

extern "C" int spanon_xvector_classify(const float* feats, const float* rel_len, const SpanonXvectorWeights* w,
                                       const float* pool_noise, float* emb_out, float* logp, void* workspace,
                                       size_t workspace_bytes, int64_t B, int64_t T, int64_t F, int32_t weights_prepared,
                                       void* stream) {
  if (!w) return SPANON_ERR_NULL_POINTER;
  if (B < 0 || T <= 0 || F != XV_CIN[0]) return SPANON_ERR_BAD_SHAPE;
  if (T < 4) return SPANON_ERR_BAD_SHAPE;     // reflect padding of 3 frames needs T > 3 (torch raises as well)
  if (B == 0) return SPANON_OK;
  if (!feats || (!logp && !emb_out) || !workspace) return SPANON_ERR_NULL_POINTER;
  if (B > 65535) return SPANON_ERR_BAD_SHAPE;
  if (workspace_bytes < spanon_xvector_workspace_bytes(B, T)) return SPANON_ERR_WORKSPACE;
  profile_entry();
  cudaStream_t st = as_stream(stream);
  const int Rp = xv_rows(T);
  XvWs ws;
  xv_carve(ws, reinterpret_cast<char*>(up256(reinterpret_cast<uintptr_t>(workspace))), B, Rp);
  int n_sm = 0;
  {
    int dev = 0;
    SPANON_CUDA_TRY(cudaGetDevice(&dev));
    SPANON_CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  }
  if (!weights_prepared) {
    for (int l = 0; l < XV_NL; ++l) {
      const int cin_pad = pad_to(XV_CIN[l], XV_KS), cout_pad = pad_to(XV_COUT[l], XV_BN);
      const int64_t n = (int64_t)cin_pad * cout_pad * XV_K[l];
      xv_pack_kernel<<<(unsigned)std::min<int64_t>(ceil_div(n, 256), 2048), 256, 0, st>>>(
          w->tdnn_w[l], ws.packed[l], XV_CIN[l], XV_COUT[l], XV_K[l], cin_pad, cout_pad);
      SPANON_TRY(check_launch("xv_pack_kernel"));
      xv_epi_const_kernel<<<(unsigned)ceil_div(cout_pad, 256), 256, 0, st>>>(w->tdnn_b[l], w->tdnn_bn[l], XV_COUT[l],
                                                                           cout_pad, ws.ec[l]);
      SPANON_TRY(check_launch("xv_epi_const_kernel"));
    }
  }
  {
    dim3 grid((unsigned)ceil_div(T + 2 * XV_H, 256), (unsigned)(F / 8), (unsigned)B);
    xv_input_kernel<<<grid, 256, 0, st>>>(feats, ws.in, (int)T, Rp, (int)F);
    SPANON_TRY(check_launch("xv_input_kernel"));
  }
  {
    static std::atomic<unsigned long long> attr_done{0};   // per device (one process may drive several GPUs)
    int dev = 0;
    SPANON_CUDA_TRY(cudaGetDevice(&dev));
    if (first_use_on_device(attr_done, dev))
      SPANON_CUDA_TRY(cudaFuncSetAttribute(xv_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)XV_SMEM_LIMIT));
  }
  const __half* cur = ws.in;
  for (int l = 0; l < XV_NL; ++l) {
    XvGemmParams p{};
    p.in = cur;
    p.out = l == XV_NL - 1 ? ws.last : ws.act[l & 1];
    p.w = ws.packed[l]; p.ec = ws.ec[l];
    p.B = (int)B; p.T = (int)T; p.Rp = Rp;
    p.cin = pad_to(XV_CIN[l], XV_KS); p.cout_pad = pad_to(XV_COUT[l], XV_BN);
    p.ntaps = XV_K[l]; p.dil = XV_DIL[l]; p.pad = XV_DIL[l] * (XV_K[l] - 1) / 2;
    p.n_rt = (int)ceil_div(T, TC_M); p.n_nt = p.cout_pad / XV_BN;
    p.slab_rows = TC_M + 2 * p.pad;
    p.lbo_a = (uint32_t)pad_to(p.slab_rows, 8) * 16u;
    p.stage_bytes = 2u * p.lbo_a + (uint32_t)p.ntaps * (2u * XV_BN * 16u);
    p.n_stages = (int)std::min<size_t>(XV_MAX_STAGES, (XV_SMEM_LIMIT - 256) / p.stage_bytes);
    if (p.n_stages < 2) return SPANON_ERR_UNSUPPORTED;
    const size_t smem = 256 + (size_t)p.n_stages * p.stage_bytes;
    const int total = p.B * p.n_rt * p.n_nt;
    xv_gemm_kernel<<<std::min(total, n_sm), XV_THREADS, smem, st>>>(p);
    SPANON_TRY(check_launch(l == 0 ? "xv_gemm:tdnn0" : l == 1 ? "xv_gemm:tdnn1" : l == 2 ? "xv_gemm:tdnn2"
                            : l == 3 ? "xv_gemm:tdnn3" : "xv_gemm:tdnn4"));
    cur = p.out;
  }
  {
    const int P = pad_to(XV_POOLC, XV_BN) / 8;
    dim3 grid((unsigned)ceil_div(XV_POOLC, 8), (unsigned)B);
    xv_pool_kernel<<<grid, 256, 0, st>>>(ws.last, rel_len, pool_noise, (int)T, Rp, P, XV_POOLC, ws.pooled);
    SPANON_TRY(check_launch("xv_pool_kernel"));
  }
  xv_head_kernel<<<(unsigned)B, XV_HEAD_NT, 0, st>>>(ws.pooled, *w, 2, emb_out, logp);
  return check_launch("xv_head_kernel");
}
