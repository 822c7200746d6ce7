// FullyConnectedAutoencoder forward, fused into one persistent kernel + a tiny head kernel.
//
// Reference: models/FullyConnected.py
//   FullyConnectedAutoencoder.forward :147-159  encoder 80-60-40-20 (ReLU), decoder 20-40-60-80
//   FullyConnSexClassifier.forward    :90-104   reshape-BN(20)-reshape, 20-40-40 (ReLU),
//                                               StatisticsPooling over T, MLP 80-40-40-20-2, log_softmax
//   StatisticsPooling                 speechbrain/speechbrain/nnet/pooling.py:263-312
//
// Kernel design: the six autoencoder layers and the two per-frame classifier layers run back to back on a 256-frame tile
// with activations in shared memory ([dim][frame]) and all 18.4 k weights resident in shared memory (transposed: [in][out]).
// fcae4_kernel (default): a thread owns 4 consecutive frames x a quarter of the layer's outputs as a register tile -- one
// 128-bit load of activations and one broadcast 128-bit load of weights per sixteen FMAs.  fcae_kernel (round 1,
// $SPANON_FCAE=1): one thread = one frame, one broadcast 128-bit weight load per four FMAs -- bound by the shared-memory
// pipe.  HBM traffic is the algorithmic minimum: 320 B in and 320 B out per frame; the pooled classifier statistics leave
// the kernel as fp64 atomics.  BatchNorm layers use running statistics (eval mode).
#include "common.cuh"
#include <algorithm>
#include <math.h>
#include <stdlib.h>

namespace spanon {
namespace {

constexpr int NTHR = 256;
constexpr int LDA = NTHR + 1;  // activation row stride (floats): conflict-free transposed access
constexpr int D_IN = 80;

struct FcLayer { int in, out, w_off, b_off; };
// smem weight offsets (floats), layers in execution order
constexpr int W_E0 = 0;                    // 80x60
constexpr int W_E1 = W_E0 + 80 * 60;       // 60x40
constexpr int W_E2 = W_E1 + 60 * 40;       // 40x20
constexpr int W_C0 = W_E2 + 40 * 20;       // 20x40 (classifier initial.0)
constexpr int W_C1 = W_C0 + 20 * 40;       // 40x40 (classifier initial.2)
constexpr int W_D0 = W_C1 + 40 * 40;       // 20x40
constexpr int W_D1 = W_D0 + 20 * 40;       // 40x60
constexpr int W_D2 = W_D1 + 40 * 60;       // 60x80
constexpr int W_END = W_D2 + 60 * 80;      // 18400
constexpr int B_E0 = 0, B_E1 = 60, B_E2 = 100, B_C0 = 120, B_C1 = 160, B_D0 = 200, B_D1 = 240, B_D2 = 300,
              B_END = 380;

struct FcSmem {
  float X[80 * LDA];
  float Y[60 * LDA];
  float W[W_END];
  float Bv[B_END];
  float bn_s[20], bn_t[20];
  float nmean[D_IN], nstd[D_IN];
};

__device__ __forceinline__ void load_wt(float* dst, const float* __restrict__ w, int out, int in, int tid) {
  // PyTorch Linear weight [out][in] -> smem [in][out]
#pragma unroll 4
  for (int e = tid; e < out * in; e += NTHR) {
    const int o = e / in, i = e - o * in;
    dst[i * out + o] = __ldg(w + e);
  }
}

template <int IN, int OUT, bool RELU>
__device__ __forceinline__ void dense(const float* __restrict__ xin, float* __restrict__ xout,
                                      const float* __restrict__ Wt, const float* __restrict__ bias, int tid) {
  constexpr int JB = 20;
  static_assert(OUT % JB == 0, "OUT must be a multiple of 20");
#pragma unroll 1
  for (int j0 = 0; j0 < OUT; j0 += JB) {
    float acc[JB];
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) acc[jj] = bias[j0 + jj];
#pragma unroll 4
    for (int i = 0; i < IN; ++i) {
      const float x = xin[i * LDA + tid];
      const float4* w4 = reinterpret_cast<const float4*>(Wt + i * OUT + j0);
#pragma unroll
      for (int q = 0; q < JB / 4; ++q) {
        const float4 w = w4[q];
        acc[4 * q + 0] = fmaf(x, w.x, acc[4 * q + 0]);
        acc[4 * q + 1] = fmaf(x, w.y, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(x, w.z, acc[4 * q + 2]);
        acc[4 * q + 3] = fmaf(x, w.w, acc[4 * q + 3]);
      }
    }
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) xout[(j0 + jj) * LDA + tid] = RELU ? fmaxf(acc[jj], 0.f) : acc[jj];
  }
}

// FullyConnSexClassifier up to the pooled sums (FullyConnected.py:93-99) for the 256 frames of one tile: bottleneck z in
// S.Y[0..20), scratch S.X; `frame`, `b`, `valid` describe this thread's frame.
__device__ __forceinline__ void fc_clf_tile(FcSmem& S, int tid, int64_t frame, int b, bool valid, int T, double* __restrict__ pool) {
    // FullyConnected.py:93-95: BatchNorm1d(20) on the [B,20,T] REINTERPRETATION of z
    const int t = valid ? (int)(frame - (int64_t)b * T) : 0;
#pragma unroll
    for (int d = 0; d < 20; ++d) {
      const int ch = (int)(((int64_t)t * 20 + d) / T);
      S.X[(60 + d) * LDA + tid] = fmaf(S.Y[d * LDA + tid], S.bn_s[ch], S.bn_t[ch]);
    }
    dense<20, 40, true>(S.X + 60 * LDA, S.X, S.W + W_C0, S.Bv + B_C0, tid);
    dense<40, 40, true>(S.X, S.X + 40 * LDA, S.W + W_C1, S.Bv + B_C1, tid);
    // StatisticsPooling sums over frames: warp reduce when the whole warp is in one utterance
    const int b0 = __shfl_sync(0xffffffffu, b, 0);
    const bool uniform = __all_sync(0xffffffffu, b == b0) && b0 >= 0;
#pragma unroll 1
    for (int c = 0; c < 40; ++c) {
      const float h = valid ? S.X[(40 + c) * LDA + tid] : 0.f;
      if (uniform) {
        const float s = warp_sum(h), q = warp_sum(h * h);
        if ((tid & 31) == 0) {
          atomicAdd(pool + ((int64_t)b0 * 40 + c) * 2 + 0, (double)s);
          atomicAdd(pool + ((int64_t)b0 * 40 + c) * 2 + 1, (double)q);
        }
      } else if (valid) {
        atomicAdd(pool + ((int64_t)b * 40 + c) * 2 + 0, (double)h);
        atomicAdd(pool + ((int64_t)b * 40 + c) * 2 + 1, (double)h * (double)h);
      }
    }
}

__global__ void __launch_bounds__(NTHR, 1)
fcae_kernel(const float* __restrict__ feats, SpanonFcaeWeights w, const float* __restrict__ norm_mean,
            const float* __restrict__ norm_std, const int32_t* __restrict__ utt_max, float top_db,
            float* __restrict__ recon, double* __restrict__ pool /*[B][40][2]*/, int64_t n_frames, int T,
            int run_clf) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FcSmem& S = *reinterpret_cast<FcSmem*>(smem_raw);
  const int tid = threadIdx.x;
  load_wt(S.W + W_E0, w.enc_w[0], 60, 80, tid);
  load_wt(S.W + W_E1, w.enc_w[1], 40, 60, tid);
  load_wt(S.W + W_E2, w.enc_w[2], 20, 40, tid);
  load_wt(S.W + W_C0, w.init_w[0], 40, 20, tid);
  load_wt(S.W + W_C1, w.init_w[1], 40, 40, tid);
  load_wt(S.W + W_D0, w.dec_w[0], 40, 20, tid);
  load_wt(S.W + W_D1, w.dec_w[1], 60, 40, tid);
  load_wt(S.W + W_D2, w.dec_w[2], 80, 60, tid);
  for (int i = tid; i < 60; i += NTHR) S.Bv[B_E0 + i] = w.enc_b[0][i];
  for (int i = tid; i < 40; i += NTHR) S.Bv[B_E1 + i] = w.enc_b[1][i];
  for (int i = tid; i < 20; i += NTHR) S.Bv[B_E2 + i] = w.enc_b[2][i];
  for (int i = tid; i < 40; i += NTHR) S.Bv[B_C0 + i] = w.init_b[0][i];
  for (int i = tid; i < 40; i += NTHR) S.Bv[B_C1 + i] = w.init_b[1][i];
  for (int i = tid; i < 40; i += NTHR) S.Bv[B_D0 + i] = w.dec_b[0][i];
  for (int i = tid; i < 60; i += NTHR) S.Bv[B_D1 + i] = w.dec_b[1][i];
  for (int i = tid; i < 80; i += NTHR) S.Bv[B_D2 + i] = w.dec_b[2][i];
  if (tid < 20) {
    const float s = w.clf_norm.weight[tid] / sqrtf(w.clf_norm.running_var[tid] + 1e-5f);
    S.bn_s[tid] = s;
    S.bn_t[tid] = w.clf_norm.bias[tid] - w.clf_norm.running_mean[tid] * s;
  }
  if (tid < D_IN) {
    S.nmean[tid] = norm_mean ? norm_mean[tid] : 0.f;
    S.nstd[tid] = norm_std ? norm_std[tid] : 1.f;
  }
  __syncthreads();

  const int64_t n_tiles = (n_frames + NTHR - 1) / NTHR;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t f0 = tile * NTHR;
    const int nf = (int)min((int64_t)NTHR, n_frames - f0);
    // coalesced load of nf*80 contiguous floats, transposed into X[d][frame]
    const float* src = feats + f0 * D_IN;
    for (int e = tid; e < NTHR * D_IN; e += NTHR) {
      const int fr = e / D_IN, d = e - fr * D_IN;
      float v = 0.f;
      if (fr < nf) {
        v = __ldg(src + e);
        if (norm_mean) {
          const int b = (int)((f0 + fr) / T);
          const float floor_db = utt_max ? key_to_float(utt_max[b]) - top_db : -INFINITY;
          v = (fmaxf(v, floor_db) - S.nmean[d]) / S.nstd[d];
        }
      }
      S.X[d * LDA + fr] = v;
    }
    __syncthreads();
    dense<80, 60, true>(S.X, S.Y, S.W + W_E0, S.Bv + B_E0, tid);
    dense<60, 40, true>(S.Y, S.X, S.W + W_E1, S.Bv + B_E1, tid);
    dense<40, 20, false>(S.X, S.Y, S.W + W_E2, S.Bv + B_E2, tid);  // bottleneck z in Y[0..20)
    const int64_t frame = f0 + tid;
    const bool valid = tid < nf;
    const int b = valid ? (int)(frame / T) : -1;
    if (run_clf) fc_clf_tile(S, tid, frame, b, valid, T, pool);
    dense<20, 40, true>(S.Y, S.X, S.W + W_D0, S.Bv + B_D0, tid);
    dense<40, 60, true>(S.X, S.Y, S.W + W_D1, S.Bv + B_D1, tid);
    dense<60, 80, false>(S.Y, S.X, S.W + W_D2, S.Bv + B_D2, tid);
    __syncthreads();
    float* dst = recon + f0 * D_IN;
    for (int e = tid; e < nf * D_IN; e += NTHR) {
      const int fr = e / D_IN, d = e - fr * D_IN;
      dst[e] = S.X[d * LDA + fr];
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// fcae4_kernel: the same forward with a REGISTER TILE of 4 frames x up to 20 outputs per thread.  fcae_kernel (one thread =
// one frame) needs one broadcast LDS.128 of weights per four FMAs and is bound by the shared-memory pipe (0.27 ms for
// 64 x 10 s = 3 % of the HBM roofline).  Here a thread owns four consecutive frames of the 256-frame tile and a quarter of
// the layer's outputs (warps 2g, 2g + 1 take output group g): per input channel one LDS.128 brings its four frames' values
// and one LDS.128 per output chunk brings four weights for sixteen FMAs -- 3.6 x fewer shared-memory instructions per FMA.
// Layers are separated by a CTA barrier (a thread no longer owns whole frames).
constexpr int LDA4 = 260;                   // activation row stride (floats): a multiple of 4 (aligned float4 over frames)
struct Fc4Smem {
  float X[80 * LDA4];
  float Y[60 * LDA4];
  float W[W_END];
  float Bv[B_END];
  float bn_s[20], bn_t[20];
  float nmean[D_IN], nstd[D_IN];
  int fb[NTHR], ft[NTHR];                    // utterance / frame-in-utterance of the tile's frames (-1 past the end)
};

// xin [IN][LDA4] -> xout [OUT][LDA4] for this thread's 4 frames (fg) and its output chunks (group og of 4)
template <int IN, int OUT, bool RELU, int NC>
__device__ __forceinline__ void dense4_body(const float* __restrict__ xp, float* __restrict__ op, const float* __restrict__ wp,
                                            const float* __restrict__ bp) {
  float acc[NC][4][4];                               // [chunk][output in chunk][frame]
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const float4 b4 = *reinterpret_cast<const float4*>(bp + 4 * c);
#pragma unroll
    for (int f = 0; f < 4; ++f) { acc[c][0][f] = b4.x; acc[c][1][f] = b4.y; acc[c][2][f] = b4.z; acc[c][3][f] = b4.w; }
  }
#pragma unroll 4
  for (int i = 0; i < IN; ++i) {
    const float4 x4 = *reinterpret_cast<const float4*>(xp + i * LDA4);
    const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const float4 w4 = *reinterpret_cast<const float4*>(wp + i * OUT + 4 * c);
      const float wv[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int f = 0; f < 4; ++f) acc[c][k][f] = fmaf(xv[f], wv[k], acc[c][k][f]);
    }
  }
#pragma unroll
  for (int c = 0; c < NC; ++c)
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float4 o = make_float4(acc[c][k][0], acc[c][k][1], acc[c][k][2], acc[c][k][3]);
      if (RELU) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
      *reinterpret_cast<float4*>(op + (4 * c + k) * LDA4) = o;
    }
}

// xin [IN][LDA4] -> xout [OUT][LDA4] for this thread's 4 frames (fg) and its output chunks (group og of 4).  The OUT / 4
// chunks of four outputs are dealt to the four groups as evenly as they go; a group's chunk count is uniform over its two
// warps, so the choice between the two instantiations is a warp-uniform branch outside the loops.
template <int IN, int OUT, bool RELU>
__device__ __forceinline__ void dense4(const float* __restrict__ xin, float* __restrict__ xout, const float* __restrict__ Wt,
                                       const float* __restrict__ bias, int fg, int og) {
  constexpr int NCH = OUT / 4;                       // float4 chunks of outputs
  constexpr int MAXC = (NCH + 3) / 4;                // chunks per thread: MAXC, or MAXC - 1 when NCH is not a multiple of 4
  static_assert(OUT % 4 == 0, "OUT must be a multiple of 4");
  const int cb = og * NCH / 4, nch = (og + 1) * NCH / 4 - cb;
  const float* xp = xin + 4 * fg;
  float* op = xout + 4 * cb * LDA4 + 4 * fg;
  if (nch == MAXC) dense4_body<IN, OUT, RELU, MAXC>(xp, op, Wt + 4 * cb, bias + 4 * cb);
  else if constexpr (MAXC > 1) dense4_body<IN, OUT, RELU, MAXC - 1>(xp, op, Wt + 4 * cb, bias + 4 * cb);
}

__global__ void __launch_bounds__(NTHR, 1)
fcae4_kernel(const float* __restrict__ feats, SpanonFcaeWeights w, const float* __restrict__ norm_mean,
             const float* __restrict__ norm_std, const int32_t* __restrict__ utt_max, float top_db,
             float* __restrict__ recon, double* __restrict__ pool /*[B][40][2]*/, int64_t n_frames, int T, int run_clf) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Fc4Smem& S = *reinterpret_cast<Fc4Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int og = warp >> 1, fg = (warp & 1) * 32 + lane;          // output group, frame group (frames 4 fg .. 4 fg + 3)
  load_wt(S.W + W_E0, w.enc_w[0], 60, 80, tid);
  load_wt(S.W + W_E1, w.enc_w[1], 40, 60, tid);
  load_wt(S.W + W_E2, w.enc_w[2], 20, 40, tid);
  load_wt(S.W + W_C0, w.init_w[0], 40, 20, tid);
  load_wt(S.W + W_C1, w.init_w[1], 40, 40, tid);
  load_wt(S.W + W_D0, w.dec_w[0], 40, 20, tid);
  load_wt(S.W + W_D1, w.dec_w[1], 60, 40, tid);
  load_wt(S.W + W_D2, w.dec_w[2], 80, 60, tid);
  for (int i = tid; i < 60; i += NTHR) S.Bv[B_E0 + i] = w.enc_b[0][i];
  for (int i = tid; i < 40; i += NTHR) S.Bv[B_E1 + i] = w.enc_b[1][i];
  for (int i = tid; i < 20; i += NTHR) S.Bv[B_E2 + i] = w.enc_b[2][i];
  for (int i = tid; i < 40; i += NTHR) S.Bv[B_C0 + i] = w.init_b[0][i];
  for (int i = tid; i < 40; i += NTHR) S.Bv[B_C1 + i] = w.init_b[1][i];
  for (int i = tid; i < 40; i += NTHR) S.Bv[B_D0 + i] = w.dec_b[0][i];
  for (int i = tid; i < 60; i += NTHR) S.Bv[B_D1 + i] = w.dec_b[1][i];
  for (int i = tid; i < 80; i += NTHR) S.Bv[B_D2 + i] = w.dec_b[2][i];
  if (tid < 20) {
    const float s = w.clf_norm.weight[tid] / sqrtf(w.clf_norm.running_var[tid] + 1e-5f);
    S.bn_s[tid] = s;
    S.bn_t[tid] = w.clf_norm.bias[tid] - w.clf_norm.running_mean[tid] * s;
  }
  if (tid < D_IN) {
    S.nmean[tid] = norm_mean ? norm_mean[tid] : 0.f;
    S.nstd[tid] = norm_std ? norm_std[tid] : 1.f;
  }
  __syncthreads();

  // every CTA takes one contiguous, equal share of the frames (tiles of up to 256 inside it): 251 whole tiles over 148 CTAs
  // would leave the second round 70 % full
  const int64_t per = (n_frames + gridDim.x - 1) / gridDim.x;
  const int64_t beg = blockIdx.x * per, end = min(n_frames, beg + per);
  for (int64_t f0 = beg; f0 < end; f0 += NTHR) {
    const int nf = (int)min((int64_t)NTHR, end - f0);
    {   // this thread's frame of the tile: utterance and position (one 64-bit division per frame and tile)
      const int64_t frame = f0 + tid;
      const int b = tid < nf ? (int)(frame / T) : -1;
      S.fb[tid] = b;
      S.ft[tid] = tid < nf ? (int)(frame - (int64_t)b * T) : 0;
    }
    // the tile's nf * 80 contiguous floats as 128-bit loads, all in flight before the first use, transposed into X[d][frame]
    const float4* src4 = reinterpret_cast<const float4*>(feats + f0 * D_IN);      // (f0 * 80 floats: 16-byte aligned)
    constexpr int NV = D_IN / 4;                       // 20 float4 per frame; 20 per thread as well
    float4 v4[NV];
#pragma unroll
    for (int r = 0; r < NV; ++r) {
      const int e4 = tid + r * NTHR;
      v4[r] = (e4 / NV) < nf ? __ldg(src4 + e4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();                                   // fb / ft visible (and the previous tile's readers are done)
#pragma unroll
    for (int r = 0; r < NV; ++r) {
      const int e4 = tid + r * NTHR;
      const int fr = e4 / NV, d = (e4 - fr * NV) * 4;
      float vv[4] = {v4[r].x, v4[r].y, v4[r].z, v4[r].w};
      if (norm_mean && fr < nf) {
        const float floor_db = utt_max ? key_to_float(utt_max[S.fb[fr]]) - top_db : -INFINITY;
#pragma unroll
        for (int q = 0; q < 4; ++q) vv[q] = (fmaxf(vv[q], floor_db) - S.nmean[d + q]) / S.nstd[d + q];
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) S.X[(d + q) * LDA4 + fr] = vv[q];
    }
    __syncthreads();
    dense4<80, 60, true>(S.X, S.Y, S.W + W_E0, S.Bv + B_E0, fg, og);
    __syncthreads();
    dense4<60, 40, true>(S.Y, S.X, S.W + W_E1, S.Bv + B_E1, fg, og);
    __syncthreads();
    dense4<40, 20, false>(S.X, S.Y, S.W + W_E2, S.Bv + B_E2, fg, og);  // bottleneck z in Y[0..20)
    __syncthreads();
    if (run_clf) {
      // FullyConnected.py:93-95: BatchNorm1d(20) on the [B,20,T] REINTERPRETATION of z -> X rows 60..79
      {
        const int t20 = S.ft[tid] * 20;                  // thread = frame: its 20 bottleneck values
        const bool valid = tid < nf;
#pragma unroll
        for (int d = 0; d < 20; ++d) {
          const int ch = (t20 + d) / T;
          S.X[(60 + d) * LDA4 + tid] = valid ? fmaf(S.Y[d * LDA4 + tid], S.bn_s[ch], S.bn_t[ch]) : 0.f;
        }
      }
      __syncthreads();
      dense4<20, 40, true>(S.X + 60 * LDA4, S.X, S.W + W_C0, S.Bv + B_C0, fg, og);
      __syncthreads();
      dense4<40, 40, true>(S.X, S.X + 40 * LDA4, S.W + W_C1, S.Bv + B_C1, fg, og);
      __syncthreads();
      // StatisticsPooling sums over frames: warp w takes channels 5w .. 5w + 4, lanes walk the tile's frames 32 at a time
      // (a group of 32 consecutive frames is reduced in the warp when it lies inside one utterance)
      const int b_first = S.fb[0], b_last = S.fb[nf - 1];
      if (b_first == b_last) {          // the whole tile lies in one utterance (the usual case: T >> 256): one reduction per channel
#pragma unroll 1
        for (int c = warp * 5; c < warp * 5 + 5; ++c) {
          const float* hrow = S.X + (40 + c) * LDA4;
          float s1 = 0.f, q1 = 0.f;
#pragma unroll
          for (int g = 0; g < NTHR; g += 32) {
            const float h = g + lane < nf ? hrow[g + lane] : 0.f;
            s1 += h; q1 = fmaf(h, h, q1);
          }
          s1 = warp_sum(s1); q1 = warp_sum(q1);
          if (lane == 0) {
            atomicAdd(pool + ((int64_t)b_first * 40 + c) * 2 + 0, (double)s1);
            atomicAdd(pool + ((int64_t)b_first * 40 + c) * 2 + 1, (double)q1);
          }
        }
      } else {
#pragma unroll 1
      for (int c = warp * 5; c < warp * 5 + 5; ++c) {
        const float* hrow = S.X + (40 + c) * LDA4;
#pragma unroll 1
        for (int g = 0; g < NTHR; g += 32) {
          const int fr = g + lane;
          const int b = S.fb[fr];
          const bool valid = b >= 0;
          const float h = valid ? hrow[fr] : 0.f;
          const int b0 = __shfl_sync(0xffffffffu, b, 0);
          const bool uniform = __all_sync(0xffffffffu, b == b0) && b0 >= 0;
          if (uniform) {
            const float s = warp_sum(h), q = warp_sum(h * h);
            if (lane == 0) {
              atomicAdd(pool + ((int64_t)b0 * 40 + c) * 2 + 0, (double)s);
              atomicAdd(pool + ((int64_t)b0 * 40 + c) * 2 + 1, (double)q);
            }
          } else if (valid) {
            atomicAdd(pool + ((int64_t)b * 40 + c) * 2 + 0, (double)h);
            atomicAdd(pool + ((int64_t)b * 40 + c) * 2 + 1, (double)h * (double)h);
          }
        }
      }
      }
      __syncthreads();
    }
    dense4<20, 40, true>(S.Y, S.X, S.W + W_D0, S.Bv + B_D0, fg, og);
    __syncthreads();
    dense4<40, 60, true>(S.X, S.Y, S.W + W_D1, S.Bv + B_D1, fg, og);
    __syncthreads();
    dense4<60, 80, false>(S.Y, S.X, S.W + W_D2, S.Bv + B_D2, fg, og);
    __syncthreads();
    float* dst = recon + f0 * D_IN;
    for (int e = tid; e < nf * D_IN; e += NTHR) {
      const int fr = e / D_IN, d = e - fr * D_IN;
      dst[e] = S.X[d * LDA4 + fr];
    }
    __syncthreads();
  }
}

// FullyConnSexClassifier.forward on its own (FullyConnected.py:88-104): the bottleneck z [B, T, 20] comes from the caller.
__global__ void __launch_bounds__(NTHR, 1)
fc_clf_kernel(const float* __restrict__ z, SpanonFcaeWeights w, double* __restrict__ pool, int64_t n_frames, int T) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FcSmem& S = *reinterpret_cast<FcSmem*>(smem_raw);
  const int tid = threadIdx.x;
  load_wt(S.W + W_C0, w.init_w[0], 40, 20, tid);
  load_wt(S.W + W_C1, w.init_w[1], 40, 40, tid);
  for (int i = tid; i < 40; i += NTHR) { S.Bv[B_C0 + i] = w.init_b[0][i]; S.Bv[B_C1 + i] = w.init_b[1][i]; }
  if (tid < 20) {
    const float s = w.clf_norm.weight[tid] / sqrtf(w.clf_norm.running_var[tid] + 1e-5f);
    S.bn_s[tid] = s;
    S.bn_t[tid] = w.clf_norm.bias[tid] - w.clf_norm.running_mean[tid] * s;
  }
  __syncthreads();
  const int64_t n_tiles = (n_frames + NTHR - 1) / NTHR;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t f0 = tile * NTHR;
    const int nf = (int)min((int64_t)NTHR, n_frames - f0);
    for (int e = tid; e < NTHR * 20; e += NTHR) {
      const int fr = e / 20, d = e - fr * 20;
      S.Y[d * LDA + fr] = fr < nf ? __ldg(z + f0 * 20 + e) : 0.f;
    }
    __syncthreads();
    const int64_t frame = f0 + tid;
    const bool valid = tid < nf;
    const int b = valid ? (int)(frame / T) : -1;
    fc_clf_tile(S, tid, frame, b, valid, T, pool);
    __syncthreads();
  }
}

__device__ __forceinline__ float bn_eval(float x, const SpanonBatchNorm& bn, int c) {
  return (x - bn.running_mean[c]) / sqrtf(bn.running_var[c] + 1e-5f) * bn.weight[c] + bn.bias[c];
}

// grid = B, block = 64
__global__ void fcae_head_kernel(const double* __restrict__ pool, const float* __restrict__ noise,
                                 SpanonFcaeWeights w, int T, float* __restrict__ logp) {
  __shared__ float s0[80], s1[40], s2[40], s3[20], s4[2];
  const int b = blockIdx.x, tid = threadIdx.x;
  if (tid < 40) {
    const double s = pool[((int64_t)b * 40 + tid) * 2 + 0], q = pool[((int64_t)b * 40 + tid) * 2 + 1];
    const double mean = s / (double)T;
    double var = T > 1 ? (q - s * s / (double)T) / (double)(T - 1) : (double)NAN;
    if (var < 0.0) var = 0.0;
    s0[tid] = (float)mean + (noise ? noise[(int64_t)b * 40 + tid] : 0.f);  // pooling.py:295-298
    s0[40 + tid] = (float)sqrt(var) + 1e-5f;                                  // pooling.py:299-300
  }
  __syncthreads();
  if (tid < 40) {
    float a = w.cls_b[0][tid];
    for (int i = 0; i < 80; ++i) a = fmaf(s0[i], w.cls_w[0][tid * 80 + i], a);
    s1[tid] = fmaxf(bn_eval(a, w.cls_bn1, tid), 0.f);
  }
  __syncthreads();
  if (tid < 40) {
    float a = w.cls_b[1][tid];
    for (int i = 0; i < 40; ++i) a = fmaf(s1[i], w.cls_w[1][tid * 40 + i], a);
    s2[tid] = fmaxf(a, 0.f);
  }
  __syncthreads();
  if (tid < 20) {
    float a = w.cls_b[2][tid];
    for (int i = 0; i < 40; ++i) a = fmaf(s2[i], w.cls_w[2][tid * 40 + i], a);
    s3[tid] = bn_eval(a, w.cls_bn6, tid);
  }
  __syncthreads();
  if (tid < 2) {
    float a = w.cls_b[3][tid];
    for (int i = 0; i < 20; ++i) a = fmaf(s3[i], w.cls_w[3][tid * 20 + i], a);
    s4[tid] = a;
  }
  __syncthreads();
  if (tid < 2) {
    const float m = fmaxf(s4[0], s4[1]);
    const float lse = m + logf(expf(s4[0] - m) + expf(s4[1] - m));
    logp[(int64_t)b * 2 + tid] = s4[tid] - lse;
  }
}

}  // namespace
}  // namespace spanon

using namespace spanon;

extern "C" size_t spanon_fcae_workspace_bytes(int64_t B, int64_t T) {
  (void)T;
  if (B < 0) return 0;
  return (size_t)B * 40 * 2 * sizeof(double) + 256;
}

extern "C" int spanon_fcae_fwd(const float* feats, const SpanonFcaeWeights* w, const float* pool_noise,
                               const float* norm_mean, const float* norm_std, const int32_t* utt_max, float top_db,
                               float* recon, float* logp, void* workspace, size_t workspace_bytes, int64_t B,
                               int64_t T, void* stream) {
  if (!w) return SPANON_ERR_NULL_POINTER;
  if (B < 0 || T < 0) return SPANON_ERR_BAD_SHAPE;
  if (w->in_dim != D_IN) return SPANON_ERR_UNSUPPORTED;
  if (B == 0 || T == 0) return SPANON_OK;
  if (!feats || !recon) return SPANON_ERR_NULL_POINTER;
  if ((norm_mean == nullptr) != (norm_std == nullptr)) return SPANON_ERR_NULL_POINTER;
  const int run_clf = logp != nullptr;
  if (run_clf && (!workspace || workspace_bytes < spanon_fcae_workspace_bytes(B, T))) return SPANON_ERR_WORKSPACE;
  if (run_clf && (reinterpret_cast<uintptr_t>(workspace) & 7)) return SPANON_ERR_MISALIGNED;
  profile_entry();
  cudaStream_t st = as_stream(stream);
  static std::atomic<unsigned long long> attr_done{0};     // per device (one process may drive several GPUs)
  int dev = 0;
  const int n_sm = current_device_sms(&dev);
  if (n_sm <= 0) return SPANON_ERR_CUDA;
  if (first_use_on_device(attr_done, dev))
  {
    SPANON_CUDA_TRY(cudaFuncSetAttribute(fcae_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(FcSmem)));
    SPANON_CUDA_TRY(cudaFuncSetAttribute(fcae4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(Fc4Smem)));
  }
  double* pool = reinterpret_cast<double*>(workspace);
  if (run_clf) SPANON_CUDA_TRY(cudaMemsetAsync(pool, 0, (size_t)B * 40 * 2 * sizeof(double), st));
  const int64_t n_frames = B * T;
  const int64_t n_tiles = ceil_div(n_frames, NTHR);
  const unsigned grid = (unsigned)std::min<int64_t>(n_tiles, n_sm);
  // $SPANON_FCAE=1: the one-frame-per-thread kernel of round 1 (A/B switch); default: the 4-frame register-tile kernel
  static const bool fcae_v1 = getenv("SPANON_FCAE") && getenv("SPANON_FCAE")[0] == '1';
  if (fcae_v1)
    fcae_kernel<<<grid, NTHR, sizeof(FcSmem), st>>>(feats, *w, norm_mean, norm_std, utt_max, top_db, recon, pool,
                                                    n_frames, (int)T, run_clf);
  else
    fcae4_kernel<<<grid, NTHR, sizeof(Fc4Smem), st>>>(feats, *w, norm_mean, norm_std, utt_max, top_db, recon, pool,
                                                      n_frames, (int)T, run_clf);
  SPANON_TRY(check_launch("fcae_kernel"));
  if (run_clf) {
    fcae_head_kernel<<<(unsigned)B, 64, 0, st>>>(pool, pool_noise, *w, (int)T, logp);
    SPANON_TRY(check_launch("fcae_head_kernel"));
  }
  return SPANON_OK;
}


// FullyConnSexClassifier.forward alone (models/FullyConnected.py:88-104, eval-mode BatchNorm): z [B, T, 20] -> log-probs [B, 2].
// Only the classifier fields of `w` are read (clf_norm, init_*, cls_*).  workspace: spanon_fcae_workspace_bytes(B, T).
extern "C" int spanon_fcae_classify(const float* z, const SpanonFcaeWeights* w, const float* pool_noise, float* logp,
                                    void* workspace, size_t workspace_bytes, int64_t B, int64_t T, void* stream) {
  if (!w) return SPANON_ERR_NULL_POINTER;
  if (B < 0 || T < 0) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (T == 0) return SPANON_ERR_BAD_SHAPE;
  if (!z || !logp || !workspace) return SPANON_ERR_NULL_POINTER;
  if (workspace_bytes < spanon_fcae_workspace_bytes(B, T)) return SPANON_ERR_WORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 7) return SPANON_ERR_MISALIGNED;
  profile_entry();
  cudaStream_t st = as_stream(stream);
  static std::atomic<unsigned long long> attr_done{0};
  int dev = 0;
  const int n_sm = current_device_sms(&dev);
  if (n_sm <= 0) return SPANON_ERR_CUDA;
  if (first_use_on_device(attr_done, dev))
    SPANON_CUDA_TRY(cudaFuncSetAttribute(fc_clf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FcSmem)));
  double* pool = reinterpret_cast<double*>(workspace);
  SPANON_CUDA_TRY(cudaMemsetAsync(pool, 0, (size_t)B * 40 * 2 * sizeof(double), st));
  const int64_t n_frames = B * T;
  const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(n_frames, NTHR), n_sm);
  fc_clf_kernel<<<grid, NTHR, sizeof(FcSmem), st>>>(z, *w, pool, n_frames, (int)T);
  SPANON_TRY(check_launch("fc_clf_kernel"));
  fcae_head_kernel<<<(unsigned)B, 64, 0, st>>>(pool, pool_noise, *w, (int)T, logp);
  return check_launch("fcae_head_kernel");
}


// ====================================================================================================================
// spanon_fcae_fwd_static -- the same forward with NO allocation anywhere on the call path (include/spanon.h).
// The pooling accumulators ([B][40][2] fp64, the only scratch memory of this path) live in a file-scope device array
// of a fixed, modest size instead of a caller-provided workspace.  Every GPU that loads the library holds the array
// until the process exits; all calls on one device share it (it keeps the last call's contents, so it is cleared on
// the stream before the kernel runs), hence two calls that overlap on one GPU race on it.
// In HOST code the array's name is only a stand-in for the symbol, not a device address: the address is taken once per
// device with cudaGetSymbolAddress while that device is current, and that pointer is what the kernels receive.
static __device__ __align__(256) double g_fcae_static_pool[(size_t)SPANON_FCAE_STATIC_MAX_B * 40 * 2];   // 160 KB

namespace spanon {
namespace {
std::atomic<double*> g_fcae_static_ptr[64];     // per device: address of g_fcae_static_pool on that device
std::atomic<unsigned long long> g_fcae_static_attr{0};

// launch bookkeeping without the profiler hook (check_launch() may create an event when a profile is being recorded)
inline int check_launch_static(const char* what) {
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_last_cuda_error(what, e); return SPANON_ERR_CUDA; }
  return SPANON_OK;
}
}  // namespace
}  // namespace spanon

extern "C" int64_t spanon_fcae_static_max_batch(void) { return SPANON_FCAE_STATIC_MAX_B; }

extern "C" int spanon_fcae_fwd_static(const float* feats, const SpanonFcaeWeights* w, const float* pool_noise,
                                      const float* norm_mean, const float* norm_std, const int32_t* utt_max,
                                      float top_db, float* recon, float* logp, int64_t B, int64_t T, void* stream) {
  if (!w) return SPANON_ERR_NULL_POINTER;
  if (B < 0 || T < 0) return SPANON_ERR_BAD_SHAPE;
  if (B > SPANON_FCAE_STATIC_MAX_B) return SPANON_ERR_STATIC_LIMIT;     // beyond the built-in scratch array
  if (w->in_dim != D_IN) return SPANON_ERR_UNSUPPORTED;
  if (B == 0 || T == 0) return SPANON_OK;
  if (!feats || !recon) return SPANON_ERR_NULL_POINTER;
  if ((norm_mean == nullptr) != (norm_std == nullptr)) return SPANON_ERR_NULL_POINTER;
  const int run_clf = logp != nullptr;
  profile_entry();
  cudaStream_t st = as_stream(stream);
  int dev = 0;
  const int n_sm = current_device_sms(&dev);
  if (n_sm <= 0 || dev < 0 || dev >= 64) return SPANON_ERR_CUDA;
  // set-up, once per device, with that device current
  double* pool = g_fcae_static_ptr[dev].load(std::memory_order_acquire);
  if (!pool) {
    void* sym = nullptr;
    SPANON_CUDA_TRY(cudaGetSymbolAddress(&sym, g_fcae_static_pool));
    if (!sym) return SPANON_ERR_CUDA;
    pool = static_cast<double*>(sym);
    g_fcae_static_ptr[dev].store(pool, std::memory_order_release);
  }
  if (first_use_on_device(g_fcae_static_attr, dev)) {
    SPANON_CUDA_TRY(cudaFuncSetAttribute(fcae_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FcSmem)));
    SPANON_CUDA_TRY(cudaFuncSetAttribute(fcae4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Fc4Smem)));
  }
  // the array holds whatever the previous call left: the accumulators of this batch must start at zero
  if (run_clf) SPANON_CUDA_TRY(cudaMemsetAsync(pool, 0, (size_t)B * 40 * 2 * sizeof(double), st));
  const int64_t n_frames = B * T;
  const int64_t n_tiles = ceil_div(n_frames, NTHR);
  const unsigned grid = (unsigned)std::min<int64_t>(n_tiles, n_sm);
  // $SPANON_FCAE=1: the one-frame-per-thread kernel of round 1 (A/B switch); default: the 4-frame register-tile kernel
  static const bool fcae_v1 = getenv("SPANON_FCAE") && getenv("SPANON_FCAE")[0] == '1';
  if (fcae_v1)
    fcae_kernel<<<grid, NTHR, sizeof(FcSmem), st>>>(feats, *w, norm_mean, norm_std, utt_max, top_db, recon, pool,
                                                    n_frames, (int)T, run_clf);
  else
    fcae4_kernel<<<grid, NTHR, sizeof(Fc4Smem), st>>>(feats, *w, norm_mean, norm_std, utt_max, top_db, recon, pool,
                                                      n_frames, (int)T, run_clf);
  SPANON_TRY(check_launch_static("fcae_kernel"));
  if (run_clf) {
    fcae_head_kernel<<<(unsigned)B, 64, 0, st>>>(pool, pool_noise, *w, (int)T, logp);
    SPANON_TRY(check_launch_static("fcae_head_kernel"));
  }
  return SPANON_OK;
}
