// Training-step kernels of the ConvAutoencoder path (reference: the autograd of
// models/ConvAutoEncoder.py:136-200 driven by speechbrain_convae_train.py:195-259 -- forward in TRAIN mode, loss.backward()).
//
// Everything is fp32 and channels-last, [utterance][row][channel]; the primitives mirror the forward lowering:
//   spanon_tapconv_fwd / _bwd     Conv1d (any dilation / padding), stride-2 Conv1d (pair view) and stride-2
//                                 ConvTranspose1d (phase view) as tap GEMMs (tapgemm.cuh):
//                                   forward   out = sum_taps in[m + shift] @ W[tap]            (conv_simt.cu kernel)
//                                   dgrad     gin = sum_taps gout[m - shift] @ W[tap]^T        (the SAME kernel: weights
//                                             packed transposed, shifts negated, input/output views swapped)
//                                   wgrad     gW[tap] = sum_{b,m} in[m + shift]^T gout[m]      (wgrad_kernel: K = B * rows)
//                                   bias grad column sums of gout
//   spanon_conv1_fwd / _bwd       Conv1d(1 -> 32, k15, p7) (encoder.0) and the C -> 1 channel output conv's gradients
//   spanon_norm_act_fwd / _bwd    InstanceNorm1d(affine) / train-mode BatchNorm1d (+ running statistics) + SiLU, from the
//                                 sums the producing conv's epilogue reduced; backward = two reductions + one pass
//   spanon_colstats               sum / sum of squares per (utterance, channel)
//   spanon_pool_reinterp_fwd/_bwd StatisticsPooling over the reshape(B, L, C) reinterpretation (ConvAutoEncoder.py:61)
// GradReverse is a sign flip of the gradient that enters the encoder from the classifier; the host does it when it adds the
// two branches' gradients.
#include "tapgemm.cuh"
#include "pack.cuh"
#include <algorithm>

namespace spanon {
namespace {

inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }

struct ConvGeom {
  int mode;                  // PackMode
  int cin_v, cout_v, ntaps;
  int shift[MAX_TAPS];
  int64_t rows_vin, rows_vout;   // view rows of the input / output
  int64_t rows_out;              // real output rows
  bool ok;
};

ConvGeom geom(const SpanonTapConv& cv, int64_t rows_in) {
  ConvGeom g{};
  g.ok = false;
  if (cv.cin <= 0 || cv.cout <= 0 || cv.k <= 0 || cv.k > MAX_TAPS || rows_in <= 0) return g;
  if (cv.kind == 0) {
    g.mode = PACK_CONV_S1;
    g.cin_v = cv.cin; g.cout_v = cv.cout == 1 ? 4 : cv.cout;
    g.ntaps = cv.k;
    for (int t = 0; t < cv.k; ++t) g.shift[t] = cv.dilation * t - cv.padding;
    g.rows_vin = rows_in;
    g.rows_out = rows_in + 2 * (int64_t)cv.padding - (int64_t)cv.dilation * (cv.k - 1);
    g.rows_vout = g.rows_out;
  } else if (cv.kind == 1) {          // Conv1d k5 s2 p2 on row pairs
    if (cv.k != 5 || cv.padding != 2 || cv.dilation != 1 || (rows_in & 1)) return g;
    g.mode = PACK_CONV_S2;
    g.cin_v = 2 * cv.cin; g.cout_v = cv.cout; g.ntaps = 3;
    g.shift[0] = -1; g.shift[1] = 0; g.shift[2] = 1;
    g.rows_vin = rows_in / 2; g.rows_out = rows_in / 2; g.rows_vout = g.rows_out;
  } else if (cv.kind == 2) {          // ConvTranspose1d k5 s2 p2 op1: the output as [even | odd] phases
    if (cv.k != 5 || cv.padding != 2 || cv.dilation != 1) return g;
    g.mode = PACK_CONVT_S2;
    g.cin_v = cv.cin; g.cout_v = 2 * cv.cout; g.ntaps = 3;
    g.shift[0] = -1; g.shift[1] = 0; g.shift[2] = 1;
    g.rows_vin = rows_in; g.rows_vout = rows_in; g.rows_out = 2 * rows_in;
  } else {
    return g;
  }
  if (g.rows_out <= 0 || g.cin_v % 32 != 0) return g;
  if (g.cout_v != 4 && g.cout_v != 64 && g.cout_v != 128) return g;
  if (g.rows_vin > (1 << 30) || g.rows_vout > (1 << 30)) return g;
  g.ok = true;
  return g;
}

// packed fp32 weights [ntaps][cin_v][cout_v]; transposed = 1: [ntaps][cout_v][cin_v] (the dgrad's weights)
__global__ void pack_train_kernel(const float* __restrict__ w, float* __restrict__ dst, int mode, int cin, int cout, int k,
                                  int ntaps, int cin_v, int cout_v, int transposed) {
  const int64_t n = (int64_t)ntaps * cin_v * cout_v;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int tap, civ, cov;
    if (!transposed) { cov = (int)(i % cout_v); civ = (int)((i / cout_v) % cin_v); }
    else { civ = (int)(i % cin_v); cov = (int)((i / cin_v) % cout_v); }
    tap = (int)(i / ((int64_t)cout_v * cin_v));
    dst[i] = packed_src(w, mode, cin, cout, k, tap, civ, cov);
  }
}

// gw[src] = gWp[tap][civ][cov] for the packed position of every weight element
__global__ void unpack_grad_kernel(const float* __restrict__ gwp, float* __restrict__ gw, int mode, int cin, int cout, int k,
                                   int ntaps, int cin_v, int cout_v) {
  const int64_t n = (int64_t)ntaps * cin_v * cout_v;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int cov = (int)(i % cout_v), civ = (int)((i / cout_v) % cin_v), tap = (int)(i / ((int64_t)cout_v * cin_v));
    const int64_t src = packed_src_index(mode, cin, cout, k, tap, civ, cov);
    if (src >= 0) gw[src] = gwp[i];
  }
}

// ------------------------------------------------------------------ wgrad
// gWp[tap][ci][co] += sum over this CTA's rows of  A[m + shift[tap]][ci] * G[m][co]   (views, fp32)
// grid = (row ranges, cin_v / 32, B), 256 threads; a thread owns CPT input channels x 4 output channels of every tap.
constexpr int WG_NT = 256, WG_SUB = 64, WG_KC = 32, WG_MAXTAPS = 5;
template <int COUT_V>
__global__ void __launch_bounds__(WG_NT)
wgrad_kernel(const float* __restrict__ a, const float* __restrict__ g, float* __restrict__ gwp, int rows_a, int rows_g,
             int cin_v, int ntaps, int s0, int s1, int s2, int s3, int s4, int rows_per_cta) {
  constexpr int CG = COUT_V / 4, RG = WG_NT / CG, CPT = WG_KC / RG;     // 128: 32 x 8, 4 ci per thread; 64: 16 x 16, 2
  extern __shared__ __align__(16) float sm[];
  const int shift[WG_MAXTAPS] = {s0, s1, s2, s3, s4};
  int lo = 0, hi = 0;
  for (int t = 0; t < ntaps; ++t) { lo = min(lo, shift[t]); hi = max(hi, shift[t]); }
  const int slab = WG_SUB - lo + hi;
  float* As = sm;                          // [slab][WG_KC + 4]
  float* Gs = sm + slab * (WG_KC + 4);     // [WG_SUB][COUT_V]
  const int tid = threadIdx.x, tx = tid % CG, ty = tid / CG;
  const int b = blockIdx.z, kc = blockIdx.y * WG_KC;
  const float* ab = a + (int64_t)b * rows_a * cin_v;
  const float* gb = g + (int64_t)b * rows_g * COUT_V;
  const int m_begin = blockIdx.x * rows_per_cta, m_end = min(rows_g, m_begin + rows_per_cta);
  float acc[WG_MAXTAPS][CPT][4];
#pragma unroll
  for (int t = 0; t < WG_MAXTAPS; ++t)
#pragma unroll
    for (int i = 0; i < CPT; ++i) { acc[t][i][0] = acc[t][i][1] = acc[t][i][2] = acc[t][i][3] = 0.f; }
  for (int m0 = m_begin; m0 < m_end; m0 += WG_SUB) {
    __syncthreads();
    for (int it = tid; it < slab * (WG_KC / 4); it += WG_NT) {
      const int r = it / (WG_KC / 4), c4 = (it - r * (WG_KC / 4)) * 4;
      const int row = m0 + lo + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row >= 0 && row < rows_a) v = __ldg(reinterpret_cast<const float4*>(ab + (int64_t)row * cin_v + kc + c4));
      *reinterpret_cast<float4*>(As + r * (WG_KC + 4) + c4) = v;
    }
    for (int it = tid; it < WG_SUB * CG; it += WG_NT) {
      const int r = it / CG, c4 = (it - r * CG) * 4;
      const int row = m0 + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row < m_end) v = __ldg(reinterpret_cast<const float4*>(gb + (int64_t)row * COUT_V + c4));
      *reinterpret_cast<float4*>(Gs + r * COUT_V + c4) = v;
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < WG_MAXTAPS; ++t) {
      if (t < ntaps) {
        const float* at = As + (shift[t] - lo) * (WG_KC + 4) + ty * CPT;
#pragma unroll 4
        for (int r = 0; r < WG_SUB; ++r) {
          const float4 gv = *reinterpret_cast<const float4*>(Gs + r * COUT_V + tx * 4);
          float avv[CPT];       // this thread's CPT consecutive input channels of row r: ONE vector load (rows are 144 B apart)
          if constexpr (CPT == 4) {
            const float4 a4 = *reinterpret_cast<const float4*>(at + r * (WG_KC + 4));
            avv[0] = a4.x; avv[1] = a4.y; avv[2] = a4.z; avv[3] = a4.w;
          } else if constexpr (CPT == 2) {
            const float2 a2 = *reinterpret_cast<const float2*>(at + r * (WG_KC + 4));
            avv[0] = a2.x; avv[1] = a2.y;
          } else {
#pragma unroll
            for (int i = 0; i < CPT; ++i) avv[i] = at[r * (WG_KC + 4) + i];
          }
#pragma unroll
          for (int i = 0; i < CPT; ++i) {
            const float av = avv[i];
            acc[t][i][0] = fmaf(av, gv.x, acc[t][i][0]); acc[t][i][1] = fmaf(av, gv.y, acc[t][i][1]);
            acc[t][i][2] = fmaf(av, gv.z, acc[t][i][2]); acc[t][i][3] = fmaf(av, gv.w, acc[t][i][3]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int t = 0; t < WG_MAXTAPS; ++t)
    if (t < ntaps)
#pragma unroll
      for (int i = 0; i < CPT; ++i) {
        float* dst = gwp + ((int64_t)t * cin_v + kc + ty * CPT + i) * COUT_V + tx * 4;
        atomicAdd(dst + 0, acc[t][i][0]); atomicAdd(dst + 1, acc[t][i][1]);
        atomicAdd(dst + 2, acc[t][i][2]); atomicAdd(dst + 3, acc[t][i][3]);
      }
}

// column sums of x [n_rows][C] into out[C % mod] (bias gradients; the phase view of a transposed conv folds 2*cout -> cout)
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ x, int64_t n_rows, int C, int mod, double* __restrict__ out) {
  // (fp64 partial sums: these are sums of ~1e5 signed values that cancel to a small remainder)
  __shared__ double red[256];
  const int tid = threadIdx.x;
  const int cpr = 256 / C > 0 ? 256 / C : 1;          // rows handled per pass (C <= 256)
  const int c = tid % C, rr = tid / C;
  double s = 0.0;
  if (rr < cpr)
    for (int64_t r = (int64_t)blockIdx.x * cpr + rr; r < n_rows; r += (int64_t)gridDim.x * cpr) s += (double)__ldg(x + r * C + c);
  red[tid] = rr < cpr ? s : 0.0;
  __syncthreads();
  if (tid < C) {
    double t = 0.0;
    for (int q = 0; q < cpr; ++q) t += red[q * C + tid];
    atomicAdd(out + tid % mod, t);
  }
}

// ------------------------------------------------------------------ 1 <-> 32 channel k15 convolutions
// out[b][l][c] = bias[c] + sum_t w[c * wc + t * wt] * x[b][l + t - 7]   (c < 32; flipped weights give the dgrad of a 32 -> 1 conv)
__global__ void __launch_bounds__(256)
conv1_k15_kernel(const float* __restrict__ x, const float* __restrict__ w, int wc, int wt, int w0, const float* __restrict__ bias,
                 float* __restrict__ out, int L) {
  __shared__ float sig[256 + 14];
  __shared__ __align__(16) float ws[15 * 32];
  __shared__ float bs[32];
  const int tid = threadIdx.x, b = blockIdx.y, l0 = blockIdx.x * 256;
  const float* xb = x + (int64_t)b * L;
  for (int i = tid; i < 256 + 14; i += 256) {
    const int l = l0 - 7 + i;
    sig[i] = (l >= 0 && l < L) ? __ldg(xb + l) : 0.f;
  }
  for (int i = tid; i < 15 * 32; i += 256) {
    const int t = i / 32, c = i - t * 32;
    ws[i] = __ldg(w + w0 + c * wc + t * wt);
  }
  if (tid < 32) bs[tid] = bias ? __ldg(bias + tid) : 0.f;
  __syncthreads();
  const int l = l0 + tid;
  if (l >= L) return;
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
      acc[4 * q + 0] = fmaf(xv, wv.x, acc[4 * q + 0]); acc[4 * q + 1] = fmaf(xv, wv.y, acc[4 * q + 1]);
      acc[4 * q + 2] = fmaf(xv, wv.z, acc[4 * q + 2]); acc[4 * q + 3] = fmaf(xv, wv.w, acc[4 * q + 3]);
    }
  }
  float4* o4 = reinterpret_cast<float4*>(out + ((int64_t)b * L + l) * 32);
#pragma unroll
  for (int q = 0; q < 8; ++q) o4[q] = make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
}

// corr[c][t] += sum_{b,l} s[b][l + sgn * (t - 7)] * v[b][l][c]   (c < 32, t < 15): the weight gradient of both thin convs
// (encoder.0: s = input signal, v = grad of its output, sgn = +1;  32 -> 1 output conv: s = grad of the output, v = its
// input, sgn = -1).  grid = (row chunks, B), 256 threads = 8 row slots x 32 channels; out index = c * oc + t * ot.
__global__ void __launch_bounds__(256)
corr15_kernel(const float* __restrict__ s, const float* __restrict__ v, int L, int sgn, float* __restrict__ out, int oc, int ot,
              int rows_per_cta) {
  __shared__ float sig[1024 + 14];
  __shared__ float red[8][32];
  const int tid = threadIdx.x, c = tid & 31, rr = tid >> 5, b = blockIdx.y;
  const int m_begin = blockIdx.x * rows_per_cta, m_end = min(L, m_begin + rows_per_cta);
  const float* sb = s + (int64_t)b * L;
  const float* vb = v + (int64_t)b * L * 32;
  float acc[15];
#pragma unroll
  for (int t = 0; t < 15; ++t) acc[t] = 0.f;
  for (int m0 = m_begin; m0 < m_end; m0 += 1024) {
    __syncthreads();
    for (int i = tid; i < 1024 + 14; i += 256) {
      const int l = m0 - 7 + i;
      sig[i] = (l >= 0 && l < L) ? __ldg(sb + l) : 0.f;
    }
    __syncthreads();
    const int n = min(1024, m_end - m0);
    for (int r = rr; r < n; r += 8) {
      const float vv = __ldg(vb + (int64_t)(m0 + r) * 32 + c);
#pragma unroll
      for (int t = 0; t < 15; ++t) acc[t] = fmaf(sig[r + 7 + sgn * (t - 7)], vv, acc[t]);
    }
  }
  for (int t = 0; t < 15; ++t) {
    __syncthreads();
    red[rr][c] = acc[t];
    __syncthreads();
    if (rr == 0) {
      float a = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) a += red[q][c];
      atomicAdd(out + c * oc + t * ot, a);
    }
  }
}

// ------------------------------------------------------------------ normalisation + activation
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// sums [B][C][2] (fp64) -> mean / rstd.  per_batch = 1: one (mean, rstd) per channel over all utterances (BatchNorm1d in
// training mode, biased variance for the normalisation, UNBIASED for the running estimate, torch semantics), also updates
// the running statistics;  per_batch = 0: per (utterance, channel) (InstanceNorm1d).
__global__ void norm_finalize_kernel(const double* __restrict__ stats, int B, int C, double rows, double count, float eps, int per_batch,
                                     float* __restrict__ mean_out, float* __restrict__ rstd_out, float* __restrict__ running_mean,
                                     float* __restrict__ running_var, float momentum) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (!per_batch) {
    if (i >= B * C) return;
    const double m = stats[2 * i] / rows;
    double var = stats[2 * i + 1] / rows - m * m;
    if (var < 0.0) var = 0.0;
    mean_out[i] = (float)m;
    rstd_out[i] = (float)(1.0 / sqrt(var + (double)eps));
  } else {
    if (i >= C) return;
    double s = 0.0, q = 0.0;
    for (int b = 0; b < B; ++b) { s += stats[2 * ((int64_t)b * C + i)]; q += stats[2 * ((int64_t)b * C + i) + 1]; }
    const double n = count > 0.0 ? count : rows * B, m = s / n;   // count: elements per channel behind the sums (SyncBN)
    double var = q / n - m * m;
    if (var < 0.0) var = 0.0;
    mean_out[i] = (float)m;
    rstd_out[i] = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean) {
      running_mean[i] = (1.f - momentum) * running_mean[i] + momentum * (float)m;
      running_var[i] = (1.f - momentum) * running_var[i] + momentum * (float)(n > 1.0 ? var * n / (n - 1.0) : var);
    }
  }
}

// out = act(gamma * (z - mean) * rstd + beta);  gamma == null: no normalisation (out = act(z));  act: 0 none, 1 SiLU
__global__ void __launch_bounds__(256)
norm_act_fwd_kernel(const float* __restrict__ z, const float* __restrict__ mean, const float* __restrict__ rstd, int stat_bstride,
                    const float* __restrict__ gamma, const float* __restrict__ beta, int act, float* __restrict__ out,
                    int64_t rows, int C, int64_t total4) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i * 4;
    const int c = (int)(e % C);
    const int64_t b = e / (rows * C);
    float4 v = __ldg(reinterpret_cast<const float4*>(z) + i);
    float* x = reinterpret_cast<float*>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float y = x[j];
      if (gamma) {
        const int64_t si = b * stat_bstride + c + j;
        y = (y - mean[si]) * rstd[si] * gamma[c + j] + beta[c + j];
      }
      if (act == 1) y = y * sigmoidf_(y);
      x[j] = y;
    }
    reinterpret_cast<float4*>(out)[i] = v;
  }
}

// dy = gout * act'(y) and the two reductions of the normalisation backward per (utterance, channel):
//   red[b][c][0] += sum dy,  red[b][c][1] += sum dy * xhat.     grid = (row chunks, B), block = 256 (C divides 256)
__global__ void __launch_bounds__(256)
norm_bwd_reduce_kernel(const float* __restrict__ z, const float* __restrict__ gout, const float* __restrict__ mean,
                       const float* __restrict__ rstd, int stat_bstride, const float* __restrict__ gamma,
                       const float* __restrict__ beta, int act, double* __restrict__ red, int64_t rows, int C, int rows_per_cta) {
  __shared__ double sh[2][256];
  const int tid = threadIdx.x, b = blockIdx.y;
  const int c = tid % C, rr = tid / C, rstep = 256 / C;
  const int64_t m_begin = (int64_t)blockIdx.x * rows_per_cta, m_end = min(rows, m_begin + rows_per_cta);
  const int64_t si = (int64_t)b * stat_bstride + c;
  const float mu = mean[si], rs = rstd[si], ga = gamma[c], be = beta[c];
  double s1 = 0.0, s2 = 0.0;
#pragma unroll 4
  for (int64_t r = m_begin + rr; r < m_end; r += rstep) {
    const int64_t idx = ((int64_t)b * rows + r) * C + c;
    const float xh = (__ldg(z + idx) - mu) * rs;
    float dy = __ldg(gout + idx);
    if (act == 1) {
      const float y = ga * xh + be, sg = sigmoidf_(y);
      dy *= sg * (1.f + y * (1.f - sg));
    }
    s1 += (double)dy; s2 += (double)dy * (double)xh;
  }
  sh[0][tid] = s1; sh[1][tid] = s2;
  __syncthreads();
  if (tid < C) {
    double a = 0.0, q = 0.0;
    for (int k = 0; k < rstep; ++k) { a += sh[0][k * C + tid]; q += sh[1][k * C + tid]; }
    atomicAdd(red + ((int64_t)b * C + tid) * 2 + 0, a);
    atomicAdd(red + ((int64_t)b * C + tid) * 2 + 1, q);
  }
}

// ggamma[c] = sum_b red[b][c][1], gbeta[c] = sum_b red[b][c][0]; per_batch: the sums over b also replace red[0][c][*]
__global__ void norm_bwd_params_kernel(double* __restrict__ red, int B, int C, int per_batch, float* __restrict__ ggamma,
                                       float* __restrict__ gbeta) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s1 = 0.0, s2 = 0.0;
  for (int b = 0; b < B; ++b) { s1 += red[((int64_t)b * C + c) * 2]; s2 += red[((int64_t)b * C + c) * 2 + 1]; }
  if (ggamma) ggamma[c] = (float)s2;
  if (gbeta) gbeta[c] = (float)s1;
  if (per_batch) { red[(int64_t)c * 2] = s1; red[(int64_t)c * 2 + 1] = s2; }
}

// gz = gamma * rstd * (dy - S1/n - xhat * S2/n), n = rows (InstanceNorm) or B * rows (BatchNorm);  gamma == null: gz = dy.
// relu_mask: z is the stored ReLU output and the gradient goes to the conv in front of the ReLU: zero where z <= 0.
__global__ void __launch_bounds__(256)
norm_bwd_apply_kernel(const float* __restrict__ z, const float* __restrict__ gout, const float* __restrict__ mean,
                      const float* __restrict__ rstd, int stat_bstride, const float* __restrict__ gamma,
                      const float* __restrict__ beta, int act, const double* __restrict__ red, int red_bstride, double n,
                      int relu_mask, float* __restrict__ gz, int64_t rows, int C, int64_t total4) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i * 4;
    const int c = (int)(e % C);
    const int64_t b = e / (rows * C);
    const float4 zv = __ldg(reinterpret_cast<const float4*>(z) + i);
    float4 gv = __ldg(reinterpret_cast<const float4*>(gout) + i);
    const float* zz = reinterpret_cast<const float*>(&zv);
    float* gg = reinterpret_cast<float*>(&gv);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float dy = gg[j];
      if (gamma) {
        const int64_t si = b * stat_bstride + c + j;
        const float xh = (zz[j] - mean[si]) * rstd[si];
        if (act == 1) {
          const float y = gamma[c + j] * xh + beta[c + j], sg = sigmoidf_(y);
          dy *= sg * (1.f + y * (1.f - sg));
        }
        const int64_t ri = (b * red_bstride + c + j) * 2;
        // (in double: the two means are subtracted from every element, and sums of gz over ~1e5 rows -- bias gradients --
        // see their rounding multiplied by the row count)
        dy = (float)((double)(gamma[c + j] * rstd[si]) * ((double)dy - red[ri] / n - (double)xh * (red[ri + 1] / n)));
      } else if (act == 1) {
        const float y = zz[j], sg = sigmoidf_(y);
        dy *= sg * (1.f + y * (1.f - sg));
      }
      if (relu_mask && !(zz[j] > 0.f)) dy = 0.f;
      gg[j] = dy;
    }
    reinterpret_cast<float4*>(gz)[i] = gv;
  }
}

// sums per (utterance, channel) of a channels-last tensor.  grid = (row chunks, B), block 256 (C divides 256)
__global__ void __launch_bounds__(256)
colstats_kernel(const float* __restrict__ x, double* __restrict__ stats, int64_t rows, int C, int rows_per_cta) {
  __shared__ double sh[2][256];
  const int tid = threadIdx.x, b = blockIdx.y, c = tid % C, rr = tid / C, rstep = 256 / C;
  const int64_t m_begin = (int64_t)blockIdx.x * rows_per_cta, m_end = min(rows, m_begin + rows_per_cta);
  double s = 0.0, q = 0.0;
#pragma unroll 4
  for (int64_t r = m_begin + rr; r < m_end; r += rstep) {
    const double v = (double)__ldg(x + ((int64_t)b * rows + r) * C + c);
    s += v; q += v * v;
  }
  sh[0][tid] = s; sh[1][tid] = q;
  __syncthreads();
  if (tid < C) {
    double a = 0.0, d = 0.0;
    for (int k = 0; k < rstep; ++k) { a += sh[0][k * C + tid]; d += sh[1][k * C + tid]; }
    atomicAdd(stats + ((int64_t)b * C + tid) * 2 + 0, a);
    atomicAdd(stats + ((int64_t)b * C + tid) * 2 + 1, d);
  }
}

// ------------------------------------------------------------------ StatisticsPooling over the reshape reinterpretation
// ConvAutoEncoder.py:61 reinterprets the channels-FIRST block [C][R] of an utterance as [R][C]: element (row r, channel c)
// has flat index c*R + r and is pooled into column (c*R + r) % C.  x is channels-last [B][R][C], C = 128.
// Threads = rows (r % 128 == tid % 128 because every chunk starts at a multiple of 128): for a fixed channel the map
// row -> column is a bijection, so a warp's shared-memory atomics never collide.
__global__ void __launch_bounds__(128)
pool_reinterp_sums_kernel(const float* __restrict__ x, double* __restrict__ pool, int R, int rows_per_cta) {
  constexpr int C = 128;
  __shared__ float bins[2][C];
  const int tid = threadIdx.x, b = blockIdx.y;
  bins[0][tid] = 0.f; bins[1][tid] = 0.f;
  __syncthreads();
  const int Rm = R % C;
  const int m_begin = blockIdx.x * rows_per_cta, m_end = min(R, m_begin + rows_per_cta);   // rows_per_cta % 128 == 0
  for (int r = m_begin + tid; r < m_end; r += C) {
    const float4* row = reinterpret_cast<const float4*>(x + ((int64_t)b * R + r) * C);
    for (int c4 = 0; c4 < C / 4; ++c4) {
      const float4 v = __ldg(row + c4);
      const float* e = reinterpret_cast<const float*>(&v);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int bin = ((4 * c4 + j) * Rm + r) & (C - 1);
        atomicAdd(&bins[0][bin], e[j]);
        atomicAdd(&bins[1][bin], e[j] * e[j]);
      }
    }
  }
  __syncthreads();
  atomicAdd(pool + ((int64_t)b * C + tid) * 2 + 0, (double)bins[0][tid]);
  atomicAdd(pool + ((int64_t)b * C + tid) * 2 + 1, (double)bins[1][tid]);
}

// out[b] = [mean + noise | std + 1e-5] (unbiased std over the R elements of a column), save[b] = [mean | std]
__global__ void pool_reinterp_finalize_kernel(const double* __restrict__ pool, const float* __restrict__ noise, int R,
                                              float* __restrict__ out, float* __restrict__ save) {
  constexpr int C = 128;
  const int b = blockIdx.x, j = threadIdx.x;
  const double s = pool[((int64_t)b * C + j) * 2], q = pool[((int64_t)b * C + j) * 2 + 1];
  const double mean = s / R;
  double var = R > 1 ? (q - s * s / R) / (R - 1) : (double)NAN;
  if (var < 0.0) var = 0.0;
  const float sd = (float)sqrt(var);
  out[(int64_t)b * 2 * C + j] = (float)mean + (noise ? noise[(int64_t)b * C + j] : 0.f);
  out[(int64_t)b * 2 * C + C + j] = sd + 1e-5f;
  save[(int64_t)b * 2 * C + j] = (float)mean;
  save[(int64_t)b * 2 * C + C + j] = sd;
}

// gx[b][r][c] = gmean[col] / R + gstd[col] * (x - mean[col]) / ((R - 1) * std[col]),  col = (c*R + r) % C
__global__ void __launch_bounds__(256)
pool_reinterp_bwd_kernel(const float* __restrict__ x, const float* __restrict__ save, const float* __restrict__ gout,
                         float* __restrict__ gx, int R, int64_t total) {
  constexpr int C = 128;
  const int Rm = R % C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int64_t br = i / C;
    const int r = (int)(br % R);
    const int64_t b = br / R;
    const int col = (c * Rm + r) & (C - 1);
    const float mean = save[b * 2 * C + col], sd = save[b * 2 * C + C + col];
    const float gm = gout[b * 2 * C + col], gs = gout[b * 2 * C + C + col];
    gx[i] = gm / (float)R + gs * (__ldg(x + i) - mean) / ((float)(R - 1) * sd);
  }
}

__global__ void d2f_kernel(const double* __restrict__ src, float* __restrict__ dst, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = (float)src[i];
}

// gb[C % mod] = column sums of x, accumulated in fp64 (scratch: >= 256 doubles)
int bias_grad(const float* x, int64_t n_rows, int C, int mod, float* gb, double* scratch, cudaStream_t st) {
  SPANON_CUDA_TRY(cudaMemsetAsync(scratch, 0, (size_t)mod * sizeof(double), st));
  colsum_kernel<<<256, 256, 0, st>>>(x, n_rows, C, mod, scratch);
  SPANON_TRY(check_launch("train:bias_grad"));
  d2f_kernel<<<(unsigned)ceil_div(mod, 128), 128, 0, st>>>(scratch, gb, mod);
  return check_launch("train:bias_grad");
}

int pack_weights(const SpanonTapConv& cv, const ConvGeom& g, const float* w, float* dst, int transposed, cudaStream_t st) {
  const int64_t n = (int64_t)g.ntaps * g.cin_v * g.cout_v;
  pack_train_kernel<<<(unsigned)std::min<int64_t>(ceil_div(n, 256), 1024), 256, 0, st>>>(w, dst, g.mode, cv.cin, cv.cout, cv.k,
                                                                                        g.ntaps, g.cin_v, g.cout_v, transposed);
  return check_launch("pack_train_kernel");
}

}  // namespace
}  // namespace spanon

using namespace spanon;

extern "C" size_t spanon_tapconv_workspace_bytes(const SpanonTapConv* cv) {
  if (!cv) return 0;
  const ConvGeom g = geom(*cv, 1024);
  if (!g.ok) return 0;
  // packed weights (forward or transposed) + packed weight gradient + the fp64 accumulators of the bias gradient
  return 2 * up256((size_t)g.ntaps * g.cin_v * g.cout_v * 4) + up256(256 * sizeof(double)) + 256;
}

extern "C" int spanon_tapconv_fwd(const float* in, const float* w, const float* bias, float* out, double* stats, int relu,
                                  void* workspace, size_t workspace_bytes, int64_t B, int64_t rows_in, const SpanonTapConv* cv,
                                  void* stream) {
  if (!cv) return SPANON_ERR_NULL_POINTER;
  if (B < 0) return SPANON_ERR_BAD_SHAPE;
  const ConvGeom g = geom(*cv, rows_in);
  if (!g.ok) return SPANON_ERR_UNSUPPORTED;
  if (B == 0) return SPANON_OK;
  if (!in || !w || !out || !workspace) return SPANON_ERR_NULL_POINTER;
  if (workspace_bytes < spanon_tapconv_workspace_bytes(cv)) return SPANON_ERR_WORKSPACE;
  cudaStream_t st = as_stream(stream);
  float* wp = reinterpret_cast<float*>(up256(reinterpret_cast<uintptr_t>(workspace)));
  SPANON_TRY(pack_weights(*cv, g, w, wp, 0, st));
  TapGemmParams p{};
  p.in = in; p.out = out;
  p.rows_in = (int)g.rows_vin; p.rows_out = (int)g.rows_vout;
  p.cin_v = g.cin_v; p.cout_v = g.cout_v; p.cin_mod = cv->cin; p.cout_mod = cv->cout;
  p.in_bstride = g.rows_vin * g.cin_v;
  p.out_mode = cv->cout == 1 ? OUT_CH0_F32 : OUT_CHANNELS_LAST;
  p.out_bstride = cv->cout == 1 ? g.rows_vout : g.rows_vout * g.cout_v;
  p.ntaps = g.ntaps;
  for (int t = 0; t < g.ntaps; ++t) { p.shift[t] = g.shift[t]; p.tap_kmax[t] = g.cin_v; p.tap_nmax[t] = g.cout_v; }
  p.w = wp;
  p.bias = bias; p.relu = relu; p.stats = stats;
  p.B = (int)B;
  p.tag = "train:conv_fwd";
  return launch_tapgemm_simt<float>(p, st);
}

extern "C" int spanon_tapconv_bwd(const float* in, const float* w, const float* gout, float* gin, float* gw, float* gb,
                                  void* workspace, size_t workspace_bytes, int64_t B, int64_t rows_in, const SpanonTapConv* cv,
                                  void* stream) {
  if (!cv) return SPANON_ERR_NULL_POINTER;
  if (B < 0) return SPANON_ERR_BAD_SHAPE;
  const ConvGeom g = geom(*cv, rows_in);
  if (!g.ok) return SPANON_ERR_UNSUPPORTED;
  if (B == 0) return SPANON_OK;
  if (!gout || !workspace || (gin && !w) || (gw && !in)) return SPANON_ERR_NULL_POINTER;
  if (workspace_bytes < spanon_tapconv_workspace_bytes(cv)) return SPANON_ERR_WORKSPACE;
  cudaStream_t st = as_stream(stream);
  char* base = reinterpret_cast<char*>(up256(reinterpret_cast<uintptr_t>(workspace)));
  const size_t wbytes = up256((size_t)g.ntaps * g.cin_v * g.cout_v * 4);
  float* wp = reinterpret_cast<float*>(base);
  float* gwp = reinterpret_cast<float*>(base + wbytes);
  if (cv->cout == 1) {
    // the C -> 1 channel output conv (decoder.8: 32 -> 1, k15): gout is [B][rows] (one channel)
    if (cv->kind != 0 || cv->cin != 32 || cv->k != 15 || cv->padding != 7 || cv->dilation != 1) return SPANON_ERR_UNSUPPORTED;
    const int L = (int)rows_in;
    if (gin) {   // gin[l][c] = sum_t gout[l - (t - 7)] w[0][c][t] = sum_t' gout[l + t' - 7] w[0][c][14 - t']
      dim3 grid((unsigned)ceil_div(L, 256), (unsigned)B);
      conv1_k15_kernel<<<grid, 256, 0, st>>>(gout, w, 15, -1, 14, nullptr, gin, L);
      SPANON_TRY(check_launch("train:conv_c1_dgrad"));
    }
    if (gw) {    // gw[0][c][t] = sum_l in[l + t - 7][c] gout[l] = sum_l' in[l'][c] gout[l' - (t - 7)]
      SPANON_CUDA_TRY(cudaMemsetAsync(gw, 0, 32 * 15 * 4, st));
      const int rpc = 2048;      // (8192 left 80 CTAs for the 8 x 10 s step)
      dim3 grid((unsigned)ceil_div(L, rpc), (unsigned)B);
      corr15_kernel<<<grid, 256, 0, st>>>(gout, in, L, -1, gw, 15, 1, rpc);
      SPANON_TRY(check_launch("train:conv_c1_wgrad"));
    }
    if (gb) SPANON_TRY(bias_grad(gout, B * (int64_t)L, 1, 1, gb, reinterpret_cast<double*>(base + 2 * wbytes), st));
    return SPANON_OK;
  }
  if (gin) {
    // dgrad: the same tap GEMM on the views swapped -- gout [rows_vout][cout_v] in, gin [rows_vin][cin_v] out, weights
    // transposed per tap, shifts negated
    if (g.cout_v % 32 != 0 || (g.cin_v != 64 && g.cin_v != 128)) return SPANON_ERR_UNSUPPORTED;
    SPANON_TRY(pack_weights(*cv, g, w, wp, 1, st));
    TapGemmParams p{};
    p.in = gout; p.out = gin;
    p.rows_in = (int)g.rows_vout; p.rows_out = (int)g.rows_vin;
    p.cin_v = g.cout_v; p.cout_v = g.cin_v; p.cin_mod = g.cout_v; p.cout_mod = g.cin_v;
    p.in_bstride = g.rows_vout * g.cout_v; p.out_bstride = g.rows_vin * g.cin_v;
    p.out_mode = OUT_CHANNELS_LAST;
    p.ntaps = g.ntaps;
    for (int t = 0; t < g.ntaps; ++t) { p.shift[t] = -g.shift[t]; p.tap_kmax[t] = g.cout_v; p.tap_nmax[t] = g.cin_v; }
    p.w = wp;
    p.B = (int)B;
    p.tag = "train:conv_dgrad";
    SPANON_TRY(launch_tapgemm_simt<float>(p, st));
  }
  if (gw) {
    if (g.ntaps > WG_MAXTAPS || (g.cout_v != 64 && g.cout_v != 128)) return SPANON_ERR_UNSUPPORTED;
    SPANON_CUDA_TRY(cudaMemsetAsync(gwp, 0, (size_t)g.ntaps * g.cin_v * g.cout_v * 4, st));
    int lo = 0, hi = 0, sh[WG_MAXTAPS] = {0, 0, 0, 0, 0};
    for (int t = 0; t < g.ntaps; ++t) { sh[t] = g.shift[t]; lo = std::min(lo, sh[t]); hi = std::max(hi, sh[t]); }
    // rows per CTA: about two waves of the co-resident CTAs (two per SM), never a nearly empty last wave -- 4096 rows gave
    // 160 CTAs on 148 SMs for every layer of the 8 x 10 s step, i.e. a second wave with 12 CTAs: half the throughput
    // (walking the A slab once for all taps instead of once per tap was tried: fewer shared-memory loads, but slower)
    const int64_t per_range = (int64_t)(g.cin_v / WG_KC) * B;
    const int n_sm_w = std::max(1, current_device_sms());
    const int64_t n_ranges = std::max<int64_t>(1, std::min<int64_t>((4 * (int64_t)n_sm_w) / per_range, ceil_div(g.rows_vout, (int64_t)WG_SUB)));
    const int rpc = (int)(ceil_div(ceil_div(g.rows_vout, n_ranges), (int64_t)WG_SUB) * WG_SUB);
    const size_t smem = ((size_t)(WG_SUB - lo + hi) * (WG_KC + 4) + (size_t)WG_SUB * g.cout_v) * sizeof(float);
    dim3 grid((unsigned)ceil_div(g.rows_vout, rpc), (unsigned)(g.cin_v / WG_KC), (unsigned)B);
    if (g.cout_v == 128) {
      SPANON_CUDA_TRY(cudaFuncSetAttribute(wgrad_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      wgrad_kernel<128><<<grid, WG_NT, smem, st>>>(in, gout, gwp, (int)g.rows_vin, (int)g.rows_vout, g.cin_v, g.ntaps, sh[0], sh[1],
                                                   sh[2], sh[3], sh[4], rpc);
    } else {
      SPANON_CUDA_TRY(cudaFuncSetAttribute(wgrad_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      wgrad_kernel<64><<<grid, WG_NT, smem, st>>>(in, gout, gwp, (int)g.rows_vin, (int)g.rows_vout, g.cin_v, g.ntaps, sh[0], sh[1],
                                                  sh[2], sh[3], sh[4], rpc);
    }
    SPANON_TRY(check_launch("train:conv_wgrad"));
    const int64_t n = (int64_t)g.ntaps * g.cin_v * g.cout_v;
    unpack_grad_kernel<<<(unsigned)std::min<int64_t>(ceil_div(n, 256), 1024), 256, 0, st>>>(gwp, gw, g.mode, cv->cin, cv->cout, cv->k,
                                                                                          g.ntaps, g.cin_v, g.cout_v);
    SPANON_TRY(check_launch("train:unpack_grad"));
  }
  if (gb) SPANON_TRY(bias_grad(gout, B * g.rows_vout, g.cout_v, cv->cout, gb, reinterpret_cast<double*>(base + 2 * wbytes), st));
  return SPANON_OK;
}

extern "C" int spanon_conv1_fwd(const float* x, const float* w, const float* bias, float* out, int64_t B, int64_t L, void* stream) {
  if (B < 0 || L <= 0 || L > (1 << 30) || B > 65535) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!x || !w || !out) return SPANON_ERR_NULL_POINTER;
  dim3 grid((unsigned)ceil_div(L, 256), (unsigned)B);
  conv1_k15_kernel<<<grid, 256, 0, as_stream(stream)>>>(x, w, 15, 1, 0, bias, out, (int)L);
  return check_launch("train:conv1_fwd");
}

extern "C" int spanon_conv1_bwd(const float* x, const float* gout, float* gw, float* gb, double* scratch, int64_t B, int64_t L,
                                void* stream) {
  if (B < 0 || L <= 0 || L > (1 << 30) || B > 65535) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!gout || (gw && !x) || (gb && !scratch)) return SPANON_ERR_NULL_POINTER;
  cudaStream_t st = as_stream(stream);
  if (gw) {      // gw[c][0][t] = sum_l x[l + t - 7] gout[l][c]
    SPANON_CUDA_TRY(cudaMemsetAsync(gw, 0, 32 * 15 * 4, st));
    const int rpc = 2048;      // (8192 left 80 CTAs for the 8 x 10 s step)
    dim3 grid((unsigned)ceil_div(L, rpc), (unsigned)B);
    corr15_kernel<<<grid, 256, 0, st>>>(x, gout, (int)L, 1, gw, 15, 1, rpc);
    SPANON_TRY(check_launch("train:conv1_wgrad"));
  }
  if (gb) SPANON_TRY(bias_grad(gout, B * L, 32, 32, gb, scratch, st));
  return SPANON_OK;
}

extern "C" int spanon_colstats(const float* x, double* stats, int64_t B, int64_t rows, int64_t C, void* stream) {
  if (B < 0 || rows <= 0 || C <= 0 || C > 256 || 256 % C != 0 || B > 65535) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!x || !stats) return SPANON_ERR_NULL_POINTER;
  cudaStream_t st = as_stream(stream);
  SPANON_CUDA_TRY(cudaMemsetAsync(stats, 0, (size_t)B * C * 2 * sizeof(double), st));
  const int rpc = 256;      // (2048 rows per CTA left 160 CTAs with two loads in flight per warp: 0.2 TB/s)
  dim3 grid((unsigned)ceil_div(rows, rpc), (unsigned)B);
  colstats_kernel<<<grid, 256, 0, st>>>(x, stats, rows, (int)C, rpc);
  return check_launch("train:colstats");
}

extern "C" int spanon_norm_act_fwd(const float* z, const double* stats, const float* gamma, const float* beta, int per_batch,
                                   int act, float eps, float* running_mean, float* running_var, float momentum, float* out,
                                   float* save_mean, float* save_rstd, int64_t B, int64_t rows, int64_t C, int64_t stats_B,
                                   double stats_count, void* stream) {
  if (B < 0 || rows <= 0 || C <= 0 || C % 4 != 0) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!z || !out) return SPANON_ERR_NULL_POINTER;
  if (gamma && (!beta || !stats || !save_mean || !save_rstd)) return SPANON_ERR_NULL_POINTER;
  cudaStream_t st = as_stream(stream);
  if (gamma) {
    const int n = per_batch ? (int)C : (int)(B * C);
    if (stats_B <= 0 || !per_batch) { stats_B = B; stats_count = 0.0; }
    norm_finalize_kernel<<<(unsigned)ceil_div(n, 128), 128, 0, st>>>(stats, (int)stats_B, (int)C, (double)rows, stats_count, eps, per_batch,
                                                                     save_mean, save_rstd, running_mean, running_var, momentum);
    SPANON_TRY(check_launch("train:norm_finalize"));
  }
  const int64_t total4 = B * rows * C / 4;
  norm_act_fwd_kernel<<<(unsigned)std::min<int64_t>(ceil_div(total4, 256), 148 * 16), 256, 0, st>>>(
      z, save_mean, save_rstd, per_batch ? 0 : (int)C, gamma, beta, act, out, rows, (int)C, total4);
  return check_launch("train:norm_act_fwd");
}

extern "C" int spanon_norm_act_bwd(const float* z, const float* gout, const float* gamma, const float* beta, const float* save_mean,
                                   const float* save_rstd, int per_batch, int act, int relu_mask, float* gz, float* ggamma,
                                   float* gbeta, double* scratch, int64_t B, int64_t rows, int64_t C, int phase, double count,
                                   void* stream) {
  if (B < 0 || rows <= 0 || C <= 0 || C % 4 != 0 || B > 65535) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!z || !gout || !gz) return SPANON_ERR_NULL_POINTER;
  if (gamma && (!beta || !save_mean || !save_rstd || !scratch || C > 256 || 256 % C != 0)) return SPANON_ERR_NULL_POINTER;
  cudaStream_t st = as_stream(stream);
  const int sb = per_batch ? 0 : (int)C;
  if (phase < 0 || phase > 2) return SPANON_ERR_UNSUPPORTED;
  if (gamma && phase != 2) {
    SPANON_CUDA_TRY(cudaMemsetAsync(scratch, 0, (size_t)B * C * 2 * sizeof(double), st));
    const int rpc = 256;    // (see colstats: enough CTAs to keep the loads of a bandwidth-bound reduction in flight)
    dim3 grid((unsigned)ceil_div(rows, rpc), (unsigned)B);
    norm_bwd_reduce_kernel<<<grid, 256, 0, st>>>(z, gout, save_mean, save_rstd, sb, gamma, beta, act, scratch, rows, (int)C, rpc);
    SPANON_TRY(check_launch("train:norm_bwd_reduce"));
    norm_bwd_params_kernel<<<(unsigned)ceil_div(C, 128), 128, 0, st>>>(scratch, (int)B, (int)C, per_batch, ggamma, gbeta);
    SPANON_TRY(check_launch("train:norm_bwd_params"));
  }
  if (phase == 1) return SPANON_OK;     // the caller all-reduces scratch[0][C][2] across ranks, then calls phase 2
  const int64_t total4 = B * rows * C / 4;
  const double n = per_batch ? (count > 0.0 ? count : (double)(rows * B)) : (double)rows;
  norm_bwd_apply_kernel<<<(unsigned)std::min<int64_t>(ceil_div(total4, 256), 148 * 16), 256, 0, st>>>(
      z, gout, save_mean, save_rstd, sb, gamma, beta, act, scratch, per_batch ? 0 : (int)C, n, relu_mask, gz, rows, (int)C, total4);
  return check_launch("train:norm_bwd_apply");
}

extern "C" int spanon_pool_reinterp_fwd(const float* x, const float* noise, float* out, float* save, double* scratch, int64_t B,
                                        int64_t rows, int64_t C, void* stream) {
  if (B < 0 || rows <= 0 || rows > (1 << 30) || B > 65535) return SPANON_ERR_BAD_SHAPE;
  if (C != 128) return SPANON_ERR_UNSUPPORTED;
  if (B == 0) return SPANON_OK;
  if (!x || !out || !save || !scratch) return SPANON_ERR_NULL_POINTER;
  cudaStream_t st = as_stream(stream);
  SPANON_CUDA_TRY(cudaMemsetAsync(scratch, 0, (size_t)B * 128 * 2 * sizeof(double), st));
  const int rpc = 256;       // (a multiple of 128, see the kernel; 1024 left 160 small CTAs for the 8 x 10 s step)
  dim3 grid((unsigned)ceil_div(rows, rpc), (unsigned)B);
  pool_reinterp_sums_kernel<<<grid, 128, 0, st>>>(x, scratch, (int)rows, rpc);
  SPANON_TRY(check_launch("train:pool_sums"));
  pool_reinterp_finalize_kernel<<<(unsigned)B, 128, 0, st>>>(scratch, noise, (int)rows, out, save);
  return check_launch("train:pool_finalize");
}

extern "C" int spanon_pool_reinterp_bwd(const float* x, const float* save, const float* gout, float* gx, int64_t B, int64_t rows,
                                        int64_t C, void* stream) {
  if (B < 0 || rows <= 0 || rows > (1 << 30)) return SPANON_ERR_BAD_SHAPE;
  if (C != 128) return SPANON_ERR_UNSUPPORTED;
  if (B == 0) return SPANON_OK;
  if (!x || !save || !gout || !gx) return SPANON_ERR_NULL_POINTER;
  const int64_t total = B * rows * 128;
  pool_reinterp_bwd_kernel<<<(unsigned)std::min<int64_t>(ceil_div(total, 256), 148 * 16), 256, 0, as_stream(stream)>>>(
      x, save, gout, gx, (int)rows, total);
  return check_launch("train:pool_bwd");
}
