// Warp-level reduction kernels for the loss statistics.
//
//   CosineSimilarityLoss.forward        utils/cosine_similarity_loss.py:53-56
//       sim = CosineSimilarity(dim=2, eps=1e-6)(x1, x2); loss = sum(1 - sim) / sim.shape[1]
//   ClusterMI._mutual_information       utils/ClusterMI.py:88-121  (Ross 2014 kNN estimator)
//       pairwise cosine distances (:12-63), same-class (k+1)-th smallest as anchor (:108-112),
//       m_i = #(d <= anchor) - 1 (:115), MI = psi(N) - <psi(N_x)> + psi(k) - <psi(m_i)>, in bits
//   GroupSamplingMI.forward             utils/GroupSamplingMI.py:49-61  -- 100 host iterations in
//       the reference; here all index sets are evaluated by ONE launch (one CTA per set).
#include "common.cuh"
#include <algorithm>
#include <math.h>

namespace spanon {
namespace {

// one warp per row; block = 256 threads
__global__ void __launch_bounds__(256)
cosine_loss_kernel(const float* __restrict__ x1, const float* __restrict__ x2, int64_t rows, int D, float eps,
                   float inv_T, float* __restrict__ loss, float* __restrict__ row_loss) {
  __shared__ float red[8];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float local = 0.f;
  for (int64_t r = (int64_t)blockIdx.x * 8 + wid; r < rows; r += (int64_t)gridDim.x * 8) {
    const float* a = x1 + r * D;
    const float* b = x2 + r * D;
    float dot = 0.f, na = 0.f, nb = 0.f;
    for (int i = lane; i < D; i += 32) {
      const float u = __ldg(a + i), v = __ldg(b + i);
      dot = fmaf(u, v, dot); na = fmaf(u, u, na); nb = fmaf(v, v, nb);
    }
    dot = warp_sum(dot); na = warp_sum(na); nb = warp_sum(nb);
    const float l = 1.f - dot / (fmaxf(sqrtf(na), eps) * fmaxf(sqrtf(nb), eps));
    if (lane == 0) {
      if (row_loss) row_loss[r] = l;
      local += l;
    }
  }
  if (lane == 0) red[wid] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[i];
    atomicAdd(loss, s * inv_T);
  }
}

__global__ void __launch_bounds__(256)
cosine_loss_bwd_kernel(const float* __restrict__ x1, const float* __restrict__ x2, const float* __restrict__ gout,
                       int64_t rows, int D, float eps, float inv_T, float* __restrict__ g1, float* __restrict__ g2) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const float g = -__ldg(gout) * inv_T;  // d loss / d cos
  for (int64_t r = (int64_t)blockIdx.x * 8 + wid; r < rows; r += (int64_t)gridDim.x * 8) {
    const float* a = x1 + r * D;
    const float* b = x2 + r * D;
    float dot = 0.f, na = 0.f, nb = 0.f;
    for (int i = lane; i < D; i += 32) {
      const float u = __ldg(a + i), v = __ldg(b + i);
      dot = fmaf(u, v, dot); na = fmaf(u, u, na); nb = fmaf(v, v, nb);
    }
    dot = warp_sum(dot); na = warp_sum(na); nb = warp_sum(nb);
    const float n1 = sqrtf(na), n2 = sqrtf(nb);
    const float m1 = fmaxf(n1, eps), m2 = fmaxf(n2, eps);
    const float inv = 1.f / (m1 * m2);
    const float c = dot * inv;
    const float k1 = n1 > eps ? c / (m1 * n1) : 0.f;
    const float k2 = n2 > eps ? c / (m2 * n2) : 0.f;
    for (int i = lane; i < D; i += 32) {
      const float u = __ldg(a + i), v = __ldg(b + i);
      if (g1) g1[r * D + i] = g * (v * inv - k1 * u);
      if (g2) g2[r * D + i] = g * (u * inv - k2 * v);
    }
  }
}

constexpr int MI_NT = 256;
constexpr int MI_WARPS = MI_NT / 32;

// psi(n) for integer n: -gamma + H_{n-1};  psi(0) = -inf (torch.digamma)
__device__ __forceinline__ double digamma_int(const double* H, int n) {
  return n <= 0 ? -INFINITY : H[n - 1] - 0.57721566490153286061;
}

// grid = n_sets ; dynamic smem: H[N+1] doubles | norm[N] | rowbuf[MI_WARPS][N] | psi_part[MI_WARPS] doubles
__global__ void __launch_bounds__(MI_NT)
knn_mi_kernel(const float* __restrict__ X, const int32_t* __restrict__ y, const int32_t* __restrict__ idx, int N, int D,
              int n_classes, int k, float* __restrict__ dists, int32_t* __restrict__ m_out, float* __restrict__ mi) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* H = reinterpret_cast<double*>(smem_raw);          // H[n] = sum_{i=1..n} 1/i
  double* psi_part = H + (N + 1);
  float* norm = reinterpret_cast<float*>(psi_part + MI_WARPS);
  int* lab = reinterpret_cast<int*>(norm + N);
  float* rowbuf = reinterpret_cast<float*>(lab + N);
  const int set = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int32_t* ids = idx ? idx + (int64_t)set * N : nullptr;

  if (tid == 0) {
    double h = 0.0;
    H[0] = 0.0;
    for (int n = 1; n <= N; ++n) { h += 1.0 / (double)n; H[n] = h; }
  }
  for (int j = tid; j < N; j += MI_NT) {
    const int64_t row = ids ? ids[j] : j;
    const float* x = X + row * D;
    float s = 0.f;
    for (int d = 0; d < D; ++d) { const float v = __ldg(x + d); s = fmaf(v, v, s); }
    norm[j] = fmaxf(sqrtf(s), 1e-8f);   // F.cosine_similarity eps (ClusterMI.py:9-10 default)
    lab[j] = y[row];
  }
  __syncthreads();

  float* drow = rowbuf + wid * N;
  double psi_acc = 0.0;
  for (int i = wid; i < N; i += MI_WARPS) {
    const int64_t ri = ids ? ids[i] : i;
    const float* xi = X + ri * D;
    const float ni = norm[i];
    const int yi = lab[i];
    for (int j = lane; j < N; j += 32) {
      float d = 0.f;  // main diagonal stays 0 (ClusterMI.py:61-62)
      if (j != i) {
        const int64_t rj = ids ? ids[j] : j;
        const float* xj = X + rj * D;
        float dot = 0.f;
        for (int q = 0; q < D; ++q) dot = fmaf(__ldg(xi + q), __ldg(xj + q), dot);
        d = 1.f - dot / (ni * norm[j]);
      }
      drow[j] = d;
      if (dists) dists[((int64_t)set * N + i) * N + j] = d;
    }
    __syncwarp();
    // (k+1)-th smallest same-class distance: k+1 passes of "smallest (value, index) above the last one"
    float last_v = -INFINITY; int last_j = -1;
    for (int pass = 0; pass <= k; ++pass) {
      float bv = INFINITY; int bj = 0x7fffffff;
      for (int j = lane; j < N; j += 32) {
        // other classes are replaced by 10e6 in the reference (:109)
        const float v = (lab[j] == yi) ? drow[j] : 10e6f;
        const bool after = (v > last_v) || (v == last_v && j > last_j);
        if (after && (v < bv || (v == bv && j < bj))) { bv = v; bj = j; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
        if (ov < bv || (ov == bv && oj < bj)) { bv = ov; bj = oj; }
      }
      last_v = bv; last_j = bj;
    }
    const float anchor = last_v;
    int cnt = 0;
    for (int j = lane; j < N; j += 32) cnt += (drow[j] <= anchor) ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    const int m = cnt - 1;
    if (lane == 0) {
      if (m_out) m_out[(int64_t)set * N + i] = m;
      psi_acc += digamma_int(H, m);
    }
    __syncwarp();
  }
  if (lane == 0) psi_part[wid] = psi_acc;
  __syncthreads();
  if (tid == 0 && mi) {
    double avg_m = 0.0;
    for (int w = 0; w < MI_WARPS; ++w) avg_m += psi_part[w];
    avg_m /= (double)N;
    double avg_nx = 0.0;
    for (int c = 0; c < n_classes; ++c) {
      int cnt = 0;
      for (int j = 0; j < N; ++j) cnt += (lab[j] == c) ? 1 : 0;
      avg_nx += ((double)cnt / (double)N) * digamma_int(H, cnt);  // 0 * -inf = NaN like torch
    }
    const double v = digamma_int(H, N) - avg_nx + digamma_int(H, k) - avg_m;
    mi[set] = (float)(v / 0.69314718055994530942);
  }
}

}  // namespace
}  // namespace spanon

using namespace spanon;

extern "C" int spanon_cosine_loss(const float* x1, const float* x2, int64_t B, int64_t T, int64_t D, float eps,
                                  float* loss, float* row_loss, void* stream) {
  if (B < 0 || T <= 0 || D <= 0 || D > INT32_MAX) return SPANON_ERR_BAD_SHAPE;
  if (!loss) return SPANON_ERR_NULL_POINTER;
  cudaStream_t st = as_stream(stream);
  SPANON_CUDA_TRY(cudaMemsetAsync(loss, 0, sizeof(float), st));
  if (B == 0) return SPANON_OK;
  if (!x1 || !x2) return SPANON_ERR_NULL_POINTER;
  const int64_t rows = B * T;
  const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(rows, 8), 148 * 8);
  cosine_loss_kernel<<<grid, 256, 0, st>>>(x1, x2, rows, (int)D, eps, 1.0f / (float)T, loss, row_loss);
  return check_launch("cosine_loss_kernel");
}

extern "C" int spanon_cosine_loss_bwd(const float* x1, const float* x2, const float* grad_out, int64_t B, int64_t T,
                                      int64_t D, float eps, float* gx1, float* gx2, void* stream) {
  if (B < 0 || T <= 0 || D <= 0 || D > INT32_MAX) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!x1 || !x2 || !grad_out) return SPANON_ERR_NULL_POINTER;
  const int64_t rows = B * T;
  const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(rows, 8), 148 * 8);
  cosine_loss_bwd_kernel<<<grid, 256, 0, as_stream(stream)>>>(x1, x2, grad_out, rows, (int)D, eps, 1.0f / (float)T,
                                                              gx1, gx2);
  return check_launch("cosine_loss_bwd_kernel");
}

extern "C" int spanon_knn_mi_stats(const float* X, const int32_t* y, const int32_t* idx, int64_t n_sets, int64_t N,
                                   int64_t D, int32_t n_classes, int32_t k, float* dists, int32_t* m_i, float* mi,
                                   void* stream) {
  if (n_sets < 0 || N <= 0 || D <= 0 || n_classes < 2 || k < 1) return SPANON_ERR_BAD_SHAPE;
  if (N > 2048 || n_sets > 65535 || k + 1 > N) return SPANON_ERR_UNSUPPORTED;
  if (n_sets == 0) return SPANON_OK;
  if (!X || !y) return SPANON_ERR_NULL_POINTER;
  const size_t smem = (size_t)(N + 1 + MI_WARPS) * sizeof(double) + (size_t)N * (sizeof(float) + sizeof(int)) +
                      (size_t)MI_WARPS * N * sizeof(float);
  SPANON_CUDA_TRY(cudaFuncSetAttribute(knn_mi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  knn_mi_kernel<<<(unsigned)n_sets, MI_NT, smem, as_stream(stream)>>>(X, y, idx, (int)N, (int)D, n_classes, k, dists,
                                                                     m_i, mi);
  return check_launch("knn_mi_kernel");
}
