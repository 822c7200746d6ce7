// SpecAugment on features (the TRAIN-stage augmentation of the anonymizer, speechbrain_convae_train.py:64-67 with
// convae.yaml:273-283): time warp by bicubic resampling, frequency masks, time masks.
// Reference: speechbrain/speechbrain/lobes/augment.py:32-200.  The random draws (warp centre / width, mask lengths and
// positions) stay on the host side of the boundary -- the module makes the same torch.randint calls in the same order as the
// reference --; these kernels apply them.
//   time_warp :127-157   left  = interpolate(x[:, :c], (w, F), bicubic, align_corners=True)
//                        right = interpolate(x[:, c:], (T - w, F), bicubic, align_corners=True)   (F unchanged: the
//                        interpolation along the feature axis is the identity)
//   mask_along_axis :159-200   x.masked_fill_(any_k(pos_k <= idx < pos_k + len_k), 0 or x.mean())
#include "common.cuh"
#include <algorithm>

namespace spanon {
namespace {

// torch's upsample_bicubic2d coefficients (A = -0.75)
__device__ __forceinline__ void cubic_coeffs(float t, float* w) {
  const float A = -0.75f;
  const float x0 = t + 1.f, x3 = 2.f - t, x2 = 1.f - t;
  w[0] = ((A * x0 - 5.f * A) * x0 + 8.f * A) * x0 - 4.f * A;
  w[1] = ((A + 2.f) * t - (A + 3.f)) * t * t + 1.f;
  w[2] = ((A + 2.f) * x2 - (A + 3.f)) * x2 * x2 + 1.f;
  w[3] = ((A * x3 - 5.f * A) * x3 + 8.f * A) * x3 - 4.f * A;
}

// out[b][t][f]: t < w resamples rows [0, c) onto w rows, t >= w resamples rows [c, T) onto T - w rows
__global__ void __launch_bounds__(256)
specaug_warp_kernel(const float* __restrict__ in, float* __restrict__ out, int T, int F, int c, int w, int64_t total) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int f = (int)(i % F);
    const int64_t bt = i / F;
    const int t = (int)(bt % T);
    const int64_t b = bt / T;
    int src0, n_in, n_out, to;
    if (t < w) { src0 = 0; n_in = c; n_out = w; to = t; }
    else { src0 = c; n_in = T - c; n_out = T - w; to = t - w; }
    const float scale = n_out > 1 ? (float)(n_in - 1) / (float)(n_out - 1) : 0.f;    // align_corners=True
    const float real = scale * (float)to;
    const int ix = (int)floorf(real);
    float cw[4];
    cubic_coeffs(real - (float)ix, cw);
    const float* row = in + (b * T + src0) * (int64_t)F + f;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int r = min(max(ix - 1 + k, 0), n_in - 1);
      acc = fmaf(cw[k], __ldg(row + (int64_t)r * F), acc);
    }
    out[i] = acc;
  }
}

__global__ void __launch_bounds__(256)
sum_kernel(const float* __restrict__ x, int64_t n, double* __restrict__ acc) {
  __shared__ double red[8];
  double s = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) s += (double)__ldg(x + i);
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < 8; ++k) t += red[k];
    atomicAdd(acc, t);
  }
}

// axis 1: time (idx = t), axis 2: feature (idx = f).  fill = sum[0] / n when sum != null (replace_with_zero = False).
__global__ void __launch_bounds__(256)
specaug_mask_kernel(float* __restrict__ x, int T, int F, int axis, int n_mask, const int32_t* __restrict__ mlen,
                    const int32_t* __restrict__ mpos, const double* __restrict__ sum, double n_total, int64_t total) {
  const float fill = sum ? (float)(sum[0] / n_total) : 0.f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int f = (int)(i % F);
    const int64_t bt = i / F;
    const int t = (int)(bt % T);
    const int64_t b = bt / T;
    const int idx = axis == 1 ? t : f;
    bool m = false;
    for (int k = 0; k < n_mask; ++k) {
      const int p = mpos[b * n_mask + k], l = mlen[b * n_mask + k];
      m |= (p <= idx) && (idx < p + l);
    }
    if (m) x[i] = fill;
  }
}

}  // namespace
}  // namespace spanon

using namespace spanon;

extern "C" int spanon_specaug_warp(const float* in, float* out, int64_t B, int64_t T, int64_t F, int32_t c, int32_t w, void* stream) {
  if (B < 0 || T <= 0 || F <= 0 || T > (1 << 24)) return SPANON_ERR_BAD_SHAPE;
  if (c < 1 || c >= T || w < 1 || w >= T) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!in || !out || in == out) return SPANON_ERR_NULL_POINTER;
  const int64_t total = B * T * F;
  specaug_warp_kernel<<<(unsigned)std::min<int64_t>(ceil_div(total, 256), 148 * 16), 256, 0, as_stream(stream)>>>(in, out, (int)T, (int)F, c, w,
                                                                                                              total);
  return check_launch("specaug_warp_kernel");
}

extern "C" int spanon_specaug_mask(float* x, int64_t B, int64_t T, int64_t F, int32_t axis, int32_t n_mask, const int32_t* mask_len,
                                   const int32_t* mask_pos, int32_t replace_with_zero, double* scratch, void* stream) {
  if (B < 0 || T <= 0 || F <= 0 || (axis != 1 && axis != 2) || n_mask < 0) return SPANON_ERR_BAD_SHAPE;
  if (B == 0 || n_mask == 0) return SPANON_OK;
  if (!x || !mask_len || !mask_pos || (!replace_with_zero && !scratch)) return SPANON_ERR_NULL_POINTER;
  cudaStream_t st = as_stream(stream);
  const int64_t total = B * T * F;
  if (!replace_with_zero) {       // val = x.mean() of the tensor as it is now (augment.py:194-197)
    SPANON_CUDA_TRY(cudaMemsetAsync(scratch, 0, sizeof(double), st));
    sum_kernel<<<(unsigned)std::min<int64_t>(ceil_div(total, 256 * 8), 1024), 256, 0, st>>>(x, total, scratch);
    SPANON_TRY(check_launch("specaug_sum_kernel"));
  }
  specaug_mask_kernel<<<(unsigned)std::min<int64_t>(ceil_div(total, 256), 148 * 16), 256, 0, st>>>(
      x, (int)T, (int)F, axis, n_mask, mask_len, mask_pos, replace_with_zero ? nullptr : scratch, (double)total, total);
  return check_launch("specaug_mask_kernel");
}
