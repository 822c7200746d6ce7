// InputNormalization statistics / application kernels.
//
// Reference (speechbrain/speechbrain/processing/features.py):
//   forward :1018-1029   actual_size = torch.round(lengths[b] * T).int(); stats on x[b, :actual_size]
//   _compute_current_stats :1133-1146  mean, UNBIASED std, std = max(std, eps)
//   forward :1119        x = (x - mean) / std
// and speechbrain_convae_train.py:62-63 (zero padding of T to a multiple of 36 AFTER
// normalisation, so padded rows are exactly 0).
//
// The reference loops over the batch in Python with a host sync per utterance; here one
// launch computes all utterances, the valid-frame count is evaluated on the device with the
// same fp32 product and round-half-even, and nothing synchronises the host.
#include "common.cuh"
#include <algorithm>
#include <math.h>

namespace spanon {
namespace {

constexpr int STAT_ROWS = 8;  // row groups per CTA

// grid = B ; block = (F_pad32 <= 128, STAT_ROWS)
__global__ void norm_stats_kernel(const float* __restrict__ feats, const float* __restrict__ rel_len, int T, int F,
                                  const int32_t* __restrict__ utt_max, float top_db, int mean_norm, int std_norm,
                                  float eps, float* __restrict__ mean_out, float* __restrict__ std_out,
                                  int32_t* __restrict__ n_valid) {
  extern __shared__ double sm[];  // [STAT_ROWS][F][2]
  const int b = blockIdx.x;
  const int f = threadIdx.x, r = threadIdx.y;
  // features.py:1021 -- fp32 product, torch.round (half to even), .int()
  int n = (int)rintf(rel_len[b] * (float)T);
  n = max(0, min(n, T));
  const float floor_db = utt_max ? key_to_float(utt_max[b]) - top_db : -INFINITY;
  const float* x = feats + (int64_t)b * T * F;
  double s = 0.0, q = 0.0;
  if (f < F) {
    for (int t = r; t < n; t += STAT_ROWS) {
      const float v = fmaxf(x[(int64_t)t * F + f], floor_db);
      s += (double)v;
      q += (double)v * (double)v;
    }
    sm[(r * F + f) * 2 + 0] = s;
    sm[(r * F + f) * 2 + 1] = q;
  }
  __syncthreads();
  if (r == 0 && f < F) {
    for (int i = 1; i < STAT_ROWS; ++i) {
      s += sm[(i * F + f) * 2 + 0];
      q += sm[(i * F + f) * 2 + 1];
    }
    // torch.mean of an empty slice is NaN; torch.std with < 2 rows is NaN; max(NaN, eps) = NaN
    const double mean = n > 0 ? s / (double)n : (double)NAN;
    double var = n > 1 ? (q - s * s / (double)n) / (double)(n - 1) : (double)NAN;
    if (var < 0.0) var = 0.0;
    float sd = (float)sqrt(var);
    sd = (sd != sd) ? sd : fmaxf(sd, eps);
    mean_out[(int64_t)b * F + f] = mean_norm ? (float)mean : 0.f;
    std_out[(int64_t)b * F + f] = std_norm ? sd : 1.f;
    if (f == 0 && n_valid) n_valid[b] = n;
  }
}

// grid = (blocks, B); each thread handles float elements of the [T_pad, F] plane
__global__ void norm_apply_kernel(const float* __restrict__ feats, float* __restrict__ out,
                                  const float* __restrict__ mean, const float* __restrict__ sd, int64_t stat_bstride,
                                  const int32_t* __restrict__ utt_max, float top_db, int T, int T_pad, int F) {
  const int b = blockIdx.y;
  const float floor_db = utt_max ? key_to_float(utt_max[b]) - top_db : -INFINITY;
  const float* x = feats + (int64_t)b * T * F;
  float* y = out + (int64_t)b * T_pad * F;
  const float* mu = mean + (int64_t)b * stat_bstride;
  const float* sg = sd + (int64_t)b * stat_bstride;
  const int64_t n_in = (int64_t)T * F, n_out = (int64_t)T_pad * F;
  // walk backwards-safe: in-place is only allowed when T_pad == T (same index)
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_out;
       i += (int64_t)gridDim.x * blockDim.x) {
    float v = 0.f;
    if (i < n_in) {
      const int f = (int)(i % F);
      v = (fmaxf(x[i], floor_db) - mu[f]) / sg[f];
    }
    y[i] = v;
  }
}

}  // namespace
}  // namespace spanon

using namespace spanon;

extern "C" int spanon_norm_stats(const float* feats, const float* rel_len, int64_t B, int64_t T, int64_t F,
                                 const int32_t* utt_max, float top_db, int mean_norm, int std_norm, float eps,
                                 float* mean_out, float* std_out, int32_t* n_valid, void* stream) {
  if (B < 0 || T < 0 || F <= 0 || F > 128) return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!feats || !rel_len || !mean_out || !std_out) return SPANON_ERR_NULL_POINTER;
  dim3 block((unsigned)((F + 31) / 32 * 32), STAT_ROWS);
  const size_t smem = (size_t)STAT_ROWS * F * 2 * sizeof(double);
  norm_stats_kernel<<<(unsigned)B, block, smem, as_stream(stream)>>>(feats, rel_len, (int)T, (int)F, utt_max, top_db,
                                                                     mean_norm, std_norm, eps, mean_out, std_out,
                                                                     n_valid);
  return check_launch("norm_stats_kernel");
}

extern "C" int spanon_norm_apply(const float* feats, float* out, const float* mean, const float* std_,
                                 int64_t stat_bstride, const int32_t* utt_max, float top_db, int64_t B, int64_t T,
                                 int64_t T_pad, int64_t F, void* stream) {
  if (B < 0 || T < 0 || T_pad < T || F <= 0) return SPANON_ERR_BAD_SHAPE;
  if (stat_bstride != 0 && stat_bstride != F) return SPANON_ERR_BAD_SHAPE;
  if (B == 0 || T_pad == 0) return SPANON_OK;
  if (!feats || !out || !mean || !std_) return SPANON_ERR_NULL_POINTER;
  if (feats == out && T_pad != T) return SPANON_ERR_UNSUPPORTED;
  if (B > 65535) return SPANON_ERR_BAD_SHAPE;
  const int64_t per = T_pad * F;
  dim3 grid((unsigned)std::min<int64_t>(ceil_div(per, 256 * 4), 1024), (unsigned)B);
  norm_apply_kernel<<<grid, 256, 0, as_stream(stream)>>>(feats, out, mean, std_, stat_bstride, utt_max, top_db,
                                                         (int)T, (int)T_pad, (int)F);
  return check_launch("norm_apply_kernel");
}
