// Host-buffer entry point: PCM on the host in, anonymized features on the host out.
// This is the composition SexAnonymizationTraining.compute_forward performs
// (speechbrain_convae_train.py:42-69): compute_features -> normalize -> pad T to a multiple
// of 36 (ConvAE only) -> anonymizer, with the top_db clamp, the normalisation and the padding
// folded into the anonymizer's first load so the features make one HBM round trip.
#include "common.cuh"
#include <algorithm>

using namespace spanon;

namespace {
inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }
struct Carve {
  float* pcm; float* fb; int32_t* utt_max; float* recon; float* logp; void* anon_ws; size_t anon_bytes; size_t total;
};
int64_t padded_T(int kind, int64_t T) { return kind == 0 ? (T + 35) / 36 * 36 : T; }  // train.py:62
Carve carve(char* base, int kind, int64_t B, int64_t n_samples, int32_t hop, int32_t n_mels,
            const SpanonConvAEOpts* opts) {
  Carve c{};
  const int64_t T = 1 + n_samples / hop, Tp = padded_T(kind, T);
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += up256(bytes); return p; };
  c.pcm = (float*)take((size_t)B * n_samples * 4);
  c.fb = (float*)take((size_t)B * T * n_mels * 4);
  c.utt_max = (int32_t*)take((size_t)B * 4);
  c.recon = (float*)take((size_t)B * Tp * n_mels * 4);
  c.logp = (float*)take((size_t)B * 2 * 4);
  c.anon_bytes = kind == 0 ? spanon_convae_workspace_bytes(B, Tp, n_mels, opts) : spanon_fcae_workspace_bytes(B, T);
  c.anon_ws = take(c.anon_bytes);
  c.total = off;
  return c;
}
}  // namespace

extern "C" size_t spanon_anonymize_workspace_bytes(int kind, int64_t B, int64_t n_samples, int32_t hop_length,
                                                   int32_t n_mels, const SpanonConvAEOpts* opts) {
  if (B <= 0 || n_samples < 0 || hop_length <= 0 || n_mels <= 0 || (kind != 0 && kind != 1)) return 0;
  return carve(nullptr, kind, B, n_samples, hop_length, n_mels, opts).total + 256;
}

extern "C" int spanon_anonymize_host(int kind, const float* pcm_host, int64_t B, int64_t n_samples,
                                     const SpanonFbankCfg* fb, const float* norm_mean, const float* norm_std,
                                     const void* weights, const float* pool_noise, const SpanonConvAEOpts* opts,
                                     float* recon_host, float* logp_host, void* dev_workspace,
                                     size_t dev_workspace_bytes, void* stream) {
  if (kind != 0 && kind != 1) return SPANON_ERR_UNSUPPORTED;
  if (!fb || !weights || !pcm_host || !recon_host || !dev_workspace || !norm_mean || !norm_std)
    return SPANON_ERR_NULL_POINTER;
  if (B <= 0 || n_samples < 0 || fb->hop_length <= 0) return SPANON_ERR_BAD_SHAPE;
  if (dev_workspace_bytes < spanon_anonymize_workspace_bytes(kind, B, n_samples, fb->hop_length, fb->n_mels, opts))
    return SPANON_ERR_WORKSPACE;
  char* base = reinterpret_cast<char*>(up256(reinterpret_cast<uintptr_t>(dev_workspace)));
  Carve c = carve(base, kind, B, n_samples, fb->hop_length, fb->n_mels, opts);
  cudaStream_t st = as_stream(stream);
  const int64_t T = 1 + n_samples / fb->hop_length, Tp = padded_T(kind, T), F = fb->n_mels;
  SPANON_CUDA_TRY(cudaMemcpyAsync(c.pcm, pcm_host, (size_t)B * n_samples * 4, cudaMemcpyHostToDevice, st));
  SPANON_TRY(spanon_fbank_f32(c.pcm, B, n_samples, n_samples, c.fb, c.utt_max, fb, stream));
  float* logp_dev = logp_host ? c.logp : nullptr;
  if (kind == 0) {
    SPANON_TRY(spanon_convae_fwd(c.fb, (const SpanonConvAEWeights*)weights, pool_noise, norm_mean, norm_std,
                                 c.utt_max, fb->top_db, T, c.recon, logp_dev, c.anon_ws, c.anon_bytes, B, Tp, F, opts,
                                 stream));
  } else {
    SPANON_TRY(spanon_fcae_fwd(c.fb, (const SpanonFcaeWeights*)weights, pool_noise, norm_mean, norm_std, c.utt_max,
                               fb->top_db, c.recon, logp_dev, c.anon_ws, c.anon_bytes, B, T, stream));
  }
  SPANON_CUDA_TRY(cudaMemcpyAsync(recon_host, c.recon, (size_t)B * Tp * F * 4, cudaMemcpyDeviceToHost, st));
  if (logp_host) SPANON_CUDA_TRY(cudaMemcpyAsync(logp_host, c.logp, (size_t)B * 2 * 4, cudaMemcpyDeviceToHost, st));
  SPANON_CUDA_TRY(cudaStreamSynchronize(st));
  return SPANON_OK;
}
