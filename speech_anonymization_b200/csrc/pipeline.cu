// Host-buffer entry point: PCM on the host in, anonymized features on the host out.
// This is the composition SexAnonymizationTraining.compute_forward performs
// (speechbrain_convae_train.py:42-69): compute_features -> normalize -> pad T to a multiple
// of 36 (ConvAE only) -> anonymizer, with the top_db clamp, the normalisation and the padding
// folded into the anonymizer's first load so the features make one HBM round trip.
#include "common.cuh"
#include <algorithm>
#include <stdlib.h>

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

namespace {
struct Lanes {   // ev_* NULL when all three are one stream; graph != NULL: replay instead of launching the kernels
  cudaStream_t h2d, compute, d2h; cudaEvent_t ev_h2d, ev_compute; cudaGraphExec_t graph = nullptr;
};

// The kernels of one batch (Fbank -> anonymizer) on `st`; everything it touches is device memory of the slot.
int enqueue_compute(int kind, int fmt, int64_t B, int64_t n_samples, const SpanonFbankCfg* fb, const float* norm_mean,
                    const float* norm_std, const void* weights, const float* pool_noise, const SpanonConvAEOpts* opts,
                    bool want_logp, const Carve& c, cudaStream_t st) {
  const int64_t T = 1 + n_samples / fb->hop_length, Tp = padded_T(kind, T), F = fb->n_mels;
  if (fmt == 2)
    SPANON_TRY(spanon_fbank_i16(reinterpret_cast<const int16_t*>(c.pcm), B, n_samples, n_samples, c.fb, c.utt_max, fb, st));
  else
    SPANON_TRY(spanon_fbank_f32(c.pcm, B, n_samples, n_samples, c.fb, c.utt_max, fb, st));
  float* logp_dev = want_logp ? c.logp : nullptr;
  if (kind == 0)
    return spanon_convae_fwd(c.fb, (const SpanonConvAEWeights*)weights, pool_noise, norm_mean, norm_std, c.utt_max,
                             fb->top_db, T, c.recon, logp_dev, c.anon_ws, c.anon_bytes, B, Tp, F, opts, st);
  return spanon_fcae_fwd(c.fb, (const SpanonFcaeWeights*)weights, pool_noise, norm_mean, norm_std, c.utt_max, fb->top_db,
                         c.recon, logp_dev, c.anon_ws, c.anon_bytes, B, T, st);
}

// Enqueues one batch: H2D on lanes.h2d, kernels on lanes.compute, D2H on lanes.d2h (ordered by the two events).
// fmt: bytes per PCM sample on the wire, 4 = fp32, 2 = int16 (half the H2D bytes; converted on load by the Fbank kernel)
int enqueue_batch(int kind, const void* pcm_host, int fmt, int64_t B, int64_t n_samples, const SpanonFbankCfg* fb,
                  const float* norm_mean, const float* norm_std, const void* weights, const float* pool_noise,
                  const SpanonConvAEOpts* opts, float* recon_host, float* logp_host, const Carve& c, const Lanes& ln) {
  const int64_t T = 1 + n_samples / fb->hop_length, Tp = padded_T(kind, T), F = fb->n_mels;
  SPANON_CUDA_TRY(cudaMemcpyAsync(c.pcm, pcm_host, (size_t)B * n_samples * fmt, cudaMemcpyHostToDevice, ln.h2d));
  if (ln.ev_h2d) {
    SPANON_CUDA_TRY(cudaEventRecord(ln.ev_h2d, ln.h2d));
    SPANON_CUDA_TRY(cudaStreamWaitEvent(ln.compute, ln.ev_h2d, 0));
  }
  if (ln.graph) {          // the kernels of this batch were captured on an earlier use of the slot: one launch
    SPANON_CUDA_TRY(cudaGraphLaunch(ln.graph, ln.compute));
  } else {
    SPANON_TRY(enqueue_compute(kind, fmt, B, n_samples, fb, norm_mean, norm_std, weights, pool_noise, opts, logp_host != nullptr,
                               c, ln.compute));
  }
  if (ln.ev_compute) {
    SPANON_CUDA_TRY(cudaEventRecord(ln.ev_compute, ln.compute));
    SPANON_CUDA_TRY(cudaStreamWaitEvent(ln.d2h, ln.ev_compute, 0));
  }
  const size_t out_elem = (kind == 0 && opts && opts->recon_fp16) ? 2 : 4;     // recon_fp16: half on the wire
  SPANON_CUDA_TRY(cudaMemcpyAsync(recon_host, c.recon, (size_t)B * Tp * F * out_elem, cudaMemcpyDeviceToHost, ln.d2h));
  if (logp_host) SPANON_CUDA_TRY(cudaMemcpyAsync(logp_host, c.logp, (size_t)B * 2 * 4, cudaMemcpyDeviceToHost, ln.d2h));
  return SPANON_OK;
}
}  // namespace

static int anonymize_host(int kind, const void* pcm_host, int fmt, int64_t B, int64_t n_samples,
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
  const Lanes ln{st, st, st, nullptr, nullptr, nullptr};
  SPANON_TRY(enqueue_batch(kind, pcm_host, fmt, B, n_samples, fb, norm_mean, norm_std, weights, pool_noise, opts, recon_host,
                           logp_host, c, ln));
  SPANON_CUDA_TRY(cudaStreamSynchronize(st));
  return SPANON_OK;
}

extern "C" int spanon_anonymize_host(int kind, const float* pcm_host, int64_t B, int64_t n_samples,
                                     const SpanonFbankCfg* fb, const float* norm_mean, const float* norm_std,
                                     const void* weights, const float* pool_noise, const SpanonConvAEOpts* opts,
                                     float* recon_host, float* logp_host, void* dev_workspace,
                                     size_t dev_workspace_bytes, void* stream) {
  return anonymize_host(kind, pcm_host, 4, B, n_samples, fb, norm_mean, norm_std, weights, pool_noise, opts, recon_host,
                        logp_host, dev_workspace, dev_workspace_bytes, stream);
}

extern "C" int spanon_anonymize_host_i16(int kind, const int16_t* pcm_host, int64_t B, int64_t n_samples,
                                         const SpanonFbankCfg* fb, const float* norm_mean, const float* norm_std,
                                         const void* weights, const float* pool_noise, const SpanonConvAEOpts* opts,
                                         float* recon_host, float* logp_host, void* dev_workspace,
                                         size_t dev_workspace_bytes, void* stream) {
  return anonymize_host(kind, pcm_host, 2, B, n_samples, fb, norm_mean, norm_std, weights, pool_noise, opts, recon_host,
                        logp_host, dev_workspace, dev_workspace_bytes, stream);
}

// ------------------------------------------------------------------ streaming pipe
// `depth` batches in flight over three streams (H2D | kernels | D2H), each batch with its own device workspace, so
// the PCIe copies of neighbouring batches hide behind the kernels of the current one.
struct SpanonPipe {
  int kind = 0, depth = 0, device = 0;
  int64_t max_B = 0, n_samples = 0;
  SpanonFbankCfg fb{};
  SpanonConvAEOpts opts{};
  union { SpanonConvAEWeights conv; SpanonFcaeWeights fc; } w;
  const float* norm_mean = nullptr; const float* norm_std = nullptr;
  cudaStream_t h2d = nullptr, compute = nullptr, d2h = nullptr;
  size_t ws_bytes = 0;
  struct Slot { void* ws = nullptr; cudaEvent_t ev_h2d = nullptr, ev_compute = nullptr, ev_done = nullptr; bool busy = false; int64_t prepared_B = -1;
                // CUDA graph of the slot's kernel sequence, valid for exactly this (batch size, sample format, noise, logp)
                cudaGraphExec_t graph = nullptr; int64_t g_B = -1; int g_fmt = 0; const float* g_noise = nullptr; bool g_logp = false; };
  Slot* slots = nullptr;
  int64_t submitted = 0, retired = 0;
};

extern "C" void spanon_pipe_destroy(SpanonPipe* p) {
  if (!p) return;
  if (p->slots) {
    for (int i = 0; i < p->depth; ++i) {
      SpanonPipe::Slot& s = p->slots[i];
      if (s.ev_done && s.busy) cudaEventSynchronize(s.ev_done);
      if (s.ev_h2d) cudaEventDestroy(s.ev_h2d);
      if (s.ev_compute) cudaEventDestroy(s.ev_compute);
      if (s.ev_done) cudaEventDestroy(s.ev_done);
      if (s.graph) cudaGraphExecDestroy(s.graph);
      if (s.ws) cudaFree(s.ws);
    }
    delete[] p->slots;
  }
  if (p->h2d) cudaStreamDestroy(p->h2d);
  if (p->compute) cudaStreamDestroy(p->compute);
  if (p->d2h) cudaStreamDestroy(p->d2h);
  delete p;
}

extern "C" int spanon_pipe_create(SpanonPipe** out, int kind, int64_t max_B, int64_t n_samples, const SpanonFbankCfg* fb,
                                  const float* norm_mean, const float* norm_std, const void* weights,
                                  const SpanonConvAEOpts* opts, int32_t depth) {
  if (!out || !fb || !weights || !norm_mean || !norm_std) return SPANON_ERR_NULL_POINTER;
  if (kind != 0 && kind != 1) return SPANON_ERR_UNSUPPORTED;
  if (max_B <= 0 || n_samples < 0 || fb->hop_length <= 0 || depth < 1 || depth > 8) return SPANON_ERR_BAD_SHAPE;
  SpanonPipe* p = new SpanonPipe();
  p->kind = kind; p->depth = depth; p->max_B = max_B; p->n_samples = n_samples; p->fb = *fb;
  if (opts) p->opts = *opts;
  if (kind == 0) p->w.conv = *(const SpanonConvAEWeights*)weights; else p->w.fc = *(const SpanonFcaeWeights*)weights;
  p->norm_mean = norm_mean; p->norm_std = norm_std;
  p->ws_bytes = spanon_anonymize_workspace_bytes(kind, max_B, n_samples, fb->hop_length, fb->n_mels, &p->opts);
  int rc = SPANON_OK;
  auto ok = [&](cudaError_t e) { if (e != cudaSuccess && rc == SPANON_OK) { set_last_cuda_error("spanon_pipe_create", e); rc = SPANON_ERR_CUDA; } return e == cudaSuccess; };
  ok(cudaGetDevice(&p->device));
  if (p->ws_bytes == 0) rc = SPANON_ERR_BAD_SHAPE;
  if (rc == SPANON_OK) {
    ok(cudaStreamCreateWithFlags(&p->h2d, cudaStreamNonBlocking));
    ok(cudaStreamCreateWithFlags(&p->compute, cudaStreamNonBlocking));
    ok(cudaStreamCreateWithFlags(&p->d2h, cudaStreamNonBlocking));
    p->slots = new SpanonPipe::Slot[depth];
    for (int i = 0; i < depth && rc == SPANON_OK; ++i) {
      SpanonPipe::Slot& s = p->slots[i];
      ok(cudaMalloc(&s.ws, p->ws_bytes));
      ok(cudaEventCreateWithFlags(&s.ev_h2d, cudaEventDisableTiming));
      ok(cudaEventCreateWithFlags(&s.ev_compute, cudaEventDisableTiming));
      ok(cudaEventCreateWithFlags(&s.ev_done, cudaEventDisableTiming));
    }
  }
  if (rc != SPANON_OK) { spanon_pipe_destroy(p); return rc; }
  *out = p;
  return SPANON_OK;
}

extern "C" int spanon_pipe_wait(SpanonPipe* p) {   // retires the oldest batch in flight (no-op when none is)
  if (!p) return SPANON_ERR_NULL_POINTER;
  if (p->retired == p->submitted) return SPANON_OK;
  SpanonPipe::Slot& s = p->slots[p->retired % p->depth];
  SPANON_CUDA_TRY(cudaEventSynchronize(s.ev_done));
  s.busy = false;
  ++p->retired;
  return SPANON_OK;
}

extern "C" int spanon_pipe_drain(SpanonPipe* p) {
  if (!p) return SPANON_ERR_NULL_POINTER;
  while (p->retired < p->submitted) SPANON_TRY(spanon_pipe_wait(p));
  return SPANON_OK;
}

static int pipe_submit(SpanonPipe* p, const void* pcm_host, int fmt, int64_t B, const float* pool_noise,
                       float* recon_host, float* logp_host) {
  if (!p || !pcm_host || !recon_host) return SPANON_ERR_NULL_POINTER;
  if (B <= 0 || B > p->max_B) return SPANON_ERR_BAD_SHAPE;
  if (p->submitted - p->retired >= p->depth) SPANON_TRY(spanon_pipe_wait(p));   // the slot we are about to reuse
  SpanonPipe::Slot& s = p->slots[p->submitted % p->depth];
  char* base = reinterpret_cast<char*>(up256(reinterpret_cast<uintptr_t>(s.ws)));
  SpanonConvAEOpts o = p->opts;
  o.weights_prepared = (p->kind == 0 && s.prepared_B == B) ? 1 : 0;   // this slot already holds the packed weights
  Carve c = carve(base, p->kind, B, p->n_samples, p->fb.hop_length, p->fb.n_mels, &o);
  const void* w = p->kind == 0 ? (const void*)&p->w.conv : (const void*)&p->w.fc;
  // From the second use of a slot at a given batch size on (packed weights in place, every attribute set) the slot's
  // ~29 launches + memsets are captured once into a CUDA graph and replayed with one call: no inter-kernel launch gaps and
  // a submit costs the host three enqueues.  $SPANON_PIPE_GRAPHS=0 turns this off.
  static const bool use_graphs = !(getenv("SPANON_PIPE_GRAPHS") && getenv("SPANON_PIPE_GRAPHS")[0] == '0') && !g_debug_sync;
  const bool want_logp = logp_host != nullptr;
  if (s.graph && !(s.g_B == B && s.g_fmt == fmt && s.g_noise == pool_noise && s.g_logp == want_logp && s.prepared_B == B)) {
    cudaGraphExecDestroy(s.graph);
    s.graph = nullptr;
  }
  if (!s.graph && use_graphs && s.prepared_B == B && !g_profile_on.load()) {
    SpanonConvAEOpts og = p->opts;
    og.weights_prepared = 1;
    cudaGraph_t g = nullptr;
    SPANON_CUDA_TRY(cudaStreamBeginCapture(p->compute, cudaStreamCaptureModeThreadLocal));
    const int rc = enqueue_compute(p->kind, fmt, B, p->n_samples, &p->fb, p->norm_mean, p->norm_std, w, pool_noise, &og,
                                   want_logp, c, p->compute);
    const cudaError_t ce = cudaStreamEndCapture(p->compute, &g);
    if (rc != SPANON_OK) { if (g) cudaGraphDestroy(g); return rc; }
    SPANON_CUDA_TRY(ce);
    const cudaError_t ci = cudaGraphInstantiate(&s.graph, g, 0);
    cudaGraphDestroy(g);
    SPANON_CUDA_TRY(ci);
    s.g_B = B; s.g_fmt = fmt; s.g_noise = pool_noise; s.g_logp = want_logp;
  }
  const Lanes ln{p->h2d, p->compute, p->d2h, s.ev_h2d, s.ev_compute, s.graph};
  SPANON_TRY(enqueue_batch(p->kind, pcm_host, fmt, B, p->n_samples, &p->fb, p->norm_mean, p->norm_std, w, pool_noise, &o,
                           recon_host, logp_host, c, ln));
  SPANON_CUDA_TRY(cudaEventRecord(s.ev_done, p->d2h));
  s.busy = true;
  s.prepared_B = B;
  ++p->submitted;
  return SPANON_OK;
}

extern "C" int spanon_pipe_submit(SpanonPipe* p, const float* pcm_host, int64_t B, const float* pool_noise,
                                  float* recon_host, float* logp_host) {
  return pipe_submit(p, pcm_host, 4, B, pool_noise, recon_host, logp_host);
}

extern "C" int spanon_pipe_submit_i16(SpanonPipe* p, const int16_t* pcm_host, int64_t B, const float* pool_noise,
                                      float* recon_host, float* logp_host) {
  return pipe_submit(p, pcm_host, 2, B, pool_noise, recon_host, logp_host);
}

extern "C" int spanon_pipe_weights_changed(SpanonPipe* p, const void* weights) {
  if (!p) return SPANON_ERR_NULL_POINTER;
  SPANON_TRY(spanon_pipe_drain(p));
  if (weights) { if (p->kind == 0) p->w.conv = *(const SpanonConvAEWeights*)weights; else p->w.fc = *(const SpanonFcaeWeights*)weights; }
  for (int i = 0; i < p->depth; ++i) {
    p->slots[i].prepared_B = -1;
    if (p->slots[i].graph) { cudaGraphExecDestroy(p->slots[i].graph); p->slots[i].graph = nullptr; }
  }
  return SPANON_OK;
}
