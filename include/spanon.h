/*
 * spanon.h -- C ABI of libspanon.so, the B200 (sm_100a) implementation of the
 * Fbank + anonymizer hot path of viswavi/speech-anonymization.
 *
 * The reference has no FFI layer: its boundary is a set of PyTorch module calls
 * (SURVEY.md section 8b).  Each entry point below replaces the library-kernel
 * sequence behind one of those calls; the reference call site is cited next to
 * it.  Conventions:
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless
 *     its name ends in _host;
 *   - the caller owns all memory (inputs, outputs, workspace); the library
 *     never frees or retains a pointer past the call;
 *   - `stream` is a cudaStream_t passed as void*; calls only enqueue work and
 *     never synchronise the host (except the *_host convenience entry points);
 *   - return value 0 = success, negative = SpanonStatus error code
 *     (spanon_strerror() gives the text).
 */
#ifndef SPANON_H_
#define SPANON_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define SPANON_API __attribute__((visibility("default")))
#else
#define SPANON_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef enum SpanonStatus {
  SPANON_OK = 0,
  SPANON_ERR_BAD_SHAPE = -1,     /* a size is <= 0, inconsistent or unsupported   */
  SPANON_ERR_NULL_POINTER = -2,  /* a required pointer is NULL                    */
  SPANON_ERR_MISALIGNED = -3,    /* a pointer breaks the stated alignment         */
  SPANON_ERR_WORKSPACE = -4,     /* workspace smaller than *_workspace_bytes()    */
  SPANON_ERR_UNSUPPORTED = -5,   /* configuration outside what the kernels cover  */
  SPANON_ERR_CUDA = -6,          /* a CUDA runtime call / kernel launch failed    */
  SPANON_ERR_ARCH = -7,          /* device is not sm_100                          */
  SPANON_ERR_STATIC_LIMIT = -8   /* size beyond the built-in scratch of a *_static entry point */
} SpanonStatus;

SPANON_API const char* spanon_strerror(int status);
/* Version of this ABI (bumped on any signature change). */
SPANON_API int spanon_abi_version(void);
/* Number of kernels launched by this library since process start (bench.py's gpu_launches). */
SPANON_API uint64_t spanon_launch_count(void);
/* Last CUDA error string seen by the library (empty if none). */
SPANON_API const char* spanon_last_cuda_error(void);

/* Kernel-time profile (bench.py's roofline leg).  Between _begin(stream) and _end() every kernel
 * this library launches must be enqueued on `stream`; an event is recorded after each launch and
 * _end() synchronises, attributes each event-to-event interval to the kernel that ended it,
 * aggregates by kernel name into `entries` and returns the number of names (or a negative status). */
typedef struct SpanonKernelTime {
  char name[56];
  uint64_t launches;
  double ms_total;
} SpanonKernelTime;
SPANON_API int spanon_profile_begin(void* stream);
SPANON_API int spanon_profile_end(SpanonKernelTime* entries, int max_entries);

/* ------------------------------------------------------------------------- *
 * Fbank: STFT -> power spectrum -> mel projection -> dB (+ per-utterance max)
 * replaces  Fbank.forward            speechbrain/speechbrain/lobes/features.py:126-143
 *           STFT.forward             speechbrain/speechbrain/processing/features.py:133-188
 *           spectral_magnitude       speechbrain/speechbrain/processing/features.py:327-356
 *           Filterbank.forward       speechbrain/speechbrain/processing/features.py:490-560
 *           Filterbank._amplitude_to_DB                                     :691-712
 * ------------------------------------------------------------------------- */
typedef struct SpanonFbankCfg {
  int32_t n_fft;        /* 400 in every reference config                              */
  int32_t win_length;   /* samples (reference: round(sample_rate/1000*25) = 400)      */
  int32_t hop_length;   /* samples (160)                                              */
  int32_t n_mels;       /* 80 (configs) or 40 (class default)                         */
  float amin;           /* 1e-10                                                      */
  float multiplier;     /* 10 (power spectrogram)                                     */
  float db_offset;      /* multiplier * log10(max(amin, ref_value)) = 0               */
  float top_db;         /* 80                                                         */
  const float* window;  /* [win_length] analysis window (periodic Hamming)            */
  const int32_t* mel_start; /* [n_mels] first STFT bin with a non-zero weight         */
  const int32_t* mel_len;   /* [n_mels] number of consecutive bins                    */
  const int32_t* mel_off;   /* [n_mels] offset of the filter's weights in mel_w       */
  const float* mel_w;       /* packed filter weights                                  */
} SpanonFbankCfg;

/* Number of frames torch.stft(center=True) produces: 1 + n_samples / hop. */
SPANON_API int64_t spanon_num_frames(int64_t n_samples, int32_t hop_length);

/* pcm [B, n_samples] fp32 (row stride pcm_stride elements) -> out [B, T, n_mels]
 * fp32 UNCLAMPED dB, and utt_max [B] int32 order-preserving keys of the per-utterance
 * maximum (over all T frames incl. the zero padded tail, like the reference).
 * Follow with spanon_fbank_clamp() (or let a fused consumer apply the floor). */
SPANON_API int spanon_fbank_f32(const float* pcm, int64_t B, int64_t n_samples, int64_t pcm_stride,
                     float* out, int32_t* utt_max, const SpanonFbankCfg* cfg, void* stream);

/* Same with int16 PCM (the FLAC/WAV payload): sample = x / 32768, exactly what the reference's read_audio
 * (torchaudio.load, normalised; speechbrain_convae_train.py:483-485, speechbrain/dataio/dataio.py read_audio) hands to
 * Fbank, so the result is bit-identical to spanon_fbank_f32 on the converted samples at half the input bytes. */
SPANON_API int spanon_fbank_i16(const int16_t* pcm, int64_t B, int64_t n_samples, int64_t pcm_stride,
                     float* out, int32_t* utt_max, const SpanonFbankCfg* cfg, void* stream);

/* x = max(x, max_b - top_db) in place on [B, T, n_mels]  (features.py:706-710). */
SPANON_API int spanon_fbank_clamp(float* feats, const int32_t* utt_max, int64_t B, int64_t T,
                       int64_t n_mels, float top_db, void* stream);

/* ------------------------------------------------------------------------- *
 * InputNormalization statistics and application
 * replaces  InputNormalization.forward / _compute_current_stats
 *           speechbrain/speechbrain/processing/features.py:998-1148
 *           and the x36 time padding of speechbrain_convae_train.py:62-63
 * ------------------------------------------------------------------------- */
/* Per utterance: n = round_half_even(rel_len[b] * T) valid frames; mean and UNBIASED
 * std over them, std floored at eps (1e-10).  n_valid [B] receives n (bit exact mask).
 * If utt_max != NULL the top_db floor is applied to the values first (fused clamp). */
SPANON_API int spanon_norm_stats(const float* feats, const float* rel_len, int64_t B, int64_t T,
                      int64_t F, const int32_t* utt_max, float top_db, int mean_norm,
                      int std_norm, float eps, float* mean_out, float* std_out,
                      int32_t* n_valid, void* stream);

/* out[b,t,:] = (x - mean)/std for t < T, 0 for T <= t < T_pad.  mean/std are [F]
 * (stat_bstride = 0) or [B, F] (stat_bstride = F).  In place allowed when T_pad == T. */
SPANON_API int spanon_norm_apply(const float* feats, float* out, const float* mean, const float* std,
                      int64_t stat_bstride, const int32_t* utt_max, float top_db,
                      int64_t B, int64_t T, int64_t T_pad, int64_t F, void* stream);

/* ------------------------------------------------------------------------- *
 * FullyConnectedAutoencoder forward (eval-mode BatchNorm)
 * replaces  FullyConnectedAutoencoder.forward  models/FullyConnected.py:147-159
 *           FullyConnSexClassifier.forward     models/FullyConnected.py:90-104
 *           StatisticsPooling.forward          speechbrain/speechbrain/nnet/pooling.py:263-312
 * Weight pointers are the module's own fp32 parameter storages (PyTorch layouts).
 * ------------------------------------------------------------------------- */
typedef struct SpanonBatchNorm {
  const float* weight; const float* bias; const float* running_mean; const float* running_var;
} SpanonBatchNorm;

typedef struct SpanonFcaeWeights {
  int32_t in_dim;                        /* mfcc_feature_dim (80) */
  const float* enc_w[3]; const float* enc_b[3];   /* in->60, 60->40, 40->20 */
  const float* dec_w[3]; const float* dec_b[3];   /* 20->40, 40->60, 60->80 */
  SpanonBatchNorm clf_norm;                        /* BatchNorm1d(20)        */
  const float* init_w[2]; const float* init_b[2]; /* 20->40, 40->40         */
  const float* cls_w[4]; const float* cls_b[4];   /* 80->40, 40->40, 40->20, 20->2 */
  SpanonBatchNorm cls_bn1;                         /* BatchNorm1d(40) after cls 0 */
  SpanonBatchNorm cls_bn6;                         /* BatchNorm1d(20) after cls 2 */
} SpanonFcaeWeights;

SPANON_API size_t spanon_fcae_workspace_bytes(int64_t B, int64_t T);
/* feats [B,T,in_dim] -> recon [B,T,80], logp [B,2].  pool_noise [B,40] is the tensor
 * StatisticsPooling adds to the mean (NULL = none).  If norm_mean != NULL the input is
 * raw Fbank dB and (clamp, (x-mean)/std) is applied on load (fused path). */
SPANON_API int spanon_fcae_fwd(const float* feats, const SpanonFcaeWeights* w, const float* pool_noise,
                    const float* norm_mean, const float* norm_std, const int32_t* utt_max,
                    float top_db, float* recon, float* logp, void* workspace,
                    size_t workspace_bytes, int64_t B, int64_t T, void* stream);

/* FullyConnSexClassifier.forward on its own (models/FullyConnected.py:88-104, eval-mode BatchNorm):
 * z [B,T,20] (the bottleneck) -> logp [B,2].  Reads only the classifier fields of `w` (clf_norm, init_*, cls_*).
 * workspace: spanon_fcae_workspace_bytes(B, T). */
SPANON_API int spanon_fcae_classify(const float* z, const SpanonFcaeWeights* w, const float* pool_noise, float* logp,
                         void* workspace, size_t workspace_bytes, int64_t B, int64_t T, void* stream);

/* spanon_fcae_fwd for callers that must not allocate and have no workspace to pass (code under CUDA-graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as spanon_fcae_fwd minus the workspace: the pooling accumulators
 * ([B][40][2] fp64, this path's only scratch) live in a file-scope `static __device__` array inside the library,
 * sized for SPANON_FCAE_STATIC_MAX_B utterances (160 KB, held by every GPU that loads the library until exit).
 * B > SPANON_FCAE_STATIC_MAX_B returns SPANON_ERR_STATIC_LIMIT.  The call path contains no cudaMalloc / cudaFree /
 * allocating library routine.  All calls on one device share the array (it is cleared on `stream` before use):
 * calls that OVERLAP on the same GPU (different streams, or threads) race on it -- serialise them, or use
 * spanon_fcae_fwd with one workspace per stream. */
#define SPANON_FCAE_STATIC_MAX_B 256
SPANON_API int64_t spanon_fcae_static_max_batch(void);
SPANON_API int spanon_fcae_fwd_static(const float* feats, const SpanonFcaeWeights* w, const float* pool_noise,
                           const float* norm_mean, const float* norm_std, const int32_t* utt_max,
                           float top_db, float* recon, float* logp, int64_t B, int64_t T, void* stream);

/* ------------------------------------------------------------------------- *
 * ConvAutoencoder forward (eval-mode BatchNorm in the TDNN sex classifier)
 * replaces  ConvAutoencoder.forward     models/ConvAutoEncoder.py:178-200
 *           TDNNSexClassifier.forward   models/ConvAutoEncoder.py:57-69
 * ------------------------------------------------------------------------- */
typedef struct SpanonConvAEWeights {
  /* encoder.{0,2,5,8,11}: Conv1d weight [Cout,Cin,k], bias [Cout] */
  const float* enc_w[5]; const float* enc_b[5];
  /* encoder.{3,6,9,12}: InstanceNorm1d affine weight/bias */
  const float* enc_in_w[4]; const float* enc_in_b[4];
  /* decoder.{0,1,4,5,8}: Conv1d [Cout,Cin,k] / ConvTranspose1d [Cin,Cout,k] */
  const float* dec_w[5]; const float* dec_b[5];
  /* decoder.{2,6}: InstanceNorm1d affine */
  const float* dec_in_w[2]; const float* dec_in_b[2];
  /* sex_classifier.norm, .tdnn.{0,3,6} convs, .tdnn.{2,5,8} BatchNorm1d(128) */
  SpanonBatchNorm clf_norm;
  const float* tdnn_w[3]; const float* tdnn_b[3];
  SpanonBatchNorm tdnn_bn[3];
  /* sex_classifier.classify.{0,3,6} Linear, .{2,5} BatchNorm1d */
  const float* cls_w[3]; const float* cls_b[3];
  SpanonBatchNorm cls_bn[2];
} SpanonConvAEWeights;

enum { SPANON_PREC_FP32 = 0, SPANON_PREC_FP16 = 1 };

typedef struct SpanonConvAEOpts {
  int32_t precision;      /* SPANON_PREC_FP32: fp32 activations + CUDA-core FMA (<=1e-4 parity);
                             SPANON_PREC_FP16: fp16 activations/operands + tcgen05 tensor cores with fp32
                             accumulation (<=1e-2 parity bar of the 16-bit mode; bf16 operands cannot hold
                             it on this network, see DESIGN.md "Precision") */
  int32_t chunk_utts;     /* utterances per L2-resident chunk; 0 = library default             */
  int32_t skip_classifier;/* 1 = do not run the sex classifier (logp untouched)                */
  int32_t weights_prepared;/* 1 = `workspace` still holds the packed weights / folded BatchNorms written by an earlier
                             call with the same workspace, weights, B, T, F and opts: skip the per-call preparation */
  int32_t recon_fp16;     /* 1 (SPANON_PREC_FP16 ConvAutoencoder only): `recon` (and `recon_host` of the host-buffer /
                             pipe entry points) is an IEEE half buffer, B*T*F*2 bytes, passed through the float* argument:
                             the 16-bit mode's own precision on the wire, half the device-to-host bytes (ABI v3) */
} SpanonConvAEOpts;

SPANON_API size_t spanon_convae_workspace_bytes(int64_t B, int64_t T, int64_t F, const SpanonConvAEOpts* opts);
/* feats [B,T,F] fp32 with (T*F) % 4 == 0 -> recon [B,T,F] fp32, logp [B,2].
 * pool_noise [B,128] as above.  Fused-normalisation inputs (all NULL for plain feats):
 * feats is then raw Fbank dB [B, T_valid, F], rows >= T_valid are the zero padding. */
SPANON_API int spanon_convae_fwd(const float* feats, const SpanonConvAEWeights* w, const float* pool_noise,
                      const float* norm_mean, const float* norm_std, const int32_t* utt_max,
                      float top_db, int64_t T_valid, float* recon, float* logp,
                      void* workspace, size_t workspace_bytes, int64_t B, int64_t T, int64_t F,
                      const SpanonConvAEOpts* opts, void* stream);

/* ------------------------------------------------------------------------- *
 * ConvReconstruction forward (the "endtoend" anonymizer variant)
 * replaces  ConvReconstruction.forward  models/EndToEnd.py:65-87 up to the reshape at :78
 *           (layers: models/EndToEnd.py:40-54; the external pretrained sex classifier of :57-61,81 is a separate module)
 * Same options / precision modes / workspace rules as spanon_convae_fwd.
 * ------------------------------------------------------------------------- */
typedef struct SpanonConvReconWeights {
  /* encoder.{0,3,6,9,12}: Conv1d weight [Cout,Cin,k] (encoder.9: ConvTranspose1d [Cin,Cout,k]), bias [Cout] */
  const float* conv_w[5]; const float* conv_b[5];
  /* encoder.{1,4,7,10}: InstanceNorm1d affine weight / bias */
  const float* in_w[4]; const float* in_b[4];
} SpanonConvReconWeights;

SPANON_API size_t spanon_convrecon_workspace_bytes(int64_t B, int64_t T, int64_t F, const SpanonConvAEOpts* opts);
/* feats [B,T,F] fp32 with (T*F) % 2 == 0 -> recon [B,T,F] fp32.  Fused-normalisation inputs as in spanon_convae_fwd. */
SPANON_API int spanon_convrecon_fwd(const float* feats, const SpanonConvReconWeights* w, const float* norm_mean,
                         const float* norm_std, const int32_t* utt_max, float top_db, int64_t T_valid,
                         float* recon, void* workspace, size_t workspace_bytes, int64_t B, int64_t T,
                         int64_t F, const SpanonConvAEOpts* opts, void* stream);

/* ------------------------------------------------------------------------- *
 * External x-vector gender classifier on (anonymized) features -- the evaluation half of
 * gender_classifier_recon.yaml / evaluator_inference.yaml
 * replaces  EncoderClassifier.classify_batch_feats  speechbrain/speechbrain/pretrained/interfaces.py:825-829 (:762-776)
 *           Xvector.forward       speechbrain/speechbrain/lobes/models/Xvector.py:95-108  (layers :60-93)
 *           Classifier            speechbrain/speechbrain/lobes/models/Xvector.py:111-172
 *           Conv1d "same"/reflect speechbrain/speechbrain/nnet/CNN.py:349-352,400-403
 *           StatisticsPooling     speechbrain/speechbrain/nnet/pooling.py:263-312
 * Architecture fixed to evaluator_inference.yaml:34-48: TDNN 80->512->512->512->512->1500 (k 5,3,3,1,1; dilation
 * 1,2,3,1,1), LeakyReLU, eval-mode BatchNorm, embedding 128, one DNN block, 2 classes.  fp16 operands on the tcgen05
 * tensor cores with fp32 accumulation.
 * ------------------------------------------------------------------------- */
typedef struct SpanonXvectorWeights {
  const float* tdnn_w[5]; const float* tdnn_b[5];   /* embedding_model blocks.{0,3,6,9,12}.conv  [Cout,Cin,k] / [Cout] */
  SpanonBatchNorm tdnn_bn[5];                        /* blocks.{2,5,8,11,14}.norm                                      */
  const float* emb_w; const float* emb_b;            /* blocks.16.w  [128,3000] / [128]                                */
  SpanonBatchNorm clf_norm;                          /* classifier norm.norm (128)                                     */
  const float* dnn_w; const float* dnn_b;            /* classifier DNN.block_0.linear.w  [128,128] / [128]             */
  SpanonBatchNorm dnn_bn;                            /* classifier DNN.block_0.norm.norm                               */
  const float* out_w; const float* out_b;            /* classifier out.w  [2,128] / [2]                                */
} SpanonXvectorWeights;

SPANON_API size_t spanon_xvector_workspace_bytes(int64_t B, int64_t T);
/* feats [B,T,80] fp32 (already normalised, like classify_batch_feats expects), rel_len [B] relative lengths (NULL = 1.0):
 * pooling runs over round(rel_len * T) frames.  pool_noise [B,1500] is the tensor StatisticsPooling adds to the mean
 * (NULL = none).  emb_out [B,128] (NULL to skip) = the x-vector; logp [B,2] = log posteriors (NULL with emb_out given:
 * Xvector.forward alone, Xvector.py:95-108 -- the classifier fields of `w` are then not read).  weights_prepared = 1: the
 * workspace still holds the packed weights of an earlier call with the same weights. */
SPANON_API int spanon_xvector_classify(const float* feats, const float* rel_len, const SpanonXvectorWeights* w,
                            const float* pool_noise, float* emb_out, float* logp, void* workspace,
                            size_t workspace_bytes, int64_t B, int64_t T, int64_t F, int32_t weights_prepared,
                            void* stream);

/* ------------------------------------------------------------------------- *
 * Training step of the ConvAutoencoder path (fp32, channels-last [utterance][row][channel] tensors)
 * replaces  the autograd graph of ConvAutoencoder.forward in TRAIN mode   models/ConvAutoEncoder.py:178-200
 *           that `(loss / gradient_accumulation).backward()` walks       speechbrain_convae_train.py:238-241
 *           nn.Conv1d / nn.ConvTranspose1d forward + backward            models/ConvAutoEncoder.py:141-172, 33-43
 *           nn.InstanceNorm1d(affine) / train-mode nn.BatchNorm1d + GLU  :146-169, :36-44, :113-120
 *           reshape + StatisticsPooling forward + backward               :61-65, speechbrain/nnet/pooling.py:263-312
 * The host (speech_anonymization_b200/train.py) chains these as torch.autograd.Function nodes, so the reference's
 * requires_grad toggling per epoch parity (speechbrain_convae_train.py:212-235) prunes the same sub-graphs.
 * ------------------------------------------------------------------------- */
typedef struct SpanonTapConv {
  int32_t kind;       /* 0: Conv1d stride 1 (any dilation / padding); 1: Conv1d k5 s2 p2; 2: ConvTranspose1d k5 s2 p2 op1 */
  int32_t cin, cout;  /* real channels: cin in {32, 64, 128}, cout in {1 (kind 0, cin 32, k15, p7), 32 (kind 2), 64, 128} */
  int32_t k, dilation, padding;
} SpanonTapConv;
SPANON_API size_t spanon_tapconv_workspace_bytes(const SpanonTapConv* cv);
/* in [B][rows_in][cin] -> out [B][rows_out][cout] ([B][rows_out] for cout == 1), + bias, optional ReLU; stats != NULL:
 * [B][cout][2] fp64 sums / sums of squares of the stored values are ADDED (clear them first): the statistics of the
 * InstanceNorm / BatchNorm that follows.  w, bias: the module's parameters in PyTorch layout. */
SPANON_API int spanon_tapconv_fwd(const float* in, const float* w, const float* bias, float* out, double* stats, int relu,
                       void* workspace, size_t workspace_bytes, int64_t B, int64_t rows_in, const SpanonTapConv* cv,
                       void* stream);
/* gout = dL/d out -> gin = dL/d in, gw = dL/d w (PyTorch layout), gb = dL/d bias; each may be NULL (frozen / not needed). */
SPANON_API int spanon_tapconv_bwd(const float* in, const float* w, const float* gout, float* gin, float* gw, float* gb,
                       void* workspace, size_t workspace_bytes, int64_t B, int64_t rows_in, const SpanonTapConv* cv,
                       void* stream);
/* encoder.0: Conv1d(1 -> 32, k15, p7) on the flattened feature signal x [B][L] -> out [B][L][32] (raw, no activation);
 * backward: gw [32][1][15], gb [32] (the input signal needs no gradient on this path); scratch: 32 doubles. */
SPANON_API int spanon_conv1_fwd(const float* x, const float* w, const float* bias, float* out, int64_t B, int64_t L, void* stream);
SPANON_API int spanon_conv1_bwd(const float* x, const float* gout, float* gw, float* gb, double* scratch, int64_t B, int64_t L,
                     void* stream);
/* stats [B][C][2] = per (utterance, channel) sum and sum of squares of x [B][rows][C] (C divides 256). */
SPANON_API int spanon_colstats(const float* x, double* stats, int64_t B, int64_t rows, int64_t C, void* stream);
/* out = act(gamma * (z - mean) * rstd + beta) from the sums in `stats`; per_batch = 0: InstanceNorm1d (statistics per
 * utterance and channel, biased variance); per_batch = 1: BatchNorm1d in training mode (statistics per channel over the
 * batch; running_mean / running_var updated in place with `momentum`, unbiased variance, when not NULL).
 * act: 0 none, 1 SiLU ("GLU" of the reference).  gamma == NULL: no normalisation, out = act(z).
 * save_mean / save_rstd ([B][C] or [C]) are kept for the backward pass.
 * SyncBatchNorm (speechbrain/core.py:1156 converts the classifier's BatchNorms under DDP): `stats` then holds the sums of
 * ALL ranks' utterances (all-gathered, stats_B rows) and stats_count their element count per channel; 0 / 0.0 = local. */
SPANON_API int spanon_norm_act_fwd(const float* z, const double* stats, const float* gamma, const float* beta, int per_batch,
                        int act, float eps, float* running_mean, float* running_var, float momentum, float* out,
                        float* save_mean, float* save_rstd, int64_t B, int64_t rows, int64_t C, int64_t stats_B,
                        double stats_count, void* stream);
/* gz = dL/dz, ggamma / gbeta (may be NULL).  relu_mask = 1: z is a stored ReLU output and gz is the gradient in FRONT of
 * that ReLU (zero where z <= 0).  scratch: [B][C][2] fp64.
 * phase 0: everything.  SyncBatchNorm: phase 1 = the two reductions (scratch[0][C][2] = this rank's sums, ggamma / gbeta
 * written), the caller all-reduces scratch[0], phase 2 = the element-wise pass with count = the global element count. */
SPANON_API int spanon_norm_act_bwd(const float* z, const float* gout, const float* gamma, const float* beta, const float* save_mean,
                        const float* save_rstd, int per_batch, int act, int relu_mask, float* gz, float* ggamma,
                        float* gbeta, double* scratch, int64_t B, int64_t rows, int64_t C, int phase, double count,
                        void* stream);
/* x [B][rows][128] channels-last -> out [B][256] = [mean + noise | unbiased std + 1e-5] over the columns of the reference's
 * reshape(B, L, C) of the channels-first tensor; save [B][256] = [mean | std] for the backward; scratch [B][128][2] fp64. */
SPANON_API int spanon_pool_reinterp_fwd(const float* x, const float* noise, float* out, float* save, double* scratch, int64_t B,
                             int64_t rows, int64_t C, void* stream);
SPANON_API int spanon_pool_reinterp_bwd(const float* x, const float* save, const float* gout, float* gx, int64_t B, int64_t rows,
                             int64_t C, void* stream);

/* ------------------------------------------------------------------------- *
 * SpecAugment on features (TRAIN stage of the anonymizer)
 * replaces  SpecAugment.time_warp / mask_along_axis   speechbrain/speechbrain/lobes/augment.py:127-200
 *           (called at speechbrain_convae_train.py:64-67, configured by convae.yaml:273-283)
 * The random draws are the caller's (the Python mirror makes the reference's torch.randint calls in the same order).
 * ------------------------------------------------------------------------- */
/* time warp: out[:, :w] = bicubic(in[:, :c] -> w rows), out[:, w:] = bicubic(in[:, c:] -> T - w rows), align_corners. */
SPANON_API int spanon_specaug_warp(const float* in, float* out, int64_t B, int64_t T, int64_t F, int32_t c, int32_t w, void* stream);
/* in place: x[b, idx in [pos, pos+len) of any of the n_mask masks] = 0 (replace_with_zero) or x.mean(); axis 1 = time,
 * 2 = feature; mask_len / mask_pos int32 [B][n_mask] on the device; scratch: one double. */
SPANON_API int spanon_specaug_mask(float* x, int64_t B, int64_t T, int64_t F, int32_t axis, int32_t n_mask, const int32_t* mask_len,
                        const int32_t* mask_pos, int32_t replace_with_zero, double* scratch, void* stream);

/* ------------------------------------------------------------------------- *
 * Loss statistics
 * replaces  CosineSimilarityLoss.forward   utils/cosine_similarity_loss.py:53-56
 *           _pairwise_dists / ClusterMI._mutual_information  utils/ClusterMI.py:12-63,88-121
 * ------------------------------------------------------------------------- */
/* x1,x2 [B,T,D] -> loss[0] = sum_{b,t}(1 - cos(x1,x2)) / T ; optional per-row 1-cos in
 * row_loss [B*T] (NULL to skip). eps = 1e-6 in the reference. */
SPANON_API int spanon_cosine_loss(const float* x1, const float* x2, int64_t B, int64_t T, int64_t D,
                       float eps, float* loss, float* row_loss, void* stream);
/* Gradients of the above wrt x1 and x2 scaled by grad_out[0] (either may be NULL). */
SPANON_API int spanon_cosine_loss_bwd(const float* x1, const float* x2, const float* grad_out, int64_t B,
                           int64_t T, int64_t D, float eps, float* gx1, float* gx2, void* stream);

/* Batched kNN mutual-information statistics.  X [Ntot, D]; for each of `n_sets` index
 * sets idx[s, 0..N) (NULL = one set 0..N-1) with labels y[Ntot] (int32):
 *   dists [n_sets, N, N] cosine distances (NULL to skip the store),
 *   m_i   [n_sets, N]    neighbour counts  (ClusterMI.py:115),
 *   mi    [n_sets]       MI in bits        (ClusterMI.py:120-121). */
SPANON_API int spanon_knn_mi_stats(const float* X, const int32_t* y, const int32_t* idx, int64_t n_sets,
                        int64_t N, int64_t D, int32_t n_classes, int32_t k, float* dists,
                        int32_t* m_i, float* mi, void* stream);

/* ------------------------------------------------------------------------- *
 * End-to-end convenience entry point with HOST buffers (bench.py's e2e leg and
 * the ctypes binding shown in INTEGRATION.md): copies pcm H2D, runs
 * Fbank -> clamp -> (x-mean)/std -> pad -> anonymizer, copies recon/logp D2H and
 * synchronises.  kind: 0 = ConvAutoencoder, 1 = FullyConnectedAutoencoder.
 * `weights` points to a SpanonConvAEWeights / SpanonFcaeWeights of DEVICE pointers.
 * dev_workspace must hold spanon_anonymize_workspace_bytes().
 * ------------------------------------------------------------------------- */
SPANON_API size_t spanon_anonymize_workspace_bytes(int kind, int64_t B, int64_t n_samples, int32_t hop_length,
                                        int32_t n_mels, const SpanonConvAEOpts* opts);
SPANON_API int spanon_anonymize_host(int kind, const float* pcm_host, int64_t B, int64_t n_samples,
                          const SpanonFbankCfg* fb, const float* norm_mean, const float* norm_std,
                          const void* weights, const float* pool_noise, const SpanonConvAEOpts* opts,
                          float* recon_host, float* logp_host, void* dev_workspace,
                          size_t dev_workspace_bytes, void* stream);

/* int16 PCM on the wire (see spanon_fbank_i16): half the host-to-device bytes. */
SPANON_API int spanon_anonymize_host_i16(int kind, const int16_t* pcm_host, int64_t B, int64_t n_samples,
                          const SpanonFbankCfg* fb, const float* norm_mean, const float* norm_std,
                          const void* weights, const float* pool_noise, const SpanonConvAEOpts* opts,
                          float* recon_host, float* logp_host, void* dev_workspace,
                          size_t dev_workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * Streaming form of the above for a long-running service: up to `depth` batches in
 * flight over three streams (H2D | kernels | D2H), each batch in its own device
 * workspace (allocated by the pipe), weights packed once per workspace.  The PCIe
 * copies of neighbouring batches hide behind the kernels of the current one.
 *   submit: enqueues one batch and returns; blocks only to retire the oldest batch
 *           when `depth` are already in flight.  Host buffers must stay valid (and
 *           should be pinned) until the batch is retired.
 *   wait:   retires the OLDEST batch in flight (its recon/logp are then complete).
 *   drain:  retires everything.
 *   weights_changed: drain, optionally replace the weight pointers, re-pack on next use.
 * norm_mean / norm_std / weight tensors are device pointers the caller keeps alive.
 * ------------------------------------------------------------------------- */
typedef struct SpanonPipe SpanonPipe;
SPANON_API int spanon_pipe_create(SpanonPipe** pipe, int kind, int64_t max_B, int64_t n_samples,
                       const SpanonFbankCfg* fb, const float* norm_mean, const float* norm_std,
                       const void* weights, const SpanonConvAEOpts* opts, int32_t depth);
SPANON_API int spanon_pipe_submit(SpanonPipe* pipe, const float* pcm_host, int64_t B, const float* pool_noise,
                       float* recon_host, float* logp_host);
SPANON_API int spanon_pipe_submit_i16(SpanonPipe* pipe, const int16_t* pcm_host, int64_t B, const float* pool_noise,
                       float* recon_host, float* logp_host);
SPANON_API int spanon_pipe_wait(SpanonPipe* pipe);
SPANON_API int spanon_pipe_drain(SpanonPipe* pipe);
SPANON_API int spanon_pipe_weights_changed(SpanonPipe* pipe, const void* weights);
SPANON_API void spanon_pipe_destroy(SpanonPipe* pipe);

#ifdef __cplusplus
}
#endif
#endif /* SPANON_H_ */
