// "Tap GEMM": the one operator every ConvAutoencoder / TDNN layer is lowered to.
//
// Activations live channels-last, [utterance][row][channel].  A layer is
//     out[b, m, :] = epilogue( sum_taps  T(in[b, m + shift[tap], :]) @ W[tap] )
// with rows outside [0, rows_in) contributing zeros (= the conv's zero padding, applied
// AFTER the input transform T like in the reference, where the padded tensor is the
// already normalised/activated one).
//   * Conv1d stride 1 (any dilation): shift = dilation*t - padding, W[tap] = w[:, :, t]^T.
//   * Conv1d stride 2 (k5, p2): the input is VIEWED as [rows/2][2*Cin] (same bytes), three
//     taps at shifts -1, 0, +1; the sixth (non-existent) tap slice is zero weights.
//   * ConvTranspose1d stride 2 (k5, p2, op1): the OUTPUT is viewed as [rows][2*Cout]
//     (even | odd phase), three taps at shifts -1, 0, +1, one zero slice.
//   (models/ConvAutoEncoder.py:141-172 for the layer list; :33-43 for the dilated TDNN.)
// Input transform T (consumer-side fusion of the previous layer's InstanceNorm + "GLU"):
//     T(x)[c] = act(x*s1[b,c] + h1[b,c]) * s2[c] + h2[c]      (each part optional)
// Epilogue: + bias, optional ReLU, per-(b,c) sum / sum-of-squares for the NEXT
// InstanceNorm (fp64 atomics), optional affine (folded eval-mode BatchNorm), store.
#pragma once
#include "common.cuh"
#include <cuda_fp16.h>

namespace spanon {

constexpr int MAX_TAPS = 16;

enum OutMode {
  OUT_CHANNELS_LAST = 0,   // [B][rows_out][cout_v] in the activation dtype
  OUT_CHANNELS_FIRST_F32,  // [B][cout_v][rows_out] fp32 (TDNN output -> pooling, reference layout)
  OUT_CH0_F32              // [B][rows_out] fp32, channel 0 only (decoder.8 -> recon)
};

// fp16 planes are chunk-planar, [utterance][channel/8][row][8] (see conv_tc.cu); how a layer's output rows/channels
// map onto the CONSUMER's planes:
enum OutFold {
  OUT_DIRECT = 0,        // plane c/8, row m
  OUT_PAIR_FOLD = 1,     // consumer views row pairs as one row of 2*C channels (stride-2 conv): plane (m&1)*C/8 + c/8, row m>>1
  OUT_PHASE_UNFOLD = 2,  // this layer is a transposed conv computing [even | odd] phases: plane (c % Cout)/8, row 2m + c/Cout
  OUT_PHASE_UNFOLD4 = 3  // the same on ROW PAIRS of the input (columns = [pair row s][phase][Cout]): plane (c % Cout)/8, row 4m + c/Cout
};

struct TapGemmParams {
  const void* in;
  void* out;
  int64_t in_bstride, out_bstride;  // elements per utterance
  int rows_in, rows_out;
  int cin_v, cout_v, cin_mod, cout_mod;
  int ntaps;
  int shift[MAX_TAPS];
  int tap_step;            // shift[t] = shift[0] + t*tap_step when the taps are equally spaced (dilation), else 0
  int tap_kmax[MAX_TAPS];  // input channels (of cin_v) with non-zero weights for this tap (stride-2 view: last tap half)
  int tap_nmax[MAX_TAPS];  // output columns (of cout_v) with non-zero weights (transposed conv: first tap half)
  const void* w;       // packed weights (layout depends on the kernel family)
  const float* t1_scale; const float* t1_shift;  // [B][cin_mod] or null
  int64_t t1_bstride;  // tcgen05 kernel: elements between utterances of t1_* (0 = per-channel constants shared by all)
  int silu;                // activation on load: 0 none, 1 SiLU, 2 ReLU (2: tcgen05 kernel only)
  const float* t2_scale; const float* t2_shift;  // [cin_mod] or null
  const float* bias;   // [cout_mod]  (bias_v != 0: [cout_v], one value per output COLUMN, tcgen05 kernel only)
  int bias_v;
  int relu;
  const float* e_scale; const float* e_shift;    // [cout_mod] or null
  double* stats;       // [B][cout_mod][2] or null
  int out_mode;
  int out_fold;            // OutFold (tcgen05 kernel, OUT_CHANNELS_LAST)
  int out_plane_rows;      // rows of one destination plane
  int in_planar;           // CUDA-core kernel: input is chunk-planar fp16 (decoder.8 reading a tcgen05 layer's output)
  int B;
  // tcgen05 kernel only: the last CTA converts `stats` into the consumer's InstanceNorm scale/shift
  // [B][cout_mod] and zeroes `stats` (fin_counter == nullptr: the caller runs instnorm_finalize_kernel)
  unsigned* fin_counter; const float* fin_gamma; const float* fin_beta; float* fin_scale; float* fin_shift;
  double fin_count;
  // tcgen05 kernel, XF_GEN launches (encoder.2 with encoder.0 fused into its operand production): the transform warps
  // GENERATE the input plane -- SiLU(Conv1d(1 -> 32, k15, p7)(x) + b) of the clamped / normalised / zero-padded feature
  // signal x -- straight into the operand ring instead of reading a stored plane (`in` is unused).
  const float* gen_feats;          // [B][gen_bstride] fp32 features (raw Fbank dB when gen_mean != null)
  int64_t gen_bstride;             // elements between utterances of gen_feats
  int64_t gen_nvalid;              // samples of an utterance that exist (T_valid * F); the rest of [0, L) is zero padding
  int gen_F;                       // mel bins (index of gen_mean / gen_std = sample % gen_F)
  const float* gen_mean; const float* gen_std;   // [F] or null (plain features)
  const int32_t* gen_umax; float gen_topdb;      // per-utterance top_db floor (with gen_mean)
  const float* gen_w; const float* gen_b;        // encoder.0 weight [32][1][15], bias [32]
  // tcgen05 kernel, dataflow (chain) launches: per-tile counters [B * tiles_per_utt] of this layer's OUTPUT (bumped once by
  // each of the 4 epilogue warps) and of its INPUT's producer (flag_in_tpu = the producer's tiles per utterance)
  unsigned* flag_out; const unsigned* flag_in; int flag_in_tpu; unsigned flag_target;
  int out_half;        // tcgen05 kernel, OUT_CH0_F32: store fp16 instead of fp32 (recon_fp16)
  int mma_group;       // tcgen05 kernel: stages per MMA issue round the caller asks for (0/1 = one; 4 where it pays)
  // pair kernel (conv_tc2.cu), tdnn.6: instead of storing the output, ReLU(acc + bias) * e_scale + e_shift is summed (value and
  // square) into the StatisticsPooling bins of the [B][cout][rows_out] -> [B][rows_out][cout] REINTERPRETATION: pool [B][128][2]
  double* pool;
  int ring_cap;        // one-SM tcgen05 kernel: cap on the operand ring's 16-channel stages (0 = TC_MAX_STAGES)
  int pdl;             // tcgen05 kernels: launch as a programmatic dependent of the previous kernel (common.cuh launch_pdl)
  unsigned* err_flag;  // set to 1 if a dataflow wait timed out (results are then invalid; the consumer kernels report it)
  long long* trace;    // debug ($SPANON_TC_TRACE=1): per-CTA clock64 stamps [grid][3 roles][TRACE_N]
  const char* tag;     // profile name of this launch ("tapgemm:<layer>")
};

// fp32 CUDA-core family (conv_fp32.cu).  Packed weights: fp32 [ntaps][cin_v][cout_v].
template <typename TAct>
int launch_tapgemm_simt(const TapGemmParams& p, cudaStream_t st);

// tcgen05 family (conv_tc.cu).  fp16 activations; packed weights: fp16 UMMA core-matrix
// layout [ntaps][cin_v/8][cout_v][8].
int launch_tapgemm_tc(const TapGemmParams& p, int n_sm, cudaStream_t st);
// A chain of up to three 128 -> 128 layers (the TDNN) in one launch, see conv_tc.cu.
int launch_tapgemm_tc_chain(const TapGemmParams* ps, int n, const int* ctas, int rr_group, cudaStream_t st);
bool tapgemm_tc_supported(const TapGemmParams& p);
// Two-SM (cta_group::2, cluster of 2) form for the 128 -> 128 layers (conv_tc2.cu).
int launch_tapgemm_tc2(const TapGemmParams& p, int n_sm, cudaStream_t st);
bool tapgemm_tc2_supported(const TapGemmParams& p);

}  // namespace spanon
