// Weight packing shared by the forward (convae.cu) and the training kernels (train.cu): how a PyTorch Conv1d /
// ConvTranspose1d weight maps onto the [tap][cin_v][cout_v] slices of a tap GEMM (tapgemm.cuh).
#pragma once
#include <stdint.h>

namespace spanon {

enum PackMode { PACK_CONV_S1 = 0, PACK_CONV_S2 = 1, PACK_CONVT_S2 = 2, PACK_TOEPLITZ = 3, PACK_COMPOSED = 4, PACK_CONV_S1_PAIR = 5 };

// Source value of packed element (tap, civ, cov); cin/cout/k are the REAL layer sizes.
__device__ __forceinline__ float packed_src(const float* __restrict__ w, int mode, int cin, int cout, int k,
                                            int tap, int civ, int cov) {
  if (mode == PACK_CONV_S1) {            // Conv1d weight [cout][cin][k]
    if (cov >= cout) return 0.f;         // padded output channels (decoder.8)
    return w[((int64_t)cov * cin + civ) * k + tap];
  } else if (mode == PACK_CONV_S2) {     // input pair rows: civ = half*cin + ci, t = 2*tap + half
    const int half = civ / cin, ci = civ - half * cin;
    const int t = 2 * tap + half;
    if (t >= k) return 0.f;
    return w[((int64_t)cov * cin + ci) * k + t];
  } else if (mode == PACK_TOEPLITZ) {    // Conv1d weight [1][cin][k] as a [cin] x [k taps (+ zero columns)] matrix
    if (cov >= k) return 0.f;
    return w[(int64_t)civ * k + cov];
  } else if (mode == PACK_CONV_S1_PAIR) {
    // stride-1 Conv1d (padding (k-1)/2) on ROW PAIRS: view row m = rows (2m, 2m+1), civ = par_i*cin + ci, cov = par_o*cout + co,
    // tap p = pair shift p - 1.  y[2m + par_o] takes x[2(m + p - 1) + par_i] with weight index kk = 2(p-1) + par_i - par_o + (k-1)/2
    const int pi = civ / cin, ci = civ - pi * cin, po = cov / cout, co = cov - po * cout;
    const int kk = 2 * (tap - 1) + pi - po + (k - 1) / 2;
    if (kk < 0 || kk >= k) return 0.f;
    return w[((int64_t)co * cin + ci) * k + kk];
  } else {                               // ConvTranspose1d weight [cin][cout][k]; cov = phase*cout + co
    const int phase = cov / cout, co = cov - phase * cout;
    const int t = 4 - 2 * tap + phase;   // shifts -1,0,+1 <-> taps (4|-), (2|3), (0|1)
    if (t >= k) return 0.f;
    return w[((int64_t)civ * cout + co) * k + t];
  }
}


// Flat index of that source element in the PyTorch weight tensor, or -1 for a structural zero (the inverse map the
// weight-gradient unpacking uses: every weight element appears in exactly one packed position).
__device__ __forceinline__ int64_t packed_src_index(int mode, int cin, int cout, int k, int tap, int civ, int cov) {
  if (mode == PACK_CONV_S1) {
    if (cov >= cout) return -1;
    return ((int64_t)cov * cin + civ) * k + tap;
  } else if (mode == PACK_CONV_S2) {
    const int half = civ / cin, ci = civ - half * cin;
    const int t = 2 * tap + half;
    if (t >= k) return -1;
    return ((int64_t)cov * cin + ci) * k + t;
  } else if (mode == PACK_CONVT_S2) {
    const int phase = cov / cout, co = cov - phase * cout;
    const int t = 4 - 2 * tap + phase;
    if (t >= k) return -1;
    return ((int64_t)civ * cout + co) * k + t;
  }
  return -1;
}

}  // namespace spanon
