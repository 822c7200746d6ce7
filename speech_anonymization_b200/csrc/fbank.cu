// Fused Fbank for B200: PCM -> frames (smem) -> Hamming window -> 400-point FFT ->
// power -> sparse triangular mel projection -> dB, plus the per-utterance maximum the
// top_db clamp needs.  The spectrum never leaves shared memory.
//
// Reference behaviour reproduced (speechbrain/speechbrain/processing/features.py):
//   STFT.forward :161-172      torch.stft(center=True, pad_mode="constant", onesided)
//   spectral_magnitude :347    re^2 + im^2 (power = 1)
//   Filterbank.forward :548    spectrogram @ fbank_matrix
//   _amplitude_to_DB :701-710  10*log10(clamp(x, amin)) - offset, max(., amax - top_db)
//
// Kernel design (fbank400_kernel): one CTA = 8 consecutive frames of one utterance (36 KB of shared memory, six
// CTAs per SM; 16 frames per CTA = 67 KB left three CTAs per SM and the issue slots half empty: 117 -> 104 us).
//   * the 1520 PCM samples the 8 frames cover are loaded once with 128-bit coalesced
//     loads (each sample is used by 2.5 frames);
//   * two real frames are packed into one complex 400-point FFT (frame t -> real part,
//     frame t+1 -> imaginary part) and separated afterwards, so the CTA runs 4 complex
//     FFTs; the FFT is a Stockham autosort radix 4,4,5,5 in shared memory with a
//     twiddle table evaluated per CTA (fp32 sincospif);
//   * power, mel (only the <= 13 non-zero taps of each filter) and log run from shared
//     memory; the only HBM traffic is 4 B/sample in and 4 B/mel-bin out.
#include "common.cuh"
#include <math.h>
#include <stdlib.h>
#include <string.h>

namespace spanon {

namespace {

constexpr int N_FFT = 400;
constexpr int HOP = 160;
constexpr int FRAMES_PER_CTA = 8;
constexpr int PAIRS = FRAMES_PER_CTA / 2;
constexpr int SEG = (FRAMES_PER_CTA - 1) * HOP + N_FFT;  // 1520 samples
constexpr int NT = 256;
constexpr int N_BINS = N_FFT / 2 + 1;  // 201
constexpr int P_STRIDE = 204;          // padded row of the power buffer

struct FbankSmem {
  float2 bufA[PAIRS * N_FFT];
  float2 bufB[PAIRS * N_FFT];
  float2 tw[N_FFT];
  float win[N_FFT];
  float seg[SEG];
  float red[NT / 32];
};

// PCM sample formats: fp32 (the reference's wire format after read_audio) or int16 (the FLAC/WAV payload itself);
// int16 -> fp32 is x / 32768, exactly what torchaudio's normalised load produces (speechbrain_convae_train.py:483-485).
__device__ __forceinline__ float load1(const float* x, int64_t i) { return __ldg(x + i); }
__device__ __forceinline__ float load1(const int16_t* x, int64_t i) { return (float)__ldg(x + i) * (1.f / 32768.f); }
__device__ __forceinline__ float4 load4(const float* x, int64_t i) { return __ldg(reinterpret_cast<const float4*>(x + i)); }
__device__ __forceinline__ float4 load4(const int16_t* x, int64_t i) {
  const uint2 u = __ldg(reinterpret_cast<const uint2*>(x + i));
  const float k = 1.f / 32768.f;
  return make_float4((float)(int16_t)(u.x & 0xffffu) * k, (float)(int16_t)(u.x >> 16) * k,
                     (float)(int16_t)(u.y & 0xffffu) * k, (float)(int16_t)(u.y >> 16) * k);
}

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

__device__ __forceinline__ void dft4(float2* v) {
  float2 t0 = make_float2(v[0].x + v[2].x, v[0].y + v[2].y);
  float2 t1 = make_float2(v[0].x - v[2].x, v[0].y - v[2].y);
  float2 t2 = make_float2(v[1].x + v[3].x, v[1].y + v[3].y);
  // (v1 - v3) * (-i) = (im, -re)
  float2 t3 = make_float2(v[1].y - v[3].y, -(v[1].x - v[3].x));
  v[0] = make_float2(t0.x + t2.x, t0.y + t2.y);
  v[1] = make_float2(t1.x + t3.x, t1.y + t3.y);
  v[2] = make_float2(t0.x - t2.x, t0.y - t2.y);
  v[3] = make_float2(t1.x - t3.x, t1.y - t3.y);
}

__device__ __forceinline__ void dft5(float2* v) {
  const float c1 = 0.30901699437494742f;   // cos(2pi/5)
  const float c2 = -0.80901699437494742f;  // cos(4pi/5)
  const float s1 = 0.95105651629515357f;   // sin(2pi/5)
  const float s2 = 0.58778525229247313f;   // sin(4pi/5)
  float2 a1 = make_float2(v[1].x + v[4].x, v[1].y + v[4].y);
  float2 a2 = make_float2(v[2].x + v[3].x, v[2].y + v[3].y);
  float2 b1 = make_float2(v[1].x - v[4].x, v[1].y - v[4].y);
  float2 b2 = make_float2(v[2].x - v[3].x, v[2].y - v[3].y);
  float2 m1 = make_float2(v[0].x + c1 * a1.x + c2 * a2.x, v[0].y + c1 * a1.y + c2 * a2.y);
  float2 m2 = make_float2(v[0].x + c2 * a1.x + c1 * a2.x, v[0].y + c2 * a1.y + c1 * a2.y);
  float2 n1 = make_float2(s1 * b1.x + s2 * b2.x, s1 * b1.y + s2 * b2.y);
  float2 n2 = make_float2(s2 * b1.x - s1 * b2.x, s2 * b1.y - s1 * b2.y);
  v[0] = make_float2(v[0].x + a1.x + a2.x, v[0].y + a1.y + a2.y);
  // y1 = m1 - i n1 ; y4 = m1 + i n1 ; y2 = m2 - i n2 ; y3 = m2 + i n2 ;  -i n = (n.y, -n.x)
  v[1] = make_float2(m1.x + n1.y, m1.y - n1.x);
  v[4] = make_float2(m1.x - n1.y, m1.y + n1.x);
  v[2] = make_float2(m2.x + n2.y, m2.y - n2.x);
  v[3] = make_float2(m2.x - n2.y, m2.y + n2.x);
}

// One Stockham stage of radix R with Ns = product of the previous radices, over PAIRS
// independent 400-point transforms.
template <int R, int NS>
__device__ __forceinline__ void stockham_stage(const float2* __restrict__ in, float2* __restrict__ out,
                                               const float2* __restrict__ tw, int tid) {
  constexpr int NB = N_FFT / R;               // butterflies per transform
  constexpr int TWS = N_FFT / (NS * R);       // twiddle index stride
  for (int w = tid; w < PAIRS * NB; w += NT) {
    const int f = w / NB;
    const int j = w - f * NB;
    const int k = j % NS;
    const float2* src = in + f * N_FFT;
    float2 v[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      v[r] = src[j + r * NB];
      if (NS > 1 && r > 0) v[r] = cmul(v[r], tw[(r * k * TWS) % N_FFT]);
    }
    if (R == 4) dft4(v); else dft5(v);
    float2* dst = out + f * N_FFT + (j - k) * R + k;
#pragma unroll
    for (int r = 0; r < R; ++r) dst[r * NS] = v[r];
  }
}

template <typename TS>
__global__ void __launch_bounds__(NT)
fbank400_kernel(const TS* __restrict__ pcm, int64_t n_samples, int64_t pcm_stride, int T,
                float* __restrict__ out, int32_t* __restrict__ utt_max, SpanonFbankCfg cfg) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FbankSmem& S = *reinterpret_cast<FbankSmem*>(smem_raw);
  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * FRAMES_PER_CTA;
  const TS* x = pcm + (int64_t)b * pcm_stride;

  // twiddles and the window, once per CTA.  fp32 sincospif (<= 1 ulp) of the correctly rounded -i/200: an fp64
  // evaluation here was ~a quarter of the kernel's instruction issue (ncu: the conversion/MUFU pipe saturated).
  for (int i = tid; i < N_FFT; i += NT) {
    float s, c;
    sincospif((float)(-i) / 200.f, &s, &c);
    S.tw[i] = make_float2(c, s);
    S.win[i] = cfg.window[i];
  }
  // PCM segment: padded sample p = t*HOP + n - N_FFT/2  (center=True, zero padding)
  const int64_t base = (int64_t)t0 * HOP - N_FFT / 2;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(x) & (4 * sizeof(TS) - 1)) == 0) && ((base & 3) == 0);
  for (int i4 = tid; i4 < SEG / 4; i4 += NT) {
    const int64_t s = base + 4 * (int64_t)i4;
    float4 v;
    if (vec_ok && s >= 0 && s + 3 < n_samples) {
      v = load4(x, s);
    } else {
      v.x = (s + 0 >= 0 && s + 0 < n_samples) ? load1(x, s + 0) : 0.f;
      v.y = (s + 1 >= 0 && s + 1 < n_samples) ? load1(x, s + 1) : 0.f;
      v.z = (s + 2 >= 0 && s + 2 < n_samples) ? load1(x, s + 2) : 0.f;
      v.w = (s + 3 >= 0 && s + 3 < n_samples) ? load1(x, s + 3) : 0.f;
    }
    *reinterpret_cast<float4*>(&S.seg[4 * i4]) = v;
  }
  __syncthreads();

  // stage 1: radix 4, Ns = 1, reads the windowed frames straight from the PCM segment
  {
    constexpr int NB = N_FFT / 4;
    for (int w = tid; w < PAIRS * NB; w += NT) {
      const int f = w / NB;
      const int j = w - f * NB;
      const float* fa = S.seg + (2 * f) * HOP;
      const float* fb = fa + HOP;
      float2 v[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int n = j + r * NB;
        const float wn = S.win[n];
        v[r] = make_float2(wn * fa[n], wn * fb[n]);
      }
      dft4(v);
      float2* dst = S.bufA + f * N_FFT + j * 4;
#pragma unroll
      for (int r = 0; r < 4; ++r) dst[r] = v[r];
    }
  }
  __syncthreads();
  stockham_stage<4, 4>(S.bufA, S.bufB, S.tw, tid);
  __syncthreads();
  stockham_stage<5, 16>(S.bufB, S.bufA, S.tw, tid);
  __syncthreads();
  stockham_stage<5, 80>(S.bufA, S.bufB, S.tw, tid);
  __syncthreads();

  // separate the two real transforms and take |.|^2 ; P aliases bufA
  float* P = reinterpret_cast<float*>(S.bufA);
  for (int w = tid; w < PAIRS * N_BINS; w += NT) {
    const int f = w / N_BINS;
    const int k = w - f * N_BINS;
    const float2 zk = S.bufB[f * N_FFT + k];
    const float2 zn = S.bufB[f * N_FFT + ((N_FFT - k) % N_FFT)];
    const float ar = zk.x + zn.x, ai = zk.y - zn.y;   // 2*X_a
    const float br = zk.y + zn.y, bi = zk.x - zn.x;   // 2*X_b (up to a unit factor)
    P[(2 * f) * P_STRIDE + k] = 0.25f * (ar * ar + ai * ai);
    P[(2 * f + 1) * P_STRIDE + k] = 0.25f * (br * br + bi * bi);
  }
  __syncthreads();

  // mel projection + dB + running max
  const int n_mels = cfg.n_mels;
  float lmax = -INFINITY;
  // amplitude -> dB in fp32 like the reference (processing/features.py:701-710); the floor value (silence: exactly
  // -100 dB in the reference's test_features.py:66-77) is evaluated once in fp64
  const float db_floor = (float)((double)cfg.multiplier * log10((double)cfg.amin)) - cfg.db_offset;
  for (int w = tid; w < FRAMES_PER_CTA * n_mels; w += NT) {
    const int fr = w / n_mels;
    const int m = w - fr * n_mels;
    const int t = t0 + fr;
    if (t >= T) continue;
    const int st = __ldg(cfg.mel_start + m);
    const int ln = __ldg(cfg.mel_len + m);
    const float* wv = cfg.mel_w + __ldg(cfg.mel_off + m);
    const float* p = P + fr * P_STRIDE + st;
    float acc = 0.f;
    for (int i = 0; i < ln; ++i) acc = fmaf(p[i], __ldg(wv + i), acc);
    const float db = acc <= cfg.amin ? db_floor : cfg.multiplier * log10f(acc) - cfg.db_offset;
    out[((int64_t)b * T + t) * n_mels + m] = db;
    lmax = fmaxf(lmax, db);
  }
  lmax = warp_max(lmax);
  if ((tid & 31) == 0) S.red[tid >> 5] = lmax;
  __syncthreads();
  if (tid == 0) {
    float m = S.red[0];
#pragma unroll
    for (int i = 1; i < NT / 32; ++i) m = fmaxf(m, S.red[i]);
    if (m > -INFINITY) atomicMax(utt_max + b, float_to_key(m));
  }
}

// ------------------------------------------------------------------ register / warp-shuffle FFT form (the one launched)
// 400 = 25 x 16: a HALF-WARP owns one complex 400-point transform (= two real frames: frame t -> real part, frame t+1 ->
// imaginary part).  Lane n2 (0..15) holds the 25 samples x[16*n1 + n2] in registers:
//   A  25-point DFT over n1 in registers (5 x 5: ten radix-5 butterflies + sixteen constant twiddles W_25^{qr}),
//   B  twiddle W_400^{n2*k1}: 24 per-lane constants evaluated once per thread (a thread runs many transforms),
//   C  16-point DFT over n2 ACROSS the half-warp's lanes: four radix-2 decimation-in-frequency stages of
//      __shfl_xor exchanges (distance 8, 4, 2, 1); lane l ends with k2 = bitrev4(l), i.e. bins k = k1 + 25*k2: a CONTIGUOUS
//      block of 25 bins per lane,
//   D  the two real spectra: X_a[k], X_b[k] from Z[k] and conj(Z[400-k]); Z[400-k] sits in lane l^15, register 25-k1
//      (k1 = 0: register 0 of lane bitrev((16 - k2) & 15)): two more shuffles per value; power into shared memory,
//   E  sparse mel projection + dB from shared memory, per-utterance maximum by warp reduce + atomicMax.
// Shared memory is touched once per sample on the way in (PCM segment) and once per bin on the way out (power); the
// transform itself never leaves the register file (the Stockham form above moves 400 complex values through shared memory
// four times per transform and synchronises the CTA five times).
constexpr int FW_NT = 256;                         // 8 warps = 16 transforms = 32 frames per CTA pass
constexpr int FW_FRAMES = 2 * (FW_NT / 16);        // 32
constexpr int FW_SEG = (FW_FRAMES - 1) * HOP + N_FFT;   // 5360 samples
struct FwSmem {
  float seg[FW_SEG];
  float win[N_FFT];
  float P[FW_FRAMES * P_STRIDE];
};

__device__ __forceinline__ void cmul_c(float2& a, float cr, float ci) {     // a *= (cr + i ci)
  const float x = a.x * cr - a.y * ci;
  a.y = a.x * ci + a.y * cr;
  a.x = x;
}

template <typename TS>
__global__ void __launch_bounds__(FW_NT, 2)
fbank400_warp_kernel(const TS* __restrict__ pcm, int64_t n_samples, int64_t pcm_stride, int T, float* __restrict__ out,
                     int32_t* __restrict__ utt_max, SpanonFbankCfg cfg) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FwSmem& S = *reinterpret_cast<FwSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n2 = lane & 15;                                  // lane inside the half-warp = input residue mod 16
  const int k2 = ((n2 & 1) << 3) | ((n2 & 2) << 1) | ((n2 & 4) >> 1) | ((n2 & 8) >> 3);   // bitrev4: output block of this lane
  const int b = blockIdx.y;
  const TS* x = pcm + (int64_t)b * pcm_stride;
  for (int i = tid; i < N_FFT; i += FW_NT) S.win[i] = cfg.window[i];
  // per-lane constants: W_400^{n2*k1} (fp32 sincospif of the exactly reduced fraction) and the three stage twiddles
  float twr[25], twi[25];
#pragma unroll
  for (int k1 = 0; k1 < 25; ++k1) sincospif(-(float)((n2 * k1) % N_FFT) / 200.f, &twi[k1], &twr[k1]);
  float swr[3], swi[3], sgn[4];
#pragma unroll
  for (int st = 0; st < 4; ++st) {
    const int h = 8 >> st;
    const bool upper = (n2 & h) != 0;
    sgn[st] = upper ? -1.f : 1.f;
    if (st < 3) {
      if (upper) sincospif(-(float)(n2 & (h - 1)) / (float)h, &swi[st], &swr[st]);     // W_{2h}^{l mod h}
      else { swr[st] = 1.f; swi[st] = 0.f; }
    }
  }
  // lane (inside the half-warp) that holds bin 25*((16 - k2) & 15): bitrev4 of that block index
  const int kz = (16 - k2) & 15;
  const int lz = (lane & 16) | ((kz & 1) << 3) | ((kz & 2) << 1) | ((kz & 4) >> 1) | ((kz & 8) >> 3);
  constexpr float C25[25][2] = {
      {1.f, 0.f}, {0.96858316112863108f, -0.24868988716485479f}, {0.87630668004386358f, -0.48175367410171532f},
      {0.72896862742141155f, -0.68454710592868862f}, {0.53582679497899666f, -0.84432792550201508f},
      {0.30901699437494742f, -0.95105651629515357f}, {0.062790519529313374f, -0.99802672842827156f},
      {-0.18738131458572463f, -0.98228725072868872f}, {-0.42577929156507272f, -0.90482705246601958f},
      {-0.63742398974868975f, -0.77051324277578925f}, {-0.80901699437494742f, -0.58778525229247313f},
      {-0.92977648588825146f, -0.36812455268467797f}, {-0.99211470131447788f, -0.12533323356430426f},
      {-0.99211470131447788f, 0.12533323356430426f}, {-0.92977648588825146f, 0.36812455268467797f},
      {-0.80901699437494742f, 0.58778525229247313f}, {-0.63742398974868975f, 0.77051324277578925f},
      {-0.42577929156507272f, 0.90482705246601958f}, {-0.18738131458572463f, 0.98228725072868872f},
      {0.062790519529313374f, 0.99802672842827156f}, {0.30901699437494742f, 0.95105651629515357f},
      {0.53582679497899666f, 0.84432792550201508f}, {0.72896862742141155f, 0.68454710592868862f},
      {0.87630668004386358f, 0.48175367410171532f}, {0.96858316112863108f, 0.24868988716485479f}};   // W_25^j = exp(-2 pi i j / 25)

  const int n_mels = cfg.n_mels;
  const float db_floor = (float)((double)cfg.multiplier * log10((double)cfg.amin)) - cfg.db_offset;
  float lmax = -INFINITY;
  for (int t0 = blockIdx.x * FW_FRAMES; t0 < T; t0 += gridDim.x * FW_FRAMES) {
    __syncthreads();                                         // (previous pass's mel reads of P / seg are done)
    // PCM segment of the pass's 32 frames: padded sample p = t*HOP + n - N_FFT/2 (center=True, zero padding)
    const int64_t base = (int64_t)t0 * HOP - N_FFT / 2;
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(x) & (4 * sizeof(TS) - 1)) == 0) && ((base & 3) == 0);
    for (int i4 = tid; i4 < FW_SEG / 4; i4 += FW_NT) {
      const int64_t sidx = base + 4 * (int64_t)i4;
      float4 v;
      if (vec_ok && sidx >= 0 && sidx + 3 < n_samples) {
        v = load4(x, sidx);
      } else {
        v.x = (sidx + 0 >= 0 && sidx + 0 < n_samples) ? load1(x, sidx + 0) : 0.f;
        v.y = (sidx + 1 >= 0 && sidx + 1 < n_samples) ? load1(x, sidx + 1) : 0.f;
        v.z = (sidx + 2 >= 0 && sidx + 2 < n_samples) ? load1(x, sidx + 2) : 0.f;
        v.w = (sidx + 3 >= 0 && sidx + 3 < n_samples) ? load1(x, sidx + 3) : 0.f;
      }
      *reinterpret_cast<float4*>(&S.seg[4 * i4]) = v;
    }
    __syncthreads();
    const int pair = warp * 2 + (lane >> 4);                 // this half-warp's transform: frames t0 + 2*pair, +1
    {   // (frames past T are computed from the zero padding and never written out: both half-warps must take part in the
        //  full-warp shuffles)
      // ---- load: a[n1] = w[n] * (x_a[n] + i x_b[n]), n = 16*n1 + n2
      float2 a[25];
      const float* fa = S.seg + (2 * pair) * HOP + n2;
#pragma unroll
      for (int n1 = 0; n1 < 25; ++n1) {
        const float wn = S.win[16 * n1 + n2];
        a[n1] = make_float2(wn * fa[16 * n1], wn * fa[16 * n1 + HOP]);
      }
      // ---- A: 25-point DFT over n1 = 5p + q  ->  Y[r + 5s]
#pragma unroll
      for (int q = 0; q < 5; ++q) {
        float2 v[5] = {a[q], a[5 + q], a[10 + q], a[15 + q], a[20 + q]};
        dft5(v);
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          if (r > 0 && q > 0) cmul_c(v[r], C25[(q * r) % 25][0], C25[(q * r) % 25][1]);
          a[5 * r + q] = v[r];                                 // T[r][q]
        }
      }
      float2 y[25];
#pragma unroll
      for (int r = 0; r < 5; ++r) {
        float2 v[5] = {a[5 * r], a[5 * r + 1], a[5 * r + 2], a[5 * r + 3], a[5 * r + 4]};
        dft5(v);
#pragma unroll
        for (int s5 = 0; s5 < 5; ++s5) y[r + 5 * s5] = v[s5];
      }
      // ---- B: twiddle W_400^{n2*k1}
#pragma unroll
      for (int k1 = 1; k1 < 25; ++k1) cmul_c(y[k1], twr[k1], twi[k1]);
      // ---- C: 16-point DFT across the half-warp (radix-2 DIF, distances 8, 4, 2, 1)
#pragma unroll
      for (int st = 0; st < 4; ++st) {
        const int h = 8 >> st;
#pragma unroll
        for (int k1 = 0; k1 < 25; ++k1) {
          const float pr = __shfl_xor_sync(0xffffffffu, y[k1].x, h), pi = __shfl_xor_sync(0xffffffffu, y[k1].y, h);
          float2 tq = make_float2(fmaf(sgn[st], y[k1].x, pr), fmaf(sgn[st], y[k1].y, pi));
          if (st < 3) cmul_c(tq, swr[st], swi[st]);
          y[k1] = tq;
        }
      }
      // ---- D: separate the two real transforms, power spectrum of bins k <= 200 into shared memory
      float* Pa = S.P + (2 * pair) * P_STRIDE;
      float* Pb = Pa + P_STRIDE;
      {   // k1 = 0: bin 25*k2, partner bin 25*((16-k2)&15) in register 0 of lane lz
        const float zr = __shfl_sync(0xffffffffu, y[0].x, lz), zi = __shfl_sync(0xffffffffu, y[0].y, lz);
        const int k = 25 * k2;
        if (k <= 200) {
          const float ar = y[0].x + zr, ai = y[0].y - zi, br = y[0].y + zi, bi = y[0].x - zr;
          Pa[k] = 0.25f * (ar * ar + ai * ai);
          Pb[k] = 0.25f * (br * br + bi * bi);
        }
      }
#pragma unroll
      for (int k1 = 1; k1 <= 12; ++k1) {
        // partner lane l^15 holds block 15-k2: its y[25-k1] = Z[400 - kA], its y[k1] = Z[400 - kB]
        const float nAr = __shfl_xor_sync(0xffffffffu, y[25 - k1].x, 15), nAi = __shfl_xor_sync(0xffffffffu, y[25 - k1].y, 15);
        const float nBr = __shfl_xor_sync(0xffffffffu, y[k1].x, 15), nBi = __shfl_xor_sync(0xffffffffu, y[k1].y, 15);
        const int kA = k1 + 25 * k2, kB = (25 - k1) + 25 * k2;
        if (kA <= 200) {
          const float ar = y[k1].x + nAr, ai = y[k1].y - nAi, br = y[k1].y + nAi, bi = y[k1].x - nAr;
          Pa[kA] = 0.25f * (ar * ar + ai * ai);
          Pb[kA] = 0.25f * (br * br + bi * bi);
        }
        if (kB <= 200) {
          const float ar = y[25 - k1].x + nBr, ai = y[25 - k1].y - nBi, br = y[25 - k1].y + nBi, bi = y[25 - k1].x - nBr;
          Pa[kB] = 0.25f * (ar * ar + ai * ai);
          Pb[kB] = 0.25f * (br * br + bi * bi);
        }
      }
    }
    __syncthreads();
    // ---- E: mel projection + dB (fp32 like the reference, processing/features.py:701-710)
    for (int w = tid; w < FW_FRAMES * n_mels; w += FW_NT) {
      const int fr = w / n_mels;
      const int m = w - fr * n_mels;
      const int t = t0 + fr;
      if (t >= T) continue;
      const int st = __ldg(cfg.mel_start + m);
      const int ln = __ldg(cfg.mel_len + m);
      const float* wv = cfg.mel_w + __ldg(cfg.mel_off + m);
      const float* pp = S.P + fr * P_STRIDE + st;
      float acc = 0.f;
      for (int i = 0; i < ln; ++i) acc = fmaf(pp[i], __ldg(wv + i), acc);
      const float db = acc <= cfg.amin ? db_floor : cfg.multiplier * log10f(acc) - cfg.db_offset;
      out[((int64_t)b * T + t) * n_mels + m] = db;
      lmax = fmaxf(lmax, db);
    }
  }
  lmax = warp_max(lmax);
  if (lane == 0 && lmax > -INFINITY) atomicMax(utt_max + b, float_to_key(lmax));
}

// ---------------------------------------------------------------------------------------------------------------------
// fbank400_rr_kernel: the 400-point transform as TWO register passes, 400 = 16 x 25 (n = 16 n1 + n2, k = k1 + 25 k2):
//   pass A  thread (transform, n2): 25-point DFT over n1 in registers (5 x 5), twiddle W_400^{n2 k1}, one store of 25 values
//   pass B  thread (transform, k1 <= 12): the 16-point DFTs (4 x 4, registers) of rows k1 AND 25 - k1 -- between them they
//           hold every pair Z[k], Z[400 - k], so the two real spectra of the frame pair separate without another exchange
// One trip through shared memory instead of the Stockham kernel's four, all butterfly indices compile-time constants
// (the Stockham form spends more issue slots on index arithmetic than on butterflies: ncu, 47 % ALU vs 28 % FMA pipe).
constexpr int FBANK_DEFAULT_FORM = 0;               // 0 rr, 1 stockham, 2 warp (see fbank_launch)
constexpr int FR_NT = 256;                          // 16 transforms x 16 residues
constexpr int FR_TR = FR_NT / 16;                   // transforms (frame pairs) per CTA pass
constexpr int FR_FRAMES = 2 * FR_TR;                // 32
constexpr int FR_SEG = (FR_FRAMES - 1) * HOP + N_FFT;   // 5360 samples
constexpr int FR_ZR = 18;                           // float2 per k1 row (16 + 2: the 128-bit row reads of pass B are conflict-free)
constexpr int FR_ZS = 25 * FR_ZR;                   // float2 per transform
constexpr int FR_PS = 205;                          // power-spectrum row (odd: lanes = frames read it conflict-free)
constexpr int FR_MAXMEL = 128, FR_MAXTAPS = 512;    // filterbanks staged in shared memory (larger ones: Stockham kernel)
struct FrSmem {
  float seg[FR_SEG];
  float win[N_FFT];
  float2 tw[25 * 16];                               // W_400^{n2 k1} at [k1][n2]
  union {
    float2 Z[FR_TR * FR_ZS];                        // pass A -> pass B: [transform][k1][n2]
    float O[FR_FRAMES * (FR_MAXMEL + 1)];           // mel stage -> coalesced store: [frame][mel] (row n_mels + 1)
  };
  float P[FR_FRAMES * FR_PS];
  float melw[FR_MAXTAPS];
  int mst[FR_MAXMEL], mln[FR_MAXMEL], moff[FR_MAXMEL];
};

// 16-point forward DFT, natural order in and out (radix 4 x 4: n = 4a + b, k = c + 4d)
__device__ __forceinline__ void dft16(float2* v) {
  constexpr float C1 = 0.92387953251128674f, S1 = 0.38268343236508977f, R2 = 0.70710678118654752f;
  float2 t[4][4];
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    float2 u[4] = {v[b], v[4 + b], v[8 + b], v[12 + b]};
    dft4(u);
#pragma unroll
    for (int c = 0; c < 4; ++c) t[b][c] = u[c];
  }
  // W_16^{b c}
  cmul_c(t[1][1], C1, -S1); cmul_c(t[1][2], R2, -R2); cmul_c(t[1][3], S1, -C1);
  cmul_c(t[2][1], R2, -R2); t[2][2] = make_float2(t[2][2].y, -t[2][2].x); cmul_c(t[2][3], -R2, -R2);
  cmul_c(t[3][1], S1, -C1); cmul_c(t[3][2], -R2, -R2); cmul_c(t[3][3], -C1, S1);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float2 u[4] = {t[0][c], t[1][c], t[2][c], t[3][c]};
    dft4(u);
#pragma unroll
    for (int d = 0; d < 4; ++d) v[c + 4 * d] = u[d];
  }
}

template <typename TS>
__global__ void __launch_bounds__(FR_NT, 2)
fbank400_rr_kernel(const TS* __restrict__ pcm, int64_t n_samples, int64_t pcm_stride, int T, float* __restrict__ out,
                   int32_t* __restrict__ utt_max, SpanonFbankCfg cfg, int passes_per_utt, int total_passes) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FrSmem& S = *reinterpret_cast<FrSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31;
  __shared__ int s_ntaps;
  if (tid == 0) s_ntaps = 0;
  const float* melw = S.melw;    // (a filterbank with more than FR_MAXTAPS weights is read from global memory instead)
  auto stage_tables = [&] {      // window, inter-pass twiddles, the sparse mel filterbank
    for (int i = tid; i < N_FFT; i += FR_NT) {
      S.win[i] = cfg.window[i];
      const int k1 = i >> 4, n2 = i & 15;
      float sn, cs;
      sincospif(-(float)(n2 * k1) / 200.f, &sn, &cs);          // n2 * k1 <= 360 < 400: already reduced
      S.tw[i] = make_float2(cs, sn);
    }
    __syncthreads();
    for (int m = tid; m < cfg.n_mels; m += FR_NT) {
      const int st = cfg.mel_start[m], ln = cfg.mel_len[m], off = cfg.mel_off[m];
      S.mst[m] = st; S.mln[m] = ln; S.moff[m] = off;
      atomicMax(&s_ntaps, off + ln);
    }
    __syncthreads();
    const int n_taps = s_ntaps;
    if (n_taps <= FR_MAXTAPS) {
      for (int i = tid; i < n_taps; i += FR_NT) S.melw[i] = cfg.mel_w[i];
    } else {
      melw = cfg.mel_w;
    }
  };
  constexpr float C25[25][2] = {
      {1.f, 0.f}, {0.96858316112863108f, -0.24868988716485479f}, {0.87630668004386358f, -0.48175367410171532f},
      {0.72896862742141155f, -0.68454710592868862f}, {0.53582679497899666f, -0.84432792550201508f},
      {0.30901699437494742f, -0.95105651629515357f}, {0.062790519529313374f, -0.99802672842827156f},
      {-0.18738131458572463f, -0.98228725072868872f}, {-0.42577929156507272f, -0.90482705246601958f},
      {-0.63742398974868975f, -0.77051324277578925f}, {-0.80901699437494742f, -0.58778525229247313f},
      {-0.92977648588825146f, -0.36812455268467797f}, {-0.99211470131447788f, -0.12533323356430426f},
      {-0.99211470131447788f, 0.12533323356430426f}, {-0.92977648588825146f, 0.36812455268467797f},
      {-0.80901699437494742f, 0.58778525229247313f}, {-0.63742398974868975f, 0.77051324277578925f},
      {-0.42577929156507272f, 0.90482705246601958f}, {-0.18738131458572463f, 0.98228725072868872f},
      {0.062790519529313374f, 0.99802672842827156f}, {0.30901699437494742f, 0.95105651629515357f},
      {0.53582679497899666f, 0.84432792550201508f}, {0.72896862742141155f, 0.68454710592868862f},
      {0.87630668004386358f, 0.48175367410171532f}, {0.96858316112863108f, 0.24868988716485479f}};   // W_25^j
  const int n_mels = cfg.n_mels;
  const float db_floor = (float)((double)cfg.multiplier * log10((double)cfg.amin)) - cfg.db_offset;
  const float l2_mult = cfg.multiplier * 0.30102999566398120f;      // multiplier * log10(x) = l2_mult * log2(x)
  bool first = true;
  // persistent CTAs (two per SM) over the flattened (utterance, 32-frame pass) list: the tables are staged once per CTA
  for (int item = blockIdx.x; item < total_passes; item += gridDim.x) {
    const int b = item / passes_per_utt;
    const int t0 = (item - b * passes_per_utt) * FR_FRAMES;
    const TS* x = pcm + (int64_t)b * pcm_stride;
    float lmax = -INFINITY;
    // PCM segment of the pass's 32 frames: padded sample p = t*HOP + n - N_FFT/2 (center=True, zero padding).  The
    // global loads are issued first and land while the tables are staged (first pass) / behind the barrier.
    const int64_t base = (int64_t)t0 * HOP - N_FFT / 2;
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(x) & (4 * sizeof(TS) - 1)) == 0) && ((base & 3) == 0);
    constexpr int NLD = (FR_SEG / 4 + FR_NT - 1) / FR_NT;      // 6
    float4 pre[NLD];
#pragma unroll
    for (int r = 0; r < NLD; ++r) {
      const int i4 = tid + r * FR_NT;
      const int64_t sidx = base + 4 * (int64_t)i4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i4 < FR_SEG / 4) {
        if (vec_ok && sidx >= 0 && sidx + 3 < n_samples) {
          v = load4(x, sidx);
        } else {
          v.x = (sidx + 0 >= 0 && sidx + 0 < n_samples) ? load1(x, sidx + 0) : 0.f;
          v.y = (sidx + 1 >= 0 && sidx + 1 < n_samples) ? load1(x, sidx + 1) : 0.f;
          v.z = (sidx + 2 >= 0 && sidx + 2 < n_samples) ? load1(x, sidx + 2) : 0.f;
          v.w = (sidx + 3 >= 0 && sidx + 3 < n_samples) ? load1(x, sidx + 3) : 0.f;
        }
      }
      pre[r] = v;
    }
    if (first) { stage_tables(); first = false; }
    __syncthreads();                                         // (previous pass's reads of O / seg are done)
#pragma unroll
    for (int r = 0; r < NLD; ++r) {
      const int i4 = tid + r * FR_NT;
      if (i4 < FR_SEG / 4) *reinterpret_cast<float4*>(&S.seg[4 * i4]) = pre[r];
    }
    __syncthreads();
    {   // ---- pass A: transform tr = frames t0 + 2 tr (real part) and + 1 (imaginary part), residue n2
      const int tr = tid >> 4, n2 = tid & 15;
      float2 a[25];
      const float* fa = S.seg + (2 * tr) * HOP + n2;
#pragma unroll
      for (int n1 = 0; n1 < 25; ++n1) {
        const float wn = S.win[16 * n1 + n2];
        a[n1] = make_float2(wn * fa[16 * n1], wn * fa[16 * n1 + HOP]);
      }
      // 25-point DFT over n1 = 5p + q -> y[r + 5s]
#pragma unroll
      for (int q = 0; q < 5; ++q) {
        float2 v[5] = {a[q], a[5 + q], a[10 + q], a[15 + q], a[20 + q]};
        dft5(v);
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          if (r > 0 && q > 0) cmul_c(v[r], C25[(q * r) % 25][0], C25[(q * r) % 25][1]);
          a[5 * r + q] = v[r];
        }
      }
      float2* zrow = S.Z + tr * FR_ZS + n2;
#pragma unroll
      for (int r = 0; r < 5; ++r) {
        float2 v[5] = {a[5 * r], a[5 * r + 1], a[5 * r + 2], a[5 * r + 3], a[5 * r + 4]};
        dft5(v);
#pragma unroll
        for (int s5 = 0; s5 < 5; ++s5) {
          const int k1 = r + 5 * s5;
          if (k1 > 0) { const float2 w = S.tw[k1 * 16 + n2]; cmul_c(v[s5], w.x, w.y); }
          zrow[k1 * FR_ZR] = v[s5];
        }
      }
    }
    __syncthreads();
    if (tid < FR_TR * 13) {   // ---- pass B: item (transform, k1 in 0..12) -> bins k1 + 25 k2 and (25 - k1) + 25 k2, k <= 200
      const int tr = tid / 13, k1 = tid - tr * 13;
      float* Pa = S.P + (2 * tr) * FR_PS;
      float* Pb = Pa + FR_PS;
      float2 A[16];
      {
        const float4* src = reinterpret_cast<const float4*>(S.Z + tr * FR_ZS + k1 * FR_ZR);
#pragma unroll
        for (int i = 0; i < 8; ++i) { const float4 u = src[i]; A[2 * i] = make_float2(u.x, u.y); A[2 * i + 1] = make_float2(u.z, u.w); }
      }
      dft16(A);
      if (k1 == 0) {          // bins 25 k2: Z[400 - 25 k2] = A[(16 - k2) & 15]
#pragma unroll
        for (int k2 = 0; k2 <= 8; ++k2) {
          const float2 y = A[k2], z = A[(16 - k2) & 15];
          const float ar = y.x + z.x, ai = y.y - z.y, br = y.y + z.y, bi = y.x - z.x;
          Pa[25 * k2] = 0.25f * (ar * ar + ai * ai);
          Pb[25 * k2] = 0.25f * (br * br + bi * bi);
        }
      } else {
        float2 Bq[16];
        const float4* src = reinterpret_cast<const float4*>(S.Z + tr * FR_ZS + (25 - k1) * FR_ZR);
#pragma unroll
        for (int i = 0; i < 8; ++i) { const float4 u = src[i]; Bq[2 * i] = make_float2(u.x, u.y); Bq[2 * i + 1] = make_float2(u.z, u.w); }
        dft16(Bq);
#pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) {
          {   // kA = k1 + 25 k2, 400 - kA = (25 - k1) + 25 (15 - k2)
            const float2 y = A[k2], z = Bq[15 - k2];
            const float ar = y.x + z.x, ai = y.y - z.y, br = y.y + z.y, bi = y.x - z.x;
            Pa[k1 + 25 * k2] = 0.25f * (ar * ar + ai * ai);
            Pb[k1 + 25 * k2] = 0.25f * (br * br + bi * bi);
          }
          {   // kB = (25 - k1) + 25 k2, 400 - kB = k1 + 25 (15 - k2)
            const float2 y = Bq[k2], z = A[15 - k2];
            const float ar = y.x + z.x, ai = y.y - z.y, br = y.y + z.y, bi = y.x - z.x;
            Pa[(25 - k1) + 25 * k2] = 0.25f * (ar * ar + ai * ai);
            Pb[(25 - k1) + 25 * k2] = 0.25f * (br * br + bi * bi);
          }
        }
      }
    }
    __syncthreads();
    // ---- mel projection + dB (fp32 like the reference, processing/features.py:701-710).  Lanes = the 32 frames, a warp
    // walks mel filters m = warp, warp + 8, ...: the filter (start, length, weights) is uniform over the warp -- no
    // divergence on the filter length, weights broadcast from shared memory -- and the results go through shared memory so
    // that the pass's 32 x n_mels outputs (contiguous in the output) leave as coalesced stores.
    {
      const int fr = lane, wq = tid >> 5;
      const float* prow = S.P + fr * FR_PS;
      for (int m = wq; m < n_mels; m += FR_NT / 32) {
        const int st = S.mst[m], ln = S.mln[m];
        const float* wv = melw + S.moff[m];
        const float* pp = prow + st;
        float acc = 0.f;
#pragma unroll 1
        for (int i = 0; i + 1 < ln; i += 2) acc = fmaf(pp[i + 1], wv[i + 1], fmaf(pp[i], wv[i], acc));
        if (ln & 1) acc = fmaf(pp[ln - 1], wv[ln - 1], acc);
        float l2;
        asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(fmaxf(acc, cfg.amin)));     // (acc >= amin: never denormal)
        S.O[fr * (n_mels + 1) + m] = acc <= cfg.amin ? db_floor : l2_mult * l2 - cfg.db_offset;
      }
    }
    __syncthreads();
    {   // rows of the staged block -> the pass's contiguous [frames][n_mels] output range, coalesced
      const int nvalid = min(FR_FRAMES, T - t0);
      float* dst = out + ((int64_t)b * T + t0) * n_mels;
      for (int fr = tid >> 5; fr < nvalid; fr += FR_NT / 32) {
        const float* srow = S.O + fr * (n_mels + 1);
        for (int m = lane; m < n_mels; m += 32) {
          const float db = srow[m];
          dst[fr * n_mels + m] = db;
          lmax = fmaxf(lmax, db);
        }
      }
    }
    lmax = warp_max(lmax);
    if (lane == 0 && lmax > -INFINITY) atomicMax(utt_max + b, float_to_key(lmax));
  }
}

// Generic (any n_fft <= 4096, win_length <= n_fft) direct-DFT path: one CTA per frame.
// O(n_fft^2) per frame -- correctness path for non-reference configurations only.
template <typename TS>
__global__ void __launch_bounds__(256)
fbank_generic_kernel(const TS* __restrict__ pcm, int64_t n_samples, int64_t pcm_stride, int T,
                     float* __restrict__ out, int32_t* __restrict__ utt_max, SpanonFbankCfg cfg) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n_fft = cfg.n_fft;
  const int n_bins = n_fft / 2 + 1;
  float2* tw = reinterpret_cast<float2*>(smem_raw);      // [n_fft]
  float* fr = reinterpret_cast<float*>(tw + n_fft);      // [n_fft] windowed frame
  float* P = fr + n_fft;                                 // [n_bins]
  __shared__ float red[8];
  const int tid = threadIdx.x;
  const int t = blockIdx.x, b = blockIdx.y;
  const TS* x = pcm + (int64_t)b * pcm_stride;
  const int wl = cfg.win_length;
  const int woff = (n_fft - wl) / 2;  // torch.stft centres a short window inside n_fft
  for (int i = tid; i < n_fft; i += blockDim.x) {
    double s, c;
    sincospi(-2.0 * (double)i / (double)n_fft, &s, &c);
    tw[i] = make_float2((float)c, (float)s);
    const int64_t p = (int64_t)t * cfg.hop_length + i - n_fft / 2;
    const float v = (p >= 0 && p < n_samples) ? load1(x, p) : 0.f;
    const float w = (i >= woff && i < woff + wl) ? cfg.window[i - woff] : 0.f;
    fr[i] = v * w;
  }
  __syncthreads();
  for (int k = tid; k < n_bins; k += blockDim.x) {
    float re = 0.f, im = 0.f;
    int idx = 0;
    for (int n = 0; n < n_fft; ++n) {
      const float2 w = tw[idx];
      re = fmaf(fr[n], w.x, re);
      im = fmaf(fr[n], w.y, im);
      idx += k; if (idx >= n_fft) idx -= n_fft;
    }
    P[k] = re * re + im * im;
  }
  __syncthreads();
  float lmax = -INFINITY;
  for (int m = tid; m < cfg.n_mels; m += blockDim.x) {
    const int st = cfg.mel_start[m], ln = cfg.mel_len[m];
    const float* wv = cfg.mel_w + cfg.mel_off[m];
    float acc = 0.f;
    for (int i = 0; i < ln; ++i) acc = fmaf(P[st + i], wv[i], acc);
    const float db = (float)((double)cfg.multiplier * log10((double)fmaxf(acc, cfg.amin))) - cfg.db_offset;
    out[((int64_t)b * T + t) * cfg.n_mels + m] = db;
    lmax = fmaxf(lmax, db);
  }
  lmax = warp_max(lmax);
  if ((tid & 31) == 0) red[tid >> 5] = lmax;
  __syncthreads();
  if (tid == 0) {
    float m = red[0];
    for (int i = 1; i < (int)(blockDim.x >> 5); ++i) m = fmaxf(m, red[i]);
    if (m > -INFINITY) atomicMax(utt_max + b, float_to_key(m));
  }
}

__global__ void fill_i32_kernel(int32_t* p, int64_t n, int32_t v) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

__global__ void fbank_clamp_kernel(float* __restrict__ x, const int32_t* __restrict__ utt_max,
                                   int64_t per_utt, float top_db) {
  const int b = blockIdx.y;
  const float floor_db = key_to_float(utt_max[b]) - top_db;
  float* p = x + (int64_t)b * per_utt;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < per_utt;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = fmaxf(p[i], floor_db);
}

}  // namespace

int fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t st) {
  if (n <= 0) return SPANON_OK;
  fill_i32_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(p, n, v);
  return check_launch("fill_i32");
}

}  // namespace spanon

using namespace spanon;

extern "C" int64_t spanon_num_frames(int64_t n_samples, int32_t hop_length) {
  if (hop_length <= 0 || n_samples < 0) return -1;
  return 1 + n_samples / hop_length;
}

namespace {
template <typename TS>
int fbank_launch(const TS* pcm, int64_t B, int64_t n_samples, int64_t pcm_stride, float* out, int32_t* utt_max,
                 const SpanonFbankCfg* cfg, void* stream) {
  if (!cfg) return SPANON_ERR_NULL_POINTER;
  if (B < 0 || n_samples < 0 || pcm_stride < n_samples) return SPANON_ERR_BAD_SHAPE;
  if (cfg->n_fft <= 0 || cfg->hop_length <= 0 || cfg->win_length <= 0 || cfg->win_length > cfg->n_fft ||
      cfg->n_mels <= 0)
    return SPANON_ERR_BAD_SHAPE;
  if (B == 0) return SPANON_OK;
  if (!pcm || !out || !utt_max || !cfg->window || !cfg->mel_start || !cfg->mel_len || !cfg->mel_off ||
      !cfg->mel_w)
    return SPANON_ERR_NULL_POINTER;
  if (B > 65535) return SPANON_ERR_BAD_SHAPE;
  profile_entry();
  cudaStream_t st = as_stream(stream);
  const int64_t T = 1 + n_samples / cfg->hop_length;
  SPANON_TRY(fill_i32(utt_max, B, INT32_MIN, st));
  if (cfg->n_fft == N_FFT && cfg->win_length == N_FFT && cfg->hop_length == HOP) {
    static std::atomic<unsigned long long> attr_done{0};   // per device: the attribute belongs to the device's context
    int dev = 0;
    SPANON_CUDA_TRY(cudaGetDevice(&dev));
    if (first_use_on_device(attr_done, dev)) {
      SPANON_CUDA_TRY(cudaFuncSetAttribute(fbank400_kernel<TS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)sizeof(FbankSmem)));
      SPANON_CUDA_TRY(cudaFuncSetAttribute(fbank400_warp_kernel<TS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)sizeof(FwSmem)));
      SPANON_CUDA_TRY(cudaFuncSetAttribute(fbank400_rr_kernel<TS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)sizeof(FrSmem)));
    }
    // $SPANON_FBANK = rr (two register passes, default) | stockham (shared-memory Stockham, round 1) | warp (register /
    // warp-shuffle FFT, measured slower: DESIGN.md 3.5)
    static const char* form_env = getenv("SPANON_FBANK");
    static const int form = !form_env ? FBANK_DEFAULT_FORM : form_env[0] == 's' ? 1 : form_env[0] == 'w' ? 2 : 0;
    if (form == 0 && cfg->n_mels <= FR_MAXMEL) {
      const int ppu = (int)ceil_div(T, FR_FRAMES);
      const int64_t total = (int64_t)ppu * B;
      const int n_sm = current_device_sms();
      if (n_sm <= 0 || total > INT32_MAX) return SPANON_ERR_CUDA;
      const unsigned grid = (unsigned)std::min<int64_t>(total, 2 * (int64_t)n_sm);
      fbank400_rr_kernel<TS><<<grid, FR_NT, sizeof(FrSmem), st>>>(pcm, n_samples, pcm_stride, (int)T, out, utt_max, *cfg, ppu,
                                                                  (int)total);
      return check_launch("fbank400_kernel");
    }
    if (form == 2) {
      dim3 grid((unsigned)ceil_div(T, FW_FRAMES), (unsigned)B);
      fbank400_warp_kernel<TS><<<grid, FW_NT, sizeof(FwSmem), st>>>(pcm, n_samples, pcm_stride, (int)T, out, utt_max, *cfg);
      return check_launch("fbank400_kernel");
    }
    dim3 grid((unsigned)ceil_div(T, FRAMES_PER_CTA), (unsigned)B);
    fbank400_kernel<TS><<<grid, NT, sizeof(FbankSmem), st>>>(pcm, n_samples, pcm_stride, (int)T, out, utt_max, *cfg);
    return check_launch("fbank400_kernel");
  }
  if (cfg->n_fft > 4096) return SPANON_ERR_UNSUPPORTED;
  const size_t smem = (size_t)cfg->n_fft * (sizeof(float2) + sizeof(float)) + (cfg->n_fft / 2 + 1) * sizeof(float);
  dim3 grid((unsigned)T, (unsigned)B);
  fbank_generic_kernel<TS><<<grid, 256, smem, st>>>(pcm, n_samples, pcm_stride, (int)T, out, utt_max, *cfg);
  return check_launch("fbank_generic_kernel");
}
}  // namespace

extern "C" int spanon_fbank_f32(const float* pcm, int64_t B, int64_t n_samples, int64_t pcm_stride,
                                float* out, int32_t* utt_max, const SpanonFbankCfg* cfg, void* stream) {
  return fbank_launch<float>(pcm, B, n_samples, pcm_stride, out, utt_max, cfg, stream);
}

extern "C" int spanon_fbank_i16(const int16_t* pcm, int64_t B, int64_t n_samples, int64_t pcm_stride,
                                float* out, int32_t* utt_max, const SpanonFbankCfg* cfg, void* stream) {
  return fbank_launch<int16_t>(pcm, B, n_samples, pcm_stride, out, utt_max, cfg, stream);
}

extern "C" int spanon_fbank_clamp(float* feats, const int32_t* utt_max, int64_t B, int64_t T, int64_t n_mels,
                                  float top_db, void* stream) {
  if (B < 0 || T < 0 || n_mels <= 0) return SPANON_ERR_BAD_SHAPE;
  if (B == 0 || T == 0) return SPANON_OK;
  if (!feats || !utt_max) return SPANON_ERR_NULL_POINTER;
  if (B > 65535) return SPANON_ERR_BAD_SHAPE;
  const int64_t per = T * n_mels;
  dim3 grid((unsigned)std::min<int64_t>(ceil_div(per, 256 * 4), 1024), (unsigned)B);
  fbank_clamp_kernel<<<grid, 256, 0, as_stream(stream)>>>(feats, utt_max, per, top_db);
  return check_launch("fbank_clamp_kernel");
}
