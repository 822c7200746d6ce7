// tcgen05 (5th-gen tensor core) implementation of the tap GEMM (tapgemm.cuh) for sm_100a.
//
// fp16 activation planes are stored CHUNK-PLANAR in global memory: [utterance][channel/8][row][8 ch], i.e. for every
// 8-channel chunk all rows are contiguous at a 16-byte pitch.  That is exactly the UMMA no-swizzle K-major operand
// layout (8 rows x 16 B core matrices, SBO = 128 B), so a (128 + halo)-row slab of one chunk is ONE contiguous
// cp.async.bulk copy into shared memory, and a conv tap is the same slab read at start_address + 16 * shift.
//
// One persistent CTA per SM (480 threads) walks a CONTIGUOUS run of (utterance, 128-row tile) pairs of one layer; the
// operand ring is cut into 16-channel stages (two 8-channel chunks x slab rows = one K=16 MMA step per tap):
//
//   warps 13,14  LOADERS     loader l owns the stages it == l (mod 2): per stage two bulk copies global -> smem issued by
//                            one elected lane of the converged warp (operands stay in uniform registers), completion on the
//                            stage's `raw` mbarrier (expect_tx).  They run as far ahead as the ring (8 or 16 stages) allows.
//   warps 4..11  TRANSFORM   warp w owns the stages it == w (mod 8), in place in shared memory, all packed fp16:
//                            InstanceNorm affine (HFMA2) -> SiLU (tanh.approx.f16x2) | ReLU -> folded BatchNorm of the
//                            PRODUCER layer; rows outside the utterance (= the conv's zero padding of the already activated
//                            tensor) are zeroed in an utterance's first / last tile; fence.proxy.async, `full` mbarrier.
//                            The per-utterance constants are reloaded only when the utterance changes.  One kernel
//                            instantiation per transform mode (XF_*): the roles share a small instruction cache.
//   warp 12      MMA ISSUER  converged warp, elect.sync-predicated tcgen05.mma (M=128, N=cout_v, K=16, kind::f16, fp32
//                            accumulate in TMEM), one per stage and tap.  All of the layer's packed weights are resident in
//                            shared memory.  (An M=128 N=128 K=16 UMMA reads 8 KB of operands from shared memory = 64 cycles
//                            at 128 B/clk, exactly its math time; N=64 takes 48: tools/ubench/umma_rate.cu.)
//   warps 0..3   EPILOGUE    tcgen05.ld of the accumulator (double-buffered in TMEM) -> + bias, ReLU, folded BN -> fp16 ->
//                            16-byte stores straight into the CONSUMER's chunk-planar view (direct, pair-folded for a
//                            stride-2 consumer, phase-unfolded after a transposed conv): a warp's 32 rows of one chunk are
//                            512 contiguous bytes.
// InstanceNorm statistics of the output (XF_STATS instantiations, eight epilogue warps): reduced in the epilogue from the
// accumulator values as they are stored (a butterfly reduce-scatter over the warp's 32 rows per 32-column block), kept
// in registers per lane = column until the utterance changes, then added to `stats` with fp64 atomics.  That costs
// +25..+48 us per producing launch on B200 against 72 us for a separate pass over the stored plane
// (instnorm_stats_kernel, still used by the SIMT path and selectable with $SPANON_FUSED_STATS=0).
//
// 16-bit operands are fp16, not bf16: tools/precision_study.py shows bf16 operands + bf16 planes cannot hold the 1e-2
// parity bar on this network (1.0-1.3e-2); fp16 gives 1.5e-3 at the same tensor-core rate.
#include "tapgemm.cuh"
#include "tc_prims.cuh"
#include <cuda_fp16.h>
#include <algorithm>
#include <stdlib.h>
#include <vector>

namespace spanon {
namespace {

using namespace tc;
constexpr int TC_XF_WARPS = 8;            // transform warps; warp w owns the stages it == w (mod 8), one stage in flight each
constexpr int TC_LOAD_WARPS = 2;          // loader l issues the stages it == l (mod 2): a lone warp needs ~300 cycles per stage
// Warp roles, E = epilogue warps (4: one per TMEM lane quarter; 8 when the epilogue also reduces the InstanceNorm
// statistics: two per quarter, each half of the columns).  A warp runs on SM sub-partition warp % 4.  The issuer's
// instruction stream (one thread, every instruction dependent on the last) is the kernel's critical path, so it gets
// sub-partition 3 almost to itself: only the two loaders (a few instructions per stage) and one epilogue warp (TMEM lane
// quarter 3 has to live there) share its scheduler; the eight transform warps run on sub-partitions 0..2.
//   warps [0, E)            epilogue
//   warps E+3, E+7          loaders
//   warp  E+11              MMA issuer
//   warp  E+10              idle
//   other warps of [E, E+10) transform (E+0,1,2,4,5,6,8,9)
__host__ __device__ constexpr int tc_epi_warps(int xf);
__host__ __device__ constexpr int tc_threads(int xf) { return (tc_epi_warps(xf) + 12) * 32; }
constexpr int TC_EPI_THREADS = 128;       // Toeplitz epilogue (always four warps)
constexpr int TC_STAGE_CH = 16;           // input channels per A stage (one K=16 MMA step per tap)
constexpr int TC_MAX_STAGES = 32;
// operand transform of a launch (compile-time: one lean instruction stream per mode, the roles share the I-cache)
enum { XF_IDENT = 0, XF_SILU = 1, XF_RELU = 2, XF_T2 = 4, XF_STATS = 8, XF_GEN = 16, XF_DYN = 32, XF_FLOW = 64, XF_EPI8 = 128 };
constexpr int TC_GEN_SIG = 320;           // fp16 samples of one tile's signal window (XF_GEN)
constexpr int TC_GEN_SLOTS = 4;           // signal windows in flight (filled by the loader warps, read by the transform warps)
constexpr int TC_GEN_MAXF = 128;          // mel bins the per-warp (mean, 1/std) table holds
__host__ __device__ constexpr int tc_epi_warps(int xf) { return (xf & (XF_STATS | XF_EPI8)) ? 8 : 4; }
constexpr size_t TC_SMEM_LIMIT = 232448;  // 227 KB opt-in maximum per CTA
constexpr int TRACE_N = 128;
constexpr int TC_TOEP_TAPS = 15;          // decoder.8 as a Toeplitz GEMM: 15 taps -> 16 output columns
constexpr int TC_TOEP_STEP = TC_M - (TC_TOEP_TAPS - 1);   // complete outputs per 128-row tile
// stages the MMA issuer handles per round (compile-time: the issue loop must stay in uniform registers); divides the
// stages per tile of every layer of that width (N = 16: 2 stages; N = 64: 4; N = 128: 8)
__host__ __device__ constexpr int tc_mma_group(int cout_v) { return cout_v == 16 ? 2 : 4; }
__host__ __device__ constexpr int TC_TMEM_COLS(int cout_v) { return 2 * cout_v < 32 ? 32 : 2 * cout_v; }

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_THREADS) : "memory"); }

// SiLU of two values given hh = x/2 as packed fp16:  x*sigmoid(x) = hh + hh*tanh(hh)  (one MUFU per pair)
__device__ __forceinline__ __half2 silu_from_half(__half2 hh) {
  uint32_t t, in = *reinterpret_cast<uint32_t*>(&hh);
  asm("tanh.approx.f16x2 %0, %1;" : "=r"(t) : "r"(in));
  return __hfma2(hh, *reinterpret_cast<__half2*>(&t), hh);
}

struct TcLayout {        // byte offsets inside dynamic shared memory
  uint32_t w, a, epi, red, sig, bars, total;
  uint32_t w_bytes, stage_bytes, lbo_a;
  int n_stages, slab;
};

constexpr int TC_ST_PITCH = 17;              // fp32 row pitch of the statistics transpose buffer ([32 rows][16 columns] per warp)
__host__ __device__ inline TcLayout tc_layout(int ntaps, int cin_v, int cout_v, int slab_used, bool stats, bool gen = false,
                                              int ring_cap = 0) {
  TcLayout L;
  L.slab = (slab_used + 7) / 8 * 8 + 1;              // == 1 (mod 8): chunks start 16 B apart modulo 128 B
  L.lbo_a = (uint32_t)L.slab * 16u;
  L.stage_bytes = (uint32_t)(TC_STAGE_CH / 8) * L.lbo_a;
  L.w_bytes = (uint32_t)ntaps * cin_v * cout_v * 2u;
  L.w = 0;
  L.epi = L.w + L.w_bytes;                            // bias | e_scale | e_shift   [3][cout_v] fp32
  L.red = L.epi + 3u * cout_v * 4u;                   // Toeplitz epilogue: 2 x [128][17] fp32 copies of the tile;
                                                      // statistics: one [32][17] fp32 transpose buffer per epilogue warp
  L.sig = L.red + (cout_v == 16 ? 4u * TC_M * 17u * 4u : stats ? 8u * 32u * TC_ST_PITCH * 4u : 0u);
  // XF_GEN: ring of TC_GEN_SLOTS signal windows (two fp16 copies each, the second shifted by one sample) + one
  // (mean, 1/std) table per loader warp
  L.bars = L.sig + (gen ? (uint32_t)TC_GEN_SLOTS * 2u * TC_GEN_SIG * 2u + (uint32_t)TC_LOAD_WARPS * 2u * TC_GEN_MAXF * 4u : 0u);
  L.a = (L.bars + 1024u + 127u) / 128u * 128u;
  const size_t avail = TC_SMEM_LIMIT > L.a ? TC_SMEM_LIMIT - L.a : 0;
  int ns = (int)(avail / L.stage_bytes);
  const int cap = ring_cap >= TC_XF_WARPS && ring_cap < TC_MAX_STAGES ? ring_cap : TC_MAX_STAGES;
  L.n_stages = ns > cap ? cap : ns;
  // a multiple of the number of transform warps: every ring slot then belongs to ONE warp, so each slot's mbarriers
  // are waited on in order by a single consumer (a waiter two phases behind would alias the parity)
  L.n_stages = L.n_stages / TC_XF_WARPS * TC_XF_WARPS;
  L.total = L.a + (uint32_t)L.n_stages * L.stage_bytes;
  return L;
}

// barrier slots (8 B each): [0,32) raw (bulk copy landed), [32,64) full (transformed), [64,96) empty (MMAs read it),
// [96,98) tmem_full, [98,100) tmem_empty, 100 weights; 102 = TMEM base pointer; 104.. XF_GEN signal ring
enum { B_RAW = 0, B_FULL = 32, B_EMPTY = 64, B_TFULL = 96, B_TEMPTY = 98, B_W = 100, B_TPTR = 102, B_SFULL = 104, B_SEMPTY = 108 };

// One CTA's work on one layer.  cta_idx / n_cta: this CTA's index among the CTAs working on the layer.  rr_group == 0: the
// CTA walks a CONTIGUOUS run of tiles; rr_group = G > 0: groups of G consecutive tiles are dealt round-robin to the
// layer's CTAs (multi-layer dataflow launches: every layer then advances through the tile sequence at the same pace).
template <int COUT_V, int XF, int MG>
__device__ __forceinline__ void tc_body(const TapGemmParams& p, int halo_lo, int slab_used, int tiles_per_utt, int total_tiles,
                                        int tile_step, int cta_idx, int n_cta, int rr_group, int mma_group) {
  extern __shared__ __align__(128) unsigned char smem[];
  const TcLayout L = tc_layout(p.ntaps, p.cin_v, COUT_V, slab_used, (XF & XF_STATS) != 0, (XF & XF_GEN) != 0, p.ring_cap);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int E = tc_epi_warps(XF), TC_MMA_WARP = E + 11;
  const int wd = warp - E;                                   // role index among the non-epilogue warps
  const bool is_loader = wd == 3 || wd == 7, is_xform = wd >= 0 && wd < 10 && !is_loader;
  const uint32_t s_base = smem_u32(smem);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars);
  const uint32_t bar0 = s_base + L.bars;
  auto BAR = [&](int i) { return bar0 + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(bars + B_TPTR);
  const int n_stage_tile = p.cin_v / TC_STAGE_CH;     // 2, 4 or 8
  // contiguous run of tiles per CTA: consecutive tiles of one utterance share halo rows (L2) and the per-utterance
  // InstanceNorm constants, which a transform warp then reloads only when the utterance changes
  // (dataflow-launch code -- round-robin tile groups, flag polls, flag publication -- exists only in XF_FLOW instantiations:
  // the four roles share a small instruction cache and the small-tile layers are sensitive to every extra instruction)
  constexpr bool flow = (XF & XF_FLOW) != 0;
  if constexpr (!flow) rr_group = 0;
  int t_begin = 0, n_my_tiles = 0;
  if (rr_group <= 0) {
    t_begin = (int)((int64_t)cta_idx * total_tiles / (int64_t)n_cta);
    n_my_tiles = (int)((int64_t)(cta_idx + 1) * total_tiles / (int64_t)n_cta) - t_begin;
  } else {
    const int n_groups = (total_tiles + rr_group - 1) / rr_group;
    const int my_groups = cta_idx < n_groups ? (n_groups - cta_idx + n_cta - 1) / n_cta : 0;
    n_my_tiles = my_groups * rr_group;
    if (my_groups > 0 && (n_groups - 1) % n_cta == cta_idx) n_my_tiles -= n_groups * rr_group - total_tiles;   // the last group is short
  }
  auto tile_of = [&](int k) {
    return rr_group <= 0 ? t_begin + k : ((k / rr_group) * n_cta + cta_idx) * rr_group + k % rr_group;
  };
  // (utterance, tile in utterance) of local tile k: one division at the start of a role's loop, then incremental in the
  // contiguous mode (the small-tile layers run ~700 cycles per tile: a division per tile and role shows)
  auto locate = [&](int k, int& b, int& tin) {
    const int t = tile_of(k);
    b = t / tiles_per_utt; tin = t - b * tiles_per_utt;
  };
  auto advance = [&](int k_next, int step, int& b, int& tin) {
    if (rr_group <= 0) { tin += step; while (tin >= tiles_per_utt) { tin -= tiles_per_utt; ++b; } }
    else if (k_next < n_my_tiles) locate(k_next, b, tin);
  };

  pdl_launch_dependents();      // the next kernel's CTAs may take this SM as soon as this CTA exits (common.cuh)
  if (warp == TC_MMA_WARP) {
    if (lane == 0) {
      for (int i = 0; i < L.n_stages; ++i) {
        mbar_init(BAR(B_RAW + i), 1);
        mbar_init(BAR(B_FULL + i), 1);
        mbar_init(BAR(B_EMPTY + i), 1);
      }
      // (Toeplitz epilogue on eight warps: each accumulator buffer belongs to ONE group of four warps)
      for (int i = 0; i < 2; ++i) { mbar_init(BAR(B_TFULL + i), 1); mbar_init(BAR(B_TEMPTY + i), (COUT_V == 16 && E == 8) ? 4 : E); }
      mbar_init(BAR(B_W), 1);
      if constexpr ((XF & XF_GEN) != 0)
        for (int i = 0; i < TC_GEN_SLOTS; ++i) { mbar_init(BAR(B_SFULL + i), 1); mbar_init(BAR(B_SEMPTY + i), (uint32_t)(p.cin_v / TC_STAGE_CH)); }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      // resident weights: one bulk async copy per tap slice
      mbar_expect_tx(BAR(B_W), L.w_bytes);
      const uint32_t per_tap = L.w_bytes / (uint32_t)p.ntaps;
      for (int t = 0; t < p.ntaps; ++t)
        bulk_g2s(s_base + L.w + (uint32_t)t * per_tap, reinterpret_cast<const char*>(p.w) + (size_t)t * per_tap,
                 per_tap, BAR(B_W));
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_ptr_smem)),
                 "r"(TC_TMEM_COLS(COUT_V))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp < E) {  // epilogue constants
    float* ec = reinterpret_cast<float*>(smem + L.epi);
    for (int c = threadIdx.x; c < COUT_V; c += E * 32) {
      const int cm = c % p.cout_mod;
      ec[c] = p.bias ? p.bias[p.bias_v ? c : cm] : 0.f;
      ec[COUT_V + c] = p.e_scale ? p.e_scale[cm] : 1.f;
      ec[2 * COUT_V + c] = p.e_shift ? p.e_shift[cm] : 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                   // everything above touched only constants (weights, biases); the planes come from the predecessor
  const uint32_t tmem_base = *tmem_ptr_smem;
  const int planes_in = p.cin_v / 8;

  if (is_loader) {
    // =========================== LOADER ===========================
    // two bulk copies per stage (one per 8-channel chunk), issued by one elected lane of the converged warp
    if constexpr ((XF & XF_GEN) != 0) {
      // XF_GEN: no plane to load.  The loader warps prepare the SIGNAL WINDOW of each tile instead (loader l: tiles
      // k == l mod 2): the clamped / normalised / zero-padded feature samples x[2*g0 - 7 ...] as fp16, twice (the second
      // copy shifted by one sample, so that the odd half of the pair view also reads aligned 32-bit words).
      const int lw = wd >> 2;
      float* tbl = reinterpret_cast<float*>(smem + L.sig + (uint32_t)TC_GEN_SLOTS * 2u * TC_GEN_SIG * 2u) + lw * 2 * TC_GEN_MAXF;
      const int F = p.gen_F;
      if (p.gen_mean)
        for (int f = lane; f < F; f += 32) { tbl[f] = __ldg(p.gen_mean + f); tbl[TC_GEN_MAXF + f] = 1.0f / __ldg(p.gen_std + f); }
      __syncwarp();
      const int Lfull = 2 * p.rows_in;
      const int nvalid = (int)min((int64_t)Lfull, p.gen_nvalid);
      int b = 0, tin = 0;
      if (lw < n_my_tiles) locate(lw, b, tin);
      for (int k = lw; k < n_my_tiles; k += TC_LOAD_WARPS, advance(k, TC_LOAD_WARPS, b, tin)) {
        const int origin = 2 * (tin * tile_step - halo_lo) - 7;       // x index of window sample 0
        const float* xb = p.gen_feats + (int64_t)b * p.gen_bstride;
        float xv[TC_GEN_SIG / 32];
#pragma unroll
        for (int q = 0; q < TC_GEN_SIG / 32; ++q) {                   // all loads in flight before any use
          const int l = origin + q * 32 + lane;
          xv[q] = (l >= 0 && l < nvalid) ? __ldg(xb + l) : 0.f;
        }
        if (p.gen_mean) {
          const float floor_db = p.gen_umax ? key_to_float(__ldg(p.gen_umax + b)) - p.gen_topdb : -INFINITY;
          int f = (origin + lane) % F;
          if (f < 0) f += F;
#pragma unroll
          for (int q = 0; q < TC_GEN_SIG / 32; ++q) {
            const int l = origin + q * 32 + lane;
            if (l >= 0 && l < nvalid) xv[q] = (fmaxf(xv[q], floor_db) - tbl[f]) * tbl[TC_GEN_MAXF + f];
            f += 32;
            if (f >= F) f -= F;
          }
        }
        const int slot = k % TC_GEN_SLOTS;
        mbar_wait(BAR(B_SEMPTY + slot), (((uint32_t)k / TC_GEN_SLOTS) & 1u) ^ 1u);
        __half* sa = reinterpret_cast<__half*>(smem + L.sig) + slot * 2 * TC_GEN_SIG;
#pragma unroll
        for (int q = 0; q < TC_GEN_SIG / 32; ++q) {
          const __half hv = __float2half_rn(xv[q]);
          const int j = q * 32 + lane;
          sa[j] = hv;
          if (j > 0) sa[TC_GEN_SIG + j - 1] = hv;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(BAR(B_SFULL + slot));
        if (p.trace && lane == 0 && k * n_stage_tile < TRACE_N)
          for (int s2 = 0; s2 < n_stage_tile; ++s2) p.trace[((int64_t)blockIdx.x * 6 + 3) * TRACE_N + k * n_stage_tile + s2] = clock64();
      }
    } else if (COUT_V == 16 && n_stage_tile == 2) {
      // decoder.8 (32 input channels = two stages per tile, 2 KB copies): with one stage per loader and tile the launch was paced
      // by the loader warps' own instruction chain (~670 cycles per iteration for ~100 dependent instructions, clock64
      // trace).  Here the two loaders take alternate TILES and issue both stages of a tile from one set of addresses.
      const __half* in_base = reinterpret_cast<const __half*>(p.in);
      const int lw = wd >> 2;
      int b = 0, tin = 0;
      if (lw < n_my_tiles) locate(lw, b, tin);
      const int64_t plane_stride = (int64_t)p.rows_in * 8;
      for (int k = lw; k < n_my_tiles; k += TC_LOAD_WARPS, advance(k, TC_LOAD_WARPS, b, tin)) {
        const int g0 = tin * tile_step - halo_lo;
        const int lo = g0 < 0 ? -g0 : 0;
        const int hi = min(slab_used, p.rows_in - g0);
        const uint32_t bytes = hi > lo ? (uint32_t)(hi - lo) * 16u : 0u;
        const __half* src = in_base + (((int64_t)b * planes_in) * p.rows_in + (g0 + lo)) * 8;
        const uint32_t it0 = (uint32_t)k * 2u;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const uint32_t it = it0 + (uint32_t)s, stage = it % (uint32_t)L.n_stages, ph = (it / (uint32_t)L.n_stages) & 1u;
          mbar_wait(BAR(B_EMPTY + stage), ph ^ 1u);
          mbar_expect_tx_elect(BAR(B_RAW + stage), bytes * (TC_STAGE_CH / 8));
          if (bytes) {
            const uint32_t dst = s_base + L.a + stage * L.stage_bytes + (uint32_t)lo * 16u;
            bulk_g2s_elect(dst, src, bytes, BAR(B_RAW + stage));
            bulk_g2s_elect(dst + L.lbo_a, src + plane_stride, bytes, BAR(B_RAW + stage));
          }
          src += (int64_t)(TC_STAGE_CH / 8) * plane_stride;
        }
      }
    } else {
      const __half* in_base = reinterpret_cast<const __half*>(p.in);
      const int lw = wd >> 2;
      uint32_t stage = (uint32_t)lw, ph = 0;
      int b = 0, tin = 0;
      if (n_my_tiles > 0) locate(0, b, tin);
      for (int k = 0; k < n_my_tiles; ++k, advance(k, 1, b, tin)) {
        const int g0 = tin * tile_step - halo_lo;                      // global row of slab row 0
        const int lo = g0 < 0 ? -g0 : 0;                               // first valid slab row
        const int hi = min(slab_used, p.rows_in - g0);                 // one past the last valid slab row
        if (flow && p.flag_in && hi > lo) {
          // dataflow launch: the rows [g0+lo, g0+hi) of this utterance come from the PRODUCER layer's tiles j0..j1, written by
          // other CTAs of the same launch; each of the producer's epilogue warps bumps the tile's counter after its
          // stores (release), the poll below acquires, and the proxy fence orders the bulk (async-proxy) reads after it
          const int j0 = (g0 + lo) / TC_M, j1 = (g0 + hi - 1) / TC_M;
          for (int j = j0; j <= j1; ++j) {
            const unsigned* f = p.flag_in + (int64_t)b * p.flag_in_tpu + j;
            unsigned v, spins = 0;
            for (;;) {
              asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
              if (v >= p.flag_target) break;
              __nanosleep(64);
              // never hang the GPU: after ~0.3 s (or as soon as any CTA has given up) flag the error and go on
              if (++spins > (1u << 22) || (p.err_flag && (spins & 1023u) == 0 && *reinterpret_cast<volatile unsigned*>(p.err_flag))) {
                if (lane == 0 && p.err_flag) atomicExch(p.err_flag, 1u);
                break;
              }
            }
          }
          asm volatile("fence.proxy.async;" ::: "memory");
        }
        const uint32_t bytes = hi > lo ? (uint32_t)(hi - lo) * 16u : 0u;
        const int64_t plane_stride = (int64_t)p.rows_in * 8;
        const __half* src = in_base + (((int64_t)b * planes_in + lw * (TC_STAGE_CH / 8)) * p.rows_in + (g0 + lo)) * 8;
        for (int s = lw; s < n_stage_tile; s += TC_LOAD_WARPS) {
          mbar_wait(BAR(B_EMPTY + stage), ph ^ 1u);
          if (p.trace && lane == 0 && k * n_stage_tile + s < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 3) * TRACE_N + k * n_stage_tile + s] = clock64();
          mbar_expect_tx_elect(BAR(B_RAW + stage), bytes * (TC_STAGE_CH / 8));
          if (bytes) {
            const uint32_t dst = s_base + L.a + stage * L.stage_bytes + (uint32_t)lo * 16u;
            bulk_g2s_elect(dst, src, bytes, BAR(B_RAW + stage));
            bulk_g2s_elect(dst + L.lbo_a, src + plane_stride, bytes, BAR(B_RAW + stage));
          }
          src += (int64_t)(TC_LOAD_WARPS * TC_STAGE_CH / 8) * plane_stride;
          stage += TC_LOAD_WARPS;
          if (stage >= (uint32_t)L.n_stages) { stage -= (uint32_t)L.n_stages; ph ^= 1u; }
        }
      }
    }
  } else if (is_xform) {
    // =========================== TRANSFORM (in place) ===========================
    // Eight warps; warp w owns the stages it == w (mod 8) (16 channels = two 8-channel chunks x slab rows), so eight
    // stages are in flight and the per-stage chain (mbarrier wait -> LDS -> math -> STS -> proxy fence -> arrive) is
    // off the critical path.  Lanes are rows (conflict-free 16 B accesses).  All arithmetic is packed fp16:
    //   y = x*s1 + h1 (InstanceNorm affine, halved when SiLU follows)  ->  SiLU: y + y*tanh(y)  |  ReLU
    //   -> y*s2 + h2 (folded eval-mode BatchNorm of the producer)
    // i.e. per 8 elements 4 HFMA2 + 4 MUFU + 4 HFMA2 (+ 4 HFMA2).  Rows outside the utterance (= the conv's zero
    // padding of the already activated tensor) exist only in an utterance's first / last tile and are zeroed there.
    const int xw = wd - (wd >> 2);
    // (XF_DYN: the activation is the layer's own, a launch may mix SiLU and ReLU layers; otherwise compile-time)
    constexpr bool dyn = (XF & XF_DYN) != 0, has_t2 = (XF & XF_T2) != 0;
    const bool do_silu = dyn ? p.silu == 1 : (XF & 3) == XF_SILU, do_relu = dyn ? p.silu == 2 : (XF & 3) == XF_RELU;
    constexpr bool ident = (XF & 7) == XF_IDENT && !dyn;
    const bool has_t1 = p.t1_scale != nullptr;
    const float pre = do_silu ? 0.5f : 1.f;                    // SiLU works on x/2: fold the 1/2 into the affine
    const int s_fix = xw % n_stage_tile;                       // stage-in-tile of this warp (n_stage_tile is 2, 4 or 8)
    const int kstep = TC_XF_WARPS / n_stage_tile;              // tiles between two stages of this warp
    int k = xw / n_stage_tile;
    uint32_t stage = (uint32_t)xw, ph = 0;                     // n_stages is 8 or 16
    int b = 0, tin = 0;
    if (k < n_my_tiles) locate(k, b, tin);
    const int cm = (s_fix * TC_STAGE_CH) % p.cin_mod;          // this warp's 16 channels
    __half2 s1[8], h1[8], s2[8], h2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s1[j] = __float2half2_rn(pre); h1[j] = __float2half2_rn(0.f); s2[j] = s1[j]; h2[j] = h1[j]; }
    if (has_t2) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s2[j] = __floats2half2_rn(__ldg(p.t2_scale + cm + 2 * j), __ldg(p.t2_scale + cm + 2 * j + 1));
        h2[j] = __floats2half2_rn(__ldg(p.t2_shift + cm + 2 * j), __ldg(p.t2_shift + cm + 2 * j + 1));
      }
    }
    int b_const = -1;                                          // utterance whose t1 constants are in registers
    if constexpr ((XF & XF_GEN) != 0) {
      // ---- encoder.0 fused into encoder.2's operand production (models/ConvAutoEncoder.py:141-144: Conv1d(1->32,k15,p7),
      // GLU, then the stride-2 conv with nothing in between).  Stage s of the pair view holds e0 rows 2r + half
      // (half = s / 2) and channels ci0 .. ci0+15 (ci0 = 16 * (s & 1)); e0[l][c] = SiLU(b[c] + sum_t w[c][t] x[l+t-7]).
      // A warp computes its stage with warp-level MMAs over a Toeplitz view of the signal, A[i][t] = x[l0 + 2i + t - 7]:
      // with sig[j] = x[l0 - 7 + j] every A fragment register is one aligned 32-bit word of sig (rows advance by two
      // samples), B = w/2 and the accumulators start at b/2, so the MMA gives h = y/2 and SiLU(y) = h + h*tanh(h).
      const int half = (s_fix * TC_STAGE_CH) / p.cin_mod, ci0 = (s_fix * TC_STAGE_CH) % p.cin_mod;
      const int g = lane >> 2, t4 = lane & 3;
      uint32_t bw[2][2];
      float bz[2][2];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        const float* wc = p.gen_w + (ci0 + 8 * nt + g) * 15;
        const int ta = 2 * t4, tb = 2 * t4 + 8;
        const __half2 wl = __floats2half2_rn(0.5f * __ldg(wc + ta), 0.5f * __ldg(wc + ta + 1));
        const __half2 wh = __floats2half2_rn(0.5f * __ldg(wc + tb), tb + 1 < 15 ? 0.5f * __ldg(wc + tb + 1) : 0.f);
        bw[nt][0] = *reinterpret_cast<const uint32_t*>(&wl);
        bw[nt][1] = *reinterpret_cast<const uint32_t*>(&wh);
        bz[nt][0] = 0.5f * __ldg(p.gen_b + ci0 + 8 * nt + 2 * t4);
        bz[nt][1] = 0.5f * __ldg(p.gen_b + ci0 + 8 * nt + 2 * t4 + 1);
      }
      for (; k < n_my_tiles; k += kstep, advance(k, kstep, b, tin)) {
        const int g0 = tin * tile_step - halo_lo;
        const int lo = g0 < 0 ? -g0 : 0;
        const int hi = min(slab_used, p.rows_in - g0);
        const int slot = k % TC_GEN_SLOTS;
        // window copy `half`: sig[j] = x[2*g0 + half - 7 + j]
        const uint32_t* sig32 = reinterpret_cast<const uint32_t*>(smem + L.sig) + (slot * 2 + half) * (TC_GEN_SIG / 2);
        mbar_wait(BAR(B_SFULL + slot), ((uint32_t)k / TC_GEN_SLOTS) & 1u);
        mbar_wait(BAR(B_EMPTY + stage), ph ^ 1u);              // the MMAs of this slot's previous use have read it
        const int it = k * n_stage_tile + s_fix;
        if (p.trace && lane == 0 && it < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 4) * TRACE_N + it] = clock64();
        unsigned char* sbase = smem + L.a + stage * L.stage_bytes;
        // 9 x 16 slab rows >= slab_used (130), three 16-row tiles per round: all loads, then the six MMAs, then the
        // twelve SiLUs and stores (the chain LDS -> HMMA -> cvt -> MUFU -> HFMA2 -> STS is latency-bound otherwise)
#pragma unroll 1
        for (int m3 = 0; m3 < 3; ++m3) {
          uint32_t a[3][4];
#pragma unroll
          for (int u = 0; u < 3; ++u) {
            const int o = g + t4 + 16 * (3 * m3 + u);
            a[u][0] = sig32[o]; a[u][1] = sig32[o + 8]; a[u][2] = sig32[o + 4]; a[u][3] = sig32[o + 12];
          }
          float acc[3][2][4];
#pragma unroll
          for (int u = 0; u < 3; ++u)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
              acc[u][nt][0] = bz[nt][0]; acc[u][nt][1] = bz[nt][1]; acc[u][nt][2] = bz[nt][0]; acc[u][nt][3] = bz[nt][1];
              asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                  : "+f"(acc[u][nt][0]), "+f"(acc[u][nt][1]), "+f"(acc[u][nt][2]), "+f"(acc[u][nt][3])
                  : "r"(a[u][0]), "r"(a[u][1]), "r"(a[u][2]), "r"(a[u][3]), "r"(bw[nt][0]), "r"(bw[nt][1]));
            }
          __half2 y[3][2][2];
#pragma unroll
          for (int u = 0; u < 3; ++u)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
              for (int hr = 0; hr < 2; ++hr)
                y[u][nt][hr] = silu_from_half(__floats2half2_rn(acc[u][nt][2 * hr], acc[u][nt][2 * hr + 1]));
#pragma unroll
          for (int u = 0; u < 3; ++u)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
              for (int hr = 0; hr < 2; ++hr) {
                const int r = 16 * (3 * m3 + u) + g + 8 * hr;
                if (r < slab_used)
                  *reinterpret_cast<__half2*>(sbase + (uint32_t)nt * L.lbo_a + (uint32_t)r * 16u + 4u * t4) = y[u][nt][hr];
              }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(BAR(B_SEMPTY + slot));      // this warp is done with the tile's signal window
        if (lo > 0 || hi < slab_used) {             // rows outside the utterance are the stride-2 conv's zero padding
          unsigned char* base = sbase + (uint32_t)lane * 16u;
          for (int r = lane; r < slab_used; r += 32)
            if (r < lo || r >= hi) {
              *reinterpret_cast<uint4*>(base + (uint32_t)(r - lane) * 16u) = make_uint4(0u, 0u, 0u, 0u);
              *reinterpret_cast<uint4*>(base + L.lbo_a + (uint32_t)(r - lane) * 16u) = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(BAR(B_FULL + stage));
        if (p.trace && lane == 0 && it < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 0) * TRACE_N + it] = clock64();
        stage += TC_XF_WARPS;
        if (stage >= (uint32_t)L.n_stages) { stage -= (uint32_t)L.n_stages; ph ^= 1u; }
      }
    } else
    for (; k < n_my_tiles; k += kstep, advance(k, kstep, b, tin)) {
      const int g0 = tin * tile_step - halo_lo;
      const int lo = g0 < 0 ? -g0 : 0;
      const int hi = min(slab_used, p.rows_in - g0);
      if (!ident && has_t1 && b_const != b && (b_const < 0 || p.t1_bstride != 0)) {
        const float4* sp = reinterpret_cast<const float4*>(p.t1_scale + (int64_t)b * p.t1_bstride + cm);
        const float4* hp = reinterpret_cast<const float4*>(p.t1_shift + (int64_t)b * p.t1_bstride + cm);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 a = __ldg(sp + q), c = __ldg(hp + q);
          s1[2 * q] = __floats2half2_rn(a.x * pre, a.y * pre); s1[2 * q + 1] = __floats2half2_rn(a.z * pre, a.w * pre);
          h1[2 * q] = __floats2half2_rn(c.x * pre, c.y * pre); h1[2 * q + 1] = __floats2half2_rn(c.z * pre, c.w * pre);
        }
        b_const = b;
      }
      mbar_wait(BAR(B_RAW + stage), ph);
      const int it = k * n_stage_tile + s_fix;
      if (p.trace && lane == 0 && it < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 4) * TRACE_N + it] = clock64();
      unsigned char* base = smem + L.a + stage * L.stage_bytes + (uint32_t)lane * 16u;
      if (!ident) {
        const bool tail = lane + 128 < slab_used;    // the halo rows beyond the first 128 (slab_used <= 160)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          unsigned char* cb = base + (uint32_t)c * L.lbo_a;
          uint4 v[5];
#pragma unroll
          for (int q = 0; q < 4; ++q) v[q] = *reinterpret_cast<const uint4*>(cb + q * 512);
          v[4] = make_uint4(0u, 0u, 0u, 0u);
          if (tail) v[4] = *reinterpret_cast<const uint4*>(cb + 4 * 512);
#pragma unroll
          for (int q = 0; q < 5; ++q) {
            __half2* hv = reinterpret_cast<__half2*>(&v[q]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              __half2 y = __hfma2(hv[j], s1[4 * c + j], h1[4 * c + j]);
              if (do_silu) y = silu_from_half(y);
              if (do_relu) y = __hmax2(y, __float2half2_rn(0.f));
              if (has_t2) y = __hfma2(y, s2[4 * c + j], h2[4 * c + j]);
              hv[j] = y;
            }
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4*>(cb + q * 512) = v[q];
          if (tail) *reinterpret_cast<uint4*>(cb + 4 * 512) = v[4];
        }
      }
      if (lo > 0 || hi < slab_used) {               // first / last tile of an utterance: rows outside it are zeros
        for (int r = lane; r < slab_used; r += 32)
          if (r < lo || r >= hi) {
            *reinterpret_cast<uint4*>(base + (uint32_t)(r - lane) * 16u) = make_uint4(0u, 0u, 0u, 0u);
            *reinterpret_cast<uint4*>(base + L.lbo_a + (uint32_t)(r - lane) * 16u) = make_uint4(0u, 0u, 0u, 0u);
          }
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(BAR(B_FULL + stage));
      if (p.trace && lane == 0 && it < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 0) * TRACE_N + it] = clock64();
      stage += TC_XF_WARPS;
      if (stage >= (uint32_t)L.n_stages) { stage -= (uint32_t)L.n_stages; ph ^= 1u; }
    }
  } else if (warp == TC_MMA_WARP) {
    // =========================== MMA ISSUER ===========================
    // One thread issues; its instruction stream is the critical path (a lone thread retires a dependent
    // instruction every ~4 cycles), so the per-MMA work is a couple of adds on the descriptors' low words:
    //   A: + dil (16 B rows) per tap;   B: + one [cin_v/8][cout_v] slice per tap, + two chunks per stage.
    // All 32 lanes run the loop converged (see umma_f16_elect); waits are warp-wide polls of the same barrier.
    {
      mbar_wait(BAR(B_W), 0);
      const uint32_t desc_hi = (uint32_t)(umma_desc(0, 0) >> 32);
      const uint32_t a_lo_const = (uint32_t)umma_desc(0, L.lbo_a);
      const uint32_t b_lo_const = (uint32_t)umma_desc(s_base + L.w, (uint32_t)COUT_V * 16u);
      const uint32_t b_tap = (uint32_t)(p.cin_v / 8) * COUT_V;      // 16 B units
      const uint32_t b_stage = (uint32_t)(TC_STAGE_CH / 8) * COUT_V;
      const uint32_t dil = (uint32_t)p.tap_step;
      const uint32_t idesc_full = umma_idesc(COUT_V), idesc_tap0 = umma_idesc(p.tap_nmax[0]);
      const int s_short = p.tap_kmax[p.ntaps - 1] / TC_STAGE_CH;  // stages >= s_short skip the last tap
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      const int last_tap = p.ntaps - 1;
      uint32_t it = 0, stage = 0, ph = 0;
      (void)mma_group;
      // The wait -> fence -> descriptor set-up -> commit sequence around a stage's MMAs costs the issuer ~140 cycles during
      // which the tensor pipe drains (a 16-channel stage is only ntaps MMAs = 190..340 cycles of work), so MG
      // consecutive stages are handled per round: one set of waits, the MMAs of all of them tap by tap, then the commits.
      // (groups never straddle the ring's wrap: ring size and stages per tile are multiples of the group)
      for (int k = 0; k < n_my_tiles; ++k) {
        const uint32_t buf = (uint32_t)k & 1u;
        mbar_wait(BAR(B_TEMPTY + buf), (((uint32_t)k >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_u + buf * (uint32_t)COUT_V;
        if constexpr (MG == 1) {
        for (int s = 0; s < n_stage_tile; ++s, ++it) {
          uint32_t a_lo = a_lo_const + ((s_base + L.a + stage * L.stage_bytes) >> 4);
          uint32_t b_lo = b_lo_const + (uint32_t)s * b_stage;
          const int nt = s >= s_short ? p.ntaps - 1 : p.ntaps;
          mbar_wait(BAR(B_FULL + stage), ph);
          tc_fence_after();
          if (p.trace && lane == 0 && it < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 5) * TRACE_N + it] = clock64();
          umma_f16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo, ((uint64_t)desc_hi << 32) | b_lo,
                         s == 0 ? idesc_full : idesc_tap0, s != 0);
#pragma unroll 1
          for (int tap = 1; tap < nt; ++tap) {
            a_lo += dil;
            b_lo += b_tap;
            umma_f16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo, ((uint64_t)desc_hi << 32) | b_lo, idesc_full, 1u);
          }
          tc_commit_elect(BAR(B_EMPTY + stage));    // slot free once these MMAs have read it
          if (p.trace && lane == 0 && it < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 1) * TRACE_N + it] = clock64();
          if (++stage == (uint32_t)L.n_stages) { stage = 0; ph ^= 1u; }
        }
        } else
        for (int s = 0; s < n_stage_tile; s += MG, it += MG) {
#pragma unroll
          for (int g = 0; g < MG; ++g) mbar_wait(BAR(B_FULL + stage + g), ph);
          tc_fence_after();
          if (p.trace && lane == 0 && it < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 5) * TRACE_N + it] = clock64();
          uint32_t a_lo[MG], b_lo[MG];
#pragma unroll
          for (int g = 0; g < MG; ++g) {
            a_lo[g] = a_lo_const + ((s_base + L.a + (stage + g) * L.stage_bytes) >> 4);
            b_lo[g] = b_lo_const + (uint32_t)(s + g) * b_stage;
          }
          // tap 0 (the first MMA of a tile initialises all columns; otherwise it may cover fewer columns)
#pragma unroll
          for (int g = 0; g < MG; ++g)
            umma_f16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo[g], ((uint64_t)desc_hi << 32) | b_lo[g],
                           (s + g) == 0 ? idesc_full : idesc_tap0, (uint32_t)((s + g) != 0));
#pragma unroll 1
          for (int tap = 1; tap < last_tap; ++tap) {
#pragma unroll
            for (int g = 0; g < MG; ++g) {
              a_lo[g] += dil;
              b_lo[g] += b_tap;
              umma_f16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo[g], ((uint64_t)desc_hi << 32) | b_lo[g], idesc_full, 1u);
            }
          }
          if (last_tap >= 1) {   // stages >= s_short skip the last tap (the non-existent sixth tap of a stride-2 k5 conv)
#pragma unroll
            for (int g = 0; g < MG; ++g)
              if (s + g < s_short)
                umma_f16_elect(d_tmem, ((uint64_t)desc_hi << 32) | (a_lo[g] + dil), ((uint64_t)desc_hi << 32) | (b_lo[g] + b_tap), idesc_full, 1u);
          }
#pragma unroll
          for (int g = 0; g < MG; ++g) tc_commit_elect(BAR(B_EMPTY + stage + g));    // slots free once these MMAs have read them
          if (p.trace && lane == 0 && it < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 1) * TRACE_N + it] = clock64();
          stage += (uint32_t)MG;
          if (stage >= (uint32_t)L.n_stages) { stage = 0; ph ^= 1u; }
        }
        tc_commit_elect(BAR(B_TFULL + buf));        // accumulator complete
      }
    }
  } else if (warp < E) {
    // =========================== EPILOGUE ===========================
    // warp w: TMEM lanes (rows) 32*(w%4).., columns [(w/4)*COUT_V/2, +COUT_V/2)
    const float* ec = reinterpret_cast<const float*>(smem + L.epi);
    const int quarter = warp & 3, chalf = warp >> 2;
    const int mloc = quarter * 32 + lane;
    constexpr int CPW = COUT_V / (E / 4);     // columns per warp
    constexpr int NCH = COUT_V / 8;     // output chunks
    const bool cl = p.out_mode == OUT_CHANNELS_LAST;
    if constexpr (COUT_V == 16) {
      // ---- Toeplitz epilogue (decoder.8, 32 -> 1 channel, 15 taps): the GEMM produced D[r][t] = sum_c a[r][c] w[c][t]
      //      for the 128 slab rows; out[l] = bias + sum_t D[l - m0 + t][t]  (slab row 0 = global row m0 - 7)
      float* dt = reinterpret_cast<float*>(smem + L.red);
      const float bias0 = ec[0];
      // Per tile the epilogue is ONE latency chain (accumulator load -> transposed store -> barrier -> 15 dependent adds ->
      // store, ~670 cycles in the clock64 trace) and paced the launch.  With eight warps (XF_EPI8) two groups of four take
      // alternate tiles -- tile k lives in accumulator buffer k & 1, so each group owns one buffer -- and two chains overlap.
      constexpr int NG = E / 4;                       // groups of four warps
      const int grp = warp >> 2;
      int b = 0, tin = 0;
      if (grp < n_my_tiles) locate(grp, b, tin);
      for (int k = grp; k < n_my_tiles; k += NG, advance(k, NG, b, tin)) {
        const int m0 = tin * tile_step;
        const uint32_t buf = (uint32_t)k & 1u;
        mbar_wait(BAR(B_TFULL + buf), ((uint32_t)k >> 1) & 1u);
        tc_fence_after();
        float v[16];
        tmem_ld16(tmem_base + buf * (uint32_t)COUT_V + ((uint32_t)(quarter * 32) << 16), v);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(BAR(B_TEMPTY + buf));
        // two transposition buffers per group: a group's buffer is rewritten two of ITS tiles later, i.e. after its next barrier
        float* d = dt + (grp * 2 + (((uint32_t)k / NG) & 1u)) * (TC_M * 17);
#pragma unroll
        for (int t = 0; t < 16; ++t) d[mloc * 17 + t] = v[t];
        if (NG == 1) epi_bar();
        else if (grp == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
        else asm volatile("bar.sync 2, 128;" ::: "memory");
        const int m = m0 + mloc;
        if (mloc < TC_TOEP_STEP && m < p.rows_out) {
          float acc = bias0;
#pragma unroll
          for (int t = 0; t < TC_TOEP_TAPS; ++t) acc += d[(mloc + t) * 17 + t];
          if (p.out_half) reinterpret_cast<__half*>(p.out)[(int64_t)b * p.out_bstride + m] = __float2half_rn(acc);
          else reinterpret_cast<float*>(p.out)[(int64_t)b * p.out_bstride + m] = acc;
        }
      }
    } else {
    // ---- InstanceNorm statistics of the output from the accumulator itself (XF_STATS): after a 32-column block is
    // stored, its column sums over the warp's 32 rows leave lane l with the sum (and the sum of squares) of column
    // cb + 2*(l & 15) + (l >> 4); they stay in registers until the utterance changes.
    constexpr bool e_stats = (XF & XF_STATS) != 0;
    // (fp64: the partial sums run over however many tiles of the utterance THIS CTA happens to own -- with fp32 the statistics,
    //  and through the fp16-rounded affine of the consumer every output, depended on the batch an utterance was in at 1e-3)
    double es[4] = {0.0, 0.0, 0.0, 0.0}, eq[4] = {0.0, 0.0, 0.0, 0.0};
    auto e_flush = [&](int bb) {
#pragma unroll
      for (int i = 0; i < CPW / 32; ++i) {
        const int ch = (chalf * CPW + 32 * i + 2 * (lane & 15) + (lane >> 4)) % p.cout_mod;
        double* dst = p.stats + ((int64_t)bb * p.cout_mod + ch) * 2;
        atomicAdd(dst, es[i]);
        atomicAdd(dst + 1, eq[i]);
        es[i] = 0.0; eq[i] = 0.0;
      }
    };
    int b_prev = -1, b = 0, tin = 0;
    if (n_my_tiles > 0) locate(0, b, tin);
    for (int k = 0; k < n_my_tiles; ++k, advance(k, 1, b, tin)) {
      if constexpr (e_stats) { if (b_prev >= 0 && b != b_prev) e_flush(b_prev); }
      b_prev = b;
      const int m0 = tin * tile_step;
      const int m = m0 + mloc;
      const bool valid = m < p.rows_out;
      const uint32_t buf = (uint32_t)k & 1u;
      mbar_wait(BAR(B_TFULL + buf), ((uint32_t)k >> 1) & 1u);
      tc_fence_after();
      if (p.trace && threadIdx.x == 0 && 2 * k < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 2) * TRACE_N + 2 * k] = clock64();
      const uint32_t taddr = tmem_base + buf * (uint32_t)COUT_V + ((uint32_t)(quarter * 32) << 16);
      __half* out_h = reinterpret_cast<__half*>(p.out) + (int64_t)b * p.out_bstride;
      // (rolled: the kernel's roles share a small instruction cache)
#pragma unroll 1
      for (int cb = chalf * CPW; cb < (chalf + 1) * CPW; cb += 32) {
        float v[32];
        tmem_ld32(taddr + (uint32_t)cb, v);
        if (cb + 32 >= (chalf + 1) * CPW) {                  // last read of this accumulator buffer: hand it back now,
          tc_fence_before();                                 // the stores (and statistics) below overlap the next MMAs
          __syncwarp();
          if (lane == 0) mbar_arrive(BAR(B_TEMPTY + buf));
        }
        if (p.bias) {   // (a bias in front of an InstanceNorm is a no-op and is dropped by the caller)
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 bz = *reinterpret_cast<const float4*>(ec + cb + j);
            v[j] += bz.x; v[j + 1] += bz.y; v[j + 2] += bz.z; v[j + 3] += bz.w;
          }
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        if (p.e_scale) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 es = *reinterpret_cast<const float4*>(ec + COUT_V + cb + j);
            const float4 eh = *reinterpret_cast<const float4*>(ec + 2 * COUT_V + cb + j);
            v[j] = fmaf(v[j], es.x, eh.x); v[j + 1] = fmaf(v[j + 1], es.y, eh.y);
            v[j + 2] = fmaf(v[j + 2], es.z, eh.z); v[j + 3] = fmaf(v[j + 3], es.w, eh.w);
          }
        }
        uint32_t hw[16];                                      // the block as stored: 16 fp16 pairs
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const __half2 h2 = __floats2half2_rn(v[2 * j], v[2 * j + 1]);
          hw[j] = *reinterpret_cast<const uint32_t*>(&h2);
        }
        if (cl) {
          if (valid) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint4 o = make_uint4(hw[4 * q], hw[4 * q + 1], hw[4 * q + 2], hw[4 * q + 3]);
              const int c8 = (cb >> 3) + q;
              int pl, row;       // destination plane / row in the consumer's view
              if (p.out_fold == OUT_PAIR_FOLD) { pl = (m & 1) * NCH + c8; row = m >> 1; }
              else if (p.out_fold == OUT_PHASE_UNFOLD) { pl = c8 % (NCH / 2); row = 2 * m + c8 / (NCH / 2); }
              else if (p.out_fold == OUT_PHASE_UNFOLD4) { pl = c8 % (NCH / 4); row = 4 * m + c8 / (NCH / 4); }
              else { pl = c8; row = m; }
              *reinterpret_cast<uint4*>(out_h + (((int64_t)pl * p.out_plane_rows + row) << 3)) = o;
            }
          }
        }
        if constexpr (e_stats) {
          // column sums over the warp's 32 rows of the values AS STORED (fp16 pairs), through a padded shared-memory
          // transpose: lane = row writes its 16 words, lane = (word w, row half) reads 16 rows of word w (both
          // conflict-free); the halves meet in one shuffle and lane l keeps column cb + 2*(l & 15) + (l >> 4)
          uint32_t* tb = reinterpret_cast<uint32_t*>(smem + L.red) + warp * (32 * TC_ST_PITCH);
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 16; ++j) tb[lane * TC_ST_PITCH + j] = valid ? hw[j] : 0u;
          __syncwarp();
          const uint32_t* col = tb + ((lane >> 4) * 16) * TC_ST_PITCH + (lane & 15);
          float a0, a1, q0, q1;
          {   // packed fp32 pairs (FADD2 / FFMA2): two issue slots per word instead of four
            float2 sa = make_float2(0.f, 0.f), sq = make_float2(0.f, 0.f);
#pragma unroll
            for (int r = 0; r < 16; ++r) {
              const uint32_t w = col[r * TC_ST_PITCH];
              const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
              sa = fadd2(sa, f);
              sq = ffma2(f, f, sq);
            }
            a0 = sa.x; a1 = sa.y; q0 = sq.x; q1 = sq.y;
          }
          a0 += __shfl_xor_sync(0xffffffffu, a0, 16); a1 += __shfl_xor_sync(0xffffffffu, a1, 16);
          q0 += __shfl_xor_sync(0xffffffffu, q0, 16); q1 += __shfl_xor_sync(0xffffffffu, q1, 16);
          const float bs = (lane >> 4) ? a1 : a0, bq = (lane >> 4) ? q1 : q0;
          const int cbi = (cb - chalf * CPW) >> 5;
#pragma unroll
          for (int i = 0; i < CPW / 32; ++i)
            if (i == cbi) { es[i] += (double)bs; eq[i] += (double)bq; }
        }
      }
      if (p.trace && threadIdx.x == 0 && 2 * k + 1 < TRACE_N) p.trace[((int64_t)blockIdx.x * 6 + 2) * TRACE_N + 2 * k + 1] = clock64();
      if (flow && p.flag_out) {      // dataflow launch: publish this warp's part of tile t to the consumer layer's loaders
        __threadfence();
        __syncwarp();
        if (lane == 0) atomicAdd(p.flag_out + (int64_t)b * tiles_per_utt + tin, 1u);
      }
    }
    if constexpr (e_stats) { if (b_prev >= 0) e_flush(b_prev); }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == TC_MMA_WARP) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TC_TMEM_COLS(COUT_V)) : "memory");
  }

}

template <int COUT_V, int XF, int MG>
__global__ void __launch_bounds__(tc_threads(XF), 1)
tapgemm_tc_kernel(const __grid_constant__ TapGemmParams p, int halo_lo, int slab_used, int tiles_per_utt, int total_tiles, int tile_step,
                  int mma_group) {
  tc_body<COUT_V, XF, MG>(p, halo_lo, slab_used, tiles_per_utt, total_tiles, tile_step, (int)blockIdx.x, (int)gridDim.x, 0, mma_group);
}

// Several layers of a chain in ONE launch (dataflow through global memory / L2): the CTAs [cta0, cta0 + n_cta) work on
// layer i (its weights resident in their shared memory); a consumer layer's loaders wait on the per-tile counters its
// producer's epilogue warps bump.  Producers have the lower block indices and never wait on consumers, and the grid is
// one CTA per SM, so every CTA a waiting CTA depends on is resident or already finished.
struct TcSeg { TapGemmParams p; int halo_lo, slab_used, tiles_per_utt, total_tiles, tile_step, cta0, n_cta; };
struct TcMulti { TcSeg seg[3]; int n_seg, rr_group, mma_group; };

template <int COUT_V, int XF, int MG>
__global__ void __launch_bounds__(tc_threads(XF), 1)
tapgemm_tc_multi_kernel(const __grid_constant__ TcMulti mp) {
  const int bx = (int)blockIdx.x;
  // (three static copies of the body: each reads its layer's parameters straight from the constant bank)
  if (mp.n_seg > 2 && bx >= mp.seg[2].cta0) {
    const TcSeg& g = mp.seg[2];
    tc_body<COUT_V, XF, MG>(g.p, g.halo_lo, g.slab_used, g.tiles_per_utt, g.total_tiles, g.tile_step, bx - g.cta0, g.n_cta, mp.rr_group, mp.mma_group);
  } else if (mp.n_seg > 1 && bx >= mp.seg[1].cta0) {
    const TcSeg& g = mp.seg[1];
    tc_body<COUT_V, XF, MG>(g.p, g.halo_lo, g.slab_used, g.tiles_per_utt, g.total_tiles, g.tile_step, bx - g.cta0, g.n_cta, mp.rr_group, mp.mma_group);
  } else {
    const TcSeg& g = mp.seg[0];
    tc_body<COUT_V, XF, MG>(g.p, g.halo_lo, g.slab_used, g.tiles_per_utt, g.total_tiles, g.tile_step, bx - g.cta0, g.n_cta, mp.rr_group, mp.mma_group);
  }
}


template <int COUT_V, int XF, int MG>
int launch_tc_mg(const TapGemmParams& p, int n_sm, cudaStream_t st) {
  int lo = 0, hi = 0;
  for (int t = 0; t < p.ntaps; ++t) { lo = std::min(lo, p.shift[t]); hi = std::max(hi, p.shift[t]); }
  const bool toeplitz = COUT_V == 16;
  const int halo_lo = -lo;
  const int slab_used = toeplitz ? TC_M : TC_M + halo_lo + hi;
  const int tile_step = toeplitz ? TC_TOEP_STEP : TC_M;
  if (slab_used > 160) return SPANON_ERR_UNSUPPORTED;
  const int nst = p.cin_v / TC_STAGE_CH;
  if (nst != 2 && nst != 4 && nst != 8) return SPANON_ERR_UNSUPPORTED;
  const TcLayout L = tc_layout(p.ntaps, p.cin_v, COUT_V, slab_used, (XF & XF_STATS) != 0, (XF & XF_GEN) != 0, p.ring_cap);
  if (L.n_stages < TC_XF_WARPS || L.total > TC_SMEM_LIMIT) return SPANON_ERR_UNSUPPORTED;
  if ((XF & XF_GEN) && (slab_used > 144 || 2 * 143 + 17 > TC_GEN_SIG || p.gen_F > TC_GEN_MAXF || p.gen_F < 32 || p.cin_mod != 32 || p.cin_v != 64 || !p.gen_feats ||
                        !p.gen_w || !p.gen_b || (p.gen_mean == nullptr) != (p.gen_std == nullptr)))
    return SPANON_ERR_UNSUPPORTED;
  auto kern = tapgemm_tc_kernel<COUT_V, XF, MG>;
  SPANON_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total));
  const int mma_group = MG;
  if (nst % MG != 0 || L.n_stages % MG != 0) return SPANON_ERR_UNSUPPORTED;
  const int tiles_per_utt = (int)ceil_div(p.rows_out, tile_step);
  const int total = tiles_per_utt * p.B;
  const int grid = std::min(total, n_sm);
  static const bool trace_on = getenv("SPANON_TC_TRACE") && getenv("SPANON_TC_TRACE")[0] == '1';
  if (!trace_on) {
    SPANON_CUDA_TRY(launch_pdl(kern, dim3((unsigned)grid), dim3((unsigned)tc_threads(XF)), L.total, st, p.pdl != 0, p, halo_lo, slab_used,
                               tiles_per_utt, total, tile_step, mma_group));
    return check_launch(p.tag ? p.tag : "tapgemm_tc_kernel");
  }
  // debug: clock64 stamps of CTA 0's three roles, printed after a synchronising launch
  static long long* dbuf = nullptr;
  const size_t tbytes = (size_t)n_sm * 6 * TRACE_N * sizeof(long long);
  if (!dbuf) SPANON_CUDA_TRY(cudaMalloc(&dbuf, tbytes));
  SPANON_CUDA_TRY(cudaMemsetAsync(dbuf, 0, tbytes, st));
  TapGemmParams q = p;
  q.trace = dbuf;
  kern<<<grid, tc_threads(XF), L.total, st>>>(q, halo_lo, slab_used, tiles_per_utt, total, tile_step, mma_group);
  int rc = check_launch(p.tag ? p.tag : "tapgemm_tc_kernel");
  SPANON_CUDA_TRY(cudaStreamSynchronize(st));
  std::vector<long long> h((size_t)6 * TRACE_N);
  SPANON_CUDA_TRY(cudaMemcpy(h.data(), dbuf, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
  const long long t_begin = h[0];
  fprintf(stderr, "[trace %s] grid %d tiles %d stages/tile %d n_stages %d smem %u (cycles rel. to first transformed stage)\n",
          p.tag ? p.tag : "?", grid, total, nst, L.n_stages, L.total);
  const int tl[] = {0, 1, 2, 3, 16, 17, 18, 19, 20, 21};
  for (int t : tl) {
    if ((t + 1) * nst > TRACE_N || t >= (total + grid - 1) / grid - 1) break;
    fprintf(stderr, "  tile %d: load", t);
    for (int s2 = 0; s2 < nst; ++s2) fprintf(stderr, " %lld", h[3 * TRACE_N + t * nst + s2] - t_begin);
    fprintf(stderr, " | xform(raw>done)");
    for (int s2 = 0; s2 < nst; ++s2) fprintf(stderr, " %lld>%lld", h[4 * TRACE_N + t * nst + s2] - t_begin, h[t * nst + s2] - t_begin);
    fprintf(stderr, " | mma(full>commit)");
    for (int s2 = 0; s2 < nst; ++s2) fprintf(stderr, " %lld>%lld", h[5 * TRACE_N + t * nst + s2] - t_begin, h[TRACE_N + t * nst + s2] - t_begin);
    fprintf(stderr, " | epi %lld..%lld\n", h[2 * TRACE_N + 2 * t] - t_begin, h[2 * TRACE_N + 2 * t + 1] - t_begin);
  }
  return rc;
}

// Stages per issue round: 4 where the ring holds two tiles or more (16 stages), fewer where the resident weights leave
// room for one tile only (the 128 -> 128 k5 layers: a round then has to wait for -- and frees -- half the ring at once).
template <int COUT_V, int XF>
int launch_tc(const TapGemmParams& p, int n_sm, cudaStream_t st) {
  if constexpr (COUT_V == 16) return launch_tc_mg<COUT_V, XF, 2>(p, n_sm, st);
  else {
    int lo = 0, hi = 0;
    for (int t = 0; t < p.ntaps; ++t) { lo = std::min(lo, p.shift[t]); hi = std::max(hi, p.shift[t]); }
    const TcLayout L = tc_layout(p.ntaps, p.cin_v, COUT_V, TC_M - lo + hi, (XF & XF_STATS) != 0, (XF & XF_GEN) != 0, p.ring_cap);
    // measured per layer on B200 (same-box A/B): rounds of 4 stages pay on decoder.4 o 5 (7 taps) and tdnn.3 / tdnn.6
    // (-5 .. -12 us), and cost 3 .. 10 us on the layers whose epilogue reduces statistics or whose ring holds one tile
    if (p.mma_group >= 4 && L.n_stages >= 16) {
      if constexpr (XF == (XF_RELU | XF_T2) || (COUT_V == 64 && XF == (XF_SILU | XF_STATS))) return launch_tc_mg<COUT_V, XF, 4>(p, n_sm, st);
    }
    return launch_tc_mg<COUT_V, XF, 1>(p, n_sm, st);
  }
}

}  // namespace

namespace {
int tc_geometry(const TapGemmParams& p, int* halo_lo, int* slab_used) {
  int lo = 0, hi = 0;
  for (int t = 0; t < p.ntaps; ++t) { lo = std::min(lo, p.shift[t]); hi = std::max(hi, p.shift[t]); }
  *halo_lo = -lo;
  *slab_used = TC_M + *halo_lo + hi;
  return *slab_used <= 160 ? SPANON_OK : SPANON_ERR_UNSUPPORTED;
}
}  // namespace

// ps[0..n): the layers of a chain (ps[i+1] reads what ps[i] writes; flag_in / flag_out / flag_target set by the caller),
// all 128 -> 128 channels with an activation + folded BatchNorm on load; ctas[i] = CTAs working on layer i.
int launch_tapgemm_tc_chain(const TapGemmParams* ps, int n, const int* ctas, int rr_group, cudaStream_t st) {
  if (n < 1 || n > 3 || rr_group < 1) return SPANON_ERR_UNSUPPORTED;
  constexpr int XF = XF_DYN | XF_T2 | XF_FLOW;
  TcMulti mp{};
  mp.n_seg = n; mp.rr_group = rr_group; mp.mma_group = tc_mma_group(128);
  int cta = 0;
  uint32_t smem_max = 0;
  for (int i = 0; i < n; ++i) {
    const TapGemmParams& p = ps[i];
    if (!tapgemm_tc_supported(p) || p.B < 1 || p.cout_v != 128 || p.stats || p.gen_feats || !p.t2_scale || (p.silu != 1 && p.silu != 2) ||
        ctas[i] < 1)
      return SPANON_ERR_UNSUPPORTED;
    TcSeg& g = mp.seg[i];
    g.p = p;
    SPANON_TRY(tc_geometry(p, &g.halo_lo, &g.slab_used));
    const int nst = p.cin_v / TC_STAGE_CH;
    if (nst != 2 && nst != 4 && nst != 8) return SPANON_ERR_UNSUPPORTED;
    const TcLayout L = tc_layout(p.ntaps, p.cin_v, 128, g.slab_used, false, false, p.ring_cap);
    if (L.n_stages < TC_XF_WARPS || L.total > TC_SMEM_LIMIT) return SPANON_ERR_UNSUPPORTED;
    smem_max = std::max(smem_max, L.total);
    g.tile_step = TC_M;
    g.tiles_per_utt = (int)ceil_div(p.rows_out, TC_M);
    g.total_tiles = g.tiles_per_utt * p.B;
    g.cta0 = cta; g.n_cta = ctas[i];
    cta += ctas[i];
  }
  auto kern = tapgemm_tc_multi_kernel<128, XF, 2>;
  SPANON_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max));
  kern<<<cta, tc_threads(XF), smem_max, st>>>(mp);
  return check_launch(ps[0].tag ? ps[0].tag : "tapgemm_tc_chain");
}

bool tapgemm_tc_supported(const TapGemmParams& p) {
  if (p.out_mode == OUT_CH0_F32)   // Toeplitz form of a C -> 1 channel conv: one "tap" at shift -(k-1)/2, 16 columns = the taps
    return p.cin_v == 32 && p.cout_v == 16 && p.ntaps == 1 && p.shift[0] == -(TC_TOEP_TAPS - 1) / 2 && !p.stats &&
           p.tap_kmax[0] == p.cin_v && p.tap_nmax[0] == 16;
  if (p.cin_v % TC_STAGE_CH != 0 || (p.cout_v != 64 && p.cout_v != 128)) return false;
  if (p.ntaps < 2 || p.ntaps > MAX_TAPS) return false;
  for (int t = 0; t < p.ntaps; ++t) {   // the issue loop assumes equally spaced taps and the two zero-slice patterns
    if (p.shift[t] != p.shift[0] + t * p.tap_step) return false;
    if (t + 1 < p.ntaps && p.tap_kmax[t] != p.cin_v) return false;
    if (t > 0 && p.tap_nmax[t] != p.cout_v) return false;
  }
  if (p.tap_step < 1 || p.tap_kmax[p.ntaps - 1] % TC_STAGE_CH != 0 || p.tap_nmax[0] % 8 != 0) return false;
  if (p.out_fold == OUT_PAIR_FOLD && (p.rows_out & 1)) return false;
  if (p.fin_counter) return false;                           // finalize is a separate launch (instnorm_finalize_kernel)
  if (p.out_mode != OUT_CHANNELS_LAST) return false;        // fp16 chunk-planar output only
  return true;
}

int launch_tapgemm_tc(const TapGemmParams& p, int n_sm, cudaStream_t st) {
  if (!tapgemm_tc_supported(p) || p.B < 1) return SPANON_ERR_UNSUPPORTED;
  // one instantiation per (columns, operand transform) the layer lists use
  const int act = p.silu == 1 ? XF_SILU : p.silu == 2 ? XF_RELU : XF_IDENT;
  const int xf = act | (p.t2_scale ? XF_T2 : 0) | (p.stats ? XF_STATS : 0);
  if (act == XF_IDENT && (p.t1_scale || p.t2_scale)) return SPANON_ERR_UNSUPPORTED;
  if (p.gen_feats) {      // encoder.0 generated inside encoder.2's transform warps
    if (p.cout_v != 64 || act != XF_IDENT) return SPANON_ERR_UNSUPPORTED;
    return p.stats ? launch_tc<64, XF_GEN | XF_STATS>(p, n_sm, st) : launch_tc<64, XF_GEN>(p, n_sm, st);
  }
  if (p.cout_v == 16 && xf == XF_SILU) {
    // $SPANON_TOEP_EPI8=0: four epilogue warps (A/B switch)
    static const bool epi8 = !(getenv("SPANON_TOEP_EPI8") && getenv("SPANON_TOEP_EPI8")[0] == '0');
    return epi8 ? launch_tc<16, XF_SILU | XF_EPI8>(p, n_sm, st) : launch_tc<16, XF_SILU>(p, n_sm, st);
  }
  if (p.cout_v == 64 && xf == XF_IDENT) return launch_tc<64, XF_IDENT>(p, n_sm, st);
  if (p.cout_v == 64 && xf == XF_SILU) return launch_tc<64, XF_SILU>(p, n_sm, st);
  if (p.cout_v == 128 && xf == XF_IDENT) return launch_tc<128, XF_IDENT>(p, n_sm, st);
  if (p.cout_v == 128 && xf == XF_SILU) return launch_tc<128, XF_SILU>(p, n_sm, st);
  if (p.cout_v == 64 && xf == (XF_IDENT | XF_STATS)) return launch_tc<64, XF_IDENT | XF_STATS>(p, n_sm, st);
  if (p.cout_v == 64 && xf == (XF_SILU | XF_STATS)) return launch_tc<64, XF_SILU | XF_STATS>(p, n_sm, st);
  if (p.cout_v == 128 && xf == (XF_IDENT | XF_STATS)) return launch_tc<128, XF_IDENT | XF_STATS>(p, n_sm, st);
  if (p.cout_v == 128 && xf == (XF_SILU | XF_STATS)) return launch_tc<128, XF_SILU | XF_STATS>(p, n_sm, st);
  if (p.cout_v == 128 && xf == (XF_SILU | XF_T2)) return launch_tc<128, XF_SILU | XF_T2>(p, n_sm, st);
  if (p.cout_v == 128 && xf == (XF_RELU | XF_T2)) return launch_tc<128, XF_RELU | XF_T2>(p, n_sm, st);
  return SPANON_ERR_UNSUPPORTED;
}

}  // namespace spanon
