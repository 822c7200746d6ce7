// Two-SM (cta_group::2) form of the tcgen05 tap GEMM for the 128-column layers (conv_tc.cu describes the one-SM kernel,
// the data layout and the warp roles; this file only states what differs).
//
// Why: an M=128 N=128 K=16 UMMA fetches 8 KB of operands from shared memory = 64 cycles at 128 B/clk, exactly its math
// time, and the bulk copies + the in-place operand transform use the same port (~100 cycles per 16-channel stage): the
// 128 -> 128 layers run at ~2x their math time on one SM.  A CTA PAIR (two SMs of a TPC, cluster of 2) executes one
// M=256 MMA over two adjacent row tiles: each CTA stages and transforms the A slab of ITS tile as before, but holds only
// HALF of the weights (64 of the 128 output columns, all taps) -- the tensor cores of both SMs read both halves -- so the
// B fetch per SM and MMA halves (6 KB -> 48 cycles < 64 of math) and 80 KB of shared memory are freed: the 128 -> 128 k5
// layers get a two-tile operand ring instead of one.
//
//   both CTAs   loaders, transform warps, epilogue warps exactly as in conv_tc.cu, on their own tile (tile 2j + rank)
//   leader      (cluster rank 0) the only MMA issuer: waits until BOTH CTAs' transform warps have filled a stage (the
//               peer's warps arrive on the leader's `full` barrier through the cluster address), issues
//               tcgen05.mma.cta_group::2 (M = 256: rows 0..127 from its own shared memory / TMEM, 128..255 from the
//               peer's at the same offsets), and commits with .multicast::cluster so that `empty` / `tmem_full` fire in
//               both CTAs.  The epilogue warps of both CTAs hand the accumulator back on the leader's `tmem_empty`.
#include <type_traits>
#include "tapgemm.cuh"
#include "tc_prims.cuh"
#include <cuda_fp16.h>
#include <algorithm>
#include <stdlib.h>

namespace spanon {
namespace {

using namespace tc;
constexpr int T2_N = 128;                   // output columns of the layer (both CTAs together)
constexpr int T2_NH = T2_N / 2;             // weight columns resident per CTA
constexpr int T2_XF_WARPS = 8, T2_LOAD_WARPS = 2, T2_STAGE_CH = 16, T2_MAX_STAGES = 32;
__host__ __device__ constexpr int t2_epi_warps(int xf) { return (xf & 24) ? 8 : 4; }     // 8 = X2_STATS, 16 = X2_POOL
__host__ __device__ constexpr int t2_threads(int xf) { return (t2_epi_warps(xf) + 12) * 32; }
#ifndef T2_MG_VALUE
#define T2_MG_VALUE 2
#endif
constexpr int T2_MG = T2_MG_VALUE;            // stages per MMA issue round (divides the 8 stages of a tile)
constexpr size_t T2_SMEM_LIMIT = 232448;
enum { X2_IDENT = 0, X2_SILU = 1, X2_RELU = 2, X2_T2 = 4, X2_STATS = 8, X2_POOL = 16 };
constexpr int T2_ST_PITCH = 17;
enum { C_RAW = 0, C_FULL = 32, C_EMPTY = 64, C_TFULL = 96, C_TEMPTY = 98, C_W = 100, C_TPTR = 102 };     // (up to 32 stages)

struct T2Layout { uint32_t w, epi, red, bars, a, total, w_bytes, stage_bytes, lbo_a; int n_stages, slab; };

__host__ __device__ inline T2Layout t2_layout(int ntaps, int cin_v, int slab_used, bool stats, bool pool, int max_stages) {
  T2Layout L;
  L.slab = (slab_used + 7) / 8 * 8 + 1;
  L.lbo_a = (uint32_t)L.slab * 16u;
  L.stage_bytes = (uint32_t)(T2_STAGE_CH / 8) * L.lbo_a;
  L.w_bytes = (uint32_t)ntaps * cin_v * T2_NH * 2u;
  L.w = 0;
  L.epi = L.w + L.w_bytes;                               // bias [128] fp32
  L.red = (L.epi + T2_N * 4u * (pool ? 3u : 1u) + 7u) / 8u * 8u;          // (pooling: bias, scale, shift)
  // statistics: one [32][17] transpose buffer per epilogue warp; pooling: the CTA's [128 bins][sum, sum of squares]
  L.bars = L.red + (stats ? 8u * 32u * T2_ST_PITCH * 4u : pool ? 2u * T2_N * 8u : 0u);
  L.a = (L.bars + 1024u + 127u) / 128u * 128u;
  const size_t avail = T2_SMEM_LIMIT > L.a ? T2_SMEM_LIMIT - L.a : 0;
  int ns = (int)(avail / L.stage_bytes);
  L.n_stages = (ns > max_stages ? max_stages : ns) / T2_XF_WARPS * T2_XF_WARPS;
  L.total = L.a + (uint32_t)L.n_stages * L.stage_bytes;
  return L;
}

__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at local offset `bar` (shared::cta address) of CTA `rank` of the cluster.  Default semantics
// (release at CTA scope, the form CUTLASS's ClusterBarrier::arrive uses): with .release.cluster every arrival cost the
// transform warps ~3 000 cycles; what the waiter's MMA reads is this SM's own shared memory through the async proxy, which
// the fence.proxy.async in front of the arrive orders.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar, uint32_t rank) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(bar), "r"(rank)
      : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {   // (plain wait, as CUTLASS's ClusterBarrier::wait)
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void umma2_f16_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc),
      "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc2_commit_elect(uint32_t bar) {     // arrives on `bar` in BOTH CTAs of the pair
  asm volatile(
      "{\n\t.reg .pred q;\n\t.reg .b16 m;\n\t"
      "mov.b16 m, 3;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t}" ::"r"(bar)
      : "memory");
}
// kind::f16 instruction descriptor: fp16 A/B, fp32 D, K-major, M = 256 (two CTAs x 128 rows)
__device__ __forceinline__ uint32_t umma2_idesc(int n) { return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(256 >> 4) << 24); }

__device__ __forceinline__ __half2 silu_from_half2(__half2 hh) {
  uint32_t t, in = *reinterpret_cast<uint32_t*>(&hh);
  asm("tanh.approx.f16x2 %0, %1;" : "=r"(t) : "r"(in));
  return __hfma2(hh, *reinterpret_cast<__half2*>(&t), hh);
}

template <int XF>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(t2_threads(XF), 1)
tapgemm_tc2_kernel(const __grid_constant__ TapGemmParams p, int halo_lo, int slab_used, int tiles_per_utt, int total_tiles, int max_stages) {
  extern __shared__ __align__(128) unsigned char smem[];
  const T2Layout L = t2_layout(p.ntaps, p.cin_v, slab_used, (XF & X2_STATS) != 0, (XF & X2_POOL) != 0, max_stages);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int E = t2_epi_warps(XF), MMA_WARP = E + 11;
  // tensor memory: two accumulator buffers
  constexpr int T2_TMEM_COLS = (XF & X2_POOL) ? 4 * T2_N : 2 * T2_N;      // (a power of two; the pooling sums use 128 more)
  const int wd = warp - E;
  const bool is_loader = wd == 3 || wd == 7, is_xform = wd >= 0 && wd < 10 && !is_loader;
  const uint32_t rank = cluster_rank();
  const bool leader = rank == 0;
  const uint32_t s_base = smem_u32(smem);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars);
  const uint32_t bar0 = s_base + L.bars;
  auto BAR = [&](int i) { return bar0 + 8u * (uint32_t)i; };
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(bars + C_TPTR);
  const int n_stage_tile = p.cin_v / T2_STAGE_CH;       // 8
  // contiguous run of tile PAIRS per cluster; this CTA works on tile 2*pair + rank (a pair's second tile may lie past
  // the end: it then has no valid rows -- nothing is loaded, everything is zero-filled, nothing is stored)
  const int n_pairs_total = (total_tiles + 1) / 2;
  const int n_clusters = (int)gridDim.x / 2, cid = (int)blockIdx.x / 2;
  const int pair_begin = (int)((int64_t)cid * n_pairs_total / n_clusters);
  const int n_my = (int)((int64_t)(cid + 1) * n_pairs_total / n_clusters) - pair_begin;
  auto locate = [&](int k, int& b, int& tin) {
    const int t = 2 * (pair_begin + k) + (int)rank;
    if (t >= total_tiles) { b = p.B - 1; tin = tiles_per_utt; }          // the empty tile
    else { b = t / tiles_per_utt; tin = t - b * tiles_per_utt; }
  };

  pdl_launch_dependents();      // (common.cuh: programmatic dependent launch)
  if (warp == MMA_WARP) {
    if (lane == 0) {
      for (int i = 0; i < L.n_stages; ++i) {
        mbar_init(BAR(C_RAW + i), 1);
        mbar_init(BAR(C_FULL + i), 2);        // one arrival per CTA of the pair (used in the leader only)
        mbar_init(BAR(C_EMPTY + i), 1);
      }
      for (int i = 0; i < 2; ++i) { mbar_init(BAR(C_TFULL + i), 1); mbar_init(BAR(C_TEMPTY + i), 2 * E); }
      mbar_init(BAR(C_W), 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      // this CTA's half of the weights: columns [64 * rank, +64) of every (tap, 8-channel chunk) slice, 1 KB each
      mbar_expect_tx(BAR(C_W), L.w_bytes);
      const int n_slices = p.ntaps * (p.cin_v / 8);
      for (int i = 0; i < n_slices; ++i)
        bulk_g2s(s_base + L.w + (uint32_t)i * (T2_NH * 16u),
                 reinterpret_cast<const char*>(p.w) + ((size_t)i * T2_N + (size_t)rank * T2_NH) * 16u, T2_NH * 16u, BAR(C_W));
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_ptr_smem)),
                 "r"(T2_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    mbar_wait(BAR(C_W), 0);                   // (before the cluster barrier below: the peer's half is then in place too)
  }
  if (warp < E) {
    float* ec = reinterpret_cast<float*>(smem + L.epi);
    for (int c = threadIdx.x; c < T2_N; c += E * 32) {
      ec[c] = p.bias ? p.bias[p.bias_v ? c : c % p.cout_mod] : 0.f;
      if constexpr ((XF & X2_POOL) != 0) { ec[T2_N + c] = p.e_scale[c]; ec[2 * T2_N + c] = p.e_shift[c]; }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                         // both CTAs' barriers initialised and weights landed before any remote access
  tc_fence_after();
  pdl_wait();                                 // only constants were touched so far; the planes come from the predecessor
  const uint32_t tmem_base = *tmem_ptr_smem;
  const int planes_in = p.cin_v / 8;

  if (is_loader) {
    const __half* in_base = reinterpret_cast<const __half*>(p.in);
    const int lw = wd >> 2;
    uint32_t stage = (uint32_t)lw, ph = 0;
    for (int k = 0; k < n_my; ++k) {
      int b, tin;
      locate(k, b, tin);
      const int g0 = tin * TC_M - halo_lo;
      const int lo = g0 < 0 ? -g0 : 0;
      const int hi = min(slab_used, p.rows_in - g0);
      const uint32_t bytes = hi > lo ? (uint32_t)(hi - lo) * 16u : 0u;
      const int64_t plane_stride = (int64_t)p.rows_in * 8;
      const __half* src = in_base + (((int64_t)b * planes_in + lw * (T2_STAGE_CH / 8)) * p.rows_in + (g0 + lo)) * 8;
      for (int s = lw; s < n_stage_tile; s += T2_LOAD_WARPS) {
        mbar_wait_cluster(BAR(C_EMPTY + stage), ph ^ 1u);
        mbar_expect_tx_elect(BAR(C_RAW + stage), bytes * (T2_STAGE_CH / 8));
        if (bytes) {
          const uint32_t dst = s_base + L.a + stage * L.stage_bytes + (uint32_t)lo * 16u;
          bulk_g2s_elect(dst, src, bytes, BAR(C_RAW + stage));
          bulk_g2s_elect(dst + L.lbo_a, src + plane_stride, bytes, BAR(C_RAW + stage));
        }
        src += (int64_t)(T2_LOAD_WARPS * T2_STAGE_CH / 8) * plane_stride;
        stage += T2_LOAD_WARPS;
        if (stage >= (uint32_t)L.n_stages) { stage -= (uint32_t)L.n_stages; ph ^= 1u; }
      }
    }
  } else if (is_xform) {
    const int xw = wd - (wd >> 2);
    constexpr bool do_silu = (XF & 3) == X2_SILU, do_relu = (XF & 3) == X2_RELU, has_t2 = (XF & X2_T2) != 0;
    constexpr bool ident = (XF & 7) == X2_IDENT;
    const bool has_t1 = p.t1_scale != nullptr;
    const float pre = do_silu ? 0.5f : 1.f;
    const int s_fix = xw % n_stage_tile, kstep = T2_XF_WARPS / n_stage_tile;
    uint32_t stage = (uint32_t)xw, ph = 0;
    const int cm = (s_fix * T2_STAGE_CH) % p.cin_mod;
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
    int b_const = -1;
    for (int k = xw / n_stage_tile; k < n_my; k += kstep) {
      int b, tin;
      locate(k, b, tin);
      const int g0 = tin * TC_M - halo_lo;
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
      mbar_wait(BAR(C_RAW + stage), ph);
      unsigned char* base = smem + L.a + stage * L.stage_bytes + (uint32_t)lane * 16u;
      if (!ident && hi > lo) {
        const bool tail = lane + 128 < slab_used;
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
              if (do_silu) y = silu_from_half2(y);
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
      if (lo > 0 || hi < slab_used) {
        for (int r = lane; r < slab_used; r += 32)
          if (r < lo || r >= hi) {
            *reinterpret_cast<uint4*>(base + (uint32_t)(r - lane) * 16u) = make_uint4(0u, 0u, 0u, 0u);
            *reinterpret_cast<uint4*>(base + L.lbo_a + (uint32_t)(r - lane) * 16u) = make_uint4(0u, 0u, 0u, 0u);
          }
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(BAR(C_FULL + stage), 0u);     // the LEADER's barrier counts both CTAs
      stage += T2_XF_WARPS;
      if (stage >= (uint32_t)L.n_stages) { stage -= (uint32_t)L.n_stages; ph ^= 1u; }
    }
  } else if (warp == MMA_WARP) {
    if (leader) {
      const uint32_t desc_hi = (uint32_t)(umma_desc(0, 0) >> 32);
      const uint32_t a_lo_const = (uint32_t)umma_desc(0, L.lbo_a);
      const uint32_t b_lo_const = (uint32_t)umma_desc(s_base + L.w, (uint32_t)T2_NH * 16u);
      const uint32_t b_tap = (uint32_t)(p.cin_v / 8) * T2_NH, b_stage = (uint32_t)(T2_STAGE_CH / 8) * T2_NH;
      const uint32_t dil = (uint32_t)p.tap_step;
      const uint32_t idesc = umma2_idesc(T2_N);
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      const int ntaps = p.ntaps;
      const int s_short = p.tap_kmax[p.ntaps - 1] / T2_STAGE_CH;
      uint32_t stage = 0, ph = 0;
      for (int k = 0; k < n_my; ++k) {
        const uint32_t buf = (uint32_t)k & 1u;
        mbar_wait_cluster(BAR(C_TEMPTY + buf), (((uint32_t)k >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_u + buf * (uint32_t)T2_N;
        // rounds of T2_MG stages: the waits (cluster-scope acquires; half of the arrivals come from the peer SM) and the
        // multicast commits cost several hundred cycles per round, a 16-channel stage is only ntaps MMAs of work
        for (int s = 0; s < n_stage_tile; s += T2_MG) {
#pragma unroll
          for (int g = 0; g < T2_MG; ++g) mbar_wait_cluster(BAR(C_FULL + stage + g), ph);
          tc_fence_after();
          uint32_t a_lo[T2_MG], b_lo[T2_MG];
#pragma unroll
          for (int g = 0; g < T2_MG; ++g) {
            a_lo[g] = a_lo_const + ((s_base + L.a + (stage + g) * L.stage_bytes) >> 4);
            b_lo[g] = b_lo_const + (uint32_t)(s + g) * b_stage;
          }
#pragma unroll
          for (int g = 0; g < T2_MG; ++g)
            umma2_f16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo[g], ((uint64_t)desc_hi << 32) | b_lo[g], idesc, (uint32_t)((s + g) != 0));
          // stages >= s_short skip the last tap (the non-existent sixth tap of a stride-2 k5 conv: pair-view layers only)
          const int nt = s >= s_short ? ntaps - 1 : ntaps;       // (s_short is a multiple of the group)
#pragma unroll 1
          for (int tap = 1; tap < nt; ++tap) {
#pragma unroll
            for (int g = 0; g < T2_MG; ++g) {
              a_lo[g] += dil;
              b_lo[g] += b_tap;
              umma2_f16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo[g], ((uint64_t)desc_hi << 32) | b_lo[g], idesc, 1u);
            }
          }
#pragma unroll
          for (int g = 0; g < T2_MG; ++g) tc2_commit_elect(BAR(C_EMPTY + stage + g));
          stage += T2_MG;
          if (stage >= (uint32_t)L.n_stages) { stage = 0; ph ^= 1u; }
        }
        tc2_commit_elect(BAR(C_TFULL + buf));
      }
    }
  } else if (warp < E) {
    const float* ec = reinterpret_cast<const float*>(smem + L.epi);
    const int quarter = warp & 3, chalf = warp >> 2;      // E = 8: two warps per TMEM lane quarter, half of the columns each
    const int mloc = quarter * 32 + lane;
    if constexpr ((XF & X2_POOL) != 0) {
      // ---- tdnn.6 + ReLU + BatchNorm + reshape + StatisticsPooling sums (ConvAutoEncoder.py:60-64): nothing is stored.
      // The reference pools the [128][rows_out] channels-first image REINTERPRETED as [rows_out][128]: value (row m,
      // channel c) lands in bin (c * rows_out + m) mod 128 = (m mod 128 + c * (rows_out mod 128)) mod 128, and m mod 128
      // is this thread's TMEM lane for every tile.  So running sums (value, square) per (lane, channel) over all tiles of one
      // utterance live in the tensor-memory columns the accumulators do not use, and only when the utterance changes are
      // they rotated into the CTA's 128 bins (shared-memory fp64 atomics), then one fp64 atomic per bin and CTA.
      // Channels c and c + 64 of one row fall into the SAME bin when rows_out is even (64 * rot = 0 mod 128), so one pair of
      // sums serves both: 2 x 64 tensor-memory columns.  The epilogue is paced by tcgen05.ld (~64 B/clk per SM): per tile it
      // reads the accumulator (64 KB) plus the sums (64 KB) -- sums per single channel (128 KB) ran at 158 us against 123 us
      // for the plain store epilogue, register-resident sums did not fit beside the transform warps' 96 registers.
      double* red = reinterpret_cast<double*>(smem + L.red);
      const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
      const int c_lo = chalf * 32;                          // eight warps: two per lane quarter, channels [c_lo, +32) and +64
      const uint32_t t_s = tmem_base + 2u * T2_N + lane_off, t_q = t_s + (uint32_t)(T2_N / 2);
      const int rot = p.rows_out & 127;
      auto pool_bar = [] { asm volatile("bar.sync 1, 256;" ::: "memory"); };
      {
        float z[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) z[j] = 0.f;
#pragma unroll
        for (int i = 0; i < 32; i += 8) { tmem_st8(t_s + (uint32_t)(c_lo + i), z); tmem_st8(t_q + (uint32_t)(c_lo + i), z); }
        tmem_st_wait();
        red[threadIdx.x] = 0.0;
        pool_bar();
      }
      auto flush = [&](int bb) {
        float z[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) z[j] = 0.f;
#pragma unroll 1
        for (int i = 0; i < 32; i += 8) {
          float s_[8], q[8];
          tmem_ld8_nowait(t_s + (uint32_t)(c_lo + i), s_);
          tmem_ld8_nowait(t_q + (uint32_t)(c_lo + i), q);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int bin = (mloc + (c_lo + i + j) * rot) & 127;
            atomicAdd(red + 2 * bin, (double)s_[j]);
            atomicAdd(red + 2 * bin + 1, (double)q[j]);
          }
          tmem_st8(t_s + (uint32_t)(c_lo + i), z);
          tmem_st8(t_q + (uint32_t)(c_lo + i), z);
        }
        tmem_st_wait();
        pool_bar();
        const int t = (int)threadIdx.x;
        atomicAdd(p.pool + (int64_t)bb * 256 + t, red[t]);
        red[t] = 0.0;
        pool_bar();
      };
      int b_prev = -1;
      for (int k = 0; k < n_my; ++k) {
        int b, tin;
        locate(k, b, tin);
        if (b_prev >= 0 && b != b_prev) flush(b_prev);
        b_prev = b;
        const float vmask = tin * TC_M + mloc < p.rows_out ? 1.f : 0.f;
        const uint32_t buf = (uint32_t)k & 1u;
        mbar_wait_cluster(BAR(C_TFULL + buf), ((uint32_t)k >> 1) & 1u);
        tc_fence_after();
        const uint32_t taddr = tmem_base + buf * (uint32_t)T2_N + lane_off;
        // four blocks of 8 channel pairs, software-pipelined over two register sets (tcgen05.wait::ld waits for ALL
        // outstanding loads, so the next block's loads are issued right after the wait for the current one)
        auto LD = [&](int c, float* va, float* vb, float* s_, float* q) {
          tmem_ld8_nowait(taddr + (uint32_t)c, va);
          tmem_ld8_nowait(taddr + (uint32_t)(c + 64), vb);
          tmem_ld8_nowait(t_s + (uint32_t)c, s_);
          tmem_ld8_nowait(t_q + (uint32_t)c, q);
        };
        auto tile = [&](auto masked_c) {
          constexpr bool masked = decltype(masked_c)::value;    // last tile of an utterance: rows past its end are masked out
          auto CP = [&](int c, float* va, float* vb, float* s_, float* q) {
#pragma unroll
            for (int j = 0; j < 8; j += 4) {
              // (the same constants through the read-only global path, __ldg, were measured 4 us slower)
              const float4 ba = *reinterpret_cast<const float4*>(ec + c + j), bb4 = *reinterpret_cast<const float4*>(ec + c + j + 64);
              const float4 sa = *reinterpret_cast<const float4*>(ec + T2_N + c + j), sb = *reinterpret_cast<const float4*>(ec + T2_N + c + j + 64);
              const float4 ha = *reinterpret_cast<const float4*>(ec + 2 * T2_N + c + j), hb = *reinterpret_cast<const float4*>(ec + 2 * T2_N + c + j + 64);
              // packed fp32 pairs (FADD2 / FFMA2): 12 issue slots per 4 values instead of 20
#pragma unroll
              for (int u = 0; u < 4; u += 2) {
                const float2 ba2 = u ? make_float2(ba.z, ba.w) : make_float2(ba.x, ba.y);
                const float2 bb2 = u ? make_float2(bb4.z, bb4.w) : make_float2(bb4.x, bb4.y);
                const float2 sa2 = u ? make_float2(sa.z, sa.w) : make_float2(sa.x, sa.y);
                const float2 sb2 = u ? make_float2(sb.z, sb.w) : make_float2(sb.x, sb.y);
                const float2 ha2 = u ? make_float2(ha.z, ha.w) : make_float2(ha.x, ha.y);
                const float2 hb2 = u ? make_float2(hb.z, hb.w) : make_float2(hb.x, hb.y);
                float2 ra = fadd2(make_float2(va[j + u], va[j + u + 1]), ba2);
                float2 rb = fadd2(make_float2(vb[j + u], vb[j + u + 1]), bb2);
                ra.x = fmaxf(ra.x, 0.f); ra.y = fmaxf(ra.y, 0.f); rb.x = fmaxf(rb.x, 0.f); rb.y = fmaxf(rb.y, 0.f);
                float2 xa = ffma2(ra, sa2, ha2), xb = ffma2(rb, sb2, hb2);
                if constexpr (masked) { const float2 m2 = make_float2(vmask, vmask); xa = fmul2(xa, m2); xb = fmul2(xb, m2); }
                const float2 sn = fadd2(make_float2(s_[j + u], s_[j + u + 1]), fadd2(xa, xb));
                const float2 qn = ffma2(xb, xb, ffma2(xa, xa, make_float2(q[j + u], q[j + u + 1])));
                s_[j + u] = sn.x; s_[j + u + 1] = sn.y; q[j + u] = qn.x; q[j + u + 1] = qn.y;
              }
            }
            tmem_st8(t_s + (uint32_t)c, s_);
            tmem_st8(t_q + (uint32_t)c, q);
          };
          float a0[8], b0[8], s0[8], q0[8], a1[8], b1[8], s1[8], q1[8];
          LD(c_lo, a0, b0, s0, q0);
#pragma unroll
          for (int i = 0; i < 4; i += 2) {
            tmem_ld_wait();
            LD(c_lo + 8 * (i + 1), a1, b1, s1, q1);
            CP(c_lo + 8 * i, a0, b0, s0, q0);
            tmem_ld_wait();
            if (i + 2 < 4) LD(c_lo + 8 * (i + 2), a0, b0, s0, q0);
            else {                                              // every accumulator column of this warp has been read
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive_cluster(BAR(C_TEMPTY + buf), 0u);
            }
            CP(c_lo + 8 * (i + 1), a1, b1, s1, q1);
          }
        };
        if ((tin + 1) * TC_M > p.rows_out) tile(std::true_type{});
        else tile(std::false_type{});
        tmem_st_wait();
      }
      if (b_prev >= 0) flush(b_prev);
    } else {
    constexpr bool e_stats = (XF & X2_STATS) != 0;
    constexpr int NCH = T2_N / 8, CPW = T2_N / (E / 4);
    // InstanceNorm statistics of the output (X2_STATS), as in conv_tc.cu: column sums over the warp's 32 rows of the values
    // AS STORED through a padded shared-memory transpose; lane l keeps column cb + 2*(l & 15) + (l >> 4) of each 32-column block
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
    int b_prev = -1;
    for (int k = 0; k < n_my; ++k) {
      int b, tin;
      locate(k, b, tin);
      if constexpr (e_stats) { if (b_prev >= 0 && b != b_prev) e_flush(b_prev); }
      b_prev = b;
      const int m = tin * TC_M + mloc;
      const bool valid = m < p.rows_out;
      const uint32_t buf = (uint32_t)k & 1u;
      mbar_wait_cluster(BAR(C_TFULL + buf), ((uint32_t)k >> 1) & 1u);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * (uint32_t)T2_N + ((uint32_t)(quarter * 32) << 16);
      __half* out_h = reinterpret_cast<__half*>(p.out) + (int64_t)b * p.out_bstride;
#pragma unroll 1
      for (int cb = chalf * CPW; cb < (chalf + 1) * CPW; cb += 32) {
        float v[32];
        tmem_ld32(taddr + (uint32_t)cb, v);
        if (cb + 32 >= (chalf + 1) * CPW) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(BAR(C_TEMPTY + buf), 0u);
        }
        if (p.bias) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 bz = *reinterpret_cast<const float4*>(ec + cb + j);
            v[j] += bz.x; v[j + 1] += bz.y; v[j + 2] += bz.z; v[j + 3] += bz.w;
          }
        }
        uint32_t hw[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const __half2 h2 = __floats2half2_rn(v[2 * j], v[2 * j + 1]);
          hw[j] = *reinterpret_cast<const uint32_t*>(&h2);
        }
        if (valid) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint4 o = make_uint4(hw[4 * q], hw[4 * q + 1], hw[4 * q + 2], hw[4 * q + 3]);
            const int c8 = (cb >> 3) + q;
            int pl = c8, row = m;          // (output views other than the direct one only occur with statistics)
            if constexpr (e_stats) {
              if (p.out_fold == OUT_PAIR_FOLD) { pl = (m & 1) * NCH + c8; row = m >> 1; }
              else if (p.out_fold == OUT_PHASE_UNFOLD) { pl = c8 % (NCH / 2); row = 2 * m + c8 / (NCH / 2); }
              else if (p.out_fold == OUT_PHASE_UNFOLD4) { pl = c8 % (NCH / 4); row = 4 * m + c8 / (NCH / 4); }
            }
            *reinterpret_cast<uint4*>(out_h + (((int64_t)pl * p.out_plane_rows + row) << 3)) = o;
          }
        }
        if constexpr (e_stats) {
          uint32_t* tb = reinterpret_cast<uint32_t*>(smem + L.red) + warp * (32 * T2_ST_PITCH);
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 16; ++j) tb[lane * T2_ST_PITCH + j] = valid ? hw[j] : 0u;
          __syncwarp();
          const uint32_t* col = tb + ((lane >> 4) * 16) * T2_ST_PITCH + (lane & 15);
          float a0, a1, q0, q1;
          {   // packed fp32 pairs (FADD2 / FFMA2): two issue slots per word instead of four
            float2 sa = make_float2(0.f, 0.f), sq = make_float2(0.f, 0.f);
#pragma unroll
            for (int r = 0; r < 16; ++r) {
              const uint32_t w = col[r * T2_ST_PITCH];
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
    }
    if constexpr (e_stats) { if (b_prev >= 0) e_flush(b_prev); }
    }   // !X2_POOL
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // neither CTA may leave (or free tensor memory) while its peer can still touch it
  if (warp == MMA_WARP) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(T2_TMEM_COLS) : "memory");
  }
}

template <int XF>
int launch_tc2(const TapGemmParams& p, int n_sm, cudaStream_t st) {
  int lo = 0, hi = 0;
  for (int t = 0; t < p.ntaps; ++t) { lo = std::min(lo, p.shift[t]); hi = std::max(hi, p.shift[t]); }
  const int halo_lo = -lo, slab_used = TC_M + halo_lo + hi;
  if (slab_used > 160 || p.cin_v != 128) return SPANON_ERR_UNSUPPORTED;
  // ring depth: up to T2_MAX_STAGES 16-channel stages (8 per tile); $SPANON_T2_STAGES caps it (A/B: 16 = two tiles)
  static const int max_stages = [] { const char* e = getenv("SPANON_T2_STAGES"); const int v = e ? atoi(e) : T2_MAX_STAGES; return v >= 8 && v <= T2_MAX_STAGES ? v : T2_MAX_STAGES; }();
  const T2Layout L = t2_layout(p.ntaps, p.cin_v, slab_used, (XF & X2_STATS) != 0, (XF & X2_POOL) != 0, max_stages);
  if (L.n_stages < T2_XF_WARPS || L.total > T2_SMEM_LIMIT) return SPANON_ERR_UNSUPPORTED;
  auto kern = tapgemm_tc2_kernel<XF>;
  SPANON_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total));
  const int tiles_per_utt = (int)ceil_div(p.rows_out, TC_M);
  const int total = tiles_per_utt * p.B;
  const int pairs = (total + 1) / 2;
  const int grid = 2 * std::max(1, std::min(pairs, n_sm / 2));
  SPANON_CUDA_TRY(launch_pdl(kern, dim3((unsigned)grid), dim3((unsigned)t2_threads(XF)), L.total, st, p.pdl != 0, p, halo_lo, slab_used, tiles_per_utt, total, max_stages));
  return check_launch(p.tag ? p.tag : "tapgemm_tc2_kernel");
}

}  // namespace

// the layers with 128 output columns and 128 (view) input channels: tdnn.0/3/6, decoder.0, and with output statistics
// encoder.8 (stride 2: pair view), encoder.11, decoder.1 (transposed: [even | odd] phase columns)
static int tc2_xf(const TapGemmParams& p) {
  const int act = p.silu == 1 ? X2_SILU : p.silu == 2 ? X2_RELU : X2_IDENT;
  return act | (p.t2_scale ? X2_T2 : 0) | (p.stats ? X2_STATS : 0) | (p.pool ? X2_POOL : 0);
}

bool tapgemm_tc2_supported(const TapGemmParams& p) {
  if (!tapgemm_tc_supported(p)) return false;
  if (p.cin_v != 128 || p.cout_v != 128 || p.gen_feats) return false;
  if (p.flag_in || p.flag_out) return false;
  if (p.pool) {         // tdnn.6 with the pooling sums in its epilogue: ReLU -> BatchNorm affine, no stored output
    if (!p.relu || !p.e_scale || !p.e_shift || p.stats || p.cout_mod != 128 || (p.rows_out & 1)) return false;
  } else if (p.relu || p.e_scale) return false;
  if (p.stats && (p.cout_mod < 32 || 128 % p.cout_mod != 0)) return false;
  static const bool unfold4_env = !(getenv("SPANON_TC2_UNFOLD4") && getenv("SPANON_TC2_UNFOLD4")[0] == '0');
  if (p.out_fold == OUT_PHASE_UNFOLD4 && !unfold4_env) return false;
  if (!p.stats && (p.cout_mod != 128 || p.out_fold != OUT_DIRECT)) return false;
  // measured per layer (same-box A/B): the pair kernel wins on the k5 layers (tdnn.0 -21, decoder.0 -20, encoder.11 -14 us)
  // and on tdnn.3 / tdnn.6 (-5 us); the 3-tap layers that also reduce statistics (encoder.8, decoder.1) are HBM- and
  // epilogue-paced and lose 12 .. 20 us to the pair's lock-step, so they stay on the one-SM kernel
  static const bool stats3_env = getenv("SPANON_TC2_STATS3") && getenv("SPANON_TC2_STATS3")[0] == '1';
  if (p.stats && p.ntaps < 5 && !stats3_env) return false;
  for (int t = 0; t + 1 < p.ntaps; ++t)
    if (p.tap_kmax[t] != p.cin_v) return false;
  if (p.tap_kmax[p.ntaps - 1] % (T2_STAGE_CH * T2_MG) != 0) return false;
  // (tap_nmax[0] < 128 -- the first shift of a transposed conv feeds one phase only -- is not exploited: the skipped
  //  columns hold zero weights)
  const int xf = tc2_xf(p);
  return xf == X2_SILU || xf == (X2_SILU | X2_T2) || xf == (X2_RELU | X2_T2) || xf == (X2_SILU | X2_STATS) ||
         xf == (X2_IDENT | X2_STATS) || xf == (X2_RELU | X2_T2 | X2_POOL);
}

int launch_tapgemm_tc2(const TapGemmParams& p, int n_sm, cudaStream_t st) {
  if (!tapgemm_tc2_supported(p) || p.B < 1 || n_sm < 2) return SPANON_ERR_UNSUPPORTED;
  const int xf = tc2_xf(p);
  if (xf == X2_SILU) return launch_tc2<X2_SILU>(p, n_sm, st);
  if (xf == (X2_SILU | X2_T2)) return launch_tc2<X2_SILU | X2_T2>(p, n_sm, st);
  if (xf == (X2_RELU | X2_T2)) return launch_tc2<X2_RELU | X2_T2>(p, n_sm, st);
  if (xf == (X2_SILU | X2_STATS)) return launch_tc2<X2_SILU | X2_STATS>(p, n_sm, st);
  if (xf == (X2_IDENT | X2_STATS)) return launch_tc2<X2_IDENT | X2_STATS>(p, n_sm, st);
  if (xf == (X2_RELU | X2_T2 | X2_POOL)) return launch_tc2<X2_RELU | X2_T2 | X2_POOL>(p, n_sm, st);
  return SPANON_ERR_UNSUPPORTED;
}

}  // namespace spanon
