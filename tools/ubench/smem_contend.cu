// Microbenchmark: does tcgen05.mma operand traffic (SS mode) contend with LDS/STS of other warps for shared memory?
// warp 0 issues MMAs (M=128,N,K=16 fp16, no-swizzle K-major) back to back; warps 1..W stream LDS.128+STS.128 in place.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory"); } while (!done);
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46);
}
__device__ __forceinline__ void umma_acc(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t idesc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.b32 p, 0, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
__global__ void __launch_bounds__(544, 1) bench(int n, int mma_iters, int ls_iters, int do_mma, int ls_warps, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == 0) {
    if (lane == 0) { mbar_init(smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = tmem_ptr;
  long long t0 = clock64();
  if (warp == 0) {
    if (lane == 0 && do_mma) {
      const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
      const uint32_t idesc = (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      uint64_t ad[8], bd[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) { ad[j] = make_desc(a0 + (j % 5) * 16u + (j & 3) * 4256u, 2128, 128); bd[j] = make_desc(b0 + (j & 3) * 4096u, n * 16, 128); }
      for (int i = 0; i < mma_iters; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) umma_acc(tm, ad[j], bd[j], idesc);
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      mbar_wait(smem_u32(&bar), 0);
      out[blockIdx.x * 4 + 0] = clock64() - t0;
    }
  } else if (warp <= ls_warps) {
    // each warp streams over its own 8 KB region: LDS.128 -> STS.128 (conflict-free, 512 B per warp instruction)
    uint4* base = reinterpret_cast<uint4*>(smem + 128 * 1024 + (warp - 1) * 4096);
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (int i = 0; i < ls_iters; ++i) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        uint4 v;
        const uint32_t a = smem_u32(base + q * 32 + lane);
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
        v.x += acc.x; acc.x ^= v.y;
        asm volatile("st.shared.v4.u32 [%4], {%0,%1,%2,%3};" :: "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(a) : "memory");
      }
    }
    if (acc.x == 0x12345) out[100000] = 1;
    if (warp == 1 && lane == 0) out[blockIdx.x * 4 + 1] = clock64() - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(256) : "memory");
}
int main() {
  long long* d; cudaMalloc(&d, 200000 * sizeof(long long));
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int mma_iters = 4000, ls_iters = 400;
  for (int n : {128, 64}) for (int do_mma : {1, 0}) for (int lw : {0, 4, 8, 16}) {
    if (!do_mma && lw == 0) continue;
    cudaMemset(d, 0, 64);
    bench<<<1, 544, 200 * 1024>>>(n, mma_iters, ls_iters, do_mma, lw, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[4]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    // LSU bytes per cycle: each iter per warp = 8 * (512 B read + 512 B write)
    double ls_bpc = lw ? (double)lw * ls_iters * 8 * 1024 / (double)h[1] : 0;
    printf("N=%3d mma=%d ls_warps=%2d : %.1f cyc/mma, LSU smem %.1f B/clk (ls done in %lld cyc)  %s\n", n, do_mma, lw, do_mma ? (double)h[0] / mma_iters : 0.0, ls_bpc, h[1], cudaGetErrorString(e));
  }
  return 0;
}
