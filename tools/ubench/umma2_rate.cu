// Microbenchmark: cycles per tcgen05.mma.cta_group::2 (M=256 over a CTA pair, N, K=16, fp16, no-swizzle K-major operands).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o umma2_rate umma2_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
__device__ __forceinline__ void umma2_acc(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t idesc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.b32 p, 0, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) bench(int n, int iters, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  uint32_t rank; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == 0) {
    if (lane == 0) { mbar_init(smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = tmem_ptr;
  if (threadIdx.x == 0 && rank == 0) {
    const uint32_t a0 = smem_u32(smem), b0 = a0 + 100 * 1024;
    const uint32_t idesc = (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
    const uint32_t b_lbo = (uint32_t)(n / 2) * 16u;                 // each CTA holds n/2 rows of B
    uint32_t phase = 0;
    for (int rep = 0; rep < 3; ++rep) {
      uint64_t ad[8], bd[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        ad[j] = make_desc(a0 + 16u * (uint32_t)(j % 5) + (uint32_t)(j & 3) * 4256u, 2128, 128);
        bd[j] = make_desc(b0 + (uint32_t)(j & 3) * 2u * b_lbo, b_lbo, 128);
      }
      long long t0 = clock64();
      for (int i = 0; i < iters; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) umma2_acc(tm, ad[j], bd[j], idesc);
      }
      long long t1 = clock64();
      asm volatile("{\n\t.reg .b16 m;\n\tmov.b16 m, 1;\n\ttcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t}" ::"r"(smem_u32(&bar)) : "memory");
      mbar_wait(smem_u32(&bar), phase); phase ^= 1;
      long long t2 = clock64();
      if (rep == 2) { out[blockIdx.x] = t1 - t0; out[blockIdx.x + 1] = t2 - t0; }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(256) : "memory");
}

int main() {
  long long* d; cudaMalloc(&d, 150 * sizeof(long long));
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int n : {64, 128, 256}) {
    for (int grid : {2, 148}) {
      const int iters = 2000;
      bench<<<grid, 128, 200 * 1024>>>(n, iters, d);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
      printf("cta_group::2 M=256 N=%3d grid %3d: issue %.1f cyc/mma, complete %.1f cyc/mma  (%s)\n", n, grid, (double)h[0] / iters, (double)h[1] / iters, cudaGetErrorString(e));
    }
  }
  return 0;
}
