// Microbenchmark: cycles per tcgen05.mma (M=128, N, K=16, fp16) as a function of the shared-memory operand layout.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o umma_rate umma_rate.cu
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
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ void umma_acc(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t idesc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.b32 p, 0, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
struct Cfg { int n; uint32_t a_lbo, a_sbo, a_layout, a_shift_rows, b_lbo, b_sbo, b_layout; int a_k_adv, b_k_adv; };

__global__ void __launch_bounds__(128, 1) bench(Cfg c, int iters, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;  // 1.0h
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
  if (threadIdx.x == 0) {
    const uint32_t a0 = smem_u32(smem), b0 = a0 + 100 * 1024;
    const uint32_t idesc = (1u << 4) | ((uint32_t)(c.n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    uint32_t phase = 0;
    for (int rep = 0; rep < 3; ++rep) {
      long long t0;
      uint64_t ad[8], bd[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        ad[j] = make_desc(a0 + c.a_shift_rows * 16u * (uint32_t)(j % 5) + (uint32_t)(j & 3) * c.a_k_adv, c.a_lbo, c.a_sbo, c.a_layout);
        bd[j] = make_desc(b0 + (uint32_t)(j & 3) * c.b_k_adv, c.b_lbo, c.b_sbo, c.b_layout);
      }
      t0 = clock64();
      for (int i = 0; i < iters; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) umma_acc(tm, ad[j], bd[j], idesc);
      }
      long long t1 = clock64();
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      mbar_wait(smem_u32(&bar), phase); phase ^= 1;
      long long t2 = clock64();
      if (rep == 2) { out[blockIdx.x * 2] = t1 - t0; out[blockIdx.x * 2 + 1] = t2 - t0; }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(256) : "memory");
}

int main() {
  long long* d; cudaMalloc(&d, 148 * 2 * sizeof(long long));
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  struct Named { const char* name; Cfg c; } cfgs[] = {
    //                                   n   a_lbo a_sbo lay shift b_lbo b_sbo lay  a_kadv b_kadv
    {"noswz A lbo2128 shift0 / B lbo2048 N128", {128, 2128, 128, 0, 0, 2048, 128, 0, 4256, 4096}},
    {"noswz A lbo2128 shift1 / B lbo2048 N128", {128, 2128, 128, 0, 1, 2048, 128, 0, 4256, 4096}},
    {"noswz A lbo2048 shift0 / B lbo2048 N128", {128, 2048, 128, 0, 0, 2048, 128, 0, 4096, 4096}},
    {"noswz A lbo2128 shift1 / B lbo1024 N64 ", {64, 2128, 128, 0, 1, 1024, 128, 0, 4256, 2048}},
    {"noswz A lbo128 sbo256 (K-contig cores) N128", {128, 128, 256, 0, 0, 128, 256, 0, 32 * 256, 32 * 256}},
    {"sw128 A/B canonical N128              ", {128, 16, 1024, 2, 0, 16, 1024, 2, 32, 32}},
    {"sw128 A/B canonical N64               ", {64, 16, 1024, 2, 0, 16, 1024, 2, 32, 32}},
    {"sw128 A shift rows(8-row = 1024B*?)   ", {128, 16, 1024, 2, 64, 16, 1024, 2, 32, 32}},
    {"sw64  A/B N128                        ", {128, 16, 512, 4, 0, 16, 512, 4, 32, 32}},
    {"sw32  A/B N128                        ", {128, 16, 256, 6, 0, 16, 256, 6, 32, 32}},
  };
  for (auto& nc : cfgs) {
    for (int grid : {1, 148}) {
      const int iters = 2000;
      bench<<<grid, 128, 200 * 1024>>>(nc.c, iters, d);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
      printf("%-46s grid %3d: issue %.1f cyc/mma, complete %.1f cyc/mma  (%s)\n", nc.name, grid, (double)h[0] / iters, (double)h[1] / iters, cudaGetErrorString(e));
    }
  }
  return 0;
}
