// mma.sync.m16n8k8 TF32 issue rate on sm_100a (legacy warp-level MMA): cycles per MMA per warp with W warps per SM.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_tf32_rate mma_tf32_rate.cu && ./mma_tf32_rate
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* out, int iters, long long* cyc) {
  float d[8][4];
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) d[i][j] = 0.f;
  unsigned a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, b0 = a0 * 3, b1 = a0 * 5;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(d[i][0]), "+f"(d[i][1]), "+f"(d[i][2]), "+f"(d[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += d[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  for (int warps : {1, 2, 4, 8, 16, 32}) {
    const int iters = 2000;
    k<<<148, warps * 32>>>(out, iters, cyc); cudaDeviceSynchronize();
    k<<<148, warps * 32>>>(out, iters, cyc); cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double mma_per_sm = (double)warps * iters * 8;
    printf("warps/SM %2d: %.2f cycles per MMA per SM  -> %.0f MAC/clk/SM (m16n8k8 = 1024 MAC)\n", warps, c / mma_per_sm, 1024.0 * mma_per_sm / c);
  }
  return 0;
}
