// Microbenchmark: per-SM throughput of MUFU variants on B200 (results per clock per SM).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
template <int OP>
__global__ void __launch_bounds__(1024, 1) k(float* out, int iters) {
  float x0 = threadIdx.x * 1e-3f, x1 = x0 + 0.1f, x2 = x0 + 0.2f, x3 = x0 + 0.3f;
  uint32_t u0 = threadIdx.x | 0x38003800u, u1 = u0 + 1, u2 = u0 + 2, u3 = u0 + 3;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    if (OP == 0) { asm volatile("tanh.approx.f32 %0, %0;" : "+f"(x0)); asm volatile("tanh.approx.f32 %0, %0;" : "+f"(x1)); asm volatile("tanh.approx.f32 %0, %0;" : "+f"(x2)); asm volatile("tanh.approx.f32 %0, %0;" : "+f"(x3)); }
    if (OP == 1) { asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(u0)); asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(u1)); asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(u2)); asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(u3)); }
    if (OP == 2) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x0)); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x1)); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x2)); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x3)); }
    if (OP == 3) { asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x0)); asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x1)); asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x2)); asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x3)); }
    if (OP == 4) { asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(u0)); asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(u1)); asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(u2)); asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(u3)); }
    if (OP == 5) { asm volatile("tanh.approx.f16 %0, %0;" : "+h"(*(unsigned short*)&u0)); asm volatile("tanh.approx.f16 %0, %0;" : "+h"(*(unsigned short*)&u1)); asm volatile("tanh.approx.f16 %0, %0;" : "+h"(*(unsigned short*)&u2)); asm volatile("tanh.approx.f16 %0, %0;" : "+h"(*(unsigned short*)&u3)); }
  }
  long long t1 = clock64();
  out[blockIdx.x * 1024 + threadIdx.x] = x0 + x1 + x2 + x3 + __uint_as_float(u0 ^ u1 ^ u2 ^ u3);
  if (threadIdx.x == 0) reinterpret_cast<long long*>(out + 200000)[blockIdx.x] = t1 - t0;
}
int main() {
  float* d; cudaMalloc(&d, 300000 * 4);
  const int iters = 2000;
  const char* names[] = {"tanh.approx.f32", "tanh.approx.f16x2 (instr)", "ex2.approx.f32", "rcp.approx.f32", "ex2.approx.f16x2 (instr)", "tanh.approx.f16"};
  for (int op = 0; op < 6; ++op) {
    switch (op) { case 0: k<0><<<1, 1024>>>(d, iters); break; case 1: k<1><<<1, 1024>>>(d, iters); break; case 2: k<2><<<1, 1024>>>(d, iters); break;
                  case 3: k<3><<<1, 1024>>>(d, iters); break; case 4: k<4><<<1, 1024>>>(d, iters); break; case 5: k<5><<<1, 1024>>>(d, iters); break; }
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, d + 200000, 8, cudaMemcpyDeviceToHost);
    printf("%-28s %.2f thread-instr / clk / SM\n", names[op], 1024.0 * 4 * iters / (double)c);
  }
  return 0;
}
