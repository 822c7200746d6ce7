// Which part of the InstanceNorm statistics pass keeps it at ~3.5 TB/s?  Same grid/data shape, pieces switched on/off.
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
constexpr int NT = 256;
__device__ __forceinline__ void mma16816(float* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float warp_sum(float v) { for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(~0u, v, o); return v; }
template <int MATH, int END, int ROWS, int BATCH>
__global__ void __launch_bounds__(NT) k(const __half* __restrict__ x, int P, int R, int C, double* __restrict__ stats) {
  __shared__ float red[NT / 32][16];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, p = blockIdx.y, b = blockIdx.z;
  const uint4* base = reinterpret_cast<const uint4*>(x) + ((int64_t)b * P + p) * R;
  const int r0 = blockIdx.x * (NT * ROWS);
  const int g = lane >> 2;
  const uint32_t one_a = g == 0 ? 0x00003C00u : g == 1 ? 0x3C000000u : 0u;
  float sq[4][4] = {}, sm[4][4] = {};
  float fs[8] = {}, fq[8] = {};
  unsigned acc = 0;
#pragma unroll 1
  for (int i0 = 0; i0 < ROWS; i0 += BATCH) {
    uint4 v[BATCH];
#pragma unroll
    for (int u = 0; u < BATCH; ++u) { const int r = r0 + (i0 + u) * NT + tid; v[u] = make_uint4(0, 0, 0, 0); if (r < R) v[u] = __ldg(base + r); }
    if (MATH == 0) {
#pragma unroll
      for (int u = 0; u < BATCH; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    } else if (MATH == 1) {
#pragma unroll
      for (int u = 0; u < BATCH; u += 2) {
        const uint32_t wa[4] = {v[u].x, v[u].y, v[u].z, v[u].w}, wb[4] = {v[u + 1].x, v[u + 1].y, v[u + 1].z, v[u + 1].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          mma16816(sq[q], wa[q] & 0xFFFFu, wa[q] & 0xFFFF0000u, wb[q] & 0xFFFFu, wb[q] & 0xFFFF0000u, wa[q], wb[q]);
          mma16816(sm[q], one_a, 0u, one_a, 0u, wa[q], wb[q]);
        }
      }
    } else {
#pragma unroll
      for (int u = 0; u < BATCH; ++u) {
        const __half2* h = reinterpret_cast<const __half2*>(&v[u]);
#pragma unroll
        for (int j = 0; j < 4; ++j) { float2 f = __half22float2(h[j]); fs[2*j] += f.x; fs[2*j+1] += f.y; fq[2*j] = fmaf(f.x, f.x, fq[2*j]); fq[2*j+1] = fmaf(f.y, f.y, fq[2*j+1]); }
      }
    }
  }
  float s[8], q[8];
  for (int kk = 0; kk < 4; ++kk) { q[2*kk] = sq[kk][0] + fq[2*kk]; q[2*kk+1] = sq[kk][2] + fq[2*kk+1]; s[2*kk] = sm[kk][0] + fs[2*kk] + (float)acc; s[2*kk+1] = sm[kk][1] + fs[2*kk+1]; }
  if (END) {
    for (int j = 0; j < 8; ++j) { s[j] = warp_sum(s[j]); q[j] = warp_sum(q[j]); }
    if (lane == 0) for (int j = 0; j < 8; ++j) { red[warp][j] = s[j]; red[warp][8 + j] = q[j]; }
    __syncthreads();
    if (tid < 16) { float a = 0.f; for (int w = 0; w < NT / 32; ++w) a += red[w][tid]; const int c = (p % (C / 8)) * 8 + (tid & 7); atomicAdd(stats + ((int64_t)b * C + c) * 2 + (tid >> 3), (double)a); }
  } else {
    float a = 0; for (int j = 0; j < 8; ++j) a += s[j] + q[j];
    if (a == 12345.678f) stats[0] = a;
  }
}
int main() {
  const int B = 60, P = 8, R = 40320, C = 64;
  const size_t bytes = (size_t)B * P * R * 16;
  __half* x; double* st; cudaMalloc(&x, bytes); cudaMalloc(&st, B * C * 2 * 8); cudaMemset(x, 0, bytes); cudaMemset(st, 0, B * C * 16);
  // a second buffer to evict L2 between runs
  char* junk; cudaMalloc(&junk, 512 << 20);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto time = [&](auto launch, const char* name) {
    float best = 1e9;
    for (int it = 0; it < 4; ++it) { cudaMemsetAsync(junk, it, 512 << 20); cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; }
    printf("%-52s %.1f us  %.0f GB/s\n", name, best * 1e3, bytes / best / 1e6);
  };
#define RUN(M, E, ROWS, BATCH) time([&] { dim3 g((R + NT * ROWS - 1) / (NT * ROWS), P, B); k<M, E, ROWS, BATCH><<<g, NT>>>(x, P, R, C, st); }, "math=" #M " end=" #E " rows/thr=" #ROWS " batch=" #BATCH)
  RUN(0, 0, 16, 4); RUN(0, 0, 32, 8); RUN(0, 1, 32, 8); RUN(2, 0, 32, 8); RUN(2, 1, 16, 4); RUN(2, 1, 32, 8); RUN(1, 0, 32, 8); RUN(1, 1, 32, 8); RUN(1, 1, 32, 4); RUN(1, 1, 64, 8); RUN(0, 0, 158, 2); RUN(1, 1, 158, 2);
  return 0;
}
