// Pure-read HBM bandwidth with LDG.128 for several launch shapes (what a reduction pass can hope for).
#include <cstdio>
#include <cuda_runtime.h>
template <int UNROLL>
__global__ void rd(const uint4* __restrict__ p, size_t n, unsigned* out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  unsigned acc = 0;
  for (; i + (UNROLL - 1) * stride < n; i += UNROLL * stride) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) v[u] = __ldg(p + i + u * stride);
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) *out = acc;
}
// block-contiguous variant: each block streams its own contiguous segment (like the stats kernel)
template <int UNROLL>
__global__ void rd_seg(const uint4* __restrict__ p, size_t per_block, unsigned* out) {
  const uint4* b = p + (size_t)blockIdx.x * per_block;
  unsigned acc = 0;
  for (size_t i = threadIdx.x; i + (UNROLL - 1) * blockDim.x < per_block; i += UNROLL * blockDim.x) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) v[u] = __ldg(b + i + u * blockDim.x);
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) *out = acc;
}
int main() {
  const size_t bytes = (size_t)1 << 30; const size_t n = bytes / 16;
  uint4* d; unsigned* o; cudaMalloc(&d, bytes); cudaMalloc(&o, 4); cudaMemset(d, 1, bytes);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto time = [&](auto launch, const char* name) {
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(e0); for (int i = 0; i < 5; ++i) launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
    printf("%-40s %.3f ms  %.0f GB/s\n", name, ms, bytes / ms / 1e6);
  };
  for (int bps : {2, 4, 8}) {
    char nm[64];
    snprintf(nm, 64, "grid-stride unroll4 %d blk/SM x256", bps);  time([&] { rd<4><<<148 * bps, 256>>>(d, n, o); }, nm);
    snprintf(nm, 64, "grid-stride unroll8 %d blk/SM x256", bps);  time([&] { rd<8><<<148 * bps, 256>>>(d, n, o); }, nm);
  }
  time([&] { rd<8><<<148 * 2, 1024>>>(d, n, o); }, "grid-stride unroll8 2 blk/SM x1024");
  for (size_t kb : {64, 128, 512, 2048}) {
    char nm[64]; size_t per = kb * 1024 / 16; unsigned blocks = (unsigned)(n / per);
    snprintf(nm, 64, "segments %zu KB/block unroll8 x256", kb); time([&] { rd_seg<8><<<blocks, 256>>>(d, per, o); }, nm);
  }
  // read + write copy for reference
  uint4* d2; cudaMalloc(&d2, bytes);
  time([&] { cudaMemcpyAsync(d2, d, bytes, cudaMemcpyDeviceToDevice); }, "cudaMemcpy D2D (bytes counted once)");
  return 0;
}
