// Shared helpers for libspanon (sm_100a only).
#pragma once
#include <stdlib.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/spanon.h"

namespace spanon {

extern std::atomic<uint64_t> g_launch_count;
extern std::atomic<int> g_profile_on;
extern const bool g_debug_sync;
void set_last_cuda_error(const char* what, cudaError_t e);
// Kernel-time profile (spanon_profile_begin/_end): an event after every launch on the profiled stream.
void profile_mark(const char* what);

// First statement of an entry point that launches kernels: while a profile is on, the time from the previous mark to this
// call (Python / host latency on an idle stream) is booked under "(host)" instead of inflating the first kernel's time.
inline void profile_entry() {
  if (g_profile_on.load(std::memory_order_relaxed)) profile_mark("(host)");
}

inline int check_launch(const char* what) {
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  if (g_profile_on.load(std::memory_order_relaxed)) profile_mark(what);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && g_debug_sync) {   // $SPANON_DEBUG_SYNC=1: synchronise and report after every launch
    fprintf(stderr, "[spanon] launched %s ... ", what);
    fflush(stderr);
    e = cudaDeviceSynchronize();
    fprintf(stderr, "%s\n", cudaGetErrorString(e));
  }
  if (e != cudaSuccess) {
    set_last_cuda_error(what, e);
    return SPANON_ERR_CUDA;
  }
  return SPANON_OK;
}

#define SPANON_CUDA_TRY(expr)                                   \
  do {                                                          \
    cudaError_t _e = (expr);                                    \
    if (_e != cudaSuccess) {                                    \
      ::spanon::set_last_cuda_error(#expr, _e);                 \
      return SPANON_ERR_CUDA;                                   \
    }                                                           \
  } while (0)

#define SPANON_TRY(expr)            \
  do {                              \
    int _s = (expr);                \
    if (_s != SPANON_OK) return _s; \
  } while (0)

// Order-preserving float <-> int32 key so that atomicMax(int) orders floats.
__host__ __device__ inline int32_t float_to_key(float f) {
#ifdef __CUDA_ARCH__
  int32_t i = __float_as_int(f);
#else
  int32_t i; memcpy(&i, &f, 4);
#endif
  return i >= 0 ? i : (i ^ 0x7FFFFFFF);
}
__host__ __device__ inline float key_to_float(int32_t k) {
  int32_t i = k >= 0 ? k : (k ^ 0x7FFFFFFF);
#ifdef __CUDA_ARCH__
  return __int_as_float(i);
#else
  float f; memcpy(&f, &i, 4); return f;
#endif
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float silu_precise(float x) { return x / (1.0f + expf(-x)); }
// x*sigmoid(x) = h + h*tanh(h), h = x/2 : one MUFU op (tanh.approx), ~2^-11 rel. error.
__device__ __forceinline__ float silu_fast(float x) {
  float h = 0.5f * x, t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  return fmaf(h, t, h);
}

__device__ __forceinline__ float to_float(float v) { return v; }
__device__ __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_float(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_float(float v);
template <> __device__ __forceinline__ float from_float<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

template <> __device__ __forceinline__ __half from_float<__half>(float v) { return __float2half_rn(v); }

// ---- programmatic dependent launch (PDL).  A kernel launched through launch_pdl() may begin while its predecessor on the
// stream is still running: its CTAs become resident as the predecessor's CTAs exit, run their prologue (barrier init, tensor
// memory allocation, weight copies -- nothing the predecessor produces) and then block in pdl_wait() until the predecessor
// grid has COMPLETED and its memory is visible.  Every kernel launched this way must call pdl_wait() before its first read or
// write of anything an earlier kernel touches; completion is transitive along the chain.  $SPANON_PDL=0: plain launches.
inline bool pdl_enabled() {
  static const bool on = !(getenv("SPANON_PDL") && getenv("SPANON_PDL")[0] == '0');
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = (pdl && pdl_enabled()) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#ifdef __CUDACC__
// (both are no-ops in a kernel launched without the attribute)
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Loads through `const T* __restrict__` (ld.global.nc) are "invariant" to the compiler, and it hoists them ABOVE the wait --
// asm memory clobber or not (seen in clf_head_kernel's SASS: the pooled sums were read before ACQBULK, i.e. before the
// producer had written them).  A pointer to data the predecessor produces must therefore be re-derived AFTER pdl_wait():
//   pdl_wait(); p = pdl_acquire(p);
template <typename T>
__device__ __forceinline__ T* pdl_acquire(T* p) {
  asm volatile("" : "+l"(p)::"memory");
  return p;
}
#endif

// Per-device one-time setup (cudaFuncSetAttribute is a per-device setting, and one process may drive several GPUs):
// returns true the first time it is called with `mask` while device `dev` is current.  Thread-safe.
inline bool first_use_on_device(std::atomic<unsigned long long>& mask, int dev) {
  const unsigned long long bit = 1ull << (dev & 63);
  return (mask.fetch_or(bit, std::memory_order_acq_rel) & bit) == 0;
}
// Multiprocessor count of the CURRENT device (cached per device).
inline int current_device_sms(int* dev_out = nullptr) {
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev_out) *dev_out = dev;
  int n = cache[dev & 63].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    cache[dev & 63].store(n, std::memory_order_relaxed);
  }
  return n;
}

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace spanon
