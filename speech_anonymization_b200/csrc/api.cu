// Library-level entry points: status strings, ABI version, launch counter.
#include "common.cuh"
#include <mutex>
#include <string>

namespace spanon {
std::atomic<uint64_t> g_launch_count{0};
static std::mutex g_err_mu;
static std::string g_last_err;
void set_last_cuda_error(const char* what, cudaError_t e) {
  std::lock_guard<std::mutex> lk(g_err_mu);
  g_last_err = std::string(what) + ": " + cudaGetErrorString(e);
}
}  // namespace spanon

extern "C" const char* spanon_strerror(int status) {
  switch (status) {
    case SPANON_OK: return "ok";
    case SPANON_ERR_BAD_SHAPE: return "bad shape: a size is non-positive, inconsistent or unsupported";
    case SPANON_ERR_NULL_POINTER: return "a required pointer is NULL";
    case SPANON_ERR_MISALIGNED: return "pointer alignment requirement violated";
    case SPANON_ERR_WORKSPACE: return "workspace too small (see *_workspace_bytes)";
    case SPANON_ERR_UNSUPPORTED: return "configuration not supported by the sm_100a kernels";
    case SPANON_ERR_CUDA: return "CUDA runtime error (see spanon_last_cuda_error)";
    case SPANON_ERR_ARCH: return "device is not sm_100";
    default: return "unknown status";
  }
}
extern "C" int spanon_abi_version(void) { return 1; }
extern "C" uint64_t spanon_launch_count(void) { return spanon::g_launch_count.load(); }
extern "C" const char* spanon_last_cuda_error(void) {
  static thread_local std::string copy;
  std::lock_guard<std::mutex> lk(spanon::g_err_mu);
  copy = spanon::g_last_err;
  return copy.c_str();
}
