// Library-level entry points: status strings, ABI version, launch counter.
#include "common.cuh"
#include <mutex>
#include <string>
#include <string.h>
#include <stdlib.h>
#include <vector>

namespace spanon {
std::atomic<uint64_t> g_launch_count{0};
static std::mutex g_err_mu;
static std::string g_last_err;
void set_last_cuda_error(const char* what, cudaError_t e) {
  std::lock_guard<std::mutex> lk(g_err_mu);
  g_last_err = std::string(what) + ": " + cudaGetErrorString(e);
}

// ---- kernel-time profile: consecutive-event deltas on one stream, aggregated by kernel name ----
std::atomic<int> g_profile_on{0};
const bool g_debug_sync = getenv("SPANON_DEBUG_SYNC") && getenv("SPANON_DEBUG_SYNC")[0] == '1';
namespace {
struct Mark { const char* what; cudaEvent_t ev; };
std::mutex g_prof_mu;
std::vector<Mark> g_marks;
cudaStream_t g_prof_stream = nullptr;
}  // namespace
void profile_mark(const char* what) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (!g_profile_on.load()) return;
  cudaEvent_t ev;
  if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaEventRecord(ev, g_prof_stream);
  g_marks.push_back({what, ev});
}
}  // namespace spanon

extern "C" int spanon_profile_begin(void* stream) {
  using namespace spanon;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (auto& m : g_marks) cudaEventDestroy(m.ev);
  g_marks.clear();
  g_prof_stream = as_stream(stream);
  cudaEvent_t ev;
  SPANON_CUDA_TRY(cudaEventCreate(&ev));
  SPANON_CUDA_TRY(cudaEventRecord(ev, g_prof_stream));
  g_marks.push_back({"<begin>", ev});
  g_profile_on.store(1);
  return SPANON_OK;
}

extern "C" int spanon_profile_end(SpanonKernelTime* entries, int max_entries) {
  using namespace spanon;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_profile_on.store(0);
  if (g_marks.empty()) return 0;
  SPANON_CUDA_TRY(cudaEventSynchronize(g_marks.back().ev));
  int n = 0;
  for (size_t i = 1; i < g_marks.size(); ++i) {
    float ms = 0.f;
    SPANON_CUDA_TRY(cudaEventElapsedTime(&ms, g_marks[i - 1].ev, g_marks[i].ev));
    int j = 0;
    for (; j < n; ++j) if (strncmp(entries[j].name, g_marks[i].what, sizeof(entries[j].name) - 1) == 0) break;
    if (j == n) {
      if (!entries || n >= max_entries) continue;
      memset(&entries[n], 0, sizeof(entries[n]));
      strncpy(entries[n].name, g_marks[i].what, sizeof(entries[n].name) - 1);
      ++n;
    }
    entries[j].launches += 1;
    entries[j].ms_total += (double)ms;
  }
  for (auto& m : g_marks) cudaEventDestroy(m.ev);
  g_marks.clear();
  return n;
}

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
    case SPANON_ERR_STATIC_LIMIT: return "size exceeds the built-in static scratch of this *_static entry point (use the workspace form)";
    default: return "unknown status";
  }
}
extern "C" int spanon_abi_version(void) { return 3; }
extern "C" uint64_t spanon_launch_count(void) { return spanon::g_launch_count.load(); }
extern "C" const char* spanon_last_cuda_error(void) {
  static thread_local std::string copy;
  std::lock_guard<std::mutex> lk(spanon::g_err_mu);
  copy = spanon::g_last_err;
  return copy.c_str();
}
