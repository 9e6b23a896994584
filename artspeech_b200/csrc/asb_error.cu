// Error reporting, launch accounting and device queries shared by all entry points.
#include <atomic>
#include <cstdarg>
#include <cstring>

#include "asb_common.cuh"

namespace {
thread_local char g_err[512] = "no error";
std::atomic<long long> g_launches{0};
}  // namespace

namespace asb {

int set_error(int status, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return status;
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return ASB_OK;
  return set_error(ASB_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace asb

#ifndef ASB_SOURCE_HASH
#define ASB_SOURCE_HASH "unknown"
#endif
extern "C" int asb_abi_version(void) { return ASB_ABI_VERSION; }
extern "C" const char* asb_build_hash(void) { return ASB_SOURCE_HASH; }
extern "C" const char* asb_last_error_string(void) { return g_err; }
extern "C" int64_t asb_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
