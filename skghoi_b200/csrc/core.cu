// Error state, device queries and library identity for the skghoi_b200 C-ABI.
#include "common.cuh"
#include "skghoi_b200.h"
#include <stdarg.h>
#include <string.h>
#include <atomic>

namespace skg {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};   // kernels enqueued through this library (skg_launch_count)

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
    return SKG_ERR_CUDA;
  }
  return SKG_OK;
}

void add_launches(long long n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
long long launch_count_now() { return g_launches.load(std::memory_order_relaxed); }

int num_sms() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return 148;
    cached = p.multiProcessorCount;
    cached_dev = dev;
  }
  return cached;
}

}  // namespace skg

extern "C" const char* skg_last_error(void) { return skg::g_err; }
extern "C" int skg_abi_version(void) { return SKG_ABI_VERSION; }
extern "C" int skg_num_sms(void) { return skg::num_sms(); }
extern "C" long long skg_launch_count(void) { return skg::g_launches.load(std::memory_order_relaxed); }
