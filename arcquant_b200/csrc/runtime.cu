// runtime.cu -- status plumbing shared by the entry points of libarcq_b200.so.
#include "common.cuh"
#include <atomic>
#include <string.h>

namespace arcq {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_error("cudaGetDevice failed: %s", cudaGetErrorString(e)); return -1; }
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int n = 0;
    e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) { set_error("cudaDeviceGetAttribute failed: %s", cudaGetErrorString(e)); return -1; }
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace arcq

extern "C" int64_t arcq_launch_count(void) { return arcq::g_launches.load(std::memory_order_relaxed); }
extern "C" const char *arcq_last_error(void) { return arcq::g_err; }
extern "C" const char *arcq_version(void) { return "arcq-b200 0.1 (sm_100a)"; }
