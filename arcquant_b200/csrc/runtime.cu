// runtime.cu -- status plumbing shared by the entry points of libarcq_b200.so.
#include "common.cuh"
#include <atomic>
#include <stdlib.h>
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

int current_device() {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_error("cudaGetDevice failed: %s", cudaGetErrorString(e)); return -1; }
  if (dev < 0 || dev >= kMaxDevices) { set_error("device ordinal %d out of range", dev); return -1; }
  return dev;
}

Options &options() {
  static Options o = {{0, 0, 0, 0, 0}, 3, -1, 0, -1, -1};
  return o;
}

int sm_count() {
  static int cached[kMaxDevices] = {0};
  const int dev = current_device();
  if (dev < 0) return -1;
  cudaError_t e;
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
extern "C" const char *arcq_version(void) { return "arcq-b200 0.2 (sm_100a)"; }

// name = value overrides of the shape-driven kernel choices (tests force every tile organisation through this).
extern "C" int arcq_set_option(const char *name, const char *value) {
  using namespace arcq;
  if (!name) { set_error("arcq_set_option: null name"); return ARCQ_EINVAL; }
  Options &o = options();
  const char *v = value ? value : "";
  if (!strcmp(name, "quant_tile")) {  // "R,threads,S[,RT[,CTAs/SM]]" or "" = shape-driven
    int f[5] = {0, 0, 0, 0, 0};
    const int nf = sscanf(v, "%d,%d,%d,%d,%d", &f[0], &f[1], &f[2], &f[3], &f[4]);
    if (nf < 3) { for (int &x : o.quant_tile) x = 0; return ARCQ_OK; }
    if (nf < 4) f[3] = f[0];
    if (nf < 5) f[4] = f[1] == 256 ? 2 : 1;
    for (int i = 0; i < 5; ++i) o.quant_tile[i] = f[i];
    return ARCQ_OK;
  }
  const bool has = *v != 0;
  const int iv = has ? atoi(v) : 0;
  if (!strcmp(name, "quant_family")) { o.quant_family = has ? iv : 3; return ARCQ_OK; }
  if (!strcmp(name, "gemm_pair")) { o.gemm_pair = has ? iv : -1; return ARCQ_OK; }
  if (!strcmp(name, "gemm_bn")) { o.gemm_bn = has ? iv : 0; return ARCQ_OK; }
  if (!strcmp(name, "gemm_splitk")) { o.gemm_splitk = has ? iv : -1; return ARCQ_OK; }
  if (!strcmp(name, "gemm_raster")) { o.gemm_raster = has ? (iv > 0 ? iv : 0) : -1; return ARCQ_OK; }
  set_error("arcq_set_option: unknown option '%s'", name);
  return ARCQ_EINVAL;
}
