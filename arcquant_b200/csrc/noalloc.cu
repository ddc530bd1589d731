// noalloc.cu -- arcq_linear_noalloc: quantise + GEMM of one linear with the intermediate NVFP4 activation in a file-scope
// device array, for callers that cannot allocate device memory (CUDA-graph capture, tight serving loops, harnesses that
// forbid cudaMalloc).  No counterpart in the reference (include/arcq.h).
//
// The two launches are exactly those of arcq_quantize_x followed by arcq_gemm_nvfp4 (same dispatch, same kernels, same
// bytes); only the home of the codes + scale factors between them differs.  Nothing here calls cudaMalloc, cudaFree or an
// allocating library routine.
#include <mutex>

#include "common.cuh"

namespace arcq {
namespace {

// The scratch: packed e2m1 codes [M, (KQ+KE)/2] at offset 0, E4M3 scale factors (arcq_sf_bytes) at the next multiple of
// 256.  Every GPU that loads the library holds these bytes until the process exits.  In HOST code this name is only a
// stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_noalloc_scratch[ARCQ_NOALLOC_SCRATCH_BYTES];

std::mutex g_scratch_mu;
unsigned char *g_scratch_ptr[kMaxDevices];  // device address of g_noalloc_scratch on device d (nullptr = not looked up yet)

int scratch_address(unsigned char **out) {
  const int dev = current_device();
  if (dev < 0) { set_error("arcq_linear_noalloc: no current CUDA device"); return ARCQ_ECUDA; }
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (!g_scratch_ptr[dev]) {
    void *ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) {
      set_error("arcq_linear_noalloc: cudaGetSymbolAddress(scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
      return ARCQ_ECUDA;
    }
    g_scratch_ptr[dev] = (unsigned char *)ptr;
  }
  *out = g_scratch_ptr[dev];
  return ARCQ_OK;
}

}  // namespace
}  // namespace arcq

using namespace arcq;

extern "C" int64_t arcq_linear_noalloc_bytes(int64_t M, int KQ, int KE) {
  if (M < 0 || KQ <= 0 || KE < 0) return -1;
  const int64_t K = (int64_t)KQ + KE;
  const int64_t codes = (M * (K / 2) + 255) / 256 * 256;
  return codes + arcq_sf_bytes(M, K);
}

extern "C" int arcq_linear_noalloc(const void *x, int64_t M, int KQ, const int16_t *idx, int KE, int flavour, const void *qw,
                                   const void *sfw, int64_t N, float alpha, const float *alpha_dev, void *y, void *stream) {
  if (M < 0 || N < 0 || KQ <= 0 || KE < 0) { set_error("arcq_linear_noalloc: bad shape (M=%lld N=%lld KQ=%d KE=%d)", (long long)M, (long long)N, KQ, KE); return ARCQ_EINVAL; }
  const int64_t need = arcq_linear_noalloc_bytes(M, KQ, KE);
  if (need < 0 || need > (int64_t)ARCQ_NOALLOC_SCRATCH_BYTES) {
    set_error("arcq_linear_noalloc: M=%lld x (KQ+KE)=%d needs %lld bytes of scratch, the built-in array holds %lld "
              "(ARCQ_NOALLOC_SCRATCH_BYTES); use arcq_quantize_x + arcq_gemm_nvfp4 with caller-owned buffers",
              (long long)M, KQ + KE, (long long)need, (long long)ARCQ_NOALLOC_SCRATCH_BYTES);
    return ARCQ_EINVAL;
  }
  if (M == 0 || N == 0) return ARCQ_OK;
  unsigned char *scratch = nullptr;
  if (int rc = scratch_address(&scratch)) return rc;
  const int64_t K = (int64_t)KQ + KE;
  unsigned char *qx = scratch;
  unsigned char *sfx = scratch + (M * (K / 2) + 255) / 256 * 256;
  // nothing in the scratch has to start at zero: the quantiser writes every code and every scale of rows < M, and the
  // GEMM never stores a row >= M (stale scale bytes of the padding rows of the last 128-row block only feed accumulator
  // rows that are dropped)
  if (int rc = arcq_quantize_x(x, M, KQ, idx, KE, flavour, qx, sfx, stream)) return rc;
  return arcq_gemm_nvfp4(qx, qw, sfx, sfw, M, N, K, alpha, alpha_dev, y, stream);
}
