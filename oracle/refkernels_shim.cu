// refkernels_shim.cu -- TEST INFRASTRUCTURE ONLY (see oracle/arcq_oracle.c header).
//
// C-ABI doorway to the reference's own quantise kernels, compiled UNCHANGED for
// sm_100a from /root/reference/kernels/src/{reorder,down,rmsnorm}.cu by
// oracle/build_ref.sh into oracle/_ref/libarcq_refkernels.so.  No reference
// source is copied: this file only re-declares the launcher templates
// (kernels/include/reorder.cuh:126-188) so it can call the explicit
// instantiations those .cu files already contain (reorder.cu:743-843,
// down.cu:405-424, rmsnorm.cu:283-314), with the same KQ dispatch table as
// kernels/src/bindings.cpp:141-160,190-207,241-251.
#include <cstdint>
#include <cuda_runtime.h>
#include "cutlass/numeric_types.h"

typedef cutlass::float_ue4m3_t sf_t;
typedef cutlass::bfloat16_t bf16_t;

template <int g, int h> void run_reorder_x_bf16_nvfp4(bf16_t *, int, int16_t *, uint8_t *, sf_t *, int, int);
template <int g, int h> void run_reorder_w_bf16_nvfp4(bf16_t *, int, int16_t *, uint8_t *, sf_t *, int, int);
template <int g, int h> void run_reorder32_x_bf16_nvfp4(bf16_t *, int, int16_t *, uint8_t *, sf_t *, int, int);
template <int g, int h> void run_reorder32_w_bf16_nvfp4(bf16_t *, int, int16_t *, uint8_t *, sf_t *, int, int);
template <int g, int h> void run_rmsnorm_x_bf16_nvfp4(bf16_t *, bf16_t *, float, int, int16_t *, uint8_t *, sf_t *, int, int);
template <int g, int h> void run_down32_x_bf16_nvfp4(bf16_t *, int, uint8_t *, sf_t *, int, int);
template <int g, int h> void run_down32_w_bf16_nvfp4(bf16_t *, int, uint8_t *, sf_t *, int, int);

#define X16(V) case V: run_reorder_x_bf16_nvfp4<16, V>(x, M, idx, q, sf, KQ, KE); break;
#define X32(V) case V: run_reorder32_x_bf16_nvfp4<32, V>(x, M, idx, q, sf, KQ, KE); break;
#define W16(V) case V: run_reorder_w_bf16_nvfp4<16, V>(x, M, idx, q, sf, KQ, KE); break;
#define W32(V) case V: run_reorder32_w_bf16_nvfp4<32, V>(x, M, idx, q, sf, KQ, KE); break;
#define D32X(V) case V: run_down32_x_bf16_nvfp4<32, V>(x, M, q, sf, KQ, KE); break;
#define D32W(V) case V: run_down32_w_bf16_nvfp4<32, V>(x, M, q, sf, KQ, KE); break;
#define RMS(V) case V: run_rmsnorm_x_bf16_nvfp4<16, V>(x, w, eps, M, idx, q, sf, KQ, KE); break;

static int finish() {
  cudaError_t e = cudaDeviceSynchronize();
  return e == cudaSuccess ? 0 : (int)e;
}

// All pointers are device pointers.  For KQ in {27648, 28672} the caller passes
// x ALREADY index_select-ed (bindings.cpp:29-38 does that with torch).
extern "C" int arcq_ref_quantize_x(void *xv, int M, int KQ, void *idxv, int KE, void *qv, void *sfv) {
  bf16_t *x = (bf16_t *)xv; int16_t *idx = (int16_t *)idxv; uint8_t *q = (uint8_t *)qv; sf_t *sf = (sf_t *)sfv;
  switch (KQ) {
    X16(2048) X16(3072) X16(4096) X16(5120) X16(8192) X16(11008) X16(13824) X16(14336)
    X32(3584) X32(18944) D32X(27648) D32X(28672)
    default: return -1;
  }
  return finish();
}

extern "C" int arcq_ref_quantize_w(void *xv, int M, int KQ, void *idxv, int KE, void *qv, void *sfv) {
  bf16_t *x = (bf16_t *)xv; int16_t *idx = (int16_t *)idxv; uint8_t *q = (uint8_t *)qv; sf_t *sf = (sf_t *)sfv;
  switch (KQ) {
    W16(2048) W16(3072) W16(4096) W16(5120) W16(8192) W16(11008) W16(13824) W16(14336)
    W32(3584) W32(18944) D32W(27648) D32W(28672)
    default: return -1;
  }
  return finish();
}

extern "C" int arcq_ref_rmsnorm_quantize_x(void *xv, void *wv, float eps, int M, int KQ, void *idxv, int KE,
                                           void *qv, void *sfv) {
  bf16_t *x = (bf16_t *)xv; bf16_t *w = (bf16_t *)wv; int16_t *idx = (int16_t *)idxv;
  uint8_t *q = (uint8_t *)qv; sf_t *sf = (sf_t *)sfv;
  switch (KQ) {
    RMS(2048) RMS(3072) RMS(3584) RMS(4096) RMS(5120) RMS(8192)
    default: return -1;
  }
  return finish();
}
