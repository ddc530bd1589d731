// common.cuh -- shared helpers for libarcq_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "arcq.h"

#if defined(__CUDA_ARCH__) && !(__CUDA_ARCH__ >= 1000)
#error "arcquant_b200 is written for sm_100a only"
#endif

namespace arcq {

// ---- host-side status plumbing --------------------------------------------------------------
void set_error(const char *fmt, ...);
void count_launch(int n = 1);
int sm_count();  // SMs of the current device (cached), <= 0 on error
int current_device();  // cudaGetDevice clamped to [0, kMaxDevices), -1 on error

constexpr int kMaxDevices = 64;

// Tuning overrides set through arcq_set_option() (include/arcq.h).  They only select among kernels / tile shapes that
// all compute the same result (the parity tests force each one through this); nothing here can skip work.  Plain
// process-wide values read at launch time: no environment lookups on the launch path.
struct Options {
  int quant_tile[5];   // {R, threads, S, RT, CTAs/SM} of the TMA-staged quantise kernel; R == 0: shape-driven choice
  int quant_family;    // 3 = TMA-staged kernel (default), 1 = single-role fallback kernel of quant.cu
  int gemm_pair;       // -1 = by shape, 0 = never, 1 = always use the cta_group::2 kernel
  int gemm_bn;         // 0 = by shape, 128 / 256 = tile width of the single-CTA kernel
  int gemm_splitk;     // -1 = by shape, 0 = never, >= 2 = forced number of K splits of the small-M kernel
  int gemm_raster;     // cta_group::2 kernel tile order: -1 = by shape, 0 = row tiles fastest, g > 0 = groups of g row tiles x all column tiles
};
Options &options();

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per (function, DEVICE): cache what was set per device so that a
// second GPU driven from the same process gets its own opt-in (and host threads racing here only repeat an idempotent call)
struct SmemAttrCache { int set[kMaxDevices]; };
template <typename Kern>
inline int ensure_dynamic_smem(Kern kern, int bytes, SmemAttrCache &cache) {
  const int dev = current_device();
  if (dev < 0) return ARCQ_ECUDA;
  if (bytes > cache.set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(%d B) failed: %s", bytes, cudaGetErrorString(e)); return ARCQ_ECUDA; }
    cache.set[dev] = bytes;
  }
  return ARCQ_OK;
}

#define ARCQ_CUDA_TRY(expr)                                                          \
  do {                                                                               \
    cudaError_t e__ = (expr);                                                        \
    if (e__ != cudaSuccess) {                                                        \
      arcq::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return ARCQ_ECUDA;                                                             \
    }                                                                                \
  } while (0)

// ---- numeric conversions (device) -----------------------------------------------------------
// The arithmetic contract is SURVEY.md Appendix A (reference kernels/src/reorder.cu:127-198).

__device__ __forceinline__ float bf16lo_to_f32(uint32_t packed) { return __uint_as_float(packed << 16); }
__device__ __forceinline__ float bf16hi_to_f32(uint32_t packed) { return __uint_as_float(packed & 0xffff0000u); }

// E4M3 RNE saturating (== cutlass NumericConverter<float_ue4m3_t,float> on device,
// third-party/cutlass/include/cutlass/float8.h:1073-1083); input is in [2^-9, 448].
__device__ __forceinline__ uint32_t f32_to_e4m3(float s) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(0.0f), "f"(s));
  return (uint32_t)(r & 0xff);
}

__device__ __forceinline__ float e4m3_to_f32(uint32_t b) {
  uint32_t h2;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)b));
  return __half2float(__ushort_as_half((unsigned short)(h2 & 0xffff)));
}

// 8 floats -> 8 e2m1 nibbles packed little-end first (element 0 in the low nibble of byte 0),
// RNE, saturating at +-6, sign kept on zero.  PTX puts the FIRST source in the HIGH nibble.
__device__ __forceinline__ uint32_t f32x8_to_e2m1(float a0, float a1, float a2, float a3, float a4,
                                                  float a5, float a6, float a7) {
  uint32_t r;
  asm("{\n\t"
      ".reg .b8 b0, b1, b2, b3;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b0, %2, %1;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b1, %4, %3;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b2, %6, %5;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b3, %8, %7;\n\t"
      "mov.b32 %0, {b0, b1, b2, b3};\n\t"
      "}"
      : "=r"(r)
      : "f"(a0), "f"(a1), "f"(a2), "f"(a3), "f"(a4), "f"(a5), "f"(a6), "f"(a7));
  return r;
}

// value of one e2m1 code (4 bits, sign | e1 e0 | m): exact.
__device__ __forceinline__ float e2m1_to_f32(uint32_t c) {
  // magnitude table {0, .5, 1, 1.5, 2, 3, 4, 6} built arithmetically: no local-memory LUT
  uint32_t e = (c >> 1) & 3u, m = c & 1u;
  // e == 0: m * 0.5 ; else (1 + m/2) * 2^(e-1)
  uint32_t bits = e ? (((126u + e) << 23) | (m << 22)) : (m ? 0x3f000000u : 0u);
  bits |= (c & 8u) << 28;
  return __uint_as_float(bits);
}

// per-column bias in the GEMM epilogues: bf16(float(y) + float(bias)) on two packed bf16 -- torch's `y + self.bias` on bf16
// tensors (model/qLinearLayer.py:74-75), applied to the ALREADY ROUNDED output so the fused result has the same bits
__device__ __forceinline__ uint32_t bias_add_bf16x2(uint32_t y, uint32_t b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(__fadd_rn(bf16lo_to_f32(y), bf16lo_to_f32(b)), __fadd_rn(bf16hi_to_f32(y), bf16hi_to_f32(b)));
  return *reinterpret_cast<uint32_t *>(&h);
}
// the eight bias values of output columns [gcol, gcol + 8) (zeros past ncols); gcol % 8 == 0 and a 16-byte aligned vector
__device__ __forceinline__ uint4 load_bias8(const __nv_bfloat16 *bias, int gcol, int ncols) {
  if (gcol + 8 <= ncols) return __ldg(reinterpret_cast<const uint4 *>(bias + gcol));
  uint32_t w[4] = {0u, 0u, 0u, 0u};
  for (int e = 0; e < 8; ++e)
    if (gcol + e < ncols) w[e >> 1] |= (uint32_t)__bfloat16_as_ushort(bias[gcol + e]) << (16 * (e & 1));
  return make_uint4(w[0], w[1], w[2], w[3]);
}

__device__ __forceinline__ uint32_t f32_to_bf16_bits(float f) {
  return (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(f));
}

}  // namespace arcq
