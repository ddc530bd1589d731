// quant_math.cuh -- the arithmetic of one 16-group shared by the quantise kernel families.
#pragma once
#include "quant.cuh"

namespace arcq {

// amax / 6 correctly rounded without the division subroutine (Markstein: q' = RN(q0 + RN(a - 6 q0) * r))
static __device__ __forceinline__ float div6(float a) {
  const float r6 = 0.16666667163372039794921875f;  // RN(1/6) = 0x3e2aaaab
  const float q0 = __fmul_rn(a, r6);
  const float rem = __fmaf_rn(-6.0f, q0, a);
  return __fmaf_rn(rem, r6, q0);
}

// eight e2m1 codes (one word, element 0 in the low nibble) -> four f16x2 (element 2b in the low half): exact
static __device__ __forceinline__ void e2m1x8_to_f16x2(uint32_t w, uint32_t (&d)[4]) {
  asm("{\n\t"
      ".reg .b8 b0, b1, b2, b3;\n\t"
      "mov.b32 {b0, b1, b2, b3}, %4;\n\t"
      "cvt.rn.f16x2.e2m1x2 %0, b0;\n\t"
      "cvt.rn.f16x2.e2m1x2 %1, b1;\n\t"
      "cvt.rn.f16x2.e2m1x2 %2, b2;\n\t"
      "cvt.rn.f16x2.e2m1x2 %3, b3;\n\t"
      "}"
      : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
      : "r"(w));
}

// 4 floats -> 4 e2m1 nibbles (element 0 in the low nibble), RNE, saturating; same conversion as f32x8_to_e2m1
static __device__ __forceinline__ uint32_t f32x4_to_e2m1(float a0, float a1, float a2, float a3) {
  uint16_t r;
  asm("{\n\t"
      ".reg .b8 b0, b1;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b0, %2, %1;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b1, %4, %3;\n\t"
      "mov.b16 %0, {b0, b1};\n\t"
      "}"
      : "=h"(r)
      : "f"(a0), "f"(a1), "f"(a2), "f"(a3));
  return (uint32_t)r;
}

// primary quantisation of 16 values (reorder.cu:127-165): returns the E4M3 byte; s = unrounded clamped scale
static __device__ __forceinline__ uint32_t quant16(const float (&x)[16], const float *rcp_lut, uint32_t &c_lo, uint32_t &c_hi,
                                            float &s) {
  float amax = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) amax = fmaxf(amax, fabsf(x[i]));
  s = fmaxf(kScaleEps, fminf(kFp8Max, div6(amax)));  // :138
  const uint32_t sfb = f32_to_e4m3(s);               // :143
  const float r = rcp_lut[sfb];                      // RN(1/sf), :146
  float t[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) t[i] = __fmul_rn(x[i], r);  // :153-156
  c_lo = f32x8_to_e2m1(t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7]);
  c_hi = f32x8_to_e2m1(t[8], t[9], t[10], t[11], t[12], t[13], t[14], t[15]);
  return sfb;
}

// (a0*b0, a1*b1) with one packed FP32 instruction (FMUL2, sm_100); each lane is an IEEE RN multiply like __fmul_rn
static __device__ __forceinline__ void fmul2(float &d0, float &d1, float a0, float a1, float b0, float b1) {
  uint64_t a, b, d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(b0), "f"(b1));
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}

// (a0*a0 + c0, a1*a1 + c1) with one packed FFMA2: two independent IEEE RN fused multiply-adds like __fmaf_rn
static __device__ __forceinline__ void fsqacc2(float &c0, float &c1, float a0, float a1) {
  uint64_t a, c;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(c0), "f"(c1));
  asm("fma.rn.f32x2 %0, %1, %1, %0;" : "+l"(c) : "l"(a));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(c0), "=f"(c1) : "l"(c));
}

// (a0*b0 + c0, a1*b1 + c1) with one packed FFMA2
static __device__ __forceinline__ void ffma2(float &d0, float &d1, float a0, float a1, float b0, float b1, float c0, float c1) {
  uint64_t a, b, c, d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(b0), "f"(b1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(c0), "f"(c1));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}

// Per-tensor pre-scale of the reference callers, fused (model/qLlamaLayer.py:73-77: `x / scale` on a bf16 tensor with a
// 0-dim fp32 scale = fp32 IEEE division, then bf16 RNE).  v = bf16x2 of two elements; returns bf16x2 of the two quotients.
// Fast path: q0 = x * RN(1/scale); r = fma(-scale, q0, x) (exact); q = fma(r, RN(1/scale), q0) is the correctly rounded
// fp32 quotient (Markstein) whenever the remainder does not underflow (|x| >= 2^-100) and the significand of scale is not
// all ones; PrescaleCtx::slow covers those scales, and quotients of |x| < 2^-100 (only possible error: a last-place
// difference of a value below 2^-48 of the tensor maximum) quantise to code +-0 with the sign forced below -- unobservable.
struct PrescaleCtx { float scale, rcp; bool slow; };
// bf16_divisor: torch's CUDA TensorIterator casts a 0-dim CUDA fp32 divisor to the common dtype (bf16) before the fp32
// division (BinaryFunctor<scalar_t, scalar_t, ...>); a 1-element 1-D divisor, a Python float or the CPU path divide by
// the fp32 value.  The caller says which expression it replaces (ARCQ_PRESCALE_* in include/arcq.h).
static __device__ __forceinline__ PrescaleCtx make_prescale(float amax, bool bf16_divisor, float &scale_fp32) {
  PrescaleCtx c;
  // torch.max(x.abs()).float() / (448.0 * 6.0): dividing a CUDA tensor by a Python scalar is computed as a * (1 / b)
  // (BinaryDivTrueKernel.cu, "if the second operand is a CPU scalar, compute a * reciprocal(b)"), both in fp32
  scale_fp32 = __fmul_rn(amax, 1.0f / 2688.0f);
  c.scale = bf16_divisor ? __bfloat162float(__float2bfloat16_rn(scale_fp32)) : scale_fp32;
  c.rcp = __frcp_rn(c.scale);
  c.slow = !(amax >= 9.094947017729282e-13f && amax <= 1.152921504606847e18f) || (__float_as_uint(c.scale) & 0x7fffffu) == 0x7fffffu;
  return c;
}
static __device__ __forceinline__ uint32_t prescale_bf16x2(uint32_t v, const PrescaleCtx &c) {
  const float a = bf16lo_to_f32(v), b = bf16hi_to_f32(v);
  float qa, qb;
  if (c.slow) {
    qa = __fdiv_rn(a, c.scale);
    qb = __fdiv_rn(b, c.scale);
  } else {
    float ra, rb;
    fmul2(qa, qb, a, b, c.rcp, c.rcp);
    ffma2(ra, rb, qa, qb, -c.scale, -c.scale, a, b);
    ffma2(qa, qb, ra, rb, c.rcp, c.rcp, qa, qb);
  }
  uint32_t o;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(o) : "f"(qb), "f"(qa));
  return (o & 0x7fff7fffu) | (v & 0x80008000u);  // the quotient has the sign of x (scale > 0), zeros included
}

// rmsnorm of two rows at once (rmsnorm.cu:165-171): v[i] = bf16x2 {row a, row b} of element i -> bf16((x * w) * rstd)
static __device__ __forceinline__ void normalise16x2(uint32_t (&v)[16], const float (&wf)[16], float rstd_a, float rstd_b) {
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    float a = bf16lo_to_f32(v[i]), b = bf16hi_to_f32(v[i]);
    fmul2(a, b, a, b, wf[i], wf[i]);
    fmul2(a, b, a, b, rstd_a, rstd_b);
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(v[i]) : "f"(b), "f"(a));
  }
}

// primary quantisation of two rows at once: v[i] = bf16x2 {row a (low half), row b (high half)} of element i.
// Same arithmetic per row as quant16; the scaling multiply is one FMUL2 per element pair.
static __device__ __forceinline__ void quant16x2(const uint32_t (&v)[16], const float *rcp_lut, uint32_t (&sfb)[2],
                                                 uint32_t (&c)[2][2], float (&s)[2]) {
  float xa[16], xb[16];
  float ama = 0.0f, amb = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    xa[i] = bf16lo_to_f32(v[i]);
    xb[i] = bf16hi_to_f32(v[i]);
    ama = fmaxf(ama, fabsf(xa[i]));
    amb = fmaxf(amb, fabsf(xb[i]));
  }
  s[0] = fmaxf(kScaleEps, fminf(kFp8Max, div6(ama)));
  s[1] = fmaxf(kScaleEps, fminf(kFp8Max, div6(amb)));
  sfb[0] = f32_to_e4m3(s[0]);
  sfb[1] = f32_to_e4m3(s[1]);
  const float ra = rcp_lut[sfb[0]], rb = rcp_lut[sfb[1]];
#pragma unroll
  for (int i = 0; i < 16; ++i) fmul2(xa[i], xb[i], xa[i], xb[i], ra, rb);
  c[0][0] = f32x8_to_e2m1(xa[0], xa[1], xa[2], xa[3], xa[4], xa[5], xa[6], xa[7]);
  c[0][1] = f32x8_to_e2m1(xa[8], xa[9], xa[10], xa[11], xa[12], xa[13], xa[14], xa[15]);
  c[1][0] = f32x8_to_e2m1(xb[0], xb[1], xb[2], xb[3], xb[4], xb[5], xb[6], xb[7]);
  c[1][1] = f32x8_to_e2m1(xb[8], xb[9], xb[10], xb[11], xb[12], xb[13], xb[14], xb[15]);
}

// residual of 16 values: e = bf16(x - deq(code) * sub) (reorder.cu:157-160 / :474), then quantised again
static __device__ __forceinline__ uint32_t residual16(const float (&x)[16], uint32_t c_lo, uint32_t c_hi, float sub,
                                               const float *rcp_lut, uint32_t &r_lo, uint32_t &r_hi) {
  float e[16];
#ifndef ARCQ_E2M1_CVT_DECODE  // cvt.rn.f16x2.e2m1x2 decoded differently from the arithmetic decode on B200 (r1c_debug.log)
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const uint32_t code = ((i < 8 ? c_lo : c_hi) >> (4 * (i & 7))) & 0xfu;
    e[i] = bf16lo_to_f32(f32_to_bf16_bits(__fmaf_rn(-e2m1_to_f32(code), sub, x[i])));
  }
#else
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint32_t d[4];
    e2m1x8_to_f16x2(h ? c_hi : c_lo, d);
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&d[b]));  // exact
      e[8 * h + 2 * b] = bf16lo_to_f32(f32_to_bf16_bits(__fmaf_rn(-f.x, sub, x[8 * h + 2 * b])));
      e[8 * h + 2 * b + 1] = bf16lo_to_f32(f32_to_bf16_bits(__fmaf_rn(-f.y, sub, x[8 * h + 2 * b + 1])));
    }
  }
#endif
  float s2;
  return quant16(e, rcp_lut, r_lo, r_hi, s2);
}

// R rows (2R bytes) of one channel chunk
template <int R> static __device__ __forceinline__ void lds_chunk(const uint8_t *src, uint32_t (&v)[(R + 1) / 2]) {
  if constexpr (R == 8) { const uint4 q = *reinterpret_cast<const uint4 *>(src); v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w; }
  else if constexpr (R == 4) { const uint2 q = *reinterpret_cast<const uint2 *>(src); v[0] = q.x; v[1] = q.y; }
  else if constexpr (R == 2) v[0] = *reinterpret_cast<const uint32_t *>(src);
  else v[0] = *reinterpret_cast<const uint16_t *>(src);
}


}  // namespace arcq
