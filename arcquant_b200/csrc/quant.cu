// quant.cu -- fused activation reorder + NVFP4 quantise + Augmented Residual Channels (sm_100a).
//
// Replaces the reference's reorder_x/w, reorder32_x/w, down32_x/w (+ torch index_select) and
// rmsnorm_x kernels (/root/reference/kernels/src/{reorder,down,rmsnorm}.cu) with ONE generic-KQ
// kernel family.  Arithmetic contract: SURVEY.md Appendix A.  Design:
//
//  * persistent CTAs; a work item ("row tile") is R rows of one 128-row scale-factor block, chosen
//    so that the R rows' scale bytes for 4 consecutive 16-groups are CONTIGUOUS in the tcgen05
//    128x4 swizzled layout (R*4 bytes, a full 32-B sector for R = 8);
//  * rows are read with 128-bit coalesced loads, transposed in registers (PRMT) and staged in
//    shared memory channel-major ("[channel][R rows]", XOR-swizzled) so the permuted gather is one
//    2R-byte shared load per channel for all R rows instead of R two-byte loads;
//  * the reorder index is read once per (group, tile) -- R-fold fewer index reads than one row per CTA;
//  * e2m1 / e4m3 conversions use the sm_100a cvt instructions; scale bytes are assembled in shared
//    memory and written as whole 4R-byte runs; codes leave as 8/16-byte coalesced stores.
#include "quant.cuh"
#include <stdlib.h>

namespace arcq {

constexpr int kThreads = 256;
// Primary quantisation of 16 values (reorder.cu:127-165): returns the E4M3 byte, the two code words,
// the decoded scale sf and the unrounded clamped scale s.
__device__ __forceinline__ uint32_t quant16(const float (&x)[16], uint32_t &c_lo, uint32_t &c_hi, float &sf,
                                            float &s) {
  float amax = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) amax = fmaxf(amax, fabsf(x[i]));
  s = fmaxf(kScaleEps, fminf(kFp8Max, __fdiv_rn(amax, kFp4Max)));  // :138
  uint32_t sfb = f32_to_e4m3(s);                                   // :143
  sf = e4m3_to_f32(sfb);
  float r = __frcp_rn(sf);  // == f32(1.0 / (double)sf) for every E4M3 value (:146)
  float t[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) t[i] = __fmul_rn(x[i], r);  // multiply by the reciprocal, :153-156
  c_lo = f32x8_to_e2m1(t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7]);
  c_hi = f32x8_to_e2m1(t[8], t[9], t[10], t[11], t[12], t[13], t[14], t[15]);
  return sfb;
}

template <int R, int FLAVOUR, bool IS_W, bool RMS>
__global__ void __launch_bounds__(kThreads, (R == 8 || RMS) ? 2 : 3) quant_kernel(const QuantParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int CB = 2 * R;            // bytes of one channel chunk
  constexpr int RP = (R + 1) / 2;      // 32-bit words per chunk (R = 1 uses half a word)
  const int KQ = p.KQ, KE = p.KE, K = KQ + KE;
  const int G = (KQ - KE) / 16;        // plain groups
  const int ngroups = KQ / 16;
  const int nunits = KQ / 8;           // 8-channel load units per row
  const int nchunks = K / 64;          // 4-slot scale chunks per row
  uint8_t *tile = smem;                                  // KQ * CB bytes
  uint8_t *sfs = tile + (size_t)KQ * CB;                 // nchunks * 4R bytes: [chunk][lr][slot%4]
  uint16_t *wsm = nullptr;
  float *part = nullptr, *rstd_s = nullptr;
  if constexpr (RMS) {
    wsm = reinterpret_cast<uint16_t *>(sfs + (size_t)nchunks * 4 * R);
    part = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(wsm) + (size_t)KQ * 2);
    rstd_s = part + R * 512;
    for (int u = threadIdx.x; u < nunits; u += kThreads)
      reinterpret_cast<uint4 *>(wsm)[u] = __ldg(reinterpret_cast<const uint4 *>(p.w) + u);
  }
  constexpr int tiles_per_block = 128 / R;

  for (int tile_id = blockIdx.x; tile_id < p.num_tiles; tile_id += gridDim.x) {
    const int blk = tile_id / tiles_per_block, t = tile_id % tiles_per_block;
    const int64_t row0 = (int64_t)blk * 128;

    // ---- phase 1: coalesced 128-bit loads -> register transpose -> channel-major shared tile ----
    for (int u = threadIdx.x; u < nunits; u += kThreads) {
      uint4 in[R];
#pragma unroll
      for (int lr = 0; lr < R; ++lr) {
        const int64_t row = row0 + TileMap<R>::row(t, lr);
        in[lr] = make_uint4(0, 0, 0, 0);
        if (row < p.M) in[lr] = __ldcs(reinterpret_cast<const uint4 *>(p.x + row * KQ) + u);
      }
      if constexpr (R == 1) {
        reinterpret_cast<uint4 *>(tile)[u] = in[0];
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint32_t out[RP];
#pragma unroll
          for (int pr = 0; pr < RP; ++pr) {
            const uint32_t a = (&in[2 * pr].x)[j >> 1], b = (&in[2 * pr + 1].x)[j >> 1];
            out[pr] = __byte_perm(a, b, (j & 1) ? 0x7632 : 0x5410);
          }
          uint8_t *dst = tile + (size_t)chan_pos<R>(8 * u + j) * CB;
          if constexpr (R == 8) *reinterpret_cast<uint4 *>(dst) = make_uint4(out[0], out[1], out[2], out[3]);
          else if constexpr (R == 4) *reinterpret_cast<uint2 *>(dst) = make_uint2(out[0], out[1]);
          else *reinterpret_cast<uint32_t *>(dst) = out[0];
        }
      }
    }
    __syncthreads();

    // ---- phase 1b (rmsnorm): sum of squares in the reference's exact order (rmsnorm.cu:113-157) ----
    if constexpr (RMS) {
      const int bdx = KQ / 16;
      for (int tt = threadIdx.x; tt < 512; tt += kThreads) {
        float acc[R];
#pragma unroll
        for (int lr = 0; lr < R; ++lr) acc[lr] = 0.0f;
        if (tt < bdx) {
#pragma unroll
          for (int half = 0; half < 2; ++half) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int c = half * 8 * bdx + 8 * tt + i;
              const uint8_t *src = tile + (size_t)chan_pos<R>(c) * CB;
              uint32_t v[RP];
              if constexpr (R == 8) { uint4 q4 = *reinterpret_cast<const uint4 *>(src); v[0] = q4.x; v[1] = q4.y; v[2] = q4.z; v[3] = q4.w; }
              else if constexpr (R == 4) { uint2 q2 = *reinterpret_cast<const uint2 *>(src); v[0] = q2.x; v[1] = q2.y; }
              else if constexpr (R == 2) v[0] = *reinterpret_cast<const uint32_t *>(src);
              else v[0] = *reinterpret_cast<const uint16_t *>(src);
#pragma unroll
              for (int lr = 0; lr < R; ++lr) {
                const float f = (lr & 1) ? bf16hi_to_f32(v[lr >> 1]) : bf16lo_to_f32(v[lr >> 1]);
                acc[lr] = __fmaf_rn(f, f, acc[lr]);
              }
            }
          }
        }
#pragma unroll
        for (int lr = 0; lr < R; ++lr) part[lr * 512 + tt] = acc[lr];
      }
      __syncthreads();
      const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
      if (warp < R) {  // one warp replays the reference's tree for one row (rmsnorm.cu:131-154)
        float *s = part + warp * 512;
        for (int k = 0; k < 8; ++k) s[lane + 32 * k] += s[lane + 32 * k + 256];
        __syncwarp();
        for (int k = 0; k < 4; ++k) s[lane + 32 * k] += s[lane + 32 * k + 128];
        __syncwarp();
        for (int k = 0; k < 2; ++k) s[lane + 32 * k] += s[lane + 32 * k + 64];
        __syncwarp();
        float v = s[lane] + s[lane + 32];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
        if (lane == 0) rstd_s[warp] = rsqrtf(__fadd_rn(__fdiv_rn(v, (float)KQ), p.eps));  // :157
      }
      __syncthreads();
    }

    // ---- phase 2: permuted gather + quantise + ARC, one 16-group per thread per step ----
    for (int g = threadIdx.x; g < ngroups; g += kThreads) {
      int16_t ci[16];
      {
        const uint4 i0 = __ldg(reinterpret_cast<const uint4 *>(p.idx + 16 * g));
        const uint4 i1 = __ldg(reinterpret_cast<const uint4 *>(p.idx + 16 * g) + 1);
        const uint32_t w[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) { ci[2 * i] = (int16_t)(w[i] & 0xffff); ci[2 * i + 1] = (int16_t)(w[i] >> 16); }
      }
      uint32_t v[16][RP];
      float wf[RMS ? 16 : 1];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int c = (int)ci[i];
        const uint8_t *src = tile + (size_t)chan_pos<R>(c) * CB;
        if constexpr (R == 8) { uint4 q4 = *reinterpret_cast<const uint4 *>(src); v[i][0] = q4.x; v[i][1] = q4.y; v[i][2] = q4.z; v[i][3] = q4.w; }
        else if constexpr (R == 4) { uint2 q2 = *reinterpret_cast<const uint2 *>(src); v[i][0] = q2.x; v[i][1] = q2.y; }
        else if constexpr (R == 2) v[i][0] = *reinterpret_cast<const uint32_t *>(src);
        else v[i][0] = *reinterpret_cast<const uint16_t *>(src);
        if constexpr (RMS) wf[i] = bf16lo_to_f32(wsm[c]);
      }
      const bool arc = g >= G;
      int slot_p = g, slot_r = g;
      if (arc) {
        const int j = g - G;
        if constexpr (FLAVOUR == 0) { slot_p = G + 2 * j; slot_r = slot_p + 1; }       // reorder.cu:139,175
        else { slot_p = G + 4 * (j >> 1) + (j & 1); slot_r = slot_p + 2; }              // reorder.cu:451-463,510-515
      }
#pragma unroll
      for (int lr = 0; lr < R; ++lr) {
        float x[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          x[i] = (lr & 1) ? bf16hi_to_f32(v[i][lr >> 1]) : bf16lo_to_f32(v[i][lr >> 1]);
          if constexpr (RMS)  // rmsnorm.cu:165-171: bf16((x * w) * rstd), left to right
            x[i] = bf16lo_to_f32(f32_to_bf16_bits(__fmul_rn(__fmul_rn(x[i], wf[i]), rstd_s[lr])));
        }
        uint32_t c_lo, c_hi, r_lo, r_hi;
        float sf, s;
        const uint32_t sfb = quant16(x, c_lo, c_hi, sf, s);
        uint32_t rsfb = sfb;
        r_lo = c_lo; r_hi = c_hi;  // weights: ARC slots duplicate the primary (reorder.cu:307-316)
        if (!IS_W && arc) {
          const float sub = FLAVOUR == 0 ? sf : s;  // 16-wide: rounded scale (:157); 32-wide: unrounded (:474)
          float e[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const uint32_t code = ((i < 8 ? c_lo : c_hi) >> (4 * (i & 7))) & 0xfu;
            e[i] = bf16lo_to_f32(f32_to_bf16_bits(__fmaf_rn(-e2m1_to_f32(code), sub, x[i])));
          }
          float sf2, s2;
          rsfb = quant16(e, r_lo, r_hi, sf2, s2);
        }
        const int64_t row = row0 + TileMap<R>::row(t, lr);
        sfs[(slot_p >> 2) * (4 * R) + lr * 4 + (slot_p & 3)] = (uint8_t)sfb;
        if (arc) sfs[(slot_r >> 2) * (4 * R) + lr * 4 + (slot_r & 3)] = (uint8_t)rsfb;
        if (row < p.M) {
          uint8_t *qrow = p.q + row * (int64_t)(K / 2);
          if (arc && FLAVOUR == 0) {
            __stcs(reinterpret_cast<uint4 *>(qrow + 8 * slot_p), make_uint4(c_lo, c_hi, r_lo, r_hi));
          } else {
            __stcs(reinterpret_cast<uint2 *>(qrow + 8 * slot_p), make_uint2(c_lo, c_hi));
            if (arc) __stcs(reinterpret_cast<uint2 *>(qrow + 8 * slot_r), make_uint2(r_lo, r_hi));
          }
        }
      }
    }
    __syncthreads();

    // ---- phase 3: scale bytes leave as contiguous 4R-byte runs of the swizzled layout ----
    {
      uint8_t *dst0 = p.sf + (int64_t)blk * nchunks * 512 + TileMap<R>::w0(t) * 4;
      for (int c = threadIdx.x; c < nchunks; c += kThreads) {
        uint8_t *dst = dst0 + (int64_t)c * 512;
        const uint8_t *src = sfs + c * (4 * R);
        if constexpr (R == 8) {
          reinterpret_cast<uint4 *>(dst)[0] = reinterpret_cast<const uint4 *>(src)[0];
          reinterpret_cast<uint4 *>(dst)[1] = reinterpret_cast<const uint4 *>(src)[1];
        } else if constexpr (R == 4) {
          *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(src);
        } else if constexpr (R == 2) {
          *reinterpret_cast<uint2 *>(dst) = *reinterpret_cast<const uint2 *>(src);
        } else {
          *reinterpret_cast<uint32_t *>(dst) = *reinterpret_cast<const uint32_t *>(src);
        }
      }
    }
    // the next iteration's phase-1 stores touch only `tile` (all gathers are behind the barrier above);
    // sfs is rewritten only after the next iteration's first barrier.
  }
}

// ---- host side ------------------------------------------------------------------------------

static size_t quant_smem_bytes(int R, int KQ, int KE, bool rms) {
  size_t b = (size_t)KQ * 2 * R + (size_t)((KQ + KE) / 64) * 4 * R;
  if (rms) b += (size_t)KQ * 2 + (size_t)R * 512 * 4 + 64;
  return b;
}

template <int R, int FLAVOUR, bool IS_W, bool RMS>
static int launch_quant(const QuantParams &p0, cudaStream_t stream) {
  QuantParams p = p0;
  auto kern = quant_kernel<R, FLAVOUR, IS_W, RMS>;
  const size_t smem = quant_smem_bytes(R, p.KQ, p.KE, RMS);
  static SmemAttrCache attr_cache = {};  // per instantiation, per device
  if (int rc = ensure_dynamic_smem(kern, (int)smem, attr_cache)) return rc;
  int occ = 0;  // fallback kernel for unusual shapes: the occupancy query (host-only, ~1 us) is not worth caching per device
  ARCQ_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, smem));
  if (occ < 1) { set_error("quantise kernel does not fit on an SM (smem %zu B)", smem); return ARCQ_ENOSUP; }
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  p.num_tiles = (int)(((int64_t)p.M + 127) / 128) * (128 / R);
  int grid = sms * occ;
  if (grid > p.num_tiles) grid = p.num_tiles;
  kern<<<grid, kThreads, smem, stream>>>(p);
  ARCQ_CUDA_TRY(cudaGetLastError());
  count_launch();
  return ARCQ_OK;
}

static int pick_rows(int KQ, int KE, bool rms) {
  for (int R = 8; R >= 1; R >>= 1)
    if (quant_smem_bytes(R, KQ, KE, rms) <= (rms ? 100u : 74u) * 1024u) return R;
  return 1;
}

static int check_common(const void *x, int64_t M, int KQ, const void *idx, int KE, int flavour, const void *q,
                        const void *sf) {
  if (M < 0 || M > 0x7fffff00) { set_error("row count %lld out of range", (long long)M); return ARCQ_EINVAL; }
  if (!idx || !sf || (M > 0 && (!x || !q))) { set_error("null pointer argument"); return ARCQ_EINVAL; }
  const int w = flavour == ARCQ_FLAVOUR_32 ? 32 : 16;
  if (KQ <= 0 || KQ > 32768 || KE < 0 || KE > KQ || KQ % w || KE % w || (KQ + KE) % 64) {
    set_error("unsupported shape KQ=%d KE=%d for the %d-wide flavour (need KQ%%%d==0, KE%%%d==0, (KQ+KE)%%64==0, KE<=KQ<=32768)",
              KQ, KE, w, w, w);
    return ARCQ_EINVAL;
  }
  if (((uintptr_t)x | (uintptr_t)idx | (uintptr_t)q | (uintptr_t)sf) & 15) {
    set_error("pointers must be 16-byte aligned");
    return ARCQ_EINVAL;
  }
  return ARCQ_OK;
}

// arcq_set_option("quant_family", "1") forces the single-role kernel of this file (its parity tests); default is the
// TMA-staged family, which hands the shapes it does not take (ARCQ_ENOSUP) to this one
static bool use_tma() { return options().quant_family == 3; }

template <bool IS_W>
static int dispatch_quant(const QuantParams &p, int flavour, cudaStream_t st) {
  if (use_tma()) {
    const int rc = launch_quant_tma(p, flavour, IS_W, false, st);
    if (rc != ARCQ_ENOSUP) return rc;
  }
  const int R = pick_rows(p.KQ, p.KE, false);
#define ARCQ_CASE(RR)                                                                      \
  case RR:                                                                                 \
    return flavour == ARCQ_FLAVOUR_32 ? launch_quant<RR, 1, IS_W, false>(p, st) : launch_quant<RR, 0, IS_W, false>(p, st);
  switch (R) { ARCQ_CASE(8) ARCQ_CASE(4) ARCQ_CASE(2) ARCQ_CASE(1) }
#undef ARCQ_CASE
  return ARCQ_EINVAL;
}

}  // namespace arcq

using namespace arcq;

extern "C" int arcq_flavour_for_kq(int KQ) {
  switch (KQ) {  // the reference's dispatch table, kernels/src/bindings.cpp:141-160
    case 2048: case 3072: case 4096: case 5120: case 8192: case 11008: case 13824: case 14336: return ARCQ_FLAVOUR_16;
    case 3584: case 18944: case 27648: case 28672: return ARCQ_FLAVOUR_32;
    default: return -1;
  }
}

extern "C" int64_t arcq_sf_bytes(int64_t rows, int64_t K) { return (rows / 128 + 1) * 128 * K / 16; }

extern "C" int64_t arcq_sf_offset(int64_t row, int64_t g, int64_t K) {
  return (row / 128) * (K / 64) * 512 + (g / 4) * 512 + (row % 32) * 16 + ((row / 32) % 4) * 4 + (g % 4);
}

static int resolve_flavour(int flavour, int KQ) {
  if (flavour == ARCQ_FLAVOUR_16 || flavour == ARCQ_FLAVOUR_32) return flavour;
  const int f = arcq_flavour_for_kq(KQ);
  return f < 0 ? ARCQ_FLAVOUR_16 : f;
}

extern "C" int arcq_quantize_x(const void *x, int64_t M, int KQ, const int16_t *idx, int KE, int flavour, void *qx,
                               void *sfx, void *stream) {
  flavour = resolve_flavour(flavour, KQ);
  if (int rc = check_common(x, M, KQ, idx, KE, flavour, qx, sfx)) return rc;
  if (M == 0) return ARCQ_OK;
  QuantParams p{(const uint16_t *)x, idx, nullptr, (uint8_t *)qx, (uint8_t *)sfx, nullptr, nullptr, nullptr, 0, 0.0f, (int)M, KQ, KE, 0, 0};
  return dispatch_quant<false>(p, flavour, (cudaStream_t)stream);
}

// ---- fused per-tensor pre-scale (reference callers model/qLlamaLayer.py:73-77) ----------------------

namespace arcq {
// max |x| over n bf16 elements as the fp32 value of the winning element; non-negative floats order like their bit patterns
__global__ void __launch_bounds__(256) absmax_kernel(const uint4 *__restrict__ x, int64_t n8, const uint16_t *tail, int ntail,
                                                      unsigned int *out_bits) {
  uint32_t m = 0;  // per-halfword running maximum of |bits|
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n8; i += 4 * stride) {  // four independent 16-byte loads in flight per thread
    const uint4 a = __ldcs(x + i), b = __ldcs(x + i + stride), c = __ldcs(x + i + 2 * stride), d = __ldcs(x + i + 3 * stride);
    m = __vmaxu2(m, __vmaxu2(__vmaxu2(a.x & 0x7fff7fffu, a.y & 0x7fff7fffu), __vmaxu2(a.z & 0x7fff7fffu, a.w & 0x7fff7fffu)));
    m = __vmaxu2(m, __vmaxu2(__vmaxu2(b.x & 0x7fff7fffu, b.y & 0x7fff7fffu), __vmaxu2(b.z & 0x7fff7fffu, b.w & 0x7fff7fffu)));
    m = __vmaxu2(m, __vmaxu2(__vmaxu2(c.x & 0x7fff7fffu, c.y & 0x7fff7fffu), __vmaxu2(c.z & 0x7fff7fffu, c.w & 0x7fff7fffu)));
    m = __vmaxu2(m, __vmaxu2(__vmaxu2(d.x & 0x7fff7fffu, d.y & 0x7fff7fffu), __vmaxu2(d.z & 0x7fff7fffu, d.w & 0x7fff7fffu)));
  }
  for (; i < n8; i += stride) {
    const uint4 v = __ldg(x + i);
    m = __vmaxu2(m, __vmaxu2(__vmaxu2(v.x & 0x7fff7fffu, v.y & 0x7fff7fffu), __vmaxu2(v.z & 0x7fff7fffu, v.w & 0x7fff7fffu)));
  }
  if (blockIdx.x == 0 && threadIdx.x < ntail) m = __vmaxu2(m, (uint32_t)(tail[threadIdx.x] & 0x7fffu));
  uint32_t h = max(m & 0xffffu, m >> 16);
  h = __reduce_max_sync(0xffffffffu, h);
  __shared__ uint32_t wmax[8];
  if ((threadIdx.x & 31) == 0) wmax[threadIdx.x >> 5] = h;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) h = max(h, wmax[w]);
    if (h) atomicMax(out_bits, h << 16);
  }
}
}  // namespace arcq

extern "C" int arcq_absmax(const void *x, int64_t n, float *amax_out, void *stream) {
  if (!amax_out || (n > 0 && !x) || n < 0) { set_error("arcq_absmax: bad argument"); return ARCQ_EINVAL; }
  if ((uintptr_t)x & 15) { set_error("arcq_absmax: x must be 16-byte aligned"); return ARCQ_EINVAL; }
  cudaStream_t st = (cudaStream_t)stream;
  ARCQ_CUDA_TRY(cudaMemsetAsync(amax_out, 0, sizeof(float), st));
  if (n == 0) return ARCQ_OK;
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  const int64_t n8 = n / 8;
  int64_t blocks = (n8 + 1023) / 1024;  // four loads per thread per round
  if (blocks > (int64_t)sms * 8) blocks = (int64_t)sms * 8;
  if (blocks < 1) blocks = 1;
  absmax_kernel<<<(int)blocks, 256, 0, st>>>((const uint4 *)x, n8, (const uint16_t *)x + n8 * 8, (int)(n - n8 * 8),
                                             (unsigned int *)amax_out);
  ARCQ_CUDA_TRY(cudaGetLastError());
  count_launch();
  return ARCQ_OK;
}

// reorder_quantize_x with the optional extras: a placement plan (plan.cu) and / or the fused per-tensor pre-scale
extern "C" int arcq_quantize_x_ex(const void *x, int64_t M, int KQ, const int16_t *idx, int KE, int flavour,
                                  const uint16_t *plan_dev, const float *amax_dev, int prescale_mode, void *qx, void *sfx,
                                  float *scale_out, void *stream) {
  flavour = resolve_flavour(flavour, KQ);
  if (int rc = check_common(x, M, KQ, idx, KE, flavour, qx, sfx)) return rc;
  if (amax_dev && prescale_mode != ARCQ_PRESCALE_FP32 && prescale_mode != ARCQ_PRESCALE_BF16_DIVISOR) {
    set_error("arcq_quantize_x_ex: unknown prescale_mode %d", prescale_mode);
    return ARCQ_EINVAL;
  }
  if (plan_dev && ((uintptr_t)plan_dev & 15)) { set_error("arcq_quantize_x_ex: the plan must be 16-byte aligned (it is fetched with bulk copies)"); return ARCQ_EINVAL; }
  if (M == 0) return ARCQ_OK;
  QuantParams p{(const uint16_t *)x, idx, nullptr, (uint8_t *)qx, (uint8_t *)sfx, plan_dev, amax_dev, scale_out, prescale_mode, 0.0f,
                (int)M, KQ, KE, 0, 0};
  if (!amax_dev) return dispatch_quant<false>(p, flavour, (cudaStream_t)stream);   // a plan is an optimisation: fallbacks ignore it
  const int rc = launch_quant_tma(p, flavour, false, false, (cudaStream_t)stream);
  if (rc == ARCQ_ENOSUP) set_error("arcq_quantize_x_ex: KQ=%d KE=%d is outside the TMA-staged kernel's shapes; pre-scale separately", KQ, KE);
  return rc;
}

extern "C" int arcq_quantize_x_scaled(const void *x, int64_t M, int KQ, const int16_t *idx, int KE, int flavour,
                                      const float *amax_dev, int prescale_mode, void *qx, void *sfx, float *scale_out,
                                      void *stream) {
  if (!amax_dev) { set_error("arcq_quantize_x_scaled: amax_dev is null"); return ARCQ_EINVAL; }
  return arcq_quantize_x_ex(x, M, KQ, idx, KE, flavour, nullptr, amax_dev, prescale_mode, qx, sfx, scale_out, stream);
}

extern "C" int arcq_quantize_w(const void *w, int64_t N, int KQ, const int16_t *idx, int KE, int flavour, void *qw,
                               void *sfw, void *stream) {
  flavour = resolve_flavour(flavour, KQ);
  if (int rc = check_common(w, N, KQ, idx, KE, flavour, qw, sfw)) return rc;
  if (N == 0) return ARCQ_OK;
  QuantParams p{(const uint16_t *)w, idx, nullptr, (uint8_t *)qw, (uint8_t *)sfw, nullptr, nullptr, nullptr, 0, 0.0f, (int)N, KQ, KE, 0, 0};
  return dispatch_quant<true>(p, flavour, (cudaStream_t)stream);
}

extern "C" int arcq_rmsnorm_quantize_x(const void *x, const void *w, float eps, int64_t M, int H, const int16_t *idx,
                                       int KE, void *qx, void *sfx, void *stream) {
  if (int rc = check_common(x, M, H, idx, KE, ARCQ_FLAVOUR_16, qx, sfx)) return rc;
  if (!w || ((uintptr_t)w & 15)) { set_error("rmsnorm weight must be a 16-byte aligned device pointer"); return ARCQ_EINVAL; }
  if (H < 2048 || H > 8192) {  // the reference's reduction tree assumes 128 <= H/16 <= 512 (rmsnorm.cu:131-146)
    set_error("rmsnorm_quantize_x supports 2048 <= H <= 8192, got %d", H);
    return ARCQ_ENOSUP;
  }
  if (M == 0) return ARCQ_OK;
  QuantParams p{(const uint16_t *)x, idx, (const uint16_t *)w, (uint8_t *)qx, (uint8_t *)sfx, nullptr, nullptr, nullptr, 0, eps, (int)M, H, KE, 0, 0};
  cudaStream_t st = (cudaStream_t)stream;
  if (use_tma()) {
    const int rc = launch_quant_tma(p, ARCQ_FLAVOUR_16, false, true, st);
    if (rc != ARCQ_ENOSUP) return rc;
  }
  switch (pick_rows(H, KE, true)) {
    case 8: return launch_quant<8, 0, false, true>(p, st);
    case 4: return launch_quant<4, 0, false, true>(p, st);
    case 2: return launch_quant<2, 0, false, true>(p, st);
    default: return launch_quant<1, 0, false, true>(p, st);
  }
}

// residual add fused in front of rmsnorm_quantize_x: h = bf16(x + res) -> sum_out, (qx, sfx) = rmsnorm_quantize_x(h)
extern "C" int arcq_add_rmsnorm_quantize_x(const void *x, const void *res, const void *w, float eps, int64_t M, int H,
                                           const int16_t *idx, int KE, void *sum_out, void *qx, void *sfx, void *stream) {
  if (int rc = check_common(x, M, H, idx, KE, ARCQ_FLAVOUR_16, qx, sfx)) return rc;
  if (!w || ((uintptr_t)w & 15)) { set_error("rmsnorm weight must be a 16-byte aligned device pointer"); return ARCQ_EINVAL; }
  if (!res || !sum_out || (((uintptr_t)res | (uintptr_t)sum_out) & 15)) {
    set_error("arcq_add_rmsnorm_quantize_x: res and sum_out must be 16-byte aligned device pointers");
    return ARCQ_EINVAL;
  }
  if (H < 2048 || H > 8192) { set_error("rmsnorm_quantize_x supports 2048 <= H <= 8192, got %d", H); return ARCQ_ENOSUP; }
  if (M == 0) return ARCQ_OK;
  QuantParams p{(const uint16_t *)x, idx, (const uint16_t *)w, (uint8_t *)qx, (uint8_t *)sfx, nullptr, nullptr, nullptr, 0, eps, (int)M, H, KE, 0, 0,
                (const uint16_t *)res, (uint16_t *)sum_out};
  const int rc = launch_quant_tma(p, ARCQ_FLAVOUR_16, false, true, (cudaStream_t)stream);
  if (rc == ARCQ_ENOSUP) set_error("arcq_add_rmsnorm_quantize_x: H=%d KE=%d is outside the TMA-staged kernel's shapes (H, KE multiples of 64); add separately", H, KE);
  return rc;
}

