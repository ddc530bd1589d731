// quant_ws.cu -- warp-specialised, persistent reorder + NVFP4 + ARC quantise kernel (sm_100a).
//
// Same arithmetic contract and output layout as quant.cu (reference
// kernels/src/{reorder,down,rmsnorm}.cu), re-organised so that HBM loads never stop:
//
//  * one (or two) CTAs per SM, NP producer warps + NC consumer warps, two channel-major tile buffers in shared memory
//    handed over with mbarriers (full/empty), so loading tile n+1 overlaps quantising tile n;
//  * producers keep a two-deep ring of LB 128-bit loads per thread in flight (registers), transpose with PRMT and
//    store channel-major, XOR-swizzled, conflict-free;
//  * the reorder index is turned ONCE per CTA into a table of shared-memory byte offsets, so the permuted gather
//    is one 2R-byte shared load per channel for all R rows with no per-element address arithmetic;
//  * consumers do the primary quantisation of every 16-group (phase A); the ARC residual re-quantisation is
//    spread over all consumer threads as (group,row) items (phase B) so no warp carries the outlier tail alone;
//  * scale = amax/6 uses an exact reciprocal-FMA division (bit-identical to __fdiv_rn for every bf16 amax,
//    tests/test_oracle.py::test_div6_exact), 1/sf and sf come from a 127-entry shared LUT;
//  * scale bytes are assembled per tile in shared memory and leave as whole 4R-byte runs of the 128x4 layout.
#include "quant.cuh"
#include "quant_math.cuh"
#include "ptx.cuh"
#include <stdlib.h>

namespace arcq {
namespace {

template <int NCT> __device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(NCT) : "memory"); }

struct Carve {  // shared-memory layout (bytes), computed identically on host and device
  int lut, sfs, tab, wtab, part, rstd, tiles, total;
};
__host__ __device__ inline Carve carve(int R, int KQ, int KE, bool rms, bool tab32) {
  Carve c;
  int o = 128;                       // mbarriers: full[2], empty[2]
  c.lut = o; o += 2 * 128 * 4;       // rcp + sf LUTs
  c.sfs = o; o += 2 * ((KQ + KE) / 64) * 4 * R;  // two scale staging buffers [chunk][lr][slot%4]
  o = (o + 15) & ~15;
  c.tab = o; o += KQ * (tab32 ? 4 : 2);
  o = (o + 15) & ~15;
  c.wtab = o; if (rms) o += KQ * 2;
  o = (o + 15) & ~15;
  c.part = o; if (rms) o += R * 512 * 4;
  c.rstd = o; if (rms) o += 64;
  o = (o + 127) & ~127;
  c.tiles = o; o += 2 * KQ * 2 * R;
  c.total = o;
  return c;
}

template <int R, int RT, int FLAVOUR, bool IS_W, bool RMS, bool TAB32, int NP, int NC, int LB, int MINB>
__global__ void __launch_bounds__(32 * (NP + NC), MINB) quant_ws_kernel(const QuantParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int NPT = 32 * NP, NCT = 32 * NC;
  constexpr int CB = 2 * R;          // bytes of one channel chunk
  constexpr int RP = (R + 1) / 2;    // 32-bit words per chunk
  constexpr int RTP = (RT + 1) / 2;  // 32-bit words of the RT rows one consumer thread handles
  constexpr int RB = R / RT;         // row blocks per tile: a consumer item is (16-group, row block)
  constexpr int UB = LB / R;         // 8-channel units per producer batch
  static_assert(LB % R == 0 && UB >= 1, "batch must hold whole units");
  static_assert(RT == R || (R == 8 && RT == 4), "row(0, rb*RT + lr) = row(0, lr) + rb must hold");
  static_assert(NC >= R, "one consumer warp per row replays the rmsnorm tree");
  constexpr int TPB = 128 / R;       // tiles per 128-row block
  const int KQ = p.KQ, KE = p.KE, K = KQ + KE;
  const int G = (KQ - KE) / 16, ngroups = KQ / 16, nunits = KQ / 8, nchunks = K / 64;
  const Carve cv = carve(R, KQ, KE, RMS, TAB32);
  uint64_t *full = reinterpret_cast<uint64_t *>(smem), *empty = full + 2;
  float *rcp_lut = reinterpret_cast<float *>(smem + cv.lut), *sf_lut = rcp_lut + 128;
  uint8_t *tiles = smem + cv.tiles;
  const int tile_bytes = KQ * CB;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    ptx::mbar_init(&full[0], NP); ptx::mbar_init(&full[1], NP);
    ptx::mbar_init(&empty[0], NC); ptx::mbar_init(&empty[1], NC);
    ptx::fence_barrier_init();
  }
  __syncthreads();  // the only CTA-wide barrier: producers start loading while consumers build their tables

  if (warp < NP) {
    // =============================== producers ===============================
    const int nbatches = (nunits + UB - 1) / UB;
    const int ipt = (nbatches + NPT - 1) / NPT;  // batches per thread per tile (uniform; ragged ones are empty)
    const uint32_t rowbytes = (uint32_t)KQ * 2;
    auto issue = [&](uint4(&v)[LB], int tile_id, int k) {
      const int ub = tid + k * NPT;
      const int blk = tile_id / TPB, t = tile_id % TPB;
      const int r0 = blk * 128 + TileMap<R>::row(t, 0);  // row(t, lr) = row(t, 0) + row(0, lr) for every R
      const uint8_t *base = reinterpret_cast<const uint8_t *>(p.x) + (int64_t)r0 * rowbytes;
      const int mrows = p.M - r0;                          // rows of this tile with row(0, lr) < mrows exist
      if (mrows > TileMap<R>::row(0, R - 1)) {             // whole tile in range (the common case): no row checks
#pragma unroll
        for (int j = 0; j < UB; ++j) {
          const int u = ub * UB + j;
          if (u < nunits) {
#pragma unroll
            for (int lr = 0; lr < R; ++lr)
              v[j * R + lr] = __ldcs(reinterpret_cast<const uint4 *>(base + ((uint32_t)u * 16 + TileMap<R>::row(0, lr) * rowbytes)));
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < UB; ++j) {
          const int u = ub * UB + j;
#pragma unroll
          for (int lr = 0; lr < R; ++lr) {
            v[j * R + lr] = make_uint4(0, 0, 0, 0);
            if (u < nunits && TileMap<R>::row(0, lr) < mrows)
              v[j * R + lr] = __ldcs(reinterpret_cast<const uint4 *>(base + ((uint32_t)u * 16 + TileMap<R>::row(0, lr) * rowbytes)));
          }
        }
      }
    };
    auto commit = [&](uint4(&v)[LB], int k, int n) {
      const int buf = n & 1;
      if (k == 0) ptx::mbar_wait(&empty[buf], ((n >> 1) & 1) ^ 1);
      uint8_t *tile = tiles + buf * tile_bytes;
      const int ub = tid + k * NPT;
#pragma unroll
      for (int j = 0; j < UB; ++j) {
        const int u = ub * UB + j;
        if (u < nunits) {
          if constexpr (R == 1) {
            reinterpret_cast<uint4 *>(tile)[u] = v[j];
          } else {
            const uint32_t u8 = (uint32_t)u * 8, um = (uint32_t)u & (64 / R - 1);  // chan_pos(8u + c) = (8u | c) ^ um
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              uint32_t out[RP];
#pragma unroll
              for (int pr = 0; pr < RP; ++pr) {
                const uint32_t a = (&v[j * R + 2 * pr].x)[c >> 1], b = (&v[j * R + 2 * pr + 1].x)[c >> 1];
                out[pr] = __byte_perm(a, b, (c & 1) ? 0x7632 : 0x5410);
              }
              uint8_t *dst = tile + ((u8 | c) ^ um) * CB;
              if constexpr (R == 8) *reinterpret_cast<uint4 *>(dst) = make_uint4(out[0], out[1], out[2], out[3]);
              else if constexpr (R == 4) *reinterpret_cast<uint2 *>(dst) = make_uint2(out[0], out[1]);
              else *reinterpret_cast<uint32_t *>(dst) = out[0];
            }
          }
        }
      }
      if (k == ipt - 1) {
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&full[buf]);
      }
    };
    int tile0 = blockIdx.x, k0 = 0, n0 = 0;
    if (tile0 < p.num_tiles) {
      uint4 A[LB], B[LB];
      issue(A, tile0, k0);
      while (true) {
        int tile1 = tile0, k1 = k0 + 1, n1 = n0;
        if (k1 == ipt) { k1 = 0; tile1 += gridDim.x; ++n1; }
        const bool v1 = tile1 < p.num_tiles;
        if (v1) issue(B, tile1, k1);
        commit(A, k0, n0);
        if (!v1) break;
        tile0 = tile1; k0 = k1 + 1; n0 = n1;
        if (k0 == ipt) { k0 = 0; tile0 += gridDim.x; ++n0; }
        const bool v0 = tile0 < p.num_tiles;
        if (v0) issue(A, tile0, k0);
        commit(B, k1, n1);
        if (!v0) break;
      }
    }
  } else {
    // =============================== consumers ===============================
    const int ctid = tid - NPT, cw = warp - NP;
    float *part = reinterpret_cast<float *>(smem + cv.part), *rstd_s = reinterpret_cast<float *>(smem + cv.rstd);
    const uint16_t *wtab = reinterpret_cast<const uint16_t *>(smem + cv.wtab);
    const int sfs_bytes = nchunks * 4 * R;

    // ---- one-time setup: LUTs and the gather-offset table (reorder index -> byte offset inside a tile) ----
    if (ctid < 128) {
      const float sf = e4m3_to_f32(ctid < 127 ? ctid : 126);
      sf_lut[ctid] = sf;
      rcp_lut[ctid] = __frcp_rn(sf);  // == f32(1.0 / (double)sf) for every E4M3 value (reorder.cu:146)
    }
    for (int u = ctid; u < nunits; u += NCT) {
      const uint4 w4 = __ldg(reinterpret_cast<const uint4 *>(p.idx) + u);
      const uint32_t ww[4] = {w4.x, w4.y, w4.z, w4.w};
      uint32_t pos[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        pos[2 * i] = (uint32_t)chan_pos<R>((int)(ww[i] & 0xffffu));
        pos[2 * i + 1] = (uint32_t)chan_pos<R>((int)(ww[i] >> 16));
      }
      if constexpr (TAB32) {
        uint4 *dst = reinterpret_cast<uint4 *>(smem + cv.tab) + 2 * u;
        dst[0] = make_uint4(pos[0] * CB, pos[1] * CB, pos[2] * CB, pos[3] * CB);
        dst[1] = make_uint4(pos[4] * CB, pos[5] * CB, pos[6] * CB, pos[7] * CB);
      } else {
        reinterpret_cast<uint4 *>(smem + cv.tab)[u] =
            make_uint4(pos[0] | (pos[1] << 16), pos[2] | (pos[3] << 16), pos[4] | (pos[5] << 16), pos[6] | (pos[7] << 16));
      }
      if constexpr (RMS) {  // rmsnorm weight in reordered order
        uint32_t wv[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          wv[2 * i] = __ldg(p.w + (ww[i] & 0xffffu));
          wv[2 * i + 1] = __ldg(p.w + (ww[i] >> 16));
        }
        reinterpret_cast<uint4 *>(smem + cv.wtab)[u] =
            make_uint4(wv[0] | (wv[1] << 16), wv[2] | (wv[3] << 16), wv[4] | (wv[5] << 16), wv[6] | (wv[7] << 16));
      }
    }
    consumer_bar<NCT>();

    // gather offsets (bytes into the tile) of the 16 channels of group g
    auto load_offsets = [&](int g, uint32_t(&off)[16]) {
      if constexpr (TAB32) {
        const uint4 *t4 = reinterpret_cast<const uint4 *>(smem + cv.tab) + 4 * g;
#pragma unroll
        for (int q = 0; q < 4; ++q) { const uint4 w = t4[q]; off[4 * q] = w.x; off[4 * q + 1] = w.y; off[4 * q + 2] = w.z; off[4 * q + 3] = w.w; }
      } else {
        const uint4 *t4 = reinterpret_cast<const uint4 *>(smem + cv.tab) + 2 * g;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const uint4 w = t4[q];
          const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
          for (int i = 0; i < 4; ++i) { off[8 * q + 2 * i] = (ww[i] & 0xffffu) * CB; off[8 * q + 2 * i + 1] = (ww[i] >> 16) * CB; }
        }
      }
    };
    auto load_w = [&](int g, float(&wf)[16]) {
      const uint4 *t4 = reinterpret_cast<const uint4 *>(wtab) + 2 * g;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const uint4 w = t4[q];
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) { wf[8 * q + 2 * i] = bf16lo_to_f32(ww[i]); wf[8 * q + 2 * i + 1] = bf16hi_to_f32(ww[i]); }
      }
    };
    auto slots = [&](int g, int &slot_p, int &slot_r) {
      slot_p = g; slot_r = g;
      if (g >= G) {
        const int j = g - G;
        if constexpr (FLAVOUR == 0) { slot_p = G + 2 * j; slot_r = slot_p + 1; }  // reorder.cu:139,175
        else { slot_p = G + 4 * (j >> 1) + (j & 1); slot_r = slot_p + 2; }        // reorder.cu:451-463,510-515
      }
    };

    int n = 0;
    for (int tile_id = blockIdx.x; tile_id < p.num_tiles; tile_id += gridDim.x, ++n) {
      const int buf = n & 1;
      const int blk = tile_id / TPB, t = tile_id % TPB;
      const int r0 = blk * 128 + TileMap<R>::row(t, 0);
      const int mrows = p.M - r0;  // row lr of this tile exists iff row(0, lr) < mrows
      const uint32_t khalf = (uint32_t)K / 2;
      uint8_t *qbase = p.q + (int64_t)r0 * khalf;
      uint8_t *sfs = smem + cv.sfs + buf * sfs_bytes;
      ptx::mbar_wait(&full[buf], (n >> 1) & 1);
      const uint8_t *tile = tiles + buf * tile_bytes;

      // ---- rmsnorm: sum of squares in the reference's exact order (rmsnorm.cu:113-157) ----
      if constexpr (RMS) {
        const int bdx = KQ / 16;
        for (int tt = ctid; tt < 512; tt += NCT) {
          float acc[R];
#pragma unroll
          for (int lr = 0; lr < R; ++lr) acc[lr] = 0.0f;
          if (tt < bdx) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const int c = half * 8 * bdx + 8 * tt + i;
                uint32_t v[RP];
                lds_chunk<R>(tile + (size_t)chan_pos<R>(c) * CB, v);
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
        consumer_bar<NCT>();
        if (cw < R) {  // one warp replays the reference's tree for one row (rmsnorm.cu:131-154)
          float *s = part + cw * 512;
          for (int k = 0; k < 8; ++k) s[lane + 32 * k] += s[lane + 32 * k + 256];
          __syncwarp();
          for (int k = 0; k < 4; ++k) s[lane + 32 * k] += s[lane + 32 * k + 128];
          __syncwarp();
          for (int k = 0; k < 2; ++k) s[lane + 32 * k] += s[lane + 32 * k + 64];
          __syncwarp();
          float v = s[lane] + s[lane + 32];
#pragma unroll
          for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
          if (lane == 0) rstd_s[cw] = rsqrtf(__fadd_rn(__fdiv_rn(v, (float)KQ), p.eps));  // :157
        }
        consumer_bar<NCT>();
      }

      // ---- phase A: permuted gather + primary quantisation; one item = (16-group, block of RT rows) ----
      for (int item = ctid; item < ngroups * RB; item += NCT) {
        const int g = item / RB, rb = item % RB;
        uint32_t off[16];
        load_offsets(g, off);
        uint32_t v[16][RTP];
#pragma unroll
        for (int i = 0; i < 16; ++i) lds_chunk<RT>(tile + off[i] + rb * (2 * RT), v[i]);
        float wf[RMS ? 16 : 1];
        if constexpr (RMS) load_w(g, wf);
        int slot_p, slot_r;
        slots(g, slot_p, slot_r);
        uint8_t *sfp = sfs + (slot_p >> 2) * (4 * R) + (slot_p & 3) + 4 * RT * rb;
        const int dslot = slot_r - slot_p;
        const bool arc = g >= G;
        uint8_t *qg = qbase + 8 * slot_p + (uint32_t)rb * khalf;  // row(0, rb*RT + lr) = row(0, lr) + rb
        const int mr = mrows - rb;
#pragma unroll
        for (int lr = 0; lr < RT; ++lr) {
          float x[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            x[i] = (lr & 1) ? bf16hi_to_f32(v[i][lr >> 1]) : bf16lo_to_f32(v[i][lr >> 1]);
            if constexpr (RMS)  // rmsnorm.cu:165-171: bf16((x * w) * rstd), left to right
              x[i] = bf16lo_to_f32(f32_to_bf16_bits(__fmul_rn(__fmul_rn(x[i], wf[i]), rstd_s[rb * RT + lr])));
          }
          uint32_t c_lo, c_hi;
          float s;
          const uint32_t sfb = quant16(x, rcp_lut, c_lo, c_hi, s);
          sfp[4 * lr] = (uint8_t)sfb;
          if (IS_W && arc) sfp[4 * lr + ((dslot >> 2) * (4 * R) + (dslot & 3))] = (uint8_t)sfb;  // weights: ARC slots duplicate the primary (reorder.cu:307-316)
          uint8_t *qp = qg + TileMap<R>::row(0, lr) * khalf;
          if (TileMap<R>::row(0, lr) < mr) {
            __stcs(reinterpret_cast<uint2 *>(qp), make_uint2(c_lo, c_hi));
            if (IS_W && arc) __stcs(reinterpret_cast<uint2 *>(qp + 8 * dslot), make_uint2(c_lo, c_hi));
          }
        }
      }

      // ---- phase B: ARC residuals, one (group, row) item per thread per step ----
      if constexpr (!IS_W) {
        const int nitems = (KE / 16) * R;
        for (int item = ctid; item < nitems; item += NCT) {
          const int g = G + item / R, lr = item % R;
          uint32_t off[16];
          load_offsets(g, off);
          float wf[RMS ? 16 : 1];
          if constexpr (RMS) load_w(g, wf);
          float x[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            x[i] = __uint_as_float((uint32_t)*reinterpret_cast<const uint16_t *>(tile + off[i] + 2 * lr) << 16);
            if constexpr (RMS) x[i] = bf16lo_to_f32(f32_to_bf16_bits(__fmul_rn(__fmul_rn(x[i], wf[i]), rstd_s[lr])));
          }
          uint32_t c_lo, c_hi, r_lo, r_hi;
          float s;
          const uint32_t sfb = quant16(x, rcp_lut, c_lo, c_hi, s);
          const float sub = FLAVOUR == 0 ? sf_lut[sfb] : s;  // 16-wide: rounded scale (:157); 32-wide: unrounded (:474)
          const uint32_t rsfb = residual16(x, c_lo, c_hi, sub, rcp_lut, r_lo, r_hi);
          int slot_p, slot_r;
          slots(g, slot_p, slot_r);
          sfs[(slot_r >> 2) * (4 * R) + 4 * lr + (slot_r & 3)] = (uint8_t)rsfb;
          const int dr = TileMap<R>::row(0, lr);
          if (dr < mrows) __stcs(reinterpret_cast<uint2 *>(qbase + (uint32_t)dr * khalf + 8 * slot_r), make_uint2(r_lo, r_hi));
        }
      }

      // the tile buffer is free again; the scale bytes of this tile are complete after the consumer barrier
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&empty[buf]);
      consumer_bar<NCT>();

      // ---- scale bytes leave as contiguous 4R-byte runs of the swizzled layout ----
      {
        uint8_t *dst0 = p.sf + (int64_t)blk * nchunks * 512 + TileMap<R>::w0(t) * 4;
        for (int c = ctid; c < nchunks; c += NCT) {
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
      // sfs[buf] is rewritten by tile n+2, i.e. after the consumer barrier of tile n+1, which every thread
      // reaches only after finishing this copy-out.
    }
  }
}

// ---- host side ------------------------------------------------------------------------------

constexpr int kMaxSmem = 227 * 1024;

template <int R, int RT, int FLAVOUR, bool IS_W, bool RMS, bool TAB32, int NP, int NC, int LB, int MINB>
int launch(const QuantParams &p0, cudaStream_t stream) {
  QuantParams p = p0;
  auto kern = quant_ws_kernel<R, RT, FLAVOUR, IS_W, RMS, TAB32, NP, NC, LB, MINB>;
  const int smem = carve(R, p.KQ, p.KE, RMS, TAB32).total;
  static int smem_set = 0;  // per instantiation
  if (smem > smem_set) {
    ARCQ_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    smem_set = smem;
  }
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  p.num_tiles = (int)(((int64_t)p.M + 127) / 128) * (128 / R);
  int per_sm = MINB;
  while (per_sm > 1 && per_sm * (smem + 1024) > 228 * 1024) --per_sm;
  const int grid = p.num_tiles < sms * per_sm ? p.num_tiles : sms * per_sm;
  kern<<<grid, 32 * (NP + NC), smem, stream>>>(p);
  ARCQ_CUDA_TRY(cudaGetLastError());
  count_launch();
  return ARCQ_OK;
}

// thread organisation: (rows per consumer thread, producer warps, consumer warps, loads per batch, CTAs/SM).
// Measured on B200 (profiles/r1_quant_notes.md): 8 producer + 8 consumer warps with 8 rows per consumer thread is the
// fastest organisation of the ones tried; tiles narrower than 8 rows are slower than the single-role kernel of
// quant.cu, so only R = 8 is instantiated and launch_quant_ws declines shapes whose 8-row tile does not fit.
template <int FLAVOUR, bool IS_W, bool RMS>
int launch_tab(const QuantParams &p, bool tab32, int cfg, cudaStream_t st) {
#define ARCQ_WS(RT, NP, NC, LB, MINB)                                                              \
  (tab32 ? launch<8, RT, FLAVOUR, IS_W, RMS, true, NP, NC, LB, MINB>(p, st)                         \
         : launch<8, RT, FLAVOUR, IS_W, RMS, false, NP, NC, LB, MINB>(p, st))
  if (cfg == 2) return ARCQ_WS(4, 8, 16, 8, 1);
  return ARCQ_WS(8, 8, 8, 8, 1);
#undef ARCQ_WS
}

}  // namespace

int launch_quant_ws(const QuantParams &p, int flavour, bool is_w, bool rms, cudaStream_t stream) {
  if (p.KQ % 64 || p.KE % 64) return ARCQ_ENOSUP;  // 4-slot scale chunks must not straddle the plain/ARC boundary
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  static const int cfg = getenv("ARCQ_QUANT_CFG") ? atoi(getenv("ARCQ_QUANT_CFG")) : 0;  // experiment knob
  const int64_t blocks = ((int64_t)p.M + 127) / 128;
  if (blocks * 16 < sms) return ARCQ_ENOSUP;  // fewer 8-row tiles than SMs: the narrow-tile kernel spreads wider
  const bool fits32 = carve(8, p.KQ, p.KE, rms, true).total <= kMaxSmem;
  const bool fits16 = carve(8, p.KQ, p.KE, rms, false).total <= kMaxSmem;
  if (!fits32 && !fits16) return ARCQ_ENOSUP;
  if (rms) return launch_tab<0, false, true>(p, fits32, cfg, stream);
  if (is_w) return flavour == ARCQ_FLAVOUR_32 ? launch_tab<1, true, false>(p, fits32, cfg, stream) : launch_tab<0, true, false>(p, fits32, cfg, stream);
  return flavour == ARCQ_FLAVOUR_32 ? launch_tab<1, false, false>(p, fits32, cfg, stream) : launch_tab<0, false, false>(p, fits32, cfg, stream);
}

}  // namespace arcq
