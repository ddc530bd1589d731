#pragma once
// quant_tma.cuh -- TMA-staged reorder + NVFP4 + ARC quantise kernel (sm_100a): kernel and launch templates.
// Included by quant_tma.cu (dispatcher) and quant_tma_cfg*.cu (one tile organisation each, so that nvcc compiles them in parallel).
//
// Same arithmetic contract and output layout as quant.cu (reference
// kernels/src/{reorder,down,rmsnorm}.cu).  Organisation:
//
//  * persistent CTAs (two per SM when the tile is small enough); a work item is R rows of one 128-row scale block;
//  * ONE thread streams whole bf16 rows into a ring of row-major staging slots with cp.async.bulk (TMA, mbarrier
//    complete_tx): the next tile(s) are in flight while the current one is quantised, with no registers tied up;
//  * all threads transpose a landed slot (LDS.128 -> PRMT -> swizzled STS) into the channel-major tile
//    ("[channel][R rows]") so the permuted gather is one 2R-byte shared load per channel for all R rows;
//  * gather offsets come from a once-per-CTA table built from reorder_index; when a thread owns the same 16-group
//    in every tile (KQ/16 <= threads) its 16 offsets live in registers for the whole kernel;
//  * phase A = primary quantisation (one group x R rows per thread), phase B = ARC residuals spread over all
//    threads as (group,row) items; amax/6 by exact FMA division, RN(1/sf) and sf from a 127-entry LUT;
//  * scale bytes are assembled per tile in shared memory and leave as whole 4R-byte runs of the 128x4 layout.
#include "quant.cuh"
#include "quant_math.cuh"
#include "ptx.cuh"

// Work-skipping experiment bits exist only in -DARCQ_EXPERIMENT builds (scripts/build_variant.py); the product library
// compiles them out, so it cannot be made to skip its work.
#ifdef ARCQ_EXPERIMENT
#include <stdlib.h>
#define ARCQ_DBG(p, bit) (((p).debug & (bit)) != 0)
#else
#define ARCQ_DBG(p, bit) false
#endif

namespace arcq {
namespace {

struct CarveT {  // shared-memory layout (bytes), computed identically on host and device
  int lut, sfs, arc, tab, tabt, wtab, part, rstd, tile, stage, total;
};
// bytes added to every staged row: the 8 row addresses of one ldmatrix must fall into 8 different 16-byte bank groups.
// Plan mode reads (two tile rows) x (four 8-channel blocks) per matrix: the odd tile row must sit 64 bytes off.
__host__ __device__ constexpr int stage_pad_of(int R, bool plan) { return plan ? 64 : stage_pad(R); }
__host__ __device__ inline CarveT carve_t(int R, int KQ, int KE, bool rms, bool tab32, int S, bool plan = false, bool own2 = false,
                                          bool add = false) {
  CarveT c;
  int o = 128;                  // mbarriers full[S] (S <= 4)
  c.lut = o; o += 2 * 128 * 4;  // rcp + sf LUTs
  c.sfs = o; o += 2 * ((KQ + KE) / 64) * 4 * R;  // two scale staging buffers [chunk][lr][slot%4]
  o = (o + 15) & ~15;
  c.arc = o; o += (KE / 16) * R * 48;            // per ARC (group,row): 16 bf16 x, 2 code words, subtrahend scale
#ifdef ARCQ_STASH_LRMAJOR
  o += R * 16;  // experiment: [row][group] stash, rows padded by 16 bytes (conflict-free STS.128 and LDS.128)
#endif
  c.tab = o; if (!own2) o += KQ * (tab32 ? 4 : 2);   // own2: the table only lives during set-up, on top of the (still empty) tile
  o = (o + 15) & ~15;
  c.tabt = o; if (plan) o += KQ * 2;   // planned chunk positions in the order the transposing lanes need them
  o = (o + 15) & ~15;
  c.wtab = o; if (rms) o += KQ * 2;
  o = (o + 15) & ~15;
  c.part = o; if (rms) o += R * 512 * 4;
  c.rstd = o; if (rms) o += 64;
  o = (o + 127) & ~127;
  c.tile = o; o += KQ * 2 * R;
  if (own2) c.tab = c.tile;
  c.stage = o; o += (add ? 2 : 1) * S * (KQ * 2 + stage_pad_of(R, plan)) * R;   // add: the residual rows land in a second set of slots
  c.total = o;
  return c;
}

// PS: fused per-tensor pre-scale (activations only): every gathered element becomes bf16(x / (amax / 2688)) first
// PLAN: chunk positions come from a conflict-free placement plan (plan.cu) instead of the arithmetic swizzle
// OWN2: every thread owns up to TWO 16-groups (KQ/16 <= 2 * THREADS) and keeps their 32 gather offsets packed in 16
// registers, so the per-tile loads of the offset table (a third of the gather's shared-memory traffic for 2-row tiles)
// disappear and the table's shared memory goes to a second staging slot
// ADD (rmsnorm, unplanned): h = bf16(x + res) replaces x -- the residual rows are staged by TMA next to the x rows, the add
// happens in the registers of the transposing pass (which already holds whole 16-byte units of every row), h leaves for
// sum_out from there and the sum of squares is taken from h: `residual + hidden_states` followed by the norm
// (benchmarks/modeling_arc.py:294-301, model/qLlamaLayer.py:162-166) in one pass over the data
template <int R, int RT, int FLAVOUR, bool IS_W, bool RMS, bool TAB32, bool OWN, int THREADS, int MINB, bool PS, bool PLAN, bool OWN2 = false,
          bool ADD = false>
__global__ void __launch_bounds__(THREADS, MINB) quant_tma_kernel(const QuantParams p, const int S) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int CB = 2 * R;        // bytes of one channel chunk
  constexpr int RP = (R + 1) / 2;  // 32-bit words per chunk
  constexpr int TPB = 128 / R;     // tiles per 128-row block
  constexpr int RTP = (RT + 1) / 2;  // 32-bit words of the RT rows one thread quantises
  constexpr int RB = R / RT;         // row blocks per tile: a phase-A item is (16-group, row block)
  static_assert(R % RT == 0, "row blocks tile the tile");
  static_assert(!(OWN && TAB32), "owned groups convert the 16-bit table once");
  const int KQ = p.KQ, KE = p.KE, K = KQ + KE;
  const int G = (KQ - KE) / 16, ngroups = KQ / 16, nunits = KQ / 8, nchunks = K / 64;
  static_assert(!PLAN || (!RMS && (R == 8 || R == 4 || R == 2)), "plans exist for the matrix-transposed tiles");
  static_assert(!OWN2 || (PLAN && !OWN && !TAB32 && R == RT), "two owned groups per thread: planned, one row block per tile");
  static_assert(!ADD || (RMS && !PLAN && !IS_W), "the residual add rides on the register transposition of the rmsnorm variant");
  const CarveT cv = carve_t(R, KQ, KE, RMS, TAB32, S, PLAN, OWN2, ADD);
  uint64_t *full = reinterpret_cast<uint64_t *>(smem);
  float *rcp_lut = reinterpret_cast<float *>(smem + cv.lut), *sf_lut = rcp_lut + 128;
  const int sfs_bytes = nchunks * 4 * R;
  uint8_t *arcs = smem + cv.arc;
  uint8_t *tile = smem + cv.tile;
  uint8_t *stage = smem + cv.stage;
  constexpr bool MATX = !RMS && (R == 8 || R == 4 || R == 2);  // transpose with ldmatrix.trans + stmatrix
  const uint32_t rowbytes = (uint32_t)KQ * 2, rowstride = rowbytes + stage_pad_of(R, PLAN);
  const int tile_bytes = KQ * CB, slot_bytes = R * (int)rowstride;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float *part = reinterpret_cast<float *>(smem + cv.part), *rstd_s = reinterpret_cast<float *>(smem + cv.rstd);
  const uint16_t *wtab = reinterpret_cast<const uint16_t *>(smem + cv.wtab);

  // called by every lane of ONE warp with warp-uniform arguments: the address arithmetic stays on the uniform datapath
  // and one elected lane streams the rows of tile `tile_id` into staging slot s
  // Per-tile side jobs are spread over DIFFERENT warps so that no warp's chain between two CTA barriers carries more than
  // one of them (thread-0 clocks of round 1: TMA refill + scale copy-out 1190 clk and the residual phase 1072 clk of a
  // 5700-clk tile all sat on warp 0): the TMA issue goes to a middle warp, the scale copy-out starts at warp 2, the ARC
  // residual items go to the LAST warps.
  constexpr int NWARPS = THREADS / 32;
  constexpr int IW = RMS ? ((NWARPS > R) ? R : 0) : NWARPS / 2;  // issuing warp: not one of the rmsnorm tree warps
  auto issue_tile = [&](int s, int tile_id) {
    const int blk = tile_id / TPB, t = tile_id % TPB;
    const int r0 = blk * 128 + TileMap<R>::row(t, 0);
    const int mrows = p.M - r0;
    int nv = 0;
#pragma unroll
    for (int lr = 0; lr < R; ++lr) nv += TileMap<R>::row(0, lr) < mrows ? 1 : 0;
    const uint8_t *src = reinterpret_cast<const uint8_t *>(p.x) + (int64_t)r0 * rowbytes;
    uint8_t *dst = stage + (size_t)s * slot_bytes;
    if (ptx::elect_one()) {
      ptx::mbar_arrive_expect_tx(&full[s], (uint32_t)nv * rowbytes * (ADD ? 2u : 1u));
#pragma unroll
      for (int lr = 0; lr < R; ++lr)
        if (TileMap<R>::row(0, lr) < mrows)
          ptx::bulk_load(dst + (size_t)lr * rowstride, src + (size_t)TileMap<R>::row(0, lr) * rowbytes, rowbytes, &full[s]);
      if constexpr (ADD) {
        const uint8_t *rsrc = reinterpret_cast<const uint8_t *>(p.res) + (int64_t)r0 * rowbytes;
        uint8_t *rdst = dst + (size_t)S * slot_bytes;
#pragma unroll
        for (int lr = 0; lr < R; ++lr)
          if (TileMap<R>::row(0, lr) < mrows)
            ptx::bulk_load(rdst + (size_t)lr * rowstride, rsrc + (size_t)TileMap<R>::row(0, lr) * rowbytes, rowbytes, &full[s]);
      }
    }
    __syncwarp();
  };

  // ---- one-time setup: barriers, LUTs, gather-offset table ----
  if (tid == 0) {
    for (int s = 0; s < S; ++s) ptx::mbar_init(&full[s], 1);
    ptx::fence_barrier_init();
  }
  uint64_t *plan_bar = full + 8;              // PLAN: completion of the two set-up bulk copies
  uint16_t *plan_pos_s = reinterpret_cast<uint16_t *>(OWN2 ? tile + 2 * KQ : tile);   // PLAN: pos[] staged on the still-empty tile
  if constexpr (PLAN) {
    if (tid == 0) { ptx::mbar_init(plan_bar, 1); ptx::fence_barrier_init(); }
  }
  if (p.M % 128) {  // a ragged last block leaves rows of some slots unwritten: give them defined (zero) content
    for (int i = tid; i < (ADD ? 2 : 1) * S * slot_bytes / 16; i += THREADS) reinterpret_cast<uint4 *>(stage)[i] = make_uint4(0, 0, 0, 0);
    ptx::fence_proxy_async();
  }
  __syncthreads();
  if constexpr (PLAN) {   // pos[KQ] -> scratch on the tile, tabt[KQ] -> its table: two bulk copies instead of 2 KQ L2 lookups
    if (tid == 0) {
      ptx::mbar_arrive_expect_tx(plan_bar, (uint32_t)KQ * 4);
      ptx::bulk_load(plan_pos_s, p.plan, (uint32_t)KQ * 2, plan_bar);
      ptx::bulk_load(smem + cv.tabt, p.plan + KQ, (uint32_t)KQ * 2, plan_bar);
    }
  }
  if (warp == IW) {
    for (int s = 0; s < S; ++s) {
      const int tl = blockIdx.x + s * gridDim.x;
      if (tl < p.num_tiles) issue_tile(s, tl);
    }
  }
  if constexpr (RMS) {  // entries [KQ/16, 512) of the partial sums are never written: the tree reads them as zero
    for (int i = tid; i < R * 512; i += THREADS) part[i] = 0.0f;
  }
  if (tid < 128) {
    const float sf = e4m3_to_f32(tid < 127 ? tid : 126);
    sf_lut[tid] = sf;
    rcp_lut[tid] = __frcp_rn(sf);  // == f32(1.0 / (double)sf) for every E4M3 value (reorder.cu:146)
  }
  if constexpr (PLAN) ptx::mbar_wait(plan_bar, 0);
  for (int u = tid; u < nunits; u += THREADS) {
    const uint4 w4 = __ldg(reinterpret_cast<const uint4 *>(p.idx) + u);
    const uint32_t ww[4] = {w4.x, w4.y, w4.z, w4.w};
    uint32_t pos[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if constexpr (PLAN) {
        pos[2 * i] = plan_pos_s[ww[i] & 0xffffu];
        pos[2 * i + 1] = plan_pos_s[ww[i] >> 16];
      } else {
        pos[2 * i] = (uint32_t)(MATX ? chan_pos_m<R>((int)(ww[i] & 0xffffu)) : chan_pos<R>((int)(ww[i] & 0xffffu)));
        pos[2 * i + 1] = (uint32_t)(MATX ? chan_pos_m<R>((int)(ww[i] >> 16)) : chan_pos<R>((int)(ww[i] >> 16)));
      }
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
  __syncthreads();

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

  uint32_t own_off[OWN ? 16 : 1];
  if constexpr (OWN) load_offsets(tid / RB < ngroups ? tid / RB : 0, own_off);
  uint32_t own2[OWN2 ? 16 : 1];   // [item][8]: two byte offsets (< 65536) per word
  if constexpr (OWN2) {
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int g = tid + it * THREADS < ngroups ? tid + it * THREADS : 0;
      const uint4 *t4 = reinterpret_cast<const uint4 *>(smem + cv.tab) + 2 * g;
      const uint4 a = t4[0], b = t4[1];
      const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
      for (int q = 0; q < 8; ++q) own2[it * 8 + q] = w[q] * CB;   // both halves: position * CB < 65536, no carry
    }
    __syncthreads();   // the table sat on top of the tile: everybody has its offsets before the first transposition writes there
  }
  PrescaleCtx psc = {1.0f, 1.0f, false};
  if constexpr (PS) {
    float scale_fp32;
    psc = make_prescale(__ldg(p.amax), p.prescale_mode == ARCQ_PRESCALE_BF16_DIVISOR, scale_fp32);
    if (p.scale_out && blockIdx.x == 0 && tid == 0) *p.scale_out = scale_fp32;
  }

  // scale bytes of one finished tile leave as contiguous 4R-byte runs of the swizzled layout
  auto copy_out_sf = [&](int buf, int blk, int t) {
    const uint8_t *sfs = smem + cv.sfs + buf * sfs_bytes;
    uint8_t *dst0 = p.sf + (int64_t)blk * nchunks * 512 + TileMap<R>::w0(t) * 4;
    for (int c = (tid + THREADS - 64) % THREADS; c < nchunks; c += THREADS) {  // starts at warp 2
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
  };

  int n = 0, prev_blk = 0, prev_t = 0;
  for (int tile_id = blockIdx.x; tile_id < p.num_tiles; tile_id += gridDim.x, ++n) {
    const int s = n % S;
    const int blk = tile_id / TPB, t = tile_id % TPB;
    const int r0 = blk * 128 + TileMap<R>::row(t, 0);
    const int mrows = p.M - r0;  // row lr of this tile exists iff row(0, lr) < mrows
    const uint32_t khalf = (uint32_t)K / 2;
    uint8_t *qbase = p.q + (int64_t)r0 * khalf;
    uint8_t *sfs = smem + cv.sfs + (n & 1) * sfs_bytes;
    ptx::mbar_wait(&full[s], (uint32_t)(n / S) & 1u);

    // ---- transpose the landed rows into the channel-major tile ----
    {
      const uint8_t *src = stage + (size_t)s * slot_bytes;
      auto load_unit = [&](int u, uint4(&in)[R]) {
#pragma unroll
        for (int lr = 0; lr < R; ++lr) in[lr] = *reinterpret_cast<const uint4 *>(src + (size_t)lr * rowstride + (size_t)u * 16);
      };
      auto store_unit = [&](int u, const uint4(&in)[R]) {
        if constexpr (R == 1) {
          reinterpret_cast<uint4 *>(tile)[u] = in[0];
        } else {
          const uint32_t u8 = (uint32_t)u * 8, um = (uint32_t)u & (64 / R - 1);  // chan_pos(8u + c) = (8u | c) ^ um
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint32_t out[RP];
#pragma unroll
            for (int pr = 0; pr < RP; ++pr) {
              const uint32_t a = (&in[2 * pr].x)[c >> 1], b = (&in[2 * pr + 1].x)[c >> 1];
              out[pr] = __byte_perm(a, b, (c & 1) ? 0x7632 : 0x5410);
            }
            uint8_t *dst = tile + ((u8 | c) ^ um) * CB;
            if constexpr (R == 8) *reinterpret_cast<uint4 *>(dst) = make_uint4(out[0], out[1], out[2], out[3]);
            else if constexpr (R == 4) *reinterpret_cast<uint2 *>(dst) = make_uint2(out[0], out[1]);
            else *reinterpret_cast<uint32_t *>(dst) = out[0];
          }
        }
      };
      if constexpr (RMS) {
        // rmsnorm: the reference's thread tt sums channels 8tt..8tt+7 and then 8(bdx+tt)..+7 with one FMA chain
        // (rmsnorm.cu:113-130) = exactly units tt and bdx+tt of this pass, so the sum of squares rides along
        const int bdx = KQ / 16;
        for (int tt = tid; tt < bdx; tt += THREADS) {
          float acc[R];
#pragma unroll
          for (int lr = 0; lr < R; ++lr) acc[lr] = 0.0f;
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint4 in[R];
            load_unit(half * bdx + tt, in);
            if constexpr (ADD) {   // h = bf16(x + res), fp32 add like torch's bf16 add; h goes to sum_out and replaces x below
              const uint8_t *rsrc = src + (size_t)S * slot_bytes;
#pragma unroll
              for (int lr = 0; lr < R; ++lr) {
                const uint4 rr = *reinterpret_cast<const uint4 *>(rsrc + (size_t)lr * rowstride + (size_t)(half * bdx + tt) * 16);
                const uint32_t xa[4] = {in[lr].x, in[lr].y, in[lr].z, in[lr].w}, ra[4] = {rr.x, rr.y, rr.z, rr.w};
                uint32_t ha[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  __nv_bfloat162 h2 = __floats2bfloat162_rn(__fadd_rn(bf16lo_to_f32(xa[k]), bf16lo_to_f32(ra[k])),
                                                            __fadd_rn(bf16hi_to_f32(xa[k]), bf16hi_to_f32(ra[k])));
                  ha[k] = *reinterpret_cast<uint32_t *>(&h2);
                }
                in[lr] = make_uint4(ha[0], ha[1], ha[2], ha[3]);
                if (TileMap<R>::row(0, lr) < mrows)
                  __stcs(reinterpret_cast<uint4 *>(p.sum_out + ((int64_t)r0 + TileMap<R>::row(0, lr)) * KQ) + (half * bdx + tt), in[lr]);
              }
            }
            if constexpr (R % 2 == 0) {  // two rows per packed FFMA2, each row's chain in the reference's order
#pragma unroll
              for (int lr = 0; lr < R; lr += 2) {
                const uint32_t wa[4] = {in[lr].x, in[lr].y, in[lr].z, in[lr].w};
                const uint32_t wb[4] = {in[lr + 1].x, in[lr + 1].y, in[lr + 1].z, in[lr + 1].w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  fsqacc2(acc[lr], acc[lr + 1], bf16lo_to_f32(wa[i]), bf16lo_to_f32(wb[i]));
                  fsqacc2(acc[lr], acc[lr + 1], bf16hi_to_f32(wa[i]), bf16hi_to_f32(wb[i]));
                }
              }
            } else {
#pragma unroll
              for (int lr = 0; lr < R; ++lr) {
                const uint32_t w[4] = {in[lr].x, in[lr].y, in[lr].z, in[lr].w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const float lo = bf16lo_to_f32(w[i]), hi = bf16hi_to_f32(w[i]);
                  acc[lr] = __fmaf_rn(lo, lo, acc[lr]);
                  acc[lr] = __fmaf_rn(hi, hi, acc[lr]);
                }
              }
            }
            store_unit(half * bdx + tt, in);
          }
#pragma unroll
          for (int lr = 0; lr < R; ++lr) part[lr * 512 + tt] = acc[lr];
        }
      } else if constexpr (PLAN) {
        // a warp step moves CW = 256/R channels x R rows with ONE ldmatrix.x4.trans whose matrices are
        // (tile rows 2h, 2h+1) x (four 8-channel blocks): transposed, a lane holds tile rows 2h, 2h+1 of channel
        // 32 j + 8 (T % 4) + T / 4, so its registers ARE whole channel chunks (R = 8: one 16-byte chunk from 4 matrices,
        // R = 4: two 8-byte chunks, R = 2: four 4-byte chunks) and go straight to their planned positions.
        constexpr int NW = THREADS / 32, CW = 256 / R, J = CW / 32;
        const int m = lane >> 3, r = lane & 7;
        const uint32_t src_s = ptx::smem_u32(src), tile_s = ptx::smem_u32(tile);
        // matrix m covers sub-step j = m / (R / 2) and row pair h = m % (R / 2)
        const int jm = m / (R / 2), hm = m % (R / 2);
        const uint32_t la0 = src_s + (uint32_t)(2 * hm + (r & 1)) * rowstride + (uint32_t)(32 * jm + 8 * (r >> 1)) * 2;
        const uint8_t *tabt = smem + cv.tabt;
        const int nsteps = ARCQ_DBG(p, 1) ? 0 : KQ / CW;
#pragma unroll 2
        for (int k = warp; k < nsteps; k += NW) {
          uint32_t f[4];
          ptx::ldsm_x4_t(la0 + (uint32_t)k * (CW * 2), f);
          if constexpr (R == 8) {
            const uint32_t pos = *reinterpret_cast<const uint16_t *>(tabt + (k * 32 + lane) * 2);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(tile_s + pos * 16), "r"(f[0]), "r"(f[1]), "r"(f[2]), "r"(f[3]) : "memory");
          } else if constexpr (R == 4) {
            const uint32_t pp = *reinterpret_cast<const uint32_t *>(tabt + (k * 32 + lane) * 4);
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(tile_s + (pp & 0xffffu) * 8), "r"(f[0]), "r"(f[1]) : "memory");
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(tile_s + (pp >> 16) * 8), "r"(f[2]), "r"(f[3]) : "memory");
          } else {
            const uint2 pp = *reinterpret_cast<const uint2 *>(tabt + (k * 32 + lane) * 8);
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile_s + (pp.x & 0xffffu) * 4), "r"(f[0]) : "memory");
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile_s + (pp.x >> 16) * 4), "r"(f[1]) : "memory");
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile_s + (pp.y & 0xffffu) * 4), "r"(f[2]) : "memory");
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile_s + (pp.y >> 16) * 4), "r"(f[3]) : "memory");
          }
        }
      } else if constexpr (MATX) {
        // a warp step moves 256/R channels x R rows: matrix m = channels [CW k + MC m, +MC); its 8 rows are the R tile
        // rows of each of the 8/R 8-channel blocks; transposed, row i of the stored matrix is the 16 bytes
        // {chunk of channel 8 hb + i : hb} = chan_pos_m.  k = warp + NW it keeps the swizzle terms per-warp constants.
        constexpr int NW = THREADS / 32, CW = 256 / R, MC = CW / 4, HB = R == 8 ? 1 : 8 / R;
        static_assert(NW % 8 == 0, "swizzle terms must not depend on the iteration");
        const int m = lane >> 3, r = lane & 7;
        const uint32_t src_s = ptx::smem_u32(src), tile_s = ptx::smem_u32(tile);
        auto ld_addr = [&](int k) {
          const int hsw = (k >> 1) & (HB - 1);
          return src_s + (uint32_t)(r % R) * rowstride + (uint32_t)(k * CW + m * MC + 8 * ((r / R) ^ hsw)) * 2;
        };
        auto st_addr = [&](int k) {
          const int b = 4 * k + m;
          return tile_s + (uint32_t)((b * 8 + (r ^ (b & 7))) * 16);
        };
        const int nch = ARCQ_DBG(p, 1) ? 0 : KQ / CW;
        uint32_t la = ld_addr(warp), sa = st_addr(warp);
#pragma unroll 4
        for (int k = warp; k < nch; k += NW) {
          uint32_t f[4];
          ptx::ldsm_x4_t(la, f);
          ptx::stsm_x4(sa, f);
          la += NW * CW * 2;
          sa += NW * CW * CB;
        }
        if constexpr (R == 2) {  // KQ = 64 (mod 128): the last half step has two matrices
          if ((KQ & 64) && warp == (KQ / CW) % NW && !ARCQ_DBG(p, 1)) {
            uint32_t f[4];
            ptx::ldsm_x2_t(ld_addr(KQ / CW), f);
            ptx::stsm_x2(st_addr(KQ / CW), f);
          }
        }
      } else {
        for (int u = tid; u < (ARCQ_DBG(p, 1) ? 0 : nunits); u += THREADS) {
          uint4 in[R];
          load_unit(u, in);
          store_unit(u, in);
        }
      }
    }
    __syncthreads();
    if (warp == IW) {  // slot s has been read by everyone: refill it with the tile S steps ahead
      const int tl = tile_id + S * gridDim.x;
      if (tl < p.num_tiles) issue_tile(s, tl);
    }
    // every thread has finished phase B of the previous tile (it passed the barrier above): its scale bytes are complete
    if (n > 0) copy_out_sf((n - 1) & 1, prev_blk, prev_t);

    // ---- rmsnorm: the reference's reduction tree over the per-thread partial sums (rmsnorm.cu:131-157) ----
    if constexpr (RMS) {
      if (warp < R) {  // one warp replays the reference's tree for one row (rmsnorm.cu:131-154)
        // strides 256, 128, 64 and the final s[t] + s[t+32] pair element t with t + 32 k: all in the same lane, so the
        // whole shared-memory tree is 15 register adds per lane in the reference's association order
        const float *sp = part + warp * 512;
        float a[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) a[k] = sp[lane + 32 * k];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = __fadd_rn(a[k], a[k + 8]);
#pragma unroll
        for (int k = 0; k < 4; ++k) a[k] = __fadd_rn(a[k], a[k + 4]);
#pragma unroll
        for (int k = 0; k < 2; ++k) a[k] = __fadd_rn(a[k], a[k + 2]);
        float v = __fadd_rn(a[0], a[1]);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
        if (lane == 0) rstd_s[warp] = rsqrtf(__fadd_rn(__fdiv_rn(v, (float)KQ), p.eps));  // :157
      }
      __syncthreads();
    }

    // ---- phase A: permuted gather + primary quantisation; one item = (16-group, block of RT rows) per thread per step ----
    constexpr int NIT = OWN2 ? 2 : 1;
    for (int item0 = tid; item0 < ngroups * RB; item0 += NIT * THREADS) {
#pragma unroll
     for (int it = 0; it < NIT; ++it) {
      const int item = item0 + it * THREADS;
      if (NIT > 1 && item >= ngroups * RB) break;
      const int g = item / RB, rb = item % RB;
      uint32_t v[16][RTP];
      if (ARCQ_DBG(p, 2)) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
#pragma unroll
          for (int w = 0; w < RTP; ++w) v[i][w] = (uint32_t)g * 0x9e3779b9u + i;
      } else if constexpr (OWN) {
#pragma unroll
        for (int i = 0; i < 16; ++i) lds_chunk<RT>(tile + own_off[i] + rb * (2 * RT), v[i]);
      } else if constexpr (OWN2) {
#pragma unroll
        for (int i = 0; i < 16; ++i) lds_chunk<RT>(tile + ((own2[it * 8 + (i >> 1)] >> (16 * (i & 1))) & 0xffffu), v[i]);
      } else {
        uint32_t off[16];
        load_offsets(g, off);
#pragma unroll
        for (int i = 0; i < 16; ++i) lds_chunk<RT>(tile + off[i] + rb * (2 * RT), v[i]);
      }
      if constexpr (PS) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
#pragma unroll
          for (int w = 0; w < RTP; ++w) v[i][w] = prescale_bf16x2(v[i][w], psc);
      }
      float wf[RMS ? 16 : 1];
      if constexpr (RMS) load_w(g, wf);
      int slot_p, slot_r;
      slots(g, slot_p, slot_r);
      uint8_t *sfp = sfs + (slot_p >> 2) * (4 * R) + (slot_p & 3) + 4 * RT * rb;
      const int dslot = slot_r - slot_p;
      const bool arc = g >= G;
      const int drb = TileMap<R>::row(0, rb * RT);  // row(0, rb*RT + lr) = row(0, rb*RT) + row(0, lr) for every TileMap
      uint8_t *qg = qbase + 8 * slot_p + (uint32_t)drb * khalf;
      const int mr = mrows - drb;
      // store one quantised row: scale byte, codes, and (activations) the residual work item of an ARC group
      auto emit_row = [&](int lr, uint32_t sfb, uint32_t c_lo, uint32_t c_hi, float sc, const uint32_t(&xb)[8]) {
        sfp[4 * lr] = (uint8_t)sfb;
        if (IS_W && arc) sfp[4 * lr + ((dslot >> 2) * (4 * R) + (dslot & 3))] = (uint8_t)sfb;  // reorder.cu:307-316
        if (!IS_W && arc) {  // hand the residual work of this (group,row) to phase B: values, codes, subtrahend scale
#ifdef ARCQ_STASH_LRMAJOR
          uint4 *a = reinterpret_cast<uint4 *>(arcs + (rb * RT + lr) * ((KE / 16) * 48 + 16) + (g - G) * 48);
#else
          uint4 *a = reinterpret_cast<uint4 *>(arcs + ((g - G) * R + rb * RT + lr) * 48);
#endif
          a[0] = make_uint4(xb[0], xb[1], xb[2], xb[3]);
          a[1] = make_uint4(xb[4], xb[5], xb[6], xb[7]);
          // 16-wide: rounded scale (reorder.cu:157); 32-wide: unrounded (reorder.cu:474)
          a[2] = make_uint4(c_lo, c_hi, __float_as_uint(FLAVOUR == 0 ? sf_lut[sfb] : sc), 0u);
        }
        uint8_t *qp = qg + TileMap<R>::row(0, lr) * khalf;
        if (TileMap<R>::row(0, lr) < mr) {
          __stcs(reinterpret_cast<uint2 *>(qp), make_uint2(c_lo, c_hi));
          if (IS_W && arc) __stcs(reinterpret_cast<uint2 *>(qp + 8 * dslot), make_uint2(c_lo, c_hi));
        }
      };
      // two rows per step with packed FP32 multiplies: measured faster for the rmsnorm variant (-6..9 %) and for 2-row
      // items (-7 %), slower for plain 4-row items (+17 %)  (profiles/r1_quant_packed_ab.log)
      constexpr bool kPacked2 = (RT % 2 == 0) && (RMS || RT == 2);
      if constexpr (kPacked2) {
        // two rows per step: the chunk word w2 holds rows 2 w2 (low halves) and 2 w2 + 1 (high halves) of every channel
#pragma unroll
        for (int w2 = 0; w2 < RT / 2; ++w2) {
          uint32_t pv[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) pv[i] = v[i][w2];
          if constexpr (RMS) normalise16x2(pv, wf, rstd_s[rb * RT + 2 * w2], rstd_s[rb * RT + 2 * w2 + 1]);
          uint32_t sfb[2], c[2][2];
          float sc[2];
          quant16x2(pv, rcp_lut, sfb, c, sc);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint32_t xb[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) xb[i] = __byte_perm(pv[2 * i], pv[2 * i + 1], h ? 0x7632 : 0x5410);
            emit_row(2 * w2 + h, sfb[h], c[h][0], c[h][1], sc[h], xb);
          }
        }
      } else {
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
          float sc;
          uint32_t sfb;
          if (ARCQ_DBG(p, 4)) { sfb = __float_as_uint(x[0]) >> 24; c_lo = __float_as_uint(x[1]); c_hi = __float_as_uint(x[15]); sc = x[2]; }
          else sfb = quant16(x, rcp_lut, c_lo, c_hi, sc);
          uint32_t xb[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) xb[i] = __byte_perm(__float_as_uint(x[2 * i]), __float_as_uint(x[2 * i + 1]), 0x7632);
          emit_row(lr, sfb, c_lo, c_hi, sc, xb);
        }
      }
     }
      if constexpr (OWN || OWN2) break;  // every item is owned: its offsets stay in registers
    }

    // ---- phase B: ARC residuals, one (group, row) item per thread per step, from the stash of phase A ----
    if constexpr (!IS_W) {
      __syncthreads();  // stash complete; also: nobody reads the tile any more (the next transpose may overwrite it)
#ifndef ARCQ_PHASEB_QUAD  // one thread per (group,row) item; the 4-lanes-per-item variant below measured 5 % slower (r1l_quant.log)
      const int nitems = ARCQ_DBG(p, 8) ? 0 : (KE / 16) * R;
      const int bwarps = (nitems + 31) / 32;                                  // residual items sit on the last warps
      const int boff = bwarps < NWARPS ? (NWARPS - bwarps) * 32 : 0;
      for (int item = tid - boff; item < nitems; item += THREADS) {
        if (item < 0) continue;
#ifdef ARCQ_STASH_LRMAJOR
        const int nper = KE / 16, lr = item / nper, g = G + item % nper;
        const uint4 *a = reinterpret_cast<const uint4 *>(arcs + lr * (nper * 48 + 16) + (g - G) * 48);
#else
        const int g = G + item / R, lr = item % R;
        const uint4 *a = reinterpret_cast<const uint4 *>(arcs + item * 48);
#endif
        const uint4 x0 = a[0], x1 = a[1], cs = a[2];
        const uint32_t xb[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
        float x[16];
#pragma unroll
        for (int i = 0; i < 8; ++i) { x[2 * i] = bf16lo_to_f32(xb[i]); x[2 * i + 1] = bf16hi_to_f32(xb[i]); }
        uint32_t r_lo, r_hi;
        const uint32_t rsfb = residual16(x, cs.x, cs.y, __uint_as_float(cs.z), rcp_lut, r_lo, r_hi);
        int slot_p, slot_r;
        slots(g, slot_p, slot_r);
        sfs[(slot_r >> 2) * (4 * R) + 4 * lr + (slot_r & 3)] = (uint8_t)rsfb;
        const int dr = TileMap<R>::row(0, lr);
        if (dr < mrows) __stcs(reinterpret_cast<uint2 *>(qbase + (uint32_t)dr * khalf + 8 * slot_r), make_uint2(r_lo, r_hi));
      }
#else
      // four lanes share one (group,row) item: each requantises 4 of its 16 residuals, the quad reduces amax
      const int nitems = ARCQ_DBG(p, 8) ? 0 : (KE / 16) * R;
      const int sub = lane & 3;
      for (int base = warp * 8; base < nitems; base += (THREADS / 32) * 8) {  // warp-uniform: shuffles below
        const int it0 = base + (lane >> 2);
        const bool active = it0 < nitems;
        const int item = active ? it0 : nitems - 1;
        const int g = G + item / R, lr = item % R;
        const uint8_t *a = arcs + item * 48;
        const uint2 xw = reinterpret_cast<const uint2 *>(a)[sub];     // x[4 sub .. 4 sub + 3] as bf16
        const uint4 cs = reinterpret_cast<const uint4 *>(a)[2];      // code words, subtrahend scale
        const uint32_t nib = ((sub < 2 ? cs.x : cs.y) >> (16 * (sub & 1))) & 0xffffu;
        const float subs = __uint_as_float(cs.z);
        const float x4[4] = {bf16lo_to_f32(xw.x), bf16hi_to_f32(xw.x), bf16lo_to_f32(xw.y), bf16hi_to_f32(xw.y)};
        float e[4];
        float amax = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {  // e = bf16(x - deq(code) * scale) (reorder.cu:157-160 / :474)
          e[j] = bf16lo_to_f32(f32_to_bf16_bits(__fmaf_rn(-e2m1_to_f32((nib >> (4 * j)) & 0xfu), subs, x4[j])));
          amax = fmaxf(amax, fabsf(e[j]));
        }
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        const float sc2 = fmaxf(kScaleEps, fminf(kFp8Max, div6(amax)));
        const uint32_t rsfb = f32_to_e4m3(sc2);
        const float r = rcp_lut[rsfb];
        const uint32_t my16 = f32x4_to_e2m1(__fmul_rn(e[0], r), __fmul_rn(e[1], r), __fmul_rn(e[2], r), __fmul_rn(e[3], r));
        const uint32_t nb16 = __shfl_xor_sync(0xffffffffu, my16, 1);
        const uint32_t word = (sub & 1) ? (nb16 | (my16 << 16)) : (my16 | (nb16 << 16));  // lanes 0,1: elements 0-7; 2,3: 8-15
        const uint32_t other = __shfl_xor_sync(0xffffffffu, word, 2);
        if (active && sub == 0) {
          int slot_p, slot_r;
          slots(g, slot_p, slot_r);
          sfs[(slot_r >> 2) * (4 * R) + 4 * lr + (slot_r & 3)] = (uint8_t)rsfb;
          const int dr = TileMap<R>::row(0, lr);
          if (dr < mrows) __stcs(reinterpret_cast<uint2 *>(qbase + (uint32_t)dr * khalf + 8 * slot_r), make_uint2(word, other));
        }
      }
#endif
    } else {
      __syncthreads();
    }
    prev_blk = blk; prev_t = t;
  }
  if (n > 0) {  // scale bytes of the last tile
    __syncthreads();
    copy_out_sf((n - 1) & 1, prev_blk, prev_t);
  }
}

// ---- host side ------------------------------------------------------------------------------

constexpr int kSmemPerSm = 227 * 1024;

template <int R, int RT, int FLAVOUR, bool IS_W, bool RMS, bool TAB32, bool OWN, int THREADS, int MINB, bool PS, bool PLAN, bool OWN2 = false,
          bool ADD = false>
int launch(const QuantParams &p0, int S, cudaStream_t stream) {
  QuantParams p = p0;
#ifdef ARCQ_EXPERIMENT
  if (const char *e = getenv("ARCQ_QUANT_DEBUG")) p.debug = atoi(e);
#endif
  auto kern = quant_tma_kernel<R, RT, FLAVOUR, IS_W, RMS, TAB32, OWN, THREADS, MINB, PS, PLAN, OWN2, ADD>;
  const int smem = carve_t(R, p.KQ, p.KE, RMS, TAB32, S, PLAN, OWN2, ADD).total;
  static SmemAttrCache attr_cache = {};  // per instantiation, per device
  if (int rc = ensure_dynamic_smem(kern, smem, attr_cache)) return rc;
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  p.num_tiles = (int)(((int64_t)p.M + 127) / 128) * (128 / R);
  const int grid = p.num_tiles < sms * MINB ? p.num_tiles : sms * MINB;
  kern<<<grid, THREADS, smem, stream>>>(p, S);
  ARCQ_CUDA_TRY(cudaGetLastError());
  count_launch();
  return ARCQ_OK;
}

template <int R, int RT, int THREADS, int MINB, int FLAVOUR, bool IS_W, bool RMS, bool PS = false>
int launch_tab(const QuantParams &p, int S, bool tab32, bool own, cudaStream_t st, bool own2 = false) {
  if constexpr (!IS_W && !RMS && (R == 8 || R == 4 || R == 2)) {
    if (p.plan) {  // planned placement: 16-bit position tables only
      if constexpr (R == RT && R <= 4) {
        if (own2) return launch<R, RT, FLAVOUR, IS_W, RMS, false, false, THREADS, MINB, PS, true, true>(p, S, st);
      }
      if (own) return launch<R, RT, FLAVOUR, IS_W, RMS, false, true, THREADS, MINB, PS, true>(p, S, st);
      return launch<R, RT, FLAVOUR, IS_W, RMS, false, false, THREADS, MINB, PS, true>(p, S, st);
    }
  }
  if constexpr (RMS) {
    if (p.res) {   // fused residual add: unplanned register transposition only (the dispatcher drops the plan)
      if (own) return launch<R, RT, FLAVOUR, IS_W, RMS, false, true, THREADS, MINB, PS, false, false, true>(p, S, st);
      if (tab32) return launch<R, RT, FLAVOUR, IS_W, RMS, true, false, THREADS, MINB, PS, false, false, true>(p, S, st);
      return launch<R, RT, FLAVOUR, IS_W, RMS, false, false, THREADS, MINB, PS, false, false, true>(p, S, st);
    }
  }
  if (own) return launch<R, RT, FLAVOUR, IS_W, RMS, false, true, THREADS, MINB, PS, false>(p, S, st);
  if (tab32) return launch<R, RT, FLAVOUR, IS_W, RMS, true, false, THREADS, MINB, PS, false>(p, S, st);
  return launch<R, RT, FLAVOUR, IS_W, RMS, false, false, THREADS, MINB, PS, false>(p, S, st);
}

template <int R, int RT, int THREADS, int MINB>
int launch_kind(const QuantParams &p, int flavour, bool is_w, bool rms, int S, bool tab32, bool own, cudaStream_t st, bool own2 = false) {
  if (rms) return launch_tab<R, RT, THREADS, MINB, 0, false, true>(p, S, tab32, own, st);
  if (is_w) return flavour == ARCQ_FLAVOUR_32 ? launch_tab<R, RT, THREADS, MINB, 1, true, false>(p, S, tab32, own, st)
                                              : launch_tab<R, RT, THREADS, MINB, 0, true, false>(p, S, tab32, own, st);
  if (p.amax)
    return flavour == ARCQ_FLAVOUR_32 ? launch_tab<R, RT, THREADS, MINB, 1, false, false, true>(p, S, tab32, own, st, own2)
                                      : launch_tab<R, RT, THREADS, MINB, 0, false, false, true>(p, S, tab32, own, st, own2);
  return flavour == ARCQ_FLAVOUR_32 ? launch_tab<R, RT, THREADS, MINB, 1, false, false>(p, S, tab32, own, st, own2)
                                    : launch_tab<R, RT, THREADS, MINB, 0, false, false>(p, S, tab32, own, st, own2);
}


}  // namespace
}  // namespace arcq
