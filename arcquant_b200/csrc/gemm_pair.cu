// gemm_pair.cu -- NVFP4 block-scaled GEMM with tcgen05 `cta_group::2`: two CTAs on the two SMs of a TPC
// compute one 256 x 256 output tile; each keeps its own 128 rows of A, HALF of the B tile (128 of the 256
// n-rows) and its own 128 x 256 fp32 accumulator in TMEM.  The tensor cores read the peer's half of B across
// the SM pair, so an SM ingests 16 KB (A) + 16 KB (B/2) + 6 KB (scale factors) per 256-deep k-block instead of
// the 54 KB of the single-CTA kernel (gemm.cu) -- measured there: the 64 B/clk/SM L2->SM ingest, not the
// tensor pipe, is what bounds a 128 x 256 single-CTA tile (profiles/r1_gemm_notes.md).
//
// Protocol (cluster = the CTA pair; rank 0 = leader):
//   full[s]   (leader's copy is the live one) count 1: the leader's producer posts expect_tx for BOTH CTAs'
//             bytes; every TMA of either CTA is issued with .cta_group::2 and credits the LEADER's barrier.
//   empty[s]  one per CTA, count 1: the leader's MMA thread commits with a multicast to both CTAs.
//   tfull     one per CTA, count 1: multicast commit after the last k-block.
//   tempty    leader's copy, count 16: all eight epilogue warps of both CTAs arrive (remote arrive from rank 1).
// Only the leader issues tcgen05.cp / tcgen05.mma (cta_group::2 makes both SMs execute them on their own
// shared memory / TMEM at the same offsets).
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"

// per-role cycle counters (scripts/gemm_prof.py): build with -DARCQ_GEMM_PROFILE, run with ARCQ_GEMM_PROF=1
#ifdef ARCQ_GEMM_PROFILE
#include <stdlib.h>
#define PROF(...) __VA_ARGS__
#else
#define PROF(...)
#endif

namespace arcq {

int cached_operand_map(CUtensorMap *out, const void *ptr, int64_t rows, int64_t row_bytes, int box_rows);
int cached_sf_map(CUtensorMap *out, const void *ptr, int64_t natoms);

namespace {

constexpr int kRows = 128;                 // rows of A (and of the B half) per CTA
constexpr int kBN = 256, kBKBytes = 128, kBK = 256;
constexpr int kThreadsP = 320;             // warp 0 TMA, warp 1 MMA (leader only issues), warps 2..9 epilogue
constexpr int kPC = 64;                    // epilogue staging pass width (columns)

constexpr int kSiluLo = 103, kSiluHi = 133;                       // biased bf16 exponents covered by the silu table: 2^-24 .. 2^7
constexpr int kSiluEntries = 2 * (kSiluHi - kSiluLo + 1) * 128;   // sign x exponent x 7-bit mantissa = 7936 bf16 values

template <int STAGES, bool LUT = false>
struct PCfg {
  static constexpr int A_BYTES = kRows * kBKBytes;     // 16 KiB
  static constexpr int B_BYTES = kRows * kBKBytes;     // this CTA's half of the B tile
  static constexpr int SFA_BYTES = 4 * 512;
  static constexpr int SFB_BYTES = 2 * 4 * 512;        // both 128-row scale blocks of the 256-wide tile
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + SFA_BYTES + SFB_BYTES;   // 38912
  static constexpr int SF_COLS = 48, SF_COL0 = kBN, TMEM_COLS = 512;
  static constexpr int EPI_WARP_BYTES = 32 * kPC * 2;
  static constexpr int EPI_BYTES = 8 * EPI_WARP_BYTES;
  static constexpr int LUT_BYTES = LUT ? kSiluEntries * 2 : 0;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + 256 + LUT_BYTES;
  static_assert(SMEM_BYTES <= 232448, "shared memory overflow");
};

struct PParams {
  __nv_bfloat16 *d;
  const float *alpha_dev;
  const float *col_scale;   // nullable fp32 [N]: column n is scaled by RN(alpha * col_scale[n]) (fused q/k/v, gate/up launches)
  float alpha;
  int M, N, K;
  int64_t ldd;              // row stride of d in elements
  int ptiles_m, ptiles_n;   // 256 x 256 pair tiles
  long long *prof;          // -DARCQ_GEMM_PROFILE builds: per-role cycle counters of block 0/1 (nullable)
  // reduce-scatter epilogue (peer.cu): rs_rows != 0 -> the 128-row block starting at global row g goes to the
  // receive buffer of its owner o = g / rs_rows, slot rs_rank: peer_d[o] + ((rs_rank * rs_rows) + g - o * rs_rows) * ldd
  __nv_bfloat16 *peer_d[ARCQ_MAX_PEERS];
  int rs_rows, rs_rank;
  const __nv_bfloat16 *col_bias;   // nullable bf16 [N] (plain epilogue only): torch's `y + bias` on the rounded output
  int raster_gm;                   // 0: tile t -> (t % ptiles_m, t / ptiles_m); g > 0: groups of g row tiles, column tiles inside a group
};

// bf16(silu(g)) * u exactly as torch computes `F.silu(g) * u` on bf16 tensors: silu in fp32 (x / (1 + exp(-x)), IEEE
// division, libdevice expf), rounded to bf16, then an fp32 multiply rounded to bf16.
// Saturated activations must not make the epilogue slow: for x < -88.7 expf(-x) is +inf (quotient exactly -0) and just
// above that the divisor sits at the top of the fp32 range, where div.rn.f32 leaves its fast path for a ~200-instruction
// subroutine (measured: 4.6x on the whole GEMM with |x| ~ 100).  Numerator and divisor are therefore brought to mid-range
// by exact power-of-two factors (x 2^32, / 2^32) and the quotient scaled back by 2^-64: all three steps are exact, the
// one rounding is the division's, so the result is bit-identical to the plain division.
__device__ __forceinline__ float silu_bf16_exact(float gb) {   // gb holds a bf16 value; returns bf16(silu(gb)) as float
  const float den = 1.0f + expf(-gb);                      // in [1, +inf]
  const bool big = den > 1.0e18f;                          // only for gb < -41: |quotient| < 4e-17, far above denormals
  const float d = fminf(den, 3.4028235e38f) * (big ? 2.3283064365386963e-10f : 1.0f);   // 2^-32
  const float n = gb * (big ? 4294967296.0f : 1.0f);                                     // 2^32
  float q = __fdiv_rn(n, d) * (big ? 5.421010862427522e-20f : 1.0f);                    // 2^-64
  if (den > 3.4028235e38f) q = -0.0f;                      // gb / +inf (gb is negative here)
  return __bfloat162float(__float2bfloat16_rn(q));
}

// The gate is rounded to bf16 BEFORE silu (torch applies silu to the bf16 GEMM output), so silu is a function of 16 bits.
// Outside 2^-24 <= |g| < 2^7 it is trivial and exact (g/2 below: 1 + exp(-g) rounds to 2; g above for g > 0: the sum
// rounds to 1; -0 above for g < 0: exp(-g) overflows), inside it comes from a 7936-entry table that every CTA fills once
// with silu_bf16_exact -- same bits as the formula, a dozen instructions and one shared load instead of ~45 per output.
__device__ __forceinline__ uint32_t silu_lut_index(uint32_t bits) {
  return ((((bits >> 15) & 1u) * (kSiluHi - kSiluLo + 1) + ((bits >> 7) & 0xffu) - kSiluLo) << 7) | (bits & 127u);
}
__device__ __forceinline__ float swiglu_bf16(float g, float u, const uint16_t *lut) {
  const uint32_t gbits = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(g));
  const float gb = __uint_as_float(gbits << 16), ub = __bfloat162float(__float2bfloat16_rn(u));
  const uint32_t e = (gbits >> 7) & 0xffu;
  const uint32_t idx = min(silu_lut_index(gbits), (uint32_t)(kSiluEntries - 1));   // clamped: the load is unconditional
  const float t = __uint_as_float((uint32_t)lut[e < kSiluLo ? 0u : idx] << 16);
  const float small = __bfloat162float(__float2bfloat16_rn(gb * 0.5f));
  const float large = (gbits & 0x8000u) ? -0.0f : gb;
  float s = e < kSiluLo ? small : (e > kSiluHi ? large : t);
  if (e == 255u && gbits != 0x7f80u) s = __uint_as_float(0x7fc00000u);   // NaN in, or -inf / (1 + inf): NaN like the formula
  return __fmul_rn(s, ub);
}

// scale of tile column c (0..HC-1 of this warp's block): lane l keeps columns l, l + 32, ... in registers (loaded once per
// tile BEFORE the accumulator wait, already multiplied by alpha); a shuffle fetches the one a store needs
template <int NREG>
__device__ __forceinline__ void load_col_scales(float (&sreg)[NREG], const float *cs, float alpha, int col0, int N, int lane) {
#pragma unroll
  for (int r = 0; r < NREG; ++r) {
    const int c = col0 + r * 32 + lane;
    sreg[r] = (cs && c < N) ? __fmul_rn(alpha, __ldg(cs + c)) : alpha;
  }
}

// persistent tile order.  Default: row tiles fastest (a wave of 74 pairs covers every row tile of 2-3 column tiles).  Grouped
// (arcq_set_option "gemm_raster" = g): g row tiles x all column tiles per group, row tiles fastest inside a group, so a
// wave touches ~g row panels of A and ~74/g column panels of B (a smaller unique footprint per wave)
__device__ __forceinline__ void tile_coords(const PParams &p, int t, int &pm, int &pn) {
  if (p.raster_gm <= 0) { pm = t % p.ptiles_m; pn = t / p.ptiles_m; return; }
  const int per_group = p.raster_gm * p.ptiles_n;
  const int group = t / per_group, first_m = group * p.raster_gm;
  const int gsz = min(p.raster_gm, p.ptiles_m - first_m);
  const int r = t - group * per_group;
  pm = first_m + r % gsz;
  pn = r / gsz;
}

// SWIGLU: the 256 columns of a pair tile are 128 "gate" columns followed by the 128 "up" columns of the SAME 128
// channels (the fused gate/up weight is interleaved in 128-row blocks, which are whole scale-factor blocks); the
// epilogue writes d[M, N/2] = bf16(silu(gate)) * up and the [M, N] intermediate never reaches HBM.
template <int STAGES, bool SWIGLU>
__global__ void __launch_bounds__(kThreadsP, 1)
gemm_nvfp4_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                       const __grid_constant__ CUtensorMap tmap_sfa, const __grid_constant__ CUtensorMap tmap_sfb,
                       const __grid_constant__ PParams p) {
  using Cfg = PCfg<STAGES, SWIGLU>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *epi = smem + STAGES * Cfg::STAGE_BYTES;
  uint64_t *bars = reinterpret_cast<uint64_t *>(epi + Cfg::EPI_BYTES);
  uint16_t *silu_lut = reinterpret_cast<uint16_t *>(epi + Cfg::EPI_BYTES + 256);
  if constexpr (SWIGLU) {   // filled before the first barrier below, read only by the epilogue warps
    for (int i = threadIdx.x; i < kSiluEntries; i += kThreadsP) {
      const uint32_t sgn = i / ((kSiluHi - kSiluLo + 1) * 128), rem = i % ((kSiluHi - kSiluLo + 1) * 128);
      const uint32_t bits = (sgn << 15) | (((rem >> 7) + kSiluLo) << 7) | (rem & 127u);
      silu_lut[i] = (uint16_t)(__float_as_uint(silu_bf16_exact(__uint_as_float(bits << 16))) >> 16);
    }
  }
  uint64_t *full = bars, *empty = bars + STAGES, *tfull = bars + 2 * STAGES, *tempty = tfull + 1;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();     // 0 = leader
  const int pair_id = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int K = p.K, nkb = (K + kBK - 1) / kBK, kc_total = K / 64;
  const int num_ptiles = p.ptiles_m * p.ptiles_n;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_a);
    ptx::prefetch_tmap(&tmap_b);
    ptx::prefetch_tmap(&tmap_sfa);
    ptx::prefetch_tmap(&tmap_sfb);
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
    ptx::mbar_init(tfull, 1);
    ptx::mbar_init(tempty, 16);
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc_pair(tmem_slot, Cfg::TMEM_COLS);
    ptx::tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs; completion bytes go to the leader's full[s]) ==========
    uint32_t it = 0;
    PROF(long long pw = 0, pw_n = 0, pt0 = clock64();)
    uint32_t ready = ptx::mbar_test(&empty[0], 1);
    for (int t = pair_id; t < num_ptiles; t += num_pairs) {
      int pm, pn;
      tile_coords(p, t, pm, pn);
      const int mblk = pm * 2 + (int)rank;        // this CTA's 128-row block of A / SFA
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        PROF(long long tp0 = clock64();)
        if (!ready) { ptx::mbar_wait(&empty[s], ph ^ 1); PROF(++pw_n;) }
        PROF(pw += clock64() - tp0;)
        ready = ptx::mbar_test(&empty[(it + 1) % STAGES], (((it + 1) / STAGES) & 1) ^ 1);
        uint8_t *st = smem + s * Cfg::STAGE_BYTES;
        if (ptx::elect_one()) {
          const uint32_t lbar = ptx::mapa(&full[s], 0);
          {
          if (rank == 0) ptx::mbar_arrive_expect_tx(&full[s], 2 * Cfg::STAGE_BYTES);
          ptx::tma_load_2d_pair(st, &tmap_a, kb * kBKBytes, mblk * kRows, lbar);
          ptx::tma_load_2d_pair(st + Cfg::A_BYTES, &tmap_b, kb * kBKBytes, pn * kBN + (int)rank * kRows, lbar);
          // scale factors: rows of the SF maps are 512-B atoms, atom index = block * (K/64) + chunk
          ptx::tma_load_2d_pair(st + Cfg::A_BYTES + Cfg::B_BYTES, &tmap_sfa, 0, mblk * kc_total + kb * 4, lbar);
          ptx::tma_load_2d_pair(st + Cfg::A_BYTES + Cfg::B_BYTES + Cfg::SFA_BYTES, &tmap_sfb, 0,
                                (pn * 2) * kc_total + kb * 4, lbar);
          ptx::tma_load_2d_pair(st + Cfg::A_BYTES + Cfg::B_BYTES + Cfg::SFA_BYTES + 2048, &tmap_sfb, 0,
                                (pn * 2 + 1) * kc_total + kb * 4, lbar);
          }
        }
        __syncwarp();
      }
    }
    PROF(if (p.prof && blockIdx.x < 2 && lane == 0) { long long *o = p.prof + blockIdx.x * 16; o[0] = clock64() - pt0; o[1] = pw; o[2] = pw_n; o[3] = it; })
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (rank == 0) {
      constexpr uint32_t idesc = ptx::make_idesc_nvfp4(256, kBN);
      const uint32_t base = ptx::smem_u32(smem);
      const uint64_t a_desc0 = ptx::make_sw128_kmajor_desc(base), b_desc0 = ptx::make_sw128_kmajor_desc(base + Cfg::A_BYTES);
      const uint64_t sfa_desc0 = ptx::make_sf_desc(base + Cfg::A_BYTES + Cfg::B_BYTES);
      const uint64_t sfb_desc0 = ptx::make_sf_desc(base + Cfg::A_BYTES + Cfg::B_BYTES + Cfg::SFA_BYTES);
      const uint32_t sf_t0 = tmem_base + Cfg::SF_COL0;
      uint32_t it = 0, tile_it = 0, ready = 0;
      PROF(long long cw = 0, cw_n = 0, ce = 0, ct0 = clock64();)
      for (int t = pair_id; t < num_ptiles; t += num_pairs, ++tile_it) {
        PROF(long long t0e = clock64();)
        ptx::mbar_wait(tempty, (tile_it & 1) ^ 1);
        PROF(ce += clock64() - t0e;)
        ptx::tc_fence_after();
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          PROF(long long t0f = clock64();)
          if (!ready) { ptx::mbar_wait(&full[s], ph); PROF(++cw_n;) }
          PROF(cw += clock64() - t0f;)
          ready = ptx::mbar_test(&full[(it + 1) % STAGES], ((it + 1) / STAGES) & 1);
          ptx::tc_fence_after();
          // descriptors of stage s: constant base + s * (stage bytes / 16) in the 14-bit address field
          const uint32_t soff = (uint32_t)s * (Cfg::STAGE_BYTES >> 4);
          const uint64_t a_desc = a_desc0 + soff, b_desc = b_desc0 + soff, sfa_desc = sfa_desc0 + soff, sfb_desc = sfb_desc0 + soff;
          const uint32_t sfa_t = sf_t0 + (it & 1) * Cfg::SF_COLS, sfb_t = sfa_t + 16;
          if (ptx::elect_one()) {
            if (kb + 1 < nkb || (kc_total & 3) == 0) {
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                ptx::tc_cp_32x128b_warpx4_pair(sfa_t + c * 4, sfa_desc + c * 32);
                ptx::tc_cp_32x128b_warpx4_pair(sfb_t + c * 8, sfb_desc + c * 32);
                ptx::tc_cp_32x128b_warpx4_pair(sfb_t + c * 8 + 4, sfb_desc + 128 + c * 32);
                ptx::tc_mma_nvfp4_pair(tmem_base, a_desc + 2 * c, b_desc + 2 * c, idesc, (kb | c) ? 1u : 0u, sfa_t + c * 4,
                                       sfb_t + c * 8);
              }
            } else {  // K tail: the last k-block holds 1..3 chunks of 64
              for (int c = 0; c < (kc_total & 3); ++c) {
                ptx::tc_cp_32x128b_warpx4_pair(sfa_t + c * 4, sfa_desc + c * 32);
                ptx::tc_cp_32x128b_warpx4_pair(sfb_t + c * 8, sfb_desc + c * 32);
                ptx::tc_cp_32x128b_warpx4_pair(sfb_t + c * 8 + 4, sfb_desc + 128 + c * 32);
                ptx::tc_mma_nvfp4_pair(tmem_base, a_desc + 2 * c, b_desc + 2 * c, idesc, (kb | c) ? 1u : 0u, sfa_t + c * 4,
                                       sfb_t + c * 8);
              }
            }
            ptx::tc_commit_pair(&empty[s], 3);   // both CTAs' stage s may be refilled
          }
          __syncwarp();
        }
        if (ptx::elect_one()) ptx::tc_commit_pair(tfull, 3);
        __syncwarp();
      }
      PROF(if (p.prof && blockIdx.x == 0 && lane == 0) { long long *o = p.prof + 4; o[0] = clock64() - ct0; o[1] = cw; o[2] = cw_n; o[3] = ce; o[4] = it; o[5] = tile_it; })
    }
  } else {
    // ===================== epilogue (warps 2..9 of both CTAs) =====================
    // plain: warp (q, half) owns lanes [32 q, +32) x columns [128 half, +128).  SWIGLU: gate columns [64 half, +64) and the
    // matching up columns [128 + 64 half, +64), so one thread holds g and u of the same (row, channel).
    constexpr int HC = kBN / 2, CH = kPC / 8, RB = kPC * 2;
    constexpr int OUTC = SWIGLU ? HC / 2 : HC;     // output columns per warp
    const int q = warp & 3, half = (warp - 2) >> 2;
    float alpha = p.alpha;
    if (p.alpha_dev) alpha = __ldg(p.alpha_dev);
    const float *cs = p.col_scale;
    uint8_t *wst = epi + (warp - 2) * Cfg::EPI_WARP_BYTES;
    const int ncols_out = SWIGLU ? p.N / 2 : p.N;
    const bool vec_ok = (ncols_out % 8) == 0 && (p.ldd % 8) == 0;
    const uint32_t tempty_leader = ptx::mapa(tempty, 0);
    uint32_t tile_it = 0;
    PROF(long long ew = 0, eld = 0, et0 = clock64();)
    for (int t = pair_id; t < num_ptiles; t += num_pairs, ++tile_it) {
      int pm, pn;
      tile_coords(p, t, pm, pn);
      // first tile column of this warp's register block (input columns of the GEMM, for the per-column scales)
      const int in_col0 = pn * kBN + (SWIGLU ? half * (HC / 2) : half * HC);
      const int out_col0 = SWIGLU ? pn * (kBN / 2) + half * (HC / 2) : in_col0;
      float sreg[HC / 32];   // plain: columns [0, 128) of the block; SWIGLU: [0, 64) gate then [0, 64) up
      if constexpr (SWIGLU) {
        float sg[HC / 64], su[HC / 64];
        load_col_scales(sg, cs, alpha, in_col0, p.N, lane);
        load_col_scales(su, cs, alpha, in_col0 + HC, p.N, lane);
#pragma unroll
        for (int r = 0; r < HC / 64; ++r) { sreg[r] = sg[r]; sreg[HC / 64 + r] = su[r]; }
      } else {
        load_col_scales(sreg, cs, alpha, in_col0, p.N, lane);
      }
      PROF(long long e0 = clock64();)
      ptx::mbar_wait(tfull, tile_it & 1);
      PROF(long long e1 = clock64(); ew += e1 - e0;)
      ptx::tc_fence_after();
      const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
      uint32_t r[HC / 32][32];
#pragma unroll
      for (int cb = 0; cb < HC / 32; ++cb) {
        const int col = SWIGLU ? (cb >> 1) * HC + half * (HC / 2) + (cb & 1) * 32 : half * HC + cb * 32;
        ptx::tmem_ld_32x32b_x32(t_lane + col, r[cb]);
      }
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(tempty_leader);
      PROF(eld += clock64() - e1;)
      const int row_base = (pm * 2 + (int)rank) * kRows + q * 32;
      __nv_bfloat16 *dbase = p.d;
      if (p.rs_rows) {   // this CTA's 128 rows belong to one owner (rs_rows is a multiple of 128): store into its slot over NVLink
        const int owner = min((pm * 2 + (int)rank) * kRows / p.rs_rows, ARCQ_MAX_PEERS - 1);
        dbase = p.peer_d[owner] + (int64_t)(p.rs_rank - owner) * p.rs_rows * p.ldd;
      }
#pragma unroll
      for (int pass = 0; pass < OUTC / kPC; ++pass) {
#pragma unroll
        for (int j = 0; j < CH; ++j) {
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int col = pass * kPC + j * 8 + 2 * e;   // column inside this warp's output block
            float v0, v1;
            if constexpr (SWIGLU) {
              float sg0 = alpha, sg1 = alpha, su0 = alpha, su1 = alpha;
              if (cs) {
                sg0 = __shfl_sync(0xffffffffu, sreg[col >> 5], col & 31);
                sg1 = __shfl_sync(0xffffffffu, sreg[col >> 5], (col & 31) + 1);
                su0 = __shfl_sync(0xffffffffu, sreg[HC / 64 + (col >> 5)], col & 31);
                su1 = __shfl_sync(0xffffffffu, sreg[HC / 64 + (col >> 5)], (col & 31) + 1);
              }
              v0 = swiglu_bf16(sg0 * __uint_as_float(r[col >> 5][col & 31]), su0 * __uint_as_float(r[2 + (col >> 5)][col & 31]), silu_lut);
              v1 = swiglu_bf16(sg1 * __uint_as_float(r[col >> 5][(col & 31) + 1]), su1 * __uint_as_float(r[2 + (col >> 5)][(col & 31) + 1]), silu_lut);
            } else {
              float s0 = alpha, s1 = alpha;
              if (cs) {
                s0 = __shfl_sync(0xffffffffu, sreg[col >> 5], col & 31);
                s1 = __shfl_sync(0xffffffffu, sreg[col >> 5], (col & 31) + 1);
              }
              v0 = s0 * __uint_as_float(r[col >> 5][col & 31]);
              v1 = s1 * __uint_as_float(r[col >> 5][(col & 31) + 1]);
            }
            __nv_bfloat162 h2 = __floats2bfloat162_rn(v0, v1);
            o[e] = *reinterpret_cast<uint32_t *>(&h2);
          }
          *reinterpret_cast<uint4 *>(wst + lane * RB + ((j ^ (lane & (CH - 1))) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
        }
        __syncwarp();
        const int j = lane & (CH - 1);
        const int gcol = out_col0 + pass * kPC + j * 8;
        uint4 bv = make_uint4(0u, 0u, 0u, 0u);
        if constexpr (!SWIGLU) {
          if (p.col_bias && gcol < ncols_out) bv = load_bias8(p.col_bias, gcol, ncols_out);
        }
#pragma unroll
        for (int i = 0; i < CH; ++i) {
          const int rr = i * (32 / CH) + lane / CH;
          uint4 v = *reinterpret_cast<const uint4 *>(wst + rr * RB + ((j ^ (rr & (CH - 1))) << 4));
          if constexpr (!SWIGLU) {
            if (p.col_bias) v = make_uint4(bias_add_bf16x2(v.x, bv.x), bias_add_bf16x2(v.y, bv.y), bias_add_bf16x2(v.z, bv.z), bias_add_bf16x2(v.w, bv.w));
          }
          const int grow = row_base + rr;
          if (grow < p.M && gcol < ncols_out) {
            __nv_bfloat16 *dst = dbase + (int64_t)grow * p.ldd + gcol;
            if (vec_ok && gcol + 8 <= ncols_out) {
              *reinterpret_cast<uint4 *>(dst) = v;
            } else {
              const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int e = 0; e < 8; ++e)
                if (gcol + e < ncols_out) dst[e] = __ushort_as_bfloat16((unsigned short)(w[e >> 1] >> (16 * (e & 1))));
            }
          }
        }
        __syncwarp();
      }
    }
    PROF(if (p.prof && blockIdx.x == 0 && warp == 2 && lane == 0) { long long *o = p.prof + 10; o[0] = clock64() - et0; o[1] = ew; o[2] = eld; })
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 1) ptx::tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
}

}  // namespace

struct ScatterArgs { void *const *peer_recv; int world, rank, rows_per_rank; };

template <bool SWIGLU>
static int launch_gemm_pair_t(const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K, float alpha,
                              const float *alpha_dev, const float *col_scale, void *d, int64_t ldd, cudaStream_t stream,
                              const ScatterArgs *sc = nullptr, const void *col_bias = nullptr) {
  constexpr int STAGES = SWIGLU ? 4 : 5;   // the silu table takes the fifth stage's place
  using Cfg = PCfg<STAGES, SWIGLU>;
  auto kern = gemm_nvfp4_pair_kernel<STAGES, SWIGLU>;
  static SmemAttrCache attr_cache = {};
  static int max_pairs[kMaxDevices] = {0};   // co-resident CTA pairs, per device
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(kThreadsP);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (int rc = ensure_dynamic_smem(kern, Cfg::SMEM_BYTES, attr_cache)) return rc;
  const int dev = current_device();
  if (dev < 0) return ARCQ_ECUDA;
  if (max_pairs[dev] == 0) {
    const int sms = sm_count();
    if (sms <= 0) return ARCQ_ECUDA;
    cfg.gridDim = dim3(2 * (sms / 2));
    int n = 0;
    ARCQ_CUDA_TRY(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
    if (n < 1) { set_error("no CTA pair fits on this device"); return ARCQ_ENOSUP; }
    max_pairs[dev] = n;
  }
  PParams p{(__nv_bfloat16 *)d, alpha_dev, col_scale, alpha, M, N, K, ldd, (M + 255) / 256, (N + 255) / 256, nullptr, {}, 0, 0, (const __nv_bfloat16 *)col_bias, 0};
  // tile order, measured inside the bench step (profiles/r2_gemm_raster_ab.log): groups of 16 row tiles are 1-3 % faster for
  // N = 4096 / 6144 (down 170.9 -> 165.6 us, o 64.5 -> 63.6, q+k+v 95.5 -> 94.7) and 2-9 % slower for N = 28672: narrow outputs only
  p.raster_gm = options().gemm_raster >= 0 ? options().gemm_raster : (p.ptiles_n <= 32 ? 16 : 0);
  if (sc) {
    for (int i = 0; i < ARCQ_MAX_PEERS; ++i) p.peer_d[i] = (__nv_bfloat16 *)sc->peer_recv[i < sc->world ? i : sc->world - 1];
    p.rs_rows = sc->rows_per_rank;
    p.rs_rank = sc->rank;
  }
#ifdef ARCQ_GEMM_PROFILE
  static long long *prof_buf = nullptr;
  const bool do_prof = getenv("ARCQ_GEMM_PROF") != nullptr;
  if (do_prof) { if (!prof_buf) cudaMalloc(&prof_buf, 64 * 8); cudaMemsetAsync(prof_buf, 0, 64 * 8, stream); p.prof = prof_buf; }
#endif
  const int64_t kc = K / 64;
  CUtensorMap ta, tb, tsa, tsb;
  if (int rc = cached_operand_map(&ta, a, M, K / 2, kRows)) return rc;
  if (int rc = cached_operand_map(&tb, b, N, K / 2, kRows)) return rc;
  if (int rc = cached_sf_map(&tsa, sfa, (int64_t)((M + 127) / 128) * kc)) return rc;
  if (int rc = cached_sf_map(&tsb, sfb, (int64_t)((N + 127) / 128) * kc)) return rc;
  int pairs = p.ptiles_m * p.ptiles_n;
  if (pairs > max_pairs[dev]) pairs = max_pairs[dev];
  cfg.gridDim = dim3(2 * pairs);
  ARCQ_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, ta, tb, tsa, tsb, p));
  count_launch();
#ifdef ARCQ_GEMM_PROFILE
  if (do_prof) {
    long long h[64];
    cudaStreamSynchronize(stream);
    cudaMemcpy(h, prof_buf, sizeof(h), cudaMemcpyDeviceToHost);
    fprintf(stderr, "[prof] producer0: total %lld clk, wait-empty %lld clk in %lld blocking waits of %lld iters | producer1: total %lld wait %lld (%lld)\n",
            h[0], h[1], h[2], h[3], h[16], h[17], h[18]);
    fprintf(stderr, "[prof] mma: total %lld clk, wait-full %lld clk in %lld blocking waits of %lld iters, wait-tempty %lld clk over %lld tiles\n",
            h[4], h[5], h[6], h[8], h[7], h[9]);
    fprintf(stderr, "[prof] epilogue warp2: total %lld clk, wait-tfull %lld clk, tmem-ld+release %lld clk\n", h[10], h[11], h[12]);
  }
#endif
  return ARCQ_OK;
}

int launch_gemm_pair(const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K, float alpha,
                     const float *alpha_dev, const float *col_scale, int swiglu, void *d, int64_t ldd, cudaStream_t stream,
                     const void *col_bias) {
  if (swiglu) return launch_gemm_pair_t<true>(a, b, sfa, sfb, M, N, K, alpha, alpha_dev, col_scale, d, ldd, stream);
  return launch_gemm_pair_t<false>(a, b, sfa, sfb, M, N, K, alpha, alpha_dev, col_scale, d, ldd, stream, nullptr, col_bias);
}

// row-parallel GEMM whose epilogue IS the first half of the reduce-scatter (always the cta_group::2 kernel: any shape works)
int launch_gemm_pair_scatter(const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K, float alpha,
                             const float *alpha_dev, const float *col_scale, void *const *peer_recv, int world, int rank,
                             int rows_per_rank, int64_t ldd, cudaStream_t stream) {
  const ScatterArgs sc{peer_recv, world, rank, rows_per_rank};
  return launch_gemm_pair_t<false>(a, b, sfa, sfb, M, N, K, alpha, alpha_dev, col_scale, peer_recv[rank], ldd, stream, &sc);
}

}  // namespace arcq
