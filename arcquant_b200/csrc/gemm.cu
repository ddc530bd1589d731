// gemm.cu -- NVFP4 block-scaled GEMM on tcgen05 (sm_100a), consuming the augmented K of quant.cu.
//
//   D[M,N] bf16 = alpha * sum_k (e2m1(A[m,k]) * e4m3(SFA[m,k/16])) * (e2m1(B[n,k]) * e4m3(SFB[n,k/16]))
//
// Replaces the reference's CUTLASS Sm120 mma.sync instantiation (/root/reference/kernels/src/nvfp4.cu:35-132).
// Hand-written: TMA (cp.async.bulk.tensor, 128-B swizzle) feeds a multi-stage shared-memory ring; scale
// factors arrive by 1-D bulk copies straight from the swizzled layout the quantise kernel wrote (one
// 512-B atom per 128 rows x 64 k) and go to TMEM with tcgen05.cp -- no relayout pass anywhere; one
// elected thread issues tcgen05.mma kind::mxf4nvf4.block_scale (M=128, N=BN, K=64) into TMEM
// accumulators.  Persistent CTAs (one per SM), warp-specialised: warp 0 = TMA producer, warp 1 = MMA
// issuer (+ TMEM allocator), warps 2..9 = epilogue.
//
// Measured design points (profiles/r1_gemm_notes.md):
//  * tcgen05.cp writes TMEM at ~64 B/clk and the x4 lane-quadrant replication makes the scale factors of a
//    128x256x256 k-block 24 KB of TMEM writes (~390 clk next to 512 clk of MMA) -> the scale-factor TMEM
//    region is double-buffered so the copies of k-block i+1 are not ordered behind the MMAs of k-block i;
//  * the epilogue pulls the whole accumulator into registers, releases TMEM at once (the next tile's MMAs
//    start while bf16 conversion and stores run) and stores through an XOR-swizzled shared staging buffer so
//    every global store instruction writes whole 128-byte lines.
#include <cuda.h>

#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "ptx.cuh"

namespace arcq {

constexpr int kBM = 128;         // tile rows (UMMA M)
constexpr int kBKBytes = 128;    // bytes of packed e2m1 per row per k-block = 256 elements = one swizzle span
constexpr int kBK = 256;
constexpr int kGemmThreads = 320;
constexpr int kEpiWarps = 8;

template <int BN, int STAGES, int PC = 64>
struct GemmCfg {
  static constexpr int NB = BN / 128;                 // 128-row scale blocks per B tile
  static constexpr int A_BYTES = kBM * kBKBytes;      // 16 KiB
  static constexpr int B_BYTES = BN * kBKBytes;
  static constexpr int SFA_BYTES = 4 * 512;           // 4 k-chunks of one 512-B atom
  static constexpr int SFB_BYTES = NB * 4 * 512;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + SFA_BYTES + SFB_BYTES;
  static constexpr int ACC_STAGES = (BN <= 128) ? 2 : 1;
  static constexpr int SF_COLS = 16 + 16 * NB;        // one scale-factor buffer: SFA (16 cols) then SFB
  static constexpr int SF_COL0 = ACC_STAGES * BN;     // TMEM columns: accumulators, then 2 scale-factor buffers
  static constexpr int TMEM_COLS = 512;
  static constexpr int EPI_WARP_BYTES = 32 * PC * 2;   // per-warp staging: 32 rows x PC bf16 columns per pass
  static constexpr int EPI_BYTES = kEpiWarps * EPI_WARP_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
  static_assert(SF_COL0 + 2 * SF_COLS <= TMEM_COLS, "TMEM overflow");
  static_assert(SMEM_BYTES <= 232448, "shared memory overflow");
};

struct GemmParams {
  const uint8_t *sfa;
  const uint8_t *sfb;
  __nv_bfloat16 *d;
  const float *alpha_dev;
  const float *col_scale;  // nullable fp32 [N]: column n is scaled by RN(alpha * col_scale[n])
  float alpha;
  int M, N, K;
  int64_t ldd;             // row stride of d in elements
  int tiles_m, tiles_n;  const __nv_bfloat16 *col_bias;  // nullable bf16 [N]: added to the rounded output like torch's `y + bias` (qLinearLayer.py:74-75)
};

template <int BN, int STAGES, int PC>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_nvfp4_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const GemmParams p) {
  using Cfg = GemmCfg<BN, STAGES, PC>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *epi = smem + STAGES * Cfg::STAGE_BYTES;
  uint64_t *bars = reinterpret_cast<uint64_t *>(epi + Cfg::EPI_BYTES);
  uint64_t *full = bars;                         // [STAGES]  TMA -> MMA
  uint64_t *empty = bars + STAGES;               // [STAGES]  MMA -> TMA
  uint64_t *tfull = bars + 2 * STAGES;           // [ACC]     MMA -> epilogue
  uint64_t *tempty = tfull + Cfg::ACC_STAGES;    // [ACC]     epilogue -> MMA
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + Cfg::ACC_STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int K = p.K;
  const int nkb = (K + kBK - 1) / kBK;
  const int kc_total = K / 64;                   // 64-element chunks (scale atoms) along K
  const int num_tiles = p.tiles_m * p.tiles_n;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_a);
    ptx::prefetch_tmap(&tmap_b);
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
    for (int a = 0; a < Cfg::ACC_STAGES; ++a) { ptx::mbar_init(&tfull[a], 1); ptx::mbar_init(&tempty[a], kEpiWarps); }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // The whole warp runs the loop and the barrier waits; only the elect.sync leader issues.  (Issuing from
    // inside an `if (lane == 0)` branch makes ptxas wrap every uniform-datapath instruction -- UTMALDG,
    // UTCCP, UTCOMMA, UTCBAR -- in an ELECT/R2UR.BROADCAST waterfall loop: ~45 ns per instruction measured.)
    {
      uint32_t it = 0;
      uint32_t ready = ptx::mbar_test(&empty[0], 1);
      const int nblk_total = (p.N + 127) / 128;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int mt = tile % p.tiles_m, nt = tile / p.tiles_m;
        const int m0 = mt * kBM, n0 = nt * BN;
        int nsfb = nblk_total - nt * Cfg::NB;
        if (nsfb > Cfg::NB) nsfb = Cfg::NB;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          if (!ready) ptx::mbar_wait(&empty[s], ph ^ 1);
          ready = ptx::mbar_test(&empty[(it + 1) % STAGES], (((it + 1) / STAGES) & 1) ^ 1);  // peek: consumed next iteration
          uint8_t *st = smem + s * Cfg::STAGE_BYTES;
          int kchunks = kc_total - kb * 4;
          if (kchunks > 4) kchunks = 4;
          const uint32_t sf_bytes = (uint32_t)kchunks * 512u;
          if (ptx::elect_one()) {
            ptx::mbar_arrive_expect_tx(&full[s], Cfg::A_BYTES + Cfg::B_BYTES + sf_bytes * (1 + nsfb));
            ptx::tma_load_2d(st, &tmap_a, kb * kBKBytes, m0, &full[s]);
            ptx::tma_load_2d(st + Cfg::A_BYTES, &tmap_b, kb * kBKBytes, n0, &full[s]);
            ptx::bulk_load(st + Cfg::A_BYTES + Cfg::B_BYTES, p.sfa + ((size_t)mt * kc_total + kb * 4) * 512, sf_bytes,
                           &full[s]);
            for (int j = 0; j < nsfb; ++j)
              ptx::bulk_load(st + Cfg::A_BYTES + Cfg::B_BYTES + Cfg::SFA_BYTES + j * 2048,
                             p.sfb + ((size_t)(nt * Cfg::NB + j) * kc_total + kb * 4) * 512, sf_bytes, &full[s]);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    {
      constexpr uint32_t idesc = ptx::make_idesc_nvfp4(kBM, BN);
      uint32_t it = 0, tile_it = 0;
      uint32_t ready = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tile_it) {
        const int acc = tile_it % Cfg::ACC_STAGES;
        const uint32_t acc_ph = (tile_it / Cfg::ACC_STAGES) & 1;
        ptx::mbar_wait(&tempty[acc], acc_ph ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          if (!ready) ptx::mbar_wait(&full[s], ph);
          ready = ptx::mbar_test(&full[(it + 1) % STAGES], ((it + 1) / STAGES) & 1);  // peek: consumed next iteration
          ptx::tc_fence_after();
          uint8_t *st = smem + s * Cfg::STAGE_BYTES;
          int kchunks = kc_total - kb * 4;
          if (kchunks > 4) kchunks = 4;
          const uint32_t a_addr = ptx::smem_u32(st), b_addr = a_addr + Cfg::A_BYTES;
          const uint32_t sfa_addr = b_addr + Cfg::B_BYTES, sfb_addr = sfa_addr + Cfg::SFA_BYTES;
          const uint32_t sfa_t = tmem_base + Cfg::SF_COL0 + (it & 1) * Cfg::SF_COLS, sfb_t = sfa_t + 16;
          const uint64_t a_desc = ptx::make_sw128_kmajor_desc(a_addr);
          const uint64_t b_desc = ptx::make_sw128_kmajor_desc(b_addr);
          if (ptx::elect_one()) {
          for (int c = 0; c < kchunks; ++c) {
            ptx::tc_cp_32x128b_warpx4(sfa_t + c * 4, ptx::make_sf_desc(sfa_addr + c * 512));
#pragma unroll
            for (int j = 0; j < Cfg::NB; ++j)
              ptx::tc_cp_32x128b_warpx4(sfb_t + c * 4 * Cfg::NB + j * 4, ptx::make_sf_desc(sfb_addr + j * 2048 + c * 512));
            // advance 32 bytes (64 e2m1) inside the 128-B swizzle span: +2 in the 16-B address field
            ptx::tc_mma_nvfp4(d_tmem, a_desc + 2 * c, b_desc + 2 * c, idesc, (kb > 0 || c > 0) ? 1u : 0u,
                              sfa_t + c * 4, sfb_t + c * 4 * Cfg::NB);
          }
          ptx::tc_commit(&empty[s]);  // frees the smem stage once these MMAs have read it
          }
          __syncwarp();
        }
        if (ptx::elect_one()) ptx::tc_commit(&tfull[acc]);  // accumulator complete -> epilogue
        __syncwarp();
      }
    }
  } else {
    // ===================== epilogue (warps 2..9): lane quadrant = warp % 4, column half = (warp - 2) / 4 ======
    constexpr int HC = BN / 2;   // columns per warp
    const int q = warp & 3, half = (warp - 2) >> 2;
    float alpha = p.alpha;
    if (p.alpha_dev) alpha = __ldg(p.alpha_dev);
    uint8_t *wst = epi + (warp - 2) * Cfg::EPI_WARP_BYTES;
    constexpr int CH = PC / 8;        // 16-byte chunks per staged row
    constexpr int RB = PC * 2;        // staged row bytes
    const bool vec_ok = (p.N % 8) == 0 && (p.ldd % 8) == 0;
    const float *cs = p.col_scale;
    uint32_t tile_it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tile_it) {
      const int mt = tile % p.tiles_m, nt = tile / p.tiles_m;
      const int acc = tile_it % Cfg::ACC_STAGES;
      const uint32_t acc_ph = (tile_it / Cfg::ACC_STAGES) & 1;
      // per-column scales of this warp's HC columns: lane l keeps columns l, l + 32, ... (already times alpha), loaded
      // before the accumulator wait; a shuffle fetches the one a store needs
      float sreg[HC / 32];
#pragma unroll
      for (int r = 0; r < HC / 32; ++r) {
        const int c = nt * BN + half * HC + r * 32 + lane;
        sreg[r] = (cs && c < p.N) ? __fmul_rn(alpha, __ldg(cs + c)) : alpha;
      }
      ptx::mbar_wait(&tfull[acc], acc_ph);
      ptx::tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + half * HC;
      uint32_t r[HC / 32][32];
#pragma unroll
      for (int cb = 0; cb < HC / 32; ++cb) ptx::tmem_ld_32x32b_x32(t_row + cb * 32, r[cb]);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty[acc]);  // accumulator is in registers: the next MMAs may start
#pragma unroll
      for (int pass = 0; pass < HC / PC; ++pass) {
        // lane = row: PC columns -> CH chunks of 8 bf16, chunk j at swizzled slot j ^ (row & (CH-1))
#pragma unroll
        for (int j = 0; j < CH; ++j) {
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int col = pass * PC + j * 8 + 2 * e;
            float s0 = alpha, s1 = alpha;
            if (cs) {
              s0 = __shfl_sync(0xffffffffu, sreg[col >> 5], col & 31);
              s1 = __shfl_sync(0xffffffffu, sreg[col >> 5], (col & 31) + 1);
            }
            __nv_bfloat162 h2 = __floats2bfloat162_rn(s0 * __uint_as_float(r[col >> 5][col & 31]),
                                                      s1 * __uint_as_float(r[col >> 5][(col & 31) + 1]));
            o[e] = *reinterpret_cast<uint32_t *>(&h2);
          }
          *reinterpret_cast<uint4 *>(wst + lane * RB + ((j ^ (lane & (CH - 1))) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
        }
        __syncwarp();
        const int j = lane & (CH - 1);
        const int gcol = nt * BN + half * HC + pass * PC + j * 8;
        uint4 bv = make_uint4(0u, 0u, 0u, 0u);
        if (p.col_bias && gcol < p.N) bv = load_bias8(p.col_bias, gcol, p.N);
#pragma unroll
        for (int i = 0; i < CH; ++i) {  // 32 / CH rows per instruction, CH instructions
          const int rr = i * (32 / CH) + lane / CH;
          uint4 v = *reinterpret_cast<const uint4 *>(wst + rr * RB + ((j ^ (rr & (CH - 1))) << 4));
          if (p.col_bias) v = make_uint4(bias_add_bf16x2(v.x, bv.x), bias_add_bf16x2(v.y, bv.y), bias_add_bf16x2(v.z, bv.z), bias_add_bf16x2(v.w, bv.w));
          const int grow = mt * kBM + q * 32 + rr;
          if (grow < p.M && gcol < p.N) {
            __nv_bfloat16 *dst = p.d + (size_t)grow * p.ldd + gcol;
            if (vec_ok && gcol + 8 <= p.N) {
              *reinterpret_cast<uint4 *>(dst) = v;
            } else {
              const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int e = 0; e < 8; ++e)
                if (gcol + e < p.N) dst[e] = __ushort_as_bfloat16((unsigned short)(w[e >> 1] >> (16 * (e & 1))));
            }
          }
        }
        __syncwarp();
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ---- host side ------------------------------------------------------------------------------

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)sym;
  });
  return fn;
}

// 2-D map over a row-major u8 matrix [rows, row_bytes]; box = 128 bytes x box_rows, 128-B swizzle,
// out-of-bounds reads (row tails, K tail) return zeros.
static int make_operand_map(CUtensorMap *map, const void *ptr, int64_t rows, int64_t row_bytes, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return ARCQ_ECUDA; }
  cuuint64_t dims[2] = {(cuuint64_t)row_bytes, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)row_bytes};
  cuuint32_t box[2] = {(cuuint32_t)kBKBytes, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return ARCQ_ECUDA; }
  return ARCQ_OK;
}

struct MapKey {
  const void *ptr; int64_t rows, row_bytes; int box_rows;
  bool operator==(const MapKey &o) const { return ptr == o.ptr && rows == o.rows && row_bytes == o.row_bytes && box_rows == o.box_rows; }
};
struct MapKeyHash {
  size_t operator()(const MapKey &k) const {
    return std::hash<const void *>()(k.ptr) ^ (std::hash<int64_t>()(k.rows) * 31) ^ (std::hash<int64_t>()(k.row_bytes) * 131) ^ k.box_rows;
  }
};

// Encoding a tensor map costs a few microseconds of host time; weights (and CUDA-graph-stable
// activation buffers) come back with the same (pointer, shape), so keep them.
int cached_operand_map(CUtensorMap *out, const void *ptr, int64_t rows, int64_t row_bytes, int box_rows) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  MapKey key{ptr, rows, row_bytes, box_rows};
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return ARCQ_OK; }
  if (int rc = make_operand_map(out, ptr, rows, row_bytes, box_rows)) return rc;
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return ARCQ_OK;
}

// scale-factor buffer as a 2-D map of 512-byte atoms (row = one 128-row x 64-k atom): box = 4 atoms = the
// 2 KB a k-block needs; out-of-range atoms (K tail, row-block tail) read as zeros
int cached_sf_map(CUtensorMap *out, const void *ptr, int64_t natoms) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  MapKey key{ptr, natoms, 512, -4};
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return ARCQ_OK; }
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return ARCQ_ECUDA; }
  cuuint64_t dims[2] = {128, (cuuint64_t)natoms};
  cuuint64_t strides[1] = {512};
  cuuint32_t box[2] = {128, 4};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (scale factors) failed with CUresult %d", (int)r); return ARCQ_ECUDA; }
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return ARCQ_OK;
}

template <int BN, int STAGES, int PC>
static int launch_gemm(const void *a, const void *b, GemmParams p, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, STAGES, PC>;
  auto kern = gemm_nvfp4_kernel<BN, STAGES, PC>;
  static SmemAttrCache attr_cache = {};  // per instantiation, per device
  if (int rc = ensure_dynamic_smem(kern, Cfg::SMEM_BYTES, attr_cache)) return rc;
  CUtensorMap ta, tb;
  if (int rc = cached_operand_map(&ta, a, p.M, p.K / 2, kBM)) return rc;
  if (int rc = cached_operand_map(&tb, b, p.N, p.K / 2, BN)) return rc;
  p.tiles_m = (p.M + kBM - 1) / kBM;
  p.tiles_n = (p.N + BN - 1) / BN;
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  int grid = p.tiles_m * p.tiles_n;
  if (grid > sms) grid = sms;
  kern<<<grid, kGemmThreads, Cfg::SMEM_BYTES, stream>>>(ta, tb, p);
  ARCQ_CUDA_TRY(cudaGetLastError());
  count_launch();
  return ARCQ_OK;
}

}  // namespace arcq

using namespace arcq;

namespace arcq {
int launch_gemm_pair(const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K, float alpha,
                     const float *alpha_dev, const float *col_scale, int swiglu, void *d, int64_t ldd, cudaStream_t stream,
                     const void *col_bias = nullptr);
int launch_gemm_splitk(const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K, float alpha,
                       const float *alpha_dev, const float *col_scale, void *d, int64_t ldd, int splits, cudaStream_t stream,
                       const void *col_bias = nullptr);
int splitk_choice(int M, int N, int K);  // 0 = do not split
}

extern "C" int arcq_gemm_nvfp4_ex2(const void *a, const void *b, const void *sfa, const void *sfb, int64_t M, int64_t N,
                                   int64_t K, float alpha, const float *alpha_dev, const float *col_scale_dev,
                                   const void *col_bias_dev, int epilogue, void *d, int64_t ldd, void *stream) {
  if (!a || !b || !sfa || !sfb || !d) { set_error("null pointer argument"); return ARCQ_EINVAL; }
  if (M < 0 || N < 0 || M > 0x7fffff00 || N > 0x7fffff00) { set_error("M/N out of range"); return ARCQ_EINVAL; }
  if (K <= 0 || K % 64 || K > (1 << 20)) { set_error("K=%lld must be a positive multiple of 64", (long long)K); return ARCQ_EINVAL; }
  if (((uintptr_t)a | (uintptr_t)b | (uintptr_t)sfa | (uintptr_t)sfb | (uintptr_t)d) & 15) {
    set_error("pointers must be 16-byte aligned");
    return ARCQ_EINVAL;
  }
  if (epilogue != ARCQ_EPILOGUE_NONE && epilogue != ARCQ_EPILOGUE_SWIGLU) { set_error("unknown epilogue %d", epilogue); return ARCQ_EINVAL; }
  const bool swiglu = epilogue == ARCQ_EPILOGUE_SWIGLU;
  if (col_bias_dev && (swiglu || ((uintptr_t)col_bias_dev & 15))) {
    set_error("col_bias_dev must be 16-byte aligned and cannot be combined with the SwiGLU epilogue");
    return ARCQ_EINVAL;
  }
  if (swiglu && N % 256) { set_error("the SwiGLU epilogue needs N %% 256 == 0 (128-row gate/up interleave), got N=%lld", (long long)N); return ARCQ_EINVAL; }
  const int64_t out_cols = swiglu ? N / 2 : N;
  if (ldd == 0) ldd = out_cols;
  if (ldd < out_cols) { set_error("ldd=%lld smaller than the %lld output columns", (long long)ldd, (long long)out_cols); return ARCQ_EINVAL; }
  if (M == 0 || N == 0) return ARCQ_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const Options &o = options();
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  if (swiglu) return launch_gemm_pair(a, b, sfa, sfb, (int)M, (int)N, (int)K, alpha, alpha_dev, col_scale_dev, 1, d, ldd, st);
  // small M (<= 1024 rows, few output tiles): split K over a thread-block cluster so every SM pulls on the weight stream (gemm_splitk.cu)
  const int splits = o.gemm_splitk >= 0 ? o.gemm_splitk : splitk_choice((int)M, (int)N, (int)K);
  if (splits >= 2 && (M <= 1024 || o.gemm_splitk >= 2))
    return launch_gemm_splitk(a, b, sfa, sfb, (int)M, (int)N, (int)K, alpha, alpha_dev, col_scale_dev, d, ldd, splits, st, col_bias_dev);
  // Kernel choice for what split-K did not take, from the measured A/B on the config-2 shapes (profiles/r2_midm_kernel_choice_ab.log):
  //  * at most one wave of 128 x 128 tiles (<= SM count): the single-CTA kernel with 128-wide tiles puts the most SMs on the
  //    problem -- q/o at M <= 512: 10.0-10.6 us against 13.1-13.7 us with 256-wide tiles, k/v up to M = 2048: 9.8-10.6 vs 13.2-13.5,
  //    fused q/k/v at M <= 384: 10.9-11.3 vs 14.1-14.7, down_proj at M = 384 / 512: 24.9 / 25.5 vs 32.0;
  //  * beyond that the cta_group::2 kernel from 40 pair tiles on (down_proj at M = 768 / 1024: 27.2 / 28.3 us against 33.0 / 33.4
  //    for the 256-wide single-CTA kernel; a tie elsewhere), the single-CTA kernel below that.
  const int64_t tiles128 = ((M + 127) / 128) * ((N + 127) / 128);
  const int64_t ptiles = ((M + 255) / 256) * ((N + 255) / 256);
  const bool one_wave_128 = o.gemm_pair < 0 && o.gemm_bn == 0 && tiles128 <= sms;
  const bool use_pair = o.gemm_pair >= 0 ? o.gemm_pair != 0 : (!one_wave_128 && ptiles >= 40);
  if (use_pair) return launch_gemm_pair(a, b, sfa, sfb, (int)M, (int)N, (int)K, alpha, alpha_dev, col_scale_dev, 0, d, ldd, st, col_bias_dev);
  GemmParams p{(const uint8_t *)sfa, (const uint8_t *)sfb, (__nv_bfloat16 *)d, alpha_dev, col_scale_dev, alpha,
               (int)M, (int)N, (int)K, ldd, 0, 0, (const __nv_bfloat16 *)col_bias_dev};
  const int64_t tiles256 = ((M + kBM - 1) / kBM) * ((N + 255) / 256);
  int bn = o.gemm_bn;
  if (bn != 128 && bn != 256) {
    // wide tiles halve A re-reads and MMA issue count; narrow tiles fill the SMs for small problems
    bn = one_wave_128 ? 128 : ((tiles256 >= 2 * (int64_t)sms || N % 256 == 0) && N > 128 ? 256 : 128);
  }
  if (bn == 256) return launch_gemm<256, 4, 16>(a, b, p, st);
  return launch_gemm<128, 5, 64>(a, b, p, st);
}

extern "C" int arcq_gemm_nvfp4_ex(const void *a, const void *b, const void *sfa, const void *sfb, int64_t M, int64_t N,
                                  int64_t K, float alpha, const float *alpha_dev, const float *col_scale_dev, int epilogue,
                                  void *d, int64_t ldd, void *stream) {
  return arcq_gemm_nvfp4_ex2(a, b, sfa, sfb, M, N, K, alpha, alpha_dev, col_scale_dev, nullptr, epilogue, d, ldd, stream);
}

extern "C" int arcq_gemm_nvfp4(const void *a, const void *b, const void *sfa, const void *sfb, int64_t M, int64_t N,
                               int64_t K, float alpha, const float *alpha_dev, void *d, void *stream) {
  return arcq_gemm_nvfp4_ex(a, b, sfa, sfb, M, N, K, alpha, alpha_dev, nullptr, ARCQ_EPILOGUE_NONE, d, N, stream);
}
