// gemm_cluster.cu -- the large-shape NVFP4 GEMM: same tcgen05 tile engine as gemm.cu (128 x 256 x 256 per
// CTA, UTCCP scale factors, TMEM accumulator) but launched as thread-block CLUSTERS of CM x CN CTAs that
// share operands through TMA multicast.
//
// Why: e2m1 operands are so dense that one 128x256 CTA tile needs 54 KB from L2 every 512 tensor-pipe
// cycles; measured on B200 the L2->SM fabric sustains ~37 B/clk/SM (10.5 TB/s), which capped the
// non-multicast kernel at ~2.4 PFLOP/s (profiles/r1_bench_v1.json).  In a CM x CN cluster the A tile
// (and SFA) of a CTA row is fetched once and multicast to its CN column peers, the B tile (and SFB) once
// per CM row peers: L2 traffic per CTA per k-block drops from 54 KB to 54/(...) e.g. 18 KB for 4 x 2.
//
// Pipeline protocol (all barriers are per CTA, same smem offset in every CTA):
//   full[s]   count 1: the CTA's own producer posts expect_tx(total bytes landing in ITS stage s); the
//             bytes arrive from its own and its peers' multicast TMA.
//   empty[s]  count CM*CN: every CTA's MMA warp multicast-commits to empty[s] of ALL cluster CTAs when its
//             MMAs on stage s have retired, so a producer only overwrites stage s (in all destination CTAs)
//             once the whole cluster has consumed it.
//   tfull / tempty: MMA <-> epilogue hand-off of the single 256-column accumulator; the eight epilogue
//             warps pull the whole accumulator into registers first and release TMEM BEFORE converting and
//             storing, so the next tile's MMAs overlap the bf16 conversion and the global stores.
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"

namespace arcq {

int cached_operand_map(CUtensorMap *out, const void *ptr, int64_t rows, int64_t row_bytes, int box_rows);

namespace {

constexpr int kBM = 128, kBN = 256, kBKBytes = 128, kBK = 256;
constexpr int kThreadsC = 320;  // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue

template <int CM, int CN, int STAGES>
struct CCfg {
  static constexpr int CSZ = CM * CN;
  static constexpr int A_BYTES = kBM * kBKBytes;   // 16 KiB
  static constexpr int B_BYTES = kBN * kBKBytes;   // 32 KiB
  static constexpr int SFA_BYTES = 4 * 512;
  static constexpr int SFB_BYTES = 2 * 4 * 512;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + SFA_BYTES + SFB_BYTES;
  static constexpr int A_SLICE_ROWS = kBM / CN;    // rows of the A tile this CTA fetches for its row peers
  static constexpr int B_SLICE_ROWS = kBN / CM;
  static constexpr int SFA_COL = kBN, SFB_COL = kBN + 16, TMEM_COLS = 512;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
  static_assert(A_SLICE_ROWS % 8 == 0 && B_SLICE_ROWS % 8 == 0, "slices must be whole swizzle atoms");
  static_assert(4 % CN == 0 && 8 % CM == 0, "scale chunks must split evenly");
};

struct CParams {
  const uint8_t *sfa;
  const uint8_t *sfb;
  __nv_bfloat16 *d;
  const float *alpha_dev;
  float alpha;
  int M, N, K;
  int tiles_m, tiles_n;      // CTA tiles
  int stiles_m, stiles_n;    // cluster super-tiles
};

template <int CM, int CN, int STAGES>
__global__ void __launch_bounds__(kThreadsC, 1)
gemm_nvfp4_cluster_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                          const CParams p) {
  using Cfg = CCfg<CM, CN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  // the dynamic smem window starts at the same CTA-relative address in every CTA of the launch, so the
  // aligned offsets below are identical across the cluster (multicast needs that)
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t *full = bars, *empty = bars + STAGES, *tfull = bars + 2 * STAGES, *tempty = tfull + 1;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t crank = ptx::cluster_ctarank();
  const int cm = crank % CM, cn = crank / CM;
  const int cluster_id = blockIdx.x / Cfg::CSZ, num_clusters = gridDim.x / Cfg::CSZ;
  const int K = p.K, nkb = (K + kBK - 1) / kBK, kc_total = K / 64;
  const int num_stiles = p.stiles_m * p.stiles_n;
  const int mblk_total = (p.M + 127) / 128, nblk_total = (p.N + 127) / 128;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_a);
    ptx::prefetch_tmap(&tmap_b);
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], Cfg::CSZ); }
    ptx::mbar_init(tfull, 1);
    ptx::mbar_init(tempty, 8);
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();  // every CTA's barriers are initialised before any peer multicasts into it
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // multicast masks: bit r = cluster rank r = cm + CM * cn
  uint16_t row_mask = 0, col_mask = 0;  // row peers share A (same cm), column peers share B (same cn)
  for (int j = 0; j < CN; ++j) row_mask |= (uint16_t)(1u << (cm + CM * j));
  for (int i = 0; i < CM; ++i) col_mask |= (uint16_t)(1u << (i + CM * cn));
  constexpr uint16_t all_mask = (uint16_t)((1u << Cfg::CSZ) - 1);

  if (warp == 0) {
    // ===================== TMA producer (whole warp loops; elect.sync leader issues) =====================
    {
      uint32_t it = 0;
      for (int st = cluster_id; st < num_stiles; st += num_clusters) {
        const int mt = (st % p.stiles_m) * CM + cm, nt = (st / p.stiles_m) * CN + cn;
        const bool a_ok = mt < mblk_total;                 // SFA block exists
        int nsfb = nblk_total - nt * 2;                    // SFB blocks of this tile that exist (0..2)
        nsfb = nsfb < 0 ? 0 : (nsfb > 2 ? 2 : nsfb);
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          ptx::mbar_wait(&empty[s], ph ^ 1);
          uint8_t *stg = smem + s * Cfg::STAGE_BYTES;
          int kchunks = kc_total - kb * 4;
          if (kchunks > 4) kchunks = 4;
          if (ptx::elect_one()) {
          ptx::mbar_arrive_expect_tx(&full[s], Cfg::A_BYTES + Cfg::B_BYTES + (a_ok ? kchunks * 512 : 0) + nsfb * kchunks * 512);
          // my slice of the A tile -> all row peers; my slice of the B tile -> all column peers
          ptx::tma_load_2d_mc(stg + cn * Cfg::A_SLICE_ROWS * kBKBytes, &tmap_a, kb * kBKBytes,
                              mt * kBM + cn * Cfg::A_SLICE_ROWS, &full[s], row_mask);
          ptx::tma_load_2d_mc(stg + Cfg::A_BYTES + cm * Cfg::B_SLICE_ROWS * kBKBytes, &tmap_b, kb * kBKBytes,
                              nt * kBN + cm * Cfg::B_SLICE_ROWS, &full[s], col_mask);
          // scale factors: SFA = 4 chunks of 512 B split over the CN row peers, SFB = 2 blocks x 4 chunks over CM
          if (a_ok) {
            constexpr int per = 4 / CN;
            int c0 = cn * per, c1 = c0 + per;
            if (c1 > kchunks) c1 = kchunks;
            if (c1 > c0)
              ptx::bulk_load_mc(stg + Cfg::A_BYTES + Cfg::B_BYTES + c0 * 512,
                                p.sfa + ((size_t)mt * kc_total + kb * 4 + c0) * 512, (c1 - c0) * 512, &full[s], row_mask);
          }
          {
            constexpr int per = 8 / CM;  // chunk slots (block j, chunk c) = j * 4 + c
            for (int q0 = cm * per; q0 < (cm + 1) * per;) {
              const int j = q0 >> 2, c0 = q0 & 3;
              int c1 = c0 + ((cm + 1) * per - q0);
              if (c1 > 4) c1 = 4;
              const int cnt = c1 - c0;
              int c1v = c1 > kchunks ? kchunks : c1;
              if (j < nsfb && c1v > c0)
                ptx::bulk_load_mc(stg + Cfg::A_BYTES + Cfg::B_BYTES + Cfg::SFA_BYTES + j * 2048 + c0 * 512,
                                  p.sfb + ((size_t)(nt * 2 + j) * kc_total + kb * 4 + c0) * 512, (c1v - c0) * 512,
                                  &full[s], col_mask);
              q0 += cnt;
            }
          }
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    {
      constexpr uint32_t idesc = ptx::make_idesc_nvfp4(kBM, kBN);
      uint32_t it = 0, tile_it = 0;
      for (int st = cluster_id; st < num_stiles; st += num_clusters, ++tile_it) {
        ptx::mbar_wait(tempty, (tile_it & 1) ^ 1);
        ptx::tc_fence_after();
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          ptx::mbar_wait(&full[s], ph);
          ptx::tc_fence_after();
          int kchunks = kc_total - kb * 4;
          if (kchunks > 4) kchunks = 4;
          const uint32_t a_addr = ptx::smem_u32(smem + s * Cfg::STAGE_BYTES), b_addr = a_addr + Cfg::A_BYTES;
          const uint32_t sfa_addr = b_addr + Cfg::B_BYTES, sfb_addr = sfa_addr + Cfg::SFA_BYTES;
          if (ptx::elect_one()) {
          for (int c = 0; c < kchunks; ++c) {
            ptx::tc_cp_32x128b_warpx4(tmem_base + Cfg::SFA_COL + c * 4, ptx::make_sf_desc(sfa_addr + c * 512));
            ptx::tc_cp_32x128b_warpx4(tmem_base + Cfg::SFB_COL + c * 8, ptx::make_sf_desc(sfb_addr + c * 512));
            ptx::tc_cp_32x128b_warpx4(tmem_base + Cfg::SFB_COL + c * 8 + 4, ptx::make_sf_desc(sfb_addr + 2048 + c * 512));
          }
          const uint64_t a_desc = ptx::make_sw128_kmajor_desc(a_addr), b_desc = ptx::make_sw128_kmajor_desc(b_addr);
          for (int c = 0; c < kchunks; ++c)
            ptx::tc_mma_nvfp4(tmem_base, a_desc + 2 * c, b_desc + 2 * c, idesc, (kb > 0 || c > 0) ? 1u : 0u,
                              tmem_base + Cfg::SFA_COL + c * 4, tmem_base + Cfg::SFB_COL + c * 8);
          ptx::tc_commit_mc(&empty[s], all_mask);  // stage s consumed here -> tell every producer in the cluster
          }
          __syncwarp();
        }
        if (ptx::elect_one()) ptx::tc_commit(tfull);
        __syncwarp();
      }
    }
  } else {
    // ===================== epilogue (warps 2..9): quadrant = warp % 4, column half = (warp - 2) / 4 ============
    const int q = warp & 3, half = (warp - 2) >> 2;
    float alpha = p.alpha;
    if (p.alpha_dev) alpha = __ldg(p.alpha_dev);
    const bool vec_ok = (p.N % 8) == 0;
    uint32_t tile_it = 0;
    for (int st = cluster_id; st < num_stiles; st += num_clusters, ++tile_it) {
      const int mt = (st % p.stiles_m) * CM + cm, nt = (st / p.stiles_m) * CN + cn;
      ptx::mbar_wait(tfull, tile_it & 1);
      ptx::tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + half * 128;
      uint32_t r[4][32];
#pragma unroll
      for (int cb = 0; cb < 4; ++cb) ptx::tmem_ld_32x32b_x32(t_row + cb * 32, r[cb]);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tempty);  // accumulator is in registers: the next tile's MMAs may start
      const int row = mt * kBM + q * 32 + lane;
      if (row < p.M) {
        __nv_bfloat16 *drow = p.d + (size_t)row * p.N;
#pragma unroll
        for (int cb = 0; cb < 4; ++cb) {
          const int col0 = nt * kBN + half * 128 + cb * 32;
          if (col0 >= p.N) break;
          if (vec_ok && col0 + 32 <= p.N) {
#pragma unroll
            for (int v = 0; v < 4; ++v) {
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                __nv_bfloat162 h2 = __floats2bfloat162_rn(alpha * __uint_as_float(r[cb][8 * v + 2 * e]),
                                                          alpha * __uint_as_float(r[cb][8 * v + 2 * e + 1]));
                o[e] = *reinterpret_cast<uint32_t *>(&h2);
              }
              *reinterpret_cast<uint4 *>(drow + col0 + 8 * v) = make_uint4(o[0], o[1], o[2], o[3]);
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (col0 + e < p.N) drow[col0 + e] = __float2bfloat16_rn(alpha * __uint_as_float(r[cb][e]));
          }
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();  // no CTA may exit while a peer can still multicast into it or signal its barriers
  if (warp == 1) ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

template <int CM, int CN, int STAGES>
int launch_cluster(const void *a, const void *b, CParams p, cudaStream_t stream) {
  using Cfg = CCfg<CM, CN, STAGES>;
  auto kern = gemm_nvfp4_cluster_kernel<CM, CN, STAGES>;
  static int max_clusters = -1;
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = Cfg::CSZ;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(kThreadsC);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (max_clusters < 0) {
    ARCQ_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    cfg.gridDim = dim3(Cfg::CSZ * 64);
    int n = 0;
    ARCQ_CUDA_TRY(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
    if (n < 1) { set_error("no %d-CTA cluster fits on this device", Cfg::CSZ); return ARCQ_ENOSUP; }
    max_clusters = n;
  }
  CUtensorMap ta, tb;
  if (int rc = cached_operand_map(&ta, a, p.M, p.K / 2, Cfg::A_SLICE_ROWS)) return rc;
  if (int rc = cached_operand_map(&tb, b, p.N, p.K / 2, Cfg::B_SLICE_ROWS)) return rc;
  p.tiles_m = (p.M + kBM - 1) / kBM;
  p.tiles_n = (p.N + kBN - 1) / kBN;
  p.stiles_m = (p.tiles_m + CM - 1) / CM;
  p.stiles_n = (p.tiles_n + CN - 1) / CN;
  int clusters = p.stiles_m * p.stiles_n;
  if (clusters > max_clusters) clusters = max_clusters;
  cfg.gridDim = dim3(clusters * Cfg::CSZ);
  ARCQ_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, ta, tb, p));
  count_launch();
  return ARCQ_OK;
}

}  // namespace

// shape = 10*CM + CN (e.g. 42 = 4 x 2 cluster); returns ARCQ_ENOSUP for unknown shapes
int launch_gemm_cluster(int shape, const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K,
                        float alpha, const float *alpha_dev, void *d, cudaStream_t stream) {
  CParams p{(const uint8_t *)sfa, (const uint8_t *)sfb, (__nv_bfloat16 *)d, alpha_dev, alpha, M, N, K, 0, 0, 0, 0};
  switch (shape) {
    case 21: return launch_cluster<2, 1, 4>(a, b, p, stream);
    case 22: return launch_cluster<2, 2, 4>(a, b, p, stream);
    case 41: return launch_cluster<4, 1, 4>(a, b, p, stream);
    case 42: return launch_cluster<4, 2, 4>(a, b, p, stream);
    case 24: return launch_cluster<2, 4, 4>(a, b, p, stream);
    default: set_error("unknown cluster shape %d", shape); return ARCQ_ENOSUP;
  }
}

}  // namespace arcq
