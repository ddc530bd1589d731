// gemm_splitk.cu -- NVFP4 block-scaled GEMM for SMALL M (decode / short prefill chunks: up to 8 row tiles of 128).
//
// At M <= 128 the GEMM is a weight stream: 0.5625 * K * N bytes of codes + scales against 2 * M * N * K FLOPs, two
// to three orders of magnitude under the tensor-core ridge (SURVEY.md §8d "min(tensor, HBM)").  One 128 x 128 output
// tile per CTA gives only N / 128 CTAs (32 for N = 4096) -- a fifth of the SMs pull on HBM.  Here a thread-block
// CLUSTER of S CTAs shares one output tile: CTA r streams k-blocks [r nkb/S, (r+1) nkb/S) of the weight through the
// same TMA -> tcgen05.mma (kind::mxf4nvf4, TMEM accumulator) pipeline as gemm.cu, writes its fp32 partial tile to its
// own shared memory, and after one cluster barrier the S CTAs each reduce a slice of the tile over DISTRIBUTED shared
// memory in a fixed order (deterministic, no workspace, no atomics, no second launch) and store bf16.
//
// Replaces (for this regime) the same reference call as gemm.cu: kernels/src/nvfp4.cu:35-132 behind bindings.cpp:99-120.
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"

namespace arcq {

int cached_operand_map(CUtensorMap *out, const void *ptr, int64_t rows, int64_t row_bytes, int box_rows);

namespace {

constexpr int kTM = 128, kTN = 128, kKBytes = 128, kKB = 256;
constexpr int kSkThreads = 192;    // warp 0 TMA, warp 1 MMA (+ TMEM allocation), warps 2..5 epilogue
constexpr int kSkStages = 3;

struct SkCfg {
  static constexpr int A_BYTES = kTM * kKBytes, B_BYTES = kTN * kKBytes, SF_BYTES = 4 * 512;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + 2 * SF_BYTES;   // 36 KiB
  static constexpr int RED_BYTES = kTM * kTN * 4;                        // fp32 partial tile, aliases the stage ring
  static constexpr int RING_BYTES = kSkStages * STAGE_BYTES;
  static constexpr int SMEM_BYTES = (RING_BYTES > RED_BYTES ? RING_BYTES : RED_BYTES) + 1024 + 256;
  static constexpr int SF_COLS = 32, SF_COL0 = kTN, TMEM_COLS = 256;     // 128 accumulator + 2 x (16 SFA + 16 SFB)
};

struct SkParams {
  const uint8_t *sfa, *sfb;
  __nv_bfloat16 *d;
  const float *alpha_dev, *col_scale;
  float alpha;
  int M, N, K;
  int64_t ldd;
  int splits;
  const __nv_bfloat16 *col_bias;   // nullable bf16 [N]
};

__device__ __forceinline__ float ld_dsmem_f32(uint32_t cluster_addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(cluster_addr) : "memory");
  return v;
}

__global__ void __launch_bounds__(kSkThreads, 2)
gemm_nvfp4_splitk_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                         const SkParams p) {
  using Cfg = SkCfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr int kBody = Cfg::RING_BYTES > Cfg::RED_BYTES ? Cfg::RING_BYTES : Cfg::RED_BYTES;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kBody);
  uint64_t *full = bars, *empty = bars + kSkStages, *tfull = bars + 2 * kSkStages;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tfull + 1);
  float *red = reinterpret_cast<float *>(smem);   // [row][col ^ (row & 31)], valid after the last MMA has read the ring

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.splits;
  const uint32_t rank = ptx::cluster_ctarank();
  const int nt = blockIdx.x / S, mt = blockIdx.y;   // output tile (mt, nt); the S CTAs of a cluster split its K range
  const int nkb = (p.K + kKB - 1) / kKB, kc_total = p.K / 64;
  const int kb0 = (int)((int64_t)nkb * rank / S), kb1 = (int)((int64_t)nkb * (rank + 1) / S);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_a);
    ptx::prefetch_tmap(&tmap_b);
    for (int s = 0; s < kSkStages; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
    ptx::mbar_init(tfull, 1);
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
    // ===================== TMA producer: this CTA's k-blocks of A (tokens), B (weights) and both scale atoms ==========
    uint32_t it = 0;
    for (int kb = kb0; kb < kb1; ++kb, ++it) {
      const int s = it % kSkStages;
      if (it >= (uint32_t)kSkStages) ptx::mbar_wait(&empty[s], ((it / kSkStages) & 1) ^ 1);
      uint8_t *st = smem + s * Cfg::STAGE_BYTES;
      int kchunks = kc_total - kb * 4;
      if (kchunks > 4) kchunks = 4;
      const uint32_t sf_bytes = (uint32_t)kchunks * 512u;
      if (ptx::elect_one()) {
        ptx::mbar_arrive_expect_tx(&full[s], Cfg::A_BYTES + Cfg::B_BYTES + 2 * sf_bytes);
        ptx::tma_load_2d(st, &tmap_a, kb * kKBytes, mt * kTM, &full[s]);
        ptx::tma_load_2d(st + Cfg::A_BYTES, &tmap_b, kb * kKBytes, nt * kTN, &full[s]);
        ptx::bulk_load(st + Cfg::A_BYTES + Cfg::B_BYTES, p.sfa + ((size_t)mt * kc_total + kb * 4) * 512, sf_bytes, &full[s]);
        ptx::bulk_load(st + Cfg::A_BYTES + Cfg::B_BYTES + Cfg::SF_BYTES, p.sfb + ((size_t)nt * kc_total + kb * 4) * 512, sf_bytes,
                       &full[s]);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = ptx::make_idesc_nvfp4(kTM, kTN);
    uint32_t it = 0;
    for (int kb = kb0; kb < kb1; ++kb, ++it) {
      const int s = it % kSkStages;
      ptx::mbar_wait(&full[s], (it / kSkStages) & 1);
      ptx::tc_fence_after();
      uint8_t *st = smem + s * Cfg::STAGE_BYTES;
      int kchunks = kc_total - kb * 4;
      if (kchunks > 4) kchunks = 4;
      const uint32_t a_addr = ptx::smem_u32(st), b_addr = a_addr + Cfg::A_BYTES;
      const uint32_t sfa_addr = b_addr + Cfg::B_BYTES, sfb_addr = sfa_addr + Cfg::SF_BYTES;
      const uint32_t sfa_t = tmem_base + Cfg::SF_COL0 + (it & 1) * Cfg::SF_COLS, sfb_t = sfa_t + 16;
      const uint64_t a_desc = ptx::make_sw128_kmajor_desc(a_addr), b_desc = ptx::make_sw128_kmajor_desc(b_addr);
      if (ptx::elect_one()) {
        for (int c = 0; c < kchunks; ++c) {
          ptx::tc_cp_32x128b_warpx4(sfa_t + c * 4, ptx::make_sf_desc(sfa_addr + c * 512));
          ptx::tc_cp_32x128b_warpx4(sfb_t + c * 4, ptx::make_sf_desc(sfb_addr + c * 512));
          ptx::tc_mma_nvfp4(tmem_base, a_desc + 2 * c, b_desc + 2 * c, idesc, (it > 0 || c > 0) ? 1u : 0u, sfa_t + c * 4,
                            sfb_t + c * 4);
        }
        ptx::tc_commit(&empty[s]);
      }
      __syncwarp();
    }
    if (ptx::elect_one()) ptx::tc_commit(tfull);
    __syncwarp();
  } else {
    // ===================== epilogue part 1 (warps 2..5): accumulator -> fp32 partial tile in OWN shared memory =======
    const int q = warp & 3;                      // TMEM lane quadrant this warp may read
    ptx::mbar_wait(tfull, 0);                    // every MMA of this CTA is complete: the stage ring is dead, `red` may alias it
    ptx::tc_fence_after();
    if (q * 32 < p.M - mt * kTM) {               // rows >= M are never reduced
      const int row = q * 32 + lane;
#pragma unroll
      for (int cb = 0; cb < kTN / 32; ++cb) {
        uint32_t r[32];
        ptx::tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + cb * 32, r);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) red[row * kTN + ((cb * 32 + j) ^ lane)] = __uint_as_float(r[j]);
      }
    }
    ptx::tc_fence_before();
  }

  // ===================== epilogue part 2 (all threads of all S CTAs): reduce over distributed shared memory ==========
  __syncthreads();
  ptx::cluster_sync();   // release/acquire: every CTA's partial tile is visible cluster-wide
  {
    float alpha = p.alpha;
    if (p.alpha_dev) alpha = __ldg(p.alpha_dev);
    const int Mv = p.M - mt * kTM < kTM ? p.M - mt * kTM : kTM;
    const int nunits = Mv * (kTN / 32);          // a unit = 32 consecutive columns of one row, one per warp step
    constexpr int NW = kSkThreads / 32;
    const uint32_t red_s = ptx::smem_u32(red);
    for (int u = (int)rank + S * warp; u < nunits; u += S * NW) {
      const int row = u >> 2, col = (u & 3) * 32 + lane;
      const uint32_t off = (uint32_t)(row * kTN + (col ^ (row & 31))) * 4u;
      float acc = 0.0f;
      for (int s = 0; s < S; ++s) {              // fixed order: bit-reproducible
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(red_s + off), "r"(s));
        acc += ld_dsmem_f32(remote);
      }
      const int gcol = nt * kTN + col;
      if (gcol < p.N) {
        float sc = alpha;
        if (p.col_scale) sc = __fmul_rn(alpha, __ldg(p.col_scale + gcol));
        __nv_bfloat16 y = __float2bfloat16_rn(sc * acc);
        if (p.col_bias) y = __float2bfloat16_rn(__fadd_rn(__bfloat162float(y), __bfloat162float(p.col_bias[gcol])));   // torch's `y + bias`
        p.d[(size_t)(mt * kTM + row) * p.ldd + gcol] = y;
      }
    }
  }
  ptx::cluster_sync();   // nobody leaves (and frees its shared memory) while a peer may still read it
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace

// Measured on the Llama-3-8B shapes (profiles/r2_midm_splitk_ab.log): the cluster kernel wins while the output has few
// 128 x 128 tiles -- up to 32 tiles with 8 splits (q/o and k/v at M <= 128 .. 512: 10.0 vs 12.1 us, 7.8 vs 11.6 us; down at
// M <= 128: 13.7 vs 29.6 us), up to 64 tiles with 4 splits when K is long (down at M = 256: 22.2 vs 30.5 us) -- and loses
// beyond that (its TMEM -> shared -> DSMEM reduction costs more than the idle SMs were worth), so: no split otherwise.
int splitk_choice(int M, int N, int K) {
  if (M > 8 * kTM) return 0;
  const int tiles = ((N + kTN - 1) / kTN) * ((M + kTM - 1) / kTM), nkb = (K + kKB - 1) / kKB;
  // (33..64 tiles with a short K: only for M <= 32, where the reduction is a few rows -- fused q/k/v at M = 8: 9.6 vs 13.6 us)
  // (with a short K the split only pays while the reduction is a few rows: q/o at M = 128 13.1 us split against ~9.9 us for one
  // 128-wide tile per CTA, profiles/r2_midm_kernel_choice_ab.log; down_proj's long K keeps its split up to M = 256)
  if (nkb < 32 && M > 64) return 0;
  const int max_s = tiles <= 32 ? 8 : (tiles <= 64 && (nkb >= 32 || M <= 32) ? 4 : 1);
  int S = 1;
  while (S * 2 <= max_s && nkb / (S * 2) >= 2) S *= 2;
  return S >= 2 ? S : 0;
}

int launch_gemm_splitk(const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K, float alpha,
                       const float *alpha_dev, const float *col_scale, void *d, int64_t ldd, int splits, cudaStream_t stream,
                       const void *col_bias) {
  if (M > 65535 * kTM) { set_error("split-K kernel: M=%d too large", M); return ARCQ_ENOSUP; }
  const int nkb = (K + kKB - 1) / kKB;
  if (splits > 8) splits = 8;
  while (splits > 1 && (nkb < splits || (splits & (splits - 1)))) --splits;   // power of two, every CTA gets a k-block
  if (splits < 1) splits = 1;
  auto kern = gemm_nvfp4_splitk_kernel;
  static SmemAttrCache attr_cache = {};
  if (int rc = ensure_dynamic_smem(kern, SkCfg::SMEM_BYTES, attr_cache)) return rc;
  CUtensorMap ta, tb;
  if (int rc = cached_operand_map(&ta, a, M, K / 2, kTM)) return rc;
  if (int rc = cached_operand_map(&tb, b, N, K / 2, kTN)) return rc;
  SkParams p{(const uint8_t *)sfa, (const uint8_t *)sfb, (__nv_bfloat16 *)d, alpha_dev, col_scale, alpha, M, N, K, ldd, splits, (const __nv_bfloat16 *)col_bias};
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = splits;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.gridDim = dim3(((N + kTN - 1) / kTN) * splits, (M + kTM - 1) / kTM);
  cfg.blockDim = dim3(kSkThreads);
  cfg.dynamicSmemBytes = SkCfg::SMEM_BYTES;
  cfg.stream = stream;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  ARCQ_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, ta, tb, p));
  count_launch();
  return ARCQ_OK;
}

}  // namespace arcq
