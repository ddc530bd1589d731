// quant_tma.cu -- shape-driven choice of the tile organisation of the TMA-staged quantise kernel (quant_tma.cuh).
#include "quant_tma.cuh"

namespace arcq {
namespace {

struct Cfg { int R, threads, minb, S, RT; };

}  // namespace

// one translation unit per tile organisation (quant_tma_cfg.cu compiled with -DARCQ_CFG_*; see build.py)
#define ARCQ_CFG_DECL(RR, RTT, TH, MB) \
  int launch_tma_cfg_##RR##_##RTT##_##TH##_##MB(const QuantParams &p, int flavour, bool is_w, bool rms, int S, bool tab32, bool own, cudaStream_t st, bool own2);
ARCQ_CFG_DECL(4, 4, 256, 2) ARCQ_CFG_DECL(4, 4, 512, 1) ARCQ_CFG_DECL(2, 2, 256, 2) ARCQ_CFG_DECL(2, 2, 512, 1) ARCQ_CFG_DECL(1, 1, 512, 1)
ARCQ_CFG_DECL(8, 8, 512, 1) ARCQ_CFG_DECL(8, 4, 512, 1)
#undef ARCQ_CFG_DECL

int launch_quant_tma(const QuantParams &p, int flavour, bool is_w, bool rms, cudaStream_t stream) {
  if (p.KQ % 64 || p.KE % 64) return ARCQ_ENOSUP;  // 4-slot scale chunks must not straddle the plain/ARC boundary
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  const int ngroups = p.KQ / 16;
  const int64_t blocks = ((int64_t)p.M + 127) / 128;
  const int wide = 512;  // 1024 threads (64 registers) measured slower than 512 on KQ=14336 (profiles/r1_quant_v3.log)
  // tile height, threads per CTA, CTAs per SM, staging slots -- first candidate that fits and gives every SM a tile
  // 8-row tiles with two threads per 16-group (16-byte chunks: 8 random chunks per shared-memory wavefront instead of
  // 16, scale bytes leave as whole 32-byte sectors) measured 5-7 % faster than 4-row tiles at two CTAs per SM where
  // every thread owns one item (profiles/r1_quant_r8_sweep.log)
  const Cfg cands[] = {{8, 512, 1, 1, 4}, {4, 256, 2, 2, 4}, {4, 256, 2, 1, 4}, {4, 512, 1, 1, 4}, {4, 512, 1, 2, 4}, {2, 256, 2, 2, 2}, {2, wide, 1, 2, 2}, {2, wide, 1, 1, 2},
                       {1, wide, 1, 2, 1}, {1, wide, 1, 1, 1}};
  Cfg c{0, 0, 0, 0, 0};
  bool tab32 = false, own = false;
  {  // arcq_set_option("quant_tile", "R,threads,S,RT,CTAs/SM"): the parity tests force every tile organisation
    const int *q = options().quant_tile;
    if (q[0] > 0) c = Cfg{q[0], q[1], q[4], q[2], q[3]};
  }
  const bool add = rms && p.res != nullptr;   // fused residual add: second set of staging slots, register transposition
  // (the rmsnorm variant on the planned path -- sum of squares in a pass of its own -- measured 36.5 us against 32.8 us
  // for the register transposition at 8192 x 4096, profiles/r2_quant_rms_plan_ab.log: it stays unplanned)
  const bool want_plan = p.plan != nullptr && !is_w && !rms;
  // with a plan the gather is conflict-free at any chunk width, which was the 8-row tile's advantage: for KQ <= 4096 the
  // 4-row tile at two CTAs per SM with ONE staging slot is then 3-4 % faster (profiles/r2_quant_tile_ab.log)
  const bool prefer4 = want_plan && ngroups <= 256 && plan_usable(4, p.KQ);
  if (c.R == 0) {
    for (const Cfg &k : cands) {
      if (prefer4 && (k.R == 8 || (k.R == 4 && k.threads == 256 && k.S == 2))) continue;
      if (rms && k.threads < 32 * k.R) continue;
      if (k.R == 8 && ngroups * (k.R / k.RT) > k.threads) continue;  // only with one owned item per thread
      if (k.R == 4 && k.threads == 256 && ngroups > 256) continue;   // two passes per tile: the wide CTA owns its items
      // a placement plan is an extra: the tile organisation is chosen as if there were none, and the plan is used only
      // if it still fits next to it (below); it must never push the choice to a smaller tile
      const int need = carve_t(k.R, p.KQ, p.KE, rms, false, k.S, false, false, add).total + 1024;
      if (need * k.minb > kSmemPerSm) continue;
      if (blocks * (128 / k.R) < (int64_t)sms * k.minb && k.R > 1) continue;
      c = k;
      break;
    }
    if (c.R == 0) return ARCQ_ENOSUP;
  }
  own = ngroups * (c.R / c.RT) <= c.threads;
  QuantParams pp = p;
  bool planned = want_plan && plan_usable(c.R, p.KQ);
  // two owned groups per thread (offsets in registers, no table in shared memory): planned 4- and 2-row tiles whose byte
  // offsets fit 16 bits
  bool own2 = planned && !own && c.R == c.RT && c.R <= 4 && ngroups <= 2 * c.threads && p.KQ * 2 * c.R <= 65536;
  if (planned && (carve_t(c.R, p.KQ, p.KE, rms, false, c.S, true, own2).total + 1024) * c.minb > kSmemPerSm) {
    if (c.S > 1 && (carve_t(c.R, p.KQ, p.KE, rms, false, 1, true, own2).total + 1024) * c.minb <= kSmemPerSm) c.S = 1;  // trade the second staging slot
    else planned = own2 = false;
  }
  pp.plan = planned ? p.plan + (int64_t)plan_section(c.R) * 2 * p.KQ : nullptr;   // this tile height's section
  tab32 = !own && !planned && (carve_t(c.R, p.KQ, p.KE, rms, true, c.S, false, false, add).total + 1024) * c.minb <= kSmemPerSm;
  if ((carve_t(c.R, p.KQ, p.KE, rms, tab32, c.S, planned, own2, add).total + 1024) * c.minb > kSmemPerSm) return ARCQ_ENOSUP;
#define ARCQ_T(RR, RTT, TH, MB) \
  if (c.R == RR && c.RT == RTT && c.threads == TH && c.minb == MB) \
    return launch_tma_cfg_##RR##_##RTT##_##TH##_##MB(pp, flavour, is_w, rms, c.S, tab32, own, stream, own2);
  ARCQ_T(4, 4, 256, 2) ARCQ_T(4, 4, 512, 1) ARCQ_T(2, 2, 256, 2) ARCQ_T(2, 2, 512, 1) ARCQ_T(1, 1, 512, 1)
  ARCQ_T(8, 8, 512, 1) ARCQ_T(8, 4, 512, 1)
#undef ARCQ_T
  return ARCQ_ENOSUP;
}

}  // namespace arcq
