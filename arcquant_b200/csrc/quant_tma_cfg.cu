// quant_tma_cfg.cu -- ONE tile organisation of the TMA-staged quantise kernel (quant_tma.cuh), selected with
// -DARCQ_CFG_R / _RT / _TH / _MB; build.py compiles this file once per organisation so that nvcc works on them in parallel.
#include "quant_tma.cuh"

#if !defined(ARCQ_CFG_R) || !defined(ARCQ_CFG_RT) || !defined(ARCQ_CFG_TH) || !defined(ARCQ_CFG_MB)
#error "compile with -DARCQ_CFG_R=.. -DARCQ_CFG_RT=.. -DARCQ_CFG_TH=.. -DARCQ_CFG_MB=.."
#endif
#define ARCQ_CFG_CAT2(a, b, c, d) launch_tma_cfg_##a##_##b##_##c##_##d
#define ARCQ_CFG_CAT(a, b, c, d) ARCQ_CFG_CAT2(a, b, c, d)

namespace arcq {

int ARCQ_CFG_CAT(ARCQ_CFG_R, ARCQ_CFG_RT, ARCQ_CFG_TH, ARCQ_CFG_MB)(const QuantParams &p, int flavour, bool is_w, bool rms, int S, bool tab32,
                                                                   bool own, cudaStream_t st, bool own2) {
  return launch_kind<ARCQ_CFG_R, ARCQ_CFG_RT, ARCQ_CFG_TH, ARCQ_CFG_MB>(p, flavour, is_w, rms, S, tab32, own, st, own2);
}

}  // namespace arcq
