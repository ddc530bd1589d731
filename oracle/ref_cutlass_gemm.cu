// ref_cutlass_gemm.cu -- TEST / MEASUREMENT INFRASTRUCTURE ONLY (never linked into the product).
//
// "The reference's GEMM, retargeted": /root/reference/kernels/src/nvfp4.cu instantiates a CUTLASS 3.x block-scaled GEMM
// with ArchTag = Sm120 (mma.sync), which cannot execute on a B200.  This shim instantiates the SAME CUTLASS collective
// builder family from the reference's own vendored CUTLASS (third-party/cutlass, v4.3.1) with ArchTag = Sm100 and the tile /
// cluster shape of CUTLASS example 72a (examples/72_blackwell_narrow_precision_gemm/72a_blackwell_nvfp4_bf16_gemm.cu:114-119:
// 256 x 256 x 256 MMA tile, cluster 2 x 4), i.e. what a maintainer would get by retargeting nvfp4.cu:28-33.  It is the
// "reference config retargeted" number to beat of BASELINE.md §1 and an independent GEMM oracle on the device.
// Built by oracle/build_ref.sh into oracle/_ref/libarcq_refgemm.so (git-ignored).  Semantics: nvfp4.cu:98-121
// (D = alpha * A.SFA * (B.SFB)^T, beta = 0, bf16 out).
#include <cuda_runtime.h>

#include "cutlass/cutlass.h"
#include "cute/tensor.hpp"
#include "cutlass/epilogue/collective/collective_builder.hpp"
#include "cutlass/gemm/collective/collective_builder.hpp"
#include "cutlass/gemm/device/gemm_universal_adapter.h"
#include "cutlass/gemm/kernel/gemm_universal.hpp"
#include "cutlass/util/packed_stride.hpp"

using namespace cute;

namespace {

template <class TileShape, class ClusterShape>
struct RefGemm {
  using Fp4 = cutlass::nv_float4_t<cutlass::float_e2m1_t>;
  using Out = cutlass::bfloat16_t;
  using Epi = typename cutlass::epilogue::collective::CollectiveBuilder<
      cutlass::arch::Sm100, cutlass::arch::OpClassBlockScaledTensorOp, TileShape, ClusterShape,
      cutlass::epilogue::collective::EpilogueTileAuto, float, float, Out, cutlass::layout::RowMajor, 8, Out,
      cutlass::layout::RowMajor, 8, cutlass::epilogue::collective::EpilogueScheduleAuto>::CollectiveOp;
  using Main = typename cutlass::gemm::collective::CollectiveBuilder<
      cutlass::arch::Sm100, cutlass::arch::OpClassBlockScaledTensorOp, Fp4, cutlass::layout::RowMajor, 32, Fp4,
      cutlass::layout::ColumnMajor, 32, float, TileShape, ClusterShape,
      cutlass::gemm::collective::StageCountAutoCarveout<static_cast<int>(sizeof(typename Epi::SharedStorage))>,
      cutlass::gemm::collective::KernelScheduleAuto>::CollectiveOp;
  using Kernel = cutlass::gemm::kernel::GemmUniversal<Shape<int, int, int, int>, Main, Epi, void>;
  using Gemm = cutlass::gemm::device::GemmUniversalAdapter<Kernel>;

  static int run(const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K, float alpha, void *d,
                 void *workspace, size_t workspace_bytes, cudaStream_t stream) {
    using SA = typename Kernel::StrideA;
    using SB = typename Kernel::StrideB;
    using SC = typename Kernel::StrideC;
    using SD = typename Kernel::StrideD;
    using Cfg = typename Kernel::CollectiveMainloop::Sm1xxBlkScaledConfig;
    auto shape = make_shape(M, N, K, 1);
    typename Gemm::Arguments args{
        cutlass::gemm::GemmUniversalMode::kGemm, shape,
        {reinterpret_cast<const cutlass::float_e2m1_t *>(a), cutlass::make_cute_packed_stride(SA{}, {M, K, 1}),
         reinterpret_cast<const cutlass::float_e2m1_t *>(b), cutlass::make_cute_packed_stride(SB{}, {N, K, 1}),
         reinterpret_cast<const cutlass::float_ue4m3_t *>(sfa), Cfg::tile_atom_to_shape_SFA(shape),
         reinterpret_cast<const cutlass::float_ue4m3_t *>(sfb), Cfg::tile_atom_to_shape_SFB(shape)},
        {{alpha, 0.0f}, reinterpret_cast<const Out *>(d), cutlass::make_cute_packed_stride(SC{}, {M, N, 1}),
         reinterpret_cast<Out *>(d), cutlass::make_cute_packed_stride(SD{}, {M, N, 1})}};
    Gemm op;
    if (Gemm::get_workspace_size(args) > workspace_bytes) return -3;
    if (op.can_implement(args) != cutlass::Status::kSuccess) return -1;
    if (op.initialize(args, workspace, stream) != cutlass::Status::kSuccess) return -2;
    return op.run(stream) == cutlass::Status::kSuccess ? 0 : -4;
  }
};

}  // namespace

// variant 0: 256 x 256 x 256, cluster 2 x 4 (example 72a); variant 1: same tile, cluster 2 x 1; variant 2: 128 x 128 x 256,
// cluster 1 x 1 (the reference's own 128 x 128 tile / 1 x 1 cluster of nvfp4.cu:32-33, on the Sm100 path)
extern "C" int arcq_refgemm(int variant, const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K,
                            float alpha, void *d, void *workspace, long long workspace_bytes, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  switch (variant) {
    case 0: return RefGemm<Shape<_256, _256, _256>, Shape<_2, _4, _1>>::run(a, b, sfa, sfb, M, N, K, alpha, d, workspace, (size_t)workspace_bytes, st);
    case 1: return RefGemm<Shape<_256, _256, _256>, Shape<_2, _1, _1>>::run(a, b, sfa, sfb, M, N, K, alpha, d, workspace, (size_t)workspace_bytes, st);
    case 2: return RefGemm<Shape<_128, _128, _256>, Shape<_1, _1, _1>>::run(a, b, sfa, sfb, M, N, K, alpha, d, workspace, (size_t)workspace_bytes, st);
    default: return -9;
  }
}
