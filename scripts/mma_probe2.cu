// mma_probe2.cu -- raw cta_group::2 tcgen05.mma / tcgen05.cp rates on one SM pair (GPU box only).
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace arcq;

// mode 0: 4 pair-MMAs per k-block; 1: + 12 pair-UTCCP; 2: 12 pair-UTCCP only; 3: 4 MMA + 8 UTCCP (SFB split)
// mode 4: mode 1 + tcgen05.commit (multicast to both CTAs) after every k-block; 5: mode 4 + mbarrier test + fence per k-block
// mode 6: mode 4 but commit to the local CTA only
__global__ void __launch_bounds__(128, 1) probe(int n_kblocks, int mode, long long *out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar, bar2[4];
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = ptx::cluster_ctarank();
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x22222222u + i * 2654435761u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); for (int i = 0; i < 4; ++i) ptx::mbar_init(&bar2[i], 1); ptx::fence_barrier_init(); }
  if (warp == 0) { ptx::tmem_alloc_pair(&tslot, 512); ptx::tmem_relinquish_pair(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tb = tslot;
  if (warp == 0 && rank == 0) {
    constexpr uint32_t idesc = ptx::make_idesc_nvfp4(256, 256);
    long long t0 = clock64();
    uint32_t acc = 0;
    for (int kb = 0; kb < n_kblocks; ++kb) {
      if (ptx::elect_one()) {
        const int stage = kb & 3;
        const uint32_t a_addr = ptx::smem_u32(smem) + stage * 38912, b_addr = a_addr + 16384;
        const uint32_t sf_addr = b_addr + 16384;
        const uint32_t sfa_t = tb + 256 + (kb & 1) * 48, sfb_t = sfa_t + 16;
        for (int c = 0; c < 4; ++c) {
          if (mode == 1 || mode == 2 || mode == 3 || mode >= 4) {
            ptx::tc_cp_32x128b_warpx4_pair(sfa_t + c * 4, ptx::make_sf_desc(sf_addr + c * 512));
            ptx::tc_cp_32x128b_warpx4_pair(sfb_t + c * 8, ptx::make_sf_desc(sf_addr + 2048 + c * 512));
            if (mode != 3) ptx::tc_cp_32x128b_warpx4_pair(sfb_t + c * 8 + 4, ptx::make_sf_desc(sf_addr + 4096 + c * 512));
          }
          if (mode != 2)
            ptx::tc_mma_nvfp4_pair(tb, ptx::make_sw128_kmajor_desc(a_addr) + 2 * c, ptx::make_sw128_kmajor_desc(b_addr) + 2 * c, idesc,
                                   (kb | c) ? 1u : 0u, sfa_t + c * 4, sfb_t + c * 8);
        }
        if (mode == 4 || mode == 5) ptx::tc_commit_pair(&bar2[kb & 3], 3);
        if (mode == 6) ptx::tc_commit_pair(&bar2[kb & 3], 1);
      }
      __syncwarp();
      if (mode == 5) { acc += ptx::mbar_test(&bar2[(kb + 1) & 3], 0); ptx::tc_fence_after(); }
    }
    long long t1 = clock64();
    if (ptx::elect_one()) ptx::tc_commit_pair(&bar, 1);
    __syncwarp();
    ptx::mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; out[2] = acc; }
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 0) ptx::tmem_dealloc_pair(tb, 512);
}

int main() {
  long long *d; cudaMalloc(&d, 64);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024);
  const char *names[] = {"4 pair-MMA 256x256x64", "4 MMA + 12 UTCCP", "12 UTCCP only", "4 MMA + 8 UTCCP", "mode1 + commit/kb (mcast)", "mode4 + test + fence", "mode1 + commit/kb (local)"};
  for (int mode = 0; mode < 7; ++mode) {
    const int nkb = 2000;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.gridDim = dim3(2); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = 210 * 1024; cfg.attrs = attr; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, probe, nkb, mode, d);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("pair mode %d %-26s: issue %.1f clk/k-block, complete %.1f clk/k-block (%s)\n", mode, names[mode], (double)h[0] / nkb,
           (double)h[1] / nkb, cudaGetErrorString(e));
  }
  return 0;
}
