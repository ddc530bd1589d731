// mma_probe.cu -- raw tcgen05.mma kind::mxf4nvf4 issue/execute rate on one SM (GPU box only).
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace arcq;

// mode 0: MMAs only.  mode 1: per 4 MMAs also 12 UTCCP (like one k-block).  mode 2: UTCCP only.
// mode 3: MMAs, each reading a different 48 KB stage (4 stages round robin) -- same as the GEMM's smem footprint
template <int BN>
__global__ void __launch_bounds__(128, 1) probe(int n_kblocks, int mode, long long *out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x22222222u + i * 2654435761u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
  if (warp == 0) { ptx::tmem_alloc(&tslot, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tb = tslot;
  if (warp == 0) {
    constexpr uint32_t idesc = ptx::make_idesc_nvfp4(128, BN);
    long long t0 = clock64();
    for (int kb = 0; kb < n_kblocks; ++kb) {
      if (ptx::elect_one()) {
        const int stage = (mode == 3) ? (kb & 3) : 0;
        const uint32_t a_addr = ptx::smem_u32(smem) + stage * 49152, b_addr = a_addr + 16384;
        const uint32_t sf_addr = ptx::smem_u32(smem) + 196608;
        const uint32_t sfa_t = tb + 256 + (kb & 1) * 48, sfb_t = sfa_t + 16;
        for (int c = 0; c < 4; ++c) {
          if (mode == 1 || mode == 2) {
            ptx::tc_cp_32x128b_warpx4(sfa_t + c * 4, ptx::make_sf_desc(sf_addr + c * 512));
            ptx::tc_cp_32x128b_warpx4(sfb_t + c * 8, ptx::make_sf_desc(sf_addr + 2048 + c * 512));
            ptx::tc_cp_32x128b_warpx4(sfb_t + c * 8 + 4, ptx::make_sf_desc(sf_addr + 4096 + c * 512));
          }
          if (mode != 2)
            ptx::tc_mma_nvfp4(tb, ptx::make_sw128_kmajor_desc(a_addr) + 2 * c, ptx::make_sw128_kmajor_desc(b_addr) + 2 * c, idesc,
                              (kb | c) ? 1u : 0u, sfa_t + c * 4, sfb_t + c * 8);
        }
      }
      __syncwarp();
    }
    long long t1 = clock64();
    if (ptx::elect_one()) ptx::tc_commit(&bar);
    __syncwarp();
    ptx::mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tb, 512);
}

template <int BN> void run(int mode, const char *name) {
  long long *d; cudaMalloc(&d, 64);
  auto k = probe<BN>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024);
  const int nkb = 2000;
  k<<<1, 128, 210 * 1024>>>(nkb, mode, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("BN=%d mode %d %-34s: issue %.1f clk/k-block, complete %.1f clk/k-block (%s)\n", BN, mode, name, (double)h[0] / nkb,
         (double)h[1] / nkb, cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  run<256>(0, "4 MMA per k-block");
  run<256>(1, "4 MMA + 12 UTCCP");
  run<256>(2, "12 UTCCP only");
  run<256>(3, "4 MMA, 4 rotating stages");
  run<128>(0, "4 MMA per k-block");
  run<128>(1, "4 MMA + 12 UTCCP");
  return 0;
}
