// sync_probe.cu -- microbenchmark of the producer/consumer mbarrier ring used by gemm.cu (GPU box only).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -Iarcquant_b200/csrc -Iinclude scripts/sync_probe.cu -o /tmp/sync_probe
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace arcq;

// mode bits: 1 = only lane 0 polls (then __syncwarp); 2 = extra spinner warps polling a never-completing barrier;
//            4 = consumer does tcgen05.fence each iteration; 8 = use elect_one instead of lane==0 for the arrive
template <int STAGES>
__global__ void ring(int iters, int mode, long long *out) {
  __shared__ uint64_t full[STAGES], empty[STAGES], never;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
    ptx::mbar_init(&never, 1);
    ptx::fence_barrier_init();
  }
  __syncthreads();
  long long t0 = clock64();
  if (warp == 0) {
    for (int it = 0; it < iters; ++it) {
      const int s = it % STAGES; const uint32_t ph = (it / STAGES) & 1;
      if (mode & 1) { if (lane == 0) ptx::mbar_wait(&empty[s], ph ^ 1); __syncwarp(); }
      else ptx::mbar_wait(&empty[s], ph ^ 1);
      if (mode & 8) { if (ptx::elect_one()) ptx::mbar_arrive(&full[s]); }
      else if (lane == 0) ptx::mbar_arrive(&full[s]);
      __syncwarp();
    }
  } else if (warp == 1) {
    for (int it = 0; it < iters; ++it) {
      const int s = it % STAGES; const uint32_t ph = (it / STAGES) & 1;
      if (mode & 1) { if (lane == 0) ptx::mbar_wait(&full[s], ph); __syncwarp(); }
      else ptx::mbar_wait(&full[s], ph);
      if (mode & 4) ptx::tc_fence_after();
      if (mode & 8) { if (ptx::elect_one()) ptx::mbar_arrive(&empty[s]); }
      else if (lane == 0) ptx::mbar_arrive(&empty[s]);
      __syncwarp();
    }
    if (lane == 0) { out[blockIdx.x] = clock64() - t0; ptx::mbar_arrive(&never); }
  } else if (mode & 2) {
    ptx::mbar_wait(&never, 0);  // spins until the consumer finishes
  }
}

template <int STAGES> void run(int threads, int mode, const char *name) {
  long long *d; cudaMalloc(&d, 148 * 8);
  const int iters = 20000;
  ring<STAGES><<<148, threads>>>(iters, mode, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("stages=%d threads=%d mode=%2d %-40s : %.1f clk/iter (%s)\n", STAGES, threads, mode, name, (double)h[0] / iters,
         cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  run<3>(64, 0, "all lanes poll");
  run<3>(64, 1, "lane0 polls");
  run<3>(64, 8, "all lanes poll, elect arrive");
  run<3>(64, 4, "all lanes poll + tcgen05.fence");
  run<3>(320, 0, "all lanes poll, 8 idle warps exited");
  run<3>(320, 2, "all lanes poll + 8 spinner warps");
  run<3>(320, 3, "lane0 polls + 8 spinner warps");
  run<1>(64, 0, "1 stage (pure 2-hop RTT)");
  run<2>(64, 0, "2 stages");
  run<4>(64, 0, "4 stages");
  run<8>(64, 0, "8 stages");
  run<8>(320, 2, "8 stages + spinners");
  return 0;
}
