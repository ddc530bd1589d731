// sync_probe2.cu -- cost of individual mbarrier operations (GPU box only).
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace arcq;

__device__ __forceinline__ bool mbar_test_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(ptx::smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}

// mode 0: try_wait on an already-complete phase, repeated;  1: test_wait same;  2: arrive repeated (count huge)
// mode 3: try_wait + arrive on own barrier (self ping: count 1)  4: same with test_wait
// mode 5: ring with test_wait polling (2 warps, 4 stages)
__global__ void probe(int iters, int mode, long long *out) {
  __shared__ uint64_t bar, full[4], empty[4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    ptx::mbar_init(&bar, mode == 2 ? (1 << 20) : 1);
    for (int s = 0; s < 4; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
    ptx::fence_barrier_init();
  }
  __syncthreads();
  long long t0 = clock64();
  if (mode <= 4 && warp == 0) {
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
      if (mode == 0) acc += ptx::mbar_try_wait(&bar, 1);       // phase "-1" counts as complete
      else if (mode == 1) acc += mbar_test_wait(&bar, 1);
      else if (mode == 2) { if (lane == 0) ptx::mbar_arrive(&bar); __syncwarp(); }
      else if (mode == 3) { if (lane == 0) ptx::mbar_arrive(&bar); __syncwarp(); while (!ptx::mbar_try_wait(&bar, it & 1)) {} }
      else if (mode == 4) { if (lane == 0) ptx::mbar_arrive(&bar); __syncwarp(); while (!mbar_test_wait(&bar, it & 1)) {} }
    }
    if (lane == 0) { out[0] = clock64() - t0; out[1] = acc; }
  } else if (mode == 5) {
    if (warp == 0) {
      for (int it = 0; it < iters; ++it) {
        const int s = it & 3; const uint32_t ph = (it >> 2) & 1;
        while (!mbar_test_wait(&empty[s], ph ^ 1)) {}
        if (lane == 0) ptx::mbar_arrive(&full[s]);
        __syncwarp();
      }
    } else if (warp == 1) {
      for (int it = 0; it < iters; ++it) {
        const int s = it & 3; const uint32_t ph = (it >> 2) & 1;
        while (!mbar_test_wait(&full[s], ph)) {}
        if (lane == 0) ptx::mbar_arrive(&empty[s]);
        __syncwarp();
      }
      if (lane == 0) out[0] = clock64() - t0;
    }
  } else if (mode == 6) {   // ring where only lane 0 of each warp participates at all (no syncwarp)
    if (warp == 0 && lane == 0) {
      for (int it = 0; it < iters; ++it) {
        const int s = it & 3; const uint32_t ph = (it >> 2) & 1;
        while (!mbar_test_wait(&empty[s], ph ^ 1)) {}
        ptx::mbar_arrive(&full[s]);
      }
    } else if (warp == 1 && lane == 0) {
      for (int it = 0; it < iters; ++it) {
        const int s = it & 3; const uint32_t ph = (it >> 2) & 1;
        while (!mbar_test_wait(&full[s], ph)) {}
        ptx::mbar_arrive(&empty[s]);
      }
      out[0] = clock64() - t0;
    }
  }
}

int main() {
  long long *d; cudaMalloc(&d, 64);
  const char *names[] = {"try_wait on complete phase", "test_wait on complete phase", "arrive only", "self arrive + try_wait",
                         "self arrive + test_wait", "2-warp ring, test_wait, 4 stages", "2-thread ring, test_wait, 4 stages"};
  for (int mode = 0; mode <= 6; ++mode) {
    const int iters = 20000;
    probe<<<1, 64>>>(iters, mode, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("mode %d %-40s: %.1f clk/iter (%s)\n", mode, names[mode], (double)h[0] / iters, cudaGetErrorString(e));
  }
  return 0;
}
