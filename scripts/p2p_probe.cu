// p2p_probe.cu -- how fast can one GPU write into a peer's memory, by instruction kind?  (measurement aid, not product)
//   st     : every thread streams 16-byte st.global to the peer (what csrc/peer.cu and the scatter epilogue do)
//   bulk   : a block stages CHUNK bytes in shared memory (cp.async.bulk global -> shared) and writes them to the peer with
//            ONE cp.async.bulk shared -> global (TMA bulk store), double buffered
// One process, devices 0 and 1, one direction and both directions at once.  nvcc -arch=sm_100a -O3 -o p2p_probe p2p_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

__global__ void st_copy(const uint4 *__restrict__ src, uint4 *__restrict__ dst, size_t n16) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = __ldcs(src + i);
}

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int CHUNK>
__global__ void bulk_copy(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, size_t nchunks) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ uint64_t bar[2];
  if (threadIdx.x == 0) {
    for (int b = 0; b < 2; ++b) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar[b])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    uint32_t ph[2] = {0, 0};
    size_t c = blockIdx.x;
    int b = 0;
    if (c < nchunks) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar[0])), "r"(CHUNK) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(sm)), "l"(src + c * CHUNK), "r"(CHUNK), "r"(s32(&bar[0])) : "memory");
    }
    for (; c < nchunks; c += gridDim.x, b ^= 1) {
      const size_t nx = c + gridDim.x;
      if (nx < nchunks) {   // the other buffer: its previous store must have READ shared memory before it is overwritten
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar[b ^ 1])), "r"(CHUNK) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(sm + (b ^ 1) * CHUNK)), "l"(src + nx * CHUNK), "r"(CHUNK), "r"(s32(&bar[b ^ 1])) : "memory");
      }
      uint32_t ok = 0;
      while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p;}" : "=r"(ok) : "r"(s32(&bar[b])), "r"(ph[b]) : "memory");
      ph[b] ^= 1;
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + c * CHUNK), "r"(s32(sm + b * CHUNK)), "r"(CHUNK) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

int main() {
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  if (n < 2) { printf("{\"error\": \"needs 2 GPUs\"}\n"); return 0; }
  const size_t nb = 256u << 20;
  uint8_t *a[2], *b[2];
  cudaStream_t st[2];
  cudaEvent_t e0[2], e1[2];
  for (int d = 0; d < 2; ++d) {
    CK(cudaSetDevice(d));
    cudaDeviceEnablePeerAccess(1 - d, 0);
    CK(cudaMalloc(&a[d], nb)); CK(cudaMalloc(&b[d], nb));
    CK(cudaMemset(a[d], d + 1, nb));
    CK(cudaStreamCreate(&st[d])); CK(cudaEventCreate(&e0[d])); CK(cudaEventCreate(&e1[d]));
  }
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  printf("{");
  for (int kind = 0; kind < 5; ++kind) {      // 0 st, 1..4 bulk 1K/4K/16K/32K
    for (int both = 0; both < 2; ++both) {
      float best = 1e30f;
      for (int rep = 0; rep < 4; ++rep) {
        for (int d = 0; d < (both ? 2 : 1); ++d) {
          CK(cudaSetDevice(d));
          CK(cudaEventRecord(e0[d], st[d]));
          if (kind == 0) st_copy<<<sms * 8, 256, 0, st[d]>>>((const uint4 *)a[d], (uint4 *)b[1 - d], nb / 16);
          else if (kind == 1) bulk_copy<1024><<<sms * 8, 32, 2 * 1024, st[d]>>>(a[d], b[1 - d], nb / 1024);
          else if (kind == 2) bulk_copy<4096><<<sms * 8, 32, 2 * 4096, st[d]>>>(a[d], b[1 - d], nb / 4096);
          else if (kind == 3) bulk_copy<16384><<<sms * 4, 32, 2 * 16384, st[d]>>>(a[d], b[1 - d], nb / 16384);
          else { cudaFuncSetAttribute(bulk_copy<32768>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536); bulk_copy<32768><<<sms * 2, 32, 2 * 32768, st[d]>>>(a[d], b[1 - d], nb / 32768); }
          CK(cudaGetLastError());
          CK(cudaEventRecord(e1[d], st[d]));
        }
        for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0[0], e1[0]));
        if (rep > 0 && ms < best) best = ms;
      }
      const char *names[5] = {"st_v4", "bulk_1K", "bulk_4K", "bulk_16K", "bulk_32K"};
      printf("%s\"%s_%s_GBps\": %.1f", (kind || both) ? ", " : "", names[kind], both ? "bidir" : "onedir", nb / best / 1e6);
    }
  }
  // sanity: the last copy arrived
  uint8_t h = 0;
  CK(cudaSetDevice(1));
  CK(cudaMemcpy(&h, b[1] + nb - 1, 1, cudaMemcpyDeviceToHost));
  printf(", \"last_byte_ok\": %s}\n", h == 1 ? "true" : "false");
  return 0;
}
