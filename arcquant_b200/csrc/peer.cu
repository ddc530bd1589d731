// peer.cu -- the exchange steps of the tensor-parallel path over PEER MEMORY (NVLink 5 / NVSwitch P2P loads and stores),
// SURVEY.md §8 f4.  The reference has no distributed code, so nothing here replaces a reference line; what
// it replaces is this repo's own NCCL data plane (bench.py `Layer`):
//
//   * the row-parallel GEMM (o_proj, down_proj: K split over ranks) no longer writes a bf16 partial that NCCL then
//     reduce-scatters: its EPILOGUE stores every 128-row block straight into the receive slot of the rank that owns those
//     tokens (gemm_pair.cu, `rs_rows`), tile by tile under the math of the next tile, and `arcq_peer_reduce` -- one
//     cross-GPU barrier + one pass over the W local slots, summed in fp32 in rank order (deterministic, one rounding) --
//     finishes the reduce-scatter.  No collective kernel, no copy of the partial through HBM;
//   * `arcq_peer_allgather` pushes this rank's NVFP4 codes + scale blocks into every rank's gathered buffer with plain
//     16-byte stores and ends with the same barrier (replaces two ncclAllGather launches per linear).
//
// Barrier protocol (flag block = ARCQ_PEER_FLAG_BYTES of zeroed, peer-mapped memory per rank; 32-bit words):
//   word  c * 16 + s      (c < 32 channels, s < 8)  arrival counter of rank s on channel c, written by rank s with
//                                                   st.release.sys, polled locally with ld.acquire.sys
//   word  512 + c         epoch of channel c (how many barriers this rank has completed on it; local use only)
//   word  544 + c         ticket counter of the running launch on channel c (local use only)
//   word  1023            sticky error: 1 = a wait timed out (peer never arrived); read by arcq_peer_status
// The epoch lives on the DEVICE, so a launch carries no host-side counter and can be replayed from a CUDA graph.  A kernel
// reads e = epoch + 1 at its start; only the last block to finish stores the new epoch, so every block of a launch sees the
// same e however the grid is scheduled.  Waits are bounded (kSpinLimitNs): a peer that never arrives sets the error word
// and lets the kernel finish with wrong data instead of hanging the GPU.
#include <cuda.h>
#include <string.h>

#include "common.cuh"

namespace arcq {
namespace {

constexpr int kFlagStride = 16;           // words per channel in the arrival area
constexpr int kEpochWord = 512, kTicketWord = 544, kErrorWord = 1023;
constexpr unsigned long long kSpinLimitNs = 4000000000ull;   // 4 s

struct PeerPtrs { void *p[ARCQ_MAX_PEERS]; };

__device__ __forceinline__ void st_release_sys(uint32_t *addr, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *addr) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_volatile(const uint32_t *addr) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// thread `src` (< world) of the calling block waits until rank src has arrived at epoch e on this channel
__device__ __forceinline__ void wait_arrival(uint32_t *my_flags, int chan, int src, uint32_t e) {
  const uint32_t *f = my_flags + chan * kFlagStride + src;
  if ((int32_t)(ld_acquire_sys(f) - e) >= 0) return;
  const unsigned long long t0 = globaltimer_ns();
  while ((int32_t)(ld_acquire_sys(f) - e) < 0) {
    __nanosleep(64);
    if (globaltimer_ns() - t0 > kSpinLimitNs) { atomicExch(my_flags + kErrorWord, 1u); break; }
  }
}

// the last block of the launch publishes the new epoch (and re-arms the ticket counter)
__device__ __forceinline__ bool take_ticket(uint32_t *my_flags, int chan) {
  __threadfence();
  const uint32_t t = atomicAdd(my_flags + kTicketWord + chan, 1u);
  return t == gridDim.x - 1;
}

// ---- reduce-scatter, second half: barrier, then out = bf16(sum over slots, fp32, rank order) ----------------------
// recv: this rank's receive buffer [world][rows][ld] bf16, slot s written by rank s's GEMM epilogue.
__global__ void __launch_bounds__(256)
peer_reduce_kernel(const __nv_bfloat16 *__restrict__ recv, const __grid_constant__ PeerPtrs flags, int world, int rank, int chan, int64_t rows, int cols,
                   int64_t ld, __nv_bfloat16 *__restrict__ out, int64_t ldo) {
  uint32_t *my_flags = (uint32_t *)flags.p[rank];
  const uint32_t e = ld_volatile(my_flags + kEpochWord + chan) + 1u;
  // every peer's partial was stored by the kernel before this one on its stream: announce it, then wait for the others
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();
    st_release_sys((uint32_t *)flags.p[threadIdx.x] + chan * kFlagStride + rank, e);
  }
  if (threadIdx.x < world) wait_arrival(my_flags, chan, threadIdx.x, e);
  __syncthreads();
  const int vec_per_row = cols / 8;
  const int64_t total = rows * vec_per_row;
  const int64_t slot = rows * ld;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / vec_per_row;
    const int c = (int)(i - r * vec_per_row) * 8;
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
    for (int s = 0; s < world; ++s) {
      // L2-only loads: the lines were written by peers over NVLink, nothing of them may come from a stale L1 line
      const uint4 v = __ldcg(reinterpret_cast<const uint4 *>(recv + s * slot + r * ld + c));
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        acc[2 * k] += bf16lo_to_f32(w[k]);
        acc[2 * k + 1] += bf16hi_to_f32(w[k]);
      }
    }
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      __nv_bfloat162 h2 = __floats2bfloat162_rn(acc[2 * k], acc[2 * k + 1]);
      o[k] = *reinterpret_cast<uint32_t *>(&h2);
    }
    *reinterpret_cast<uint4 *>(out + r * ldo + c) = make_uint4(o[0], o[1], o[2], o[3]);
  }
  __syncthreads();
  if (threadIdx.x == 0 && take_ticket(my_flags, chan)) {
    my_flags[kTicketWord + chan] = 0u;
    my_flags[kEpochWord + chan] = e;
    __threadfence();
  }
}

// ---- all-gather by stores: every block copies its share of up to two local segments to all ranks, then barrier ----
struct Segment { const uint4 *src; int64_t n16; int64_t dst_off16; };   // n16 / dst_off16 in 16-byte units

__global__ void __launch_bounds__(256)
peer_allgather_kernel(const __grid_constant__ Segment s0, const __grid_constant__ Segment s1, const __grid_constant__ PeerPtrs dst0,
                      const __grid_constant__ PeerPtrs dst1, const __grid_constant__ PeerPtrs flags, int world, int rank, int chan) {
  uint32_t *my_flags = (uint32_t *)flags.p[rank];
  const uint32_t e = ld_volatile(my_flags + kEpochWord + chan) + 1u;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x, tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll 1
  for (int seg = 0; seg < 2; ++seg) {
    const Segment &s = seg ? s1 : s0;
    const PeerPtrs &d = seg ? dst1 : dst0;
    for (int64_t i = tid; i < s.n16; i += stride) {
      const uint4 v = __ldcs(s.src + i);
      for (int k = 0; k < world; ++k) {
        const int pr = (rank + k) % world;   // start with the own copy, then walk the ring so peers are hit evenly
        reinterpret_cast<uint4 *>(d.p[pr])[s.dst_off16 + i] = v;
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const bool last = take_ticket(my_flags, chan);
    if (last) {
      // every block's stores are ordered before its ticket; this block has seen all tickets: announce, then wait
      __threadfence_system();
      for (int pr = 0; pr < world; ++pr) st_release_sys((uint32_t *)flags.p[pr] + chan * kFlagStride + rank, e);
      for (int src = 0; src < world; ++src) wait_arrival(my_flags, chan, src, e);
      my_flags[kTicketWord + chan] = 0u;
      my_flags[kEpochWord + chan] = e;
      __threadfence();
    }
  }
}

int check_peers(const char *who, void *const *ptrs, int world, int rank) {
  if (!ptrs || world < 1 || world > ARCQ_MAX_PEERS || rank < 0 || rank >= world) {
    set_error("%s: world=%d rank=%d out of range (at most %d peers)", who, world, rank, ARCQ_MAX_PEERS);
    return ARCQ_EINVAL;
  }
  for (int i = 0; i < world; ++i)
    if (!ptrs[i] || ((uintptr_t)ptrs[i] & 15)) { set_error("%s: peer pointer %d is null or not 16-byte aligned", who, i); return ARCQ_EINVAL; }
  return ARCQ_OK;
}

}  // namespace

int launch_gemm_pair_scatter(const void *a, const void *b, const void *sfa, const void *sfb, int M, int N, int K, float alpha,
                             const float *alpha_dev, const float *col_scale, void *const *peer_recv, int world, int rank,
                             int rows_per_rank, int64_t ldd, cudaStream_t stream);

}  // namespace arcq

using namespace arcq;

// ---- peer-mappable memory: plain cudaMalloc (exportable with CUDA IPC; torch's caching allocator sub-allocates) --------
extern "C" int arcq_peer_alloc(int64_t bytes, void **ptr_out) {
  if (!ptr_out || bytes <= 0) { set_error("arcq_peer_alloc: bad arguments"); return ARCQ_EINVAL; }
  void *p = nullptr;
  ARCQ_CUDA_TRY(cudaMalloc(&p, (size_t)bytes));
  cudaError_t e = cudaMemset(p, 0, (size_t)bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { cudaFree(p); set_error("arcq_peer_alloc: clearing failed: %s", cudaGetErrorString(e)); return ARCQ_ECUDA; }
  *ptr_out = p;
  return ARCQ_OK;
}

extern "C" int arcq_peer_free(void *ptr) {
  if (ptr) ARCQ_CUDA_TRY(cudaFree(ptr));
  return ARCQ_OK;
}

extern "C" int arcq_peer_export(void *ptr, void *handle_out) {
  static_assert(sizeof(cudaIpcMemHandle_t) == ARCQ_PEER_HANDLE_BYTES, "IPC handle size");
  if (!ptr || !handle_out) { set_error("arcq_peer_export: null argument"); return ARCQ_EINVAL; }
  cudaIpcMemHandle_t h;
  ARCQ_CUDA_TRY(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle_out, &h, sizeof(h));
  return ARCQ_OK;
}

extern "C" int arcq_peer_open(const void *handle, void **ptr_out) {
  if (!handle || !ptr_out) { set_error("arcq_peer_open: null argument"); return ARCQ_EINVAL; }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void *p = nullptr;
  ARCQ_CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr_out = p;
  return ARCQ_OK;
}

extern "C" int arcq_peer_close(void *ptr) {
  if (ptr) ARCQ_CUDA_TRY(cudaIpcCloseMemHandle(ptr));
  return ARCQ_OK;
}

// one process driving several GPUs: map `peer_device`'s memory into the current device (idempotent)
extern "C" int arcq_peer_enable(int peer_device) {
  const int dev = current_device();
  if (dev < 0) return ARCQ_ECUDA;
  if (peer_device == dev) return ARCQ_OK;
  int can = 0;
  ARCQ_CUDA_TRY(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  if (!can) { set_error("arcq_peer_enable: device %d cannot map the memory of device %d", dev, peer_device); return ARCQ_ENOSUP; }
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return ARCQ_OK; }
  ARCQ_CUDA_TRY(e);
  return ARCQ_OK;
}

extern "C" int arcq_peer_status(const void *flags_local, void *stream) {
  if (!flags_local) { set_error("arcq_peer_status: null flag block"); return ARCQ_EINVAL; }
  uint32_t err = 0;
  ARCQ_CUDA_TRY(cudaMemcpyAsync(&err, (const uint32_t *)flags_local + kErrorWord, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  ARCQ_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  if (err) { set_error("a peer barrier timed out after %llu ms: some rank never arrived (results of that step are invalid)", kSpinLimitNs / 1000000ull); return ARCQ_ECUDA; }
  return ARCQ_OK;
}

extern "C" int arcq_gemm_nvfp4_scatter(const void *a, const void *b, const void *sfa, const void *sfb, int64_t M, int64_t N,
                                       int64_t K, float alpha, const float *alpha_dev, const float *col_scale_dev,
                                       void *const *peer_recv, int world, int rank, int64_t rows_per_rank, int64_t ldd,
                                       void *stream) {
  if (!a || !b || !sfa || !sfb) { set_error("null pointer argument"); return ARCQ_EINVAL; }
  if (int rc = check_peers("arcq_gemm_nvfp4_scatter", peer_recv, world, rank)) return rc;
  if (M <= 0 || N <= 0 || M > 0x7fffff00 || N > 0x7fffff00) { set_error("M/N out of range"); return ARCQ_EINVAL; }
  if (K <= 0 || K % 64 || K > (1 << 20)) { set_error("K=%lld must be a positive multiple of 64", (long long)K); return ARCQ_EINVAL; }
  if (((uintptr_t)a | (uintptr_t)b | (uintptr_t)sfa | (uintptr_t)sfb) & 15) { set_error("pointers must be 16-byte aligned"); return ARCQ_EINVAL; }
  if (rows_per_rank <= 0 || rows_per_rank % 128 || rows_per_rank * world < M) {
    set_error("arcq_gemm_nvfp4_scatter: rows_per_rank=%lld must be a multiple of 128 with rows_per_rank * world >= M=%lld",
              (long long)rows_per_rank, (long long)M);
    return ARCQ_EINVAL;
  }
  if (ldd == 0) ldd = N;
  if (ldd < N || ldd % 8) { set_error("ldd=%lld must be >= N and a multiple of 8", (long long)ldd); return ARCQ_EINVAL; }
  return launch_gemm_pair_scatter(a, b, sfa, sfb, (int)M, (int)N, (int)K, alpha, alpha_dev, col_scale_dev, peer_recv, world, rank,
                                  (int)rows_per_rank, ldd, (cudaStream_t)stream);
}

extern "C" int arcq_peer_reduce(const void *recv_local, void *const *peer_flags, int world, int rank, int channel, int64_t rows,
                                int64_t cols, int64_t ld, void *out, int64_t ldo, void *stream) {
  if (!recv_local || !out) { set_error("arcq_peer_reduce: null pointer argument"); return ARCQ_EINVAL; }
  if (int rc = check_peers("arcq_peer_reduce", peer_flags, world, rank)) return rc;
  if (channel < 0 || channel >= ARCQ_PEER_CHANNELS) { set_error("arcq_peer_reduce: channel %d out of range", channel); return ARCQ_EINVAL; }
  if (ld == 0) ld = cols;
  if (ldo == 0) ldo = cols;
  if (rows < 0 || cols <= 0 || cols % 8 || ld % 8 || ldo % 8 || ld < cols || ldo < cols || cols > 0x7fffff00 ||
      (((uintptr_t)recv_local | (uintptr_t)out) & 15)) {
    set_error("arcq_peer_reduce: rows=%lld cols=%lld ld=%lld ldo=%lld: cols and strides must be multiples of 8, buffers 16-byte aligned",
              (long long)rows, (long long)cols, (long long)ld, (long long)ldo);
    return ARCQ_EINVAL;
  }
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  PeerPtrs f = {};
  for (int i = 0; i < world; ++i) f.p[i] = peer_flags[i];
  const int64_t vecs = rows * (cols / 8);
  int grid = (int)((vecs + 255) / 256 < (int64_t)sms * 8 ? (vecs + 255) / 256 : (int64_t)sms * 8);
  if (grid < 1) grid = 1;
  peer_reduce_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16 *)recv_local, f, world, rank, channel, rows, (int)cols, ld,
                                                            (__nv_bfloat16 *)out, ldo);
  ARCQ_CUDA_TRY(cudaGetLastError());
  count_launch();
  return ARCQ_OK;
}

extern "C" int arcq_peer_allgather(const void *src0, int64_t bytes0, void *const *peer_dst0, int64_t dst_off0, const void *src1,
                                   int64_t bytes1, void *const *peer_dst1, int64_t dst_off1, void *const *peer_flags, int world,
                                   int rank, int channel, void *stream) {
  if (int rc = check_peers("arcq_peer_allgather", peer_flags, world, rank)) return rc;
  if (channel < 0 || channel >= ARCQ_PEER_CHANNELS) { set_error("arcq_peer_allgather: channel %d out of range", channel); return ARCQ_EINVAL; }
  if (!src0 || bytes0 <= 0) { set_error("arcq_peer_allgather: the first segment is empty"); return ARCQ_EINVAL; }
  if (int rc = check_peers("arcq_peer_allgather", peer_dst0, world, rank)) return rc;
  if (bytes1 > 0) {
    if (!src1) { set_error("arcq_peer_allgather: null second segment"); return ARCQ_EINVAL; }
    if (int rc = check_peers("arcq_peer_allgather", peer_dst1, world, rank)) return rc;
  }
  if (((uintptr_t)src0 | (uintptr_t)src1 | (uint64_t)bytes0 | (uint64_t)(bytes1 > 0 ? bytes1 : 0) | (uint64_t)dst_off0 | (uint64_t)dst_off1) & 15 ||
      dst_off0 < 0 || dst_off1 < 0) {
    set_error("arcq_peer_allgather: sources, sizes and destination offsets must be non-negative multiples of 16 bytes");
    return ARCQ_EINVAL;
  }
  const int sms = sm_count();
  if (sms <= 0) return ARCQ_ECUDA;
  PeerPtrs f = {}, d0 = {}, d1 = {};
  for (int i = 0; i < world; ++i) { f.p[i] = peer_flags[i]; d0.p[i] = peer_dst0[i]; d1.p[i] = bytes1 > 0 ? peer_dst1[i] : nullptr; }
  const Segment s0{(const uint4 *)src0, bytes0 / 16, dst_off0 / 16}, s1{(const uint4 *)src1, bytes1 > 0 ? bytes1 / 16 : 0, dst_off1 / 16};
  const int64_t n16 = s0.n16 + s1.n16;
  int grid = (int)((n16 + 255) / 256 < (int64_t)sms * 2 ? (n16 + 255) / 256 : (int64_t)sms * 2);
  if (grid < 1) grid = 1;
  peer_allgather_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(s0, s1, d0, d1, f, world, rank, channel);
  ARCQ_CUDA_TRY(cudaGetLastError());
  count_launch();
  return ARCQ_OK;
}
