/*
 * arcq.h -- C ABI of the B200-native NVFP4 + ARC hot path (libarcq_b200.so).
 *
 * This is the drop-in boundary: every entry point below is what the reference's
 * pybind11 module `agemm` (kernels/src/bindings.cpp, /root/reference) would bind
 * for its four hot functions.  Plain pointers and sizes only; all data pointers
 * are DEVICE pointers unless the name ends in _host; `stream` is a cudaStream_t
 * passed as void* (NULL = legacy default stream, which is what the reference
 * launches on, reorder.cu:344 / nvfp4.cu:124).
 *
 * All functions return 0 on success and a negative ARCQ_E* code otherwise; they
 * never throw, never allocate output buffers, never synchronise the device.
 * arcq_last_error() returns a thread-local human-readable message for the last
 * non-zero status.  There is NO CPU fallback: without a CUDA device every
 * compute entry point fails with ARCQ_ECUDA.
 */
#ifndef ARCQ_H_
#define ARCQ_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ARCQ_OK 0
#define ARCQ_EINVAL (-1)  /* bad shape / argument (reference: std::runtime_error, bindings.cpp:157-160) */
#define ARCQ_ECUDA (-2)   /* CUDA runtime / driver error */
#define ARCQ_ENOSUP (-3)  /* shape outside what the kernels support */

/* Which residual/interleave flavour the reference uses for a given KQ
 * (bindings.cpp:141-160): 0 = 16-wide (reorder_x_kernel), 1 = 32-wide
 * (reorder32_x_kernel / down32_x_kernel), -1 = KQ not in the reference's table
 * (the 16-wide flavour is then canonical). */
#define ARCQ_FLAVOUR_16 0
#define ARCQ_FLAVOUR_32 1
int arcq_flavour_for_kq(int KQ);

/* Scale-factor buffer size in bytes for `rows` x K (K = KQ + KE):
 * (rows/128 + 1) * 128 * K/16 -- replaces get_sfa_buffer_size_in_bytes /
 * get_sfb_buffer_size_in_bytes, bindings.cpp:83-95. */
int64_t arcq_sf_bytes(int64_t rows, int64_t K);

/* Byte offset of the scale factor of (row, 16-group g): the tcgen05 "128x4"
 * swizzle = CUTLASS Sm1xxBlockScaledConfig<16> (reorder.cu:140-143,343). */
int64_t arcq_sf_offset(int64_t row, int64_t g, int64_t K);

/* reorder_quantize_x -- replaces bindings.cpp:122-163 (+ reorder.cu:68-203,
 * :380-555, down.cu:71-233 and the torch index_select of bindings.cpp:29-38).
 *   x    bf16 [M, KQ] row-major        idx  int16 [KQ] permutation
 *   qx   u8   [M, (KQ+KE)/2]           sfx  u8 [arcq_sf_bytes(M, KQ+KE)]
 * flavour: ARCQ_FLAVOUR_16 / _32, or -1 = pick by the reference's KQ table. */
int arcq_quantize_x(const void *x, int64_t M, int KQ, const int16_t *idx, int KE, int flavour,
                    void *qx, void *sfx, void *stream);

/* reorder_quantize_w -- replaces bindings.cpp:170-210 (+ reorder.cu:210-330,
 * :562-696, down.cu:240-361).  Same shapes with N rows; ARC slots hold a copy
 * of the primary codes and scale. */
int arcq_quantize_w(const void *w, int64_t N, int KQ, const int16_t *idx, int KE, int flavour,
                    void *qw, void *sfw, void *stream);

/* rmsnorm_quantize_x -- replaces bindings.cpp:216-254 (+ rmsnorm.cu:68-255).
 * Always the 16-wide flavour (bindings.cpp:243).  w: bf16 [H]. */
int arcq_rmsnorm_quantize_x(const void *x, const void *w, float eps, int64_t M, int H,
                            const int16_t *idx, int KE, void *qx, void *sfx, void *stream);

/* matmul -- replaces bindings.cpp:99-120 + nvfp4.cu:35-132.
 *   a u8 [M, K/2], b u8 [N, K/2] (both K-major packed e2m1, low nibble = even k)
 *   sfa / sfb: E4M3 bytes in the layout of arcq_sf_offset
 *   d bf16 [M, N] row-major = alpha * (a.sfa) (b.sfb)^T, fp32 accumulate.
 * alpha_dev (nullable): device pointer to one float; when non-NULL it is read by
 * the kernel instead of `alpha` (removes the reference's D2H sync on `scale`,
 * bindings.cpp:104). */
int arcq_gemm_nvfp4(const void *a, const void *b, const void *sfa, const void *sfb, int64_t M,
                    int64_t N, int64_t K, float alpha, const float *alpha_dev, void *d,
                    void *stream);

/* matmul with a fused epilogue and per-column scales -- the same kernels as arcq_gemm_nvfp4 (which is this call with
 * col_scale_dev = NULL, epilogue = NONE, ldd = N).  SURVEY.md §8 f2/f3; reference call sites that share one quantised
 * input: model/qLlamaLayer.py:301-303 (q/k/v), :432 (gate/up), benchmarks/modeling_arc.py:109-111,165-167.
 *   col_scale_dev (nullable) fp32 [N]: column n is scaled by RN(alpha * col_scale_dev[n]) -- several weights, each with
 *       its own per-tensor scale (qLinearLayer.py:26), concatenated along N (rows a multiple of 128) run as ONE launch;
 *   epilogue ARCQ_EPILOGUE_SWIGLU: b is the gate and up weights interleaved in 128-row blocks (rows [256j, 256j+128) =
 *       gate rows [128j, +128), rows [256j+128, 256j+256) = up rows [128j, +128)); d is bf16 [M, N/2] =
 *       bf16(silu(bf16 gate)) * bf16 up, rounded exactly like torch's `F.silu(g) * u` on bf16 tensors
 *       (modeling_arc.py:110); N % 256 == 0;
 *   ldd: row stride of d in elements (0 = dense). */
#define ARCQ_EPILOGUE_NONE 0
#define ARCQ_EPILOGUE_SWIGLU 1
int arcq_gemm_nvfp4_ex(const void *a, const void *b, const void *sfa, const void *sfb, int64_t M,
                       int64_t N, int64_t K, float alpha, const float *alpha_dev,
                       const float *col_scale_dev, int epilogue, void *d, int64_t ldd, void *stream);

/* arcq_gemm_nvfp4_ex plus a per-column bias: col_bias_dev (nullable) bf16 [N], 16-byte aligned; the epilogue computes
 * bf16(float(y) + float(bias[n])) on the ALREADY ROUNDED bf16 output y -- exactly torch's `y = y + self.bias` of the
 * reference's linear layer (model/qLinearLayer.py:74-75; Qwen's q/k/v projections) without its extra pass over [M, N].
 * Not combinable with ARCQ_EPILOGUE_SWIGLU. */
int arcq_gemm_nvfp4_ex2(const void *a, const void *b, const void *sfa, const void *sfb, int64_t M,
                        int64_t N, int64_t K, float alpha, const float *alpha_dev,
                        const float *col_scale_dev, const void *col_bias_dev, int epilogue, void *d,
                        int64_t ldd, void *stream);

/* Fused per-tensor pre-scale of the reference callers (NVFP4_reorder_quantize_x, model/qLlamaLayer.py:73-77,
 * qQwenLayer.py:72-75, benchmarks/modeling_arc.py:58-61: `scale = max|x| / 2688; agemm.reorder_quantize_x(x / scale)`).
 *   arcq_absmax: amax_out[0] = max |x| over n bf16 elements, as fp32 (one memset + one kernel, stream ordered);
 *   arcq_quantize_x_scaled: arcq_quantize_x applied to bf16(x / scale), scale = amax / 2688, computed in-kernel with
 *       torch-identical rounding (fp32 IEEE division, bf16 RNE); scale_out[0] (nullable) receives the fp32 scale for the
 *       GEMM alpha.  prescale_mode names the torch expression being replaced:
 *         ARCQ_PRESCALE_FP32          divisor used as fp32: `(x / scale.reshape(1)).to(bf16)`, CPU tensors, Python floats;
 *         ARCQ_PRESCALE_BF16_DIVISOR  divisor rounded to bf16 first: `x / scale` with x a bf16 CUDA tensor and scale a
 *                                     0-dim CUDA fp32 tensor (TensorIterator casts the operand to the common dtype) --
 *                                     the literal expression of qLlamaLayer.py:76.
 * Bytes equal "torch pre-scale, then arcq_quantize_x".  ARCQ_ENOSUP for shapes only the fallback kernel takes. */
#define ARCQ_PRESCALE_FP32 0
#define ARCQ_PRESCALE_BF16_DIVISOR 1
int arcq_absmax(const void *x, int64_t n, float *amax_out, void *stream);
int arcq_quantize_x_scaled(const void *x, int64_t M, int KQ, const int16_t *idx, int KE, int flavour,
                           const float *amax_dev, int prescale_mode, void *qx, void *sfx, float *scale_out,
                           void *stream);

/* Conflict-free shared-memory placement for one reorder_index ("plan"; csrc/plan.cu) -- computed ONCE per layer on the
 * host (index arithmetic only, O(KQ log 32)), uploaded by the caller, and passed to arcq_quantize_x_ex.  It removes the
 * bank conflicts of the permuted gather (reorder.cu:115-118) for any permutation.  A plan is only ever an optimisation:
 * the kernel requires it to be a bijection (arcq_plan_build guarantees it) but derives its gather offsets from
 * plan[reorder_index[k]] itself, so a plan built for a DIFFERENT index still gives correct bytes.
 *   arcq_plan_entries(KQ)  number of uint16 entries (6 * KQ: per 8-, 4- and 2-row tile the position of every channel and
 *                          the same positions in the order the transposing lanes consume them);
 *   arcq_plan_build        host pointers; ARCQ_EINVAL if idx is not a permutation of [0, KQ);
 *   arcq_plan_max_conflict worst bank-slot multiplicity of any wavefront under a plan for tile height R (1 = none). */
int64_t arcq_plan_entries(int KQ);
int arcq_plan_build(const int16_t *idx_host, int KQ, uint16_t *plan_host);
int arcq_plan_max_conflict(const int16_t *idx_host, int KQ, int R, const uint16_t *plan_host);

/* arcq_quantize_x with the optional extras: plan_dev (nullable, device copy of a plan) and the fused pre-scale
 * (amax_dev nullable; prescale_mode / scale_out as in arcq_quantize_x_scaled). */
int arcq_quantize_x_ex(const void *x, int64_t M, int KQ, const int16_t *idx, int KE, int flavour,
                       const uint16_t *plan_dev, const float *amax_dev, int prescale_mode, void *qx,
                       void *sfx, float *scale_out, void *stream);

/* The residual add in front of the norm, fused: `hidden_states = residual + hidden_states` followed by
 * the rmsnorm_quantize_x of the next norm (benchmarks/modeling_arc.py:294-301, model/qLlamaLayer.py:162-166,295-300):
 *   sum_out bf16 [M, H] = bf16(x + res)   (fp32 add, RNE -- what torch's bf16 add computes; may alias neither input)
 *   (qx, sfx)           = arcq_rmsnorm_quantize_x(sum_out, w, eps, ...), byte for byte.
 * One pass: x and res are read once, the sum is written once; the separate torch add (2 reads + 1 write) and the norm's
 * re-read of its result disappear.  H, KE multiples of 64, 2048 <= H <= 8192, else ARCQ_ENOSUP. */
int arcq_add_rmsnorm_quantize_x(const void *x, const void *res, const void *w, float eps, int64_t M, int H,
                                const int16_t *idx, int KE, void *sum_out, void *qx, void *sfx, void *stream);

/* One linear (arcq_quantize_x followed by arcq_gemm_nvfp4: y bf16 [M, N] = alpha * Q(x[:, idx]) . (qw.sfw)^T) for callers
 * that cannot allocate device memory -- code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc.  NO COUNTERPART IN THE REFERENCE.  The quantised activation
 * (codes + scale factors) that the two kernels hand to each other lives in a file-scope `static __device__` array of
 * ARCQ_NOALLOC_SCRATCH_BYTES inside the library (csrc/noalloc.cu; every GPU that loads the library holds it until
 * exit) instead of caller buffers; the call makes no cudaMalloc / cudaFree / allocating library call.
 *   arcq_linear_noalloc_bytes(M, KQ, KE)  scratch bytes such a call needs (-1 on a bad shape);
 *   arcq_linear_noalloc                   ARCQ_EINVAL with a message naming both numbers when that exceeds
 *                                         ARCQ_NOALLOC_SCRATCH_BYTES (16 MiB: M <= 6 846 at KQ+KE = 4 352,
 *                                         M <= 2 003 at KQ+KE = 14 848).
 * Results are byte-identical to the two separate calls (same kernels, same dispatch).  ALL calls on one GPU share the
 * one array: calls that may overlap on the GPU (different streams without an ordering between them) race on it; calls
 * on one stream, or otherwise ordered, are safe.  The array keeps the last call's contents; nothing in it has to start
 * at zero (every byte the GEMM's stored rows depend on is rewritten by the quantiser of the same call). */
#define ARCQ_NOALLOC_SCRATCH_BYTES (16 * 1024 * 1024)
int64_t arcq_linear_noalloc_bytes(int64_t M, int KQ, int KE);
int arcq_linear_noalloc(const void *x, int64_t M, int KQ, const int16_t *idx, int KE, int flavour,
                        const void *qw, const void *sfw, int64_t N, float alpha, const float *alpha_dev,
                        void *y, void *stream);

/* ---- Tensor-parallel exchange over peer memory (NVLink / NVSwitch P2P), SURVEY.md §8 f4 ---------------------------------
 * NO COUNTERPART IN THE REFERENCE (it has no distributed code): these replace this repo's own NCCL data plane
 * for the sharding the north star names (o_proj / down_proj split K over ranks, tokens reduce-scattered; the inputs of
 * the column-parallel linears all-gathered as NVFP4 codes + scale blocks).  csrc/peer.cu, csrc/gemm_pair.cu.
 *
 * Memory: arcq_peer_alloc = zeroed cudaMalloc memory (exportable; a caching allocator's sub-allocations are not);
 * arcq_peer_export / _open / _close move it between the processes of one node as a 64-byte CUDA IPC handle (the caller
 * carries the bytes, e.g. torch.distributed.all_gather_object); arcq_peer_enable is the single-process-several-GPUs
 * equivalent (cudaDeviceEnablePeerAccess).  Every `peer_*` argument below is a HOST array of `world` DEVICE pointers,
 * entry r = rank r's buffer as mapped into THIS process (entry `rank` = the local buffer).
 *
 * Flag block: ARCQ_PEER_FLAG_BYTES of zeroed peer memory per rank, shared by all calls of one communicator; `channel`
 * (< ARCQ_PEER_CHANNELS) separates barriers that may be in flight at the same time (one per stream is enough).  The
 * barrier epoch lives in the flag block, not in a host counter, so launches replay from CUDA graphs.  Waits are bounded
 * (4 s): a rank that never arrives makes the kernel finish with a sticky error that arcq_peer_status reports
 * (ARCQ_ECUDA) -- it does not hang the GPU.
 *
 *   arcq_gemm_nvfp4_scatter  arcq_gemm_nvfp4_ex (EPILOGUE_NONE) whose epilogue stores the 128-row block starting at row g
 *       into peer_recv[g / rows_per_rank] at row (rank * rows_per_rank + g % rows_per_rank): rank o's receive buffer is
 *       bf16 [world][rows_per_rank][ldd], slot s = rank s's partial of the rows o owns.  The stores leave tile by tile
 *       under the math of the next tile; there is no partial-sum buffer in local HBM and no collective kernel.
 *       rows_per_rank % 128 == 0, rows_per_rank * world >= M.
 *   arcq_peer_reduce         cross-GPU barrier (every rank's scatter GEMM, launched before it on the same stream, has
 *       finished), then out[r, c] = bf16(sum_s fp32(recv_local[s][r][c])), s ascending: deterministic, one rounding
 *       (NCCL's ring rounds to bf16 per hop).  cols % 8 == 0.
 *   arcq_peer_allgather      copies up to two local segments (bytes1 = 0: one) to byte offset dst_off of every rank's
 *       destination buffer, then the same barrier: when the kernel completes, every rank's segments have landed here.
 *       Sizes / offsets multiples of 16.  (codes: offset rank * Mx * K/2; scale blocks: rank * Mx * K/16.)
 * Buffer reuse is the caller's contract: a receive / gather buffer may be overwritten by a peer as soon as that peer has
 * passed a barrier that this rank entered AFTER its last read of the buffer (in a layer that alternates these calls this
 * always holds; otherwise alternate two buffers). */
#define ARCQ_MAX_PEERS 8
#define ARCQ_PEER_HANDLE_BYTES 64
#define ARCQ_PEER_FLAG_BYTES 4096
#define ARCQ_PEER_CHANNELS 32
int arcq_peer_alloc(int64_t bytes, void **ptr_out);
int arcq_peer_free(void *ptr);
int arcq_peer_export(void *ptr, void *handle_out);
int arcq_peer_open(const void *handle, void **ptr_out);
int arcq_peer_close(void *ptr);
int arcq_peer_enable(int peer_device);
int arcq_peer_status(const void *flags_local, void *stream);
int arcq_gemm_nvfp4_scatter(const void *a, const void *b, const void *sfa, const void *sfb, int64_t M, int64_t N,
                            int64_t K, float alpha, const float *alpha_dev, const float *col_scale_dev,
                            void *const *peer_recv, int world, int rank, int64_t rows_per_rank, int64_t ldd,
                            void *stream);
int arcq_peer_reduce(const void *recv_local, void *const *peer_flags, int world, int rank, int channel, int64_t rows,
                     int64_t cols, int64_t ld, void *out, int64_t ldo, void *stream);
int arcq_peer_allgather(const void *src0, int64_t bytes0, void *const *peer_dst0, int64_t dst_off0, const void *src1,
                        int64_t bytes1, void *const *peer_dst1, int64_t dst_off1, void *const *peer_flags, int world,
                        int rank, int channel, void *stream);

/* name = value override of a shape-driven kernel choice ("" = back to the default).  Every choice computes the same
 * result; the parity tests force each tile organisation through this.  Options: "quant_tile" = "R,threads,S,RT,CTAs/SM",
 * "quant_family" = 3 | 1, "gemm_pair" = 0 | 1, "gemm_bn" = 128 | 256, "gemm_splitk" = 0 | 2 | 4 | 8, "gemm_raster" = 0 | g (tile order of the
 * cta_group::2 kernel: row tiles fastest, or groups of g row tiles; default by shape).  The product
 * library reads NO environment variables and has no switch that skips work. */
int arcq_set_option(const char *name, const char *value);

/* Number of kernels this library has launched since load (all entry points). */
int64_t arcq_launch_count(void);

const char *arcq_last_error(void);
const char *arcq_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ARCQ_H_ */
