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

/* Number of kernels this library has launched since load (all entry points). */
int64_t arcq_launch_count(void);

const char *arcq_last_error(void);
const char *arcq_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ARCQ_H_ */
