// quant.cuh -- parameters shared by the quantise kernel families (quant.cu, quant_tma.cu).
#pragma once
#include "common.cuh"

namespace arcq {

struct QuantParams {
  const uint16_t *x;     // bf16 [M, KQ]
  const int16_t *idx;    // [KQ]
  const uint16_t *w;     // rmsnorm weight bf16 [KQ] (RMS only)
  uint8_t *q;            // [M, K/2]
  uint8_t *sf;           // swizzled scale bytes
  const uint16_t *plan;  // nullable: conflict-free chunk positions (plan.cu); the launcher narrows it to one tile height's section
  const float *amax;     // PRESCALE only: device fp32 max|x| of the whole tensor (arcq_absmax)
  float *scale_out;      // PRESCALE only (nullable): receives amax / 2688, the per-tensor scale the GEMM alpha needs
  int prescale_mode;     // PRESCALE only: ARCQ_PRESCALE_FP32 / ARCQ_PRESCALE_BF16_DIVISOR
  float eps;
  int M, KQ, KE;
  int num_tiles;
  int debug;  // -DARCQ_EXPERIMENT builds only: 1 skip transpose, 2 skip gather, 4 skip quantise math, 8 skip residuals
  // ADD (rmsnorm only): the normalised input is h = bf16(x + res); h is also written to sum_out (the new residual stream)
  const uint16_t *res;   // bf16 [M, KQ] (nullable = no add)
  uint16_t *sum_out;     // bf16 [M, KQ]
};

constexpr float kFp4Max = 6.0f;            // reorder.cu:17
constexpr float kFp8Max = 448.0f;          // reorder.cu:18
constexpr float kScaleEps = 0.001953125f;  // reorder.cu:19

// local row lr of tile t -> row inside the 128-row scale-factor block; w0 = first 32-bit SF word of the tile's
// 4R-byte run inside each 512-byte atom (the tcgen05 128x4 layout: byte = (row%32)*16 + (row/32)*4 + g%4)
template <int R> struct TileMap;
template <> struct TileMap<8> {
  static __device__ __forceinline__ int row(int t, int lr) { return 32 * (lr & 3) + 2 * t + (lr >> 2); }
  static __device__ __forceinline__ int w0(int t) { return 8 * t; }
};
template <> struct TileMap<4> {
  static __device__ __forceinline__ int row(int t, int lr) { return 32 * lr + t; }
  static __device__ __forceinline__ int w0(int t) { return 4 * t; }
};
template <> struct TileMap<2> {
  static __device__ __forceinline__ int row(int t, int lr) { return 32 * (2 * (t >> 5) + lr) + (t & 31); }
  static __device__ __forceinline__ int w0(int t) { return 4 * (t & 31) + 2 * (t >> 5); }
};
template <> struct TileMap<1> {
  static __device__ __forceinline__ int row(int t, int) { return t; }
  static __device__ __forceinline__ int w0(int t) { return 4 * (t & 31) + (t >> 5); }
};

// position (in units of one channel chunk = 2R bytes) of channel c inside the staged channel-major tile
template <int R> __device__ __forceinline__ int chan_pos(int c) {
  if constexpr (R == 1) return c;
  else return c ^ ((c >> 3) & (64 / R - 1));
}

// channel position for the matrix-transposed tile (quant_tma.cu): one stmatrix row (16 bytes) holds the chunks of the
// 8/R channels {c, c+8, ...} of a 16- (R=4) or 32-channel (R=2) block; the row index is XOR-ed with the block number
// and the position inside the row with the next block bits, so identity-like permutations gather without conflicts
template <int R> __device__ __forceinline__ int chan_pos_m(int c) {
  if constexpr (R == 4) { const int b = c >> 4; return (((b << 3) | ((c & 7) ^ (b & 7))) << 1) | (((c >> 3) ^ (c >> 7)) & 1); }
  else if constexpr (R == 2) { const int b = c >> 5; return (((b << 3) | ((c & 7) ^ (b & 7))) << 2) | (((c >> 3) ^ (c >> 8)) & 3); }
  else if constexpr (R == 8) { const int b = c >> 3; return (b << 3) | ((c & 7) ^ (b & 7)); }  // one 16-byte chunk per stmatrix row
  else return chan_pos<R>(c);
}
// bytes added to every staged row so that the 8 row addresses of one ldmatrix fall into 8 different 16-byte bank groups
__host__ __device__ constexpr int stage_pad(int R) { return R == 8 ? 16 : R == 4 ? 32 : R == 2 ? 64 : 0; }

// placement plans (plan.cu)
bool plan_usable(int R, int KQ);
int plan_section(int R);

// TMA-staged kernel family (quant_tma.cu); returns ARCQ_ENOSUP for shapes it does not take
int launch_quant_tma(const QuantParams &p, int flavour, bool is_w, bool rms, cudaStream_t stream);

}  // namespace arcq
