// ref_probe.cu -- TEST INFRASTRUCTURE ONLY (see oracle/arcq_oracle.c header).
//
// Host-run probe of the reference's OWN arithmetic library: the CUTLASS v4.3.1 headers vendored at
// /root/reference/third-party/cutlass (the converters reorder.cu:103-109 instantiates, and the CuTe
// scale-factor layout reorder.cu:343 / reorder.cuh:118-123 uses).  Compiled by oracle/build_ref.sh with
// nvcc from where those headers lie; writes raw tables to the directory given as argv[1]; those are
// turned into the committed fixtures under tests/golden/ by tests/golden/make_golden.py.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "cute/tensor.hpp"
#include "cutlass/detail/sm100_blockscaled_layout.hpp"
#include "cutlass/numeric_conversion.h"
#include "cutlass/numeric_types.h"

using fp4_t = cutlass::float_e2m1_t;
using sf_t = cutlass::float_ue4m3_t;
using bf16_t = cutlass::bfloat16_t;

static void dump(const std::string &path, const void *p, size_t n) {
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) { perror(path.c_str()); exit(1); }
  fwrite(p, 1, n, f);
  fclose(f);
}

static float bits2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static uint32_t f2bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

int main(int argc, char **argv) {
  std::string dir = argc > 1 ? argv[1] : ".";
  cutlass::NumericConverter<fp4_t, float, cutlass::FloatRoundStyle::round_to_nearest> f2e2m1;
  cutlass::NumericConverter<float, fp4_t, cutlass::FloatRoundStyle::round_to_nearest> e2m12f;
  cutlass::NumericConverter<sf_t, float, cutlass::FloatRoundStyle::round_to_nearest> f2ue4m3;
  cutlass::NumericConverter<float, sf_t, cutlass::FloatRoundStyle::round_to_nearest> ue4m32f;
  cutlass::NumericConverter<bf16_t, float, cutlass::FloatRoundStyle::round_to_nearest> f2bf16;

  // (1) e2m1 code of every fp32 value whose low 16 bits are zero (= every bf16), clamped to [-6, 6] as the
  //     kernels do (reorder.cu:153-156); NaNs skipped (code 0xff).
  std::vector<uint8_t> e2m1(65536);
  for (uint32_t h = 0; h < 65536; ++h) {
    float x = bits2f(h << 16);
    if (x != x) { e2m1[h] = 0xff; continue; }
    float t = x > 6.0f ? 6.0f : (x < -6.0f ? -6.0f : x);
    e2m1[h] = f2e2m1(t).storage & 0xf;
  }
  dump(dir + "/e2m1_of_bf16.bin", e2m1.data(), e2m1.size());

  // (2) e2m1 around every rounding boundary in fp32: boundary, one ulp below, one ulp above, both signs
  const float bnd[7] = {0.25f, 0.75f, 1.25f, 1.75f, 2.5f, 3.5f, 5.0f};
  std::vector<uint32_t> e2m1_in;
  std::vector<uint8_t> e2m1_out;
  for (int s = 0; s < 2; ++s)
    for (float b : bnd)
      for (int d = -1; d <= 1; ++d) {
        uint32_t u = f2bits(b) + d;
        if (s) u |= 0x80000000u;
        e2m1_in.push_back(u);
        e2m1_out.push_back(f2e2m1(bits2f(u)).storage & 0xf);
      }
  for (uint32_t u : {0x00000000u, 0x80000000u, 0x00000001u, 0x80000001u, 0x40c00000u, 0xc0c00000u}) {
    e2m1_in.push_back(u);
    e2m1_out.push_back(f2e2m1(bits2f(u)).storage & 0xf);
  }
  dump(dir + "/e2m1_boundary_in.bin", e2m1_in.data(), e2m1_in.size() * 4);
  dump(dir + "/e2m1_boundary_out.bin", e2m1_out.data(), e2m1_out.size());

  // (3) e2m1 decode table
  float dec[16];
  for (int c = 0; c < 16; ++c) { fp4_t v; v.storage = (uint8_t)c; dec[c] = e2m12f(v); }
  dump(dir + "/e2m1_decode.bin", dec, sizeof(dec));

  // (4) ue4m3: every bf16-valued input in [2^-9, 448] plus fp32 neighbours of every midpoint between
  //     consecutive representable values; and the decode table of bytes 0..0x7e.
  std::vector<uint32_t> sf_in;
  for (uint32_t h = 0; h < 32768; ++h) {
    float x = bits2f(h << 16);
    if (x >= 0.001953125f && x <= 448.0f) sf_in.push_back(h << 16);
  }
  float decsf[127];
  for (int b = 0; b < 127; ++b) { sf_t v; v.storage = (uint8_t)b; decsf[b] = ue4m32f(v); }
  for (int b = 1; b < 126; ++b) {
    float mid = 0.5f * (decsf[b] + decsf[b + 1]);
    for (int d = -2; d <= 2; ++d) sf_in.push_back(f2bits(mid) + d);
    sf_in.push_back(f2bits(decsf[b]));
  }
  uint64_t lcg = 0x9e3779b97f4a7c15ull;  // pseudo-random fp32 in [2^-9, 448)
  for (int i = 0; i < 20000; ++i) {
    lcg = lcg * 6364136223846793005ull + 1442695040888963407ull;
    uint32_t mant = (uint32_t)(lcg >> 41);
    uint32_t expo = 118 + (uint32_t)((lcg >> 20) % 18);  // 2^-9 .. 2^8
    float x = bits2f((expo << 23) | mant);
    if (x >= 0.001953125f && x <= 448.0f) sf_in.push_back(f2bits(x));
  }
  std::vector<uint8_t> sf_out(sf_in.size());
  for (size_t i = 0; i < sf_in.size(); ++i) sf_out[i] = f2ue4m3(bits2f(sf_in[i])).storage;
  dump(dir + "/ue4m3_in.bin", sf_in.data(), sf_in.size() * 4);
  dump(dir + "/ue4m3_out.bin", sf_out.data(), sf_out.size());
  dump(dir + "/ue4m3_decode.bin", decsf, sizeof(decsf));

  // (5) bf16 RNE of pseudo-random fp32 values (the residual rounding, reorder.cu:157-160)
  std::vector<uint32_t> bf_in;
  std::vector<uint16_t> bf_out;
  for (int i = 0; i < 20000; ++i) {
    lcg = lcg * 6364136223846793005ull + 1442695040888963407ull;
    uint32_t u = (uint32_t)(lcg >> 32);
    if (((u >> 23) & 0xff) == 0xff) continue;
    if (i % 10 == 0) u = (u & 0xffff0000u) | 0x8000u;  // exact ties
    bf_in.push_back(u);
    bf_out.push_back(f2bf16(bits2f(u)).storage);
  }
  dump(dir + "/bf16_in.bin", bf_in.data(), bf_in.size() * 4);
  dump(dir + "/bf16_out.bin", bf_out.data(), bf_out.size() * 2);

  // (6) scale-factor byte offsets from CuTe for (rows, K) problems: every (row, 16-group)
  using Cfg = cutlass::detail::Sm1xxBlockScaledConfig<16>;
  const int probs[3][2] = {{300, 4160}, {128, 64}, {257, 1088}};
  for (auto &pr : probs) {
    int rows = pr[0], K = pr[1];
    // the unfiltered layout takes ELEMENT coordinates (the 16 elements of a group share one byte: stride 0);
    // filter_zeros is only for the size of the buffer
    auto layout = Cfg::tile_atom_to_shape_SFA(cute::make_shape(rows, 128, K, 1));
    std::vector<int32_t> off((size_t)rows * (K / 16));
    for (int r = 0; r < rows; ++r)
      for (int g = 0; g < K / 16; ++g) off[(size_t)r * (K / 16) + g] = (int32_t)layout(cute::make_coord(r, g * 16, 0));
    dump(dir + "/sf_offsets_" + std::to_string(rows) + "x" + std::to_string(K) + ".bin", off.data(), off.size() * 4);
    int32_t cs = (int32_t)cute::cosize(cute::filter_zeros(layout));
    dump(dir + "/sf_cosize_" + std::to_string(rows) + "x" + std::to_string(K) + ".bin", &cs, 4);
  }
  printf("ref_probe: wrote tables to %s\n", dir.c_str());
  return 0;
}
