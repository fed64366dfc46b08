// philox.cuh — Philox4x32 counter-based RNG (Salmon et al., SC'11) for the sampler kernels,
// with 7 rounds: the smallest round count that paper reports as passing BigCrush (10 is its
// safety-margin default).  The samplers are bound by these integer rounds, not by HBM
// (profiles/r1_paths.jsonl), so the three rounds are worth ~25 % of the sampler forward.
// Stateless: the 128-bit counter is (element index, stream id) and the key is the 64-bit seed,
// so forward and backward kernels regenerate identical noise without storing it.
#pragma once
#include <stdint.h>

namespace mm25 {
namespace rng {

constexpr int PHILOX_ROUNDS = 7;

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}

// 4 x 32 random bits for counter (idx, stream) under key `seed`
__device__ __forceinline__ void philox4(uint64_t seed, uint64_t idx, uint64_t stream,
                                        uint32_t (&out)[4]) {
  uint32_t c[4] = {(uint32_t)idx, (uint32_t)(idx >> 32), (uint32_t)stream,
                   (uint32_t)(stream >> 32)};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < PHILOX_ROUNDS; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) out[i] = c[i];
}

// uniform in (0, 1]: never 0, so log() is finite
__device__ __forceinline__ float u01(uint32_t x) {
  return ((float)(x >> 8) + 1.0f) * (1.0f / 16777216.0f);
}

// 4 standard normals (Box-Muller) for counter (idx, stream).  The samplers are bound by integer
// and special-function work, not by HBM (profiles/r2_paths.jsonl), so the transform is kept to
// the minimum: uniforms from the mantissa trick (no I2F: conversions share the MUFU pipe),
// rad = sqrt(-2 ln2 lg2 u1) with lg2.approx / sqrt.approx, and the angle 2 pi u2' with u2' in
// [1, 2) fed to sin.approx / cos.approx unreduced (one period further along: same values).
__device__ __forceinline__ void normal4(uint64_t seed, uint64_t idx, uint64_t stream,
                                        float (&z)[4]) {
  uint32_t r[4];
  philox4(seed, idx, stream, r);
#pragma unroll
  for (int p = 0; p < 2; ++p) {
    // u1 in (0, 1]: 2 - [1, 2); u2' in [1, 2)
    const float u1 = 2.0f - __uint_as_float(0x3f800000u | (r[2 * p] >> 9));
    const float t2 = __uint_as_float(0x3f800000u | (r[2 * p + 1] >> 9));
    float l2, rad, sn, cs;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(u1));
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(rad) : "f"(-1.3862943611198906f * l2));
    const float ang = 6.283185307179586f * t2;
    asm("sin.approx.ftz.f32 %0, %1;" : "=f"(sn) : "f"(ang));
    asm("cos.approx.ftz.f32 %0, %1;" : "=f"(cs) : "f"(ang));
    z[2 * p] = rad * cs;
    z[2 * p + 1] = rad * sn;
  }
}

// element k of a row of standard normals: rows are keyed by `row`, 4 elements per counter
__device__ __forceinline__ float normal_at(uint64_t seed, uint64_t row, int k, uint64_t stream) {
  float z[4];
  normal4(seed, row * 0x100000ull + (uint64_t)(k >> 2), stream, z);
  return z[k & 3];
}

}  // namespace rng
}  // namespace mm25
