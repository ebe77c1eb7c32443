// Philox4x32-10 counter-based generator shared by every dropout in the library (head Dropout, BERT hidden /
// attention-probability dropout): one call yields four 32-bit words for four consecutive elements, keyed by
// (seed, offset + element_index/4), so a backward kernel regenerates the forward mask instead of reading it.
#pragma once
#include <stdint.h>

namespace b200 {

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  const uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
  const uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
__device__ __forceinline__ void philox4x32_10(uint64_t seed, uint64_t ctr, uint32_t (&out)[4]) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  uint32_t c[4] = {(uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u};
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
}

// uniform in [0,1) from the top 24 bits
__device__ __forceinline__ float philox_uniform(uint32_t r) { return (float)(r >> 8) * (1.0f / 16777216.0f); }

// dropout multipliers (0 or 1/(1-p)) for the four consecutive elements starting at flat index `base` (base % 4 == 0)
__device__ __forceinline__ void dropout_scale4(uint64_t seed, uint64_t offset, long long base, float p,
                                               float inv_keep, float (&m)[4]) {
  uint32_t r[4];
  philox4x32_10(seed, offset + (uint64_t)(base / 4), r);
#pragma unroll
  for (int e = 0; e < 4; ++e) m[e] = philox_uniform(r[e]) >= p ? inv_keep : 0.f;
}

}  // namespace b200
