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

// ---------------------------------------------------------------- dropout masks (BERT sites)
// One Philox call serves EIGHT consecutive elements: each 32-bit word is split into two 16-bit uniforms, compared
// against thresh = round(p * 65536) (p is honoured to 1.5e-5; the scale is 65536 / (65536 - thresh), so the mask
// stays exactly unbiased).  Halving the generator work matters: the LayerNorm and attention kernels were
// instruction-bound on Philox, not bandwidth-bound (profiles/r01m_bert_ncu_full_summary.txt).
__host__ __device__ __forceinline__ uint32_t dropout_thresh16(float p) {
  const float t = p * 65536.f + 0.5f;
  return t <= 0.f ? 0u : (t >= 65535.f ? 65535u : (uint32_t)t);
}
inline float dropout_inv_keep(float p) {
  return p > 0.f ? 65536.f / (65536.f - (float)dropout_thresh16(p)) : 1.f;
}
// bit e of the result = element base+e is kept (base % 8 == 0)
__device__ __forceinline__ uint32_t dropout_keep8(uint64_t seed, uint64_t offset, long long base, uint32_t thresh) {
  uint32_t r[4];
  philox4x32_10(seed, offset + (uint64_t)(base >> 3), r);
  uint32_t bits = 0u;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    bits |= ((r[w] & 0xFFFFu) >= thresh ? 1u : 0u) << (2 * w);
    bits |= ((r[w] >> 16) >= thresh ? 1u : 0u) << (2 * w + 1);
  }
  return bits;
}
// multipliers (0 or inv_keep) for the four consecutive elements starting at `base` (base % 4 == 0): the matching
// half of the eight-element group
__device__ __forceinline__ void dropout_scale4(uint64_t seed, uint64_t offset, long long base, float p,
                                               float inv_keep, float (&m)[4]) {
  const uint32_t bits = dropout_keep8(seed, offset, base & ~7LL, dropout_thresh16(p)) >> (base & 4);
#pragma unroll
  for (int e = 0; e < 4; ++e) m[e] = ((bits >> e) & 1u) ? inv_keep : 0.f;
}

}  // namespace b200
