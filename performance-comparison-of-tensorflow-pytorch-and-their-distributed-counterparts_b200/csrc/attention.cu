// Multi-head self-attention of BertSelfAttention (the 12 `attention.self` modules inside the call at
// pytorch_on_language_distr.py:258-261; HF source not in the reference tree): scores = Q K^T / sqrt(D) with padded
// keys masked out, softmax, dropout on the probabilities, context = P V.  Head dimension D = 64.
//
// qkv is the output of the fused QKV Linear: [N*S][3*heads*D] rows = (q | k | v), head h at columns h*D of each
// third.  ctx is [N*S][heads*D].  probs [N][heads][S][S] keeps the softmax output BEFORE dropout for the backward;
// the dropout mask is regenerated from Philox (seed, offset, flat probs index).
//
// This file is the fp32-arithmetic path (CUDA cores, shared-memory tiles) used by the exact-parity mode and as the
// reference the tensor-core kernel is tested against.
#include "common.cuh"
#include "philox.cuh"

namespace b200 {

constexpr int ATT_D = 64;     // head dimension
constexpr int ATT_QT = 32;    // query rows (or key rows in the dK/dV kernel) per CTA
constexpr int ATT_KC = 64;    // rows per staged chunk
constexpr int ATT_KQ = 32;    // query rows per staged chunk of the dK/dV kernel (static smem <= 48 KB)
constexpr int ATT_THREADS = 256;
constexpr float ATT_MASKED = -3.0e38f;

template <typename T>
__device__ __forceinline__ void att_load_chunk(float (*dst)[ATT_D + 1], const T* __restrict__ src, long long ld,
                                               int row0, int nrows, int S) {
  // dst[r][d] = src[(row0 + r) * ld + d] for r < nrows, zero past S
  for (int i = threadIdx.x; i < nrows * ATT_D; i += ATT_THREADS) {
    const int r = i / ATT_D, d = i % ATT_D;
    dst[r][d] = (row0 + r < S) ? to_f(src[(long long)(row0 + r) * ld + d]) : 0.f;
  }
}

// ---------------------------------------------------------------- forward
template <typename T>
__global__ void __launch_bounds__(ATT_THREADS)
attention_fwd_kernel(const T* __restrict__ qkv, const long long* __restrict__ mask, T* __restrict__ probs,
                     T* __restrict__ ctx, int S, int heads, float scale, float p, float inv_keep, uint64_t seed,
                     uint64_t offset, const unsigned long long* __restrict__ ctr) {
  extern __shared__ float smem[];
  if (ctr != nullptr) offset += *ctr << 40;
  const int Sp = (S + 3) / 4 * 4 + 4;                          // padded row of the score tile
  float (*Qs)[ATT_D + 1] = reinterpret_cast<float (*)[ATT_D + 1]>(smem);
  float (*KVs)[ATT_D + 1] = reinterpret_cast<float (*)[ATT_D + 1]>(smem + ATT_QT * (ATT_D + 1));
  float* Ps = smem + (ATT_QT + ATT_KC) * (ATT_D + 1);
  const int q0 = blockIdx.x * ATT_QT, h = blockIdx.y, n = blockIdx.z;
  const int HD = heads * ATT_D;
  const long long ld = 3LL * HD;
  const T* qb = qkv + (long long)n * S * ld + h * ATT_D;
  const T* kb = qb + HD;
  const T* vb = qb + 2 * HD;
  att_load_chunk<T>(Qs, qb, ld, q0, ATT_QT, S);
  const int q = threadIdx.x >> 3, g = threadIdx.x & 7;
  // phase 1: scores
  for (int k0 = 0; k0 < S; k0 += ATT_KC) {
    __syncthreads();
    att_load_chunk<T>(KVs, kb, ld, k0, ATT_KC, S);
    __syncthreads();
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll 8
    for (int d = 0; d < ATT_D; ++d) {
      const float qv = Qs[q][d];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += qv * KVs[g + 8 * j][d];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = k0 + g + 8 * j;
      if (k < S) {
        const bool keep = mask == nullptr || mask[(long long)n * S + k] != 0;
        Ps[q * Sp + k] = keep ? acc[j] * scale : ATT_MASKED;
      }
    }
  }
  __syncthreads();
  // phase 2: softmax over each row (warp w owns rows 4w..4w+3), probs out, dropout into the smem copy
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = warp * 4; r < warp * 4 + 4; ++r) {
    if (q0 + r >= S) break;
    float* pr = Ps + r * Sp;
    float m = ATT_MASKED;
    for (int k = lane; k < S; k += 32) m = fmaxf(m, pr[k]);
    m = warp_max(m);
    float s = 0.f;
    for (int k = lane; k < S; k += 32) {
      const float e = __expf(pr[k] - m);
      pr[k] = e;
      s += e;
    }
    s = 1.f / warp_sum(s);
    __syncwarp();
    const long long pbase = (((long long)n * heads + h) * S + (q0 + r)) * S;
    for (int k4 = lane * 4; k4 < S; k4 += 128) {
      float mk[4] = {1.f, 1.f, 1.f, 1.f};
      if (p > 0.f) dropout_scale4(seed, offset, pbase + k4, p, inv_keep, mk);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int k = k4 + e;
        if (k < S) {
          const T pv = from_f<T>(pr[k] * s);
          probs[pbase + k] = pv;
          pr[k] = to_f(pv) * mk[e];
        }
      }
    }
  }
  // phase 3: context = P V
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int k0 = 0; k0 < S; k0 += ATT_KC) {
    __syncthreads();
    att_load_chunk<T>(KVs, vb, ld, k0, ATT_KC, S);
    __syncthreads();
    const int kn = S - k0 < ATT_KC ? S - k0 : ATT_KC;
    for (int k = 0; k < kn; ++k) {
      const float pv = Ps[q * Sp + k0 + k];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += pv * KVs[k][g + 8 * j];
    }
  }
  if (q0 + q < S) {
    T* o = ctx + ((long long)n * S + q0 + q) * HD + h * ATT_D;
#pragma unroll
    for (int j = 0; j < 8; ++j) o[g + 8 * j] = from_f<T>(acc[j]);
  }
}

// ---------------------------------------------------------------- backward, query side
// dP = dO V^T ; t = dP * dropmask ; Dq = sum_k P t ; dS = P (t - Dq) ; dQ = scale * dS K.  dS is also written to
// `ds_out` [N][heads][S][S] for the key-side kernel.
template <typename T>
__global__ void __launch_bounds__(ATT_THREADS)
attention_bwd_q_kernel(const T* __restrict__ qkv, const T* __restrict__ probs, const T* __restrict__ dctx,
                       T* __restrict__ ds_out, T* __restrict__ dqkv, int S, int heads, float scale, float p,
                       float inv_keep, uint64_t seed, uint64_t offset, const unsigned long long* __restrict__ ctr) {
  extern __shared__ float smem[];
  if (ctr != nullptr) offset += *ctr << 40;
  const int Sp = (S + 3) / 4 * 4 + 4;
  float (*dOs)[ATT_D + 1] = reinterpret_cast<float (*)[ATT_D + 1]>(smem);
  float (*KVs)[ATT_D + 1] = reinterpret_cast<float (*)[ATT_D + 1]>(smem + ATT_QT * (ATT_D + 1));
  float* Ps = smem + (ATT_QT + ATT_KC) * (ATT_D + 1);
  const int q0 = blockIdx.x * ATT_QT, h = blockIdx.y, n = blockIdx.z;
  const int HD = heads * ATT_D;
  const long long ld = 3LL * HD;
  const T* kb = qkv + (long long)n * S * ld + h * ATT_D + HD;
  const T* vb = kb + HD;
  att_load_chunk<T>(dOs, dctx + (long long)n * S * HD + h * ATT_D, HD, q0, ATT_QT, S);
  const int q = threadIdx.x >> 3, g = threadIdx.x & 7;
  for (int k0 = 0; k0 < S; k0 += ATT_KC) {
    __syncthreads();
    att_load_chunk<T>(KVs, vb, ld, k0, ATT_KC, S);
    __syncthreads();
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll 8
    for (int d = 0; d < ATT_D; ++d) {
      const float ov = dOs[q][d];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += ov * KVs[g + 8 * j][d];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = k0 + g + 8 * j;
      if (k < S) Ps[q * Sp + k] = acc[j];
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = warp * 4; r < warp * 4 + 4; ++r) {
    if (q0 + r >= S) break;
    float* pr = Ps + r * Sp;
    const long long pbase = (((long long)n * heads + h) * S + (q0 + r)) * S;
    float dsum = 0.f;
    float pk[16], tk[16];   // S <= 512: at most 4 groups of 4 per lane
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int k4 = lane * 4 + c * 128;
      float mk[4] = {1.f, 1.f, 1.f, 1.f};
      if (p > 0.f && k4 < S) dropout_scale4(seed, offset, pbase + k4, p, inv_keep, mk);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int k = k4 + e;
        const float pv = k < S ? to_f(probs[pbase + k]) : 0.f;
        const float t = k < S ? pr[k] * mk[e] : 0.f;
        pk[c * 4 + e] = pv;
        tk[c * 4 + e] = t;
        dsum += pv * t;
      }
    }
    dsum = warp_sum(dsum);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int k4 = lane * 4 + c * 128;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int k = k4 + e;
        if (k < S) {
          const T dsv = from_f<T>(pk[c * 4 + e] * (tk[c * 4 + e] - dsum));
          ds_out[pbase + k] = dsv;
          pr[k] = to_f(dsv);
        }
      }
    }
  }
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int k0 = 0; k0 < S; k0 += ATT_KC) {
    __syncthreads();
    att_load_chunk<T>(KVs, kb, ld, k0, ATT_KC, S);
    __syncthreads();
    const int kn = S - k0 < ATT_KC ? S - k0 : ATT_KC;
    for (int k = 0; k < kn; ++k) {
      const float dv = Ps[q * Sp + k0 + k];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += dv * KVs[k][g + 8 * j];
    }
  }
  if (q0 + q < S) {
    T* o = dqkv + ((long long)n * S + q0 + q) * ld + h * ATT_D;
#pragma unroll
    for (int j = 0; j < 8; ++j) o[g + 8 * j] = from_f<T>(acc[j] * scale);
  }
}

// ---------------------------------------------------------------- backward, key side
// dV[k] = sum_q (P dropmask)[q][k] dO[q] ; dK[k] = scale * sum_q dS[q][k] Q[q]
template <typename T>
__global__ void __launch_bounds__(ATT_THREADS)
attention_bwd_kv_kernel(const T* __restrict__ qkv, const T* __restrict__ probs, const T* __restrict__ ds_in,
                        const T* __restrict__ dctx, T* __restrict__ dqkv, int S, int heads, float scale, float p,
                        float inv_keep, uint64_t seed, uint64_t offset, const unsigned long long* __restrict__ ctr) {
  __shared__ float Qs[ATT_KQ][ATT_D + 1];
  if (ctr != nullptr) offset += *ctr << 40;
  __shared__ float dOs[ATT_KQ][ATT_D + 1];
  __shared__ float Pt[ATT_KQ][ATT_QT + 1];    // [query in chunk][key in tile]
  __shared__ float dSt[ATT_KQ][ATT_QT + 1];
  const int k0 = blockIdx.x * ATT_QT, h = blockIdx.y, n = blockIdx.z;
  const int HD = heads * ATT_D;
  const long long ld = 3LL * HD;
  const T* qb = qkv + (long long)n * S * ld + h * ATT_D;
  const T* dob = dctx + (long long)n * S * HD + h * ATT_D;
  const int kk = threadIdx.x >> 3, g = threadIdx.x & 7;
  float av[8], ak[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) av[j] = ak[j] = 0.f;
  for (int qc = 0; qc < S; qc += ATT_KQ) {
    __syncthreads();
    att_load_chunk<T>(Qs, qb, ld, qc, ATT_KQ, S);
    att_load_chunk<T>(dOs, dob, HD, qc, ATT_KQ, S);
    // P / dS tile: ATT_KQ query rows x ATT_QT keys, four consecutive keys per thread
    for (int i = threadIdx.x; i < ATT_KQ * (ATT_QT / 4); i += ATT_THREADS) {
      const int r = i / (ATT_QT / 4), c4 = (i % (ATT_QT / 4)) * 4;
      const long long pbase = (((long long)n * heads + h) * S + (qc + r)) * S;
      float mk[4] = {1.f, 1.f, 1.f, 1.f};
      const bool row_ok = qc + r < S;
      if (p > 0.f && row_ok && k0 + c4 < S) dropout_scale4(seed, offset, pbase + k0 + c4, p, inv_keep, mk);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int k = k0 + c4 + e;
        const bool ok = row_ok && k < S;
        Pt[r][c4 + e] = ok ? to_f(probs[pbase + k]) * mk[e] : 0.f;
        dSt[r][c4 + e] = ok ? to_f(ds_in[pbase + k]) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll 4
    for (int r = 0; r < ATT_KQ; ++r) {
      const float pv = Pt[r][kk], dv = dSt[r][kk];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        av[j] += pv * dOs[r][g + 8 * j];
        ak[j] += dv * Qs[r][g + 8 * j];
      }
    }
  }
  if (k0 + kk < S) {
    T* ok_ = dqkv + ((long long)n * S + k0 + kk) * ld + h * ATT_D + HD;
    T* ov_ = ok_ + HD;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      ok_[g + 8 * j] = from_f<T>(ak[j] * scale);
      ov_[g + 8 * j] = from_f<T>(av[j]);
    }
  }
}

static inline size_t att_smem_bytes(int S) {
  const int Sp = (S + 3) / 4 * 4 + 4;
  return ((size_t)(ATT_QT + ATT_KC) * (ATT_D + 1) + (size_t)ATT_QT * Sp) * sizeof(float);
}

template <typename T>
static int32_t attention_fwd_t(const T* qkv, const long long* mask, T* probs, T* ctx, int N, int S, int heads,
                               float p, uint64_t seed, uint64_t offset, cudaStream_t st) {
  const size_t smem = att_smem_bytes(S);
  static int attr_state[B200_MAX_DEVICES] = {0};   // per device; raised once per size, outside any capture replay
  ensure_dyn_smem(attention_fwd_kernel<T>, (int)smem, attr_state);
  dim3 grid((S + ATT_QT - 1) / ATT_QT, heads, N);
  attention_fwd_kernel<T><<<grid, ATT_THREADS, smem, st>>>(qkv, mask, probs, ctx, S, heads, 0.125f, p,
                                                          dropout_inv_keep(p), seed, offset,
                                                          dropout_ctr());
  return check_launch("attention_fwd");
}
template <typename T>
static int32_t attention_bwd_t(const T* qkv, const T* probs, const T* dctx, T* ds, T* dqkv, int N, int S, int heads,
                               float p, uint64_t seed, uint64_t offset, cudaStream_t st) {
  const size_t smem = att_smem_bytes(S);
  static int attr_state[B200_MAX_DEVICES] = {0};
  ensure_dyn_smem(attention_bwd_q_kernel<T>, (int)smem, attr_state);
  dim3 grid((S + ATT_QT - 1) / ATT_QT, heads, N);
  const float ik = dropout_inv_keep(p);
  attention_bwd_q_kernel<T><<<grid, ATT_THREADS, smem, st>>>(qkv, probs, dctx, ds, dqkv, S, heads, 0.125f, p, ik,
                                                            seed, offset, dropout_ctr());
  attention_bwd_kv_kernel<T><<<grid, ATT_THREADS, 0, st>>>(qkv, probs, ds, dctx, dqkv, S, heads, 0.125f, p, ik, seed,
                                                          offset, dropout_ctr());
  return check_launch("attention_bwd", 2);
}

}  // namespace b200

using namespace b200;

extern "C" {

int32_t b200_attention_fwd(const void* qkv, const int64_t* mask, void* probs, void* ctx, int32_t N, int32_t S,
                           int32_t heads, int32_t D, float p_drop, uint64_t seed, uint64_t offset, int32_t dtype,
                           void* stream) {
  if (!qkv || !probs || !ctx || N <= 0 || S <= 0 || heads <= 0)
    return fail(B200_INVALID_ARG, "attention_fwd: bad argument");
  if (D != ATT_D || S > 512 || S % 4 != 0 || p_drop < 0.f || p_drop >= 1.f)
    return fail(B200_UNSUPPORTED, "attention_fwd: needs head dim 64, S <= 512, S %% 4 == 0 (got D=%d S=%d)", D, S);
  if (dtype == B200_F32)
    return attention_fwd_t<float>((const float*)qkv, (const long long*)mask, (float*)probs, (float*)ctx, N, S, heads,
                                  p_drop, seed, offset, as_stream(stream));
  if (dtype == B200_BF16)
    return attention_fwd_t<__nv_bfloat16>((const __nv_bfloat16*)qkv, (const long long*)mask, (__nv_bfloat16*)probs,
                                          (__nv_bfloat16*)ctx, N, S, heads, p_drop, seed, offset, as_stream(stream));
  return fail(B200_INVALID_ARG, "attention_fwd: bad dtype");
}

int32_t b200_attention_bwd(const void* qkv, const void* probs, const void* dctx, void* dscores, void* dqkv, int32_t N,
                           int32_t S, int32_t heads, int32_t D, float p_drop, uint64_t seed, uint64_t offset,
                           int32_t dtype, void* stream) {
  if (!qkv || !probs || !dctx || !dscores || !dqkv || N <= 0 || S <= 0 || heads <= 0)
    return fail(B200_INVALID_ARG, "attention_bwd: bad argument");
  if (D != ATT_D || S > 512 || S % 4 != 0 || p_drop < 0.f || p_drop >= 1.f)
    return fail(B200_UNSUPPORTED, "attention_bwd: needs head dim 64, S <= 512, S %% 4 == 0 (got D=%d S=%d)", D, S);
  if (dtype == B200_F32)
    return attention_bwd_t<float>((const float*)qkv, (const float*)probs, (const float*)dctx, (float*)dscores,
                                  (float*)dqkv, N, S, heads, p_drop, seed, offset, as_stream(stream));
  if (dtype == B200_BF16)
    return attention_bwd_t<__nv_bfloat16>((const __nv_bfloat16*)qkv, (const __nv_bfloat16*)probs,
                                          (const __nv_bfloat16*)dctx, (__nv_bfloat16*)dscores, (__nv_bfloat16*)dqkv,
                                          N, S, heads, p_drop, seed, offset, as_stream(stream));
  return fail(B200_INVALID_ARG, "attention_bwd: bad dtype");
}

}  // extern "C"
