// BERT-base encoder kernels other than the Linear GEMMs (those run on the tcgen05 implicit-GEMM kernel as flat 1x1
// problems): embedding gather + LayerNorm, (dropout +) residual + LayerNorm forward / backward, erf-GELU, column
// sums for bias gradients, tanh for the pooler.  Replaces what `model(b_input_ids, token_type_ids=None,
// attention_mask=b_input_mask, labels=b_labels)` (pytorch_on_language_distr.py:258-261) and `loss.backward()` (:268)
// dispatch to ATen for BertEmbeddings / BertSelfOutput / BertIntermediate / BertOutput / BertPooler.
// All of these are HBM-bound row kernels: one warp owns a token row (hidden = NV*32 vectors of 16 bytes), statistics
// and every reduction in fp32, activations in the storage dtype (bf16 or fp32).  Dropout masks are never stored: the
// backward kernels regenerate them from (seed, offset, element index) with Philox.
#include "common.cuh"
#include "philox.cuh"

namespace b200 {

// multipliers (0 or inv_keep) for W consecutive elements starting at flat index `base` (base % W == 0)
template <int W>
__device__ __forceinline__ void dropout_scale_vec(uint64_t seed, uint64_t offset, long long base, float p,
                                                  float inv_keep, float (&m)[W]) {
  if (W == 8) {
    const uint32_t bits = dropout_keep8(seed, offset, base, dropout_thresh16(p));
#pragma unroll
    for (int e = 0; e < W; ++e) m[e] = ((bits >> e) & 1u) ? inv_keep : 0.f;
  } else {
    float t[4];
    dropout_scale4(seed, offset, base, p, inv_keep, t);
#pragma unroll
    for (int e = 0; e < 4; ++e) m[e] = t[e];
  }
}

template <typename T> __device__ __forceinline__ float round_to(float v) { return to_f(from_f<T>(v)); }

// raw 16-byte vectors (kept packed while they wait in registers as the prefetch of the next row)
__device__ __forceinline__ void unpack_vec(const float4& v, float (&f)[4]) { f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w; }
__device__ __forceinline__ void unpack_vec(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

// row statistics of v[NV][W] held by one warp: mean and 1/sqrt(var + eps) (two-pass, in registers)
template <int NV, int W>
__device__ __forceinline__ void warp_row_stats(const float (&v)[NV][W], int H, float eps, float& mean, float& rstd) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int e = 0; e < W; ++e) s += v[i][e];
  mean = warp_sum(s) / (float)H;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int e = 0; e < W; ++e) {
      const float d = v[i][e] - mean;
      q += d * d;
    }
  rstd = rsqrtf(warp_sum(q) / (float)H + eps);
}

// ---------------------------------------------------------------- LayerNorm forward
// s = dropout(x) + residual ; y = LN(s).   s (rounded to the storage dtype) is what backward needs: it is written to
// sum_out (may alias x).
template <typename T, int NV>
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res, T* sum_out, T* __restrict__ y,
              float* __restrict__ mean_out, float* __restrict__ rstd_out, const float* __restrict__ gamma,
              const float* __restrict__ beta, float eps, long long rows, float p, float inv_keep, uint64_t seed,
              uint64_t offset, const unsigned long long* __restrict__ ctr) {
  constexpr int W = VecTraits<T>::W;
  constexpr int H = NV * 32 * W;
  __shared__ __align__(16) float sg[H], sb[H];   // scalar __ldg of gamma/beta was 48 load instructions per lane and row
  for (int i = threadIdx.x; i < H; i += blockDim.x) {
    sg[i] = gamma[i];
    sb[i] = beta[i];
  }
  __syncthreads();
  if (ctr != nullptr) offset += *ctr << 40;
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp; row < rows; row += nwarps) {
    float v[NV][W];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const long long o = row * H + (lane + 32 * i) * W;
      load_vec(x + o, v[i]);
      if (p > 0.f) {
        float m[W];
        dropout_scale_vec<W>(seed, offset, o, p, inv_keep, m);
#pragma unroll
        for (int e = 0; e < W; ++e) v[i][e] *= m[e];
      }
      if (res != nullptr) {
        float r[W];
        load_vec(res + o, r);
#pragma unroll
        for (int e = 0; e < W; ++e) v[i][e] += r[e];
      }
#pragma unroll
      for (int e = 0; e < W; ++e) v[i][e] = round_to<T>(v[i][e]);
      if (sum_out != nullptr) store_vec(sum_out + o, v[i]);
    }
    float mean, rstd;
    warp_row_stats<NV, W>(v, H, eps, mean, rstd);
    if (lane == 0) {
      mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = (lane + 32 * i) * W;
      float o_[W];
#pragma unroll
      for (int q = 0; q < W / 4; ++q) {
        const float4 gq = *reinterpret_cast<const float4*>(sg + col + 4 * q);
        const float4 bq = *reinterpret_cast<const float4*>(sb + col + 4 * q);
        o_[4 * q + 0] = (v[i][4 * q + 0] - mean) * rstd * gq.x + bq.x;
        o_[4 * q + 1] = (v[i][4 * q + 1] - mean) * rstd * gq.y + bq.y;
        o_[4 * q + 2] = (v[i][4 * q + 2] - mean) * rstd * gq.z + bq.z;
        o_[4 * q + 3] = (v[i][4 * q + 3] - mean) * rstd * gq.w + bq.w;
      }
      store_vec(y + row * H + col, o_);
    }
  }
}

// ---------------------------------------------------------------- embeddings: gather + LayerNorm + dropout
// s = (word[id] + type[t]) + pos[position] (the association order of BertEmbeddings.forward), y = dropout(LN(s)).
template <typename T, int NV>
__global__ void __launch_bounds__(256)
embed_ln_fwd_kernel(const long long* __restrict__ ids, const long long* __restrict__ type_ids,
                    const float* __restrict__ word, const float* __restrict__ pos, const float* __restrict__ type,
                    T* __restrict__ sum_out, T* __restrict__ y, float* __restrict__ mean_out,
                    float* __restrict__ rstd_out, const float* __restrict__ gamma, const float* __restrict__ beta,
                    float eps, long long rows, int S, int vocab, float p, float inv_keep, uint64_t seed,
                    uint64_t offset, const unsigned long long* __restrict__ ctr, int* __restrict__ bad_index) {
  constexpr int W = VecTraits<T>::W;
  constexpr int H = NV * 32 * W;
  if (ctr != nullptr) offset += *ctr << 40;
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp; row < rows; row += nwarps) {
    long long id = ids[row];
    const long long tt = type_ids != nullptr ? type_ids[row] : 0;
    if (id < 0 || id >= vocab || tt < 0 || tt > 1) {   // torch raises a device assert here; we flag and clamp
      if (lane == 0) atomicExch(bad_index, 1);
      id = 0;
    }
    const int s = (int)(row % S);
    float v[NV][W];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = (lane + 32 * i) * W;
#pragma unroll
      for (int q = 0; q < W / 4; ++q) {
        const float4 a = *reinterpret_cast<const float4*>(word + id * H + col + 4 * q);
        const float4 t = *reinterpret_cast<const float4*>(type + (tt > 0 ? H : 0) + col + 4 * q);
        const float4 c = *reinterpret_cast<const float4*>(pos + (long long)s * H + col + 4 * q);
        v[i][4 * q + 0] = round_to<T>((a.x + t.x) + c.x);
        v[i][4 * q + 1] = round_to<T>((a.y + t.y) + c.y);
        v[i][4 * q + 2] = round_to<T>((a.z + t.z) + c.z);
        v[i][4 * q + 3] = round_to<T>((a.w + t.w) + c.w);
      }
      store_vec(sum_out + row * H + col, v[i]);
    }
    float mean, rstd;
    warp_row_stats<NV, W>(v, H, eps, mean, rstd);
    if (lane == 0) {
      mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = (lane + 32 * i) * W;
      float o_[W];
#pragma unroll
      for (int e = 0; e < W; ++e) o_[e] = (v[i][e] - mean) * rstd * gamma[col + e] + beta[col + e];
      if (p > 0.f) {
        float m[W];
        dropout_scale_vec<W>(seed, offset, row * H + col, p, inv_keep, m);
#pragma unroll
        for (int e = 0; e < W; ++e) o_[e] *= m[e];
      }
      store_vec(y + row * H + col, o_);
    }
  }
}

// ---------------------------------------------------------------- LayerNorm backward
// dy = dy_a (+ dy_b)            [drop_after: dy *= dropout mask, the embedding's y = dropout(LN(s))]
// xhat = (s - mean) * rstd ; g = dy * gamma ; ds = rstd * (g - mean_row(g) - xhat * mean_row(g * xhat))
// dres = ds ; dx = ds * dropout mask (the branch that went through dropout; same buffer as dres when p == 0)
// per-block partial column sums: [block][3][H] = (sum dy*xhat, sum dy, sum dx)  -> dgamma, dbeta, bias grad of the
// Linear that produced x.  Lanes own the same columns for every row, so the sums accumulate in registers.
// One row is spread over TPR = H / (16-byte vector) threads (3 warps at H = 768 bf16), G rows per block: every thread
// keeps only its own vector's three column accumulators, so the kernel stays at ~60 registers and ~24 warps per SM.
// The two row sums cross warps through shared memory and one named barrier per row (slots alternate between rows).
template <typename T, int TPR, int G>
__global__ void __launch_bounds__(TPR * G, (TPR * G <= 384 ? 2 : 1))
ln_bwd_kernel(const T* __restrict__ dy_a, const T* __restrict__ dy_b, const T* __restrict__ s,
              const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const float* __restrict__ gamma,
              T* __restrict__ dres, T* __restrict__ dx, float* __restrict__ partials, long long rows, float p,
              float inv_keep, uint64_t seed, uint64_t offset, const unsigned long long* __restrict__ ctr,
              int drop_after) {
  constexpr int W = VecTraits<T>::W;
  constexpr int H = TPR * W;
  if (ctr != nullptr) offset += *ctr << 40;
  constexpr int WPR = TPR / 32;
  __shared__ float red[G][2][WPR][2];
  __shared__ float fin[3 * H];
  const int grp = threadIdx.x / TPR, t = threadIdx.x % TPR, lane = threadIdx.x & 31, wr = t >> 5;
  const int col = t * W;
  float gm[W], ag[W], ab[W], ad[W];
#pragma unroll
  for (int e = 0; e < W; ++e) {
    gm[e] = gamma[col + e];
    ag[e] = ab[e] = ad[e] = 0.f;
  }
  // the next row's three vectors (and its statistics) are fetched before this row's reduction barrier, so the
  // dependent load -> reduce -> store chain of one row overlaps the memory latency of the next
  using V = typename VecTraits<T>::V;
  const long long stride = (long long)gridDim.x * G;
  long long row = (long long)blockIdx.x * G + grp;
  V na = V(), nb = V(), ns = V();
  float nmean = 0.f, nrstd = 0.f;
  if (row < rows) {
    const long long o = row * H + col;
    na = *reinterpret_cast<const V*>(dy_a + o);
    if (dy_b != nullptr) nb = *reinterpret_cast<const V*>(dy_b + o);
    ns = *reinterpret_cast<const V*>(s + o);
    nmean = mean_in[row];
    nrstd = rstd_in[row];
  }
  int it = 0;
  for (; row < rows; row += stride, ++it) {
    const float mean = nmean, rstd = nrstd;
    const long long o = row * H + col;
    float d[W], xh[W];
    unpack_vec(na, d);
    if (dy_b != nullptr) {
      float tb[W];
      unpack_vec(nb, tb);
#pragma unroll
      for (int e = 0; e < W; ++e) d[e] += tb[e];
    }
    unpack_vec(ns, xh);
    if (row + stride < rows) {
      const long long o2 = (row + stride) * H + col;
      na = *reinterpret_cast<const V*>(dy_a + o2);
      if (dy_b != nullptr) nb = *reinterpret_cast<const V*>(dy_b + o2);
      ns = *reinterpret_cast<const V*>(s + o2);
      nmean = mean_in[row + stride];
      nrstd = rstd_in[row + stride];
    }
    if (drop_after && p > 0.f) {
      float m[W];
      dropout_scale_vec<W>(seed, offset, o, p, inv_keep, m);
#pragma unroll
      for (int e = 0; e < W; ++e) d[e] *= m[e];
    }
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int e = 0; e < W; ++e) {
      xh[e] = (xh[e] - mean) * rstd;
      ab[e] += d[e];
      ag[e] += d[e] * xh[e];
      d[e] *= gm[e];
      c1 += d[e];
      c2 += d[e] * xh[e];
    }
    c1 = warp_sum(c1);
    c2 = warp_sum(c2);
    float (*slot)[2] = red[grp][it & 1];
    if (lane == 0) {
      slot[wr][0] = c1;
      slot[wr][1] = c2;
    }
    asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "n"(TPR) : "memory");
    c1 = 0.f;
    c2 = 0.f;
#pragma unroll
    for (int w = 0; w < WPR; ++w) {
      c1 += slot[w][0];
      c2 += slot[w][1];
    }
    c1 *= 1.f / (float)H;
    c2 *= 1.f / (float)H;
#pragma unroll
    for (int e = 0; e < W; ++e) d[e] = rstd * (d[e] - c1 - xh[e] * c2);
    store_vec(dres + o, d);
    if (!drop_after) {
      if (p > 0.f) {
        float m[W];
        dropout_scale_vec<W>(seed, offset, o, p, inv_keep, m);
#pragma unroll
        for (int e = 0; e < W; ++e) d[e] *= m[e];
        store_vec(dx + o, d);
      }
#pragma unroll
      for (int e = 0; e < W; ++e) ad[e] += round_to<T>(d[e]);
    }
  }
  // deterministic cross-group sum: the row groups add into shared memory one after the other
  for (int g = 0; g < G; ++g) {
    if (grp == g) {
#pragma unroll
      for (int e = 0; e < W; ++e) {
        if (g == 0) {
          fin[col + e] = ag[e];
          fin[H + col + e] = ab[e];
          fin[2 * H + col + e] = ad[e];
        } else {
          fin[col + e] += ag[e];
          fin[H + col + e] += ab[e];
          fin[2 * H + col + e] += ad[e];
        }
      }
    }
    __syncthreads();
  }
  float* out = partials + (long long)blockIdx.x * 3 * H;
  for (int i = threadIdx.x; i < 3 * H; i += TPR * G) out[i] = fin[i];
}

// out[c] = sum_b part[b][c]   (fixed order: deterministic)
__global__ void __launch_bounds__(256)
sum_partials_kernel(const float* __restrict__ part, int nparts, int cols, float* __restrict__ o0,
                    float* __restrict__ o1, float* __restrict__ o2, int H) {
  __shared__ float red[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float a0 = 0.f, a1 = 0.f;
  if (c < cols) {
    int b = ry;
    for (; b + 8 < nparts; b += 16) {
      a0 += part[(long long)b * cols + c];
      a1 += part[(long long)(b + 8) * cols + c];
    }
    if (b < nparts) a0 += part[(long long)b * cols + c];
  }
  red[ry][cx] = a0 + a1;
  __syncthreads();
  if (ry == 0 && c < cols) {
    float v = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) v += red[j][cx];
    // cols = 3*H laid out [3][H] (LayerNorm backward) or a single vector (H == cols, o1 == o2 == NULL)
    float* dst = c < H ? o0 : (c < 2 * H ? o1 : o2);
    if (dst != nullptr) dst[c % H] = v;
  }
}

// ---------------------------------------------------------------- embedding backward
// dword[id] += ds (fp32 atomics: the only non-deterministic summation order in the library, as in ATen's
// embedding_dense_backward without sorting); dtype[t] likewise after a per-block pre-reduction is not worth it for
// two rows, so the type table takes the same path; dpos[s] = sum_n ds[n, s] in fixed order.
template <typename T>
__global__ void embed_bwd_scatter_kernel(const long long* __restrict__ ids, const long long* __restrict__ type_ids,
                                         const T* __restrict__ ds, float* __restrict__ dword,
                                         float* __restrict__ dtype_tab, long long rows, int H, int vocab) {
  constexpr int W = VecTraits<T>::W;
  const int nvec = H / W;
  const long long total = rows * nvec;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / nvec;
    const int col = (int)(i % nvec) * W;
    long long id = ids[row];
    if (id < 0 || id >= vocab) continue;
    const long long tt = type_ids != nullptr ? type_ids[row] : 0;
    float v[W];
    load_vec(ds + row * H + col, v);
    // 16-byte vector reductions (red.global.add.v4.f32, sm_90+): a quarter of the atomic instructions
    float* wp = dword + id * H + col;
    if ((reinterpret_cast<uintptr_t>(wp) & 15) == 0) {
#pragma unroll
      for (int e = 0; e < W; e += 4) atomicAdd(reinterpret_cast<float4*>(wp + e), make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]));
    } else {
#pragma unroll
      for (int e = 0; e < W; ++e) atomicAdd(wp + e, v[e]);
    }
    if (type_ids != nullptr) {
#pragma unroll
      for (int e = 0; e < W; ++e) atomicAdd(dtype_tab + (tt > 0 ? H : 0) + col + e, v[e]);
    }
  }
}
// dpos[s][c] = sum_n ds[n][s][c]; rows s >= S of the table get zero.  One thread per (s, column vector).
template <typename T>
__global__ void embed_bwd_pos_kernel(const T* __restrict__ ds, float* __restrict__ dpos, int N, int S, int H,
                                     int max_pos) {
  constexpr int W = VecTraits<T>::W;
  const int nvec = H / W;
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= (long long)max_pos * nvec) return;
  const int s = (int)(i / nvec), col = (int)(i % nvec) * W;
  float acc[W];
#pragma unroll
  for (int e = 0; e < W; ++e) acc[e] = 0.f;
  if (s < S)
    for (int n = 0; n < N; ++n) {
      float v[W];
      load_vec(ds + ((long long)n * S + s) * H + col, v);
#pragma unroll
      for (int e = 0; e < W; ++e) acc[e] += v[e];
    }
#pragma unroll
  for (int e = 0; e < W; ++e) dpos[(long long)s * H + col + e] = acc[e];
}

// ---------------------------------------------------------------- erf-GELU (config hidden_act = "gelu")
// erf through Abramowitz-Stegun 7.1.26 (|error| < 1.5e-7, below fp32 rounding of the surrounding arithmetic): one
// exponential, shared between Phi(x) and the density term of the derivative (exp(-x^2/2) is the same number).
// erff + expf made the backward kernel ALU-bound (43 us for 150 MB).
__device__ __forceinline__ void gelu_parts(float x, float& cdf, float& pdf) {
  const float u = fabsf(x) * 0.70710678118654752f;
  const float e = __expf(-u * u);
  const float t = __fdividef(1.f, fmaf(0.3275911f, u, 1.f));
  const float poly = t * fmaf(t, fmaf(t, fmaf(t, fmaf(t, 1.061405429f, -1.453152027f), 1.421413741f), -0.284496736f),
                              0.254829592f);
  const float erf_abs = fmaf(-poly, e, 1.f);
  cdf = 0.5f * (1.f + copysignf(erf_abs, x));
  pdf = 0.3989422804014327f * e;
}
__device__ __forceinline__ float gelu_f(float x) {
  float cdf, pdf;
  gelu_parts(x, cdf, pdf);
  return x * cdf;
}
__device__ __forceinline__ float gelu_grad_f(float x) {
  float cdf, pdf;
  gelu_parts(x, cdf, pdf);
  return fmaf(x, pdf, cdf);
}
template <typename T>
__global__ void gelu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long long nvec) {
  constexpr int W = VecTraits<T>::W;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    float v[W];
    load_vec(x + i * W, v);
#pragma unroll
    for (int e = 0; e < W; ++e) v[e] = gelu_f(v[e]);
    store_vec(y + i * W, v);
  }
}
template <typename T>
__global__ void gelu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx,
                                long long nvec) {
  constexpr int W = VecTraits<T>::W;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    float v[W], d[W];
    load_vec(x + i * W, v);
    load_vec(dy + i * W, d);
#pragma unroll
    for (int e = 0; e < W; ++e) d[e] *= gelu_grad_f(v[e]);
    store_vec(dx + i * W, d);
  }
}

// ---------------------------------------------------------------- column sums (bias gradients), two-stage
// block = 32 column vectors x 8 row lanes over one row chunk; partial[chunk][cols]
template <typename T>
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const T* __restrict__ x, float* __restrict__ part, long long rows, int cols,
                      long long rows_per_chunk) {
  constexpr int W = VecTraits<T>::W;
  __shared__ float red[8][32 * W + 1];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + cx) * W;
  const long long r0 = blockIdx.y * rows_per_chunk;
  long long r1 = r0 + rows_per_chunk;
  if (r1 > rows) r1 = rows;
  float acc[W];
#pragma unroll
  for (int e = 0; e < W; ++e) acc[e] = 0.f;
  if (col < cols)
    for (long long r = r0 + ry; r < r1; r += 8) {
      float v[W];
      load_vec(x + r * cols + col, v);
#pragma unroll
      for (int e = 0; e < W; ++e) acc[e] += v[e];
    }
#pragma unroll
  for (int e = 0; e < W; ++e) red[ry][cx * W + e] = acc[e];
  __syncthreads();
  for (int c = threadIdx.x; c < 32 * W; c += 256) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += red[j][c];
    const int gc = blockIdx.x * 32 * W + c;
    if (gc < cols) part[(long long)blockIdx.y * cols + gc] = s;
  }
}

// dx = dy * gelu'(x) and, in the same pass, partial column sums of dx (the bias gradient of BertIntermediate.dense)
template <typename T>
__global__ void __launch_bounds__(256)
gelu_bwd_colsum_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, float* __restrict__ part,
                       long long rows, int cols, long long rows_per_chunk) {
  constexpr int W = VecTraits<T>::W;
  __shared__ float red[8][32 * W + 1];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + cx) * W;
  const long long r0 = blockIdx.y * rows_per_chunk;
  long long r1 = r0 + rows_per_chunk;
  if (r1 > rows) r1 = rows;
  float acc[W];
#pragma unroll
  for (int e = 0; e < W; ++e) acc[e] = 0.f;
  if (col < cols)
    for (long long r = r0 + ry; r < r1; r += 8) {
      float v[W], d[W];
      load_vec(x + r * cols + col, v);
      load_vec(dy + r * cols + col, d);
#pragma unroll
      for (int e = 0; e < W; ++e) {
        d[e] *= gelu_grad_f(v[e]);
        acc[e] += round_to<T>(d[e]);
      }
      store_vec(dx + r * cols + col, d);
    }
#pragma unroll
  for (int e = 0; e < W; ++e) red[ry][cx * W + e] = acc[e];
  __syncthreads();
  for (int c = threadIdx.x; c < 32 * W; c += 256) {
    float sm = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) sm += red[j][c];
    const int gc = blockIdx.x * 32 * W + c;
    if (gc < cols) part[(long long)blockIdx.y * cols + gc] = sm;
  }
}

__global__ void tanh_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = tanhf(x[i]);
}
__global__ void tanh_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ dx,
                                long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dx[i] = dy[i] * (1.f - y[i] * y[i]);
}
// rows[n][:] = x[n*S][:] as fp32 ([CLS] token of every sequence) and its transpose-scatter for the backward
template <typename T>
__global__ void gather_rows_kernel(const T* __restrict__ x, float* __restrict__ out, int N, long long row_stride,
                                   int H) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= (long long)N * H) return;
  out[i] = to_f(x[(i / H) * row_stride + (i % H)]);
}
template <typename T>
__global__ void scatter_rows_kernel(const float* __restrict__ d, T* __restrict__ out, int N, long long row_stride,
                                    int H) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= (long long)N * H) return;
  out[(i / H) * row_stride + (i % H)] = from_f<T>(d[i]);
}

// ---------------------------------------------------------------- host-side helpers
static inline int row_grid(long long rows, int warps_per_block, int blocks_per_sm) {
  const long long want = (rows + warps_per_block - 1) / warps_per_block;
  const long long cap = (long long)num_sms() * blocks_per_sm;
  return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
}
static inline int ew_grid(long long n, int threads) {
  const long long want = (n + threads - 1) / threads;
  const long long cap = (long long)num_sms() * 8;
  return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
}
constexpr int LN_BWD_BLOCKS_PER_SM = 2;
static inline int ln_bwd_grid(long long rows) { return row_grid(rows, 2, LN_BWD_BLOCKS_PER_SM); }
static inline int ln_nv(int H, int W) { return (H % (32 * W) == 0) ? H / (32 * W) : 0; }

}  // namespace b200

using namespace b200;

#define B200_LN_DISPATCH_NV(NV_, CALL)                 \
  switch (NV_) {                                       \
    case 1: { constexpr int NV = 1; CALL; } break;     \
    case 2: { constexpr int NV = 2; CALL; } break;     \
    case 3: { constexpr int NV = 3; CALL; } break;     \
    case 4: { constexpr int NV = 4; CALL; } break;     \
    case 6: { constexpr int NV = 6; CALL; } break;     \
    case 8: { constexpr int NV = 8; CALL; } break;     \
    default: return fail(B200_UNSUPPORTED, "layernorm: hidden size %d not supported (needs 32*vec*{1,2,3,4,6,8})", H); \
  }

template <typename T>
static int32_t ln_fwd_t(const T* x, const T* res, T* sum_out, T* y, float* mean, float* rstd, const float* gamma,
                        const float* beta, float eps, long long rows, int H, float p, uint64_t seed, uint64_t offset,
                        cudaStream_t st) {
  const int nv = ln_nv(H, VecTraits<T>::W);
  const float ik = dropout_inv_keep(p);
  const int grid = row_grid(rows, 8, 8);
  B200_LN_DISPATCH_NV(nv, (ln_fwd_kernel<T, NV><<<grid, 256, 0, st>>>(x, res, sum_out, y, mean, rstd, gamma, beta, eps,
                                                                    rows, p, ik, seed, offset, dropout_ctr())));
  return check_launch("ln_fwd");
}

template <typename T>
static int32_t embed_fwd_t(const long long* ids, const long long* tt, const float* word, const float* pos,
                           const float* type, T* sum_out, T* y, float* mean, float* rstd, const float* gamma,
                           const float* beta, float eps, long long rows, int S, int H, int vocab, float p,
                           uint64_t seed, uint64_t offset, int* status, cudaStream_t st) {
  const int nv = ln_nv(H, VecTraits<T>::W);
  const float ik = dropout_inv_keep(p);
  const int grid = row_grid(rows, 8, 8);
  B200_LN_DISPATCH_NV(nv, (embed_ln_fwd_kernel<T, NV><<<grid, 256, 0, st>>>(ids, tt, word, pos, type, sum_out, y, mean,
                                                                          rstd, gamma, beta, eps, rows, S, vocab, p,
                                                                          ik, seed, offset, dropout_ctr(), status)));
  return check_launch("embed_ln_fwd");
}

template <typename T>
static int32_t ln_bwd_t(const T* dy_a, const T* dy_b, const T* s, const float* mean, const float* rstd,
                        const float* gamma, long long rows, int H, float p, uint64_t seed, uint64_t offset,
                        int drop_after, T* dres, T* dx, float* dgamma, float* dbeta, float* dbias, float* ws,
                        cudaStream_t st) {
  constexpr int W = VecTraits<T>::W;
  const int tpr = (H % (32 * W) == 0) ? H / W : 0;
  const float ik = dropout_inv_keep(p);
  const int grid = ln_bwd_grid(rows);
#define B200_LN_BWD_CASE(TPR_, G_)                                                                              \
  case TPR_:                                                                                                     \
    ln_bwd_kernel<T, TPR_, G_><<<grid, TPR_ * G_, 0, st>>>(dy_a, dy_b, s, mean, rstd, gamma, dres, dx, ws, rows, p, \
                                                         ik, seed, offset, dropout_ctr(), drop_after);           \
    break;
  switch (tpr) {
    B200_LN_BWD_CASE(32, 8)
    B200_LN_BWD_CASE(64, 8)
    B200_LN_BWD_CASE(96, 4)
    B200_LN_BWD_CASE(128, 4)
    B200_LN_BWD_CASE(192, 2)
    B200_LN_BWD_CASE(256, 2)
    default:
      return fail(B200_UNSUPPORTED, "layernorm backward: hidden size %d not supported", H);
  }
#undef B200_LN_BWD_CASE
  sum_partials_kernel<<<(3 * H + 31) / 32, 256, 0, st>>>(ws, grid, 3 * H, dgamma, dbeta, dbias, H);
  return check_launch("ln_bwd", 2);
}

static inline int colsum_chunks(long long rows, int cols, int W) {
  const int cblocks = (cols + 32 * W - 1) / (32 * W);
  long long chunks = (8LL * num_sms() + cblocks - 1) / cblocks;   // ~8 blocks per SM: enough loads in flight
  if (chunks > (rows + 63) / 64) chunks = (rows + 63) / 64;
  return (int)(chunks < 1 ? 1 : chunks);
}

template <typename T>
static int32_t gelu_bwd_colsum_t(const T* dy, const T* x, T* dx, float* dbias, long long rows, int cols, float* ws,
                                 long long ws_floats, cudaStream_t st) {
  constexpr int W = VecTraits<T>::W;
  const int cblocks = (cols + 32 * W - 1) / (32 * W);
  const int chunks = colsum_chunks(rows, cols, W);
  if ((long long)chunks * cols > ws_floats) return fail(B200_WORKSPACE_TOO_SMALL, "gelu_bwd_bias: workspace too small");
  const long long rpc = (rows + chunks - 1) / chunks;
  gelu_bwd_colsum_kernel<T><<<dim3(cblocks, chunks), 256, 0, st>>>(dy, x, dx, ws, rows, cols, rpc);
  sum_partials_kernel<<<(cols + 31) / 32, 256, 0, st>>>(ws, chunks, cols, dbias, nullptr, nullptr, cols);
  return check_launch("gelu_bwd_bias", 2);
}

template <typename T>
static int32_t colsum_t(const T* x, float* out, long long rows, int cols, float* ws, long long ws_floats,
                        cudaStream_t st) {
  constexpr int W = VecTraits<T>::W;
  const int cblocks = (cols + 32 * W - 1) / (32 * W);
  const int chunks = colsum_chunks(rows, cols, W);
  if ((long long)chunks * cols > ws_floats) return fail(B200_WORKSPACE_TOO_SMALL, "colsum: workspace too small");
  const long long rpc = (rows + chunks - 1) / chunks;
  colsum_partial_kernel<T><<<dim3(cblocks, chunks), 256, 0, st>>>(x, ws, rows, cols, rpc);
  sum_partials_kernel<<<(cols + 31) / 32, 256, 0, st>>>(ws, chunks, cols, out, nullptr, nullptr, cols);
  return check_launch("colsum", 2);
}

extern "C" {

int32_t b200_embed_ln_fwd(const int64_t* ids, const int64_t* type_ids, const float* word, const float* pos,
                          const float* type, const float* gamma, const float* beta, float eps, int32_t N, int32_t S,
                          int32_t H, int32_t vocab, int32_t max_pos, float p_drop, uint64_t seed, uint64_t offset,
                          void* sum_out, void* y, float* mean, float* rstd, int32_t* status, int32_t dtype,
                          void* stream) {
  if (!ids || !word || !pos || !type || !gamma || !beta || !sum_out || !y || !mean || !rstd || !status)
    return fail(B200_INVALID_ARG, "embed_ln_fwd: null pointer");
  if (N <= 0 || S <= 0 || S > max_pos) return fail(B200_INVALID_ARG, "embed_ln_fwd: sequence length %d exceeds the position table (%d)", S, max_pos);
  if (p_drop < 0.f || p_drop >= 1.f) return fail(B200_INVALID_ARG, "embed_ln_fwd: bad dropout probability");
  const long long rows = (long long)N * S;
  if (dtype == B200_F32)
    return embed_fwd_t<float>((const long long*)ids, (const long long*)type_ids, word, pos, type, (float*)sum_out,
                              (float*)y, mean, rstd, gamma, beta, eps, rows, S, H, vocab, p_drop, seed, offset, status,
                              as_stream(stream));
  if (dtype == B200_BF16)
    return embed_fwd_t<__nv_bfloat16>((const long long*)ids, (const long long*)type_ids, word, pos, type,
                                      (__nv_bfloat16*)sum_out, (__nv_bfloat16*)y, mean, rstd, gamma, beta, eps, rows,
                                      S, H, vocab, p_drop, seed, offset, status, as_stream(stream));
  return fail(B200_INVALID_ARG, "embed_ln_fwd: bad dtype");
}

int32_t b200_embed_bwd(const int64_t* ids, const int64_t* type_ids, const void* dsum, int32_t N, int32_t S, int32_t H,
                       int32_t vocab, int32_t max_pos, float* dword, float* dpos, float* dtype_tab, void* ws,
                       int64_t ws_bytes, int32_t dtype, void* stream) {
  if (!ids || !dsum || !dword || !dpos || !dtype_tab) return fail(B200_INVALID_ARG, "embed_bwd: null pointer");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "embed_bwd: bad dtype");
  const int W = dtype == B200_F32 ? 4 : 8;
  if (H % W != 0) return fail(B200_UNSUPPORTED, "embed_bwd: hidden size %d", H);
  cudaStream_t st = as_stream(stream);
  const long long rows = (long long)N * S;
  cudaMemsetAsync(dword, 0, (size_t)vocab * H * sizeof(float), st);
  cudaMemsetAsync(dtype_tab, 0, (size_t)2 * H * sizeof(float), st);
  const long long total = rows * (H / W);
  const int grid = ew_grid(total, 256);
  const int pgrid = (int)(((long long)max_pos * (H / W) + 255) / 256);
  int32_t rc;
  if (dtype == B200_F32) {
    embed_bwd_scatter_kernel<float><<<grid, 256, 0, st>>>((const long long*)ids, (const long long*)type_ids,
                                                         (const float*)dsum, dword, dtype_tab, rows, H, vocab);
    embed_bwd_pos_kernel<float><<<pgrid, 256, 0, st>>>((const float*)dsum, dpos, N, S, H, max_pos);
    rc = check_launch("embed_bwd", 2);
    if (rc == B200_OK && type_ids == nullptr)
      rc = colsum_t<float>((const float*)dsum, dtype_tab, rows, H, (float*)ws, ws_bytes / 4, st);
  } else {
    embed_bwd_scatter_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const long long*)ids, (const long long*)type_ids,
                                                                 (const __nv_bfloat16*)dsum, dword, dtype_tab, rows,
                                                                 H, vocab);
    embed_bwd_pos_kernel<__nv_bfloat16><<<pgrid, 256, 0, st>>>((const __nv_bfloat16*)dsum, dpos, N, S, H, max_pos);
    rc = check_launch("embed_bwd", 2);
    if (rc == B200_OK && type_ids == nullptr)
      rc = colsum_t<__nv_bfloat16>((const __nv_bfloat16*)dsum, dtype_tab, rows, H, (float*)ws, ws_bytes / 4, st);
  }
  return rc;
}

int32_t b200_ln_fwd(const void* x, const void* residual, void* sum_out, void* y, float* mean, float* rstd,
                    const float* gamma, const float* beta, float eps, int64_t rows, int32_t H, float p_drop,
                    uint64_t seed, uint64_t offset, int32_t dtype, void* stream) {
  if (!x || !y || !mean || !rstd || !gamma || !beta || rows <= 0) return fail(B200_INVALID_ARG, "ln_fwd: bad argument");
  if (p_drop < 0.f || p_drop >= 1.f) return fail(B200_INVALID_ARG, "ln_fwd: bad dropout probability");
  if (dtype == B200_F32)
    return ln_fwd_t<float>((const float*)x, (const float*)residual, (float*)sum_out, (float*)y, mean, rstd, gamma, beta,
                           eps, rows, H, p_drop, seed, offset, as_stream(stream));
  if (dtype == B200_BF16)
    return ln_fwd_t<__nv_bfloat16>((const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, (__nv_bfloat16*)sum_out,
                                   (__nv_bfloat16*)y, mean, rstd, gamma, beta, eps, rows, H, p_drop, seed, offset,
                                   as_stream(stream));
  return fail(B200_INVALID_ARG, "ln_fwd: bad dtype");
}

int64_t b200_ln_bwd_workspace_bytes(int64_t rows, int32_t H) {
  return (int64_t)ln_bwd_grid(rows) * 3 * H * (int64_t)sizeof(float);
}

int32_t b200_ln_bwd(const void* dy_a, const void* dy_b, const void* sum, const float* mean, const float* rstd,
                    const float* gamma, int64_t rows, int32_t H, float p_drop, uint64_t seed, uint64_t offset,
                    int32_t drop_after, void* dres, void* dx, float* dgamma, float* dbeta, float* dbias, void* ws,
                    int64_t ws_bytes, int32_t dtype, void* stream) {
  if (!dy_a || !sum || !mean || !rstd || !gamma || !dres || !dgamma || !dbeta || rows <= 0)
    return fail(B200_INVALID_ARG, "ln_bwd: bad argument");
  if (p_drop < 0.f || p_drop >= 1.f) return fail(B200_INVALID_ARG, "ln_bwd: bad dropout probability");
  if (!drop_after && p_drop > 0.f && !dx) return fail(B200_INVALID_ARG, "ln_bwd: dx required when dropout is on");
  if (!ws || ws_bytes < b200_ln_bwd_workspace_bytes(rows, H))
    return fail(B200_WORKSPACE_TOO_SMALL, "ln_bwd: workspace too small");
  if (dtype == B200_F32)
    return ln_bwd_t<float>((const float*)dy_a, (const float*)dy_b, (const float*)sum, mean, rstd, gamma, rows, H,
                           p_drop, seed, offset, drop_after, (float*)dres, (float*)dx, dgamma, dbeta, dbias, (float*)ws,
                           as_stream(stream));
  if (dtype == B200_BF16)
    return ln_bwd_t<__nv_bfloat16>((const __nv_bfloat16*)dy_a, (const __nv_bfloat16*)dy_b, (const __nv_bfloat16*)sum,
                                   mean, rstd, gamma, rows, H, p_drop, seed, offset, drop_after, (__nv_bfloat16*)dres,
                                   (__nv_bfloat16*)dx, dgamma, dbeta, dbias, (float*)ws, as_stream(stream));
  return fail(B200_INVALID_ARG, "ln_bwd: bad dtype");
}

int32_t b200_gelu_fwd(const void* x, void* y, int64_t n, int32_t dtype, void* stream) {
  if (!x || !y || n < 0) return fail(B200_INVALID_ARG, "gelu_fwd: bad argument");
  const int W = dtype == B200_F32 ? 4 : 8;
  if (n % W != 0) return fail(B200_UNSUPPORTED, "gelu_fwd: element count must be a multiple of %d", W);
  if (n == 0) return B200_OK;
  const int grid = ew_grid(n / W, 256);
  if (dtype == B200_F32)
    gelu_fwd_kernel<float><<<grid, 256, 0, as_stream(stream)>>>((const float*)x, (float*)y, n / W);
  else if (dtype == B200_BF16)
    gelu_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, as_stream(stream)>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y,
                                                                       n / W);
  else
    return fail(B200_INVALID_ARG, "gelu_fwd: bad dtype");
  return check_launch("gelu_fwd");
}

int32_t b200_gelu_bwd(const void* dy, const void* x, void* dx, int64_t n, int32_t dtype, void* stream) {
  if (!dy || !x || !dx || n < 0) return fail(B200_INVALID_ARG, "gelu_bwd: bad argument");
  const int W = dtype == B200_F32 ? 4 : 8;
  if (n % W != 0) return fail(B200_UNSUPPORTED, "gelu_bwd: element count must be a multiple of %d", W);
  if (n == 0) return B200_OK;
  const int grid = ew_grid(n / W, 256);
  if (dtype == B200_F32)
    gelu_bwd_kernel<float><<<grid, 256, 0, as_stream(stream)>>>((const float*)dy, (const float*)x, (float*)dx, n / W);
  else if (dtype == B200_BF16)
    gelu_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, as_stream(stream)>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (__nv_bfloat16*)dx, n / W);
  else
    return fail(B200_INVALID_ARG, "gelu_bwd: bad dtype");
  return check_launch("gelu_bwd");
}

int32_t b200_gelu_bwd_bias(const void* dy, const void* x, void* dx, float* dbias, int64_t rows, int32_t cols,
                           int32_t dtype, void* ws, int64_t ws_bytes, void* stream) {
  if (!dy || !x || !dx || !dbias || !ws || rows <= 0 || cols <= 0) return fail(B200_INVALID_ARG, "gelu_bwd_bias: bad argument");
  const int W = dtype == B200_F32 ? 4 : 8;
  if (cols % W != 0) return fail(B200_UNSUPPORTED, "gelu_bwd_bias: column count must be a multiple of %d", W);
  if (dtype == B200_F32)
    return gelu_bwd_colsum_t<float>((const float*)dy, (const float*)x, (float*)dx, dbias, rows, cols, (float*)ws,
                                    ws_bytes / 4, as_stream(stream));
  if (dtype == B200_BF16)
    return gelu_bwd_colsum_t<__nv_bfloat16>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (__nv_bfloat16*)dx,
                                            dbias, rows, cols, (float*)ws, ws_bytes / 4, as_stream(stream));
  return fail(B200_INVALID_ARG, "gelu_bwd_bias: bad dtype");
}

int64_t b200_colsum_workspace_bytes(int64_t rows, int32_t cols) {
  (void)rows;
  return (int64_t)(8 * num_sms() + 8) * cols * (int64_t)sizeof(float);
}

int32_t b200_colsum(const void* x, float* out, int64_t rows, int32_t cols, int32_t dtype, void* ws, int64_t ws_bytes,
                    void* stream) {
  if (!x || !out || !ws || rows <= 0 || cols <= 0) return fail(B200_INVALID_ARG, "colsum: bad argument");
  const int W = dtype == B200_F32 ? 4 : 8;
  if (cols % W != 0) return fail(B200_UNSUPPORTED, "colsum: column count must be a multiple of %d", W);
  if (dtype == B200_F32) return colsum_t<float>((const float*)x, out, rows, cols, (float*)ws, ws_bytes / 4, as_stream(stream));
  if (dtype == B200_BF16)
    return colsum_t<__nv_bfloat16>((const __nv_bfloat16*)x, out, rows, cols, (float*)ws, ws_bytes / 4, as_stream(stream));
  return fail(B200_INVALID_ARG, "colsum: bad dtype");
}

int32_t b200_dropout_counter(const uint64_t* device_counter) {
  dropout_ctr() = reinterpret_cast<const unsigned long long*>(device_counter);
  return B200_OK;
}

int32_t b200_tanh_fwd_f32(const float* x, float* y, int64_t n, void* stream) {
  if (!x || !y || n < 0) return fail(B200_INVALID_ARG, "tanh_fwd: bad argument");
  if (n == 0) return B200_OK;
  tanh_fwd_kernel<<<ew_grid(n, 256), 256, 0, as_stream(stream)>>>(x, y, n);
  return check_launch("tanh_fwd");
}
int32_t b200_tanh_bwd_f32(const float* dy, const float* y, float* dx, int64_t n, void* stream) {
  if (!dy || !y || !dx || n < 0) return fail(B200_INVALID_ARG, "tanh_bwd: bad argument");
  if (n == 0) return B200_OK;
  tanh_bwd_kernel<<<ew_grid(n, 256), 256, 0, as_stream(stream)>>>(dy, y, dx, n);
  return check_launch("tanh_bwd");
}

int32_t b200_gather_rows(const void* x, float* out, int32_t N, int64_t row_stride, int32_t H, int32_t dtype,
                         void* stream) {
  if (!x || !out || N <= 0 || H <= 0) return fail(B200_INVALID_ARG, "gather_rows: bad argument");
  const int grid = (int)(((long long)N * H + 255) / 256);
  if (dtype == B200_F32)
    gather_rows_kernel<float><<<grid, 256, 0, as_stream(stream)>>>((const float*)x, out, N, row_stride, H);
  else if (dtype == B200_BF16)
    gather_rows_kernel<__nv_bfloat16><<<grid, 256, 0, as_stream(stream)>>>((const __nv_bfloat16*)x, out, N, row_stride, H);
  else
    return fail(B200_INVALID_ARG, "gather_rows: bad dtype");
  return check_launch("gather_rows");
}
int32_t b200_scatter_rows(const float* d, void* out, int32_t N, int64_t row_stride, int32_t H, int32_t dtype,
                          void* stream) {
  if (!d || !out || N <= 0 || H <= 0) return fail(B200_INVALID_ARG, "scatter_rows: bad argument");
  const int grid = (int)(((long long)N * H + 255) / 256);
  if (dtype == B200_F32)
    scatter_rows_kernel<float><<<grid, 256, 0, as_stream(stream)>>>(d, (float*)out, N, row_stride, H);
  else if (dtype == B200_BF16)
    scatter_rows_kernel<__nv_bfloat16><<<grid, 256, 0, as_stream(stream)>>>(d, (__nv_bfloat16*)out, N, row_stride, H);
  else
    return fail(B200_INVALID_ARG, "scatter_rows: bad dtype");
  return check_launch("scatter_rows");
}

}  // extern "C"
