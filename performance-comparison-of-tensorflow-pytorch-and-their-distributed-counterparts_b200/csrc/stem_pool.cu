// The stem's BatchNorm -> ReLU -> MaxPool2d(3, stride 2, pad 1) as ONE pass each way (torchvision ResNet.forward:
// bn1 / relu / maxpool, the model the reference builds in pytorch_training_inference_on_image.ipynb).
//
// Unfused, the stem's tail moves the largest activation of the network (N x 80 x 80 x 64 at the benchmark shape,
// 210 MB in bf16) through HBM seven times: apply (read y, write z), pool (read z), pool backward (write dz),
// BatchNorm backward reduction (read dz, y) and dx pass (read dz, y; write dy).  Here
//   forward : pooled = maxpool(relu(y * scale + shift)) straight from the convolution output y -- z never exists
//             (window re-reads are L1/L2 hits), argmax as one byte per pooled element;
//   backward: the gradient of y's element (h, w) is gathered from the <= 4 windows that cover it,
//             g = sum over windows [argmax == this element] * [pooled > 0] * dpool   (pooled > 0 <=> the winner passed the
//             ReLU), once in the reduction pass (sum g, sum g * (y - mean) -> the usual partial rows -> finalize) and once
//             in the dx pass -- dz never exists either.
// One block per image row, 32-bit index arithmetic, and every gather load issued before the first one is consumed (the
// older pooling kernels in pool.cu interleave load and compare, which leaves one 16-byte load in flight per thread:
// 127 / 256 us forward / backward at the benchmark shape against ~45 us of memory time).
#include "common.cuh"

namespace b200 {

int32_t bn_finalize_bwd_launch(const float* part, int nblk, int C, double n, const float* gamma, const float* invstd,
                               float* dgamma, float* dbeta, float* coef, cudaStream_t st);   // bn.cu
int bn_max_partial_blocks();                                                                 // bn.cu

constexpr int SP_THREADS = 256;

// raw 16-byte vector -> floats
__device__ __forceinline__ void unpack16(const uint4& v, float (&f)[4]) {
  f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y); f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
}
__device__ __forceinline__ void unpack16(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

// ---- lane-parallel helpers on ONE 32-bit register of a 16-byte vector: two bf16 lanes or one fp32 lane.
// These kernels are instruction-bound when written element by element (ncu: ~1000 instructions per pooled vector,
// issue slots 70-80 % busy at a third of the memory rate), so the window maximum, the arg-max search and the
// backward's winner / ReLU masks are done on packed registers.
template <typename T> struct Lanes;
template <> struct Lanes<__nv_bfloat16> {
  static constexpr uint32_t SIGN = 0x80008000u, NEG_INF = 0xFF80FF80u, KSPLAT = 0x00010001u;
  static __device__ __forceinline__ __nv_bfloat162 as2(uint32_t v) { return *reinterpret_cast<__nv_bfloat162*>(&v); }
  static __device__ __forceinline__ uint32_t vmax(uint32_t a, uint32_t b) {
    const __nv_bfloat162 r = __hmax2_nan(as2(a), as2(b));     // NaN propagates, as in torch's max_pool2d
    return *reinterpret_cast<const uint32_t*>(&r);
  }
  static __device__ __forceinline__ uint32_t eq_mask(uint32_t a, uint32_t b) { return __heq2_mask(as2(a), as2(b)); }
  static __device__ __forceinline__ uint32_t gt0_mask(uint32_t a) { return __hgt2_mask(as2(a), as2(0u)); }
  static __device__ __forceinline__ uint32_t add(uint32_t a, uint32_t b) {
    const __nv_bfloat162 r = __hadd2(as2(a), as2(b));
    return *reinterpret_cast<const uint32_t*>(&r);
  }
  // byte masks of eight winners (two words) -> lane mask of register j (elements 2j, 2j+1)
  static __device__ __forceinline__ uint32_t lane_mask(uint32_t m0, uint32_t m1, int j) {
    return __byte_perm(j < 2 ? m0 : m1, 0u, (j & 1) ? 0x3322u : 0x1100u);
  }
  // arg-max lanes of the four registers -> one byte per element
  static __device__ __forceinline__ void pack_index(const uint32_t (&ix)[4], uint32_t (&out)[2]) {
    out[0] = __byte_perm(ix[0], ix[1], 0x6420u);
    out[1] = __byte_perm(ix[2], ix[3], 0x6420u);
  }
};
template <> struct Lanes<float> {
  static constexpr uint32_t SIGN = 0x80000000u, NEG_INF = 0xFF800000u, KSPLAT = 1u;
  static __device__ __forceinline__ uint32_t vmax(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("max.NaN.f32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
  }
  static __device__ __forceinline__ uint32_t eq_mask(uint32_t a, uint32_t b) {
    return __uint_as_float(a) == __uint_as_float(b) ? 0xffffffffu : 0u;
  }
  static __device__ __forceinline__ uint32_t gt0_mask(uint32_t a) { return __uint_as_float(a) > 0.f ? 0xffffffffu : 0u; }
  static __device__ __forceinline__ uint32_t add(uint32_t a, uint32_t b) {
    return __float_as_uint(__uint_as_float(a) + __uint_as_float(b));
  }
  static __device__ __forceinline__ uint32_t lane_mask(uint32_t m0, uint32_t, int j) {
    return __byte_perm(m0, 0u, 0x1111u * (uint32_t)j);
  }
  static __device__ __forceinline__ void pack_index(const uint32_t (&ix)[4], uint32_t (&out)[2]) {
    out[0] = ix[0] | (ix[1] << 8) | (ix[2] << 16) | (ix[3] << 24);
    out[1] = 0u;
  }
};

// one block = one pooled row (n, p); thread -> (q, channel vector); SP_THREADS % (C / VEC) == 0, so a thread keeps its
// channel vector (and that vector's scale / shift) for all of its pixels.
// relu(y * scale + shift) is monotone in y for each channel, so the window maximum is taken on the raw convolution
// output (with the sign flipped where scale < 0: max of -y = -min of y) and the affine map + ReLU + rounding applied to
// the one winner: bit-identical pooled values to apply-then-pool; the arg-max is the first element holding the extreme
// raw value (apply-then-pool breaks ties between DIFFERENT raw values that round to the same output by position; here
// the larger raw value wins -- the two differ only where the gradient is split between equal outputs anyway).
template <typename T, bool AFFINE>
__global__ void __launch_bounds__(SP_THREADS, 3)
bn_relu_maxpool_fwd_kernel(const T* __restrict__ y, const float* __restrict__ scale, const float* __restrict__ shift,
                           T* __restrict__ pooled, uint8_t* __restrict__ amax, int H, int W, int C, int Ho, int Wo) {
  constexpr int VEC = VecTraits<T>::W;
  using L = Lanes<T>;
  const int cv = C / VEC;
  const int n = blockIdx.x / Ho, p = blockIdx.x - n * Ho;
  const T* img = y + (size_t)n * H * W * C;
  const int c = (threadIdx.x % cv) * VEC;
  float sc[VEC], sh[VEC];
  uint32_t flip[4] = {0u, 0u, 0u, 0u};
  if (AFFINE) {
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      sc[e] = scale[c + e];
      sh[e] = shift[c + e];
      if (sc[e] < 0.f) flip[e * 4 / VEC] |= (VEC == 8 ? (0x8000u << ((e & 1) * 16)) : 0x80000000u);
    }
  }
  for (int q = threadIdx.x / cv; q < Wo; q += SP_THREADS / cv) {
    // all nine window loads are issued before any of them is consumed (clamped coordinates; out-of-image taps are
    // replaced by -inf afterwards)
    uint4 raw[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hc = min(max(2 * p - 1 + r, 0), H - 1);
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int wc = min(max(2 * q - 1 + s, 0), W - 1);
        raw[r * 3 + s] = __ldg(reinterpret_cast<const uint4*>(img + ((size_t)hc * W + wc) * C + c));
      }
    }
    uint32_t v[9][4], m[4];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int h = 2 * p - 1 + k / 3, w = 2 * q - 1 + k % 3;
      const bool ok = h >= 0 && h < H && w >= 0 && w < W;
      v[k][0] = ok ? (raw[k].x ^ flip[0]) : L::NEG_INF;
      v[k][1] = ok ? (raw[k].y ^ flip[1]) : L::NEG_INF;
      v[k][2] = ok ? (raw[k].z ^ flip[2]) : L::NEG_INF;
      v[k][3] = ok ? (raw[k].w ^ flip[3]) : L::NEG_INF;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      m[j] = v[0][j];
#pragma unroll
      for (int k = 1; k < 9; ++k) m[j] = L::vmax(m[j], v[k][j]);
    }
    const size_t o = (((size_t)n * Ho + p) * Wo + q) * C + c;
    if (amax) {
      uint32_t ix[4], pk[2];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        ix[j] = 0u;
#pragma unroll
        for (int k = 8; k >= 0; --k) {     // descending: the FIRST element equal to the maximum is what remains
          const uint32_t eq = L::eq_mask(v[k][j], m[j]);
          ix[j] = (ix[j] & ~eq) | ((L::KSPLAT * (uint32_t)k) & eq);
        }
      }
      L::pack_index(ix, pk);
      if (VEC == 8) *reinterpret_cast<uint2*>(amax + o) = make_uint2(pk[0], pk[1]);
      else *reinterpret_cast<uint32_t*>(amax + o) = pk[0];
    }
    float best[VEC];
    unpack16(make_uint4(m[0] ^ flip[0], m[1] ^ flip[1], m[2] ^ flip[2], m[3] ^ flip[3]), best);
    if (AFFINE) {
#pragma unroll
      for (int e = 0; e < VEC; ++e) best[e] = fmaxf(fmaf(best[e], sc[e], sh[e]), 0.f);
    }
    store_vec(pooled + o, best);
  }
}

// Backward gather, one 2 x 2 input patch (rows 2p, 2p+1; columns 2q, 2q+1 -- the stride cell) per thread and channel
// vector.  The patch is covered by the four windows (p, q), (p, q+1), (p+1, q), (p+1, q+1) and by no other, and which
// window element each of its pixels is, is a compile-time constant:
//     pixel (0,0): W00 #4           pixel (0,1): W00 #5, W01 #3
//     pixel (1,0): W00 #7, W10 #1   pixel (1,1): W00 #8, W01 #6, W10 #2, W11 #0
// so twelve loads (arg-max bytes, pooled gradient, pooled value of four windows) serve four output vectors, against
// thirteen per output vector when every pixel gathers for itself.  g[dh][dw] = packed gradient of the pooling input
// (bf16: contributions are added in bf16, as the separate pooling-backward kernel stores them).
// RELU: a window passes its gradient on only when its pooled value is positive.
template <typename T, bool RELU>
__device__ __forceinline__ void patch_gather(const T* __restrict__ dpool, const uint8_t* __restrict__ amax,
                                             const T* __restrict__ pooled, int p, int q, int c, int C, int Ho, int Wo,
                                             uint32_t (&g)[2][2][4]) {
  constexpr int VEC = VecTraits<T>::W;
  using L = Lanes<T>;
  const bool vp = p + 1 < Ho, vq = q + 1 < Wo;
  const int o00 = (p * Wo + q) * C + c;                       // element offsets inside this image (< 2^31: launcher)
  const int off[4] = {o00, o00 + (vq ? C : 0), o00 + (vp ? Wo * C : 0), o00 + (vp ? Wo * C : 0) + (vq ? C : 0)};
  uint32_t idx[4][2];
  uint4 dv[4], pv[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {       // every load is issued before the first use; invalid windows re-read a valid one
    if (VEC == 8) { const uint2 t = __ldg(reinterpret_cast<const uint2*>(amax + off[k])); idx[k][0] = t.x; idx[k][1] = t.y; }
    else { idx[k][0] = __ldg(reinterpret_cast<const uint32_t*>(amax + off[k])); idx[k][1] = 0xffffffffu; }
    dv[k] = __ldg(reinterpret_cast<const uint4*>(dpool + off[k]));
    if (RELU) pv[k] = __ldg(reinterpret_cast<const uint4*>(pooled + off[k]));
  }
  // ... and then take no part: no element code matches 0xff
  if (!vq) { idx[1][0] = idx[1][1] = idx[3][0] = idx[3][1] = 0xffffffffu; }
  if (!vp) { idx[2][0] = idx[2][1] = idx[3][0] = idx[3][1] = 0xffffffffu; }
  uint32_t d[4][4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    d[k][0] = dv[k].x; d[k][1] = dv[k].y; d[k][2] = dv[k].z; d[k][3] = dv[k].w;
    if (RELU) {
      d[k][0] &= L::gt0_mask(pv[k].x); d[k][1] &= L::gt0_mask(pv[k].y);
      d[k][2] &= L::gt0_mask(pv[k].z); d[k][3] &= L::gt0_mask(pv[k].w);
    }
  }
  auto sel = [&](int k, uint32_t code, uint32_t (&out)[4], bool first) {
    const uint32_t m0 = __vcmpeq4(idx[k][0], code * 0x01010101u), m1 = __vcmpeq4(idx[k][1], code * 0x01010101u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t t = d[k][j] & L::lane_mask(m0, m1, j);
      out[j] = first ? t : L::add(out[j], t);
    }
  };
  sel(0, 4u, g[0][0], true);
  sel(0, 5u, g[0][1], true);  sel(1, 3u, g[0][1], false);
  sel(0, 7u, g[1][0], true);  sel(2, 1u, g[1][0], false);
  sel(0, 8u, g[1][1], true);  sel(1, 6u, g[1][1], false);  sel(2, 2u, g[1][1], false);  sel(3, 0u, g[1][1], false);
}

// reduction pass: block b owns the patch rows (n, p) [b * rows_per_block, ...) of the N*Ho; thread = (channel vector,
// patch lane), patches of the block's rows are walked as one flat index.  part layout [nblk][2][C] as in bn.cu.
template <typename T>
__global__ void __launch_bounds__(SP_THREADS, 3)
maxpool_bn_bwd_reduce_kernel(const T* __restrict__ dpool, const uint8_t* __restrict__ amax, const T* __restrict__ pooled,
                             const T* __restrict__ y, const float* __restrict__ mean, int NP, int H, int W, int C, int Ho,
                             int Wo, int rows_per_block, float* __restrict__ part) {
  constexpr int VEC = VecTraits<T>::W;
  __shared__ float red[SP_THREADS][2 * VEC + 1];
  const int cv = C / VEC;                    // divides SP_THREADS (checked by the launcher)
  const int ry = SP_THREADS / cv;
  const int tx = threadIdx.x % cv, ty = threadIdx.x / cv;
  const int c = tx * VEC;
  float s0[VEC], s1[VEC], mu[VEC];
#pragma unroll
  for (int e = 0; e < VEC; ++e) { s0[e] = 0.f; s1[e] = 0.f; mu[e] = mean[c + e]; }
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(r0 + rows_per_block, NP);
  for (int t = ty; t < (r1 - r0) * Wo; t += ry) {
    const int r = r0 + t / Wo, q = t % Wo;
    const int n = r / Ho, p = r - n * Ho;
    const size_t po = (size_t)n * Ho * Wo * C;
    const T* yimg = y + (size_t)n * H * W * C;
    uint4 xr[2][2];
#pragma unroll
    for (int dh = 0; dh < 2; ++dh)
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        const int h = min(2 * p + dh, H - 1), w = min(2 * q + dw, W - 1);
        xr[dh][dw] = __ldg(reinterpret_cast<const uint4*>(yimg + (h * W + w) * C + c));
      }
    uint32_t g[2][2][4];
    patch_gather<T, true>(dpool + po, amax + po, pooled + po, p, q, c, C, Ho, Wo, g);
#pragma unroll
    for (int dh = 0; dh < 2; ++dh)
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        if (2 * p + dh >= H || 2 * q + dw >= W) continue;
        float gv[VEC], xv[VEC];
        unpack16(make_uint4(g[dh][dw][0], g[dh][dw][1], g[dh][dw][2], g[dh][dw][3]), gv);
        unpack16(xr[dh][dw], xv);
#pragma unroll
        for (int e = 0; e < VEC; ++e) { s0[e] += gv[e]; s1[e] = fmaf(gv[e], xv[e] - mu[e], s1[e]); }
      }
  }
#pragma unroll
  for (int e = 0; e < VEC; ++e) { red[threadIdx.x][e] = s0[e]; red[threadIdx.x][VEC + e] = s1[e]; }
  __syncthreads();
  if (ty == 0) {
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      float a = 0.f, b = 0.f;
      for (int yy = 0; yy < ry; ++yy) { a += red[yy * cv + tx][e]; b += red[yy * cv + tx][VEC + e]; }
      part[((size_t)blockIdx.x * 2 + 0) * C + c + e] = a;
      part[((size_t)blockIdx.x * 2 + 1) * C + c + e] = b;
    }
  }
}

// dx pass: block = SP_ROWS patch rows of one image; dy = cA * g - cB * y + cK,  cK = cB * mean - Cc  (coef layout
// [3][C] of bn.cu: A, B, Cc).  HAS_BN = false: plain pooling backward, dy = g.
constexpr int SP_ROWS = 4;
template <typename T, bool HAS_BN>
__global__ void __launch_bounds__(SP_THREADS, 3)
maxpool_bn_bwd_dx_kernel(const T* __restrict__ dpool, const uint8_t* __restrict__ amax, const T* __restrict__ pooled,
                         const T* __restrict__ y, T* __restrict__ dy, const float* __restrict__ mean,
                         const float* __restrict__ coef, int H, int W, int C, int Ho, int Wo, int blocks_per_image) {
  constexpr int VEC = VecTraits<T>::W;
  const int cv = C / VEC;
  const int n = blockIdx.x / blocks_per_image, pb = (blockIdx.x - n * blocks_per_image) * SP_ROWS;
  const int rows = min(SP_ROWS, Ho - pb);
  const int c = (threadIdx.x % cv) * VEC;        // SP_THREADS % cv == 0: one channel vector per thread
  float cA[VEC], cB[VEC], cK[VEC];
  if (HAS_BN) {
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      cA[e] = coef[c + e];
      cB[e] = coef[C + c + e];
      cK[e] = fmaf(cB[e], mean[c + e], -coef[2 * C + c + e]);
    }
  }
  const size_t po = (size_t)n * Ho * Wo * C, io = (size_t)n * H * W * C;
  for (int t = threadIdx.x / cv; t < rows * Wo; t += SP_THREADS / cv) {
    const int p = pb + t / Wo, q = t % Wo;
    uint4 xr[2][2];
    if (HAS_BN) {
#pragma unroll
      for (int dh = 0; dh < 2; ++dh)
#pragma unroll
        for (int dw = 0; dw < 2; ++dw) {
          const int h = min(2 * p + dh, H - 1), w = min(2 * q + dw, W - 1);
          xr[dh][dw] = __ldg(reinterpret_cast<const uint4*>(y + io + (h * W + w) * C + c));
        }
    }
    uint32_t g[2][2][4];
    patch_gather<T, HAS_BN>(dpool + po, amax + po, pooled + po, p, q, c, C, Ho, Wo, g);
#pragma unroll
    for (int dh = 0; dh < 2; ++dh)
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        const int h = 2 * p + dh, w = 2 * q + dw;
        if (h >= H || w >= W) continue;
        const uint4 gp = make_uint4(g[dh][dw][0], g[dh][dw][1], g[dh][dw][2], g[dh][dw][3]);
        T* dst = dy + io + (h * W + w) * C + c;
        if (HAS_BN) {
          float gv[VEC], xv[VEC];
          unpack16(gp, gv);
          unpack16(xr[dh][dw], xv);
#pragma unroll
          for (int e = 0; e < VEC; ++e) xv[e] = fmaf(cA[e], gv[e], fmaf(-cB[e], xv[e], cK[e]));
          store_vec(dst, xv);
        } else {
          *reinterpret_cast<uint4*>(dst) = gp;
        }
      }
  }
}

static inline bool sp_shapes_ok(int N, int H, int W, int C, int vec) {
  if (N <= 0 || H <= 0 || W <= 0 || C <= 0 || C % vec != 0) return false;
  if (C / vec > SP_THREADS || SP_THREADS % (C / vec) != 0) return false;      // a thread keeps one channel vector
  const long long rows = (long long)N * H;
  return rows < (1LL << 30) && (long long)H * W * C < (1LL << 31);       // per-image offsets are 32-bit
}

template <typename T>
static int32_t stem_fwd_impl(const T* y, const float* scale, const float* shift, T* pooled, uint8_t* amax, int N, int H,
                             int W, int C, cudaStream_t st) {
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  if (scale) bn_relu_maxpool_fwd_kernel<T, true><<<N * Ho, SP_THREADS, 0, st>>>(y, scale, shift, pooled, amax, H, W, C, Ho, Wo);
  else bn_relu_maxpool_fwd_kernel<T, false><<<N * Ho, SP_THREADS, 0, st>>>(y, nullptr, nullptr, pooled, amax, H, W, C, Ho, Wo);
  return check_launch("bn_relu_maxpool_fwd");
}

template <typename T>
static int32_t stem_bwd_impl(const T* dpool, const uint8_t* amax, const T* pooled, const T* y, T* dy, int N, int H, int W,
                             int C, const float* gamma, const float* mean, const float* invstd, float* dgamma,
                             float* dbeta, float* ws, cudaStream_t st) {
  constexpr int VEC = VecTraits<T>::W;
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  const int NP = N * Ho;
  const int maxb = bn_max_partial_blocks();
  const int rpb = (NP + maxb - 1) / maxb;
  const int nblk = (NP + rpb - 1) / rpb;
  float* part = ws;
  float* coef = ws + (size_t)maxb * 2 * C;
  maxpool_bn_bwd_reduce_kernel<T><<<nblk, SP_THREADS, 0, st>>>(dpool, amax, pooled, y, mean, NP, H, W, C, Ho, Wo, rpb, part);
  if (int32_t rc = bn_finalize_bwd_launch(part, nblk, C, (double)N * H * W, gamma, invstd, dgamma, dbeta, coef, st)) return rc;
  const int bpi = (Ho + SP_ROWS - 1) / SP_ROWS;
  maxpool_bn_bwd_dx_kernel<T, true><<<N * bpi, SP_THREADS, 0, st>>>(dpool, amax, pooled, y, dy, mean, coef, H, W, C, Ho, Wo, bpi);
  return check_launch("maxpool_bn_bwd", 2);      // reduce + dx (the finalize launch counted itself)
}

}  // namespace b200

using namespace b200;

// y [N][H][W][C] (convolution output) -> pooled [N][Ho][Wo][C] = maxpool3x3s2p1(relu(y * scale + shift)), argmax one
// byte per pooled element (may be NULL).  scale == NULL: plain pooling of y.  C % 8 == 0 (bf16) / C % 4 == 0 (fp32).
extern "C" int32_t b200_bn_relu_maxpool_fwd(const void* y, const float* scale, const float* shift, void* pooled,
                                            uint8_t* argmax, int32_t N, int32_t H, int32_t W, int32_t C, int32_t dtype,
                                            void* stream) {
  if (!y || !pooled || (scale != nullptr) != (shift != nullptr)) return fail(B200_INVALID_ARG, "bn_relu_maxpool_fwd: bad arguments");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "bn_relu_maxpool_fwd: bad dtype");
  if (!sp_shapes_ok(N, H, W, C, dtype == B200_BF16 ? 8 : 4)) return fail(B200_UNSUPPORTED, "bn_relu_maxpool_fwd: unsupported shape");
  if ((reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(pooled) | reinterpret_cast<uintptr_t>(argmax)) & 15)
    return fail(B200_INVALID_ARG, "bn_relu_maxpool_fwd: tensors must be 16-byte aligned");
  if (dtype == B200_F32) return stem_fwd_impl<float>((const float*)y, scale, shift, (float*)pooled, argmax, N, H, W, C, as_stream(stream));
  return stem_fwd_impl<__nv_bfloat16>((const __nv_bfloat16*)y, scale, shift, (__nv_bfloat16*)pooled, argmax, N, H, W, C, as_stream(stream));
}

// Backward of the same: dpool, argmax, pooled [N][Ho][Wo][C]; y, dy [N][H][W][C]; BatchNorm batch statistics mean /
// invstd and gamma (NULL = 1); dgamma / dbeta may be NULL.  ws: b200_bn_workspace_bytes(N*H*W, C).
extern "C" int32_t b200_maxpool_bn_bwd(const void* dpool, const uint8_t* argmax, const void* pooled, const void* y, void* dy,
                                       int32_t N, int32_t H, int32_t W, int32_t C, int32_t dtype, const float* gamma,
                                       const float* mean, const float* invstd, float* dgamma, float* dbeta, void* ws,
                                       int64_t ws_bytes, void* stream) {
  if (!dpool || !argmax || !pooled || !y || !dy || !mean || !invstd) return fail(B200_INVALID_ARG, "maxpool_bn_bwd: bad arguments");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "maxpool_bn_bwd: bad dtype");
  if (!sp_shapes_ok(N, H, W, C, dtype == B200_BF16 ? 8 : 4)) return fail(B200_UNSUPPORTED, "maxpool_bn_bwd: unsupported shape");
  if (!ws || ws_bytes < b200_bn_workspace_bytes((int64_t)N * H * W, C)) return fail(B200_WORKSPACE_TOO_SMALL, "maxpool_bn_bwd: workspace too small");
  if ((reinterpret_cast<uintptr_t>(dpool) | reinterpret_cast<uintptr_t>(pooled) | reinterpret_cast<uintptr_t>(argmax) |
       reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(dy)) & 15)
    return fail(B200_INVALID_ARG, "maxpool_bn_bwd: tensors must be 16-byte aligned");
  if (dtype == B200_F32)
    return stem_bwd_impl<float>((const float*)dpool, argmax, (const float*)pooled, (const float*)y, (float*)dy, N, H, W, C, gamma, mean, invstd, dgamma, dbeta, (float*)ws, as_stream(stream));
  return stem_bwd_impl<__nv_bfloat16>((const __nv_bfloat16*)dpool, argmax, (const __nv_bfloat16*)pooled, (const __nv_bfloat16*)y, (__nv_bfloat16*)dy, N, H, W, C, gamma, mean, invstd, dgamma, dbeta, (float*)ws, as_stream(stream));
}

// MaxPool2d(3, 2, 1) backward alone, row-per-block indexing (replaces the 64-bit-index kernel of pool.cu on the
// vectorisable shapes).  Returns B200_UNSUPPORTED for shapes it does not take so that the caller can fall back.
extern "C" int32_t b200_maxpool3x3s2_bwd_rows(const void* dy, const uint8_t* argmax, void* dx, int32_t N, int32_t H,
                                              int32_t W, int32_t C, int32_t dtype, void* stream) {
  if (!dy || !argmax || !dx) return fail(B200_INVALID_ARG, "maxpool_bwd_rows: bad arguments");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "maxpool_bwd_rows: bad dtype");
  if (!sp_shapes_ok(N, H, W, C, dtype == B200_BF16 ? 8 : 4)) return B200_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx) | reinterpret_cast<uintptr_t>(argmax)) & 15) return B200_UNSUPPORTED;
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  const int bpi = (Ho + SP_ROWS - 1) / SP_ROWS;
  if (dtype == B200_F32)
    maxpool_bn_bwd_dx_kernel<float, false><<<N * bpi, SP_THREADS, 0, as_stream(stream)>>>((const float*)dy, argmax, nullptr, nullptr, (float*)dx, nullptr, nullptr, H, W, C, Ho, Wo, bpi);
  else
    maxpool_bn_bwd_dx_kernel<__nv_bfloat16, false><<<N * bpi, SP_THREADS, 0, as_stream(stream)>>>((const __nv_bfloat16*)dy, argmax, nullptr, nullptr, (__nv_bfloat16*)dx, nullptr, nullptr, H, W, C, Ho, Wo, bpi);
  return check_launch("maxpool_bwd_rows");
}
