// MaxPool2d(3, stride 2, pad 1) and AdaptiveAvgPool2d((1,1)) over NHWC activations.
// max-pool keeps a uint8 window index per output element (first maximum in scan order, like ATen) so the backward
// is a pure gather with no atomics.
#include "common.cuh"
#include <math_constants.h>

namespace b200 {

template <typename T>
__global__ void maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ amax, int N,
                                   int H, int W, int C, int Ho, int Wo) {
  const long long total = (long long)N * Ho * Wo * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long t = i / C;
    const int q = (int)(t % Wo);
    t /= Wo;
    const int p = (int)(t % Ho);
    const int n = (int)(t / Ho);
    float best = -CUDART_INF_F;
    int bi = 0;
    bool any = false;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = 2 * p - 1 + r;
      if (h < 0 || h >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int w = 2 * q - 1 + s;
        if (w < 0 || w >= W) continue;
        const float v = to_f(x[(((long long)n * H + h) * W + w) * C + c]);
        if (!any || v > best || v != v) {
          best = v;
          bi = r * 3 + s;
          any = true;
        }
      }
    }
    y[i] = from_f<T>(best);
    if (amax) amax[i] = (uint8_t)bi;
  }
}

template <typename T>
__global__ void maxpool_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ amax, T* __restrict__ dx,
                                   int N, int H, int W, int C, int Ho, int Wo) {
  const long long total = (long long)N * H * W * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long t = i / C;
    const int w = (int)(t % W);
    t /= W;
    const int h = (int)(t % H);
    const int n = (int)(t / H);
    float acc = 0.f;
    // windows p with 2p-1 <= h <= 2p+1
    const int p_lo = h / 2, p_hi = (h + 1) / 2;
    const int q_lo = w / 2, q_hi = (w + 1) / 2;
    for (int p = p_lo; p <= p_hi; ++p) {
      if (p >= Ho) continue;
      const int r = h - (2 * p - 1);
      for (int q = q_lo; q <= q_hi; ++q) {
        if (q >= Wo) continue;
        const int s = w - (2 * q - 1);
        const long long o = (((long long)n * Ho + p) * Wo + q) * C + c;
        if (amax[o] == r * 3 + s) acc += to_f(dy[o]);
      }
    }
    dx[i] = from_f<T>(acc);
  }
}

// vectorised variants (C % VEC == 0, 16-byte aligned): one thread = VEC channels of one pixel
template <typename T, int VEC>
__global__ void maxpool_fwd_vec_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ amax, int N,
                                       int H, int W, int C, int Ho, int Wo) {
  const int cv = C / VEC;
  const long long total = (long long)N * Ho * Wo * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * VEC;
    long long t = i / cv;
    const int q = (int)(t % Wo);
    t /= Wo;
    const int p = (int)(t % Ho);
    const int n = (int)(t / Ho);
    float best[VEC];
    int bi[VEC];
    bool any = false;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = 2 * p - 1 + r;
      if (h < 0 || h >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int w = 2 * q - 1 + s;
        if (w < 0 || w >= W) continue;
        float v[VEC];
        load_vec(x + (((long long)n * H + h) * W + w) * C + c, v);
#pragma unroll
        for (int e = 0; e < VEC; ++e)
          if (!any || v[e] > best[e] || v[e] != v[e]) { best[e] = v[e]; bi[e] = r * 3 + s; }
        any = true;
      }
    }
    const long long o = (((long long)n * Ho + p) * Wo + q) * C + c;
    store_vec(y + o, best);
    if (amax) {
#pragma unroll
      for (int e = 0; e < VEC; ++e) amax[o + e] = (uint8_t)bi[e];
    }
  }
}

template <typename T, int VEC>
__global__ void maxpool_bwd_vec_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ amax,
                                       T* __restrict__ dx, int N, int H, int W, int C, int Ho, int Wo) {
  const int cv = C / VEC;
  const long long total = (long long)N * H * W * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * VEC;
    long long t = i / cv;
    const int w = (int)(t % W);
    t /= W;
    const int h = (int)(t % H);
    const int n = (int)(t / H);
    float acc[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] = 0.f;
    const int p_lo = h / 2, p_hi = (h + 1) / 2, q_lo = w / 2, q_hi = (w + 1) / 2;
    for (int p = p_lo; p <= p_hi; ++p) {
      if (p >= Ho) continue;
      const int r = h - (2 * p - 1);
      for (int q = q_lo; q <= q_hi; ++q) {
        if (q >= Wo) continue;
        const int want = r * 3 + (w - (2 * q - 1));
        const long long o = (((long long)n * Ho + p) * Wo + q) * C + c;
        uint8_t idx[VEC];
        if (VEC == 8) *reinterpret_cast<uint2*>(idx) = *reinterpret_cast<const uint2*>(amax + o);
        else *reinterpret_cast<uint32_t*>(idx) = *reinterpret_cast<const uint32_t*>(amax + o);
        float g[VEC];
        load_vec(dy + o, g);
#pragma unroll
        for (int e = 0; e < VEC; ++e)
          if (idx[e] == want) acc[e] += g[e];
      }
    }
    store_vec(dx + (((long long)n * H + h) * W + w) * C + c, acc);
  }
}

template <typename T>
__global__ void avgpool_fwd_kernel(const T* __restrict__ x, float* __restrict__ y, int N, int HW, int C) {
  const long long total = (long long)N * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long n = i / C;
    float acc = 0.f;
    for (int j = 0; j < HW; ++j) acc += to_f(x[(n * HW + j) * C + c]);
    y[i] = acc / (float)HW;
  }
}
template <typename T>
__global__ void avgpool_bwd_kernel(const float* __restrict__ dy, T* __restrict__ dx, int N, int HW, int C) {
  const long long total = (long long)N * HW * C;
  const float inv = 1.f / (float)HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long n = i / ((long long)HW * C);
    dx[i] = from_f<T>(dy[n * C + c] * inv);
  }
}

// eight channels per thread, 16-byte stores, 32-bit indices (the scalar kernel above: 41 us for the 26 MB gradient of the
// N = 256 step)
__global__ void __launch_bounds__(256) avgpool_bwd_bf16x8_kernel(const float* __restrict__ dy, uint4* __restrict__ dx, int N,
                                                                 int HW, int C8) {
  const int total = N * HW * C8;
  const float inv = 1.f / (float)HW;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c8 = i % C8, n = i / (HW * C8);
    const float4 a = reinterpret_cast<const float4*>(dy)[(n * C8 + c8) * 2];
    const float4 b = reinterpret_cast<const float4*>(dy)[(n * C8 + c8) * 2 + 1];
    uint4 v;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
    h[0] = __floats2bfloat162_rn(a.x * inv, a.y * inv); h[1] = __floats2bfloat162_rn(a.z * inv, a.w * inv);
    h[2] = __floats2bfloat162_rn(b.x * inv, b.y * inv); h[3] = __floats2bfloat162_rn(b.z * inv, b.w * inv);
    dx[i] = v;
  }
}

static inline int pgrid(long long total) {
  long long b = (total + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (b > cap) b = cap;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace b200

using namespace b200;

extern "C" int32_t b200_maxpool3x3s2_fwd(const void* x, void* y, uint8_t* argmax, int32_t N, int32_t H, int32_t W,
                                         int32_t C, int32_t dtype, void* stream) {
  if (!x || !y || N <= 0 || H <= 0 || W <= 0 || C <= 0) return fail(B200_INVALID_ARG, "maxpool_fwd: bad arguments");
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  const long long total = (long long)N * Ho * Wo * C;
  cudaStream_t st = as_stream(stream);
  const bool al = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(argmax)) & 15) == 0;
  // vectorisable shapes: the row-per-block kernel of stem_pool.cu (32-bit indexing)
  const int vecw = dtype == B200_BF16 ? 8 : 4;
  if (al && (dtype == B200_F32 || dtype == B200_BF16) && C % vecw == 0 && C / vecw <= 256 && 256 % (C / vecw) == 0 &&
      (long long)N * H < (1LL << 30))
    return b200_bn_relu_maxpool_fwd(x, nullptr, nullptr, y, argmax, N, H, W, C, dtype, stream);
  if (dtype == B200_F32 && al && C % 4 == 0) maxpool_fwd_vec_kernel<float, 4><<<pgrid(total / 4), 256, 0, st>>>((const float*)x, (float*)y, argmax, N, H, W, C, Ho, Wo);
  else if (dtype == B200_BF16 && al && C % 8 == 0) maxpool_fwd_vec_kernel<__nv_bfloat16, 8><<<pgrid(total / 8), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, argmax, N, H, W, C, Ho, Wo);
  else if (dtype == B200_F32) maxpool_fwd_kernel<float><<<pgrid(total), 256, 0, st>>>((const float*)x, (float*)y, argmax, N, H, W, C, Ho, Wo);
  else if (dtype == B200_BF16) maxpool_fwd_kernel<__nv_bfloat16><<<pgrid(total), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, argmax, N, H, W, C, Ho, Wo);
  else return fail(B200_INVALID_ARG, "maxpool_fwd: bad dtype");
  return check_launch("maxpool_fwd");
}

extern "C" int32_t b200_maxpool3x3s2_bwd(const void* dy, const uint8_t* argmax, void* dx, int32_t N, int32_t H,
                                         int32_t W, int32_t C, int32_t dtype, void* stream) {
  if (!dy || !argmax || !dx || N <= 0 || H <= 0 || W <= 0 || C <= 0) return fail(B200_INVALID_ARG, "maxpool_bwd: bad arguments");
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  const long long total = (long long)N * H * W * C;
  cudaStream_t st = as_stream(stream);
  const bool al = ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx) | reinterpret_cast<uintptr_t>(argmax)) & 15) == 0;
  {
    const int32_t rc = b200_maxpool3x3s2_bwd_rows(dy, argmax, dx, N, H, W, C, dtype, stream);   // stem_pool.cu
    if (rc != B200_UNSUPPORTED) return rc;
  }
  if (dtype == B200_F32 && al && C % 4 == 0) maxpool_bwd_vec_kernel<float, 4><<<pgrid(total / 4), 256, 0, st>>>((const float*)dy, argmax, (float*)dx, N, H, W, C, Ho, Wo);
  else if (dtype == B200_BF16 && al && C % 8 == 0) maxpool_bwd_vec_kernel<__nv_bfloat16, 8><<<pgrid(total / 8), 256, 0, st>>>((const __nv_bfloat16*)dy, argmax, (__nv_bfloat16*)dx, N, H, W, C, Ho, Wo);
  else if (dtype == B200_F32) maxpool_bwd_kernel<float><<<pgrid(total), 256, 0, st>>>((const float*)dy, argmax, (float*)dx, N, H, W, C, Ho, Wo);
  else if (dtype == B200_BF16) maxpool_bwd_kernel<__nv_bfloat16><<<pgrid(total), 256, 0, st>>>((const __nv_bfloat16*)dy, argmax, (__nv_bfloat16*)dx, N, H, W, C, Ho, Wo);
  else return fail(B200_INVALID_ARG, "maxpool_bwd: bad dtype");
  return check_launch("maxpool_bwd");
}

extern "C" int32_t b200_avgpool_fwd(const void* x, float* y, int32_t N, int32_t HW, int32_t C, int32_t dtype,
                                    void* stream) {
  if (!x || !y || N <= 0 || HW <= 0 || C <= 0) return fail(B200_INVALID_ARG, "avgpool_fwd: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (dtype == B200_F32) avgpool_fwd_kernel<float><<<pgrid((long long)N * C), 256, 0, st>>>((const float*)x, y, N, HW, C);
  else if (dtype == B200_BF16) avgpool_fwd_kernel<__nv_bfloat16><<<pgrid((long long)N * C), 256, 0, st>>>((const __nv_bfloat16*)x, y, N, HW, C);
  else return fail(B200_INVALID_ARG, "avgpool_fwd: bad dtype");
  return check_launch("avgpool_fwd");
}

extern "C" int32_t b200_avgpool_bwd(const float* dy, void* dx, int32_t N, int32_t HW, int32_t C, int32_t dtype,
                                    void* stream) {
  if (!dy || !dx || N <= 0 || HW <= 0 || C <= 0) return fail(B200_INVALID_ARG, "avgpool_bwd: bad arguments");
  cudaStream_t st = as_stream(stream);
  const long long total = (long long)N * HW * C;
  if (dtype == B200_BF16 && C % 8 == 0 && total < (1LL << 31) &&
      ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx)) & 15) == 0) {
    avgpool_bwd_bf16x8_kernel<<<pgrid(total / 8), 256, 0, st>>>(dy, (uint4*)dx, N, HW, C / 8);
    return check_launch("avgpool_bwd");
  }
  if (dtype == B200_F32) avgpool_bwd_kernel<float><<<pgrid(total), 256, 0, st>>>(dy, (float*)dx, N, HW, C);
  else if (dtype == B200_BF16) avgpool_bwd_kernel<__nv_bfloat16><<<pgrid(total), 256, 0, st>>>(dy, (__nv_bfloat16*)dx, N, HW, C);
  else return fail(B200_INVALID_ARG, "avgpool_bwd: bad dtype");
  return check_launch("avgpool_bwd");
}
