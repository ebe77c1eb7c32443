// VGG16 path (`models.vgg16` at pytorch_training_inference_on_image.ipynb:1968-2041,
// another_neural_net.py:219-381, Standalone_Inference_Imagenette_trial.ipynb:355-374) -- the kernels it needs beyond
// the ResNet / BERT set: MaxPool2d(2, 2), AdaptiveAvgPool2d((7, 7)) and Dropout on activation matrices in the storage
// dtype.  The 3x3 convolutions (+bias +ReLU in the epilogue) and the classifier's Linears run on the tcgen05
// implicit-GEMM kernel; these are the HBM-bound leftovers, NHWC, 16-byte vectors.
#include "common.cuh"
#include "philox.cuh"

namespace b200 {

static inline int vgrid(long long n, int threads = 256) {
  const long long want = (n + threads - 1) / threads;
  const long long cap = (long long)num_sms() * 16;
  return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
}

// ---------------------------------------------------------------- MaxPool2d(kernel 2, stride 2), floor mode
// y[n,p,q,c] = max over the 2x2 window; argmax byte = window position (first maximum wins, as in ATen: NaN-free data)
template <typename T>
__global__ void maxpool2_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ arg, int N, int H,
                                    int W, int C, int Ho, int Wo) {
  constexpr int VW = VecTraits<T>::W;
  const int cv = C / VW;
  const long long total = (long long)N * Ho * Wo * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * VW;
    long long r = i / cv;
    const int q = (int)(r % Wo);
    r /= Wo;
    const int p = (int)(r % Ho);
    const int n = (int)(r / Ho);
    float best[VW];
    uint8_t idx[VW];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float v[VW];
      load_vec(x + (((long long)n * H + 2 * p + (k >> 1)) * W + 2 * q + (k & 1)) * C + c, v);
#pragma unroll
      for (int e = 0; e < VW; ++e)
        if (k == 0 || v[e] > best[e]) {
          best[e] = v[e];
          idx[e] = (uint8_t)k;
        }
    }
    const long long o = (((long long)n * Ho + p) * Wo + q) * C + c;
    store_vec(y + o, best);
    if (arg != nullptr) {
#pragma unroll
      for (int e = 0; e < VW; ++e) arg[o + e] = idx[e];
    }
  }
}
// every input pixel belongs to at most one window: no atomics; rows / columns past 2*Ho, 2*Wo get zero
template <typename T>
__global__ void maxpool2_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ arg, T* __restrict__ dx, int N,
                                    int H, int W, int C, int Ho, int Wo) {
  constexpr int VW = VecTraits<T>::W;
  const int cv = C / VW;
  const long long total = (long long)N * H * W * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * VW;
    long long r = i / cv;
    const int w = (int)(r % W);
    r /= W;
    const int h = (int)(r % H);
    const int n = (int)(r / H);
    float g[VW];
#pragma unroll
    for (int e = 0; e < VW; ++e) g[e] = 0.f;
    const int p = h >> 1, q = w >> 1;
    if (p < Ho && q < Wo) {
      const long long o = (((long long)n * Ho + p) * Wo + q) * C + c;
      const int k = ((h & 1) << 1) | (w & 1);
      float d[VW];
      load_vec(dy + o, d);
#pragma unroll
      for (int e = 0; e < VW; ++e) g[e] = arg[o + e] == k ? d[e] : 0.f;
    }
    store_vec(dx + i * VW, g);
  }
}

// ---------------------------------------------------------------- AdaptiveAvgPool2d((OH, OW))
// window of output i along a dimension of size L: [floor(i*L/O), ceil((i+1)*L/O))  (ATen's start/end index rule)
__device__ __forceinline__ int ap_start(int i, int L, int O) { return (i * L) / O; }
__device__ __forceinline__ int ap_end(int i, int L, int O) { return ((i + 1) * L + O - 1) / O; }

template <typename T>
__global__ void adaptive_avgpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int H, int W, int C,
                                            int OH, int OW) {
  const long long total = (long long)N * OH * OW * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long r = i / C;
    const int ow = (int)(r % OW);
    r /= OW;
    const int oh = (int)(r % OH);
    const int n = (int)(r / OH);
    const int h0 = ap_start(oh, H, OH), h1 = ap_end(oh, H, OH), w0 = ap_start(ow, W, OW), w1 = ap_end(ow, W, OW);
    float s = 0.f;
    for (int h = h0; h < h1; ++h)
      for (int w = w0; w < w1; ++w) s += to_f(x[(((long long)n * H + h) * W + w) * C + c]);
    y[i] = from_f<T>(s / (float)((h1 - h0) * (w1 - w0)));
  }
}
template <typename T>
__global__ void adaptive_avgpool_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int N, int H, int W, int C,
                                            int OH, int OW) {
  const long long total = (long long)N * H * W * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long r = i / C;
    const int w = (int)(r % W);
    r /= W;
    const int h = (int)(r % H);
    const int n = (int)(r / H);
    float s = 0.f;
    for (int oh = 0; oh < OH; ++oh) {
      const int h0 = ap_start(oh, H, OH), h1 = ap_end(oh, H, OH);
      if (h < h0 || h >= h1) continue;
      for (int ow = 0; ow < OW; ++ow) {
        const int w0 = ap_start(ow, W, OW), w1 = ap_end(ow, W, OW);
        if (w < w0 || w >= w1) continue;
        s += to_f(dy[(((long long)n * OH + oh) * OW + ow) * C + c]) / (float)((h1 - h0) * (w1 - w0));
      }
    }
    dx[i] = from_f<T>(s);
  }
}

// ---------------------------------------------------------------- Dropout on an activation matrix (storage dtype)
// mask from Philox (seed, offset, element index), eight elements per call; the backward regenerates it.
template <typename T>
__global__ void dropout_act_kernel(const T* __restrict__ x, T* __restrict__ y, long long nvec8, float p, float inv_keep,
                                   uint64_t seed, uint64_t offset, const unsigned long long* __restrict__ ctr) {
  if (ctr != nullptr) offset += *ctr << 40;
  const uint32_t thresh = dropout_thresh16(p);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec8;
       i += (long long)gridDim.x * blockDim.x) {
    const uint32_t bits = dropout_keep8(seed, offset, i * 8, thresh);
    constexpr int VW = VecTraits<T>::W;
#pragma unroll
    for (int h = 0; h < 8 / VW; ++h) {
      float v[VW];
      load_vec(x + i * 8 + h * VW, v);
#pragma unroll
      for (int e = 0; e < VW; ++e) v[e] = ((bits >> (h * VW + e)) & 1u) ? v[e] * inv_keep : 0.f;
      store_vec(y + i * 8 + h * VW, v);
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

int32_t b200_maxpool2x2_fwd(const void* x, void* y, uint8_t* argmax, int32_t N, int32_t H, int32_t W, int32_t C,
                            int32_t dtype, void* stream) {
  if (!x || !y || N <= 0 || H < 2 || W < 2 || C <= 0) return fail(B200_INVALID_ARG, "maxpool2x2_fwd: bad arguments");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "maxpool2x2_fwd: bad dtype");
  const int VW = dtype == B200_F32 ? 4 : 8;
  if (C % VW != 0) return fail(B200_UNSUPPORTED, "maxpool2x2_fwd: channels must be a multiple of %d", VW);
  cudaStream_t st = as_stream(stream);
  const int Ho = H / 2, Wo = W / 2;
  const int grid = vgrid((long long)N * Ho * Wo * (C / VW));
  if (dtype == B200_F32)
    maxpool2_fwd_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (float*)y, argmax, N, H, W, C, Ho, Wo);
  else
    maxpool2_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, argmax, N, H,
                                                            W, C, Ho, Wo);
  return check_launch("maxpool2x2_fwd");
}

int32_t b200_maxpool2x2_bwd(const void* dy, const uint8_t* argmax, void* dx, int32_t N, int32_t H, int32_t W,
                            int32_t C, int32_t dtype, void* stream) {
  if (!dy || !argmax || !dx || N <= 0 || H < 2 || W < 2 || C <= 0)
    return fail(B200_INVALID_ARG, "maxpool2x2_bwd: bad arguments");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "maxpool2x2_bwd: bad dtype");
  const int VW = dtype == B200_F32 ? 4 : 8;
  if (C % VW != 0) return fail(B200_UNSUPPORTED, "maxpool2x2_bwd: channels must be a multiple of %d", VW);
  cudaStream_t st = as_stream(stream);
  const int grid = vgrid((long long)N * H * W * (C / VW));
  if (dtype == B200_F32)
    maxpool2_bwd_kernel<float><<<grid, 256, 0, st>>>((const float*)dy, argmax, (float*)dx, N, H, W, C, H / 2, W / 2);
  else
    maxpool2_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dy, argmax, (__nv_bfloat16*)dx, N, H,
                                                            W, C, H / 2, W / 2);
  return check_launch("maxpool2x2_bwd");
}

int32_t b200_adaptive_avgpool_fwd(const void* x, void* y, int32_t N, int32_t H, int32_t W, int32_t C, int32_t OH,
                                  int32_t OW, int32_t dtype, void* stream) {
  if (!x || !y || N <= 0 || H <= 0 || W <= 0 || C <= 0 || OH <= 0 || OW <= 0)
    return fail(B200_INVALID_ARG, "adaptive_avgpool_fwd: bad arguments");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "adaptive_avgpool_fwd: bad dtype");
  cudaStream_t st = as_stream(stream);
  const int grid = vgrid((long long)N * OH * OW * C);
  if (dtype == B200_F32)
    adaptive_avgpool_fwd_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (float*)y, N, H, W, C, OH, OW);
  else
    adaptive_avgpool_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, N, H, W,
                                                                    C, OH, OW);
  return check_launch("adaptive_avgpool_fwd");
}

int32_t b200_adaptive_avgpool_bwd(const void* dy, void* dx, int32_t N, int32_t H, int32_t W, int32_t C, int32_t OH,
                                  int32_t OW, int32_t dtype, void* stream) {
  if (!dy || !dx || N <= 0 || H <= 0 || W <= 0 || C <= 0 || OH <= 0 || OW <= 0)
    return fail(B200_INVALID_ARG, "adaptive_avgpool_bwd: bad arguments");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "adaptive_avgpool_bwd: bad dtype");
  cudaStream_t st = as_stream(stream);
  const int grid = vgrid((long long)N * H * W * C);
  if (dtype == B200_F32)
    adaptive_avgpool_bwd_kernel<float><<<grid, 256, 0, st>>>((const float*)dy, (float*)dx, N, H, W, C, OH, OW);
  else
    adaptive_avgpool_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, N, H,
                                                                    W, C, OH, OW);
  return check_launch("adaptive_avgpool_bwd");
}

int32_t b200_dropout_act(const void* x, void* y, int64_t n, float p_drop, uint64_t seed, uint64_t offset,
                         int32_t dtype, void* stream) {
  if (!x || !y || n < 0 || p_drop < 0.f || p_drop >= 1.f) return fail(B200_INVALID_ARG, "dropout_act: bad arguments");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "dropout_act: bad dtype");
  if (n % 8 != 0) return fail(B200_UNSUPPORTED, "dropout_act: element count must be a multiple of 8");
  if (n == 0) return B200_OK;
  cudaStream_t st = as_stream(stream);
  const int grid = vgrid(n / 8);
  const float ik = dropout_inv_keep(p_drop);
  if (dtype == B200_F32)
    dropout_act_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (float*)y, n / 8, p_drop, ik, seed, offset,
                                                   dropout_ctr());
  else
    dropout_act_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n / 8, p_drop,
                                                           ik, seed, offset, dropout_ctr());
  return check_launch("dropout_act");
}

}  // extern "C"
