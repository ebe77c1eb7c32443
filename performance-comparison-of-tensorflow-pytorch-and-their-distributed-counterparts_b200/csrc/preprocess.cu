// Input pipeline on the GPU: the reference's per-image transform
//     transforms.Compose([transforms.Resize((224, 224)), transforms.ToTensor()])
// (pytorch_training_inference_on_image.ipynb:325-326; another_neural_net.py:38-46) for a batch of decoded images:
// uint8 HWC pixels -> bilinear resize with PIL's antialiasing (triangle filter widened by the down-scale factor,
// pixel centres at i + 0.5, PIL's own fixed-point arithmetic) -> x / 255 -> either
//   * the stem's own input layout, NHWC with 4 channels (8-byte pixels) in bf16 -- nothing left for the first
//     kernel of the model to convert, and the host -> device copy carries 3 bytes per source pixel instead of 12 per
//     resized one, or
//   * NCHW fp32, exactly the tensor ToTensor() hands the reference's loop.
// JPEG entropy decoding stays on the host (PIL); everything after it runs here.  Images of one batch may have
// different sizes (device tables of pointers / heights / widths / row strides).  One thread per output pixel.
#include "common.cuh"

namespace b200 {

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// Pillow's 8-bit resampling (src/libImaging/Resample.c), restated: per axis, the triangle filter widened by the
// down-scale factor, weights normalised in double precision and quantised to 22-bit fixed point, products summed in
// an int that starts at one half, then shifted down and clipped to a byte -- horizontal pass first, a byte image in
// between, vertical pass second.  Restating the integer arithmetic (rather than "the same filter in fp32") is what
// makes the output agree with the reference transform byte for byte.
constexpr int PRE_BITS = 22;
constexpr int PRE_MAX_TAPS = 96;      // taps per axis held per thread: down-scale factors up to ~47

struct AxisTaps {
  int lo, n;
};
__device__ __forceinline__ AxisTaps axis_coeffs(int o, int in, int out, int antialias, int* __restrict__ k) {
  const double scale = (double)in / (double)out;
  const double fscale = (antialias && scale > 1.0) ? scale : 1.0;
  const double support = 1.0 * fscale;                 // bilinear: support 1
  const double center = ((double)o + 0.5) * scale;
  const double ss = 1.0 / fscale;
  int lo = (int)(center - support + 0.5);
  if (lo < 0) lo = 0;
  int hi = (int)(center + support + 0.5);
  if (hi > in) hi = in;
  int n = hi - lo;
  if (n > PRE_MAX_TAPS) n = PRE_MAX_TAPS;
  double ww = 0.0;
  for (int j = 0; j < n; ++j) {
    double x = ((double)(j + lo) - center + 0.5) * ss;
    x = x < 0 ? -x : x;
    const double w = x < 1.0 ? 1.0 - x : 0.0;
    ww += w;
  }
  for (int j = 0; j < n; ++j) {
    double x = ((double)(j + lo) - center + 0.5) * ss;
    x = x < 0 ? -x : x;
    const double w = ww != 0.0 ? (x < 1.0 ? 1.0 - x : 0.0) / ww : 0.0;    // a true division, as Pillow does
    k[j] = (int)(0.5 + w * (double)(1 << PRE_BITS));   // weights are non-negative for this filter
  }
  AxisTaps t;
  t.lo = lo;
  t.n = n;
  return t;
}
__device__ __forceinline__ int clip8(int v) {
  v >>= PRE_BITS;
  return v < 0 ? 0 : (v > 255 ? 255 : v);
}

template <int MODE>   // 0: NHWC4 bf16, 1: NCHW fp32
__global__ void __launch_bounds__(128)
preprocess_kernel(const uint8_t* const* __restrict__ ptrs, const int* __restrict__ hs, const int* __restrict__ ws,
                  const long long* __restrict__ strides, int N, int Ho, int Wo, void* __restrict__ out, int antialias) {
  const long long total = (long long)N * Ho * Wo;
  int kx[PRE_MAX_TAPS], ky[PRE_MAX_TAPS];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int ox = (int)(i % Wo);
    const long long t = i / Wo;
    const int oy = (int)(t % Ho);
    const int n = (int)(t / Ho);
    const uint8_t* src = ptrs[n];
    const int H = hs[n], W = ws[n];
    const long long rs = strides[n];
    const AxisTaps tx = axis_coeffs(ox, W, Wo, antialias, kx), ty = axis_coeffs(oy, H, Ho, antialias, ky);
    int a0 = 1 << (PRE_BITS - 1), a1 = a0, a2 = a0;
    for (int jy = 0; jy < ty.n; ++jy) {
      const uint8_t* row = src + (long long)(jy + ty.lo) * rs + (long long)tx.lo * 3;
      int r0 = 1 << (PRE_BITS - 1), r1 = r0, r2 = r0;
      for (int jx = 0; jx < tx.n; ++jx) {
        const int k = kx[jx];
        r0 += (int)row[3 * jx + 0] * k;
        r1 += (int)row[3 * jx + 1] * k;
        r2 += (int)row[3 * jx + 2] * k;
      }
      const int k = ky[jy];
      a0 += clip8(r0) * k;       // the byte image between the two passes
      a1 += clip8(r1) * k;
      a2 += clip8(r2) * k;
    }
    const float v0 = __fdiv_rn((float)clip8(a0), 255.f);    // ToTensor(): byte / 255 (a division, not a reciprocal)
    const float v1 = __fdiv_rn((float)clip8(a1), 255.f);
    const float v2 = __fdiv_rn((float)clip8(a2), 255.f);
    if (MODE == 0) {
      uint2 px;
      px.x = pack_bf16(v0, v1);
      px.y = pack_bf16(v2, 0.f);
      reinterpret_cast<uint2*>(out)[i] = px;
    } else {
      float* o = reinterpret_cast<float*>(out);
      const long long plane = (long long)Ho * Wo;
      const long long base = (long long)n * 3 * plane + (long long)oy * Wo + ox;
      o[base] = v0;
      o[base + plane] = v1;
      o[base + 2 * plane] = v2;
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" int32_t b200_preprocess_u8(const void* const* image_ptrs, const int32_t* heights, const int32_t* widths,
                                      const int64_t* row_strides, int32_t N, int32_t Ho, int32_t Wo, void* out,
                                      int32_t out_layout, int32_t antialias, void* stream) {
  if (!image_ptrs || !heights || !widths || !row_strides || !out || N <= 0 || Ho <= 0 || Wo <= 0)
    return fail(B200_INVALID_ARG, "preprocess_u8: bad arguments");
  if (out_layout != 0 && out_layout != 1) return fail(B200_INVALID_ARG, "preprocess_u8: out_layout 0 (NHWC4 bf16) or 1 (NCHW fp32)");
  const long long total = (long long)N * Ho * Wo;
  long long blocks = (total + 127) / 128;
  const long long cap = (long long)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = as_stream(stream);
  if (out_layout == 0)
    preprocess_kernel<0><<<(int)blocks, 128, 0, st>>>((const uint8_t* const*)image_ptrs, heights, widths, (const long long*)row_strides, N, Ho, Wo, out, antialias);
  else
    preprocess_kernel<1><<<(int)blocks, 128, 0, st>>>((const uint8_t* const*)image_ptrs, heights, widths, (const long long*)row_strides, N, Ho, Wo, out, antialias);
  return check_launch("preprocess_u8");
}
