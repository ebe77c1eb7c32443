// Shared helpers for the b200net C-ABI library: status codes, last-error string, typed vector access.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/b200net.h"
#include "device_state.cuh"

namespace b200 {

// thread-local: torch's autograd engine calls us from its own worker thread
inline char* last_error_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}
inline int32_t fail(int32_t code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}
// number of kernels this library has launched (bench.py reports the delta over the timed region)
long long& launch_counter();
inline int32_t check_launch(const char* what, int kernels = 1) {
  launch_counter() += kernels;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(B200_CUDA_ERROR, "%s: %s", what, cudaGetErrorString(e));
  return B200_OK;
}
inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
// Optional device-resident dropout call counter of the calling thread (b200_dropout_counter): when set, every
// dropout site adds (*ctr) << 40 to its Philox offset at EXECUTION time, so a captured CUDA graph draws a fresh
// mask at every replay.
inline const unsigned long long*& dropout_ctr() {
  static thread_local const unsigned long long* p = nullptr;
  return p;
}

template <typename T> struct VecTraits;
template <> struct VecTraits<float> { static constexpr int W = 4; using V = float4; };
template <> struct VecTraits<__nv_bfloat16> { static constexpr int W = 8; using V = uint4; };

// load W consecutive elements as floats
__device__ __forceinline__ void load_vec(const float* p, float (&f)[4]) {
  float4 v = *reinterpret_cast<const float4*>(p);
  f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
}
__device__ __forceinline__ void load_vec(const __nv_bfloat16* p, float (&f)[8]) {
  uint4 v = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ void store_vec(float* p, const float (&f)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
}
__device__ __forceinline__ void store_vec(__nv_bfloat16* p, const float (&f)[8]) {
  uint4 v;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = v;
}
__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace b200

