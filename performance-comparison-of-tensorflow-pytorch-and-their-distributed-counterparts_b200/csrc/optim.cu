// Multi-tensor optimizer / gradient utilities (fp32): SGD, Adam/AdamW, L2-norm, clip scale, zero, bucket
// pack/unpack.  Tensor pointer tables are host arrays; each launch carries up to MT_MAX tensors BY VALUE in the
// kernel parameter block (no device-side table, no allocation), grid = (blocks, tensors).
#include "common.cuh"

namespace b200 {

constexpr int MT_MAX = 48;
constexpr int MT_THREADS = 256;

struct MtArgs {
  void* a[MT_MAX];
  void* b[MT_MAX];
  void* c[MT_MAX];
  void* d[MT_MAX];
  long long n[MT_MAX];
};

struct SgdHyper { float lr, momentum, dampening, wd, grad_scale; int nesterov, first_step, has_mom; };
__global__ void __launch_bounds__(MT_THREADS) mt_sgd_kernel(MtArgs t, SgdHyper h) {
  const int ti = blockIdx.y;
  float* __restrict__ p = (float*)t.a[ti];
  const float* __restrict__ g = (const float*)t.b[ti];
  float* __restrict__ m = (float*)t.c[ti];
  const long long n = t.n[ti];
  auto update = [&](float& pi, float gi, float& mi) {
    gi *= h.grad_scale;
    if (h.wd != 0.f) gi = fmaf(h.wd, pi, gi);
    if (h.has_mom) {
      mi = h.first_step ? gi : fmaf(h.momentum, mi, (1.f - h.dampening) * gi);
      gi = h.nesterov ? fmaf(h.momentum, mi, gi) : mi;
    }
    pi = fmaf(-h.lr, gi, pi);
  };
  long long done = 0;
  if (((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | (h.has_mom ? reinterpret_cast<uintptr_t>(m) : 0)) & 15) == 0) {
    const long long nv = n >> 2;       // 16-byte accesses
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
      float4 p4 = reinterpret_cast<float4*>(p)[i];
      const float4 g4 = reinterpret_cast<const float4*>(g)[i];
      float4 m4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (h.has_mom) m4 = reinterpret_cast<float4*>(m)[i];
      update(p4.x, g4.x, m4.x);
      update(p4.y, g4.y, m4.y);
      update(p4.z, g4.z, m4.z);
      update(p4.w, g4.w, m4.w);
      if (h.has_mom) reinterpret_cast<float4*>(m)[i] = m4;
      reinterpret_cast<float4*>(p)[i] = p4;
    }
    done = nv << 2;
  }
  for (long long i = done + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float pi = p[i], mi = h.has_mom ? m[i] : 0.f;
    update(pi, g[i], mi);
    if (h.has_mom) m[i] = mi;
    p[i] = pi;
  }
}

struct AdamHyper { float lr, beta1, beta2, eps, wd, step_size, bc2_sqrt, grad_scale; int decoupled; };
__global__ void __launch_bounds__(MT_THREADS) mt_adam_kernel(MtArgs t, AdamHyper h) {
  const int ti = blockIdx.y;
  float* __restrict__ p = (float*)t.a[ti];
  const float* __restrict__ g = (const float*)t.b[ti];
  float* __restrict__ m = (float*)t.c[ti];
  float* __restrict__ v = (float*)t.d[ti];
  const long long n = t.n[ti];
  auto update = [&](float& pi, float gi, float& mi, float& vi) {
    gi *= h.grad_scale;
    if (h.wd != 0.f) {
      if (h.decoupled) pi *= (1.f - h.lr * h.wd);
      else gi = fmaf(h.wd, pi, gi);
    }
    // same operation order as torch.optim.adam._single_tensor_adam
    mi = mi + (gi - mi) * (1.f - h.beta1);                  // exp_avg.lerp_(grad, 1-beta1)
    vi = vi * h.beta2 + (1.f - h.beta2) * gi * gi;           // mul_(beta2).addcmul_(g, g, 1-beta2)
    const float denom = sqrtf(vi) / h.bc2_sqrt + h.eps;
    pi = pi - h.step_size * (mi / denom);
  };
  long long done = 0;
  if (((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
        reinterpret_cast<uintptr_t>(v)) & 15) == 0) {
    // 16-byte accesses: four streams of float4 in flight per thread (same arithmetic per element)
    const long long nv = n >> 2;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
      float4 p4 = reinterpret_cast<float4*>(p)[i], m4 = reinterpret_cast<float4*>(m)[i], v4 = reinterpret_cast<float4*>(v)[i];
      const float4 g4 = reinterpret_cast<const float4*>(g)[i];
      update(p4.x, g4.x, m4.x, v4.x);
      update(p4.y, g4.y, m4.y, v4.y);
      update(p4.z, g4.z, m4.z, v4.z);
      update(p4.w, g4.w, m4.w, v4.w);
      reinterpret_cast<float4*>(m)[i] = m4;
      reinterpret_cast<float4*>(v)[i] = v4;
      reinterpret_cast<float4*>(p)[i] = p4;
    }
    done = nv << 2;
  }
  for (long long i = done + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float pi = p[i], mi = m[i], vi = v[i];
    update(pi, g[i], mi, vi);
    m[i] = mi;
    v[i] = vi;
    p[i] = pi;
  }
}

// Sum of squares, flat work split: block b handles one chunk of `chunk` elements of the tensor whose cumulative
// chunk range contains b (t.b[i] carries the first chunk index of tensor i), so one 23M-element embedding table and a
// hundred 768-element vectors spread evenly over the SMs.  Eight independent 16-byte loads per thread.
__global__ void __launch_bounds__(MT_THREADS) mt_sumsq_kernel(MtArgs t, int cnt, long long chunk,
                                                              float* __restrict__ partial, int base) {
  __shared__ float red[MT_THREADS / 32];
  int ti = 0;
  while (ti + 1 < cnt && (long long)(intptr_t)t.b[ti + 1] <= (long long)blockIdx.x) ++ti;
  const float* __restrict__ x = (const float*)t.a[ti];
  const long long n = t.n[ti];
  const long long lo = ((long long)blockIdx.x - (long long)(intptr_t)t.b[ti]) * chunk;
  const long long hi = lo + chunk < n ? lo + chunk : n;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if ((reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    const long long v0 = lo >> 2, v1 = hi >> 2;    // lo is a multiple of chunk (a multiple of 4)
    for (long long i = v0 + threadIdx.x; i < v1; i += MT_THREADS) {
      const float4 v = reinterpret_cast<const float4*>(x)[i];
      s0 = fmaf(v.x, v.x, s0);
      s1 = fmaf(v.y, v.y, s1);
      s2 = fmaf(v.z, v.z, s2);
      s3 = fmaf(v.w, v.w, s3);
    }
    for (long long i = (v1 << 2) + threadIdx.x; i < hi; i += MT_THREADS) s0 = fmaf(x[i], x[i], s0);
  } else {
    for (long long i = lo + threadIdx.x; i < hi; i += MT_THREADS) s0 = fmaf(x[i], x[i], s0);
  }
  float s = warp_sum((s0 + s1) + (s2 + s3));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < MT_THREADS / 32 ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) partial[base + blockIdx.x] = v;
  }
}
__global__ void sum_partials_kernel(const float* __restrict__ partial, int count, float* __restrict__ out) {
  __shared__ double red[32];
  double s = 0;
  for (int i = threadIdx.x; i < count; i += blockDim.x) s += partial[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    double v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) out[0] = (float)v;
  }
}

__global__ void __launch_bounds__(MT_THREADS) mt_clip_kernel(MtArgs t, const float* __restrict__ norm_sq, float max_norm) {
  const float coef = fminf(1.f, max_norm / (sqrtf(norm_sq[0]) + 1e-6f));
  if (coef >= 1.f) return;
  const int ti = blockIdx.y;
  float* __restrict__ x = (float*)t.a[ti];
  const long long n = t.n[ti];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    x[i] *= coef;
}
__global__ void __launch_bounds__(MT_THREADS) mt_zero_kernel(MtArgs t) {
  const int ti = blockIdx.y;
  float* __restrict__ x = (float*)t.a[ti];
  const long long n = t.n[ti];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    x[i] = 0.f;
}
// dst(a) = src(b) * scale
__global__ void __launch_bounds__(MT_THREADS) mt_copy_scale_kernel(MtArgs t, float scale) {
  const int ti = blockIdx.y;
  float* __restrict__ dst = (float*)t.a[ti];
  const float* __restrict__ src = (const float*)t.b[ti];
  const long long n = t.n[ti];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dst[i] = src[i] * scale;
}

static inline int mt_blocks(const int64_t* sizes, int lo, int hi, int cap) {
  long long mx = 1;
  for (int i = lo; i < hi; ++i) if (sizes[i] > mx) mx = sizes[i];
  long long b = (mx + MT_THREADS * 4 - 1) / (MT_THREADS * 4);
  if (b > cap) b = cap;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace b200

using namespace b200;

#define MT_FOREACH_CHUNK(n_, ...)                          \
  for (int lo = 0; lo < (n_); lo += MT_MAX) {              \
    const int hi = lo + MT_MAX < (n_) ? lo + MT_MAX : (n_); \
    const int cnt = hi - lo;                               \
    ++nlaunch;                                             \
    MtArgs t;                                              \
    __VA_ARGS__                                            \
  }

extern "C" int32_t b200_multi_tensor_sgd(int32_t n, float* const* params, const float* const* grads,
                                         float* const* momentum_bufs, const int64_t* sizes, float lr, float momentum,
                                         float dampening, float weight_decay, int32_t nesterov, int32_t first_step,
                                         float grad_scale, void* stream) {
  if (n < 0 || (n > 0 && (!params || !grads || !sizes))) return fail(B200_INVALID_ARG, "sgd: bad arguments");
  if (momentum != 0.f && !momentum_bufs) return fail(B200_INVALID_ARG, "sgd: momentum needs buffers");
  cudaStream_t st = as_stream(stream);
  SgdHyper h{lr, momentum, dampening, weight_decay, grad_scale, nesterov, first_step, momentum != 0.f};
  int nlaunch = 0;
  MT_FOREACH_CHUNK(n, {
    for (int i = 0; i < cnt; ++i) {
      t.a[i] = params[lo + i]; t.b[i] = (void*)grads[lo + i]; t.c[i] = h.has_mom ? momentum_bufs[lo + i] : nullptr;
      t.d[i] = nullptr; t.n[i] = sizes[lo + i];
    }
    mt_sgd_kernel<<<dim3(mt_blocks(sizes, lo, hi, 2 * num_sms()), cnt), MT_THREADS, 0, st>>>(t, h);
  })
  return check_launch("multi_tensor_sgd", nlaunch);
}

extern "C" int32_t b200_multi_tensor_adam(int32_t n, float* const* params, const float* const* grads,
                                          float* const* exp_avg, float* const* exp_avg_sq, const int64_t* sizes,
                                          float lr, float beta1, float beta2, float eps, float weight_decay,
                                          int32_t decoupled_wd, int64_t step, float grad_scale, void* stream) {
  if (n < 0 || (n > 0 && (!params || !grads || !exp_avg || !exp_avg_sq || !sizes)) || step < 1)
    return fail(B200_INVALID_ARG, "adam: bad arguments");
  cudaStream_t st = as_stream(stream);
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  AdamHyper h{lr, beta1, beta2, eps, weight_decay, (float)((double)lr / bc1), (float)sqrt(bc2), grad_scale, decoupled_wd};
  int nlaunch = 0;
  MT_FOREACH_CHUNK(n, {
    for (int i = 0; i < cnt; ++i) {
      t.a[i] = params[lo + i]; t.b[i] = (void*)grads[lo + i]; t.c[i] = exp_avg[lo + i]; t.d[i] = exp_avg_sq[lo + i];
      t.n[i] = sizes[lo + i];
    }
    mt_adam_kernel<<<dim3(mt_blocks(sizes, lo, hi, 2 * num_sms()), cnt), MT_THREADS, 0, st>>>(t, h);
  })
  return check_launch("multi_tensor_adam", nlaunch);
}

constexpr int L2_MAX_PARTIALS = 16384;
constexpr long long L2_CHUNK = 8192;     // elements per block at least (32 per thread)
extern "C" int64_t b200_l2norm_workspace_bytes(void) { return (int64_t)L2_MAX_PARTIALS * 4; }

extern "C" int32_t b200_multi_tensor_l2norm_sq(int32_t n, const float* const* tensors, const int64_t* sizes, float* out,
                                               void* ws, int64_t ws_bytes, void* stream) {
  if (n <= 0 || !tensors || !sizes || !out) return fail(B200_INVALID_ARG, "l2norm: bad arguments");
  if (!ws || ws_bytes < b200_l2norm_workspace_bytes()) return fail(B200_WORKSPACE_TOO_SMALL, "l2norm: workspace too small");
  if (n > L2_MAX_PARTIALS / 2) return fail(B200_UNSUPPORTED, "l2norm: more than %d tensors", L2_MAX_PARTIALS / 2);
  // chunk size: at most L2_MAX_PARTIALS partial sums in total (every tensor needs at least one)
  long long total = 0;
  for (int i = 0; i < n; ++i) total += sizes[i];
  long long chunk = L2_CHUNK;
  while (total / chunk + n > L2_MAX_PARTIALS) chunk *= 2;
  cudaStream_t st = as_stream(stream);
  int base = 0;
  int nlaunch = 0;
  MT_FOREACH_CHUNK(n, {
    long long cum = 0;
    for (int i = 0; i < cnt; ++i) {
      t.a[i] = (void*)tensors[lo + i];
      t.b[i] = (void*)(intptr_t)cum;
      t.c[i] = t.d[i] = nullptr;
      t.n[i] = sizes[lo + i];
      const long long c = (sizes[lo + i] + chunk - 1) / chunk;
      cum += c < 1 ? 1 : c;
    }
    mt_sumsq_kernel<<<(unsigned)cum, MT_THREADS, 0, st>>>(t, cnt, chunk, (float*)ws, base);
    base += (int)cum;
  })
  sum_partials_kernel<<<1, 1024, 0, st>>>((const float*)ws, base, out);
  return check_launch("multi_tensor_l2norm_sq", nlaunch + 1);
}

extern "C" int32_t b200_multi_tensor_clip_scale(int32_t n, float* const* tensors, const int64_t* sizes,
                                                const float* norm_sq, float max_norm, void* stream) {
  if (n <= 0 || !tensors || !sizes || !norm_sq) return fail(B200_INVALID_ARG, "clip_scale: bad arguments");
  cudaStream_t st = as_stream(stream);
  int nlaunch = 0;
  MT_FOREACH_CHUNK(n, {
    for (int i = 0; i < cnt; ++i) { t.a[i] = tensors[lo + i]; t.b[i] = t.c[i] = t.d[i] = nullptr; t.n[i] = sizes[lo + i]; }
    mt_clip_kernel<<<dim3(mt_blocks(sizes, lo, hi, 2 * num_sms()), cnt), MT_THREADS, 0, st>>>(t, norm_sq, max_norm);
  })
  return check_launch("multi_tensor_clip_scale", nlaunch);
}

extern "C" int32_t b200_multi_tensor_zero(int32_t n, float* const* tensors, const int64_t* sizes, void* stream) {
  if (n <= 0 || !tensors || !sizes) return fail(B200_INVALID_ARG, "zero: bad arguments");
  cudaStream_t st = as_stream(stream);
  int nlaunch = 0;
  MT_FOREACH_CHUNK(n, {
    for (int i = 0; i < cnt; ++i) { t.a[i] = tensors[lo + i]; t.b[i] = t.c[i] = t.d[i] = nullptr; t.n[i] = sizes[lo + i]; }
    mt_zero_kernel<<<dim3(mt_blocks(sizes, lo, hi, 2 * num_sms()), cnt), MT_THREADS, 0, st>>>(t);
  })
  return check_launch("multi_tensor_zero", nlaunch);
}

extern "C" int32_t b200_bucket_pack(int32_t n, const float* const* srcs, const int64_t* sizes, float* flat, float scale,
                                    void* stream) {
  if (n <= 0 || !srcs || !sizes || !flat) return fail(B200_INVALID_ARG, "bucket_pack: bad arguments");
  cudaStream_t st = as_stream(stream);
  long long off = 0;
  int nlaunch = 0;
  MT_FOREACH_CHUNK(n, {
    for (int i = 0; i < cnt; ++i) {
      t.a[i] = flat + off; t.b[i] = (void*)srcs[lo + i]; t.c[i] = t.d[i] = nullptr; t.n[i] = sizes[lo + i];
      off += sizes[lo + i];
    }
    mt_copy_scale_kernel<<<dim3(mt_blocks(sizes, lo, hi, 2 * num_sms()), cnt), MT_THREADS, 0, st>>>(t, scale);
  })
  return check_launch("bucket_pack", nlaunch);
}

extern "C" int32_t b200_bucket_unpack(int32_t n, float* const* dsts, const int64_t* sizes, const float* flat,
                                      float scale, void* stream) {
  if (n <= 0 || !dsts || !sizes || !flat) return fail(B200_INVALID_ARG, "bucket_unpack: bad arguments");
  cudaStream_t st = as_stream(stream);
  long long off = 0;
  int nlaunch = 0;
  MT_FOREACH_CHUNK(n, {
    for (int i = 0; i < cnt; ++i) {
      t.a[i] = dsts[lo + i]; t.b[i] = (void*)(flat + off); t.c[i] = t.d[i] = nullptr; t.n[i] = sizes[lo + i];
      off += sizes[lo + i];
    }
    mt_copy_scale_kernel<<<dim3(mt_blocks(sizes, lo, hi, 2 * num_sms()), cnt), MT_THREADS, 0, st>>>(t, scale);
  })
  return check_launch("bucket_unpack", nlaunch);
}

extern "C" int32_t b200_multi_tensor_copy(int32_t n, float* const* dsts, const float* const* srcs,
                                          const int64_t* sizes, float scale, void* stream) {
  if (n <= 0 || !dsts || !srcs || !sizes) return fail(B200_INVALID_ARG, "multi_tensor_copy: bad arguments");
  cudaStream_t st = as_stream(stream);
  int nlaunch = 0;
  MT_FOREACH_CHUNK(n, {
    for (int i = 0; i < cnt; ++i) {
      t.a[i] = dsts[lo + i]; t.b[i] = (void*)srcs[lo + i]; t.c[i] = t.d[i] = nullptr; t.n[i] = sizes[lo + i];
    }
    mt_copy_scale_kernel<<<dim3(mt_blocks(sizes, lo, hi, 2 * num_sms()), cnt), MT_THREADS, 0, st>>>(t, scale);
  })
  return check_launch("multi_tensor_copy", nlaunch);
}

namespace b200 {
long long& launch_counter() {
  static long long n = 0;
  return n;
}
}  // namespace b200
extern "C" int64_t b200_launch_count(void) { return b200::launch_counter(); }
extern "C" int32_t b200_version(void) { return 100; }
extern "C" const char* b200_last_error(void) { return last_error_buf(); }
