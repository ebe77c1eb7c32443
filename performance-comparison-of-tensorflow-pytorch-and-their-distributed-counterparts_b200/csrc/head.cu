// Classifier-head kernels (fp32): ReLU, Dropout (Philox4x32-10, mask replay in backward), LogSoftmax, NLLLoss,
// bias-gradient column sums.  Small tensors: these are launch-latency kernels, written for determinism.
#include "common.cuh"
#include "philox.cuh"

namespace b200 {

__global__ void relu_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = fmaxf(x[i], 0.f);
}

__global__ void dropout_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, uint8_t* __restrict__ mask,
                                   long long n, float p, float scale, uint64_t seed, uint64_t offset) {
  const long long groups = (n + 3) / 4;
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < groups;
       g += (long long)gridDim.x * blockDim.x) {
    uint32_t r[4];
    philox4x32_10(seed, offset + (uint64_t)g, r);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const long long i = g * 4 + e;
      if (i < n) {
        const float u = (float)(r[e] >> 8) * (1.0f / 16777216.0f);  // [0,1)
        const uint8_t keep = u >= p;
        mask[i] = keep;
        y[i] = keep ? x[i] * scale : 0.f;
      }
    }
  }
}
__global__ void dropout_bwd_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ mask,
                                   float* __restrict__ dx, long long n, float scale) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dx[i] = mask[i] ? dy[i] * scale : 0.f;
}

// ---------------------------------------------------------------- log-softmax: one warp per row
__global__ void log_softmax_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int rows, int cols) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float* xr = x + (long long)warp * cols;
  float m = -3.402823466e38f;
  for (int c = lane; c < cols; c += 32) m = fmaxf(m, xr[c]);
  m = warp_max(m);
  float s = 0.f;
  for (int c = lane; c < cols; c += 32) s += expf(xr[c] - m);
  s = warp_sum(s);
  const float lse = m + logf(s);
  for (int c = lane; c < cols; c += 32) y[(long long)warp * cols + c] = xr[c] - lse;
}
// dx = dy - exp(y) * sum(dy)
__global__ void log_softmax_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                       float* __restrict__ dx, int rows, int cols) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const long long o = (long long)warp * cols;
  float s = 0.f;
  for (int c = lane; c < cols; c += 32) s += dy[o + c];
  s = warp_sum(s);
  for (int c = lane; c < cols; c += 32) dx[o + c] = dy[o + c] - expf(y[o + c]) * s;
}

// ---------------------------------------------------------------- NLL loss (mean reduction), single block
__global__ void nll_fwd_kernel(const float* __restrict__ logp, const long long* __restrict__ labels,
                               float* __restrict__ loss, int rows, int cols) {
  __shared__ float red[32];
  float s = 0.f;
  for (int r = threadIdx.x; r < rows; r += blockDim.x) {
    const long long l = labels[r];
    if (l >= 0 && l < cols) s -= logp[(long long)r * cols + l];
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) loss[0] = v / (float)rows;
  }
}
__global__ void nll_bwd_kernel(const float* __restrict__ dloss, const long long* __restrict__ labels,
                               float* __restrict__ dlogp, int rows, int cols) {
  const long long total = (long long)rows * cols;
  const float g = -dloss[0] / (float)rows;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), c = (int)(i - (long long)r * cols);
    dlogp[i] = (labels[r] == c) ? g : 0.f;
  }
}

// db[c] = sum_r dy[r][c] : one block per 32 columns, 8 row-lanes, fixed-order reduction
__global__ void colsum_kernel(const float* __restrict__ dy, float* __restrict__ db, int rows, int cols) {
  __shared__ float red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (c < cols)
    for (int r = threadIdx.y; r < rows; r += 8) s += dy[(long long)r * cols + c];
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += red[j][threadIdx.x];
    db[c] = t;
  }
}

static inline int hgrid(long long total) {
  long long b = (total + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (b > cap) b = cap;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace b200

using namespace b200;

extern "C" int32_t b200_relu_fwd_f32(const float* x, float* y, int64_t n, void* stream) {
  if (!x || !y || n <= 0) return fail(B200_INVALID_ARG, "relu_fwd: bad arguments");
  relu_fwd_kernel<<<hgrid(n), 256, 0, as_stream(stream)>>>(x, y, n);
  return check_launch("relu_fwd");
}
extern "C" int32_t b200_dropout_fwd(const float* x, float* y, uint8_t* mask, int64_t n, float p, uint64_t seed,
                                    uint64_t offset, void* stream) {
  if (!x || !y || !mask || n <= 0 || p < 0.f || p >= 1.f) return fail(B200_INVALID_ARG, "dropout_fwd: bad arguments");
  dropout_fwd_kernel<<<hgrid((n + 3) / 4), 256, 0, as_stream(stream)>>>(x, y, mask, n, p, 1.f / (1.f - p), seed, offset);
  return check_launch("dropout_fwd");
}
extern "C" int32_t b200_dropout_bwd(const float* dy, const uint8_t* mask, float* dx, int64_t n, float p, void* stream) {
  if (!dy || !dx || !mask || n <= 0 || p < 0.f || p >= 1.f) return fail(B200_INVALID_ARG, "dropout_bwd: bad arguments");
  dropout_bwd_kernel<<<hgrid(n), 256, 0, as_stream(stream)>>>(dy, mask, dx, n, 1.f / (1.f - p));
  return check_launch("dropout_bwd");
}
extern "C" int32_t b200_log_softmax_fwd(const float* x, float* y, int32_t rows, int32_t cols, void* stream) {
  if (!x || !y || rows <= 0 || cols <= 0) return fail(B200_INVALID_ARG, "log_softmax_fwd: bad arguments");
  log_softmax_fwd_kernel<<<(rows + 7) / 8, 256, 0, as_stream(stream)>>>(x, y, rows, cols);
  return check_launch("log_softmax_fwd");
}
extern "C" int32_t b200_log_softmax_bwd(const float* dy, const float* y, float* dx, int32_t rows, int32_t cols,
                                        void* stream) {
  if (!dy || !y || !dx || rows <= 0 || cols <= 0) return fail(B200_INVALID_ARG, "log_softmax_bwd: bad arguments");
  log_softmax_bwd_kernel<<<(rows + 7) / 8, 256, 0, as_stream(stream)>>>(dy, y, dx, rows, cols);
  return check_launch("log_softmax_bwd");
}
extern "C" int32_t b200_nll_loss_fwd(const float* logp, const int64_t* labels, float* loss, int32_t rows, int32_t cols,
                                     void* stream) {
  if (!logp || !labels || !loss || rows <= 0 || cols <= 0) return fail(B200_INVALID_ARG, "nll_loss_fwd: bad arguments");
  nll_fwd_kernel<<<1, 256, 0, as_stream(stream)>>>(logp, (const long long*)labels, loss, rows, cols);
  return check_launch("nll_loss_fwd");
}
extern "C" int32_t b200_nll_loss_bwd(const float* dloss, const int64_t* labels, float* dlogp, int32_t rows,
                                     int32_t cols, void* stream) {
  if (!dloss || !labels || !dlogp || rows <= 0 || cols <= 0) return fail(B200_INVALID_ARG, "nll_loss_bwd: bad arguments");
  nll_bwd_kernel<<<hgrid((long long)rows * cols), 256, 0, as_stream(stream)>>>(dloss, (const long long*)labels, dlogp, rows, cols);
  return check_launch("nll_loss_bwd");
}
extern "C" int32_t b200_colsum_f32(const float* dy, float* db, int32_t rows, int32_t cols, void* stream) {
  if (!dy || !db || rows <= 0 || cols <= 0) return fail(B200_INVALID_ARG, "colsum: bad arguments");
  colsum_kernel<<<(cols + 31) / 32, dim3(32, 8), 0, as_stream(stream)>>>(dy, db, rows, cols);
  return check_launch("colsum");
}
