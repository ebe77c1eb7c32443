// Layout / packing kernels at the PyTorch boundary: NCHW fp32 <-> NHWC (fp32|bf16), OIHW fp32 -> KRSC, casts,
// im2col / col2im for the explicit-GEMM convolution paths.
#include <vector>
#include "common.cuh"

namespace b200 {

// ---------------------------------------------------------------- NCHW <-> NHWC
// small C (images, C<=4): one thread per pixel, coalesced plane reads, contiguous pixel writes
template <typename T>
__global__ void nchw_to_nhwc_smallc(const float* __restrict__ x, T* __restrict__ y, int C, int Cpad, long long HW,
                                    long long total_pix) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_pix;
       i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / HW, hw = i - n * HW;
    const float* src = x + n * C * HW + hw;
    T* dst = y + i * Cpad;
    for (int c = 0; c < Cpad; ++c) dst[c] = from_f<T>(c < C ? src[c * HW] : 0.f);
  }
}
// general: per image transpose [C][HW] -> [HW][C] through a 32x33 smem tile
template <typename TI, typename TO>
__global__ void transpose_batched(const TI* __restrict__ x, TO* __restrict__ y, int rows, int cols) {
  // x: [batch][rows][cols] -> y: [batch][cols][rows]
  __shared__ float tile[32][33];
  const long long b = blockIdx.z;
  const TI* xb = x + b * (long long)rows * cols;
  TO* yb = y + b * (long long)rows * cols;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    int r = r0 + j, c = c0 + threadIdx.x;
    if (r < rows && c < cols) tile[j][threadIdx.x] = to_f(xb[(long long)r * cols + c]);
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    int c = c0 + j, r = r0 + threadIdx.x;
    if (r < rows && c < cols) yb[(long long)c * rows + r] = from_f<TO>(tile[threadIdx.x][j]);
  }
}

// ---------------------------------------------------------------- weights: OIHW fp32 -> [K][ld] with (r,s,c) order
template <typename T>
__global__ void pack_weight_kernel(const float* __restrict__ w, T* __restrict__ out, int K, int Cw, int C, int R,
                                   int S, int ld) {
  // Cw = channels of the OIHW parameter, C >= Cw = channels of the packed operand (extra channels are zero)
  const long long total = (long long)K * ld;
  const int RSC = R * S * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / ld), j = (int)(i - (long long)k * ld);
    float v = 0.f;
    if (j < RSC) {
      const int c = j % C, rs = j / C;
      if (c < Cw) v = w[((long long)k * Cw + c) * (R * S) + rs];
    }
    out[i] = from_f<T>(v);
  }
}

// Several weights in one launch (the training graph re-packs all 53 convolution weights every step: one launch per
// 48 tensors instead of one each).  The table travels by value in the kernel parameter block.
constexpr int PACK_MAX = 48;
struct PackArgs {
  const float* w[PACK_MAX];
  void* out[PACK_MAX];
  int K[PACK_MAX], Cw[PACK_MAX], C[PACK_MAX], RS[PACK_MAX], ld[PACK_MAX];
};
template <typename T>
__global__ void __launch_bounds__(256) pack_weight_multi_kernel(PackArgs a) {
  const int ti = blockIdx.y;
  const float* __restrict__ w = a.w[ti];
  T* __restrict__ out = reinterpret_cast<T*>(a.out[ti]);
  const int K = a.K[ti], Cw = a.Cw[ti], C = a.C[ti], RS = a.RS[ti], ld = a.ld[ti];
  const long long total = (long long)K * ld;
  const int RSC = RS * C;
  if (RS == 1 && Cw == C && ld == C && (total & 7) == 0 &&
      ((reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
    // 1x1 weights and Linear matrices: the packed form is the matrix itself -- eight elements per thread, 16-byte
    // accesses (the element-wise path below spends a 64-bit division per element: 1.9 TB/s on the ResNet weights)
    const long long nv = total >> 3;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
      const float4 lo = reinterpret_cast<const float4*>(w)[2 * i], hi = reinterpret_cast<const float4*>(w)[2 * i + 1];
      if constexpr (sizeof(T) == 4) {
        reinterpret_cast<float4*>(out)[2 * i] = lo;
        reinterpret_cast<float4*>(out)[2 * i + 1] = hi;
      } else {
        uint4 v;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
        h[0] = __floats2bfloat162_rn(lo.x, lo.y); h[1] = __floats2bfloat162_rn(lo.z, lo.w);
        h[2] = __floats2bfloat162_rn(hi.x, hi.y); h[3] = __floats2bfloat162_rn(hi.z, hi.w);
        reinterpret_cast<uint4*>(out)[i] = v;
      }
    }
    return;
  }
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / ld), j = (int)(i - (long long)k * ld);
    float v = 0.f;
    if (j < RSC) {
      const int c = j % C, rs = j / C;
      if (c < Cw) v = w[((long long)k * Cw + c) * RS + rs];
    }
    out[i] = from_f<T>(v);
  }
}

template <typename TI, typename TO>
__global__ void cast_kernel(const TI* __restrict__ x, TO* __restrict__ y, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = from_f<TO>(to_f(x[i]));
}

// ---------------------------------------------------------------- im2col (NHWC -> [M][ld], column = (r*S+s)*C + c)
template <typename T, int VEC>
__global__ void im2col_kernel(const T* __restrict__ x, T* __restrict__ col, int N, int H, int W, int C, int R, int S,
                              int stride, int pad, int Ho, int Wo, int ld) {
  const int cv = C / VEC;               // channel vectors per tap
  const int per_row = R * S * cv;
  const long long M = (long long)N * Ho * Wo;
  const long long total = M * per_row;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / per_row;
    const int j = (int)(i - m * per_row);
    const int tap = j / cv, c = (j - tap * cv) * VEC;
    const int r = tap / S, s = tap - r * S;
    const int q = (int)(m % Wo);
    const long long t = m / Wo;
    const int p = (int)(t % Ho);
    const int n = (int)(t / Ho);
    const int h = p * stride + r - pad, w = q * stride + s - pad;
    T* dst = col + m * ld + (long long)tap * C + c;
    if (h >= 0 && h < H && w >= 0 && w < W) {
      const T* src = x + (((long long)n * H + h) * W + w) * C + c;
      if constexpr (VEC == 1) dst[0] = src[0];
      else if constexpr (VEC * sizeof(T) == 8) *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(src);
      else *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
    } else {
      if constexpr (VEC == 1) dst[0] = from_f<T>(0.f);
      else if constexpr (VEC * sizeof(T) == 8) *reinterpret_cast<uint2*>(dst) = make_uint2(0u, 0u);
      else *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
    }
  }
}
// 8-byte pixels (bf16, C = 4: the channel-padded image): one warp per output pixel, lane l writes the 8-byte slots
// l, l+32, ... of the pixel's row (slot t < R*S is filter tap t, the remaining slots up to ld are zero).  Writes are
// fully coalesced, reads hit L1/L2 (neighbouring pixels share most of their window), index math is per warp.
__global__ void __launch_bounds__(256)
im2col_px8_kernel(const uint2* __restrict__ x, uint2* __restrict__ col, int N, int H, int W, int R, int S, int stride,
                  int pad, int Ho, int Wo, int slots) {
  const int lane = threadIdx.x & 31;
  const long long M = (long long)N * Ho * Wo;
  const int taps = R * S;
  for (long long m = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5); m < M;
       m += (long long)gridDim.x * (blockDim.x >> 5)) {
    const int q = (int)(m % Wo);
    const long long t2 = m / Wo;
    const int p = (int)(t2 % Ho);
    const int n = (int)(t2 / Ho);
    const int h0 = p * stride - pad, w0 = q * stride - pad;
    const uint2* xn = x + (long long)n * H * W;
    uint2* dst = col + m * slots;
    for (int t = lane; t < slots; t += 32) {
      uint2 v = make_uint2(0u, 0u);
      if (t < taps) {
        const int r = t / S, s2 = t - r * S;
        const int h = h0 + r, w = w0 + s2;
        if (h >= 0 && h < H && w >= 0 && w < W) v = xn[(long long)h * W + w];
      }
      dst[t] = v;
    }
  }
}

// Row-pair window staging for a stride-2 convolution over a 4-channel (8-byte pixel) image -- the ResNet stem,
// 7x7 / 2 over the channel-padded RGB batch.  out[n][j][q][r2*32 + s*4 + c] = x[n][2j + r2 - pad][2q + s - pad][c]
// (s = 0..7, zero outside the image): one 128-byte row holds, for output column q, the 8-pixel windows of the two
// input rows of row pair j.  The convolution then is an R' = ceil(R/2) x 1, stride-1 convolution over 64 "channels"
// on the implicit tcgen05 path (filter-row pair j' is a coordinate shift along j), and the staged tensor is
// (Ho + R' - 1) * Wo * 128 bytes per image -- a quarter of a full im2col matrix, small enough to stay in L2 between
// the four taps.  One thread moves 16 bytes (two pixels).
__global__ void __launch_bounds__(256)
stem_stage_kernel(const uint2* __restrict__ x, uint4* __restrict__ out, int N, int H, int W, int pad, int Hs, int Wo) {
  const long long total = (long long)N * Hs * Wo * 8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(i & 7);            // 16-byte chunk of the row: r2 = t >> 2, pixels 2*(t&3), 2*(t&3)+1
    long long rest = i >> 3;
    const int q = (int)(rest % Wo);
    rest /= Wo;
    const int j = (int)(rest % Hs);
    const int n = (int)(rest / Hs);
    const int h = 2 * j + (t >> 2) - pad;
    const int w = 2 * q + 2 * (t & 3) - pad;
    uint2 a = make_uint2(0u, 0u), b = make_uint2(0u, 0u);
    if (h >= 0 && h < H) {
      const uint2* row = x + ((long long)n * H + h) * W;
      if (w >= 0 && w < W) a = row[w];
      if (w + 1 >= 0 && w + 1 < W) b = row[w + 1];
    }
    out[i] = make_uint4(a.x, a.y, b.x, b.y);
  }
}
// The same staged tensor straight from the fp32 NCHW batch the loaders produce (Cw <= 4 channels, the rest zero): the
// layout conversion to 8-byte pixels and the staging are one pass -- the intermediate NHWC4 image is never written.
__global__ void __launch_bounds__(256)
stem_stage_nchw_kernel(const float* __restrict__ x, uint4* __restrict__ out, int N, int Cw, int H, int W, int pad, int Hs,
                       int Wo) {
  const long long total = (long long)N * Hs * Wo * 8;
  const long long plane = (long long)H * W;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(i & 7);
    long long rest = i >> 3;
    const int q = (int)(rest % Wo);
    rest /= Wo;
    const int j = (int)(rest % Hs);
    const int n = (int)(rest / Hs);
    const int h = 2 * j + (t >> 2) - pad;
    const int w = 2 * q + 2 * (t & 3) - pad;
    float v[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    if (h >= 0 && h < H) {
      const float* row = x + (long long)n * Cw * plane + (long long)h * W;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (c >= Cw) break;
        if (w >= 0 && w < W) v[0][c] = row[c * plane + w];
        if (w + 1 >= 0 && w + 1 < W) v[1][c] = row[c * plane + w + 1];
      }
    }
    uint4 o;
    __nv_bfloat162* hp = reinterpret_cast<__nv_bfloat162*>(&o);
    hp[0] = __floats2bfloat162_rn(v[0][0], v[0][1]); hp[1] = __floats2bfloat162_rn(v[0][2], v[0][3]);
    hp[2] = __floats2bfloat162_rn(v[1][0], v[1][1]); hp[3] = __floats2bfloat162_rn(v[1][2], v[1][3]);
    out[i] = o;
  }
}
// OIHW fp32 stem filter [K][Cw <= 4][R][S <= 8] -> [K][R' * 64] bf16 in the column order of stem_stage_kernel:
// column j*64 + r2*32 + s*4 + c holds w[k][c][2j + r2][s] (zero where r >= R, s >= S or c >= Cw).
__global__ void stem_pack_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int K, int Cw, int R,
                                        int S, int ld) {
  const long long total = (long long)K * ld;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / ld), col = (int)(i - (long long)k * ld);
    const int j = col >> 6, r2 = (col >> 5) & 1, s = (col >> 2) & 7, c = col & 3;
    const int r = 2 * j + r2;
    float v = 0.f;
    if (r < R && s < S && c < Cw) v = w[(((long long)k * Cw + c) * R + r) * S + s];
    out[i] = __float2bfloat16_rn(v);
  }
}

template <typename T>
__global__ void im2col_pad_kernel(T* __restrict__ col, long long M, int used, int ld) {
  const int padw = ld - used;
  const long long total = M * padw;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / padw;
    const int j = (int)(i - m * padw);
    col[m * ld + used + j] = from_f<T>(0.f);
  }
}

// ---------------------------------------------------------------- col2im: dx[n,h,w,c] (+)= sum over taps of dcol
// dcol is fp32 [M][ld] with column (r*S+s)*C + c.
template <typename T>
__global__ void col2im_kernel(const float* __restrict__ dcol, T* __restrict__ dx, int N, int H, int W, int C, int R,
                              int S, int stride, int pad, int Ho, int Wo, int ld, int accumulate) {
  const long long total = (long long)N * H * W * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long t = i / C;
    const int w = (int)(t % W);
    t /= W;
    const int h = (int)(t % H);
    const int n = (int)(t / H);
    float acc = accumulate ? to_f(dx[i]) : 0.f;
    for (int r = 0; r < R; ++r) {
      const int hp = h + pad - r;
      if (hp < 0 || hp % stride) continue;
      const int p = hp / stride;
      if (p >= Ho) continue;
      for (int s = 0; s < S; ++s) {
        const int wq = w + pad - s;
        if (wq < 0 || wq % stride) continue;
        const int q = wq / stride;
        if (q >= Wo) continue;
        acc += dcol[(((long long)n * Ho + p) * Wo + q) * ld + (long long)(r * S + s) * C + c];
      }
    }
    dx[i] = from_f<T>(acc);
  }
}

// ---------------------------------------------------------------- split-K reduce
// partial slabs [S][rows][ld] fp32 -> wgrad in OIHW fp32: row k, column (r*S+s)*C+c  ->  dw[((k*C+c)*R+r)*S+s]
__global__ void reduce_wgrad_kernel(const float* __restrict__ part, int slabs, long long slab_stride, int K, int C,
                                    int RS, int ld, float* __restrict__ dw) {
  const int RSC = RS * C;
  const long long total = (long long)K * RSC;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / RSC), j = (int)(i - (long long)k * RSC);
    const int rs = j / C, c = j - rs * C;
    float acc = 0.f;
    for (int s = 0; s < slabs; ++s) acc += part[s * slab_stride + (long long)k * ld + j];
    dw[((long long)k * C + c) * RS + rs] = acc;
  }
}
// The same reduction for up to RW_MAX layers in ONE launch.  A training step runs 53 of these tiny reductions (one per
// convolution, 0.41 ms of a 14 ms step when launched one by one: each is a few microseconds of work behind a kernel
// boundary); b200_wgrad_defer(1) makes launch_reduce_wgrad queue them (the caller gives every layer its own slab
// buffer) and b200_wgrad_flush launches them together.  Block b works on job j with blk0[j] <= b < blk0[j + 1].
constexpr int RW_MAX = 48;
struct ReduceBatch {
  const float* part[RW_MAX];
  float* dw[RW_MAX];
  long long stride[RW_MAX];
  int slabs[RW_MAX], K[RW_MAX], C[RW_MAX], RS[RW_MAX], ld[RW_MAX];
  int blk0[RW_MAX + 1];
};
__global__ void __launch_bounds__(256) reduce_wgrad_multi_kernel(const __grid_constant__ ReduceBatch b) {
  int j = 0;
  while ((int)blockIdx.x >= b.blk0[j + 1]) ++j;
  const int lb = (int)blockIdx.x - b.blk0[j], nb = b.blk0[j + 1] - b.blk0[j];
  const float* __restrict__ part = b.part[j];
  float* __restrict__ dw = b.dw[j];
  const int K = b.K[j], C = b.C[j], RS = b.RS[j], ld = b.ld[j], slabs = b.slabs[j];
  const long long slab_stride = b.stride[j];
  const int RSC = RS * C;
  const int total = K * RSC;                     // < 2^31: a convolution weight
  for (int i = lb * 256 + (int)threadIdx.x; i < total; i += nb * 256) {
    const int k = i / RSC, jj = i - k * RSC;
    const int rs = jj / C, c = jj - rs * C;
    float acc = 0.f;
    for (int s = 0; s < slabs; ++s) acc += part[s * slab_stride + (long long)k * ld + jj];
    dw[((long long)k * C + c) * RS + rs] = acc;
  }
}
// partial slabs -> plain [rows][ldo] (fp32 or bf16), optional accumulate
template <typename T>
__global__ void reduce_plain_kernel(const float* __restrict__ part, int slabs, long long slab_stride, int rows,
                                    int cols, long long ldp, T* __restrict__ out, long long ldo, int accumulate) {
  const long long total = (long long)rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = (int)(i - r * cols);
    float acc = accumulate ? to_f(out[r * ldo + c]) : 0.f;
    for (int s = 0; s < slabs; ++s) acc += part[s * slab_stride + r * ldp + c];
    out[r * ldo + c] = from_f<T>(acc);
  }
}

static inline int grid_for(long long total, int threads = 256, int max_waves = 8) {
  long long b = (total + threads - 1) / threads;
  long long cap = (long long)num_sms() * max_waves;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// host-side helpers used by conv.cu
int32_t launch_im2col(const void* x, void* col, int dtype, int N, int H, int W, int C, int R, int S, int stride,
                      int pad, int Ho, int Wo, int ld, cudaStream_t st) {
  const long long M = (long long)N * Ho * Wo;
  if (dtype == B200_F32) {
    if (C % 4 == 0 && ld % 4 == 0)
      im2col_kernel<float, 4><<<grid_for(M * R * S * (C / 4)), 256, 0, st>>>((const float*)x, (float*)col, N, H, W, C, R, S, stride, pad, Ho, Wo, ld);
    else
      im2col_kernel<float, 1><<<grid_for(M * R * S * C), 256, 0, st>>>((const float*)x, (float*)col, N, H, W, C, R, S, stride, pad, Ho, Wo, ld);
    if (ld > R * S * C) im2col_pad_kernel<float><<<grid_for(M * (ld - R * S * C)), 256, 0, st>>>((float*)col, M, R * S * C, ld);
  } else {
    if (C == 4 && ld % 4 == 0 && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(col)) & 7) == 0) {
      const int warps_per_block = 8;
      long long blocks = (M + warps_per_block - 1) / warps_per_block;
      if (blocks > (long long)num_sms() * 32) blocks = (long long)num_sms() * 32;
      im2col_px8_kernel<<<(int)blocks, 256, 0, st>>>((const uint2*)x, (uint2*)col, N, H, W, R, S, stride, pad, Ho, Wo, ld / 4);
      return check_launch("im2col");
    }
    if (C % 8 == 0 && ld % 8 == 0)
      im2col_kernel<__nv_bfloat16, 8><<<grid_for(M * R * S * (C / 8)), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)col, N, H, W, C, R, S, stride, pad, Ho, Wo, ld);
    else if (C % 4 == 0 && ld % 4 == 0)   // the channel-padded stem: 8-byte pixels
      im2col_kernel<__nv_bfloat16, 4><<<grid_for(M * R * S * (C / 4), 256, 32), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)col, N, H, W, C, R, S, stride, pad, Ho, Wo, ld);
    else
      im2col_kernel<__nv_bfloat16, 1><<<grid_for(M * R * S * C), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)col, N, H, W, C, R, S, stride, pad, Ho, Wo, ld);
    if (ld > R * S * C) im2col_pad_kernel<__nv_bfloat16><<<grid_for(M * (ld - R * S * C)), 256, 0, st>>>((__nv_bfloat16*)col, M, R * S * C, ld);
  }
  return check_launch("im2col", ld > R * S * C ? 2 : 1);
}
int32_t launch_col2im(const float* dcol, void* dx, int dtype, int N, int H, int W, int C, int R, int S, int stride,
                      int pad, int Ho, int Wo, int ld, int accumulate, cudaStream_t st) {
  const long long total = (long long)N * H * W * C;
  if (dtype == B200_F32)
    col2im_kernel<float><<<grid_for(total), 256, 0, st>>>(dcol, (float*)dx, N, H, W, C, R, S, stride, pad, Ho, Wo, ld, accumulate);
  else
    col2im_kernel<__nv_bfloat16><<<grid_for(total), 256, 0, st>>>(dcol, (__nv_bfloat16*)dx, N, H, W, C, R, S, stride, pad, Ho, Wo, ld, accumulate);
  return check_launch("col2im");
}
struct ReduceJob {
  const float* part;
  float* dw;
  long long stride;
  int slabs, K, C, RS, ld;
};
struct ReduceQueue {
  bool on = false;
  std::vector<ReduceJob> jobs;
};
static ReduceQueue& reduce_queue() {      // per host thread: autograd runs the backward on its own worker thread
  static thread_local ReduceQueue q;
  return q;
}
int32_t launch_reduce_wgrad(const float* part, int slabs, long long slab_stride, int K, int C, int RS, int ld,
                            float* dw, cudaStream_t st) {
  ReduceQueue& q = reduce_queue();
  if (q.on && (long long)K * RS * C < (1LL << 31)) {
    q.jobs.push_back(ReduceJob{part, dw, slab_stride, slabs, K, C, RS, ld});
    return B200_OK;
  }
  reduce_wgrad_kernel<<<grid_for((long long)K * RS * C), 256, 0, st>>>(part, slabs, slab_stride, K, C, RS, ld, dw);
  return check_launch("reduce_wgrad");
}
int32_t reduce_wgrad_flush(cudaStream_t st) {
  ReduceQueue& q = reduce_queue();
  size_t at = 0;
  int launches = 0;
  while (at < q.jobs.size()) {
    ReduceBatch b{};
    int n = 0, blocks = 0;
    for (; at < q.jobs.size() && n < RW_MAX; ++at, ++n) {
      const ReduceJob& jb = q.jobs[at];
      b.part[n] = jb.part; b.dw[n] = jb.dw; b.stride[n] = jb.stride;
      b.slabs[n] = jb.slabs; b.K[n] = jb.K; b.C[n] = jb.C; b.RS[n] = jb.RS; b.ld[n] = jb.ld;
      const long long total = (long long)jb.K * jb.RS * jb.C;
      long long nb = (total + 1023) / 1024;           // four elements per thread
      if (nb > 2 * num_sms()) nb = 2 * num_sms();
      if (nb < 1) nb = 1;
      b.blk0[n] = blocks;
      blocks += (int)nb;
    }
    for (int i = n; i <= RW_MAX; ++i) b.blk0[i] = blocks;
    reduce_wgrad_multi_kernel<<<blocks, 256, 0, st>>>(b);
    ++launches;
  }
  q.jobs.clear();
  return launches ? check_launch("reduce_wgrad_multi", launches) : B200_OK;
}
void reduce_wgrad_defer(bool on) {
  ReduceQueue& q = reduce_queue();
  q.on = on;
  if (!on) q.jobs.clear();
}
int32_t launch_reduce_plain(const float* part, int slabs, long long slab_stride, int rows, int cols, long long ldp,
                            void* out, long long ldo, int out_dtype, int accumulate, cudaStream_t st) {
  if (out_dtype == B200_F32)
    reduce_plain_kernel<float><<<grid_for((long long)rows * cols), 256, 0, st>>>(part, slabs, slab_stride, rows, cols, ldp, (float*)out, ldo, accumulate);
  else
    reduce_plain_kernel<__nv_bfloat16><<<grid_for((long long)rows * cols), 256, 0, st>>>(part, slabs, slab_stride, rows, cols, ldp, (__nv_bfloat16*)out, ldo, accumulate);
  return check_launch("reduce_plain");
}

}  // namespace b200

using namespace b200;

extern "C" int32_t b200_nchw_to_nhwc(const float* x, void* y, int32_t N, int32_t C, int32_t H, int32_t W,
                                     int32_t Cpad, int32_t out_dtype, void* stream) {
  if (!x || !y || N <= 0 || C <= 0 || H <= 0 || W <= 0 || Cpad < C) return fail(B200_INVALID_ARG, "nchw_to_nhwc: bad args");
  if (Cpad != C && Cpad > 4) return fail(B200_UNSUPPORTED, "nchw_to_nhwc: channel padding is supported up to 4 channels");
  cudaStream_t st = as_stream(stream);
  const long long HW = (long long)H * W;
  if (Cpad <= 4) {
    if (out_dtype == B200_F32) nchw_to_nhwc_smallc<float><<<grid_for(N * HW), 256, 0, st>>>(x, (float*)y, C, Cpad, HW, N * HW);
    else if (out_dtype == B200_BF16) nchw_to_nhwc_smallc<__nv_bfloat16><<<grid_for(N * HW), 256, 0, st>>>(x, (__nv_bfloat16*)y, C, Cpad, HW, N * HW);
    else return fail(B200_INVALID_ARG, "bad dtype");
  } else {
    dim3 grid((unsigned)((HW + 31) / 32), (C + 31) / 32, N), block(32, 8);
    if (out_dtype == B200_F32) transpose_batched<float, float><<<grid, block, 0, st>>>(x, (float*)y, C, (int)HW);
    else if (out_dtype == B200_BF16) transpose_batched<float, __nv_bfloat16><<<grid, block, 0, st>>>(x, (__nv_bfloat16*)y, C, (int)HW);
    else return fail(B200_INVALID_ARG, "bad dtype");
  }
  return check_launch("nchw_to_nhwc");
}

extern "C" int32_t b200_nhwc_to_nchw(const void* x, float* y, int32_t N, int32_t C, int32_t H, int32_t W,
                                     int32_t in_dtype, void* stream) {
  if (!x || !y || N <= 0 || C <= 0 || H <= 0 || W <= 0) return fail(B200_INVALID_ARG, "nhwc_to_nchw: bad args");
  cudaStream_t st = as_stream(stream);
  const long long HW = (long long)H * W;
  dim3 grid((C + 31) / 32, (unsigned)((HW + 31) / 32), N), block(32, 8);
  if (in_dtype == B200_F32) transpose_batched<float, float><<<grid, block, 0, st>>>((const float*)x, y, (int)HW, C);
  else if (in_dtype == B200_BF16) transpose_batched<__nv_bfloat16, float><<<grid, block, 0, st>>>((const __nv_bfloat16*)x, y, (int)HW, C);
  else return fail(B200_INVALID_ARG, "bad dtype");
  return check_launch("nhwc_to_nchw");
}

extern "C" int32_t b200_pack_weight(const float* w, void* out, int32_t K, int32_t Cw, int32_t R, int32_t S,
                                    int32_t C, int32_t ld, int32_t out_dtype, void* stream) {
  if (!w || !out || C < Cw || ld < R * S * C) return fail(B200_INVALID_ARG, "pack_weight: bad args");
  cudaStream_t st = as_stream(stream);
  if (out_dtype == B200_F32) pack_weight_kernel<float><<<grid_for((long long)K * ld), 256, 0, st>>>(w, (float*)out, K, Cw, C, R, S, ld);
  else if (out_dtype == B200_BF16) pack_weight_kernel<__nv_bfloat16><<<grid_for((long long)K * ld), 256, 0, st>>>(w, (__nv_bfloat16*)out, K, Cw, C, R, S, ld);
  else return fail(B200_INVALID_ARG, "bad dtype");
  return check_launch("pack_weight");
}

extern "C" int32_t b200_pack_weight_multi(int32_t count, const void* const* w_oihw, void* const* w_krsc,
                                          const int32_t* K, const int32_t* Cw, const int32_t* C, const int32_t* RS,
                                          const int32_t* ld, int32_t out_dtype, void* stream) {
  if (count <= 0) return B200_OK;
  if (!w_oihw || !w_krsc || !K || !Cw || !C || !RS || !ld) return fail(B200_INVALID_ARG, "pack_weight_multi: null table");
  if (out_dtype != B200_F32 && out_dtype != B200_BF16) return fail(B200_INVALID_ARG, "pack_weight_multi: bad dtype");
  cudaStream_t st = as_stream(stream);
  for (int lo = 0; lo < count; lo += PACK_MAX) {
    const int cnt = count - lo < PACK_MAX ? count - lo : PACK_MAX;
    PackArgs a;
    long long biggest = 0;
    for (int i = 0; i < cnt; ++i) {
      const int t = lo + i;
      if (!w_oihw[t] || !w_krsc[t] || C[t] < Cw[t] || ld[t] < RS[t] * C[t] || K[t] <= 0)
        return fail(B200_INVALID_ARG, "pack_weight_multi: bad entry %d", t);
      a.w[i] = (const float*)w_oihw[t]; a.out[i] = w_krsc[t];
      a.K[i] = K[t]; a.Cw[i] = Cw[t]; a.C[i] = C[t]; a.RS[i] = RS[t]; a.ld[i] = ld[t];
      const long long n = (long long)K[t] * ld[t];
      if (n > biggest) biggest = n;
    }
    const int bx = grid_for(biggest, 256, 2);
    if (out_dtype == B200_F32) pack_weight_multi_kernel<float><<<dim3(bx, cnt), 256, 0, st>>>(a);
    else pack_weight_multi_kernel<__nv_bfloat16><<<dim3(bx, cnt), 256, 0, st>>>(a);
    if (int32_t rc = check_launch("pack_weight_multi")) return rc;
  }
  return B200_OK;
}

extern "C" int32_t b200_cast(const void* x, void* y, int64_t n, int32_t in_dtype, int32_t out_dtype, void* stream) {
  cudaStream_t st = as_stream(stream);
  if (n <= 0) return B200_OK;
  if (in_dtype == B200_F32 && out_dtype == B200_BF16)
    cast_kernel<float, __nv_bfloat16><<<grid_for(n), 256, 0, st>>>((const float*)x, (__nv_bfloat16*)y, n);
  else if (in_dtype == B200_BF16 && out_dtype == B200_F32)
    cast_kernel<__nv_bfloat16, float><<<grid_for(n), 256, 0, st>>>((const __nv_bfloat16*)x, (float*)y, n);
  else if (in_dtype == B200_F32 && out_dtype == B200_F32)
    cast_kernel<float, float><<<grid_for(n), 256, 0, st>>>((const float*)x, (float*)y, n);
  else if (in_dtype == B200_BF16 && out_dtype == B200_BF16)
    cast_kernel<__nv_bfloat16, __nv_bfloat16><<<grid_for(n), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n);
  else return fail(B200_INVALID_ARG, "cast: bad dtypes");
  return check_launch("cast");
}

/* explicit staging of a convolution's A operand: col[m][(r*S+s)*C + c], row stride ld >= R*S*C, zero-filled tail */
extern "C" int32_t b200_stem_stage(const void* x, void* out, int32_t N, int32_t H, int32_t W, int32_t R, int32_t S,
                                   int32_t pad, void* stream) {
  if (!x || !out || N <= 0 || H <= 0 || W <= 0 || R <= 0 || S <= 0 || S > 8 || pad < 0)
    return fail(B200_INVALID_ARG, "stem_stage: bad arguments (8-byte pixels, stride 2, S <= 8)");
  if ((reinterpret_cast<uintptr_t>(x) & 7) || (reinterpret_cast<uintptr_t>(out) & 15))
    return fail(B200_INVALID_ARG, "stem_stage: misaligned buffers");
  const int Ho = (H + 2 * pad - R) / 2 + 1, Wo = (W + 2 * pad - S) / 2 + 1;
  if (Ho <= 0 || Wo <= 0) return fail(B200_INVALID_ARG, "stem_stage: filter larger than input");
  const int Hs = Ho + (R + 1) / 2 - 1;
  const long long total = (long long)N * Hs * Wo * 8;
  stem_stage_kernel<<<grid_for(total, 256, 16), 256, 0, as_stream(stream)>>>((const uint2*)x, (uint4*)out, N, H, W, pad, Hs, Wo);
  return check_launch("stem_stage");
}

extern "C" int32_t b200_stem_stage_nchw(const float* x_nchw, void* out, int32_t N, int32_t Cw, int32_t H, int32_t W,
                                        int32_t R, int32_t S, int32_t pad, void* stream) {
  if (!x_nchw || !out || N <= 0 || Cw <= 0 || Cw > 4 || H <= 0 || W <= 0 || R <= 0 || S <= 0 || S > 8 || pad < 0)
    return fail(B200_INVALID_ARG, "stem_stage_nchw: bad arguments (at most 4 channels, stride 2, S <= 8)");
  if (reinterpret_cast<uintptr_t>(out) & 15) return fail(B200_INVALID_ARG, "stem_stage_nchw: misaligned output");
  const int Ho = (H + 2 * pad - R) / 2 + 1, Wo = (W + 2 * pad - S) / 2 + 1;
  if (Ho <= 0 || Wo <= 0) return fail(B200_INVALID_ARG, "stem_stage_nchw: filter larger than input");
  const int Hs = Ho + (R + 1) / 2 - 1;
  const long long total = (long long)N * Hs * Wo * 8;
  stem_stage_nchw_kernel<<<grid_for(total, 256, 16), 256, 0, as_stream(stream)>>>(x_nchw, (uint4*)out, N, Cw, H, W, pad, Hs, Wo);
  return check_launch("stem_stage_nchw");
}

extern "C" int32_t b200_stem_pack_weight(const float* w_oihw, void* out, int32_t K, int32_t Cw, int32_t R, int32_t S,
                                         void* stream) {
  if (!w_oihw || !out || K <= 0 || Cw <= 0 || Cw > 4 || R <= 0 || S <= 0 || S > 8)
    return fail(B200_INVALID_ARG, "stem_pack_weight: bad arguments (at most 4 channels, S <= 8)");
  const int ld = ((R + 1) / 2) * 64;
  stem_pack_weight_kernel<<<grid_for((long long)K * ld), 256, 0, as_stream(stream)>>>(w_oihw, (__nv_bfloat16*)out, K, Cw, R, S, ld);
  return check_launch("stem_pack_weight");
}

extern "C" int32_t b200_im2col(const void* x, void* col, int32_t N, int32_t H, int32_t W, int32_t C, int32_t R,
                               int32_t S, int32_t stride, int32_t pad, int32_t ld, int32_t dtype, void* stream) {
  if (!x || !col || N <= 0 || H <= 0 || W <= 0 || C <= 0 || R <= 0 || S <= 0 || stride <= 0 || pad < 0 || ld < R * S * C)
    return fail(B200_INVALID_ARG, "im2col: bad arguments");
  if (dtype != B200_F32 && dtype != B200_BF16) return fail(B200_INVALID_ARG, "im2col: bad dtype");
  const int Ho = (H + 2 * pad - R) / stride + 1, Wo = (W + 2 * pad - S) / stride + 1;
  if (Ho <= 0 || Wo <= 0) return fail(B200_INVALID_ARG, "im2col: filter larger than input");
  return launch_im2col(x, col, dtype, N, H, W, C, R, S, stride, pad, Ho, Wo, ld, as_stream(stream));
}

// Deferred weight-gradient reductions (see reduce_wgrad_multi_kernel): between b200_wgrad_defer(1) and
// b200_wgrad_flush on the SAME host thread, b200_conv_wgrad leaves its split-K slabs in the workspace it was given
// (so every call needs its own) and dw is written by the flush.  b200_wgrad_defer(0) ends the mode and drops anything
// still queued.
extern "C" int32_t b200_wgrad_defer(int32_t on) {
  reduce_wgrad_defer(on != 0);
  return B200_OK;
}
extern "C" int32_t b200_wgrad_flush(void* stream) { return reduce_wgrad_flush(as_stream(stream)); }

