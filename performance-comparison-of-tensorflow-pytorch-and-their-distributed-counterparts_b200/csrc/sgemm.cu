// fp32 CUDA-core GEMM (FFMA, fp32 accumulate): the exact-arithmetic path used in fp32 parity mode and for the
// small Linear layers of the head.  128x128x8 tiles, 256 threads, 8x8 outputs per thread, register-prefetched.
// Both operands may be K-major or MN-major (see include/b200net.h), split-K writes fp32 slabs.
#include "common.cuh"

namespace b200 {

constexpr int SG_BM = 128, SG_BN = 128, SG_BK = 8;

// loads 4 elements of operand X for tile row-range [x0, x0+128) and k-range [k0, k0+8) into regs
// K-major : thread -> (row = t/2, kk = (t%2)*4), 4 consecutive k
// MN-major: thread -> (kk = t/32, row = (t%32)*4), 4 consecutive rows
template <bool MN>
__device__ __forceinline__ void sg_load(const float* __restrict__ X, long long ld, int rows, int kend, int x0, int k0,
                                        int t, bool vec_ok, float (&v)[4]) {
  if (!MN) {
    const int r = x0 + (t >> 1), k = k0 + (t & 1) * 4;
    if (r < rows && vec_ok && k + 3 < kend) {
      float4 f = *reinterpret_cast<const float4*>(X + (long long)r * ld + k);
      v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) v[i] = (r < rows && k + i < kend) ? X[(long long)r * ld + k + i] : 0.f;
    }
  } else {
    const int k = k0 + (t >> 5), r = x0 + (t & 31) * 4;
    if (k < kend && vec_ok && r + 3 < rows) {
      float4 f = *reinterpret_cast<const float4*>(X + (long long)k * ld + r);
      v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) v[i] = (k < kend && r + i < rows) ? X[(long long)k * ld + r + i] : 0.f;
    }
  }
}
template <bool MN>
__device__ __forceinline__ void sg_store(float (*S)[SG_BM], int t, const float (&v)[4]) {
  if (!MN) {
    const int r = t >> 1, k = (t & 1) * 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) S[k + i][r] = v[i];
  } else {
    const int k = t >> 5, r = (t & 31) * 4;
    *reinterpret_cast<float4*>(&S[k][r]) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(256)
sgemm_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ C, int M, int N, int K,
             long long lda, long long ldb, long long ldc, const float* __restrict__ bias, int relu, int accumulate,
             int k_per_split, long long split_stride, int a_vec, int b_vec) {
  __shared__ __align__(16) float As[2][SG_BK][SG_BM];
  __shared__ __align__(16) float Bs[2][SG_BK][SG_BN];
  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.y * SG_BM, n0 = blockIdx.x * SG_BN;
  const int kbeg = blockIdx.z * k_per_split;
  int kend = kbeg + k_per_split;
  if (kend > K) kend = K;

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float ra[4], rb[4];
  sg_load<A_MN>(A, lda, M, kend, m0, kbeg, t, a_vec, ra);
  sg_load<B_MN>(B, ldb, N, kend, n0, kbeg, t, b_vec, rb);
  sg_store<A_MN>(As[0], t, ra);
  sg_store<B_MN>(Bs[0], t, rb);
  __syncthreads();
  int buf = 0;
  for (int k0 = kbeg; k0 < kend; k0 += SG_BK) {
    const bool has_next = k0 + SG_BK < kend;
    if (has_next) {
      sg_load<A_MN>(A, lda, M, kend, m0, k0 + SG_BK, t, a_vec, ra);
      sg_load<B_MN>(B, ldb, N, kend, n0, k0 + SG_BK, t, b_vec, rb);
    }
#pragma unroll
    for (int kk = 0; kk < SG_BK; ++kk) {
      float a[8], b[8];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (has_next) {
      sg_store<A_MN>(As[buf ^ 1], t, ra);
      sg_store<B_MN>(Bs[buf ^ 1], t, rb);
      __syncthreads();
      buf ^= 1;
    }
  }

  float* Cz = C + (long long)blockIdx.z * split_stride;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (n >= N) continue;
      float v = acc[i][j];
      if (bias) v += bias[n];
      float* p = Cz + (long long)m * ldc + n;
      if (accumulate) v += *p;
      if (relu) v = fmaxf(v, 0.f);
      *p = v;
    }
  }
}

// Small-problem variant: 32x32 output tiles (2x2 per thread), BK = 32.  The head's GEMMs ([256 x 2048] x [512 x 2048]^T
// and friends) give only a handful of 128x128 tiles; 32x32 tiles spread them over every SM.
template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(256)
sgemm_small_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ C, int M, int N, int K,
                   long long lda, long long ldb, long long ldc, const float* __restrict__ bias, int relu,
                   int accumulate) {
  __shared__ float As[32][33];  // [k][m]
  __shared__ float Bs[32][33];  // [k][n]
  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
  float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
  // the next k-tile's eight global loads are in flight while the current one is multiplied (the loop used to be
  // load -> sync -> multiply -> sync: one global-memory latency per 32 k, 82 us for the head's [256 x 2048] x [512 x 2048]^T)
  float ra[4], rb[4];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int idx = t + j * 256;
      int m, k;
      if (!A_MN) { k = idx & 31; m = idx >> 5; } else { m = idx & 31; k = idx >> 5; }
      const int gm = m0 + m, gk = k0 + k;
      ra[j] = (gm < M && gk < K) ? (A_MN ? A[(long long)gk * lda + gm] : A[(long long)gm * lda + gk]) : 0.f;
      int n;
      if (!B_MN) { k = idx & 31; n = idx >> 5; } else { n = idx & 31; k = idx >> 5; }
      const int gn = n0 + n, gk2 = k0 + k;
      rb[j] = (gn < N && gk2 < K) ? (B_MN ? B[(long long)gk2 * ldb + gn] : B[(long long)gn * ldb + gk2]) : 0.f;
    }
  };
  fetch(0);
  for (int k0 = 0; k0 < K; k0 += 32) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int idx = t + j * 256;
      if (!A_MN) As[idx & 31][idx >> 5] = ra[j]; else As[idx >> 5][idx & 31] = ra[j];
      if (!B_MN) Bs[idx & 31][idx >> 5] = rb[j]; else Bs[idx >> 5][idx & 31] = rb[j];
    }
    __syncthreads();
    if (k0 + 32 < K) fetch(k0 + 32);
#pragma unroll
    for (int kk = 0; kk < 32; ++kk) {
      const float a0 = As[kk][ty], a1 = As[kk][ty + 16];
      const float b0 = Bs[kk][tx], b1 = Bs[kk][tx + 16];
      acc[0][0] = fmaf(a0, b0, acc[0][0]);
      acc[0][1] = fmaf(a0, b1, acc[0][1]);
      acc[1][0] = fmaf(a1, b0, acc[1][0]);
      acc[1][1] = fmaf(a1, b1, acc[1][1]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int m = m0 + ty + i * 16;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int n = n0 + tx + j * 16;
      if (n >= N) continue;
      float v = acc[i][j];
      if (bias) v += bias[n];
      float* p = C + (long long)m * ldc + n;
      if (accumulate) v += *p;
      if (relu) v = fmaxf(v, 0.f);
      *p = v;
    }
  }
}

int32_t launch_sgemm(const float* A, const float* B, float* C, int M, int N, int K, long long lda, long long ldb,
                     long long ldc, bool a_mn, bool b_mn, const float* bias, int relu, int accumulate, int splits,
                     long long split_stride, cudaStream_t st) {
  if (M <= 0 || N <= 0 || K <= 0) return fail(B200_INVALID_ARG, "sgemm: empty problem");
  if (splits < 1) splits = 1;
  if (splits == 1 && (long long)((M + SG_BM - 1) / SG_BM) * ((N + SG_BN - 1) / SG_BN) < num_sms() / 2) {
    dim3 g((N + 31) / 32, (M + 31) / 32);
    if (!a_mn && !b_mn) sgemm_small_kernel<false, false><<<g, 256, 0, st>>>(A, B, C, M, N, K, lda, ldb, ldc, bias, relu, accumulate);
    else if (!a_mn && b_mn) sgemm_small_kernel<false, true><<<g, 256, 0, st>>>(A, B, C, M, N, K, lda, ldb, ldc, bias, relu, accumulate);
    else if (a_mn && !b_mn) sgemm_small_kernel<true, false><<<g, 256, 0, st>>>(A, B, C, M, N, K, lda, ldb, ldc, bias, relu, accumulate);
    else sgemm_small_kernel<true, true><<<g, 256, 0, st>>>(A, B, C, M, N, K, lda, ldb, ldc, bias, relu, accumulate);
    return check_launch("sgemm_small");
  }
  int k_per = ((K + splits - 1) / splits + SG_BK - 1) / SG_BK * SG_BK;
  splits = (K + k_per - 1) / k_per;
  const int a_vec = ((reinterpret_cast<uintptr_t>(A) & 15) == 0) && (lda % 4 == 0);
  const int b_vec = ((reinterpret_cast<uintptr_t>(B) & 15) == 0) && (ldb % 4 == 0);
  dim3 grid((N + SG_BN - 1) / SG_BN, (M + SG_BM - 1) / SG_BM, splits);
#define SG_LAUNCH(AM, BMJ)                                                                                         \
  sgemm_kernel<AM, BMJ><<<grid, 256, 0, st>>>(A, B, C, M, N, K, lda, ldb, ldc, bias, relu, accumulate, k_per,     \
                                              split_stride, a_vec, b_vec)
  if (!a_mn && !b_mn) SG_LAUNCH(false, false);
  else if (!a_mn && b_mn) SG_LAUNCH(false, true);
  else if (a_mn && !b_mn) SG_LAUNCH(true, false);
  else SG_LAUNCH(true, true);
#undef SG_LAUNCH
  return check_launch("sgemm");
}
// number of slabs launch_sgemm will write for (K, splits)
int sgemm_effective_splits(int K, int splits) {
  if (splits < 1) splits = 1;
  int k_per = ((K + splits - 1) / splits + SG_BK - 1) / SG_BK * SG_BK;
  return (K + k_per - 1) / k_per;
}

}  // namespace b200

using namespace b200;

extern "C" int32_t b200_gemm_f32(const float* A, const float* B, float* C, int32_t M, int32_t N, int32_t K,
                                 int64_t lda, int64_t ldb, int64_t ldc, int32_t a_mn, int32_t b_mn, const float* bias,
                                 int32_t relu, int32_t accumulate, void* stream) {
  if (!A || !B || !C) return fail(B200_INVALID_ARG, "gemm_f32: null pointer");
  return launch_sgemm(A, B, C, M, N, K, lda, ldb, ldc, a_mn != 0, b_mn != 0, bias, relu, accumulate, 1, 0,
                      as_stream(stream));
}
