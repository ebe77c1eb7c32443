// Convolution fprop / dgrad / wgrad as GEMMs over NHWC activations.
//   fp32  : exact FFMA GEMM (sgemm.cu) -- the parity mode
//   bf16  : tcgen05 GEMM (gemm_sm100.cuh), fp32 accumulation in TMEM
// 1x1 / stride-1 convolutions run directly on the activation matrix [N*H*W][C]; everything else goes through an
// explicit im2col / col2im staging buffer in the caller-provided workspace (round-1 path; the 4-D TMA implicit
// path validated by tools/umma_probe.cu replaces the staging for the bf16 3x3 convolutions).
#include "common.cuh"
#include "gemm_sm100.cuh"
#include <vector>
#include <utility>

namespace b200 {
int32_t launch_sgemm(const float*, const float*, float*, int, int, int, long long, long long, long long, bool, bool,
                     const float*, int, int, int, long long, cudaStream_t);
int sgemm_effective_splits(int K, int splits);
int32_t launch_im2col(const void*, void*, int, int, int, int, int, int, int, int, int, int, int, int, cudaStream_t);
int32_t launch_col2im(const float*, void*, int, int, int, int, int, int, int, int, int, int, int, int, int,
                      cudaStream_t);
int32_t launch_reduce_wgrad(const float*, int, long long, int, int, int, int, float*, cudaStream_t);
int32_t launch_reduce_plain(const float*, int, long long, int, int, long long, void*, long long, int, int,
                            cudaStream_t);

struct ConvGeom {
  int Ho, Wo, RSC, ldw;
  long long M;
  bool direct;
  size_t esz;
};
static inline ConvGeom geom(const b200_conv_desc* d) {
  ConvGeom g;
  g.Ho = (d->H + 2 * d->pad - d->R) / d->stride + 1;
  g.Wo = (d->W + 2 * d->pad - d->S) / d->stride + 1;
  g.RSC = d->R * d->S * d->C;
  g.ldw = (g.RSC + 7) / 8 * 8;
  g.M = (long long)d->N * g.Ho * g.Wo;
  g.direct = d->R == 1 && d->S == 1 && d->stride == 1 && d->pad == 0 && (d->C % 8 == 0);
  g.esz = d->dtype == B200_BF16 ? 2 : 4;
  return g;
}
static inline int32_t check_desc(const b200_conv_desc* d) {
  if (!d) return fail(B200_INVALID_ARG, "conv: null descriptor");
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->C <= 0 || d->K <= 0 || d->R <= 0 || d->S <= 0 || d->stride <= 0 ||
      d->pad < 0)
    return fail(B200_INVALID_ARG, "conv: non-positive dimension");
  if (d->dtype != B200_F32 && d->dtype != B200_BF16) return fail(B200_INVALID_ARG, "conv: unknown dtype");
  if (d->H + 2 * d->pad < d->R || d->W + 2 * d->pad < d->S) return fail(B200_INVALID_ARG, "conv: filter larger than input");
  if (d->dtype == B200_BF16 && d->K % 8 != 0)
    return fail(B200_UNSUPPORTED, "conv(bf16): output channels must be a multiple of 8 (got %d)", d->K);
  return B200_OK;
}

static inline int wgrad_splits(const b200_conv_desc* d, const ConvGeom& g) {
  const int bn = d->dtype == B200_BF16 ? (g.RSC > 128 ? 256 : (g.RSC > 64 ? 128 : 64)) : 128;
  const long long tiles = (long long)((d->K + 127) / 128) * ((g.RSC + bn - 1) / bn);
  long long s = (2LL * num_sms() + tiles - 1) / tiles;
  const long long kb = (g.M + 63) / 64;
  if (s > kb) s = kb;
  if (s > 512) s = 512;
  const long long slab = (long long)d->K * g.ldw * 4;
  while (s > 1 && s * slab > (512LL << 20)) --s;
  if (s < 1) s = 1;
  return (int)s;
}

// optional per-launch timing of the tcgen05 GEMM (bench.py roofline leg): CUDA events on the launching stream
struct GemmTiming {
  bool on = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev;
  double flops = 0;
};
static GemmTiming& gemm_timing() {
  static GemmTiming t;
  return t;
}

static int32_t gemm_bf16_run(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                             long long ldc, bool a_mn, bool b_mn, int out_dtype, int accumulate, int splits,
                             long long split_stride, cudaStream_t st) {
  GemmProblem p;
  p.A = A; p.B = B; p.lda = lda; p.ldb = ldb; p.M = M; p.N = N; p.K = K; p.a_mn = a_mn; p.b_mn = b_mn;
  p.splits = splits;
  p.out.ptr = C; p.out.ld = ldc; p.out.split_stride = split_stride; p.out.out_fp32 = out_dtype == B200_F32;
  p.out.accumulate = accumulate;
  p.out.vec_ok = ((reinterpret_cast<uintptr_t>(C) & 15) == 0) && (ldc % (out_dtype == B200_F32 ? 4 : 8) == 0);
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) || (lda % 8) || (ldb % 8))
    return fail(B200_UNSUPPORTED, "gemm_bf16: operands must be 16-byte aligned with leading dimensions multiple of 8");
  GemmTiming& gt = gemm_timing();
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (gt.on) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, st);
  }
  int rc = launch_gemm_sm100(p, st);
  if (gt.on) {
    cudaEventRecord(e1, st);
    gt.ev.emplace_back(e0, e1);
    gt.flops += 2.0 * M * N * K;
  }
  launch_counter() += 1;
  if (rc) return fail(B200_CUDA_ERROR, "gemm_bf16: launch failed (code %d: %s)", rc, cudaGetErrorString(cudaGetLastError()));
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int32_t b200_conv_weight_ld(int32_t R, int32_t S, int32_t C) { return (R * S * C + 7) / 8 * 8; }

extern "C" int64_t b200_conv_workspace_bytes(const b200_conv_desc* d, int32_t which) {
  if (check_desc(d)) return -1;
  const ConvGeom g = geom(d);
  size_t col = g.direct ? 0 : align_up((size_t)g.M * g.ldw * g.esz, 256);
  if (which == B200_CONV_FPROP) return (int64_t)col;
  if (which == B200_CONV_DGRAD) return g.direct ? 0 : (int64_t)align_up((size_t)g.M * g.ldw * 4, 256);
  if (which == B200_CONV_WGRAD) {
    const int s = wgrad_splits(d, g);
    return (int64_t)(col + align_up((size_t)s * d->K * g.ldw * 4, 256));
  }
  return -1;
}

extern "C" int32_t b200_conv_fprop(const b200_conv_desc* d, const void* x, const void* w, void* y, void* ws,
                                   int64_t ws_bytes, void* stream) {
  if (int32_t rc = check_desc(d)) return rc;
  if (!x || !w || !y) return fail(B200_INVALID_ARG, "conv_fprop: null pointer");
  const ConvGeom g = geom(d);
  cudaStream_t st = as_stream(stream);
  const void* A = x;
  long long lda = d->C;
  int Kred = d->C;
  if (!g.direct) {
    if (ws_bytes < b200_conv_workspace_bytes(d, B200_CONV_FPROP) || !ws)
      return fail(B200_WORKSPACE_TOO_SMALL, "conv_fprop: workspace too small");
    if (int32_t rc = launch_im2col(x, ws, d->dtype, d->N, d->H, d->W, d->C, d->R, d->S, d->stride, d->pad, g.Ho, g.Wo,
                                   g.ldw, st))
      return rc;
    A = ws;
    lda = g.ldw;
    Kred = g.ldw;
  }
  if (d->dtype == B200_F32)
    return launch_sgemm((const float*)A, (const float*)w, (float*)y, (int)g.M, d->K, Kred, lda, g.ldw, d->K, false,
                        false, nullptr, 0, 0, 1, 0, st);
  return gemm_bf16_run(A, w, y, (int)g.M, d->K, Kred, lda, g.ldw, d->K, false, false, B200_BF16, 0, 1, 0, st);
}

extern "C" int32_t b200_conv_dgrad(const b200_conv_desc* d, const void* dy, const void* w, void* dx,
                                   int32_t accumulate, void* ws, int64_t ws_bytes, void* stream) {
  if (int32_t rc = check_desc(d)) return rc;
  if (!dy || !w || !dx) return fail(B200_INVALID_ARG, "conv_dgrad: null pointer");
  const ConvGeom g = geom(d);
  cudaStream_t st = as_stream(stream);
  if (g.direct) {
    // dx[M][C] (+)= dy[M][K] * W[K][C]
    if (d->dtype == B200_F32)
      return launch_sgemm((const float*)dy, (const float*)w, (float*)dx, (int)g.M, d->C, d->K, d->K, g.ldw, d->C,
                          false, true, nullptr, 0, accumulate, 1, 0, st);
    return gemm_bf16_run(dy, w, dx, (int)g.M, d->C, d->K, d->K, g.ldw, d->C, false, true, B200_BF16, accumulate, 1, 0,
                         st);
  }
  if (ws_bytes < b200_conv_workspace_bytes(d, B200_CONV_DGRAD) || !ws)
    return fail(B200_WORKSPACE_TOO_SMALL, "conv_dgrad: workspace too small");
  float* dcol = (float*)ws;  // [M][ldw] fp32 = dy[M][K] * W[K][RSC]
  int32_t rc;
  if (d->dtype == B200_F32)
    rc = launch_sgemm((const float*)dy, (const float*)w, dcol, (int)g.M, g.RSC, d->K, d->K, g.ldw, g.ldw, false, true,
                      nullptr, 0, 0, 1, 0, st);
  else
    rc = gemm_bf16_run(dy, w, dcol, (int)g.M, g.RSC, d->K, d->K, g.ldw, g.ldw, false, true, B200_F32, 0, 1, 0, st);
  if (rc) return rc;
  return launch_col2im(dcol, dx, d->dtype, d->N, d->H, d->W, d->C, d->R, d->S, d->stride, d->pad, g.Ho, g.Wo, g.ldw,
                       accumulate, st);
}

extern "C" int32_t b200_conv_wgrad(const b200_conv_desc* d, const void* x, const void* dy, float* dw, void* ws,
                                   int64_t ws_bytes, void* stream) {
  if (int32_t rc = check_desc(d)) return rc;
  if (!x || !dy || !dw) return fail(B200_INVALID_ARG, "conv_wgrad: null pointer");
  const ConvGeom g = geom(d);
  cudaStream_t st = as_stream(stream);
  if (ws_bytes < b200_conv_workspace_bytes(d, B200_CONV_WGRAD) || !ws)
    return fail(B200_WORKSPACE_TOO_SMALL, "conv_wgrad: workspace too small");
  const size_t col_bytes = g.direct ? 0 : align_up((size_t)g.M * g.ldw * g.esz, 256);
  const void* Bm = x;
  long long ldb = d->C;
  int Ncols = d->C;
  if (!g.direct) {
    if (int32_t rc = launch_im2col(x, ws, d->dtype, d->N, d->H, d->W, d->C, d->R, d->S, d->stride, d->pad, g.Ho, g.Wo,
                                   g.ldw, st))
      return rc;
    Bm = ws;
    ldb = g.ldw;
    Ncols = g.RSC;
  }
  float* slabs = reinterpret_cast<float*>(static_cast<uint8_t*>(ws) + col_bytes);
  const int want = wgrad_splits(d, g);
  const long long slab_stride = (long long)d->K * g.ldw;
  int eff;
  int32_t rc;
  // dWp[K][RSC] = dy^T[K][M] * col[M][RSC] : both operands have the reduction (pixel) index outermost
  if (d->dtype == B200_F32) {
    eff = sgemm_effective_splits((int)g.M, want);
    rc = launch_sgemm((const float*)dy, (const float*)Bm, slabs, d->K, Ncols, (int)g.M, d->K, ldb, g.ldw, true, true,
                      nullptr, 0, 0, want, slab_stride, st);
  } else {
    eff = gemm_effective_splits((int)g.M, want);
    rc = gemm_bf16_run(dy, Bm, slabs, d->K, Ncols, (int)g.M, d->K, ldb, g.ldw, true, true, B200_F32, 0, want,
                       slab_stride, st);
  }
  if (rc) return rc;
  return launch_reduce_wgrad(slabs, eff, slab_stride, d->K, d->C, d->R * d->S, g.ldw, dw, st);
}

extern "C" int32_t b200_gemm_bf16(const void* A, const void* B, void* C, int32_t M, int32_t N, int32_t K, int64_t lda,
                                  int64_t ldb, int64_t ldc, int32_t a_mn, int32_t b_mn, int32_t out_dtype,
                                  int32_t accumulate, int32_t splits, void* ws, int64_t ws_bytes, void* stream) {
  if (!A || !B || !C || M <= 0 || N <= 0 || K <= 0) return fail(B200_INVALID_ARG, "gemm_bf16: bad arguments");
  cudaStream_t st = as_stream(stream);
  const int eff = gemm_effective_splits(K, splits);
  if (eff <= 1) return gemm_bf16_run(A, B, C, M, N, K, lda, ldb, ldc, a_mn != 0, b_mn != 0, out_dtype, accumulate, 1, 0, st);
  const long long slab = (long long)M * ldc;
  if (!ws || ws_bytes < (int64_t)eff * slab * 4) return fail(B200_WORKSPACE_TOO_SMALL, "gemm_bf16: split-K workspace too small");
  if (int32_t rc = gemm_bf16_run(A, B, ws, M, N, K, lda, ldb, ldc, a_mn != 0, b_mn != 0, B200_F32, 0, splits, slab, st))
    return rc;
  return launch_reduce_plain((const float*)ws, eff, slab, M, N, ldc, C, ldc, out_dtype, accumulate, st);
}

extern "C" int32_t b200_gemm_timing_enable(int32_t on) {
  GemmTiming& gt = gemm_timing();
  for (auto& pr : gt.ev) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  gt.ev.clear();
  gt.flops = 0;
  gt.on = on != 0;
  return B200_OK;
}
extern "C" int32_t b200_gemm_timing_read(double* total_ms, int64_t* launches, double* flops) {
  GemmTiming& gt = gemm_timing();
  double ms = 0;
  for (auto& pr : gt.ev) {
    if (cudaEventSynchronize(pr.second) != cudaSuccess) return fail(B200_CUDA_ERROR, "gemm_timing_read: event sync failed");
    float t = 0;
    cudaEventElapsedTime(&t, pr.first, pr.second);
    ms += t;
  }
  if (total_ms) *total_ms = ms;
  if (launches) *launches = (int64_t)gt.ev.size();
  if (flops) *flops = gt.flops;
  return B200_OK;
}
