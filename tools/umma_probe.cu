// Standalone GPU probe (no torch): validates the tcgen05 GEMM (all operand major-ness combinations, tails,
// split-K, accumulate) against a CPU reference, and checks the 4-D tiled-TMA behaviours the implicit-GEMM
// convolution relies on (negative / out-of-bounds coordinates -> zero fill, elementStrides traversal).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I<pkg>/csrc tools/umma_probe.cu -o tools/umma_probe
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <string>
#include "gemm_sm100.cuh"

using namespace b200;

#define CK(x)                                                                            \
  do {                                                                                   \
    cudaError_t e_ = (x);                                                                \
    if (e_ != cudaSuccess) {                                                             \
      printf("CUDA error %s at %s:%d: %s\n", #x, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(2);                                                                           \
    }                                                                                    \
  } while (0)

static float bf16_round(float f) { return __bfloat162float(__float2bfloat16_rn(f)); }

static uint32_t rng_state = 12345u;
static float frand() {
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xFFFF) / 65536.0f - 0.5f;
}

static int run_gemm_case(int M, int N, int K, bool a_mn, bool b_mn, int bn, int splits, bool out_fp32,
                         bool accumulate) {
  // storage: A K-major [M][K] or MN-major [K][M]; B K-major [N][K] or MN-major [K][N]
  const int Kp = (K + 7) / 8 * 8, Mp = (M + 7) / 8 * 8, Np = (N + 7) / 8 * 8;
  const long long lda = a_mn ? Mp : Kp, ldb = b_mn ? Np : Kp;
  std::vector<float> A((size_t)M * K), B((size_t)N * K);
  for (auto& v : A) v = bf16_round(frand());
  for (auto& v : B) v = bf16_round(frand());
  std::vector<__nv_bfloat16> hA((size_t)(a_mn ? K : M) * lda, __float2bfloat16(0.f));
  std::vector<__nv_bfloat16> hB((size_t)(b_mn ? K : N) * ldb, __float2bfloat16(0.f));
  for (int m = 0; m < M; ++m)
    for (int k = 0; k < K; ++k) hA[a_mn ? (size_t)k * lda + m : (size_t)m * lda + k] = __float2bfloat16(A[(size_t)m * K + k]);
  for (int n = 0; n < N; ++n)
    for (int k = 0; k < K; ++k) hB[b_mn ? (size_t)k * ldb + n : (size_t)n * ldb + k] = __float2bfloat16(B[(size_t)n * K + k]);
  std::vector<double> ref((size_t)M * N);
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double s = 0;
      for (int k = 0; k < K; ++k) s += (double)A[(size_t)m * K + k] * B[(size_t)n * K + k];
      ref[(size_t)m * N + n] = s;
    }
  const int eff = gemm_effective_splits(K, splits);
  const long long ldo = Np;
  void *dA, *dB, *dO;
  CK(cudaMalloc(&dA, hA.size() * 2));
  CK(cudaMalloc(&dB, hB.size() * 2));
  size_t out_elems = (size_t)M * ldo * (out_fp32 ? eff : 1);
  size_t out_bytes = out_elems * (out_fp32 ? 4 : 2);
  CK(cudaMalloc(&dO, out_bytes));
  CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
  std::vector<float> init((size_t)M * ldo, 0.f);
  if (accumulate) {
    for (auto& v : init) v = bf16_round(frand());
    if (out_fp32) CK(cudaMemcpy(dO, init.data(), init.size() * 4, cudaMemcpyHostToDevice));
    else {
      std::vector<__nv_bfloat16> t(init.size());
      for (size_t i = 0; i < t.size(); ++i) t[i] = __float2bfloat16(init[i]);
      CK(cudaMemcpy(dO, t.data(), t.size() * 2, cudaMemcpyHostToDevice));
    }
  } else {
    CK(cudaMemset(dO, 0xFF, out_bytes));  // NaN pattern: unwritten outputs show up
  }
  GemmProblem p;
  p.A = dA; p.B = dB; p.lda = lda; p.ldb = ldb; p.M = M; p.N = N; p.K = K; p.a_mn = a_mn; p.b_mn = b_mn;
  p.splits = splits;
  p.out.ptr = dO; p.out.ld = ldo; p.out.split_stride = (long long)M * ldo; p.out.out_fp32 = out_fp32;
  p.out.accumulate = accumulate; p.out.vec_ok = 1;
  int rc = launch_gemm_sm100(p, 0, bn);
  cudaError_t e = cudaDeviceSynchronize();
  if (rc || e != cudaSuccess) {
    printf("  launch rc=%d sync=%s\n", rc, cudaGetErrorString(e));
    return 1;
  }
  std::vector<float> got((size_t)M * ldo, 0.f);
  if (out_fp32) {
    std::vector<float> t(out_elems);
    CK(cudaMemcpy(t.data(), dO, out_bytes, cudaMemcpyDeviceToHost));
    for (int s = 0; s < eff; ++s)
      for (size_t i = 0; i < got.size(); ++i) got[i] += t[(size_t)s * M * ldo + i];
  } else {
    std::vector<__nv_bfloat16> t(out_elems);
    CK(cudaMemcpy(t.data(), dO, out_bytes, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < got.size(); ++i) got[i] = __bfloat162float(t[i]);
  }
  double max_err = 0, max_ref = 0;
  int bad = 0;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double r = ref[(size_t)m * N + n] + (accumulate ? init[(size_t)m * ldo + n] : 0.0);
      double g = got[(size_t)m * ldo + n];
      double err = fabs(r - g);
      double tol = out_fp32 ? 1e-3 + 1e-4 * fabs(r) : 2e-2 + 1e-2 * fabs(r);
      if (!(err <= tol)) {
        if (bad < 4) printf("    mismatch m=%d n=%d ref=%g got=%g\n", m, n, r, g);
        ++bad;
      }
      if (err > max_err) max_err = err;
      if (fabs(r) > max_ref) max_ref = fabs(r);
    }
  CK(cudaFree(dA)); CK(cudaFree(dB)); CK(cudaFree(dO));
  printf("  M=%d N=%d K=%d a_mn=%d b_mn=%d bn=%d splits=%d(eff %d) fp32=%d acc=%d : %s max_err=%.3g max_ref=%.3g bad=%d\n",
         M, N, K, a_mn, b_mn, bn, splits, eff, out_fp32, accumulate, bad ? "FAIL" : "PASS", max_err, max_ref, bad);
  return bad ? 1 : 0;
}

// ------------------------------------------------------------------------ 4-D TMA probe
__global__ void tma4d_probe_kernel(const __grid_constant__ CUtensorMap tm, int c0, int c1, int c2, int c3,
                                   uint32_t bytes, uint8_t* dump) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  for (uint32_t i = threadIdx.x; i < bytes; i += blockDim.x) smem[i] = 0xEE;  // sentinel
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar, bytes);
    tma_load_4d(smem, &tm, &bar, c0, c1, c2, c3);
  }
  mbar_wait(&bar, 0);
  for (uint32_t i = threadIdx.x; i < bytes; i += blockDim.x) dump[i] = smem[i];
}

static int run_tma4d_case(int C, int W, int H, int NB, int bc, int bw, int bh, int bn, int es_w, int es_h, int c0,
                          int w0, int h0, int n0) {
  // tensor NHWC bf16, value encodes its own coordinates
  std::vector<__nv_bfloat16> h((size_t)NB * H * W * C);
  auto val = [&](int n, int y, int x, int c) { return (float)(((n * 7 + y) * 5 + x) % 61 + 1) + (c % 4) * 0.25f; };
  for (int n = 0; n < NB; ++n)
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x)
        for (int c = 0; c < C; ++c) h[(((size_t)n * H + y) * W + x) * C + c] = __float2bfloat16(val(n, y, x, c));
  void* d;
  CK(cudaMalloc(&d, h.size() * 2));
  CK(cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  CUtensorMap tm;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)NB};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {(cuuint32_t)bc, (cuuint32_t)(bw * es_w), (cuuint32_t)(bh * es_h), (cuuint32_t)bn};
  cuuint32_t es[4] = {1, (cuuint32_t)es_w, (cuuint32_t)es_h, 1};
  CUresult r = tma_encode_fn()(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, d, dims, strides, box, es,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    printf("  tma4d encode failed rc=%d\n", (int)r);
    return 1;
  }
  const uint32_t rows = bw * bh * bn;
  const uint32_t bytes = rows * bc * 2;
  uint8_t* dd;
  CK(cudaMalloc(&dd, bytes));
  tma4d_probe_kernel<<<1, 128, bytes + 1024>>>(tm, c0, w0, h0, n0, bytes, dd);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("  tma4d kernel failed: %s\n", cudaGetErrorString(e));
    return 1;
  }
  std::vector<uint8_t> dump(bytes);
  CK(cudaMemcpy(dump.data(), dd, bytes, cudaMemcpyDeviceToHost));
  // hypothesis: smem row index = (in*bh + ih)*bw + iw, 128 B per row when bc==64, 16-byte chunk j stored at j ^ (row & 7)
  int bad = 0;
  const int row_bytes = bc * 2;
  for (int in = 0; in < bn; ++in)
    for (int ih = 0; ih < bh; ++ih)
      for (int iw = 0; iw < bw; ++iw) {
        const int row = (in * bh + ih) * bw + iw;
        const int n = n0 + in, y = h0 + ih * es_h, x = w0 + iw * es_w;
        for (int c = 0; c < bc; ++c) {
          const int chunk = (c * 2) / 16, within = (c * 2) % 16;
          const int pchunk = row_bytes == 128 ? (chunk ^ (row & 7)) : chunk;
          const size_t off = (size_t)row * row_bytes + pchunk * 16 + within;
          __nv_bfloat16 g;
          memcpy(&g, &dump[off], 2);
          const bool inb = n >= 0 && n < NB && y >= 0 && y < H && x >= 0 && x < W && (c0 + c) < C;
          const float expect = inb ? bf16_round(val(n, y, x, c0 + c)) : 0.f;
          if (__bfloat162float(g) != expect) {
            if (bad < 4) printf("    tma4d mismatch row=%d c=%d expect=%g got=%g\n", row, c, expect, __bfloat162float(g));
            ++bad;
          }
        }
      }
  CK(cudaFree(d)); CK(cudaFree(dd));
  printf("  tma4d C=%d W=%d H=%d N=%d box=(%d,%d,%d,%d) es=(%d,%d) at (%d,%d,%d,%d): %s bad=%d\n", C, W, H, NB, bc, bw,
         bh, bn, es_w, es_h, c0, w0, h0, n0, bad ? "FAIL" : "PASS", bad);
  return bad ? 1 : 0;
}

int main(int argc, char** argv) {
  std::string which = argc > 1 ? argv[1] : "all";
  int fails = 0;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("device: %s sm_%d%d SMs=%d\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
  if (which == "all" || which == "nt") {
    printf("[NT: A K-major, B K-major]\n");
    fails += run_gemm_case(128, 64, 64, false, false, 64, 1, true, false);
    fails += run_gemm_case(128, 128, 256, false, false, 128, 1, true, false);
    fails += run_gemm_case(300, 200, 200, false, false, 128, 1, true, false);
    fails += run_gemm_case(300, 520, 1000, false, false, 256, 1, false, false);
    fails += run_gemm_case(1000, 72, 136, false, false, 64, 1, false, true);
    fails += run_gemm_case(256, 256, 4096, false, false, 256, 5, true, false);
  }
  if (which == "all" || which == "tn") {
    printf("[TN: A MN-major, B MN-major]\n");
    fails += run_gemm_case(128, 64, 64, true, true, 64, 1, true, false);
    fails += run_gemm_case(128, 128, 256, true, true, 128, 1, true, false);
    fails += run_gemm_case(200, 264, 1000, true, true, 256, 1, true, false);
    fails += run_gemm_case(64, 576, 5000, true, true, 256, 7, true, false);
  }
  if (which == "all" || which == "mixed") {
    printf("[mixed: A K-major, B MN-major]\n");
    fails += run_gemm_case(128, 128, 128, false, true, 128, 1, true, false);
    fails += run_gemm_case(333, 264, 520, false, true, 256, 1, false, false);
    printf("[mixed: A MN-major, B K-major]\n");
    fails += run_gemm_case(128, 128, 128, true, false, 128, 1, true, false);
    fails += run_gemm_case(333, 72, 520, true, false, 64, 1, true, true);
  }
  if (which == "all" || which == "tma") {
    printf("[4-D tiled TMA]\n");
    fails += run_tma4d_case(64, 20, 20, 3, 64, 8, 8, 2, 1, 1, 0, 0, 0, 0);     // interior
    fails += run_tma4d_case(64, 20, 20, 3, 64, 8, 8, 2, 1, 1, 0, -1, -1, 1);   // negative start (padding)
    fails += run_tma4d_case(128, 20, 20, 3, 64, 8, 8, 2, 1, 1, 64, 15, 16, 2); // high OOB in w,h,n
    fails += run_tma4d_case(64, 21, 21, 2, 64, 8, 8, 2, 2, 2, 0, -1, -1, 0);   // stride 2 with padding
    fails += run_tma4d_case(64, 10, 10, 4, 64, 10, 6, 2, 1, 1, 0, 0, 6, 1);    // non-pow2 box
    fails += run_tma4d_case(64, 5, 5, 8, 64, 5, 5, 5, 1, 1, 0, 0, 0, 2);       // 5x5x5 box
  }
  printf("PROBE %s (%d failing cases)\n", fails ? "FAILED" : "OK", fails);
  return fails ? 1 : 0;
}
