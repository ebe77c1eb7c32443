// Standalone GPU probe (no torch) for the stem convolution's operand fetch: a 5-D tiled tensor map with OVERLAPPING
// strides over a zero-padded NHWC4 bf16 image reads, for every output pixel q of a 7x7 / stride-2 convolution, the
// two input rows of a row pair as one 128-byte shared-memory row
//     [row 2j : 8 pixels x 4 channels = 32 elements][row 2j+1 : 32 elements]
// i.e. the im2col row of that pixel for two filter rows, without any staging pass in global memory.
//   dims    {32 elements, 2 rows of the pair, Wo, pairs, N}
//   strides {            row pitch,          16 B (= 2 pixels), 2 * pitch, image}
// The probe checks that the driver accepts the map (the strides are not monotonic) and that a box lands in shared
// memory in the order the UMMA K-major 128-byte-swizzled layout expects.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I<pkg>/csrc tools/stem_tma_probe.cu -o tools/stem_tma_probe
#include <cstdio>
#include <cstdlib>
#include <vector>
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

__device__ __forceinline__ void tma_load_5d_probe(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                                  int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], "
      "[%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

constexpr int BW = 8, BH = 4;   // box: 8 output columns x 4 row pairs = 32 rows of 128 bytes

__global__ void probe_kernel(const __grid_constant__ CUtensorMap tm, uint16_t* out, int q0, int j0, int n0, int swz,
                             int order) {
  __shared__ __align__(1024) uint8_t tile[BW * BH * 128];
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar, BW * BH * 128);
    if (order == 0) tma_load_5d_probe(tile, &tm, &bar, 0, 0, q0, j0, n0);   // {el, r2, q, pair, n}
    else tma_load_5d_probe(tile, &tm, &bar, 0, q0, 0, j0, n0);              // {el, q, r2, pair, n}
  }
  mbar_wait(&bar, 0);
  // un-swizzle (swz != 0): logical 16-byte chunk c of 128-byte row r sits at chunk c ^ (r & 7)
  for (int i = threadIdx.x; i < BW * BH * 64; i += blockDim.x) {
    const int r = i / 64, e = i % 64;
    const int chunk = swz ? ((e / 8) ^ (r & 7)) : (e / 8);
    out[i] = *reinterpret_cast<const uint16_t*>(tile + r * 128 + chunk * 16 + (e % 8) * 2);
  }
}

int main() {
  const int N = 3, H = 21, W = 18;                 // odd height on purpose
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  const int Hp = 2 * Ho + 6, Wp = 2 * Wo + 6;
  const size_t pitch = (size_t)Wp * 4;             // elements per padded row
  std::vector<__nv_bfloat16> h((size_t)N * Hp * pitch, __float2bfloat16(0.f));
  auto val = [&](int n, int y, int x, int c) { return (float)((n * 131 + y * 17 + x * 3 + c) % 251) - 125.f; };
  for (int n = 0; n < N; ++n)
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x)
        for (int c = 0; c < 3; ++c) h[((size_t)(n * Hp + y + 3) * Wp + x + 3) * 4 + c] = __float2bfloat16(val(n, y, x, c));
  void* d;
  CK(cudaMalloc(&d, h.size() * 2));
  CK(cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  EncodeTiledFn enc = tma_encode_fn();
  if (!enc) { printf("no encoder\n"); return 2; }
  uint16_t* dout;
  CK(cudaMalloc(&dout, BW * BH * 64 * 2));
  int bad = 0;
  for (int mode = 0; mode < 4; ++mode) {
  const int order = mode >> 1, swz = mode & 1;
  CUtensorMap tm;
  cuuint64_t dimsA[5] = {32, 2, (cuuint64_t)Wo, (cuuint64_t)(Ho + 3), (cuuint64_t)N};
  cuuint64_t stridesA[4] = {pitch * 2, 16, 2 * pitch * 2, (cuuint64_t)Hp * pitch * 2};
  cuuint32_t boxA[5] = {32, 2, BW, BH, 1};
  cuuint64_t dimsB[5] = {32, (cuuint64_t)Wo, 2, (cuuint64_t)(Ho + 3), (cuuint64_t)N};
  cuuint64_t stridesB[4] = {16, pitch * 2, 2 * pitch * 2, (cuuint64_t)Hp * pitch * 2};
  cuuint32_t boxB[5] = {32, BW, 2, BH, 1};
  cuuint32_t es[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, d, order ? dimsB : dimsA, order ? stridesB : stridesA,
                   order ? boxB : boxA, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("mode %d (order %d, swizzle %d): cuTensorMapEncodeTiled -> %d\n", mode, order, swz, (int)r);
  if (r != CUDA_SUCCESS) { ++bad; continue; }
  const int cases[3][3] = {{0, 0, 0}, {8, 4, 1}, {2, 9, 2}};   // second / third: boxes that run past Wo / the pair count
  for (auto& cs : cases) {
    const int q0 = cs[0], j0 = cs[1], n0 = cs[2];
    CK(cudaMemset(dout, 0xff, BW * BH * 64 * 2));
    probe_kernel<<<1, 128>>>(tm, dout, q0, j0, n0, swz, order);
    cudaError_t ke = cudaDeviceSynchronize();
    if (ke != cudaSuccess) { printf("  kernel failed: %s\n", cudaGetErrorString(ke)); return 3; }
    std::vector<uint16_t> o(BW * BH * 64);
    CK(cudaMemcpy(o.data(), dout, o.size() * 2, cudaMemcpyDeviceToHost));
    int cbad = 0;
    for (int jj = 0; jj < BH; ++jj)
      for (int qq = 0; qq < BW; ++qq)
        for (int e = 0; e < 64; ++e) {
          const int r2 = e / 32, px = (e % 32) / 4, c = e % 4;
          // linear position in the box (32-element units): order 0 = {r2, q, pair}, order 1 = {q, r2, pair}
          const int unit = order == 0 ? (jj * BW + qq) * 2 + r2 : (jj * 2 + r2) * BW + qq;
          const int row = unit / 2;
          const int eo = (unit % 2) * 32 + e % 32;
          const int q = q0 + qq, j = j0 + jj;
          float want = 0.f;
          if (q < Wo && j < Ho + 3) {                    // out-of-bounds box elements are zero-filled
            const int py = 2 * j + r2, pxx = 2 * q + px; // padded coordinates
            const int y = py - 3, x = pxx - 3;
            if (y >= 0 && y < H && x >= 0 && x < W && c < 3) want = val(n0, y, x, c);
          }
          __nv_bfloat16 got;
          *reinterpret_cast<uint16_t*>(&got) = o[row * 64 + eo];
          if (__bfloat162float(got) != want) {
            if (cbad < 5) printf("  mismatch box(q0=%d,j0=%d,n=%d) row %d e %d: got %f want %f\n", q0, j0, n0, row, e, __bfloat162float(got), want);
            ++cbad;
          }
        }
    printf("box (q0=%d, j0=%d, n=%d): %s\n", q0, j0, n0, cbad ? "FAIL" : "ok");
    bad += cbad;
  }
  }
  printf(bad ? "STEM TMA PROBE FAILED\n" : "STEM TMA PROBE PASSED\n");
  return bad ? 1 : 0;
}
