// Standalone GPU probe (no torch): can a tcgen05.mma operand descriptor start at an ARBITRARY ROW of a 128-byte
// swizzled shared-memory tile?  If yes, a 3x3 convolution can load one halo tile of activations per (pixel tile,
// channel chunk) and feed all nine filter taps from row-shifted views of it instead of fetching nine boxes from L2.
//
//   K-major A : tile [RT rows][64 k] bf16 (128-byte rows, TMA SWIZZLE_128B, 1024-byte aligned); the MMA's A operand
//               starts at row r0: D[i][n] = sum_k T[r0 + i][k] * B[n][k],  i < 128
//   MN-major B: tile [RT rows = k][64 n]; the MMA's B operand starts at k-row r0 (the weight-gradient form):
//               D[m][n] = sum_{k<64} A[k][m] * T[r0 + k][n]
// Each case is tried with the descriptor's base-offset field = 0 and = (r0 & 7).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I<pkg>/csrc tools/umma_shift_probe.cu -o tools/umma_shift_probe
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

constexpr int RT = 192;   // rows of the big tile

__device__ __forceinline__ uint64_t desc_with_base(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t base_off) {
  uint64_t d = umma_smem_desc(saddr, lbo, sbo);
  d |= static_cast<uint64_t>(base_off & 7u) << 49;
  return d;
}

// mode 0: K-major A shifted by r0 rows.  mode 1: MN-major B shifted by r0 k-rows.
__global__ void __launch_bounds__(128) shift_probe_kernel(const __grid_constant__ CUtensorMap tmT,
                                                          const __grid_constant__ CUtensorMap tmO, float* out, int mode,
                                                          int r0, int use_base) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);
  uint8_t* tileT = smem;                       // RT x 128 B
  uint8_t* tileO = smem + RT * 128;            // other operand: 128 x 128 B (mode 0: B [64 n][64 k]; mode 1: A [64 k][128 m] as 2 boxes)
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + RT * 128 + 128 * 128 * 2);
  uint64_t* done = bar + 1;
  uint32_t* slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(done, 1);
    mbar_fence_init();
  }
  if (warp == 0) {
    tmem_alloc(slot, 64);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  if (threadIdx.x == 0) {
    if (mode == 0) {
      mbar_expect_tx(bar, RT * 128 + 64 * 128);
      tma_load_2d(tileT, &tmT, bar, 0, 0);          // box {64 k, RT rows}
      tma_load_2d(tileO, &tmO, bar, 0, 0);          // box {64 k, 64 n}
    } else {
      mbar_expect_tx(bar, RT * 128 + 2 * 64 * 128);
      tma_load_2d(tileT, &tmT, bar, 0, 0);          // box {64 n, RT k-rows}
      tma_load_2d(tileO, &tmO, bar, 0, 0);          // A: box {64 m, 64 k} for m 0..63
      tma_load_2d(tileO + 8192, &tmO, bar, 64, 0);  //    and m 64..127
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    const uint32_t base_off = use_base ? static_cast<uint32_t>(r0 & 7) : 0u;
    if (mode == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, 64, false, false);
      const uint32_t a_addr = smem_u32(tileT) + r0 * 128, b_addr = smem_u32(tileO);
      for (int k = 0; k < 4; ++k)
        umma_f16(tmem, desc_with_base(a_addr + k * 32, 16, 1024, base_off), umma_smem_desc(b_addr + k * 32, 16, 1024), idesc, k > 0);
    } else {
      const uint32_t idesc = umma_idesc_bf16(128, 64, true, true);
      const uint32_t a_addr = smem_u32(tileO), b_addr = smem_u32(tileT) + r0 * 128;
      for (int k = 0; k < 4; ++k)     // 16 k-rows per step = 2048 bytes
        umma_f16(tmem, umma_smem_desc(a_addr + k * 2048, 8192, 1024), desc_with_base(b_addr + k * 2048, 8192, 1024, base_off), idesc, k > 0);
    }
    umma_commit(done);
  }
  mbar_wait(done, 0);
  tc_fence_after();
  uint32_t r[32];
  for (int c = 0; c < 2; ++c) {
    tmem_ld_32x32b_x32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + c * 32, r);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) out[(threadIdx.x) * 64 + c * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 64);
  }
}

static float bf(float f) { return __bfloat162float(__float2bfloat16_rn(f)); }

int main() {
  // T: [RT][64]; O (mode 0): B [64 n][64 k]; O (mode 1): A [64 k][128 m]
  std::vector<float> T((size_t)RT * 64), B(64 * 64), A(64 * 128);
  uint32_t s = 777u;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return ((s >> 8) & 0xFFFF) / 65536.0f - 0.5f; };
  for (auto& v : T) v = bf(rnd());
  for (auto& v : B) v = bf(rnd());
  for (auto& v : A) v = bf(rnd());
  auto up = [&](const std::vector<float>& h) {
    std::vector<__nv_bfloat16> t(h.size());
    for (size_t i = 0; i < h.size(); ++i) t[i] = __float2bfloat16(h[i]);
    void* d;
    CK(cudaMalloc(&d, t.size() * 2));
    CK(cudaMemcpy(d, t.data(), t.size() * 2, cudaMemcpyHostToDevice));
    return d;
  };
  void *dT = up(T), *dB = up(B), *dA = up(A);
  float* dout;
  CK(cudaMalloc(&dout, 128 * 64 * 4));
  CUtensorMap tmT, tmB, tmA;
  if (make_tmap_bf16_2d(&tmT, dT, 64, RT, 64, 64, RT) || make_tmap_bf16_2d(&tmB, dB, 64, 64, 64, 64, 64) ||
      make_tmap_bf16_2d(&tmA, dA, 128, 64, 128, 64, 64)) {
    printf("tensor map encode failed\n");
    return 2;
  }
  const int smem = RT * 128 + 128 * 128 * 2 + 1024 + 64;
  CK(cudaFuncSetAttribute(shift_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int fails = 0;
  const int shifts[] = {0, 8, 1, 3, 7, 9, 21, 42, 43};
  for (int mode = 0; mode < 2; ++mode)
    for (int r0 : shifts)
      for (int use_base = 0; use_base < 2; ++use_base) {
        if (use_base && (r0 & 7) == 0) continue;
        CK(cudaMemset(dout, 0, 128 * 64 * 4));
        shift_probe_kernel<<<1, 128, smem>>>(mode ? tmT : tmT, mode ? tmA : tmB, dout, mode, r0, use_base);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(e)); return 3; }
        std::vector<float> o(128 * 64);
        CK(cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost));
        int bad = 0;
        double worst = 0;
        for (int i = 0; i < 128; ++i)
          for (int n = 0; n < 64; ++n) {
            double ref = 0;
            if (mode == 0) for (int k = 0; k < 64; ++k) ref += (double)T[(size_t)(r0 + i) * 64 + k] * B[n * 64 + k];
            else for (int k = 0; k < 64; ++k) ref += (double)A[k * 128 + i] * T[(size_t)(r0 + k) * 64 + n];
            const double d = fabs(ref - o[i * 64 + n]);
            if (d > worst) worst = d;
            if (d > 1e-3) ++bad;
          }
        printf("mode %d (%s) shift %2d base_offset %d: %s (max err %.3g, bad %d)\n", mode, mode ? "MN-major B, k shift" : "K-major A, row shift",
               r0, use_base ? (r0 & 7) : 0, bad ? "FAIL" : "ok", worst, bad);
        if (bad && !use_base && (r0 & 7) == 0) ++fails;     // aligned shifts must work
      }
  printf(fails ? "UMMA SHIFT PROBE: aligned case FAILED\n" : "UMMA SHIFT PROBE DONE\n");
  return fails ? 1 : 0;
}
