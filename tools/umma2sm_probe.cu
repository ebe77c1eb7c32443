// Standalone probe of the 2-SM tensor-core path (tcgen05 cta_group::2) before it goes into the GEMM kernels:
// one cluster of two CTAs computes D[256][256] = A[256][K] * B[256][K]^T (bf16 -> fp32).  Each CTA loads ITS 128 rows
// of A and ITS 128 rows of B by TMA (cta_group::2 form: the bytes are counted on the leader CTA's mbarrier), the
// leader issues tcgen05.mma.cta_group::2 (M = 256, N = 256) reading both CTAs' shared memory, the commit is
// multicast to both CTAs, and each CTA drains its own 128 accumulator rows from its TMEM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I<pkg>/csrc tools/umma2sm_probe.cu -o tools/umma2sm_probe
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "gemm_sm100.cuh"

using namespace b200;

#define CK(x)                                                                                 \
  do {                                                                                        \
    cudaError_t e_ = (x);                                                                     \
    if (e_ != cudaSuccess) {                                                                  \
      printf("CUDA error %s at %s:%d: %s\n", #x, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(2);                                                                                \
    }                                                                                         \
  } while (0)

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA load whose completion bytes are counted on the LEADER CTA's barrier (same smem offset, peer bit cleared)
__device__ __forceinline__ void tma_load_2d_2sm(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], "
      "[%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_2sm() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_f16_2sm(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once all earlier MMAs of this thread are done) on the barrier at this smem offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}

constexpr int KB = 64;   // k-block: 64 bf16 = one 128-byte swizzled row

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, float* __restrict__ D,
             int nkb) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);
  uint8_t* smA = smem;
  uint8_t* smB = smem + nkb * 16384;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + 2 * nkb * 16384);
  uint64_t* done = full + 1;
  uint32_t* slot = reinterpret_cast<uint32_t*>(done + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t rank = cluster_ctarank();
  if (tid == 0) {
    mbar_init(full, 1);
    mbar_init(done, 1);
    mbar_fence_init();
  }
  cluster_sync_all();
  if (warp == 0) {
    tmem_alloc_2sm(slot, 256);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *slot;
  if (tid == 0) {
    if (rank == 0) mbar_expect_tx(full, (uint32_t)nkb * 4u * 16384u);   // A and B tiles of BOTH CTAs
    for (int kb = 0; kb < nkb; ++kb) {
      tma_load_2d_2sm(smA + kb * 16384, &tmA, full, kb * KB, (int)rank * 128);
      tma_load_2d_2sm(smB + kb * 16384, &tmB, full, kb * KB, (int)rank * 128);
    }
  }
  if (rank == 0 && warp == 0) {
    if (tid == 0) {
      mbar_wait(full, 0);
      tc_fence_after();
      const uint32_t idesc = umma_idesc_bf16(256, 256, false, false);
      for (int kb = 0; kb < nkb; ++kb) {
        const uint32_t a = smem_u32(smA + kb * 16384), b = smem_u32(smB + kb * 16384);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_f16_2sm(tmem_base, umma_smem_desc(a + k * 32, 16, 1024), umma_smem_desc(b + k * 32, 16, 1024), idesc,
                       (kb > 0 || k > 0) ? 1u : 0u);
      }
      umma_commit_2sm(done);
    }
    __syncwarp();
  }
  mbar_wait(done, 0);
  tc_fence_after();
  const uint32_t trow = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
  float* out = D + ((size_t)rank * 128 + tid) * 256;
  for (int ch = 0; ch < 8; ++ch) {
    uint32_t r[32];
    tmem_ld_32x32b_x32(trow + ch * 32, r);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) out[ch * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc_2sm(tmem_base, 256);
}

int main() {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("device: %s sm_%d%d SMs=%d\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
  int fails = 0;
  for (int nkb = 1; nkb <= 4; nkb += 3) {
    const int K = nkb * KB, M = 256, N = 256;
    std::vector<float> A((size_t)M * K), B((size_t)N * K);
    uint32_t st = 777u + nkb;
    auto rnd = [&]() { st = st * 1664525u + 1013904223u; return ((st >> 8) & 0xFFFF) / 65536.0f - 0.5f; };
    std::vector<__nv_bfloat16> hA(A.size()), hB(B.size());
    for (size_t i = 0; i < A.size(); ++i) { hA[i] = __float2bfloat16(rnd()); A[i] = __bfloat162float(hA[i]); }
    for (size_t i = 0; i < B.size(); ++i) { hB[i] = __float2bfloat16(rnd()); B[i] = __bfloat162float(hB[i]); }
    void *dA, *dB;
    float* dD;
    CK(cudaMalloc(&dA, hA.size() * 2));
    CK(cudaMalloc(&dB, hB.size() * 2));
    CK(cudaMalloc(&dD, (size_t)M * N * 4));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0xFF, (size_t)M * N * 4));
    CUtensorMap ta, tb;
    if (make_tmap_bf16_2d(&ta, dA, K, M, K, KB, 128) || make_tmap_bf16_2d(&tb, dB, K, N, K, KB, 128)) {
      printf("tensor map failed\n");
      return 2;
    }
    const int smem = 2 * nkb * 16384 + 64 + 1024;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    probe_kernel<<<2, 128, smem>>>(ta, tb, dD, nkb);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    std::vector<float> D((size_t)M * N);
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double maxerr = 0, maxref = 0;
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) {
        double s = 0;
        for (int k = 0; k < K; ++k) s += (double)A[(size_t)m * K + k] * B[(size_t)n * K + k];
        maxerr = fmax(maxerr, fabs(s - D[(size_t)m * N + n]));
        maxref = fmax(maxref, fabs(s));
      }
    const bool ok = maxerr <= 1e-3 * maxref;
    printf("2-SM MMA M=256 N=256 K=%d: max err %.3e (ref scale %.3e) %s\n", K, maxerr, maxref, ok ? "ok" : "FAIL");
    fails += !ok;
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
  }
  printf("PROBE2SM %s\n", fails ? "FAILED" : "OK");
  return fails ? 1 : 0;
}
