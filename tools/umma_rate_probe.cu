// Standalone GPU probe (no torch): how many cycles does one tcgen05.mma (M = 128, K = 16, bf16) cost on B200 as a
// function of N, of whether consecutive MMAs accumulate into the SAME TMEM columns, of how often the issuing thread
// commits, and of where the operands' shared-memory rows start?  One CTA per SM; operands are zero-filled shared memory
// (the values do not matter), one elected thread issues `iters` batches of four k-steps and the CTA clocks the span
// from the first issue to the arrival of the final commit.  `delay` adds that many cycles of unrelated work to the
// issuing thread after every batch: if the total does not grow, the MMA queue is deep enough to hide it.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I<pkg>/csrc tools/umma_rate_probe.cu -o tools/umma_rate_probe
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

// nacc: accumulators the batches rotate over; commit_every: batches between commits (0 = only at the end);
// bmn: B operand MN-major (the dgrad form); a_shift: A operand starts this many 128-byte rows into its tile;
// stages: operand tiles the batches rotate over (different shared-memory addresses, as a ring would give)
template <int N>
__global__ void __launch_bounds__(128) rate_kernel(long long* out, int iters, int nacc, int commit_every, int bmn,
                                                   int a_shift, int stages, int delay, int kind, int batch) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t s0 = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (s0 & 1023u)) & 1023u);
  constexpr int A_BYTES = 256 * 128, B_BYTES = N * 128, ST = A_BYTES + B_BYTES;
  uint64_t* done = reinterpret_cast<uint64_t*>(smem + 3 * ST);
  uint64_t* side = done + 1;
  uint64_t* ready = side + 1;
  uint32_t* slot = reinterpret_cast<uint32_t*>(ready + 1);
  for (int i = threadIdx.x; i < 3 * ST / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(done, 1);
    mbar_init(side, 1);
    mbar_init(ready, 1);
    mbar_arrive(ready);      // phase 0 of `ready` is complete from the start
    mbar_fence_init();
  }
  if (warp == 0) {
    tmem_alloc(slot, 512);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  long long t0 = 0, t1 = 0, t_issue = 0;
  if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(128, N, false, bmn != 0);
      float chain = 1.f;
      t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        const int st = i % stages;
        const uint32_t a_addr = smem_u32(smem + st * ST) + a_shift * 128, b_addr = smem_u32(smem + st * ST + A_BYTES);
        const uint32_t d = tmem + (i % nacc) * N;
        umma_f16_x4(d, umma_smem_desc(a_addr, 16, 1024), bmn ? umma_smem_desc(b_addr, 8192, 1024) : umma_smem_desc(b_addr, 16, 1024),
                    2u, bmn ? 128u : 2u, idesc, i >= nacc ? 1u : 0u);
        if ((i + 1) % batch != 0) continue;       // `batch` groups of four go out back to back
        if (commit_every && ((i + 1) / batch) % commit_every == 0) umma_commit(side);
        // the issuing thread's other work: kind 0 = spin on the clock, 1 = a chain of `delay` dependent FMAs (4 cycles
        // each), 2 = `delay` try_waits on a barrier whose phase has long completed (what a ring's full-barrier costs)
        if (delay && kind == 0) { const long long c0 = clock64(); while (clock64() - c0 < delay) {} }
        if (delay && kind == 1) { for (int j = 0; j < delay; ++j) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(chain) : "f"(1.0001f)); }
        if (delay && kind == 2) { for (int j = 0; j < delay; ++j) mbar_wait(ready, 0); }
      }
      umma_commit(done);
      t_issue = clock64();
      mbar_wait(done, 0);
      t1 = clock64();
      out[blockIdx.x * 2] = t1 - t0;
      out[blockIdx.x * 2 + 1] = t_issue - t0 + (chain == 12345.f ? 1 : 0);
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

template <int N>
static void run(long long* dout, int grid, int iters, int nacc, int commit_every, int bmn, int a_shift, int stages, int delay = 0, int kind = 0, int batch = 1) {
  const int smem = 3 * (256 * 128 + N * 128) + 1024 + 64;
  CK(cudaFuncSetAttribute(rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  if (nacc * N > 512) return;
  rate_kernel<N><<<grid, 128, smem>>>(dout, iters, nacc, commit_every, bmn, a_shift, stages, delay, kind, batch);
  CK(cudaDeviceSynchronize());
  rate_kernel<N><<<grid, 128, smem>>>(dout, iters, nacc, commit_every, bmn, a_shift, stages, delay, kind, batch);
  CK(cudaDeviceSynchronize());
  std::vector<long long> h(grid * 2);
  CK(cudaMemcpy(h.data(), dout, h.size() * 8, cudaMemcpyDeviceToHost));
  long long worst = 0, issue = 0;
  for (int i = 0; i < grid; ++i) { if (h[2 * i] > worst) worst = h[2 * i]; if (h[2 * i + 1] > issue) issue = h[2 * i + 1]; }
  printf("N %3d grid %3d nacc %d commit_every %d B %s a_shift %2d stages %d delay %3d kind %d batch %d: %7.1f cyc / MMA (issue side %6.1f); floor 128*N/256 = %d\n", N, grid,
         nacc, commit_every, bmn ? "MN" : "K ", a_shift, stages, delay, kind, batch, (double)worst / (4.0 * iters), (double)issue / (4.0 * iters), N / 2);
}

int main() {
  long long* dout;
  CK(cudaMalloc(&dout, 148 * 2 * 8));
  const int iters = 2000;
  for (int grid : {1, 148}) {
    for (int nacc : {1, 2}) {
      run<64>(dout, grid, iters, nacc, 0, 0, 0, 1);
      run<128>(dout, grid, iters, nacc, 0, 0, 0, 1);
      run<256>(dout, grid, iters, nacc, 0, 0, 0, 1);
    }
  }
  // commits after every batch (what the convolution's ring does), operand rotation, MN-major B, shifted A
  run<64>(dout, 148, iters, 1, 1, 0, 0, 1);
  run<64>(dout, 148, iters, 1, 1, 0, 0, 3);
  run<64>(dout, 148, iters, 2, 1, 0, 0, 3);
  run<64>(dout, 148, iters, 1, 1, 1, 0, 3);
  run<64>(dout, 148, iters, 1, 1, 0, 43, 3);
  run<64>(dout, 148, iters, 1, 1, 0, 1, 3);
  run<128>(dout, 148, iters, 1, 1, 0, 0, 3);
  run<128>(dout, 148, iters, 1, 1, 1, 0, 3);
  run<256>(dout, 148, iters, 1, 1, 0, 0, 3);
  run<256>(dout, 148, iters, 1, 1, 1, 0, 3);
  run<256>(dout, 148, iters, 1, 1, 0, 21, 3);
  // does the MMA queue hide the issuing thread's per-batch work?  `delay` cycles of it after every batch of four
  for (int delay : {50, 200, 400}) run<64>(dout, 148, iters, 1, 1, 0, 0, 3, delay, 0);
  for (int delay : {100, 300, 500}) run<256>(dout, 148, iters, 1, 1, 0, 0, 3, delay, 0);
  for (int delay : {10, 25, 50, 75, 100}) run<64>(dout, 148, iters, 1, 1, 0, 0, 3, delay, 1);
  for (int delay : {25, 50, 100, 150}) run<256>(dout, 148, iters, 1, 1, 0, 0, 3, delay, 1);
  for (int delay : {1, 2, 4}) run<64>(dout, 148, iters, 1, 1, 0, 0, 3, delay, 2);
  for (int delay : {1, 2, 4}) run<256>(dout, 148, iters, 1, 1, 0, 0, 3, delay, 2);
  // larger batches between the issuing thread's other work, with and without a commit per batch
  for (int batch : {1, 2, 4, 9, 18})
    for (int ce : {0, 1})
      for (int delay : {0, 25}) run<64>(dout, 148, 1800, 1, ce, 0, 0, 3, delay, 1, batch);
  for (int batch : {1, 2, 4, 9})
    for (int ce : {0, 1})
      for (int delay : {0, 25}) run<256>(dout, 148, 1800, 1, ce, 0, 0, 3, delay, 1, batch);
  printf("UMMA RATE PROBE DONE\n");
  return 0;
}
