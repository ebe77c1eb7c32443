// Standalone GPU probe (no torch): what does a kernel boundary cost inside a replayed CUDA graph, and how much of it does
// programmatic dependent launch (PDL: cudaLaunchAttributeProgrammaticStreamSerialization + griddepcontrol) remove?
// A chain of `n` dependent kernels is captured from one stream and replayed; each kernel is `blocks` CTAs of 256
// threads that spin for `work` clock cycles.  Reported: replay time / n minus the kernel's own duration.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/graph_gap_probe.cu -o tools/graph_gap_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x)                                                                            \
  do {                                                                                   \
    cudaError_t e_ = (x);                                                                \
    if (e_ != cudaSuccess) {                                                             \
      printf("CUDA error %s at %s:%d: %s\n", #x, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(2);                                                                           \
    }                                                                                    \
  } while (0)

template <bool PDL>
__global__ void __launch_bounds__(256) spin_kernel(float* buf, int work) {
  if (PDL) {
    asm volatile("griddepcontrol.launch_dependents;");   // the next kernel's CTAs may be scheduled as ours retire
    asm volatile("griddepcontrol.wait;" ::: "memory");   // ... and we touch memory only after our predecessor is complete
  }
  const long long t0 = clock64();
  float v = buf[blockIdx.x * 256 + threadIdx.x];
  while (clock64() - t0 < work) v = v * 1.0001f + 1.0f;
  buf[blockIdx.x * 256 + threadIdx.x] = v;
}

static float run(int n, int blocks, int work, bool pdl, float* buf, cudaStream_t st) {
  cudaGraph_t graph;
  cudaGraphExec_t exec;
  CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
  for (int i = 0; i < n; ++i) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    if (pdl) CK(cudaLaunchKernelEx(&cfg, spin_kernel<true>, buf, work));
    else CK(cudaLaunchKernelEx(&cfg, spin_kernel<false>, buf, work));
  }
  CK(cudaStreamEndCapture(st, &graph));
  CK(cudaGraphInstantiate(&exec, graph, 0));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  for (int i = 0; i < 3; ++i) CK(cudaGraphLaunch(exec, st));
  CK(cudaEventRecord(a, st));
  for (int i = 0; i < 10; ++i) CK(cudaGraphLaunch(exec, st));
  CK(cudaEventRecord(b, st));
  CK(cudaStreamSynchronize(st));
  float ms;
  CK(cudaEventElapsedTime(&ms, a, b));
  CK(cudaGraphExecDestroy(exec));
  CK(cudaGraphDestroy(graph));
  return ms * 1000.f / (10.f * n);
}

int main() {
  float* buf;
  CK(cudaMalloc(&buf, 4096 * 256 * 4));
  CK(cudaMemset(buf, 0, 4096 * 256 * 4));
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  const int n = 500;
  for (int blocks : {1, 148, 1184}) {
    for (int work : {0, 10000, 40000}) {       // ~0, ~5 us, ~20 us of spinning at ~1.9 GHz
      const float plain = run(n, blocks, work, false, buf, st);
      const float pdl = run(n, blocks, work, true, buf, st);
      printf("blocks %4d work %5d cycles: %7.2f us per kernel in a graph chain, %7.2f us with PDL  (saves %.2f us)\n", blocks,
             work, plain, pdl, plain - pdl);
    }
  }
  printf("GRAPH GAP PROBE DONE\n");
  return 0;
}
