// Per-device cached state shared by every launcher of the library (and by the stand-alone probes under tools/).
#pragma once
#include <cuda_runtime.h>

namespace b200 {

// Per-DEVICE cached state.  Function attributes (the dynamic shared-memory opt-in) and the SM count belong to a
// device, not to the process: a host that drives cuda:1 after cuda:0 from one process needs them set / read again.
constexpr int B200_MAX_DEVICES = 64;
inline int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= B200_MAX_DEVICES) dev = 0;
  return dev;
}
inline int num_sms() {
  static int n[B200_MAX_DEVICES] = {0};
  const int dev = current_device();
  if (!n[dev]) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    n[dev] = v > 0 ? v : 148;
  }
  return n[dev];
}
// Raises a kernel's dynamic shared-memory limit to `bytes` on the current device if this call site has not yet done
// so for that device (`state` = one slot per device, zero-initialised, owned by the call site).  Not a stream
// operation: the first launch per device must happen outside stream capture (the warm-up calls do that).
template <typename Kern>
inline cudaError_t ensure_dyn_smem(Kern kern, int bytes, int (&state)[B200_MAX_DEVICES]) {
  const int dev = current_device();
  if (state[dev] >= bytes) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) state[dev] = bytes;
  return e;
}

}  // namespace b200
