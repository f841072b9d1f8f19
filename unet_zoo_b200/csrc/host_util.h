// Internal host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>

namespace uz {
int set_error(int code, const char* fmt, ...);
int cuda_error(cudaError_t e, const char* what);
int sm_count();  // SM count of the current device; -1 no device, -2 not compute capability 10.x
// Per-device one-time state (cudaFuncSetAttribute applies to ONE device; a process may drive several): index of the
// current device clamped into [0, kMaxDevices)
constexpr int kMaxDevices = 64;
inline int device_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) return 0;
  return dev < kMaxDevices ? dev : kMaxDevices - 1;
}
}  // namespace uz
