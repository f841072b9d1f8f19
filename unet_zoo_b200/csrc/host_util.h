// Internal host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>

namespace uz {
int set_error(int code, const char* fmt, ...);
int cuda_error(cudaError_t e, const char* what);
int sm_count();  // SM count of the current device; -1 no device, -2 not compute capability 10.x
}  // namespace uz
