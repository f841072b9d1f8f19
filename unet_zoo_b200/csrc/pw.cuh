// Device helpers shared by the bandwidth kernels added after pointwise.cu (gate.cu, resample.cu): 16-byte bf16x8
// accesses, warp / block reductions, launch-error macros.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/unet_zoo_b200.h"
#include "host_util.h"

namespace uz {

__device__ __forceinline__ void pw_unpack8(const uint4& v, float (&f)[8]) {
  f[0] = __uint_as_float(v.x << 16);
  f[1] = __uint_as_float(v.x & 0xFFFF0000u);
  f[2] = __uint_as_float(v.y << 16);
  f[3] = __uint_as_float(v.y & 0xFFFF0000u);
  f[4] = __uint_as_float(v.z << 16);
  f[5] = __uint_as_float(v.z & 0xFFFF0000u);
  f[6] = __uint_as_float(v.w << 16);
  f[7] = __uint_as_float(v.w & 0xFFFF0000u);
}
__device__ __forceinline__ uint32_t pw_pack2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ uint4 pw_pack8(const float (&f)[8]) {
  return make_uint4(pw_pack2(f[0], f[1]), pw_pack2(f[2], f[3]), pw_pack2(f[4], f[5]), pw_pack2(f[6], f[7]));
}
__device__ __forceinline__ uint4 pw_ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void pw_load8f(const float* p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
  f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ float pw_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float pw_sigmoid(float v) { return 1.f / (1.f + __expf(-v)); }

// launchers exported by pointwise.cu for reuse by the other translation units
cudaError_t launch_bn_bwd_finalize(const float* partial, int rows, int C, double count, const float* bn, float* dgamma,
                                   float* dbeta, int accumulate, float* coef, cudaStream_t st, int q_is_gz = 0,
                                   float* dbias = nullptr);
cudaError_t launch_reduce_rows(const float* partial, int rows, int C, float* out, int accumulate, cudaStream_t st);

}  // namespace uz

#define PW_REQUIRE_DEVICE()                \
  do {                                     \
    const int sms_ = uz_device_sm_count(); \
    if (sms_ < 0) return sms_;             \
  } while (0)
#define PW_CHECK_LAUNCH(what)                                       \
  do {                                                              \
    cudaError_t e_ = cudaGetLastError();                            \
    if (e_ != cudaSuccess) return uz::cuda_error(e_, what);         \
  } while (0)
