// Residual add of the RSU blocks of U^2-Net (reference unet_zoo/models/u2net.py:78,127,171,210,
// 237: `return hx1d + hxin`), NHWC bf16, 8 channels (16 bytes) per thread:
//   uz_residual_add_fwd   y = a + b, optionally with the MaxPool2d(2, stride 2) of the sum as a second output (the
//                         stage pooling that follows every encoder RSU, u2net.py:222-232; sizes are even here, so
//                         ceil_mode changes nothing)
//   uz_residual_add_bwd   g = dy + route(dpooled) (the pooled gradient goes to the FIRST maximum of the stored y in
//                         row-major window order, as torch's max_pool2d backward does); da (+)= g, db (+)= g
#include <cstdint>

#include "pw.cuh"

namespace uz {
namespace {

__device__ __forceinline__ void ra_ld8(const __nv_bfloat16* p, float (&f)[8]) { pw_unpack8(pw_ldg16(p), f); }
__device__ __forceinline__ void ra_st8(__nv_bfloat16* p, const float (&f)[8]) { *reinterpret_cast<uint4*>(p) = pw_pack8(f); }
__device__ __forceinline__ float ra_bf16r(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// one thread per (2x2 quad, 8-channel group); partial quads of odd sizes get their sum but no pooling window
__global__ void residual_add_fwd_kernel(const __nv_bfloat16* __restrict__ a, int a_ld, const __nv_bfloat16* __restrict__ b,
                                        int b_ld, int N, int H, int W, int C, __nv_bfloat16* __restrict__ y, int y_ld,
                                        __nv_bfloat16* __restrict__ pooled, int pooled_ld) {
  const int cv = C / 8;
  const int Hq = (H + 1) / 2, Wq = (W + 1) / 2, Hp = H / 2, Wp = W / 2;
  const long long total = static_cast<long long>(N) * Hq * Wq * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long q = i / cv;
    const int qx = static_cast<int>(q % Wq);
    q /= Wq;
    const int qy = static_cast<int>(q % Hq);
    const long long n = q / Hq;
    float m[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int yy = 2 * qy + (k >> 1), xx = 2 * qx + (k & 1);
      if (yy >= H || xx >= W) continue;
      const long long p = (n * H + yy) * W + xx;
      float f[8], r[8];
      ra_ld8(a + p * a_ld + g * 8, f);
      ra_ld8(b + p * b_ld + g * 8, r);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = ra_bf16r(f[j] + r[j]);
      ra_st8(y + p * y_ld + g * 8, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) m[j] = (k == 0 || f[j] > m[j]) ? f[j] : m[j];
    }
    if (pooled != nullptr && qy < Hp && qx < Wp) ra_st8(pooled + ((n * Hp + qy) * Wp + qx) * pooled_ld + g * 8, m);
  }
}

__global__ void residual_add_bwd_kernel(const __nv_bfloat16* __restrict__ y, int y_ld, int N, int H, int W, int C,
                                        const __nv_bfloat16* __restrict__ dy, int dy_ld,
                                        const __nv_bfloat16* __restrict__ dpooled, int dpooled_ld,
                                        __nv_bfloat16* __restrict__ da, int da_ld, int acc_a,
                                        __nv_bfloat16* __restrict__ db, int db_ld, int acc_b) {
  const int cv = C / 8;
  const int Hq = (H + 1) / 2, Wq = (W + 1) / 2, Hp = H / 2, Wp = W / 2;
  const long long total = static_cast<long long>(N) * Hq * Wq * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long q = i / cv;
    const int qx = static_cast<int>(q % Wq);
    q /= Wq;
    const int qy = static_cast<int>(q % Hq);
    const long long n = q / Hq;
    const bool full = dpooled != nullptr && qy < Hp && qx < Wp;
    float dp[8];
    int am[8];
    if (full) {
      float s[4][8];
      ra_ld8(dpooled + ((n * Hp + qy) * Wp + qx) * dpooled_ld + g * 8, dp);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        ra_ld8(y + ((n * H + 2 * qy + (k >> 1)) * W + 2 * qx + (k & 1)) * y_ld + g * 8, s[k]);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        am[j] = 0;
        float mv = s[0][j];
#pragma unroll
        for (int k = 1; k < 4; ++k)
          if (s[k][j] > mv) {
            mv = s[k][j];
            am[j] = k;
          }
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int yy = 2 * qy + (k >> 1), xx = 2 * qx + (k & 1);
      if (yy >= H || xx >= W) continue;
      const long long p = (n * H + yy) * W + xx;
      float o[8];
      if (dy) {
        ra_ld8(dy + p * dy_ld + g * 8, o);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = 0.f;
      }
      if (full) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (am[j] == k) o[j] += dp[j];
      }
      __nv_bfloat16* dst[2] = {da + p * da_ld + g * 8, db + p * db_ld + g * 8};
      const int acc[2] = {acc_a, acc_b};
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = o[j];
        if (acc[t]) {
          float old[8];
          ra_ld8(dst[t], old);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] += old[j];
        }
        ra_st8(dst[t], v);
      }
    }
  }
}

int ra_grid(long long total) {
  const int sms = uz_device_sm_count();
  long long g = (total + 255) / 256, cap = static_cast<long long>(sms > 0 ? sms : 148) * 16;
  return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
}
int ra_chk(const char* what, const void* p, int C, int ld) {
  if (p == nullptr || C < 8 || (C & 7) || (ld & 7) || ld < C || (reinterpret_cast<uintptr_t>(p) & 15u))
    return set_error(UZ_ERR_INVALID, "%s: channels=%d ld=%d must be multiples of 8, pointer 16-byte aligned", what, C, ld);
  return UZ_OK;
}

}  // namespace
}  // namespace uz

using namespace uz;

extern "C" int uz_residual_add_fwd(const void* a, int a_ld, const void* b, int b_ld, int N, int H, int W, int C, void* y,
                                   int y_ld, void* pooled, int pooled_ld, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = ra_chk("residual_add_fwd(a)", a, C, a_ld);
  if (rc || (rc = ra_chk("residual_add_fwd(b)", b, C, b_ld)) || (rc = ra_chk("residual_add_fwd(y)", y, C, y_ld))) return rc;
  if (pooled && (rc = ra_chk("residual_add_fwd(pooled)", pooled, C, pooled_ld))) return rc;
  if (N < 1 || H < 1 || W < 1 || (pooled && (H < 2 || W < 2))) return set_error(UZ_ERR_INVALID, "residual_add_fwd: bad extents");
  const long long total = static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
  residual_add_fwd_kernel<<<ra_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(a), a_ld, static_cast<const __nv_bfloat16*>(b), b_ld, N, H, W, C,
      static_cast<__nv_bfloat16*>(y), y_ld, static_cast<__nv_bfloat16*>(pooled), pooled_ld);
  PW_CHECK_LAUNCH("residual_add_fwd launch");
  return UZ_OK;
}

extern "C" int uz_residual_add_bwd(const void* y, int y_ld, int N, int H, int W, int C, const void* dy, int dy_ld,
                                   const void* dpooled, int dpooled_ld, void* da, int da_ld, int acc_a, void* db, int db_ld,
                                   int acc_b, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = ra_chk("residual_add_bwd(da)", da, C, da_ld);
  if (rc || (rc = ra_chk("residual_add_bwd(db)", db, C, db_ld))) return rc;
  if (!dy && !dpooled) return set_error(UZ_ERR_INVALID, "residual_add_bwd: no incoming gradient");
  if (dy && (rc = ra_chk("residual_add_bwd(dy)", dy, C, dy_ld))) return rc;
  if (dpooled && ((rc = ra_chk("residual_add_bwd(dpooled)", dpooled, C, dpooled_ld)) || (rc = ra_chk("residual_add_bwd(y)", y, C, y_ld))))
    return rc;
  if (N < 1 || H < 1 || W < 1) return set_error(UZ_ERR_INVALID, "residual_add_bwd: bad extents");
  if (da == db) return set_error(UZ_ERR_INVALID, "residual_add_bwd: da and db must be different buffers");
  const long long total = static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
  residual_add_bwd_kernel<<<ra_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(y), y_ld, N, H, W, C, static_cast<const __nv_bfloat16*>(dy), dy_ld,
      static_cast<const __nv_bfloat16*>(dpooled), dpooled_ld, static_cast<__nv_bfloat16*>(da), da_ld, acc_a,
      static_cast<__nv_bfloat16*>(db), db_ld, acc_b);
  PW_CHECK_LAUNCH("residual_add_bwd launch");
  return UZ_OK;
}
