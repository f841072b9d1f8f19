// Residual add of the RSU blocks of U^2-Net (reference unet_zoo/models/u2net.py:78,127,171,210,
// 237: `return hx1d + hxin`), NHWC bf16, 8 channels (16 bytes) per thread:
//   uz_residual_add_fwd   y = a + b, optionally with the MaxPool2d(2, stride 2) of the sum as a second output (the
//                         stage pooling that follows every encoder RSU, u2net.py:222-232; sizes are even here, so
//                         ceil_mode changes nothing)
//   uz_residual_add_bwd   g = dy + route(dpooled) (the pooled gradient goes to the FIRST maximum of the stored y in
//                         row-major window order, as torch's max_pool2d backward does); da (+)= g, db (+)= g
//   uz_side_head_fwd/_bwd the output stage: outconv over the six upsampled side maps + the side maps themselves as fp32
//                         NCHW logits (u2net.py:273-295) in one pass over the slot buffer
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
                                        __nv_bfloat16* __restrict__ pooled, int pooled_ld, int relu) {
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
    uint4 va[4], vb[4];      // all eight loads of the quad in flight before any arithmetic
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int yy = 2 * qy + (k >> 1), xx = 2 * qx + (k & 1);
      va[k] = vb[k] = make_uint4(0u, 0u, 0u, 0u);
      if (yy >= H || xx >= W) continue;
      const long long p = (n * H + yy) * W + xx;
      va[k] = pw_ldg16(a + p * a_ld + g * 8);
      vb[k] = pw_ldg16(b + p * b_ld + g * 8);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int yy = 2 * qy + (k >> 1), xx = 2 * qx + (k & 1);
      if (yy >= H || xx >= W) continue;
      const long long p = (n * H + yy) * W + xx;
      float f[8], r[8];
      pw_unpack8(va[k], f);
      pw_unpack8(vb[k], r);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = ra_bf16r(relu ? fmaxf(f[j] + r[j], 0.f) : f[j] + r[j]);
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
                                        __nv_bfloat16* __restrict__ db, int db_ld, int acc_b, int relu) {
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
    // every load of the quad is issued before any arithmetic (up to 17 16-byte words in flight per thread): the
    // straight-line form of this loop ran four dependent load -> store chains and reached 0.57 of the HBM roof
    long long pix[4];
    bool ok[4];
    uint4 ry[4], rd[4], roa[4], rob[4], rp = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int yy = 2 * qy + (k >> 1), xx = 2 * qx + (k & 1);
      ok[k] = yy < H && xx < W;
      pix[k] = (n * H + yy) * W + xx;
      ry[k] = rd[k] = roa[k] = rob[k] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (full || relu) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (ok[k]) ry[k] = pw_ldg16(y + pix[k] * y_ld + g * 8);
    }
    if (full) rp = pw_ldg16(dpooled + ((n * Hp + qy) * Wp + qx) * dpooled_ld + g * 8);
    if (dy) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (ok[k]) rd[k] = pw_ldg16(dy + pix[k] * dy_ld + g * 8);
    }
    if (acc_a) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (ok[k]) roa[k] = pw_ldg16(da + pix[k] * da_ld + g * 8);
    }
    if (acc_b) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (ok[k]) rob[k] = pw_ldg16(db + pix[k] * db_ld + g * 8);
    }
    float dp[8], s[4][8];
    int am[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) pw_unpack8(ry[k], s[k]);
    pw_unpack8(rp, dp);
    if (full) {
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
      if (!ok[k]) continue;
      float o[8], old[8];
      pw_unpack8(rd[k], o);
      if (full) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (am[j] == k) o[j] += dp[j];
      }
      if (relu) {      // ReLU after the add: the gradient passes where the stored sum is positive
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (!(s[k][j] > 0.f)) o[j] = 0.f;
      }
      float v[8];
      pw_unpack8(roa[k], old);      // zeros unless accumulating
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = o[j] + old[j];
      ra_st8(da + pix[k] * da_ld + g * 8, v);
      pw_unpack8(rob[k], old);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = o[j] + old[j];
      ra_st8(db + pix[k] * db_ld + g * 8, v);
    }
  }
}

// Without pooling the quad structure buys nothing: one thread per (pixel, 8-channel group), consecutive threads on
// consecutive 16-byte words of a pixel row (every residual add of MultiResUnet; the decoder RSU blocks of U^2-Net).
__global__ void residual_add_lin_fwd_kernel(const __nv_bfloat16* __restrict__ a, int a_ld, const __nv_bfloat16* __restrict__ b,
                                            int b_ld, long long npix, int C, __nv_bfloat16* __restrict__ y, int y_ld, int relu) {
  const int cv = C / 8;
  const long long total = npix * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long p = i / cv;
    const int g = static_cast<int>(i - p * cv);
    float f[8], r[8];
    ra_ld8(a + p * a_ld + g * 8, f);
    ra_ld8(b + p * b_ld + g * 8, r);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = relu ? fmaxf(f[j] + r[j], 0.f) : f[j] + r[j];
    ra_st8(y + p * y_ld + g * 8, f);
  }
}

__global__ void residual_add_lin_bwd_kernel(const __nv_bfloat16* __restrict__ y, int y_ld, long long npix, int C,
                                            const __nv_bfloat16* __restrict__ dy, int dy_ld, __nv_bfloat16* __restrict__ da,
                                            int da_ld, int acc_a, __nv_bfloat16* __restrict__ db, int db_ld, int acc_b,
                                            int relu) {
  const int cv = C / 8;
  const long long total = npix * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long p = i / cv;
    const int g = static_cast<int>(i - p * cv);
    float o[8];
    ra_ld8(dy + p * dy_ld + g * 8, o);
    if (relu) {
      float s[8];
      ra_ld8(y + p * y_ld + g * 8, s);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (!(s[j] > 0.f)) o[j] = 0.f;
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

// ------------------------------------------------------------------------------------------------ logits extraction
// The first K (<= 8) channels of an NHWC bf16 map as fp32 NCHW planes: MultiResUnet's conv_final (Conv2d 1x1 ->
// BatchNorm2d(affine=False), multiresunet.py:200-201,238) runs as a conv-BN pair whose single class is padded to 8
// channels.  One thread per pixel: one 16-byte load, K coalesced fp32 stores.
__global__ void nhwc_to_nchw_f32_kernel(const __nv_bfloat16* __restrict__ x, int ld, long long npix, long long plane, int K,
                                        float* __restrict__ out) {
  for (long long p = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; p < npix;
       p += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long n = p / plane, hw = p - n * plane;
    float f[8];
    ra_ld8(x + p * ld, f);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (k < K) out[(n * K + k) * plane + hw] = f[k];
  }
}

// ------------------------------------------------------------------------------------------------ fused output stage
// U2NET's last lines (u2net.py:273-295): d0 = outconv(cat(d1..d6)), returned together with d1..d6.  The upsampled side
// maps live in 8-channel slots of one NHWC bf16 buffer (slot i = channels [8 i, 8 i + K), the rest of a slot is zero).
// One thread per pixel: nsides 16-byte loads, (1 + nsides) * K fp32 stores into NCHW planes (coalesced over pixels).
constexpr int kSideMax = 8;    // slots
constexpr int kSideK = 4;      // classes per map handled by these kernels

template <int K>
__global__ void __launch_bounds__(256) side_head_fwd_kernel(const __nv_bfloat16* __restrict__ fuse, int ld, long long npix,
                                                            long long plane, int nsides,
                                                            const float* __restrict__ w /* [K][nsides*K] */,
                                                            const float* __restrict__ b, float* __restrict__ logits) {
  __shared__ float ws[K * kSideMax * K + K];
  const int nw = K * nsides * K;
  for (int i = threadIdx.x; i < nw + K; i += blockDim.x) ws[i] = i < nw ? w[i] : (b ? b[i - nw] : 0.f);
  __syncthreads();
  const int rows = (1 + nsides) * K;
  for (long long p = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; p < npix;
       p += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long n = p / plane, hw = p - n * plane;
    float* out = logits + n * rows * plane + hw;
    float mainv[K];
#pragma unroll
    for (int j = 0; j < K; ++j) mainv[j] = ws[nw + j];
#pragma unroll
    for (int i = 0; i < kSideMax; ++i) {
      if (i < nsides) {
        float f[8];
        ra_ld8(fuse + p * ld + i * 8, f);
#pragma unroll
        for (int k = 0; k < K; ++k) {
          out[(static_cast<long long>(1 + i) * K + k) * plane] = f[k];
#pragma unroll
          for (int j = 0; j < K; ++j) mainv[j] = fmaf(ws[j * nsides * K + i * K + k], f[k], mainv[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < K; ++j) out[j * plane] = mainv[j];
  }
}

// dfuse[p][slot i][k] = dl[side i][k] + sum_j w[j][i K + k] dl[main j]   (channels K..7 of a slot: 0)
// partial[block][(j * nsides + i) * K + k] = sum_p dl[main j] * side[i][k]  ;  partial[block][K*nsides*K + j] = sum_p dl[main j]
template <int K>
__global__ void __launch_bounds__(256) side_head_bwd_kernel(const __nv_bfloat16* __restrict__ fuse, int ld, long long npix,
                                                            long long plane, int nsides, const float* __restrict__ w,
                                                            const float* __restrict__ dl, __nv_bfloat16* __restrict__ dfuse,
                                                            int dld, float* __restrict__ partial) {
  __shared__ float ws[K * kSideMax * K];
  __shared__ float red[8][K * kSideMax * K + K];
  const int nw = K * nsides * K;
  for (int i = threadIdx.x; i < nw; i += blockDim.x) ws[i] = w[i];
  __syncthreads();
  const int rows = (1 + nsides) * K;
  float acc[K][kSideMax][K], accb[K];
#pragma unroll
  for (int j = 0; j < K; ++j) {
    accb[j] = 0.f;
#pragma unroll
    for (int i = 0; i < kSideMax; ++i)
#pragma unroll
      for (int k = 0; k < K; ++k) acc[j][i][k] = 0.f;
  }
  for (long long p = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; p < npix;
       p += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long n = p / plane, hw = p - n * plane;
    const float* g = dl + n * rows * plane + hw;
    float gm[K];
#pragma unroll
    for (int j = 0; j < K; ++j) {
      gm[j] = __ldg(g + j * plane);
      accb[j] += gm[j];
    }
#pragma unroll
    for (int i = 0; i < kSideMax; ++i) {
      if (i < nsides) {
        float f[8], o[8];
        ra_ld8(fuse + p * ld + i * 8, f);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          float v = __ldg(g + (static_cast<long long>(1 + i) * K + k) * plane);
#pragma unroll
          for (int j = 0; j < K; ++j) {
            v = fmaf(ws[j * nsides * K + i * K + k], gm[j], v);
            acc[j][i][k] = fmaf(gm[j], f[k], acc[j][i][k]);
          }
          o[k] = v;
        }
        ra_st8(dfuse + p * dld + i * 8, o);
      }
    }
  }
  // block reduction in a fixed order: warp shuffles, then the eight warps through shared memory
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < K; ++j) {
    const float vb = pw_warp_sum(accb[j]);
    if (lane == 0) red[warp][K * kSideMax * K + j] = vb;
#pragma unroll
    for (int i = 0; i < kSideMax; ++i)
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const float v = pw_warp_sum(acc[j][i][k]);
        if (lane == 0) red[warp][(j * kSideMax + i) * K + k] = v;
      }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < nw + K; t += blockDim.x) {
    int src;
    if (t < nw) {
      const int j = t / (nsides * K), r = t - j * nsides * K;
      src = (j * kSideMax + r / K) * K + r % K;
    } else {
      src = K * kSideMax * K + (t - nw);
    }
    float v = 0.f;
#pragma unroll
    for (int wq = 0; wq < 8; ++wq) v += red[wq][src];
    partial[static_cast<size_t>(blockIdx.x) * (nw + K) + t] = v;
  }
}

// out[c] = sum_r partial[r][c0 + c] for c < n (rows of `pitch` floats), fixed order, one warp per column
__global__ void side_reduce_kernel(const float* __restrict__ partial, int rows, int pitch, int c0, int n,
                                   float* __restrict__ out) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= n) return;
  double s = 0.0;
  for (int r = lane; r < rows; r += 32) s += static_cast<double>(partial[static_cast<size_t>(r) * pitch + c0 + c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) out[c] = static_cast<float>(s);
}
cudaError_t launch_reduce_rows_strided(const float* partial, int rows, int pitch, int c0, int n, float* out,
                                       cudaStream_t st) {
  side_reduce_kernel<<<(n + 3) / 4, 128, 0, st>>>(partial, rows, pitch, c0, n, out);
  return cudaGetLastError();
}

int side_rows() {
  const int sms = uz_device_sm_count();
  return sms < 0 ? sms : sms * 4;
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
                                   int y_ld, void* pooled, int pooled_ld, int relu, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = ra_chk("residual_add_fwd(a)", a, C, a_ld);
  if (rc || (rc = ra_chk("residual_add_fwd(b)", b, C, b_ld)) || (rc = ra_chk("residual_add_fwd(y)", y, C, y_ld))) return rc;
  if (pooled && (rc = ra_chk("residual_add_fwd(pooled)", pooled, C, pooled_ld))) return rc;
  if (N < 1 || H < 1 || W < 1 || (pooled && (H < 2 || W < 2))) return set_error(UZ_ERR_INVALID, "residual_add_fwd: bad extents");
  if (!pooled) {
    const long long npix = static_cast<long long>(N) * H * W;
    residual_add_lin_fwd_kernel<<<ra_grid(npix * (C / 8)), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(a), a_ld, static_cast<const __nv_bfloat16*>(b), b_ld, npix, C,
        static_cast<__nv_bfloat16*>(y), y_ld, relu);
    PW_CHECK_LAUNCH("residual_add_fwd launch");
    return UZ_OK;
  }
  const long long total = static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
  residual_add_fwd_kernel<<<ra_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(a), a_ld, static_cast<const __nv_bfloat16*>(b), b_ld, N, H, W, C,
      static_cast<__nv_bfloat16*>(y), y_ld, static_cast<__nv_bfloat16*>(pooled), pooled_ld, relu);
  PW_CHECK_LAUNCH("residual_add_fwd launch");
  return UZ_OK;
}

extern "C" int uz_residual_add_bwd(const void* y, int y_ld, int N, int H, int W, int C, const void* dy, int dy_ld,
                                   const void* dpooled, int dpooled_ld, void* da, int da_ld, int acc_a, void* db, int db_ld,
                                   int acc_b, int relu, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = ra_chk("residual_add_bwd(da)", da, C, da_ld);
  if (rc || (rc = ra_chk("residual_add_bwd(db)", db, C, db_ld))) return rc;
  if (!dy && !dpooled) return set_error(UZ_ERR_INVALID, "residual_add_bwd: no incoming gradient");
  if (dy && (rc = ra_chk("residual_add_bwd(dy)", dy, C, dy_ld))) return rc;
  if (dpooled && (rc = ra_chk("residual_add_bwd(dpooled)", dpooled, C, dpooled_ld))) return rc;
  if ((dpooled || relu) && (rc = ra_chk("residual_add_bwd(y)", y, C, y_ld))) return rc;
  if (N < 1 || H < 1 || W < 1) return set_error(UZ_ERR_INVALID, "residual_add_bwd: bad extents");
  if (da == db) return set_error(UZ_ERR_INVALID, "residual_add_bwd: da and db must be different buffers");
  if (!dpooled) {
    const long long npix = static_cast<long long>(N) * H * W;
    residual_add_lin_bwd_kernel<<<ra_grid(npix * (C / 8)), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(y), y_ld, npix, C, static_cast<const __nv_bfloat16*>(dy), dy_ld,
        static_cast<__nv_bfloat16*>(da), da_ld, acc_a, static_cast<__nv_bfloat16*>(db), db_ld, acc_b, relu);
    PW_CHECK_LAUNCH("residual_add_bwd launch");
    return UZ_OK;
  }
  const long long total = static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
  residual_add_bwd_kernel<<<ra_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(y), y_ld, N, H, W, C, static_cast<const __nv_bfloat16*>(dy), dy_ld,
      static_cast<const __nv_bfloat16*>(dpooled), dpooled_ld, static_cast<__nv_bfloat16*>(da), da_ld, acc_a,
      static_cast<__nv_bfloat16*>(db), db_ld, acc_b, relu);
  PW_CHECK_LAUNCH("residual_add_bwd launch");
  return UZ_OK;
}

extern "C" int uz_nhwc_to_nchw_f32(const void* x, int ld, int N, int H, int W, int K, float* out, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = ra_chk("nhwc_to_nchw_f32(x)", x, 8, ld);
  if (rc) return rc;
  if (N < 1 || H < 1 || W < 1 || K < 1 || K > 8 || out == nullptr)
    return set_error(UZ_ERR_INVALID, "nhwc_to_nchw_f32: bad extents (N=%d H=%d W=%d K=%d, K <= 8)", N, H, W, K);
  const long long plane = static_cast<long long>(H) * W, npix = plane * N;
  nhwc_to_nchw_f32_kernel<<<ra_grid(npix), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), ld, npix, plane, K, out);
  PW_CHECK_LAUNCH("nhwc_to_nchw_f32 launch");
  return UZ_OK;
}

extern "C" size_t uz_side_head_bwd_workspace_bytes(int nsides, int K) {
  const int rows = side_rows();
  return rows < 0 ? 0 : sizeof(float) * static_cast<size_t>(rows) * (K * nsides * K + K);
}

static int side_check(const char* who, const void* fuse, int ld, int N, int H, int W, int nsides, int K) {
  if (nsides < 1 || nsides > kSideMax || K < 1 || K > kSideK)
    return set_error(UZ_ERR_UNSUPPORTED, "%s: needs 1..%d side maps of 1..%d classes (got %d x %d)", who, kSideMax, kSideK,
                     nsides, K);
  if (N < 1 || H < 1 || W < 1) return set_error(UZ_ERR_INVALID, "%s: bad extents", who);
  return ra_chk(who, fuse, 8 * nsides, ld);
}

extern "C" int uz_side_head_fwd(const void* fuse, int ld, int N, int H, int W, int nsides, int K, const float* w,
                                const float* b, float* logits, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = side_check("side_head_fwd", fuse, ld, N, H, W, nsides, K);
  if (rc) return rc;
  if (!w || !logits) return set_error(UZ_ERR_INVALID, "side_head_fwd: null weights / output");
  const long long plane = static_cast<long long>(H) * W, npix = plane * N;
  const long long need = (npix + 255) / 256;
  const int grid = static_cast<int>(need < side_rows() * 2 ? need : side_rows() * 2);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* fp = static_cast<const __nv_bfloat16*>(fuse);
  switch (K) {
    case 1: side_head_fwd_kernel<1><<<grid, 256, 0, st>>>(fp, ld, npix, plane, nsides, w, b, logits); break;
    case 2: side_head_fwd_kernel<2><<<grid, 256, 0, st>>>(fp, ld, npix, plane, nsides, w, b, logits); break;
    case 3: side_head_fwd_kernel<3><<<grid, 256, 0, st>>>(fp, ld, npix, plane, nsides, w, b, logits); break;
    default: side_head_fwd_kernel<4><<<grid, 256, 0, st>>>(fp, ld, npix, plane, nsides, w, b, logits); break;
  }
  PW_CHECK_LAUNCH("side_head_fwd launch");
  return UZ_OK;
}

extern "C" int uz_side_head_bwd(const void* fuse, int ld, const float* dlogits, int N, int H, int W, int nsides, int K,
                                const float* w, void* dfuse, int dfuse_ld, float* dw, float* db, void* workspace,
                                size_t workspace_bytes, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = side_check("side_head_bwd", fuse, ld, N, H, W, nsides, K);
  if (rc || (rc = ra_chk("side_head_bwd(dfuse)", dfuse, 8 * nsides, dfuse_ld))) return rc;
  if (!w || !dlogits || !dw || !db) return set_error(UZ_ERR_INVALID, "side_head_bwd: null argument");
  if (!workspace || workspace_bytes < uz_side_head_bwd_workspace_bytes(nsides, K))
    return set_error(UZ_ERR_WORKSPACE, "side_head_bwd: workspace too small");
  const long long plane = static_cast<long long>(H) * W, npix = plane * N;
  const long long need = (npix + 255) / 256;
  const int grid = static_cast<int>(need < side_rows() ? need : side_rows());
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* partial = static_cast<float*>(workspace);
  const int nw = K * nsides * K;
  const __nv_bfloat16* fp = static_cast<const __nv_bfloat16*>(fuse);
  __nv_bfloat16* dfp = static_cast<__nv_bfloat16*>(dfuse);
  switch (K) {
    case 1: side_head_bwd_kernel<1><<<grid, 256, 0, st>>>(fp, ld, npix, plane, nsides, w, dlogits, dfp, dfuse_ld, partial); break;
    case 2: side_head_bwd_kernel<2><<<grid, 256, 0, st>>>(fp, ld, npix, plane, nsides, w, dlogits, dfp, dfuse_ld, partial); break;
    case 3: side_head_bwd_kernel<3><<<grid, 256, 0, st>>>(fp, ld, npix, plane, nsides, w, dlogits, dfp, dfuse_ld, partial); break;
    default: side_head_bwd_kernel<4><<<grid, 256, 0, st>>>(fp, ld, npix, plane, nsides, w, dlogits, dfp, dfuse_ld, partial); break;
  }
  PW_CHECK_LAUNCH("side_head_bwd launch");
  // partial rows are [nw weights | K biases]: reduce the two column ranges into dw and db
  cudaError_t e = launch_reduce_rows_strided(partial, grid, nw + K, 0, nw, dw, st);
  if (e == cudaSuccess) e = launch_reduce_rows_strided(partial, grid, nw + K, nw, K, db, st);
  if (e != cudaSuccess) return cuda_error(e, "side_head_bwd reduce launch");
  return UZ_OK;
}
