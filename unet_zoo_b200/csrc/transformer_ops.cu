// Bandwidth kernels around the attention blocks of unet_transformer (reference
// unet_zoo/models/unet_transformer.py:68-229), NHWC bf16, 8 channels (16 bytes) per thread:
//   uz_addpe_fwd / _bwd           x + positional encoding (a fixed [H, W, C] fp32 table, :68-113), optionally followed by
//                                 MaxPool2d(2) of the sum (Sconv_process[0], :153) without storing the sum
//   uz_adaptive_avgpool_fwd / _bwd  F.adaptive_avg_pool2d to the attention resolution (:201-203), ATen window arithmetic
//   uz_bilinear_resize_fwd / _bwd   F.interpolate(size=..., mode="bilinear", align_corners=True) (:222), ATen index
//                                 arithmetic; the backward is the exact adjoint in gather form (deterministic, no atomics)
#include <cstdint>
#include <cstdlib>

#include "pw.cuh"

namespace uz {
namespace {

__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float (&f)[8]) { *reinterpret_cast<uint4*>(p) = pw_pack8(f); }
__device__ __forceinline__ float bf16r(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
// eight consecutive fp32 (32-byte aligned: the positional-encoding table is [H][W][C] with C a multiple of 8) as two
// 16-byte loads - eight scalar loads per 16 bytes of activation made these kernels LSU-issue bound
__device__ __forceinline__ void ld8f(const float* p, float (&e)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  e[0] = a.x; e[1] = a.y; e[2] = a.z; e[3] = a.w;
  e[4] = b.x; e[5] = b.y; e[6] = b.z; e[7] = b.w;
}

// y[n, h, w, :] = bf16(x + pe[h, w, :]); pooled = 2x2 max of y (floor), either output may be null
__global__ void addpe_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, const float* __restrict__ pe, int N, int H,
                                 int W, int C, __nv_bfloat16* __restrict__ y, int y_ld, __nv_bfloat16* __restrict__ pooled,
                                 int pooled_ld) {
  const int cv = C / 8;
  if (pooled == nullptr) {
    const long long total = static_cast<long long>(N) * H * W * cv;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
      const int g = static_cast<int>(i % cv);
      const long long p = i / cv;
      const long long hw = p % (static_cast<long long>(H) * W);
      float f[8];
      ld8(x + p * x_ld + g * 8, f);
      float ev[8];
      ld8f(pe + hw * C + g * 8, ev);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] += ev[j];
      st8(y + p * y_ld + g * 8, f);
    }
    return;
  }
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
      float f[8];
      ld8(x + p * x_ld + g * 8, f);
      float ev[8];
      ld8f(pe + (static_cast<long long>(yy) * W + xx) * C + g * 8, ev);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = bf16r(f[j] + ev[j]);
      if (y) st8(y + p * y_ld + g * 8, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) m[j] = (k == 0 || f[j] > m[j]) ? f[j] : m[j];
    }
    if (qy < Hp && qx < Wp) st8(pooled + ((n * Hp + qy) * Wp + qx) * pooled_ld + g * 8, m);
  }
}

// dx (+)= dy + route(dpooled): the pooled gradient goes to the FIRST maximum of bf16(x + pe) in row-major window order
__global__ void addpe_bwd_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, const float* __restrict__ pe, int N, int H,
                                 int W, int C, const __nv_bfloat16* __restrict__ dy, int dy_ld,
                                 const __nv_bfloat16* __restrict__ dpooled, int dpooled_ld, __nv_bfloat16* __restrict__ dx,
                                 int dx_ld, int accumulate) {
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
    float s[4][8], dp[8];
    int am[8];
    if (full) {
      ld8(dpooled + ((n * Hp + qy) * Wp + qx) * dpooled_ld + g * 8, dp);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const long long p = (n * H + 2 * qy + (k >> 1)) * W + 2 * qx + (k & 1);
        ld8(x + p * x_ld + g * 8, s[k]);
        float ev[8];
        ld8f(pe + (static_cast<long long>(2 * qy + (k >> 1)) * W + 2 * qx + (k & 1)) * C + g * 8, ev);
#pragma unroll
        for (int j = 0; j < 8; ++j) s[k][j] = bf16r(s[k][j] + ev[j]);
      }
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
        ld8(dy + p * dy_ld + g * 8, o);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = 0.f;
      }
      if (full) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (am[j] == k) o[j] += dp[j];
      }
      __nv_bfloat16* d = dx + p * dx_ld + g * 8;
      if (accumulate) {
        float old[8];
        ld8(d, old);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] += old[j];
      }
      st8(d, o);
    }
  }
}

// ATen adaptive pooling windows: [floor(o * I / O), ceil((o + 1) * I / O))
__device__ __forceinline__ int ap_start(int o, int I, int O) { return static_cast<int>((static_cast<long long>(o) * I) / O); }
__device__ __forceinline__ int ap_end(int o, int I, int O) {
  return static_cast<int>((static_cast<long long>(o + 1) * I + O - 1) / O);
}

__global__ void adaptive_avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, int N, int H, int W, int C,
                                            __nv_bfloat16* __restrict__ out, int out_ld, int OH, int OW) {
  const int cv = C / 8;
  const long long total = static_cast<long long>(N) * OH * OW * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long q = i / cv;
    const int ox = static_cast<int>(q % OW);
    q /= OW;
    const int oy = static_cast<int>(q % OH);
    const long long n = q / OH;
    const int y0 = ap_start(oy, H, OH), y1 = ap_end(oy, H, OH), x0 = ap_start(ox, W, OW), x1 = ap_end(ox, W, OW);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int yy = y0; yy < y1; ++yy)
      for (int xx = x0; xx < x1; ++xx) {
        float f[8];
        ld8(x + ((n * H + yy) * W + xx) * x_ld + g * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    const float inv = 1.f / static_cast<float>((y1 - y0) * (x1 - x0));
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= inv;
    st8(out + ((n * OH + oy) * OW + ox) * out_ld + g * 8, acc);
  }
}

// dx[iy, ix] (+)= sum over the output cells whose window contains (iy, ix) of dout / window area
__global__ void adaptive_avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dout, int dout_ld, int N, int H, int W, int C,
                                            int OH, int OW, __nv_bfloat16* __restrict__ dx, int dx_ld, int accumulate) {
  const int cv = C / 8;
  const long long total = static_cast<long long>(N) * H * W * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long q = i / cv;
    const int ix = static_cast<int>(q % W);
    q /= W;
    const int iy = static_cast<int>(q % H);
    const long long n = q / H;
    int oy_lo = static_cast<int>((static_cast<long long>(iy) * OH) / H) - 1;
    int oy_hi = static_cast<int>((static_cast<long long>(iy + 1) * OH + H - 1) / H);
    int ox_lo = static_cast<int>((static_cast<long long>(ix) * OW) / W) - 1;
    int ox_hi = static_cast<int>((static_cast<long long>(ix + 1) * OW + W - 1) / W);
    oy_lo = oy_lo < 0 ? 0 : oy_lo;
    ox_lo = ox_lo < 0 ? 0 : ox_lo;
    oy_hi = oy_hi > OH - 1 ? OH - 1 : oy_hi;
    ox_hi = ox_hi > OW - 1 ? OW - 1 : ox_hi;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int oy = oy_lo; oy <= oy_hi; ++oy) {
      const int y0 = ap_start(oy, H, OH), y1 = ap_end(oy, H, OH);
      if (iy < y0 || iy >= y1) continue;
      for (int ox = ox_lo; ox <= ox_hi; ++ox) {
        const int x0 = ap_start(ox, W, OW), x1 = ap_end(ox, W, OW);
        if (ix < x0 || ix >= x1) continue;
        float f[8];
        ld8(dout + ((n * OH + oy) * OW + ox) * dout_ld + g * 8, f);
        const float inv = 1.f / static_cast<float>((y1 - y0) * (x1 - x0));
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(f[j], inv, acc[j]);
      }
    }
    __nv_bfloat16* d = dx + ((n * H + iy) * W + ix) * dx_ld + g * 8;
    if (accumulate) {
      float old[8];
      ld8(d, old);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += old[j];
    }
    st8(d, acc);
  }
}

// ATen upsample_bilinear2d source index in fp32: align_corners = True: src = dst * (I - 1) / (O - 1) (off = 0);
// align_corners = False: src = max((dst + 0.5) * I / O - 0.5, 0) (off = 0.5 * scale - 0.5); i0 = floor, i1 = min(i0 + 1, I - 1)
__device__ __forceinline__ void bl_src(int dst, float scale, float off, int I, int* i0, int* i1, float* l1) {
  const float src = fmaxf(fmaf(scale, static_cast<float>(dst), off), 0.f);
  int a = static_cast<int>(src);
  a = a > I - 1 ? I - 1 : a;
  *i0 = a;
  *i1 = a + (a < I - 1 ? 1 : 0);
  *l1 = src - static_cast<float>(a);
}

__global__ void bilinear_resize_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, int N, int H, int W, int C,
                                           __nv_bfloat16* __restrict__ out, int out_ld, int OH, int OW, float sy,
                                           float sx, float offy, float offx) {
  const int cv = C / 8;
  const long long total = static_cast<long long>(N) * OH * OW * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long q = i / cv;
    const int ox = static_cast<int>(q % OW);
    q /= OW;
    const int oy = static_cast<int>(q % OH);
    const long long n = q / OH;
    int y0, y1, x0, x1;
    float ly, lx;
    bl_src(oy, sy, offy, H, &y0, &y1, &ly);
    bl_src(ox, sx, offx, W, &x0, &x1, &lx);
    float a[8], b[8], c[8], d[8], o[8];
    ld8(x + ((n * H + y0) * W + x0) * x_ld + g * 8, a);
    ld8(x + ((n * H + y0) * W + x1) * x_ld + g * 8, b);
    ld8(x + ((n * H + y1) * W + x0) * x_ld + g * 8, c);
    ld8(x + ((n * H + y1) * W + x1) * x_ld + g * 8, d);
    const float hy = 1.f - ly, hx = 1.f - lx;
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = hy * (hx * a[j] + lx * b[j]) + ly * (hx * c[j] + lx * d[j]);
    st8(out + ((n * OH + oy) * OW + ox) * out_ld + g * 8, o);
  }
}

// exact adjoint in gather form: input pixel (iy, ix) collects every output whose (i0, i1) pair contains it
__global__ void bilinear_resize_bwd_kernel(const __nv_bfloat16* __restrict__ dout, int dout_ld, int N, int H, int W, int C,
                                           int OH, int OW, float sy, float sx, float offy, float offx,
                                           __nv_bfloat16* __restrict__ dx, int dx_ld, int accumulate) {
  const int cv = C / 8;
  const long long total = static_cast<long long>(N) * H * W * cv;
  const float ry = sy > 0.f ? 1.f / sy : 0.f, rx = sx > 0.f ? 1.f / sx : 0.f;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long q = i / cv;
    const int ix = static_cast<int>(q % W);
    q /= W;
    const int iy = static_cast<int>(q % H);
    const long long n = q / H;
    // candidate outputs: src in (i - 1, i + 1)  ->  dst in ((i - 1) / scale, (i + 1) / scale), widened by one
    int oy_lo = sy > 0.f ? static_cast<int>(floorf((iy - 1 - offy) * ry)) - 1 : 0;
    int oy_hi = sy > 0.f ? static_cast<int>(ceilf((iy + 1 - offy) * ry)) + 1 : OH - 1;
    int ox_lo = sx > 0.f ? static_cast<int>(floorf((ix - 1 - offx) * rx)) - 1 : 0;
    int ox_hi = sx > 0.f ? static_cast<int>(ceilf((ix + 1 - offx) * rx)) + 1 : OW - 1;
    oy_lo = oy_lo < 0 ? 0 : oy_lo;
    ox_lo = ox_lo < 0 ? 0 : ox_lo;
    oy_hi = oy_hi > OH - 1 ? OH - 1 : oy_hi;
    ox_hi = ox_hi > OW - 1 ? OW - 1 : ox_hi;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int oy = oy_lo; oy <= oy_hi; ++oy) {
      int y0, y1;
      float ly;
      bl_src(oy, sy, offy, H, &y0, &y1, &ly);
      const float wy = (y0 == iy ? 1.f - ly : 0.f) + (y1 == iy ? ly : 0.f);
      if (wy == 0.f) continue;
      for (int ox = ox_lo; ox <= ox_hi; ++ox) {
        int x0, x1;
        float lx;
        bl_src(ox, sx, offx, W, &x0, &x1, &lx);
        const float wx = (x0 == ix ? 1.f - lx : 0.f) + (x1 == ix ? lx : 0.f);
        if (wx == 0.f) continue;
        float f[8];
        ld8(dout + ((n * OH + oy) * OW + ox) * dout_ld + g * 8, f);
        const float w = wy * wx;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(f[j], w, acc[j]);
      }
    }
    __nv_bfloat16* d = dx + ((n * H + iy) * W + ix) * dx_ld + g * 8;
    if (accumulate) {
      float old[8];
      ld8(d, old);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += old[j];
    }
    st8(d, acc);
  }
}

// Exact 2x magnification with align_corners = False (every _upsample_like inside the RSU blocks and between the stages of
// U^2-Net, u2net.py:19-23): out[2 i] = 0.25 in[i - 1] + 0.75 in[i], out[2 i + 1] = 0.75 in[i] + 0.25 in[i + 1] per axis,
// indices clamped at the borders - the same fp32 weights ATen computes (src = i -+ 0.25).  One thread per INPUT pixel and
// 8-channel group: 3 x 3 neighbourhood in, 2 x 2 quad out; 32-bit index arithmetic.
__global__ void bilinear2x_half_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, int N, int H, int W, int C,
                                           __nv_bfloat16* __restrict__ out, int out_ld) {
  const int cv = C / 8;
  const unsigned total = static_cast<unsigned>(N) * H * W * cv;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int g = i % cv;
    unsigned q = i / cv;
    const int ix = q % W;
    q /= W;
    const int iy = q % H;
    const int n = q / H;
    const int ys[3] = {iy > 0 ? iy - 1 : 0, iy, iy < H - 1 ? iy + 1 : H - 1};
    const int xs[3] = {ix > 0 ? ix - 1 : 0, ix, ix < W - 1 ? ix + 1 : W - 1};
    float v[3][3][8];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int b = 0; b < 3; ++b)
        ld8(x + (static_cast<long long>(n * H + ys[a]) * W + xs[b]) * x_ld + g * 8, v[a][b]);
    // along x first: h[a][px] = 0.25 * (left | right neighbour) + 0.75 * centre; then the same along y
    float h[3][2][8];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        h[a][0][j] = 0.25f * v[a][0][j] + 0.75f * v[a][1][j];
        h[a][1][j] = 0.75f * v[a][1][j] + 0.25f * v[a][2][j];
      }
#pragma unroll
    for (int py = 0; py < 2; ++py)
#pragma unroll
      for (int px = 0; px < 2; ++px) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
          o[j] = py == 0 ? 0.25f * h[0][px][j] + 0.75f * h[1][px][j] : 0.75f * h[1][px][j] + 0.25f * h[2][px][j];
        st8(out + (static_cast<long long>(n * 2 * H + 2 * iy + py) * (2 * W) + 2 * ix + px) * out_ld + g * 8, o);
      }
  }
}

// exact adjoint: din[i] = wa d[2 i - 1] + wb d[2 i] + wc d[2 i + 1] + wd d[2 i + 2] per axis with (0.25, 0.75, 0.75, 0.25),
// border pixels take the weight of the clamped neighbour as well (wb = 1 at i = 0, wc = 1 at i = I - 1)
__global__ void bilinear2x_half_bwd_kernel(const __nv_bfloat16* __restrict__ dout, int dout_ld, int N, int H, int W, int C,
                                           __nv_bfloat16* __restrict__ dx, int dx_ld, int accumulate) {
  const int cv = C / 8;
  const unsigned total = static_cast<unsigned>(N) * H * W * cv;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int g = i % cv;
    unsigned q = i / cv;
    const int ix = q % W;
    q /= W;
    const int iy = q % H;
    const int n = q / H;
    const float wy[4] = {iy > 0 ? 0.25f : 0.f, iy > 0 ? 0.75f : 1.f, iy < H - 1 ? 0.75f : 1.f, iy < H - 1 ? 0.25f : 0.f};
    const float wx[4] = {ix > 0 ? 0.25f : 0.f, ix > 0 ? 0.75f : 1.f, ix < W - 1 ? 0.75f : 1.f, ix < W - 1 ? 0.25f : 0.f};
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      if (wy[a] == 0.f) continue;
      const int oy = 2 * iy - 1 + a;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        if (wx[b] == 0.f) continue;
        const int ox = 2 * ix - 1 + b;
        float f[8];
        ld8(dout + (static_cast<long long>(n * 2 * H + oy) * (2 * W) + ox) * dout_ld + g * 8, f);
        const float w = wy[a] * wx[b];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(f[j], w, acc[j]);
      }
    }
    __nv_bfloat16* d = dx + (static_cast<long long>(n * H + iy) * W + ix) * dx_ld + g * 8;
    if (accumulate) {
      float old[8];
      ld8(d, old);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += old[j];
    }
    st8(d, acc);
  }
}

// Separable form of the same adjoint for large magnifications (the side maps of U^2-Net grow 4-32x, u2net.py:276-284: in
// the one-pass gather every input pixel would walk (2 s + 2)^2 candidate outputs serially with a few thousand threads in
// flight).  Bilinear weights factor as wy * wx, so: pass X reduces every output row onto the W input columns (fp32
// intermediate [N][OH][W][C]), pass Y reduces the OH rows of the intermediate onto the H input rows.
__global__ void bilinear_bwd_x_kernel(const __nv_bfloat16* __restrict__ dout, int dout_ld, int N, int W, int C, int OH,
                                      int OW, float sx, float offx, float* __restrict__ tmp) {
  const int cv = C / 8;
  const long long total = static_cast<long long>(N) * OH * W * cv;
  const float rx = sx > 0.f ? 1.f / sx : 0.f;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long q = i / cv;
    const int ix = static_cast<int>(q % W);
    q /= W;                                   // q = n * OH + oy
    int ox_lo = sx > 0.f ? static_cast<int>(floorf((ix - 1 - offx) * rx)) - 1 : 0;
    int ox_hi = sx > 0.f ? static_cast<int>(ceilf((ix + 1 - offx) * rx)) + 1 : OW - 1;
    ox_lo = ox_lo < 0 ? 0 : ox_lo;
    ox_hi = ox_hi > OW - 1 ? OW - 1 : ox_hi;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int ox = ox_lo; ox <= ox_hi; ++ox) {
      int x0, x1;
      float lx;
      bl_src(ox, sx, offx, W, &x0, &x1, &lx);
      const float wx = (x0 == ix ? 1.f - lx : 0.f) + (x1 == ix ? lx : 0.f);
      if (wx == 0.f) continue;
      float f[8];
      ld8(dout + (q * OW + ox) * dout_ld + g * 8, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = fmaf(f[j], wx, acc[j]);
    }
    float4* t = reinterpret_cast<float4*>(tmp + (q * W + ix) * C + g * 8);
    t[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    t[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}

__global__ void bilinear_bwd_y_kernel(const float* __restrict__ tmp, int N, int H, int W, int C, int OH, float sy, float offy,
                                      __nv_bfloat16* __restrict__ dx, int dx_ld, int accumulate) {
  const int cv = C / 8;
  const long long total = static_cast<long long>(N) * H * W * cv;
  const float ry = sy > 0.f ? 1.f / sy : 0.f;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long q = i / cv;
    const int ix = static_cast<int>(q % W);
    q /= W;
    const int iy = static_cast<int>(q % H);
    const long long n = q / H;
    int oy_lo = sy > 0.f ? static_cast<int>(floorf((iy - 1 - offy) * ry)) - 1 : 0;
    int oy_hi = sy > 0.f ? static_cast<int>(ceilf((iy + 1 - offy) * ry)) + 1 : OH - 1;
    oy_lo = oy_lo < 0 ? 0 : oy_lo;
    oy_hi = oy_hi > OH - 1 ? OH - 1 : oy_hi;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int oy = oy_lo; oy <= oy_hi; ++oy) {
      int y0, y1;
      float ly;
      bl_src(oy, sy, offy, H, &y0, &y1, &ly);
      const float wy = (y0 == iy ? 1.f - ly : 0.f) + (y1 == iy ? ly : 0.f);
      if (wy == 0.f) continue;
      float f[8];
      pw_load8f(tmp + ((n * OH + oy) * W + ix) * C + g * 8, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = fmaf(f[j], wy, acc[j]);
    }
    __nv_bfloat16* d = dx + ((n * H + iy) * W + ix) * dx_ld + g * 8;
    if (accumulate) {
      float old[8];
      ld8(d, old);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += old[j];
    }
    st8(d, acc);
  }
}

int ew_grid(long long total) {
  const int sms = uz_device_sm_count();
  long long g = (total + 255) / 256, cap = static_cast<long long>(sms > 0 ? sms : 148) * 16;
  return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
}
int chk(const char* what, const void* p, int C, int ld) {
  if (p == nullptr || C < 8 || (C & 7) || (ld & 7) || ld < C || (reinterpret_cast<uintptr_t>(p) & 15u))
    return set_error(UZ_ERR_INVALID, "%s: channels=%d ld=%d must be multiples of 8, pointer 16-byte aligned", what, C, ld);
  return UZ_OK;
}

}  // namespace
}  // namespace uz

using namespace uz;

extern "C" int uz_addpe_fwd(const void* x, int x_ld, const float* pe, int N, int H, int W, int C, void* y, int y_ld,
                            void* pooled, int pooled_ld, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = chk("addpe_fwd(x)", x, C, x_ld);
  if (rc) return rc;
  if (!pe || (!y && !pooled) || N < 1 || H < 1 || W < 1) return set_error(UZ_ERR_INVALID, "addpe_fwd: bad arguments");
  if (reinterpret_cast<uintptr_t>(pe) & 15u) return set_error(UZ_ERR_INVALID, "addpe_fwd: pe must be 16-byte aligned");
  if (y && (rc = chk("addpe_fwd(y)", y, C, y_ld))) return rc;
  if (pooled && (rc = chk("addpe_fwd(pooled)", pooled, C, pooled_ld))) return rc;
  if (pooled && (H < 2 || W < 2)) return set_error(UZ_ERR_INVALID, "addpe_fwd: pooling needs H, W >= 2");
  const long long total = pooled ? static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8)
                                 : static_cast<long long>(N) * H * W * (C / 8);
  addpe_fwd_kernel<<<ew_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), x_ld, pe, N, H, W, C, static_cast<__nv_bfloat16*>(y), y_ld,
      static_cast<__nv_bfloat16*>(pooled), pooled_ld);
  PW_CHECK_LAUNCH("addpe_fwd launch");
  return UZ_OK;
}

extern "C" int uz_addpe_bwd(const void* x, int x_ld, const float* pe, int N, int H, int W, int C, const void* dy,
                            int dy_ld, const void* dpooled, int dpooled_ld, void* dx, int dx_ld, int accumulate,
                            void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = chk("addpe_bwd(dx)", dx, C, dx_ld);
  if (rc) return rc;
  if (!dy && !dpooled) return set_error(UZ_ERR_INVALID, "addpe_bwd: no incoming gradient");
  if (dy && (rc = chk("addpe_bwd(dy)", dy, C, dy_ld))) return rc;
  if (dpooled && ((rc = chk("addpe_bwd(dpooled)", dpooled, C, dpooled_ld)) || (rc = chk("addpe_bwd(x)", x, C, x_ld)))) return rc;
  if (dpooled && !pe) return set_error(UZ_ERR_INVALID, "addpe_bwd: the pooled path needs x and pe");
  if (pe && (reinterpret_cast<uintptr_t>(pe) & 15u)) return set_error(UZ_ERR_INVALID, "addpe_bwd: pe must be 16-byte aligned");
  const long long total = static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
  addpe_bwd_kernel<<<ew_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), x_ld, pe, N, H, W, C, static_cast<const __nv_bfloat16*>(dy), dy_ld,
      static_cast<const __nv_bfloat16*>(dpooled), dpooled_ld, static_cast<__nv_bfloat16*>(dx), dx_ld, accumulate);
  PW_CHECK_LAUNCH("addpe_bwd launch");
  return UZ_OK;
}

extern "C" int uz_adaptive_avgpool_fwd(const void* x, int x_ld, int N, int H, int W, int C, void* out, int out_ld,
                                       int OH, int OW, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = chk("adaptive_avgpool_fwd(x)", x, C, x_ld);
  if (rc || (rc = chk("adaptive_avgpool_fwd(out)", out, C, out_ld))) return rc;
  if (N < 1 || H < 1 || W < 1 || OH < 1 || OW < 1) return set_error(UZ_ERR_INVALID, "adaptive_avgpool_fwd: bad extents");
  adaptive_avgpool_fwd_kernel<<<ew_grid(static_cast<long long>(N) * OH * OW * (C / 8)), 256, 0,
                                static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(x), x_ld, N, H, W,
                                                                     C, static_cast<__nv_bfloat16*>(out), out_ld, OH, OW);
  PW_CHECK_LAUNCH("adaptive_avgpool_fwd launch");
  return UZ_OK;
}

extern "C" int uz_adaptive_avgpool_bwd(const void* dout, int dout_ld, int N, int H, int W, int C, int OH, int OW,
                                       void* dx, int dx_ld, int accumulate, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = chk("adaptive_avgpool_bwd(dout)", dout, C, dout_ld);
  if (rc || (rc = chk("adaptive_avgpool_bwd(dx)", dx, C, dx_ld))) return rc;
  if (N < 1 || H < 1 || W < 1 || OH < 1 || OW < 1) return set_error(UZ_ERR_INVALID, "adaptive_avgpool_bwd: bad extents");
  adaptive_avgpool_bwd_kernel<<<ew_grid(static_cast<long long>(N) * H * W * (C / 8)), 256, 0,
                                static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(dout), dout_ld, N, H,
                                                                     W, C, OH, OW, static_cast<__nv_bfloat16*>(dx), dx_ld,
                                                                     accumulate);
  PW_CHECK_LAUNCH("adaptive_avgpool_bwd launch");
  return UZ_OK;
}

// ATen area_pixel_compute_scale: align_corners -> (I - 1) / (O - 1) (0 for O = 1); else I / O (no scale_factor given)
static float bl_scale(int in, int out, int align) {
  if (align) return out > 1 ? static_cast<float>(in - 1) / static_cast<float>(out - 1) : 0.f;
  return static_cast<float>(in) / static_cast<float>(out);
}
static float bl_off(float scale, int align) { return align ? 0.f : 0.5f * scale - 0.5f; }
// the specialised 2x kernels: align_corners = False, exact doubling, index space below 2^31 (32-bit arithmetic)
static bool bl_exact2x(int N, int H, int W, int C, int OH, int OW, int align) {
  return !align && OH == 2 * H && OW == 2 * W && H >= 2 && W >= 2 &&
         static_cast<long long>(N) * OH * OW * (C / 8) < 0x7fffffffLL && getenv("UZ_BILINEAR_GENERIC") == nullptr;
}

extern "C" int uz_bilinear_resize_fwd(const void* x, int x_ld, int N, int H, int W, int C, void* out, int out_ld, int OH,
                                      int OW, int align_corners, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = chk("bilinear_resize_fwd(x)", x, C, x_ld);
  if (rc || (rc = chk("bilinear_resize_fwd(out)", out, C, out_ld))) return rc;
  if (N < 1 || H < 1 || W < 1 || OH < 1 || OW < 1) return set_error(UZ_ERR_INVALID, "bilinear_resize_fwd: bad extents");
  if (bl_exact2x(N, H, W, C, OH, OW, align_corners)) {
    bilinear2x_half_fwd_kernel<<<ew_grid(static_cast<long long>(N) * H * W * (C / 8)), 256, 0,
                                 static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(x), x_ld, N, H, W, C,
                                                                      static_cast<__nv_bfloat16*>(out), out_ld);
    PW_CHECK_LAUNCH("bilinear_resize_fwd (2x) launch");
    return UZ_OK;
  }
  bilinear_resize_fwd_kernel<<<ew_grid(static_cast<long long>(N) * OH * OW * (C / 8)), 256, 0,
                               static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(x), x_ld, N, H, W, C,
                                                                    static_cast<__nv_bfloat16*>(out), out_ld, OH, OW,
                                                                    bl_scale(H, OH, align_corners), bl_scale(W, OW, align_corners),
                                                                    bl_off(bl_scale(H, OH, align_corners), align_corners),
                                                                    bl_off(bl_scale(W, OW, align_corners), align_corners));
  PW_CHECK_LAUNCH("bilinear_resize_fwd launch");
  return UZ_OK;
}

// separable two-pass backward from 4x magnification on (needs the fp32 intermediate [N][OH][W][C])
static bool bl_separable(int H, int W, int OH, int OW) { return OH >= 4 * H && OW >= 4 * W; }

extern "C" size_t uz_bilinear_resize_bwd_workspace_bytes(int N, int H, int W, int C, int OH, int OW) {
  return bl_separable(H, W, OH, OW) ? sizeof(float) * static_cast<size_t>(N) * OH * W * C : 0;
}

extern "C" int uz_bilinear_resize_bwd(const void* dout, int dout_ld, int N, int H, int W, int C, int OH, int OW, void* dx,
                                      int dx_ld, int accumulate, int align_corners, void* workspace, size_t workspace_bytes,
                                      void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = chk("bilinear_resize_bwd(dout)", dout, C, dout_ld);
  if (rc || (rc = chk("bilinear_resize_bwd(dx)", dx, C, dx_ld))) return rc;
  if (N < 1 || H < 1 || W < 1 || OH < 1 || OW < 1) return set_error(UZ_ERR_INVALID, "bilinear_resize_bwd: bad extents");
  if (bl_exact2x(N, H, W, C, OH, OW, align_corners)) {
    bilinear2x_half_bwd_kernel<<<ew_grid(static_cast<long long>(N) * H * W * (C / 8)), 256, 0,
                                 static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(dout), dout_ld, N, H,
                                                                      W, C, static_cast<__nv_bfloat16*>(dx), dx_ld,
                                                                      accumulate);
    PW_CHECK_LAUNCH("bilinear_resize_bwd (2x) launch");
    return UZ_OK;
  }
  if (bl_separable(H, W, OH, OW)) {
    const size_t need = uz_bilinear_resize_bwd_workspace_bytes(N, H, W, C, OH, OW);
    if (!workspace || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 15u))
      return set_error(UZ_ERR_WORKSPACE, "bilinear_resize_bwd: workspace %zu < required %zu bytes (16-byte aligned)",
                       workspace_bytes, need);
    const float sy = bl_scale(H, OH, align_corners), sx = bl_scale(W, OW, align_corners);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* tmp = static_cast<float*>(workspace);
    bilinear_bwd_x_kernel<<<ew_grid(static_cast<long long>(N) * OH * W * (C / 8)), 256, 0, st>>>(
        static_cast<const __nv_bfloat16*>(dout), dout_ld, N, W, C, OH, OW, sx, bl_off(sx, align_corners), tmp);
    PW_CHECK_LAUNCH("bilinear_resize_bwd (x pass) launch");
    bilinear_bwd_y_kernel<<<ew_grid(static_cast<long long>(N) * H * W * (C / 8)), 256, 0, st>>>(
        tmp, N, H, W, C, OH, sy, bl_off(sy, align_corners), static_cast<__nv_bfloat16*>(dx), dx_ld, accumulate);
    PW_CHECK_LAUNCH("bilinear_resize_bwd (y pass) launch");
    return UZ_OK;
  }
  bilinear_resize_bwd_kernel<<<ew_grid(static_cast<long long>(N) * H * W * (C / 8)), 256, 0,
                               static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(dout), dout_ld, N, H,
                                                                    W, C, OH, OW, bl_scale(H, OH, align_corners),
                                                                    bl_scale(W, OW, align_corners),
                                                                    bl_off(bl_scale(H, OH, align_corners), align_corners),
                                                                    bl_off(bl_scale(W, OW, align_corners), align_corners),
                                                                    static_cast<__nv_bfloat16*>(dx), dx_ld, accumulate);
  PW_CHECK_LAUNCH("bilinear_resize_bwd launch");
  return UZ_OK;
}
