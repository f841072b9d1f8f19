// Bilinear x2 upsampling with align_corners=True (reference: nn.Upsample(scale_factor=2, mode='bilinear',
// align_corners=True), nested_unet.py:33, used ten times per forward, nested_unet.py:81-93) - pure HBM bandwidth.
//
// Forward reads the low-resolution NHWC bf16 tensor and writes straight into a channel slice of the destination
// level's concat buffer (pointer + pixel stride), so neither the upsampled tensor nor torch.cat exist as separate
// passes.  Source index arithmetic follows ATen exactly: scale = (in-1)/(out-1), src = scale*dst (fp32),
// i0 = (int)src, i1 = i0 + (i0 < in-1), l1 = src - i0, l0 = 1 - l1.
// Backward is written in gather form (no atomics, deterministic): every low-resolution pixel sums the contributions
// of the few high-resolution pixels whose (i0, i1) touch it, recomputing their weights with the forward formula.
#include <cstdlib>

#include "pw.cuh"

namespace uz {

__device__ __forceinline__ void bl_src(int dst, float scale, int in, int* i0, int* i1, float* l0, float* l1) {
  const float src = scale * static_cast<float>(dst);
  const int a = static_cast<int>(src);
  *i0 = a;
  *i1 = a + (a < in - 1 ? 1 : 0);
  *l1 = src - static_cast<float>(a);
  *l0 = 1.f - *l1;
}

__global__ void bilinear2x_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, int N, int H, int W, int cv,
                                      __nv_bfloat16* __restrict__ out, int out_ld, float sy, float sx) {
  const int Ho = 2 * H, Wo = 2 * W;
  const long long total = static_cast<long long>(N) * Ho * Wo * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long r = i / cv;
    const int ox = static_cast<int>(r % Wo);
    r /= Wo;
    const int oy = static_cast<int>(r % Ho);
    const int n = static_cast<int>(r / Ho);
    int y0, y1, x0, x1;
    float ly0, ly1, lx0, lx1;
    bl_src(oy, sy, H, &y0, &y1, &ly0, &ly1);
    bl_src(ox, sx, W, &x0, &x1, &lx0, &lx1);
    const __nv_bfloat16* base = x + static_cast<long long>(n) * H * W * x_ld + g * 8;
    float a[8], b[8], c[8], d[8], o[8];
    pw_unpack8(pw_ldg16(base + (static_cast<long long>(y0) * W + x0) * x_ld), a);
    pw_unpack8(pw_ldg16(base + (static_cast<long long>(y0) * W + x1) * x_ld), b);
    pw_unpack8(pw_ldg16(base + (static_cast<long long>(y1) * W + x0) * x_ld), c);
    pw_unpack8(pw_ldg16(base + (static_cast<long long>(y1) * W + x1) * x_ld), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = ly0 * (lx0 * a[j] + lx1 * b[j]) + ly1 * (lx0 * c[j] + lx1 * d[j]);
    *reinterpret_cast<uint4*>(out + ((static_cast<long long>(n) * Ho + oy) * Wo + ox) * out_ld + g * 8) = pw_pack8(o);
  }
}

// weight with which high-res index `o` contributes to low-res index `i` along one axis (0 when it does not)
__device__ __forceinline__ float bl_weight(int o, int i, float scale, int in) {
  int i0, i1;
  float l0, l1;
  bl_src(o, scale, in, &i0, &i1, &l0, &l1);
  float w = 0.f;
  if (i0 == i) w += l0;
  if (i1 == i) w += l1;
  return w;
}

constexpr int kBlCand = 8;  // candidate high-res rows / columns per low-res index: [2i-3, 2i+4]

__global__ void bilinear2x_bwd_kernel(const __nv_bfloat16* __restrict__ dout, int dout_ld, int N, int H, int W, int cv,
                                      __nv_bfloat16* __restrict__ dx, int dx_ld, int accumulate, float sy, float sx) {
  const int Ho = 2 * H, Wo = 2 * W;
  const long long total = static_cast<long long>(N) * H * W * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long r = i / cv;
    const int xl = static_cast<int>(r % W);
    r /= W;
    const int yl = static_cast<int>(r % H);
    const int n = static_cast<int>(r / H);
    float wy[kBlCand], wx[kBlCand];
#pragma unroll
    for (int k = 0; k < kBlCand; ++k) {
      const int oy = 2 * yl - 3 + k, ox = 2 * xl - 3 + k;
      wy[k] = (oy >= 0 && oy < Ho) ? bl_weight(oy, yl, sy, H) : 0.f;
      wx[k] = (ox >= 0 && ox < Wo) ? bl_weight(ox, xl, sx, W) : 0.f;
    }
    float acc[8];
    __nv_bfloat16* dst = dx + ((static_cast<long long>(n) * H + yl) * W + xl) * dx_ld + g * 8;
    if (accumulate) {
      pw_unpack8(*reinterpret_cast<const uint4*>(dst), acc);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    }
    const __nv_bfloat16* base = dout + static_cast<long long>(n) * Ho * Wo * dout_ld + g * 8;
#pragma unroll
    for (int ky = 0; ky < kBlCand; ++ky) {
      if (wy[ky] == 0.f) continue;
      const int oy = 2 * yl - 3 + ky;
#pragma unroll
      for (int kx = 0; kx < kBlCand; ++kx) {
        if (wx[kx] == 0.f) continue;
        const int ox = 2 * xl - 3 + kx;
        float f[8];
        pw_unpack8(pw_ldg16(base + (static_cast<long long>(oy) * Wo + ox) * dout_ld), f);
        const float w = wy[ky] * wx[kx];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(w, f[j], acc[j]);
      }
    }
    *reinterpret_cast<uint4*>(dst) = pw_pack8(acc);
  }
}

// acc += w * f on two fp32 lanes with one sm_100 packed FMA (fma.rn.f32x2); used by the forward (measured 0.437 ->
// 0.411 ms at 32 x 256^2 x 64 together with the four-product form), not by the backward (0.532 -> 0.658 ms there)
__device__ __forceinline__ void bl_fma2(float& a0, float& a1, float f0, float f1, float w) {
  asm("{\n\t.reg .b64 ra, rw, rc;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rw, {%4, %4};\n\tmov.b64 rc, {%0, %1};\n\t"
      "fma.rn.f32x2 rc, ra, rw, rc;\n\tmov.b64 {%0, %1}, rc;\n\t}"
      : "+f"(a0), "+f"(a1)
      : "f"(f0), "f"(f1), "f"(w));
}

// ---- row-per-block forms (the default): the element-per-thread kernels above spend ~3 64-bit divisions per 16 bytes
// moved and run instruction-bound at ~2 TB/s (ncu / torch.profiler, transatt_unet batch 32: 1.2 ms forward, 2.35 ms
// backward for 2.5 GB).  Here a block owns one output row: the row decomposition (n, y) and the vertical taps are
// computed once per row, the sweep over (x, channel group) uses 32-bit arithmetic, and the backward visits the six
// candidate high-resolution columns [2x-2, 2x+3] unconditionally (weights are zero where a candidate does not
// contribute) while skipping non-contributing candidate rows, which is a block-uniform branch.
__global__ void __launch_bounds__(256)
bilinear2x_fwd_rows_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, int N, int H, int W, int cv,
                           __nv_bfloat16* __restrict__ out, int out_ld, float sy, float sx) {
  const int Ho = 2 * H, Wo = 2 * W;
  const int rows = N * Ho;
  const int per_row = Wo * cv;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const int n = row / Ho, oy = row - n * Ho;
    int y0, y1;
    float ly0, ly1;
    bl_src(oy, sy, H, &y0, &y1, &ly0, &ly1);
    const __nv_bfloat16* r0 = x + (static_cast<size_t>(n) * H + y0) * W * x_ld;
    const __nv_bfloat16* r1 = x + (static_cast<size_t>(n) * H + y1) * W * x_ld;
    __nv_bfloat16* orow = out + static_cast<size_t>(row) * Wo * out_ld;
#pragma unroll 2
    for (int e = threadIdx.x; e < per_row; e += 256) {
      const int ox = e / cv, g = e - ox * cv;
      int x0, x1;
      float lx0, lx1;
      bl_src(ox, sx, W, &x0, &x1, &lx0, &lx1);
      const int o0 = x0 * x_ld + g * 8, o1 = x1 * x_ld + g * 8;
      float a[8], b[8], c[8], d[8], o[8];
      pw_unpack8(pw_ldg16(r0 + o0), a);
      pw_unpack8(pw_ldg16(r0 + o1), b);
      pw_unpack8(pw_ldg16(r1 + o0), c);
      pw_unpack8(pw_ldg16(r1 + o1), d);
      const float w00 = ly0 * lx0, w01 = ly0 * lx1, w10 = ly1 * lx0, w11 = ly1 * lx1;
#pragma unroll
      for (int j = 0; j < 8; j += 2) {
        o[j] = w00 * a[j];
        o[j + 1] = w00 * a[j + 1];
        bl_fma2(o[j], o[j + 1], b[j], b[j + 1], w01);
        bl_fma2(o[j], o[j + 1], c[j], c[j + 1], w10);
        bl_fma2(o[j], o[j + 1], d[j], d[j + 1], w11);
      }
      *reinterpret_cast<uint4*>(orow + static_cast<size_t>(ox) * out_ld + g * 8) = pw_pack8(o);
    }
  }
}

constexpr int kBlWin = 6;  // candidate high-res rows / columns of low-res index i when the axis has >= 3 samples:
                           // floor(s*o) in {i-1, i} with s = (in-1)/(2in-1)  =>  o in [2i-2, 2i+3]

__global__ void __launch_bounds__(256)
bilinear2x_bwd_rows_kernel(const __nv_bfloat16* __restrict__ dout, int dout_ld, int N, int H, int W, int cv,
                           __nv_bfloat16* __restrict__ dx, int dx_ld, int accumulate, float sy, float sx) {
  const int Ho = 2 * H, Wo = 2 * W;
  const int rows = N * H;
  const int per_row = W * cv;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const int n = row / H, yl = row - n * H;
    float wy[kBlWin];
#pragma unroll
    for (int k = 0; k < kBlWin; ++k) {
      const int oy = 2 * yl - 2 + k;
      wy[k] = (oy >= 0 && oy < Ho) ? bl_weight(oy, yl, sy, H) : 0.f;
    }
    const __nv_bfloat16* img = dout + static_cast<size_t>(n) * Ho * Wo * dout_ld;
    __nv_bfloat16* drow = dx + static_cast<size_t>(row) * W * dx_ld;
    for (int e = threadIdx.x; e < per_row; e += 256) {
      const int xl = e / cv, g = e - xl * cv;
      float wx[kBlWin];
      int off[kBlWin];
#pragma unroll
      for (int k = 0; k < kBlWin; ++k) {
        const int ox = 2 * xl - 2 + k;
        const bool in = ox >= 0 && ox < Wo;
        wx[k] = in ? bl_weight(ox, xl, sx, W) : 0.f;
        off[k] = (in ? ox : 2 * xl) * dout_ld + g * 8;       // out-of-image candidates re-read a valid pixel (weight 0)
      }
      float acc[8];
      __nv_bfloat16* dst = drow + static_cast<size_t>(xl) * dx_ld + g * 8;
      if (accumulate) {
        pw_unpack8(*reinterpret_cast<const uint4*>(dst), acc);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
      }
#pragma unroll
      for (int ky = 0; ky < kBlWin; ++ky) {
        if (wy[ky] == 0.f) continue;                            // same for every thread of the block
        const __nv_bfloat16* hrow = img + static_cast<size_t>(2 * yl - 2 + ky) * Wo * dout_ld;
        uint4 v[kBlWin];
#pragma unroll
        for (int kx = 0; kx < kBlWin; ++kx) v[kx] = pw_ldg16(hrow + off[kx]);
#pragma unroll
        for (int kx = 0; kx < kBlWin; ++kx) {
          float f[8];
          pw_unpack8(v[kx], f);
          const float w = wy[ky] * wx[kx];
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = fmaf(w, f[j], acc[j]);   // (fma.rn.f32x2 here measured 24 % SLOWER)
        }
      }
      *reinterpret_cast<uint4*>(dst) = pw_pack8(acc);
    }
  }
}

static bool rs_rows_ok(int N, int H, int W, int C, int ld_lo, int ld_hi) {
  // 32-bit in-row offsets and row counts; the six-candidate window needs at least three samples per axis
  return H >= 3 && W >= 3 && static_cast<long long>(N) * 2 * H < (1ll << 30) &&
         static_cast<long long>(2 * W) * (ld_hi > ld_lo ? ld_hi : ld_lo) < (1ll << 30) && C >= 8 &&
         std::getenv("UZ_BILINEAR_GEN1") == nullptr;
}
static int rs_row_grid(long long rows) {
  const long long cap = static_cast<long long>(uz_device_sm_count()) * 8;   // 8 x 256 threads = one SM's worth of warps
  return static_cast<int>(rows < cap ? (rows < 1 ? 1 : rows) : cap);
}

static int rs_check(const char* who, const void* p, int C, int ld) {
  if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15u) || C < 8 || (C & 7) || (ld & 7) || ld < C)
    return set_error(UZ_ERR_INVALID, "%s: needs a 16-B aligned pointer, channels (=%d) and ld (=%d) multiples of 8", who, C, ld);
  return UZ_OK;
}
static int rs_grid(long long total) {
  const int sms = uz_device_sm_count();
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(sms) * 16;
  if (blocks > cap) blocks = cap;
  return static_cast<int>(blocks < 1 ? 1 : blocks);
}
static float rs_scale(int in) { return in > 0 ? static_cast<float>(in - 1) / static_cast<float>(2 * in - 1) : 0.f; }

}  // namespace uz

using namespace uz;

// nearest x2 (nn.Upsample(scale_factor=2), common_layers.py:70): out[n, oy, ox, :] = x[n, oy >> 1, ox >> 1, :]
// Only the weight gradient of the up-conv needs the upsampled tensor (it runs the 3x3 halo-box wgrad kernel on it);
// the forward and the data gradient fold the upsampling into the convolution's gather.
namespace uz {
__global__ void nearest2x_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, int N, int H, int W, int cv,
                                     __nv_bfloat16* __restrict__ out, int out_ld) {
  const int Ho = 2 * H, Wo = 2 * W;
  const long long total = static_cast<long long>(N) * Ho * Wo * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % cv);
    long long r = i / cv;
    const int ox = static_cast<int>(r % Wo);
    r /= Wo;
    const int oy = static_cast<int>(r % Ho);
    const int n = static_cast<int>(r / Ho);
    const long long src = (static_cast<long long>(n) * H + (oy >> 1)) * W + (ox >> 1);
    const long long dst = (static_cast<long long>(n) * Ho + oy) * Wo + ox;
    *reinterpret_cast<uint4*>(out + dst * out_ld + g * 8) = pw_ldg16(x + src * x_ld + g * 8);
  }
}
}  // namespace uz

namespace uz {
// row-per-block form of the copy above (same reasoning as bilinear2x_fwd_rows_kernel: no 64-bit division per 16 bytes)
__global__ void __launch_bounds__(256)
nearest2x_fwd_rows_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, int N, int H, int W, int cv,
                          __nv_bfloat16* __restrict__ out, int out_ld) {
  const int Ho = 2 * H, Wo = 2 * W;
  const int rows = N * Ho;
  const int per_row = Wo * cv;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const int n = row / Ho, oy = row - n * Ho;
    const __nv_bfloat16* src = x + (static_cast<size_t>(n) * H + (oy >> 1)) * W * x_ld;
    __nv_bfloat16* orow = out + static_cast<size_t>(row) * Wo * out_ld;
#pragma unroll 4
    for (int e = threadIdx.x; e < per_row; e += 256) {
      const int ox = e / cv, g = e - ox * cv;
      *reinterpret_cast<uint4*>(orow + static_cast<size_t>(ox) * out_ld + g * 8) =
          pw_ldg16(src + static_cast<size_t>(ox >> 1) * x_ld + g * 8);
    }
  }
}
}  // namespace uz

namespace uz {
// dst[n, y, x, :] = (y < Hs && x < Ws) ? src[n, y, x, :] : 0 over the dst extent, 16 bytes per thread
__global__ void pad_copy_kernel(const uint8_t* __restrict__ src, long long src_px, int Hs, int Ws,
                                uint8_t* __restrict__ dst, long long dst_px, int Hd, int Wd, int N, int vecs) {
  const long long total = static_cast<long long>(N) * Hd * Wd * vecs;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int v = static_cast<int>(i % vecs);
    long long p = i / vecs;
    const int x = static_cast<int>(p % Wd);
    p /= Wd;
    const int y = static_cast<int>(p % Hd);
    const long long n = p / Hd;
    uint4 val = make_uint4(0u, 0u, 0u, 0u);
    if (y < Hs && x < Ws)
      val = __ldg(reinterpret_cast<const uint4*>(src + ((n * Hs + y) * Ws + x) * src_px) + v);
    reinterpret_cast<uint4*>(dst + ((n * Hd + y) * Wd + x) * dst_px)[v] = val;
  }
}
}  // namespace uz

// F.pad (bottom / right, zeros) or crop of an NHWC tensor between two pixel grids: the reference centre-pads the
// upsampled tensor to the skip tensor's size when H or W is odd (common_layers.py:110-113, :173-178: pad = [0, dx, 0, dy]
// for dx, dy in {0, 1}); its backward is the crop.  Element-type agnostic: `row_bytes` bytes per pixel (multiple of
// 16) are moved, pixel strides are given in bytes.
extern "C" int uz_pad_copy(const void* src, long long src_pixel_bytes, int Hs, int Ws, void* dst,
                           long long dst_pixel_bytes, int Hd, int Wd, int N, int row_bytes, void* stream) {
  PW_REQUIRE_DEVICE();
  if (!src || !dst || N < 1 || Hs < 1 || Ws < 1 || Hd < 1 || Wd < 1)
    return set_error(UZ_ERR_INVALID, "pad_copy: bad extents");
  if (row_bytes < 16 || (row_bytes & 15) || (src_pixel_bytes & 15) || (dst_pixel_bytes & 15) ||
      src_pixel_bytes < row_bytes || dst_pixel_bytes < row_bytes ||
      (reinterpret_cast<uintptr_t>(src) & 15u) || (reinterpret_cast<uintptr_t>(dst) & 15u))
    return set_error(UZ_ERR_INVALID, "pad_copy: rows / strides / pointers must be multiples of 16 bytes");
  const int vecs = row_bytes / 16;
  const long long total = static_cast<long long>(N) * Hd * Wd * vecs;
  pad_copy_kernel<<<rs_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(src), src_pixel_bytes, Hs, Ws, static_cast<uint8_t*>(dst), dst_pixel_bytes, Hd, Wd, N,
      vecs);
  PW_CHECK_LAUNCH("pad_copy launch");
  return UZ_OK;
}

extern "C" int uz_nearest2x_fwd(const void* x, int x_ld, int N, int H, int W, int C, void* out, int out_ld,
                                void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = rs_check("nearest2x_fwd(x)", x, C, x_ld);
  if (rc) return rc;
  if ((rc = rs_check("nearest2x_fwd(out)", out, C, out_ld))) return rc;
  if (N < 1 || H < 1 || W < 1) return set_error(UZ_ERR_INVALID, "nearest2x_fwd: bad size N=%d H=%d W=%d", N, H, W);
  if (static_cast<long long>(N) * 2 * H < (1ll << 30) && static_cast<long long>(2 * W) * (C / 8) < (1ll << 30) &&
      std::getenv("UZ_BILINEAR_GEN1") == nullptr) {
    nearest2x_fwd_rows_kernel<<<rs_row_grid(static_cast<long long>(N) * 2 * H), 256, 0,
                                static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), x_ld, N, H, W, C / 8, static_cast<__nv_bfloat16*>(out), out_ld);
    PW_CHECK_LAUNCH("nearest2x_fwd launch");
    return UZ_OK;
  }
  const long long total = static_cast<long long>(N) * 4 * H * W * (C / 8);
  nearest2x_fwd_kernel<<<rs_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), x_ld, N, H, W, C / 8, static_cast<__nv_bfloat16*>(out), out_ld);
  PW_CHECK_LAUNCH("nearest2x_fwd launch");
  return UZ_OK;
}

extern "C" int uz_bilinear2x_fwd(const void* x, int x_ld, int N, int H, int W, int C, void* out, int out_ld,
                                 void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = rs_check("bilinear2x_fwd(x)", x, C, x_ld);
  if (rc) return rc;
  if ((rc = rs_check("bilinear2x_fwd(out)", out, C, out_ld))) return rc;
  if (N < 1 || H < 1 || W < 1) return set_error(UZ_ERR_INVALID, "bilinear2x_fwd: bad size N=%d H=%d W=%d", N, H, W);
  if (rs_rows_ok(N, H, W, C, x_ld, out_ld)) {
    bilinear2x_fwd_rows_kernel<<<rs_row_grid(static_cast<long long>(N) * 2 * H), 256, 0,
                                 static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), x_ld, N, H, W, C / 8, static_cast<__nv_bfloat16*>(out), out_ld,
        rs_scale(H), rs_scale(W));
    PW_CHECK_LAUNCH("bilinear2x_fwd launch");
    return UZ_OK;
  }
  const long long total = static_cast<long long>(N) * 4 * H * W * (C / 8);
  bilinear2x_fwd_kernel<<<rs_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), x_ld, N, H, W, C / 8, static_cast<__nv_bfloat16*>(out), out_ld, rs_scale(H),
      rs_scale(W));
  PW_CHECK_LAUNCH("bilinear2x_fwd launch");
  return UZ_OK;
}

extern "C" int uz_bilinear2x_bwd(const void* dout, int dout_ld, int N, int H, int W, int C, void* dx, int dx_ld,
                                 int accumulate, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = rs_check("bilinear2x_bwd(dout)", dout, C, dout_ld);
  if (rc) return rc;
  if ((rc = rs_check("bilinear2x_bwd(dx)", dx, C, dx_ld))) return rc;
  if (N < 1 || H < 1 || W < 1) return set_error(UZ_ERR_INVALID, "bilinear2x_bwd: bad size N=%d H=%d W=%d", N, H, W);
  if (rs_rows_ok(N, H, W, C, dx_ld, dout_ld)) {
    bilinear2x_bwd_rows_kernel<<<rs_row_grid(static_cast<long long>(N) * H), 256, 0,
                                 static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(dout), dout_ld, N, H, W, C / 8, static_cast<__nv_bfloat16*>(dx), dx_ld,
        accumulate, rs_scale(H), rs_scale(W));
    PW_CHECK_LAUNCH("bilinear2x_bwd launch");
    return UZ_OK;
  }
  const long long total = static_cast<long long>(N) * H * W * (C / 8);
  bilinear2x_bwd_kernel<<<rs_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dout), dout_ld, N, H, W, C / 8, static_cast<__nv_bfloat16*>(dx), dx_ld, accumulate,
      rs_scale(H), rs_scale(W));
  PW_CHECK_LAUNCH("bilinear2x_bwd launch");
  return UZ_OK;
}
