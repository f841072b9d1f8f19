// Input pipeline tail fused into the stem operand.
//
// The reference's dataset turns every decoded image into a normalised fp32 CHW tensor on the CPU
// (transforms.ToTensor + transforms.Normalize, unet_zoo/data/datasets.py:40-44), the loop then ships 12 bytes per
// pixel to the device (`img.float().to(device)`, training_loop.py:109), and the mask becomes `(m > 0.5).float()`
// (datasets.py:59) - 4 more bytes per pixel.  Here the DECODED bytes travel (3 + 1 bytes per pixel) and the device does
// the rest:
//   uz_stem_im2col_u8 : NHWC uint8 image -> /255 -> (v - mean[c]) / std[c] -> bf16 -> 3x3 im2col operand of the stem
//                       GEMM, in one pass (same output as uz_stem_im2col on the normalised fp32 tensor).
//   uz_mask_u8_to_f32 : uint8 mask -> (m / 255 > 0.5) ? 1 : 0 as fp32 (the target of uz_bce_dice).
// The arithmetic is the reference's, operation for operation, in fp32 with IEEE division.
#include <cstdlib>

#include "pw.cuh"

namespace uz {
namespace {

struct Norm {
  float mean[8];
  float std[8];
};

template <int KP>
__global__ void stem_im2col_u8_kernel(const uint8_t* __restrict__ img, __nv_bfloat16* __restrict__ out, int N, int cin,
                                      int H, int W, const Norm nm) {
  const long long npix = static_cast<long long>(N) * H * W;
  for (long long p = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; p < npix;
       p += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int xw = static_cast<int>(p % W);
    const long long r0 = p / W;
    const int yh = static_cast<int>(r0 % H);
    const long long n = r0 / H;
    float v[KP];
#pragma unroll
    for (int k = 0; k < KP; ++k) v[k] = 0.f;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int yy = yh + t / 3 - 1, xx = xw + t % 3 - 1;
      const bool ok = yy >= 0 && yy < H && xx >= 0 && xx < W;
      const uint8_t* src = img + ((n * H + yy) * W + xx) * cin;
      for (int ci = 0; ci < cin; ++ci) {
        const int k = t * cin + ci;
        if (k < KP) {
          float f = 0.f;                       // zero padding applies to the NORMALISED image (Conv2d padding=1)
          if (ok) f = __fdiv_rn(__fdiv_rn(static_cast<float>(__ldg(src + ci)), 255.f) - nm.mean[ci], nm.std[ci]);
          v[k] = f;
        }
      }
    }
    uint4* o = reinterpret_cast<uint4*>(out + p * KP);
#pragma unroll
    for (int g = 0; g < KP / 8; ++g) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = v[g * 8 + j];
      o[g] = pw_pack8(f);
    }
  }
}

// Tiled form (the default; same scheme as stem_im2col_tile_kernel in pointwise.cu): a block normalises the 10 x 66
// pixel halo of its 8 x 64 tile ONCE into shared memory (the per-pixel kernel above divides every input value nine
// times), then each thread assembles one 16-byte channel group of one pixel - a warp stores 512 contiguous bytes.
constexpr int kU8TH = 8, kU8TW = 64;
template <int KP>
__global__ void __launch_bounds__(256)
stem_im2col_u8_tile_kernel(const uint8_t* __restrict__ img, __nv_bfloat16* __restrict__ out, int N, int cin, int H, int W,
                           int tiles_x, int tiles_y, const Norm nm) {
  extern __shared__ float u8_tile[];     // [cin][kU8TH + 2][kU8TW + 2], normalised fp32
  __shared__ int tab[KP];
  __shared__ float s_mean[8], s_std[8];
  constexpr int tw2 = kU8TW + 2, th2 = kU8TH + 2, per = th2 * tw2;
  constexpr int GV = KP / 8;
  if (threadIdx.x < KP) {
    const int k = threadIdx.x;
    int off = -1;
    if (k < 9 * cin) {
      const int t = k / cin, ci = k - t * cin;
      off = ci * per + (t / 3) * tw2 + (t % 3);
    }
    tab[k] = off;
  }
  if (threadIdx.x < 8) {
    s_mean[threadIdx.x] = nm.mean[threadIdx.x];
    s_std[threadIdx.x] = nm.std[threadIdx.x];
  }
  const long long tiles = static_cast<long long>(N) * tiles_y * tiles_x;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int tx = static_cast<int>(t % tiles_x);
    const long long r0 = t / tiles_x;
    const int ty = static_cast<int>(r0 % tiles_y);
    const long long n = r0 / tiles_y;
    const int x0 = tx * kU8TW, y0 = ty * kU8TH;
    __syncthreads();
    for (int i = threadIdx.x; i < cin * per; i += 256) {       // pixel-major sweep: the NHWC bytes coalesce
      const int r = i / cin, ci = i - r * cin;
      const int yy = y0 - 1 + r / tw2, xx = x0 - 1 + r % tw2;
      float f = 0.f;                       // zero padding applies to the NORMALISED image (Conv2d padding=1)
      if (yy >= 0 && yy < H && xx >= 0 && xx < W)
        f = __fdiv_rn(__fdiv_rn(static_cast<float>(__ldg(img + ((n * H + yy) * W + xx) * cin + ci)), 255.f) - s_mean[ci],
                      s_std[ci]);
      u8_tile[ci * per + r] = f;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kU8TH * kU8TW * GV; i += 256) {
      const int g = i % GV, px = (i / GV) % kU8TW, py = i / (GV * kU8TW);
      const int yy = y0 + py, xx = x0 + px;
      if (yy >= H || xx >= W) continue;
      const int at = py * tw2 + px;
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int off = tab[g * 8 + j];
        f[j] = off >= 0 ? u8_tile[off + at] : 0.f;
      }
      *reinterpret_cast<uint4*>(out + ((n * H + yy) * W + xx) * KP + g * 8) = pw_pack8(f);
    }
  }
}

__global__ void mask_u8_to_f32_kernel(const uint8_t* __restrict__ m, long long n, float* __restrict__ out) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = __fdiv_rn(static_cast<float>(__ldg(m + i)), 255.f) > 0.5f ? 1.f : 0.f;
}


// ---------------------------------------------------------------------------------------------- PIL-exact resize
// Pillow's ImagingResample for 8-bit images (libImaging/Resample.c), which is what transforms.Resize((512, 512)) runs
// on the decoded PIL image (datasets.py:41,47): a horizontal pass into a uint8 intermediate, then a vertical pass, each
// a fixed-point convolution  out = clip8((2^21 + sum_k in[x0 + k] * kk[k]) >> 22)  with per-output-pixel windows
// [x0, x0 + n) and integer coefficients kk (built on the host exactly as Pillow builds them, data.py).  Integer
// arithmetic: the device result is BIT-EXACT with Pillow.
constexpr int kResizePrecisionBits = 32 - 8 - 2;

__device__ __forceinline__ uint8_t clip8(int v) {
  v >>= kResizePrecisionBits;
  return static_cast<uint8_t>(v < 0 ? 0 : (v > 255 ? 255 : v));
}

// horizontal: src [N][H][Win][C] -> dst [N][H][Wout][C]; one thread per output (n, y, x), all C channels (C <= 4)
__global__ void resize_h_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, long long rows, int Win,
                                   int Wout, int C, const int* __restrict__ bounds, const int* __restrict__ kk,
                                   int ksize) {
  const long long total = rows * Wout;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int x = static_cast<int>(i % Wout);
    const long long r = i / Wout;
    const int x0 = __ldg(bounds + 2 * x), n = __ldg(bounds + 2 * x + 1);
    const int* k = kk + static_cast<long long>(x) * ksize;
    int acc[4] = {1 << (kResizePrecisionBits - 1), 1 << (kResizePrecisionBits - 1), 1 << (kResizePrecisionBits - 1),
                  1 << (kResizePrecisionBits - 1)};
    const uint8_t* p = src + (r * Win + x0) * C;
    for (int t = 0; t < n; ++t) {
      const int w = __ldg(k + t);
      for (int c = 0; c < C; ++c) acc[c] += static_cast<int>(__ldg(p + t * C + c)) * w;
    }
    uint8_t* o = dst + i * C;
    for (int c = 0; c < C; ++c) o[c] = clip8(acc[c]);
  }
}

// vertical: src [N][Hin][W][C] -> dst [N][Hout][W][C]; one thread per output byte (coalesced along the row)
__global__ void resize_v_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int N, int Hin, int Hout,
                                   long long rowbytes, const int* __restrict__ bounds, const int* __restrict__ kk,
                                   int ksize) {
  const long long total = static_cast<long long>(N) * Hout * rowbytes;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long b = i % rowbytes;
    const long long r = i / rowbytes;
    const int y = static_cast<int>(r % Hout);
    const long long n = r / Hout;
    const int y0 = __ldg(bounds + 2 * y), cnt = __ldg(bounds + 2 * y + 1);
    const int* k = kk + static_cast<long long>(y) * ksize;
    int acc = 1 << (kResizePrecisionBits - 1);
    const uint8_t* p = src + (n * Hin + y0) * rowbytes + b;
    for (int t = 0; t < cnt; ++t) acc += static_cast<int>(__ldg(p + t * rowbytes)) * __ldg(k + t);
    dst[i] = clip8(acc);
  }
}

}  // namespace
}  // namespace uz

using namespace uz;

extern "C" int uz_stem_im2col_u8(const uint8_t* img, int N, int cin, int H, int W, const float* mean_host,
                                 const float* std_host, void* out, int kp, void* stream) {
  PW_REQUIRE_DEVICE();
  if (!img || !out || !mean_host || !std_host || cin < 1 || cin > 7 || 9 * cin > kp ||
      (reinterpret_cast<uintptr_t>(out) & 15))
    return set_error(UZ_ERR_INVALID, "stem_im2col_u8: 1 <= cin <= 7, 9*cin=%d must fit kp=%d, output 16-B aligned",
                     9 * cin, kp);
  Norm nm{};
  for (int c = 0; c < cin; ++c) {
    if (!(std_host[c] > 0.f)) return set_error(UZ_ERR_INVALID, "stem_im2col_u8: std[%d] must be positive", c);
    nm.mean[c] = mean_host[c];
    nm.std[c] = std_host[c];
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  __nv_bfloat16* o = static_cast<__nv_bfloat16*>(out);
  if (std::getenv("UZ_STEM_GEN1") == nullptr && (kp == 16 || kp == 32 || kp == 64)) {
    const int tiles_x = (W + kU8TW - 1) / kU8TW, tiles_y = (H + kU8TH - 1) / kU8TH;
    const long long tiles = static_cast<long long>(N) * tiles_x * tiles_y;
    const int grid = static_cast<int>(tiles < 148 * 8 ? tiles : 148 * 8);
    const size_t smem = sizeof(float) * cin * (kU8TH + 2) * (kU8TW + 2);
    switch (kp) {
      case 16: stem_im2col_u8_tile_kernel<16><<<grid, 256, smem, st>>>(img, o, N, cin, H, W, tiles_x, tiles_y, nm); break;
      case 32: stem_im2col_u8_tile_kernel<32><<<grid, 256, smem, st>>>(img, o, N, cin, H, W, tiles_x, tiles_y, nm); break;
      default: stem_im2col_u8_tile_kernel<64><<<grid, 256, smem, st>>>(img, o, N, cin, H, W, tiles_x, tiles_y, nm); break;
    }
    PW_CHECK_LAUNCH("stem_im2col_u8 launch");
    return UZ_OK;
  }
  const long long npix = static_cast<long long>(N) * H * W;
  int blocks = static_cast<int>((npix + 127) / 128);
  if (blocks > 148 * 16) blocks = 148 * 16;
  switch (kp) {
    case 16: stem_im2col_u8_kernel<16><<<blocks, 128, 0, st>>>(img, o, N, cin, H, W, nm); break;
    case 32: stem_im2col_u8_kernel<32><<<blocks, 128, 0, st>>>(img, o, N, cin, H, W, nm); break;
    case 64: stem_im2col_u8_kernel<64><<<blocks, 128, 0, st>>>(img, o, N, cin, H, W, nm); break;
    default: return set_error(UZ_ERR_UNSUPPORTED, "stem_im2col_u8: kp must be 16, 32 or 64 (got %d)", kp);
  }
  PW_CHECK_LAUNCH("stem_im2col_u8 launch");
  return UZ_OK;
}

extern "C" int uz_mask_u8_to_f32(const uint8_t* mask, long long n, float* out, void* stream) {
  PW_REQUIRE_DEVICE();
  if (!mask || !out || n < 1) return set_error(UZ_ERR_INVALID, "mask_u8_to_f32: bad arguments");
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  mask_u8_to_f32_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(mask, n, out);
  PW_CHECK_LAUNCH("mask_u8_to_f32 launch");
  return UZ_OK;
}


// transforms.Resize on a decoded uint8 image (datasets.py:41,47), bit-exact with Pillow: horizontal pass (skipped when
// Wout == Win) into `tmp` [N][Hin][Wout][C], vertical pass (skipped when Hout == Hin) into `dst` [N][Hout][Wout][C].
// bounds_* [out][2] = {first input index, count}, kk_* [out][ksize_*] fixed-point coefficients (22 fractional bits) as
// built by unet_zoo_b200.data.resample_coeffs (Pillow's precompute_coeffs + normalize_coeffs_8bpc), device memory.
extern "C" int uz_resize_bilinear_u8(const uint8_t* src, int N, int Hin, int Win, int C, uint8_t* tmp, uint8_t* dst,
                                     int Hout, int Wout, const int* bounds_h, const int* kk_h, int ksize_h,
                                     const int* bounds_v, const int* kk_v, int ksize_v, void* stream) {
  PW_REQUIRE_DEVICE();
  if (!src || !dst || N < 1 || Hin < 1 || Win < 1 || Hout < 1 || Wout < 1 || C < 1 || C > 4)
    return set_error(UZ_ERR_INVALID, "resize_bilinear_u8: bad extents (1 <= C <= 4)");
  const bool do_h = Wout != Win, do_v = Hout != Hin;
  if ((do_h && (!bounds_h || !kk_h || ksize_h < 1)) || (do_v && (!bounds_v || !kk_v || ksize_v < 1)))
    return set_error(UZ_ERR_INVALID, "resize_bilinear_u8: missing coefficient tables");
  if (do_h && do_v && !tmp) return set_error(UZ_ERR_INVALID, "resize_bilinear_u8: two passes need the intermediate");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int sms_ = uz_device_sm_count();
  auto grid = [&](long long total) {
    long long g = (total + 255) / 256, cap = static_cast<long long>(sms_ > 0 ? sms_ : 148) * 16;
    return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
  };
  const uint8_t* vin = src;
  if (do_h) {
    uint8_t* hout = do_v ? tmp : dst;
    const long long rows = static_cast<long long>(N) * Hin;
    resize_h_u8_kernel<<<grid(rows * Wout), 256, 0, st>>>(src, hout, rows, Win, Wout, C, bounds_h, kk_h, ksize_h);
    PW_CHECK_LAUNCH("resize_h launch");
    vin = hout;
  }
  if (do_v) {
    const long long rowbytes = static_cast<long long>(Wout) * C;
    resize_v_u8_kernel<<<grid(static_cast<long long>(N) * Hout * rowbytes), 256, 0, st>>>(vin, dst, N, Hin, Hout, rowbytes,
                                                                                          bounds_v, kk_v, ksize_v);
    PW_CHECK_LAUNCH("resize_v launch");
  } else if (!do_h) {
    cudaError_t e = cudaMemcpyAsync(dst, src, static_cast<size_t>(N) * Hin * Win * C, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return cuda_error(e, "resize copy");
  }
  return UZ_OK;
}
