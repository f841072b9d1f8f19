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
