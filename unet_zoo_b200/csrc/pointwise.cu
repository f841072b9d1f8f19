// Bandwidth-bound kernels of the UNet hot path (sm_100a): weight packing, stem im2col, BatchNorm statistics /
// finalize / apply (+ReLU, +2x2 max-pool, +residual), BatchNorm+ReLU(+max-pool) backward, per-channel sums.
// All activations are NHWC bf16 addressed as (pointer, pixel stride `ld`), 8 channels (16 bytes) per access.
// Every thread keeps one fixed group of 8 channels for its whole lifetime so the per-channel constants live in
// registers; reductions are two-stage (per-block partials, then a deterministic finalize) - no float atomics.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdlib>

#include "../../include/unet_zoo_b200.h"
#include "host_util.h"

namespace uz {

// ---------------------------------------------------------------------------------------------- small device helpers
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = __uint_as_float(v.x << 16);
  f[1] = __uint_as_float(v.x & 0xFFFF0000u);
  f[2] = __uint_as_float(v.y << 16);
  f[3] = __uint_as_float(v.y & 0xFFFF0000u);
  f[4] = __uint_as_float(v.z << 16);
  f[5] = __uint_as_float(v.z & 0xFFFF0000u);
  f[6] = __uint_as_float(v.w << 16);
  f[7] = __uint_as_float(v.w & 0xFFFF0000u);
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack2(f[0], f[1]), pack2(f[2], f[3]), pack2(f[4], f[5]), pack2(f[6], f[7]));
}
__device__ __forceinline__ float bf16_round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void load8f(const float* p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
  f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

// launch geometry shared by the "fixed channel group per thread" kernels
struct Geo {
  int cv;      // channel groups (C / 8)
  int block;   // threads per block, multiple of cv
  int ppb;     // items (pixels or 2x2 quads) per block iteration
  int grid;
};
static Geo make_geo(int C, long long items, int blocks_per_sm) {
  Geo g;
  g.cv = C / 8;
  g.ppb = 256 / g.cv;
  if (g.ppb < 1) g.ppb = 1;
  g.block = g.ppb * g.cv;
  int sms = sm_count();
  if (sms < 1) sms = 148;
  long long need = (items + g.ppb - 1) / g.ppb;
  long long cap = static_cast<long long>(sms) * blocks_per_sm;
  g.grid = static_cast<int>(need < cap ? need : cap);
  if (g.grid < 1) g.grid = 1;
  return g;
}

// grid of a persistent grid-stride kernel: exactly the number of blocks that are resident at once (no partial second
// wave at reduced occupancy), capped by the number of partial-sum rows the workspace holds
template <typename K>
static int resident_grid(K kernel, int block, size_t smem, long long need, int cap_per_sm) {
  int per_sm = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
  if (per_sm > cap_per_sm) per_sm = cap_per_sm;
  int sms = sm_count();
  if (sms < 1) sms = 148;
  const long long cap = static_cast<long long>(sms) * per_sm;
  return static_cast<int>(need < cap ? (need < 1 ? 1 : need) : cap);
}

// ---------------------------------------------------------------------------------------------- weight packing
// dst[a*dst_pitch + t*B + b] = bf16(src[a*sa + b*sb + t*st] * (row_scale ? row_scale[a] : 1))
__global__ void pack_weight_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int A, int T, int B,
                                   long long sa, long long sb, long long st, long long dst_pitch,
                                   const float* __restrict__ row_scale) {
  const long long total = static_cast<long long>(A) * B;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int b = static_cast<int>(i % B);
    const int a = static_cast<int>(i / B);
    const float sc = row_scale ? __ldg(row_scale + a) : 1.f;
    const float* s = src + a * sa + b * sb;
    __nv_bfloat16* d = dst + a * dst_pitch + b;
    for (int t = 0; t < T; ++t) d[static_cast<long long>(t) * B] = __float2bfloat16_rn(__ldg(s + t * st) * sc);
  }
}

// Every weight of a model in ONE launch (a training step re-packs each convolution's fp32 master weights twice - forward
// and data-gradient layout - and every pack_weight launch costs ~5 us of a CUDA-graph step whatever its size: 235 launches
// in u2net).  jobs[j] = {src, dst, A, T, B, sa, sb, st, dst_pitch, first}: `first` = number of (a, b) pairs of the jobs
// before j; a thread finds its job by binary search.
constexpr int kPackJobFields = 10;
__global__ void pack_weights_batched_kernel(const long long* __restrict__ jobs, int njobs, long long total) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    int lo = 0, hi = njobs - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (__ldg(jobs + static_cast<size_t>(mid) * kPackJobFields + 9) <= i) lo = mid; else hi = mid - 1;
    }
    const long long* j = jobs + static_cast<size_t>(lo) * kPackJobFields;
    const float* src = reinterpret_cast<const float*>(__ldg(j + 0));
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(__ldg(j + 1));
    const int T = static_cast<int>(__ldg(j + 3)), B = static_cast<int>(__ldg(j + 4));
    const long long sa = __ldg(j + 5), sb = __ldg(j + 6), st = __ldg(j + 7), pitch = __ldg(j + 8);
    const long long e = i - __ldg(j + 9);
    const int b = static_cast<int>(e % B);
    const long long a = e / B;
    const float* sp = src + a * sa + b * sb;
    __nv_bfloat16* d = dst + a * pitch + b;
    for (int t = 0; t < T; ++t) d[static_cast<long long>(t) * B] = __float2bfloat16_rn(__ldg(sp + t * st));
  }
}

// Pre-summed weights of the nearest-x2 conv in phase form (host_conv.cu, UZ_UP2X_CONV3X3): for output phase (py, px) the
// rows r of the 3x3 filter that read low-resolution row offset a - 1 + py are  R(0,0) = {0}, R(0,1) = {1,2},
// R(1,0) = {0,1}, R(1,1) = {2}  (same for columns):  dst[x][ph*4 + a*2 + b][y] = bf16(sum_{r in R(py,a), s in R(px,b)} w[co][ci][r][s])
// with (x, y) = (co, ci) for the forward layout and (ci, co) for the data-gradient layout.  Sums in fp32, one rounding.
__global__ void pack_up2x_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ dst, int Co, int Ci,
                                 int transpose) {
  const long long total = static_cast<long long>(Co) * Ci;
  const int X = transpose ? Ci : Co, Y = transpose ? Co : Ci;
  (void)X;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int y = static_cast<int>(i % Y);
    const int x = static_cast<int>(i / Y);
    const int co = transpose ? y : x, ci = transpose ? x : y;
    float t[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) t[k] = __ldg(w + (static_cast<long long>(co) * Ci + ci) * 9 + k);
    __nv_bfloat16* d = dst + static_cast<long long>(x) * 16 * Y + y;
#pragma unroll
    for (int ph = 0; ph < 4; ++ph) {
      const int py = ph >> 1, px = ph & 1;
#pragma unroll
      for (int ab = 0; ab < 4; ++ab) {
        const int a = ab >> 1, b = ab & 1;
        const int r0 = py == 0 ? (a == 0 ? 0 : 1) : (a == 0 ? 0 : 2), r1 = py == 0 ? (a == 0 ? 0 : 2) : (a == 0 ? 1 : 2);
        const int s0 = px == 0 ? (b == 0 ? 0 : 1) : (b == 0 ? 0 : 2), s1 = px == 0 ? (b == 0 ? 0 : 2) : (b == 0 ? 1 : 2);
        float acc = 0.f;
        for (int r = r0; r <= r1; ++r)
          for (int c = s0; c <= s1; ++c) acc += t[r * 3 + c];
        d[static_cast<long long>(ph * 4 + ab) * Y] = __float2bfloat16_rn(acc);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------- stem im2col
// out[n, y, x, t*cin + ci] = x[n, ci, y + r - 1, x + s - 1] (t = r*3 + s), zero outside the image and for columns
// >= 9*cin.  One thread per pixel; reads are coalesced along W and served mostly by L1 (each input is used 9 times).
template <int KP>
__global__ void stem_im2col_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int N, int cin, int H,
                                   int W) {
  const long long npix = static_cast<long long>(N) * H * W;
  const long long plane = static_cast<long long>(H) * W;
  for (long long p = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; p < npix;
       p += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int xw = static_cast<int>(p % W);
    const long long r0 = p / W;
    const int yh = static_cast<int>(r0 % H);
    const int n = static_cast<int>(r0 / H);
    const float* xb = x + static_cast<long long>(n) * cin * plane;
    float v[KP];
#pragma unroll
    for (int k = 0; k < KP; ++k) v[k] = 0.f;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int yy = yh + t / 3 - 1, xx = xw + t % 3 - 1;
      const bool ok = yy >= 0 && yy < H && xx >= 0 && xx < W;
      for (int ci = 0; ci < cin; ++ci) {
        const int k = t * cin + ci;
        if (k < KP) v[k] = ok ? __ldg(xb + ci * plane + static_cast<long long>(yy) * W + xx) : 0.f;
      }
    }
    uint4* o = reinterpret_cast<uint4*>(out + p * KP);
#pragma unroll
    for (int g = 0; g < KP / 8; ++g) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = v[g * 8 + j];
      o[g] = pack8(f);
    }
  }
}

// Tiled form (the default): the one-thread-per-pixel kernel above writes 64 B per thread with 16-byte stores at a
// 64-byte lane stride and walks 27 scalar loads through 64-bit index arithmetic - 0.525 ms for 32 x 512^2 (1.2 TB/s,
// torch.profiler / ncu launch lists) where its bytes (100 MB in, 537 MB out) are worth ~0.12 ms.  Here a block stages
// the (8+2) x (64+2) halo of every input plane in shared memory with row-coalesced loads, then each thread assembles
// one 16-byte channel group of one pixel, so a warp stores 512 contiguous bytes.
constexpr int kStemTH = 8, kStemTW = 64;
template <int KP>
__global__ void __launch_bounds__(256)
stem_im2col_tile_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int N, int cin, int H, int W,
                        int tiles_x, int tiles_y) {
  extern __shared__ float stem_tile[];   // [cin][kStemTH + 2][kStemTW + 2]
  __shared__ int tab[KP];                // column k = t*cin + ci -> offset of (ci, r, s) inside the halo tile, -1 = zero column
  constexpr int tw2 = kStemTW + 2, th2 = kStemTH + 2, per = th2 * tw2;
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
  const long long tiles = static_cast<long long>(N) * tiles_y * tiles_x;
  for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int tx = static_cast<int>(t % tiles_x);
    const long long r0 = t / tiles_x;
    const int ty = static_cast<int>(r0 % tiles_y);
    const int n = static_cast<int>(r0 / tiles_y);
    const int x0 = tx * kStemTW, y0 = ty * kStemTH;
    __syncthreads();   // the previous tile's readers are done (and `tab` is visible on the first trip)
    for (int i = threadIdx.x; i < cin * per; i += 256) {
      const int ci = i / per, r = i - ci * per;
      const int yy = y0 - 1 + r / tw2, xx = x0 - 1 + r % tw2;
      stem_tile[i] = (yy >= 0 && yy < H && xx >= 0 && xx < W)
                         ? __ldg(x + (static_cast<long long>(n) * cin + ci) * H * W + static_cast<long long>(yy) * W + xx)
                         : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kStemTH * kStemTW * GV; i += 256) {
      const int g = i % GV, px = (i / GV) % kStemTW, py = i / (GV * kStemTW);
      const int yy = y0 + py, xx = x0 + px;
      if (yy >= H || xx >= W) continue;
      const int at = py * tw2 + px;
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int off = tab[g * 8 + j];
        f[j] = off >= 0 ? stem_tile[off + at] : 0.f;
      }
      *reinterpret_cast<uint4*>(out + ((static_cast<long long>(n) * H + yy) * W + xx) * KP + g * 8) = pack8(f);
    }
  }
}

// plain NCHW fp32 -> NHWC bf16 with channel padding (columns >= C are zero); one thread per (pixel, 8 channels)
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int N, int C, int H,
                                    int W, int out_ld, int cpad) {
  const long long plane = static_cast<long long>(H) * W;
  const int cv = cpad / 8;
  const long long total = static_cast<long long>(N) * plane * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    // pixel is the fast index so that the fp32 reads of one channel plane coalesce
    const long long pix = i % (static_cast<long long>(N) * plane);
    const int g = static_cast<int>(i / (static_cast<long long>(N) * plane));
    const int n = static_cast<int>(pix / plane);
    const long long hw = pix % plane;
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = g * 8 + j;
      f[j] = c < C ? __ldg(x + (static_cast<long long>(n) * C + c) * plane + hw) : 0.f;
    }
    *reinterpret_cast<uint4*>(out + pix * out_ld + g * 8) = pack8(f);
  }
}

// ---------------------------------------------------------------------------------------------- BatchNorm statistics
// per-channel sum / sum of squares of an arbitrary NHWC bf16 tensor -> partial[block][2][C]
__global__ void bn_stats_kernel(const __nv_bfloat16* __restrict__ x, int ld, long long npix, int C, int cv, int ppb,
                                float* __restrict__ partial) {
  extern __shared__ float red[];  // [ppb][cv][16]
  const int g = threadIdx.x % cv;
  const int pl = threadIdx.x / cv;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  for (long long p = static_cast<long long>(blockIdx.x) * ppb + pl; p < npix; p += static_cast<long long>(gridDim.x) * ppb) {
    float f[8];
    unpack8(ldg16(x + p * ld + g * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[j] += f[j];
      q[j] = fmaf(f[j], f[j], q[j]);
    }
  }
  float* mine = red + static_cast<size_t>(threadIdx.x) * 16;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mine[j] = s[j];
    mine[8 + j] = q[j];
  }
  __syncthreads();
  if (pl == 0) {
    for (int r = 1; r < ppb; ++r) {
      const float* o = red + (static_cast<size_t>(r) * cv + g) * 16;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s[j] += o[j];
        q[j] += o[8 + j];
      }
    }
    float* dst = partial + static_cast<size_t>(blockIdx.x) * 2 * C;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      dst[g * 8 + j] = s[j];
      dst[C + g * 8 + j] = q[j];
    }
  }
}

// lane-strided fp64 sums of one channel's {sum, sum-of-squares} partial rows; eight rows are loaded before any is added
// (the kernel is a chain of dependent L2 round trips otherwise: ~15 us for 1184 rows)
__device__ __forceinline__ void partial_column_sums(const float* __restrict__ partial, int rows, int C, int c, int lane,
                                                    double& s, double& q) {
  constexpr int U = 8;
  for (int r0 = lane; r0 < rows; r0 += 32 * U) {
    float a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int r = r0 + 32 * u;
      a[u] = r < rows ? __ldg(partial + static_cast<size_t>(r) * 2 * C + c) : 0.f;
      b[u] = r < rows ? __ldg(partial + static_cast<size_t>(r) * 2 * C + C + c) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      s += static_cast<double>(a[u]);
      q += static_cast<double>(b[u]);
    }
  }
}

// partial[rows][2][C] -> bn[4][C] = {mean, invstd, scale, shift}; running statistics updated like nn.BatchNorm2d
// (momentum EMA, unbiased variance for running_var).  One thread per channel, fp64 accumulation.
__global__ void bn_finalize_kernel(const float* __restrict__ partial, int rows, int C, double count,
                                   const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                   float momentum, float* __restrict__ running_mean, float* __restrict__ running_var,
                                   float* __restrict__ bn, long long* __restrict__ num_batches_tracked) {
  // one warp per channel: lanes stride over the partial rows, fp64 shuffle reduction (deterministic)
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (num_batches_tracked != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *num_batches_tracked += 1;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  partial_column_sums(partial, rows, C, c, lane, s, q);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane != 0) return;
  const double mean = s / count;
  double var = q / count - mean * mean;
  if (var < 0.0) var = 0.0;
  const float invstd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
  const float ga = gamma ? gamma[c] : 1.f;
  const float be = beta ? beta[c] : 0.f;
  const float scale = ga * invstd;
  bn[c] = static_cast<float>(mean);
  bn[C + c] = invstd;
  bn[2 * C + c] = scale;
  bn[3 * C + c] = be - static_cast<float>(mean) * scale;
  if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * static_cast<float>(mean);
  if (running_var) {
    const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * static_cast<float>(unbiased);
  }
}

// ---------------------------------------------------------------------------------------------- BatchNorm apply
// y = act(z*scale + shift [+ residual]) ; POOL additionally writes the 2x2 max of y.  item = pixel or 2x2 quad.
constexpr int kApplyUnroll = 4;
template <bool POOL>
__global__ void bn_apply_kernel(const __nv_bfloat16* __restrict__ z, int z_ld, int N, int H, int W, int C, int cv,
                                int ppb, const float* __restrict__ scale, const float* __restrict__ shift, int relu,
                                const __nv_bfloat16* __restrict__ res, int res_ld, __nv_bfloat16* __restrict__ y,
                                int y_ld, __nv_bfloat16* __restrict__ pooled, int pooled_ld) {
  const int g = threadIdx.x % cv;
  const int pl = threadIdx.x / cv;
  float sc[8], sh[8];
  if (scale) {
    load8f(scale + g * 8, sc);
    load8f(shift + g * 8, sh);
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sc[j] = 1.f;
      sh[j] = 0.f;
    }
  }
  // POOL: items are 2x2 quads INCLUDING the partial ones of an odd last row / column (MaxPool2d(2,2) floors: those
  // pixels get their activation but belong to no pooling window - common_layers.py:94 on odd sizes)
  const int Hq = POOL ? (H + 1) / 2 : H, Wq = POOL ? (W + 1) / 2 : W;
  const int Hp = H / 2, Wp = W / 2;   // pooled extent
  const long long items = static_cast<long long>(N) * Hq * Wq;
  for (long long it = static_cast<long long>(blockIdx.x) * ppb + pl; it < items;
       it += static_cast<long long>(gridDim.x) * ppb * (POOL ? 1 : kApplyUnroll)) {
    if (!POOL) {
      // kApplyUnroll pixels per thread iteration: all loads are issued before the arithmetic (bytes in flight)
      const long long stride = static_cast<long long>(gridDim.x) * ppb;
      uint4 zr[kApplyUnroll], rr[kApplyUnroll];
#pragma unroll
      for (int u = 0; u < kApplyUnroll; ++u) {
        const long long p = it + u * stride;
        if (p < items) {
          zr[u] = ldg16(z + p * z_ld + g * 8);
          if (res) rr[u] = ldg16(res + p * res_ld + g * 8);
        }
      }
#pragma unroll
      for (int u = 0; u < kApplyUnroll; ++u) {
        const long long p = it + u * stride;
        if (p < items) {
          float f[8];
          unpack8(zr[u], f);
          if (res) {
            float r[8];
            unpack8(rr[u], r);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmaf(f[j], sc[j], sh[j]) + r[j];
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmaf(f[j], sc[j], sh[j]);
          }
          if (relu) {
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
          }
          *reinterpret_cast<uint4*>(y + p * y_ld + g * 8) = pack8(f);
        }
      }
    } else {
      const int qx = static_cast<int>(it % Wq);
      const long long r0 = it / Wq;
      const int qy = static_cast<int>(r0 % Hq);
      const int n = static_cast<int>(r0 / Hq);
      const long long p00 = (static_cast<long long>(n) * H + 2 * qy) * W + 2 * qx;
      const long long pix[4] = {p00, p00 + 1, p00 + W, p00 + W + 1};
      const bool full = qy < Hp && qx < Wp;
      const bool ok[4] = {true, 2 * qx + 1 < W, 2 * qy + 1 < H, full};
      uint4 raw[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) raw[i] = ok[i] ? ldg16(z + pix[i] * z_ld + g * 8) : make_uint4(0u, 0u, 0u, 0u);
      float m[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float f[8];
        unpack8(raw[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          f[j] = fmaf(f[j], sc[j], sh[j]);
          if (relu) f[j] = fmaxf(f[j], 0.f);
        }
        const uint4 o = pack8(f);
        if (y && ok[i]) *reinterpret_cast<uint4*>(y + pix[i] * y_ld + g * 8) = o;
        unpack8(o, f);  // pool the values as stored
#pragma unroll
        for (int j = 0; j < 8; ++j) m[j] = i == 0 ? f[j] : fmaxf(m[j], f[j]);
      }
      if (full)
        *reinterpret_cast<uint4*>(pooled + ((static_cast<long long>(n) * Hp + qy) * Wp + qx) * pooled_ld + g * 8) = pack8(m);
    }
  }
}

// ---------------------------------------------------------------------------------------------- BatchNorm(+ReLU)(+pool) backward
// g = (dy + routed(dpool)) * [relu ? (z*scale+shift > 0) : 1] ;  x^ = (z - mean) * invstd
// reduce:  partial[block][2][C] = { sum g, sum g*x^ }
// apply :  dz = scale*g + A + B*z   with A,B from bn_bwd_finalize   (== scale*(g - mean(g) - x^*mean(g*x^)))
struct BwdArgs {
  const __nv_bfloat16* z;
  int z_ld;
  const __nv_bfloat16* dy;     // may be null
  int dy_ld;
  const __nv_bfloat16* dpool;  // may be null (POOL only)
  int dpool_ld;
  int N, H, W, C, cv, ppb;
  const float* bn;             // [4][C] mean, invstd, scale, shift
  int relu;
};

// Pooled path, register-lean form: the nine 16-byte words of one 2x2 quad (4 z, 4 dy, 1 dpool) stay raw; the gradient
// is rebuilt two channels at a time and handed to `sink(word, pixel, z_lo, z_hi, g_lo, g_hi)`.
struct QuadRaw {
  uint32_t z[4][4], d[4][4], dp[4];
  long long pix[4];
};
// Quads include the partial ones of an odd last row / column: their missing pixels read as z = dy = 0 and are marked
// pix = -1 (never stored), and they receive no pooled gradient - MaxPool2d(2,2) floors, so those pixels are in no window.
__device__ __forceinline__ void quad_load(const BwdArgs& a, long long it, int g, QuadRaw& r) {
  const int Hq = (a.H + 1) / 2, Wq = (a.W + 1) / 2, Hp = a.H / 2, Wp = a.W / 2;
  const int qx = static_cast<int>(it % Wq);
  const long long r0 = it / Wq;
  const int qy = static_cast<int>(r0 % Hq);
  const int n = static_cast<int>(r0 / Hq);
  const long long p00 = (static_cast<long long>(n) * a.H + 2 * qy) * a.W + 2 * qx;
  const bool full = qy < Hp && qx < Wp;
  const bool okx = 2 * qx + 1 < a.W, oky = 2 * qy + 1 < a.H;
  r.pix[0] = p00;
  r.pix[1] = okx ? p00 + 1 : -1;
  r.pix[2] = oky ? p00 + a.W : -1;
  r.pix[3] = full ? p00 + a.W + 1 : -1;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const bool ok = r.pix[i] >= 0;
    const uint4 v = ok ? ldg16(a.z + r.pix[i] * a.z_ld + g * 8) : make_uint4(0u, 0u, 0u, 0u);
    r.z[i][0] = v.x; r.z[i][1] = v.y; r.z[i][2] = v.z; r.z[i][3] = v.w;
    const uint4 d = (a.dy && ok) ? ldg16(a.dy + r.pix[i] * a.dy_ld + g * 8) : make_uint4(0u, 0u, 0u, 0u);
    r.d[i][0] = d.x; r.d[i][1] = d.y; r.d[i][2] = d.z; r.d[i][3] = d.w;
  }
  const uint4 p = (a.dpool && full)
                      ? ldg16(a.dpool + ((static_cast<long long>(n) * Hp + qy) * Wp + qx) * a.dpool_ld + g * 8)
                      : make_uint4(0u, 0u, 0u, 0u);
  r.dp[0] = p.x; r.dp[1] = p.y; r.dp[2] = p.z; r.dp[3] = p.w;
}
template <typename Sink>
__device__ __forceinline__ void quad_grad(const BwdArgs& a, const QuadRaw& r, const float (&sc)[8], const float (&sh)[8],
                                          Sink&& sink) {
#pragma unroll
  for (int w = 0; w < 4; ++w) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int j = 2 * w + h;
      float zv[4], yv[4], gv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        zv[i] = h ? __uint_as_float(r.z[i][w] & 0xFFFF0000u) : __uint_as_float(r.z[i][w] << 16);
        gv[i] = h ? __uint_as_float(r.d[i][w] & 0xFFFF0000u) : __uint_as_float(r.d[i][w] << 16);
        float u = fmaf(zv[i], sc[j], sh[j]);
        if (a.relu) u = fmaxf(u, 0.f);
        yv[i] = bf16_round(u);  // the forward pooled the stored (bf16) activations
      }
      const float dp = h ? __uint_as_float(r.dp[w] & 0xFFFF0000u) : __uint_as_float(r.dp[w] << 16);
      int am = 0;               // first maximum in row-major window order wins (nn.MaxPool2d index semantics)
      float mv = yv[0];
#pragma unroll
      for (int i = 1; i < 4; ++i)
        if (yv[i] > mv) {
          mv = yv[i];
          am = i;
        }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (i == am) gv[i] += dp;
        if (a.relu && !(yv[i] > 0.f)) gv[i] = 0.f;
        sink(w, h, i, zv[i], gv[i]);
      }
    }
  }
}

constexpr int kBwdUnrollReduce = 4;
constexpr int kBwdUnrollApply = 4;
// loads of one pixel on the no-pool path (issued for two pixels before any arithmetic: memory-level parallelism)
struct BwdLoad {
  uint4 z, d;
};
__device__ __forceinline__ void bwd_load(const BwdArgs& a, long long p, int g, BwdLoad& l) {
  l.z = ldg16(a.z + p * a.z_ld + g * 8);
  l.d = a.dy ? ldg16(a.dy + p * a.dy_ld + g * 8) : make_uint4(0u, 0u, 0u, 0u);
}
__device__ __forceinline__ void bwd_grad(const BwdArgs& a, const BwdLoad& l, const float (&sc)[8], const float (&sh)[8],
                                         float (&zv)[8], float (&gv)[8]) {
  unpack8(l.z, zv);
  unpack8(l.d, gv);
  if (a.relu) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (!(fmaf(zv[j], sc[j], sh[j]) > 0.f)) gv[j] = 0.f;
  }
}

// partial[block][2][C] = { sum g, sum g*z }  (x^ = (z - mean)*invstd is applied in the finalize kernel)
template <bool POOL>
__global__ void __launch_bounds__(256, POOL ? 2 : 3) bn_bwd_reduce_kernel(const BwdArgs a, float* __restrict__ partial) {
  extern __shared__ float red[];
  const int g = threadIdx.x % a.cv;
  const int pl = threadIdx.x / a.cv;
  float sc[8], sh[8];
  load8f(a.bn + 2 * a.C + g * 8, sc);
  load8f(a.bn + 3 * a.C + g * 8, sh);
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  const long long items = POOL ? static_cast<long long>(a.N) * ((a.H + 1) / 2) * ((a.W + 1) / 2)
                               : static_cast<long long>(a.N) * a.H * a.W;
  const long long stride = static_cast<long long>(gridDim.x) * a.ppb;
  if (POOL) {
    for (long long it = static_cast<long long>(blockIdx.x) * a.ppb + pl; it < items; it += stride) {
      QuadRaw r;
      quad_load(a, it, g, r);
      quad_grad(a, r, sc, sh, [&](int w, int h, int, float zv, float gv) {
        s[2 * w + h] += gv;
        q[2 * w + h] = fmaf(gv, zv, q[2 * w + h]);
      });
    }
  } else {
    constexpr int U = kBwdUnrollReduce;   // pixels in flight per thread: 2*U 16-byte loads issued before any arithmetic
    for (long long it = static_cast<long long>(blockIdx.x) * a.ppb + pl; it < items; it += U * stride) {
      BwdLoad l[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (it + u * stride < items) bwd_load(a, it + u * stride, g, l[u]);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (it + u * stride < items) {
          float zv[8], gv[8];
          bwd_grad(a, l[u], sc, sh, zv, gv);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            s[j] += gv[j];
            q[j] = fmaf(gv[j], zv[j], q[j]);
          }
        }
      }
    }
  }
  float* mine = red + static_cast<size_t>(threadIdx.x) * 16;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mine[j] = s[j];
    mine[8 + j] = q[j];
  }
  __syncthreads();
  if (pl == 0) {
    for (int r = 1; r < a.ppb; ++r) {
      const float* o = red + (static_cast<size_t>(r) * a.cv + g) * 16;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s[j] += o[j];
        q[j] += o[8 + j];
      }
    }
    float* dst = partial + static_cast<size_t>(blockIdx.x) * 2 * a.C;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      dst[g * 8 + j] = s[j];
      dst[a.C + g * 8 + j] = q[j];
    }
  }
}

// partial[rows][2][C] -> dgamma, dbeta (+=) and coef[2][C] = {A, B}
// q_is_gz: the second partial is sum g*z (not sum g*x^): sum g*x^ = invstd * (sum g*z - mean * sum g), in fp64
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ partial, int rows, int C, double count,
                                       const float* __restrict__ bn, float* __restrict__ dgamma,
                                       float* __restrict__ dbeta, int accumulate, float* __restrict__ coef, int q_is_gz,
                                       float* __restrict__ dbias) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  partial_column_sums(partial, rows, C, c, lane, s, q);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane != 0) return;
  const double mean = bn[c], invstd = bn[C + c], scale = bn[2 * C + c];
  if (q_is_gz) q = invstd * (q - mean * s);
  if (dgamma) dgamma[c] = (accumulate ? dgamma[c] : 0.f) + static_cast<float>(q);
  if (dbeta) dbeta[c] = (accumulate ? dbeta[c] : 0.f) + static_cast<float>(s);
  const double c1 = s / count, c2 = q / count;
  const double ca = -scale * c1 + scale * c2 * invstd * mean, cb = -scale * c2 * invstd;
  coef[c] = static_cast<float>(ca);
  coef[C + c] = static_cast<float>(cb);
  // gradient of a conv bias in front of a train-mode BatchNorm = sum dz = scale*sum g + n*A + B*sum z, which is
  // identically zero (sum z = n*mean); evaluated here in fp64 instead of summing the bf16-rounded dz tensor
  if (dbias) dbias[c] = (accumulate ? dbias[c] : 0.f) + static_cast<float>(scale * s + count * ca + cb * count * mean);
}

template <bool POOL>
__global__ void __launch_bounds__(256, 2) bn_bwd_apply_kernel(const BwdArgs a, const float* __restrict__ coef, __nv_bfloat16* __restrict__ dz,
                                    int dz_ld, float* __restrict__ dbias_partial, int dz_accumulate) {
  const int g = threadIdx.x % a.cv;
  const int pl = threadIdx.x / a.cv;
  float sc[8], sh[8], ca[8], cb[8];
  load8f(a.bn + 2 * a.C + g * 8, sc);
  load8f(a.bn + 3 * a.C + g * 8, sh);
  load8f(coef + g * 8, ca);
  load8f(coef + a.C + g * 8, cb);
  const long long items = POOL ? static_cast<long long>(a.N) * ((a.H + 1) / 2) * ((a.W + 1) / 2)
                               : static_cast<long long>(a.N) * a.H * a.W;
  const long long stride = static_cast<long long>(gridDim.x) * a.ppb;
  auto emit = [&](long long p, const float (&zv)[8], const float (&gv)[8]) {
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaf(sc[j], gv[j], fmaf(cb[j], zv[j], ca[j]));
    __nv_bfloat16* dst = dz + p * dz_ld + g * 8;
    if (dz_accumulate) {  // the destination already holds a gradient from another consumer of the tensor
      float old[8];
      unpack8(*reinterpret_cast<const uint4*>(dst), old);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += old[j];
    }
    *reinterpret_cast<uint4*>(dst) = pack8(o);
  };
  if (POOL) {
    for (long long it = static_cast<long long>(blockIdx.x) * a.ppb + pl; it < items; it += stride) {
      QuadRaw r;
      quad_load(a, it, g, r);
      float lo[4];               // low half of the output word being assembled, per pixel
      uint32_t out[4][4];
      quad_grad(a, r, sc, sh, [&](int w, int h, int i, float zv, float gv) {
        const int j = 2 * w + h;
        const float o = fmaf(sc[j], gv, fmaf(cb[j], zv, ca[j]));
        if (h == 0) lo[i] = o;
        else out[i][w] = pack2(lo[i], o);
      });
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (r.pix[i] < 0) continue;     // outside the image (partial quad of an odd last row / column)
        __nv_bfloat16* dst = dz + r.pix[i] * dz_ld + g * 8;
        uint4 v = make_uint4(out[i][0], out[i][1], out[i][2], out[i][3]);
        if (dz_accumulate) {
          float old[8], cur[8];
          unpack8(*reinterpret_cast<const uint4*>(dst), old);
          unpack8(v, cur);
#pragma unroll
          for (int j = 0; j < 8; ++j) cur[j] += old[j];
          v = pack8(cur);
        }
        *reinterpret_cast<uint4*>(dst) = v;
      }
    }
  } else {
    constexpr int U = kBwdUnrollApply;   // 2*U 16-byte loads in flight per thread; only the raw words stay live
    for (long long it = static_cast<long long>(blockIdx.x) * a.ppb + pl; it < items; it += U * stride) {
      BwdLoad l[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (it + u * stride < items) bwd_load(a, it + u * stride, g, l[u]);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (it + u * stride < items) {
          float zv[8], gv[8];
          bwd_grad(a, l[u], sc, sh, zv, gv);
          emit(it + u * stride, zv, gv);
        }
      }
    }
  }
  (void)dbias_partial;
}

// ---------------------------------------------------------------------------------------------- two BatchNorms + add + ReLU
// y = relu( actA(a*scaleA + shiftA) + (b*scaleB + shiftB) ),  actA = ReLU when relu_a, straight from the two raw
// (pre-BatchNorm) tensors: `x = F.relu(BN(cat) + BN(conv1x1(x)))` of Multiresblock (multiresunet.py:79-83) and
// `F.relu(F.relu(BN(conv3x3(x))) + BN(conv1x1(x)))` of Respath (multiresunet.py:123-126,130-133).  Neither normalised
// branch is ever stored.  Backward: g = dy * [y > 0] (recomputed from a and b), gA = g * [actA'], gB = g; one reduce pass
// feeds both BatchNorm backward finalizes, one apply pass writes both input gradients.
struct DualArgs {
  const __nv_bfloat16* a;
  const __nv_bfloat16* b;
  int a_ld, b_ld;
  const float* bnA;   // [4][C] mean, invstd, scale, shift
  const float* bnB;
  // backward only, optional: the BatchNorm that FOLLOWS y (out = BN_o(y), multiresunet.py:84,127,134).  With bnO / coefO
  // ([4][C] statistics, [2][C] coefficients of uz_bn_bwd_coef) the incoming gradient is d(out) and
  // dy = scale_o * d(out) + coefO[0] + coefO[1] * y is formed on the fly from the recomputed y - the gradient of y is
  // never written or read.
  const float* bnO;
  const float* coefO;
  int relu_a;
  long long npix;
  int C, cv, ppb;
};

constexpr int kDualUnroll = 2;
// `partial` (optional): [gridDim.x][2][C] per-channel sum / sum of squares of the STORED (bf16) output - the statistics of
// the BatchNorm that follows (multiresunet.py:84,127,134) without a pass of their own.
__global__ void __launch_bounds__(256, 2) dual_bn_fwd_kernel(const DualArgs d, __nv_bfloat16* __restrict__ y, int y_ld,
                                                             float* __restrict__ partial) {
  extern __shared__ float red[];
  const int g = threadIdx.x % d.cv;
  const int pl = threadIdx.x / d.cv;
  float sa[8], ha[8], sb[8], hb[8];
  load8f(d.bnA + 2 * d.C + g * 8, sa);
  load8f(d.bnA + 3 * d.C + g * 8, ha);
  load8f(d.bnB + 2 * d.C + g * 8, sb);
  load8f(d.bnB + 3 * d.C + g * 8, hb);
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  constexpr int U = 4;
  const long long stride = static_cast<long long>(gridDim.x) * d.ppb;
  for (long long it = static_cast<long long>(blockIdx.x) * d.ppb + pl; it < d.npix; it += U * stride) {
    uint4 ra[U], rb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long p = it + u * stride;
      if (p < d.npix) {
        ra[u] = ldg16(d.a + p * d.a_ld + g * 8);
        rb[u] = ldg16(d.b + p * d.b_ld + g * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long p = it + u * stride;
      if (p < d.npix) {
        float fa[8], fb[8];
        unpack8(ra[u], fa);
        unpack8(rb[u], fb);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float ua = fmaf(fa[j], sa[j], ha[j]);
          if (d.relu_a) ua = fmaxf(ua, 0.f);
          fa[j] = fmaxf(ua + fmaf(fb[j], sb[j], hb[j]), 0.f);
        }
        const uint4 o = pack8(fa);
        *reinterpret_cast<uint4*>(y + p * y_ld + g * 8) = o;
        if (partial) {
          unpack8(o, fa);   // statistics of the values as stored
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            s[j] += fa[j];
            q[j] = fmaf(fa[j], fa[j], q[j]);
          }
        }
      }
    }
  }
  if (!partial) return;
  float* mine = red + static_cast<size_t>(threadIdx.x) * 16;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mine[j] = s[j];
    mine[8 + j] = q[j];
  }
  __syncthreads();
  if (pl == 0) {
    for (int r = 1; r < d.ppb; ++r) {
      const float* o = red + (static_cast<size_t>(r) * d.cv + g) * 16;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s[j] += o[j];
        q[j] += o[8 + j];
      }
    }
    float* dst = partial + static_cast<size_t>(blockIdx.x) * 2 * d.C;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      dst[g * 8 + j] = s[j];
      dst[d.C + g * 8 + j] = q[j];
    }
  }
}

// gradient of one pixel group w.r.t. the two normalised branches (before the BatchNorm backward proper)
template <bool OUTER>
__device__ __forceinline__ void dual_grad(const DualArgs& d, int g, const uint4& ra, const uint4& rb, const uint4& rd,
                                          const float (&sa)[8], const float (&ha)[8], const float (&sb)[8],
                                          const float (&hb)[8], float (&fa)[8], float (&fb)[8], float (&ga)[8],
                                          float (&gb)[8]) {
  unpack8(ra, fa);
  unpack8(rb, fb);
  unpack8(rd, gb);
  float so[8], c0[8], c1[8];
  if (OUTER) {        // L1-resident after the first pixel; kept out of the persistent registers
    load8f(d.bnO + 2 * d.C + g * 8, so);
    load8f(d.coefO + g * 8, c0);
    load8f(d.coefO + d.C + g * 8, c1);
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float ua = fmaf(fa[j], sa[j], ha[j]);
    const float ya = d.relu_a ? fmaxf(ua, 0.f) : ua;
    const float sum = ya + fmaf(fb[j], sb[j], hb[j]);
    if (OUTER) gb[j] = fmaf(so[j], gb[j], fmaf(c1[j], bf16_round(fmaxf(sum, 0.f)), c0[j]));   // y as it was stored
    if (!(sum > 0.f)) gb[j] = 0.f;
    ga[j] = (d.relu_a && !(ua > 0.f)) ? 0.f : gb[j];
  }
}

// partialA[block][2][C] = { sum gA, sum gA*a } ; partialB[block][2][C] = { sum gB, sum gB*b }
template <bool OUTER>
__global__ void __launch_bounds__(256, 2) dual_bn_bwd_reduce_kernel(const DualArgs d, const __nv_bfloat16* __restrict__ dy,
                                                                    int dy_ld, float* __restrict__ partialA,
                                                                    float* __restrict__ partialB) {
  extern __shared__ float red[];
  const int g = threadIdx.x % d.cv;
  const int pl = threadIdx.x / d.cv;
  float sa[8], ha[8], sb[8], hb[8];
  load8f(d.bnA + 2 * d.C + g * 8, sa);
  load8f(d.bnA + 3 * d.C + g * 8, ha);
  load8f(d.bnB + 2 * d.C + g * 8, sb);
  load8f(d.bnB + 3 * d.C + g * 8, hb);
  float s1[8], q1[8], s2[8], q2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s1[j] = q1[j] = s2[j] = q2[j] = 0.f;
  constexpr int U = OUTER ? 1 : kDualUnroll;
  const long long stride = static_cast<long long>(gridDim.x) * d.ppb;
  for (long long it = static_cast<long long>(blockIdx.x) * d.ppb + pl; it < d.npix; it += U * stride) {
    uint4 ra[U], rb[U], rd[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long p = it + u * stride;
      if (p < d.npix) {
        ra[u] = ldg16(d.a + p * d.a_ld + g * 8);
        rb[u] = ldg16(d.b + p * d.b_ld + g * 8);
        rd[u] = ldg16(dy + p * dy_ld + g * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (it + u * stride < d.npix) {
        float fa[8], fb[8], ga[8], gb[8];
        dual_grad<OUTER>(d, g, ra[u], rb[u], rd[u], sa, ha, sb, hb, fa, fb, ga, gb);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          s1[j] += ga[j];
          q1[j] = fmaf(ga[j], fa[j], q1[j]);
          s2[j] += gb[j];
          q2[j] = fmaf(gb[j], fb[j], q2[j]);
        }
      }
    }
  }
  float* mine = red + static_cast<size_t>(threadIdx.x) * 32;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mine[j] = s1[j];
    mine[8 + j] = q1[j];
    mine[16 + j] = s2[j];
    mine[24 + j] = q2[j];
  }
  __syncthreads();
  if (pl == 0) {
    for (int r = 1; r < d.ppb; ++r) {
      const float* o = red + (static_cast<size_t>(r) * d.cv + g) * 32;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s1[j] += o[j];
        q1[j] += o[8 + j];
        s2[j] += o[16 + j];
        q2[j] += o[24 + j];
      }
    }
    float* da = partialA + static_cast<size_t>(blockIdx.x) * 2 * d.C;
    float* db = partialB + static_cast<size_t>(blockIdx.x) * 2 * d.C;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      da[g * 8 + j] = s1[j];
      da[d.C + g * 8 + j] = q1[j];
      db[g * 8 + j] = s2[j];
      db[d.C + g * 8 + j] = q2[j];
    }
  }
}

// dA = scaleA*gA + coefA[0] + coefA[1]*a ; dB likewise (coef from bn_bwd_finalize_kernel); dA may alias a, dB may alias b
template <bool OUTER>
__global__ void __launch_bounds__(256, 2) dual_bn_bwd_apply_kernel(const DualArgs d, const __nv_bfloat16* __restrict__ dy,
                                                                   int dy_ld, const float* __restrict__ coefA,
                                                                   const float* __restrict__ coefB,
                                                                   __nv_bfloat16* __restrict__ dA, int dA_ld, int accA,
                                                                   __nv_bfloat16* __restrict__ dB, int dB_ld, int accB) {
  const int g = threadIdx.x % d.cv;
  const int pl = threadIdx.x / d.cv;
  float sa[8], ha[8], sb[8], hb[8];
  load8f(d.bnA + 2 * d.C + g * 8, sa);
  load8f(d.bnA + 3 * d.C + g * 8, ha);
  load8f(d.bnB + 2 * d.C + g * 8, sb);
  load8f(d.bnB + 3 * d.C + g * 8, hb);
  constexpr int U = OUTER ? 1 : kDualUnroll;
  const long long stride = static_cast<long long>(gridDim.x) * d.ppb;
  for (long long it = static_cast<long long>(blockIdx.x) * d.ppb + pl; it < d.npix; it += U * stride) {
    uint4 ra[U], rb[U], rd[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long p = it + u * stride;
      if (p < d.npix) {
        ra[u] = ldg16(d.a + p * d.a_ld + g * 8);
        rb[u] = ldg16(d.b + p * d.b_ld + g * 8);
        rd[u] = ldg16(dy + p * dy_ld + g * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long p = it + u * stride;
      if (p < d.npix) {
        float fa[8], fb[8], ga[8], gb[8], c0[8], c1[8];
        dual_grad<OUTER>(d, g, ra[u], rb[u], rd[u], sa, ha, sb, hb, fa, fb, ga, gb);
        load8f(coefA + g * 8, c0);
        load8f(coefA + d.C + g * 8, c1);
#pragma unroll
        for (int j = 0; j < 8; ++j) ga[j] = fmaf(sa[j], ga[j], fmaf(c1[j], fa[j], c0[j]));
        load8f(coefB + g * 8, c0);
        load8f(coefB + d.C + g * 8, c1);
#pragma unroll
        for (int j = 0; j < 8; ++j) gb[j] = fmaf(sb[j], gb[j], fmaf(c1[j], fb[j], c0[j]));
        __nv_bfloat16* pa = dA + p * dA_ld + g * 8;
        __nv_bfloat16* pb = dB + p * dB_ld + g * 8;
        if (accA) {
          float old[8];
          unpack8(*reinterpret_cast<const uint4*>(pa), old);
#pragma unroll
          for (int j = 0; j < 8; ++j) ga[j] += old[j];
        }
        if (accB) {
          float old[8];
          unpack8(*reinterpret_cast<const uint4*>(pb), old);
#pragma unroll
          for (int j = 0; j < 8; ++j) gb[j] += old[j];
        }
        *reinterpret_cast<uint4*>(pa) = pack8(ga);
        *reinterpret_cast<uint4*>(pb) = pack8(gb);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------- column sums
// partial[block][C] = per-channel sum over this block's pixels
__global__ void colsum_kernel(const __nv_bfloat16* __restrict__ x, int ld, long long npix, int C, int cv, int ppb,
                              float* __restrict__ partial) {
  extern __shared__ float red[];
  const int g = threadIdx.x % cv;
  const int pl = threadIdx.x / cv;
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  for (long long p = static_cast<long long>(blockIdx.x) * ppb + pl; p < npix; p += static_cast<long long>(gridDim.x) * ppb) {
    float f[8];
    unpack8(ldg16(x + p * ld + g * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] += f[j];
  }
  float* mine = red + static_cast<size_t>(threadIdx.x) * 8;
#pragma unroll
  for (int j = 0; j < 8; ++j) mine[j] = s[j];
  __syncthreads();
  if (pl == 0) {
    for (int r = 1; r < ppb; ++r) {
      const float* o = red + (static_cast<size_t>(r) * cv + g) * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += o[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) partial[static_cast<size_t>(blockIdx.x) * C + g * 8 + j] = s[j];
  }
}

// out[c] (+)= sum_r partial[r][c]
__global__ void reduce_rows_kernel(const float* __restrict__ partial, int rows, int C, float* __restrict__ out,
                                   int accumulate) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= C) return;
  double s = 0.0;
  for (int r = lane; r < rows; r += 32) s += static_cast<double>(partial[static_cast<size_t>(r) * C + c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane != 0) return;
  out[c] = (accumulate ? out[c] : 0.f) + static_cast<float>(s);
}

cudaError_t launch_bn_bwd_finalize(const float* partial, int rows, int C, double count, const float* bn, float* dgamma,
                                   float* dbeta, int accumulate, float* coef, cudaStream_t st, int q_is_gz,
                                   float* dbias) {
  bn_bwd_finalize_kernel<<<(C + 3) / 4, 128, 0, st>>>(partial, rows, C, count, bn, dgamma, dbeta, accumulate, coef,
                                                         q_is_gz, dbias);
  return cudaGetLastError();
}
cudaError_t launch_reduce_rows(const float* partial, int rows, int C, float* out, int accumulate, cudaStream_t st) {
  reduce_rows_kernel<<<(C + 3) / 4, 128, 0, st>>>(partial, rows, C, out, accumulate);
  return cudaGetLastError();
}

}  // namespace uz

using namespace uz;

// ================================================================================================ C ABI
static bool a16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
#define UZ_REQUIRE_DEVICE()                \
  do {                                     \
    const int sms_ = uz_device_sm_count(); \
    if (sms_ < 0) return sms_;             \
  } while (0)
#define UZ_CHECK_LAUNCH(what)                                   \
  do {                                                          \
    cudaError_t e_ = cudaGetLastError();                        \
    if (e_ != cudaSuccess) return cuda_error(e_, what);         \
  } while (0)

extern "C" int uz_pack_weight(const float* src, void* dst_bf16, int A, int T, int B, long long sa, long long sb,
                              long long st, long long dst_pitch, const float* row_scale, void* stream) {
  UZ_REQUIRE_DEVICE();
  if (A < 1 || T < 1 || B < 1 || dst_pitch < static_cast<long long>(T) * B)
    return set_error(UZ_ERR_INVALID, "pack_weight: bad extents A=%d T=%d B=%d pitch=%lld", A, T, B, dst_pitch);
  const long long total = static_cast<long long>(A) * B;
  int blocks = static_cast<int>((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  pack_weight_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst_bf16), A, T, B, sa, sb, st, dst_pitch, row_scale);
  UZ_CHECK_LAUNCH("pack_weight launch");
  return UZ_OK;
}

extern "C" int uz_pack_weights_batched(const long long* jobs_device, int njobs, long long total_pairs, void* stream) {
  UZ_REQUIRE_DEVICE();
  if (!jobs_device || njobs < 1 || total_pairs < 1) return set_error(UZ_ERR_INVALID, "pack_weights_batched: bad arguments");
  long long blocks = (total_pairs + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  pack_weights_batched_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(jobs_device, njobs,
                                                                                                      total_pairs);
  UZ_CHECK_LAUNCH("pack_weights_batched launch");
  return UZ_OK;
}

extern "C" int uz_up2x_phase_form(int H_low, int W_low, int cout) {
  static const bool gen1 = std::getenv("UZ_UP2X_GEN1") != nullptr && std::atoi(std::getenv("UZ_UP2X_GEN1")) != 0;
  return (!gen1 && H_low >= 16 && W_low >= 8 && cout >= 64 && (cout % 64) == 0) ? 1 : 0;
}

extern "C" int uz_pack_up2x(const float* w, void* dst_bf16, int Co, int Ci, int transpose, void* stream) {
  UZ_REQUIRE_DEVICE();
  if (!w || !dst_bf16 || Co < 1 || Ci < 1) return set_error(UZ_ERR_INVALID, "pack_up2x: bad extents");
  const long long total = static_cast<long long>(Co) * Ci;
  int blocks = static_cast<int>((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  pack_up2x_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(w, static_cast<__nv_bfloat16*>(dst_bf16), Co,
                                                                           Ci, transpose);
  UZ_CHECK_LAUNCH("pack_up2x launch");
  return UZ_OK;
}

extern "C" int uz_stem_im2col(const float* x, int N, int cin, int H, int W, void* out, int kp, void* stream) {
  UZ_REQUIRE_DEVICE();
  if (cin < 1 || 9 * cin > kp || !a16(out))
    return set_error(UZ_ERR_INVALID, "stem_im2col: 9*cin=%d must fit kp=%d, output 16-B aligned", 9 * cin, kp);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  __nv_bfloat16* o = static_cast<__nv_bfloat16*>(out);
  if (std::getenv("UZ_STEM_GEN1") == nullptr && (kp == 16 || kp == 32 || kp == 64)) {
    const int tiles_x = (W + kStemTW - 1) / kStemTW, tiles_y = (H + kStemTH - 1) / kStemTH;
    const long long tiles = static_cast<long long>(N) * tiles_x * tiles_y;
    const int grid = static_cast<int>(tiles < 148 * 8 ? tiles : 148 * 8);
    const size_t smem = sizeof(float) * cin * (kStemTH + 2) * (kStemTW + 2);   // cin <= 7: at most 18.5 KB
    switch (kp) {
      case 16: stem_im2col_tile_kernel<16><<<grid, 256, smem, st>>>(x, o, N, cin, H, W, tiles_x, tiles_y); break;
      case 32: stem_im2col_tile_kernel<32><<<grid, 256, smem, st>>>(x, o, N, cin, H, W, tiles_x, tiles_y); break;
      default: stem_im2col_tile_kernel<64><<<grid, 256, smem, st>>>(x, o, N, cin, H, W, tiles_x, tiles_y); break;
    }
    UZ_CHECK_LAUNCH("stem_im2col launch");
    return UZ_OK;
  }
  const long long npix = static_cast<long long>(N) * H * W;
  int blocks = static_cast<int>((npix + 127) / 128);
  if (blocks > 148 * 16) blocks = 148 * 16;
  switch (kp) {
    case 16: stem_im2col_kernel<16><<<blocks, 128, 0, st>>>(x, o, N, cin, H, W); break;
    case 32: stem_im2col_kernel<32><<<blocks, 128, 0, st>>>(x, o, N, cin, H, W); break;
    case 64: stem_im2col_kernel<64><<<blocks, 128, 0, st>>>(x, o, N, cin, H, W); break;
    default: return set_error(UZ_ERR_UNSUPPORTED, "stem_im2col: kp must be 16, 32 or 64 (got %d)", kp);
  }
  UZ_CHECK_LAUNCH("stem_im2col launch");
  return UZ_OK;
}

extern "C" int uz_nchw_to_nhwc(const float* x, int N, int C, int H, int W, void* out, int out_ld, void* stream) {
  UZ_REQUIRE_DEVICE();
  const int cpad = (C + 7) / 8 * 8;
  if (out_ld < cpad || (out_ld & 7) || !a16(out))
    return set_error(UZ_ERR_INVALID, "nchw_to_nhwc: out_ld=%d must be a multiple of 8 >= %d", out_ld, cpad);
  const long long total = static_cast<long long>(N) * H * W * (cpad / 8);
  int blocks = static_cast<int>((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  nchw_to_nhwc_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, static_cast<__nv_bfloat16*>(out), N, C,
                                                                            H, W, out_ld, cpad);
  UZ_CHECK_LAUNCH("nchw_to_nhwc launch");
  return UZ_OK;
}

static int check_act(const char* who, const void* p, int C, int ld) {
  if (p == nullptr || !a16(p) || C < 8 || (C & 7) || (ld & 7) || ld < C || C > 2048)
    return set_error(UZ_ERR_INVALID, "%s: activation needs 16-B aligned pointer, C (=%d) and ld (=%d) multiples of 8, 8 <= C <= 2048",
                     who, C, ld);
  return UZ_OK;
}

extern "C" int uz_reduce_rows() { return uz_device_sm_count() * 4; }

extern "C" int uz_bn_stats(const void* x, int ld, long long npix, int C, float* partial, void* stream) {
  UZ_REQUIRE_DEVICE();
  int rc = check_act("bn_stats", x, C, ld);
  if (rc) return rc;
  const Geo g = make_geo(C, npix, 4);
  // always fill uz_reduce_rows() rows so the consumer can use a fixed row count
  const int rows = uz_reduce_rows();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemsetAsync(partial, 0, sizeof(float) * 2 * C * rows, st);
  if (e != cudaSuccess) return cuda_error(e, "bn_stats memset");
  bn_stats_kernel<<<g.grid, g.block, g.block * 16 * sizeof(float), st>>>(static_cast<const __nv_bfloat16*>(x), ld, npix,
                                                                         C, g.cv, g.ppb, partial);
  UZ_CHECK_LAUNCH("bn_stats launch");
  return UZ_OK;
}

extern "C" int uz_bn_finalize(const float* partial, int rows, int C, long long count, const float* gamma,
                              const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                              float* bn, long long* num_batches_tracked, void* stream) {
  UZ_REQUIRE_DEVICE();
  if (rows < 1 || C < 1 || count < 1) return set_error(UZ_ERR_INVALID, "bn_finalize: bad extents");
  bn_finalize_kernel<<<(C + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      partial, rows, C, static_cast<double>(count), gamma, beta, eps, momentum, running_mean, running_var, bn,
      num_batches_tracked);
  UZ_CHECK_LAUNCH("bn_finalize launch");
  return UZ_OK;
}

extern "C" int uz_bn_apply(const void* z, int z_ld, int N, int H, int W, int C, const float* scale, const float* shift,
                           int relu, const void* residual, int res_ld, void* y, int y_ld, void* pooled, int pooled_ld,
                           void* stream) {
  UZ_REQUIRE_DEVICE();
  int rc = check_act("bn_apply(z)", z, C, z_ld);
  if (rc) return rc;
  if (y && (rc = check_act("bn_apply(y)", y, C, y_ld))) return rc;
  if (residual && (rc = check_act("bn_apply(residual)", residual, C, res_ld))) return rc;
  if ((scale == nullptr) != (shift == nullptr)) return set_error(UZ_ERR_INVALID, "bn_apply: scale/shift must both be given");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (pooled) {
    if ((rc = check_act("bn_apply(pooled)", pooled, C, pooled_ld))) return rc;
    if (H < 2 || W < 2) return set_error(UZ_ERR_INVALID, "bn_apply: pooling needs H, W >= 2 (got %dx%d)", H, W);
    if (residual) return set_error(UZ_ERR_UNSUPPORTED, "bn_apply: residual + pool not supported");
    const Geo g = make_geo(C, static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2), 8);
    bn_apply_kernel<true><<<g.grid, g.block, 0, st>>>(
        static_cast<const __nv_bfloat16*>(z), z_ld, N, H, W, C, g.cv, g.ppb, scale, shift, relu, nullptr, 0,
        static_cast<__nv_bfloat16*>(y), y_ld, static_cast<__nv_bfloat16*>(pooled), pooled_ld);
  } else {
    if (!y) return set_error(UZ_ERR_INVALID, "bn_apply: no output");
    const Geo g = make_geo(C, static_cast<long long>(N) * H * W, 8);
    const long long items = static_cast<long long>(N) * H * W;
    const int grid = resident_grid(bn_apply_kernel<false>, g.block, 0, (items + g.ppb - 1) / g.ppb, 8);
    bn_apply_kernel<false><<<grid, g.block, 0, st>>>(
        static_cast<const __nv_bfloat16*>(z), z_ld, N, H, W, C, g.cv, g.ppb, scale, shift, relu,
        static_cast<const __nv_bfloat16*>(residual), res_ld, static_cast<__nv_bfloat16*>(y), y_ld, nullptr, 0);
  }
  UZ_CHECK_LAUNCH("bn_apply launch");
  return UZ_OK;
}

extern "C" size_t uz_bn_bwd_workspace_bytes(int C) {
  const int rows = uz_reduce_rows();
  if (rows < 0) return 0;
  return sizeof(float) * static_cast<size_t>(rows) * 3 * C + sizeof(float) * 2 * C;
}

extern "C" int uz_bn_bwd(const void* z, int z_ld, const void* dy, int dy_ld, const void* dpool, int dpool_ld, int N,
                         int H, int W, int C, const float* bn, int relu, void* dz, int dz_ld, float* dgamma,
                         float* dbeta, float* dbias, int accumulate, void* workspace, size_t workspace_bytes,
                         void* stream) {
  UZ_REQUIRE_DEVICE();
  int rc = check_act("bn_bwd(z)", z, C, z_ld);
  if (rc) return rc;
  if ((rc = check_act("bn_bwd(dz)", dz, C, dz_ld))) return rc;
  if (dy && (rc = check_act("bn_bwd(dy)", dy, C, dy_ld))) return rc;
  if (dpool && (rc = check_act("bn_bwd(dpool)", dpool, C, dpool_ld))) return rc;
  if (!dy && !dpool) return set_error(UZ_ERR_INVALID, "bn_bwd: no incoming gradient");
  if (dpool && (H < 2 || W < 2)) return set_error(UZ_ERR_INVALID, "bn_bwd: pooling needs H, W >= 2");
  if (workspace_bytes < uz_bn_bwd_workspace_bytes(C) || !a16(workspace))
    return set_error(UZ_ERR_WORKSPACE, "bn_bwd: workspace too small (%zu < %zu)", workspace_bytes,
                     uz_bn_bwd_workspace_bytes(C));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int dz_acc = (accumulate & 2) ? 1 : 0;
  accumulate &= 1;
  const bool pool = dpool != nullptr;
  const long long items = pool ? static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2)
                               : static_cast<long long>(N) * H * W;
  Geo g = make_geo(C, items, 4);
  const size_t smem = static_cast<size_t>(g.block) * 16 * sizeof(float);
  const long long need = (items + g.ppb - 1) / g.ppb;
  const int grid_r = pool ? resident_grid(bn_bwd_reduce_kernel<true>, g.block, smem, need, 4)
                          : resident_grid(bn_bwd_reduce_kernel<false>, g.block, smem, need, 4);
  const int grid_a = pool ? resident_grid(bn_bwd_apply_kernel<true>, g.block, 0, need, 4)
                          : resident_grid(bn_bwd_apply_kernel<false>, g.block, 0, need, 4);
  const int rows = grid_r;
  float* partial = static_cast<float*>(workspace);
  float* dbias_partial = partial + static_cast<size_t>(uz_reduce_rows()) * 2 * C;
  float* coef = dbias_partial + static_cast<size_t>(uz_reduce_rows()) * C;
  BwdArgs a;
  a.z = static_cast<const __nv_bfloat16*>(z);
  a.z_ld = z_ld;
  a.dy = static_cast<const __nv_bfloat16*>(dy);
  a.dy_ld = dy_ld;
  a.dpool = static_cast<const __nv_bfloat16*>(dpool);
  a.dpool_ld = dpool_ld;
  a.N = N; a.H = H; a.W = W; a.C = C; a.cv = g.cv; a.ppb = g.ppb;
  a.bn = bn;
  a.relu = relu;
  if (pool)
    bn_bwd_reduce_kernel<true><<<grid_r, g.block, smem, st>>>(a, partial);
  else
    bn_bwd_reduce_kernel<false><<<grid_r, g.block, smem, st>>>(a, partial);
  UZ_CHECK_LAUNCH("bn_bwd_reduce launch");
  bn_bwd_finalize_kernel<<<(C + 3) / 4, 128, 0, st>>>(partial, rows, C, static_cast<double>(N) * H * W, bn, dgamma,
                                                         dbeta, accumulate, coef, 1, dbias);
  UZ_CHECK_LAUNCH("bn_bwd_finalize launch");
  float* dbp = nullptr;  // the bias gradient is produced analytically by the finalize kernel
  if (pool)
    bn_bwd_apply_kernel<true><<<grid_a, g.block, 0, st>>>(a, coef, static_cast<__nv_bfloat16*>(dz), dz_ld, dbp,
                                                             dz_acc);
  else
    bn_bwd_apply_kernel<false><<<grid_a, g.block, 0, st>>>(a, coef, static_cast<__nv_bfloat16*>(dz), dz_ld, dbp,
                                                              dz_acc);
  UZ_CHECK_LAUNCH("bn_bwd_apply launch");
  (void)dbias_partial;
  return UZ_OK;
}

extern "C" int uz_bn_bwd_coef(const void* z, int z_ld, const void* dy, int dy_ld, int N, int H, int W, int C, const float* bn,
                              int relu, float* coef, float* dgamma, float* dbeta, float* dbias, int accumulate,
                              void* workspace, size_t workspace_bytes, void* stream) {
  UZ_REQUIRE_DEVICE();
  int rc = check_act("bn_bwd_coef(z)", z, C, z_ld);
  if (rc || (rc = check_act("bn_bwd_coef(dy)", dy, C, dy_ld))) return rc;
  if (coef == nullptr || bn == nullptr) return set_error(UZ_ERR_INVALID, "bn_bwd_coef: missing statistics / output");
  if (workspace_bytes < uz_bn_bwd_workspace_bytes(C) || !a16(workspace))
    return set_error(UZ_ERR_WORKSPACE, "bn_bwd_coef: workspace too small (%zu < %zu)", workspace_bytes,
                     uz_bn_bwd_workspace_bytes(C));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long items = static_cast<long long>(N) * H * W;
  Geo g = make_geo(C, items, 4);
  const size_t smem = static_cast<size_t>(g.block) * 16 * sizeof(float);
  const int grid_r = resident_grid(bn_bwd_reduce_kernel<false>, g.block, smem, (items + g.ppb - 1) / g.ppb, 4);
  BwdArgs a;
  a.z = static_cast<const __nv_bfloat16*>(z);
  a.z_ld = z_ld;
  a.dy = static_cast<const __nv_bfloat16*>(dy);
  a.dy_ld = dy_ld;
  a.dpool = nullptr;
  a.dpool_ld = 0;
  a.N = N; a.H = H; a.W = W; a.C = C; a.cv = g.cv; a.ppb = g.ppb;
  a.bn = bn;
  a.relu = relu;
  float* partial = static_cast<float*>(workspace);
  bn_bwd_reduce_kernel<false><<<grid_r, g.block, smem, st>>>(a, partial);
  UZ_CHECK_LAUNCH("bn_bwd_coef reduce launch");
  bn_bwd_finalize_kernel<<<(C + 3) / 4, 128, 0, st>>>(partial, grid_r, C, static_cast<double>(items), bn, dgamma, dbeta,
                                                         accumulate & 1, coef, 1, dbias);
  UZ_CHECK_LAUNCH("bn_bwd_coef finalize launch");
  return UZ_OK;
}

extern "C" size_t uz_dual_bn_bwd_workspace_bytes(int C) {
  const int rows = uz_reduce_rows();
  if (rows < 0) return 0;
  return sizeof(float) * static_cast<size_t>(rows) * 4 * C + sizeof(float) * 4 * C;
}

static int dual_args(DualArgs& d, Geo& g, const void* a, int a_ld, const float* bnA, int relu_a, const void* b, int b_ld,
                     const float* bnB, int N, int H, int W, int C, int blocks_per_sm) {
  int rc = check_act("dual_bn(a)", a, C, a_ld);
  if (rc || (rc = check_act("dual_bn(b)", b, C, b_ld))) return rc;
  if (!bnA || !bnB || N < 1 || H < 1 || W < 1) return set_error(UZ_ERR_INVALID, "dual_bn: missing statistics or bad extents");
  d.a = static_cast<const __nv_bfloat16*>(a);
  d.b = static_cast<const __nv_bfloat16*>(b);
  d.a_ld = a_ld;
  d.b_ld = b_ld;
  d.bnA = bnA;
  d.bnB = bnB;
  d.bnO = nullptr;
  d.coefO = nullptr;
  d.relu_a = relu_a;
  d.npix = static_cast<long long>(N) * H * W;
  d.C = C;
  g = make_geo(C, d.npix, blocks_per_sm);
  d.cv = g.cv;
  d.ppb = g.ppb;
  return UZ_OK;
}

extern "C" int uz_dual_bn_add_relu_fwd(const void* a, int a_ld, const float* bnA, int relu_a, const void* b, int b_ld,
                                       const float* bnB, int N, int H, int W, int C, void* y, int y_ld, float* y_stats,
                                       void* stream) {
  UZ_REQUIRE_DEVICE();
  DualArgs d;
  Geo g;
  int rc = dual_args(d, g, a, a_ld, bnA, relu_a, b, b_ld, bnB, N, H, W, C, 4);   // grid <= uz_reduce_rows()
  if (rc || (rc = check_act("dual_bn(y)", y, C, y_ld))) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (y_stats) {   // every one of the uz_reduce_rows() rows is defined: the consumer (uz_bn_finalize) takes a fixed count
    cudaError_t e = cudaMemsetAsync(y_stats, 0, sizeof(float) * 2 * C * uz_reduce_rows(), st);
    if (e != cudaSuccess) return cuda_error(e, "dual_bn_fwd memset");
  }
  dual_bn_fwd_kernel<<<g.grid, g.block, y_stats ? g.block * 16 * sizeof(float) : 0, st>>>(
      d, static_cast<__nv_bfloat16*>(y), y_ld, y_stats);
  UZ_CHECK_LAUNCH("dual_bn_fwd launch");
  return UZ_OK;
}

extern "C" int uz_dual_bn_add_relu_bwd(const void* a, int a_ld, const float* bnA, int relu_a, const void* b, int b_ld,
                                       const float* bnB, int N, int H, int W, int C, const void* dy, int dy_ld,
                                       const float* bn_outer, const float* coef_outer, void* dA, int dA_ld, int accA,
                                       float* dbiasA, void* dB, int dB_ld, int accB, float* dbiasB, void* workspace,
                                       size_t workspace_bytes, void* stream) {
  UZ_REQUIRE_DEVICE();
  DualArgs d;
  Geo g;
  int rc = dual_args(d, g, a, a_ld, bnA, relu_a, b, b_ld, bnB, N, H, W, C, 4);
  if ((bn_outer == nullptr) != (coef_outer == nullptr))
    return set_error(UZ_ERR_INVALID, "dual_bn_bwd: bn_outer and coef_outer must be given together");
  d.bnO = bn_outer;
  d.coefO = coef_outer;
  if (rc || (rc = check_act("dual_bn_bwd(dy)", dy, C, dy_ld)) || (rc = check_act("dual_bn_bwd(dA)", dA, C, dA_ld)) ||
      (rc = check_act("dual_bn_bwd(dB)", dB, C, dB_ld)))
    return rc;
  if (workspace_bytes < uz_dual_bn_bwd_workspace_bytes(C) || !a16(workspace))
    return set_error(UZ_ERR_WORKSPACE, "dual_bn_bwd: workspace too small (%zu < %zu)", workspace_bytes,
                     uz_dual_bn_bwd_workspace_bytes(C));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = static_cast<size_t>(g.block) * 32 * sizeof(float);
  const long long need = (d.npix + g.ppb - 1) / g.ppb;
  const bool outer = coef_outer != nullptr;
  const int grid_r = outer ? resident_grid(dual_bn_bwd_reduce_kernel<true>, g.block, smem, need, 4)
                           : resident_grid(dual_bn_bwd_reduce_kernel<false>, g.block, smem, need, 4);
  const int grid_a = outer ? resident_grid(dual_bn_bwd_apply_kernel<true>, g.block, 0, need, 4)
                           : resident_grid(dual_bn_bwd_apply_kernel<false>, g.block, 0, need, 4);
  const size_t max_rows = static_cast<size_t>(uz_reduce_rows());
  float* partialA = static_cast<float*>(workspace);
  float* partialB = partialA + max_rows * 2 * C;
  float* coefA = partialB + max_rows * 2 * C;
  float* coefB = coefA + 2 * C;
  const __nv_bfloat16* g_dy = static_cast<const __nv_bfloat16*>(dy);
  if (outer)
    dual_bn_bwd_reduce_kernel<true><<<grid_r, g.block, smem, st>>>(d, g_dy, dy_ld, partialA, partialB);
  else
    dual_bn_bwd_reduce_kernel<false><<<grid_r, g.block, smem, st>>>(d, g_dy, dy_ld, partialA, partialB);
  UZ_CHECK_LAUNCH("dual_bn_bwd_reduce launch");
  const double count = static_cast<double>(d.npix);
  bn_bwd_finalize_kernel<<<(C + 3) / 4, 128, 0, st>>>(partialA, grid_r, C, count, bnA, nullptr, nullptr, 0, coefA, 1, dbiasA);
  UZ_CHECK_LAUNCH("dual_bn_bwd finalize A launch");
  bn_bwd_finalize_kernel<<<(C + 3) / 4, 128, 0, st>>>(partialB, grid_r, C, count, bnB, nullptr, nullptr, 0, coefB, 1, dbiasB);
  UZ_CHECK_LAUNCH("dual_bn_bwd finalize B launch");
  if (outer)
    dual_bn_bwd_apply_kernel<true><<<grid_a, g.block, 0, st>>>(d, g_dy, dy_ld, coefA, coefB, static_cast<__nv_bfloat16*>(dA),
                                                               dA_ld, accA, static_cast<__nv_bfloat16*>(dB), dB_ld, accB);
  else
    dual_bn_bwd_apply_kernel<false><<<grid_a, g.block, 0, st>>>(d, g_dy, dy_ld, coefA, coefB, static_cast<__nv_bfloat16*>(dA),
                                                                dA_ld, accA, static_cast<__nv_bfloat16*>(dB), dB_ld, accB);
  UZ_CHECK_LAUNCH("dual_bn_bwd_apply launch");
  return UZ_OK;
}

extern "C" size_t uz_colsum_workspace_bytes(int C) {
  const int rows = uz_reduce_rows();
  return rows < 0 ? 0 : sizeof(float) * static_cast<size_t>(rows) * C;
}

extern "C" int uz_colsum(const void* x, int ld, long long npix, int C, float* out, int accumulate, void* workspace,
                         size_t workspace_bytes, void* stream) {
  UZ_REQUIRE_DEVICE();
  int rc = check_act("colsum", x, C, ld);
  if (rc) return rc;
  if (workspace_bytes < uz_colsum_workspace_bytes(C)) return set_error(UZ_ERR_WORKSPACE, "colsum: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const Geo g = make_geo(C, npix, 4);
  colsum_kernel<<<g.grid, g.block, g.block * 8 * sizeof(float), st>>>(static_cast<const __nv_bfloat16*>(x), ld, npix, C,
                                                                      g.cv, g.ppb, static_cast<float*>(workspace));
  UZ_CHECK_LAUNCH("colsum launch");
  reduce_rows_kernel<<<(C + 3) / 4, 128, 0, st>>>(static_cast<const float*>(workspace), g.grid, C, out, accumulate);
  UZ_CHECK_LAUNCH("colsum reduce launch");
  return UZ_OK;
}

// ---------------------------------------------------------------------------------- uz_colsum without a workspace
// Same two kernels as uz_colsum, with the per-block partial sums in a file-scope device array instead of a caller
// buffer: for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
// Every device that loads the library holds the array (kNoAllocRows x kNoAllocMaxC fp32 = 2.5 MiB) until exit.
// All calls on a device share it; colsum_kernel writes every row reduce_rows_kernel reads (rows = the launch's grid),
// so nothing has to start at zero, and the previous call's contents are never read.
namespace {
constexpr int kNoAllocMaxC = 1024;   // widest channel count of the hot-path models (bottle_neck)
constexpr int kNoAllocRows = 640;    // >= 4 blocks per SM on 148 SMs (592); larger devices get a clamped grid
static __device__ __align__(256) float g_colsum_scratch[static_cast<size_t>(kNoAllocRows) * kNoAllocMaxC];
float* g_colsum_scratch_ptr[uz::kMaxDevices] = {};   // device address of the array above, per device (host-side cache)

// `g_colsum_scratch` in HOST code is only a handle: its device address comes from cudaGetSymbolAddress, once per device,
// with that device current.
int colsum_scratch(float** out) {
  const int slot = uz::device_slot();
  if (!g_colsum_scratch_ptr[slot]) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_scratch);
    if (e != cudaSuccess) return uz::cuda_error(e, "colsum_static: cudaGetSymbolAddress");
    if (!p) return uz::set_error(UZ_ERR_CUDA, "colsum_static: null scratch address");
    g_colsum_scratch_ptr[slot] = p;
  }
  *out = g_colsum_scratch_ptr[slot];
  return UZ_OK;
}
}  // namespace

extern "C" int uz_colsum_static_max_channels(void) { return kNoAllocMaxC; }

extern "C" int uz_colsum_static(const void* x, int ld, long long npix, int C, float* out, int accumulate, void* stream) {
  UZ_REQUIRE_DEVICE();
  int rc = check_act("colsum_static", x, C, ld);
  if (rc) return rc;
  if (!out || npix < 1) return set_error(UZ_ERR_INVALID, "colsum_static: bad arguments");
  if (C > kNoAllocMaxC)
    return set_error(UZ_ERR_WORKSPACE, "colsum_static: C=%d exceeds the built-in scratch (max %d channels); use uz_colsum",
                     C, kNoAllocMaxC);
  float* scratch = nullptr;
  rc = colsum_scratch(&scratch);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  Geo g = make_geo(C, npix, 4);
  if (g.grid > kNoAllocRows) g.grid = kNoAllocRows;   // grid-stride kernel: any grid covers all pixels
  colsum_kernel<<<g.grid, g.block, g.block * 8 * sizeof(float), st>>>(static_cast<const __nv_bfloat16*>(x), ld, npix, C,
                                                                      g.cv, g.ppb, scratch);
  UZ_CHECK_LAUNCH("colsum_static launch");
  reduce_rows_kernel<<<(C + 3) / 4, 128, 0, st>>>(scratch, g.grid, C, out, accumulate);
  UZ_CHECK_LAUNCH("colsum_static reduce launch");
  return UZ_OK;
}
