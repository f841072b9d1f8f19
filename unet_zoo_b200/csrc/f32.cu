// fp32 parity mode of the hot path (BASELINE.json north_star: "fp32 mode within 1e-4 relative on logits";
// configs[0]: unet batch 1, 256x256 fp32 forward).  Forward only.
//
// The tensor-core kernels compute in bf16 with fp32 accumulation and cannot meet 1e-4; this translation unit carries
// the same operator set on NHWC fp32 activations with CUDA-core FFMA arithmetic: one generic implicit-GEMM
// convolution (3x3 / 1x1, stride 1 / 2, virtual channel concat, nearest-x2 gather, ConvTranspose2d k2s2 phases,
// bias / ReLU / accumulate epilogue, channel-slice output) and the bandwidth passes around it (BatchNorm statistics
// with fp64 accumulators, affine + ReLU + residual + 2x2 max-pool, bilinear x2, attention-gate apply, layout
// changes).  The reference call sites are the same as for the bf16 entry points (include/unet_zoo_b200.h).
//
// Tiling of the convolution: 256 threads compute 64 pixels x 64 output channels, K in chunks of 16 input channels
// of one tap; every thread owns a 4x4 register tile (16 FFMA per two 16-byte shared-memory loads).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/unet_zoo_b200.h"
#include "host_util.h"

namespace uz {
namespace {

constexpr int kTP = 64;    // pixels per CTA
constexpr int kTN = 64;    // output channels per CTA
constexpr int kTK = 16;    // input channels per K chunk
constexpr int kLdA = 68;   // padded row of the transposed A tile (floats; 16-byte multiple)

struct F32Seg {
  const float* p;
  int C;
  int ld;
};

struct F32Conv {
  F32Seg seg[UZ_F32_MAX_SEGS];
  int nseg, cin_total;
  int N, H, W;     // input grid (before the nearest-x2 gather)
  int Ho, Wo;      // grid the CTA pixels enumerate (ConvTranspose2d: the INPUT grid, outputs are its 2x2 phases)
  int taps, stride, up2x, convt;
  const float* w;  // [taps or 4 phases][cin_total][coutp]
  int cout, coutp;
  const float* bias;
  float* out;
  int out_ld;
  int relu, accumulate;
};

__global__ void __launch_bounds__(256) f32_conv_kernel(const F32Conv a) {
  __shared__ __align__(16) float As[kTK][kLdA];
  __shared__ __align__(16) float Bs[kTK][kTN];
  const int tid = threadIdx.x;
  const long long P = static_cast<long long>(a.N) * a.Ho * a.Wo;
  const long long p0 = static_cast<long long>(blockIdx.x) * kTP;
  const int n0 = blockIdx.y * kTN;
  const int phase = blockIdx.z;

  // loader role: pixel li, channel quad lq of the chunk
  const int li = tid >> 2, lq = tid & 3;
  const long long lp = p0 + li;
  const bool lvalid = lp < P;
  int ln = 0, loy = 0, lox = 0;
  if (lvalid) {
    lox = static_cast<int>(lp % a.Wo);
    const long long r = lp / a.Wo;
    loy = static_cast<int>(r % a.Ho);
    ln = static_cast<int>(r / a.Ho);
  }
  const int brow = tid >> 4, bcol = (tid & 15) * 4;
  const int ty = tid >> 4, tx = tid & 15;
  const int pad = a.taps == 9 ? 1 : 0;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int t = 0; t < a.taps; ++t) {
    const int dy = a.taps == 9 ? t / 3 : 0, dx = a.taps == 9 ? t % 3 : 0;
    int iy, ix;
    bool inb = lvalid;
    if (a.up2x) {
      const int uy = loy + dy - 1, ux = lox + dx - 1;
      inb = inb && uy >= 0 && uy < 2 * a.H && ux >= 0 && ux < 2 * a.W;
      iy = uy >> 1;
      ix = ux >> 1;
    } else {
      iy = loy * a.stride + dy - pad;
      ix = lox * a.stride + dx - pad;
      inb = inb && iy >= 0 && iy < a.H && ix >= 0 && ix < a.W;
    }
    const long long ipix = (static_cast<long long>(ln) * a.H + iy) * a.W + ix;
    const int wt = a.convt ? phase : t;
    int coff = 0;
    for (int s = 0; s < a.nseg; ++s) {
      const F32Seg sg = a.seg[s];
      const float* src = sg.p + ipix * sg.ld;
      const bool vec = ((sg.ld & 3) == 0) && ((reinterpret_cast<uintptr_t>(sg.p) & 15) == 0);
      for (int c0 = 0; c0 < sg.C; c0 += kTK) {
        // A tile: 64 pixels x 16 channels, stored transposed [k][pixel]
        const int c = c0 + lq * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (inb) {
          if (vec && c + 3 < sg.C) {
            v = __ldg(reinterpret_cast<const float4*>(src + c));
          } else {
            if (c < sg.C) v.x = __ldg(src + c);
            if (c + 1 < sg.C) v.y = __ldg(src + c + 1);
            if (c + 2 < sg.C) v.z = __ldg(src + c + 2);
            if (c + 3 < sg.C) v.w = __ldg(src + c + 3);
          }
        }
        // B tile: 16 input channels x 64 output channels
        float4 wv = make_float4(0.f, 0.f, 0.f, 0.f);
        {
          const int ck = c0 + brow;
          const int co = n0 + bcol;
          if (ck < sg.C && co < a.coutp)
            wv = __ldg(reinterpret_cast<const float4*>(
                a.w + (static_cast<long long>(wt) * a.cin_total + coff + ck) * a.coutp + co));
        }
        __syncthreads();   // previous chunk fully consumed
        As[lq * 4 + 0][li] = v.x;
        As[lq * 4 + 1][li] = v.y;
        As[lq * 4 + 2][li] = v.z;
        As[lq * 4 + 3][li] = v.w;
        *reinterpret_cast<float4*>(&Bs[brow][bcol]) = wv;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kTK; ++k) {
          const float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
          const float4 bv = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
          const float ar[4] = {av.x, av.y, av.z, av.w};
          const float br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
        }
      }
      coff += sg.C;
    }
  }

  // epilogue
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long p = p0 + ty * 4 + i;
    if (p >= P) continue;
    long long opix = p;
    if (a.convt) {
      const int ox = static_cast<int>(p % a.Wo);
      const long long r = p / a.Wo;
      const int oy = static_cast<int>(r % a.Ho);
      const long long n = r / a.Ho;
      opix = (n * (2 * a.Ho) + (2 * oy + (phase >> 1))) * (2 * a.Wo) + (2 * ox + (phase & 1));
    }
    float* o = a.out + opix * a.out_ld;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co >= a.cout) continue;
      float v = acc[i][j];
      if (a.bias) v += __ldg(a.bias + co);
      if (a.accumulate) v += o[co];
      if (a.relu) v = fmaxf(v, 0.f);
      o[co] = v;
    }
  }
}

// dst[t][ci][co] (cin_total x coutp, zero padded) = src[co*s_co + ci*s_ci + t] * row_scale[co]
__global__ void f32_pack_weight_kernel(const float* __restrict__ src, float* __restrict__ dst, int T, int cin,
                                       int cin_total, int cout, int coutp, long long s_co, long long s_ci,
                                       const float* __restrict__ row_scale) {
  const long long total = static_cast<long long>(T) * cin_total * coutp;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int co = static_cast<int>(i % coutp);
    const long long r = i / coutp;
    const int ci = static_cast<int>(r % cin_total);
    const int t = static_cast<int>(r / cin_total);
    float v = 0.f;
    if (co < cout && ci < cin) {
      v = __ldg(src + co * s_co + ci * s_ci + t);
      if (row_scale) v *= __ldg(row_scale + co);
    }
    dst[i] = v;
  }
}

__global__ void f32_nchw_to_nhwc_kernel(const float* __restrict__ x, int N, int C, int HW, float* __restrict__ out,
                                        int ld) {
  const long long total = static_cast<long long>(N) * HW * ld;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % ld);
    const long long p = i / ld;
    const long long n = p / HW, hw = p % HW;
    out[i] = c < C ? __ldg(x + (n * C + c) * HW + hw) : 0.f;
  }
}

__global__ void f32_nhwc_to_nchw_kernel(const float* __restrict__ in, int ld, int N, int C, int HW,
                                        float* __restrict__ out) {
  const long long total = static_cast<long long>(N) * C * HW;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long hw = i % HW;
    const long long r = i / HW;
    const int c = static_cast<int>(r % C);
    const long long n = r / C;
    out[i] = __ldg(in + (n * HW + hw) * ld + c);
  }
}

// partial[row][0][c] = sum, partial[row][1][c] = sum of squares over the row's pixel range (fp64 accumulation)
__global__ void __launch_bounds__(256) f32_bn_stats_kernel(const float* __restrict__ x, int ld, long long npix, int C,
                                                           int cw, float* __restrict__ partial) {
  __shared__ double ss[256], sq[256];
  const int rows = gridDim.x;
  const long long per = (npix + rows - 1) / rows;
  const long long lo = blockIdx.x * per;
  long long hi = lo + per;
  if (hi > npix) hi = npix;
  const int cl = threadIdx.x % cw, pl = threadIdx.x / cw, npl = 256 / cw;
  float* prow = partial + static_cast<long long>(blockIdx.x) * 2 * C;
  for (int cb = 0; cb < C; cb += cw) {
    const int c = cb + cl;
    double s = 0.0, q = 0.0;
    if (c < C) {
      for (long long p = lo + pl; p < hi; p += npl) {
        const double v = static_cast<double>(__ldg(x + p * ld + c));
        s += v;
        q += v * v;
      }
    }
    ss[threadIdx.x] = s;
    sq[threadIdx.x] = q;
    __syncthreads();
    if (pl == 0 && c < C) {
      for (int k = 1; k < npl; ++k) {
        s += ss[k * cw + cl];
        q += sq[k * cw + cl];
      }
      prow[c] = static_cast<float>(s);
      prow[C + c] = static_cast<float>(q);
    }
    __syncthreads();
  }
}

__device__ __forceinline__ float f32_act(float z, float sc, float sh, float r, int relu) {
  float v = fmaf(z, sc, sh) + r;
  return relu ? fmaxf(v, 0.f) : v;
}

// y = act(z*scale + shift + residual); POOL: items are 2x2 quads, pooled = max of the four y (MaxPool2d(2,2))
template <bool POOL>
__global__ void f32_affine_act_kernel(const float* __restrict__ z, int z_ld, int N, int H, int W, int C,
                                      const float* __restrict__ scale, const float* __restrict__ shift, int relu,
                                      const float* __restrict__ res, int res_ld, float* __restrict__ y, int y_ld,
                                      float* __restrict__ pooled, int pooled_ld) {
  // POOL: quads incl. the partial ones of an odd last row / column (written to y, part of no pooling window)
  const int Hi = POOL ? (H + 1) / 2 : H, Wi = POOL ? (W + 1) / 2 : W;
  const int Hp = H / 2, Wp = W / 2;
  const long long total = static_cast<long long>(N) * Hi * Wi * C;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    const long long p = i / C;
    const float sc = scale ? __ldg(scale + c) : 1.f;
    const float sh = shift ? __ldg(shift + c) : 0.f;
    if (!POOL) {
      const float r = res ? __ldg(res + p * res_ld + c) : 0.f;
      y[p * y_ld + c] = f32_act(__ldg(z + p * z_ld + c), sc, sh, r, relu);
    } else {
      const int px = static_cast<int>(p % Wi);
      const long long rr = p / Wi;
      const int py = static_cast<int>(rr % Hi);
      const long long n = rr / Hi;
      float m = 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (2 * py + (k >> 1) >= H || 2 * px + (k & 1) >= W) continue;
        const long long q = (n * H + 2 * py + (k >> 1)) * W + 2 * px + (k & 1);
        const float v = f32_act(__ldg(z + q * z_ld + c), sc, sh, 0.f, relu);
        if (y) y[q * y_ld + c] = v;
        m = (k == 0 || v > m) ? v : m;
      }
      if (py < Hp && px < Wp) pooled[((n * Hp + py) * Wp + px) * pooled_ld + c] = m;
    }
  }
}

// nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True) - ATen index arithmetic (see resample.cu)
__device__ __forceinline__ void f32_bl_src(int dst, float scale, int in, int* i0, int* i1, float* l0, float* l1) {
  const float src = scale * static_cast<float>(dst);
  const int a = static_cast<int>(src);
  *i0 = a;
  *i1 = a + (a < in - 1 ? 1 : 0);
  *l1 = src - static_cast<float>(a);
  *l0 = 1.f - *l1;
}

__global__ void f32_bilinear2x_kernel(const float* __restrict__ x, int x_ld, int N, int H, int W, int C,
                                      float* __restrict__ out, int out_ld, float sy, float sx) {
  const int Ho = 2 * H, Wo = 2 * W;
  const long long total = static_cast<long long>(N) * Ho * Wo * C;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    long long r = i / C;
    const int ox = static_cast<int>(r % Wo);
    r /= Wo;
    const int oy = static_cast<int>(r % Ho);
    const long long n = r / Ho;
    int y0, y1, x0, x1;
    float ly0, ly1, lx0, lx1;
    f32_bl_src(oy, sy, H, &y0, &y1, &ly0, &ly1);
    f32_bl_src(ox, sx, W, &x0, &x1, &lx0, &lx1);
    const float* base = x + n * H * W * x_ld + c;
    const float v00 = __ldg(base + (static_cast<long long>(y0) * W + x0) * x_ld);
    const float v01 = __ldg(base + (static_cast<long long>(y0) * W + x1) * x_ld);
    const float v10 = __ldg(base + (static_cast<long long>(y1) * W + x0) * x_ld);
    const float v11 = __ldg(base + (static_cast<long long>(y1) * W + x1) * x_ld);
    out[((n * Ho + oy) * Wo + ox) * out_ld + c] = ly0 * (lx0 * v00 + lx1 * v01) + ly1 * (lx0 * v10 + lx1 * v11);
  }
}

// out = x * sigmoid(psi_raw*scale + shift)   (AttentionBlock, attention_unet.py:38-40; bn = {mean, invstd, scale, shift})
__global__ void f32_gate_apply_kernel(const float* __restrict__ x, int x_ld, const float* __restrict__ psi_raw,
                                      const float* __restrict__ bn, long long npix, int C, float* __restrict__ out,
                                      int out_ld) {
  const float sc = __ldg(bn + 2), sh = __ldg(bn + 3);
  const long long total = npix * C;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    const long long p = i / C;
    const float v = fmaf(__ldg(psi_raw + p), sc, sh);
    const float s = 1.f / (1.f + expf(-v));
    out[p * out_ld + c] = __ldg(x + p * x_ld + c) * s;
  }
}

int ew_grid(long long total) {
  const int sms = sm_count();
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(sms > 0 ? sms : 148) * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

#define F32_REQUIRE_DEVICE()               \
  do {                                     \
    const int sms_ = uz_device_sm_count(); \
    if (sms_ < 0) return sms_;             \
  } while (0)
#define F32_CHECK_LAUNCH(what)                              \
  do {                                                      \
    cudaError_t e_ = cudaGetLastError();                    \
    if (e_ != cudaSuccess) return uz::cuda_error(e_, what); \
  } while (0)

}  // namespace
}  // namespace uz

using namespace uz;

extern "C" int uz_f32_pack_weight(const float* src, float* dst, int T, int cin, int cin_total, int cout, int coutp,
                                  long long s_co, long long s_ci, const float* row_scale, void* stream) {
  F32_REQUIRE_DEVICE();
  if (!src || !dst || T < 1 || cin < 1 || cin_total < cin || cout < 1 || coutp < cout || (coutp & 3))
    return set_error(UZ_ERR_INVALID, "f32_pack_weight: bad arguments (T=%d cin=%d/%d cout=%d/%d)", T, cin, cin_total,
                     cout, coutp);
  const long long total = static_cast<long long>(T) * cin_total * coutp;
  f32_pack_weight_kernel<<<ew_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, dst, T, cin, cin_total, cout, coutp, s_co, s_ci, row_scale);
  F32_CHECK_LAUNCH("f32_pack_weight launch");
  return UZ_OK;
}

extern "C" int uz_f32_conv(int mode, int N, int H, int W, int nseg, const void* const* seg_ptr, const int* seg_c,
                           const int* seg_ld, const float* w_packed, int cout, int coutp, const float* bias, float* out,
                           int out_ld, int flags, void* stream) {
  F32_REQUIRE_DEVICE();
  if (nseg < 1 || nseg > UZ_F32_MAX_SEGS) return set_error(UZ_ERR_INVALID, "f32_conv: 1..%d segments", UZ_F32_MAX_SEGS);
  if (!w_packed || !out || N < 1 || H < 1 || W < 1 || cout < 1 || coutp < cout || (coutp & 3) || out_ld < 1)
    return set_error(UZ_ERR_INVALID, "f32_conv: bad arguments");
  if (flags & UZ_FLAG_STATS) return set_error(UZ_ERR_UNSUPPORTED, "f32_conv: statistics come from uz_f32_bn_stats");
  F32Conv a{};
  a.nseg = nseg;
  a.cin_total = 0;
  for (int s = 0; s < nseg; ++s) {
    if (!seg_ptr[s] || seg_c[s] < 1 || seg_ld[s] < seg_c[s])
      return set_error(UZ_ERR_INVALID, "f32_conv: bad segment %d (C=%d ld=%d)", s, seg_c[s], seg_ld[s]);
    a.seg[s] = {static_cast<const float*>(seg_ptr[s]), seg_c[s], seg_ld[s]};
    a.cin_total += seg_c[s];
  }
  a.N = N; a.H = H; a.W = W;
  a.Ho = H; a.Wo = W;
  a.taps = 9; a.stride = 1; a.up2x = 0; a.convt = 0;
  int phases = 1;
  switch (mode) {
    case UZ_CONV3X3: break;
    case UZ_CONV1X1: a.taps = 1; break;
    case UZ_CONV3X3_S2: a.stride = 2; a.Ho = (H - 1) / 2 + 1; a.Wo = (W - 1) / 2 + 1; break;
    case UZ_CONV1X1_S2: a.taps = 1; a.stride = 2; a.Ho = (H - 1) / 2 + 1; a.Wo = (W - 1) / 2 + 1; break;
    case UZ_UP2X_CONV3X3: a.up2x = 1; a.Ho = 2 * H; a.Wo = 2 * W; break;
    case UZ_CONVT2X2: a.taps = 1; a.convt = 1; phases = 4; break;
    default: return set_error(UZ_ERR_UNSUPPORTED, "f32_conv: mode %d has no fp32 (forward-only) kernel", mode);
  }
  a.w = w_packed; a.cout = cout; a.coutp = coutp; a.bias = bias; a.out = out; a.out_ld = out_ld;
  a.relu = (flags & UZ_FLAG_RELU) ? 1 : 0;
  a.accumulate = (flags & UZ_FLAG_ACCUMULATE) ? 1 : 0;
  const long long P = static_cast<long long>(N) * a.Ho * a.Wo;
  const long long gx = (P + kTP - 1) / kTP;
  if (gx > 0x7fffffffLL) return set_error(UZ_ERR_UNSUPPORTED, "f32_conv: too many pixels");
  dim3 grid(static_cast<unsigned>(gx), static_cast<unsigned>((cout + kTN - 1) / kTN), phases);
  f32_conv_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(a);
  F32_CHECK_LAUNCH("f32_conv launch");
  return UZ_OK;
}

extern "C" int uz_f32_nchw_to_nhwc(const float* x, int N, int C, int H, int W, float* out, int out_ld, void* stream) {
  F32_REQUIRE_DEVICE();
  if (!x || !out || out_ld < C) return set_error(UZ_ERR_INVALID, "f32_nchw_to_nhwc: bad arguments");
  const long long total = static_cast<long long>(N) * H * W * out_ld;
  f32_nchw_to_nhwc_kernel<<<ew_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, N, C, H * W, out, out_ld);
  F32_CHECK_LAUNCH("f32_nchw_to_nhwc launch");
  return UZ_OK;
}

extern "C" int uz_f32_nhwc_to_nchw(const float* in, int in_ld, int N, int C, int H, int W, float* out, void* stream) {
  F32_REQUIRE_DEVICE();
  if (!in || !out || in_ld < C) return set_error(UZ_ERR_INVALID, "f32_nhwc_to_nchw: bad arguments");
  const long long total = static_cast<long long>(N) * H * W * C;
  f32_nhwc_to_nchw_kernel<<<ew_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(in, in_ld, N, C, H * W, out);
  F32_CHECK_LAUNCH("f32_nhwc_to_nchw launch");
  return UZ_OK;
}

extern "C" int uz_f32_bn_stats(const float* x, int ld, long long npix, int C, float* partial, void* stream) {
  F32_REQUIRE_DEVICE();
  if (!x || !partial || C < 1 || ld < C || npix < 1) return set_error(UZ_ERR_INVALID, "f32_bn_stats: bad arguments");
  const int rows = uz_reduce_rows();
  int cw = 1;
  while (cw < C && cw < 256) cw <<= 1;
  f32_bn_stats_kernel<<<rows, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, ld, npix, C, cw, partial);
  F32_CHECK_LAUNCH("f32_bn_stats launch");
  return UZ_OK;
}

extern "C" int uz_f32_affine_act(const float* z, int z_ld, int N, int H, int W, int C, const float* scale,
                                 const float* shift, int relu, const float* residual, int res_ld, float* y, int y_ld,
                                 float* pooled, int pooled_ld, void* stream) {
  F32_REQUIRE_DEVICE();
  if (!z || C < 1 || z_ld < C || (!y && !pooled)) return set_error(UZ_ERR_INVALID, "f32_affine_act: bad arguments");
  if ((scale == nullptr) != (shift == nullptr))
    return set_error(UZ_ERR_INVALID, "f32_affine_act: scale/shift must both be given");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (pooled) {
    if (H < 2 || W < 2) return set_error(UZ_ERR_INVALID, "f32_affine_act: pooling needs H, W >= 2");
    if (residual) return set_error(UZ_ERR_UNSUPPORTED, "f32_affine_act: residual + pool not supported");
    const long long total = static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2) * C;
    f32_affine_act_kernel<true><<<ew_grid(total), 256, 0, st>>>(z, z_ld, N, H, W, C, scale, shift, relu, nullptr, 0, y,
                                                                y_ld, pooled, pooled_ld);
  } else {
    const long long total = static_cast<long long>(N) * H * W * C;
    f32_affine_act_kernel<false><<<ew_grid(total), 256, 0, st>>>(z, z_ld, N, H, W, C, scale, shift, relu, residual,
                                                                 res_ld, y, y_ld, nullptr, 0);
  }
  F32_CHECK_LAUNCH("f32_affine_act launch");
  return UZ_OK;
}

extern "C" int uz_f32_bilinear2x(const float* x, int x_ld, int N, int H, int W, int C, float* out, int out_ld,
                                 void* stream) {
  F32_REQUIRE_DEVICE();
  if (!x || !out || C < 1 || x_ld < C || out_ld < C) return set_error(UZ_ERR_INVALID, "f32_bilinear2x: bad arguments");
  const float sy = H > 1 ? static_cast<float>(H - 1) / static_cast<float>(2 * H - 1) : 0.f;
  const float sx = W > 1 ? static_cast<float>(W - 1) / static_cast<float>(2 * W - 1) : 0.f;
  const long long total = static_cast<long long>(N) * 4 * H * W * C;
  f32_bilinear2x_kernel<<<ew_grid(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, x_ld, N, H, W, C, out, out_ld,
                                                                                      sy, sx);
  F32_CHECK_LAUNCH("f32_bilinear2x launch");
  return UZ_OK;
}

extern "C" int uz_f32_gate_apply(const float* x, int x_ld, const float* psi_raw, const float* bn_psi, long long npix,
                                 int C, float* out, int out_ld, void* stream) {
  F32_REQUIRE_DEVICE();
  if (!x || !psi_raw || !bn_psi || !out || C < 1) return set_error(UZ_ERR_INVALID, "f32_gate_apply: bad arguments");
  f32_gate_apply_kernel<<<ew_grid(npix * C), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, x_ld, psi_raw, bn_psi, npix,
                                                                                         C, out, out_ld);
  F32_CHECK_LAUNCH("f32_gate_apply launch");
  return UZ_OK;
}
