// Host side of the tcgen05 convolution entry points: argument checking, TMA tensor-map construction,
// tap / phase tables per convolution flavour, grid sizing, launch.  C ABI declared in include/unet_zoo_b200.h.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/unet_zoo_b200.h"
#include "host_util.h"
#include "igemm.cuh"

namespace uz {
cudaError_t launch_igemm_fwd(const IgemmParams& p, int bn, int grid, cudaStream_t stream);
cudaError_t launch_igemm_fwd2(const Igemm2Params& p, int bn, int grid, int smem_bytes, cudaStream_t stream);
cudaError_t launch_igemm_wgrad(const WgradParams& p, int grid, cudaStream_t stream);
cudaError_t launch_igemm_wgrad2(const Wgrad2Params& p, int grid, cudaStream_t stream);
cudaError_t launch_wgrad2_reduce(const Wgrad2Params& p, float* dst, long long s_row, int accumulate, cudaStream_t stream);
cudaError_t launch_bmm_tn_small(const void* A, long long lda, long long a_img, const void* B, long long ldb,
                                long long b_img, void* C, long long ldc, long long c_img, int N, int M, int K, int J,
                                cudaStream_t st);
cudaError_t launch_wgrad_reduce(const float* partial, float* dst, int splits, int rows, int taps, int cols,
                                long long s_row, long long s_col, long long s_tap, int accumulate,
                                cudaStream_t stream);

// ------------------------------------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";
int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
int cuda_error(cudaError_t e, const char* what) {
  return set_error(UZ_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

int sm_count() {
  static int cached[kMaxDevices] = {0};   // per device: a process may drive several GPUs (cuda:0 then cuda:1)
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  const bool slot_ok = dev >= 0 && dev < kMaxDevices;
  if (slot_ok && cached[dev] != 0) return cached[dev];
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  int major = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (major != 10) n = -2;
  if (slot_ok) cached[dev] = n;
  return n;
}

// ------------------------------------------------------------------------------------------------ tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

struct Lattice {
  int oy, ox, step;
};

// 4-D (C, W, H, N) bf16 view of an NHWC activation (pixel stride ld elements), restricted to the pixel lattice
// (oy + step*i, ox + step*j); box = (64 channels, tw, th, tn), 128-byte swizzle, zero fill outside.
static int make_act_map(CUtensorMap* m, const void* ptr, int C, int ld, int N, int H, int W, Lattice lat, int tw,
                        int th, int tn, int box_c = 64) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return set_error(UZ_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  const int Hl = (H - lat.oy + lat.step - 1) / lat.step;
  const int Wl = (W - lat.ox + lat.step - 1) / lat.step;
  if (Hl < 1 || Wl < 1) return set_error(UZ_ERR_INVALID, "empty activation lattice (H=%d W=%d)", H, W);
  const char* base = static_cast<const char*>(ptr) + (static_cast<size_t>(lat.oy) * W + lat.ox) * ld * 2;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)Wl, (cuuint64_t)Hl, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)ld * 2 * lat.step, (cuuint64_t)W * ld * 2 * lat.step,
                           (cuuint64_t)H * W * ld * 2};
  cuuint32_t box[4] = {(cuuint32_t)box_c, (cuuint32_t)tw, (cuuint32_t)th, (cuuint32_t)tn};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  // L2 promotion widens every 128-byte request to 256 bytes.  On a channel SLICE of a wider buffer the other 128 bytes
  // belong to a different tensor (the skip half of a concat buffer): measured 2x DRAM reads for the transposed
  // convolution's data / weight gradient (profiles/r02_g_unet_per_launch.md).  Promote only where the neighbour is ours.
  const bool whole_rows = C == ld || ((C * 2) % 256 == 0 && (reinterpret_cast<uintptr_t>(base) & 255u) == 0);
  static const int promo_env = getenv("UZ_TMA_PROMO") ? atoi(getenv("UZ_TMA_PROMO")) : -1;
  const CUtensorMapL2promotion promo = (promo_env == 256 || (promo_env < 0 && whole_rows))
                                           ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<char*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, box_c == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                  promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_error(UZ_ERR_CUDA, "cuTensorMapEncodeTiled(activation C=%d ld=%d N=%d H=%d W=%d) failed: %d", C, ld,
                     N, H, W, (int)r);
  return UZ_OK;
}

// 3-D (K, taps, rows) bf16 packed weights, box (64, 1, bn).
static int make_weight_map(CUtensorMap* m, const void* ptr, int K, int taps, int rows, int bn) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return set_error(UZ_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)taps, (cuuint64_t)rows};
  cuuint64_t strides[2] = {(cuuint64_t)K * 2, (cuuint64_t)K * 2 * taps};
  cuuint32_t box[3] = {64, 1, (cuuint32_t)bn};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_error(UZ_ERR_CUDA, "cuTensorMapEncodeTiled(weights K=%d taps=%d rows=%d) failed: %d", K, taps, rows,
                     (int)r);
  return UZ_OK;
}

static int pow2_ceil(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}
static int ilog2(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}
// pick (tw, th, tn) with tw*th*tn == pixels, preferring 16-wide, square-ish patches inside one image
static void pick_tile(int pixels, int H, int W, int* tw, int* th, int* tn) {
  int w = pow2_ceil(W);
  if (w > 16) w = 16;
  if (w > pixels) w = pixels;
  int h = pow2_ceil(H);
  if (h > pixels / w) h = pixels / w;
  *tw = w;
  *th = h;
  *tn = pixels / (w * h);
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// row/column decomposition of the stride-2 3x3 conv: input row 2y + r - 1 lives on lattice parity `par`, index y + d
static void s2_tap(int r, int* par, int* d) {
  if (r == 0) {
    *par = 1;
    *d = -1;
  } else if (r == 1) {
    *par = 0;
    *d = 0;
  } else {
    *par = 1;
    *d = 0;
  }
}

struct ModeInfo {
  int Hm, Wm;        // M index space (per image)
  int Ho, Wo;        // output spatial size
  int oscale;
  int num_lat;
  Lattice lat[4];
  int num_phases;
  int ntaps[4];
  int btaps;         // tap extent of the packed weight tensor
};

static int build_mode(int mode, int H, int W, int out_h, int out_w, IgemmParams* p, ModeInfo* mi, int dil = 1) {
  memset(mi, 0, sizeof(*mi));
  mi->oscale = 1;
  mi->num_lat = 1;
  mi->lat[0] = {0, 0, 1};
  mi->num_phases = 1;
  mi->Hm = H;
  mi->Wm = W;
  mi->Ho = H;
  mi->Wo = W;
  for (int ph = 0; ph < kMaxPhases; ++ph) {
    p->phase_oy[ph] = 0;
    p->phase_ox[ph] = 0;
  }
  auto set_tap = [&](int ph, int i, int dy, int dx, int btap, int lat) {
    p->taps[ph].dy[i] = (int8_t)dy;
    p->taps[ph].dx[i] = (int8_t)dx;
    p->taps[ph].btap[i] = (int8_t)btap;
    p->taps[ph].lat[i] = (int8_t)lat;
  };
  switch (mode) {
    case UZ_CONV3X3:
    case UZ_CONV3X3_DGRAD: {
      const int sgn = (mode == UZ_CONV3X3) ? 1 : -1;
      for (int r = 0; r < 3; ++r)
        for (int s = 0; s < 3; ++s) set_tap(0, r * 3 + s, sgn * dil * (r - 1), sgn * dil * (s - 1), r * 3 + s, 0);
      mi->ntaps[0] = 9;
      mi->btaps = 9;
      break;
    }
    case UZ_CONV1X1:
      set_tap(0, 0, 0, 0, 0, 0);
      mi->ntaps[0] = 1;
      mi->btaps = 1;
      break;
    case UZ_CONV3X3_S2: {
      mi->Hm = mi->Ho = (H - 1) / 2 + 1;
      mi->Wm = mi->Wo = (W - 1) / 2 + 1;
      mi->num_lat = 4;
      for (int l = 0; l < 4; ++l) mi->lat[l] = {l >> 1, l & 1, 2};
      for (int r = 0; r < 3; ++r)
        for (int s = 0; s < 3; ++s) {
          int pr, dr, pc, dc;
          s2_tap(r, &pr, &dr);
          s2_tap(s, &pc, &dc);
          set_tap(0, r * 3 + s, dr, dc, r * 3 + s, pr * 2 + pc);
        }
      mi->ntaps[0] = 9;
      mi->btaps = 9;
      break;
    }
    case UZ_CONV3X3_S2_DGRAD: {
      // input = dY (H x W, low res); output dX (out_h x out_w) with (out_h-1)/2+1 == H
      if ((out_h - 1) / 2 + 1 != H || (out_w - 1) / 2 + 1 != W)
        return set_error(UZ_ERR_INVALID, "stride-2 dgrad: output %dx%d inconsistent with dY %dx%d", out_h, out_w, H,
                         W);
      mi->Ho = out_h;
      mi->Wo = out_w;
      mi->oscale = 2;
      mi->num_phases = 4;
      mi->btaps = 9;
      for (int ph = 0; ph < 4; ++ph) {
        const int py = ph >> 1, px = ph & 1;
        p->phase_oy[ph] = (int8_t)py;
        p->phase_ox[ph] = (int8_t)px;
        // rows: py==0 -> (r=1, dy=0); py==1 -> (r=0, dy=+1), (r=2, dy=0)
        int rr[2], rd[2], nr, cc[2], cd[2], nc;
        if (py == 0) { nr = 1; rr[0] = 1; rd[0] = 0; } else { nr = 2; rr[0] = 0; rd[0] = 1; rr[1] = 2; rd[1] = 0; }
        if (px == 0) { nc = 1; cc[0] = 1; cd[0] = 0; } else { nc = 2; cc[0] = 0; cd[0] = 1; cc[1] = 2; cd[1] = 0; }
        int i = 0;
        for (int a = 0; a < nr; ++a)
          for (int b = 0; b < nc; ++b) set_tap(ph, i++, rd[a], cd[b], rr[a] * 3 + cc[b], 0);
        mi->ntaps[ph] = i;
      }
      break;
    }
    case UZ_CONV1X1_S2:
      mi->Hm = mi->Ho = (H - 1) / 2 + 1;
      mi->Wm = mi->Wo = (W - 1) / 2 + 1;
      mi->lat[0] = {0, 0, 2};
      set_tap(0, 0, 0, 0, 0, 0);
      mi->ntaps[0] = 1;
      mi->btaps = 1;
      break;
    case UZ_CONV1X1_S2_DGRAD:
      // only even output pixels receive gradient; caller zero-fills / accumulates the rest
      if ((out_h - 1) / 2 + 1 != H || (out_w - 1) / 2 + 1 != W)
        return set_error(UZ_ERR_INVALID, "stride-2 1x1 dgrad: output %dx%d inconsistent with dY %dx%d", out_h, out_w,
                         H, W);
      mi->Ho = out_h;
      mi->Wo = out_w;
      mi->oscale = 2;
      set_tap(0, 0, 0, 0, 0, 0);
      mi->ntaps[0] = 1;
      mi->btaps = 1;
      break;
    case UZ_CONVT2X2:
      mi->Ho = 2 * H;
      mi->Wo = 2 * W;
      mi->oscale = 2;
      mi->num_phases = 4;
      mi->btaps = 4;
      for (int ph = 0; ph < 4; ++ph) {
        p->phase_oy[ph] = (int8_t)(ph >> 1);
        p->phase_ox[ph] = (int8_t)(ph & 1);
        set_tap(ph, 0, 0, 0, ph, 0);
        mi->ntaps[ph] = 1;
      }
      break;
    case UZ_CONVT2X2_DGRAD:
      if ((H & 1) || (W & 1)) return set_error(UZ_ERR_INVALID, "convT dgrad needs even dY size, got %dx%d", H, W);
      mi->Hm = mi->Ho = H / 2;
      mi->Wm = mi->Wo = W / 2;
      mi->num_lat = 4;
      for (int l = 0; l < 4; ++l) {
        mi->lat[l] = {l >> 1, l & 1, 2};
        set_tap(0, l, 0, 0, l, l);
      }
      mi->ntaps[0] = 4;
      mi->btaps = 4;
      break;
    case UZ_UP2X_CONV3X3:
      mi->Ho = 2 * H;
      mi->Wo = 2 * W;
      mi->oscale = 2;
      mi->num_phases = 4;
      mi->btaps = 9;
      for (int ph = 0; ph < 4; ++ph) {
        const int py = ph >> 1, px = ph & 1;
        p->phase_oy[ph] = (int8_t)py;
        p->phase_ox[ph] = (int8_t)px;
        for (int r = 0; r < 3; ++r)
          for (int s = 0; s < 3; ++s) {
            // source row of upsampled row (2y + py + r - 1) is floor((2y + py + r - 1) / 2) = y + floor((py+r-1)/2)
            const int dy = (py + r - 1 + 2) / 2 - 1;
            const int dx = (px + s - 1 + 2) / 2 - 1;
            set_tap(ph, r * 3 + s, dy, dx, r * 3 + s, 0);
          }
        mi->ntaps[ph] = 9;
      }
      break;
    case UZ_UP2X_CONV3X3_DGRAD: {
      // input = dZ of the up2x conv (H x W, high res); output = gradient of the LOW-res tensor (H/2 x W/2).
      // Forward tap (r,s) of output phase (py,px) reads low-res row y' + floor((py+r-1)/2) for output row 2y'+py, so
      // low-res pixel y collects dZ lattice (py,px) at index y - floor((py+r-1)/2): 36 taps over 4 strided views.
      if ((H & 1) || (W & 1)) return set_error(UZ_ERR_INVALID, "up2x dgrad needs even dZ size, got %dx%d", H, W);
      mi->Hm = mi->Ho = H / 2;
      mi->Wm = mi->Wo = W / 2;
      mi->num_lat = 4;
      int i = 0;
      for (int l = 0; l < 4; ++l) {
        const int py = l >> 1, px = l & 1;
        mi->lat[l] = {py, px, 2};
        for (int r = 0; r < 3; ++r)
          for (int sx = 0; sx < 3; ++sx) {
            const int dyr = (py + r - 1 + 2) / 2 - 1;
            const int dxs = (px + sx - 1 + 2) / 2 - 1;
            set_tap(0, i++, -dyr, -dxs, r * 3 + sx, l);
          }
      }
      mi->ntaps[0] = 36;
      mi->btaps = 9;
      break;
    }
    default:
      return set_error(UZ_ERR_INVALID, "unknown conv mode %d", mode);
  }
  if (out_h != mi->Ho || out_w != mi->Wo)
    return set_error(UZ_ERR_INVALID, "mode %d: expected output %dx%d, got %dx%d", mode, mi->Ho, mi->Wo, out_h, out_w);
  return UZ_OK;
}

}  // namespace uz

using namespace uz;

extern "C" const char* uz_last_error(void) { return g_err; }
extern "C" int uz_version(void) { return 100; }
extern "C" int uz_device_sm_count(void) {
  int n = sm_count();
  if (n == -2) return set_error(UZ_ERR_UNSUPPORTED, "current device is not compute capability 10.x (sm_100a required)");
  if (n < 0) return set_error(UZ_ERR_CUDA, "no CUDA device");
  return n;
}
// rows of the per-channel statistics partials: one per CTA of the persistent conv kernels
extern "C" int uz_conv_stats_rows(void) {
  const int n = uz_device_sm_count();
  return n < 0 ? n : 2 * n;   // two epilogue groups per CTA, each flushes its own row
}

namespace uz {
static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}

// Second-generation path (igemm_fwd2.cu).  Returns 1 when it handled the call, 0 when the first-generation kernel
// should run, negative on error.
static int try_fwd2(int mode, int N, int H, int W, int nseg, const void* const* seg_ptr, const int* seg_c,
                    const int* seg_ld, const void* w_packed, int cout, const float* bias, void* out, int out_ld,
                    float* stats, int flags, int sms, cudaStream_t st) {
  static const int disabled = env_int("UZ_DISABLE_FWD2", 0);
  if (disabled) return 0;
  if (!(mode == UZ_CONV3X3 || mode == UZ_CONV3X3_DGRAD || mode == UZ_CONV1X1 || mode == UZ_CONVT2X2 ||
        mode == UZ_CONVT2X2_DGRAD || mode == UZ_CONV1X1_S2 || mode == UZ_CONV3X3_S2 || mode == UZ_UP2X_CONV3X3 ||
        mode == UZ_UP2X_CONV3X3_DGRAD))
    return 0;
  // nearest-x2 upsample + conv3x3 (UpConvBlock, common_layers.py:69-73) in PHASE form: output pixel (2y+py, 2x+px) reads
  // rows y + floor((py + r - 1) / 2), r = 0..2, of the low-resolution input - two distinct rows per phase, likewise two
  // columns - so each of the four output phases is a 2x2 convolution of the low-resolution image whose weights are sums
  // of the 3x3 taps (uz_pack_up2x): 16 tap-pixels per low-resolution pixel instead of 36.  The packed weights carry
  // 16 "taps" = phase * 4 + a * 2 + b, tap (a, b) of phase (py, px) at low-resolution offset (a - 1 + py, b - 1 + px).
  const bool up2x = mode == UZ_UP2X_CONV3X3, up2x_d = mode == UZ_UP2X_CONV3X3_DGRAD;
  // (the caller packs 16-tap phase weights exactly when uz_up2x_phase_form() says so - same predicate here)
  if (up2x && (nseg != 1 || !uz_up2x_phase_form(H, W, cout))) return 0;
  if (up2x_d && (nseg != 1 || (H & 1) || (W & 1) || !uz_up2x_phase_form(H / 2, W / 2, seg_c[0]))) return 0;
  // 1x1 stride 2 (ResidualConv.conv_skip, common_layers.py:194) = a 1x1 GEMM over the even-pixel lattice of the input
  // 3x3 stride 2 (ResidualConv.conv_block, common_layers.py:188): the input splits into its four parity lattices; tap
  // (r, s) reads lattice (r != 1, s != 1) shifted by (r == 0 ? -1 : 0, s == 0 ? -1 : 0) - four segments over the SAME
  // channels, each with its own subset of the nine taps (1 + 2 + 2 + 4), one halo box per lattice and channel block
  const bool s2 = mode == UZ_CONV1X1_S2, s2c = mode == UZ_CONV3X3_S2;
  // UZ_FLAG_RES1X1 / _S2 (inference form of ResidualConv, common_layers.py:182-199): the LAST segment is the operand of the
  // block's 1x1 skip convolution - it contributes through the centre tap only (its other eight weight taps are never
  // read), and with _S2 it is the even-pixel lattice of a 2H x 2W tensor (the stride-2 skip of the encoder blocks).  The
  // residual sum then forms in ONE accumulator: no separate skip launch, no read-modify-write of the output.
  const bool res1 = (flags & (UZ_FLAG_RES1X1 | UZ_FLAG_RES1X1_S2)) != 0;
  if (res1 && (mode != UZ_CONV3X3 || nseg < 2)) return set_error(UZ_ERR_INVALID, "UZ_FLAG_RES1X1 needs UZ_CONV3X3 and >= 2 segments");
  const int Hin = H, Win = W;
  if (s2c && nseg != 1) return 0;
  if (s2 || s2c) {
    H = (H - 1) / 2 + 1;
    W = (W - 1) / 2 + 1;  // tile space = the output
  }
  // ConvTranspose2d k2s2 forward = four 1x1 GEMMs (one per output phase) stored through strided output views;
  // its data gradient = one 1x1 GEMM whose K runs over the four strided views of dY (K = 4*Cout).
  const bool convt = mode == UZ_CONVT2X2, convt_d = mode == UZ_CONVT2X2_DGRAD;
  if (convt_d || up2x_d) {
    if ((H & 1) || (W & 1) || nseg != 1) return 0;
    H /= 2;
    W /= 2;  // tile space = the low-resolution gradient
  }
  if (H < 16 || W < 8 || nseg > kMaxSegs) return 0;
  if ((convt || up2x) && (cout % 64) != 0) return 0;
  Igemm2Params p;
  memset(&p, 0, sizeof(p));
  const int pad = (mode == UZ_CONV3X3 || mode == UZ_CONV3X3_DGRAD || s2c || up2x || up2x_d) ? 0 + 1 : 0;
  p.pad = pad;
  p.halo_w = 8 + 2 * pad;
  p.halo_bytes = p.halo_w * (16 + 2 * pad) * 128;
  p.num_taps = up2x ? 4 : pad ? 9 : 1;
  if (up2x) {
    p.ph_ntaps = 4;
    for (int ph = 0; ph < 4; ++ph)
      for (int t = 0; t < 4; ++t)   // tap (a, b) = (t >> 1, t & 1) of phase (py, px): halo position (a + py, b + px)
        p.ph_row[ph][t] = (int8_t)(((t >> 1) + (ph >> 1)) * p.halo_w + (t & 1) + (ph & 1));
  }
  for (int t = 0; t < p.num_taps; ++t) {
    const int r = t / 3, s = t % 3;
    if (mode == UZ_CONV3X3 || up2x_d) p.tap_row[t] = r * p.halo_w + s;
    else if (mode == UZ_CONV3X3_DGRAD) p.tap_row[t] = (2 - r) * p.halo_w + (2 - s);
    else if (s2c) p.tap_row[t] = (r == 0 ? 0 : 1) * p.halo_w + (s == 0 ? 0 : 1);
    else p.tap_row[t] = 0;
    p.tap_b[t] = t;
  }
  int bn = cout <= 64 ? 64 : cout <= 128 ? 128 : 256;
  int ctot = 0, kblocks = 0;
  const Lattice full = {0, 0, 1};
  if (up2x_d) {
    // four segments = the four output phases of dZ (strided views with a halo), all over the SAME K range (the conv's
    // output channels); phase (py, px) contributes through its taps (a, b) at halo position (2 - a - py, 2 - b - px)
    for (int l = 0; l < 4; ++l) {
      p.seg_cblocks[l] = (seg_c[0] + 63) / 64;
      p.seg_k0[l] = 0;
      kblocks += p.seg_cblocks[l];
      int rc = make_act_map(&p.amap[l], seg_ptr[0], seg_c[0], seg_ld[0], N, 2 * H, 2 * W, Lattice{l >> 1, l & 1, 2}, p.halo_w,
                            16 + 2 * pad, 1);
      if (rc != UZ_OK) return rc;
      for (int t4 = 0; t4 < 4; ++t4) {
        const int hp = (2 - (t4 >> 1) - (l >> 1)) * 3 + (2 - (t4 & 1) - (l & 1));
        p.seg_taps[l] |= 1 << hp;
        p.seg_tap_b[l][hp] = (int8_t)(l * 4 + t4);
      }
    }
    ctot = seg_c[0];
    nseg = 4;
  } else if (convt_d) {
    for (int l = 0; l < 4; ++l) {
      p.seg_cblocks[l] = (seg_c[0] + 63) / 64;
      p.seg_k0[l] = ctot;
      ctot += seg_c[0];
      kblocks += p.seg_cblocks[l];
      int rc = make_act_map(&p.amap[l], seg_ptr[0], seg_c[0], seg_ld[0], N, 2 * H, 2 * W, Lattice{l >> 1, l & 1, 2}, 8, 16, 1);
      if (rc != UZ_OK) return rc;
    }
    nseg = 4;
  } else if (s2c) {
    for (int l = 0; l < 4; ++l) {
      p.seg_cblocks[l] = (seg_c[0] + 63) / 64;
      p.seg_k0[l] = 0;
      kblocks += p.seg_cblocks[l];
      int rc = make_act_map(&p.amap[l], seg_ptr[0], seg_c[0], seg_ld[0], N, Hin, Win, Lattice{l >> 1, l & 1, 2}, p.halo_w,
                            16 + 2 * pad, 1);
      if (rc != UZ_OK) return rc;
      for (int t = 0; t < 9; ++t)
        if (((t / 3 != 1) ? 1 : 0) == (l >> 1) && ((t % 3 != 1) ? 1 : 0) == (l & 1)) p.seg_taps[l] |= 1 << t;
    }
    ctot = seg_c[0];
    nseg = 4;
  } else {
    for (int s = 0; s < nseg; ++s) {
      p.seg_cblocks[s] = (seg_c[s] + 63) / 64;
      p.seg_k0[s] = ctot;
      ctot += seg_c[s];
      kblocks += p.seg_cblocks[s];
      int rc = s2 ? make_act_map(&p.amap[s], seg_ptr[s], seg_c[s], seg_ld[s], N, Hin, Win, Lattice{0, 0, 2}, 8, 16, 1)
               : (res1 && s == nseg - 1 && (flags & UZ_FLAG_RES1X1_S2))
                   ? make_act_map(&p.amap[s], seg_ptr[s], seg_c[s], seg_ld[s], N, 2 * H, 2 * W, Lattice{0, 0, 2}, p.halo_w,
                                  16 + 2 * pad, 1)
                   : make_act_map(&p.amap[s], seg_ptr[s], seg_c[s], seg_ld[s], N, H, W, full, p.halo_w, 16 + 2 * pad, 1);
      if (rc != UZ_OK) return rc;
    }
  }
  p.num_segs = nseg;
  p.kblocks = kblocks;
  for (int sg = 0; sg < nseg; ++sg) {
    const int c = (convt_d || s2c || up2x_d) ? seg_c[0] : seg_c[sg];
    const int rem = c - (p.seg_cblocks[sg] - 1) * 64;   // channels of the last block: 8..64
    p.seg_klast[sg] = env_int("UZ_FWD2_KSKIP", 1) ? (rem + 15) / 16 : 4;
  }
  p.b_items = 0;
  for (int sg = 0; sg < nseg; ++sg) {
    if (!s2c && !up2x_d) {
      p.seg_taps[sg] = (1 << p.num_taps) - 1;
      for (int t = 0; t < p.num_taps; ++t) p.seg_tap_b[sg][t] = (int8_t)p.tap_b[t];
    } else if (s2c) {
      for (int t = 0; t < 9; ++t) p.seg_tap_b[sg][t] = (int8_t)p.tap_b[t];
    }
    if (res1 && sg == nseg - 1) p.seg_taps[sg] = 1 << 4;   // centre tap only
    p.b_items += p.seg_cblocks[sg] * __builtin_popcount(p.seg_taps[sg]);
  }
  p.N = N; p.H = H; p.W = W;
  p.xb = (W + 7) / 8;
  p.yb = (H + 15) / 16;
  p.m_tiles_real = N * p.xb * p.yb;
  p.num_n_tiles = (cout + bn - 1) / bn;
  // ConvTranspose2d: all four output phases of a 64-column slice in ONE 256-wide item (Igemm2Params::ph_major) - the input
  // tile is loaded once per slice instead of once per (phase, N tile), and a 128->64 layer issues N = 256 MMAs instead
  // of four times N = 64
  // (measured, tools/bench_convt.py, batch 32: 128->64 @256^2 0.561 -> 0.331 ms = 0.74 of the HBM roof; 256->128 @128^2
  // 0.217 -> 0.234 ms and wider layers unchanged - their N = 128 / 256 tiles were fine already: 64 output channels only)
  const int ph_env = env_int("UZ_CONVT_PH_MAJOR", 1);
  const bool ph_major = convt && (ph_env == 2 || (ph_env == 1 && cout == 64));
  if (ph_major) {
    bn = 256;
    p.ph_major = 1;
    p.num_n_tiles = cout / 64;
  } else if (convt || up2x) {
    p.nt_per_phase = p.num_n_tiles;
    p.num_n_tiles *= 4;
  }
  // shared-memory plan: [A ring][B ring or resident weights][1-2 staging tiles][barriers + bias]
  const int kMaxSmem = 232448 - 1024;  // 227 KB minus the alignment slack
  int a_slot = (p.halo_bytes + 1023) & ~1023;
  // CTA pairs (cta_group::2): 1 = layers with <= 128 output channels (their N = 64 / 128 MMAs are bound by shared-memory
  // read bandwidth: 192 / 128 B/clk against the 128 B/clk port; the pair halves the B bytes per CTA), 2 = all layers
  const int pair_env = env_int("UZ_FWD2_PAIR", 1);
  // measured per layer (tools/bench_conv.py): the pair wins when a tile carries >= 18 (cblock, tap) steps (+4..12 %, up
  // to 1690 TFLOP/s); with fewer the tile is epilogue bound and pair mode adds ~1 us per tile pair
  const bool pair_auto = p.b_items >= env_int("UZ_FWD2_PAIR_MIN_STEPS", 18);
  p.pair = (pair_env == 2 || (pair_env == 1 && pair_auto)) && sms >= 2 && !ph_major ? 1 : 0;
  // 3x3 + 1x1 in one GEMM with 65..128 output channels: the 1x1 segment's halo boxes feed four MMAs each instead of 36, and
  // the pair's cross-SM handshake per box then dominates (384->128 @256^2: 0.916 ms as a pair, 0.615 ms as one CTA with
  // two pixel tiles per weight tile, 0.622 ms as two separate launches - tools/bench_res_tail.py)
  if (res1 && bn == 128 && !env_int("UZ_FWD2_RES_PAIR", 0)) p.pair = 0;
  int b_item = bn * 128 / (p.pair ? 2 : 1);     // bytes of one weight tile per CTA
  const int kTail = 2560;              // barriers (512) + one bias vector per epilogue group (2 x 1024)
  const long long b_all = (long long)p.b_items * b_item;
  int cluster = p.pair ? 2 : 1;
  bool stationary = false;
  if (p.num_n_tiles == 1 && !ph_major && !env_int("UZ_FWD2_NO_STATIONARY", 0)) {
    // weights resident for the whole launch, provided at least four halo stages still fit (TMA latency cover)
    for (int nbuf = 2; nbuf >= 2 && !stationary; --nbuf) {   // one staging tile per epilogue group
      const long long left = (long long)kMaxSmem - kTail - nbuf * 16384 - b_all;
      if (left >= (long long)env_int("UZ_FWD2_MIN_A", 2) * a_slot) {
        stationary = true;
        p.stg_bufs = nbuf;
        p.b_stages = 0;
        const int as = (int)(left / a_slot);
        p.a_stages = as > 6 ? 6 : as;
      }
    }
  }
  p.mp = 1;
  if (!stationary && !p.pair && bn == 128 && env_int("UZ_FWD2_MP", 2) == 2) {
    // streamed weights, 65..128 output channels: one 128-wide weight tile against TWO pixel tiles (4 x 128 TMEM columns
    // with double buffering).  Measured: 128->128 @256^2 0.496 -> 0.455 ms, 256->128 0.921 -> 0.854 ms.  Wider layers
    // stay on BN = 256 (one A read per 256 columns is cheaper in shared-memory bandwidth than two tiles x 128)
    p.mp = 2;
    bn = 128;
    b_item = bn * 128;
    a_slot *= 2;
    p.num_n_tiles = (cout + bn - 1) / bn;
    if (convt || up2x) {
      p.nt_per_phase = p.num_n_tiles;
      p.num_n_tiles *= 4;
    }
  }
  if (!stationary) {
    if (!p.pair && !ph_major) {
      cluster = env_int("UZ_FWD2_CLUSTER", 1);
      if (cluster != 1 && cluster != 2 && cluster != 4) cluster = 1;
      if (bn / cluster < 16) cluster = 1;
    }
    p.stg_bufs = 2;                      // one staging tile per epilogue group
    p.a_stages = env_int("UZ_FWD2_A_STAGES", p.mp == 2 ? (a_slot > 40000 ? 2 : 3) : bn == 256 ? 2 : bn == 128 ? 3 : 4);
    if (p.a_stages < 1 || p.a_stages > 6) p.a_stages = 2;
    int bs = (kMaxSmem - kTail - p.stg_bufs * 16384 - p.a_stages * a_slot) / b_item;
    p.b_stages = bs > 12 ? 12 : bs;
    if (p.b_stages < 2) return 0;
  }
  p.cluster = cluster;
  p.n_minor = ((convt || up2x) && env_int("UZ_FWD2_N_MINOR", 1)) ? 1 : 0;
  p.m_tiles = ((p.m_tiles_real + p.mp - 1) / p.mp + cluster - 1) / cluster * cluster;
  p.total_tiles = p.m_tiles * p.num_n_tiles;
  p.ragged = ((H % 16) != 0 || (W % 8) != 0 || p.m_tiles * p.mp != p.m_tiles_real) ? 1 : 0;
  p.a_off = 0;
  p.b_off = p.a_stages * a_slot;
  const long long b_bytes = p.b_stages ? (long long)p.b_stages * b_item : b_all;
  p.stg_off = (int)(p.b_off + b_bytes);
  p.bar_off = p.stg_off + p.stg_bufs * 16384;
  const int smem_bytes = p.bar_off + kTail + 1024;
  // packed weights: conv [Cout][taps][Ctot]; ConvT fwd [Cout][4 phases][Cin] (the phase is the "tap" coordinate);
  // ConvT dgrad [Cin][4][Cout] read as rows of K = 4*Cout
  int rc = make_weight_map(&p.bmap, w_packed, ctot, convt ? 4 : (up2x || up2x_d) ? 16 : p.num_taps, cout,
                           ph_major ? 64 : bn / cluster);
  if (rc != UZ_OK) return rc;
  if (convt || up2x) {
    for (int l = 0; l < 4 && rc == UZ_OK; ++l)
      rc = make_act_map(&p.omap[l], out, cout, out_ld, N, 2 * H, 2 * W, Lattice{l >> 1, l & 1, 2}, 8, 16, 1);
  } else {
    rc = make_act_map(&p.omap[0], out, cout, out_ld, N, H, W, full, 8, 16, 1);
  }
  if (rc != UZ_OK) return rc;
  p.ncols = cout;
  p.bias = bias;
  p.stats = stats;
  p.flags = flags & (UZ_FLAG_RELU | UZ_FLAG_ACCUMULATE | UZ_FLAG_STATS);
  int grid = p.total_tiles < sms ? p.total_tiles : sms;
  grid = grid / cluster * cluster;
  if (grid < cluster) return 0;
  cudaError_t e = launch_igemm_fwd2(p, bn, grid, smem_bytes, st);
  if (e != cudaSuccess) return cuda_error(e, "igemm_fwd2 launch");
  return 1;
}
}  // namespace uz

extern "C" int uz_conv_igemm(int mode, int N, int H, int W, int nseg, const void* const* seg_ptr, const int* seg_c,
                             const int* seg_ld, const void* w_packed, int cout, const float* bias, void* out,
                             int out_ld, int out_h, int out_w, float* stats, int flags, void* stream) {
  const int sms = uz_device_sm_count();
  if (sms < 0) return sms;
  // dilated 3x3 convolution (u2net.py:10: padding = dilation = dirate): UZ_CONV3X3 | UZ_DILATION(d).  Only the tap shifts
  // change, so it runs on the per-tap kernel (one shifted TMA box per tap, zero fill outside the image); the halo-box
  // kernel would need a (8 + 2d) x (16 + 2d) box per 64 channels - 98 KB at d = 8
  const int dil = ((mode >> 8) & 0xff) > 1 ? ((mode >> 8) & 0xff) : 1;
  mode &= 0xff;
  if (dil > 1 && mode != UZ_CONV3X3 && mode != UZ_CONV3X3_DGRAD)
    return set_error(UZ_ERR_UNSUPPORTED, "dilation %d is only supported for UZ_CONV3X3 / UZ_CONV3X3_DGRAD (mode %d)", dil, mode);
  if (dil > 64) return set_error(UZ_ERR_UNSUPPORTED, "dilation %d > 64", dil);
  if (N < 1 || H < 1 || W < 1) return set_error(UZ_ERR_INVALID, "bad activation size N=%d H=%d W=%d", N, H, W);
  if (nseg < 1 || nseg > kMaxSegs) return set_error(UZ_ERR_INVALID, "nseg=%d out of range [1,%d]", nseg, kMaxSegs);
  if (cout < 8 || (cout & 7) || (out_ld & 7) || out_ld < cout)
    return set_error(UZ_ERR_INVALID, "cout=%d / out_ld=%d must be multiples of 8 with out_ld >= cout", cout, out_ld);
  if (!aligned16(out) || !aligned16(w_packed)) return set_error(UZ_ERR_INVALID, "out / weights must be 16-byte aligned");
  if ((flags & UZ_FLAG_STATS) && stats == nullptr) return set_error(UZ_ERR_INVALID, "UZ_FLAG_STATS without stats buffer");

  IgemmParams p;
  memset(&p, 0, sizeof(p));
  ModeInfo mi;
  int rc = build_mode(mode, H, W, out_h, out_w, &p, &mi, dil);
  if (rc != UZ_OK) return rc;
  if (mi.num_phases > 1) {
    for (int ph = 1; ph < mi.num_phases; ++ph)
      if (mi.ntaps[ph] != mi.ntaps[0] && mode != UZ_CONV3X3_S2_DGRAD)
        return set_error(UZ_ERR_INVALID, "internal: ragged tap table");
  }
  if (nseg * mi.num_lat > kMaxAMaps)
    return set_error(UZ_ERR_UNSUPPORTED, "%d segments x %d lattices exceeds %d tensor maps", nseg, mi.num_lat, kMaxAMaps);

  int tw, th, tn;
  pick_tile(128, mi.Hm, mi.Wm, &tw, &th, &tn);
  int ctot = 0;
  for (int s = 0; s < nseg; ++s) {
    if (seg_c[s] < 8 || (seg_c[s] & 7) || (seg_ld[s] & 7) || seg_ld[s] < seg_c[s] || !aligned16(seg_ptr[s]))
      return set_error(UZ_ERR_INVALID, "segment %d: channels=%d ld=%d must be multiples of 8, pointer 16-B aligned", s,
                       seg_c[s], seg_ld[s]);
    p.seg_cblocks[s] = (seg_c[s] + 63) / 64;
    p.seg_k0[s] = ctot;
    ctot += seg_c[s];
    for (int l = 0; l < mi.num_lat; ++l) {
      rc = make_act_map(&p.amap[s * mi.num_lat + l], seg_ptr[s], seg_c[s], seg_ld[s], N, H, W, mi.lat[l], tw, th, tn);
      if (rc != UZ_OK) return rc;
    }
  }
  const int bn = cout <= 32 ? 32 : cout <= 64 ? 64 : cout <= 128 ? 128 : 256;
  rc = make_weight_map(&p.bmap, w_packed, ctot, mi.btaps, cout, bn);
  if (rc != UZ_OK) return rc;

  p.num_segs = nseg;
  p.num_lat = mi.num_lat;
  p.num_phases = mi.num_phases;
  p.N = N;
  p.H = mi.Hm;
  p.W = mi.Wm;
  p.tw_log2 = ilog2(tw);
  p.th_log2 = ilog2(th);
  p.xb = (mi.Wm + tw - 1) / tw;
  p.yb = (mi.Hm + th - 1) / th;
  p.nb = (N + tn - 1) / tn;
  p.num_n_tiles = (cout + bn - 1) / bn;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.Ho = mi.Ho;
  p.Wo = mi.Wo;
  p.out_ld = out_ld;
  p.oscale = mi.oscale;
  p.ncols = cout;
  p.bias = bias;
  p.stats = stats;
  p.flags = flags;
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  if (flags & UZ_FLAG_STATS) {
    cudaError_t e = cudaMemsetAsync(stats, 0, sizeof(float) * 2 * cout * uz_conv_stats_rows(), st);
    if (e != cudaSuccess) return cuda_error(e, "cudaMemsetAsync(stats)");
  }
  if (dil > 1) {
    // per-tap kernel below
  } else if ((mi.num_phases == 1 && mi.oscale == 1) || mode == UZ_CONVT2X2 || mode == UZ_CONV1X1_S2 || mode == UZ_CONV3X3_S2 ||
      mode == UZ_UP2X_CONV3X3 || mode == UZ_UP2X_CONV3X3_DGRAD) {
    rc = try_fwd2(mode, N, H, W, nseg, seg_ptr, seg_c, seg_ld, w_packed, cout, bias, out, out_ld, stats, flags, sms, st);
    if (rc != 0) return rc < 0 ? rc : UZ_OK;
  }
  if (flags & (UZ_FLAG_RES1X1 | UZ_FLAG_RES1X1_S2))
    return set_error(UZ_ERR_UNSUPPORTED, "UZ_FLAG_RES1X1 needs the halo-box kernel (output of at least 16 x 8 pixels, no dilation)");
  if (mode == UZ_CONV3X3_S2_DGRAD) {
    // ragged tap counts per phase: one launch per phase (each is a 1-phase problem with its own tap table)
    if (flags & UZ_FLAG_STATS) return set_error(UZ_ERR_UNSUPPORTED, "stats not supported for stride-2 dgrad");
    for (int ph = 0; ph < 4; ++ph) {
      IgemmParams q = p;
      q.num_phases = 1;
      q.taps[0] = p.taps[ph];
      q.phase_oy[0] = p.phase_oy[ph];
      q.phase_ox[0] = p.phase_ox[ph];
      q.num_taps = mi.ntaps[ph];
      q.total_tiles = q.xb * q.yb * q.nb * q.num_n_tiles;
      const int grid = q.total_tiles < sms ? q.total_tiles : sms;
      cudaError_t e = launch_igemm_fwd(q, bn, grid, st);
      if (e != cudaSuccess) return cuda_error(e, "igemm_fwd launch");
    }
    return UZ_OK;
  }
  p.num_taps = mi.ntaps[0];
  p.total_tiles = p.xb * p.yb * p.nb * p.num_n_tiles * p.num_phases;
  const int grid = p.total_tiles < sms ? p.total_tiles : sms;
  cudaError_t e = launch_igemm_fwd(p, bn, grid, st);
  if (e != cudaSuccess) return cuda_error(e, "igemm_fwd launch");
  return UZ_OK;
}

// ------------------------------------------------------------------------------------------------ batched TN GEMM
// C[n][m][j] = sum_k A[n][m][k] * B[n][j][k]  (bf16, fp32 accumulation) - the matrix products of the attention
// bottlenecks (torch.bmm / matmul in transatt_unet.py:49-53,99-105, unet_transformer.py:131-135,214-217).  A is read like
// an NHWC activation whose "image" n has M pixels of K channels (W = 8, H = M / 8: a 128-pixel tile is 128 consecutive
// rows), B is the 1x1 kernel's weight operand with one matrix PER IMAGE (Igemm2Params::b_batched), C its output.
// Shapes below one tile, or not multiples of 8, run on a plain CUDA-core kernel (csrc/attention.cu).
extern "C" int uz_bmm_tn(const void* A, long long lda, long long a_img, const void* B, long long b_img, void* C,
                         long long ldc, long long c_img, int N, int M, int K, int J, void* stream) {
  const int sms = uz_device_sm_count();
  if (sms < 0) return sms;
  if (!A || !B || !C || N < 1 || M < 1 || K < 1 || J < 1 || lda < K || ldc < J)
    return set_error(UZ_ERR_INVALID, "bmm_tn: bad extents N=%d M=%d K=%d J=%d", N, M, K, J);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool fast = M >= env_int("UZ_BMM_MIN_M", 8) && (M % 8) == 0 && (K % 8) == 0 && (J % 8) == 0 && (lda % 8) == 0 && (ldc % 8) == 0 &&
                    a_img == static_cast<long long>(M) * lda && c_img == static_cast<long long>(M) * ldc &&
                    (b_img % 8) == 0 && b_img >= static_cast<long long>(J) * K && aligned16(A) && aligned16(B) &&
                    aligned16(C) && N < 65536 && !env_int("UZ_BMM_SMALL", 0);
  if (!fast) {
    cudaError_t e = launch_bmm_tn_small(A, lda, a_img, B, K, b_img, C, ldc, c_img, N, M, K, J, st);
    if (e != cudaSuccess) return cuda_error(e, "bmm_tn_small launch");
    return UZ_OK;
  }
  Igemm2Params p;
  memset(&p, 0, sizeof(p));
  const int H = M / 8, W = 8;
  p.pad = 0;
  p.halo_w = 8;
  p.halo_bytes = 8 * 16 * 128;
  p.num_taps = 1;
  p.num_segs = 1;
  p.seg_cblocks[0] = (K + 63) / 64;
  p.seg_klast[0] = (K - (p.seg_cblocks[0] - 1) * 64 + 15) / 16;
  p.seg_taps[0] = 1;
  p.kblocks = p.b_items = p.seg_cblocks[0];
  p.N = N; p.H = H; p.W = W;
  p.xb = 1;
  p.yb = (H + 15) / 16;
  p.m_tiles_real = N * p.yb;
  const int bn = J <= 64 ? 64 : J <= 128 ? 128 : 256;
  p.num_n_tiles = (J + bn - 1) / bn;
  p.mp = 1;
  p.cluster = 1;
  p.b_batched = 1;
  p.m_tiles = p.m_tiles_real;
  p.total_tiles = p.m_tiles * p.num_n_tiles;
  p.ragged = (H % 16) != 0 ? 1 : 0;
  const int kMaxSmem = 232448 - 1024, kTail = 2560;
  const int a_slot = p.halo_bytes, b_item = bn * 128;
  p.stg_bufs = 2;
  p.a_stages = 4;
  int bs = (kMaxSmem - kTail - p.stg_bufs * 16384 - p.a_stages * a_slot) / b_item;
  p.b_stages = bs > 8 ? 8 : bs;
  p.a_off = 0;
  p.b_off = p.a_stages * a_slot;
  p.stg_off = p.b_off + p.b_stages * b_item;
  p.bar_off = p.stg_off + p.stg_bufs * 16384;
  const int smem_bytes = p.bar_off + kTail + 1024;
  const Lattice full = {0, 0, 1};
  int rc = make_act_map(&p.amap[0], A, K, (int)lda, N, H, W, full, 8, 16, 1);
  if (rc != UZ_OK) return rc;
  rc = make_act_map(&p.omap[0], C, J, (int)ldc, N, H, W, full, 8, 16, 1);
  if (rc != UZ_OK) return rc;
  {   // B: 3-D (K, image, row) - the image index takes the place of the filter tap
    EncodeTiledFn fn = encode_fn();
    if (!fn) return set_error(UZ_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)N, (cuuint64_t)J};
    cuuint64_t strides[2] = {(cuuint64_t)b_img * 2, (cuuint64_t)K * 2};
    cuuint32_t box[3] = {64, 1, (cuuint32_t)bn};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(&p.bmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(B), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(UZ_ERR_CUDA, "cuTensorMapEncodeTiled(bmm B K=%d N=%d J=%d) failed: %d", K, N, J, (int)r);
  }
  p.ncols = J;
  int grid = p.total_tiles < sms ? p.total_tiles : sms;
  cudaError_t e = launch_igemm_fwd2(p, bn, grid, smem_bytes, st);
  if (e != cudaSuccess) return cuda_error(e, "bmm_tn (igemm_fwd2) launch");
  return UZ_OK;
}

// ------------------------------------------------------------------------------------------------ wgrad
namespace uz {
struct WgPlan {
  int rc;
  int cp, cq, taps, phases;
  int Hr, Wr;  // reduction pixel space per image
  int tw, th, tn, xb, yb, nb, chunks;
  int m_tiles, q_cblocks, atoms, n_slabs, splits;
  size_t ws_bytes;
};
static WgPlan plan_wgrad(int mode, int N, int H, int W, int cin, int cout) {
  WgPlan pl;
  memset(&pl, 0, sizeof(pl));
  pl.rc = UZ_OK;
  pl.phases = 1;
  pl.Hr = H;
  pl.Wr = W;
  pl.cp = cout;
  pl.cq = cin;
  switch (mode) {
    case UZ_CONV3X3: pl.taps = 9; break;
    case UZ_CONV1X1: pl.taps = 1; break;
    case UZ_CONV3X3_S2: pl.taps = 9; pl.Hr = (H - 1) / 2 + 1; pl.Wr = (W - 1) / 2 + 1; break;
    case UZ_CONV1X1_S2: pl.taps = 1; pl.Hr = (H - 1) / 2 + 1; pl.Wr = (W - 1) / 2 + 1; break;
    case UZ_UP2X_CONV3X3: pl.taps = 9; pl.phases = 4; break;
    case UZ_CONVT2X2: pl.taps = 4; pl.cp = cin; pl.cq = cout; break;
    default: pl.rc = set_error(UZ_ERR_INVALID, "wgrad: unsupported mode %d", mode); return pl;
  }
  pick_tile(64, pl.Hr, pl.Wr, &pl.tw, &pl.th, &pl.tn);
  pl.xb = (pl.Wr + pl.tw - 1) / pl.tw;
  pl.yb = (pl.Hr + pl.th - 1) / pl.th;
  pl.nb = (N + pl.tn - 1) / pl.tn;
  pl.chunks = pl.xb * pl.yb * pl.nb;
  pl.m_tiles = (pl.cp + 127) / 128;
  pl.q_cblocks = (pl.cq + 63) / 64;
  pl.atoms = pl.taps * pl.q_cblocks;
  pl.n_slabs = (pl.atoms + 3) / 4;
  const int items0 = pl.m_tiles * pl.n_slabs;
  const int total_chunks = pl.chunks * pl.phases;
  int sms = sm_count();
  if (sms < 1) sms = 148;
  int splits = 1;
  if (items0 < 2 * sms) splits = (2 * sms + items0 - 1) / items0;
  int max_splits = total_chunks / 8;
  if (max_splits < 1) max_splits = 1;
  if (splits > max_splits) splits = max_splits;
  pl.splits = splits;
  pl.ws_bytes = sizeof(float) * (size_t)splits * pl.cp * pl.taps * pl.cq;
  return pl;
}
}  // namespace uz

namespace uz {
// second-generation kernel: 3x3 stride-1 on images of at least one 8x16 chunk
static bool use_wgrad2(int mode, int H, int W, int cout = 64) {
  static const bool disabled = getenv("UZ_DISABLE_WGRAD2") != nullptr;
  if (disabled || H < 16 || W < 8) return false;
  // nearest-x2 conv (H, W = the LOW-resolution image): phase form under the predicate the forward pass uses
  return mode == UZ_CONV3X3 || (mode == UZ_UP2X_CONV3X3 && uz_up2x_phase_form(H, W, cout));
}
struct Wg2Plan {
  int xb, yb, chunks, co_blocks, ci_blocks, splits, pair, ci_units, wide;
  size_t ws_bytes;
};
static Wg2Plan plan_wgrad2(int N, int H, int W, int cin, int cout, bool up2x = false) {
  Wg2Plan pl;
  pl.xb = (W + 7) / 8;
  pl.yb = (H + 15) / 16;
  pl.chunks = N * pl.xb * pl.yb;
  // WIDE: 128 output channels per work unit, N = 128 MMAs in two passes over the chunk range (see igemm_wgrad2.cu):
  // the N = 64 tile is capped by the shared-memory port at 67 % (single CTA) / 80 % (pair) of the tensor rate
  pl.ci_blocks = (cin + 63) / 64;
  // measured per layer (tools/bench_wgrad.py): WIDE wins 2-19 % wherever the CTA pair applies; a single WIDE CTA
  // (64 input channels) is 4 % slower than the N = 64 tile (its 4-stage ring of 55 KB stages is shallower)
  pl.wide = (!up2x && cout >= 128 && pl.ci_blocks >= 2 && env_int("UZ_WGRAD2_WIDE", 1)) ? 1 : 0;
  pl.co_blocks = pl.wide ? (cout + 127) / 128 : (cout + 63) / 64;
  int sms = sm_count();
  if (sms < 1) sms = 148;
  // CTA pairs (cta_group::2) over two adjacent input-channel blocks: the M = 128, N = 64 MMA of a single CTA reads
  // 192 B/clk of shared memory (67 % ceiling, measured 59-64 % tensor-active); the pair halves the dZ bytes per CTA
  pl.pair = (pl.ci_blocks >= 2 && sms >= 2 && env_int("UZ_WGRAD2_PAIR", 1)) ? 1 : 0;
  pl.ci_units = pl.pair ? (pl.ci_blocks + 1) / 2 : pl.ci_blocks;
  const int units = pl.co_blocks * pl.ci_units;
  if (pl.pair) sms /= 2;   // work is scheduled over CTA pairs
  // split-K factor: minimise the makespan  waves(items) x (chunks per item + epilogue)  over the persistent grid -
  // an item count just above a multiple of the SM count would run a nearly empty extra wave (measured: 256->256 @128^2
  // took 1.43x its ideal time with 304 items on 148 SMs).  The epilogue (147 KB of fp32 partials, not overlapped with
  // the main loop because the accumulators fill TMEM) is charged as 4 chunk times.
  int max_splits = pl.chunks / 4;
  if (max_splits < 1) max_splits = 1;
  if (max_splits > 4 * sms) max_splits = 4 * sms;
  int splits = 1;
  long long best = -1;
  for (int sp = 1; sp <= max_splits; ++sp) {
    const long long items = (long long)units * sp;
    const long long waves = (items + sms - 1) / sms;
    // WIDE: per chunk 3 + 2 MMA groups of 64 cycles against 5 of 32, two epilogues of 384 + 256 columns against one
    // of 320 - same relative weight of main loop and epilogue, so the same formula ranks the split factors
    const long long cost = waves * ((pl.chunks + sp - 1) / sp + 4);
    if (best < 0 || cost < best) {
      best = cost;
      splits = sp;
    }
  }
  pl.splits = splits;
  // tile-major partials: one contiguous 9 x (64 | 128) x 64 fp32 region per (split, co block, ci block)
  pl.ws_bytes = sizeof(float) * (size_t)splits * pl.co_blocks * pl.ci_blocks * (up2x ? 16 : 9) * (pl.wide ? 128 : 64) * 64;
  return pl;
}
}  // namespace uz

extern "C" size_t uz_conv_wgrad_workspace_bytes(int mode, int N, int H, int W, int cin, int cout) {
  const int dil = ((mode >> 8) & 0xff) > 1 ? ((mode >> 8) & 0xff) : 1;
  mode &= 0xff;
  if (dil == 1 && use_wgrad2(mode, H, W, cout)) return plan_wgrad2(N, H, W, cin, cout, mode == UZ_UP2X_CONV3X3).ws_bytes;
  WgPlan pl = plan_wgrad(mode, N, H, W, cin, cout);
  return pl.rc == UZ_OK ? pl.ws_bytes : 0;
}

extern "C" int uz_conv_wgrad(int mode, int N, int H, int W, const void* x, int cin, int x_ld, const void* dy,
                             int cout, int dy_ld, float* dw, int cin_total, int c_off, int accumulate,
                             void* workspace, size_t workspace_bytes, void* stream) {
  const int sms = uz_device_sm_count();
  if (sms < 0) return sms;
  const int dil = ((mode >> 8) & 0xff) > 1 ? ((mode >> 8) & 0xff) : 1;   // UZ_CONV3X3 | UZ_DILATION(d), see uz_conv_igemm
  mode &= 0xff;
  if (dil > 1 && mode != UZ_CONV3X3) return set_error(UZ_ERR_UNSUPPORTED, "wgrad: dilation needs UZ_CONV3X3 (mode %d)", mode);
  if (dil > 64) return set_error(UZ_ERR_UNSUPPORTED, "wgrad: dilation %d > 64", dil);
  if ((cin & 7) || (cout & 7) || (x_ld & 7) || (dy_ld & 7) || cin < 8 || cout < 8)
    return set_error(UZ_ERR_INVALID, "wgrad: channel counts / strides must be multiples of 8");
  if (!aligned16(x) || !aligned16(dy) || !aligned16(workspace))
    return set_error(UZ_ERR_INVALID, "wgrad: pointers must be 16-byte aligned");
  if (dil == 1 && use_wgrad2(mode, H, W, cout)) {
    const bool up2x = mode == UZ_UP2X_CONV3X3;
    const Wg2Plan p2 = plan_wgrad2(N, H, W, cin, cout, up2x);
    if (workspace_bytes < p2.ws_bytes)
      return set_error(UZ_ERR_WORKSPACE, "wgrad: workspace %zu < required %zu bytes", workspace_bytes, p2.ws_bytes);
    Wgrad2Params q;
    memset(&q, 0, sizeof(q));
    const Lattice full2 = {0, 0, 1};
    int rc2 = make_act_map(&q.xmap, x, cin, x_ld, N, H, W, full2, 10, 18, 1);
    if (up2x) {   // dy is the HIGH-resolution gradient (2H x 2W): one strided view per output phase
      for (int ph = 0; ph < 4 && rc2 == UZ_OK; ++ph) {
        const Lattice lat = {ph >> 1, ph & 1, 2};
        rc2 = make_act_map(&q.dzmap_ph[ph], dy, cout, dy_ld, N, 2 * H, 2 * W, lat, 8, 16, 1);
        if (rc2 == UZ_OK && p2.pair) rc2 = make_act_map(&q.dzmap_half_ph[ph], dy, cout, dy_ld, N, 2 * H, 2 * W, lat, 8, 16, 1, 32);
      }
      q.up2x = 1;
    } else {
      if (rc2 == UZ_OK) rc2 = make_act_map(&q.dzmap, dy, cout, dy_ld, N, H, W, full2, 8, 16, 1);
      if (rc2 == UZ_OK && p2.pair) rc2 = make_act_map(&q.dzmap_half, dy, cout, dy_ld, N, H, W, full2, 8, 16, 1, 32);
    }
    if (rc2 != UZ_OK) return rc2;
    q.pair = p2.pair;
    q.wide = p2.wide;
    q.ci_units = p2.ci_units;
    q.xb = p2.xb;
    q.yb = p2.yb;
    q.total_chunks = p2.chunks;
    q.cp = cout;
    q.cq = cin;
    q.co_blocks = p2.co_blocks;
    q.ci_blocks = p2.ci_blocks;
    q.splits = p2.splits;
    q.total_items = p2.co_blocks * p2.ci_units * p2.splits;
    q.partial = static_cast<float*>(workspace);
    cudaStream_t st2 = static_cast<cudaStream_t>(stream);
    const int grid2 = p2.pair ? 2 * (q.total_items < sms / 2 ? q.total_items : sms / 2)
                              : (q.total_items < sms ? q.total_items : sms);
    cudaError_t e2 = launch_igemm_wgrad2(q, grid2, st2);
    if (e2 != cudaSuccess) return cuda_error(e2, "igemm_wgrad2 launch");
    e2 = launch_wgrad2_reduce(q, dw + (long long)c_off * 9, (long long)cin_total * 9, accumulate, st2);
    if (e2 != cudaSuccess) return cuda_error(e2, "wgrad_reduce launch");
    return UZ_OK;
  }
  WgPlan pl = plan_wgrad(mode, N, H, W, cin, cout);
  if (pl.rc != UZ_OK) return pl.rc;
  if (workspace_bytes < pl.ws_bytes)
    return set_error(UZ_ERR_WORKSPACE, "wgrad: workspace %zu < required %zu bytes", workspace_bytes, pl.ws_bytes);

  WgradParams p;
  memset(&p, 0, sizeof(p));
  auto set_tap = [&](int ph, int i, int dy_, int dx_, int lat) {
    p.taps[ph].dy[i] = (int8_t)dy_;
    p.taps[ph].dx[i] = (int8_t)dx_;
    p.taps[ph].btap[i] = (int8_t)i;
    p.taps[ph].lat[i] = (int8_t)lat;
  };
  int rc = UZ_OK;
  const Lattice full = {0, 0, 1};
  switch (mode) {
    case UZ_CONV3X3:
      rc = make_act_map(&p.pmap[0], dy, cout, dy_ld, N, H, W, full, pl.tw, pl.th, pl.tn);
      if (rc == UZ_OK) rc = make_act_map(&p.qmap[0], x, cin, x_ld, N, H, W, full, pl.tw, pl.th, pl.tn);
      for (int r = 0; r < 3; ++r)
        for (int s = 0; s < 3; ++s) set_tap(0, r * 3 + s, dil * (r - 1), dil * (s - 1), 0);
      break;
    case UZ_CONV1X1:
      rc = make_act_map(&p.pmap[0], dy, cout, dy_ld, N, H, W, full, pl.tw, pl.th, pl.tn);
      if (rc == UZ_OK) rc = make_act_map(&p.qmap[0], x, cin, x_ld, N, H, W, full, pl.tw, pl.th, pl.tn);
      set_tap(0, 0, 0, 0, 0);
      break;
    case UZ_CONV3X3_S2:
      rc = make_act_map(&p.pmap[0], dy, cout, dy_ld, N, pl.Hr, pl.Wr, full, pl.tw, pl.th, pl.tn);
      for (int l = 0; l < 4 && rc == UZ_OK; ++l)
        rc = make_act_map(&p.qmap[l], x, cin, x_ld, N, H, W, Lattice{l >> 1, l & 1, 2}, pl.tw, pl.th, pl.tn);
      for (int r = 0; r < 3; ++r)
        for (int s = 0; s < 3; ++s) {
          int pr, dr, pc, dc;
          s2_tap(r, &pr, &dr);
          s2_tap(s, &pc, &dc);
          set_tap(0, r * 3 + s, dr, dc, pr * 2 + pc);
        }
      break;
    case UZ_CONV1X1_S2:
      rc = make_act_map(&p.pmap[0], dy, cout, dy_ld, N, pl.Hr, pl.Wr, full, pl.tw, pl.th, pl.tn);
      if (rc == UZ_OK) rc = make_act_map(&p.qmap[0], x, cin, x_ld, N, H, W, Lattice{0, 0, 2}, pl.tw, pl.th, pl.tn);
      set_tap(0, 0, 0, 0, 0);
      break;
    case UZ_UP2X_CONV3X3:
      // pixel space = low-res (N,H,W); phase (py,px) pairs dY[2y+py, 2x+px] with x_low[y+dy, x+dx]
      rc = make_act_map(&p.qmap[0], x, cin, x_ld, N, H, W, full, pl.tw, pl.th, pl.tn);
      for (int ph = 0; ph < 4 && rc == UZ_OK; ++ph) {
        const int py = ph >> 1, px = ph & 1;
        rc = make_act_map(&p.pmap[ph], dy, cout, dy_ld, N, 2 * H, 2 * W, Lattice{py, px, 2}, pl.tw, pl.th, pl.tn);
        for (int r = 0; r < 3; ++r)
          for (int s = 0; s < 3; ++s)
            set_tap(ph, r * 3 + s, (py + r - 1 + 2) / 2 - 1, (px + s - 1 + 2) / 2 - 1, 0);
      }
      break;
    case UZ_CONVT2X2:
      // rows = input channels (P = x, low res), cols = output channels (Q = dY lattices of the 2H x 2W gradient)
      rc = make_act_map(&p.pmap[0], x, cin, x_ld, N, H, W, full, pl.tw, pl.th, pl.tn);
      for (int l = 0; l < 4 && rc == UZ_OK; ++l) {
        rc = make_act_map(&p.qmap[l], dy, cout, dy_ld, N, 2 * H, 2 * W, Lattice{l >> 1, l & 1, 2}, pl.tw, pl.th, pl.tn);
        set_tap(0, l, 0, 0, l);
      }
      break;
    default:
      return set_error(UZ_ERR_INVALID, "wgrad: unsupported mode %d", mode);
  }
  if (rc != UZ_OK) return rc;

  p.num_phases = pl.phases;
  p.num_taps = pl.taps;
  p.N = N;
  p.H = pl.Hr;
  p.W = pl.Wr;
  p.tw_log2 = ilog2(pl.tw);
  p.th_log2 = ilog2(pl.th);
  p.xb = pl.xb;
  p.yb = pl.yb;
  p.nb = pl.nb;
  p.chunks_per_phase = pl.chunks;
  p.cp = pl.cp;
  p.cq = pl.cq;
  p.m_tiles = pl.m_tiles;
  p.q_cblocks = pl.q_cblocks;
  p.num_atoms = pl.atoms;
  p.n_slabs = pl.n_slabs;
  p.splits = pl.splits;
  p.total_items = pl.m_tiles * pl.n_slabs * pl.splits;
  p.partial = static_cast<float*>(workspace);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = p.total_items < sms ? p.total_items : sms;
  cudaError_t e = launch_igemm_wgrad(p, grid, st);
  if (e != cudaSuccess) return cuda_error(e, "igemm_wgrad launch");

  long long s_row, s_col, s_tap = 1;
  float* dst = dw;
  if (mode == UZ_CONVT2X2) {
    s_row = (long long)cout * 4;
    s_col = 4;
  } else {
    s_row = (long long)cin_total * pl.taps;
    s_col = pl.taps;
    dst = dw + (long long)c_off * pl.taps;
  }
  e = launch_wgrad_reduce(p.partial, dst, pl.splits, pl.cp, pl.taps, pl.cq, s_row, s_col, s_tap, accumulate, st);
  if (e != cudaSuccess) return cuda_error(e, "wgrad_reduce launch");
  return UZ_OK;
}
