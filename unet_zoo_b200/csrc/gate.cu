// Additive attention gate of attention_unet (reference: AttentionBlock, attention_unet.py:6-40) - the bandwidth part.
//
//   g1 = BN_g(W_g * g + b_g)      x1 = BN_x(W_x * x + b_x)        (1x1 convs: tcgen05 GEMMs, uz_conv_igemm)
//   r  = ReLU(g1 + x1)            psi_raw = w_psi . r + b_psi       (per pixel dot over F = f_int channels)
//   out = x * sigmoid(BN_psi(psi_raw))
//
// The two GEMM outputs zg, zx (bf16 [npix, F]) are the only F-wide tensors that touch HBM: r is rebuilt in registers
// wherever it is needed and psi_raw is one fp32 per pixel.  LPP = F/8 lanes cooperate on one pixel (16-byte loads,
// xor-shuffle reduction); per-channel constants stay in registers because a lane keeps its channel group.
// Statistics for BN_psi (a 1-channel BatchNorm) are two-stage: per-block partials, then the generic finalize kernels.
#include "pw.cuh"

namespace uz {

constexpr int kGateBlock = 256;

// ---------------------------------------------------------------------------------------------- forward: psi_raw + stats
// bn_g / bn_x: [4][F] = {mean, invstd, scale, shift} (only scale / shift are read here)
template <int LPP>
__global__ void gate_psi_fwd_kernel(const __nv_bfloat16* __restrict__ zg, int zg_ld, const __nv_bfloat16* __restrict__ zx,
                                    int zx_ld, long long npix, int F, const float* __restrict__ bn_g,
                                    const float* __restrict__ bn_x, const float* __restrict__ w_psi,
                                    const float* __restrict__ b_psi, float* __restrict__ psi_raw,
                                    float* __restrict__ partial) {
  __shared__ float red[2][kGateBlock / 32];
  const int g = threadIdx.x % LPP;
  const int slot = threadIdx.x / LPP;
  constexpr int spb = kGateBlock / LPP;
  const bool active = g * 8 < F;
  float sg[8], hg[8], sx[8], hx[8], w[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) sg[j] = hg[j] = sx[j] = hx[j] = w[j] = 0.f;
  if (active) {
    pw_load8f(bn_g + 2 * F + g * 8, sg);
    pw_load8f(bn_g + 3 * F + g * 8, hg);
    pw_load8f(bn_x + 2 * F + g * 8, sx);
    pw_load8f(bn_x + 3 * F + g * 8, hx);
    pw_load8f(w_psi + g * 8, w);
  }
  const float bias = b_psi ? __ldg(b_psi) : 0.f;
  float s = 0.f, q = 0.f;
  const long long stride = static_cast<long long>(gridDim.x) * spb;
  const long long iters = (npix + stride - 1) / stride;
  for (long long i = 0; i < iters; ++i) {  // uniform trip count for the shuffles
    const long long p = i * stride + static_cast<long long>(blockIdx.x) * spb + slot;
    const bool ok = p < npix;
    float dot = 0.f;
    if (ok && active) {
      float a[8], b[8];
      pw_unpack8(pw_ldg16(zg + p * zg_ld + g * 8), a);
      pw_unpack8(pw_ldg16(zx + p * zx_ld + g * 8), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float r = fmaxf(fmaf(a[j], sg[j], hg[j]) + fmaf(b[j], sx[j], hx[j]), 0.f);
        dot = fmaf(w[j], r, dot);
      }
    }
#pragma unroll
    for (int o = LPP / 2; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (ok && g == 0) {
      const float v = dot + bias;
      psi_raw[p] = v;
      s += v;
      q = fmaf(v, v, q);
    }
  }
  s = pw_warp_sum(s);
  q = pw_warp_sum(q);
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = s;
    red[1][threadIdx.x >> 5] = q;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float ts = 0.f, tq = 0.f;
#pragma unroll
    for (int i = 0; i < kGateBlock / 32; ++i) {
      ts += red[0][i];
      tq += red[1][i];
    }
    partial[blockIdx.x * 2] = ts;
    partial[blockIdx.x * 2 + 1] = tq;
  }
}

// ---------------------------------------------------------------------------------------------- forward: out = x * sigmoid(BN(psi_raw))
__global__ void gate_apply_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, const float* __restrict__ psi_raw,
                                  const float* __restrict__ psi_bn, long long npix, int cv, __nv_bfloat16* __restrict__ out,
                                  int out_ld) {
  const float sc = __ldg(psi_bn + 2), sh = __ldg(psi_bn + 3);
  const long long total = npix * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long p = i / cv;
    const int g = static_cast<int>(i - p * cv);
    const float a = pw_sigmoid(fmaf(__ldg(psi_raw + p), sc, sh));
    float f[8];
    pw_unpack8(pw_ldg16(x + p * x_ld + g * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= a;
    *reinterpret_cast<uint4*>(out + p * out_ld + g * 8) = pw_pack8(f);
  }
}

// ---------------------------------------------------------------------------------------------- backward stage 1
// a = sigmoid(BN(psi_raw)); dx (+)= a * dout; dpsi = sum_c dout*x; dsb = dpsi * a * (1-a)  (gradient at the BN_psi output)
// partial[block][2] = { sum dsb, sum dsb * xhat }   with xhat = (psi_raw - mean) * invstd
// LPP lanes share a pixel (LPP = channel groups of 8 rounded up to a power of two, at most 32), so a warp works on
// 32 / LPP pixels at once - with one warp per pixel (first version) only 8 of 32 lanes were busy at C = 64, the
// largest gate of attention_unet (1.72 ms per step for four gates, ~3x the HBM time of their bytes).  Two pixels per
// lane group are in flight.
template <int LPP>
__global__ void __launch_bounds__(kGateBlock)
gate_bwd_psi_kernel(const __nv_bfloat16* __restrict__ x, int x_ld, const __nv_bfloat16* __restrict__ dout,
                    int dout_ld, const float* __restrict__ psi_raw, const float* __restrict__ psi_bn,
                    long long npix, int cv, __nv_bfloat16* __restrict__ dx, int dx_ld, int accumulate,
                    float* __restrict__ dsb, float* __restrict__ partial) {
  __shared__ float red[2][kGateBlock / 32];
  constexpr int PPW = 32 / LPP;   // pixels per warp and step
  constexpr int U = 2;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int sub = lane / LPP, gl = lane % LPP;
  const float mean = __ldg(psi_bn), invstd = __ldg(psi_bn + 1), sc = __ldg(psi_bn + 2), sh = __ldg(psi_bn + 3);
  float s = 0.f, q = 0.f;
  constexpr int wpb = kGateBlock / 32;
  const long long stride = static_cast<long long>(gridDim.x) * wpb * PPW;
  // the loop bound is warp-uniform: every lane takes part in the shuffles below
  for (long long base = (static_cast<long long>(blockIdx.x) * wpb + warp) * PPW; base < npix; base += U * stride) {
    float dot[U], a[U], pr[U];
    bool valid[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long p = base + u * stride + sub;
      valid[u] = p < npix;
      dot[u] = 0.f;
      pr[u] = valid[u] ? __ldg(psi_raw + p) : 0.f;
      a[u] = pw_sigmoid(fmaf(pr[u], sc, sh));
    }
    for (int g = gl; g < cv; g += LPP) {
      uint4 xr[U], dr[U], orr[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!valid[u]) continue;
        const long long p = base + u * stride + sub;
        xr[u] = pw_ldg16(x + p * x_ld + g * 8);
        dr[u] = pw_ldg16(dout + p * dout_ld + g * 8);
        if (accumulate) orr[u] = *reinterpret_cast<const uint4*>(dx + p * dx_ld + g * 8);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!valid[u]) continue;
        const long long p = base + u * stride + sub;
        float xv[8], dv[8], o[8];
        pw_unpack8(xr[u], xv);
        pw_unpack8(dr[u], dv);
        if (accumulate) {
          pw_unpack8(orr[u], o);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          dot[u] = fmaf(dv[j], xv[j], dot[u]);
          o[j] = fmaf(a[u], dv[j], o[j]);
        }
        *reinterpret_cast<uint4*>(dx + p * dx_ld + g * 8) = pw_pack8(o);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float d = dot[u];
#pragma unroll
      for (int o = LPP / 2; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
      if (gl == 0 && valid[u]) {
        d = d * a[u] * (1.f - a[u]);
        dsb[base + u * stride + sub] = d;
        s += d;
        q = fmaf(d, (pr[u] - mean) * invstd, q);
      }
    }
  }
  s = pw_warp_sum(s);   // lanes with gl != 0 hold zeros
  q = pw_warp_sum(q);
  if (lane == 0) {
    red[0][warp] = s;
    red[1][warp] = q;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float ts = 0.f, tq = 0.f;
#pragma unroll
    for (int i = 0; i < wpb; ++i) {
      ts += red[0][i];
      tq += red[1][i];
    }
    partial[blockIdx.x * 2] = ts;
    partial[blockIdx.x * 2 + 1] = tq;
  }
}

// ---------------------------------------------------------------------------------------------- backward stage 2
// dpr = scale_psi * dsb + A + B * psi_raw   (BatchNorm backward of the 1-channel BN_psi, coef = {A, B})
// r rebuilt from zg, zx;  dr[p, c] = (r > 0) ? dpr * w_psi[c] : 0  (bf16, the "dy" of BN_g and BN_x)
// wpart[block][F] = sum_p dpr * r[p, c] ;  bpart[block] = sum_p dpr
template <int LPP>
__global__ void gate_bwd_r_kernel(const __nv_bfloat16* __restrict__ zg, int zg_ld, const __nv_bfloat16* __restrict__ zx,
                                  int zx_ld, long long npix, int F, const float* __restrict__ bn_g,
                                  const float* __restrict__ bn_x, const float* __restrict__ w_psi,
                                  const float* __restrict__ psi_raw, const float* __restrict__ dsb,
                                  const float* __restrict__ psi_bn, const float* __restrict__ coef,
                                  __nv_bfloat16* __restrict__ dr, int dr_ld, float* __restrict__ wpart,
                                  float* __restrict__ bpart) {
  extern __shared__ float red[];  // [kGateBlock][9]
  const int g = threadIdx.x % LPP;
  const int slot = threadIdx.x / LPP;
  constexpr int spb = kGateBlock / LPP;
  const bool active = g * 8 < F;
  float sg[8], hg[8], sx[8], hx[8], w[8], aw[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) sg[j] = hg[j] = sx[j] = hx[j] = w[j] = aw[j] = 0.f;
  if (active) {
    pw_load8f(bn_g + 2 * F + g * 8, sg);
    pw_load8f(bn_g + 3 * F + g * 8, hg);
    pw_load8f(bn_x + 2 * F + g * 8, sx);
    pw_load8f(bn_x + 3 * F + g * 8, hx);
    pw_load8f(w_psi + g * 8, w);
  }
  const float sc = __ldg(psi_bn + 2), ca = __ldg(coef), cb = __ldg(coef + 1);
  float ab = 0.f;
  if (active) {
    for (long long p = static_cast<long long>(blockIdx.x) * spb + slot; p < npix; p += static_cast<long long>(gridDim.x) * spb) {
      const float pr = __ldg(psi_raw + p);
      const float dpr = fmaf(sc, __ldg(dsb + p), fmaf(cb, pr, ca));
      float a[8], b[8], o[8];
      pw_unpack8(pw_ldg16(zg + p * zg_ld + g * 8), a);
      pw_unpack8(pw_ldg16(zx + p * zx_ld + g * 8), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float r = fmaxf(fmaf(a[j], sg[j], hg[j]) + fmaf(b[j], sx[j], hx[j]), 0.f);
        aw[j] = fmaf(dpr, r, aw[j]);
        o[j] = r > 0.f ? dpr * w[j] : 0.f;
      }
      *reinterpret_cast<uint4*>(dr + p * dr_ld + g * 8) = pw_pack8(o);
      if (g == 0) ab += dpr;
    }
  }
  float* mine = red + static_cast<size_t>(threadIdx.x) * 9;
#pragma unroll
  for (int j = 0; j < 8; ++j) mine[j] = aw[j];
  mine[8] = ab;
  __syncthreads();
  if (slot == 0 && active) {
    for (int r = 1; r < spb; ++r) {
      const float* o = red + (static_cast<size_t>(r) * LPP + g) * 9;
#pragma unroll
      for (int j = 0; j < 8; ++j) aw[j] += o[j];
      ab += o[8];
    }
    float* dst = wpart + static_cast<size_t>(blockIdx.x) * F;
#pragma unroll
    for (int j = 0; j < 8; ++j) dst[g * 8 + j] = aw[j];
    if (g == 0) bpart[blockIdx.x] = ab;
  }
}

static int gate_lpp(int F) {
  const int cv = F / 8;
  int l = 1;
  while (l < cv) l <<= 1;
  return l;
}
static int gate_rows() {
  const int sms = uz_device_sm_count();
  return sms < 0 ? sms : sms * 4;
}
static int gate_check(const char* who, const void* p, int C, int ld, int cmax) {
  if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15u) || C < 8 || (C & 7) || (ld & 7) || ld < C || C > cmax)
    return set_error(UZ_ERR_INVALID, "%s: needs a 16-B aligned pointer, channels (=%d) and ld (=%d) multiples of 8, channels <= %d",
                     who, C, ld, cmax);
  return UZ_OK;
}

}  // namespace uz

using namespace uz;

extern "C" int uz_gate_psi_fwd(const void* zg, int zg_ld, const void* zx, int zx_ld, long long npix, int F,
                               const float* bn_g, const float* bn_x, const float* w_psi, const float* b_psi,
                               float* psi_raw, float* partial, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = gate_check("gate_psi_fwd(zg)", zg, F, zg_ld, 256);
  if (rc) return rc;
  if ((rc = gate_check("gate_psi_fwd(zx)", zx, F, zx_ld, 256))) return rc;
  if (npix < 1 || !bn_g || !bn_x || !w_psi || !psi_raw || !partial) return set_error(UZ_ERR_INVALID, "gate_psi_fwd: null argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int rows = gate_rows();
  cudaError_t e = cudaMemsetAsync(partial, 0, sizeof(float) * 2 * rows, st);
  if (e != cudaSuccess) return cuda_error(e, "gate_psi_fwd memset");
  const int lpp = gate_lpp(F), spb = kGateBlock / lpp;
  const long long need = (npix + spb - 1) / spb;
  const int grid = static_cast<int>(need < rows ? need : rows);
  const __nv_bfloat16* a = static_cast<const __nv_bfloat16*>(zg);
  const __nv_bfloat16* b = static_cast<const __nv_bfloat16*>(zx);
#define UZ_GATE_FWD(L) \
  gate_psi_fwd_kernel<L><<<grid, kGateBlock, 0, st>>>(a, zg_ld, b, zx_ld, npix, F, bn_g, bn_x, w_psi, b_psi, psi_raw, partial)
  switch (lpp) {
    case 1: UZ_GATE_FWD(1); break;
    case 2: UZ_GATE_FWD(2); break;
    case 4: UZ_GATE_FWD(4); break;
    case 8: UZ_GATE_FWD(8); break;
    case 16: UZ_GATE_FWD(16); break;
    default: UZ_GATE_FWD(32); break;
  }
#undef UZ_GATE_FWD
  PW_CHECK_LAUNCH("gate_psi_fwd launch");
  return UZ_OK;
}

extern "C" int uz_gate_apply(const void* x, int x_ld, const float* psi_raw, const float* psi_bn, long long npix, int C,
                             void* out, int out_ld, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = gate_check("gate_apply(x)", x, C, x_ld, 1 << 20);
  if (rc) return rc;
  if ((rc = gate_check("gate_apply(out)", out, C, out_ld, 1 << 20))) return rc;
  if (npix < 1 || !psi_raw || !psi_bn) return set_error(UZ_ERR_INVALID, "gate_apply: null argument");
  const int cv = C / 8;
  const long long total = npix * cv;
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(gate_rows()) * 4;
  if (blocks > cap) blocks = cap;
  gate_apply_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), x_ld, psi_raw, psi_bn, npix, cv, static_cast<__nv_bfloat16*>(out), out_ld);
  PW_CHECK_LAUNCH("gate_apply launch");
  return UZ_OK;
}

extern "C" size_t uz_gate_bwd_workspace_bytes(long long npix, int F) {
  const int rows = gate_rows();
  if (rows < 0 || npix < 1) return 0;
  // dsb[npix] | partial[rows][2] | coef[2] (padded to 4) | wpart[rows][F] | bpart[rows]
  return sizeof(float) * (static_cast<size_t>((npix + 3) / 4 * 4) + static_cast<size_t>(rows) * 2 + 4 +
                          static_cast<size_t>(rows) * (F + 1));
}

extern "C" int uz_gate_bwd(const void* x, int x_ld, const void* dout, int dout_ld, const float* psi_raw,
                           const float* psi_bn, long long npix, int C, void* dx, int dx_ld, int accumulate_dx,
                           const void* zg, int zg_ld, const void* zx, int zx_ld, int F, const float* bn_g,
                           const float* bn_x, const float* w_psi, void* dr, int dr_ld, float* dw_psi, float* db_psi,
                           float* dgamma_psi, float* dbeta_psi, void* workspace, size_t workspace_bytes, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = gate_check("gate_bwd(x)", x, C, x_ld, 1 << 20);
  if (rc) return rc;
  if ((rc = gate_check("gate_bwd(dout)", dout, C, dout_ld, 1 << 20))) return rc;
  if ((rc = gate_check("gate_bwd(dx)", dx, C, dx_ld, 1 << 20))) return rc;
  if ((rc = gate_check("gate_bwd(zg)", zg, F, zg_ld, 256))) return rc;
  if ((rc = gate_check("gate_bwd(zx)", zx, F, zx_ld, 256))) return rc;
  if ((rc = gate_check("gate_bwd(dr)", dr, F, dr_ld, 256))) return rc;
  if (!psi_raw || !psi_bn || !bn_g || !bn_x || !w_psi || !dw_psi) return set_error(UZ_ERR_INVALID, "gate_bwd: null argument");
  if (workspace_bytes < uz_gate_bwd_workspace_bytes(npix, F) || (reinterpret_cast<uintptr_t>(workspace) & 15u))
    return set_error(UZ_ERR_WORKSPACE, "gate_bwd: workspace too small (%zu < %zu)", workspace_bytes,
                     uz_gate_bwd_workspace_bytes(npix, F));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int rows = gate_rows();
  float* dsb = static_cast<float*>(workspace);
  float* partial = dsb + (npix + 3) / 4 * 4;
  float* coef = partial + static_cast<size_t>(rows) * 2;
  float* wpart = coef + 4;
  float* bpart = wpart + static_cast<size_t>(rows) * F;

  // stage 1: gate product backward + BN_psi reduction
  {
    const int lpp = gate_lpp(C) > 32 ? 32 : gate_lpp(C);
    const long long ppb = static_cast<long long>(kGateBlock / 32) * (32 / lpp);   // pixels per block and step
    const long long need = (npix + ppb - 1) / ppb;
    const int grid = static_cast<int>(need < rows ? need : rows);
#define UZ_GATE_PSI(L)                                                                                           \
  gate_bwd_psi_kernel<L><<<grid, kGateBlock, 0, st>>>(static_cast<const __nv_bfloat16*>(x), x_ld,                \
                                                      static_cast<const __nv_bfloat16*>(dout), dout_ld, psi_raw, \
                                                      psi_bn, npix, C / 8, static_cast<__nv_bfloat16*>(dx), dx_ld, \
                                                      accumulate_dx, dsb, partial)
    switch (lpp) {
      case 1: UZ_GATE_PSI(1); break;
      case 2: UZ_GATE_PSI(2); break;
      case 4: UZ_GATE_PSI(4); break;
      case 8: UZ_GATE_PSI(8); break;
      case 16: UZ_GATE_PSI(16); break;
      default: UZ_GATE_PSI(32); break;
    }
#undef UZ_GATE_PSI
    PW_CHECK_LAUNCH("gate_bwd_psi launch");
    cudaError_t e = launch_bn_bwd_finalize(partial, grid, 1, static_cast<double>(npix), psi_bn, dgamma_psi, dbeta_psi, 0,
                                           coef, st);
    if (e != cudaSuccess) return cuda_error(e, "gate_bwd finalize launch");
  }
  // stage 2: through w_psi and the ReLU into dr; dw_psi / db_psi partials
  {
    const int lpp = gate_lpp(F), spb = kGateBlock / lpp;
    const long long need = (npix + spb - 1) / spb;
    const int grid = static_cast<int>(need < rows ? need : rows);
    const size_t smem = sizeof(float) * kGateBlock * 9;
    const __nv_bfloat16* a = static_cast<const __nv_bfloat16*>(zg);
    const __nv_bfloat16* b = static_cast<const __nv_bfloat16*>(zx);
    __nv_bfloat16* d = static_cast<__nv_bfloat16*>(dr);
#define UZ_GATE_BWD(L)                                                                                              \
  gate_bwd_r_kernel<L><<<grid, kGateBlock, smem, st>>>(a, zg_ld, b, zx_ld, npix, F, bn_g, bn_x, w_psi, psi_raw, dsb, \
                                                      psi_bn, coef, d, dr_ld, wpart, bpart)
    switch (lpp) {
      case 1: UZ_GATE_BWD(1); break;
      case 2: UZ_GATE_BWD(2); break;
      case 4: UZ_GATE_BWD(4); break;
      case 8: UZ_GATE_BWD(8); break;
      case 16: UZ_GATE_BWD(16); break;
      default: UZ_GATE_BWD(32); break;
    }
#undef UZ_GATE_BWD
    PW_CHECK_LAUNCH("gate_bwd_r launch");
    cudaError_t e = launch_reduce_rows(wpart, grid, F, dw_psi, 0, st);
    if (e == cudaSuccess && db_psi) e = launch_reduce_rows(bpart, grid, 1, db_psi, 0, st);
    if (e != cudaSuccess) return cuda_error(e, "gate_bwd reduce launch");
  }
  return UZ_OK;
}
