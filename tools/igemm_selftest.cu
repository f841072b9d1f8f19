// Standalone bring-up harness for the tcgen05 implicit-GEMM kernels (no torch): random bf16 operands, naive
// CPU restatement of each convolution flavour, max-error report per case.  Usage: igemm_selftest [case_id]
// Each case runs in its own process invocation from tools/run_selftest.sh so a hang only loses one case.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../include/unet_zoo_b200.h"

#define CK(x)                                                                   \
  do {                                                                          \
    cudaError_t e_ = (x);                                                       \
    if (e_ != cudaSuccess) {                                                    \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(2);                                                                  \
    }                                                                           \
  } while (0)

static uint32_t rng_state = 12345u;
static float frand() {  // uniform in [-1, 1)
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xFFFF) / 32768.0f - 1.0f;
}
static float bf16r(float v) { return __bfloat162float(__float2bfloat16(v)); }

struct Act {  // NHWC bf16 host+device
  int N, H, W, C, ld;
  std::vector<float> h;  // values (already bf16-representable), dense [N][H][W][ld]
  __nv_bfloat16* d = nullptr;
  void alloc(int n, int hh, int w, int c, int ld_ = 0) {
    N = n; H = hh; W = w; C = c; ld = ld_ ? ld_ : c;
    h.assign((size_t)N * H * W * ld, 0.f);
    CK(cudaMalloc(&d, h.size() * 2));
  }
  void randomize(float scale = 1.f) { for (auto& v : h) v = bf16r(frand() * scale); }
  void upload() {
    std::vector<__nv_bfloat16> t(h.size());
    for (size_t i = 0; i < h.size(); ++i) t[i] = __float2bfloat16(h[i]);
    CK(cudaMemcpy(d, t.data(), t.size() * 2, cudaMemcpyHostToDevice));
  }
  void download() {
    std::vector<__nv_bfloat16> t(h.size());
    CK(cudaMemcpy(t.data(), d, t.size() * 2, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < h.size(); ++i) h[i] = __bfloat162float(t[i]);
  }
  float at(int n, int y, int x, int c) const {
    if (y < 0 || y >= H || x < 0 || x >= W) return 0.f;
    return h[(((size_t)n * H + y) * W + x) * ld + c];
  }
  float& ref(int n, int y, int x, int c) { return h[(((size_t)n * H + y) * W + x) * ld + c]; }
};

static __nv_bfloat16* upload_bf16(const std::vector<float>& v) {
  std::vector<__nv_bfloat16> t(v.size());
  for (size_t i = 0; i < v.size(); ++i) t[i] = __float2bfloat16(v[i]);
  __nv_bfloat16* d;
  CK(cudaMalloc(&d, t.size() * 2));
  CK(cudaMemcpy(d, t.data(), t.size() * 2, cudaMemcpyHostToDevice));
  return d;
}

static int report(const char* name, double max_err, double max_ref, double tol) {
  const bool ok = max_err <= tol * (max_ref > 1.0 ? max_ref : 1.0);
  printf("[%s] %-44s max_err=%.5f max_ref=%.3f\n", ok ? "PASS" : "FAIL", name, max_err, max_ref);
  fflush(stdout);
  return ok ? 0 : 1;
}

// ---------------------------------------------------------------------------------------------- forward-type cases
// weights in "parameter" form W[co][ci][kh][kw] (conv) for all conv modes; convT uses Wt[ci][co][2][2].
struct FwdCase {
  const char* name;
  int mode, N, H, W;
  int nseg, segc[3];
  int cout;
  int flags;
  bool bias;
  int out_ld_extra;
};

static int run_fwd_case(const FwdCase& c) {
  printf("... %s\n", c.name);
  fflush(stdout);
  int cin = 0;
  for (int s = 0; s < c.nseg; ++s) cin += c.segc[s];
  const bool is_convt = (c.mode == UZ_CONVT2X2 || c.mode == UZ_CONVT2X2_DGRAD);
  int kh = 3;
  if (c.mode == UZ_CONV1X1 || c.mode == UZ_CONV1X1_S2 || c.mode == UZ_CONV1X1_S2_DGRAD) kh = 1;
  if (is_convt) kh = 2;
  const int T = kh * kh;
  // "forward" parameter orientation: for dgrad modes the op maps (channels = fwd Cout) -> (fwd Cin)
  const bool dgrad = (c.mode == UZ_CONV3X3_DGRAD || c.mode == UZ_CONV3X3_S2_DGRAD || c.mode == UZ_CONV1X1_S2_DGRAD ||
                      c.mode == UZ_CONVT2X2_DGRAD);
  // Wf[o][i][t]: o = op output channel, i = op input channel (already arranged for this op's direction)
  std::vector<float> Wf((size_t)c.cout * cin * T);
  for (auto& v : Wf) v = bf16r(frand() * 0.25f);
  // packed [cout][T][cin]
  std::vector<float> Wp((size_t)c.cout * T * cin);
  for (int o = 0; o < c.cout; ++o)
    for (int t = 0; t < T; ++t)
      for (int i = 0; i < cin; ++i) Wp[((size_t)o * T + t) * cin + i] = Wf[((size_t)o * cin + i) * T + t];
  __nv_bfloat16* dW = upload_bf16(Wp);

  std::vector<Act> segs(c.nseg);
  std::vector<const void*> ptrs(c.nseg);
  std::vector<int> sc(c.nseg), sld(c.nseg);
  for (int s = 0; s < c.nseg; ++s) {
    segs[s].alloc(c.N, c.H, c.W, c.segc[s], c.segc[s] + (s == 0 ? 8 : 0));  // first segment strided (slice view)
    segs[s].randomize();
    segs[s].upload();
    ptrs[s] = segs[s].d;
    sc[s] = c.segc[s];
    sld[s] = segs[s].ld;
  }
  auto in_at = [&](int n, int y, int x, int ci) -> float {
    int s = 0;
    while (ci >= c.segc[s]) { ci -= c.segc[s]; ++s; }
    return segs[s].at(n, y, x, ci);
  };
  int Ho = c.H, Wo = c.W;
  switch (c.mode) {
    case UZ_CONV3X3_S2: case UZ_CONV1X1_S2: Ho = (c.H - 1) / 2 + 1; Wo = (c.W - 1) / 2 + 1; break;
    case UZ_CONV3X3_S2_DGRAD: case UZ_CONV1X1_S2_DGRAD: case UZ_CONVT2X2: case UZ_UP2X_CONV3X3: Ho = 2 * c.H; Wo = 2 * c.W; break;
    case UZ_CONVT2X2_DGRAD: Ho = c.H / 2; Wo = c.W / 2; break;
    default: break;
  }
  Act out;
  out.alloc(c.N, Ho, Wo, c.cout, c.cout + c.out_ld_extra);
  out.randomize();  // pre-existing content (matters for ACCUMULATE and for untouched padding channels)
  out.upload();
  std::vector<float> before = out.h;
  std::vector<float> bias(c.cout, 0.f);
  float* dbias = nullptr;
  if (c.bias) {
    for (auto& v : bias) v = frand();
    CK(cudaMalloc(&dbias, c.cout * 4));
    CK(cudaMemcpy(dbias, bias.data(), c.cout * 4, cudaMemcpyHostToDevice));
  }
  float* dstats = nullptr;
  const int rows = uz_conv_stats_rows();
  if (c.flags & UZ_FLAG_STATS) CK(cudaMalloc(&dstats, (size_t)rows * 2 * c.cout * 4));

  int rc = uz_conv_igemm(c.mode, c.N, c.H, c.W, c.nseg, ptrs.data(), sc.data(), sld.data(), dW, c.cout, dbias, out.d,
                         out.ld, Ho, Wo, dstats, c.flags, nullptr);
  if (rc != 0) {
    printf("[FAIL] %s: uz_conv_igemm rc=%d (%s)\n", c.name, rc, uz_last_error());
    return 1;
  }
  CK(cudaDeviceSynchronize());
  out.download();

  // CPU restatement
  double max_err = 0, max_ref = 0;
  std::vector<double> ssum(c.cout, 0.0), ssq(c.cout, 0.0);
  for (int n = 0; n < c.N; ++n)
    for (int oy = 0; oy < Ho; ++oy)
      for (int ox = 0; ox < Wo; ++ox)
        for (int o = 0; o < c.cout; ++o) {
          double acc = 0;
          bool written = true;
          auto W_at = [&](int i, int r, int s) { return Wf[((size_t)o * cin + i) * T + r * kh + s]; };
          switch (c.mode) {
            case UZ_CONV3X3:
              for (int r = 0; r < 3; ++r) for (int s = 0; s < 3; ++s) for (int i = 0; i < cin; ++i)
                acc += (double)in_at(n, oy + r - 1, ox + s - 1, i) * W_at(i, r, s);
              break;
            case UZ_CONV3X3_DGRAD:
              for (int r = 0; r < 3; ++r) for (int s = 0; s < 3; ++s) for (int i = 0; i < cin; ++i)
                acc += (double)in_at(n, oy + 1 - r, ox + 1 - s, i) * W_at(i, r, s);
              break;
            case UZ_CONV1X1:
              for (int i = 0; i < cin; ++i) acc += (double)in_at(n, oy, ox, i) * W_at(i, 0, 0);
              break;
            case UZ_CONV3X3_S2:
              for (int r = 0; r < 3; ++r) for (int s = 0; s < 3; ++s) for (int i = 0; i < cin; ++i)
                acc += (double)in_at(n, 2 * oy + r - 1, 2 * ox + s - 1, i) * W_at(i, r, s);
              break;
            case UZ_CONV3X3_S2_DGRAD:
              // dX[oy,ox] = sum_{r,s} dY[(oy+1-r)/2, (ox+1-s)/2] * W[r,s] when divisible
              for (int r = 0; r < 3; ++r) for (int s = 0; s < 3; ++s) {
                const int a = oy + 1 - r, b = ox + 1 - s;
                if (a < 0 || b < 0 || (a & 1) || (b & 1)) continue;
                for (int i = 0; i < cin; ++i) acc += (double)in_at(n, a / 2, b / 2, i) * W_at(i, r, s);
              }
              break;
            case UZ_CONV1X1_S2:
              for (int i = 0; i < cin; ++i) acc += (double)in_at(n, 2 * oy, 2 * ox, i) * W_at(i, 0, 0);
              break;
            case UZ_CONV1X1_S2_DGRAD:
              if ((oy & 1) || (ox & 1)) { written = false; break; }
              for (int i = 0; i < cin; ++i) acc += (double)in_at(n, oy / 2, ox / 2, i) * W_at(i, 0, 0);
              break;
            case UZ_CONVT2X2:
              for (int i = 0; i < cin; ++i) acc += (double)in_at(n, oy / 2, ox / 2, i) * W_at(i, oy & 1, ox & 1);
              break;
            case UZ_CONVT2X2_DGRAD:
              for (int r = 0; r < 2; ++r) for (int s = 0; s < 2; ++s) for (int i = 0; i < cin; ++i)
                acc += (double)in_at(n, 2 * oy + r, 2 * ox + s, i) * W_at(i, r, s);
              break;
            case UZ_UP2X_CONV3X3:
              for (int r = 0; r < 3; ++r) for (int s = 0; s < 3; ++s) {
                const int uy = oy + r - 1, ux = ox + s - 1;
                if (uy < 0 || ux < 0 || uy >= Ho || ux >= Wo) continue;
                for (int i = 0; i < cin; ++i) acc += (double)in_at(n, uy >> 1, ux >> 1, i) * W_at(i, r, s);
              }
              break;
          }
          const size_t oi = (((size_t)n * Ho + oy) * Wo + ox) * out.ld + o;
          double ref;
          if (!written) {
            ref = before[oi];
          } else {
            ref = acc + (c.bias ? bias[o] : 0.0);
            if (c.flags & UZ_FLAG_ACCUMULATE) ref += before[oi];
            if (c.flags & UZ_FLAG_RELU) ref = ref > 0 ? ref : 0;
          }
          const double got = out.h[oi];
          ssum[o] += got;
          ssq[o] += got * got;
          max_err = fmax(max_err, fabs(got - ref) - 0.008 * fabs(ref));  // bf16 output rounding allowance
          max_ref = fmax(max_ref, fabs(ref));
        }
  // padding channels must be untouched
  int bad_pad = 0;
  for (size_t px = 0; px < (size_t)c.N * Ho * Wo; ++px)
    for (int o = c.cout; o < out.ld; ++o)
      if (out.h[px * out.ld + o] != before[px * out.ld + o]) ++bad_pad;
  int fails = report(c.name, max_err, max_ref, 0.01);
  if (bad_pad) { printf("[FAIL] %s: %d padding elements overwritten\n", c.name, bad_pad); ++fails; }
  if (c.flags & UZ_FLAG_STATS) {
    std::vector<float> st((size_t)rows * 2 * c.cout);
    CK(cudaMemcpy(st.data(), dstats, st.size() * 4, cudaMemcpyDeviceToHost));
    double e = 0, m = 0;
    for (int o = 0; o < c.cout; ++o) {
      double s = 0, q = 0;
      for (int r = 0; r < rows; ++r) { s += st[(size_t)r * 2 * c.cout + o]; q += st[(size_t)r * 2 * c.cout + c.cout + o]; }
      e = fmax(e, fabs(s - ssum[o]) / (1.0 + fabs(ssum[o])));
      e = fmax(e, fabs(q - ssq[o]) / (1.0 + fabs(ssq[o])));
      m = fmax(m, fabs(ssq[o]));
    }
    char nm[128];
    snprintf(nm, sizeof(nm), "%s [stats]", c.name);
    fails += report(nm, e, 1.0, 1e-3);
  }
  return fails;
}

// ---------------------------------------------------------------------------------------------- wgrad cases
struct WgCase {
  const char* name;
  int mode, N, H, W, cin, cout;
};

static int run_wg_case(const WgCase& c) {
  printf("... %s\n", c.name);
  fflush(stdout);
  int Hy = c.H, Wy = c.W;  // dY spatial size
  int taps = 9;
  switch (c.mode) {
    case UZ_CONV1X1: taps = 1; break;
    case UZ_CONV3X3_S2: Hy = (c.H - 1) / 2 + 1; Wy = (c.W - 1) / 2 + 1; break;
    case UZ_CONV1X1_S2: taps = 1; Hy = (c.H - 1) / 2 + 1; Wy = (c.W - 1) / 2 + 1; break;
    case UZ_UP2X_CONV3X3: Hy = 2 * c.H; Wy = 2 * c.W; break;
    case UZ_CONVT2X2: taps = 4; Hy = 2 * c.H; Wy = 2 * c.W; break;
    default: break;
  }
  Act x, dy;
  x.alloc(c.N, c.H, c.W, c.cin, c.cin + 8);
  dy.alloc(c.N, Hy, Wy, c.cout);
  x.randomize();
  dy.randomize(0.5f);
  x.upload();
  dy.upload();
  const size_t ndw = (size_t)c.cin * c.cout * taps;
  float* ddw;
  CK(cudaMalloc(&ddw, ndw * 4));
  CK(cudaMemset(ddw, 0xff, ndw * 4));
  const size_t ws = uz_conv_wgrad_workspace_bytes(c.mode, c.N, c.H, c.W, c.cin, c.cout);
  void* dws;
  CK(cudaMalloc(&dws, ws ? ws : 16));
  int rc = uz_conv_wgrad(c.mode, c.N, c.H, c.W, x.d, c.cin, x.ld, dy.d, c.cout, dy.ld, ddw, c.cin, 0, 0, dws, ws, nullptr);
  if (rc != 0) {
    printf("[FAIL] %s: uz_conv_wgrad rc=%d (%s)\n", c.name, rc, uz_last_error());
    return 1;
  }
  CK(cudaDeviceSynchronize());
  std::vector<float> dw(ndw);
  CK(cudaMemcpy(dw.data(), ddw, ndw * 4, cudaMemcpyDeviceToHost));
  double max_err = 0, max_ref = 0;
  const int kh = taps == 9 ? 3 : taps == 4 ? 2 : 1;
  for (int co = 0; co < c.cout; ++co)
    for (int ci = 0; ci < c.cin; ++ci)
      for (int r = 0; r < kh; ++r)
        for (int s = 0; s < kh; ++s) {
          double acc = 0;
          for (int n = 0; n < c.N; ++n)
            for (int oy = 0; oy < Hy; ++oy)
              for (int ox = 0; ox < Wy; ++ox) {
                float xv = 0.f;
                switch (c.mode) {
                  case UZ_CONV3X3: xv = x.at(n, oy + r - 1, ox + s - 1, ci); break;
                  case UZ_CONV1X1: xv = x.at(n, oy, ox, ci); break;
                  case UZ_CONV3X3_S2: xv = x.at(n, 2 * oy + r - 1, 2 * ox + s - 1, ci); break;
                  case UZ_CONV1X1_S2: xv = x.at(n, 2 * oy, 2 * ox, ci); break;
                  case UZ_UP2X_CONV3X3: {
                    const int uy = oy + r - 1, ux = ox + s - 1;
                    xv = (uy < 0 || ux < 0 || uy >= Hy || ux >= Wy) ? 0.f : x.at(n, uy >> 1, ux >> 1, ci);
                    break;
                  }
                  case UZ_CONVT2X2: xv = ((oy & 1) == r && (ox & 1) == s) ? x.at(n, oy >> 1, ox >> 1, ci) : 0.f; break;
                }
                acc += (double)xv * dy.at(n, oy, ox, co);
              }
          const size_t idx = (c.mode == UZ_CONVT2X2) ? (((size_t)ci * c.cout + co) * 4 + r * 2 + s)
                                                     : (((size_t)co * c.cin + ci) * taps + r * kh + s);
          max_err = fmax(max_err, fabs(dw[idx] - acc));
          max_ref = fmax(max_ref, fabs(acc));
        }
  return report(c.name, max_err, max_ref, 2e-3);
}

int main(int argc, char** argv) {
  const int sms = uz_device_sm_count();
  if (sms < 0) {
    printf("no sm_100 device: %s\n", uz_last_error());
    return 3;
  }
  const FwdCase fwd[] = {
      {"conv3x3 64->64 32x32 N2", UZ_CONV3X3, 2, 32, 32, 1, {64}, 64, 0, false, 0},
      {"conv3x3 128->128 20x24 N3 bias relu", UZ_CONV3X3, 3, 20, 24, 1, {128}, 128, UZ_FLAG_RELU, true, 8},
      {"conv3x3 (64+32)->256 16x16 N2 stats", UZ_CONV3X3, 2, 16, 16, 2, {64, 32}, 256, UZ_FLAG_STATS, false, 0},
      {"conv3x3 64->512 8x8 N2 (2 n-tiles) stats", UZ_CONV3X3, 2, 8, 8, 1, {64}, 512, UZ_FLAG_STATS, false, 0},
      {"conv3x3 32->32 4x4 N5 (multi-image tile)", UZ_CONV3X3, 5, 4, 4, 1, {32}, 32, 0, false, 0},
      {"conv3x3 dgrad 128->64 32x32 N2 accumulate", UZ_CONV3X3_DGRAD, 2, 32, 32, 1, {128}, 64, UZ_FLAG_ACCUMULATE, false, 64},
      {"conv1x1 64->32 16x16 N2 bias", UZ_CONV1X1, 2, 16, 16, 1, {64}, 32, 0, true, 0},
      {"conv3x3 s2 64->128 32x32 N2", UZ_CONV3X3_S2, 2, 32, 32, 1, {64}, 128, 0, false, 0},
      {"conv3x3 s2 dgrad 128->64 16x16->32x32 N2", UZ_CONV3X3_S2_DGRAD, 2, 16, 16, 1, {128}, 64, 0, false, 0},
      {"conv1x1 s2 64->128 32x32 N2", UZ_CONV1X1_S2, 2, 32, 32, 1, {64}, 128, 0, false, 0},
      {"conv1x1 s2 dgrad 128->64 N2 accumulate", UZ_CONV1X1_S2_DGRAD, 2, 16, 16, 1, {128}, 64, UZ_FLAG_ACCUMULATE, false, 0},
      {"convT2x2 128->64 16x16 N2 bias (-> cat slice)", UZ_CONVT2X2, 2, 16, 16, 1, {128}, 64, 0, true, 64},
      {"convT2x2 dgrad 64->128 32x32 N2", UZ_CONVT2X2_DGRAD, 2, 32, 32, 1, {64}, 128, 0, false, 0},
      {"up2x+conv3x3 128->64 16x16 N2", UZ_UP2X_CONV3X3, 2, 16, 16, 1, {128}, 64, 0, false, 0},
      {"conv3x3 64->64 64x64 N5 (multi-wave) stats", UZ_CONV3X3, 5, 64, 64, 1, {64}, 64, UZ_FLAG_STATS, false, 0},
      {"fwd2 conv3x3 256->256 32x32 N2 ring stats bias", UZ_CONV3X3, 2, 32, 32, 1, {256}, 256, UZ_FLAG_STATS, true, 0},
      {"fwd2 conv3x3 64->512 16x24 N3 (2 n-tiles) stats relu", UZ_CONV3X3, 3, 16, 24, 1, {64}, 512, UZ_FLAG_STATS | UZ_FLAG_RELU, false, 64},
      {"fwd2 conv3x3 dgrad 256->128 24x40 N2 ragged", UZ_CONV3X3_DGRAD, 2, 24, 40, 1, {256}, 128, 0, false, 0},
      {"fwd2 conv3x3 (128+128)->128 64x64 N4 ring", UZ_CONV3X3, 4, 64, 64, 2, {128, 128}, 128, UZ_FLAG_STATS, false, 0},
  };
  const WgCase wg[] = {
      {"wgrad conv3x3 64->64 32x32 N2", UZ_CONV3X3, 2, 32, 32, 64, 64},
      {"wgrad conv3x3 128->256 20x24 N3", UZ_CONV3X3, 3, 20, 24, 128, 256},
      {"wgrad conv3x3 32->32 8x8 N3", UZ_CONV3X3, 3, 8, 8, 32, 32},
      {"wgrad conv1x1 64->32 16x16 N2", UZ_CONV1X1, 2, 16, 16, 64, 32},
      {"wgrad conv3x3 s2 64->128 32x32 N2", UZ_CONV3X3_S2, 2, 32, 32, 64, 128},
      {"wgrad conv1x1 s2 64->128 32x32 N2", UZ_CONV1X1_S2, 2, 32, 32, 64, 128},
      {"wgrad convT2x2 128->64 16x16 N2", UZ_CONVT2X2, 2, 16, 16, 128, 64},
      {"wgrad up2x+conv3x3 128->64 16x16 N2", UZ_UP2X_CONV3X3, 2, 16, 16, 128, 64},
      {"wgrad2 conv3x3 64->64 16x8 N1 (one chunk)", UZ_CONV3X3, 1, 16, 8, 64, 64},
      {"wgrad2 conv3x3 192->96 24x40 N2 (ragged)", UZ_CONV3X3, 2, 24, 40, 192, 96},
      {"wgrad2 conv3x3 64->128 64x64 N3 (many chunks)", UZ_CONV3X3, 3, 64, 64, 64, 128},
  };
  const int nf = sizeof(fwd) / sizeof(fwd[0]), nw = sizeof(wg) / sizeof(wg[0]);
  if (argc > 1 && strcmp(argv[1], "count") == 0) {
    printf("%d\n", nf + nw);
    return 0;
  }
  int fails = 0;
  const int only = argc > 1 ? atoi(argv[1]) : -1;
  for (int i = 0; i < nf + nw; ++i) {
    if (only >= 0 && i != only) continue;
    rng_state = 777u + i;
    fails += i < nf ? run_fwd_case(fwd[i]) : run_wg_case(wg[i - nf]);
  }
  return fails ? 1 : 0;
}
