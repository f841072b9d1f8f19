// Last DoubleConv half + classifier head without materialising the activation between them (sm_100a).
//
//   y = ReLU(BN(z)),  logits[n,k,h,w] = b[k] + sum_c w[k,c] y[n,h,w,c]          (common_layers.py:28-33 + 125-128)
//
// The widest activation of the network (64 channels at full resolution) is consumed only by the 1x1 head, and the
// head's data gradient dy[p,c] = sum_k dlogits[p,k] w[k,c] is a rank-K outer product.  So neither y nor dy has to
// exist in HBM: the forward reads z once and writes the fp32 logits; the backward recomputes y from z, forms dy on the
// fly from dlogits (4 B per pixel and class) and feeds the BatchNorm backward directly.  Bytes per pixel-channel:
// 2 (fwd) + 2 (reduce) + 4 (apply) = 8 instead of 20 (bn_apply 4, head_fwd 2, head_bwd 4, bn_bwd 10).
//
// Thread layout as in the other bandwidth kernels: a thread owns 8 consecutive channels (one 16-byte word) of a pixel,
// the cv = C/8 threads of a pixel are adjacent lanes of one warp (cv is a power of two <= 32), U pixels in flight per
// thread.
#include "pw.cuh"

namespace uz {
namespace {

constexpr int kBhThreads = 256;
constexpr int kBhUnroll = 4;

struct BhArgs {
  const __nv_bfloat16* z;
  int z_ld;
  long long npix;
  long long hw;        // H*W: logits are NCHW, logits[(n*K + k)*hw + pixel-in-image]
  int C, cv, ppb, relu;
  const float* scale;  // [C]
  const float* shift;  // [C]
  const float* w;      // [K][C]
};

template <int K>
__device__ __forceinline__ void bh_load_w(const BhArgs& a, int g, float (&w)[K][8]) {
#pragma unroll
  for (int k = 0; k < K; ++k) pw_load8f(a.w + static_cast<size_t>(k) * a.C + g * 8, w[k]);
}

// y (fp32, not rounded: it never leaves the registers) of one 16-byte word of z
__device__ __forceinline__ void bh_act(const uint4& raw, const float (&sc)[8], const float (&sh)[8], int relu,
                                       float (&zv)[8], float (&yv)[8]) {
  pw_unpack8(raw, zv);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float t = fmaf(zv[j], sc[j], sh[j]);
    yv[j] = relu ? fmaxf(t, 0.f) : t;
  }
}

template <int K>
__global__ void __launch_bounds__(kBhThreads, K <= 2 ? 3 : 2) bn_head_fwd_kernel(const BhArgs a, const float* __restrict__ bias,
                                                                    float* __restrict__ logits) {
  const int g = threadIdx.x % a.cv;
  const int pl = threadIdx.x / a.cv;
  float sc[8], sh[8], w[K][8];
  pw_load8f(a.scale + g * 8, sc);
  pw_load8f(a.shift + g * 8, sh);
  bh_load_w<K>(a, g, w);
  float b[K];
#pragma unroll
  for (int k = 0; k < K; ++k) b[k] = bias ? __ldg(bias + k) : 0.f;
  const long long stride = static_cast<long long>(gridDim.x) * a.ppb;
  constexpr int U = kBhUnroll;
  // block-uniform trip count (the loop bound is the block's first pixel), so the shuffles below are convergent
  for (long long base = static_cast<long long>(blockIdx.x) * a.ppb; base < a.npix; base += U * stride) {
    const long long it = base + pl;
    uint4 raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (it + u * stride < a.npix) raw[u] = pw_ldg16(a.z + (it + u * stride) * a.z_ld + g * 8);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long p = it + u * stride;
      const bool live = p < a.npix;
      float zv[8], yv[8];
      if (live) bh_act(raw[u], sc, sh, a.relu, zv, yv);
      float dot[K];
#pragma unroll
      for (int k = 0; k < K; ++k) {
        float d = 0.f;
        if (live) {
#pragma unroll
          for (int j = 0; j < 8; ++j) d = fmaf(yv[j], w[k][j], d);
        }
        for (int o = a.cv >> 1; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        dot[k] = d;
      }
      if (live && g == 0) {
        const long long n = p / a.hw, r = p - n * a.hw;
#pragma unroll
        for (int k = 0; k < K; ++k) logits[(n * K + k) * a.hw + r] = dot[k] + b[k];
      }
    }
  }
}

template <int K>
__device__ __forceinline__ void bh_load_dl(const float* __restrict__ dlogits, long long p, long long hw, float (&dl)[K]) {
  if (K == 1) {
    dl[0] = __ldg(dlogits + p);
  } else {
    const long long n = p / hw, r = p - n * hw;
#pragma unroll
    for (int k = 0; k < K; ++k) dl[k] = __ldg(dlogits + (n * K + k) * hw + r);
  }
}

// gradient w.r.t. y of one word: gy[j] = sum_k dl[k] w[k][j], masked by the ReLU
template <int K>
__device__ __forceinline__ void bh_grad(const float (&dl)[K], const float (&w)[K][8], const float (&yv)[8], int relu,
                                        float (&gv)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) t = fmaf(dl[k], w[k][j], t);
    gv[j] = (relu && !(yv[j] > 0.f)) ? 0.f : t;
  }
}

// partial[block][2][C] = {sum g, sum g*z};  hwpart[block][K][C] = sum dl[k]*y;  dbpart[block][K] = sum dl[k]
template <int K>
__global__ void __launch_bounds__(kBhThreads, 2) bn_head_bwd_reduce_kernel(const BhArgs a,
                                                                           const float* __restrict__ dlogits,
                                                                           float* __restrict__ partial,
                                                                           float* __restrict__ hwpart,
                                                                           float* __restrict__ dbpart) {
  extern __shared__ float red[];
  constexpr int kPer = 16 + 8 * K + K;   // floats per thread in the block reduction
  const int g = threadIdx.x % a.cv;
  const int pl = threadIdx.x / a.cv;
  float sc[8], sh[8], w[K][8];
  pw_load8f(a.scale + g * 8, sc);
  pw_load8f(a.shift + g * 8, sh);
  bh_load_w<K>(a, g, w);
  float s[8], q[8], hwacc[K][8], db[K];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    db[k] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) hwacc[k][j] = 0.f;
  }
  const long long stride = static_cast<long long>(gridDim.x) * a.ppb;
  constexpr int U = kBhUnroll;
  for (long long it = static_cast<long long>(blockIdx.x) * a.ppb + pl; it < a.npix; it += U * stride) {
    uint4 raw[U];
    float dl[U][K];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (it + u * stride < a.npix) {
        raw[u] = pw_ldg16(a.z + (it + u * stride) * a.z_ld + g * 8);
        bh_load_dl<K>(dlogits, it + u * stride, a.hw, dl[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (it + u * stride < a.npix) {
        float zv[8], yv[8], gv[8];
        bh_act(raw[u], sc, sh, a.relu, zv, yv);
        bh_grad<K>(dl[u], w, yv, a.relu, gv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          s[j] += gv[j];
          q[j] = fmaf(gv[j], zv[j], q[j]);
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
          db[k] += dl[u][k];
#pragma unroll
          for (int j = 0; j < 8; ++j) hwacc[k][j] = fmaf(dl[u][k], yv[j], hwacc[k][j]);
        }
      }
    }
  }
  float* mine = red + static_cast<size_t>(threadIdx.x) * kPer;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mine[j] = s[j];
    mine[8 + j] = q[j];
  }
#pragma unroll
  for (int k = 0; k < K; ++k) {
    mine[16 + 8 * K + k] = db[k];
#pragma unroll
    for (int j = 0; j < 8; ++j) mine[16 + 8 * k + j] = hwacc[k][j];
  }
  __syncthreads();
  if (pl == 0) {
    for (int r = 1; r < a.ppb; ++r) {
      const float* o = red + (static_cast<size_t>(r) * a.cv + g) * kPer;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s[j] += o[j];
        q[j] += o[8 + j];
      }
#pragma unroll
      for (int k = 0; k < K; ++k) {
        db[k] += o[16 + 8 * K + k];
#pragma unroll
        for (int j = 0; j < 8; ++j) hwacc[k][j] += o[16 + 8 * k + j];
      }
    }
    float* dst = partial + static_cast<size_t>(blockIdx.x) * 2 * a.C;
    float* hdst = hwpart + static_cast<size_t>(blockIdx.x) * K * a.C;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      dst[g * 8 + j] = s[j];
      dst[a.C + g * 8 + j] = q[j];
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
#pragma unroll
      for (int j = 0; j < 8; ++j) hdst[k * a.C + g * 8 + j] = hwacc[k][j];
      if (g == 0) dbpart[static_cast<size_t>(blockIdx.x) * K + k] = db[k];   // every channel group saw every pixel
    }
  }
}

// dz = scale*g + A + B*z  with g rebuilt from dlogits (same coefficients as bn_bwd_apply_kernel)
template <int K>
__global__ void __launch_bounds__(kBhThreads, 2) bn_head_bwd_apply_kernel(const BhArgs a,
                                                                          const float* __restrict__ dlogits,
                                                                          const float* __restrict__ coef,
                                                                          __nv_bfloat16* __restrict__ dz, int dz_ld) {
  const int g = threadIdx.x % a.cv;
  const int pl = threadIdx.x / a.cv;
  float sc[8], sh[8], ca[8], cb[8], w[K][8];
  pw_load8f(a.scale + g * 8, sc);
  pw_load8f(a.shift + g * 8, sh);
  pw_load8f(coef + g * 8, ca);
  pw_load8f(coef + a.C + g * 8, cb);
  bh_load_w<K>(a, g, w);
  const long long stride = static_cast<long long>(gridDim.x) * a.ppb;
  constexpr int U = kBhUnroll;
  for (long long it = static_cast<long long>(blockIdx.x) * a.ppb + pl; it < a.npix; it += U * stride) {
    uint4 raw[U];
    float dl[U][K];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (it + u * stride < a.npix) {
        raw[u] = pw_ldg16(a.z + (it + u * stride) * a.z_ld + g * 8);
        bh_load_dl<K>(dlogits, it + u * stride, a.hw, dl[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (it + u * stride < a.npix) {
        float zv[8], yv[8], gv[8], o[8];
        bh_act(raw[u], sc, sh, a.relu, zv, yv);
        bh_grad<K>(dl[u], w, yv, a.relu, gv);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = fmaf(sc[j], gv[j], fmaf(cb[j], zv[j], ca[j]));
        *reinterpret_cast<uint4*>(dz + (it + u * stride) * dz_ld + g * 8) = pw_pack8(o);
      }
    }
  }
}

template <typename Kern>
int bh_grid(Kern kernel, size_t smem, long long need, int cap_per_sm) {
  int per_sm = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kBhThreads, smem) != cudaSuccess || per_sm < 1)
    per_sm = 1;
  if (per_sm > cap_per_sm) per_sm = cap_per_sm;
  const long long cap = static_cast<long long>(uz_device_sm_count()) * per_sm;
  return static_cast<int>(need < cap ? (need < 1 ? 1 : need) : cap);
}

int bh_check(const char* what, const void* z, int z_ld, int C, int K) {
  if (C < 8 || C > 256 || (C & (C - 1)) != 0)
    return set_error(UZ_ERR_INVALID, "%s: C = %d must be a power of two in [8, 256]", what, C);
  if (K < 1 || K > 4) return set_error(UZ_ERR_INVALID, "%s: K = %d classes (1..4 supported)", what, K);
  if (z == nullptr || (reinterpret_cast<uintptr_t>(z) & 15) || (z_ld & 7) || z_ld < C)
    return set_error(UZ_ERR_INVALID, "%s: activation needs a 16-byte aligned pointer and ld %% 8 == 0, ld >= C", what);
  return UZ_OK;
}

BhArgs bh_args(const void* z, int z_ld, int N, int H, int W, int C, const float* scale, const float* shift, int relu,
               const float* w) {
  BhArgs a;
  a.z = static_cast<const __nv_bfloat16*>(z);
  a.z_ld = z_ld;
  a.npix = static_cast<long long>(N) * H * W;
  a.hw = static_cast<long long>(H) * W;
  a.C = C;
  a.cv = C / 8;
  a.ppb = kBhThreads / a.cv;
  a.relu = relu;
  a.scale = scale;
  a.shift = shift;
  a.w = w;
  return a;
}

template <int K>
int bh_fwd(const BhArgs& a, const float* b, float* logits, cudaStream_t st) {
  const long long need = (a.npix + a.ppb - 1) / a.ppb;
  const int grid = bh_grid(bn_head_fwd_kernel<K>, 0, need, 8);
  bn_head_fwd_kernel<K><<<grid, kBhThreads, 0, st>>>(a, b, logits);
  PW_CHECK_LAUNCH("bn_head_fwd launch");
  return UZ_OK;
}

template <int K>
int bh_bwd(const BhArgs& a, const float* bn, const float* dlogits, __nv_bfloat16* dz, int dz_ld, float* dgamma,
           float* dbeta, float* dbias, float* dw, float* db, int accumulate, float* ws, cudaStream_t st) {
  const int max_rows = uz_reduce_rows();
  float* partial = ws;
  float* hwpart = partial + static_cast<size_t>(max_rows) * 2 * a.C;
  float* dbpart = hwpart + static_cast<size_t>(max_rows) * K * a.C;
  float* coef = dbpart + static_cast<size_t>(max_rows) * K;
  const size_t smem = sizeof(float) * kBhThreads * (16 + 8 * K + K);
  static bool configured_dev[kMaxDevices] = {false};   // the attribute is per device
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(bn_head_bwd_reduce_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return cuda_error(e, "bn_head_bwd_reduce attribute");
    configured = true;
  }
  const long long need = (a.npix + a.ppb - 1) / a.ppb;
  const int rows = bh_grid(bn_head_bwd_reduce_kernel<K>, smem, need, 4);
  bn_head_bwd_reduce_kernel<K><<<rows, kBhThreads, smem, st>>>(a, dlogits, partial, hwpart, dbpart);
  PW_CHECK_LAUNCH("bn_head_bwd_reduce launch");
  cudaError_t e = launch_bn_bwd_finalize(partial, rows, a.C, static_cast<double>(a.npix), bn, dgamma, dbeta, accumulate,
                                         coef, st, 1, dbias);
  if (e != cudaSuccess) return cuda_error(e, "bn_head finalize launch");
  if (dw && (e = launch_reduce_rows(hwpart, rows, K * a.C, dw, accumulate, st)) != cudaSuccess)
    return cuda_error(e, "bn_head dw reduce launch");
  if (db && (e = launch_reduce_rows(dbpart, rows, K, db, accumulate, st)) != cudaSuccess)
    return cuda_error(e, "bn_head db reduce launch");
  const int grid = bh_grid(bn_head_bwd_apply_kernel<K>, 0, need, 8);
  bn_head_bwd_apply_kernel<K><<<grid, kBhThreads, 0, st>>>(a, dlogits, coef, dz, dz_ld);
  PW_CHECK_LAUNCH("bn_head_bwd_apply launch");
  return UZ_OK;
}

}  // namespace
}  // namespace uz

using namespace uz;

extern "C" size_t uz_bn_head_workspace_bytes(int C, int K) {
  const int rows = uz_reduce_rows();
  if (rows < 0) return 0;
  return sizeof(float) * (static_cast<size_t>(rows) * ((2 + K) * C + K) + 2 * C);
}

extern "C" int uz_bn_head_fwd(const void* z, int z_ld, int N, int H, int W, int C, const float* scale,
                              const float* shift, int relu, int K, const float* w, const float* b, float* logits,
                              void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = bh_check("bn_head_fwd", z, z_ld, C, K);
  if (rc) return rc;
  if (!scale || !shift || !w || !logits) return set_error(UZ_ERR_INVALID, "bn_head_fwd: null argument");
  const BhArgs a = bh_args(z, z_ld, N, H, W, C, scale, shift, relu, w);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (K) {
    case 1: return bh_fwd<1>(a, b, logits, st);
    case 2: return bh_fwd<2>(a, b, logits, st);
    case 3: return bh_fwd<3>(a, b, logits, st);
    default: return bh_fwd<4>(a, b, logits, st);
  }
}

extern "C" int uz_bn_head_bwd(const void* z, int z_ld, int N, int H, int W, int C, const float* bn, int relu, int K,
                              const float* w, const float* dlogits, void* dz, int dz_ld, float* dgamma, float* dbeta,
                              float* dbias, float* dw, float* db, int accumulate, void* workspace,
                              size_t workspace_bytes, void* stream) {
  PW_REQUIRE_DEVICE();
  int rc = bh_check("bn_head_bwd", z, z_ld, C, K);
  if (rc) return rc;
  if ((rc = bh_check("bn_head_bwd(dz)", dz, dz_ld, C, K))) return rc;
  if (!bn || !w || !dlogits) return set_error(UZ_ERR_INVALID, "bn_head_bwd: null argument");
  if (workspace_bytes < uz_bn_head_workspace_bytes(C, K) || (reinterpret_cast<uintptr_t>(workspace) & 15))
    return set_error(UZ_ERR_WORKSPACE, "bn_head_bwd: workspace too small (%zu < %zu)", workspace_bytes,
                     uz_bn_head_workspace_bytes(C, K));
  const BhArgs a = bh_args(z, z_ld, N, H, W, C, bn + 2 * C, bn + 3 * C, relu, w);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* ws = static_cast<float*>(workspace);
  __nv_bfloat16* dzp = static_cast<__nv_bfloat16*>(dz);
  switch (K) {
    case 1: return bh_bwd<1>(a, bn, dlogits, dzp, dz_ld, dgamma, dbeta, dbias, dw, db, accumulate & 1, ws, st);
    case 2: return bh_bwd<2>(a, bn, dlogits, dzp, dz_ld, dgamma, dbeta, dbias, dw, db, accumulate & 1, ws, st);
    case 3: return bh_bwd<3>(a, bn, dlogits, dzp, dz_ld, dgamma, dbeta, dbias, dw, db, accumulate & 1, ws, st);
    default: return bh_bwd<4>(a, bn, dlogits, dzp, dz_ld, dgamma, dbeta, dbias, dw, db, accumulate & 1, ws, st);
  }
}
