// 1x1 classifier head (OutConv / conv_1x1 / final / output_layer of the reference models): a per-pixel dot product
// over C channels for a handful of classes - pure HBM bandwidth.  CV = C/8 lanes cooperate on one pixel (each lane
// owns 8 channels = one 16-byte load), partial dots are combined with warp shuffles; logits are fp32 NCHW.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/unet_zoo_b200.h"
#include "host_util.h"

namespace uz {

constexpr int kHeadKT = 4;  // classes handled per pass

__device__ __forceinline__ void h_unpack8(const uint4& v, float (&f)[8]) {
  f[0] = __uint_as_float(v.x << 16);
  f[1] = __uint_as_float(v.x & 0xFFFF0000u);
  f[2] = __uint_as_float(v.y << 16);
  f[3] = __uint_as_float(v.y & 0xFFFF0000u);
  f[4] = __uint_as_float(v.z << 16);
  f[5] = __uint_as_float(v.z & 0xFFFF0000u);
  f[6] = __uint_as_float(v.w << 16);
  f[7] = __uint_as_float(v.w & 0xFFFF0000u);
}
__device__ __forceinline__ uint32_t h_pack2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

template <int CV, int KT>
__global__ void head_fwd_kernel(const __nv_bfloat16* __restrict__ y, int ld, long long npix, long long plane, int C,
                                int K, int k0, int kn, const float* __restrict__ w, const float* __restrict__ b,
                                float* __restrict__ logits) {
  const int g = threadIdx.x % CV;
  const int slot = threadIdx.x / CV;
  const int spb = blockDim.x / CV;
  float wr[KT][8];
#pragma unroll
  for (int k = 0; k < KT; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) wr[k][j] = k < kn ? __ldg(w + static_cast<size_t>(k0 + k) * C + g * 8 + j) : 0.f;
  const long long iters = (npix + static_cast<long long>(gridDim.x) * spb - 1) / (static_cast<long long>(gridDim.x) * spb);
  constexpr int U = 4;   // pixels in flight per thread: one 16-byte load per trip left the kernel latency bound (0.39 ms
                         // for 64 channels x 32 x 512^2 where its bytes are worth 0.19 ms)
  for (long long i = 0; i < iters; i += U) {  // uniform trip count: the shuffles below need the whole warp
    uint4 raw[U];
    long long pp[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      pp[u] = ((i + u) * gridDim.x + blockIdx.x) * spb + slot;
      ok[u] = i + u < iters && pp[u] < npix;
      raw[u] = ok[u] ? __ldg(reinterpret_cast<const uint4*>(y + pp[u] * ld + g * 8)) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float f[8];
      h_unpack8(raw[u], f);
      float acc[KT];
#pragma unroll
      for (int k = 0; k < KT; ++k) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s = fmaf(wr[k][j], f[j], s);
#pragma unroll
        for (int o = CV / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        acc[k] = s;
      }
      if (ok[u] && g == 0) {
        const long long p = pp[u];
        const long long n = p / plane, hw = p - n * plane;
#pragma unroll
        for (int k = 0; k < KT; ++k)
          if (k < kn) logits[(n * K + k0 + k) * plane + hw] = acc[k] + (b ? __ldg(b + k0 + k) : 0.f);
      }
    }
  }
}

// dy[p, c] (+)= sum_k dl[p,k] w[k,c] ;  partial[block][kn][C] = sum_p dl[p,k] y[p,c] ; bpartial[block][kn] = sum_p dl[p,k]
template <int CV, int KT>
__global__ void __launch_bounds__(256, KT == 1 ? 4 : 2) head_bwd_kernel(const __nv_bfloat16* __restrict__ y, int ld, const float* __restrict__ dlogits,
                                long long npix, long long plane, int C, int K, int k0, int kn,
                                const float* __restrict__ w, __nv_bfloat16* __restrict__ dy, int dy_ld, int dy_accumulate,
                                float* __restrict__ partial, float* __restrict__ bpartial) {
  extern __shared__ float red[];  // [blockDim][KT*8 + KT]
  const int g = threadIdx.x % CV;
  const int slot = threadIdx.x / CV;
  const int spb = blockDim.x / CV;
  float wr[KT][8], aw[KT][8], ab[KT];
#pragma unroll
  for (int k = 0; k < KT; ++k) {
    ab[k] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      wr[k][j] = k < kn ? __ldg(w + static_cast<size_t>(k0 + k) * C + g * 8 + j) : 0.f;
      aw[k][j] = 0.f;
    }
  }
  for (long long p = static_cast<long long>(blockIdx.x) * spb + slot; p < npix; p += static_cast<long long>(gridDim.x) * spb) {
    float f[8];
    h_unpack8(__ldg(reinterpret_cast<const uint4*>(y + p * ld + g * 8)), f);
    const long long n = p / plane, hw = p - n * plane;
    float d[8];
    if (dy_accumulate) {
      h_unpack8(*reinterpret_cast<const uint4*>(dy + p * dy_ld + g * 8), d);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = 0.f;
    }
#pragma unroll
    for (int k = 0; k < KT; ++k) {
      if (k < kn) {
        const float dl = __ldg(dlogits + (n * K + k0 + k) * plane + hw);
        ab[k] += dl;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          aw[k][j] = fmaf(dl, f[j], aw[k][j]);
          d[j] = fmaf(dl, wr[k][j], d[j]);
        }
      }
    }
    *reinterpret_cast<uint4*>(dy + p * dy_ld + g * 8) =
        make_uint4(h_pack2(d[0], d[1]), h_pack2(d[2], d[3]), h_pack2(d[4], d[5]), h_pack2(d[6], d[7]));
  }
  constexpr int kPer = KT * 8 + KT;
  float* mine = red + static_cast<size_t>(threadIdx.x) * kPer;
#pragma unroll
  for (int k = 0; k < KT; ++k) {
#pragma unroll
    for (int j = 0; j < 8; ++j) mine[k * 8 + j] = aw[k][j];
    mine[KT * 8 + k] = ab[k];
  }
  __syncthreads();
  if (slot == 0) {
    for (int r = 1; r < spb; ++r) {
      const float* o = red + (static_cast<size_t>(r) * CV + g) * kPer;
#pragma unroll
      for (int k = 0; k < KT; ++k) {
#pragma unroll
        for (int j = 0; j < 8; ++j) aw[k][j] += o[k * 8 + j];
        ab[k] += o[KT * 8 + k];
      }
    }
#pragma unroll
    for (int k = 0; k < KT; ++k)
      if (k < kn) {
#pragma unroll
        for (int j = 0; j < 8; ++j) partial[(static_cast<size_t>(blockIdx.x) * kn + k) * C + g * 8 + j] = aw[k][j];
        if (g == 0) bpartial[static_cast<size_t>(blockIdx.x) * kn + k] = ab[k];
      }
  }
}

__global__ void head_reduce_kernel(const float* __restrict__ partial, int rows, int cols, float* __restrict__ out,
                                   int accumulate) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= cols) return;
  double s = 0.0;
  for (int r = lane; r < rows; r += 32) s += static_cast<double>(partial[static_cast<size_t>(r) * cols + c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane != 0) return;
  out[c] = (accumulate ? out[c] : 0.f) + static_cast<float>(s);
}

template <int CV>
static void launch_head_fwd(int grid, int block, cudaStream_t st, const __nv_bfloat16* y, int ld, long long npix,
                            long long plane, int C, int K, int k0, int kn, const float* w, const float* b, float* lg) {
  if (kn == 1)
    head_fwd_kernel<CV, 1><<<grid, block, 0, st>>>(y, ld, npix, plane, C, K, k0, kn, w, b, lg);
  else
    head_fwd_kernel<CV, kHeadKT><<<grid, block, 0, st>>>(y, ld, npix, plane, C, K, k0, kn, w, b, lg);
}
template <int CV>
static void launch_head_bwd(int grid, int block, cudaStream_t st, const __nv_bfloat16* y, int ld, const float* dl,
                            long long npix, long long plane, int C, int K, int k0, int kn, const float* w,
                            __nv_bfloat16* dy, int dy_ld, int acc, float* partial, float* bpartial) {
  if (kn == 1) {
    const size_t smem = static_cast<size_t>(block) * 9 * sizeof(float);
    head_bwd_kernel<CV, 1><<<grid, block, smem, st>>>(y, ld, dl, npix, plane, C, K, k0, kn, w, dy, dy_ld, acc, partial,
                                                      bpartial);
  } else {
    const size_t smem = static_cast<size_t>(block) * (kHeadKT * 8 + kHeadKT) * sizeof(float);
    head_bwd_kernel<CV, kHeadKT><<<grid, block, smem, st>>>(y, ld, dl, npix, plane, C, K, k0, kn, w, dy, dy_ld, acc,
                                                            partial, bpartial);
  }
}

}  // namespace uz

using namespace uz;

static int head_check(const char* who, const void* y, int C, int ld) {
  const int cv = C / 8;
  if (y == nullptr || (reinterpret_cast<uintptr_t>(y) & 15u) || (C & 7) || (ld & 7) || ld < C || cv < 1 || cv > 32 ||
      (cv & (cv - 1)))
    return set_error(UZ_ERR_UNSUPPORTED, "%s: needs C in {8,16,32,64,128,256} (got %d), ld multiple of 8, aligned pointer",
                     who, C);
  return UZ_OK;
}

static int head_rows() {
  const int sms = uz_device_sm_count();
  return sms < 0 ? sms : sms * 4;
}

extern "C" int uz_head_fwd(const void* y, int ld, int N, int H, int W, int C, int K, const float* w, const float* b,
                           float* logits, void* stream) {
  const int rows = head_rows();
  if (rows < 0) return rows;
  int rc = head_check("head_fwd", y, C, ld);
  if (rc) return rc;
  if (K < 1) return set_error(UZ_ERR_INVALID, "head_fwd: K=%d", K);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long plane = static_cast<long long>(H) * W, npix = plane * N;
  const int cv = C / 8, block = 256, spb = block / cv;
  long long need = (npix + spb - 1) / spb;
  const int grid = static_cast<int>(need < rows * 2 ? need : rows * 2);
  const __nv_bfloat16* yp = static_cast<const __nv_bfloat16*>(y);
  for (int k0 = 0; k0 < K; k0 += kHeadKT) {
    const int kn = K - k0 < kHeadKT ? K - k0 : kHeadKT;
    switch (cv) {
      case 1: launch_head_fwd<1>(grid, block, st, yp, ld, npix, plane, C, K, k0, kn, w, b, logits); break;
      case 2: launch_head_fwd<2>(grid, block, st, yp, ld, npix, plane, C, K, k0, kn, w, b, logits); break;
      case 4: launch_head_fwd<4>(grid, block, st, yp, ld, npix, plane, C, K, k0, kn, w, b, logits); break;
      case 8: launch_head_fwd<8>(grid, block, st, yp, ld, npix, plane, C, K, k0, kn, w, b, logits); break;
      case 16: launch_head_fwd<16>(grid, block, st, yp, ld, npix, plane, C, K, k0, kn, w, b, logits); break;
      default: launch_head_fwd<32>(grid, block, st, yp, ld, npix, plane, C, K, k0, kn, w, b, logits); break;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_error(e, "head_fwd launch");
  }
  return UZ_OK;
}

extern "C" size_t uz_head_bwd_workspace_bytes(int C) {
  const int rows = head_rows();
  return rows < 0 ? 0 : sizeof(float) * static_cast<size_t>(rows) * kHeadKT * (C + 1);
}

extern "C" int uz_head_bwd(const void* y, int ld, const float* dlogits, int N, int H, int W, int C, int K,
                           const float* w, void* dy, int dy_ld, float* dw, float* db, int accumulate, void* workspace,
                           size_t workspace_bytes, void* stream) {
  const int rows = head_rows();
  if (rows < 0) return rows;
  int rc = head_check("head_bwd(y)", y, C, ld);
  if (rc) return rc;
  if ((rc = head_check("head_bwd(dy)", dy, C, dy_ld))) return rc;
  if (workspace_bytes < uz_head_bwd_workspace_bytes(C)) return set_error(UZ_ERR_WORKSPACE, "head_bwd: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long plane = static_cast<long long>(H) * W, npix = plane * N;
  const int cv = C / 8, block = 256, spb = block / cv;
  long long need = (npix + spb - 1) / spb;
  const int grid = static_cast<int>(need < rows ? need : rows);
  float* partial = static_cast<float*>(workspace);
  float* bpartial = partial + static_cast<size_t>(rows) * kHeadKT * C;
  const __nv_bfloat16* yp = static_cast<const __nv_bfloat16*>(y);
  __nv_bfloat16* dyp = static_cast<__nv_bfloat16*>(dy);
  for (int k0 = 0; k0 < K; k0 += kHeadKT) {
    const int kn = K - k0 < kHeadKT ? K - k0 : kHeadKT;
    const int acc = k0 > 0 ? 1 : 0;
    switch (cv) {
      case 1: launch_head_bwd<1>(grid, block, st, yp, ld, dlogits, npix, plane, C, K, k0, kn, w, dyp, dy_ld, acc, partial, bpartial); break;
      case 2: launch_head_bwd<2>(grid, block, st, yp, ld, dlogits, npix, plane, C, K, k0, kn, w, dyp, dy_ld, acc, partial, bpartial); break;
      case 4: launch_head_bwd<4>(grid, block, st, yp, ld, dlogits, npix, plane, C, K, k0, kn, w, dyp, dy_ld, acc, partial, bpartial); break;
      case 8: launch_head_bwd<8>(grid, block, st, yp, ld, dlogits, npix, plane, C, K, k0, kn, w, dyp, dy_ld, acc, partial, bpartial); break;
      case 16: launch_head_bwd<16>(grid, block, st, yp, ld, dlogits, npix, plane, C, K, k0, kn, w, dyp, dy_ld, acc, partial, bpartial); break;
      default: launch_head_bwd<32>(grid, block, st, yp, ld, dlogits, npix, plane, C, K, k0, kn, w, dyp, dy_ld, acc, partial, bpartial); break;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_error(e, "head_bwd launch");
    head_reduce_kernel<<<(kn * C + 3) / 4, 128, 0, st>>>(partial, grid, kn * C, dw + static_cast<size_t>(k0) * C, accumulate);
    head_reduce_kernel<<<(kn + 3) / 4, 128, 0, st>>>(bpartial, grid, kn, db + k0, accumulate);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_error(e, "head_bwd reduce launch");
  }
  return UZ_OK;
}
