// Attention bottleneck pieces of the neighbouring models: transatt_unet's PAM / channel
// self-attention (transatt_unet.py:26-107) and unet_transformer's MHSA / MHCA (unet_transformer.py:115-229) are batched
// matrix products with a softmax in between.  The products run on the tcgen05 1x1 kernel in its batched-B mode
// (uz_bmm_tn in host_conv.cu); this file holds what surrounds them: the row softmax and its backward, a batched
// transpose (the TN kernel wants both operands K-contiguous), and a plain CUDA-core batched GEMM for shapes below one
// 128-row tile (bottlenecks of small test images) - never a library call.
#include <cstdint>

#include <cstdlib>

#include "pw.cuh"

namespace uz {
namespace {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float block_reduce(float v, bool is_max, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v = is_max ? warp_max(v) : warp_sum(v);
  __syncthreads();                 // red may still be read by the previous reduction
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = red[0];
  for (int i = 1; i < nw; ++i) r = is_max ? fmaxf(r, red[i]) : r + red[i];
  return r;
}

// y[r][c] = softmax_c(scale * x[r][c]); one block per row, fp32 arithmetic, bf16 storage
__global__ void softmax_rows_kernel(const __nv_bfloat16* __restrict__ x, long long x_ld, __nv_bfloat16* __restrict__ y,
                                    long long y_ld, int cols, float scale) {
  __shared__ float red[32];
  const __nv_bfloat16* xr = x + static_cast<long long>(blockIdx.x) * x_ld;
  __nv_bfloat16* yr = y + static_cast<long long>(blockIdx.x) * y_ld;
  float m = -INFINITY;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) m = fmaxf(m, scale * __bfloat162float(xr[c]));
  m = block_reduce(m, true, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) s += __expf(scale * __bfloat162float(xr[c]) - m);
  s = block_reduce(s, false, red);
  const float inv = 1.f / s;
  for (int c = threadIdx.x; c < cols; c += blockDim.x)
    yr[c] = __float2bfloat16_rn(__expf(scale * __bfloat162float(xr[c]) - m) * inv);
}

// dx[r][c] = scale * y[r][c] * (dy[r][c] - sum_j dy[r][j] * y[r][j])
__global__ void softmax_rows_bwd_kernel(const __nv_bfloat16* __restrict__ y, long long y_ld,
                                        const __nv_bfloat16* __restrict__ dy, long long dy_ld,
                                        __nv_bfloat16* __restrict__ dx, long long dx_ld, int cols, float scale) {
  __shared__ float red[32];
  const __nv_bfloat16* yr = y + static_cast<long long>(blockIdx.x) * y_ld;
  const __nv_bfloat16* gr = dy + static_cast<long long>(blockIdx.x) * dy_ld;
  __nv_bfloat16* dr = dx + static_cast<long long>(blockIdx.x) * dx_ld;
  float s = 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) s += __bfloat162float(yr[c]) * __bfloat162float(gr[c]);
  s = block_reduce(s, false, red);
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    const float yv = __bfloat162float(yr[c]);
    dr[c] = __float2bfloat16_rn(scale * yv * (__bfloat162float(gr[c]) - s));
  }
}

// Softmax over the ROW index of each column (nn.Softmax(dim=1) on a [b, rows, cols] score matrix, unet_transformer.py:
// 122,130): y[r][c] = exp(scale * x[r][c] - m_c) / s_c.  Block = 64 columns x 4 row groups, three sweeps over the rows
// (max, sum, write); consecutive threads touch consecutive columns (coalesced 128-byte rows).
constexpr int kColTile = 64, kColGroups = 4;
__global__ void __launch_bounds__(kColTile * kColGroups) softmax_cols_kernel(const __nv_bfloat16* __restrict__ x,
                                                                            __nv_bfloat16* __restrict__ y, int rows,
                                                                            int cols, long long img, float scale) {
  __shared__ float red[kColGroups][kColTile];
  const int cl = threadIdx.x % kColTile, grp = threadIdx.x / kColTile;
  const int c = blockIdx.x * kColTile + cl;
  const bool ok = c < cols;
  const __nv_bfloat16* xi = x + blockIdx.y * img + c;
  __nv_bfloat16* yi = y + blockIdx.y * img + c;
  float m = -INFINITY;
  if (ok)
    for (int r = grp; r < rows; r += kColGroups) m = fmaxf(m, scale * __bfloat162float(xi[static_cast<long long>(r) * cols]));
  red[grp][cl] = m;
  __syncthreads();
  m = fmaxf(fmaxf(red[0][cl], red[1][cl]), fmaxf(red[2][cl], red[3][cl]));
  __syncthreads();
  float s = 0.f;
  if (ok)
    for (int r = grp; r < rows; r += kColGroups) s += __expf(scale * __bfloat162float(xi[static_cast<long long>(r) * cols]) - m);
  red[grp][cl] = s;
  __syncthreads();
  s = (red[0][cl] + red[1][cl]) + (red[2][cl] + red[3][cl]);
  const float inv = 1.f / s;
  if (ok)
    for (int r = grp; r < rows; r += kColGroups)
      yi[static_cast<long long>(r) * cols] =
          __float2bfloat16_rn(__expf(scale * __bfloat162float(xi[static_cast<long long>(r) * cols]) - m) * inv);
}

// dx[r][c] = scale * y[r][c] * (dy[r][c] - sum_r' dy[r'][c] * y[r'][c])
__global__ void __launch_bounds__(kColTile * kColGroups) softmax_cols_bwd_kernel(const __nv_bfloat16* __restrict__ y,
                                                                                const __nv_bfloat16* __restrict__ dy,
                                                                                __nv_bfloat16* __restrict__ dx, int rows,
                                                                                int cols, long long img, float scale) {
  __shared__ float red[kColGroups][kColTile];
  const int cl = threadIdx.x % kColTile, grp = threadIdx.x / kColTile;
  const int c = blockIdx.x * kColTile + cl;
  const bool ok = c < cols;
  const long long base = blockIdx.y * img + c;
  float s = 0.f;
  if (ok)
    for (int r = grp; r < rows; r += kColGroups) {
      const long long o = base + static_cast<long long>(r) * cols;
      s += __bfloat162float(y[o]) * __bfloat162float(dy[o]);
    }
  red[grp][cl] = s;
  __syncthreads();
  s = (red[0][cl] + red[1][cl]) + (red[2][cl] + red[3][cl]);
  if (ok)
    for (int r = grp; r < rows; r += kColGroups) {
      const long long o = base + static_cast<long long>(r) * cols;
      dx[o] = __float2bfloat16_rn(scale * __bfloat162float(y[o]) * (__bfloat162float(dy[o]) - s));
    }
}

// The same two operators with 16-byte accesses (cols a multiple of 8, 16-byte aligned images): 8 threads cover the 64
// columns of a row, 32 row groups per block, four rows in flight per thread.  The scalar kernels above move 2 bytes per
// thread and access and reached 2.5 TB/s on the 4096 x 4096 score matrices of unet_transformer (1.7 ms per softmax).
// Forward: ONE sweep finds each column's maximum and sum (running maximum, the partial sum is rescaled only when the
// maximum moves - a handful of times per column), a second sweep writes: 2 reads + 1 write instead of 3 + 1.
constexpr int kColV8Groups = 32, kColV8Unroll = 4;
static bool env_is_one(const char* name) {      // "=1" selects the scalar kernels (A/B); unset or empty does not
  const char* v = getenv(name);
  return v != nullptr && v[0] == '1';
}
__global__ void __launch_bounds__(256) softmax_cols_v8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                              int rows, int cols, long long img, float scale) {
  __shared__ float red_m[kColV8Groups][kColTile], red_s[kColV8Groups][kColTile];
  const int vt = threadIdx.x & 7, grp = threadIdx.x >> 3;
  const int c = blockIdx.x * kColTile + vt * 8;
  const bool ok = c < cols;
  const __nv_bfloat16* xi = x + blockIdx.y * img + c;
  __nv_bfloat16* yi = y + blockIdx.y * img + c;
  float m[8], sum[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    m[j] = -INFINITY;
    sum[j] = 0.f;
  }
  constexpr int U = kColV8Unroll;
  if (ok) {
    for (int r = grp; r < rows; r += kColV8Groups * U) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (r + u * kColV8Groups < rows) v[u] = pw_ldg16(xi + static_cast<long long>(r + u * kColV8Groups) * cols);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (r + u * kColV8Groups >= rows) continue;
        float f[8];
        pw_unpack8(v[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float t = scale * f[j];
          if (t > m[j]) {
            sum[j] *= __expf(m[j] - t);    // exp(-inf) = 0 on the first row
            m[j] = t;
          }
          sum[j] += __expf(t - m[j]);
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    red_m[grp][vt * 8 + j] = m[j];
    red_s[grp][vt * 8 + j] = sum[j];
  }
  __syncthreads();
  float inv[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float mm = -INFINITY;
    for (int g = 0; g < kColV8Groups; ++g) mm = fmaxf(mm, red_m[g][vt * 8 + j]);
    float ss = 0.f;
    for (int g = 0; g < kColV8Groups; ++g) {
      const float mg = red_m[g][vt * 8 + j];
      if (mg > -INFINITY) ss += red_s[g][vt * 8 + j] * __expf(mg - mm);
    }
    m[j] = mm;
    inv[j] = 1.f / ss;
  }
  if (!ok) return;
  for (int r = grp; r < rows; r += kColV8Groups * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (r + u * kColV8Groups < rows) v[u] = pw_ldg16(xi + static_cast<long long>(r + u * kColV8Groups) * cols);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (r + u * kColV8Groups >= rows) continue;
      float f[8];
      pw_unpack8(v[u], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = __expf(scale * f[j] - m[j]) * inv[j];
      *reinterpret_cast<uint4*>(yi + static_cast<long long>(r + u * kColV8Groups) * cols) = pw_pack8(f);
    }
  }
}

__global__ void __launch_bounds__(256) softmax_cols_bwd_v8_kernel(const __nv_bfloat16* __restrict__ y,
                                                                  const __nv_bfloat16* __restrict__ dy,
                                                                  __nv_bfloat16* __restrict__ dx, int rows, int cols,
                                                                  long long img, float scale) {
  __shared__ float red[kColV8Groups][kColTile];
  const int vt = threadIdx.x & 7, grp = threadIdx.x >> 3;
  const int c = blockIdx.x * kColTile + vt * 8;
  const bool ok = c < cols;
  const long long base = blockIdx.y * img + c;
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  constexpr int U = kColV8Unroll;
  if (ok) {
    for (int r = grp; r < rows; r += kColV8Groups * U) {
      uint4 a[U], b[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (r + u * kColV8Groups < rows) {
          const long long o = base + static_cast<long long>(r + u * kColV8Groups) * cols;
          a[u] = pw_ldg16(y + o);
          b[u] = pw_ldg16(dy + o);
        }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (r + u * kColV8Groups >= rows) continue;
        float fy[8], fg[8];
        pw_unpack8(a[u], fy);
        pw_unpack8(b[u], fg);
#pragma unroll
        for (int j = 0; j < 8; ++j) s[j] = fmaf(fy[j], fg[j], s[j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[grp][vt * 8 + j] = s[j];
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float t = 0.f;
    for (int g = 0; g < kColV8Groups; ++g) t += red[g][vt * 8 + j];
    s[j] = t;
  }
  if (!ok) return;
  for (int r = grp; r < rows; r += kColV8Groups * U) {
    uint4 a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (r + u * kColV8Groups < rows) {
        const long long o = base + static_cast<long long>(r + u * kColV8Groups) * cols;
        a[u] = pw_ldg16(y + o);
        b[u] = pw_ldg16(dy + o);
      }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (r + u * kColV8Groups >= rows) continue;
      float fy[8], fg[8];
      pw_unpack8(a[u], fy);
      pw_unpack8(b[u], fg);
#pragma unroll
      for (int j = 0; j < 8; ++j) fy[j] = scale * fy[j] * (fg[j] - s[j]);
      *reinterpret_cast<uint4*>(dx + base + static_cast<long long>(r + u * kColV8Groups) * cols) = pw_pack8(fy);
    }
  }
}

// dst[n][c][r] = src[n][r][c]  (32 x 32 tiles through shared memory)
__global__ void transpose_bf16_kernel(const __nv_bfloat16* __restrict__ src, long long src_ld, long long src_img,
                                      __nv_bfloat16* __restrict__ dst, long long dst_ld, long long dst_img, int R, int C) {
  __shared__ __nv_bfloat16 tile[32][33];
  const int n = blockIdx.z, r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  const __nv_bfloat16* s = src + n * src_img;
  __nv_bfloat16* d = dst + n * dst_img;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    if (r < R && c < C) tile[i][threadIdx.x] = s[static_cast<long long>(r) * src_ld + c];
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (r < R && c < C) d[static_cast<long long>(c) * dst_ld + r] = tile[threadIdx.x][i];
  }
}

// Same transpose on 64 x 64 tiles with 16-byte global accesses (rows, columns, pitches multiples of 8 elements, 16-byte
// aligned bases): the backward of the attention products transposes 4096 x 4096 score matrices (33.5 MB per image) -
// the 32 x 32 kernel above moves 2 bytes per thread and access and reached 1.3 TB/s there.
__global__ void __launch_bounds__(256) transpose64_bf16_kernel(const __nv_bfloat16* __restrict__ src, long long src_ld,
                                                               long long src_img, __nv_bfloat16* __restrict__ dst,
                                                               long long dst_ld, long long dst_img, int R, int C) {
  __shared__ uint32_t tile[64 * 33];   // [row][33 words]: 64 bf16 per row + one word of padding
  const int n = blockIdx.z, r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
  const __nv_bfloat16* s = src + n * src_img;
  __nv_bfloat16* d = dst + n * dst_img;
  const int t = threadIdx.x, v = t & 7;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int r = (t >> 3) + 32 * i;
    uint4 q = make_uint4(0u, 0u, 0u, 0u);
    if (r0 + r < R && c0 + v * 8 < C) q = __ldg(reinterpret_cast<const uint4*>(s + static_cast<long long>(r0 + r) * src_ld + c0 + v * 8));
    uint32_t* w = tile + r * 33 + v * 4;
    w[0] = q.x; w[1] = q.y; w[2] = q.z; w[3] = q.w;
  }
  __syncthreads();
  const uint16_t* th = reinterpret_cast<const uint16_t*>(tile);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int c = (t >> 3) + 32 * i;            // output row = source column
    if (c0 + c >= C || r0 + v * 8 >= R) continue;
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t lo = th[(v * 8 + 2 * k) * 66 + c], hi = th[(v * 8 + 2 * k + 1) * 66 + c];
      o[k] = lo | (hi << 16);
    }
    *reinterpret_cast<uint4*>(d + static_cast<long long>(c0 + c) * dst_ld + r0 + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// C[n][m][j] = sum_k A[n][m][k] * B[n][j][k]   (small shapes only: 32 x 32 output tile per block, K in steps of 32)
__global__ void bmm_tn_small_kernel(const __nv_bfloat16* __restrict__ A, long long lda, long long a_img,
                                    const __nv_bfloat16* __restrict__ B, long long ldb, long long b_img,
                                    __nv_bfloat16* __restrict__ Cm, long long ldc, long long c_img, int M, int K, int J) {
  __shared__ float as[32][33], bs[32][33];
  const int n = blockIdx.z, m0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
  const __nv_bfloat16* a = A + n * a_img;
  const __nv_bfloat16* b = B + n * b_img;
  const int tx = threadIdx.x, ty = threadIdx.y;   // 32 x 8
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k0 = 0; k0 < K; k0 += 32) {
    for (int i = ty; i < 32; i += 8) {
      const int k = k0 + tx;
      as[i][tx] = (m0 + i < M && k < K) ? __bfloat162float(a[static_cast<long long>(m0 + i) * lda + k]) : 0.f;
      bs[i][tx] = (j0 + i < J && k < K) ? __bfloat162float(b[static_cast<long long>(j0 + i) * ldb + k]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int mi = ty + 8 * q;
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < 32; ++k) s = fmaf(as[mi][k], bs[tx][k], s);
      acc[q] += s;
    }
    __syncthreads();
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int m = m0 + ty + 8 * q, j = j0 + tx;
    if (m < M && j < J) Cm[n * c_img + static_cast<long long>(m) * ldc + j] = __float2bfloat16_rn(acc[q]);
  }
}

}  // namespace

cudaError_t launch_bmm_tn_small(const void* A, long long lda, long long a_img, const void* B, long long ldb,
                                long long b_img, void* C, long long ldc, long long c_img, int N, int M, int K, int J,
                                cudaStream_t st) {
  dim3 grid((J + 31) / 32, (M + 31) / 32, N), block(32, 8);
  bmm_tn_small_kernel<<<grid, block, 0, st>>>(static_cast<const __nv_bfloat16*>(A), lda, a_img,
                                              static_cast<const __nv_bfloat16*>(B), ldb, b_img,
                                              static_cast<__nv_bfloat16*>(C), ldc, c_img, M, K, J);
  return cudaGetLastError();
}
}  // namespace uz

using namespace uz;

extern "C" int uz_softmax_rows(const void* x, long long x_ld, void* y, long long y_ld, long long rows, int cols,
                               float scale, void* stream) {
  PW_REQUIRE_DEVICE();
  if (!x || !y || rows < 1 || cols < 1 || rows > 0x7fffffffLL || x_ld < cols || y_ld < cols)
    return set_error(UZ_ERR_INVALID, "softmax_rows: bad extents rows=%lld cols=%d", rows, cols);
  const int threads = cols >= 1024 ? 256 : cols >= 256 ? 128 : 64;
  softmax_rows_kernel<<<static_cast<int>(rows), threads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), x_ld, static_cast<__nv_bfloat16*>(y), y_ld, cols, scale);
  PW_CHECK_LAUNCH("softmax_rows launch");
  return UZ_OK;
}

extern "C" int uz_softmax_rows_bwd(const void* y, long long y_ld, const void* dy, long long dy_ld, void* dx,
                                   long long dx_ld, long long rows, int cols, float scale, void* stream) {
  PW_REQUIRE_DEVICE();
  if (!y || !dy || !dx || rows < 1 || cols < 1 || rows > 0x7fffffffLL || y_ld < cols || dy_ld < cols || dx_ld < cols)
    return set_error(UZ_ERR_INVALID, "softmax_rows_bwd: bad extents rows=%lld cols=%d", rows, cols);
  const int threads = cols >= 1024 ? 256 : cols >= 256 ? 128 : 64;
  softmax_rows_bwd_kernel<<<static_cast<int>(rows), threads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(y), y_ld, static_cast<const __nv_bfloat16*>(dy), dy_ld,
      static_cast<__nv_bfloat16*>(dx), dx_ld, cols, scale);
  PW_CHECK_LAUNCH("softmax_rows_bwd launch");
  return UZ_OK;
}

extern "C" int uz_transpose_bf16(const void* src, long long src_ld, long long src_img, void* dst, long long dst_ld,
                                 long long dst_img, int N, int R, int C, void* stream) {
  PW_REQUIRE_DEVICE();
  if (!src || !dst || N < 1 || R < 1 || C < 1 || src_ld < C || dst_ld < R || N > 65535)
    return set_error(UZ_ERR_INVALID, "transpose_bf16: bad extents N=%d R=%d C=%d", N, R, C);
  const bool vec = ((R | C) & 7) == 0 && ((src_ld | dst_ld | src_img | dst_img) & 7) == 0 &&
                   ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15u) == 0;
  if (vec) {
    dim3 grid64((C + 63) / 64, (R + 63) / 64, N);
    transpose64_bf16_kernel<<<grid64, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(src), src_ld, src_img, static_cast<__nv_bfloat16*>(dst), dst_ld, dst_img, R, C);
    PW_CHECK_LAUNCH("transpose_bf16 (64 x 64) launch");
    return UZ_OK;
  }
  dim3 grid((C + 31) / 32, (R + 31) / 32, N), block(32, 8);
  transpose_bf16_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(src), src_ld, src_img, static_cast<__nv_bfloat16*>(dst), dst_ld, dst_img, R, C);
  PW_CHECK_LAUNCH("transpose_bf16 launch");
  return UZ_OK;
}

extern "C" int uz_softmax_cols(const void* x, void* y, int N, int rows, int cols, float scale, void* stream) {
  PW_REQUIRE_DEVICE();
  if (!x || !y || N < 1 || N > 65535 || rows < 1 || cols < 1) return set_error(UZ_ERR_INVALID, "softmax_cols: bad extents");
  dim3 grid((cols + kColTile - 1) / kColTile, N);
  const bool v8 = (cols % 8) == 0 && (reinterpret_cast<uintptr_t>(x) & 15u) == 0 && (reinterpret_cast<uintptr_t>(y) & 15u) == 0 &&
                  !env_is_one("UZ_SOFTMAX_COLS_SCALAR");
  if (v8) {
    softmax_cols_v8_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), rows, cols, static_cast<long long>(rows) * cols, scale);
    PW_CHECK_LAUNCH("softmax_cols launch");
    return UZ_OK;
  }
  softmax_cols_kernel<<<grid, kColTile * kColGroups, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), rows, cols, static_cast<long long>(rows) * cols, scale);
  PW_CHECK_LAUNCH("softmax_cols launch");
  return UZ_OK;
}

extern "C" int uz_softmax_cols_bwd(const void* y, const void* dy, void* dx, int N, int rows, int cols, float scale,
                                   void* stream) {
  PW_REQUIRE_DEVICE();
  if (!y || !dy || !dx || N < 1 || N > 65535 || rows < 1 || cols < 1)
    return set_error(UZ_ERR_INVALID, "softmax_cols_bwd: bad extents");
  dim3 grid((cols + kColTile - 1) / kColTile, N);
  const bool v8 = (cols % 8) == 0 && ((reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx)) & 15u) == 0 &&
                  !env_is_one("UZ_SOFTMAX_COLS_SCALAR");
  if (v8) {
    softmax_cols_bwd_v8_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), rows, cols,
        static_cast<long long>(rows) * cols, scale);
    PW_CHECK_LAUNCH("softmax_cols_bwd launch");
    return UZ_OK;
  }
  softmax_cols_bwd_kernel<<<grid, kColTile * kColGroups, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), rows, cols,
      static_cast<long long>(rows) * cols, scale);
  PW_CHECK_LAUNCH("softmax_cols_bwd launch");
  return UZ_OK;
}
