// Tail of the training step: everything the reference's loop body does around
// forward / backward, as device-side bandwidth kernels with no host synchronisation.
//
//  uz_bce_dice   : nn.BCEWithLogitsLoss() (mean; scripts/train.py:135, training_loop.py:78) + its gradient
//                  dL/dlogits = (sigmoid(z) - y) / n + the Dice statistics of metrics.py:7-24 (sigmoid -> > 0.5 ->
//                  intersection / prediction / target sums, "union == 0 -> 1.0" resolved on the device) in ONE pass
//                  over logits and mask.  The reference runs ~12 ATen kernels and two host syncs (`if union == 0`,
//                  `.item()`) for the same result.
//  uz_sumsq      : partial sums of squares of the flat fp32 gradient buffer (clip_grad_norm_, training_loop.py:120)
//  uz_adamw_flat : clip coefficient + bias corrections from the partials (one thread), then one AdamW pass over the
//                  flat parameter / gradient / moment buffers (optim.AdamW(lr, weight_decay), scripts/train.py:156;
//                  training_loop.py:121).  The step counter lives in device memory, so the call is CUDA-graph
//                  capturable; hyper-parameters are read from a device array for the same reason.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/unet_zoo_b200.h"
#include "host_util.h"

namespace uz {
namespace {

constexpr int kTailThreads = 256;

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// block-wide sum of K doubles per thread; result valid in thread 0
template <int K>
__device__ __forceinline__ void block_sum(double (&v)[K], double* smem /* [K][8] */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    v[k] = warp_sum_d(v[k]);
    if (lane == 0) smem[k * 8 + warp] = v[k];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      double s = 0.0;
      for (int w = 0; w < kTailThreads / 32; ++w) s += smem[k * 8 + w];
      v[k] = s;
    }
  }
}

__device__ __forceinline__ void bce_elem(float z, float y, float gscale, float& loss, float& inter, float& pred,
                                         float& targ, float& dl) {
  // softplus form of -[y log s + (1-y) log(1-s)]:  max(z,0) - z*y + log1p(exp(-|z|))
  loss += fmaxf(z, 0.f) - z * y + log1pf(expf(-fabsf(z)));
  const float s = 1.f / (1.f + expf(-z));          // torch.sigmoid's formulation: the mask below is bit-exact with it
  const float m = s > 0.5f ? 1.f : 0.f;
  inter += m * y;
  pred += m;
  targ += y;
  dl = (s - y) * gscale;
}

__global__ void __launch_bounds__(kTailThreads) bce_dice_kernel(const float* __restrict__ logits,
                                                                const float* __restrict__ target, long long n,
                                                                float gscale, float* __restrict__ dlogits,
                                                                double* __restrict__ partial /* [grid][4] */) {
  __shared__ double smem[4 * 8];
  float loss = 0.f, inter = 0.f, pred = 0.f, targ = 0.f;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  const long long n4 = n >> 2;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  int since = 0;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    const float4 z = __ldg(reinterpret_cast<const float4*>(logits) + i);
    const float4 y = __ldg(reinterpret_cast<const float4*>(target) + i);
    float4 d;
    bce_elem(z.x, y.x, gscale, loss, inter, pred, targ, d.x);
    bce_elem(z.y, y.y, gscale, loss, inter, pred, targ, d.y);
    bce_elem(z.z, y.z, gscale, loss, inter, pred, targ, d.z);
    bce_elem(z.w, y.w, gscale, loss, inter, pred, targ, d.w);
    if (dlogits) reinterpret_cast<float4*>(dlogits)[i] = d;
    if (++since == 64) {   // fold the fp32 running sums into fp64 every 256 elements
      acc[0] += loss; acc[1] += inter; acc[2] += pred; acc[3] += targ;
      loss = inter = pred = targ = 0.f;
      since = 0;
    }
  }
  // tail (n not a multiple of 4)
  for (long long i = (n4 << 2) + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride) {
    float d;
    bce_elem(__ldg(logits + i), __ldg(target + i), gscale, loss, inter, pred, targ, d);
    if (dlogits) dlogits[i] = d;
  }
  acc[0] += loss; acc[1] += inter; acc[2] += pred; acc[3] += targ;
  block_sum<4>(acc, smem);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 4; ++k) partial[blockIdx.x * 4 + k] = acc[k];
  }
}

// result[0] = mean loss, [1] = dice, [2] = intersection, [3] = prediction sum, [4] = target sum
__global__ void bce_dice_finalize_kernel(const double* __restrict__ partial, int rows, double n, float eps,
                                         float* __restrict__ result) {
  __shared__ double smem[4 * 8];
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (int r = threadIdx.x; r < rows; r += blockDim.x)
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[k] += partial[r * 4 + k];
  block_sum<4>(acc, smem);
  if (threadIdx.x == 0) {
    result[0] = static_cast<float>(acc[0] / n);
    // metrics.py:15-22 in fp32 like the reference: union == 0 -> 1.0, else (2I + eps) / (U + eps)
    const float inter = static_cast<float>(acc[1]);
    const float uni = static_cast<float>(acc[2]) + static_cast<float>(acc[3]);
    result[1] = uni == 0.f ? 1.f : (2.f * inter + eps) / (uni + eps);
    result[2] = inter;
    result[3] = static_cast<float>(acc[2]);
    result[4] = static_cast<float>(acc[3]);
  }
}

__global__ void __launch_bounds__(kTailThreads) sumsq_kernel(const float* __restrict__ g, long long n,
                                                             double* __restrict__ partial) {
  __shared__ double smem[8];
  double acc[1] = {0.0};
  float s = 0.f;
  int since = 0;
  const long long n4 = n >> 2;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(g) + i);
    s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    if (++since == 64) {
      acc[0] += s;
      s = 0.f;
      since = 0;
    }
  }
  for (long long i = (n4 << 2) + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride) {
    const float v = __ldg(g + i);
    s += v * v;
  }
  acc[0] += s;
  block_sum<1>(acc, smem);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc[0];
}

// hyper = {lr, beta1, beta2, eps, weight_decay, max_norm (<= 0: no clipping)}
// state = {step, total_norm, clip_coef, bias_correction1, sqrt(bias_correction2)}   (state[0] persists across steps)
__global__ void adamw_prepare_kernel(const double* __restrict__ partial, int rows, const float* __restrict__ hyper,
                                     float* __restrict__ state) {
  __shared__ double smem[8];
  double acc[1] = {0.0};
  for (int r = threadIdx.x; r < rows; r += blockDim.x) acc[0] += partial[r];
  block_sum<1>(acc, smem);
  if (threadIdx.x == 0) {
    const float step = state[0] + 1.f;
    const float norm = static_cast<float>(sqrt(acc[0]));
    const float max_norm = hyper[5];
    float coef = 1.f;
    if (max_norm > 0.f) coef = fminf(max_norm / (norm + 1e-6f), 1.f);   // torch.nn.utils.clip_grad_norm_
    state[0] = step;
    state[1] = norm;
    state[2] = coef;
    state[3] = static_cast<float>(1.0 - pow(static_cast<double>(hyper[1]), static_cast<double>(step)));
    state[4] = static_cast<float>(sqrt(1.0 - pow(static_cast<double>(hyper[2]), static_cast<double>(step))));
  }
}

__device__ __forceinline__ void adamw_elem(float& p, float& g, float& m, float& v, float lr, float b1, float b2,
                                           float eps, float wd, float coef, float bc1, float bc2s) {
  g *= coef;                         // the clipped gradient is what the caller sees afterwards, as with clip_grad_norm_
  p *= 1.f - lr * wd;                // decoupled weight decay
  m = b1 * m + (1.f - b1) * g;
  v = b2 * v + (1.f - b2) * g * g;
  const float denom = sqrtf(v) / bc2s + eps;
  p -= (lr / bc1) * (m / denom);
}

__global__ void __launch_bounds__(kTailThreads) adamw_flat_kernel(float* __restrict__ p, float* __restrict__ g,
                                                                  float* __restrict__ m, float* __restrict__ v,
                                                                  long long n, const float* __restrict__ hyper,
                                                                  const float* __restrict__ state) {
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4];
  const float coef = state[2], bc1 = state[3], bc2s = state[4];
  const long long n4 = n >> 2;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i], gg = reinterpret_cast<float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    adamw_elem(pp.x, gg.x, mm.x, vv.x, lr, b1, b2, eps, wd, coef, bc1, bc2s);
    adamw_elem(pp.y, gg.y, mm.y, vv.y, lr, b1, b2, eps, wd, coef, bc1, bc2s);
    adamw_elem(pp.z, gg.z, mm.z, vv.z, lr, b1, b2, eps, wd, coef, bc1, bc2s);
    adamw_elem(pp.w, gg.w, mm.w, vv.w, lr, b1, b2, eps, wd, coef, bc1, bc2s);
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(g)[i] = gg;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  for (long long i = (n4 << 2) + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride)
    adamw_elem(p[i], g[i], m[i], v[i], lr, b1, b2, eps, wd, coef, bc1, bc2s);
}

int tail_rows() {
  const int sms = sm_count();
  return sms > 0 ? sms * 4 : -1;
}

#define TAIL_REQUIRE_DEVICE()              \
  do {                                     \
    const int sms_ = uz_device_sm_count(); \
    if (sms_ < 0) return sms_;             \
  } while (0)
#define TAIL_CHECK_LAUNCH(what)                             \
  do {                                                      \
    cudaError_t e_ = cudaGetLastError();                    \
    if (e_ != cudaSuccess) return uz::cuda_error(e_, what); \
  } while (0)

}  // namespace
}  // namespace uz

using namespace uz;

extern "C" size_t uz_tail_workspace_bytes(void) {
  const int rows = tail_rows();
  return rows < 0 ? 0 : sizeof(double) * 4 * static_cast<size_t>(rows);
}

extern "C" int uz_bce_dice(const float* logits, const float* target, long long n, float grad_scale, float dice_eps,
                           float* dlogits, float* result, void* workspace, size_t workspace_bytes, void* stream) {
  TAIL_REQUIRE_DEVICE();
  if (!logits || !target || !result || n < 1) return set_error(UZ_ERR_INVALID, "bce_dice: bad arguments");
  if ((reinterpret_cast<uintptr_t>(logits) | reinterpret_cast<uintptr_t>(target) |
       reinterpret_cast<uintptr_t>(dlogits)) & 15)
    return set_error(UZ_ERR_INVALID, "bce_dice: logits / target / dlogits must be 16-byte aligned");
  const int rows = tail_rows();
  if (!workspace || workspace_bytes < uz_tail_workspace_bytes())
    return set_error(UZ_ERR_WORKSPACE, "bce_dice: workspace too small (%zu < %zu)", workspace_bytes,
                     uz_tail_workspace_bytes());
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  double* partial = static_cast<double*>(workspace);
  bce_dice_kernel<<<rows, kTailThreads, 0, st>>>(logits, target, n, grad_scale, dlogits, partial);
  TAIL_CHECK_LAUNCH("bce_dice launch");
  bce_dice_finalize_kernel<<<1, kTailThreads, 0, st>>>(partial, rows, static_cast<double>(n), dice_eps, result);
  TAIL_CHECK_LAUNCH("bce_dice finalize launch");
  return UZ_OK;
}

extern "C" int uz_adamw_flat(float* params, float* grads, float* exp_avg, float* exp_avg_sq, long long n,
                             const float* hyper, float* state, void* workspace, size_t workspace_bytes, void* stream) {
  TAIL_REQUIRE_DEVICE();
  if (!params || !grads || !exp_avg || !exp_avg_sq || !hyper || !state || n < 1)
    return set_error(UZ_ERR_INVALID, "adamw_flat: bad arguments");
  if ((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) | reinterpret_cast<uintptr_t>(exp_avg) |
       reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15)
    return set_error(UZ_ERR_INVALID, "adamw_flat: buffers must be 16-byte aligned");
  const int rows = tail_rows();
  if (!workspace || workspace_bytes < uz_tail_workspace_bytes())
    return set_error(UZ_ERR_WORKSPACE, "adamw_flat: workspace too small (%zu < %zu)", workspace_bytes,
                     uz_tail_workspace_bytes());
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  double* partial = static_cast<double*>(workspace);
  sumsq_kernel<<<rows, kTailThreads, 0, st>>>(grads, n, partial);
  TAIL_CHECK_LAUNCH("sumsq launch");
  adamw_prepare_kernel<<<1, kTailThreads, 0, st>>>(partial, rows, hyper, state);
  TAIL_CHECK_LAUNCH("adamw_prepare launch");
  adamw_flat_kernel<<<rows * 2, kTailThreads, 0, st>>>(params, grads, exp_avg, exp_avg_sq, n, hyper, state);
  TAIL_CHECK_LAUNCH("adamw_flat launch");
  return UZ_OK;
}
