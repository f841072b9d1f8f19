// Hardware probe: can tcgen05.mma read the nine 3x3-tap views of ONE TMA-loaded halo tile (64 ch x 10 x 18 pixels,
// SWIZZLE_128B) through smem descriptors whose start address is shifted by whole 128-byte rows and whose 8-row group
// stride (SBO) is 1280 B (= 10 pixels), i.e. not a multiple of the 1024-B swizzle period?
// mode 0: base_offset field = 0;  mode 1: base_offset = (start >> 7) & 7.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../unet_zoo_b200/csrc/ptx.cuh"
using namespace uz;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s line %d\n", cudaGetErrorString(e_), __LINE__); exit(2);} } while (0)

constexpr int HW = 10, HH = 18;                 // halo extent (x, y)
constexpr int A_BYTES = HW * HH * 128;          // 23040
constexpr int B_BYTES = 9 * 64 * 128;           // 9 taps x 64 couts x 64 cin

struct Params { CUtensorMap amap, bmap; float* out; int x0, y0, mode; };

__global__ void __launch_bounds__(128, 1) probe_kernel(const __grid_constant__ Params p) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t a_addr = base, b_addr = base + 23552 /* 23 KiB, 1024-aligned */, bar = b_addr + B_BYTES, bar2 = bar + 8,
                 slot = bar + 16;
  uint8_t* sm = raw + (base - smem_u32(raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init(bar2, 1); mbar_fence_init(); }
  if (warp == 0) { tmem_alloc(slot, 64); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(sm + (slot - base));
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, A_BYTES + B_BYTES);
    tma_load_3d(a_addr, &p.amap, bar, 0, p.x0 - 1, p.y0 - 1);
    for (int t = 0; t < 9; ++t) tma_load_3d(b_addr + t * 8192, &p.bmap, bar, 0, 0, t);
    mbar_wait(bar, 0);
    tc_fence_after();
    constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
    for (int t = 0; t < 9; ++t) {
      const int dy = t / 3, dx = t % 3;  // already +1 shifted
      for (int k = 0; k < 4; ++k) {
        const uint32_t a_start = a_addr + (dy * HW + dx) * 128 + k * 32;
        uint64_t ad = umma_smem_desc(a_start, 16, HW * 128);
        if (p.mode == 1) ad |= static_cast<uint64_t>((a_start >> 7) & 7) << 49;
        const uint64_t bd = umma_smem_desc(b_addr + t * 8192 + k * 32, 16, 1024);
        umma_bf16(tmem, ad, bd, idesc, (t | k) != 0);
      }
    }
    umma_commit(bar2);
  }
  __syncthreads();
  mbar_wait(bar2, 0);
  tc_fence_after();
  uint32_t v[32];
  for (int h = 0; h < 2; ++h) {
    tmem_ld32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + h * 32, v);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) p.out[(warp * 32 + lane) * 64 + h * 32 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 64); }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int H = 24, W = 16, C = 64;
  std::vector<float> X((size_t)H * W * C), Wt((size_t)9 * 64 * 64);
  uint32_t s = 12345;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return ((s >> 8) & 0xFFFF) / 32768.0f - 1.0f; };
  for (auto& v : X) v = __bfloat162float(__float2bfloat16(rnd()));
  for (auto& v : Wt) v = __bfloat162float(__float2bfloat16(rnd() * 0.25f));
  std::vector<__nv_bfloat16> xb(X.size()), wb(Wt.size());
  for (size_t i = 0; i < X.size(); ++i) xb[i] = __float2bfloat16(X[i]);
  for (size_t i = 0; i < Wt.size(); ++i) wb[i] = __float2bfloat16(Wt[i]);
  __nv_bfloat16 *dx, *dw; float* dout;
  CK(cudaMalloc(&dx, xb.size() * 2)); CK(cudaMalloc(&dw, wb.size() * 2)); CK(cudaMalloc(&dout, 128 * 64 * 4));
  CK(cudaMemcpy(dx, xb.data(), xb.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dw, wb.data(), wb.size() * 2, cudaMemcpyHostToDevice));
  void* fp = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  EncodeTiledFn enc = (EncodeTiledFn)fp;
  Params p;
  {
    cuuint64_t dims[3] = {64, (cuuint64_t)W, (cuuint64_t)H}; cuuint64_t str[2] = {128, (cuuint64_t)W * 128};
    cuuint32_t box[3] = {64, HW, HH}; cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(&p.amap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, dx, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r) { printf("amap encode failed %d\n", (int)r); return 2; }
    cuuint64_t d2[3] = {64, 64, 9}; cuuint64_t s2[2] = {128, 64 * 128}; cuuint32_t b2[3] = {64, 64, 1};
    r = enc(&p.bmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, dw, d2, s2, b2, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r) { printf("bmap encode failed %d\n", (int)r); return 2; }
  }
  p.out = dout; p.x0 = 4; p.y0 = 3;
  const int smem = 1024 + 23552 + B_BYTES + 64;
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  for (int mode = 0; mode < 2; ++mode) {
    p.mode = mode;
    probe_kernel<<<1, 128, smem>>>(p);
    CK(cudaDeviceSynchronize());
    std::vector<float> out(128 * 64);
    CK(cudaMemcpy(out.data(), dout, out.size() * 4, cudaMemcpyDeviceToHost));
    double maxerr = 0, maxref = 0;
    for (int r = 0; r < 128; ++r) {
      const int y = p.y0 + r / 8, x = p.x0 + r % 8;
      for (int co = 0; co < 64; ++co) {
        double acc = 0;
        for (int t = 0; t < 9; ++t) {
          const int yy = y + t / 3 - 1, xx = x + t % 3 - 1;
          if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
          for (int ci = 0; ci < 64; ++ci) acc += (double)X[((size_t)yy * W + xx) * C + ci] * Wt[((size_t)t * 64 + co) * 64 + ci];
        }
        maxerr = fmax(maxerr, fabs(acc - out[r * 64 + co]));
        maxref = fmax(maxref, fabs(acc));
      }
    }
    printf("halo probe mode %d (base_offset %s): max_err=%.5f max_ref=%.3f -> %s\n", mode, mode ? "=(addr>>7)&7" : "=0",
           maxerr, maxref, maxerr < 1e-2 ? "MATCH" : "MISMATCH");
  }
  return 0;
}
