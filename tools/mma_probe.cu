// Hardware probe: cycles per tcgen05.mma (kind::f16, cta_group::1, both operands in shared memory, K-major
// SWIZZLE_128B) for M in {64,128} and N in {64,128,256}, issued back to back by one thread on every SM.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include "../unet_zoo_b200/csrc/ptx.cuh"
using namespace uz;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s line %d\n", cudaGetErrorString(e_), __LINE__); exit(2);} } while (0)

__global__ void __launch_bounds__(128, 1) probe(int M, int N, int iters, int a_sbo, int a_off, long long* out) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t a_addr = base + a_off, b_addr = base + 65536, bar = base + 65536 + 32768, slot = bar + 16;
  uint8_t* sm = raw + (base - smem_u32(raw));
  for (int i = threadIdx.x; i < (65536 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0x3c003c00u + i;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_fence_init(); }
  if (warp == 0) { tmem_alloc(slot, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(sm + (slot - base));
  if (threadIdx.x == 0) {
    const uint32_t idesc = umma_idesc_bf16(M, N, 0, 0);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t ad = umma_smem_desc(a_addr + k * 32, 16, a_sbo);
        const uint64_t bd = umma_smem_desc(b_addr + k * 32, 16, 1024);
        umma_bf16(tmem + (it & 1) * 256, ad, bd, idesc, 1);
      }
    }
    umma_commit(bar);
    mbar_wait(bar, 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* d; CK(cudaMalloc(&d, 8));
  const int smem = 1024 + 65536 + 32768 + 64;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int iters = 2000;
  int Ms[2] = {128, 64}; int Ns[4] = {32, 64, 128, 256};
  for (int mi = 0; mi < 2; ++mi) for (int ni = 0; ni < 4; ++ni) for (int v = 0; v < 2; ++v) {
    const int sbo = v ? 1280 : 1024, off = v ? 11 * 128 : 0;
    probe<<<148, 128, smem>>>(Ms[mi], Ns[ni], iters, sbo, off, d);
    CK(cudaDeviceSynchronize());
    long long c; CK(cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost));
    printf("M=%3d N=%3d %s: %.1f cycles/MMA  (ideal %d)\n", Ms[mi], Ns[ni], v ? "halo(SBO=1280,+11 rows)" : "dense(SBO=1024)        ",
           (double)c / (iters * 4), 128 * Ns[ni] / 256);
  }
  return 0;
}
