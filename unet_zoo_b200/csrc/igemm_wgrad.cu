// Weight-gradient-type GEMM on tcgen05 / TMEM (sm_100a only).
//
//   D[t][i, j] = sum over pixels  P[pixel, i] * Q[pixel + shift(t), j]
//
// P is the un-shifted operand (dY for conv3x3, X for ConvTranspose2d), Q the shifted one.  The reduction
// index (pixels) is the slow axis of both NHWC tensors, so both operands are "MN-major" for the tensor core:
// TMA drops a (64-pixel x 64-channel) box into smem as 64 rows of 128 B (SWIZZLE_128B) and the UMMA
// descriptors describe that as an MN-major atom (LBO = next 64-channel atom, SBO = next 8-pixel group).
// A work item is (128 rows of P-channels) x (slab of up to 4 (tap, 64-channel) atoms of Q) x (split-K slice);
// fp32 partials go to a workspace and a small reduce kernel sums the splits into the parameter-gradient layout.
#include "host_util.h"
#include "igemm.cuh"
#include "ptx.cuh"

namespace uz {

constexpr int kWgThreads = 192;
constexpr int kChunkPix = 64;                       // reduction chunk (pixels per pipeline stage)
constexpr int kAtomBytes = kChunkPix * 128;         // 64 pixels x 64 channels bf16 = 8 KiB
constexpr int kWgPBytes = 2 * kAtomBytes;           // 128 P-channels
constexpr int kWgQBytes = 4 * kAtomBytes;           // up to 256 Q-columns
constexpr int kWgStageBytes = kWgPBytes + kWgQBytes;  // 48 KiB
constexpr int kWgStages = 4;
constexpr int kWgTmemCols = 512;
constexpr int kWgSmemBytes = kWgStages * kWgStageBytes + 256 + 1024;

struct WgItem {
  int m_tile, slab, split, atom0, natoms, chunk_begin, chunk_end;
};

__device__ __forceinline__ WgItem decode_item(const WgradParams& p, int item) {
  WgItem w;
  w.split = item % p.splits;
  int r = item / p.splits;
  w.slab = r % p.n_slabs;
  w.m_tile = r / p.n_slabs;
  w.atom0 = w.slab * 4;
  w.natoms = min(4, p.num_atoms - w.atom0);
  const int total = p.num_phases * p.chunks_per_phase;
  const int base = total / p.splits, rem = total % p.splits;
  w.chunk_begin = w.split * base + min(w.split, rem);
  w.chunk_end = w.chunk_begin + base + (w.split < rem ? 1 : 0);
  return w;
}

__global__ void __launch_bounds__(kWgThreads, 1) igemm_wgrad_kernel(const __grid_constant__ WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_addr);
  const uint32_t bar_base = base + kWgStages * kWgStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kWgStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kWgStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kWgStages + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kWgStages + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem + kWgStages * kWgStageBytes + 8 * (2 * kWgStages + 4));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kWgStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 4);
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, kWgTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
        const WgItem w = decode_item(p, item);
        const int p_atoms = min(2, (p.cp - w.m_tile * 128 + 63) / 64);
        const uint32_t tx_bytes = static_cast<uint32_t>(p_atoms + w.natoms) * kAtomBytes;
        for (int chunk = w.chunk_begin; chunk < w.chunk_end; ++chunk) {
          const int ph = chunk / p.chunks_per_phase;
          int r = chunk - ph * p.chunks_per_phase;
          const int xb = r % p.xb;
          r /= p.xb;
          const int yb = r % p.yb;
          const int nb = r / p.yb;
          const int x0 = xb << p.tw_log2, y0 = yb << p.th_log2, n0 = nb << (6 - p.tw_log2 - p.th_log2);
          const TapTable& tt = p.taps[ph];
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t p_dst = base + stage * kWgStageBytes;
          const uint32_t q_dst = p_dst + kWgPBytes;
          mbar_expect_tx(full_bar(stage), tx_bytes);
          for (int a = 0; a < p_atoms; ++a)
            tma_load_4d(p_dst + a * kAtomBytes, &p.pmap[ph], full_bar(stage), w.m_tile * 128 + a * 64, x0, y0, n0);
          for (int g = 0; g < w.natoms; ++g) {
            const int atom = w.atom0 + g;
            const int tap = atom / p.q_cblocks;
            const int cb = atom - tap * p.q_cblocks;
            tma_load_4d(q_dst + g * kAtomBytes, &p.qmap[tt.lat[tap]], full_bar(stage), cb * 64, x0 + tt.dx[tap],
                        y0 + tt.dy[tap], n0);
          }
          if (++stage == kWgStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
        const WgItem w = decode_item(p, item);
        const uint32_t idesc = umma_idesc_bf16(128, 64 * w.natoms, 1, 1);  // both operands MN-major
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
        bool first = true;
        for (int chunk = w.chunk_begin; chunk < w.chunk_end; ++chunk) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t p_addr = base + stage * kWgStageBytes;
          const uint32_t q_addr = p_addr + kWgPBytes;
#pragma unroll
          for (int k = 0; k < kChunkPix / 16; ++k) {
            // 16 pixels = two 8-row groups of 1024 B; next 64-channel atom is kAtomBytes away
            const uint64_t adesc = umma_smem_desc(p_addr + k * 2048, kAtomBytes, 1024);
            const uint64_t bdesc = umma_smem_desc(q_addr + k * 2048, kAtomBytes, 1024);
            umma_bf16(d_tmem, adesc, bdesc, idesc, (first && k == 0) ? 0u : 1u);
          }
          first = false;
          umma_commit(empty_bar(stage));
          if (++stage == kWgStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit(tfull_bar(acc));
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else {
    const int q = warp & 3;
    const int row_in_tile = q * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
      const WgItem w = decode_item(p, item);
      const int row = w.m_tile * 128 + row_in_tile;
      const bool row_ok = row < p.cp;
      const bool has_work = w.chunk_end > w.chunk_begin;
      if (has_work) {
        mbar_wait(tfull_bar(acc), acc_phase);
        tc_fence_after();
      }
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 256;
      float* prow = p.partial + (static_cast<size_t>(w.split) * p.cp + row) * p.num_taps * p.cq;
      for (int g = 0; g < w.natoms; ++g) {
        const int atom = w.atom0 + g;
        const int tap = atom / p.q_cblocks;
        const int cb = atom - tap * p.q_cblocks;
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
          uint32_t v[32];
          if (has_work) {
            tmem_ld32(t_addr + g * 64 + h * 32, v);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0u;
          }
          const int c0 = cb * 64 + h * 32;
          if (row_ok) {
            float* dst = prow + static_cast<size_t>(tap) * p.cq + c0;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (c0 + j < p.cq)
                *reinterpret_cast<float4*>(dst + j) =
                    make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                __uint_as_float(v[j + 3]));
            }
          }
        }
      }
      if (has_work) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(acc));
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kWgTmemCols);
  }
}

cudaError_t launch_igemm_wgrad(const WgradParams& p, int grid, cudaStream_t stream) {
  static bool configured_dev[kMaxDevices] = {false};   // the attribute is per device
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    cudaError_t e =
        cudaFuncSetAttribute(igemm_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmemBytes);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  igemm_wgrad_kernel<<<grid, kWgThreads, kWgSmemBytes, stream>>>(p);
  return cudaGetLastError();
}

// Sum split-K partials [splits][rows][taps][cols] into a strided fp32 destination (the parameter-gradient layout):
//   dst[row*s_row + col*s_col + tap*s_tap] (+)= sum_s partial[s][row][tap][col]
__global__ void wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dst, int splits, int rows,
                                    int taps, int cols, long long s_row, long long s_col, long long s_tap,
                                    int accumulate) {
  // block = (one row, 32 columns); thread = (tap, column).  Partial reads coalesce over the 32 columns of each tap; the
  // sums are transposed through shared memory so that the block writes ONE contiguous run of 32*taps floats of the
  // OIHW / IOHW destination (s_col == taps, s_tap == 1) instead of one scattered sector per element.
  __shared__ float stage[9 * 33];
  const int cchunks = (cols + 31) / 32;
  const int row = blockIdx.x / cchunks;
  const int c0 = (blockIdx.x % cchunks) * 32;
  const int tap = threadIdx.x / 32, cl = threadIdx.x % 32;
  const int col = c0 + cl;
  const long long total = static_cast<long long>(rows) * taps * cols;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (col < cols) {
    const float* q = partial + (static_cast<long long>(row) * taps + tap) * cols + col;
    int k = 0;
    for (; k + 4 <= splits; k += 4) {
      s0 += __ldg(q + (k + 0) * total);
      s1 += __ldg(q + (k + 1) * total);
      s2 += __ldg(q + (k + 2) * total);
      s3 += __ldg(q + (k + 3) * total);
    }
    for (; k < splits; ++k) s0 += __ldg(q + k * total);
  }
  const float s = (s0 + s1) + (s2 + s3);
  const bool dense = s_tap == 1 && s_col == taps;
  if (!dense) {
    if (col < cols) {
      float* o = dst + row * s_row + col * s_col + tap * s_tap;
      *o = accumulate ? *o + s : s;
    }
    return;
  }
  stage[tap * 33 + cl] = s;
  __syncthreads();
  const int ncol = min(32, cols - c0);
  const int i = threadIdx.x;                 // position in the contiguous run: i = column_local * taps + tap
  if (i < ncol * taps) {
    const int c = i / taps, t = i - c * taps;
    float* o = dst + row * s_row + static_cast<long long>(c0) * taps + i;
    const float v = stage[t * 33 + c];
    *o = accumulate ? *o + v : v;
  }
}

cudaError_t launch_wgrad_reduce(const float* partial, float* dst, int splits, int rows, int taps, int cols,
                                long long s_row, long long s_col, long long s_tap, int accumulate,
                                cudaStream_t stream) {
  if (taps < 1 || taps > 9) return cudaErrorInvalidValue;
  const long long blocks = static_cast<long long>(rows) * ((cols + 31) / 32);
  if (blocks < 1 || blocks > 0x7fffffffLL) return cudaErrorInvalidValue;
  wgrad_reduce_kernel<<<static_cast<int>(blocks), taps * 32, 0, stream>>>(partial, dst, splits, rows, taps, cols, s_row,
                                                                          s_col, s_tap, accumulate);
  return cudaGetLastError();
}

}  // namespace uz
