// Forward-type implicit-GEMM convolution on tcgen05 / TMEM, fed by TMA (sm_100a only).
//
// One kernel serves every "pixels x channels" contraction of the UNet hot path:
//   conv3x3 (stride 1/2), its data-gradient (same kernel, flipped weights), conv1x1, ConvTranspose2d k2s2
//   forward (4 output phases) and data-gradient (4 taps on strided views), and nearest-up2x + conv3x3.
// The im2col gather is done by TMA itself: an M tile is a (tn x th x tw) = 128-pixel patch, and filter tap
// (dy, dx) is the same 4-D box shifted by (dx, dy); out-of-image rows/columns are zero-filled by the TMA unit,
// which is exactly the conv zero padding.  A virtual channel concat is a list of segments (one tensor map each).
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer,
// warps 2..5 = epilogue (TMEM -> registers -> bias/ReLU/stats -> bf16 NHWC stores).  Accumulators are
// double-buffered in TMEM so the epilogue of tile i overlaps the main loop of tile i+1.
#include "host_util.h"
#include "igemm.cuh"
#include "ptx.cuh"

namespace uz {

constexpr int kTileM = 128;
constexpr int kBlockK = 64;                     // 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int kABytes = kTileM * kBlockK * 2;   // 16 KiB
constexpr int kFwdThreads = 192;

template <int BN>
struct FwdCfg {
  static constexpr int kBBytes = BN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128) ? 6 : 8;
  static constexpr int kTmemCols = (2 * BN < 32) ? 32 : 2 * BN;  // power of two for BN in {16,32,64,128,256}
  static constexpr int kStatBytes = 4 * 2 * BN * 4 + 4 * 32 * 33 * 4;  // per-warp column sums + transpose scratch
  static constexpr int kSmemBytes = kStages * kStageBytes + kStatBytes + 256 /*barriers*/ + 1024 /*align slack*/;
};

struct TileCoord {
  int n0, y0, x0, n_tile, phase;
};

__device__ __forceinline__ TileCoord decode_tile(const IgemmParams& p, int tile) {
  TileCoord t;
  t.n_tile = tile % p.num_n_tiles;
  int r = tile / p.num_n_tiles;
  int xb = r % p.xb;
  r /= p.xb;
  int yb = r % p.yb;
  r /= p.yb;
  int nb = r % p.nb;
  t.phase = r / p.nb;
  t.x0 = xb << p.tw_log2;
  t.y0 = yb << p.th_log2;
  t.n0 = nb << (7 - p.tw_log2 - p.th_log2);
  return t;
}

template <int BN>
__global__ void __launch_bounds__(kFwdThreads, 1) igemm_fwd_kernel(const __grid_constant__ IgemmParams p) {
  using Cfg = FwdCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;  // SWIZZLE_128B atoms need 1024-B alignment
  uint8_t* smem = smem_raw + (base - raw_addr);

  const uint32_t stage_base = base;
  float* stat_smem = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes);
  const uint32_t bar_base = base + Cfg::kStages * Cfg::kStageBytes + Cfg::kStatBytes;
  // barrier layout: full[kStages], empty[kStages], tfull[2], tempty[2], tmem ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * Cfg::kStages + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem + Cfg::kStages * Cfg::kStageBytes + Cfg::kStatBytes +
                                           8 * (2 * Cfg::kStages + 4));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 4);  // one arrive per epilogue warp
    }
    mbar_fence_init();
    for (int i = 0; i < p.num_segs * p.num_lat; ++i) tma_prefetch_desc(&p.amap[i]);
    tma_prefetch_desc(&p.bmap);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  int kb_per_tap = 0;
  for (int s = 0; s < p.num_segs; ++s) kb_per_tap += p.seg_cblocks[s];
  const int num_kb = kb_per_tap * p.num_taps;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const TileCoord t = decode_tile(p, tile);
        const TapTable& tt = p.taps[t.phase];
        for (int tap = 0; tap < p.num_taps; ++tap) {
          const int ax = t.x0 + tt.dx[tap];
          const int ay = t.y0 + tt.dy[tap];
          const int btap = tt.btap[tap];
          const int lat = tt.lat[tap];
          for (int seg = 0; seg < p.num_segs; ++seg) {
            const CUtensorMap* am = &p.amap[seg * p.num_lat + lat];
            const int ncb = p.seg_cblocks[seg];
            const int k0 = p.seg_k0[seg];
            for (int cb = 0; cb < ncb; ++cb) {
              mbar_wait(empty_bar(stage), phase ^ 1u);
              const uint32_t a_dst = stage_base + stage * Cfg::kStageBytes;
              const uint32_t b_dst = a_dst + kABytes;
              mbar_expect_tx(full_bar(stage), Cfg::kStageBytes);
              tma_load_4d(a_dst, am, full_bar(stage), cb * kBlockK, ax, ay, t.n0);
              tma_load_3d(b_dst, &p.bmap, full_bar(stage), k0 + cb * kBlockK, btap, t.n_tile * BN);
              if (++stage == Cfg::kStages) {
                stage = 0;
                phase ^= 1u;
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(kTileM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t a_addr = stage_base + stage * Cfg::kStageBytes;
          const uint32_t b_addr = a_addr + kABytes;
          const uint64_t adesc = umma_smem_desc(a_addr, 16, 1024);
          const uint64_t bdesc = umma_smem_desc(b_addr, 16, 1024);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            // advance 16 bf16 = 32 B along K inside the swizzle atom: +2 in the (addr >> 4) field
            umma_bf16(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));  // frees the smem slot once these MMAs have read it
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit(tfull_bar(acc));  // accumulator complete -> epilogue
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (2..5)
    const int q = warp & 3;             // TMEM lane quarter this warp may read
    const int ew = warp - 2;            // 0..3, index for smem scratch
    const int et = threadIdx.x - 64;    // 0..127 epilogue thread id
    const int row = q * 32 + lane;      // accumulator row = pixel index inside the tile
    const bool do_stats = (p.flags & kFlagStats) != 0;
    const bool do_relu = (p.flags & kFlagRelu) != 0;
    const bool do_acc = (p.flags & kFlagAccumulate) != 0;
    float* warp_sums = stat_smem + ew * 2 * BN;              // [2][BN]
    float* scratch = stat_smem + 4 * 2 * BN + ew * 32 * 33;  // [32][33]
    // running per-CTA column sums live in registers of epilogue thread `et` (columns et, et+128 per n tile);
    // flushed whenever the n tile changes (static schedule => deterministic order).
    float run_s[(BN + 127) / 128], run_q[(BN + 127) / 128];
#pragma unroll
    for (int i = 0; i < (BN + 127) / 128; ++i) run_s[i] = run_q[i] = 0.f;
    int run_ntile = -1;
    auto flush_stats = [&]() {
      if (run_ntile < 0) return;
#pragma unroll
      for (int i = 0; i < (BN + 127) / 128; ++i) {
        const int c = et + i * 128;
        if (c < BN) {
          const int col = run_ntile * BN + c;
          if (col < p.ncols) {
            float* dst = p.stats + static_cast<size_t>(blockIdx.x) * 2 * p.ncols;
            dst[col] += run_s[i];
            dst[p.ncols + col] += run_q[i];
          }
        }
        run_s[i] = run_q[i] = 0.f;
      }
    };

    int acc = 0;
    uint32_t acc_phase = 0;
    const int tw = 1 << p.tw_log2;
    const int th = 1 << p.th_log2;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const TileCoord t = decode_tile(p, tile);
      const int xi = row & (tw - 1);
      const int yi = (row >> p.tw_log2) & (th - 1);
      const int ni = row >> (p.tw_log2 + p.th_log2);
      const int x = t.x0 + xi, y = t.y0 + yi, n = t.n0 + ni;
      const bool valid = (x < p.W) && (y < p.H) && (n < p.N);
      const int oy = y * p.oscale + p.phase_oy[t.phase];
      const int ox = x * p.oscale + p.phase_ox[t.phase];
      __nv_bfloat16* orow =
          p.out + (static_cast<size_t>(n) * p.Ho + oy) * static_cast<size_t>(p.Wo) * p.out_ld +
          static_cast<size_t>(ox) * p.out_ld;
      if (do_stats && t.n_tile != run_ntile) {
        flush_stats();
        run_ntile = t.n_tile;
      }

      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int ch = 0; ch < BN / 32; ++ch) {
        uint32_t v[32];
        tmem_ld32(t_addr + ch * 32, v);
        tmem_ld_wait();
        const int col0 = t.n_tile * BN + ch * 32;
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < p.ncols) f[j] += __ldg(p.bias + col0 + j);
        }
        if (do_acc && valid) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            if (col0 + g * 8 < p.ncols) {
              const uint4 old = *reinterpret_cast<const uint4*>(orow + col0 + g * 8);
              f[g * 8 + 0] += bf16_lo(old.x);
              f[g * 8 + 1] += bf16_hi(old.x);
              f[g * 8 + 2] += bf16_lo(old.y);
              f[g * 8 + 3] += bf16_hi(old.y);
              f[g * 8 + 4] += bf16_lo(old.z);
              f[g * 8 + 5] += bf16_hi(old.z);
              f[g * 8 + 6] += bf16_lo(old.w);
              f[g * 8 + 7] += bf16_hi(old.w);
            }
          }
        }
        if (do_relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        uint32_t w[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) w[j] = pack_bf16x2(f[2 * j], f[2 * j + 1]);
        if (valid) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            if (col0 + g * 8 < p.ncols) {
              *reinterpret_cast<uint4*>(orow + col0 + g * 8) =
                  make_uint4(w[g * 4 + 0], w[g * 4 + 1], w[g * 4 + 2], w[g * 4 + 3]);
            }
          }
        }
        if (do_stats) {
          // column sums over this warp's 32 pixels of the values as stored (bf16-rounded)
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            scratch[lane * 33 + 2 * j] = valid ? bf16_lo(w[j]) : 0.f;
            scratch[lane * 33 + 2 * j + 1] = valid ? bf16_hi(w[j]) : 0.f;
          }
          __syncwarp();
          float s = 0.f, sq = 0.f;
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float xv = scratch[i * 33 + lane];
            s += xv;
            sq = fmaf(xv, xv, sq);
          }
          warp_sums[ch * 32 + lane] = s;
          warp_sums[BN + ch * 32 + lane] = sq;
          __syncwarp();
        }
      }
      // release the accumulator buffer to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;

      if (do_stats) {
        asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
        for (int i = 0; i < (BN + 127) / 128; ++i) {
          const int c = et + i * 128;
          if (c < BN) {
            float s = 0.f, sq = 0.f;
#pragma unroll
            for (int wq = 0; wq < 4; ++wq) {
              s += stat_smem[wq * 2 * BN + c];
              sq += stat_smem[wq * 2 * BN + BN + c];
            }
            run_s[i] += s;
            run_q[i] += sq;
          }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
    }
    if (do_stats) flush_stats();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

template <int BN>
static cudaError_t launch_fwd_bn(const IgemmParams& p, int grid, cudaStream_t stream) {
  using Cfg = FwdCfg<BN>;
  static bool configured_dev[kMaxDevices] = {false};   // the attribute is per device
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(igemm_fwd_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  igemm_fwd_kernel<BN><<<grid, kFwdThreads, Cfg::kSmemBytes, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_igemm_fwd(const IgemmParams& p, int bn, int grid, cudaStream_t stream) {
  switch (bn) {
    case 32: return launch_fwd_bn<32>(p, grid, stream);
    case 64: return launch_fwd_bn<64>(p, grid, stream);
    case 128: return launch_fwd_bn<128>(p, grid, stream);
    case 256: return launch_fwd_bn<256>(p, grid, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace uz
