// Weight gradient of the 3x3 / pad-1 / stride-1 convolution, second generation (sm_100a only).
//
//   dW[co, tap, ci] = sum over pixels  dZ[p, co] * X[p + shift(tap), ci]
//
// The first-generation kernel (igemm_wgrad.cu) loads one shifted X tile per filter tap, i.e. moves the activation nine
// times through L2 and is L2->SM bandwidth bound (ncu, profiles/r01_a_*).  Here ONE TMA halo box of X
// (64 channels x 10 x 18 pixels, SWIZZLE_128B) feeds all nine taps: tcgen05 shared-memory descriptors may start at any
// 128-byte row of the box and step between 8-row groups with SBO = 1280 B (10 pixels) - verified on hardware by
// tools/halo_probe.cu.  A work unit is (64 output channels) x (64 input channels) x (9 taps) x (a range of 8x16-pixel
// chunks); per chunk it pulls 16 KB of dZ + 22.5 KB of X through L2 for 40 MMAs.
//
// GEMM orientation: both operands are "MN-major" (pixels = K are the slow axis of NHWC).  A = X halo with M = 128 =
// TWO taps x 64 input channels: the descriptor's leading-dimension byte offset (distance between the two 64-element MN
// atoms) is simply the address shift between the two taps (128 B for dx+1, 1280 B for dy+1), so five MMAs cover the nine
// taps (the tenth half is discarded).  B = dZ chunk, N = 64 output channels.  D blocks live in 5 x 64 TMEM columns.
//
// WIDE mode (output-channel blocks of 128): the M = 128 (x2 in a pair), N = 64, K = 16 MMA above reads its 4 KB A operand
// (two taps x 64 channels x 16 pixels) in 32 tensor cycles = 128 B/clk, i.e. the A side alone saturates the 128 B/clk
// shared-memory port whatever happens to B (ncu: tensor-memory path 81-85 % busy, tensor pipe 68-72 %).  Only a wider N
// buys cycles per A byte, but five accumulators of 128 columns do not fit the 512 TMEM columns - so a work unit makes
// TWO passes over its chunk range: tap pairs {0,1} {3,4} {6,7} (3 x 128 columns), then {2,5} {8,-} (2 x 128).  The X halo
// and dZ are loaded once per pass (L2 hits); per CTA and MMA 4 KB (A) + 4 KB (B; 2 KB in a pair) leave shared memory in
// 64 cycles = 128 (96) B/clk.
#include <string.h>

#include "host_util.h"
#include "igemm.cuh"
#include "ptx.cuh"

namespace uz {

constexpr int kW2Threads = 192;
constexpr int kW2HaloW = 10, kW2HaloH = 18;
constexpr int kW2HaloBytes = kW2HaloW * kW2HaloH * 128;  // 23040
constexpr int kW2HaloSlot = 23552;                       // rounded up to 1024
constexpr int kW2DzBytes = 128 * 128;                    // 8x16 pixels x 64 channels bf16
constexpr int kW2TmemCols = 512;
// per (CG, WIDE): bytes of dZ one CTA keeps per stage, stage size, ring depth
// (UP: four dZ phase boxes per stage - 8 KB halves in a pair, 16 KB otherwise)
__host__ __device__ constexpr int w2_dz_phase(int cg) { return cg == 2 ? kW2DzBytes / 2 : kW2DzBytes; }
__host__ __device__ constexpr int w2_dz_slot(int cg, int wide, int up = 0) {
  return up ? 4 * w2_dz_phase(cg) : (wide && cg == 1) ? 2 * kW2DzBytes : kW2DzBytes;
}
__host__ __device__ constexpr int w2_stage_bytes(int cg, int wide, int up = 0) { return kW2HaloSlot + w2_dz_slot(cg, wide, up); }
__host__ __device__ constexpr int w2_stages(int cg, int wide, int up = 0) {
  return up ? (cg == 2 ? 4 : 2) : (wide && cg == 1) ? 4 : 5;
}
__host__ __device__ constexpr int w2_smem_bytes(int cg, int wide, int up = 0) {
  return w2_stages(cg, wide, up) * w2_stage_bytes(cg, wide, up) + 256 + 1024;
}

struct W2Item {
  int co_blk, ci_blk, split, chunk_begin, chunk_end;
};

// `rank` = CTA rank inside the pair (0 when single): the pair's CTAs take adjacent input-channel blocks
__device__ __forceinline__ W2Item w2_decode(const Wgrad2Params& p, int item, int cg, int rank) {
  W2Item w;
  const int units = p.co_blocks * p.ci_units;
  w.split = item / units;
  const int u = item - w.split * units;
  w.co_blk = u / p.ci_units;
  w.ci_blk = (u - w.co_blk * p.ci_units) * cg + rank;
  const int base = p.total_chunks / p.splits, rem = p.total_chunks % p.splits;
  w.chunk_begin = w.split * base + min(w.split, rem);
  w.chunk_end = w.chunk_begin + base + (w.split < rem ? 1 : 0);
  return w;
}

template <int CG, int WIDE, int UP = 0>
__global__ void __launch_bounds__(kW2Threads, 1) igemm_wgrad2_kernel(const __grid_constant__ Wgrad2Params p) {
  static_assert(!(UP && WIDE), "the phase form of the nearest-x2 conv uses narrow tiles (8 x 64 TMEM columns)");
  constexpr bool kPair = CG == 2;
  constexpr int kW2StageBytes = w2_stage_bytes(CG, WIDE, UP);
  constexpr int kW2Stages = w2_stages(CG, WIDE, UP);
  constexpr int kDzPh = w2_dz_phase(CG);      // UP: bytes of one phase's dZ box in this CTA
  constexpr int kTaps = UP ? 16 : 9;          // taps of the partial layout
  constexpr int kBN = WIDE ? 128 : 64;        // output channels per work unit = accumulator columns per tap pair
  constexpr int kPasses = WIDE ? 2 : 1;       // WIDE: tap pairs 0..2 in the first pass over the chunks, 3..4 in the second
  constexpr int kCoBlk = kBN;
  const int rank = kPair ? static_cast<int>(cluster_ctarank()) : 0;
  const bool leader = rank == 0;
  const int first_item = blockIdx.x / CG, item_step = gridDim.x / CG;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_addr);
  const uint32_t bar_base = base + kW2Stages * kW2StageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kW2Stages + s); };
  const uint32_t tfull_bar = bar_base + 8u * (2 * kW2Stages);
  const uint32_t tempty_bar = bar_base + 8u * (2 * kW2Stages + 1);
  const uint32_t tmem_slot = bar_base + 8u * (2 * kW2Stages + 2);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem + kW2Stages * kW2StageBytes + 8 * (2 * kW2Stages + 2));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kW2Stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tfull_bar, 1);
    mbar_init(tempty_bar, 4 * CG);   // pair: the leader's barrier collects both CTAs' epilogue warps
    mbar_fence_init();
    tma_prefetch_desc(&p.xmap);
    if (UP) {
      for (int ph = 0; ph < 4; ++ph) tma_prefetch_desc(kPair ? &p.dzmap_half_ph[ph] : &p.dzmap_ph[ph]);
    } else {
      tma_prefetch_desc((kPair && !WIDE) ? &p.dzmap_half : &p.dzmap);
    }
  }
  if (warp == 1) {
    if (kPair) {
      tmem_alloc2(tmem_slot, kW2TmemCols);
      tmem_relinquish2();
    } else {
      tmem_alloc(tmem_slot, kW2TmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (kPair) cluster_sync_all();   // the peer's barriers exist before a remote completion can reach them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = first_item; item < p.total_items; item += item_step) {
        const W2Item w = w2_decode(p, item, CG, rank);
        for (int pass = 0; pass < kPasses; ++pass)
        for (int chunk = w.chunk_begin; chunk < w.chunk_end; ++chunk) {
          const int xb = chunk % p.xb;
          int r = chunk / p.xb;
          const int yb = r % p.yb;
          const int n = r / p.yb;
          const int x0 = xb * 8, y0 = yb * 16;
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t halo_dst = base + stage * kW2StageBytes;
          const uint32_t dz_dst = halo_dst + kW2HaloSlot;
          if (UP) {   // one X halo + the four dZ phase boxes of the chunk
            if (kPair) {
              if (leader) mbar_expect_tx(full_bar(stage), 2 * kW2HaloBytes + 4 * kW2DzBytes);
              tma_load_4d_pair(halo_dst, &p.xmap, full_bar(stage), w.ci_blk * 64, x0 - 1, y0 - 1, n);
#pragma unroll
              for (int ph = 0; ph < 4; ++ph)
                tma_load_4d_pair(dz_dst + ph * kDzPh, &p.dzmap_half_ph[ph], full_bar(stage), w.co_blk * 64 + rank * 32, x0, y0, n);
            } else {
              mbar_expect_tx(full_bar(stage), kW2HaloBytes + 4 * kW2DzBytes);
              tma_load_4d(halo_dst, &p.xmap, full_bar(stage), w.ci_blk * 64, x0 - 1, y0 - 1, n);
#pragma unroll
              for (int ph = 0; ph < 4; ++ph)
                tma_load_4d(dz_dst + ph * kDzPh, &p.dzmap_ph[ph], full_bar(stage), w.co_blk * 64, x0, y0, n);
            }
          } else if (kPair && WIDE) {   // own halo + own 64 of the 128 dZ channels; completes on the leader's barrier
            if (leader) mbar_expect_tx(full_bar(stage), 2 * kW2HaloBytes + 2 * kW2DzBytes);
            tma_load_4d_pair(halo_dst, &p.xmap, full_bar(stage), w.ci_blk * 64, x0 - 1, y0 - 1, n);
            tma_load_4d_pair(dz_dst, &p.dzmap, full_bar(stage), w.co_blk * kCoBlk + rank * 64, x0, y0, n);
          } else if (kPair) {   // own halo + own half of the dZ chunk; everything completes on the leader's barrier
            if (leader) mbar_expect_tx(full_bar(stage), 2 * kW2HaloBytes + kW2DzBytes);
            tma_load_4d_pair(halo_dst, &p.xmap, full_bar(stage), w.ci_blk * 64, x0 - 1, y0 - 1, n);
            tma_load_4d_pair(dz_dst, &p.dzmap_half, full_bar(stage), w.co_blk * 64 + rank * 32, x0, y0, n);
          } else if (WIDE) {    // two 64-channel dZ boxes side by side (the B descriptor's LBO spans them)
            mbar_expect_tx(full_bar(stage), kW2HaloBytes + 2 * kW2DzBytes);
            tma_load_4d(halo_dst, &p.xmap, full_bar(stage), w.ci_blk * 64, x0 - 1, y0 - 1, n);
            tma_load_4d(dz_dst, &p.dzmap, full_bar(stage), w.co_blk * kCoBlk, x0, y0, n);
            tma_load_4d(dz_dst + kW2DzBytes, &p.dzmap, full_bar(stage), w.co_blk * kCoBlk + 64, x0, y0, n);
          } else {
            mbar_expect_tx(full_bar(stage), kW2HaloBytes + kW2DzBytes);
            tma_load_4d(halo_dst, &p.xmap, full_bar(stage), w.ci_blk * 64, x0 - 1, y0 - 1, n);
            tma_load_4d(dz_dst, &p.dzmap, full_bar(stage), w.co_blk * 64, x0, y0, n);
          }
          if (++stage == kW2Stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (leader && elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128 * CG, kBN, 1, 1);  // both operands MN-major
      // tap pairs stacked along M: {first tap row offset (in halo pixels), byte distance to the second tap}
      //   taps are (dy, dx) in {-1,0,1}^2 with tap index (dy+1)*3 + (dx+1); halo row of tap = (dy+1)*10 + (dx+1)
      constexpr int pair_row[5] = {0 * 10 + 0, 1 * 10 + 0, 2 * 10 + 0, 0 * 10 + 2, 2 * 10 + 2};
      constexpr int pair_lbo[5] = {128, 128, 128, 1280, 128};
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc_phase = 0;
      for (int item = first_item; item < p.total_items; item += item_step) {
        const W2Item w = w2_decode(p, item, CG, rank);
        if (w.chunk_end <= w.chunk_begin) continue;
        for (int pass = 0; pass < kPasses; ++pass) {
        const int b_begin = WIDE ? (pass == 0 ? 0 : 3) : 0;
        const int b_end = WIDE ? (pass == 0 ? 3 : 5) : 5;
        mbar_wait(tempty_bar, acc_phase ^ 1u);
        tc_fence_after();
        bool first = true;
        for (int chunk = w.chunk_begin; chunk < w.chunk_end; ++chunk) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t halo = base + stage * kW2StageBytes;
          const uint32_t dz = halo + kW2HaloSlot;
          const uint32_t halo_lo = umma_desc_lo(halo, 0);
          // WIDE single CTA: N = 128 = two 64-channel boxes, LBO = distance between them
          const uint32_t dz_lo = umma_desc_lo(dz, (WIDE && !kPair) ? kW2DzBytes : 1024);
          constexpr uint32_t a_hi = umma_desc_hi(kW2HaloW * 128);
          // dZ: 128-byte pixel rows (SWIZZLE_128B), or the narrow pair's 64-byte half rows (SWIZZLE_64B = layout code 4)
          constexpr bool kHalfRows = kPair && !WIDE;
          constexpr uint32_t b_hi = kHalfRows ? (((512u >> 4) & 0x3FFFu) | (1u << 14) | (4u << 29)) : umma_desc_hi(1024);
          constexpr uint32_t b_kstep = (kHalfRows ? 1024u : 2048u) >> 4;   // 16 pixels further, in 16-byte units
          if (UP) {
            uint32_t bz[4];
#pragma unroll
            for (int ph = 0; ph < 4; ++ph) bz[ph] = dz_lo + ((ph * kDzPh) >> 4);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
#pragma unroll
              for (int blk = 0; blk < 8; ++blk) {   // (phase, a): taps (a, 0), (a, 1) of the phase stacked along M
                constexpr uint32_t kLbo = (128u >> 4) << 16;
                const int ph = blk >> 1, a = blk & 1;
                const uint32_t a_lo = halo_lo + ((((a + (ph >> 1)) * kW2HaloW + (ph & 1)) * 128 + j * 2 * kW2HaloW * 128) >> 4) + kLbo;
                const uint32_t bz_lo = bz[ph];
                const uint32_t d_col = tmem_base + blk * 64;
                if (kPair)
                  umma2_bf16_words(d_col, a_lo, a_hi, bz_lo + j * b_kstep, b_hi, idesc, (first && j == 0) ? 0u : 1u);
                else
                  umma_bf16_words(d_col, a_lo, a_hi, bz_lo + j * b_kstep, b_hi, idesc, (first && j == 0) ? 0u : 1u);
              }
            }
          } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {  // 16 pixels = image rows 2j, 2j+1 of the chunk
#pragma unroll
            for (int b = 0; b < 5; ++b) {
              if (b < b_begin || b >= b_end) continue;
              // start row of the tap pair (+ two halo rows per K step), LBO = byte distance to the pair's second tap
              const uint32_t a_lo = halo_lo + ((pair_row[b] * 128 + j * 2 * kW2HaloW * 128) >> 4) + ((pair_lbo[b] >> 4) << 16);
              const uint32_t d_col = tmem_base + (b - b_begin) * kBN;
              if (kPair)
                umma2_bf16_words(d_col, a_lo, a_hi, dz_lo + j * b_kstep, b_hi, idesc, (first && j == 0) ? 0u : 1u);
              else
                umma_bf16_words(d_col, a_lo, a_hi, dz_lo + j * b_kstep, b_hi, idesc, (first && j == 0) ? 0u : 1u);
            }
          }
          }
          first = false;
          if (kPair)
            umma2_commit_mc(empty_bar(stage));
          else
            umma_commit(empty_bar(stage));
          if (++stage == kW2Stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (kPair)
          umma2_commit_mc(tfull_bar);
        else
          umma_commit(tfull_bar);
        acc_phase ^= 1u;
        }  // pass
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (2..5): TMEM -> fp32 partials
    const int q = warp & 3;
    const int row = q * 32 + lane;   // accumulator row: half (tap of the pair) and input channel
    const int half = row >> 6;
    const int ci_l = row & 63;
    constexpr int pair_tap[5][2] = {{0, 1}, {3, 4}, {6, 7}, {2, 5}, {8, -1}};
    uint32_t acc_phase = 0;
    for (int item = first_item; item < p.total_items; item += item_step) {
      const W2Item w = w2_decode(p, item, CG, rank);
      const bool has_work = w.chunk_end > w.chunk_begin;
      const int ci = w.ci_blk * 64 + ci_l;
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
      // partials are TILE-major: [split][co block][ci block][tap][co_l][ci_l] - one contiguous 9 x kBN x 64 fp32 region
      // per work unit.  (With the parameter layout [co][tap][ci] a unit scattered 256-byte pieces over a 9*cq*4-byte
      // pitch: the epilogue of the weight-heavy layers ran at < 1 TB/s, 23-28 us per unit.)
      float* pbase = p.partial + ((static_cast<size_t>(w.split) * p.co_blocks + w.co_blk) * p.ci_blocks + w.ci_blk) *
                                     (kTaps * kBN * 64) + ci_l;
#pragma unroll 1
      for (int pass = 0; pass < kPasses; ++pass) {
      const int b_begin = WIDE ? (pass == 0 ? 0 : 3) : 0;
      const int b_end = UP ? 8 : WIDE ? (pass == 0 ? 3 : 5) : 5;
      if (has_work) {
        mbar_wait(tfull_bar, acc_phase);
        tc_fence_after();
      }
#pragma unroll 1
      for (int b = b_begin; b < b_end; ++b) {
        const int tap = UP ? (b >> 1) * 4 + (b & 1) * 2 + half : pair_tap[b < 5 ? b : 0][half];
#pragma unroll 1
        for (int h = 0; h < kBN / 32; ++h) {
          uint32_t v[32];
          if (has_work) {
            tmem_ld32(t_addr + (b - b_begin) * kBN + h * 32, v);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0u;
          }
          if (tap >= 0 && ci < p.cq) {
            float* o = pbase + (tap * kBN + h * 32) * 64;
#pragma unroll
            for (int j = 0; j < 32; ++j) o[j * 64] = __uint_as_float(v[j]);
          }
        }
      }
      if (has_work) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (kPair)
            mbar_arrive_cluster(tempty_bar, 0);
          else
            mbar_arrive(tempty_bar);
        }
        acc_phase ^= 1u;
      }
      }  // pass
    }
  }

  tc_fence_before();
  __syncthreads();
  if (kPair) cluster_sync_all();   // nobody leaves while the peer may still signal this CTA
  if (warp == 1) {
    tc_fence_after();
    if (kPair)
      tmem_dealloc2(tmem_base, kW2TmemCols);
    else
      tmem_dealloc(tmem_base, kW2TmemCols);
  }
}

// dst[co][c_off + ci][tap] (OIHW, row pitch s_row floats) (+)= sum over splits of the tile-major partials.
// block = (output channel, input-channel block): thread = (split group, tap, lane) sums two columns over every
// kRedGroups-th split (256-byte contiguous reads per tap, four loads in flight), the groups are combined through shared
// memory in a fixed order (deterministic), the sums are transposed there and leave as ONE contiguous run of 64 * 9 floats.
// (Three split groups per block: the layers with few work units - nested_unet's 32-channel convolutions, 32 blocks -
// take up to ~300 splits and were bound by the serial chain of dependent loads: 24 us per launch, 5 % of that model's step.)
constexpr int kRedGroups = 3;
__global__ void __launch_bounds__(288 * kRedGroups) wgrad2_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dst,
                                                            int splits, int cp, int cq, int co_blocks, int ci_blocks,
                                                            int bn, long long s_row, int accumulate) {
  __shared__ float stage[kRedGroups][9 * 65];
  const int co = blockIdx.x / ci_blocks, ci_blk = blockIdx.x % ci_blocks;
  const int grp = threadIdx.x / 288, t288 = threadIdx.x - grp * 288;
  const int tap = t288 >> 5, lane = t288 & 31;
  const int co_blk = co / bn, co_l = co % bn;
  const size_t tile = static_cast<size_t>(9) * bn * 64;
  const size_t split_stride = static_cast<size_t>(co_blocks) * ci_blocks * tile;
  const float* q = partial + (static_cast<size_t>(co_blk) * ci_blocks + ci_blk) * tile +
                   (static_cast<size_t>(tap) * bn + co_l) * 64 + lane;
  float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
  int k = grp;
  for (; k + kRedGroups < splits; k += 2 * kRedGroups) {   // four independent loads in flight per thread
    a0 += __ldg(q + k * split_stride);
    b0 += __ldg(q + k * split_stride + 32);
    a1 += __ldg(q + (k + kRedGroups) * split_stride);
    b1 += __ldg(q + (k + kRedGroups) * split_stride + 32);
  }
  if (k < splits) {
    a0 += __ldg(q + k * split_stride);
    b0 += __ldg(q + k * split_stride + 32);
  }
  stage[grp][tap * 65 + lane] = a0 + a1;
  stage[grp][tap * 65 + lane + 32] = b0 + b1;
  __syncthreads();
  const int ncol = min(64, cq - ci_blk * 64);
  float* o = dst + co * s_row + static_cast<long long>(ci_blk) * 64 * 9;
  for (int i = threadIdx.x; i < ncol * 9; i += 288 * kRedGroups) {
    const int c = i / 9, t = i - c * 9;
    float v = stage[0][t * 65 + c];
#pragma unroll
    for (int g = 1; g < kRedGroups; ++g) v += stage[g][t * 65 + c];
    o[i] = accumulate ? o[i] + v : v;
  }
}

// Phase form of the nearest-x2 conv: partial[split][co block][ci block][16 = phase*4 + a*2 + b][64 co][64 ci] -> OIHW.
// Filter row r of output phase py reads low-resolution row offset a - 1 + py with a = a_of(py, r): a_of(0, .) = {0, 1, 1},
// a_of(1, .) = {0, 0, 1} (same for columns), so dW[r][s] = sum over the four phases of partial tap (py, px, a_of(py, r), a_of(px, s)).
__global__ void __launch_bounds__(288) wgrad2_up2x_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dst,
                                                                 int splits, int cp, int cq, int co_blocks, int ci_blocks,
                                                                 long long s_row, int accumulate) {
  __shared__ float stage[9 * 65];
  const int co = blockIdx.x / ci_blocks, ci_blk = blockIdx.x % ci_blocks;
  const int tap = threadIdx.x >> 5, lane = threadIdx.x & 31;   // tap = r * 3 + s of the 3x3 filter
  const int r = tap / 3, sx = tap - r * 3;
  const int co_blk = co / 64, co_l = co % 64;
  const size_t tile = static_cast<size_t>(16) * 64 * 64;
  const size_t split_stride = static_cast<size_t>(co_blocks) * ci_blocks * tile;
  const float* base = partial + (static_cast<size_t>(co_blk) * ci_blocks + ci_blk) * tile + static_cast<size_t>(co_l) * 64 + lane;
  float a0 = 0.f, b0 = 0.f;
#pragma unroll
  for (int ph = 0; ph < 4; ++ph) {
    const int py = ph >> 1, px = ph & 1;
    const int a = py == 0 ? (r == 0 ? 0 : 1) : (r == 2 ? 1 : 0);
    const int b = px == 0 ? (sx == 0 ? 0 : 1) : (sx == 2 ? 1 : 0);
    const float* q = base + static_cast<size_t>(ph * 4 + a * 2 + b) * 64 * 64;
    for (int k = 0; k < splits; ++k) {
      a0 += __ldg(q + k * split_stride);
      b0 += __ldg(q + k * split_stride + 32);
    }
  }
  stage[tap * 65 + lane] = a0;
  stage[tap * 65 + lane + 32] = b0;
  __syncthreads();
  const int ncol = min(64, cq - ci_blk * 64);
  float* o = dst + co * s_row + static_cast<long long>(ci_blk) * 64 * 9;
  for (int i = threadIdx.x; i < ncol * 9; i += 288) {
    const int c = i / 9, t = i - c * 9;
    const float v = stage[t * 65 + c];
    o[i] = accumulate ? o[i] + v : v;
  }
}

cudaError_t launch_wgrad2_reduce(const Wgrad2Params& p, float* dst, long long s_row, int accumulate,
                                 cudaStream_t stream) {
  if (p.up2x) {
    const long long blocks_u = static_cast<long long>(p.cp) * p.ci_blocks;
    if (blocks_u < 1 || blocks_u > 0x7fffffffLL) return cudaErrorInvalidValue;
    wgrad2_up2x_reduce_kernel<<<static_cast<int>(blocks_u), 288, 0, stream>>>(p.partial, dst, p.splits, p.cp, p.cq,
                                                                             p.co_blocks, p.ci_blocks, s_row, accumulate);
    return cudaGetLastError();
  }
  const long long blocks = static_cast<long long>(p.cp) * p.ci_blocks;
  if (blocks < 1 || blocks > 0x7fffffffLL) return cudaErrorInvalidValue;
  wgrad2_reduce_kernel<<<static_cast<int>(blocks), 288 * kRedGroups, 0, stream>>>(p.partial, dst, p.splits, p.cp, p.cq, p.co_blocks,
                                                                     p.ci_blocks, p.wide ? 128 : 64, s_row, accumulate);
  return cudaGetLastError();
}

template <int CG, int WIDE, int UP = 0>
static cudaError_t launch_wgrad2_cg(const Wgrad2Params& p, int grid, cudaStream_t stream) {
  constexpr int kW2SmemBytes = w2_smem_bytes(CG, WIDE, UP);
  static bool configured_dev[kMaxDevices] = {false};   // the attribute is per device
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(igemm_wgrad2_kernel<CG, WIDE, UP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kW2SmemBytes);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kW2Threads);
  cfg.dynamicSmemBytes = kW2SmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = CG > 1 ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, igemm_wgrad2_kernel<CG, WIDE, UP>, p);
}

cudaError_t launch_igemm_wgrad2(const Wgrad2Params& p, int grid, cudaStream_t stream) {
  if (p.up2x) return p.pair ? launch_wgrad2_cg<2, 0, 1>(p, grid, stream) : launch_wgrad2_cg<1, 0, 1>(p, grid, stream);
  if (p.wide) return p.pair ? launch_wgrad2_cg<2, 1>(p, grid, stream) : launch_wgrad2_cg<1, 1>(p, grid, stream);
  return p.pair ? launch_wgrad2_cg<2, 0>(p, grid, stream) : launch_wgrad2_cg<1, 0>(p, grid, stream);
}

}  // namespace uz
