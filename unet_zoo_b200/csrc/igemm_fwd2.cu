// Forward-type implicit-GEMM convolution, second generation (sm_100a only): conv3x3 (stride 1, pad 1), its data
// gradient and conv1x1 on 8x16-pixel output tiles.
//
// Why: ncu showed the first-generation kernel (igemm_fwd.cu) bound by L2->SM bandwidth - every filter tap re-loads
// its shifted A tile, the weight tile is re-loaded per output tile, and the epilogue scatters 16-byte stores.  Here
//   * ONE TMA halo box (64 ch x 10 x 18 px, SWIZZLE_128B) per 64-channel block feeds all nine taps: the tcgen05
//     descriptor of tap (dy,dx) starts (dy+1)*10+(dx+1) rows into the box with SBO = 1280 B (tools/halo_probe.cu);
//   * weights are either resident in shared memory for the whole launch (layers whose packed weights fit) or streamed
//     through their own ring by a dedicated producer warp, multicast across a thread-block cluster when one is used;
//   * the epilogue writes TMEM -> registers -> swizzled staging tile -> TMA store (or TMA reduce-add when the output
//     accumulates), and takes the BatchNorm column sums from the staged bf16 tile (conflict-free column reads).
//
// Warp roles (384 threads): 0 = A (halo) producer, 1 = TMEM allocator + MMA issuer, 2 = B producer, 3 = idle,
// 4..11 = epilogue: two groups of four warps (one per TMEM lane quarter), the groups take alternate tiles.
#include <stdlib.h>

#include <type_traits>
#include <string.h>

#include "host_util.h"
#include "igemm.cuh"
#include "ptx.cuh"

namespace uz {

// Diagnostic knock-outs (compile-time, tools/build_variants.sh ... -DUZ_F2_KNOCK=<mask>; results are garbage, only the
// timing is meaningful): 1 = the A producer stops issuing TMA loads once the ring has been filled once (barriers still
// complete), 2 = the issuer skips the UTCHMMA instructions (commits stay), 4 = the epilogue only hands the accumulator
// back, 8 = the epilogue skips the TMA store, 16 = the epilogue skips the BatchNorm column sums, 32 = no aempty commits
// (the producer does not wait for free stages either; may deadlock - run under `timeout`), 64 = no tcgen05 fences in the
// issuer, 128 = one EXTRA tcgen05.commit per tile onto a dummy barrier (prices a commit), 256 = the issuer does not wait
// for a drained accumulator, 512 / 1024 = the tfull / aempty commit is issued BEFORE the tile's MMAs (the handshake then
// does not wait for their completion).  0 (the product build)
// compiles to exactly the code without them.
#ifndef UZ_F2_KNOCK
#define UZ_F2_KNOCK 0
#endif
constexpr int kF2Knock = UZ_F2_KNOCK;
constexpr int kF2Threads = 384;
constexpr int kF2EpiThreads = 256;
constexpr int kF2StgBytes = 128 * 128;  // one staged 128-pixel x 64-channel bf16 chunk

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_shared_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_shared_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}

struct F2Tile {
  int n_tile, n, x0, y0;
  bool dummy;
};
// work item -> (N tile, first pixel tile).  n_minor: items of one cluster-sized group of pixel tiles are adjacent
// (group-major, N tile, rank) so that the CTAs of a cluster / pair still share their N tile.
__device__ __forceinline__ void f2_item(const Igemm2Params& p, int tile, int* n_tile, int* m_item) {
  if (p.n_minor) {
    const int grp = tile / p.cluster, r = tile - grp * p.cluster;
    *n_tile = grp % p.num_n_tiles;
    *m_item = (grp / p.num_n_tiles) * p.cluster + r;
  } else {
    *n_tile = tile / p.m_tiles;
    *m_item = tile - *n_tile * p.m_tiles;
  }
}
__device__ __forceinline__ F2Tile f2_decode(const Igemm2Params& p, int tile, int sub = 0) {
  F2Tile t;
  int m_item;
  f2_item(p, tile, &t.n_tile, &m_item);
  const int m = m_item * p.mp + sub;
  t.dummy = m >= p.m_tiles_real;
  const int xb = m % p.xb;
  const int r = m / p.xb;
  const int yb = r % p.yb;
  t.n = t.dummy ? p.N : r / p.yb;  // n == N is outside every tensor map: loads are zero-filled, stores clipped
  t.x0 = xb * 8;
  t.y0 = yb * 16;
  return t;
}

// FAST: stationary weights, all nine taps of every segment, one CTA, one pixel tile per item - the issue loop of these
// (short-K, <= 128 output channels) layers then carries no ring / tap-mask logic: the generic loop spends ~35 instructions
// (six R2UR) per tap, ~150 cycles, where the four N = 64 MMAs of a tap last 128 - measured by knock-outs (UZ_F2_KNOCK): the
// per-tile time did not move without TMA loads or without the epilogue, only without the MMAs.
// FAST = 4: all four K steps; 1..3: EVERY channel block of the launch holds exactly FAST x 16 channels (single-block segments
// of 16 / 32 / 48 channels: nested_unet level 0, the RSU interiors of u2net) - the K-step count is then a compile-time
// constant and the MMAs of a tile are one straight line of UIADD3 / UTCHMMA (the dynamic K-skip loop needs an R2UR and a
// branch per MMA: ~74 cycles per MMA where an N = 64 MMA lasts 32); 5: FAST loop with the dynamic K-skip; 0: generic loop.
template <int BN, int MP, int CG, bool KSKIP, int FAST>
__global__ void __launch_bounds__(kF2Threads, 1) igemm_fwd2_kernel(const __grid_constant__ Igemm2Params p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_addr);
  const uint32_t a_base = base + p.a_off, b_base = base + p.b_off, stg_base = base + p.stg_off;
  const uint32_t bars = base + p.bar_off;
  auto afull = [&](int s) { return bars + 8u * s; };
  auto aempty = [&](int s) { return bars + 64u + 8u * s; };
  auto bfull = [&](int s) { return bars + 128u + 8u * s; };
  auto bempty = [&](int s) { return bars + 256u + 8u * s; };
  auto tfull = [&](int s) { return bars + 384u + 8u * s; };     // up to 8 accumulator stages
  auto tempty = [&](int s) { return bars + 448u + 8u * s; };
  const uint32_t tmem_slot = bars + 56u;                        // the eighth afull slot (a_stages <= 6)
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem + p.bar_off + 56);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int cs = p.cluster;
  const uint32_t rank = cs > 1 ? cluster_ctarank() : 0u;
  const uint16_t mc_mask = static_cast<uint16_t>((1u << cs) - 1u);
  const bool ring = p.b_stages > 0;
  const uint32_t a_half = static_cast<uint32_t>(p.halo_bytes + 1023) & ~1023u;  // one pixel tile's halo box
  const uint32_t a_slot = a_half * MP;
  constexpr bool kPair = CG == 2;                // CTA pair: see Igemm2Params::pair
  static_assert(!kPair || MP == 1, "pair mode has one pixel tile per CTA");
  constexpr uint32_t kBItem = BN * 128 / CG;     // bytes of one (cblock, tap) weight tile held by THIS CTA
  const bool leader = !kPair || rank == 0;
  // accumulator stages: as many as TMEM holds, at most 4 (short-K layers are bound by the tfull -> epilogue -> tempty ->
  // MMA round trip, which crosses SMs in pair mode)
#ifndef UZ_F2_MAX_ACC
#define UZ_F2_MAX_ACC 8
#endif
  constexpr int kAcc = 512 / (MP * BN) > UZ_F2_MAX_ACC ? UZ_F2_MAX_ACC : 512 / (MP * BN);
  constexpr uint32_t kTmemCols = kAcc * MP * BN;
  static_assert(kTmemCols <= 512, "accumulators exceed TMEM");

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < p.a_stages; ++s) {
      mbar_init(afull(s), 1);
      mbar_init(aempty(s), 1);
    }
    const int nb = ring ? p.b_stages : 1;
    for (int s = 0; s < nb; ++s) {
      mbar_init(bfull(s), 1);
      mbar_init(bempty(s), kPair ? 1 : cs);
    }
    for (int s = 0; s < kAcc; ++s) {
      mbar_init(tfull(s), 1);
      mbar_init(tempty(s), 4 * CG);   // one epilogue group (four warps) drains a tile; pair: the leader's barrier
                                      // collects both CTAs' groups
    }
    if (kF2Knock & 128) mbar_init(bars + 48u, 1);   // dummy barrier (seventh afull slot, unused: a_stages <= 6)
    mbar_fence_init();
    for (int s = 0; s < p.num_segs; ++s) tma_prefetch_desc(&p.amap[s]);
    tma_prefetch_desc(&p.bmap);
    for (int i = 0; i < ((p.nt_per_phase || p.ph_major) ? 4 : 1); ++i) tma_prefetch_desc(&p.omap[i]);
  }
  if (warp == 1) {
    if (kPair) {
      tmem_alloc2(tmem_slot, kTmemCols);
      tmem_relinquish2();
    } else {
      tmem_alloc(tmem_slot, kTmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (cs > 1) cluster_sync_all();  // every CTA's barriers exist before any multicast / remote commit can arrive
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  const int cidx = blockIdx.x / cs;
  const int nclusters = gridDim.x / cs;
  const int ngroups = p.total_tiles / cs;

  if (warp == 0) {
    // ------------------------------------------------------------------ A producer: one halo box per (segment, cblock)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      [[maybe_unused]] int issued = 0;
      for (int g = cidx; g < ngroups; g += nclusters) {
        F2Tile t[MP];
#pragma unroll
        for (int sub = 0; sub < MP; ++sub) t[sub] = f2_decode(p, g * cs + rank, sub);
        for (int seg = 0; seg < p.num_segs; ++seg) {
          const int ncb = p.seg_cblocks[seg];
          for (int cb = 0; cb < ncb; ++cb) {
            if (!(kF2Knock & 32)) mbar_wait(aempty(stage), phase ^ 1u);
            if (kPair) {   // both halos complete on the leader's barrier
              if (leader) mbar_expect_tx(afull(stage), p.halo_bytes * 2);
              tma_load_4d_pair(a_base + stage * a_slot, &p.amap[seg], afull(stage), cb * 64, t[0].x0 - p.pad,
                               t[0].y0 - p.pad, t[0].n);
            } else if ((kF2Knock & 1) && issued >= p.a_stages) {
              mbar_arrive(afull(stage));
            } else {
              if (kF2Knock & 1) ++issued;
              mbar_expect_tx(afull(stage), p.halo_bytes * MP);
#pragma unroll
              for (int sub = 0; sub < MP; ++sub)
                tma_load_4d(a_base + stage * a_slot + sub * a_half, &p.amap[seg], afull(stage), cb * 64,
                            t[sub].x0 - p.pad, t[sub].y0 - p.pad, t[sub].n);
            }
            if (++stage == p.a_stages) {
              stage = 0;
              phase ^= 1u;
            }
          }
        }
      }
    }
  } else if (warp == 2) {
    // ------------------------------------------------------------------ B producer: weight tiles
    if (elect_one()) {
      if (!ring) {
        // stationary: every (cblock, tap) tile is loaded once and stays for the whole launch (num_n_tiles == 1)
        if (leader) mbar_expect_tx(bfull(0), static_cast<uint32_t>(p.b_items) * kBItem * CG);
        int item = 0;
        for (int seg = 0; seg < p.num_segs; ++seg)
          for (int cb = 0; cb < p.seg_cblocks[seg]; ++cb)
            for (int tap = 0; tap < p.num_taps; ++tap) {
              if (!((p.seg_taps[seg] >> tap) & 1)) continue;
              if (kPair)   // this CTA's half of the weight rows
                tma_load_3d_pair(b_base + item * kBItem, &p.bmap, bfull(0), p.seg_k0[seg] + cb * 64,
                                 p.seg_tap_b[seg][tap], rank * (BN / 2));
              else
                tma_load_3d(b_base + item * kBItem, &p.bmap, bfull(0), p.seg_k0[seg] + cb * 64, p.seg_tap_b[seg][tap], 0);
              ++item;
            }
      } else {
        int stage = 0;
        uint32_t phase = 0;
        const int rows = BN / cs;
        for (int g = cidx; g < ngroups; g += nclusters) {
          int n_tile, m_unused;
          f2_item(p, g * cs, &n_tile, &m_unused);
          const int b_image = p.b_batched ? f2_decode(p, g * cs, 0).n : -1;   // batched GEMM: this image's B matrix
          const int b_phase = p.nt_per_phase ? n_tile / p.nt_per_phase : -1;
          const int b_row0 = (p.nt_per_phase ? n_tile % p.nt_per_phase : n_tile) * BN;
          for (int seg = 0; seg < p.num_segs; ++seg) {
            const int ncb = p.seg_cblocks[seg];
            for (int cb = 0; cb < ncb; ++cb) {
              const int k = p.seg_k0[seg] + cb * 64;
              for (int tap = 0; tap < p.num_taps; ++tap) {
                if (!((p.seg_taps[seg] >> tap) & 1)) continue;
                mbar_wait(bempty(stage), phase ^ 1u);
                if (leader) mbar_expect_tx(bfull(stage), kBItem * CG);
                const uint32_t dst = b_base + stage * kBItem;
                const int btap = b_image >= 0 ? b_image
                                 : b_phase >= 0 ? (p.ph_ntaps ? b_phase * 4 + tap : b_phase) : p.seg_tap_b[seg][tap];
                if (p.ph_major) {   // four 64-row boxes (one per output phase = weight "tap") stacked into the 256-row tile
#pragma unroll
                  for (int ph = 0; ph < 4; ++ph) tma_load_3d(dst + ph * 64 * 128, &p.bmap, bfull(stage), k, ph, n_tile * 64);
                } else if (kPair)
                  tma_load_3d_pair(dst, &p.bmap, bfull(stage), k, btap, b_row0 + rank * rows);
                else if (cs == 1)
                  tma_load_3d(dst, &p.bmap, bfull(stage), k, btap, b_row0);
                else
                  tma_load_3d_mc(dst + rank * rows * 128, &p.bmap, bfull(stage), k, btap, b_row0 + rank * rows, mc_mask);
                if (++stage == p.b_stages) {
                  stage = 0;
                  phase ^= 1u;
                }
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    // The issue loop must stay ahead of the tensor core (64 cycles per MMA at BN = 128): descriptors are kept as
    // (lo, hi) words, hi is loop invariant, lo advances by plain 32-bit adds (tap row offset, K step of 32 bytes).
    if (leader && elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128 * CG, BN, 0, 0);
      const uint32_t a_hi = ((static_cast<uint32_t>(p.halo_w * 128) >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
      const uint32_t b_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
      uint32_t tap_lo[9];
#pragma unroll
      for (int t = 0; t < 9; ++t) tap_lo[t] = static_cast<uint32_t>(p.tap_row[t]) * 8u;  // rows of 128 B, in 16-B units
      const int ntaps = p.num_taps;
      int astage = 0, bstage = 0, acc = 0;
      uint32_t aphase = 0, bphase = 0, acc_phase = 0;
      if (!ring) {
        mbar_wait(bfull(0), 0);
        tc_fence_after();
      }
      for (int g = cidx; g < ngroups; g += nclusters) {
        if (p.ph_ntaps) {   // nearest-x2 conv: the tile's output phase selects its 2x2 window of the halo
          int n_tile, m_unused;
          f2_item(p, g * cs, &n_tile, &m_unused);
          const int ph = n_tile / p.nt_per_phase;
#pragma unroll
          for (int t = 0; t < 4; ++t) tap_lo[t] = static_cast<uint32_t>(p.ph_row[ph][t]) * 8u;
        }
        if (!(kF2Knock & 256)) mbar_wait(tempty(acc), acc_phase ^ 1u);
        if (!(kF2Knock & 64)) tc_fence_after();
        if ((kF2Knock & 512) && !kPair) umma_commit(tfull(acc));
        const uint32_t d_tmem = tmem_base + acc * (MP * BN);
        uint32_t b_item = (b_base >> 4) | (1u << 16);   // stationary weights: walks through the resident tiles
        uint32_t fresh = 0;   // 0 for the first MMA of the tile (overwrite the accumulator), 1 afterwards
        for (int seg = 0; seg < p.num_segs; ++seg) {
        const uint32_t seg_mask = static_cast<uint32_t>(p.seg_taps[seg]);
        for (int cb = 0; cb < p.seg_cblocks[seg]; ++cb) {
          const int ksteps = (KSKIP && cb + 1 == p.seg_cblocks[seg]) ? p.seg_klast[seg] : 4;
          mbar_wait(afull(astage), aphase);
          if (!(kF2Knock & 64)) tc_fence_after();
          const uint32_t a_lo = ((a_base + astage * a_slot) >> 4) | (1u << 16);
          if ((kF2Knock & 1024) && !kPair) umma_commit(aempty(astage));
          if constexpr (FAST != 0) {
            // nine taps x KS K-steps of one channel block as one straight line (KS is a compile-time constant here)
            auto issue_block = [&](auto ks_tag) {
              constexpr int KS = decltype(ks_tag)::value;
#pragma unroll
              for (int tap = 0; tap < 9; ++tap) {
                const uint32_t a_tap = a_lo + tap_lo[tap];
#pragma unroll
                for (int k = 0; k < KS; ++k) {
                  if (kF2Knock & 2) continue;
                  if (kPair)
                    umma2_bf16_words(d_tmem, a_tap + 2u * k, a_hi, b_item + 2u * k, b_hi, idesc,
                                     (tap == 0 && k == 0) ? fresh : 1u);
                  else
                    umma_bf16_words(d_tmem, a_tap + 2u * k, a_hi, b_item + 2u * k, b_hi, idesc,
                                    (tap == 0 && k == 0) ? fresh : 1u);
                }
                b_item += kBItem >> 4;
              }
            };
            if constexpr (FAST < 5) {
              issue_block(std::integral_constant<int, FAST>{});
            } else {   // some segment ends in a partial block: ONE branch per channel block picks its unrolled form
              switch (ksteps) {
                case 1: issue_block(std::integral_constant<int, 1>{}); break;
                case 2: issue_block(std::integral_constant<int, 2>{}); break;
                case 3: issue_block(std::integral_constant<int, 3>{}); break;
                default: issue_block(std::integral_constant<int, 4>{}); break;
              }
            }
            fresh = 1u;
          } else {
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            if (tap < ntaps && ((seg_mask >> tap) & 1u)) {
              uint32_t b_lo;
              if (ring) {
                mbar_wait(bfull(bstage), bphase);
                tc_fence_after();
                b_lo = ((b_base + bstage * kBItem) >> 4) | (1u << 16);
              } else {
                b_lo = b_item;
                b_item += kBItem >> 4;
              }
#pragma unroll
              for (int sub = 0; sub < MP; ++sub) {      // the same weight tile against each pixel tile of the item
                const uint32_t a_tap = a_lo + sub * (a_half >> 4) + tap_lo[tap];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  // KSKIP instantiation only (some segment ends in a channel block of < 64 channels): stop after the
                  // block's last 16-channel step.  The common instantiation has NO test here - the issue loop of an
                  // N = 64 tile has 32 cycles per MMA, a predicate + branch per UTCHMMA cost 14 % on 64->64 @512^2,
                  // and any less regular loop shape made the compiler drop the uniform-register descriptors (+65 %).
                  if (KSKIP && k >= ksteps) break;
                  if (kF2Knock & 2) continue;
                  if (kPair)
                    umma2_bf16_words(d_tmem, a_tap + 2u * k, a_hi, b_lo + 2u * k, b_hi, idesc, k == 0 ? fresh : 1u);
                  else
                    umma_bf16_words(d_tmem + sub * BN, a_tap + 2u * k, a_hi, b_lo + 2u * k, b_hi, idesc,
                                    k == 0 ? fresh : 1u);
                }
              }
              fresh = 1u;
              if (ring) {
                if (kPair)
                  umma2_commit_mc(bempty(bstage));
                else if (cs == 1)
                  umma_commit(bempty(bstage));
                else
                  umma_commit_mc(bempty(bstage), mc_mask);
                if (++bstage == p.b_stages) {
                  bstage = 0;
                  bphase ^= 1u;
                }
              }
            }
          }
          }
          if (kF2Knock & (32 | 1024)) {
          } else if (kPair)
            umma2_commit_mc(aempty(astage));
          else
            umma_commit(aempty(astage));
          if (++astage == p.a_stages) {
            astage = 0;
            aphase ^= 1u;
          }
        }
        }
        if ((kF2Knock & 128) && !kPair) umma_commit(bars + 48u);
        if (kF2Knock & 512) {
        } else if (kPair)
          umma2_commit_mc(tfull(acc));
        else
          umma_commit(tfull(acc));
        if (++acc == kAcc) {
          acc = 0;
          acc_phase ^= 1u;
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ epilogue warps (4..11): TWO groups of four
    // Group g (warps 4+4g .. 7+4g, one per TMEM lane quarter) drains the CTA's tiles with local index parity g, so two
    // tiles are in flight: ncu showed the single 8-warp epilogue latency-bound (IPC 1.7 of 4, 17 % of the samples in the
    // 256-thread barrier) and the 64-channel 512^2 layers capped by it at half of the HBM roof.  Each group owns one
    // staging tile, one bias vector, one named barrier and one row of statistics partials.
    const int ew = warp - 4;           // 0..7
    const int grp = ew >> 2;           // epilogue group
    const int gw = ew & 3;             // warp inside the group
    const int q = warp & 3;            // TMEM lane quarter of this warp (== gw)
    const int gt = threadIdx.x - 128 - grp * 128;   // 0..127 inside the group
    const int row = q * 32 + lane;     // accumulator row = pixel (y = row >> 3, x = row & 7) of the tile
    const bool do_stats = (p.flags & kFlagStats) != 0;
    const bool do_relu = (p.flags & kFlagRelu) != 0;
    const bool do_acc = (p.flags & kFlagAccumulate) != 0;
    const uint32_t sbuf = stg_base + grp * kF2StgBytes;
    const uint32_t bias_s = bars + 512u + grp * (BN * 4);  // BN floats (zero beyond ncols / without bias)
    constexpr int kChunks = BN / 64;
    float run_s[kChunks][2], run_q[kChunks][2];
#pragma unroll
    for (int c = 0; c < kChunks; ++c) run_s[c][0] = run_s[c][1] = run_q[c][0] = run_q[c][1] = 0.f;
    int run_row0 = -1;
    auto gbar = [&]() {
      if (grp == 0)
        asm volatile("bar.sync 1, 128;" ::: "memory");
      else
        asm volatile("bar.sync 2, 128;" ::: "memory");
    };
    // column-sum reads: this warp sums rows [gw*32, gw*32+32) of the staged chunk, lane owns columns 2*lane, 2*lane+1
    uint32_t st_off[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) st_off[j] = ((((lane >> 2) ^ j) << 4) + (lane & 3) * 4) + (gw * 32 + j) * 128;
    // One row of partials per (CTA, group): the four warps (32 tile rows each, all columns) are summed through the
    // group's (idle) staging tile before the flush - rare (n-tile change / end of kernel), deterministic order.
    auto flush_stats = [&]() {   // called by all 128 threads of the group
      if (run_row0 < 0) return;
      if (gt == 0) tma_store_wait_read0();   // a TMA store may still be reading the staging tile
      gbar();
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int col = c * 64 + 2 * lane + h;   // each warp holds all columns of its 32 tile rows
          st_shared_f32(sbuf + ((gw * 2 + 0) * BN + col) * 4, run_s[c][h]);
          st_shared_f32(sbuf + ((gw * 2 + 1) * BN + col) * 4, run_q[c][h]);
          run_s[c][h] = run_q[c][h] = 0.f;
        }
      }
      gbar();
      float* dst = p.stats + (static_cast<size_t>(blockIdx.x) * 2 + grp) * 2 * p.ncols;
      for (int i = gt; i < 2 * BN; i += 128) {
        const int which = i / BN, c = i - which * BN;
        const int col = run_row0 + c;
        if (col < p.ncols) {
          float v = 0.f;
#pragma unroll
          for (int w4 = 0; w4 < 4; ++w4) v += __uint_as_float(ld_shared_u32(sbuf + ((w4 * 2 + which) * BN + c) * 4));
          dst[which * p.ncols + col] += v;
        }
      }
      gbar();   // the staging tile is free for the next chunk
    };
    int lt = 0;   // local tile counter of this CTA (both groups count every tile)
    for (int g = cidx; g < ngroups; g += nclusters, ++lt) {
      if ((lt & 1) != grp) continue;
      const int acc = lt % kAcc;
      const uint32_t acc_phase = static_cast<uint32_t>(lt / kAcc) & 1u;
      int item_ntile = 0, m_unused;
      if (p.num_n_tiles > 1) f2_item(p, g * cs + rank, &item_ntile, &m_unused);
      const int o_phase = p.nt_per_phase ? item_ntile / p.nt_per_phase : 0;
      const int row0 = p.ph_major ? item_ntile * 64                                          // first output column of the tile
                                  : (p.nt_per_phase ? item_ntile % p.nt_per_phase : item_ntile) * BN;
      if (row0 != run_row0) {   // statistics and bias belong to the column range (not to the output phase)
        if (do_stats) flush_stats();
        run_row0 = row0;
        gbar();
        for (int c = gt; c < BN; c += 128) {
          const int col = row0 + (p.ph_major ? (c & 63) : c);   // phase-major: the same 64 biases for every phase
          st_shared_f32(bias_s + c * 4, (p.bias != nullptr && col < p.ncols) ? __ldg(p.bias + col) : 0.f);
        }
        gbar();
      }
      mbar_wait(tfull(acc), acc_phase);
      tc_fence_after();
#pragma unroll 1
      for (int sub = 0; sub < ((kF2Knock & 4) ? 0 : MP); ++sub) {
      // only the thread that issues the TMA store needs the tile coordinates when every tile is full
      F2Tile t;
      t.n = t.x0 = t.y0 = 0;
      bool tile_full = true, valid = true;
      if (p.ragged || gt == 0) {
        t = f2_decode(p, g * cs + rank, sub);
        tile_full = !t.dummy && (t.x0 + 8 <= p.W) && (t.y0 + 16 <= p.H);
        valid = !t.dummy && (t.x0 + (row & 7) < p.W) && (t.y0 + (row >> 3) < p.H);
      }
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * (MP * BN) + sub * BN;
#pragma unroll 1
      for (int ch = 0; ch < kChunks; ++ch) {
        const int col0 = p.ph_major ? row0 : row0 + ch * 64;   // phase-major: chunk ch = phase ch of the same column slice
        const int c_phase = p.ph_major ? ch : o_phase;
        if (col0 >= p.ncols) break;  // uniform over the CTA
        // 32-channel outputs (nested_unet level 0): the chunk's upper 32 columns are padding that the TMA store clips -
        // skip their TMEM load, conversion and staging (a third of this warp's instructions per tile)
        const bool upper = col0 + 32 < p.ncols;
        uint32_t v[2][32];
        tmem_ld32(t_addr + ch * 64, v[0]);
        if (upper) tmem_ld32(t_addr + ch * 64 + 32, v[1]);
        if (gt == 0) tma_store_wait_read0();  // the group's staging tile is free again
        gbar();                               // (also: every thread has finished the column sums of the previous chunk)
        tmem_ld_wait();
#pragma unroll
        for (int hcol = 0; hcol < 2; ++hcol) {
          if (hcol == 1 && !upper) break;
          float f[32];
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 b4 = ld_shared_f4(bias_s + (ch * 64 + hcol * 32 + j4 * 4) * 4);
            f[j4 * 4 + 0] = __uint_as_float(v[hcol][j4 * 4 + 0]) + b4.x;
            f[j4 * 4 + 1] = __uint_as_float(v[hcol][j4 * 4 + 1]) + b4.y;
            f[j4 * 4 + 2] = __uint_as_float(v[hcol][j4 * 4 + 2]) + b4.z;
            f[j4 * 4 + 3] = __uint_as_float(v[hcol][j4 * 4 + 3]) + b4.w;
          }
          if (do_relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
          }
          if (!tile_full && !valid) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = 0.f;  // keeps out-of-image rows out of the column sums
          }
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int chunk16 = hcol * 4 + c;  // 16-byte chunk index inside the 128-byte row
            st_shared_v4(sbuf + row * 128 + ((chunk16 ^ (row & 7)) << 4), pack_bf16x2(f[c * 8 + 0], f[c * 8 + 1]),
                         pack_bf16x2(f[c * 8 + 2], f[c * 8 + 3]), pack_bf16x2(f[c * 8 + 4], f[c * 8 + 5]),
                         pack_bf16x2(f[c * 8 + 6], f[c * 8 + 7]));
          }
        }
        fence_proxy_async_smem();
        gbar();
        if (gt == 0 && !(kF2Knock & 8)) {
          if (do_acc)
            tma_reduce_add_4d(&p.omap[c_phase], sbuf, col0, t.x0, t.y0, t.n);
          else
            tma_store_4d(&p.omap[c_phase], sbuf, col0, t.x0, t.y0, t.n);
          tma_store_commit();
        }
        if (do_stats && !(kF2Knock & 16)) {
          float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
          for (int rr = 0; rr < 32; ++rr) {
            const uint32_t w = ld_shared_u32(sbuf + st_off[rr & 7] + (rr >> 3) * 1024);
            const float lo = bf16_lo(w), hi = bf16_hi(w);
            s0 += lo;
            s1 += hi;
            q0 = fmaf(lo, lo, q0);
            q1 = fmaf(hi, hi, q1);
          }
          run_s[ch][0] += s0;
          run_s[ch][1] += s1;
          run_q[ch][0] += q0;
          run_q[ch][1] += q1;
        }
      }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (kPair)
          mbar_arrive_cluster(tempty(acc), 0);   // the accumulator pair is free when BOTH CTAs have drained theirs
        else
          mbar_arrive(tempty(acc));
      }
    }
    if (do_stats) flush_stats();
    if (gt == 0) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (cs > 1) cluster_sync_all();  // nobody leaves while a peer may still multicast into / signal this CTA
  if (warp == 1) {
    tc_fence_after();
    if (kPair)
      tmem_dealloc2(tmem_base, kTmemCols);
    else
      tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <int BN, int MP, int CG, bool KSKIP, int FAST>
static cudaError_t launch_fwd2_fast(const Igemm2Params& p, int grid, int smem_bytes, cudaStream_t stream) {
  static int configured[kMaxDevices] = {0};   // the attribute is per device
  int& conf = configured[device_slot()];
  if (conf < smem_bytes) {
    cudaError_t e =
        cudaFuncSetAttribute(igemm_fwd2_kernel<BN, MP, CG, KSKIP, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    conf = smem_bytes;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kF2Threads);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  int nattr = 0;
  if (p.cluster > 1) {
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = p.cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    nattr = 1;
  }
  cfg.attrs = attr;
  cfg.numAttrs = nattr;
  return cudaLaunchKernelEx(&cfg, igemm_fwd2_kernel<BN, MP, CG, KSKIP, FAST>, p);
}

template <int BN, int MP, int CG, bool KSKIP>
static cudaError_t launch_fwd2_ks(const Igemm2Params& p, int grid, int smem_bytes, cudaStream_t stream) {
  if constexpr (MP == 1 && BN <= 128) {
    static const bool off = getenv("UZ_FWD2_NO_FAST") != nullptr && getenv("UZ_FWD2_NO_FAST")[0] == '1';
    bool fast = !off && p.b_stages == 0 && p.num_taps == 9 && p.ph_ntaps == 0;
    for (int sg = 0; sg < p.num_segs; ++sg) fast = fast && p.seg_taps[sg] == 0x1ff;
    if (fast) {
      if constexpr (!KSKIP) {
        return launch_fwd2_fast<BN, MP, CG, KSKIP, 4>(p, grid, smem_bytes, stream);
      } else {
        int ks = p.seg_klast[0];     // static K-step count: every segment is ONE channel block of the same size
        for (int sg = 0; sg < p.num_segs; ++sg)
          if (p.seg_cblocks[sg] != 1 || p.seg_klast[sg] != ks) ks = 0;
        switch (ks) {
          case 1: return launch_fwd2_fast<BN, MP, CG, KSKIP, 1>(p, grid, smem_bytes, stream);
          case 2: return launch_fwd2_fast<BN, MP, CG, KSKIP, 2>(p, grid, smem_bytes, stream);
          case 3: return launch_fwd2_fast<BN, MP, CG, KSKIP, 3>(p, grid, smem_bytes, stream);
          default: return launch_fwd2_fast<BN, MP, CG, KSKIP, 5>(p, grid, smem_bytes, stream);
        }
      }
    }
  }
  return launch_fwd2_fast<BN, MP, CG, KSKIP, 0>(p, grid, smem_bytes, stream);
}

template <int BN, int MP, int CG>
static cudaError_t launch_fwd2_bn(const Igemm2Params& p, int grid, int smem_bytes, cudaStream_t stream) {
  bool kskip = false;   // any segment whose last channel block holds fewer than 64 channels
  for (int s = 0; s < p.num_segs; ++s) kskip = kskip || p.seg_klast[s] < 4;
  return kskip ? launch_fwd2_ks<BN, MP, CG, true>(p, grid, smem_bytes, stream)
               : launch_fwd2_ks<BN, MP, CG, false>(p, grid, smem_bytes, stream);
}

cudaError_t launch_igemm_fwd2(const Igemm2Params& p, int bn, int grid, int smem_bytes, cudaStream_t stream) {
  switch (bn) {
    case 64:
      return p.pair ? launch_fwd2_bn<64, 1, 2>(p, grid, smem_bytes, stream)
                    : launch_fwd2_bn<64, 1, 1>(p, grid, smem_bytes, stream);
    case 128:
      return p.pair      ? launch_fwd2_bn<128, 1, 2>(p, grid, smem_bytes, stream)
             : p.mp == 2 ? launch_fwd2_bn<128, 2, 1>(p, grid, smem_bytes, stream)
                         : launch_fwd2_bn<128, 1, 1>(p, grid, smem_bytes, stream);
    case 256:
      return p.pair ? launch_fwd2_bn<256, 1, 2>(p, grid, smem_bytes, stream)
                    : launch_fwd2_bn<256, 1, 1>(p, grid, smem_bytes, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace uz
