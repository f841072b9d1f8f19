// Shared declarations for the two tcgen05 implicit-GEMM kernels (forward-type and weight-gradient-type).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace uz {

constexpr int kMaxAMaps = 8;    // A-operand tensor maps: (segment, lattice) pairs
constexpr int kMaxSegs = 6;     // channel segments of a virtual concat
constexpr int kMaxTaps = 36;    // filter taps per phase (36 = 4 phases x 9 for the fused up2x dgrad)
constexpr int kMaxPhases = 4;   // output sub-lattices (2x2 scatter for transposed conv / nearest-up2x conv)

struct TapTable {
  int8_t dy[kMaxTaps];    // A-tile row shift (in A-lattice pixels)
  int8_t dx[kMaxTaps];    // A-tile column shift
  int8_t btap[kMaxTaps];  // tap coordinate in the packed weight tensor
  int8_t lat[kMaxTaps];   // which lattice map of the segment this tap reads
};

enum IgemmFlags : int {
  kFlagRelu = 1,
  kFlagAccumulate = 2,  // out += result (bf16 read-modify-write)
  kFlagStats = 4,       // per-channel sum / sum-of-squares of the stored (bf16-rounded) values
};

// Forward-type implicit GEMM:  out[pixel, col] = sum_{seg,tap,c} A_seg[pixel + shift(tap), c] * Wp[col, tap, k0(seg)+c]
struct alignas(64) IgemmParams {
  CUtensorMap amap[kMaxAMaps];  // 4-D (C, W, H, N) bf16 views, box (64, tw, th, tn), SWIZZLE_128B
  CUtensorMap bmap;             // 3-D (Ctot, taps, cols) bf16 packed weights, box (64, 1, BN)
  TapTable taps[kMaxPhases];
  int8_t phase_oy[kMaxPhases];
  int8_t phase_ox[kMaxPhases];
  int num_segs;
  int seg_cblocks[kMaxSegs];  // ceil(C_seg / 64)
  int seg_k0[kMaxSegs];       // first weight-channel index of the segment
  int num_lat;
  int num_taps;
  int num_phases;
  int N, H, W;                 // extent of the M index space (A-lattice pixels)
  int tw_log2, th_log2;        // M tile = tn x th x tw pixels = 128
  int xb, yb, nb;              // tile counts along x, y, n
  int num_n_tiles;
  int total_tiles;             // xb*yb*nb*num_n_tiles*num_phases
  // output: pixel (n, y*oscale + phase_oy, x*oscale + phase_ox) in an (N, Ho, Wo, out_ld) bf16 tensor
  __nv_bfloat16* out;
  int Ho, Wo, out_ld, oscale;
  int ncols;                   // valid output columns (Cout)
  const float* bias;           // optional, per column
  float* stats;                // optional [gridDim.x][2][ncols] partial sums (kFlagStats)
  int flags;
};

// Weight-gradient-type GEMM:  D[t][i, j] = sum_{phase, pixel} P[phase][pixel, i] * Q[pixel + shift(phase, t), j]
struct alignas(64) WgradParams {
  CUtensorMap pmap[kMaxPhases];  // un-shifted operand, one lattice view per phase; box (64, tw, th, tn)
  CUtensorMap qmap[kMaxPhases];  // shifted operand, one view per lattice
  TapTable taps[kMaxPhases];
  int num_phases, num_taps;
  int N, H, W;                   // pixel index space of the reduction
  int tw_log2, th_log2;          // K chunk = tn x th x tw pixels = 64
  int xb, yb, nb;
  int chunks_per_phase;          // xb*yb*nb
  int cp, cq;                    // channels of P (rows) and Q (cols)
  int m_tiles;                   // ceil(cp / 128)
  int q_cblocks;                 // ceil(cq / 64)
  int num_atoms;                 // num_taps * q_cblocks
  int n_slabs;                   // ceil(num_atoms / 4)
  int splits;                    // split-K factor
  int total_items;               // m_tiles * n_slabs * splits
  float* partial;                // [splits][cp][num_taps][cq] fp32
};

// Second-generation 3x3 weight gradient (igemm_wgrad2.cu): one X halo box per 8x16-pixel chunk feeds all nine taps.
struct alignas(64) Wgrad2Params {
  CUtensorMap xmap;    // X:  4-D (C, W, H, N), box (64, 10, 18, 1), SWIZZLE_128B, zero fill outside the image
  CUtensorMap dzmap;   // dZ: 4-D (C, W, H, N), box (64, 8, 16, 1)
  int xb, yb;          // chunks along x (W/8) and y (H/16), rounded up
  int total_chunks;    // N * yb * xb
  int cp, cq;          // output channels (dZ) and input channels (X)
  int co_blocks, ci_blocks;  // ceil(cp/64), ceil(cq/64)
  int splits;          // split-K factor over the chunk range
  int total_items;     // co_blocks * ci_blocks * splits
  float* partial;      // tile-major fp32: [splits][co_blocks][ci_blocks][9][64 or 128 (wide)][64]
  // CTA-pair mode (cta_group::2): the two CTAs of a cluster take two adjacent input-channel blocks of the same
  // (output-channel block, chunk range); each loads its own X halo and HALF of the dZ chunk (32 channels), one
  // M = 256, N = 64 MMA serves both.  Per CTA and MMA 5 KB instead of 6 KB leave shared memory (160 vs 192 B/clk).
  CUtensorMap dzmap_half;  // dZ: box (32, 8, 16, 1), SWIZZLE_64B
  int pair;                // 0 / 1
  int ci_units;            // ci_blocks (single CTA) or ceil(ci_blocks / 2) (pair)
  // WIDE mode: output-channel blocks of 128 (co_blocks = ceil(cp/128)), N = 128 MMAs, two passes over the chunk range
  // (tap pairs 0..2, then 3..4) because five 128-column accumulators exceed the 512 TMEM columns.
  int wide;                // 0 / 1
  // Nearest-x2 conv weight gradient in PHASE form (see host_conv.cu, UZ_UP2X_CONV3X3): the pixel space is the
  // LOW-resolution image; per 8x16 chunk one X halo box and FOUR dZ boxes (the four output phases = strided views of the
  // high-resolution gradient); phase (py, px) pairs its dZ with the 2x2 window of the halo at (a + py, b + px) - two tap
  // pairs per phase, eight 64-column accumulators = all 512 TMEM columns in ONE pass.  Partials carry 16 "taps"
  // (phase * 4 + a * 2 + b); the reduce kernel folds them onto the 3x3 filter.
  int up2x;                // 0 / 1 (narrow tiles only)
  CUtensorMap dzmap_ph[4];       // dZ phase views, box (64, 8, 16, 1)
  CUtensorMap dzmap_half_ph[4];  // pair mode: box (32, 8, 16, 1), SWIZZLE_64B
};

// Second-generation forward-type implicit GEMM (igemm_fwd2.cu): conv3x3 / its data gradient / conv1x1 on 8x16-pixel
// tiles.  One TMA halo box per (segment, 64-channel block) feeds every filter tap; the weight operand is either kept
// resident in shared memory for the whole launch (small layers) or streamed through a ring, optionally multicast
// across a thread-block cluster; the epilogue goes TMEM -> registers -> swizzled smem -> TMA store.
struct alignas(64) Igemm2Params {
  CUtensorMap amap[kMaxSegs];  // 4-D (C, W, H, N) bf16, box (64, 8+2*pad, 16+2*pad, 1), SWIZZLE_128B
  CUtensorMap bmap;            // 3-D (Ctot, taps, cols) packed weights, box (64, 1, BN / cluster)
  CUtensorMap omap[4];         // 4-D (cout, W, H, N) output view(s), box (64, 8, 16, 1), SWIZZLE_128B (TMA store);
                               // transposed conv: one strided view per output phase (dy, dx) of the 2x2 scatter
  int num_segs;
  int seg_cblocks[kMaxSegs];
  int seg_k0[kMaxSegs];
  int num_taps;
  int seg_klast[kMaxSegs];     // K steps (of 16 channels) in the LAST 64-channel block of the segment: 1..4.  A 32-channel
                               // segment (nested_unet level 0) would otherwise spend half of its MMAs on zero padding
  int seg_taps[kMaxSegs];      // bit t set: segment uses tap t (all taps except for the stride-2 3x3 convolution, whose
                               // four segments are the four parity lattices of ONE input, each with its own taps)
  int b_items;                 // weight tiles per output tile = sum over segments of cblocks x taps used
  int tap_row[9];              // halo row (pixel index in the box) of each tap's first pixel
  int tap_b[9];                // tap coordinate in the packed weights
  int8_t seg_tap_b[kMaxSegs][9];  // tap coordinate in the packed weights per (segment, halo position).  Equal to tap_b
                               // for every mode but the nearest-x2 conv's data gradient, whose four segments (the output
                               // phases of dZ) each use four of the nine halo positions with their own pre-summed weights
  int ph_ntaps;                // 0, or 4: nearest-x2 conv forward - N tile = (output phase, column tile) like the transposed
                               // conv, and each phase reads a 2x2 window of the LOW-resolution halo (the nine taps of the
                               // 3x3 filter on the upsampled image collapse onto it; weights pre-summed per phase):
  int8_t ph_row[4][4];         //   halo row (pixel index in the box) of tap t of phase p; weight tap coordinate p*4 + t
  int pad;                     // 1 for 3x3, 0 for 1x1
  int halo_w;                  // 8 + 2*pad
  int halo_bytes;              // halo_w * (16 + 2*pad) * 128
  int N, H, W;
  int xb, yb;                  // tiles along x, y
  int m_tiles;                 // work items along M: ceil(N*yb*xb / mp) rounded up to a multiple of the cluster size
  int m_tiles_real;
  int num_n_tiles;
  int total_tiles;             // m_tiles * num_n_tiles (tile = n_tile * m_tiles + m)
  int kblocks;                 // sum of seg_cblocks
  int a_stages, b_stages;      // ring depths; b_stages == 0 -> weights stationary (all kblocks*taps items resident)
  int a_off, b_off, stg_off, bar_off;  // byte offsets from the 1024-aligned smem base
  int stg_bufs;                // staging tiles for the TMA-store epilogue (1 or 2)
  int cluster;                 // thread-block cluster size (1, 2 or 4): weight tiles are multicast when > 1
  int pair;                    // 1: the 2-cluster is a tcgen05 CTA pair (cta_group::2): ONE M = 256 MMA covers both CTAs'
                               // pixel tiles, each CTA holds half of the weight rows (halves the B bytes read from shared
                               // memory per MMA: N = 64 / 128 tiles are shared-memory-bandwidth bound otherwise)
  int mp;                      // M tiles per work item (1 or 2).  2: every streamed weight tile feeds TWO pixel tiles (two
                               // TMEM accumulators), halving the weight bytes an SM has to ingest per MMA cycle - the
                               // streamed-weight layers were capped at ~72 % tensor-active by SM ingest (~50 B/clk)
  int nt_per_phase;            // 0: N tiles walk the output columns.  > 0 (ConvTranspose2d k2s2 as four 1x1 GEMMs): N tile
                               // = (phase, column tile); the phase selects the weight "tap" coordinate and the output view
  int ph_major;                // 1 (ConvTranspose2d with cout a multiple of 64, BN = 256): ONE item covers all four output
                               // phases of a 64-column slice - N tile = column slice, accumulator columns [64 ph, 64 ph + 64)
                               // belong to phase ph: the input tile is loaded once instead of four times, the MMAs are 256 wide,
                               // each epilogue chunk goes to its phase's output view.  nt_per_phase is 0 then
  int b_batched;               // 1: batched GEMM (uz_bmm_tn): the "weight" operand differs per image - the B tensor map's tap
                               // coordinate is the image index of the tile (ring mode, no CTA pair, no stationary weights)
  int ragged;                  // 1: some tile is partial (H % 16, W % 8) or a dummy (M tiles rounded up to the work-item
                               // size).  0: every tile is full - the epilogue threads then skip the tile decode (five
                               // integer divisions per tile; ncu showed the epilogue of the 64-channel 512^2 layers
                               // instruction-bound with ~30 % of its instructions in division sequences)
  int n_minor;                 // 1: consecutive work items walk the N tiles of ONE pixel tile (transposed conv: the four
                               // output phases re-read the same input tile - adjacent CTAs then share it through L2
                               // instead of streaming the input four times from HBM); 0: all pixel tiles of an N tile first
  int ncols;                   // valid output columns (Cout)
  const float* bias;
  float* stats;                // [8*gridDim.x][2][ncols] partial sums (kFlagStats), zeroed by the host
  int flags;
};

}  // namespace uz
