// 3x3 conv forward / dgrad with shared-memory halo re-use.
//
// A CTA works on a super-tile of 16 rows x 8*J columns of one image (J = 1, 2 or 4 accumulators of 128 pixels each
// in TMEM). For every 64- (or 32-) channel block it loads ONE halo tile of 18 x (8J+2) pixels with a single TMA box
// and forms the A operand of all nine taps and all J pixel tiles by pointing UMMA descriptors into that tile:
// tap (r,s) of pixel tile j starts at row r*(8J+2) + s + 8j of the halo tile, eight consecutive pixels of an image
// row are eight consecutive smem rows, and the stride between image rows (8J+2 smem rows) is the descriptor's
// stride-byte-offset. UMMA swizzles on absolute address bits (csrc/umma_probe.cu), so row-shifted starts and a
// stride that is not a multiple of the swizzle period read exactly the intended rows.
// Compared with igemm.cu (one TMA box per tap per 128 pixels) this cuts the L2->smem traffic of the activations by
// ~9x and issues 9*J*kc/16 MMAs per barrier round trip instead of kc/16, which is what the 32/64-channel layers at
// 128x128 and 64x64 need: they are HBM-bound, not tensor-bound, and were issue/L2-bound before.
#pragma once
#include <cuda.h>

#include "igemm.cuh"

namespace ub {

struct HaloParams {
  int N, H, W;
  int J;                  // 8-wide column strips per super-tile
  int tiles_h, tiles_w;   // super-tiles per image
  int n_tiles, block_n;   // GEMM-N tiling
  int num_src;
  int src_kblocks[2];
  int src_ci_off[2];
  int kc;                 // 64 (SWIZZLE_128B) or 32 (SWIZZLE_64B)
  int tpb;                // taps per weight stage: 9, 3 or 1
  int b_stages;
  int h_stages;           // halo ring depth (2..4)
  int timing;             // debug: accumulate per-role wait cycles (UNETB200_HALO_TIMING)
  int acc_double;         // two accumulator sets: epilogue overlaps the next super-tile's MMAs
  int fast_epi;           // set by halo_launch: 0 generic epilogue, 1 forward, 2 dgrad, 3 dgrad split at block_n/2,
                          // 4 = 3 with destination 0 summed 2x2 onto the coarse grid
  int rp;                 // row-pair variant (conv_halo_rp.cu): 32-row super-tiles, accumulators 2 * block_n wide
  IgemmEpilogue epi;      // d2s_cout must be 0
};

// Can the halo kernel run this shape, and with which parameters? Returns false to fall back to igemm.
bool halo_plan(HaloParams& p, int N, int H, int W, int kc, int ncols);

int halo_launch(const HaloParams& p, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                cudaStream_t stream);

// Row-pair variant for GEMM N = 32 / 64 (conv_halo_rp.cu): plan, which lean epilogue (0 = not covered) and launch.
// The activation tensor map must have a box of 34 rows x (8J+2) columns.
bool halo_rp_plan(HaloParams& p, int N, int H, int W, int kc, int ncols);
int halo_rp_epilogue_code(const HaloParams& p);
int halo_rp_launch(const HaloParams& p, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                   cudaStream_t stream);

}  // namespace ub
