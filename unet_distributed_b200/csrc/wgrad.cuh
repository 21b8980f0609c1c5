// Weight-gradient tile engine: dW[tap][ci][co] = sum over pixels  dZ[pixel][co] * X[pixel + tap][ci].
//
// The reduction (GEMM K) dimension is the pixel dimension, which is the *slow* dimension of both NHWC operands, so
// both are fed to tcgen05.mma as MN-major tiles: a TMA box of KP pixels x 64 (or 32) channels lands in shared memory
// exactly in the canonical MN-major swizzled layout. The tap shift of X is a TMA coordinate shift with zero fill.
// One CTA owns one (tap, 128 output channels, <=128 input channels, pixel range) and accumulates in TMEM; pixel
// ranges (split-K) go to separate slices of the output that a second kernel sums in a fixed order (deterministic).
#pragma once
#include <cstdint>
#include <cuda.h>

namespace ub {

struct WgradParams {
  int N, H, W;              // pixel grid walked by the reduction (conv: activation grid; tconv: the coarse grid)
  int TN, TH, TW;           // pixel box per K block, TN*TH*TW == 64
  int blocks_n, blocks_h, blocks_w;
  int splits;               // pixel-range splits; blocks_per_split = ceil(total/splits)
  int taps_r, taps_s, pad;
  int z_mode;               // 0: dZ is a plain 4-D map (c,w,h,n); 1: 5-D tconv view (o,b,j,a,n*h), tap = (a,b)
  int x_shift;              // 1: X box is shifted by the tap (conv); 0: not shifted (tconv)
  int z_C;                  // channels of dZ (GEMM M extent)
  int z_atom;               // 64 (SWIZZLE_128B) or 32 (SWIZZLE_64B)
  int m_tiles;              // ceil(z_C / 128), or ceil(taps * z_C / 128) with tap_in_m
  int tap_in_m;             // z_mode 1 only: stack the taps along GEMM M (row = tap * z_C + channel), one CTA for all taps
  int z_pair;               // tap_in_m with z_C = 32 (bf16): one TMA box takes BOTH column taps of a filter row -- the two
                            // fine pixels are adjacent in memory, 2 x 64 B = one 128-byte swizzle row = a 64-wide M atom
  int num_src;              // X sources (concat inputs)
  int x_C[2];               // channels per source
  int x_atom[2];            // 64 or 32
  int x_block_n[2];         // GEMM N per tile for that source: min(x_C, 128)
  int x_ntiles[2];
  int x_col_off[2];         // column offset of the source in the output (channel offset in the concat)
  int stages;
  int elt;                  // element bytes of dZ and X: 2 = bf16 (0 means 2), 4 = fp32 read as tf32 (atoms of 32 channels)
  float* out;
  long long out_split_stride, out_tap_stride, out_m_stride, out_n_stride;
  // optional (bf16 only): bias_out[(split * taps + tap) * z_C + co] = sum of dZ[.., co] over the pixels this (split, tap)
  // walks -- the CTAs of the first X tile sum their dZ tiles in shared memory while the tensor core works on them
  float* bias_out;
};
constexpr int kWgradBiasScratch = 128 * 8 * 4;   // bytes of shared memory for the column sums (128 threads x 8 floats)

int wgrad_launch(const WgradParams& p, const CUtensorMap& z, const CUtensorMap& x0, const CUtensorMap& x1,
                 cudaStream_t stream);

// Halo variant for 32/64-channel inputs and gradients (wgrad_halo.cu): all nine taps per CTA, one halo tile per
// super-tile, filter columns stacked along GEMM M. One launch per source of a concatenated input.
bool wgrad_halo_supported(int H, int W, int Cx, int Cz);
int wgrad_halo_pick_J(int Cx, int W);
size_t wgrad_halo_ws_bytes(int N, int H, int W, int Cx, int Cz);
int wgrad_halo_launch(const CUtensorMap& x_map, const CUtensorMap& z_map, int N, int H, int W, int Cx, int Cz, int J,
                      float* dw, int cin_total, int ci_off, float* ws, cudaStream_t stream, float* db = nullptr);

// Row variant for wide layers (wgrad_row.cu): one CTA = one filter row (three taps) of a 128 x 128 tile.
struct WgradRowParams {
  int N, H, W;
  int blocks_h, blocks_w;   // 8 x 8 pixel blocks per image
  int splits;
  int z_C, m_tiles;
  int num_src;
  int x_block_n[2], x_ntiles[2], x_col_off[2];
  int stages;
  float* out;
  long long out_split_stride, out_tap_stride, out_n_stride;   // output channel (M) stride is 1
  float* bias_out;          // optional: bias_out[split * bias_split_stride + co] = sum of dZ over the split's pixels
  long long bias_split_stride;
};
bool wgrad_row_supported(int H, int W, int Cz, int C0, int C1);
int wgrad_row_plan(WgradRowParams& p, int N, int H, int W, int Cz, int C0, int C1);
int wgrad_row_launch(const WgradRowParams& p, const CUtensorMap& z, const CUtensorMap& x0, const CUtensorMap& x1,
                     cudaStream_t stream);

// out[i] = sum_s ws[s*stride + i] (fixed order), i < n; outputs [n_a, n) go to out_b[i - n_a] when out_b is given.
int splitk_reduce_launch(const float* ws, long long split_stride, int splits, float* out, long long n,
                         cudaStream_t stream, float* out_b = nullptr, long long n_a = 0);

}  // namespace ub
