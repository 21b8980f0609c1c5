#include "ops.cuh"

#include <climits>
#include <cstdlib>
#include <cstring>

#include "common.h"
#include "conv_halo.cuh"
#include "elementwise.cuh"
#include "igemm.cuh"
#include "wgrad.cuh"

namespace ub {


namespace {

struct Box {
  int TN, TH, TW;
};

// 128-pixel output tile (igemm) or 64-pixel reduction block (wgrad) for an NHWC grid.
// Small grids pack several images into one box. `image_aligned` (needed by the 5-D transposed-conv view, whose
// last dimension is the merged n*H+i index) forbids boxes that straddle images unless they cover whole images.
Box pick_box(int pixels, int H, int W, bool image_aligned = false) {
  Box b;
  b.TW = W > 8 ? 16 : 8;
  const int max_th = pixels / b.TW;
  if (H >= max_th) {
    b.TH = max_th;
    b.TN = 1;
  } else {
    int th = 1;
    while (th < H) th <<= 1;
    b.TH = th;
    b.TN = max_th / th;
    if (image_aligned && b.TN > 1 && b.TH != H) {
      b.TW = 16;
      b.TH = pixels / 16;
      b.TN = 1;
    }
  }
  return b;
}

int cdiv(int a, int b) { return (a + b - 1) / b; }

int pick_kc(int C0, int C1, int elt = 2) {
  if (elt == 4) return (C0 % 32 == 0 && (C1 == 0 || C1 % 32 == 0)) ? 32 : 0;   // 32 fp32 = one 128-byte swizzle row
  if (C0 % 64 == 0 && (C1 == 0 || C1 % 64 == 0)) return 64;
  if (C0 % 32 == 0 && (C1 == 0 || C1 % 32 == 0)) return 32;
  return 0;
}

int nhwc_map(CUtensorMap* m, const void* p, int N, int H, int W, int C, int box_c, Box b, int elt = 2, int swizzle = 0) {
  const uint64_t e = static_cast<uint64_t>(elt);
  const uint64_t dims[4] = {static_cast<uint64_t>(C), static_cast<uint64_t>(W), static_cast<uint64_t>(H),
                            static_cast<uint64_t>(N)};
  const uint64_t str[3] = {static_cast<uint64_t>(C) * e, static_cast<uint64_t>(W) * C * e,
                           static_cast<uint64_t>(H) * W * C * e};
  const uint32_t box[4] = {static_cast<uint32_t>(box_c), static_cast<uint32_t>(b.TW), static_cast<uint32_t>(b.TH),
                           static_cast<uint32_t>(b.TN)};
  return make_tmap(m, p, 4, dims, str, box, swizzle ? swizzle : box_c * elt, elt);
}

// 5-D view of the fine tensor y[N, 2H, 2W, C] as (c, b, j, a, n*H+i) with y[n, 2i+a, 2j+b, c]; (H, W) = coarse grid.
int fine_view_map(CUtensorMap* m, const void* p, int N, int H, int W, int C, int box_c, Box b, int elt = 2,
                  int swizzle = 0) {
  const uint64_t e = static_cast<uint64_t>(elt);
  const uint64_t dims[5] = {static_cast<uint64_t>(C), 2, static_cast<uint64_t>(W), 2,
                            static_cast<uint64_t>(N) * H};
  const uint64_t str[4] = {static_cast<uint64_t>(C) * e, static_cast<uint64_t>(2) * C * e,
                           static_cast<uint64_t>(2) * W * C * e, static_cast<uint64_t>(4) * W * C * e};
  const uint32_t box[5] = {static_cast<uint32_t>(box_c), 1, static_cast<uint32_t>(b.TW), 1,
                           static_cast<uint32_t>(b.TN * b.TH)};
  return make_tmap(m, p, 5, dims, str, box, swizzle ? swizzle : box_c * elt, elt);
}

// The same fine tensor with both column taps of a coarse pixel in one row: [N*H][a][W][2*C] -- the fine pixels (2j, 2j+1)
// are adjacent in memory, so a coarse pixel's (b, channel) pair is 2*C contiguous elements. For C = 32 (bf16) that is one
// 128-byte swizzle row. (A 5-D box [C, 2, ...] does NOT give this: with the inner box narrower than the swizzle span the
// rows land 128 bytes apart in shared memory.) Coordinates: (0, j, a, n*H + h).
int fine_pair_map(CUtensorMap* m, const void* p, int N, int H, int W, int C, Box b, int elt = 2) {
  const uint64_t e = static_cast<uint64_t>(elt);
  const uint64_t dims[4] = {static_cast<uint64_t>(2 * C), static_cast<uint64_t>(W), 2, static_cast<uint64_t>(N) * H};
  const uint64_t str[3] = {static_cast<uint64_t>(2) * C * e, static_cast<uint64_t>(2) * W * C * e,
                           static_cast<uint64_t>(4) * W * C * e};
  const uint32_t box[4] = {static_cast<uint32_t>(2 * C), static_cast<uint32_t>(b.TW), 1,
                           static_cast<uint32_t>(b.TN * b.TH)};
  return make_tmap(m, p, 4, dims, str, box, 2 * C * static_cast<int>(e), elt);
}

// Packed weights [ncols][taps][ktot] bf16.
int weight_map(CUtensorMap* m, const void* p, int ncols, int taps, int ktot, int kc, int block_n, int elt = 2) {
  const uint64_t e = static_cast<uint64_t>(elt);
  const uint64_t dims[3] = {static_cast<uint64_t>(ktot), static_cast<uint64_t>(taps), static_cast<uint64_t>(ncols)};
  const uint64_t str[2] = {static_cast<uint64_t>(ktot) * e, static_cast<uint64_t>(taps) * ktot * e};
  const uint32_t box[3] = {static_cast<uint32_t>(kc), 1, static_cast<uint32_t>(block_n)};
  return make_tmap(m, p, 3, dims, str, box, kc * elt, elt);
}

void fill_grid(IgemmParams& p, int N, int H, int W, Box b) {
  p.N = N; p.H = H; p.W = W;
  p.TN = b.TN; p.TH = b.TH; p.TW = b.TW;
  p.tiles_n = cdiv(N, b.TN);
  p.tiles_h = cdiv(H, b.TH);
  p.tiles_w = cdiv(W, b.TW);
}

// Fewer tiles than half the SMs (the 8x8 level at bs128: 64 pixel tiles): halve a 256-wide N tile so that twice as
// many SMs work (ncu r02: conv5b forward / dgrad ran on 64 of 148 SMs at 25 % tensor-pipe activity).
int widen_grid(int block_n, int m_tiles, int ncols) {
  static const bool off = getenv("UNETB200_NO_WIDEN_GRID") != nullptr;
  if (off) return block_n;
  while (block_n > 128 && m_tiles * (ncols / block_n) * 2 <= device_sm_count()) block_n >>= 1;
  return block_n;
}

int pick_block_n(int ncols_per_dest) {
  if (ncols_per_dest % 256 == 0) return 256;
  if (ncols_per_dest % 128 == 0) return 128;
  if (ncols_per_dest % 64 == 0) return 64;
  if (ncols_per_dest % 32 == 0) return 32;
  return 0;
}

// Run a 3x3 conv (forward or dgrad, described by the generic params `p`) on the halo kernel when the shape allows.
// Returns true when it took the launch (status in *rc).
bool try_halo(const IgemmParams& p, const void* src0, int C0, const void* src1, int C1, const void* w_packed,
              int ncols, int ktot, cudaStream_t s, int* rc) {
  {   // row-pair variant first: twice the GEMM N per A-operand read on the 32/64-column layers
    HaloParams rp;
    std::memset(&rp, 0, sizeof(rp));
    if (halo_rp_plan(rp, p.N, p.H, p.W, p.kc, ncols)) {
      rp.num_src = p.num_src;
      for (int i = 0; i < 2; ++i) {
        rp.src_kblocks[i] = p.src_kblocks[i];
        rp.src_ci_off[i] = p.src_ci_off[i];
      }
      rp.epi = p.epi;
      rp.fast_epi = halo_rp_epilogue_code(rp);
      if (rp.fast_epi) {
        const Box hb = {1, 34, 8 * rp.J + 2};
        CUtensorMap a0, a1, bm;
        *rc = nhwc_map(&a0, src0, p.N, p.H, p.W, C0, p.kc, hb);
        if (*rc) return true;
        a1 = a0;
        if (C1) {
          *rc = nhwc_map(&a1, src1, p.N, p.H, p.W, C1, p.kc, hb);
          if (*rc) return true;
        }
        *rc = weight_map(&bm, w_packed, ncols, 9, ktot, p.kc, rp.block_n);
        if (*rc) return true;
        *rc = halo_rp_launch(rp, a0, a1, bm, s);
        return true;
      }
    }
  }
  if (p.epi.maxpool_out) return false;   // only the row-pair kernel fuses MaxPool2D (the caller reports the error)
  HaloParams hp;
  std::memset(&hp, 0, sizeof(hp));
  if (!halo_plan(hp, p.N, p.H, p.W, p.kc, ncols)) return false;
  hp.num_src = p.num_src;
  for (int i = 0; i < 2; ++i) {
    hp.src_kblocks[i] = p.src_kblocks[i];
    hp.src_ci_off[i] = p.src_ci_off[i];
  }
  hp.epi = p.epi;
  const Box hb = {1, 18, 8 * hp.J + 2};
  CUtensorMap a0, a1, bm;
  *rc = nhwc_map(&a0, src0, p.N, p.H, p.W, C0, p.kc, hb);
  if (*rc) return true;
  a1 = a0;
  if (C1) {
    *rc = nhwc_map(&a1, src1, p.N, p.H, p.W, C1, p.kc, hb);
    if (*rc) return true;
  }
  *rc = weight_map(&bm, w_packed, ncols, 9, ktot, p.kc, hp.block_n);
  if (*rc) return true;
  *rc = halo_launch(hp, a0, a1, bm, s);
  return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ conv3x3 forward
bool conv3x3_fwd_maxpool_ok(int N, int H, int W, int C0, int C1, int Cout) {
  static const bool off = getenv("UNETB200_NO_FUSED_MAXPOOL") != nullptr;
  const int kc = pick_kc(C0, C1);
  if (off || kc == 0 || H % 2 != 0 || W % 2 != 0) return false;
  HaloParams rp;
  std::memset(&rp, 0, sizeof(rp));
  return halo_rp_plan(rp, N, H, W, kc, Cout);
}

int conv3x3_fwd(ConvSrc x0, ConvSrc x1, const void* w_packed, const float* bias, void* y, int N, int H, int W, int Cout,
                int relu, DropoutCfg drop, cudaStream_t s, unsigned short* gate_out, void* maxpool_out, int dt) {
  const int C1 = x1.ptr ? x1.C : 0;
  const int elt = dt_bytes(dt);
  const int kc = pick_kc(x0.C, C1, elt);
  UB_REQUIRE(kc != 0, "conv3x3_fwd: input channels %d/%d must be multiples of 32", x0.C, C1);
  const int block_n = pick_block_n(Cout);
  UB_REQUIRE(block_n != 0, "conv3x3_fwd: Cout %d must be a multiple of 32", Cout);
  IgemmParams p;
  std::memset(&p, 0, sizeof(p));
  const Box b = pick_box(128, H, W);
  fill_grid(p, N, H, W, b);
  const int block_n_igemm = widen_grid(block_n, p.tiles_n * p.tiles_h * p.tiles_w, Cout);
  p.n_tiles = Cout / block_n;
  p.block_n = block_n;
  p.a_mode = IG_A_SHIFT4D;
  p.taps_r = 3; p.taps_s = 3; p.pad = 1;
  p.num_src = C1 ? 2 : 1;
  p.src_kblocks[0] = x0.C / kc;
  p.src_kblocks[1] = C1 / kc;
  p.src_ci_off[0] = 0;
  p.src_ci_off[1] = x0.C;
  p.kc = kc;
  p.elt = elt;
  p.stages = igemm_pick_stages(block_n, kc, elt);
  p.epi.dst[0] = y; p.epi.dst_C[0] = Cout; p.epi.dst_coff[0] = 0;
  p.epi.split_col = INT_MAX;
  p.epi.bias = bias;
  p.epi.relu = relu;
  p.epi.mask_scale = 1.0f;
  p.epi.drop_keep = drop.keep;
  p.epi.drop_seed = drop.seed;
  p.epi.drop_stream = drop.stream;
  p.epi.drop_seed_add = drop.seed_add;
  p.epi.drop_bits = drop.keep < 1.0f ? drop.bits : nullptr;
  p.epi.gate_out = gate_out;
  p.epi.maxpool_out = maxpool_out;
  int rc = UB_OK;
  // the halo kernels are bf16 engines; the fp32/tf32 mode runs every conv on the generic implicit-GEMM kernel
  if (elt == 2 && try_halo(p, x0.ptr, x0.C, x1.ptr, C1, w_packed, Cout, x0.C + C1, s, &rc)) return rc;
  UB_REQUIRE(!maxpool_out, "conv3x3_fwd: fused max pooling is not available for this shape (conv3x3_fwd_maxpool_ok)");
  if (block_n_igemm != block_n) {
    p.block_n = block_n_igemm;
    p.n_tiles = Cout / block_n_igemm;
    p.stages = igemm_pick_stages(block_n_igemm, kc, elt);
  }
  const int bn = p.block_n;
  CUtensorMap a0, a1, bm;
  rc = nhwc_map(&a0, x0.ptr, N, H, W, x0.C, kc, b, elt);
  if (rc) return rc;
  a1 = a0;
  if (C1) {
    rc = nhwc_map(&a1, x1.ptr, N, H, W, C1, kc, b, elt);
    if (rc) return rc;
  }
  rc = weight_map(&bm, w_packed, Cout, 9, x0.C + C1, kc, bn, elt);
  if (rc) return rc;
  return igemm_launch(p, a0, a1, bm, s);
}

// ------------------------------------------------------------------------------------------------ conv3x3 dgrad
bool conv3x3_dgrad_pool2_ok(int N, int H, int W, int Cout, int C0, int C1) {
  static const bool no_fast = getenv("UNETB200_NO_FAST_EPI") != nullptr || getenv("UNETB200_NO_POOL2") != nullptr;
  if (no_fast || C0 != C1 || C0 % 32 != 0) return false;
  const int kc = pick_kc(Cout, 0);
  if (kc == 0) return false;
  HaloParams hp;
  std::memset(&hp, 0, sizeof(hp));
  if (!halo_plan(hp, N, H, W, kc, C0 + C1)) return false;
  return hp.n_tiles == 1 && hp.block_n == C0 + C1 && hp.block_n >= 64 && H % 16 == 0 && W % (8 * hp.J) == 0;
}

int conv3x3_dgrad(const void* dz, int Cout, const void* w_packed_d, DgradDst d0, DgradDst d1, int N, int H, int W,
                  cudaStream_t s, int dt) {
  const int C1 = d1.ptr ? d1.C : 0;
  const int elt = dt_bytes(dt);
  const int kc = pick_kc(Cout, 0, elt);
  UB_REQUIRE(kc != 0, "conv3x3_dgrad: Cout %d must be a multiple of 32", Cout);
  int block_n = pick_block_n(d0.C);
  if (C1) {
    const int b1 = pick_block_n(C1);
    block_n = b1 < block_n ? b1 : block_n;
  }
  UB_REQUIRE(block_n != 0, "conv3x3_dgrad: destination channels %d/%d must be multiples of 32", d0.C, C1);
  IgemmParams p;
  std::memset(&p, 0, sizeof(p));
  const Box b = pick_box(128, H, W);
  fill_grid(p, N, H, W, b);
  p.n_tiles = (d0.C + C1) / block_n;
  p.block_n = block_n;
  p.a_mode = IG_A_SHIFT4D;
  p.taps_r = 3; p.taps_s = 3; p.pad = 1;
  p.num_src = 1;
  p.src_kblocks[0] = Cout / kc;
  p.kc = kc;
  p.elt = elt;
  p.stages = igemm_pick_stages(block_n, kc, elt);
  p.epi.dst[0] = d0.ptr; p.epi.dst_C[0] = d0.C; p.epi.aux[0] = d0.gate ? nullptr : d0.mask; p.epi.aux_C[0] = d0.C;
  p.epi.dst[1] = d1.ptr; p.epi.dst_C[1] = C1; p.epi.aux[1] = d1.gate ? nullptr : d1.mask; p.epi.aux_C[1] = C1;
  p.epi.gate_in[0] = d0.gate;
  p.epi.pool2_slot0 = d0.pool2;
  p.epi.gate_in[1] = C1 ? d1.gate : nullptr;
  p.epi.split_col = C1 ? d0.C : INT_MAX;
  p.epi.mask_scale = d0.scale;
  UB_REQUIRE(!C1 || !d0.mask || !d1.mask || d0.scale == d1.scale,
             "conv3x3_dgrad: both destinations must share the mask scale");
  if (!(d0.mask || d0.gate) && (d1.mask || d1.gate)) p.epi.mask_scale = d1.scale;
  p.epi.drop_keep = 1.0f;
  int rc = UB_OK;
  if (elt == 2 && try_halo(p, dz, Cout, nullptr, 0, w_packed_d, d0.C + C1, Cout, s, &rc)) return rc;
  UB_REQUIRE(!d0.pool2, "conv3x3_dgrad: fused 2x2 pooling is not available for this shape (conv3x3_dgrad_pool2_ok)");
  {
    const int bn2 = widen_grid(block_n, p.tiles_n * p.tiles_h * p.tiles_w, d0.C + C1);
    if (bn2 != block_n && (!C1 || d0.C % bn2 == 0)) {
      block_n = bn2;
      p.block_n = bn2;
      p.n_tiles = (d0.C + C1) / bn2;
      p.stages = igemm_pick_stages(bn2, kc, elt);
    }
  }
  CUtensorMap a0, bm;
  rc = nhwc_map(&a0, dz, N, H, W, Cout, kc, b, elt);
  if (rc) return rc;
  rc = weight_map(&bm, w_packed_d, d0.C + C1, 9, Cout, kc, block_n, elt);
  if (rc) return rc;
  return igemm_launch(p, a0, a0, bm, s);
}

// ------------------------------------------------------------------------------------------------ wgrad helpers
namespace {

struct WgradPlan {
  WgradParams p;
  long long nweights;
};

int plan_wgrad(WgradPlan& wp, int N, int H, int W, int taps_r, int taps_s, int pad, int z_C, int C0, int C1,
               int elt = 2) {
  WgradParams& p = wp.p;
  std::memset(&p, 0, sizeof(p));
  p.elt = elt;
  const Box b = pick_box(64, H, W, taps_r == 2);
  p.N = N; p.H = H; p.W = W;
  p.TN = b.TN; p.TH = b.TH; p.TW = b.TW;
  p.blocks_n = cdiv(N, b.TN);
  p.blocks_h = cdiv(H, b.TH);
  p.blocks_w = cdiv(W, b.TW);
  p.taps_r = taps_r; p.taps_s = taps_s; p.pad = pad;
  UB_REQUIRE(z_C % 32 == 0, "wgrad: gradient channels %d must be a multiple of 32", z_C);
  p.z_C = z_C;
  p.z_atom = (elt == 2 && z_C % 64 == 0) ? 64 : 32;   // atom = one swizzle row: <= 128 bytes of channels
  p.m_tiles = cdiv(z_C, 128);
  static const bool no_tap_in_m = getenv("UNETB200_NO_TAP_IN_M") != nullptr;
  // transposed conv with few channels (transConv9: z_C = 32): the four taps stacked along GEMM M fill one 128-row tile
  // and the X tile is read once (0.118 -> 0.091 ms); wider layers already fill M per tap and were slower stacked
  p.tap_in_m = (taps_r == 2 && taps_r * taps_s * z_C <= 128 && !no_tap_in_m) ? 1 : 0;
  if (p.tap_in_m) p.m_tiles = cdiv(taps_r * taps_s * z_C, 128);
  // ... and with 32 channels the two column taps of a filter row are one 128-byte row in memory: one TMA box of 128-byte
  // rows per filter row instead of two boxes of 64-byte pieces at a 128-byte stride
  static const bool no_z_pair = getenv("UNETB200_NO_Z_PAIR") != nullptr;
  p.z_pair = (p.tap_in_m && elt == 2 && z_C == 32 && taps_s == 2 && !no_z_pair) ? 1 : 0;
  if (p.z_pair) p.z_atom = 64;
  p.num_src = C1 ? 2 : 1;
  const int Cs[2] = {C0, C1};
  int off = 0;
  int total_nt = 0;
  for (int s = 0; s < p.num_src; ++s) {
    UB_REQUIRE(Cs[s] % 32 == 0, "wgrad: input channels %d must be a multiple of 32", Cs[s]);
    p.x_C[s] = Cs[s];
    p.x_atom[s] = (elt == 2 && Cs[s] % 64 == 0) ? 64 : 32;
    p.x_block_n[s] = Cs[s] >= 128 ? 128 : Cs[s];
    UB_REQUIRE(Cs[s] % p.x_block_n[s] == 0, "wgrad: input channels %d not tileable", Cs[s]);
    p.x_ntiles[s] = Cs[s] / p.x_block_n[s];
    p.x_col_off[s] = off;
    off += Cs[s];
    total_nt += p.x_ntiles[s];
  }
  const int total_pb = p.blocks_n * p.blocks_h * p.blocks_w;
  const int base = (p.tap_in_m ? 1 : taps_r * taps_s) * p.m_tiles * total_nt;
  static const int waves = getenv("UNETB200_WGRAD_WAVES") ? atoi(getenv("UNETB200_WGRAD_WAVES")) : 1;
  int splits = waves == 1 ? device_sm_count() / base : cdiv(waves * device_sm_count(), base);
  const int max_splits = total_pb / 8 > 1 ? total_pb / 8 : 1;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  const int per = cdiv(total_pb, splits);
  p.splits = cdiv(total_pb, per);
  p.stages = elt == 4 ? 3 : 6;
  wp.nweights = static_cast<long long>(taps_r) * taps_s * (C0 + C1) * z_C;
  return UB_OK;
}

}  // namespace

size_t conv3x3_wgrad_ws_bytes(int N, int H, int W, int Cin_total, int Cout, bool with_bias) {
  WgradPlan wp;
  // the split count does not depend on how Cin is divided between sources beyond the tile count; use the
  // single-source plan as an upper bound (two sources never have fewer tiles, hence never more splits).
  if (plan_wgrad(wp, N, H, W, 3, 3, 1, Cout, Cin_total, 0) != UB_OK) return 0;
  size_t generic = 0;
  if (wp.p.splits > 1) generic = static_cast<size_t>(wp.p.splits) * wp.nweights * sizeof(float);
  // the halo variant (32/64-channel sources) keeps one partial gradient per CTA; Cin_total may be one source or
  // the concat of two, so take the larger per-source case
  size_t halo = 0;
  const int cx = Cin_total >= 64 ? 64 : 32;
  if (Cin_total <= 128 && wgrad_halo_supported(H, W, cx, Cout)) halo = wgrad_halo_ws_bytes(N, H, W, cx, Cout);
  // the row variant (wide layers) splits the pixel range further because it has three times fewer CTAs per range
  size_t row = 0;
  if (wgrad_row_supported(H, W, Cout, Cin_total, 0)) {
    WgradRowParams rp;
    std::memset(&rp, 0, sizeof(rp));
    wgrad_row_plan(rp, N, H, W, Cout, Cin_total, 0);
    if (rp.splits > 1) row = static_cast<size_t>(rp.splits) * (wp.nweights + (with_bias ? Cout : 0)) * sizeof(float);
  }
  size_t need = generic > halo ? generic : halo;
  if (row > need) need = row;
  // the optional bias gradient of the non-halo paths reuses the workspace after the split-K reduction
  const size_t bias = with_bias ? static_cast<size_t>(colsum_ws_floats(Cout)) * sizeof(float) : 0;
  return need > bias ? need : bias;
}

// Bias gradient for the wgrad paths that do not produce it themselves.
static int bias_grad_after(const void* dz, int Cout, float* db, void* ws, size_t ws_bytes, int N, int H, int W,
                           cudaStream_t s, int dt = UB_DT_BF16) {
  const size_t need = static_cast<size_t>(colsum_ws_floats(Cout)) * sizeof(float);
  UB_REQUIRE(ws != nullptr && ws_bytes >= need, "conv3x3_wgrad: workspace of %zu bytes needed for the bias gradient, %zu given",
             need, ws_bytes);
  return colsum_act(dz, db, reinterpret_cast<float*>(ws), static_cast<long long>(N) * H * W, Cout, s, dt);
}

int conv3x3_wgrad(const void* dz, int Cout, ConvSrc x0, ConvSrc x1, float* dw, void* ws, size_t ws_bytes, int N, int H,
                  int W, cudaStream_t s, float* db, int dt) {
  const int C1 = x1.ptr ? x1.C : 0;
  const int elt = dt_bytes(dt);   // the halo / row engines are bf16 only; fp32/tf32 goes to the generic per-tap kernel
  if (elt == 2 && wgrad_halo_supported(H, W, x0.C, Cout) && (!C1 || wgrad_halo_supported(H, W, C1, Cout))) {
    const ConvSrc srcs[2] = {x0, x1};
    int off = 0;
    for (int i = 0; i < (C1 ? 2 : 1); ++i) {
      const int Cx = srcs[i].C;
      const size_t need = wgrad_halo_ws_bytes(N, H, W, Cx, Cout);
      UB_REQUIRE(ws != nullptr && ws_bytes >= need, "conv3x3_wgrad: workspace of %zu bytes needed, %zu given", need,
                 ws_bytes);
      const int J = wgrad_halo_pick_J(Cx, W);
      CUtensorMap xm, zm;
      int rc = nhwc_map(&xm, srcs[i].ptr, N, H, W, Cx, Cx, Box{1, 16, 8 * J + 2});   // column halo
      if (rc) return rc;
      rc = nhwc_map(&zm, dz, N, H, W, Cout, Cout, Box{1, 18, 8 * J});              // row halo
      if (rc) return rc;
      rc = wgrad_halo_launch(xm, zm, N, H, W, Cx, Cout, J, dw, x0.C + C1, off, reinterpret_cast<float*>(ws), s,
                             i == 0 ? db : nullptr);   // fused: one extra MMA per K step against a tile of ones
      if (rc) return rc;
      off += Cx;
    }
    return UB_OK;
  }
  if (elt == 2 && wgrad_row_supported(H, W, Cout, x0.C, C1)) {
    WgradRowParams rp;
    std::memset(&rp, 0, sizeof(rp));
    int rc = wgrad_row_plan(rp, N, H, W, Cout, x0.C, C1);
    if (rc) return rc;
    const long long cin = x0.C + C1;
    const long long nweights = 9 * cin * Cout;
    rp.out_tap_stride = cin * Cout;
    rp.out_n_stride = Cout;
    // with a bias gradient every split's partial is [nweights | Cout]: one reduction produces dw and db
    const long long per_split = nweights + (db ? Cout : 0);
    rp.out_split_stride = per_split;
    if (rp.splits > 1) {
      const size_t need = static_cast<size_t>(rp.splits) * per_split * sizeof(float);
      UB_REQUIRE(ws != nullptr && ws_bytes >= need, "conv3x3_wgrad: workspace of %zu bytes needed, %zu given", need,
                 ws_bytes);
      rp.out = reinterpret_cast<float*>(ws);
      rp.bias_out = db ? rp.out + nweights : nullptr;
      rp.bias_split_stride = per_split;
    } else {
      rp.out = dw;
      rp.bias_out = db;
      rp.bias_split_stride = 0;
    }
    CUtensorMap zm, xm0, xm1;
    rc = nhwc_map(&zm, dz, N, H, W, Cout, 64, Box{1, 8, 8});
    if (rc) return rc;
    rc = nhwc_map(&xm0, x0.ptr, N, H, W, x0.C, 64, Box{1, 8, 10});
    if (rc) return rc;
    xm1 = xm0;
    if (C1) {
      rc = nhwc_map(&xm1, x1.ptr, N, H, W, C1, 64, Box{1, 8, 10});
      if (rc) return rc;
    }
    rc = wgrad_row_launch(rp, zm, xm0, xm1, s);
    if (rc) return rc;
    if (rp.splits > 1) {
      rc = splitk_reduce_launch(rp.out, per_split, rp.splits, dw, per_split, s, db, nweights);
      if (rc) return rc;
    }
    return UB_OK;
  }
  WgradPlan wp;
  int rc = plan_wgrad(wp, N, H, W, 3, 3, 1, Cout, x0.C, C1, elt);
  if (rc) return rc;
  WgradParams& p = wp.p;
  p.z_mode = 0;
  p.x_shift = 1;
  const long long cin = x0.C + C1;
  p.out_tap_stride = cin * Cout;
  p.out_n_stride = Cout;
  p.out_m_stride = 1;
  p.out_split_stride = wp.nweights;
  if (p.splits > 1) {
    UB_REQUIRE(ws != nullptr && ws_bytes >= static_cast<size_t>(p.splits) * wp.nweights * sizeof(float),
               "conv3x3_wgrad: workspace of %zu bytes needed, %zu given",
               static_cast<size_t>(p.splits) * wp.nweights * sizeof(float), ws_bytes);
    p.out = reinterpret_cast<float*>(ws);
  } else {
    p.out = dw;
  }
  const Box b = {p.TN, p.TH, p.TW};
  CUtensorMap zm, xm0, xm1;
  const int mn_sw = elt == 4 ? UB_SWIZZLE_128B_ATOM32 : 0;   // MN-major tf32 tiles (see wgrad.cu)
  rc = nhwc_map(&zm, dz, N, H, W, Cout, p.z_atom, b, elt, mn_sw);
  if (rc) return rc;
  rc = nhwc_map(&xm0, x0.ptr, N, H, W, x0.C, p.x_atom[0], b, elt, mn_sw);
  if (rc) return rc;
  xm1 = xm0;
  if (C1) {
    rc = nhwc_map(&xm1, x1.ptr, N, H, W, C1, p.x_atom[1], b, elt, mn_sw);
    if (rc) return rc;
  }
  rc = wgrad_launch(p, zm, xm0, xm1, s);
  if (rc) return rc;
  if (p.splits > 1) {
    rc = splitk_reduce_launch(p.out, wp.nweights, p.splits, dw, wp.nweights, s);
    if (rc) return rc;
  }
  return db ? bias_grad_after(dz, Cout, db, ws, ws_bytes, N, H, W, s, dt) : UB_OK;
}

// ------------------------------------------------------------------------------------------------ transposed conv
int tconv2x2_fwd(const void* x, int Cin, const void* w_packed, const float* bias, void* y, int Cout, int N, int H, int W,
                 cudaStream_t s, int dt) {
  const int elt = dt_bytes(dt);
  const int kc = pick_kc(Cin, 0, elt);
  UB_REQUIRE(kc != 0, "tconv2x2_fwd: Cin %d must be a multiple of 32", Cin);
  UB_REQUIRE(Cout % 32 == 0, "tconv2x2_fwd: Cout %d must be a multiple of 32", Cout);
  // GEMM N = 4 * Cout (the four sub-pixels side by side): one tile may span several of them, so the input tile is
  // loaded once and the MMAs are 128 / 256 wide even for Cout = 32 (four N = 32 tiles ran at 70 TFLOP/s)
  const int block_n = pick_block_n(4 * Cout);
  IgemmParams p;
  std::memset(&p, 0, sizeof(p));
  const Box b = pick_box(128, H, W);
  fill_grid(p, N, H, W, b);
  p.n_tiles = 4 * Cout / block_n;
  p.block_n = block_n;
  p.a_mode = IG_A_SHIFT4D;
  p.taps_r = 1; p.taps_s = 1; p.pad = 0;
  p.num_src = 1;
  p.src_kblocks[0] = Cin / kc;
  p.kc = kc;
  p.elt = elt;
  p.stages = igemm_pick_stages(block_n, kc, elt);
  p.epi.dst[0] = y; p.epi.dst_C[0] = Cout;
  p.epi.split_col = INT_MAX;
  p.epi.bias = bias;
  p.epi.mask_scale = 1.0f;
  p.epi.d2s_cout = Cout;
  p.epi.drop_keep = 1.0f;
  CUtensorMap a0, bm;
  int rc = nhwc_map(&a0, x, N, H, W, Cin, kc, b, elt);
  if (rc) return rc;
  rc = weight_map(&bm, w_packed, 4 * Cout, 1, Cin, kc, block_n, elt);
  if (rc) return rc;
  return igemm_launch(p, a0, a0, bm, s);
}

int tconv2x2_dgrad(const void* dy, int Cout, const void* w_packed_d, void* dx, int Cin, const void* mask, float scale,
                   int N, int H, int W, cudaStream_t s, const unsigned short* gate, int dt) {
  const int elt = dt_bytes(dt);
  int kc = pick_kc(Cout, 0, elt);
  UB_REQUIRE(kc != 0, "tconv2x2_dgrad: Cout %d must be a multiple of 32", Cout);
  const int block_n = pick_block_n(Cin);
  UB_REQUIRE(block_n != 0, "tconv2x2_dgrad: Cin %d must be a multiple of 32", Cin);
  // 32 output channels (transConv9): the two column taps of a filter row are adjacent in dY (2 x 64 B) and adjacent in
  // the packed weights' K order (tap * Cout + co), so one 128-byte K block per filter row replaces two 64-byte ones
  static const bool no_pair = getenv("UNETB200_NO_Z_PAIR") != nullptr;
  const bool pair = elt == 2 && Cout == 32 && !no_pair;
  if (pair) kc = 64;
  IgemmParams p;
  std::memset(&p, 0, sizeof(p));
  const Box b = pick_box(128, H, W, true);
  UB_REQUIRE(b.TN == 1 || b.TH == H, "tconv2x2_dgrad: %dx%d grid needs an image-aligned tile", H, W);
  fill_grid(p, N, H, W, b);
  p.n_tiles = Cin / block_n;
  p.block_n = block_n;
  p.a_mode = IG_A_TCONV5D;
  p.taps_r = 2; p.taps_s = pair ? 1 : 2; p.pad = 0;
  p.num_src = 1;
  p.src_kblocks[0] = pair ? 1 : Cout / kc;
  p.kc = kc;
  p.elt = elt;
  p.stages = igemm_pick_stages(block_n, kc, elt);
  p.epi.dst[0] = dx; p.epi.dst_C[0] = Cin; p.epi.aux[0] = gate ? nullptr : mask; p.epi.aux_C[0] = Cin;
  p.epi.gate_in[0] = gate;
  p.epi.split_col = INT_MAX;
  p.epi.mask_scale = scale;
  p.epi.drop_keep = 1.0f;
  CUtensorMap a0, bm;
  if (pair) p.a_mode = IG_A_TCONV_PAIR;
  int rc = pair ? fine_pair_map(&a0, dy, N, H, W, Cout, b, elt) : fine_view_map(&a0, dy, N, H, W, Cout, kc, b, elt);
  if (rc) return rc;
  rc = pair ? weight_map(&bm, w_packed_d, Cin, 2, 2 * Cout, kc, block_n, elt)
            : weight_map(&bm, w_packed_d, Cin, 4, Cout, kc, block_n, elt);
  if (rc) return rc;
  return igemm_launch(p, a0, a0, bm, s);
}

size_t tconv2x2_wgrad_ws_bytes(int N, int H, int W, int Cin, int Cout) {
  WgradPlan wp;
  if (plan_wgrad(wp, N, H, W, 2, 2, 0, Cout, Cin, 0) != UB_OK) return 0;
  // split-K partials of the weights (a single pixel range writes the gradient directly), then the bias partials of
  // every (split, tap); the tf32 path sums the bias with colsum_act in the same workspace afterwards
  const size_t wpart = wp.p.splits > 1 ? static_cast<size_t>(wp.p.splits) * wp.nweights * sizeof(float) : 0;
  const size_t bpart = static_cast<size_t>(wp.p.splits) * 4 * Cout * sizeof(float);
  const size_t colsum = static_cast<size_t>(colsum_ws_floats(Cout)) * sizeof(float);
  return wpart + bpart > colsum ? wpart + bpart : colsum;
}

int tconv2x2_wgrad(const void* dy, int Cout, const void* x, int Cin, float* dw, void* ws, size_t ws_bytes, int N, int H,
                   int W, cudaStream_t s, int dt, float* db) {
  const int elt = dt_bytes(dt);
  WgradPlan wp;
  int rc = plan_wgrad(wp, N, H, W, 2, 2, 0, Cout, Cin, 0, elt);
  if (rc) return rc;
  WgradParams& p = wp.p;
  UB_REQUIRE(p.TN == 1 || p.TH == H, "tconv2x2_wgrad: %dx%d grid needs an image-aligned block", H, W);
  p.z_mode = 1;
  p.x_shift = 0;
  p.out_tap_stride = static_cast<long long>(Cout) * Cin;
  p.out_m_stride = Cin;
  p.out_n_stride = 1;
  p.out_split_stride = wp.nweights;
  const size_t wpart = p.splits > 1 ? static_cast<size_t>(p.splits) * wp.nweights * sizeof(float) : 0;
  const bool fused_bias = db != nullptr && elt == 2;
  const size_t bpart = fused_bias ? static_cast<size_t>(p.splits) * 4 * Cout * sizeof(float) : 0;
  UB_REQUIRE(wpart + bpart == 0 || (ws != nullptr && ws_bytes >= wpart + bpart),
             "tconv2x2_wgrad: workspace of %zu bytes needed, %zu given", wpart + bpart, ws_bytes);
  p.out = p.splits > 1 ? reinterpret_cast<float*>(ws) : dw;
  // the bias gradient rides along: the CTAs of the first input-channel tile sum the dY tiles they stage
  float* bias_part = fused_bias ? reinterpret_cast<float*>(static_cast<char*>(ws) + wpart) : nullptr;
  p.bias_out = bias_part;
  const Box b = {p.TN, p.TH, p.TW};
  CUtensorMap zm, xm;
  const int mn_sw = elt == 4 ? UB_SWIZZLE_128B_ATOM32 : 0;
  if (p.z_pair) rc = fine_pair_map(&zm, dy, N, H, W, Cout, b, elt);
  else rc = fine_view_map(&zm, dy, N, H, W, Cout, p.z_atom, b, elt, mn_sw);
  if (rc) return rc;
  rc = nhwc_map(&xm, x, N, H, W, Cin, p.x_atom[0], b, elt, mn_sw);
  if (rc) return rc;
  rc = wgrad_launch(p, zm, xm, xm, s);
  if (rc) return rc;
  if (p.splits > 1) {
    rc = splitk_reduce_launch(p.out, wp.nweights, p.splits, dw, wp.nweights, s);
    if (rc) return rc;
  }
  if (fused_bias) return rows_sum(bias_part, p.splits * 4, Cout, db, s);
  if (db) {   // tf32: a separate pass over dY
    const size_t need = static_cast<size_t>(colsum_ws_floats(Cout)) * sizeof(float);
    UB_REQUIRE(ws != nullptr && ws_bytes >= need, "tconv2x2_wgrad: workspace of %zu bytes needed for the bias gradient", need);
    return colsum_act(dy, db, reinterpret_cast<float*>(ws), static_cast<long long>(N) * 4 * H * W, Cout, s, dt);
  }
  return UB_OK;
}

}  // namespace ub
