#include "conv_halo.cuh"

#include <cstdlib>

#include "common.h"
#include "ptx.cuh"

namespace ub {

namespace {

constexpr int kThreads = 384;  // warp0 halo TMA, warp1 MMA + TMEM, warp2 weight TMA, warp3 idle, warps4..11 epilogue
constexpr int kMaxHaloStages = 4;
constexpr int kMaxBStages = 4;
constexpr int kTH = 16;        // image rows per super-tile
constexpr int kSmemBudget = 225 * 1024;

struct __align__(8) HaloBarriers {
  uint64_t halo_full[kMaxHaloStages];
  uint64_t halo_empty[kMaxHaloStages];
  uint64_t b_full[kMaxBStages];
  uint64_t b_empty[kMaxBStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

__host__ __device__ inline int round_up_1k(int x) { return (x + 1023) & ~1023; }

// Optional wait-time accounting (UNETB200_HALO_TIMING=1): cycles each role spends blocked, summed over CTAs.
// [0] MMA on tmem_empty, [1] MMA on halo_full, [2] MMA on b_full, [3] MMA total, [4] epilogue(warp4) on tmem_full,
// [5] epilogue(warp4) total, [6] halo producer on halo_empty, [7] CTAs
__device__ unsigned long long g_halo_timing[8];

__device__ __forceinline__ void timed_wait(uint64_t* bar, uint32_t parity, bool on, unsigned long long& acc) {
  if (!on) { mbar_wait(bar, parity); return; }
  const long long t0 = clock64();
  mbar_wait(bar, parity);
  acc += static_cast<unsigned long long>(clock64() - t0);
}

__device__ __forceinline__ void decode_tile(const HaloParams& p, int tile, int& n, int& h0, int& w0, int& nt) {
  nt = tile % p.n_tiles;
  int t = tile / p.n_tiles;
  const int tw_i = t % p.tiles_w;
  t /= p.tiles_w;
  const int th_i = t % p.tiles_h;
  n = t / p.tiles_h;
  h0 = th_i * kTH;
  w0 = tw_i * 8 * p.J;
}

__device__ __forceinline__ uint64_t desc_pair(uint32_t lo, uint32_t hi) {
  return (static_cast<uint64_t>(hi) << 32) | lo;
}

// MMA issuer of one CTA, specialised on the plan (J pixel strips, KC channels per block, BN columns) so that every
// operand descriptor of a k-block is "stage base + compile-time constant": measured with csrc/mma_rate.cu, an SS-mode
// 128xNx16 MMA costs 32 + N/4 cycles (40 / 48 / 64 for N = 32 / 64 / 128) when the issuing lane keeps up, and each
// extra scalar instruction between two MMAs costs ~4 cycles of a single warp's dependent-issue latency -- the
// run-time-generic loop (tap / 3, 64-bit descriptor adds, R2UR moves) ran at 62 / 60 / 92 cycles per MMA.
// The whole warp walks the (uniform) control flow and waits; one elected lane issues.
template <int J, int KC, int BN>
__device__ __forceinline__ void halo_mma_role(const HaloParams& p, HaloBarriers* bars, uint32_t halo_addr,
                                              uint32_t halo_stride, uint32_t b_addr, uint32_t tmem_base,
                                              int num_tiles, int lane) {
  constexpr int P = 8 * J + 2;                      // halo pitch in pixels
  constexpr uint32_t ROW16 = KC * 2 / 16;           // one halo row, in 16-byte units
  constexpr int KSTEPS = KC / 16;
  constexpr int BTILE = BN * KC * 2;
  constexpr int TPB = (9 * BTILE <= 36 * 1024) ? 9 : 3;   // must match halo_plan()
  constexpr int TAP_GROUPS = 9 / TPB;
  constexpr uint32_t BSTAGE = TPB * BTILE;
  constexpr uint32_t SW = KC == 64 ? UMMA_SW_128B : UMMA_SW_64B;
  const uint32_t idesc = umma_idesc_bf16(128, BN, 0, 0);
  const uint32_t a_hi = static_cast<uint32_t>(umma_smem_desc(0, 0, P * KC * 2, SW) >> 32);   // SBO = next image row
  const uint32_t b_hi = static_cast<uint32_t>(umma_smem_desc(0, 0, 8 * KC * 2, SW) >> 32);
  const int total_kb = p.src_kblocks[0] + (p.num_src > 1 ? p.src_kblocks[1] : 0);
  int hs = 0, bs = 0, it = 0;
  uint32_t hphase = 0, bphase = 0;
  unsigned long long t_e = 0, t_h = 0, t_b = 0;
  const long long t_mma0 = clock64();
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const int a = p.acc_double ? (it & 1) : 0;
    const uint32_t eparity = p.acc_double ? (((it >> 1) & 1) ^ 1u) : ((it & 1) ^ 1u);
    timed_wait(&bars->tmem_empty[a], eparity, p.timing != 0, t_e);
    tc_fence_after();
    const uint32_t d0 = tmem_base + static_cast<uint32_t>(a * J * BN);
    for (int kb = 0; kb < total_kb; ++kb) {
      timed_wait(&bars->halo_full[hs], hphase, p.timing != 0, t_h);
      const uint32_t a_lo = (halo_addr + static_cast<uint32_t>(hs) * halo_stride) >> 4;
      const uint32_t fresh = kb == 0 ? 0u : 1u;   // accumulate flag of the very first MMA of each accumulator
#pragma unroll
      for (int tg = 0; tg < TAP_GROUPS; ++tg) {
        timed_wait(&bars->b_full[bs], bphase, p.timing != 0, t_b);
        tc_fence_after();
        const uint32_t b_lo = (b_addr + static_cast<uint32_t>(bs) * BSTAGE) >> 4;
        if (elect_one()) {
#pragma unroll
          for (int t = 0; t < TPB; ++t) {
            const int tap = tg * TPB + t;
            const int r = tap / 3, s = tap % 3;
#pragma unroll
            for (int j = 0; j < J; ++j) {
#pragma unroll
              for (int k = 0; k < KSTEPS; ++k) {
                umma_f16(d0 + static_cast<uint32_t>(j * BN),
                         desc_pair(a_lo + static_cast<uint32_t>((r * P + s + 8 * j) * ROW16 + 2 * k), a_hi),
                         desc_pair(b_lo + static_cast<uint32_t>(t * (BTILE >> 4) + 2 * k), b_hi), idesc,
                         (tap == 0 && k == 0) ? fresh : 1u);
              }
            }
          }
          umma_commit(&bars->b_empty[bs]);
          if (tg == TAP_GROUPS - 1) umma_commit(&bars->halo_empty[hs]);
        }
        __syncwarp();
        if (++bs == p.b_stages) { bs = 0; bphase ^= 1u; }
      }
      if (++hs == p.h_stages) { hs = 0; hphase ^= 1u; }
    }
    if (elect_one()) umma_commit(&bars->tmem_full[a]);
    __syncwarp();
  }
  if (p.timing && lane == 0) {
    atomicAdd(&g_halo_timing[0], t_e);
    atomicAdd(&g_halo_timing[1], t_h);
    atomicAdd(&g_halo_timing[2], t_b);
    atomicAdd(&g_halo_timing[3], static_cast<unsigned long long>(clock64() - t_mma0));
  }
}

// Epilogue of one CTA's tiles, specialised at compile time so that a 16-column chunk costs tens of instructions, not
// hundreds (ncu: the run-time-generic version executed 33 M warp instructions for conv1b and was the bottleneck of
// every 32/64-channel layer). Work items are (pixel tile j, 16-column chunk); the two warps of a TMEM lane quarter
// take alternate items. BIAS_RELU: forward conv (bias + ReLU, bias may be null); DROP: Philox dropout after ReLU;
// AUX: dgrad, gate by the producer's stored output (two destinations possible). The gate of the NEXT item is
// requested before the current accumulators are read so that its latency hides behind the TMEM load and the stores.
template <int BN, bool BIAS_RELU, bool DROP, bool AUX, bool GIN>
__device__ __forceinline__ void halo_epilogue(const HaloParams& p, HaloBarriers* bars, uint32_t tmem_base, int warp,
                                              int lane, int num_tiles) {
  constexpr int kChunks = BN / 16;
  constexpr int kMaxItems = 8;   // per warp: J * kChunks / 2 <= 8 for every plan halo_plan() produces
  const IgemmEpilogue& e = p.epi;
  const int quarter = warp & 3;
  const int half = (warp - 4) >> 2;
  const int m = quarter * 32 + lane;
  const int th = m >> 3, twl = m & 7;
  const int items = p.J * kChunks;
  const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
  unsigned long long seed = 0;
  if (DROP) seed = e.drop_seed + (e.drop_seed_add ? static_cast<unsigned long long>(__ldg(e.drop_seed_add)) : 0ull);
  const float inv_keep = DROP ? 1.0f / e.drop_keep : 1.0f;
  int it = 0;
  unsigned long long t_wait = 0;
  const long long t_begin = clock64();
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const int a = p.acc_double ? (it & 1) : 0;
    const uint32_t fparity = p.acc_double ? ((it >> 1) & 1) : (it & 1);
    int n, h0, w0, nt;
    decode_tile(p, tile, n, h0, w0, nt);
    const int h = h0 + th;
    const int wbase = w0 + twl;
    const bool row_ok = h < p.H;
    const size_t pix0 = (static_cast<size_t>(n) * p.H + h) * p.W + wbase;   // pixel of strip j = pix0 + 8 j
    const int col_tile = nt * BN;
    // per-destination byte bases of this thread's pixel row (strip 0)
    const char* aux_b[2] = {nullptr, nullptr};
    char* dst_b[2];
#pragma unroll
    for (int sl = 0; sl < 2; ++sl) {
      dst_b[sl] = reinterpret_cast<char*>(e.dst[sl]) + (pix0 * e.dst_C[sl] + e.dst_coff[sl]) * 2;
      if (AUX && e.aux[sl]) {
        aux_b[sl] = reinterpret_cast<const char*>(e.aux[sl]) + (pix0 * e.aux_C[sl] + e.aux_coff[sl]) * 2;
      }
    }
    uint4 ax0 = make_uint4(0u, 0u, 0u, 0u), ax1 = ax0;
    auto load_aux = [&](int item) {
      const int j = item / kChunks, c = (item % kChunks) * 16;
      const int col = col_tile + c;
      const int sl = col >= e.split_col ? 1 : 0;
      if (row_ok && (wbase + 8 * j) < p.W && aux_b[sl]) {
        const int ch = sl ? col - e.split_col : col;
        const uint4* ap = reinterpret_cast<const uint4*>(aux_b[sl] + (static_cast<size_t>(8 * j) * e.aux_C[sl] + ch) * 2);
        ax0 = __ldg(ap);
        ax1 = __ldg(ap + 1);
      }
    };
    if (AUX && half < items) load_aux(half);
    // GIN: the producer's ReLU gates as bits -- all of this warp's items of the tile are requested up front (one
    // 2-byte load each), so their latency overlaps the MMAs of this very tile
    uint32_t gb[kMaxItems];
    if (GIN) {
#pragma unroll
      for (int k = 0; k < kMaxItems; ++k) {
        const int item = half + 2 * k;
        gb[k] = 0xFFFFu;
        if (item < items) {
          const int j = item / kChunks, c = (item % kChunks) * 16;
          const int col = col_tile + c;
          const int sl = col >= e.split_col ? 1 : 0;
          if (row_ok && (wbase + 8 * j) < p.W && e.gate_in[sl]) {
            const int ch = sl ? col - e.split_col : col;
            gb[k] = __ldg(e.gate_in[sl] + (((pix0 + 8 * j) * e.aux_C[sl] + e.aux_coff[sl] + ch) >> 4));
          }
        }
      }
    }
    timed_wait(&bars->tmem_full[a], fparity, p.timing != 0, t_wait);
    tc_fence_after();
#pragma unroll
    for (int k = 0; k < kMaxItems; ++k) {
      const int item = half + 2 * k;
      if (item >= items) break;
      const int j = item / kChunks, c = (item % kChunks) * 16;
      const bool valid = row_ok && (wbase + 8 * j) < p.W;
      const uint4 cur0 = ax0, cur1 = ax1;
      if (AUX && item + 2 < items) load_aux(item + 2);
      uint32_t r[16];
      tmem_ld_32x32b_x16(t_lane + static_cast<uint32_t>((a * p.J + j) * BN + c), r);
      tmem_ld_wait();
      if (!valid) continue;
      const int col = col_tile + c;
      const int sl = ((AUX || GIN) && col >= e.split_col) ? 1 : 0;
      const int ch = sl ? col - e.split_col : col;
      float v[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
      if (BIAS_RELU) {
        if (e.bias) {
          const float4* bp = reinterpret_cast<const float4*>(e.bias + col);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 bq = __ldg(bp + q);
            v[4 * q + 0] += bq.x; v[4 * q + 1] += bq.y; v[4 * q + 2] += bq.z; v[4 * q + 3] += bq.w;
          }
        }
        if (e.relu) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.0f);
        }
      }
      const size_t off = (static_cast<size_t>(8 * j) * e.dst_C[sl] + ch) * 2;   // bytes from dst_b[sl]
      if (DROP) {
        const size_t elem0 = (pix0 + 8 * j) * e.dst_C[sl] + e.dst_coff[sl] + ch;
        if (e.drop_bits) {
          const uint32_t bits = __ldg(e.drop_bits + (elem0 >> 4));
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = ((bits >> i) & 1u) ? v[i] * inv_keep : 0.0f;
        } else {
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const uint32_t bits = dropout_keep_bits8(seed, e.drop_stream, elem0 / 8 + g, e.drop_keep);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[8 * g + i] = ((bits >> i) & 1u) ? v[8 * g + i] * inv_keep : 0.0f;
        }
        }
      }
      if (BIAS_RELU) {
        if (e.gate_out) {
          uint32_t bits = 0;
#pragma unroll
          for (int i = 0; i < 16; ++i) bits |= (v[i] > 0.0f ? 1u : 0u) << i;
          e.gate_out[((pix0 + 8 * j) * e.dst_C[0] + e.dst_coff[0] + ch) >> 4] = static_cast<unsigned short>(bits);
        }
      }
      if (GIN) {
        if (e.gate_in[sl]) {
          const uint32_t b = gb[k];
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = ((b >> i) & 1u) ? v[i] * e.mask_scale : 0.0f;
        }
      }
      if (AUX) {
        if (aux_b[sl]) {
          const uint32_t aw[8] = {cur0.x, cur0.y, cur0.z, cur0.w, cur1.x, cur1.y, cur1.z, cur1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[2 * i] = bf16_lo(aw[i]) > 0.0f ? v[2 * i] * e.mask_scale : 0.0f;
            v[2 * i + 1] = bf16_hi(aw[i]) > 0.0f ? v[2 * i + 1] * e.mask_scale : 0.0f;
          }
        }
      }
      uint4* dp = reinterpret_cast<uint4*>(dst_b[sl] + off);
      uint4 o0, o1;
      o0.x = pack_bf16x2(v[0], v[1]);   o0.y = pack_bf16x2(v[2], v[3]);
      o0.z = pack_bf16x2(v[4], v[5]);   o0.w = pack_bf16x2(v[6], v[7]);
      o1.x = pack_bf16x2(v[8], v[9]);   o1.y = pack_bf16x2(v[10], v[11]);
      o1.z = pack_bf16x2(v[12], v[13]); o1.w = pack_bf16x2(v[14], v[15]);
      dp[0] = o0;
      dp[1] = o1;
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&bars->tmem_empty[a]);
  }
  if (p.timing && warp == 4 && lane == 0) {
    atomicAdd(&g_halo_timing[4], t_wait);
    atomicAdd(&g_halo_timing[5], static_cast<unsigned long long>(clock64() - t_begin));
  }
}

// Lean epilogue for the shapes the UNet step actually runs (full super-tiles, one N tile, no dropout, gates as bits):
// everything that does not depend on the item is hoisted out of the item loop or made a compile-time constant.
// ncu on conv1b forward (128 x 128x128 x 32 -> 32) showed the generic epilogue above executing 235 warp
// instructions per (32 pixels x 16 columns) item -- 46 IMAD, 25 ISETP, 20 IADD3, 19 BRA, 19 R2UR -- with only two
// epilogue warps per scheduler, i.e. latency-bound at ~1100 cycles per item while the MMAs of a super-tile need ~720.
//   FWD   : v + bias -> cvt.rn.relu.bf16x2 (ReLU inside the conversion) -> store; the ReLU gates come from the
//           packed words: w + 0x7fff7fff sets bit 15 / 31 iff the low / high bf16 is non-zero.
//   !FWD  : dgrad; per slot optional gate bits (prefetched for the whole tile before the accumulators are ready).
//   SPLIT : columns [BN/2, BN) go to destination 1 (the skip connection's gradient).
__device__ __forceinline__ uint32_t pack_bf16x2_relu(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

//   POOL  : (with SPLIT) destination 0 is the coarse tensor: each 2x2 block of the gradient is summed across lanes
//           (lane ^ 1 = next column, lane ^ 8 = next row) in fp32 and the even-row/even-column lane stores it.
//   DROP  : (forward) Philox dropout after the ReLU, same element numbering as the generic epilogue.
// With several N tiles (dgrad whose destinations split at a tile boundary) the destination is chosen per tile.
template <int J, int BN, bool FWD, bool SPLIT, bool POOL = false, bool DROP = false>
__device__ __forceinline__ void halo_epilogue_fast(const HaloParams& p, HaloBarriers* bars, uint32_t tmem_base,
                                                   int warp, int lane, int num_tiles) {
  static_assert(!SPLIT || BN >= 64, "split destinations need 32-column halves");
  static_assert(!POOL || (SPLIT && !FWD), "2x2 pooling of destination 0 is a split-dgrad feature");
  static_assert(!DROP || FWD, "dropout belongs to the forward epilogue");
  constexpr int kCl = BN / 32;              // 16-column chunks per warp (the two warps of a quarter interleave)
  constexpr int kItems = J * kCl;
  constexpr bool kBiasRegs = FWD && BN <= 64;
  const IgemmEpilogue& e = p.epi;
  const int quarter = warp & 3;
  const int half = (warp - 4) >> 2;
  const int m = quarter * 32 + lane;
  const int th = m >> 3, twl = m & 7;
  const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(half * 16);
  float bias_r[kBiasRegs ? kCl : 1][16];
  if (kBiasRegs) {
#pragma unroll
    for (int cl = 0; cl < kCl; ++cl) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
        if (e.bias) b = __ldg(reinterpret_cast<const float4*>(e.bias + half * 16 + 32 * cl) + q);
        bias_r[cl][4 * q + 0] = b.x; bias_r[cl][4 * q + 1] = b.y; bias_r[cl][4 * q + 2] = b.z; bias_r[cl][4 * q + 3] = b.w;
      }
    }
  }
  const float scale = e.mask_scale;
  const bool scaled = scale != 1.0f;
  unsigned long long seed = 0;
  if (DROP) seed = e.drop_seed + (e.drop_seed_add ? static_cast<unsigned long long>(__ldg(e.drop_seed_add)) : 0ull);
  const float inv_keep = DROP ? 1.0f / e.drop_keep : 1.0f;
  constexpr int kSlots = SPLIT ? 2 : 1;
  uint32_t strip_b[kSlots], gstrip[kSlots];   // bytes / gate words between two 8-pixel strips
#pragma unroll
  for (int sl = 0; sl < kSlots; ++sl) {
    strip_b[sl] = static_cast<uint32_t>(16 * e.dst_C[sl]);
    gstrip[sl] = static_cast<uint32_t>(((FWD || (POOL && sl == 0)) ? e.dst_C[sl] : e.aux_C[sl]) >> 1);
  }
  int it = 0;
  unsigned long long t_wait = 0;
  const long long t_begin = clock64();
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const int a = p.acc_double ? (it & 1) : 0;
    const uint32_t fparity = p.acc_double ? ((it >> 1) & 1) : (it & 1);
    int n, h0, w0, nt;
    decode_tile(p, tile, n, h0, w0, nt);
    const size_t pix0 = (static_cast<size_t>(n) * p.H + (h0 + th)) * p.W + (w0 + twl);
    char* dst_b[kSlots];
    const unsigned short* gin_b[kSlots];
    unsigned short* gout_b = nullptr;
    if (SPLIT) {
#pragma unroll
      for (int sl = 0; sl < kSlots; ++sl) {
        dst_b[sl] = reinterpret_cast<char*>(e.dst[sl]) + (pix0 * e.dst_C[sl] + e.dst_coff[sl] + half * 16) * 2;
        gin_b[sl] = nullptr;
        if (!FWD && e.gate_in[sl]) gin_b[sl] = e.gate_in[sl] + ((pix0 * e.aux_C[sl] + e.aux_coff[sl]) >> 4) + half;
      }
    } else {
      // one destination per N tile: the tile's first column decides which (warp-uniform), channels count from there
      const int col0 = nt * BN;
      const int sr = (!FWD && col0 >= e.split_col) ? 1 : 0;
      const int ch0 = col0 - (sr ? e.split_col : 0) + half * 16;
      dst_b[0] = reinterpret_cast<char*>(e.dst[sr]) + (pix0 * e.dst_C[sr] + e.dst_coff[sr] + ch0) * 2;
      strip_b[0] = static_cast<uint32_t>(16 * e.dst_C[sr]);
      gin_b[0] = nullptr;
      if (!FWD && e.gate_in[sr]) {
        gin_b[0] = e.gate_in[sr] + ((pix0 * e.aux_C[sr] + e.aux_coff[sr] + ch0) >> 4);
        gstrip[0] = static_cast<uint32_t>(e.aux_C[sr] >> 1);
      }
    }
    if (FWD && e.gate_out) gout_b = e.gate_out + ((pix0 * e.dst_C[0] + e.dst_coff[0]) >> 4) + half;
    const bool pool_lane = (lane & 9) == 0;   // even row and even column of the tile
    if (POOL) {   // destination 0 addressed on the coarse grid (this lane's 2x2 block)
      const size_t cpix0 = (static_cast<size_t>(n) * (p.H >> 1) + ((h0 + th) >> 1)) * (p.W >> 1) + ((w0 + twl) >> 1);
      dst_b[0] = reinterpret_cast<char*>(e.dst[0]) + (cpix0 * e.dst_C[0] + e.dst_coff[0] + half * 16) * 2;
      gin_b[0] = nullptr;
      if (e.gate_in[0] && pool_lane) gin_b[0] = e.gate_in[0] + ((cpix0 * e.dst_C[0] + e.dst_coff[0]) >> 4) + half;
    }
    uint32_t gb[kItems];
    if (!FWD) {
#pragma unroll
      for (int k = 0; k < kItems; ++k) {
        const int j = k / kCl, cl = k % kCl;
        const int sl = (SPLIT && 32 * cl >= BN / 2) ? 1 : 0;
        gb[k] = 0xFFFFu;
        const uint32_t gs = (POOL && sl == 0) ? (gstrip[0] >> 1) : gstrip[sl];   // coarse strips are 4 pixels apart
        if (gin_b[sl]) gb[k] = __ldg(gin_b[sl] + j * gs + (2 * cl - sl * (BN / 32)));
      }
    }
    timed_wait(&bars->tmem_full[a], fparity, p.timing != 0, t_wait);
    tc_fence_after();
    const uint32_t t_acc = t_lane + static_cast<uint32_t>(a * J * BN);
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
      const int j = k / kCl, cl = k % kCl;
      const int sl = (SPLIT && 32 * cl >= BN / 2) ? 1 : 0;
      uint32_t r[16];
      tmem_ld_32x32b_x16(t_acc + static_cast<uint32_t>(j * BN + 32 * cl), r);
      tmem_ld_wait();
      uint32_t w[8];
      if (FWD) {
        float v[16];
        if (kBiasRegs) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]) + bias_r[kBiasRegs ? cl : 0][i];
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
          if (e.bias) {
            const float4* bp = reinterpret_cast<const float4*>(e.bias + half * 16 + 32 * cl);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float4 bq = __ldg(bp + q);
              v[4 * q + 0] += bq.x; v[4 * q + 1] += bq.y; v[4 * q + 2] += bq.z; v[4 * q + 3] += bq.w;
            }
          }
        }
        if (DROP) {   // (dropout of a negative value stays negative or zero: the ReLU in the conversion still applies)
          const size_t elem0 = (pix0 + 8 * j) * e.dst_C[0] + e.dst_coff[0] + half * 16 + 32 * cl;
          if (e.drop_bits) {   // precomputed keep bits: one 2-byte load per item
            const uint32_t bits = __ldg(e.drop_bits + (elem0 >> 4));
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = ((bits >> i) & 1u) ? v[i] * inv_keep : 0.0f;
          } else {
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            const uint32_t bits = dropout_keep_bits8(seed, e.drop_stream, elem0 / 8 + g, e.drop_keep);
#pragma unroll
            for (int i = 0; i < 8; ++i) v[8 * g + i] = ((bits >> i) & 1u) ? v[8 * g + i] * inv_keep : 0.0f;
          }
          }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = pack_bf16x2_relu(v[2 * i], v[2 * i + 1]);
        if (gout_b) {
          uint32_t acc = 0;
#pragma unroll
          for (int i = 0; i < 8; ++i) acc |= ((w[i] + 0x7FFF7FFFu) >> (15 - 2 * i)) & (0x00010001u << (2 * i));
          gout_b[j * gstrip[0] + 2 * cl] = static_cast<unsigned short>((acc & 0xFFFFu) | (acc >> 15));
        }
      } else {
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
        if (POOL && sl == 0) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            v[i] += __shfl_xor_sync(0xffffffffu, v[i], 1);
            v[i] += __shfl_xor_sync(0xffffffffu, v[i], 8);
          }
        }
        if (gin_b[sl]) {
          const uint32_t b = gb[k];
          if (scaled) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] *= scale;
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = ((b >> i) & 1u) ? v[i] : 0.0f;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
      }
      U32x8 o;
#pragma unroll
      for (int i = 0; i < 8; ++i) o.v[i] = w[i];
      if (POOL && sl == 0) {   // coarse strips are 4 pixels apart; one lane per 2x2 block stores
        if (pool_lane) st_global_v8(dst_b[0] + j * (strip_b[0] >> 1) + 32 * cl * 2, o);
      } else {
        st_global_v8(dst_b[sl] + j * strip_b[sl] + (32 * cl - sl * (BN / 2)) * 2, o);   // one full sector per lane
      }
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&bars->tmem_empty[a]);
  }
  if (p.timing && warp == 4 && lane == 0) {
    atomicAdd(&g_halo_timing[4], t_wait);
    atomicAdd(&g_halo_timing[5], static_cast<unsigned long long>(clock64() - t_begin));
  }
}

__global__ void __launch_bounds__(kThreads, 1)
conv_halo_kernel(const __grid_constant__ HaloParams p, const __grid_constant__ CUtensorMap tmA0,
                 const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB) {
  pdl_trigger();     // the next kernel of the stream may take SMs as this grid's CTAs leave them
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

  const int P = 8 * p.J + 2;                       // halo pitch in pixels
  const int rowbytes = p.kc * 2;
  const int halo_bytes = (kTH + 2) * P * rowbytes; // exactly one TMA box
  const int halo_stride = round_up_1k(halo_bytes);
  const int btile_bytes = p.block_n * rowbytes;
  const int bstage_bytes = p.tpb * btile_bytes;    // multiple of 1024 (block_n >= 32, kc >= 32 -> >= 2048)
  uint8_t* halo_base = smem;
  uint8_t* b_base = smem + p.h_stages * halo_stride;
  HaloBarriers* bars = reinterpret_cast<HaloBarriers*>(b_base + p.b_stages * bstage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.N * p.tiles_h * p.tiles_w * p.n_tiles;
  const int tap_groups = 9 / p.tpb;
  const int acc_sets = p.acc_double ? 2 : 1;
  uint32_t tmem_cols = 32;
  while (tmem_cols < static_cast<uint32_t>(acc_sets * p.J * p.block_n)) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.h_stages; ++s) {
      mbar_init(&bars->halo_full[s], 1);
      mbar_init(&bars->halo_empty[s], 1);
    }
    for (int s = 0; s < p.b_stages; ++s) {
      mbar_init(&bars->b_full[s], 1);
      mbar_init(&bars->b_empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&bars->tmem_full[a], 1);
      mbar_init(&bars->tmem_empty[a], 8);
    }
    mbar_fence_init();
    tma_prefetch_desc(&tmA0);
    if (p.num_src > 1) tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    tmem_alloc(&bars->tmem_base, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();        // everything above overlapped the previous kernel's tail; its results are visible from here
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------ halo producer
    if (lane == 0) {
      int hs = 0;
      uint32_t hphase = 0;
      unsigned long long t_prod = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int n, h0, w0, nt;
        decode_tile(p, tile, n, h0, w0, nt);
        for (int src = 0; src < p.num_src; ++src) {
          const CUtensorMap* tmA = src == 0 ? &tmA0 : &tmA1;
          for (int cb = 0; cb < p.src_kblocks[src]; ++cb) {
            timed_wait(&bars->halo_empty[hs], hphase ^ 1u, p.timing != 0, t_prod);
            mbar_arrive_expect_tx(&bars->halo_full[hs], static_cast<uint32_t>(halo_bytes));
            tma_load_4d(halo_base + hs * halo_stride, tmA, &bars->halo_full[hs], cb * p.kc, w0 - 1, h0 - 1, n);
            if (++hs == p.h_stages) { hs = 0; hphase ^= 1u; }
          }
        }
      }
      if (p.timing) {
        atomicAdd(&g_halo_timing[6], t_prod);
        atomicAdd(&g_halo_timing[7], 1ull);
      }
    }
  } else if (warp == 2) {
    // ------------------------------------------------------------ weight producer
    if (lane == 0) {
      int bs = 0;
      uint32_t bphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int nt = tile % p.n_tiles;
        for (int src = 0; src < p.num_src; ++src) {
          for (int cb = 0; cb < p.src_kblocks[src]; ++cb) {
            for (int tg = 0; tg < tap_groups; ++tg) {
              mbar_wait(&bars->b_empty[bs], bphase ^ 1u);
              mbar_arrive_expect_tx(&bars->b_full[bs], static_cast<uint32_t>(bstage_bytes));
              uint8_t* dst = b_base + bs * bstage_bytes;
              for (int t = 0; t < p.tpb; ++t) {
                tma_load_3d(dst + t * btile_bytes, &tmB, &bars->b_full[bs], p.src_ci_off[src] + cb * p.kc,
                            tg * p.tpb + t, nt * p.block_n);
              }
              if (++bs == p.b_stages) { bs = 0; bphase ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (see halo_mma_role)
    const uint32_t ha = smem_u32(halo_base), ba = smem_u32(b_base);
    const uint32_t hst = static_cast<uint32_t>(halo_stride);
#define UB_MMA(JJ, KK, NN) halo_mma_role<JJ, KK, NN>(p, bars, ha, hst, ba, tmem_base, num_tiles, lane)
#define UB_MMA_J(KK, NN)                        \
    do {                                        \
      if (p.J == 4) UB_MMA(4, KK, NN);          \
      else if (p.J == 2) UB_MMA(2, KK, NN);     \
      else UB_MMA(1, KK, NN);                   \
    } while (0)
    if (p.kc == 64) {
      if (p.block_n == 32) UB_MMA_J(64, 32);
      else if (p.block_n == 64) UB_MMA_J(64, 64);
      else UB_MMA_J(64, 128);
    } else {
      if (p.block_n == 32) UB_MMA_J(32, 32);
      else if (p.block_n == 64) UB_MMA_J(32, 64);
      else UB_MMA_J(32, 128);
    }
#undef UB_MMA_J
#undef UB_MMA
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue: 8 warps, two per TMEM lane quarter
    const IgemmEpilogue& e = p.epi;
    const bool aux = (e.aux[0] != nullptr) || (e.aux[1] != nullptr);
    const bool gin = (e.gate_in[0] != nullptr) || (e.gate_in[1] != nullptr);
    const bool drop = e.drop_keep < 1.0f;
#define UB_FAST(JJ, BN)                                                                                    \
    do {                                                                                                   \
      if (p.fast_epi == 1) halo_epilogue_fast<JJ, BN, true, false>(p, bars, tmem_base, warp, lane, num_tiles);       \
      else if (p.fast_epi == 5) halo_epilogue_fast<JJ, BN, true, false, false, true>(p, bars, tmem_base, warp, lane, num_tiles); \
      else if (p.fast_epi == 2) halo_epilogue_fast<JJ, BN, false, false>(p, bars, tmem_base, warp, lane, num_tiles); \
      else if (p.fast_epi == 3) halo_epilogue_fast<JJ, (BN >= 64 ? BN : 64), false, true>(p, bars, tmem_base, warp, lane, num_tiles); \
      else halo_epilogue_fast<JJ, (BN >= 64 ? BN : 64), false, true, true>(p, bars, tmem_base, warp, lane, num_tiles); \
    } while (0)
#define UB_EPI(BN)                                                                                         \
    do {                                                                                                   \
      if (p.fast_epi) {                                                                                    \
        if (p.J == 4) UB_FAST(4, BN);                                                                      \
        else if (p.J == 2) UB_FAST(2, BN);                                                                 \
        else UB_FAST(1, BN);                                                                               \
      }                                                                                                    \
      else if (gin && aux) halo_epilogue<BN, false, false, true, true>(p, bars, tmem_base, warp, lane, num_tiles); \
      else if (gin) halo_epilogue<BN, false, false, false, true>(p, bars, tmem_base, warp, lane, num_tiles);    \
      else if (aux) halo_epilogue<BN, false, false, true, false>(p, bars, tmem_base, warp, lane, num_tiles); \
      else if (drop) halo_epilogue<BN, true, true, false, false>(p, bars, tmem_base, warp, lane, num_tiles); \
      else if (e.bias || e.relu) halo_epilogue<BN, true, false, false, false>(p, bars, tmem_base, warp, lane, num_tiles); \
      else halo_epilogue<BN, false, false, false, false>(p, bars, tmem_base, warp, lane, num_tiles);       \
    } while (0)
    if (p.block_n == 32) UB_EPI(32);
    else if (p.block_n == 64) UB_EPI(64);
    else UB_EPI(128);
#undef UB_EPI
#undef UB_FAST
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

size_t halo_smem_bytes(const HaloParams& p) {
  const int P = 8 * p.J + 2;
  const int halo_stride = round_up_1k((kTH + 2) * P * p.kc * 2);
  return static_cast<size_t>(p.h_stages) * halo_stride + static_cast<size_t>(p.b_stages) * p.tpb * p.block_n * p.kc * 2 +
         sizeof(HaloBarriers) + 1024;
}

}  // namespace

bool halo_plan(HaloParams& p, int N, int H, int W, int kc, int ncols) {
  static const bool disabled = getenv("UNETB200_NO_HALO") != nullptr;
  if (disabled) return false;
  if (H < kTH || W < 8) return false;   // small grids: several images per tile, handled by igemm
  int block_n = ncols % 256 == 0 ? 256 : ncols;
  // measured: with 256-wide tiles (16x16 level) the per-tap igemm path is as fast or faster (few super-tiles,
  // single-buffered accumulators); the halo path pays off where the GEMM N is small. 256 columns on the 32x32 level
  // and above (conv7a's dgrad: 128 + 128) run as two 128-wide N tiles: the per-tap path streams 48 KB per 512 MMA
  // cycles there, twice what L2 delivers per SM.
  static const int nsplit_min_h = getenv("UNETB200_HALO_NSPLIT_MINH") ? atoi(getenv("UNETB200_HALO_NSPLIT_MINH")) : 32;
  static const int nsplit_max_cols = getenv("UNETB200_HALO_NSPLIT_MAXC") ? atoi(getenv("UNETB200_HALO_NSPLIT_MAXC")) : 256;
  if (ncols % 128 == 0 && ncols >= 256 && ncols <= nsplit_max_cols && H >= nsplit_min_h) block_n = 128;
  if (block_n != 32 && block_n != 64 && block_n != 128) return false;
  int J = W >= 25 ? 4 : (W >= 9 ? 2 : 1);
  // prefer strips that tile the width exactly (e.g. 240 = 15 x 16): full tiles take the lean epilogue
  while (J > 1 && W % (8 * J) != 0 && W % (4 * J) == 0) J >>= 1;
  while (J > 1 && J * block_n > 512) J >>= 1;
  p.N = N; p.H = H; p.W = W;
  p.block_n = block_n;
  p.n_tiles = ncols / block_n;
  p.kc = kc;
  const int btile = block_n * kc * 2;
  for (;; J >>= 1) {
    p.J = J;
    p.acc_double = (2 * J * block_n <= 512) ? 1 : 0;
    p.tpb = (9 * btile <= 36 * 1024) ? 9 : ((3 * btile <= 48 * 1024) ? 3 : 1);
    const int P = 8 * J + 2;
    const int halo_stride = round_up_1k((kTH + 2) * P * kc * 2);
    const int budget = kSmemBudget - 1024 - static_cast<int>(sizeof(HaloBarriers));
    // at least two weight stages; then as many halo stages as fit (the halo load is the long-latency one: a box of
    // up to 612 rows), then the rest to the weight ring
    int hst = (budget - 2 * p.tpb * btile) / halo_stride;
    if (hst > kMaxHaloStages) hst = kMaxHaloStages;
    if (hst >= 2) {
      int stages = (budget - hst * halo_stride) / (p.tpb * btile);
      if (stages > kMaxBStages) stages = kMaxBStages;
      p.h_stages = hst;
      p.b_stages = stages;
      break;
    }
    if (J == 1) return false;
  }
  p.tiles_h = (H + kTH - 1) / kTH;
  p.tiles_w = (W + 8 * p.J - 1) / (8 * p.J);
  return true;
}

int halo_launch(const HaloParams& p_in, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                cudaStream_t stream) {
  UB_REQUIRE(p_in.epi.d2s_cout == 0, "conv_halo: depth-to-space epilogue is not supported here");
  const size_t smem = halo_smem_bytes(p_in);
  UB_REQUIRE(smem <= 227 * 1024, "conv_halo: %zu bytes of shared memory requested", smem);
  static bool attr_set[64] = {false};
  int dev = 0;
  UB_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    UB_CHECK_CUDA(cudaFuncSetAttribute(conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev] = true;
  }
  const int num_tiles = p_in.N * p_in.tiles_h * p_in.tiles_w * p_in.n_tiles;
  UB_REQUIRE(num_tiles > 0, "conv_halo: empty problem");
  HaloParams p = p_in;
  {
    // the lean epilogue covers full super-tiles of one N tile without dropout or tensor masks (see halo_epilogue_fast)
    static const bool no_fast = getenv("UNETB200_NO_FAST_EPI") != nullptr;
    const IgemmEpilogue& e = p.epi;
    const bool aligned = reinterpret_cast<uintptr_t>(e.dst[0]) % 32 == 0 && reinterpret_cast<uintptr_t>(e.dst[1]) % 32 == 0 &&
                         e.dst_coff[0] % 16 == 0 && e.dst_coff[1] % 16 == 0 && e.dst_C[0] % 16 == 0 && e.dst_C[1] % 16 == 0;
    const bool full = aligned && p.H % kTH == 0 && p.W % (8 * p.J) == 0;
    const bool plain = !e.aux[0] && !e.aux[1];
    const bool dropped = e.drop_keep < 1.0f;
    const bool gates_in = e.gate_in[0] || e.gate_in[1];
    const int ncols = p.n_tiles * p.block_n;
    p.fast_epi = 0;
    if (!no_fast && full && plain) {
      if (e.relu && !gates_in && e.split_col >= ncols && p.n_tiles == 1) p.fast_epi = dropped ? 5 : 1;
      else if (!e.relu && !e.bias && !e.gate_out && !dropped) {
        // one destination, or two that meet at an N-tile boundary: the destination is picked per tile
        if (e.split_col >= ncols || e.split_col % p.block_n == 0) p.fast_epi = 2;
        else if (p.n_tiles == 1 && e.split_col == p.block_n / 2 && p.block_n >= 64) p.fast_epi = e.pool2_slot0 ? 4 : 3;
      }
    }
    UB_REQUIRE(!e.pool2_slot0 || p.fast_epi == 4,
               "conv_halo: the fused 2x2 gradient pooling needs full tiles, two equal aligned destinations and no dropout");
  }
  int grid = device_sm_count();
  if (grid > num_tiles) grid = num_tiles;
  if (profiling_enabled()) {
    const double pixels = static_cast<double>(p.N) * p.H * p.W;
    const double kdim = 9.0 * p.kc * (p.src_kblocks[0] + (p.num_src > 1 ? p.src_kblocks[1] : 0));
    const double ncols = static_cast<double>(p.n_tiles) * p.block_n;
    prof_begin(UB_KIND_IGEMM, 2.0 * pixels * kdim * ncols, 2.0 * pixels * (kdim / 9.0 + ncols), stream);
  }
  static const bool timing = getenv("UNETB200_HALO_TIMING") != nullptr;
  if (timing) {
    HaloParams pt = p;
    pt.timing = 1;
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaMemcpyToSymbol(g_halo_timing, z, sizeof(z));
    UB_CHECK_CUDA(launch_pdl(conv_halo_kernel, dim3(grid), dim3(kThreads), smem, stream, pt, a0, a1, b));
    cudaStreamSynchronize(stream);
    cudaMemcpyFromSymbol(z, g_halo_timing, sizeof(z));
    const double c = static_cast<double>(z[7] ? z[7] : 1);
    fprintf(stderr, "[halo timing] N%d %dx%d kc%d bn%d J%d srcs%d tiles%d grid%d | per CTA cycles: mma_total %.0f (wait tmem_empty %.0f, "
            "halo_full %.0f, b_full %.0f) epi_total %.0f (wait tmem_full %.0f) halo_producer wait %.0f\n",
            p.N, p.H, p.W, p.kc, p.block_n, p.J, p.num_src, num_tiles, grid, z[3] / c, z[0] / c, z[1] / c, z[2] / c, z[5] / c,
            z[4] / c, z[6] / c);
    return UB_OK;
  }
  UB_CHECK_CUDA(launch_pdl(conv_halo_kernel, dim3(grid), dim3(kThreads), smem, stream, p, a0, a1, b));
  UB_CHECK_CUDA(cudaGetLastError());
  prof_end(stream);
  return UB_OK;
}

}  // namespace ub
