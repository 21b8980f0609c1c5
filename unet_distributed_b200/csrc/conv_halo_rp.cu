// Row-pair variant of the halo conv (forward / dgrad) for GEMM N = 32 or 64 -- the 128x128 and 64x64 levels.
//
// With both operands in shared memory a 128 x N x 16 MMA costs 32 + N/4 cycles (csrc/mma_rate.cu): the 4 KB A tile
// dominates, so an N = 32 layer runs the tensor core at 40 % at best. Here one accumulator row is an EVEN image row
// h and the accumulator is 2N columns wide: columns [0, N) are the outputs of row h, columns [N, 2N) those of row
// h + 1. The A operand read at row offset o (rows h + o, o = -1..2) then serves both:
//     o = -1 : [W(r=0)        ]  -> row h          (N columns,  left half)
//     o =  0 : [W(r=1) | W(r=0)]  -> rows h | h+1   (2N columns)
//     o =  1 : [W(r=2) | W(r=1)]  -> rows h | h+1   (2N columns)
//     o =  2 : [        W(r=2)]  -> row h + 1      (N columns,  right half)
// i.e. per 256 output pixels, filter column s and 16 input channels: 2 MMAs of N + 2 MMAs of 2N instead of 6 of N
// (N = 32: 176 cycles instead of 240; N = 64: 224 instead of 288). The weight tiles of one filter column sit in
// shared memory in the order r = 2, 1, 0, so "[W(r+1) | W(r)]" is one B descriptor over 2N consecutive rows.
// A rows of one MMA are image rows two apart: the descriptor's SBO is two rows of the halo tile (34 x (8J+2) pixels).
//
// Only the shapes and epilogues the UNet step uses are covered (full 32-row tiles, one N tile, forward with bias +
// ReLU + optional gate bits, dgrad with optional gate bits and one or two destinations); everything else stays on
// conv_halo.cu. Same warp roles: warp 0 halo TMA, warp 1 MMA + TMEM, warp 2 weight TMA, warps 4..11 epilogue.
#include "conv_halo.cuh"

#include <cstdlib>

#include "common.h"
#include "ptx.cuh"

namespace ub {

namespace {

constexpr int kThreads = 384;
constexpr int kTHR = 32;            // image rows per super-tile
constexpr int kHaloRows = kTHR + 2;
constexpr int kSmemBudget = 225 * 1024;

struct __align__(8) RpBarriers {
  uint64_t halo_full[2];
  uint64_t halo_empty[2];
  uint64_t b_full[4];
  uint64_t b_empty[4];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

__host__ __device__ inline int round_up_1k(int x) { return (x + 1023) & ~1023; }

__device__ __forceinline__ uint64_t desc_pair(uint32_t lo, uint32_t hi) {
  return (static_cast<uint64_t>(hi) << 32) | lo;
}

__device__ __forceinline__ uint32_t pack_bf16x2_relu(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

__host__ __device__ inline int rp_tpb(int block_n, int kc) { return 9 * block_n * kc * 2 <= 18 * 1024 ? 9 : 3; }

__device__ __forceinline__ void decode_tile(const HaloParams& p, int tile, int& n, int& h0, int& w0) {
  const int tw_i = tile % p.tiles_w;
  int t = tile / p.tiles_w;
  const int th_i = t % p.tiles_h;
  n = t / p.tiles_h;
  h0 = th_i * kTHR;
  w0 = tw_i * 8 * p.J;
}

// ------------------------------------------------------------------------------------------------ MMA issuer
template <int J, int KC, int BN>
__device__ __forceinline__ void rp_mma_role(const HaloParams& p, RpBarriers* bars, uint32_t halo_addr,
                                            uint32_t halo_stride, uint32_t b_addr, uint32_t tmem_base, int num_tiles) {
  constexpr int P = 8 * J + 2;
  constexpr uint32_t ROW16 = KC * 2 / 16;
  constexpr int KSTEPS = KC / 16;
  constexpr int BTILE = BN * KC * 2;
  constexpr int TPB = (9 * BTILE <= 18 * 1024) ? 9 : 3;   // == rp_tpb(): whole filter columns per weight stage
  constexpr int SGROUPS = 9 / TPB;                        // weight stages per k block
  constexpr int S_PER = TPB / 3;                          // filter columns per weight stage
  constexpr uint32_t BSTAGE = TPB * BTILE;
  constexpr uint32_t SW = KC == 64 ? UMMA_SW_128B : UMMA_SW_64B;
  const uint32_t idesc1 = umma_idesc_bf16(128, BN, 0, 0);
  const uint32_t idesc2 = umma_idesc_bf16(128, 2 * BN, 0, 0);
  const uint32_t a_hi = static_cast<uint32_t>(umma_smem_desc(0, 0, 2 * P * KC * 2, SW) >> 32);   // SBO = two image rows
  const uint32_t b_hi = static_cast<uint32_t>(umma_smem_desc(0, 0, 8 * KC * 2, SW) >> 32);
  const int total_kb = p.src_kblocks[0] + (p.num_src > 1 ? p.src_kblocks[1] : 0);
  int hs = 0, bs = 0, it = 0;
  uint32_t hphase = 0, bphase = 0;
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const int a = p.acc_double ? (it & 1) : 0;
    const uint32_t eparity = p.acc_double ? (((it >> 1) & 1) ^ 1u) : ((it & 1) ^ 1u);
    mbar_wait(&bars->tmem_empty[a], eparity);
    tc_fence_after();
    const uint32_t d0 = tmem_base + static_cast<uint32_t>(a * J * 2 * BN);
    for (int kb = 0; kb < total_kb; ++kb) {
      mbar_wait(&bars->halo_full[hs], hphase);
      const uint32_t a_lo = (halo_addr + static_cast<uint32_t>(hs) * halo_stride) >> 4;
      const uint32_t fresh = kb == 0 ? 0u : 1u;
#pragma unroll
      for (int sg = 0; sg < SGROUPS; ++sg) {
        mbar_wait(&bars->b_full[bs], bphase);
        tc_fence_after();
        const uint32_t b_lo = (b_addr + static_cast<uint32_t>(bs) * BSTAGE) >> 4;
        if (elect_one()) {
#pragma unroll
          for (int sp = 0; sp < S_PER; ++sp) {
            const int s = sg * S_PER + sp;
            // weight tiles of this filter column: r = 2, 1, 0 at tile slots 3 sp + 0, 1, 2
            const uint32_t w2 = b_lo + static_cast<uint32_t>((3 * sp + 0) * (BTILE >> 4));
            const uint32_t w1 = b_lo + static_cast<uint32_t>((3 * sp + 1) * (BTILE >> 4));
            const uint32_t w0 = b_lo + static_cast<uint32_t>((3 * sp + 2) * (BTILE >> 4));
#pragma unroll
            for (int j = 0; j < J; ++j) {
              const uint32_t dj = d0 + static_cast<uint32_t>(j * 2 * BN);
#pragma unroll
              for (int k = 0; k < KSTEPS; ++k) {
                // halo tile row of (even row 2t, offset o) = 2t + o + 1
                const uint32_t ao = a_lo + static_cast<uint32_t>((s + 8 * j) * ROW16 + 2 * k);
                const uint32_t first = (s == 0 && k == 0) ? fresh : 1u;
                // the two half-width MMAs come first: on the first k block they initialise the two halves
                umma_f16(dj, desc_pair(ao + 0 * P * ROW16, a_hi), desc_pair(w0 + 2 * k, b_hi), idesc1, first);
                umma_f16(dj + BN, desc_pair(ao + 3 * P * ROW16, a_hi), desc_pair(w2 + 2 * k, b_hi), idesc1, first);
                umma_f16(dj, desc_pair(ao + 1 * P * ROW16, a_hi), desc_pair(w1 + 2 * k, b_hi), idesc2, 1u);
                umma_f16(dj, desc_pair(ao + 2 * P * ROW16, a_hi), desc_pair(w2 + 2 * k, b_hi), idesc2, 1u);
              }
            }
          }
          umma_commit(&bars->b_empty[bs]);
          if (sg == SGROUPS - 1) umma_commit(&bars->halo_empty[hs]);
        }
        __syncwarp();
        if (++bs == p.b_stages) { bs = 0; bphase ^= 1u; }
      }
      if (++hs == p.h_stages) { hs = 0; hphase ^= 1u; }
    }
    if (elect_one()) umma_commit(&bars->tmem_full[a]);
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------------ epilogue
// Accumulator row m = 8 t + column: pixel (h0 + 2t, w0 + 8j + column); accumulator columns [0, BN) belong to that
// pixel, [BN, 2BN) to the pixel one image row below. Items are 16-column chunks; the two warps of a TMEM lane quarter
// take alternate chunks. SPLIT: channels [BN/2, BN) go to destination 1. POOL (with SPLIT): destination 0 is the coarse
// tensor and receives the 2x2 block sums -- the two rows of a block are the two halves of this lane's accumulator row,
// the two columns are lanes l and l ^ 1; the even-column lane stores.
// FPOOL (forward): MaxPool2D fused -- the max of the block's two rows is taken between the two halves of the lane's
// accumulator row (packed bf16x2 max), the two columns between lanes l and l ^ 1; the even-column lane stores.
template <int J, int BN, bool FWD, bool SPLIT, bool POOL = false, bool FPOOL = false>
__device__ __forceinline__ void rp_epilogue(const HaloParams& p, RpBarriers* bars, uint32_t tmem_base, int warp, int lane,
                                            int num_tiles) {
  static_assert(!SPLIT || (BN == 64 && !FWD), "split destinations: dgrad into two 32-channel tensors");
  static_assert(!POOL || SPLIT, "2x2 pooling of destination 0 is a split-dgrad feature");
  static_assert(!FPOOL || FWD, "max pooling belongs to the forward epilogue");
  constexpr int kCl = 2 * BN / 32;          // 16-column chunks per warp and strip
  constexpr int kItems = J * kCl;
  constexpr int kBiasCl = BN / 32;
  const IgemmEpilogue& e = p.epi;
  const int quarter = warp & 3;
  const int half = (warp - 4) >> 2;
  const int m = quarter * 32 + lane;
  const int th = m >> 3, twl = m & 7;
  const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(half * 16);
  float bias_r[FWD ? kBiasCl : 1][16];
  if (FWD) {
#pragma unroll
    for (int cl = 0; cl < kBiasCl; ++cl) {
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
  constexpr int kSlots = SPLIT ? 2 : 1;
  uint32_t strip_b[kSlots], row_b[kSlots], gstrip[kSlots], grow[kSlots];
#pragma unroll
  for (int sl = 0; sl < kSlots; ++sl) {
    strip_b[sl] = static_cast<uint32_t>(16 * e.dst_C[sl]);                 // 8 pixels, bytes
    row_b[sl] = static_cast<uint32_t>(2 * p.W * e.dst_C[sl]);             // one image row, bytes
    const int gc = (FWD || (POOL && sl == 0)) ? e.dst_C[sl] : e.aux_C[sl];
    gstrip[sl] = static_cast<uint32_t>(gc >> 1);                          // 8 pixels, gate words
    grow[sl] = static_cast<uint32_t>((p.W * gc) >> 4);                    // one image row, gate words
  }
  int it = 0;
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const int a = p.acc_double ? (it & 1) : 0;
    const uint32_t fparity = p.acc_double ? ((it >> 1) & 1) : (it & 1);
    int n, h0, w0;
    decode_tile(p, tile, n, h0, w0);
    const size_t pix0 = (static_cast<size_t>(n) * p.H + (h0 + 2 * th)) * p.W + (w0 + twl);
    const bool row_ok = (h0 + 2 * th) < p.H;   // H is even: both rows of the pair are inside or outside together
    char* dst_b[kSlots];
    const unsigned short* gin_b[kSlots];
    unsigned short* gout_b = nullptr;
#pragma unroll
    for (int sl = 0; sl < kSlots; ++sl) {
      dst_b[sl] = reinterpret_cast<char*>(e.dst[sl]) + (pix0 * e.dst_C[sl] + e.dst_coff[sl] + half * 16) * 2;
      gin_b[sl] = nullptr;
      if (!FWD && e.gate_in[sl]) gin_b[sl] = e.gate_in[sl] + ((pix0 * e.aux_C[sl] + e.aux_coff[sl]) >> 4) + half;
    }
    if (FWD && e.gate_out) gout_b = e.gate_out + ((pix0 * e.dst_C[0] + e.dst_coff[0]) >> 4) + half;
    const bool pool_lane = (lane & 1) == 0;   // even column (every accumulator row is an even image row)
    char* mp_b = nullptr;
    if (FPOOL) {
      const size_t cpix0 = (static_cast<size_t>(n) * (p.H >> 1) + ((h0 >> 1) + th)) * (p.W >> 1) + ((w0 + twl) >> 1);
      mp_b = reinterpret_cast<char*>(e.maxpool_out) + (cpix0 * e.dst_C[0] + half * 16) * 2;
    }
    if (POOL) {   // destination 0 on the coarse grid
      const size_t cpix0 = (static_cast<size_t>(n) * (p.H >> 1) + ((h0 >> 1) + th)) * (p.W >> 1) + ((w0 + twl) >> 1);
      dst_b[0] = reinterpret_cast<char*>(e.dst[0]) + (cpix0 * e.dst_C[0] + e.dst_coff[0] + half * 16) * 2;
      gin_b[0] = nullptr;
      if (e.gate_in[0] && pool_lane) gin_b[0] = e.gate_in[0] + ((cpix0 * e.dst_C[0] + e.dst_coff[0]) >> 4) + half;
    }
    uint32_t gb[kItems];
    if (!FWD) {
#pragma unroll
      for (int k = 0; k < kItems; ++k) {
        const int j = k / kCl, cl = k % kCl;
        const int par = (32 * cl >= BN) ? 1 : 0;            // 0: row h, 1: row h + 1
        const int chl = 32 * cl - par * BN;                 // first channel of the chunk pair (before + half * 16)
        const int sl = (SPLIT && chl >= BN / 2) ? 1 : 0;
        gb[k] = 0xFFFFu;
        if (POOL && sl == 0) {   // one coarse gate word per block, fetched with the lower-row item
          if (par == 1 && gin_b[0] && row_ok) gb[k] = __ldg(gin_b[0] + j * (gstrip[0] >> 1) + (chl >> 4));
        } else if (gin_b[sl] && row_ok) {
          gb[k] = __ldg(gin_b[sl] + j * gstrip[sl] + par * grow[sl] + ((chl - sl * (BN / 2)) >> 4));
        }
      }
    }
    mbar_wait(&bars->tmem_full[a], fparity);
    tc_fence_after();
    const uint32_t t_acc = t_lane + static_cast<uint32_t>(a * J * 2 * BN);
    float vs[POOL ? 16 : 1];   // upper-row values of the block waiting for the lower row
    uint32_t wsave[FPOOL ? kBiasCl : 1][8];   // forward pooling: packed upper-row outputs per channel chunk
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
      const int j = k / kCl, cl = k % kCl;
      const int par = (32 * cl >= BN) ? 1 : 0;
      const int chl = 32 * cl - par * BN;
      const int sl = (SPLIT && chl >= BN / 2) ? 1 : 0;
      uint32_t r[16];
      tmem_ld_32x32b_x16(t_acc + static_cast<uint32_t>(j * 2 * BN + 32 * cl), r);
      tmem_ld_wait();
      uint32_t w[8];
      if (FWD) {
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]) + bias_r[FWD ? chl / 32 : 0][i];
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = pack_bf16x2_relu(v[2 * i], v[2 * i + 1]);
        if (FPOOL) {
          if (par == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) wsave[FPOOL ? chl / 32 : 0][i] = w[i];
          } else {
            U32x8 mx;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              __nv_bfloat162 a2 = *reinterpret_cast<__nv_bfloat162*>(&w[i]);
              __nv_bfloat162 b2 = *reinterpret_cast<__nv_bfloat162*>(&wsave[FPOOL ? chl / 32 : 0][i]);
              __nv_bfloat162 m2 = __hmax2(a2, b2);
              uint32_t mw = *reinterpret_cast<uint32_t*>(&m2);
              const uint32_t ow = __shfl_xor_sync(0xffffffffu, mw, 1);
              __nv_bfloat162 o2 = *reinterpret_cast<const __nv_bfloat162*>(&ow);
              m2 = __hmax2(m2, o2);
              mx.v[i] = *reinterpret_cast<uint32_t*>(&m2);
            }
            if (pool_lane && row_ok) st_global_v8(mp_b + j * (strip_b[0] >> 1) + chl * 2, mx);
          }
        }
        if (gout_b && row_ok) {
          uint32_t acc = 0;
#pragma unroll
          for (int i = 0; i < 8; ++i) acc |= ((w[i] + 0x7FFF7FFFu) >> (15 - 2 * i)) & (0x00010001u << (2 * i));
          gout_b[j * gstrip[0] + par * grow[0] + (chl >> 4)] = static_cast<unsigned short>((acc & 0xFFFFu) | (acc >> 15));
        }
      } else {
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
        if (POOL && sl == 0) {
          if (par == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) vs[POOL ? i : 0] = v[i];
            continue;
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            v[i] += vs[POOL ? i : 0];
            v[i] += __shfl_xor_sync(0xffffffffu, v[i], 1);
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
      if (POOL && sl == 0) {   // coarse strips are 4 pixels apart
        if (pool_lane && row_ok) st_global_v8(dst_b[0] + j * (strip_b[0] >> 1) + chl * 2, o);
      } else {
        if (row_ok) st_global_v8(dst_b[sl] + j * strip_b[sl] + par * row_b[sl] + (chl - sl * (BN / 2)) * 2, o);
      }
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&bars->tmem_empty[a]);
  }
}

__global__ void __launch_bounds__(kThreads, 1)
conv_halo_rp_kernel(const __grid_constant__ HaloParams p, const __grid_constant__ CUtensorMap tmA0,
                    const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB) {
  pdl_trigger();     // the next kernel of the stream may take SMs as this grid's CTAs leave them
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

  const int P = 8 * p.J + 2;
  const int rowbytes = p.kc * 2;
  const int halo_bytes = kHaloRows * P * rowbytes;       // exactly one TMA box
  const int halo_stride = round_up_1k(halo_bytes);
  const int btile_bytes = p.block_n * rowbytes;
  const int bstage_bytes = p.tpb * btile_bytes;
  uint8_t* halo_base = smem;
  uint8_t* b_base = smem + p.h_stages * halo_stride;
  RpBarriers* bars = reinterpret_cast<RpBarriers*>(b_base + p.b_stages * bstage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.N * p.tiles_h * p.tiles_w;
  const int acc_sets = p.acc_double ? 2 : 1;
  uint32_t tmem_cols = 32;
  while (tmem_cols < static_cast<uint32_t>(acc_sets * p.J * 2 * p.block_n)) tmem_cols <<= 1;

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
    // ------------------------------------------------------------ halo producer: rows h0-1 .. h0+32
    if (lane == 0) {
      int hs = 0;
      uint32_t hphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int n, h0, w0;
        decode_tile(p, tile, n, h0, w0);
        for (int src = 0; src < p.num_src; ++src) {
          const CUtensorMap* tmA = src == 0 ? &tmA0 : &tmA1;
          for (int cb = 0; cb < p.src_kblocks[src]; ++cb) {
            mbar_wait(&bars->halo_empty[hs], hphase ^ 1u);
            mbar_arrive_expect_tx(&bars->halo_full[hs], static_cast<uint32_t>(halo_bytes));
            tma_load_4d(halo_base + hs * halo_stride, tmA, &bars->halo_full[hs], cb * p.kc, w0 - 1, h0 - 1, n);
            if (++hs == p.h_stages) { hs = 0; hphase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 2) {
    // ------------------------------------------------------------ weight producer: filter columns, rows 2, 1, 0
    if (lane == 0) {
      int bs = 0;
      uint32_t bphase = 0;
      const int sgroups = 9 / p.tpb;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        for (int src = 0; src < p.num_src; ++src) {
          for (int cb = 0; cb < p.src_kblocks[src]; ++cb) {
            for (int sg = 0; sg < sgroups; ++sg) {
              mbar_wait(&bars->b_empty[bs], bphase ^ 1u);
              mbar_arrive_expect_tx(&bars->b_full[bs], static_cast<uint32_t>(bstage_bytes));
              uint8_t* dst = b_base + bs * bstage_bytes;
              for (int t = 0; t < p.tpb; ++t) {
                const int s = sg * (p.tpb / 3) + t / 3;
                const int r = 2 - (t % 3);
                tma_load_3d(dst + t * btile_bytes, &tmB, &bars->b_full[bs], p.src_ci_off[src] + cb * p.kc, r * 3 + s, 0);
              }
              if (++bs == p.b_stages) { bs = 0; bphase ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t ha = smem_u32(halo_base), ba = smem_u32(b_base);
    const uint32_t hst = static_cast<uint32_t>(halo_stride);
#define UB_RP_MMA(JJ, KK, NN) rp_mma_role<JJ, KK, NN>(p, bars, ha, hst, ba, tmem_base, num_tiles)
#define UB_RP_MMA_J(KK, NN)                      \
    do {                                         \
      if (p.J == 4) UB_RP_MMA(4, KK, NN);        \
      else if (p.J == 2) UB_RP_MMA(2, KK, NN);   \
      else UB_RP_MMA(1, KK, NN);                 \
    } while (0)
    if (p.kc == 64) {
      if (p.block_n == 32) UB_RP_MMA_J(64, 32);
      else UB_RP_MMA_J(64, 64);
    } else {
      if (p.block_n == 32) UB_RP_MMA_J(32, 32);
      else UB_RP_MMA_J(32, 64);
    }
#undef UB_RP_MMA_J
#undef UB_RP_MMA
  } else if (warp >= 4) {
#define UB_RP_EPI(JJ, NN)                                                                               \
    do {                                                                                                \
      if (p.fast_epi == 1) rp_epilogue<JJ, NN, true, false>(p, bars, tmem_base, warp, lane, num_tiles);  \
      else if (p.fast_epi == 5) rp_epilogue<JJ, NN, true, false, false, true>(p, bars, tmem_base, warp, lane, num_tiles); \
      else if (p.fast_epi == 2) rp_epilogue<JJ, NN, false, false>(p, bars, tmem_base, warp, lane, num_tiles); \
      else if (p.fast_epi == 3) rp_epilogue<JJ, 64, false, true>(p, bars, tmem_base, warp, lane, num_tiles); \
      else rp_epilogue<JJ, 64, false, true, true>(p, bars, tmem_base, warp, lane, num_tiles);            \
    } while (0)
#define UB_RP_EPI_J(NN)                      \
    do {                                     \
      if (p.J == 4) UB_RP_EPI(4, NN);        \
      else if (p.J == 2) UB_RP_EPI(2, NN);   \
      else UB_RP_EPI(1, NN);                 \
    } while (0)
    if (p.block_n == 32) UB_RP_EPI_J(32);
    else UB_RP_EPI_J(64);
#undef UB_RP_EPI_J
#undef UB_RP_EPI
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

size_t rp_smem_bytes(const HaloParams& p) {
  const int P = 8 * p.J + 2;
  const int halo_stride = round_up_1k(kHaloRows * P * p.kc * 2);
  return static_cast<size_t>(p.h_stages) * halo_stride + static_cast<size_t>(p.b_stages) * p.tpb * p.block_n * p.kc * 2 +
         sizeof(RpBarriers) + 1024;
}

}  // namespace

// Row-pair plan for (N, H, W, kc, ncols); false when the shape is not covered. The caller still has to check that the
// epilogue it needs is one of the lean ones (halo_rp_epilogue_ok).
bool halo_rp_plan(HaloParams& p, int N, int H, int W, int kc, int ncols) {
  static const bool disabled = getenv("UNETB200_NO_HALO") != nullptr || getenv("UNETB200_NO_RP") != nullptr;
  if (disabled) return false;
  if (ncols != 32 && ncols != 64) return false;
  if (H < kTHR || H % 2 != 0 || W % 8 != 0) return false;   // the last row tile may be partial (rows are masked)
  int J = ncols == 32 ? 4 : 2;                 // TMEM: two sets of J accumulators of 2 * ncols columns
  if (kc == 64 && J > 2) J = 2;                // shared memory: a 34-row halo tile of 128-byte pixels
  while (J > 1 && W % (8 * J) != 0) J >>= 1;
  p.N = N; p.H = H; p.W = W;
  p.block_n = ncols;
  p.n_tiles = 1;
  p.kc = kc;
  p.J = J;
  p.acc_double = (2 * J * 2 * ncols <= 512) ? 1 : 0;
  p.tpb = rp_tpb(ncols, kc);
  p.h_stages = 2;
  const int P = 8 * J + 2;
  const int halo_stride = round_up_1k(kHaloRows * P * kc * 2);
  const int budget = kSmemBudget - 1024 - static_cast<int>(sizeof(RpBarriers)) - 2 * halo_stride;
  int stages = budget / (p.tpb * ncols * kc * 2);
  if (stages > 4) stages = 4;
  if (stages < 2) return false;
  p.b_stages = stages;
  p.tiles_h = (H + kTHR - 1) / kTHR;
  p.tiles_w = W / (8 * J);
  p.rp = 1;
  return true;
}

// fast_epi code (1 forward, 2 dgrad, 3 dgrad split, 4 dgrad split + 2x2 pooling of destination 0, 5 forward + fused
// MaxPool2D) if the lean row-pair epilogue covers this epilogue, else 0.
int halo_rp_epilogue_code(const HaloParams& p) {
  const IgemmEpilogue& e = p.epi;
  const bool aligned = reinterpret_cast<uintptr_t>(e.dst[0]) % 32 == 0 && reinterpret_cast<uintptr_t>(e.dst[1]) % 32 == 0 &&
                       e.dst_coff[0] % 16 == 0 && e.dst_coff[1] % 16 == 0 && e.dst_C[0] % 16 == 0 && e.dst_C[1] % 16 == 0;
  if (!aligned || e.aux[0] || e.aux[1] || e.drop_keep < 1.0f || e.d2s_cout) return 0;
  if (e.pool2_slot0 && !(p.block_n == 64 && e.split_col == 32 && !e.relu && !e.bias && !e.gate_out)) return 0;
  const bool gates_in = e.gate_in[0] || e.gate_in[1];
  if (e.relu && !gates_in && e.split_col >= p.block_n) {
    if (!e.maxpool_out) return 1;
    return (reinterpret_cast<uintptr_t>(e.maxpool_out) % 32 == 0 && e.dst_coff[0] == 0) ? 5 : 0;
  }
  if (e.maxpool_out) return 0;
  if (!e.relu && !e.bias && !e.gate_out) {
    if (e.split_col >= p.block_n) return 2;
    if (p.block_n == 64 && e.split_col == 32) return e.pool2_slot0 ? 4 : 3;
  }
  return 0;
}

int halo_rp_launch(const HaloParams& p, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                   cudaStream_t stream) {
  UB_REQUIRE(p.rp == 1 && p.fast_epi >= 1 && p.fast_epi <= 5, "conv_halo_rp: not a row-pair plan");
  const size_t smem = rp_smem_bytes(p);
  UB_REQUIRE(smem <= 227 * 1024, "conv_halo_rp: %zu bytes of shared memory requested", smem);
  static bool attr_set[64] = {false};
  int dev = 0;
  UB_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    UB_CHECK_CUDA(cudaFuncSetAttribute(conv_halo_rp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev] = true;
  }
  const int num_tiles = p.N * p.tiles_h * p.tiles_w;
  UB_REQUIRE(num_tiles > 0, "conv_halo_rp: empty problem");
  int grid = device_sm_count();
  if (grid > num_tiles) grid = num_tiles;
  if (profiling_enabled()) {
    const double pixels = static_cast<double>(p.N) * p.H * p.W;
    const double kdim = 9.0 * p.kc * (p.src_kblocks[0] + (p.num_src > 1 ? p.src_kblocks[1] : 0));
    prof_begin(UB_KIND_IGEMM, 2.0 * pixels * kdim * p.block_n, 2.0 * pixels * (kdim / 9.0 + p.block_n), stream);
  }
  UB_CHECK_CUDA(launch_pdl(conv_halo_rp_kernel, dim3(grid), dim3(kThreads), smem, stream, p, a0, a1, b));
  UB_CHECK_CUDA(cudaGetLastError());
  prof_end(stream);
  return UB_OK;
}

}  // namespace ub
