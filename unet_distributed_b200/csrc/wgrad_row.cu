// Weight gradient of a 3x3 conv for wide layers (64+ channels on both sides): one CTA accumulates the THREE taps of one
// filter row for a 128 x 128 (Cout x Cin) tile.
//
// wgrad.cu gives every tap its own CTA, so each 64-pixel reduction block costs 32 KB of operand traffic for 256
// tensor-core cycles (128 B/cycle/SM -- about three times what L2 can deliver; measured 350-450 TFLOP/s). Here the
// gradient tile (dZ, 64 pixels x 128 channels) is loaded once for three taps, and the three shifted input tiles are
// three views of ONE strip of 8 x 10 pixels: tap column s is the strip shifted by s pixel rows in shared memory
// (descriptor start + s*128 B), image rows are 10 smem rows apart (the descriptor's SBO). 36 KB per 768 cycles.
#include "wgrad.cuh"

#include <cstdlib>

#include "common.h"
#include "ptx.cuh"

namespace ub {

namespace {

constexpr int kThreads = 192;                 // warp0 TMA, warp1 MMA + TMEM, warps 2..5 epilogue
constexpr int kMaxStages = 6;
constexpr int kZBytes = 2 * 64 * 128;         // two 64-channel atoms x 64 pixels
constexpr int kXAtomBytes = 8 * 10 * 128;     // one 64-channel atom x (8 rows x 10 columns)
constexpr int kStageBytes = kZBytes + 2 * kXAtomBytes;   // 36,864
constexpr int kBiasScratch = 4096;            // [128 threads][8] fp32 column sums of dZ

struct __align__(8) RowBarriers {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(kThreads, 1)
wgrad_row_kernel(const __grid_constant__ WgradRowParams p, const __grid_constant__ CUtensorMap tmZ,
                 const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1) {
  pdl_trigger();     // the next kernel of the stream may take SMs as this grid's CTAs leave them
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  float* bias_scratch = reinterpret_cast<float*>(smem + p.stages * kStageBytes);
  RowBarriers* bars = reinterpret_cast<RowBarriers*>(smem + p.stages * kStageBytes + kBiasScratch);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int total_nt = p.x_ntiles[0] + (p.num_src > 1 ? p.x_ntiles[1] : 0);
  int bid = blockIdx.x;
  const int ntg = bid % total_nt;
  bid /= total_nt;
  const int mt = bid % p.m_tiles;
  bid /= p.m_tiles;
  const int r = bid % 3;
  const int split = bid / 3;
  const int src = (ntg >= p.x_ntiles[0]) ? 1 : 0;
  const int nt = src ? ntg - p.x_ntiles[0] : ntg;
  const int block_n = p.x_block_n[src];
  const int m0 = mt * 128, c0 = nt * block_n;

  const int total_pb = p.N * p.blocks_h * p.blocks_w;
  const int per_split = (total_pb + p.splits - 1) / p.splits;
  const int pb_begin = split * per_split;
  const int pb_end = (pb_begin + per_split < total_pb) ? pb_begin + per_split : total_pb;
  const int nkb = pb_end > pb_begin ? pb_end - pb_begin : 0;

  int m_atoms = (p.z_C - m0 + 63) / 64;
  if (m_atoms > 2) m_atoms = 2;
  const int n_atoms = block_n / 64;
  const uint32_t tmem_cols = 512;   // three accumulators of up to 128 columns
  // one CTA per (Cout tile, pixel split) also sums its dZ tiles over pixels: the bias gradient
  const bool do_bias = p.bias_out != nullptr && ntg == 0 && r == 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], do_bias ? 5 : 1);   // MMA commit (+ the four column-summing warps)
    }
    mbar_init(&bars->tmem_full, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tmZ);
    tma_prefetch_desc(src ? &tmX1 : &tmX0);
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
    if (lane == 0) {
      const CUtensorMap* tmX = src ? &tmX1 : &tmX0;
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx = static_cast<uint32_t>(m_atoms * (kZBytes / 2) + n_atoms * kXAtomBytes);
      for (int pb = pb_begin; pb < pb_end; ++pb) {
        int t = pb;
        const int bw = t % p.blocks_w;
        t /= p.blocks_w;
        const int bh = t % p.blocks_h;
        const int n = t / p.blocks_h;
        const int h0 = bh * 8, w0 = bw * 8;
        mbar_wait(&bars->empty[stage], phase ^ 1u);
        uint8_t* sz = smem + stage * kStageBytes;
        uint8_t* sx = sz + kZBytes;
        mbar_arrive_expect_tx(&bars->full[stage], tx);
        for (int a = 0; a < m_atoms; ++a) {
          tma_load_4d(sz + a * (kZBytes / 2), &tmZ, &bars->full[stage], m0 + a * 64, w0, h0, n);
        }
        for (int a = 0; a < n_atoms; ++a) {
          tma_load_4d(sx + a * kXAtomBytes, tmX, &bars->full[stage], c0 + a * 64, w0 - 1, h0 + r - 1, n);
        }
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    if (nkb > 0) {
      const uint32_t idesc = umma_idesc_bf16(128, static_cast<uint32_t>(block_n), 1, 1);
      // A (dZ): 64-channel atoms 8 KB apart, 8-pixel groups 1 KB apart. B (X strip): atoms 10 KB apart, image rows
      // (8-pixel groups) 10 smem rows apart.
      const uint64_t a_hi = umma_smem_desc(0, kZBytes / 2, 1024, UMMA_SW_128B);
      const uint64_t b_hi = umma_smem_desc(0, kXAtomBytes, 1280, UMMA_SW_128B);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&bars->full[stage], phase);
        tc_fence_after();
        const uint32_t sz = smem_u32(smem + stage * kStageBytes);
        const uint64_t adesc = a_hi + (sz >> 4);
        const uint64_t bdesc = b_hi + ((sz + kZBytes) >> 4);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {          // 16 pixels = two image rows per MMA
#pragma unroll
            for (int s = 0; s < 3; ++s) {        // tap column = strip shifted by s pixel rows
              umma_f16(tmem_base + static_cast<uint32_t>(s * block_n), adesc + static_cast<uint64_t>(k * (2048 >> 4)),
                       bdesc + static_cast<uint64_t>(k * (2560 >> 4) + s * (128 >> 4)), idesc, (kb | k) ? 1u : 0u);
            }
          }
          umma_commit(&bars->empty[stage]);
          if (kb == nkb - 1) umma_commit(&bars->tmem_full);
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else {
    if (do_bias) {
      // While the tensor core works on a stage, warps 2..5 sum its dZ tile (2 atoms x 64 pixels x 128 B, 128-byte
      // swizzle: chunk ^= pixel & 7) over pixels. Thread t: atom t/64, 16-byte chunk t%8, pixels (t/8)%8 + 8k.
      const int t = threadIdx.x - 64;
      const int atom = t >> 6, q = t & 7, rg = (t >> 3) & 7;
      float bsum[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) bsum[j] = 0.0f;
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&bars->full[stage], phase);
        if (atom < m_atoms) {
          const uint8_t* zt = smem + stage * kStageBytes + atom * (kZBytes / 2);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int px = rg + 8 * k;
            const uint4 v = *reinterpret_cast<const uint4*>(zt + px * 128 + ((q ^ (px & 7)) << 4));
            bsum[0] += bf16_lo(v.x); bsum[1] += bf16_hi(v.x); bsum[2] += bf16_lo(v.y); bsum[3] += bf16_hi(v.y);
            bsum[4] += bf16_lo(v.z); bsum[5] += bf16_hi(v.z); bsum[6] += bf16_lo(v.w); bsum[7] += bf16_hi(v.w);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->empty[stage]);
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) bias_scratch[t * 8 + j] = bsum[j];
      asm volatile("bar.sync 1, 128;" ::: "memory");
      // channel c of the tile: atom c/64, chunk (c%64)/8, element c%8; its eight row-group threads in a fixed order
      const int c = t;
      if (m0 + c < p.z_C) {
        const int u0 = (c >> 6) * 64 + ((c & 63) >> 3);
        float acc = 0.0f;
#pragma unroll
        for (int g = 0; g < 8; ++g) acc += bias_scratch[(u0 + 8 * g) * 8 + (c & 7)];
        p.bias_out[split * p.bias_split_stride + m0 + c] = acc;
      }
    }
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int co = m0 + row;
    const bool valid = co < p.z_C;
    if (nkb > 0) {
      mbar_wait(&bars->tmem_full, 0);
      tc_fence_after();
    }
    for (int s = 0; s < 3; ++s) {
      float* out = p.out + split * p.out_split_stride + (r * 3 + s) * p.out_tap_stride + co +
                   static_cast<long long>(p.x_col_off[src] + c0) * p.out_n_stride;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                             static_cast<uint32_t>(s * block_n);
      for (int c = 0; c < block_n; c += 16) {
        uint32_t v[16];
        if (nkb > 0) {
          tmem_ld_32x32b_x16(t_row + static_cast<uint32_t>(c), v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = 0u;
        }
        if (valid) {
#pragma unroll
          for (int i = 0; i < 16; ++i) out[static_cast<long long>(c + i) * p.out_n_stride] = __uint_as_float(v[i]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

}  // namespace

bool wgrad_row_supported(int H, int W, int Cz, int C0, int C1) {
  static const bool disabled = getenv("UNETB200_NO_ROW") != nullptr;
  if (disabled) return false;
  if (Cz % 64 != 0 || C0 % 64 != 0 || (C1 && C1 % 64 != 0)) return false;
  if (C0 > 64 && C0 % 128 != 0) return false;
  if (C1 > 64 && C1 % 128 != 0) return false;
  return H >= 8 && W >= 8;
}

int wgrad_row_plan(WgradRowParams& p, int N, int H, int W, int Cz, int C0, int C1) {
  p.N = N; p.H = H; p.W = W;
  p.blocks_h = (H + 7) / 8;
  p.blocks_w = (W + 7) / 8;
  p.z_C = Cz;
  p.m_tiles = (Cz + 127) / 128;
  p.num_src = C1 ? 2 : 1;
  const int Cs[2] = {C0, C1};
  int off = 0, total_nt = 0;
  for (int s = 0; s < 2; ++s) {
    p.x_block_n[s] = Cs[s] >= 128 ? 128 : 64;
    p.x_ntiles[s] = Cs[s] ? Cs[s] / p.x_block_n[s] : 0;
    p.x_col_off[s] = off;
    off += Cs[s];
    total_nt += p.x_ntiles[s];
  }
  const int total_pb = N * p.blocks_h * p.blocks_w;
  const int base = 3 * p.m_tiles * total_nt;
  // one CTA per SM (216 KB of shared memory): a second wave pays prologue, epilogue and a second set of partials
  // (measured: 1 wave 1.01 ms, 2 waves 1.13 ms, 3 waves 1.16 ms of wgrad time per step)
  static const int waves = getenv("UNETB200_ROW_WAVES") ? atoi(getenv("UNETB200_ROW_WAVES")) : 1;
  int splits = waves == 1 ? device_sm_count() / base : (waves * device_sm_count() + base - 1) / base;
  const int max_splits = total_pb / 8 > 1 ? total_pb / 8 : 1;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  const int per = (total_pb + splits - 1) / splits;
  p.splits = (total_pb + per - 1) / per;
  p.stages = kMaxStages;
  return UB_OK;
}

int wgrad_row_launch(const WgradRowParams& p, const CUtensorMap& z, const CUtensorMap& x0, const CUtensorMap& x1,
                     cudaStream_t stream) {
  const size_t smem = static_cast<size_t>(p.stages) * kStageBytes + kBiasScratch + sizeof(RowBarriers) + 1024;
  static bool attr_set[64] = {false};
  int dev = 0;
  UB_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    UB_CHECK_CUDA(cudaFuncSetAttribute(wgrad_row_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev] = true;
  }
  const int total_nt = p.x_ntiles[0] + (p.num_src > 1 ? p.x_ntiles[1] : 0);
  const long long grid = static_cast<long long>(p.splits) * 3 * p.m_tiles * total_nt;
  UB_REQUIRE(grid > 0 && grid < (1ll << 30), "wgrad_row: grid %lld", grid);
  if (profiling_enabled()) {
    const double pixels = static_cast<double>(p.N) * p.H * p.W;
    const double cin = static_cast<double>(p.x_ntiles[0]) * p.x_block_n[0] +
                       (p.num_src > 1 ? static_cast<double>(p.x_ntiles[1]) * p.x_block_n[1] : 0.0);
    prof_begin(UB_KIND_WGRAD, 2.0 * pixels * 9 * cin * p.z_C, 2.0 * pixels * (cin + p.z_C), stream);
  }
  UB_CHECK_CUDA(launch_pdl(wgrad_row_kernel, dim3(static_cast<unsigned>(grid)), dim3(kThreads), smem, stream, p, z, x0, x1));
  UB_CHECK_CUDA(cudaGetLastError());
  prof_end(stream);
  return UB_OK;
}

}  // namespace ub
