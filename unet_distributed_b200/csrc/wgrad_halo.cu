// Weight gradient of a 3x3 conv for 32- and 64-channel inputs (the 128x128 and 64x64 levels of the UNet), where the
// per-tap wgrad engine (wgrad.cu) is issue- and L2-bound: C is so small that one (tap, Cout, Cin) tile is a sliver of
// the tensor core and every tap re-reads both operands.
//
// Here a CTA walks super-tiles of 16 rows x 8J columns. Per super-tile it loads ONE tile of the input with a column
// halo (16 x (8J+2) pixels) and ONE tile of the output gradient with a row halo (18 x 8J pixels) and accumulates ALL
// nine taps in TMEM:
//   dW[r, s, ci, co] = sum_q X[q + (0, s-1), ci] * dZ[q - (r-1) rows, co]        (q = p + (r-1) rows)
//   * the reduction (GEMM K) runs over pixels q: 16 pixels per MMA = rows 2i and 2i+1 of an 8-wide strip; the strip's
//     8 pixels are 8 consecutive smem rows and the stride to the next image row is the descriptor's SBO;
//   * filter COLUMNS are stacked along GEMM M: both operands are MN-major, and the "next channel atom along M" (the
//     A descriptor's LBO) is ONE pixel row, so atom a of A is the X tile shifted by a pixels = tap column s0 + a;
//   * filter ROWS are stacked along GEMM N the same way: the B descriptor's LBO is one IMAGE row of the dZ tile, so
//     atom a of B is dZ shifted by a rows = filter row r = 2 - a. One MMA (M = 128, N = 3*Cz) therefore produces up to
//     4 x 3 taps: D[(s, ci), (a, co)] -- a 128 x 96 x 16 MMA costs 56 cycles (32 + N/4, csrc/mma_rate.cu) where the
//     three 128 x 32 x 16 MMAs of the earlier row-by-row version cost 120. Operand reads from shared memory, not math,
//     bound these layers.
// Partial results of the CTAs are summed in a fixed order by wgrad_halo_reduce_kernel (bit-reproducible).
#include "wgrad.cuh"

#include <cstdlib>

#include "common.h"
#include "ptx.cuh"

namespace ub {

namespace {

constexpr int kThreads = 256;   // warp0 TMA producer, warp1 MMA + TMEM, warps 4..7 epilogue
constexpr int kMaxStages = 3;
constexpr int kTH = 16;
constexpr int kBiasScratch = 4096; // [128 threads][8] fp32: per-thread column sums of dZ before the CTA-level sum

struct WgradHaloParams {
  int N, H, W;
  int J;                    // 8-wide strips per super-tile
  int tiles_h, tiles_w;
  int Cx, Cz;               // 32 or 64 each
  int stages;
  float* ws;                // [gridDim.x][9*Cx*Cz (+ Cz)] partial gradients (+ partial bias gradient)
  int with_bias;            // also reduce dZ over pixels (the otherwise idle epilogue warps sum the smem tile)
};

struct __align__(8) WgradHaloBarriers {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

__host__ __device__ inline int round_up_1k(int x) { return (x + 1023) & ~1023; }

__device__ __forceinline__ uint64_t desc_pair(uint32_t lo, uint32_t hi) {
  return (static_cast<uint64_t>(hi) << 32) | lo;
}

// MMA issuer, specialised on (J, Cx, Cz): every descriptor of a super-tile is "stage base + compile-time constant"
// (csrc/mma_rate.cu: a 128 x 32 x 16 SS MMA retires every 40 cycles only if the issuing lane spends no more than a
// few scalar instructions per MMA; the run-time-generic loop with 64-bit descriptor arithmetic ran at ~75).
template <int J, int CX, int CZ>
__device__ __forceinline__ void wgrad_halo_mma_role(WgradHaloBarriers* bars, uint32_t smem_addr, uint32_t stage_bytes,
                                                    uint32_t x_stride, int stages, uint32_t tmem_base, int my_tiles) {
  constexpr int P = 8 * J + 2;                 // X tile pitch (column halo)
  constexpr int PZ = 8 * J;                    // dZ tile pitch
  constexpr uint32_t XROW = CX * 2, ZROW = CZ * 2;
  constexpr int GROUPS = CX == 32 ? 1 : 2;     // MMAs per K block: 4 stacked filter columns (Cx=32) or 2 x 2 (Cx=64)
  constexpr uint32_t XSW = CX == 64 ? UMMA_SW_128B : UMMA_SW_64B;
  constexpr uint32_t ZSW = CZ == 64 ? UMMA_SW_128B : UMMA_SW_64B;
  const uint32_t idesc = umma_idesc_bf16(128, 3 * CZ, 1, 1);
  // A: atoms along M one pixel row apart (LBO = row bytes); 8-pixel groups along K one image row apart (SBO).
  // B: atoms along N one IMAGE row apart (LBO = PZ rows); 8-pixel groups along K one image row apart (SBO).
  const uint32_t x_hi = static_cast<uint32_t>(umma_smem_desc(0, XROW, P * XROW, XSW) >> 32);
  const uint32_t z_hi = static_cast<uint32_t>(umma_smem_desc(0, PZ * ZROW, PZ * ZROW, ZSW) >> 32);
  // LBO lives in bits [16,30) of the low word
  const uint32_t x_lo_const = static_cast<uint32_t>(umma_smem_desc(0, XROW, P * XROW, UMMA_SW_NONE) & 0xFFFFFFFFu);
  const uint32_t z_lo_const =
      static_cast<uint32_t>(umma_smem_desc(0, PZ * ZROW, PZ * ZROW, UMMA_SW_NONE) & 0xFFFFFFFFu);
  int stage = 0;
  uint32_t phase = 0;
  for (int k = 0; k < my_tiles; ++k) {
    mbar_wait(&bars->full[stage], phase);
    tc_fence_after();
    const uint32_t sx = smem_addr + static_cast<uint32_t>(stage) * stage_bytes;
    const uint32_t x_lo = x_lo_const + (sx >> 4);
    const uint32_t z_lo = z_lo_const + ((sx + x_stride) >> 4);
    if (elect_one()) {
      const uint32_t first = k ? 1u : 0u;
      // i: pairs of image rows (16 pixels of reduction), j: strips, g: groups of stacked filter columns.
      // X rows 2i, 2i+1 of the tile meet dZ tile rows 2i + a, 2i + 1 + a (the dZ tile starts one row above).
#pragma unroll
      for (int i = 0; i < kTH / 2; ++i) {
#pragma unroll
        for (int j = 0; j < J; ++j) {
#pragma unroll
          for (int g = 0; g < GROUPS; ++g) {
            umma_f16(tmem_base + static_cast<uint32_t>(g * 3 * CZ),
                     desc_pair(x_lo + static_cast<uint32_t>((2 * i * P + 8 * j + 2 * g) * (XROW >> 4)), x_hi),
                     desc_pair(z_lo + static_cast<uint32_t>((2 * i * PZ + 8 * j) * (ZROW >> 4)), z_hi), idesc,
                     (i == 0 && j == 0) ? first : 1u);
          }
        }
      }
      umma_commit(&bars->empty[stage]);
      if (k == my_tiles - 1) umma_commit(&bars->tmem_full);
    }
    __syncwarp();
    if (++stage == stages) { stage = 0; phase ^= 1u; }
  }
}

__global__ void __launch_bounds__(kThreads, 1)
wgrad_halo_kernel(const __grid_constant__ WgradHaloParams p, const __grid_constant__ CUtensorMap tmX,
                  const __grid_constant__ CUtensorMap tmZ) {
  pdl_trigger();     // the next kernel of the stream may take SMs as this grid's CTAs leave them
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

  const int P = 8 * p.J + 2;
  const int xrow = p.Cx * 2, zrow = p.Cz * 2;
  const int halo_bytes = kTH * P * xrow;                   // X: 16 rows x (8J+2) columns
  const int z_bytes = (kTH + 2) * 8 * p.J * zrow;          // dZ: 18 rows x 8J columns
  // slack after the X tile: the junk filter column (s = 3) reads up to 3 pixel rows past its end
  const int halo_stride = round_up_1k(halo_bytes + 4 * xrow);
  const int stage_bytes = halo_stride + round_up_1k(z_bytes);
  float* bias_scratch = reinterpret_cast<float*>(smem + p.stages * stage_bytes);
  WgradHaloBarriers* bars = reinterpret_cast<WgradHaloBarriers*>(smem + p.stages * stage_bytes + kBiasScratch);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.N * p.tiles_h * p.tiles_w;
  const int groups = p.Cx == 32 ? 1 : 2;          // stacked filter columns per MMA: 4 (Cx=32) or 2 (Cx=64)
  const int n_acc = 3 * groups;                   // accumulator g holds filter rows 2, 1, 0 side by side (3*Cz columns)
  uint32_t tmem_cols = 32;
  while (tmem_cols < static_cast<uint32_t>(n_acc * p.Cz)) tmem_cols <<= 1;
  const int my_tiles = (num_tiles > static_cast<int>(blockIdx.x))
                           ? (num_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                                 static_cast<int>(gridDim.x)
                           : 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], p.with_bias ? 5 : 1);   // MMA commit (+ the four column-summing warps)
    }
    mbar_init(&bars->tmem_full, 1);
    mbar_fence_init();
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmZ);
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
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int t = tile;
        const int tw_i = t % p.tiles_w;
        t /= p.tiles_w;
        const int th_i = t % p.tiles_h;
        const int n = t / p.tiles_h;
        const int h0 = th_i * kTH, w0 = tw_i * 8 * p.J;
        mbar_wait(&bars->empty[stage], phase ^ 1u);
        uint8_t* sx = smem + stage * stage_bytes;
        mbar_arrive_expect_tx(&bars->full[stage], static_cast<uint32_t>(halo_bytes + z_bytes));
        tma_load_4d(sx, &tmX, &bars->full[stage], 0, w0 - 1, h0, n);
        tma_load_4d(sx + halo_stride, &tmZ, &bars->full[stage], 0, w0, h0 - 1, n);
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    if (my_tiles > 0) {
      const uint32_t sa = smem_u32(smem);
#define UB_WH(JJ, CX, CZ)                                                                                 \
      wgrad_halo_mma_role<JJ, CX, CZ>(bars, sa, static_cast<uint32_t>(stage_bytes),                       \
                                      static_cast<uint32_t>(halo_stride), p.stages, tmem_base, my_tiles)
#define UB_WH_J(CX, CZ)                        \
      do {                                     \
        if (p.J == 4) UB_WH(4, CX, CZ);        \
        else if (p.J == 2) UB_WH(2, CX, CZ);   \
        else UB_WH(1, CX, CZ);                 \
      } while (0)
      if (p.Cx == 32) {
        if (p.Cz == 32) UB_WH_J(32, 32);
        else UB_WH_J(32, 64);
      } else {
        if (p.Cz == 32) UB_WH_J(64, 32);
        else UB_WH_J(64, 64);
      }
#undef UB_WH_J
#undef UB_WH
    }
  } else if (warp >= 4) {
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    const int a = m / p.Cx;              // which stacked filter column this TMEM lane belongs to
    const int ci = m % p.Cx;
    const int n_out = 9 * p.Cx * p.Cz + (p.with_bias ? p.Cz : 0);
    float* out = p.ws + static_cast<size_t>(blockIdx.x) * n_out;
    if (p.with_bias) {
      // Bias gradient: while the tensor core works on a stage, these four warps sum its dZ tile over pixels.
      // Thread t owns the 16-byte channel chunk q = t % chunks of every (128 / chunks)-th pixel row; the TMA swizzle
      // (64-byte rows: chunk ^= (row >> 1) & 3, 128-byte rows: chunk ^= row & 7) is undone per row.
      const int t = threadIdx.x - 128;
      const int chunks = p.Cz >> 3;
      const int q = t & (chunks - 1);
      const int row_begin = 8 * p.J, row_end = (kTH + 1) * 8 * p.J;   // skip the dZ tile's halo rows
      float bsum[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) bsum[j] = 0.0f;
      int stage = 0;
      uint32_t phase = 0;
      for (int k = 0; k < my_tiles; ++k) {
        mbar_wait(&bars->full[stage], phase);
        const uint8_t* zt = smem + stage * stage_bytes + halo_stride;
        for (int row = row_begin + t / chunks; row < row_end; row += 128 / chunks) {
          const int phys = p.Cz == 64 ? (q ^ (row & 7)) : (q ^ ((row >> 1) & 3));
          const uint4 v = *reinterpret_cast<const uint4*>(zt + row * zrow + phys * 16);
          bsum[0] += bf16_lo(v.x); bsum[1] += bf16_hi(v.x); bsum[2] += bf16_lo(v.y); bsum[3] += bf16_hi(v.y);
          bsum[4] += bf16_lo(v.z); bsum[5] += bf16_hi(v.z); bsum[6] += bf16_lo(v.w); bsum[7] += bf16_hi(v.w);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->empty[stage]);
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) bias_scratch[t * 8 + j] = bsum[j];
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (t < p.Cz) {   // channel t = chunk (t / 8), element (t % 8): sum the threads that own that chunk in a fixed order
        float acc = 0.0f;
        for (int u = t >> 3; u < 128; u += chunks) acc += bias_scratch[u * 8 + (t & 7)];
        out[9 * p.Cx * p.Cz + t] = acc;
      }
    }
    if (my_tiles > 0) {
      mbar_wait(&bars->tmem_full, 0);
      tc_fence_after();
    }
    for (int r = 0; r < 3; ++r) {
      for (int g = 0; g < groups; ++g) {
        const int s = (p.Cx == 32 ? 0 : 2 * g) + a;
        // filter row r is B atom 2 - r of accumulator g
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                               static_cast<uint32_t>((g * 3 + (2 - r)) * p.Cz);
        for (int c = 0; c < p.Cz; c += 16) {
          uint32_t v[16];
          if (my_tiles > 0) {
            tmem_ld_32x32b_x16(t_row + static_cast<uint32_t>(c), v);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = 0u;
          }
          if (s < 3) {
            float4* o4 = reinterpret_cast<float4*>(out + (static_cast<size_t>(r * 3 + s) * p.Cx + ci) * p.Cz + c);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              o4[q] = make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]),
                                  __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3]));
            }
          }
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

// dw[tap][ci_off + ci][co] = sum over parts of ws[part][tap][ci][co]; 8 interleaved partial sums then a fixed 8-way sum.
__global__ void wgrad_halo_reduce_kernel(const float* __restrict__ ws, int nparts, int Cx, int Cz, float* __restrict__ dw,
                                         int cin_total, int ci_off, float* __restrict__ db) {
  __shared__ float part[32][32];
  const int lanes = blockDim.x >> 5;
  const int nw = 9 * Cx * Cz;
  const int n = nw + (db ? Cz : 0);
  const int c = threadIdx.x & 31, q = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + c;
  float acc = 0.0f;
  if (i < n) {
    for (int r = q; r < nparts; r += lanes) acc += ws[static_cast<size_t>(r) * n + i];
  }
  part[q][c] = acc;
  __syncthreads();
  if (q == 0 && i < n) {
    float s = part[0][c];
    for (int k = 1; k < lanes; ++k) s += part[k][c];
    if (i >= nw) {
      db[i - nw] = s;
    } else {
      const int co = i % Cz;
      const int ci = (i / Cz) % Cx;
      const int tap = i / (Cz * Cx);
      dw[(static_cast<size_t>(tap) * cin_total + ci_off + ci) * Cz + co] = s;
    }
  }
}

int pick_J(int Cx, int W) {
  int J = Cx == 32 ? 4 : 2;
  while (J > 1 && 8 * (J / 2) >= W) J >>= 1;   // do not make strips that lie entirely outside the image
  return J;
}

}  // namespace

bool wgrad_halo_supported(int H, int W, int Cx, int Cz) {
  static const bool disabled = getenv("UNETB200_NO_HALO") != nullptr;
  if (disabled) return false;
  return (Cx == 32 || Cx == 64) && (Cz == 32 || Cz == 64) && H >= kTH && W >= 8;
}

static int halo_grid(int N, int H, int W, int J) {
  const int tiles = N * ((H + kTH - 1) / kTH) * ((W + 8 * J - 1) / (8 * J));
  const int sms = device_sm_count();
  return tiles < sms ? tiles : sms;
}

size_t wgrad_halo_ws_bytes(int N, int H, int W, int Cx, int Cz) {
  const int J = pick_J(Cx, W);
  return static_cast<size_t>(halo_grid(N, H, W, J)) * (9 * Cx * Cz + Cz) * sizeof(float);
}

// One source of the (possibly concatenated) input: writes dw[:, ci_off : ci_off + Cx, :].
int wgrad_halo_launch(const CUtensorMap& x_map, const CUtensorMap& z_map, int N, int H, int W, int Cx, int Cz, int J,
                      float* dw, int cin_total, int ci_off, float* ws, cudaStream_t stream, float* db) {
  WgradHaloParams p;
  p.N = N; p.H = H; p.W = W;
  p.J = J;
  p.tiles_h = (H + kTH - 1) / kTH;
  p.tiles_w = (W + 8 * J - 1) / (8 * J);
  p.Cx = Cx; p.Cz = Cz;
  p.ws = ws;
  p.with_bias = db ? 1 : 0;
  const int P = 8 * J + 2;
  const int stage_bytes = round_up_1k(kTH * P * Cx * 2 + 4 * Cx * 2) + round_up_1k((kTH + 2) * 8 * J * Cz * 2);
  int stages = (225 * 1024 - 1024 - kBiasScratch - static_cast<int>(sizeof(WgradHaloBarriers))) / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  UB_REQUIRE(stages >= 2, "wgrad_halo: stage of %d bytes does not fit twice in shared memory", stage_bytes);
  p.stages = stages;
  const size_t smem = static_cast<size_t>(stages) * stage_bytes + kBiasScratch + sizeof(WgradHaloBarriers) + 1024;
  static bool attr_set[64] = {false};
  int dev = 0;
  UB_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    UB_CHECK_CUDA(cudaFuncSetAttribute(wgrad_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev] = true;
  }
  const int grid = halo_grid(N, H, W, J);
  UB_REQUIRE(grid > 0, "wgrad_halo: empty problem");
  if (profiling_enabled()) {
    const double pixels = static_cast<double>(N) * H * W;
    prof_begin(UB_KIND_WGRAD, 2.0 * pixels * 9 * Cx * Cz, 2.0 * pixels * (Cx + Cz), stream);
  }
  UB_CHECK_CUDA(launch_pdl(wgrad_halo_kernel, dim3(grid), dim3(kThreads), smem, stream, p, x_map, z_map));
  UB_CHECK_CUDA(cudaGetLastError());
  prof_end(stream);
  const int n = 9 * Cx * Cz + (db ? Cz : 0);
  wgrad_halo_reduce_kernel<<<(n + 31) / 32, 1024, 0, stream>>>(ws, grid, Cx, Cz, dw, cin_total, ci_off, db);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int wgrad_halo_pick_J(int Cx, int W) { return pick_J(Cx, W); }

}  // namespace ub
