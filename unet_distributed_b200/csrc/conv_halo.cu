#include "conv_halo.cuh"

#include <cstdlib>

#include "common.h"
#include "ptx.cuh"

namespace ub {

namespace {

constexpr int kThreads = 384;  // warp0 halo TMA, warp1 MMA + TMEM, warp2 weight TMA, warp3 idle, warps4..11 epilogue
constexpr int kHaloStages = 2;
constexpr int kMaxBStages = 4;
constexpr int kTH = 16;        // image rows per super-tile
constexpr int kSmemBudget = 225 * 1024;

struct __align__(8) HaloBarriers {
  uint64_t halo_full[kHaloStages];
  uint64_t halo_empty[kHaloStages];
  uint64_t b_full[kMaxBStages];
  uint64_t b_empty[kMaxBStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

__host__ __device__ inline int round_up_1k(int x) { return (x + 1023) & ~1023; }

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

// All MMAs of one filter tap: J pixel tiles x KSTEPS K slices, fully unrolled (descriptor = base + constant).
template <int J, int KSTEPS>
__device__ __forceinline__ void issue_tap(uint32_t d0, int block_n, uint64_t adesc0, uint32_t strip16, uint64_t bdesc,
                                          uint32_t idesc, uint32_t started) {
#pragma unroll
  for (int j = 0; j < J; ++j) {
#pragma unroll
    for (int k = 0; k < KSTEPS; ++k) {
      umma_f16(d0 + static_cast<uint32_t>(j * block_n), adesc0 + static_cast<uint64_t>(j * strip16 + 2 * k),
               bdesc + static_cast<uint64_t>(2 * k), idesc, (started | static_cast<uint32_t>(k)) ? 1u : 0u);
    }
  }
}

__global__ void __launch_bounds__(kThreads, 1)
conv_halo_kernel(const __grid_constant__ HaloParams p, const __grid_constant__ CUtensorMap tmA0,
                 const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB) {
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
  uint8_t* b_base = smem + kHaloStages * halo_stride;
  HaloBarriers* bars = reinterpret_cast<HaloBarriers*>(b_base + p.b_stages * bstage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.N * p.tiles_h * p.tiles_w * p.n_tiles;
  const int tap_groups = 9 / p.tpb;
  const int acc_sets = p.acc_double ? 2 : 1;
  uint32_t tmem_cols = 32;
  while (tmem_cols < static_cast<uint32_t>(acc_sets * p.J * p.block_n)) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kHaloStages; ++s) {
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
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------ halo producer
    if (lane == 0) {
      int hs = 0;
      uint32_t hphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int n, h0, w0, nt;
        decode_tile(p, tile, n, h0, w0, nt);
        for (int src = 0; src < p.num_src; ++src) {
          const CUtensorMap* tmA = src == 0 ? &tmA0 : &tmA1;
          for (int cb = 0; cb < p.src_kblocks[src]; ++cb) {
            mbar_wait(&bars->halo_empty[hs], hphase ^ 1u);
            mbar_arrive_expect_tx(&bars->halo_full[hs], static_cast<uint32_t>(halo_bytes));
            tma_load_4d(halo_base + hs * halo_stride, tmA, &bars->halo_full[hs], cb * p.kc, w0 - 1, h0 - 1, n);
            if (++hs == kHaloStages) { hs = 0; hphase ^= 1u; }
          }
        }
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
    // ------------------------------------------------------------ MMA issuer
    // The whole warp walks the (uniform) control flow and waits; one elected lane issues, so the compiler keeps
    // descriptors in uniform registers and emits bare UTCHMMAs. The 32/64-channel layers issue N=32/64 MMAs that
    // last 16-48 cycles each: the issue loop has to stay at a handful of instructions per MMA.
    const uint32_t idesc = umma_idesc_bf16(128, static_cast<uint32_t>(p.block_n), 0, 0);
    const uint32_t sw = p.kc == 64 ? UMMA_SW_128B : UMMA_SW_64B;
    const uint32_t a_sbo = static_cast<uint32_t>(P * rowbytes);   // next image row of the halo tile
    const uint32_t b_sbo = static_cast<uint32_t>(8 * rowbytes);
    const uint64_t desc_hi_a = umma_smem_desc(0, 0, a_sbo, sw);   // everything but the start address
    const uint64_t desc_hi_b = umma_smem_desc(0, 0, b_sbo, sw);
    const uint32_t row16 = static_cast<uint32_t>(rowbytes >> 4);  // one halo row, in 16-byte units
    int hs = 0, bs = 0, it = 0;
    uint32_t hphase = 0, bphase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int a = p.acc_double ? (it & 1) : 0;
      const uint32_t eparity = p.acc_double ? (((it >> 1) & 1) ^ 1u) : ((it & 1) ^ 1u);
      mbar_wait(&bars->tmem_empty[a], eparity);
      tc_fence_after();
      const uint32_t d0 = tmem_base + static_cast<uint32_t>(a * p.J * p.block_n);
      uint32_t started = 0;  // 0 until the first MMA of every accumulator has been issued
      for (int src = 0; src < p.num_src; ++src) {
        for (int cb = 0; cb < p.src_kblocks[src]; ++cb) {
          mbar_wait(&bars->halo_full[hs], hphase);
          const uint64_t a_base = desc_hi_a + (smem_u32(halo_base + hs * halo_stride) >> 4);
          for (int tg = 0; tg < tap_groups; ++tg) {
            mbar_wait(&bars->b_full[bs], bphase);
            tc_fence_after();
            const uint64_t b_stage = desc_hi_b + (smem_u32(b_base + bs * bstage_bytes) >> 4);
            if (elect_one()) {
              for (int t = 0; t < p.tpb; ++t) {
                const int tap = tg * p.tpb + t;
                const int r = tap / 3, s = tap - 3 * r;
                const uint64_t bdesc = b_stage + static_cast<uint32_t>(t * (btile_bytes >> 4));
                const uint64_t adesc0 = a_base + static_cast<uint32_t>((r * P + s)) * row16;
                if (p.kc == 64) {
                  if (p.J == 4) issue_tap<4, 4>(d0, p.block_n, adesc0, 8 * row16, bdesc, idesc, started);
                  else if (p.J == 2) issue_tap<2, 4>(d0, p.block_n, adesc0, 8 * row16, bdesc, idesc, started);
                  else issue_tap<1, 4>(d0, p.block_n, adesc0, 8 * row16, bdesc, idesc, started);
                } else {
                  if (p.J == 4) issue_tap<4, 2>(d0, p.block_n, adesc0, 8 * row16, bdesc, idesc, started);
                  else if (p.J == 2) issue_tap<2, 2>(d0, p.block_n, adesc0, 8 * row16, bdesc, idesc, started);
                  else issue_tap<1, 2>(d0, p.block_n, adesc0, 8 * row16, bdesc, idesc, started);
                }
                started = 1;
              }
              umma_commit(&bars->b_empty[bs]);
              if (tg == tap_groups - 1) umma_commit(&bars->halo_empty[hs]);
            }
            __syncwarp();
            started = 1;
            if (++bs == p.b_stages) { bs = 0; bphase ^= 1u; }
          }
          if (++hs == kHaloStages) { hs = 0; hphase ^= 1u; }
        }
      }
      if (elect_one()) umma_commit(&bars->tmem_full[a]);
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue: 8 warps, two per TMEM lane quarter
    // Work items are (pixel tile j, 16-column chunk); the two warps of a quarter take alternate items. The gate
    // tensor of the NEXT item (dgrad: the producer's forward output) is requested before the current item's
    // accumulators are read, so its latency hides behind the TMEM load and the stores.
    const int quarter = warp & 3;
    const int half = (warp - 4) >> 2;
    const int m = quarter * 32 + lane;
    const int th = m >> 3, twl = m & 7;
    const IgemmEpilogue& e = p.epi;
    const int chunks = p.block_n >> 4;
    const int items = p.J * chunks;
    const bool has_aux = (e.aux[0] != nullptr) || (e.aux[1] != nullptr);
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int a = p.acc_double ? (it & 1) : 0;
      const uint32_t fparity = p.acc_double ? ((it >> 1) & 1) : (it & 1);
      int n, h0, w0, nt;
      decode_tile(p, tile, n, h0, w0, nt);
      const int h = h0 + th;
      const size_t row_pix = (static_cast<size_t>(n) * p.H + h) * p.W;
      uint4 ax0 = make_uint4(0u, 0u, 0u, 0u), ax1 = ax0;
      auto aux_ptr = [&](int item) -> const uint4* {
        const int j = item / chunks, c = (item - j * chunks) << 4;
        const int w = w0 + 8 * j + twl;
        const int col = nt * p.block_n + c;
        const int slot = col >= e.split_col ? 1 : 0;
        if (!(h < p.H && w < p.W) || e.aux[slot] == nullptr) return nullptr;
        const int ch = slot ? col - e.split_col : col;
        return reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(e.aux[slot]) +
                                              (row_pix + w) * e.aux_C[slot] + e.aux_coff[slot] + ch);
      };
      if (has_aux && half < items) {
        const uint4* ap = aux_ptr(half);
        if (ap) { ax0 = __ldg(ap); ax1 = __ldg(ap + 1); }
      }
      mbar_wait(&bars->tmem_full[a], fparity);
      tc_fence_after();
      for (int item = half; item < items; item += 2) {
        const int j = item / chunks, c = (item - j * chunks) << 4;
        const int w = w0 + 8 * j + twl;
        const bool valid = (h < p.H) && (w < p.W);
        const size_t pix = row_pix + w;
        const uint4 cur0 = ax0, cur1 = ax1;
        if (has_aux && item + 2 < items) {
          const uint4* ap = aux_ptr(item + 2);
          if (ap) { ax0 = __ldg(ap); ax1 = __ldg(ap + 1); }
        }
        uint32_t r[16];
        tmem_ld_32x32b_x16(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                               static_cast<uint32_t>((a * p.J + j) * p.block_n + c), r);
        tmem_ld_wait();
        if (!valid) continue;
        const int col = nt * p.block_n + c;
        const int slot = col >= e.split_col ? 1 : 0;
        const int ch = slot ? col - e.split_col : col;
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
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
        const size_t elem0 = pix * e.dst_C[slot] + e.dst_coff[slot] + ch;
        if (e.drop_keep < 1.0f) {
          const float inv = 1.0f / e.drop_keep;
          const unsigned long long seed =
              e.drop_seed + (e.drop_seed_add ? static_cast<unsigned long long>(__ldg(e.drop_seed_add)) : 0ull);
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            const uint32_t bits = dropout_keep_bits8(seed, e.drop_stream, elem0 / 8 + g, e.drop_keep);
#pragma unroll
            for (int i = 0; i < 8; ++i) v[8 * g + i] = ((bits >> i) & 1u) ? v[8 * g + i] * inv : 0.0f;
          }
        }
        if (e.aux[slot]) {
          const uint32_t aw[8] = {cur0.x, cur0.y, cur0.z, cur0.w, cur1.x, cur1.y, cur1.z, cur1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[2 * i] = bf16_lo(aw[i]) > 0.0f ? v[2 * i] * e.mask_scale : 0.0f;
            v[2 * i + 1] = bf16_hi(aw[i]) > 0.0f ? v[2 * i + 1] * e.mask_scale : 0.0f;
          }
        }
        uint4* dp = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(e.dst[slot]) + elem0);
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          uint4 o;
          o.x = pack_bf16x2(v[8 * g + 0], v[8 * g + 1]);
          o.y = pack_bf16x2(v[8 * g + 2], v[8 * g + 3]);
          o.z = pack_bf16x2(v[8 * g + 4], v[8 * g + 5]);
          o.w = pack_bf16x2(v[8 * g + 6], v[8 * g + 7]);
          dp[g] = o;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->tmem_empty[a]);
    }
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
  return static_cast<size_t>(kHaloStages) * halo_stride + static_cast<size_t>(p.b_stages) * p.tpb * p.block_n * p.kc * 2 +
         sizeof(HaloBarriers) + 1024;
}

}  // namespace

bool halo_plan(HaloParams& p, int N, int H, int W, int kc, int ncols) {
  static const bool disabled = getenv("UNETB200_NO_HALO") != nullptr;
  if (disabled) return false;
  if (H < kTH || W < 8) return false;   // small grids: several images per tile, handled by igemm
  int block_n = ncols % 256 == 0 ? 256 : ncols;
  // measured: with 256-wide tiles (16x16 level) the per-tap igemm path is as fast or faster (few super-tiles,
  // single-buffered accumulators); the halo path pays off where the GEMM N is small
  if (block_n != 32 && block_n != 64 && block_n != 128) return false;
  int J = W >= 25 ? 4 : (W >= 9 ? 2 : 1);
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
    const int halo = kHaloStages * round_up_1k((kTH + 2) * P * kc * 2);
    int stages = (kSmemBudget - halo - 1024 - static_cast<int>(sizeof(HaloBarriers))) / (p.tpb * btile);
    if (stages > kMaxBStages) stages = kMaxBStages;
    if (stages >= 2) {
      p.b_stages = stages;
      break;
    }
    if (J == 1) return false;
  }
  p.tiles_h = (H + kTH - 1) / kTH;
  p.tiles_w = (W + 8 * p.J - 1) / (8 * p.J);
  return true;
}

int halo_launch(const HaloParams& p, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                cudaStream_t stream) {
  UB_REQUIRE(p.epi.d2s_cout == 0, "conv_halo: depth-to-space epilogue is not supported here");
  const size_t smem = halo_smem_bytes(p);
  UB_REQUIRE(smem <= 227 * 1024, "conv_halo: %zu bytes of shared memory requested", smem);
  static bool attr_set[64] = {false};
  int dev = 0;
  UB_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    UB_CHECK_CUDA(cudaFuncSetAttribute(conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev] = true;
  }
  const int num_tiles = p.N * p.tiles_h * p.tiles_w * p.n_tiles;
  UB_REQUIRE(num_tiles > 0, "conv_halo: empty problem");
  int grid = device_sm_count();
  if (grid > num_tiles) grid = num_tiles;
  if (profiling_enabled()) {
    const double pixels = static_cast<double>(p.N) * p.H * p.W;
    const double kdim = 9.0 * p.kc * (p.src_kblocks[0] + (p.num_src > 1 ? p.src_kblocks[1] : 0));
    const double ncols = static_cast<double>(p.n_tiles) * p.block_n;
    prof_begin(UB_KIND_IGEMM, 2.0 * pixels * kdim * ncols, 2.0 * pixels * (kdim / 9.0 + ncols), stream);
  }
  conv_halo_kernel<<<grid, kThreads, smem, stream>>>(p, a0, a1, b);
  UB_CHECK_CUDA(cudaGetLastError());
  prof_end(stream);
  return UB_OK;
}

}  // namespace ub
