#include "wgrad.cuh"

#include "common.h"
#include "ptx.cuh"

namespace ub {

namespace {

constexpr int kThreads = 192;
constexpr int kMaxStages = 6;
constexpr int kKP = 64;                    // pixels per K block
// smem reserved per operand per stage: 128 channels x 64 pixels of elt-byte elements (16 KB bf16, 32 KB fp32/tf32)
__host__ __device__ constexpr int operand_bytes(int elt) { return 128 * kKP * elt; }

struct __align__(8) WgradBarriers {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(kThreads, 1)
wgrad_kernel(const __grid_constant__ WgradParams p, const __grid_constant__ CUtensorMap tmZ,
             const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1) {
  pdl_trigger();     // the next kernel of the stream may take SMs as this grid's CTAs leave them
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const int elt = p.elt == 4 ? 4 : 2;
  const int kOperandBytes = operand_bytes(elt);
  const int kStageBytes = 2 * kOperandBytes;
  float* bias_scratch = reinterpret_cast<float*>(smem + p.stages * kStageBytes);
  WgradBarriers* bars = reinterpret_cast<WgradBarriers*>(smem + p.stages * kStageBytes + kWgradBiasScratch);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // ---- which output tile / pixel range does this CTA own?
  const int total_nt = p.x_ntiles[0] + (p.num_src > 1 ? p.x_ntiles[1] : 0);
  int bid = blockIdx.x;
  int ntg = bid % total_nt;
  bid /= total_nt;
  const int mt = bid % p.m_tiles;
  bid /= p.m_tiles;
  // tap_in_m (transposed conv): the four taps are stacked along GEMM M (row = tap * z_C + channel), so one CTA reads
  // the X tile once for all of them and the MMAs use all 128 rows even when z_C is 32
  const int ntaps = p.tap_in_m ? 1 : p.taps_r * p.taps_s;
  const int tap = bid % ntaps;
  const int split = bid / ntaps;
  const int m_total = p.tap_in_m ? p.taps_r * p.taps_s * p.z_C : p.z_C;
  const int atoms_per_tap = p.z_pair ? 1 : p.z_C / p.z_atom;
  const int src = (ntg >= p.x_ntiles[0]) ? 1 : 0;
  const int nt = src ? ntg - p.x_ntiles[0] : ntg;
  const int block_n = p.x_block_n[src];
  const int x_atom = p.x_atom[src];
  const int tr = tap / p.taps_s, ts = tap % p.taps_s;
  const int m0 = mt * 128;
  const int c0 = nt * block_n;

  const int total_pb = p.blocks_n * p.blocks_h * p.blocks_w;
  const int per_split = (total_pb + p.splits - 1) / p.splits;
  const int pb_begin = split * per_split;
  const int pb_end = (pb_begin + per_split < total_pb) ? pb_begin + per_split : total_pb;
  const int nkb = pb_end > pb_begin ? pb_end - pb_begin : 0;

  int m_atoms = (m_total - m0 + p.z_atom - 1) / p.z_atom;
  if (m_atoms > 128 / p.z_atom) m_atoms = 128 / p.z_atom;
  const int n_atoms = block_n / x_atom;
  const uint32_t z_atom_bytes = static_cast<uint32_t>(kKP * p.z_atom * elt);
  const uint32_t x_atom_bytes = static_cast<uint32_t>(kKP * x_atom * elt);
  const uint32_t tmem_cols = block_n < 32 ? 32u : static_cast<uint32_t>(block_n);
  // the CTAs of the first X tile together read every dZ element once: they also sum it over pixels (bias gradient)
  const bool do_bias = p.bias_out != nullptr && ntg == 0;

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
      const uint32_t tx = m_atoms * z_atom_bytes + n_atoms * x_atom_bytes;
      for (int pb = pb_begin; pb < pb_end; ++pb) {
        int t = pb;
        const int bw = t % p.blocks_w;
        t /= p.blocks_w;
        const int bh = t % p.blocks_h;
        const int bn = t / p.blocks_h;
        const int n0 = bn * p.TN, h0 = bh * p.TH, w0 = bw * p.TW;
        mbar_wait(&bars->empty[stage], phase ^ 1u);
        uint8_t* sa = smem + stage * kStageBytes;
        uint8_t* sb = sa + kOperandBytes;
        mbar_arrive_expect_tx(&bars->full[stage], tx);
        for (int a = 0; a < m_atoms; ++a) {
          if (p.z_mode == 0) {
            tma_load_4d(sa + a * z_atom_bytes, &tmZ, &bars->full[stage], m0 + a * p.z_atom, w0, h0, n0);
          } else if (p.z_pair) {
            // atom = filter row m0 / 64 + a of the pair view (ops.cu fine_pair_map): 128-byte rows of (column tap, channel)
            tma_load_4d(sa + a * z_atom_bytes, &tmZ, &bars->full[stage], 0, w0, m0 / p.z_atom + a, n0 * p.H + h0);
          } else if (p.tap_in_m) {
            const int ga = m0 / p.z_atom + a;            // atom index over all taps
            const int atap = ga / atoms_per_tap;
            tma_load_5d(sa + a * z_atom_bytes, &tmZ, &bars->full[stage], (ga % atoms_per_tap) * p.z_atom,
                        atap % p.taps_s, w0, atap / p.taps_s, n0 * p.H + h0);
          } else {
            tma_load_5d(sa + a * z_atom_bytes, &tmZ, &bars->full[stage], m0 + a * p.z_atom, ts, w0, tr,
                        n0 * p.H + h0);
          }
        }
        const int dw = p.x_shift ? ts - p.pad : 0;
        const int dh = p.x_shift ? tr - p.pad : 0;
        for (int a = 0; a < n_atoms; ++a) {
          tma_load_4d(sb + a * x_atom_bytes, tmX, &bars->full[stage], c0 + a * x_atom, w0 + dw, h0 + dh, n0);
        }
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    if (nkb > 0) {
      // warp-uniform control flow, one elected lane issues
      const uint32_t idesc = elt == 4 ? umma_idesc_tf32(128, static_cast<uint32_t>(block_n), 1, 1)
                                      : umma_idesc_bf16(128, static_cast<uint32_t>(block_n), 1, 1);
      const int kmma = 32 / elt;   // pixels (GEMM K) per instruction: 16 bf16 or 8 tf32
      // MN-major tf32 operands exist in one layout only: 128-byte rows swizzled in 32-byte chunks with a period of
      // FOUR pixel rows (SWIZZLE_128B_BASE32B; the tensor maps use CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B), so the
      // stride between K groups is 4 rows; bf16 uses the ordinary 128 B / 64 B swizzles with 8-row groups
      const uint32_t z_sw = elt == 4 ? UMMA_SW_128B_BASE32B : (p.z_atom * elt == 128 ? UMMA_SW_128B : UMMA_SW_64B);
      const uint32_t x_sw = elt == 4 ? UMMA_SW_128B_BASE32B : (x_atom * elt == 128 ? UMMA_SW_128B : UMMA_SW_64B);
      const int krows = elt == 4 ? 4 : 8;
      const uint32_t z_sbo = static_cast<uint32_t>(krows * p.z_atom * elt);
      const uint32_t x_sbo = static_cast<uint32_t>(krows * x_atom * elt);
      const uint32_t z_kstep = static_cast<uint32_t>(kmma * p.z_atom * elt) >> 4;  // kmma pixels, in 16-byte units
      const uint32_t x_kstep = static_cast<uint32_t>(kmma * x_atom * elt) >> 4;
      const uint64_t a_hi = umma_smem_desc(0, z_atom_bytes, z_sbo, z_sw);
      const uint64_t b_hi = umma_smem_desc(0, x_atom_bytes, x_sbo, x_sw);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&bars->full[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * kStageBytes);
        const uint64_t adesc = a_hi + (sa >> 4);
        const uint64_t bdesc = b_hi + ((sa + kOperandBytes) >> 4);
        if (elect_one()) {
          if (elt == 4) {
#pragma unroll
            for (int k = 0; k < kKP / 8; ++k) {
              umma_tf32(tmem_base, adesc + static_cast<uint64_t>(k * z_kstep),
                        bdesc + static_cast<uint64_t>(k * x_kstep), idesc, (kb | k) ? 1u : 0u);
            }
          } else {
#pragma unroll
          for (int k = 0; k < kKP / 16; ++k) {
            umma_f16(tmem_base, adesc + static_cast<uint64_t>(k * z_kstep), bdesc + static_cast<uint64_t>(k * x_kstep),
                     idesc, (kb | k) ? 1u : 0u);
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
      // While the tensor core works on a stage, warps 2..5 sum its dZ tile over the 64 pixels. The tile is m_atoms atoms
      // of 64 pixel rows x z_atom channels (128-byte rows: 16-byte chunk ^= pixel & 7; 64-byte rows: chunk ^=
      // (pixel >> 1) & 3). Thread t: 16-byte chunk t % 16 of the tile's channels, pixels t / 16 + 8k.
      const int t = threadIdx.x - 64;
      const int cpa = p.z_atom >> 3;                 // 16-byte chunks per atom row: 8 or 4
      const int q = t & 15, rg = t >> 4;
      const int atom = q / cpa, cq = q % cpa;
      const int rowb = p.z_atom * 2;
      float bsum[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) bsum[j] = 0.0f;
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&bars->full[stage], phase);
        if (atom < m_atoms) {
          const uint8_t* zt = smem + stage * kStageBytes + atom * z_atom_bytes;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int px = rg + 8 * k;
            const int sw = cpa == 8 ? (px & 7) : ((px >> 1) & 3);
            const uint4 v = *reinterpret_cast<const uint4*>(zt + px * rowb + ((cq ^ sw) << 4));
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
      // GEMM row c of the tile = chunk c / 8, element c % 8; its eight pixel-group threads in a fixed order
      const int c = t;
      if (m0 + c < m_total) {
        float acc = 0.0f;
#pragma unroll
        for (int g = 0; g < 8; ++g) acc += bias_scratch[((c >> 3) + 16 * g) * 8 + (c & 7)];
        const int btap = p.tap_in_m ? (m0 + c) / p.z_C : tap;
        const int bco = p.tap_in_m ? (m0 + c) % p.z_C : m0 + c;
        p.bias_out[(static_cast<long long>(split) * p.taps_r * p.taps_s + btap) * p.z_C + bco] = acc;
      }
    }
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int mrow = m0 + row;
    const bool valid = mrow < m_total;
    const int otap = p.tap_in_m ? mrow / p.z_C : tap;
    const int co = p.tap_in_m ? mrow % p.z_C : mrow;
    float* out = p.out + split * p.out_split_stride + otap * p.out_tap_stride + co * p.out_m_stride +
                 static_cast<long long>(p.x_col_off[src] + c0) * p.out_n_stride;
    if (nkb > 0) {
      mbar_wait(&bars->tmem_full, 0);
      tc_fence_after();
    }
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    for (int c = 0; c < block_n; c += 16) {
      uint32_t r[16];
      if (nkb > 0) {
        tmem_ld_32x32b_x16(t_row + static_cast<uint32_t>(c), r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) r[i] = 0u;
      }
      if (valid) {
        if (p.out_n_stride == 1) {
          float4* o4 = reinterpret_cast<float4*>(out + c);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            o4[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) out[static_cast<long long>(c + i) * p.out_n_stride] = __uint_as_float(r[i]);
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

// out[i] = sum over splits of ws[s * split_stride + i], in a fixed order: R row lanes per column walk the splits
// s = p, p + R, p + 2R, ... (four loads in flight), then the R partial sums are added in lane order. R depends only on
// the split count, so a layer's gradient is reduced in the same order on every run (bit-reproducible). A block is 256
// threads = R lanes x 256/R consecutive columns (coalesced rows). The one-thread-per-column loop this replaces ran at
// 2 TB/s with as few as 32 blocks (ncu r02: 149 us per step for 298 MB).
// fewer than eight partials per column: one thread per column, loads independent of the running sum
__global__ void splitk_reduce_few_kernel(const float* __restrict__ ws, long long split_stride, int splits,
                                         float* __restrict__ out, long long n, float* __restrict__ out_b, long long n_a) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float acc = 0.0f;
  for (int s = 0; s < splits; ++s) acc += ws[s * split_stride + i];
  if (out_b && i >= n_a) out_b[i - n_a] = acc;
  else out[i] = acc;
}

template <int R>
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ ws, long long split_stride, int splits, float* __restrict__ out,
                     long long n, float* __restrict__ out_b, long long n_a) {
  constexpr int C = 256 / R;
  __shared__ float part[R][C];
  const int c = threadIdx.x % C, p = threadIdx.x / C;
  const long long i = static_cast<long long>(blockIdx.x) * C + c;
  float acc = 0.0f;
  if (i < n) {
    const float* col = ws + i;
    int sidx = p;
    for (; sidx + 3 * R < splits; sidx += 4 * R) {
      const float v0 = col[static_cast<long long>(sidx) * split_stride];
      const float v1 = col[static_cast<long long>(sidx + R) * split_stride];
      const float v2 = col[static_cast<long long>(sidx + 2 * R) * split_stride];
      const float v3 = col[static_cast<long long>(sidx + 3 * R) * split_stride];
      acc += v0; acc += v1; acc += v2; acc += v3;
    }
    for (; sidx < splits; sidx += R) acc += col[static_cast<long long>(sidx) * split_stride];
  }
  part[p][c] = acc;
  __syncthreads();
  if (p == 0 && i < n) {
    float t = part[0][c];
#pragma unroll
    for (int q = 1; q < R; ++q) t += part[q][c];
    if (out_b && i >= n_a) out_b[i - n_a] = t;
    else out[i] = t;
  }
}

}  // namespace

int wgrad_launch(const WgradParams& p, const CUtensorMap& z, const CUtensorMap& x0, const CUtensorMap& x1,
                 cudaStream_t stream) {
  UB_REQUIRE(p.TN * p.TH * p.TW == kKP, "wgrad: pixel box %dx%dx%d is not %d pixels", p.TN, p.TH, p.TW, kKP);
  UB_REQUIRE(p.stages >= 2 && p.stages <= kMaxStages, "wgrad: stages %d out of range", p.stages);
  const int elt = p.elt == 4 ? 4 : 2;
  UB_REQUIRE(p.z_atom * elt == 128 || (elt == 2 && p.z_atom == 32), "wgrad: z_atom %d x %d bytes", p.z_atom, elt);
  for (int s = 0; s < p.num_src; ++s) {
    UB_REQUIRE(p.x_atom[s] * elt == 128 || (elt == 2 && p.x_atom[s] == 32), "wgrad: x_atom %d x %d bytes", p.x_atom[s],
               elt);
    UB_REQUIRE(p.x_block_n[s] >= 32 && p.x_block_n[s] <= 128 && p.x_block_n[s] % 16 == 0, "wgrad: block_n %d",
               p.x_block_n[s]);
  }
  const int total_pb = p.blocks_n * p.blocks_h * p.blocks_w;
  UB_REQUIRE(p.splits >= 1 && p.splits <= total_pb, "wgrad: splits %d vs %d pixel blocks", p.splits, total_pb);
  UB_REQUIRE(!p.bias_out || elt == 2, "wgrad: the fused bias gradient is a bf16 feature");
  const size_t smem = static_cast<size_t>(p.stages) * 2 * operand_bytes(elt) + kWgradBiasScratch + sizeof(WgradBarriers) + 1024;
  UB_REQUIRE(smem <= 227 * 1024, "wgrad: %zu bytes of shared memory requested", smem);
  static bool attr_set[64] = {false};
  int dev = 0;
  UB_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    UB_CHECK_CUDA(cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev] = true;
  }
  const int total_nt = p.x_ntiles[0] + (p.num_src > 1 ? p.x_ntiles[1] : 0);
  const long long grid = static_cast<long long>(p.splits) * (p.tap_in_m ? 1 : p.taps_r * p.taps_s) * p.m_tiles * total_nt;
  UB_REQUIRE(grid > 0 && grid < (1ll << 30), "wgrad: grid %lld", grid);
  if (profiling_enabled()) {
    const double pixels = static_cast<double>(p.N) * p.H * p.W;
    const double cin = static_cast<double>(p.x_C[0]) + (p.num_src > 1 ? p.x_C[1] : 0);
    prof_begin(UB_KIND_WGRAD, 2.0 * pixels * p.taps_r * p.taps_s * cin * p.z_C,
               static_cast<double>(elt) * pixels * (cin + p.z_C), stream);
  }
  UB_CHECK_CUDA(launch_pdl(wgrad_kernel, dim3(static_cast<unsigned>(grid)), dim3(kThreads), smem, stream, p, z, x0, x1));
  UB_CHECK_CUDA(cudaGetLastError());
  prof_end(stream);
  return UB_OK;
}

int splitk_reduce_launch(const float* ws, long long split_stride, int splits, float* out, long long n,
                         cudaStream_t stream, float* out_b, long long n_a) {
  if (n <= 0) return UB_OK;
  // at least four splits per lane (four loads in flight): with few splits one thread per column is fastest
  // (ncu r02: R = 4 on 4..7 splits ran at 1 TB/s, twice as long as the one-thread loop it replaced)
  const int R = splits >= 32 ? 8 : (splits >= 16 ? 4 : (splits >= 8 ? 2 : 1));
  const long long blocks = (n + 256 / R - 1) / (256 / R);
  const unsigned g = static_cast<unsigned>(blocks);
  if (R == 1) {
    splitk_reduce_few_kernel<<<g, 256, 0, stream>>>(ws, split_stride, splits, out, n, out_b, n_a);
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  switch (R) {
    case 8: splitk_reduce_kernel<8><<<g, 256, 0, stream>>>(ws, split_stride, splits, out, n, out_b, n_a); break;
    case 4: splitk_reduce_kernel<4><<<g, 256, 0, stream>>>(ws, split_stride, splits, out, n, out_b, n_a); break;
    case 2: splitk_reduce_kernel<2><<<g, 256, 0, stream>>>(ws, split_stride, splits, out, n, out_b, n_a); break;
    default: splitk_reduce_kernel<1><<<g, 256, 0, stream>>>(ws, split_stride, splits, out, n, out_b, n_a); break;
  }
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

}  // namespace ub
