#include "igemm.cuh"

#include <cstdlib>

#include "common.h"
#include "ptx.cuh"

namespace ub {

namespace {

constexpr int kThreads = 320;  // warp0: TMA producer, warp1: MMA issuer + TMEM owner, warps 2..9: epilogue (two per
                               // TMEM lane quarter, alternating column groups: layers with a short K are epilogue-bound)
constexpr int kMaxStages = 8;

struct __align__(8) IgemmBarriers {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ void decode_tile(const IgemmParams& p, int tile, int& n0, int& h0, int& w0, int& nt) {
  nt = tile % p.n_tiles;
  int mt = tile / p.n_tiles;
  int tw_i = mt % p.tiles_w;
  mt /= p.tiles_w;
  int th_i = mt % p.tiles_h;
  int tn_i = mt / p.tiles_h;
  n0 = tn_i * p.TN;
  h0 = th_i * p.TH;
  w0 = tw_i * p.TW;
}

__global__ void __launch_bounds__(kThreads, 1)
igemm_kernel(const __grid_constant__ IgemmParams p, const __grid_constant__ CUtensorMap tmA0,
             const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB) {
  pdl_trigger();     // the next kernel of the stream may take SMs as this grid's CTAs leave them
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

  const int elt = p.elt == 4 ? 4 : 2;        // operand / activation element: bf16, or fp32 read as tf32
  const int a_bytes = 128 * p.kc * elt;
  const int b_bytes = p.block_n * p.kc * elt;
  const int stage_bytes = a_bytes + b_bytes;
  IgemmBarriers* bars = reinterpret_cast<IgemmBarriers*>(smem + p.stages * stage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_n * p.tiles_h * p.tiles_w * p.n_tiles;
  const int kb_per_tap = p.src_kblocks[0] + (p.num_src > 1 ? p.src_kblocks[1] : 0);
  const int num_kb = p.taps_r * p.taps_s * kb_per_tap;
  const uint32_t tmem_cols = (2 * p.block_n) < 32 ? 32u : static_cast<uint32_t>(2 * p.block_n);

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], 1);
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
    // ------------------------------------------------------------ TMA producer (one thread)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int n0, h0, w0, nt;
        decode_tile(p, tile, n0, h0, w0, nt);
        for (int tap = 0; tap < p.taps_r * p.taps_s; ++tap) {
          const int tr = tap / p.taps_s, ts = tap % p.taps_s;
          for (int src = 0; src < p.num_src; ++src) {
            const CUtensorMap* tmA = src == 0 ? &tmA0 : &tmA1;
            for (int cb = 0; cb < p.src_kblocks[src]; ++cb) {
              mbar_wait(&bars->empty[stage], phase ^ 1u);
              uint8_t* sa = smem + stage * stage_bytes;
              uint8_t* sb = sa + a_bytes;
              mbar_arrive_expect_tx(&bars->full[stage], static_cast<uint32_t>(stage_bytes));
              if (p.a_mode == IG_A_SHIFT4D) {
                tma_load_4d(sa, tmA, &bars->full[stage], cb * p.kc, w0 + ts - p.pad, h0 + tr - p.pad, n0);
              } else if (p.a_mode == IG_A_TCONV_PAIR) {
                tma_load_4d(sa, tmA, &bars->full[stage], cb * p.kc, w0, tr, n0 * p.H + h0);
              } else {
                tma_load_5d(sa, tmA, &bars->full[stage], cb * p.kc, ts, w0, tr, n0 * p.H + h0);
              }
              tma_load_3d(sb, &tmB, &bars->full[stage], p.src_ci_off[src] + cb * p.kc, tap, nt * p.block_n);
              if (++stage == p.stages) { stage = 0; phase ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    // The warp walks the uniform control flow together; one elected lane issues (bare UTCHMMAs, uniform registers).
    const uint32_t idesc = elt == 4 ? umma_idesc_tf32(128, static_cast<uint32_t>(p.block_n), 0, 0)
                                    : umma_idesc_bf16(128, static_cast<uint32_t>(p.block_n), 0, 0);
    const int row_bytes = p.kc * elt;                          // one pixel's K block: 128 B (SWIZZLE_128B) or 64 B
    const uint32_t sw = row_bytes == 128 ? UMMA_SW_128B : UMMA_SW_64B;
    const uint32_t sbo = static_cast<uint32_t>(8 * row_bytes);  // 8 rows of kc elements
    const uint64_t desc_hi = umma_smem_desc(0, 0, sbo, sw);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      mbar_wait(&bars->tmem_empty[acc], (((it >> 1) & 1) ^ 1u));
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * p.block_n);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&bars->full[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * stage_bytes);
        const uint64_t adesc = desc_hi + (sa >> 4);
        const uint64_t bdesc = desc_hi + ((sa + static_cast<uint32_t>(a_bytes)) >> 4);
        if (elect_one()) {
          const uint32_t acc_flag = kb ? 1u : 0u;
          if (elt == 4) {   // kind::tf32: K = 8 fp32 words = 32 bytes per instruction, like 16 bf16
            umma_tf32(d_tmem, adesc, bdesc, idesc, acc_flag);
            umma_tf32(d_tmem, adesc + 2, bdesc + 2, idesc, 1u);
            if (row_bytes == 128) {
              umma_tf32(d_tmem, adesc + 4, bdesc + 4, idesc, 1u);
              umma_tf32(d_tmem, adesc + 6, bdesc + 6, idesc, 1u);
            }
          } else {
          umma_f16(d_tmem, adesc, bdesc, idesc, acc_flag);
          umma_f16(d_tmem, adesc + 2, bdesc + 2, idesc, 1u);
          if (p.kc == 64) {
            umma_f16(d_tmem, adesc + 4, bdesc + 4, idesc, 1u);
            umma_f16(d_tmem, adesc + 6, bdesc + 6, idesc, 1u);
          }
          }
          umma_commit(&bars->empty[stage]);
          if (kb == num_kb - 1) umma_commit(&bars->tmem_full[acc]);
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue (4 warps, one TMEM lane quarter each)
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;   // which of the two warps of this quarter
    const int m = quarter * 32 + lane;
    const int tw = m % p.TW;
    const int th = (m / p.TW) % p.TH;
    const int tn = m / (p.TW * p.TH);
    const IgemmEpilogue& e = p.epi;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      int n0, h0, w0, nt;
      decode_tile(p, tile, n0, h0, w0, nt);
      const int n = n0 + tn, h = h0 + th, w = w0 + tw;
      const bool valid = (n < p.N) && (h < p.H) && (w < p.W);
      const int col0 = nt * p.block_n;  // first GEMM column of this tile

      if (p.lean) {
        // Lean epilogue: full tiles, 32-byte aligned destinations, no dropout, no tensor masks. Everything that does
        // not depend on the 16-column chunk is computed once per tile, two chunks are in flight per TMEM wait and a
        // lane stores one full 32-byte sector per chunk. (ncu: the general code below spends ~300 instructions per
        // chunk with one warp per scheduler -- ~1000 cycles per chunk; the transposed-conv layers, whose K is 64,
        // were bound by it: 0.125 ms for the 64x64 -> 128x128 layer.)
        int slot = 0, chb = col0;
        if (e.d2s_cout == 0 && col0 >= e.split_col) { slot = 1; chb = col0 - e.split_col; }
        const size_t pix = (static_cast<size_t>(n) * p.H + h) * p.W + w;
        char* dbase = reinterpret_cast<char*>(e.dst[slot]) + (pix * e.dst_C[slot] + e.dst_coff[slot] + chb) * 2;
        // depth-to-space: byte offsets of the four sub-pixels (a, b) from sub-pixel (0, 0) of this lane's pixel
        char* d2s_base = nullptr;
        size_t sub_off[3] = {0, 0, 0};   // [1]: next fine column, [2]: next fine row
        if (e.d2s_cout > 0) {
          const size_t fpix = (static_cast<size_t>(n) * (2 * p.H) + 2 * h) * (2 * p.W) + 2 * w;
          d2s_base = reinterpret_cast<char*>(e.dst[0]) + (fpix * e.dst_C[0] + e.dst_coff[0]) * 2;
          const size_t px = static_cast<size_t>(e.dst_C[0]) * 2;
          sub_off[1] = px; sub_off[2] = px * 2 * p.W;
        }
        const unsigned short* gin = nullptr;
        if (e.gate_in[slot]) gin = e.gate_in[slot] + ((pix * e.aux_C[slot] + e.aux_coff[slot] + chb) >> 4);
        unsigned short* gout = nullptr;
        if (e.gate_out) gout = e.gate_out + ((pix * e.dst_C[0] + e.dst_coff[0] + col0) >> 4);
        const float scale = e.mask_scale;
        const unsigned short* dbits = nullptr;   // precomputed dropout keep bits (forward, single plain destination)
        if (e.drop_keep < 1.0f) dbits = e.drop_bits + ((pix * e.dst_C[0] + e.dst_coff[0] + col0) >> 4);
        const float inv_keep = e.drop_keep < 1.0f ? 1.0f / e.drop_keep : 1.0f;
        const int d2s_shift = e.d2s_cout > 0 ? 31 - __clz(e.d2s_cout) : 0;
        mbar_wait(&bars->tmem_full[acc], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                               static_cast<uint32_t>(acc * p.block_n);
        for (int c = half * 32; c < p.block_n; c += 64) {
          uint32_t r[2][16];
          uint32_t gbits[2] = {0xFFFFu, 0xFFFFu};
          tmem_ld_32x32b_x16(t_row + static_cast<uint32_t>(c), r[0]);
          tmem_ld_32x32b_x16(t_row + static_cast<uint32_t>(c + 16), r[1]);
          if (gin) {
            gbits[0] = __ldg(gin + (c >> 4));
            gbits[1] = __ldg(gin + (c >> 4) + 1);
          }
          float4 bq[2][4];   // bias of both chunks, requested before the accumulators are waited for
          if (e.bias) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int cc = c + 16 * u;
              const int bcol = (e.d2s_cout > 0) ? ((col0 + cc) & (e.d2s_cout - 1)) : col0 + cc;   // Cout is a power of 2
              const float4* bp = reinterpret_cast<const float4*>(e.bias + bcol);
#pragma unroll
              for (int q = 0; q < 4; ++q) bq[u][q] = __ldg(bp + q);
            }
          }
          tmem_ld_wait();
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int cc = c + 16 * u;
            float v[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[u][i]);
            if (e.bias) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                v[4 * q + 0] += bq[u][q].x; v[4 * q + 1] += bq[u][q].y; v[4 * q + 2] += bq[u][q].z; v[4 * q + 3] += bq[u][q].w;
              }
            }
            if (gin) {
#pragma unroll
              for (int i = 0; i < 16; ++i) v[i] = ((gbits[u] >> i) & 1u) ? v[i] * scale : 0.0f;
            }
            if (dbits) {
              const uint32_t kb = __ldg(dbits + (cc >> 4));
#pragma unroll
              for (int i = 0; i < 16; ++i) v[i] = ((kb >> i) & 1u) ? v[i] * inv_keep : 0.0f;
            }
            U32x8 o;
            if (e.relu) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(o.v[i]) : "f"(v[2 * i + 1]), "f"(v[2 * i]));
              }
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i) o.v[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
            }
            if (gout) {
              uint32_t accb = 0;
#pragma unroll
              for (int i = 0; i < 8; ++i) accb |= ((o.v[i] + 0x7FFF7FFFu) >> (15 - 2 * i)) & (0x00010001u << (2 * i));
              gout[cc >> 4] = static_cast<unsigned short>((accb & 0xFFFFu) | (accb >> 15));
            }
            if (e.d2s_cout > 0) {
              const int col = col0 + cc;
              const int ab = col >> d2s_shift;
              const size_t so = ((ab & 1) ? sub_off[1] : 0) + ((ab & 2) ? sub_off[2] : 0);
              st_global_v8(d2s_base + so + (col & (e.d2s_cout - 1)) * 2, o);
            } else {
              st_global_v8(dbase + cc * 2, o);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
        continue;
      }

      // Destination slot, pixel index and channel origin.
      int slot = 0, ch0 = col0;
      size_t pix;
      if (e.d2s_cout > 0) {
        const int ab = col0 / e.d2s_cout;
        ch0 = col0 % e.d2s_cout;
        pix = (static_cast<size_t>(n) * (2 * p.H) + (2 * h + (ab >> 1))) * (2 * p.W) + (2 * w + (ab & 1));
      } else {
        if (col0 >= e.split_col) { slot = 1; ch0 = col0 - e.split_col; }
        pix = (static_cast<size_t>(n) * p.H + h) * p.W + w;
      }
      // byte pointers: elements are bf16 or (fp32/tf32 mode) fp32
      char* dst = reinterpret_cast<char*>(e.dst[slot]) + (pix * e.dst_C[slot] + e.dst_coff[slot] + ch0) * elt;
      const char* aux = nullptr;
      if (e.aux[slot]) {
        aux = reinterpret_cast<const char*>(e.aux[slot]) + (pix * e.aux_C[slot] + e.aux_coff[slot] + ch0) * elt;
      }
      const size_t drop_elem0 = pix * e.dst_C[slot] + e.dst_coff[slot] + ch0;
      // (a depth-to-space tile wider than Cout starts at ab = 0, ch0 = 0: bias index = column mod Cout)
      const int bias0 = (e.d2s_cout > 0) ? ch0 : col0;

      mbar_wait(&bars->tmem_full[acc], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                             static_cast<uint32_t>(acc * p.block_n);
      for (int c = half * 16; c < p.block_n; c += 32) {
        uint32_t r[16];
        tmem_ld_32x32b_x16(t_row + static_cast<uint32_t>(c), r);
        tmem_ld_wait();
        if (e.d2s_cout > 0 && p.block_n > e.d2s_cout) {
          // depth-to-space tile spanning several (a, b) sub-pixels: destination of this 16-column chunk
          const int col = col0 + c;
          const int ab = col / e.d2s_cout;
          const size_t px = (static_cast<size_t>(n) * (2 * p.H) + (2 * h + (ab >> 1))) * (2 * p.W) + (2 * w + (ab & 1));
          dst = reinterpret_cast<char*>(e.dst[0]) +
                (static_cast<long long>(px * e.dst_C[0] + e.dst_coff[0] + (col % e.d2s_cout)) - c) * elt;
        }
        if (valid) {
          float v[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
          if (e.bias) {
            const int bcol = (e.d2s_cout > 0) ? (bias0 + c) % e.d2s_cout : bias0 + c;
            const float4* bp = reinterpret_cast<const float4*>(e.bias + bcol);
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
          if (e.drop_keep < 1.0f) {
            const float inv = 1.0f / e.drop_keep;
            const unsigned long long seed =
                e.drop_seed + (e.drop_seed_add ? static_cast<unsigned long long>(__ldg(e.drop_seed_add)) : 0ull);
            if (e.drop_bits) {
              const uint32_t bits = __ldg(e.drop_bits + ((drop_elem0 + c) >> 4));
#pragma unroll
              for (int i = 0; i < 16; ++i) v[i] = ((bits >> i) & 1u) ? v[i] * inv : 0.0f;
            }
#pragma unroll
            for (int g = 0; g < 2 && !e.drop_bits; ++g) {
              const uint32_t bits = dropout_keep_bits8(seed, e.drop_stream, (drop_elem0 + c) / 8 + g, e.drop_keep);
#pragma unroll
              for (int i = 0; i < 8; ++i) v[8 * g + i] = ((bits >> i) & 1u) ? v[8 * g + i] * inv : 0.0f;
            }
          }
          if (e.gate_out && e.d2s_cout == 0) {   // ReLU gates of this output as bits (see igemm.cuh)
            uint32_t bits = 0;
#pragma unroll
            for (int i = 0; i < 16; ++i) bits |= (v[i] > 0.0f ? 1u : 0u) << i;
            e.gate_out[(drop_elem0 + c) >> 4] = static_cast<unsigned short>(bits);
          }
          if (e.gate_in[slot]) {
            const uint32_t b = __ldg(e.gate_in[slot] + ((pix * e.aux_C[slot] + e.aux_coff[slot] + ch0 + c) >> 4));
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = ((b >> i) & 1u) ? v[i] * e.mask_scale : 0.0f;
          }
          if (aux && elt == 4) {
            const float4* ap = reinterpret_cast<const float4*>(aux + c * 4);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float4 a = __ldg(ap + q);
              v[4 * q + 0] = a.x > 0.0f ? v[4 * q + 0] * e.mask_scale : 0.0f;
              v[4 * q + 1] = a.y > 0.0f ? v[4 * q + 1] * e.mask_scale : 0.0f;
              v[4 * q + 2] = a.z > 0.0f ? v[4 * q + 2] * e.mask_scale : 0.0f;
              v[4 * q + 3] = a.w > 0.0f ? v[4 * q + 3] * e.mask_scale : 0.0f;
            }
          } else if (aux) {
            const uint4* ap = reinterpret_cast<const uint4*>(aux + c * 2);
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              const uint4 a = __ldg(ap + g);
              const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                v[8 * g + 2 * i] = bf16_lo(aw[i]) > 0.0f ? v[8 * g + 2 * i] * e.mask_scale : 0.0f;
                v[8 * g + 2 * i + 1] = bf16_hi(aw[i]) > 0.0f ? v[8 * g + 2 * i + 1] * e.mask_scale : 0.0f;
              }
            }
          }
          if (elt == 4) {   // fp32 output, rounded to tf32 for the MMAs that read it next
            float4* dp4 = reinterpret_cast<float4*>(dst + c * 4);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              dp4[q] = make_float4(round_tf32(v[4 * q]), round_tf32(v[4 * q + 1]), round_tf32(v[4 * q + 2]),
                                   round_tf32(v[4 * q + 3]));
            }
          }
          uint4* dp = reinterpret_cast<uint4*>(dst + c * 2);
#pragma unroll
          for (int g = 0; g < 2 && elt == 2; ++g) {
            uint4 o;
            o.x = pack_bf16x2(v[8 * g + 0], v[8 * g + 1]);
            o.y = pack_bf16x2(v[8 * g + 2], v[8 * g + 3]);
            o.z = pack_bf16x2(v[8 * g + 4], v[8 * g + 5]);
            o.w = pack_bf16x2(v[8 * g + 6], v[8 * g + 7]);
            dp[g] = o;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
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

int igemm_launch(const IgemmParams& p_in, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                 cudaStream_t stream) {
  IgemmParams p = p_in;
  UB_REQUIRE(p.TN * p.TH * p.TW == 128, "igemm: tile box %dx%dx%d is not 128 pixels", p.TN, p.TH, p.TW);
  UB_REQUIRE(p.block_n == 32 || p.block_n == 64 || p.block_n == 128 || p.block_n == 256,
             "igemm: block_n %d unsupported", p.block_n);
  UB_REQUIRE(p.kc == 16 || p.kc == 32 || p.kc == 64, "igemm: kc %d unsupported", p.kc);
  UB_REQUIRE(p.stages >= 2 && p.stages <= kMaxStages, "igemm: stages %d out of range", p.stages);
  UB_REQUIRE(p.elt == 0 || p.elt == 2 || p.elt == 4, "igemm: element size %d", p.elt);
  if (p.elt == 0) p.elt = 2;
  UB_REQUIRE(p.kc * p.elt == 128 || p.kc * p.elt == 64, "igemm: K block of %d bytes (kc %d)", p.kc * p.elt, p.kc);
  const int stage_bytes = (128 + p.block_n) * p.kc * p.elt;
  const size_t smem = static_cast<size_t>(p.stages) * stage_bytes + sizeof(IgemmBarriers) + 1024;
  UB_REQUIRE(smem <= 227 * 1024, "igemm: %zu bytes of shared memory requested", smem);
  static bool attr_set[64] = {false};
  int dev = 0;
  UB_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    UB_CHECK_CUDA(cudaFuncSetAttribute(igemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev] = true;
  }
  const int num_tiles = p.tiles_n * p.tiles_h * p.tiles_w * p.n_tiles;
  UB_REQUIRE(num_tiles > 0, "igemm: empty problem");
  {
    static const bool no_lean = getenv("UNETB200_NO_LEAN_IGEMM") != nullptr;
    const IgemmEpilogue& e = p.epi;
    const bool full = p.N % p.TN == 0 && p.H % p.TH == 0 && p.W % p.TW == 0;
    bool aligned = reinterpret_cast<uintptr_t>(e.dst[0]) % 32 == 0 && reinterpret_cast<uintptr_t>(e.dst[1]) % 32 == 0 &&
                   e.dst_coff[0] % 16 == 0 && e.dst_coff[1] % 16 == 0 && e.dst_C[0] % 16 == 0 && e.dst_C[1] % 16 == 0;
    // two destinations must meet at a tile boundary; gate bits are only written for a single plain destination
    const bool split_ok = e.split_col >= p.n_tiles * p.block_n || e.split_col % p.block_n == 0;
    const bool gate_ok = !e.gate_out || (e.d2s_cout == 0 && e.split_col >= p.n_tiles * p.block_n);
    p.lean = (!no_lean && p.elt == 2 && full && aligned && split_ok && gate_ok && p.block_n >= 32 && !e.aux[0] && !e.aux[1] &&
              (e.drop_keep >= 1.0f || (e.drop_bits && e.d2s_cout == 0 && e.split_col >= p.n_tiles * p.block_n)) &&
              (e.d2s_cout == 0 || (e.d2s_cout >= 32 && (e.d2s_cout & (e.d2s_cout - 1)) == 0))) ? 1 : 0;
  }
  int grid = device_sm_count();
  if (grid > num_tiles) grid = num_tiles;
  if (profiling_enabled()) {
    const double pixels = static_cast<double>(p.N) * p.H * p.W;
    const double kdim = static_cast<double>(p.taps_r) * p.taps_s * p.kc *
                        (p.src_kblocks[0] + (p.num_src > 1 ? p.src_kblocks[1] : 0));
    const double ncols = static_cast<double>(p.n_tiles) * p.block_n;
    prof_begin(UB_KIND_IGEMM, 2.0 * pixels * kdim * ncols,
               static_cast<double>(p.elt) * pixels * (kdim / (p.taps_r * p.taps_s) + ncols), stream);
  }
  UB_CHECK_CUDA(launch_pdl(igemm_kernel, dim3(grid), dim3(kThreads), smem, stream, p, a0, a1, b));
  UB_CHECK_CUDA(cudaGetLastError());
  prof_end(stream);
  return UB_OK;
}

int igemm_pick_stages(int block_n, int kc, int elt) {
  const int stage_bytes = (128 + block_n) * kc * elt;
  int s = (200 * 1024) / stage_bytes;
  if (s > kMaxStages) s = kMaxStages;
  if (s < 2) s = 2;
  return s;
}

}  // namespace ub
