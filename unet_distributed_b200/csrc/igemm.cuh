// Implicit-GEMM tile engine used for: 3x3 conv forward, 3x3 conv dgrad, 2x2/s2 transposed-conv forward and dgrad.
//
//   D[128 pixels, BLOCK_N channels] = sum over K-blocks  A[128 pixels, kc] * B[BLOCK_N, kc]^T
//
// A (activations, NHWC bf16) is fetched by TMA straight from the activation tensor: one box of TNxTHxTW pixels x kc
// channels per (tap, source, channel-block); the tap is a coordinate shift and image borders are TMA zero fill, so no
// im2col buffer and no concat buffer ever exists (two sources = two tensor maps walked by the same K loop).
// B (packed bf16 weights, [Cout][tap][Cin]) is a 3-D tensor map. Accumulators live in TMEM (double buffered so the
// epilogue of tile i overlaps the MMAs of tile i+1); tcgen05.mma is issued by one thread.
#pragma once
#include <cstdint>
#include <cuda.h>

namespace ub {

enum : int { IG_A_SHIFT4D = 0, IG_A_TCONV5D = 1, IG_A_TCONV_PAIR = 2 };   // PAIR: fine_pair_map (ops.cu), taps_s = 1

struct IgemmEpilogue {
  // Up to two destinations (concat dgrad writes the `up` half and the `skip` half to different tensors).
  void* dst[2];
  const void* aux[2];  // optional bf16 tensors, same pixel grid as dst: out = aux>0 ? acc*mask_scale : 0
  int dst_C[2];
  int dst_coff[2];
  int aux_C[2];
  int aux_coff[2];
  // ReLU gates as bits: one uint16 per (pixel, 16-channel group), bit i = channel 16g+i is > 0. A forward epilogue
  // can write them next to its output (gate_out); a dgrad epilogue reads them (gate_in) instead of re-reading the
  // producer's whole activation tensor through `aux` -- 1/16 of the bytes and one 2-byte load per 16 columns.
  unsigned short* gate_out;
  const unsigned short* gate_in[2];
  // dgrad, conv_halo fast path only: destination 0 is the COARSE tensor [N, H/2, W/2, dst_C[0]] and receives the sum
  // of each 2x2 block of the gradient (ResizeNearestNeighborGrad fused into the producer); gate_in[0] then holds the
  // coarse tensor's ReLU gates
  int pool2_slot0;
  // forward, conv_halo_rp only: also write max over each 2x2 block of the (ReLU'd) output to this coarse tensor
  // [N, H/2, W/2, dst_C[0]] (MaxPool2D fused into the producer)
  void* maxpool_out;
  int split_col;       // GEMM columns >= split_col go to slot 1
  const float* bias;   // fp32 per GEMM column (tconv: per column modulo d2s_cout), or null
  int relu;
  float mask_scale;
  int d2s_cout;        // >0: depth-to-space store; GEMM column = (a*2+b)*d2s_cout + o
  float drop_keep;     // <1: dropout with keep probability drop_keep (scaled by 1/keep)
  unsigned long long drop_seed;
  const long long* drop_seed_add;  // optional device counter added to drop_seed (graph-replayable per-step masks)
  unsigned int drop_stream;
  // Optional: the keep decisions of this dropout layer precomputed by dropout_keep_bits16() -- one uint16 per 16
  // consecutive output elements (bit i = keep element 16g + i), the same element numbering as the Philox stream.
  // The epilogue then loads two bytes per 16 columns instead of running Philox4x32-10 twice (ncu r02: conv3a's forward
  // epilogue spent more cycles on Philox than the tile's MMAs took: 53 us at 17 % tensor-pipe activity).
  const unsigned short* drop_bits;
};

struct IgemmParams {
  int N, H, W;                      // pixel grid the M dimension walks
  int TN, TH, TW;                   // tile box, TN*TH*TW == 128
  int tiles_n, tiles_h, tiles_w;
  int n_tiles;                      // tiles along GEMM N
  int block_n;                      // 32/64/128/256
  int a_mode;
  int taps_r, taps_s, pad;          // tap grid and the shift origin (3,3,1 for conv; 1,1,0 plain; 2,2,0 tconv dgrad)
  int num_src;
  int src_kblocks[2];               // channel blocks per source
  int src_ci_off[2];                // channel offset of the source inside the packed weight row
  int kc;                           // K block in elements: 128 B (SWIZZLE_128B) or 64 B (SWIZZLE_64B) of elt-byte elements
  int elt;                          // element bytes of A, B and the outputs: 2 = bf16 (0 means 2), 4 = fp32 as tf32
  int stages;
  int lean;                         // set by igemm_launch: the lean epilogue applies (see igemm.cu)
  IgemmEpilogue epi;
};

// Launch on `stream`; tensor maps are passed by value (kernel parameter space).
int igemm_launch(const IgemmParams& p, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                 cudaStream_t stream);

int igemm_pick_stages(int block_n, int kc, int elt = 2);

// Philox-based keep decision shared by the conv epilogue and the stand-alone mask generator.
__host__ __device__ inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = static_cast<uint64_t>(M0) * c0;
    uint64_t p1 = static_cast<uint64_t>(M1) * c2;
    uint32_t hi0 = static_cast<uint32_t>(p0 >> 32), lo0 = static_cast<uint32_t>(p0);
    uint32_t hi1 = static_cast<uint32_t>(p1 >> 32), lo1 = static_cast<uint32_t>(p1);
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n1 = lo1;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    uint32_t n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// Keep bits for the 8 consecutive elements of group `g` (element index = 8*g + i): bit i set = keep.
__host__ __device__ inline uint32_t dropout_keep_bits8(unsigned long long seed, uint32_t stream, unsigned long long g,
                                                       float keep) {
  uint32_t r[4];
  philox4x32_10(static_cast<uint32_t>(g), static_cast<uint32_t>(g >> 32), stream, 0x5eedu,
                static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32), r);
  const uint32_t thr = static_cast<uint32_t>(keep * 65536.0f + 0.5f);
  uint32_t bits = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    bits |= ((r[i] & 0xFFFFu) < thr ? 1u : 0u) << (2 * i);
    bits |= ((r[i] >> 16) < thr ? 1u : 0u) << (2 * i + 1);
  }
  return bits;
}

}  // namespace ub
