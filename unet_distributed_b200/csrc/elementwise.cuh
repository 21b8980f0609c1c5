// `dt` (UB_DT_BF16 = 0 default, UB_DT_TF32 = 1): activations are bf16, or fp32 holding tf32-rounded values (the
// fp32/tf32 precision mode; there the generic 8-channel kernels run, as templates over the 8-element vector type).
// HBM-bound kernels of the UNet step: first-layer direct conv, pooling, upsampling, the fused 1x1-sigmoid-dice head,
// bias gradients, weight packing and the fused TF-style Adam update. All activations are NHWC bf16 unless noted.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ub {

// conv1a: direct 3x3 'same' conv on the fp32 network input (Cin in 1..8), fp32 HWIO weights, bias, ReLU -> bf16.
// gate_out (optional, only for Cout == 32 and Cin in {1, 4}): ReLU gates as bits, uint16 per 16 channels.
int conv_first_fwd(const float* x, const float* w_hwio, const float* bias, void* y_bf16, int N, int H, int W, int Cin,
                   int Cout, cudaStream_t s, unsigned short* gate_out = nullptr, int dt = 0);
bool conv_first_fwd_writes_gates(int Cin, int Cout);
// dW[3][3][Cin][Cout] and db[Cout] from fp32 x and bf16 dz; ws needs conv_first_wgrad_ws_floats() floats.
long long conv_first_wgrad_ws_floats(int Cin, int Cout);
int conv_first_wgrad(const float* x, const void* dz_bf16, float* dw, float* db, float* ws, int N, int H, int W, int Cin,
                     int Cout, cudaStream_t s, int dt = 0);

int maxpool2x2_fwd(const void* x, void* y, int N, int H, int W, int C, cudaStream_t s, int dt = 0);
// dz_full = (skip_grad or 0) + route(dpool) * (x > 0); skip_grad may alias dz_full; may be null.
int maxpool2x2_bwd(const void* x, const void* dpool, const void* skip_grad, void* dz_full, int N, int H, int W, int C,
                   cudaStream_t s, int dt = 0);

int upsample2x2_fwd(const void* x, void* y, int N, int H, int W, int C, cudaStream_t s, int dt = 0);
// dx = (sum of the 2x2 block of dy) * (x > 0 ? 1 : 0) when relu_mask != 0.
int upsample2x2_bwd(const void* dy, const void* x, void* dx, int N, int H, int W, int C, int relu_mask, cudaStream_t s,
                    int dt = 0);

// Head: p = sigmoid(x . w + b) for n_cl classes (x bf16 [P][C], w fp32 [C][n_cl]); p stored fp32 [P][n_cl].
// Accumulates the three dice sums (sum y*p, sum y, sum p) over all P*n_cl elements into sums[3] (fp32, device),
// through per-block partials in ws (head_ws_floats() floats) reduced in a fixed order.
long long head_ws_floats(long long P, int C, int n_cl);
int head_fwd(const void* x, const float* w, const float* b, const float* y_true, float* p_out, float* sums, float* ws,
             long long P, int C, int n_cl, cudaStream_t s, int dt = 0);
// out[0] = loss, out[1] = dice, out[2] = sensitivity, out[3] = specificity (reference formulas, smooth).
int dice_finalize(const float* sums, float smooth, float* out4, cudaStream_t s);
// dlogit = (-2 y/(2I+s) + 1/(Sy+Sp+s)) * p (1-p) * grad_scale; dx = (dlogit . w^T) * (x>0) -> bf16;
// dw[C][n_cl], db[n_cl] reduced through ws.
int head_bwd(const void* x, const float* w, const float* y_true, const float* p, const float* sums, float smooth,
             float grad_scale, void* dx_bf16, float* dw, float* db, float* ws, long long P, int C, int n_cl,
             cudaStream_t s, int dt = 0);

// out[i] = sum over r < nrows of ws[r * n + i], rows added in a fixed order (partial sums a kernel left per CTA).
int rows_sum(const float* ws, int nrows, int n, float* out, cudaStream_t s);
// db[c] = sum over pixels of dz[pixel][c] (bf16 in, fp32 out); ws needs colsum_ws_floats(C) floats.
long long colsum_ws_floats(int C);
int colsum_bf16(const void* dz, float* db, float* ws, long long P, int C, cudaStream_t s);
int colsum_act(const void* dz, float* db, float* ws, long long P, int C, cudaStream_t s, int dt);   // bf16 or fp32 dz
// Same result without a workspace argument: partial sums go to a fixed device array inside the library (C <= 512).
int colsum_bf16_static(const void* dz, float* db, long long P, int C, cudaStream_t s);

// Weight packing fp32 master (TF layout) -> bf16 GEMM operand layouts.
//   conv fwd  : out[co][tap][ci] = w[tap][ci][co]
//   conv dgrad: out[ci][tap][co] = w[T-1-tap][ci][co]
//   tconv fwd : out[(ab*Cout+o)][c] = w[ab][o][c]            (cast only)
//   tconv dgrad: out[c][ab][o] = w[ab][o][c]
// dt = UB_DT_TF32: same layouts with fp32 elements rounded to tf32.
int pack_conv_fwd(const float* w, void* out, int taps, int Cin, int Cout, cudaStream_t s, int dt = 0);
int pack_conv_dgrad(const float* w, void* out, int taps, int Cin, int Cout, cudaStream_t s, int dt = 0);
int pack_tconv_fwd(const float* w, void* out, int Cin, int Cout, cudaStream_t s, int dt = 0);
int pack_tconv_dgrad(const float* w, void* out, int Cin, int Cout, cudaStream_t s, int dt = 0);

// All repacks of a network in one launch. A block owns one 32 x 32 tile of one tap of one layer: it reads the fp32
// master once (coalesced), writes the operand whose layout matches the master's straight away and the transposed
// one through shared memory, so both global sides move 64-128 contiguous bytes per warp row.
struct PackLayer {
  long long src;        // float offset of the TF-layout master inside the flat parameter buffer
  long long dst_fwd;    // byte offsets of the packed bf16 operands inside the arena
  long long dst_dgrad;
  int kind;             // 0 conv3x3 (w[tap][ci][co]), 1 transposed conv (w[ab][o][c])
  int taps, cin, cout;  // multiples of 32 channels
  int tile_start;       // first block of this layer (layers sorted by tile_start)
};
// layers_dev may point into the middle of the table: tile_base is then the tile_start of its first entry.
int pack_all(const PackLayer* layers_dev, int nlayers, int total_tiles, const float* params, void* arena,
             cudaStream_t s, int tile_base = 0, int dt = 0);

// TF-1.4 AdamOptimizer on a flat fp32 buffer. step_ptr holds the number of updates applied so far (device int64).
//   lr = const_lr ? lr0 : lr0 * lr_fraction^(step/decay_steps);  t = step+1
//   lr_t = lr*sqrt(1-b2^t)/(1-b1^t); m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; theta -= lr_t m/(sqrt(v)+eps)
// g is multiplied by grad_scale first (1/world_size for averaged data-parallel gradients).
int adam_step(float* theta, float* m, float* v, const float* g, long long n, const long long* step_ptr, float lr0,
              int const_lr, float lr_fraction, float decay_steps, float beta1, float beta2, float eps,
              float grad_scale, cudaStream_t s);
int sgd_step(float* theta, const float* g, long long n, const long long* step_ptr, float lr0, int const_lr,
             float lr_fraction, float decay_steps, float grad_scale, cudaStream_t s);
int step_increment(long long* step_ptr, cudaStream_t s);

// keep[i] = 1/0 for the dropout stream (same Philox function as the conv epilogue); n multiple of 8.
int dropout_mask(unsigned char* keep, long long n, float keep_prob, unsigned long long seed, unsigned int stream_id,
                 cudaStream_t s);

// dst[r][0..row_elems) = float(src[perm ? perm[first + r] : first + r][...]) for r in [0, rows): one shuffled batch cut
// out of a device-resident training set (data.py:29-54). src rows are fp32 or float64; perm is a device int64 array.
int gather_rows(const void* src, int src_is_f64, long long src_rows, const long long* perm, long long first, int rows,
                long long row_elems, float* dst, cudaStream_t s);

// bits[g] (uint16) = keep decisions of elements 16g .. 16g+15 of a dropout layer's output (n elements, multiple of
// 16): Philox stream (seed + *seed_add, stream_id), the numbering dropout_mask and the conv epilogues use.
int dropout_keep_bits16(unsigned short* bits, long long n, float keep_prob, unsigned long long seed,
                        const long long* seed_add, unsigned int stream_id, cudaStream_t s);

int cast_f32_to_bf16(const float* x, void* y, long long n, cudaStream_t s);
int cast_bf16_to_f32(const void* x, float* y, long long n, cudaStream_t s);

}  // namespace ub
