// Layer-level operations built on the igemm / wgrad tile engines: shape -> tile choice -> tensor maps -> launch.
// `dt` (UB_DT_BF16 = 0 default, UB_DT_TF32 = 1) is the element type of activations, gradients and packed weights:
// bf16, or fp32 holding tf32-rounded values that the tensor cores read with kind::tf32 (the fp32/tf32 precision mode;
// there every conv runs on the generic implicit-GEMM / per-tap weight-gradient kernels).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace ub {

struct ConvSrc {
  const void* ptr;  // NHWC bf16
  int C;
};

struct DropoutCfg {
  float keep;  // 1.0 = off
  unsigned long long seed;
  unsigned int stream;
  const long long* seed_add;  // optional device counter added to seed
  // optional: keep bits of this layer's output precomputed with dropout_keep_bits16 (same seed / stream / numbering);
  // the conv epilogue then loads them instead of running Philox (elementwise.cuh)
  const unsigned short* bits = nullptr;
};

// y = relu?(conv3x3_same([x0, x1]) + bias) (* dropout); w_packed = bf16 [Cout][9][C0+C1].
// maxpool_out (optional): also write MaxPool2D(y) to this [N, H/2, W/2, Cout] tensor from the conv's epilogue; only where
// conv3x3_fwd_maxpool_ok() says so (the row-pair kernel).
bool conv3x3_fwd_maxpool_ok(int N, int H, int W, int C0, int C1, int Cout);
int conv3x3_fwd(ConvSrc x0, ConvSrc x1, const void* w_packed, const float* bias, void* y, int N, int H, int W, int Cout,
                int relu, DropoutCfg drop, cudaStream_t s, unsigned short* gate_out = nullptr, void* maxpool_out = nullptr,
                int dt = 0);

// dx_k = conv3x3_same(dz, flipped weights)[channels of source k] * (mask_k > 0 ? scale_k : 0)   (mask optional)
// w_packed_d = bf16 [C0+C1][9][Cout] (tap-flipped).
struct DgradDst {
  void* ptr;         // NHWC bf16, C channels
  int C;
  const void* mask;  // same shape as ptr, or null
  float scale;
  const unsigned short* gate = nullptr;  // ReLU gates of the producer as bits (takes precedence over mask)
  // destination 0 only: ptr is the COARSE tensor [N, H/2, W/2, C]; it receives the 2x2 block sums of the gradient
  // (ResizeNearestNeighborGrad fused into the dgrad), gated by `gate` (coarse). Needs conv3x3_dgrad_pool2_ok().
  int pool2 = 0;
};
bool conv3x3_dgrad_pool2_ok(int N, int H, int W, int Cout, int C0, int C1);
int conv3x3_dgrad(const void* dz, int Cout, const void* w_packed_d, DgradDst d0, DgradDst d1, int N, int H, int W,
                  cudaStream_t s, int dt = 0);

// dw (fp32 [3][3][C0+C1][Cout]) = correlation of [x0,x1] with dz over all pixels. ws: conv3x3_wgrad_ws_bytes().
size_t conv3x3_wgrad_ws_bytes(int N, int H, int W, int Cin_total, int Cout, bool with_bias = false);
// db (optional): also the bias gradient sum_pixels dz, fused into the halo kernel where that one runs.
int conv3x3_wgrad(const void* dz, int Cout, ConvSrc x0, ConvSrc x1, float* dw, void* ws, size_t ws_bytes, int N, int H,
                  int W, cudaStream_t s, float* db = nullptr, int dt = 0);

// Conv2DTranspose 2x2 stride 2: y[n,2i+a,2j+b,o] = sum_c x[n,i,j,c] w[a,b,o,c] + bias[o].
// w_packed = bf16 [4*Cout][Cin] (cast of the TF layout); (N,H,W) is the input grid.
int tconv2x2_fwd(const void* x, int Cin, const void* w_packed, const float* bias, void* y, int Cout, int N, int H, int W,
                 cudaStream_t s, int dt = 0);
// dx[n,i,j,c] = sum_{a,b,o} dy[n,2i+a,2j+b,o] w[a,b,o,c], masked by (mask>0). w_packed_d = bf16 [Cin][4][Cout].
int tconv2x2_dgrad(const void* dy, int Cout, const void* w_packed_d, void* dx, int Cin, const void* mask, float scale,
                   int N, int H, int W, cudaStream_t s, const unsigned short* gate = nullptr, int dt = 0);
// dw fp32 [2][2][Cout][Cin].
size_t tconv2x2_wgrad_ws_bytes(int N, int H, int W, int Cin, int Cout);
// db (optional): also the bias gradient, summed from the dY tiles the kernel stages (bf16) or by a second pass (tf32).
int tconv2x2_wgrad(const void* dy, int Cout, const void* x, int Cin, float* dw, void* ws, size_t ws_bytes, int N, int H,
                   int W, cudaStream_t s, int dt = 0, float* db = nullptr);

}  // namespace ub
