// Stand-alone dice reductions for the eager API (model.dice_coef / dice_coef_loss on arbitrary fp32 tensors).
// The training plan uses the fused head kernels instead; these exist so that the reference's call pattern
// `dice_coef_loss(msks, preds)` (test_dist.py:211-212) never touches a CPU or a generic tensor library.
#include "../../include/unetb200.h"
#include "common.h"
#include "ptx.cuh"

using namespace ub;

namespace {

constexpr int kBlocks = 296;
constexpr int kThreads = 256;

__global__ void dice_sums_kernel(const float* __restrict__ y, const float* __restrict__ p, long long n,
                                 float* __restrict__ ws) {
  float sI = 0.f, sY = 0.f, sP = 0.f;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float yv = __ldg(y + i), pv = __ldg(p + i);
    sI = fmaf(yv, pv, sI);
    sY += yv;
    sP += pv;
  }
  sI = warp_sum(sI); sY = warp_sum(sY); sP = warp_sum(sP);
  __shared__ float red[3 * (kThreads / 32)];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { red[warp * 3] = sI; red[warp * 3 + 1] = sY; red[warp * 3 + 2] = sP; }
  __syncthreads();
  if (threadIdx.x < 3) {
    float a = 0.f;
    for (int w = 0; w < kThreads / 32; ++w) a += red[w * 3 + threadIdx.x];
    ws[blockIdx.x * 3 + threadIdx.x] = a;
  }
}

__global__ void dice_sums_final_kernel(const float* __restrict__ ws, float* __restrict__ sums) {
  if (threadIdx.x < 3) {
    float a = 0.f;
    for (int b = 0; b < kBlocks; ++b) a += ws[b * 3 + threadIdx.x];
    sums[threadIdx.x] = a;
  }
}

// which: 0 loss, 1 dice, 2 sensitivity, 3 specificity. dp = upstream * d(metric)/dp (elementwise, closed form).
__global__ void dice_grad_kernel(const float* __restrict__ y, const float* __restrict__ sums, float smooth, int which,
                                 const float* __restrict__ upstream, float* __restrict__ dp, long long n) {
  const float I = sums[0], Sy = sums[1], Sp = sums[2];
  const float up = upstream ? *upstream : 1.0f;
  float a, b;  // d/dp = a*y + b
  if (which == 0) {
    a = -2.0f / (2.0f * I + smooth);
    b = 1.0f / (Sy + Sp + smooth);
  } else if (which == 1) {
    const float den = Sy + Sp + smooth;
    a = 2.0f / den;
    b = -(2.0f * I + smooth) / (den * den);
  } else if (which == 2) {
    a = 1.0f / (Sy + smooth);
    b = 0.0f;
  } else {
    const float den = Sp + smooth;
    a = 1.0f / den;
    b = -(I + smooth) / (den * den);
  }
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    dp[i] = up * fmaf(a, __ldg(y + i), b);
  }
}

}  // namespace

extern "C" {

size_t unetb200_dice_sums_ws_bytes(void) { return static_cast<size_t>(kBlocks) * 3 * sizeof(float); }

int unetb200_dice_sums(const float* y_true, const float* y_pred, long long n, float* sums, void* ws, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  dice_sums_kernel<<<kBlocks, kThreads, 0, s>>>(y_true, y_pred, n, reinterpret_cast<float*>(ws));
  UB_CHECK_CUDA(cudaGetLastError());
  dice_sums_final_kernel<<<1, 32, 0, s>>>(reinterpret_cast<const float*>(ws), sums);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int unetb200_dice_grad(const float* y_true, const float* sums, float smooth, int which, const float* upstream,
                       float* dp, long long n, void* stream) {
  UB_REQUIRE(which >= 0 && which <= 3, "dice_grad: metric %d outside 0..3", which);
  dice_grad_kernel<<<kBlocks, kThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(y_true, sums, smooth, which,
                                                                                      upstream, dp, n);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

}  // extern "C"
