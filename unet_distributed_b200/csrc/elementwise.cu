#include "elementwise.cuh"

#include <cstdlib>
#include <initializer_list>

#include <atomic>
#include <mutex>

#include "common.h"
#include "igemm.cuh"
#include "ptx.cuh"

namespace ub {

namespace {

constexpr int kRedBlocks = 592;  // 4 x 148 SMs: fixed so per-block partial layouts are shape independent
constexpr int kThreads = 256;

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
  o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
  return o;
}

// Eight consecutive channels of an activation tensor: 16 bytes of bf16 (uint4) or, in the fp32/tf32 mode, 32 bytes of
// fp32 (F32x8). The generic kernels below are templates over this vector type; stores of F32x8 round to tf32 because
// the tensors they write are read by kind::tf32 MMAs (exact only on tf32-representable values).
struct __align__(32) F32x8 {
  float4 lo, hi;
};
__device__ __forceinline__ void unpackv(const uint4& v, float (&f)[8]) { unpack8(v, f); }
__device__ __forceinline__ void unpackv(const F32x8& v, float (&f)[8]) {
  f[0] = v.lo.x; f[1] = v.lo.y; f[2] = v.lo.z; f[3] = v.lo.w; f[4] = v.hi.x; f[5] = v.hi.y; f[6] = v.hi.z; f[7] = v.hi.w;
}
__device__ __forceinline__ uint4 ldv(const uint4* p) { return __ldg(p); }
__device__ __forceinline__ F32x8 ldv(const F32x8* p) {
  F32x8 v;
  v.lo = __ldg(reinterpret_cast<const float4*>(p));
  v.hi = __ldg(reinterpret_cast<const float4*>(p) + 1);
  return v;
}
template <typename V>
__device__ __forceinline__ V zerov() {
  V v;
  memset(&v, 0, sizeof(V));
  return v;
}
template <typename V>
__device__ __forceinline__ void ld8(const V* p, float (&f)[8]) { unpackv(ldv(p), f); }
// plain (coherent) load: for tensors the same kernel also writes (in-place accumulation)
__device__ __forceinline__ void ld8_rw(const uint4* p, float (&f)[8]) { unpack8(*p, f); }
__device__ __forceinline__ void ld8_rw(const F32x8* p, float (&f)[8]) { unpackv(*p, f); }
__device__ __forceinline__ void st8(uint4* p, const float (&f)[8]) { *p = pack8(f); }
__device__ __forceinline__ void st8(F32x8* p, const float (&f)[8]) {
  F32x8 v;
  v.lo = make_float4(round_tf32(f[0]), round_tf32(f[1]), round_tf32(f[2]), round_tf32(f[3]));
  v.hi = make_float4(round_tf32(f[4]), round_tf32(f[5]), round_tf32(f[6]), round_tf32(f[7]));
  *p = v;
}

inline unsigned grid_for(long long items, int threads) {
  long long b = (items + threads - 1) / threads;
  if (b < 1) b = 1;
  return static_cast<unsigned>(b);
}

// ------------------------------------------------------------------------------------------ conv1a forward
template <typename V>
__global__ void conv_first_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                      const float* __restrict__ bias, V* __restrict__ y, int N, int H, int W,
                                      int Cin, int Cout) {
  extern __shared__ float sw[];  // [9*Cin*Cout] weights then [Cout] bias
  const int nw = 9 * Cin * Cout;
  for (int i = threadIdx.x; i < nw; i += blockDim.x) sw[i] = w[i];
  for (int i = threadIdx.x; i < Cout; i += blockDim.x) sw[nw + i] = bias ? bias[i] : 0.0f;
  __syncthreads();
  const int groups = Cout / 8;
  const unsigned total = static_cast<unsigned>(N) * H * W * groups;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const int g = static_cast<int>(idx % groups);
    const unsigned pix = idx / groups;
    const int wq = static_cast<int>(pix % W);
    const unsigned row = pix / W;              // n*H + h
    const int hq = static_cast<int>(row % H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = sw[nw + g * 8 + j];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hh = hq + r - 1;
      if (hh < 0 || hh >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int ww = wq + s - 1;
        if (ww < 0 || ww >= W) continue;
        const float* xp = x + (static_cast<size_t>(row + r - 1) * W + ww) * Cin;
        const float* wp = sw + ((r * 3 + s) * Cin) * Cout + g * 8;
        for (int ci = 0; ci < Cin; ++ci) {
          const float xv = __ldg(xp + ci);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = fmaf(xv, wp[ci * Cout + j], acc[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = fmaxf(acc[j], 0.0f);
    st8(y + idx, acc);
  }
}

// ------------------------------------------------------------------------------------------ conv1a wgrad (+ bias grad)
// Each thread owns one 8-channel group of dz and walks pixels with a stride; partial sums go block -> ws -> final.
template <typename V>
__global__ void conv_first_wgrad_kernel(const float* __restrict__ x, const V* __restrict__ dz,
                                        float* __restrict__ ws, int N, int H, int W, int Cin, int Cout) {
  // every warp owns one row of partial sums in shared memory, the block sums its warps in a fixed order and writes
  // one row to ws (no atomics: results are bit-reproducible); rows are summed by rows_reduce_kernel
  extern __shared__ float red[];  // [warps][nout]
  const int nout = 9 * Cin * Cout + Cout;
  const int groups = Cout / 8;
  const int g = threadIdx.x % groups;
  float* wout = red + (threadIdx.x >> 5) * nout;
  const unsigned total = static_cast<unsigned>(N) * H * W * groups;
  const unsigned stride = gridDim.x * blockDim.x;
  for (int ci = 0; ci < Cin; ++ci) {
    float acc[9][8];
    float accb[8];
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[t][j] = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) accb[j] = 0.0f;
    for (unsigned idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
      const unsigned pix = idx / groups;
      const int wq = static_cast<int>(pix % W);
      const unsigned row = pix / W;            // n*H + h
      const int hq = static_cast<int>(row % H);
      float d[8];
      ld8(dz + idx, d);
      if (ci == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) accb[j] += d[j];
      }
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int hh = hq + r - 1;
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          const int ww = wq + s - 1;
          float xv = 0.0f;
          if (hh >= 0 && hh < H && ww >= 0 && ww < W)
            xv = __ldg(x + (static_cast<size_t>(row + r - 1) * W + ww) * Cin + ci);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[r * 3 + s][j] = fmaf(xv, d[j], acc[r * 3 + s][j]);
        }
      }
    }
    // lanes with the same channel group: xor-shuffle over offsets >= groups
#pragma unroll
    for (int t = 0; t < 9; ++t) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v = acc[t][j];
        for (int o = 16; o >= groups; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) < groups) wout[(t * Cin + ci) * Cout + g * 8 + j] = v;
      }
    }
    if (ci == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v = accb[j];
        for (int o = 16; o >= groups; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) < groups) wout[9 * Cin * Cout + g * 8 + j] = v;
      }
    }
  }
  __syncthreads();
  const int warps = blockDim.x >> 5;
  for (int i = threadIdx.x; i < nout; i += blockDim.x) {
    float s = red[i];
    for (int wv = 1; wv < warps; ++wv) s += red[wv * nout + i];
    ws[static_cast<size_t>(blockIdx.x) * nout + i] = s;
  }
}

// out[i] = sum_r ws[r*n + i] over nrows rows in a fixed order (8 interleaved partial sums per column, then a fixed
// 8-way sum); first n_a outputs go to out_a, the rest to out_b. Launch with 256 threads = 32 columns x 8 lanes.
__global__ void rows_reduce_kernel(const float* __restrict__ ws, int nrows, int n, float* __restrict__ out_a, int n_a,
                                   float* __restrict__ out_b) {
  pdl_trigger();
  pdl_wait();
  __shared__ float part[32][32];
  const int lanes = blockDim.x >> 5;   // row lanes per column (any block of 32*k threads, k <= 32)
  const int c = threadIdx.x & 31, p = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + c;
  float acc = 0.0f;
  if (i < n) {
    // each lane walks its rows in order; four loads are in flight, summed in the same order every time
    int r = p;
    for (; r + 3 * lanes < nrows; r += 4 * lanes) {
      const float v0 = ws[static_cast<long long>(r) * n + i];
      const float v1 = ws[static_cast<long long>(r + lanes) * n + i];
      const float v2 = ws[static_cast<long long>(r + 2 * lanes) * n + i];
      const float v3 = ws[static_cast<long long>(r + 3 * lanes) * n + i];
      acc += v0; acc += v1; acc += v2; acc += v3;
    }
    for (; r < nrows; r += lanes) acc += ws[static_cast<long long>(r) * n + i];
  }
  part[p][c] = acc;
  __syncthreads();
  if (p == 0 && i < n) {
    float s = part[0][c];
    for (int q = 1; q < lanes; ++q) s += part[q][c];
    if (i < n_a) out_a[i] = s;
    else if (out_b) out_b[i - n_a] = s;
  }
}

// ------------------------------------------------------------------------------------------ pooling
template <typename V>
__global__ void maxpool_fwd_kernel(const V* __restrict__ x, V* __restrict__ y, int N, int H, int W, int C) {
  const int groups = C / 8;
  const int Ho = H / 2, Wo = W / 2;
  const long long total = static_cast<long long>(N) * Ho * Wo * groups;
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int g = static_cast<int>(idx % groups);
  long long t = idx / groups;
  const int wo = static_cast<int>(t % Wo);
  t /= Wo;
  const int ho = static_cast<int>(t % Ho);
  const long long n = t / Ho;
  const long long base = ((n * H + 2 * ho) * W + 2 * wo) * groups + g;
  float fa[8], fb[8], fc[8], fd[8], o[8];
  ld8(x + base, fa); ld8(x + base + groups, fb);
  ld8(x + base + static_cast<long long>(W) * groups, fc); ld8(x + base + static_cast<long long>(W) * groups + groups, fd);
#pragma unroll
  for (int j = 0; j < 8; ++j) o[j] = fmaxf(fmaxf(fa[j], fb[j]), fmaxf(fc[j], fd[j]));
  st8(y + idx, o);
}

template <typename V>
__global__ void maxpool_bwd_kernel(const V* __restrict__ x, const V* __restrict__ dp,
                                   const V* skip, V* dz, int N, int H, int W, int C) {
  pdl_trigger();
  pdl_wait();
  const int groups = C / 8;
  const int Ho = H / 2, Wo = W / 2;
  const long long total = static_cast<long long>(N) * Ho * Wo * groups;
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int g = static_cast<int>(idx % groups);
  long long t = idx / groups;
  const int wo = static_cast<int>(t % Wo);
  t /= Wo;
  const int ho = static_cast<int>(t % Ho);
  const long long n = t / Ho;
  const long long base = ((n * H + 2 * ho) * W + 2 * wo) * groups + g;
  const long long off[4] = {base, base + groups, base + static_cast<long long>(W) * groups,
                            base + static_cast<long long>(W) * groups + groups};
  float xv[4][8], gv[4][8], d[8];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    ld8(x + off[k], xv[k]);
    if (skip) {
      ld8_rw(skip + off[k], gv[k]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) gv[k][j] = 0.0f;
    }
  }
  ld8(dp + idx, d);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int arg = 0;
    float m = xv[0][j];
#pragma unroll
    for (int k = 1; k < 4; ++k) {
      if (xv[k][j] > m) { m = xv[k][j]; arg = k; }
    }
    // ReLU of the producer: the pooled tensor is a ReLU output, so a zero maximum carries no gradient.
    if (m > 0.0f) {
#pragma unroll
      for (int k = 0; k < 4; ++k) gv[k][j] += (k == arg) ? d[j] : 0.0f;
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) st8(dz + off[k], gv[k]);
}

// ------------------------------------------------------------------------------------------ nearest 2x upsampling
template <typename V>
__global__ void upsample_fwd_kernel(const V* __restrict__ x, V* __restrict__ y, int N, int H, int W, int C) {
  const int groups = C / 8;
  const long long total = static_cast<long long>(N) * (2 * H) * (2 * W) * groups;
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int g = static_cast<int>(idx % groups);
  long long t = idx / groups;
  const int wo = static_cast<int>(t % (2 * W));
  t /= (2 * W);
  const int ho = static_cast<int>(t % (2 * H));
  const long long n = t / (2 * H);
  y[idx] = ldv(x + ((n * H + ho / 2) * W + wo / 2) * groups + g);
}

template <typename V>
__global__ void upsample_bwd_kernel(const V* __restrict__ dy, const V* __restrict__ x, V* __restrict__ dx,
                                    int N, int H, int W, int C, int relu_mask) {
  const int groups = C / 8;
  const long long total = static_cast<long long>(N) * H * W * groups;
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int g = static_cast<int>(idx % groups);
  long long t = idx / groups;
  const int w = static_cast<int>(t % W);
  t /= W;
  const int h = static_cast<int>(t % H);
  const long long n = t / H;
  const long long base = ((n * 2 * H + 2 * h) * (2 * W) + 2 * w) * groups + g;
  float a[8], b[8], c[8], d[8], o[8];
  ld8(dy + base, a);
  ld8(dy + base + groups, b);
  ld8(dy + base + static_cast<long long>(2 * W) * groups, c);
  ld8(dy + base + static_cast<long long>(2 * W) * groups + groups, d);
#pragma unroll
  for (int j = 0; j < 8; ++j) o[j] = (a[j] + b[j]) + (c[j] + d[j]);
  if (relu_mask) {
    float xv[8];
    ld8(x + idx, xv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = xv[j] > 0.0f ? o[j] : 0.0f;
  }
  st8(dx + idx, o);
}

// ------------------------------------------------------------------------------------------ head: 1x1 conv + sigmoid + dice sums
constexpr int kMaxCl = 4;

template <typename V>
__global__ void head_fwd_kernel(const V* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                                const float* __restrict__ y_true, float* __restrict__ p_out, float* __restrict__ ws,
                                long long P, int C, int n_cl) {
  extern __shared__ float sh[];  // [C*n_cl] weights, [n_cl] bias, then 3 floats per warp
  const int subs = C / 8;
  for (int i = threadIdx.x; i < C * n_cl; i += blockDim.x) sh[i] = w[i];
  for (int i = threadIdx.x; i < n_cl; i += blockDim.x) sh[C * n_cl + i] = b ? b[i] : 0.0f;
  __syncthreads();
  const int sub = threadIdx.x % subs;
  const long long total = P * subs;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  float sI = 0.0f, sY = 0.0f, sP = 0.0f;
  // every lane of a pixel group must stay in the loop for the shuffles: iterate on a uniform bound
  const long long iters = (total + stride - 1) / stride;
  long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  V nxt = zerov<V>();
  if (idx < total) nxt = ldv(x + idx);
  for (long long itn = 0; itn < iters; ++itn, idx += stride) {
    const bool live = idx < total;
    const V cur = nxt;
    if (idx + stride < total) nxt = ldv(x + idx + stride);   // next iteration's load is already in flight
    float acc[kMaxCl] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (live) {
      float f[8];
      unpackv(cur, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        for (int k = 0; k < n_cl; ++k) acc[k] = fmaf(f[j], sh[(sub * 8 + j) * n_cl + k], acc[k]);
      }
    }
    for (int k = 0; k < n_cl; ++k) {
      for (int o = subs >> 1; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
    }
    if (live && sub == 0) {
      const long long pix = idx / subs;
      for (int k = 0; k < n_cl; ++k) {
        const float z = acc[k] + sh[C * n_cl + k];
        const float p = 1.0f / (1.0f + expf(-z));
        const float yt = y_true ? __ldg(y_true + pix * n_cl + k) : 0.0f;
        p_out[pix * n_cl + k] = p;
        sI = fmaf(yt, p, sI);
        sY += yt;
        sP += p;
      }
    }
  }
  sI = warp_sum(sI); sY = warp_sum(sY); sP = warp_sum(sP);
  float* wred = sh + C * n_cl + n_cl;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { wred[warp * 3 + 0] = sI; wred[warp * 3 + 1] = sY; wred[warp * 3 + 2] = sP; }
  __syncthreads();
  if (threadIdx.x < 3) {
    float a = 0.0f;
    for (int wv = 0; wv < static_cast<int>(blockDim.x >> 5); ++wv) a += wred[wv * 3 + threadIdx.x];
    ws[blockIdx.x * 3 + threadIdx.x] = a;
  }
}

__global__ void dice_finalize_kernel(const float* __restrict__ sums, float smooth, float* __restrict__ out4) {
  pdl_trigger();
  pdl_wait();
  const float I = sums[0], Sy = sums[1], Sp = sums[2];
  out4[0] = -logf(2.0f * I + smooth) + logf(Sy + Sp + smooth);
  out4[1] = (2.0f * I + smooth) / (Sy + Sp + smooth);
  out4[2] = (I + smooth) / (Sy + smooth);
  out4[3] = (I + smooth) / (Sp + smooth);
}

template <typename V>
__global__ void head_bwd_kernel(const V* __restrict__ x, const float* __restrict__ w,
                                const float* __restrict__ y_true, const float* __restrict__ p,
                                const float* __restrict__ sums, float smooth, float grad_scale, V* __restrict__ dx,
                                float* __restrict__ ws, long long P, int C, int n_cl) {
  extern __shared__ float sh[];  // [C*n_cl] weights
  const int subs = C / 8;
  for (int i = threadIdx.x; i < C * n_cl; i += blockDim.x) sh[i] = w[i];
  __syncthreads();
  // sums == nullptr: y_true holds an explicit dL/dp (generic autograd path) instead of the labels
  float ca = 1.0f, cb = 0.0f;
  if (sums) {
    const float I = sums[0], Sy = sums[1], Sp = sums[2];
    ca = -2.0f / (2.0f * I + smooth);
    cb = 1.0f / (Sy + Sp + smooth);
  }
  const int sub = threadIdx.x % subs;
  const long long total = P * subs;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  float adw[8][kMaxCl];
  float adb[kMaxCl];
#pragma unroll
  for (int j = 0; j < 8; ++j)
#pragma unroll
    for (int k = 0; k < kMaxCl; ++k) adw[j][k] = 0.0f;
#pragma unroll
  for (int k = 0; k < kMaxCl; ++k) adb[k] = 0.0f;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const long long pix = idx / subs;
    float f[8], o[8];
    ld8(x + idx, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = 0.0f;
#pragma unroll
    for (int k = 0; k < kMaxCl; ++k) {
      if (k < n_cl) {
        const float pv = __ldg(p + pix * n_cl + k);
        const float yt = __ldg(y_true + pix * n_cl + k);
        const float dl = (ca * yt + cb) * pv * (1.0f - pv) * grad_scale;
        if (sub == 0) adb[k] += dl;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          adw[j][k] = fmaf(f[j], dl, adw[j][k]);
          o[j] = fmaf(dl, sh[(sub * 8 + j) * n_cl + k], o[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = f[j] > 0.0f ? o[j] : 0.0f;
    st8(dx + idx, o);
  }
  // reduce over lanes sharing `sub`; every warp owns one row in shared memory, the block sums its warps in a fixed
  // order and writes one row of partials to ws (no atomics)
  const int lane = threadIdx.x & 31;
  const int nout = C * n_cl + n_cl;
  float* red = sh + C * n_cl;   // [warps][nout]
  float* wout = red + (threadIdx.x >> 5) * nout;
#pragma unroll
  for (int k = 0; k < kMaxCl; ++k) {
    if (k < n_cl) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v = adw[j][k];
        for (int o2 = 16; o2 >= subs; o2 >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o2);
        if (lane < subs) wout[(sub * 8 + j) * n_cl + k] = v;
      }
      float vb = adb[k];
      vb = warp_sum(vb);
      if (lane == 0) wout[C * n_cl + k] = vb;
    }
  }
  __syncthreads();
  const int warps = blockDim.x >> 5;
  for (int i = threadIdx.x; i < nout; i += blockDim.x) {
    float s = red[i];
    for (int wv = 1; wv < warps; ++wv) s += red[wv * nout + i];
    ws[static_cast<size_t>(blockIdx.x) * nout + i] = s;
  }
}

// ------------------------------------------------------------------------------------------ bias gradient (column sums)
template <typename V>
__global__ void colsum_kernel(const V* __restrict__ dz, float* __restrict__ ws, long long P, int C) {
  // one row of partial sums per block in ws; inside the block: lanes -> warps (or runs of C/8 threads) -> fixed-order sum
  extern __shared__ float red[];  // [blockDim/span][C]
  const int groups = C / 8;
  const int g = threadIdx.x % groups;
  const int span = groups > 32 ? groups : 32;
  const int row = threadIdx.x / span;
  const int rows = blockDim.x / span;
  const long long total = P * groups;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; idx + 3 * stride < total; idx += 4 * stride) {   // four independent 16-byte loads in flight
    const V v0 = ldv(dz + idx), v1 = ldv(dz + idx + stride), v2 = ldv(dz + idx + 2 * stride),
            v3 = ldv(dz + idx + 3 * stride);
    float f0[8], f1[8], f2[8], f3[8];
    unpackv(v0, f0); unpackv(v1, f1); unpackv(v2, f2); unpackv(v3, f3);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += (f0[j] + f1[j]) + (f2[j] + f3[j]);
  }
  for (; idx < total; idx += stride) {
    float f[8];
    ld8(dz + idx, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += f[j];
  }
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float v = acc[j];
    for (int o = 16; o >= groups; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (groups >= 32 || lane < groups) red[row * C + g * 8 + j] = v;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = red[c];
    for (int r = 1; r < rows; ++r) s += red[r * C + c];
    ws[static_cast<long long>(blockIdx.x) * C + c] = s;
  }
}

// ------------------------------------------------------------------------------------------ weight packing
// GEMM operand element from an fp32 master: bf16, or (fp32/tf32 mode) the nearest tf32 value kept in an fp32 word
__device__ __forceinline__ void cvt_operand(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }
__device__ __forceinline__ void cvt_operand(float* p, float v) { *p = round_tf32(v); }

template <typename T>
__global__ void pack_conv_fwd_kernel(const float* __restrict__ w, T* __restrict__ out, int taps, int Cin,
                                     int Cout) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long n = static_cast<long long>(taps) * Cin * Cout;
  if (i >= n) return;
  const int ci = static_cast<int>(i % Cin);
  const int tap = static_cast<int>((i / Cin) % taps);
  const int co = static_cast<int>(i / (static_cast<long long>(Cin) * taps));
  cvt_operand(out + i, w[(static_cast<long long>(tap) * Cin + ci) * Cout + co]);
}
template <typename T>
__global__ void pack_conv_dgrad_kernel(const float* __restrict__ w, T* __restrict__ out, int taps, int Cin,
                                       int Cout) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long n = static_cast<long long>(taps) * Cin * Cout;
  if (i >= n) return;
  const int co = static_cast<int>(i % Cout);
  const int tap = static_cast<int>((i / Cout) % taps);
  const int ci = static_cast<int>(i / (static_cast<long long>(Cout) * taps));
  cvt_operand(out + i, w[(static_cast<long long>(taps - 1 - tap) * Cin + ci) * Cout + co]);
}
template <typename T>
__global__ void pack_tconv_dgrad_kernel(const float* __restrict__ w, T* __restrict__ out, int Cin, int Cout) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long n = 4ll * Cin * Cout;
  if (i >= n) return;
  const int o = static_cast<int>(i % Cout);
  const int ab = static_cast<int>((i / Cout) % 4);
  const int c = static_cast<int>(i / (4ll * Cout));
  cvt_operand(out + i, w[(static_cast<long long>(ab) * Cout + o) * Cin + c]);
}
template <typename T>
__global__ void __launch_bounds__(256)
pack_all_kernel(const PackLayer* __restrict__ layers, int nlayers, const float* __restrict__ params,
                uint8_t* __restrict__ arena, int tile_base) {
  __shared__ float tile[32][33];
  int li = 0;
  const int tile_id = static_cast<int>(blockIdx.x) + tile_base;
  while (li + 1 < nlayers && layers[li + 1].tile_start <= tile_id) ++li;   // block-uniform
  const PackLayer L = layers[li];
  const int t = tile_id - L.tile_start;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float* w = params + L.src;
  T* out_f = reinterpret_cast<T*>(arena + L.dst_fwd);
  T* out_d = reinterpret_cast<T*>(arena + L.dst_dgrad);
  if (L.kind == 0) {
    // master w[tap][ci][co]; fwd out[co][tap][ci] (transposed), dgrad out[ci][T-1-tap][co] (straight)
    const int tiles_co = L.cout >> 5, tiles_ci = L.cin >> 5;
    const int co0 = (t % tiles_co) << 5;
    const int ci0 = ((t / tiles_co) % tiles_ci) << 5;
    const int tap = t / (tiles_co * tiles_ci);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int ci = ci0 + ty + 8 * r;
      const float v = w[(static_cast<size_t>(tap) * L.cin + ci) * L.cout + co0 + tx];
      tile[ty + 8 * r][tx] = v;
      cvt_operand(out_d + (static_cast<size_t>(ci) * L.taps + (L.taps - 1 - tap)) * L.cout + co0 + tx, v);
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int co = co0 + ty + 8 * r;
      cvt_operand(out_f + (static_cast<size_t>(co) * L.taps + tap) * L.cin + ci0 + tx, tile[tx][ty + 8 * r]);
    }
  } else {
    // master w[ab][o][c]; fwd out[ab*Cout + o][c] (straight), dgrad out[c][ab][o] (transposed)
    const int tiles_c = L.cin >> 5, tiles_o = L.cout >> 5;
    const int c0 = (t % tiles_c) << 5;
    const int o0 = ((t / tiles_c) % tiles_o) << 5;
    const int ab = t / (tiles_c * tiles_o);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int o = o0 + ty + 8 * r;
      const size_t idx = (static_cast<size_t>(ab) * L.cout + o) * L.cin + c0 + tx;
      const float v = w[idx];
      tile[ty + 8 * r][tx] = v;
      cvt_operand(out_f + idx, v);
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int c = c0 + ty + 8 * r;
      cvt_operand(out_d + (static_cast<size_t>(c) * 4 + ab) * L.cout + o0 + tx, tile[tx][ty + 8 * r]);
    }
  }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) y[i] = __float2bfloat16(x[i]);
}
__global__ void round_tf32_kernel(const float* __restrict__ x, float* __restrict__ y, long long n) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) y[i] = round_tf32(x[i]);
}
__global__ void cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ y, long long n) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) y[i] = __bfloat162float(x[i]);
}

// ------------------------------------------------------------------------------------------ batch gather (data.py:29-54)
// dst[r][:] = float(src[perm[first + r]][:]): one shuffled batch (or one replica's shard of it) cut out of a training
// set that stays in HBM. A thread moves four elements: one 16-byte load (fp32 rows) or two (float64 rows, the dtype
// update_channels hands out, preprocess.py:301-302) and one 16-byte store; the row index is read once per thread.
template <typename T>
__global__ void gather_rows_kernel(const T* __restrict__ src, const long long* __restrict__ perm, long long first,
                                   long long row_elems, long long quads_per_row, long long total_quads,
                                   long long src_rows, float* __restrict__ dst) {
  for (long long q = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; q < total_quads;
       q += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = q / quads_per_row;
    const long long e = (q - r * quads_per_row) * 4;
    long long row = perm ? perm[first + r] : first + r;
    row = row < 0 ? 0 : (row >= src_rows ? src_rows - 1 : row);   // a bad index must not read outside the set
    const T* p = src + row * row_elems + e;
    float4 o;
    if constexpr (sizeof(T) == 8) {
      const double2 a = *reinterpret_cast<const double2*>(p);
      const double2 b = *reinterpret_cast<const double2*>(p + 2);
      o = make_float4(static_cast<float>(a.x), static_cast<float>(a.y), static_cast<float>(b.x),
                      static_cast<float>(b.y));
    } else {
      o = *reinterpret_cast<const float4*>(p);
    }
    *reinterpret_cast<float4*>(dst + r * row_elems + e) = o;
  }
}

// ------------------------------------------------------------------------------------------ optimizers
__device__ __forceinline__ float schedule_lr(long long step, float lr0, int const_lr, float lr_fraction,
                                             float decay_steps) {
  if (const_lr) return lr0;
  return static_cast<float>(static_cast<double>(lr0) *
                            pow(static_cast<double>(lr_fraction), static_cast<double>(step) / decay_steps));
}

__global__ void adam_kernel(float* __restrict__ theta, float* __restrict__ m, float* __restrict__ v,
                            const float* __restrict__ g, long long n, const long long* __restrict__ step_ptr, float lr0,
                            int const_lr, float lr_fraction, float decay_steps, float beta1, float beta2, float eps,
                            float grad_scale) {
  __shared__ float s_lrt;
  if (threadIdx.x == 0) {
    const long long step = *step_ptr;
    const double t = static_cast<double>(step + 1);
    const double lr = schedule_lr(step, lr0, const_lr, lr_fraction, decay_steps);
    s_lrt = static_cast<float>(lr * sqrt(1.0 - pow(static_cast<double>(beta2), t)) /
                               (1.0 - pow(static_cast<double>(beta1), t)));
  }
  __syncthreads();
  const float lrt = s_lrt;
  const float omb1 = 1.0f - beta1, omb2 = 1.0f - beta2;
  const long long n4 = n / 4;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 gt = reinterpret_cast<const float4*>(g)[i];
    float4 mt = reinterpret_cast<float4*>(m)[i];
    float4 vt = reinterpret_cast<float4*>(v)[i];
    float4 th = reinterpret_cast<float4*>(theta)[i];
    float* gp = &gt.x; float* mp = &mt.x; float* vp = &vt.x; float* tp = &th.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gk = gp[k] * grad_scale;
      mp[k] = beta1 * mp[k] + omb1 * gk;
      vp[k] = beta2 * vp[k] + omb2 * gk * gk;
      tp[k] = tp[k] - lrt * mp[k] / (sqrtf(vp[k]) + eps);
    }
    reinterpret_cast<float4*>(m)[i] = mt;
    reinterpret_cast<float4*>(v)[i] = vt;
    reinterpret_cast<float4*>(theta)[i] = th;
  }
  if (blockIdx.x == 0) {
    for (long long i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) {
      const float gk = g[i] * grad_scale;
      const float mk = beta1 * m[i] + omb1 * gk;
      const float vk = beta2 * v[i] + omb2 * gk * gk;
      m[i] = mk;
      v[i] = vk;
      theta[i] = theta[i] - lrt * mk / (sqrtf(vk) + eps);
    }
  }
}

__global__ void sgd_kernel(float* __restrict__ theta, const float* __restrict__ g, long long n,
                           const long long* __restrict__ step_ptr, float lr0, int const_lr, float lr_fraction,
                           float decay_steps, float grad_scale) {
  __shared__ float s_lr;
  if (threadIdx.x == 0) s_lr = schedule_lr(*step_ptr, lr0, const_lr, lr_fraction, decay_steps);
  __syncthreads();
  const float lr = s_lr * grad_scale;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    theta[i] = theta[i] - lr * g[i];
  }
}

__global__ void step_increment_kernel(long long* step_ptr) { *step_ptr += 1; }

__global__ void dropout_mask_kernel(unsigned char* __restrict__ keep, long long ngroups, float keep_prob,
                                    unsigned long long seed, unsigned int stream_id) {
  const long long g = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (g >= ngroups) return;
  const uint32_t bits = dropout_keep_bits8(seed, stream_id, static_cast<unsigned long long>(g), keep_prob);
#pragma unroll
  for (int i = 0; i < 8; ++i) keep[g * 8 + i] = (bits >> i) & 1u;
}

// Keep bits of a whole dropout layer: bits[g] (uint16) = keep decisions of elements 16g .. 16g+15, from the same Philox
// stream and element numbering as the conv epilogue's in-line generator and dropout_mask_kernel.
__global__ void dropout_bits16_kernel(unsigned short* __restrict__ bits, long long ngroups16, float keep_prob,
                                      unsigned long long seed, const long long* __restrict__ seed_add,
                                      unsigned int stream_id) {
  const long long g = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (g >= ngroups16) return;
  const unsigned long long sd = seed + (seed_add ? static_cast<unsigned long long>(__ldg(seed_add)) : 0ull);
  const uint32_t lo = dropout_keep_bits8(sd, stream_id, static_cast<unsigned long long>(2 * g), keep_prob);
  const uint32_t hi = dropout_keep_bits8(sd, stream_id, static_cast<unsigned long long>(2 * g + 1), keep_prob);
  bits[g] = static_cast<unsigned short>(lo | (hi << 8));
}

#include "level0.cuh"
#include "pool.cuh"

int check_groups(int C, const char* what) {
  const int groups = C / 8;
  if (C % 8 != 0 || groups < 1 || groups > 256 || (groups & (groups - 1)) != 0) {
    set_error("%s: channel count %d must be 8 * power of two (<= 2048)", what, C);
    return UB_ERR_INVALID;
  }
  return UB_OK;
}

}  // namespace

// ============================================================================================ host launchers
bool conv_first_fwd_writes_gates(int Cin, int Cout) { return Cout == 32 && (Cin == 1 || Cin == 4); }

int conv_first_fwd(const float* x, const float* w, const float* bias, void* y, int N, int H, int W, int Cin, int Cout,
                   cudaStream_t s, unsigned short* gate_out, int dt) {
  UB_REQUIRE(Cin >= 1 && Cin <= 8, "conv_first_fwd: Cin %d outside 1..8", Cin);
  UB_REQUIRE(gate_out == nullptr || conv_first_fwd_writes_gates(Cin, Cout),
             "conv_first_fwd: gate bits are only written for Cout 32 and Cin 1 or 4");
  UB_REQUIRE(Cout % 8 == 0 && Cout <= 64, "conv_first_fwd: Cout %d unsupported", Cout);
  const long long total = static_cast<long long>(N) * H * W * (Cout / 8);
  UB_REQUIRE(total < (1ll << 31), "conv_first_fwd: tensor too large");
  const bool y32 = reinterpret_cast<uintptr_t>(y) % 32 == 0;   // the fast path stores 32 bytes per instruction
  UB_REQUIRE(gate_out == nullptr || (y32 && W % 2 == 0),
             "conv_first_fwd: gate bits need a 32-byte aligned output tensor and an even width");
  UB_REQUIRE(dt == UB_DT_BF16 || gate_out == nullptr, "conv_first_fwd: gate bits are a bf16-mode feature");
  static const bool no_strip = getenv("UNETB200_NO_CONV1A_STRIP") != nullptr;
  if (dt == UB_DT_BF16 && !no_strip && Cin == 1 && Cout == 32 && W % 4 == 0 &&
      reinterpret_cast<uintptr_t>(y) % 16 == 0 && reinterpret_cast<uintptr_t>(w) % 16 == 0) {
    // register-weight strip kernel (level0.cuh): one resident wave of 128-thread blocks, grid-stride over strips
    const unsigned n_strips = static_cast<unsigned>(static_cast<long long>(N) * H * (W / 4));
    unsigned blocks = grid_for(static_cast<long long>(n_strips) * 4, 128);
    const unsigned cap = 12u * static_cast<unsigned>(device_sm_count());
    if (blocks > cap) blocks = cap;
    UB_CHECK_CUDA(launch_pdl(conv_first32_cin1_fwd_kernel, dim3(blocks), dim3(128), 0, s, x, w, bias, reinterpret_cast<uint4*>(y),
                                                        reinterpret_cast<unsigned int*>(gate_out), H, W, n_strips));
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  if (dt == UB_DT_BF16 && y32 && W % 2 == 0 && Cout == 32 && (Cin == 1 || Cin == 4)) {   // two-pixels-per-thread fast path (the network's conv1a)
    const unsigned P = static_cast<unsigned>(static_cast<long long>(N) * H * W);
    unsigned blocks = grid_for(P / 2, kThreads);
    const unsigned cap = 8u * static_cast<unsigned>(device_sm_count());
    if (blocks > cap) blocks = cap;
    if (Cin == 1) {
      conv_first32_fwd_kernel<1><<<blocks, kThreads, 0, s>>>(x, w, bias, reinterpret_cast<uint4*>(y),
                                                             reinterpret_cast<unsigned int*>(gate_out), H, W, P);
    } else {
      conv_first32_fwd_kernel<4><<<blocks, kThreads, 0, s>>>(x, w, bias, reinterpret_cast<uint4*>(y),
                                                             reinterpret_cast<unsigned int*>(gate_out), H, W, P);
    }
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  const size_t smem = (9 * Cin * Cout + Cout) * sizeof(float);
  unsigned blocks = grid_for(total, kThreads);
  const unsigned cap = 8u * static_cast<unsigned>(device_sm_count());   // weights are staged once per block
  if (blocks > cap) blocks = cap;
  if (dt == UB_DT_TF32) {
    conv_first_fwd_kernel<<<blocks, kThreads, smem, s>>>(x, w, bias, reinterpret_cast<F32x8*>(y), N, H, W, Cin, Cout);
  } else {
    conv_first_fwd_kernel<<<blocks, kThreads, smem, s>>>(x, w, bias, reinterpret_cast<uint4*>(y), N, H, W, Cin, Cout);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

long long conv_first_wgrad_ws_floats(int Cin, int Cout) {
  return static_cast<long long>(kRedBlocks) * (kThreads / 32) * (9 * Cin * Cout + Cout);
}

int conv_first_wgrad(const float* x, const void* dz, float* dw, float* db, float* ws, int N, int H, int W, int Cin,
                     int Cout, cudaStream_t s, int dt) {
  UB_REQUIRE(Cin >= 1 && Cin <= 8, "conv_first_wgrad: Cin %d outside 1..8", Cin);
  const int groups = Cout / 8;
  UB_REQUIRE(Cout % 8 == 0 && groups >= 1 && groups <= 8 && (groups & (groups - 1)) == 0,
             "conv_first_wgrad: Cout %d unsupported", Cout);
  const int nout = 9 * Cin * Cout + Cout;
  UB_REQUIRE(static_cast<long long>(N) * H * W * groups < (1ll << 31), "conv_first_wgrad: tensor too large");
  if (dt == UB_DT_BF16 && Cout == 32 && (Cin == 1 || Cin == 4) && W % 8 == 0) {   // strip kernel (level0.cuh)
    const unsigned n_strips = static_cast<unsigned>(N) * H * (W / 8);
    const uint4* dz4 = reinterpret_cast<const uint4*>(dz);
    if (Cin == 1) conv_first32_wgrad_kernel<1><<<kRedBlocks, 128, 0, s>>>(x, dz4, ws, H, W, n_strips);
    else conv_first32_wgrad_kernel<4><<<kRedBlocks, 128, 0, s>>>(x, dz4, ws, H, W, n_strips);
    UB_CHECK_CUDA(cudaGetLastError());
    rows_reduce_kernel<<<grid_for(nout, 32), 1024, 0, s>>>(ws, kRedBlocks, nout, dw, 9 * Cin * Cout, db);
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  const size_t smem = static_cast<size_t>(kThreads / 32) * nout * sizeof(float);
  if (smem > 48 * 1024) {
    UB_CHECK_CUDA(cudaFuncSetAttribute(conv_first_wgrad_kernel<uint4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       static_cast<int>(smem)));
    UB_CHECK_CUDA(cudaFuncSetAttribute(conv_first_wgrad_kernel<F32x8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       static_cast<int>(smem)));
  }
  if (dt == UB_DT_TF32) {
    conv_first_wgrad_kernel<<<kRedBlocks, kThreads, smem, s>>>(x, reinterpret_cast<const F32x8*>(dz), ws, N, H, W,
                                                              Cin, Cout);
  } else {
    conv_first_wgrad_kernel<<<kRedBlocks, kThreads, smem, s>>>(x, reinterpret_cast<const uint4*>(dz), ws, N, H, W,
                                                              Cin, Cout);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  rows_reduce_kernel<<<grid_for(nout, 32), 1024, 0, s>>>(ws, kRedBlocks, nout, dw, 9 * Cin * Cout, db);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

// The 16-channel / 256-bit kernels (pool.cuh) apply when every tensor is 32-byte aligned, C % 16 == 0 and the fine
// tensor has fewer than 2^32 32-byte units.
static bool v8_ok(int C, long long fine_pixels, std::initializer_list<const void*> ptrs, int which = 0) {
  static const int off_mask = getenv("UNETB200_V8_OFF") ? atoi(getenv("UNETB200_V8_OFF")) : 0;
  if (off_mask & which) return false;
  if (C % 16 != 0 || fine_pixels * (C / 16) >= (1ll << 32)) return false;
  for (const void* p : ptrs) {
    if (reinterpret_cast<uintptr_t>(p) % 32 != 0) return false;
  }
  return true;
}

int maxpool2x2_fwd(const void* x, void* y, int N, int H, int W, int C, cudaStream_t s, int dt) {
  UB_REQUIRE(C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "maxpool2x2_fwd: shape %dx%dx%d unsupported", H, W, C);
  if (dt == UB_DT_TF32) {
    const long long total = static_cast<long long>(N) * (H / 2) * (W / 2) * (C / 8);
    maxpool_fwd_kernel<<<grid_for(total, kThreads), kThreads, 0, s>>>(reinterpret_cast<const F32x8*>(x),
                                                                      reinterpret_cast<F32x8*>(y), N, H, W, C);
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  if (v8_ok(C, static_cast<long long>(N) * H * W, {x, y}, 1)) {
    const unsigned groups = C / 16, total16 = static_cast<unsigned>(N) * (H / 2) * (W / 2) * groups;
    UB_CHECK_CUDA(launch_pdl(maxpool_fwd_v8_kernel, dim3(grid_for(total16, 256)), dim3(256), 0, s, reinterpret_cast<const U32x8*>(x),
                                                                 reinterpret_cast<U32x8*>(y), total16, groups, W / 2, H / 2));
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  const long long total = static_cast<long long>(N) * (H / 2) * (W / 2) * (C / 8);
  maxpool_fwd_kernel<<<grid_for(total, kThreads), kThreads, 0, s>>>(reinterpret_cast<const uint4*>(x),
                                                                    reinterpret_cast<uint4*>(y), N, H, W, C);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int maxpool2x2_bwd(const void* x, const void* dpool, const void* skip_grad, void* dz_full, int N, int H, int W, int C,
                   cudaStream_t s, int dt) {
  UB_REQUIRE(C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "maxpool2x2_bwd: shape %dx%dx%d unsupported", H, W, C);
  const long long total = static_cast<long long>(N) * (H / 2) * (W / 2) * (C / 8);
  if (dt == UB_DT_TF32) {
    maxpool_bwd_kernel<<<grid_for(total, kThreads), kThreads, 0, s>>>(
        reinterpret_cast<const F32x8*>(x), reinterpret_cast<const F32x8*>(dpool),
        reinterpret_cast<const F32x8*>(skip_grad), reinterpret_cast<F32x8*>(dz_full), N, H, W, C);
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  UB_CHECK_CUDA(launch_pdl(maxpool_bwd_kernel<uint4>, dim3(grid_for(total, kThreads)), dim3(kThreads), 0, s,
                           reinterpret_cast<const uint4*>(x), reinterpret_cast<const uint4*>(dpool),
                           reinterpret_cast<const uint4*>(skip_grad), reinterpret_cast<uint4*>(dz_full), N, H, W, C));
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int upsample2x2_fwd(const void* x, void* y, int N, int H, int W, int C, cudaStream_t s, int dt) {
  UB_REQUIRE(C % 8 == 0, "upsample2x2_fwd: C %d", C);
  if (dt == UB_DT_TF32) {
    const long long total = static_cast<long long>(N) * (2 * H) * (2 * W) * (C / 8);
    upsample_fwd_kernel<<<grid_for(total, kThreads), kThreads, 0, s>>>(reinterpret_cast<const F32x8*>(x),
                                                                       reinterpret_cast<F32x8*>(y), N, H, W, C);
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  if (v8_ok(C, static_cast<long long>(N) * 4 * H * W, {x, y}, 4)) {
    const unsigned groups = C / 16, total16 = static_cast<unsigned>(N) * H * W * groups;
    UB_CHECK_CUDA(launch_pdl(upsample_fwd_v8_kernel, dim3(grid_for(total16, 256)), dim3(256), 0, s, reinterpret_cast<const U32x8*>(x),
                                                                  reinterpret_cast<U32x8*>(y), total16, groups, W, H));
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  const long long total = static_cast<long long>(N) * (2 * H) * (2 * W) * (C / 8);
  upsample_fwd_kernel<<<grid_for(total, kThreads), kThreads, 0, s>>>(reinterpret_cast<const uint4*>(x),
                                                                     reinterpret_cast<uint4*>(y), N, H, W, C);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int upsample2x2_bwd(const void* dy, const void* x, void* dx, int N, int H, int W, int C, int relu_mask, cudaStream_t s,
                    int dt) {
  UB_REQUIRE(C % 8 == 0, "upsample2x2_bwd: C %d", C);
  if (dt == UB_DT_TF32) {
    const long long total = static_cast<long long>(N) * H * W * (C / 8);
    upsample_bwd_kernel<<<grid_for(total, kThreads), kThreads, 0, s>>>(
        reinterpret_cast<const F32x8*>(dy), reinterpret_cast<const F32x8*>(x), reinterpret_cast<F32x8*>(dx), N, H, W, C,
        relu_mask);
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  if (v8_ok(C, static_cast<long long>(N) * 4 * H * W, {dy, x, dx}, 8)) {
    const unsigned groups = C / 16, total16 = static_cast<unsigned>(N) * H * W * groups;
    UB_CHECK_CUDA(launch_pdl(upsample_bwd_v8_kernel, dim3(grid_for(total16, 256)), dim3(256), 0, s, 
        reinterpret_cast<const U32x8*>(dy), reinterpret_cast<const U32x8*>(x), reinterpret_cast<U32x8*>(dx), total16,
        groups, W, H, relu_mask));
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  const long long total = static_cast<long long>(N) * H * W * (C / 8);
  upsample_bwd_kernel<<<grid_for(total, kThreads), kThreads, 0, s>>>(
      reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(x), reinterpret_cast<uint4*>(dx), N, H, W, C,
      relu_mask);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

long long head_ws_floats(long long, int C, int n_cl) {
  const long long fwd = 3ll * kRedBlocks;
  const long long bwd = static_cast<long long>(kRedBlocks) * (kThreads / 32) * (C * n_cl + n_cl);
  return fwd > bwd ? fwd : bwd;
}

int head_fwd(const void* x, const float* w, const float* b, const float* y_true, float* p_out, float* sums, float* ws,
             long long P, int C, int n_cl, cudaStream_t s, int dt) {
  UB_REQUIRE(n_cl >= 1 && n_cl <= kMaxCl, "head_fwd: n_cl_out %d outside 1..%d", n_cl, kMaxCl);
  const int subs = C / 8;
  UB_REQUIRE(C % 8 == 0 && subs >= 1 && subs <= 32 && (subs & (subs - 1)) == 0, "head_fwd: C %d unsupported", C);
  if (dt == UB_DT_TF32) {
    const size_t smem = (C * n_cl + n_cl + 3 * (kThreads / 32)) * sizeof(float);
    head_fwd_kernel<<<kRedBlocks, kThreads, smem, s>>>(reinterpret_cast<const F32x8*>(x), w, b, y_true, p_out, ws, P, C,
                                                       n_cl);
  } else if (C == 32 && P < (1ll << 31) && reinterpret_cast<uintptr_t>(x) % 32 == 0) {   // thread-per-pixel fast path (the network's Mask layer)
    const uint4* xv = reinterpret_cast<const uint4*>(x);
    const unsigned Pu = static_cast<unsigned>(P);
    switch (n_cl) {
      case 1: UB_CHECK_CUDA(launch_pdl(head32_fwd_kernel<1>, dim3(kRedBlocks), dim3(kThreads), 0, s, xv, w, b, y_true, p_out, ws, Pu)); break;
      case 2: UB_CHECK_CUDA(launch_pdl(head32_fwd_kernel<2>, dim3(kRedBlocks), dim3(kThreads), 0, s, xv, w, b, y_true, p_out, ws, Pu)); break;
      case 3: UB_CHECK_CUDA(launch_pdl(head32_fwd_kernel<3>, dim3(kRedBlocks), dim3(kThreads), 0, s, xv, w, b, y_true, p_out, ws, Pu)); break;
      default: UB_CHECK_CUDA(launch_pdl(head32_fwd_kernel<4>, dim3(kRedBlocks), dim3(kThreads), 0, s, xv, w, b, y_true, p_out, ws, Pu)); break;
    }
  } else {
    const size_t smem = (C * n_cl + n_cl + 3 * (kThreads / 32)) * sizeof(float);
    head_fwd_kernel<<<kRedBlocks, kThreads, smem, s>>>(reinterpret_cast<const uint4*>(x), w, b, y_true, p_out, ws, P, C,
                                                       n_cl);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  UB_CHECK_CUDA(launch_pdl(rows_reduce_kernel, dim3(1), dim3(1024), 0, s, ws, kRedBlocks, 3, sums, 3, nullptr));
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int dice_finalize(const float* sums, float smooth, float* out4, cudaStream_t s) {
  UB_CHECK_CUDA(launch_pdl(dice_finalize_kernel, dim3(1), dim3(1), 0, s, sums, smooth, out4));
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int head_bwd(const void* x, const float* w, const float* y_true, const float* p, const float* sums, float smooth,
             float grad_scale, void* dx, float* dw, float* db, float* ws, long long P, int C, int n_cl, cudaStream_t s,
             int dt) {
  UB_REQUIRE(n_cl >= 1 && n_cl <= kMaxCl, "head_bwd: n_cl_out %d outside 1..%d", n_cl, kMaxCl);
  const int subs = C / 8;
  UB_REQUIRE(C % 8 == 0 && subs >= 1 && subs <= 32 && (subs & (subs - 1)) == 0, "head_bwd: C %d unsupported", C);
  const int nout = C * n_cl + n_cl;
  int rows = kRedBlocks;   // partial rows in ws = blocks launched
  if (dt == UB_DT_TF32) {
    const size_t smem = (C * n_cl + (kThreads / 32) * nout) * sizeof(float);
    head_bwd_kernel<<<kRedBlocks, kThreads, smem, s>>>(reinterpret_cast<const F32x8*>(x), w, y_true, p, sums, smooth,
                                                       grad_scale, reinterpret_cast<F32x8*>(dx), ws, P, C, n_cl);
  } else if (C == 32 && n_cl == 1 && P < (1ll << 31) && reinterpret_cast<uintptr_t>(x) % 32 == 0 &&
      reinterpret_cast<uintptr_t>(dx) % 32 == 0) {   // thread-per-pixel fast path
    // one wave: 78 registers x 256 threads -> three resident blocks per SM (592 blocks would run as 444 + 148)
    static std::atomic<int> blocks_per_sm_of[64];   // per device (zero-initialised), like device_sm_count()
    int dev = 0;
    UB_CHECK_CUDA(cudaGetDevice(&dev));
    UB_REQUIRE(dev >= 0 && dev < 64, "head_bwd: device index %d unsupported", dev);
    int blocks_per_sm = blocks_per_sm_of[dev].load(std::memory_order_relaxed);
    if (blocks_per_sm == 0) {
      UB_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, head32_bwd_kernel, kThreads, 0));
      if (blocks_per_sm < 1) blocks_per_sm = 1;
      blocks_per_sm_of[dev].store(blocks_per_sm, std::memory_order_relaxed);
    }
    rows = blocks_per_sm * device_sm_count();
    if (rows > kRedBlocks) rows = kRedBlocks;
    UB_CHECK_CUDA(launch_pdl(head32_bwd_kernel, dim3(rows), dim3(kThreads), 0, s, reinterpret_cast<const uint4*>(x), w, y_true, p, sums, smooth,
                                                grad_scale, reinterpret_cast<uint4*>(dx), ws, static_cast<unsigned>(P)));
  } else {
    const size_t smem = (C * n_cl + (kThreads / 32) * nout) * sizeof(float);
    head_bwd_kernel<<<kRedBlocks, kThreads, smem, s>>>(reinterpret_cast<const uint4*>(x), w, y_true, p, sums, smooth,
                                                       grad_scale, reinterpret_cast<uint4*>(dx), ws, P, C, n_cl);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  UB_CHECK_CUDA(launch_pdl(rows_reduce_kernel, dim3(grid_for(nout, 32)), dim3(1024), 0, s, ws, rows, nout, dw, C * n_cl, db));
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int rows_sum(const float* ws, int nrows, int n, float* out, cudaStream_t s) {
  if (n <= 0) return UB_OK;
  rows_reduce_kernel<<<grid_for(n, 32), 1024, 0, s>>>(ws, nrows, n, out, n, nullptr);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

long long colsum_ws_floats(int C) { return static_cast<long long>(kRedBlocks) * (kThreads / 32) * C; }

int colsum_bf16(const void* dz, float* db, float* ws, long long P, int C, cudaStream_t s) {
  return colsum_act(dz, db, ws, P, C, s, UB_DT_BF16);
}

int colsum_act(const void* dz, float* db, float* ws, long long P, int C, cudaStream_t s, int dt) {
  int rc = check_groups(C, "colsum");
  if (rc) return rc;
  const size_t smem = (kThreads / 32) * C * sizeof(float);
  if (dt == UB_DT_TF32) {
    colsum_kernel<<<kRedBlocks, kThreads, smem, s>>>(reinterpret_cast<const F32x8*>(dz), ws, P, C);
  } else {
    colsum_kernel<<<kRedBlocks, kThreads, smem, s>>>(reinterpret_cast<const uint4*>(dz), ws, P, C);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  rows_reduce_kernel<<<grid_for(C, 32), 1024, 0, s>>>(ws, kRedBlocks, C, db, C, nullptr);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

// ------------------------------------------------------------------------------------------ bias gradient, no allocation
// Same two kernels as colsum_bf16, but the partial-sum rows live in a fixed device array that is part of the
// loaded library image instead of a caller-provided workspace.  Meant for callers that have nowhere to get a
// workspace from: plug-ins of host applications that own all device memory and forbid cudaMalloc on their worker
// threads, loops run under a fixed memory budget with a no-allocation rule, embedders that only see given tensors.
// Every GPU that loads the library carries these kStaticScratchFloats * 4 bytes (about 2.3 MiB) until the process
// exits.  All calls on a device share the array: two calls that overlap on the GPU (different streams) race on it.
// Nothing in it has to start at zero -- the first kernel overwrites every row it later reads -- so stale contents
// of an earlier call are harmless.
constexpr int kStaticMaxC = 512;   // widest layer of the UNet
constexpr long long kStaticScratchFloats = 606208;  // 2,424,832 bytes; kRedBlocks rows x kStaticMaxC fit twice
static __device__ __align__(256) float g_bias_grad_scratch[kStaticScratchFloats];

int colsum_bf16_static(const void* dz, float* db, long long P, int C, cudaStream_t s) {
  UB_REQUIRE(C >= 8 && C <= kStaticMaxC,
             "bias_grad_static: C = %d is beyond the built-in scratch (supports 8 <= C <= %d); use "
             "unetb200_bias_grad with a caller workspace for wider tensors", C, kStaticMaxC);
  int rc = check_groups(C, "bias_grad_static");
  if (rc) return rc;
  const int rows = kRedBlocks;
  UB_REQUIRE(static_cast<long long>(rows) * C <= kStaticScratchFloats, "bias_grad_static: internal scratch too small");
  // The array's name is only a host-side handle; the device address has to be asked for, once per device, with
  // that device current.  This first-call set-up is host work outside any stream, so make the first call on a
  // device before starting a stream capture.
  static float* scratch_of_device[64] = {nullptr};
  static std::mutex mu;
  int dev = 0;
  UB_CHECK_CUDA(cudaGetDevice(&dev));
  UB_REQUIRE(dev >= 0 && dev < 64, "bias_grad_static: device index %d unsupported", dev);
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!scratch_of_device[dev]) {
      void* p = nullptr;
      UB_CHECK_CUDA(cudaGetSymbolAddress(&p, g_bias_grad_scratch));
      scratch_of_device[dev] = static_cast<float*>(p);
    }
    scratch = scratch_of_device[dev];
  }
  colsum_kernel<<<kRedBlocks, kThreads, (kThreads / 32) * C * sizeof(float), s>>>(reinterpret_cast<const uint4*>(dz),
                                                                                 scratch, P, C);
  UB_CHECK_CUDA(cudaGetLastError());
  rows_reduce_kernel<<<grid_for(C, 32), 1024, 0, s>>>(scratch, rows, C, db, C, nullptr);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int pack_all(const PackLayer* layers_dev, int nlayers, int total_tiles, const float* params, void* arena,
             cudaStream_t s, int tile_base, int dt) {
  if (nlayers <= 0 || total_tiles <= 0) return UB_OK;
  if (dt == UB_DT_TF32) {
    pack_all_kernel<float><<<static_cast<unsigned>(total_tiles), 256, 0, s>>>(
        layers_dev, nlayers, params, reinterpret_cast<uint8_t*>(arena), tile_base);
  } else {
    pack_all_kernel<__nv_bfloat16><<<static_cast<unsigned>(total_tiles), 256, 0, s>>>(
        layers_dev, nlayers, params, reinterpret_cast<uint8_t*>(arena), tile_base);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int pack_conv_fwd(const float* w, void* out, int taps, int Cin, int Cout, cudaStream_t s, int dt) {
  const long long n = static_cast<long long>(taps) * Cin * Cout;
  if (dt == UB_DT_TF32) {
    pack_conv_fwd_kernel<<<grid_for(n, kThreads), kThreads, 0, s>>>(w, reinterpret_cast<float*>(out), taps, Cin, Cout);
  } else {
    pack_conv_fwd_kernel<<<grid_for(n, kThreads), kThreads, 0, s>>>(w, reinterpret_cast<__nv_bfloat16*>(out), taps, Cin,
                                                                    Cout);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}
int pack_conv_dgrad(const float* w, void* out, int taps, int Cin, int Cout, cudaStream_t s, int dt) {
  const long long n = static_cast<long long>(taps) * Cin * Cout;
  if (dt == UB_DT_TF32) {
    pack_conv_dgrad_kernel<<<grid_for(n, kThreads), kThreads, 0, s>>>(w, reinterpret_cast<float*>(out), taps, Cin,
                                                                      Cout);
  } else {
    pack_conv_dgrad_kernel<<<grid_for(n, kThreads), kThreads, 0, s>>>(w, reinterpret_cast<__nv_bfloat16*>(out), taps,
                                                                      Cin, Cout);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}
int pack_tconv_fwd(const float* w, void* out, int Cin, int Cout, cudaStream_t s, int dt) {
  if (dt == UB_DT_TF32) {
    const long long n = 4ll * Cin * Cout;
    round_tf32_kernel<<<grid_for(n, kThreads), kThreads, 0, s>>>(w, reinterpret_cast<float*>(out), n);
    UB_CHECK_CUDA(cudaGetLastError());
    return UB_OK;
  }
  return cast_f32_to_bf16(w, out, 4ll * Cin * Cout, s);
}
int pack_tconv_dgrad(const float* w, void* out, int Cin, int Cout, cudaStream_t s, int dt) {
  const long long n = 4ll * Cin * Cout;
  if (dt == UB_DT_TF32) {
    pack_tconv_dgrad_kernel<<<grid_for(n, kThreads), kThreads, 0, s>>>(w, reinterpret_cast<float*>(out), Cin, Cout);
  } else {
    pack_tconv_dgrad_kernel<<<grid_for(n, kThreads), kThreads, 0, s>>>(w, reinterpret_cast<__nv_bfloat16*>(out), Cin,
                                                                       Cout);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int adam_step(float* theta, float* m, float* v, const float* g, long long n, const long long* step_ptr, float lr0,
              int const_lr, float lr_fraction, float decay_steps, float beta1, float beta2, float eps, float grad_scale,
              cudaStream_t s) {
  UB_REQUIRE((reinterpret_cast<uintptr_t>(theta) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
              reinterpret_cast<uintptr_t>(g)) % 16 == 0,
             "adam_step: buffers must be 16-byte aligned");
  const int blocks = 4 * device_sm_count();
  adam_kernel<<<blocks, kThreads, 0, s>>>(theta, m, v, g, n, step_ptr, lr0, const_lr, lr_fraction, decay_steps, beta1,
                                          beta2, eps, grad_scale);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int sgd_step(float* theta, const float* g, long long n, const long long* step_ptr, float lr0, int const_lr,
             float lr_fraction, float decay_steps, float grad_scale, cudaStream_t s) {
  const int blocks = 4 * device_sm_count();
  sgd_kernel<<<blocks, kThreads, 0, s>>>(theta, g, n, step_ptr, lr0, const_lr, lr_fraction, decay_steps, grad_scale);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int step_increment(long long* step_ptr, cudaStream_t s) {
  step_increment_kernel<<<1, 1, 0, s>>>(step_ptr);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int dropout_mask(unsigned char* keep, long long n, float keep_prob, unsigned long long seed, unsigned int stream_id,
                 cudaStream_t s) {
  UB_REQUIRE(n % 8 == 0, "dropout_mask: n %lld must be a multiple of 8", n);
  const long long ng = n / 8;
  dropout_mask_kernel<<<grid_for(ng, kThreads), kThreads, 0, s>>>(keep, ng, keep_prob, seed, stream_id);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int dropout_keep_bits16(unsigned short* bits, long long n, float keep_prob, unsigned long long seed,
                        const long long* seed_add, unsigned int stream_id, cudaStream_t s) {
  UB_REQUIRE(n % 16 == 0, "dropout_keep_bits16: n %lld must be a multiple of 16", n);
  const long long ng = n / 16;
  dropout_bits16_kernel<<<grid_for(ng, kThreads), kThreads, 0, s>>>(bits, ng, keep_prob, seed, seed_add, stream_id);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int cast_f32_to_bf16(const float* x, void* y, long long n, cudaStream_t s) {
  cast_f32_bf16_kernel<<<grid_for(n, kThreads), kThreads, 0, s>>>(x, reinterpret_cast<__nv_bfloat16*>(y), n);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}
int cast_bf16_to_f32(const void* x, float* y, long long n, cudaStream_t s) {
  cast_bf16_f32_kernel<<<grid_for(n, kThreads), kThreads, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(x), y, n);
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

int gather_rows(const void* src, int src_is_f64, long long src_rows, const long long* perm, long long first, int rows,
                long long row_elems, float* dst, cudaStream_t s) {
  UB_REQUIRE(src && dst && rows >= 0 && row_elems > 0 && src_rows > 0, "gather_batch: null buffer or empty shape");
  UB_REQUIRE(row_elems % 4 == 0, "gather_batch: row of %lld elements is not a multiple of 4", row_elems);
  UB_REQUIRE(reinterpret_cast<uintptr_t>(src) % 16 == 0 && reinterpret_cast<uintptr_t>(dst) % 16 == 0,
             "gather_batch: buffers must be 16-byte aligned");
  UB_REQUIRE(perm || first + rows <= src_rows, "gather_batch: rows [%lld, %lld) outside the set of %lld", first,
             first + rows, src_rows);
  if (rows == 0) return UB_OK;
  const long long qpr = row_elems / 4, total = qpr * rows;
  long long blocks = (total + kThreads - 1) / kThreads;
  const long long cap = static_cast<long long>(device_sm_count()) * 16;   // grid-stride: a few waves at most
  if (blocks > cap) blocks = cap;
  if (src_is_f64) {
    gather_rows_kernel<double><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(
        reinterpret_cast<const double*>(src), perm, first, row_elems, qpr, total, src_rows, dst);
  } else {
    gather_rows_kernel<float><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(
        reinterpret_cast<const float*>(src), perm, first, row_elems, qpr, total, src_rows, dst);
  }
  UB_CHECK_CUDA(cudaGetLastError());
  return UB_OK;
}

}  // namespace ub
