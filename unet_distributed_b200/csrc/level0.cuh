// Thread-per-pixel kernels for the 32-channel full-resolution level (conv1a forward, the 1x1 Mask head forward and
// backward). The generic kernels in elementwise.cu split a pixel over four threads and carry run-time channel
// loops; at 128x128x128 images these ops each touch a 134 MB tensor once and were instruction-bound (ncu: 150 M and
// 90 M warp instructions for conv1a / head forward). Here every thread owns one pixel's 32 channels (64 contiguous
// bytes = four 16-byte accesses), all loops are compile-time, indices are 32-bit.
// Included inside elementwise.cu's anonymous namespace (uses its helpers).
#pragma once

// ---------------------------------------------------------------------------------------------- conv1a forward, Cout = 32
template <int CIN>
__global__ void __launch_bounds__(256)
conv_first32_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                        uint4* __restrict__ y, unsigned int* __restrict__ gate, int H, int W, unsigned P) {
  // A thread owns TWO horizontally adjacent pixels: every weight vector read from shared memory feeds two FMAs groups
  // (ncu on the one-pixel version: 72 LDS.128 per pixel, short-scoreboard + MIO-throttle stalls dominate) and the
  // 3 x 4 input window is shared. W is even (checked by the launcher), so a pair never straddles an image row.
  __shared__ float4 sw4[9 * CIN * 8 + 8];   // [tap][ci][32 channels] then bias, as float4
  float* sw = reinterpret_cast<float*>(sw4);
  for (int i = threadIdx.x; i < 9 * CIN * 32; i += blockDim.x) sw[i] = w[i];
  for (int i = threadIdx.x; i < 32; i += blockDim.x) sw[9 * CIN * 32 + i] = bias ? bias[i] : 0.0f;
  __syncthreads();
  const unsigned stride = gridDim.x * blockDim.x;
  const unsigned pairs = P >> 1;
  for (unsigned pp = blockIdx.x * blockDim.x + threadIdx.x; pp < pairs; pp += stride) {
    const unsigned pix = pp * 2;
    const int wq = static_cast<int>(pix % W);
    const unsigned row = pix / W;                 // n*H + h
    const int hq = static_cast<int>(row % H);
    float xv[3][4][CIN];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hh = hq + r - 1;
      const bool row_in = hh >= 0 && hh < H;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int ww = wq + c - 1;
        const bool in = row_in && ww >= 0 && ww < W;
        const float* xp = x + (static_cast<size_t>(row + r - 1) * W + ww) * CIN;
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) xv[r][c][ci] = in ? __ldg(xp + ci) : 0.0f;
      }
    }
    float4 acc[2][8];
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[0][q] = acc[1][q] = sw4[9 * CIN * 8 + q];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int s3 = 0; s3 < 3; ++s3) {
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) {
          const float v0 = xv[r][s3][ci], v1 = xv[r][s3 + 1][ci];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 wv = sw4[((r * 3 + s3) * CIN + ci) * 8 + q];
            acc[0][q].x = fmaf(v0, wv.x, acc[0][q].x); acc[0][q].y = fmaf(v0, wv.y, acc[0][q].y);
            acc[0][q].z = fmaf(v0, wv.z, acc[0][q].z); acc[0][q].w = fmaf(v0, wv.w, acc[0][q].w);
            acc[1][q].x = fmaf(v1, wv.x, acc[1][q].x); acc[1][q].y = fmaf(v1, wv.y, acc[1][q].y);
            acc[1][q].z = fmaf(v1, wv.z, acc[1][q].z); acc[1][q].w = fmaf(v1, wv.w, acc[1][q].w);
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (gate) {   // ReLU gates as bits: two uint16 (16 channels each) per pixel = one 32-bit word
        unsigned int bits = 0;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          bits |= (acc[k][q].x > 0.f ? 1u : 0u) << (4 * q);
          bits |= (acc[k][q].y > 0.f ? 1u : 0u) << (4 * q + 1);
          bits |= (acc[k][q].z > 0.f ? 1u : 0u) << (4 * q + 2);
          bits |= (acc[k][q].w > 0.f ? 1u : 0u) << (4 * q + 3);
        }
        gate[pix + k] = bits;
      }
      uint4* yp = y + static_cast<size_t>(pix + k) * 4;
#pragma unroll
      for (int q2 = 0; q2 < 2; ++q2) {   // two 32-byte stores per pixel
        U32x8 o;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const float4 a = acc[k][4 * q2 + 2 * q], b = acc[k][4 * q2 + 2 * q + 1];
          o.v[4 * q + 0] = pack_bf16x2(fmaxf(a.x, 0.f), fmaxf(a.y, 0.f));
          o.v[4 * q + 1] = pack_bf16x2(fmaxf(a.z, 0.f), fmaxf(a.w, 0.f));
          o.v[4 * q + 2] = pack_bf16x2(fmaxf(b.x, 0.f), fmaxf(b.y, 0.f));
          o.v[4 * q + 3] = pack_bf16x2(fmaxf(b.z, 0.f), fmaxf(b.w, 0.f));
        }
        st_global_v8(yp + 2 * q2, o);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------- conv1a forward, Cin = 1, Cout = 32
// The two-pixels-per-thread kernel above reads every weight vector from shared memory (72 LDS.128 per pixel pair) and
// ran at 1.4 TB/s, a fifth of the HBM rate this layer is bound by (ncu, r01: 66 us for a 134 MB write). Here the
// weights live in REGISTERS: a thread owns 8 of the 32 output channels (72 weights + 8 biases, loaded once) and walks
// strips of 4 consecutive pixels of an image row; the 3 x 6 input window of a strip is loaded once (the four threads
// of a strip share it through L1) and feeds 288 FMAs. The four threads of a pixel store its 64 bytes as four adjacent
// 16-byte pieces, and build the pixel's 32 ReLU gate bits with two shuffles.
__global__ void __launch_bounds__(128, 3)
conv_first32_cin1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                             uint4* __restrict__ y, unsigned int* __restrict__ gate, int H, int W, unsigned n_strips) {
  pdl_trigger();
  pdl_wait();
  const int g = threadIdx.x & 3;
  float wr[9][8], br[8];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(w + t * 32 + g * 8));
    const float4 b = __ldg(reinterpret_cast<const float4*>(w + t * 32 + g * 8 + 4));
    wr[t][0] = a.x; wr[t][1] = a.y; wr[t][2] = a.z; wr[t][3] = a.w;
    wr[t][4] = b.x; wr[t][5] = b.y; wr[t][6] = b.z; wr[t][7] = b.w;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) br[j] = bias ? __ldg(bias + g * 8 + j) : 0.0f;
  const unsigned strips_w = static_cast<unsigned>(W) >> 2;
  const unsigned step = (gridDim.x * blockDim.x) >> 2;
  // warp-uniform trip count (the gate-bit shuffles need every lane): lanes past the end compute on strip 0, store nothing
  const unsigned warp_first = ((blockIdx.x * blockDim.x + threadIdx.x) & ~31u) >> 2;
  for (unsigned base = warp_first; base < n_strips; base += step) {
    const unsigned mine = base + ((threadIdx.x & 31u) >> 2);
    const bool live = mine < n_strips;
    const unsigned sidx = live ? mine : 0u;
    const unsigned row = sidx / strips_w;                        // n*H + h
    const int w0 = static_cast<int>(sidx - row * strips_w) * 4;
    const int hq = static_cast<int>(row % static_cast<unsigned>(H));
    float xr[3][6];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hh = hq + r - 1;
      const bool row_in = hh >= 0 && hh < H;
      const float* xp = x + static_cast<size_t>(row + r - 1) * W + w0 - 1;
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        const int ww = w0 - 1 + c;
        xr[r][c] = (row_in && ww >= 0 && ww < W) ? __ldg(xp + c) : 0.0f;
      }
    }
    const size_t pix0 = static_cast<size_t>(row) * W + w0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = br[j];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int sx = 0; sx < 3; ++sx) {
          const float xv = xr[r][k + sx];
#pragma unroll
          for (int j = 0; j < 8; j += 2) fma2(acc[j], acc[j + 1], xv, wr[r * 3 + sx][j], wr[r * 3 + sx][j + 1]);
        }
      }
      uint4 o;   // ReLU inside the conversion (ncu r02: the kernel is issue-bound, every instruction counts)
      o.x = pack_bf16x2_relu_rn(acc[0], acc[1]);
      o.y = pack_bf16x2_relu_rn(acc[2], acc[3]);
      o.z = pack_bf16x2_relu_rn(acc[4], acc[5]);
      o.w = pack_bf16x2_relu_rn(acc[6], acc[7]);
      if (gate) {   // 32 gate bits per pixel: this thread's 8 at bit 8g, merged over the pixel's four threads.
        // A bf16 half is non-zero iff adding 0x7fff carries into its sign bit (outputs are >= 0 after the ReLU).
        const unsigned int ow[4] = {o.x, o.y, o.z, o.w};
        unsigned int acc2 = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) acc2 |= ((ow[i] + 0x7FFF7FFFu) >> (15 - 2 * i)) & (0x00010001u << (2 * i));
        unsigned int bits = ((acc2 & 0xFFu) | ((acc2 >> 15) & 0xFFu)) << (8 * g);
        bits |= __shfl_xor_sync(0xffffffffu, bits, 1);
        bits |= __shfl_xor_sync(0xffffffffu, bits, 2);
        if (g == 0 && live) gate[pix0 + k] = bits;
      }
      if (live) y[(pix0 + k) * 4 + g] = o;
    }
  }
}

// ---------------------------------------------------------------------------------------------- head forward, C = 32
template <int NCL>
__global__ void __launch_bounds__(256)
head32_fwd_kernel(const uint4* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                  const float* __restrict__ y_true, float* __restrict__ p_out, float* __restrict__ ws, unsigned P) {
  pdl_trigger();
  pdl_wait();
  __shared__ float sw[32 * NCL + NCL];
  __shared__ float wred[8 * 3];
  for (int i = threadIdx.x; i < 32 * NCL; i += blockDim.x) sw[i] = w[i];
  for (int i = threadIdx.x; i < NCL; i += blockDim.x) sw[32 * NCL + i] = b ? b[i] : 0.0f;
  __syncthreads();
  float sI = 0.0f, sY = 0.0f, sP = 0.0f;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned pix = blockIdx.x * blockDim.x + threadIdx.x; pix < P; pix += stride) {
    const uint4* xp = x + static_cast<size_t>(pix) * 4;
    const U32x8 v0 = ld_global_nc_v8(xp), v1 = ld_global_nc_v8(xp + 2);   // 64 bytes per pixel, two sector loads
    float f[32];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      f[2 * j] = bf16_lo(v0.v[j]); f[2 * j + 1] = bf16_hi(v0.v[j]);
      f[16 + 2 * j] = bf16_lo(v1.v[j]); f[16 + 2 * j + 1] = bf16_hi(v1.v[j]);
    }
#pragma unroll
    for (int k = 0; k < NCL; ++k) {
      float z = sw[32 * NCL + k];
#pragma unroll
      for (int c = 0; c < 32; ++c) z = fmaf(f[c], sw[c * NCL + k], z);
      const float p = 1.0f / (1.0f + expf(-z));
      const float yt = y_true ? __ldg(y_true + static_cast<size_t>(pix) * NCL + k) : 0.0f;
      p_out[static_cast<size_t>(pix) * NCL + k] = p;
      sI = fmaf(yt, p, sI);
      sY += yt;
      sP += p;
    }
  }
  sI = warp_sum(sI); sY = warp_sum(sY); sP = warp_sum(sP);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { wred[warp * 3 + 0] = sI; wred[warp * 3 + 1] = sY; wred[warp * 3 + 2] = sP; }
  __syncthreads();
  if (threadIdx.x < 3) {
    float a = 0.0f;
    for (int wv = 0; wv < static_cast<int>(blockDim.x >> 5); ++wv) a += wred[wv * 3 + threadIdx.x];
    ws[blockIdx.x * 3 + threadIdx.x] = a;
  }
}

// ---------------------------------------------------------------------------------------------- head backward, C = 32, one class
// dl = (ca*y + cb) * p(1-p) * scale (dice) or y * p(1-p) * scale (explicit dL/dp in y when sums == nullptr);
// dx = dl * w masked by x > 0; dW[c] += x[c]*dl, db += dl -> ws[block][33].
__global__ void __launch_bounds__(256)
head32_bwd_kernel(const uint4* __restrict__ x, const float* __restrict__ w, const float* __restrict__ y_true,
                  const float* __restrict__ p, const float* __restrict__ sums, float smooth, float grad_scale,
                  uint4* __restrict__ dx, float* __restrict__ ws, unsigned P) {
  pdl_trigger();
  pdl_wait();
  __shared__ float sw[32];
  __shared__ float red[8][33];
  if (threadIdx.x < 32) sw[threadIdx.x] = w[threadIdx.x];
  __syncthreads();
  float ca = 1.0f, cb = 0.0f;
  if (sums) {
    const float I = sums[0], Sy = sums[1], Sp = sums[2];
    ca = -2.0f / (2.0f * I + smooth);
    cb = 1.0f / (Sy + Sp + smooth);
  }
  float adw[32];
#pragma unroll
  for (int c = 0; c < 32; ++c) adw[c] = 0.0f;
  float adb = 0.0f;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned pix = blockIdx.x * blockDim.x + threadIdx.x; pix < P; pix += stride) {
    const uint4* xp = x + static_cast<size_t>(pix) * 4;
    const U32x8 v[2] = {ld_global_nc_v8(xp), ld_global_nc_v8(xp + 2)};
    const float pv = __ldg(p + pix);
    const float yt = __ldg(y_true + pix);
    const float dl = (ca * yt + cb) * pv * (1.0f - pv) * grad_scale;
    adb += dl;
    uint4* dp = dx + static_cast<size_t>(pix) * 4;
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      U32x8 o;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float f0 = bf16_lo(v[q].v[j]), f1 = bf16_hi(v[q].v[j]);
        const int c = 16 * q + 2 * j;
        adw[c] = fmaf(f0, dl, adw[c]);
        adw[c + 1] = fmaf(f1, dl, adw[c + 1]);
        o.v[j] = pack_bf16x2(f0 > 0.0f ? dl * sw[c] : 0.0f, f1 > 0.0f ? dl * sw[c + 1] : 0.0f);
      }
      st_global_v8(dp + 2 * q, o);
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int c = 0; c < 32; ++c) {
    const float s = warp_sum(adw[c]);
    if (lane == 0) red[warp][c] = s;
  }
  adb = warp_sum(adb);
  if (lane == 0) red[warp][32] = adb;
  __syncthreads();
  if (threadIdx.x < 33) {
    float s = red[0][threadIdx.x];
    for (int wv = 1; wv < static_cast<int>(blockDim.x >> 5); ++wv) s += red[wv][threadIdx.x];
    ws[static_cast<size_t>(blockIdx.x) * 33 + threadIdx.x] = s;
  }
}

// ---------------------------------------------------------------------------------------------- conv1a wgrad, Cout = 32
// dW[3][3][CIN][32] and db[32] from the fp32 network input and bf16 dz. The generic kernel spends ~170 instructions
// per (pixel, 8-channel group) on index arithmetic and nine boundary-checked input loads (ncu: 137 us for a 134 MB
// read). Here a thread owns one 8-channel group of a strip of 8 consecutive pixels of an image row: the 3 x 10 input
// window is loaded once per strip, the eight 16-byte dz loads are issued back to back, and the 72 + 8 accumulators
// stay in registers across strips. Lanes l and l^4, l^8, l^16 hold the same channel group; warps, then blocks, are
// summed in a fixed order (bit-reproducible): block -> ws[block][nout] -> rows_reduce_kernel.
template <int CIN>
__global__ void __launch_bounds__(128, 4)
conv_first32_wgrad_kernel(const float* __restrict__ x, const uint4* __restrict__ dz, float* __restrict__ ws, int H,
                          int W, unsigned n_strips) {
  constexpr int kNout = 9 * CIN * 32 + 32;
  __shared__ float red[4][kNout];
  const int g = threadIdx.x & 3;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned strips_w = static_cast<unsigned>(W) >> 3;
  const unsigned first = (blockIdx.x * blockDim.x + threadIdx.x) >> 2;
  const unsigned step = (gridDim.x * blockDim.x) >> 2;
  for (int ci = 0; ci < CIN; ++ci) {
    float acc[9][8];
    float accb[8];
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[t][j] = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) accb[j] = 0.0f;
    for (unsigned sidx = first; sidx < n_strips; sidx += step) {
      const unsigned row = sidx / strips_w;                       // n*H + h
      const int w0 = static_cast<int>(sidx - row * strips_w) * 8;
      const int hq = static_cast<int>(row % static_cast<unsigned>(H));
      uint4 q[8];
      const uint4* dp = dz + (static_cast<size_t>(row) * W + w0) * 4 + g;
#pragma unroll
      for (int k = 0; k < 8; ++k) q[k] = __ldg(dp + 4 * k);
      float xr[3][10];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int hh = hq + r - 1;
        const bool row_in = hh >= 0 && hh < H;
        const float* xp = x + (static_cast<size_t>(row + r - 1) * W + w0 - 1) * CIN + ci;
#pragma unroll
        for (int c = 0; c < 10; ++c) {
          const int ww = w0 - 1 + c;
          xr[r][c] = (row_in && ww >= 0 && ww < W) ? __ldg(xp + c * CIN) : 0.0f;
        }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float d[8];
        unpack8(q[k], d);
        if (ci == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) accb[j] += d[j];
        }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
#pragma unroll
          for (int sx = 0; sx < 3; ++sx) {
            const float xv = xr[r][k + sx];
#pragma unroll
            for (int j = 0; j < 8; j += 2) fma2(acc[r * 3 + sx][j], acc[r * 3 + sx][j + 1], xv, d[j], d[j + 1]);
          }
        }
      }
    }
#pragma unroll
    for (int t = 0; t < 9; ++t) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v = acc[t][j];
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        if (lane < 4) red[warp][(t * CIN + ci) * 32 + g * 8 + j] = v;
      }
    }
    if (ci == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v = accb[j];
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        if (lane < 4) red[warp][9 * CIN * 32 + g * 8 + j] = v;
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kNout; i += blockDim.x) {
    ws[static_cast<size_t>(blockIdx.x) * kNout + i] = (red[0][i] + red[1][i]) + (red[2][i] + red[3][i]);
  }
}
