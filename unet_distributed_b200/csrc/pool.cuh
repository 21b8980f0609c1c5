// 2x2 max pooling and nearest-neighbour 2x upsampling, forward and backward, for NHWC bf16 tensors whose channel
// count is a multiple of 16 and whose base is 32-byte aligned: one thread moves 16 channels (32 bytes = one DRAM
// sector) per access with 256-bit loads/stores and 32-bit index arithmetic. The generic 8-channel kernels in
// elementwise.cu remain the fallback for other shapes. Included inside elementwise.cu's anonymous namespace.
#pragma once

__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  __nv_bfloat162 x = *reinterpret_cast<__nv_bfloat162*>(&a), y = *reinterpret_cast<__nv_bfloat162*>(&b);
  __nv_bfloat162 m = __hmax2(x, y);
  return *reinterpret_cast<uint32_t*>(&m);
}

// (pixel index of the top-left fine pixel, in units of 16-channel groups) for coarse element idx
struct Quad {
  unsigned base, right, down;   // base, +1 pixel, +1 row (all in 32-byte units)
};
__device__ __forceinline__ Quad quad_of(unsigned idx, unsigned groups, unsigned Wc, unsigned Hc) {
  const unsigned g = idx % groups;
  unsigned t = idx / groups;
  const unsigned wc = t % Wc;
  t /= Wc;                       // n * Hc + hc
  Quad q;
  q.base = ((2u * t) * (2u * Wc) + 2u * wc) * groups + g;
  q.right = groups;
  q.down = 2u * Wc * groups;
  return q;
}

__global__ void __launch_bounds__(256)
maxpool_fwd_v8_kernel(const U32x8* __restrict__ x, U32x8* __restrict__ y, unsigned total, unsigned groups, unsigned Wc,
                      unsigned Hc) {
  pdl_trigger();
  pdl_wait();
  const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const Quad q = quad_of(idx, groups, Wc, Hc);
  const U32x8 a = ld_global_nc_v8(x + q.base), b = ld_global_nc_v8(x + q.base + q.right);
  const U32x8 c = ld_global_nc_v8(x + q.base + q.down), d = ld_global_nc_v8(x + q.base + q.down + q.right);
  U32x8 o;
#pragma unroll
  for (int i = 0; i < 8; ++i) o.v[i] = bf16x2_max(bf16x2_max(a.v[i], b.v[i]), bf16x2_max(c.v[i], d.v[i]));
  st_global_v8(y + idx, o);
}

// (MaxPoolGrad stays on the 8-channel kernel in elementwise.cu: a 16-channel variant needs 70+ live registers per
// thread and measured 2-3% slower per step.)

__global__ void __launch_bounds__(256)
upsample_fwd_v8_kernel(const U32x8* __restrict__ x, U32x8* __restrict__ y, unsigned total, unsigned groups, unsigned Wc,
                       unsigned Hc) {
  pdl_trigger();
  pdl_wait();
  const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;   // coarse element
  if (idx >= total) return;
  const Quad q = quad_of(idx, groups, Wc, Hc);
  const U32x8 v = ld_global_nc_v8(x + idx);
  st_global_v8(y + q.base, v);
  st_global_v8(y + q.base + q.right, v);
  st_global_v8(y + q.base + q.down, v);
  st_global_v8(y + q.base + q.down + q.right, v);
}

__global__ void __launch_bounds__(256)
upsample_bwd_v8_kernel(const U32x8* __restrict__ dy, const U32x8* __restrict__ x, U32x8* __restrict__ dx,
                       unsigned total, unsigned groups, unsigned Wc, unsigned Hc, int relu_mask) {
  pdl_trigger();
  pdl_wait();
  const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;   // coarse element
  if (idx >= total) return;
  const Quad q = quad_of(idx, groups, Wc, Hc);
  const U32x8 a = ld_global_nc_v8(dy + q.base), b = ld_global_nc_v8(dy + q.base + q.right);
  const U32x8 c = ld_global_nc_v8(dy + q.base + q.down), d = ld_global_nc_v8(dy + q.base + q.down + q.right);
  U32x8 xv;
  if (relu_mask) xv = ld_global_nc_v8(x + idx);
  U32x8 o;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float lo = (bf16_lo(a.v[i]) + bf16_lo(b.v[i])) + (bf16_lo(c.v[i]) + bf16_lo(d.v[i]));
    float hi = (bf16_hi(a.v[i]) + bf16_hi(b.v[i])) + (bf16_hi(c.v[i]) + bf16_hi(d.v[i]));
    if (relu_mask) {
      lo = bf16_lo(xv.v[i]) > 0.0f ? lo : 0.0f;
      hi = bf16_hi(xv.v[i]) > 0.0f ? hi : 0.0f;
    }
    o.v[i] = pack_bf16x2(lo, hi);
  }
  st_global_v8(dx + idx, o);
}
