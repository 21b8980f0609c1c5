// extern "C" boundary of libunetb200 (declarations and reference citations: include/unetb200.h).
#include "../../include/unetb200.h"

#include "common.h"
#include "elementwise.cuh"
#include "ops.cuh"

using namespace ub;

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

extern "C" {

int unetb200_version(void) { return 100; }
const char* unetb200_last_error_string(void) { return last_error(); }

int unetb200_pack_conv3x3_fwd(const float* w, void* out, int Cin, int Cout, void* stream) {
  return pack_conv_fwd(w, out, 9, Cin, Cout, S(stream));
}
int unetb200_pack_conv3x3_dgrad(const float* w, void* out, int Cin, int Cout, void* stream) {
  return pack_conv_dgrad(w, out, 9, Cin, Cout, S(stream));
}
int unetb200_pack_tconv2x2_fwd(const float* w, void* out, int Cin, int Cout, void* stream) {
  return pack_tconv_fwd(w, out, Cin, Cout, S(stream));
}
int unetb200_pack_tconv2x2_dgrad(const float* w, void* out, int Cin, int Cout, void* stream) {
  return pack_tconv_dgrad(w, out, Cin, Cout, S(stream));
}

int unetb200_conv3x3_fwd(const void* x0, int C0, const void* x1, int C1, const void* w_packed, const float* bias,
                         void* y, int N, int H, int W, int Cout, int relu, float drop_keep, uint64_t drop_seed,
                         uint32_t drop_stream, void* stream) {
  ConvSrc s0{x0, C0}, s1{x1, x1 ? C1 : 0};
  DropoutCfg d{drop_keep <= 0.0f ? 1.0f : drop_keep, drop_seed, drop_stream, nullptr};
  return conv3x3_fwd(s0, s1, w_packed, bias, y, N, H, W, Cout, relu, d, S(stream));
}

int unetb200_conv3x3_dgrad(const void* dz, int Cout, const void* w_packed_d, void* dx0, int C0, const void* mask0,
                           void* dx1, int C1, const void* mask1, float scale, int N, int H, int W, void* stream) {
  DgradDst d0{dx0, C0, mask0, scale}, d1{dx1, dx1 ? C1 : 0, mask1, scale};
  return conv3x3_dgrad(dz, Cout, w_packed_d, d0, d1, N, H, W, S(stream));
}

size_t unetb200_conv3x3_wgrad_ws_bytes(int N, int H, int W, int Cin_total, int Cout) {
  return conv3x3_wgrad_ws_bytes(N, H, W, Cin_total, Cout);
}
int unetb200_conv3x3_wgrad(const void* dz, int Cout, const void* x0, int C0, const void* x1, int C1, float* dw,
                           void* ws, size_t ws_bytes, int N, int H, int W, void* stream) {
  ConvSrc s0{x0, C0}, s1{x1, x1 ? C1 : 0};
  return conv3x3_wgrad(dz, Cout, s0, s1, dw, ws, ws_bytes, N, H, W, S(stream));
}

int unetb200_conv_first_fwd(const float* x, const float* w, const float* bias, void* y, int N, int H, int W, int Cin,
                            int Cout, void* stream) {
  return conv_first_fwd(x, w, bias, y, N, H, W, Cin, Cout, S(stream));
}
size_t unetb200_conv_first_wgrad_ws_bytes(int Cin, int Cout) {
  return static_cast<size_t>(conv_first_wgrad_ws_floats(Cin, Cout)) * sizeof(float);
}
int unetb200_conv_first_wgrad(const float* x, const void* dz, float* dw, float* db, void* ws, int N, int H, int W,
                              int Cin, int Cout, void* stream) {
  return conv_first_wgrad(x, dz, dw, db, reinterpret_cast<float*>(ws), N, H, W, Cin, Cout, S(stream));
}
size_t unetb200_bias_grad_ws_bytes(int C) { return static_cast<size_t>(colsum_ws_floats(C)) * sizeof(float); }
int unetb200_bias_grad(const void* dz, float* db, void* ws, long long pixels, int C, void* stream) {
  return colsum_bf16(dz, db, reinterpret_cast<float*>(ws), pixels, C, S(stream));
}
int unetb200_bias_grad_static(const void* dz, float* db, long long pixels, int C, void* stream) {
  return colsum_bf16_static(dz, db, pixels, C, S(stream));
}

int unetb200_maxpool2x2_fwd(const void* x, void* y, int N, int H, int W, int C, void* stream) {
  return maxpool2x2_fwd(x, y, N, H, W, C, S(stream));
}
int unetb200_maxpool2x2_bwd(const void* x, const void* dpool, const void* skip_grad, void* dz, int N, int H, int W,
                            int C, void* stream) {
  return maxpool2x2_bwd(x, dpool, skip_grad, dz, N, H, W, C, S(stream));
}

int unetb200_tconv2x2_fwd(const void* x, int Cin, const void* w_packed, const float* bias, void* y, int Cout, int N,
                          int H, int W, void* stream) {
  return tconv2x2_fwd(x, Cin, w_packed, bias, y, Cout, N, H, W, S(stream));
}
int unetb200_tconv2x2_dgrad(const void* dy, int Cout, const void* w_packed_d, void* dx, int Cin, const void* mask,
                            float scale, int N, int H, int W, void* stream) {
  return tconv2x2_dgrad(dy, Cout, w_packed_d, dx, Cin, mask, scale, N, H, W, S(stream));
}
size_t unetb200_tconv2x2_wgrad_ws_bytes(int N, int H, int W, int Cin, int Cout) {
  return tconv2x2_wgrad_ws_bytes(N, H, W, Cin, Cout);
}
int unetb200_tconv2x2_wgrad(const void* dy, int Cout, const void* x, int Cin, float* dw, void* ws, size_t ws_bytes,
                            int N, int H, int W, void* stream) {
  return tconv2x2_wgrad(dy, Cout, x, Cin, dw, ws, ws_bytes, N, H, W, S(stream));
}
int unetb200_upsample2x2_fwd(const void* x, void* y, int N, int H, int W, int C, void* stream) {
  return upsample2x2_fwd(x, y, N, H, W, C, S(stream));
}
int unetb200_upsample2x2_bwd(const void* dy, const void* x, void* dx, int N, int H, int W, int C, int relu_mask,
                             void* stream) {
  return upsample2x2_bwd(dy, x, dx, N, H, W, C, relu_mask, S(stream));
}

size_t unetb200_head_ws_bytes(long long pixels, int C, int n_cl) {
  return static_cast<size_t>(head_ws_floats(pixels, C, n_cl)) * sizeof(float);
}
int unetb200_head_fwd(const void* x, const float* w, const float* b, const float* y_true, float* p, float* sums,
                      void* ws, long long pixels, int C, int n_cl, void* stream) {
  return head_fwd(x, w, b, y_true, p, sums, reinterpret_cast<float*>(ws), pixels, C, n_cl, S(stream));
}
int unetb200_dice_finalize(const float* sums, float smooth, float* out4, void* stream) {
  return dice_finalize(sums, smooth, out4, S(stream));
}
int unetb200_head_bwd(const void* x, const float* w, const float* y_true, const float* p, const float* sums,
                      float smooth, float grad_scale, void* dx, float* dw, float* db, void* ws, long long pixels,
                      int C, int n_cl, void* stream) {
  return head_bwd(x, w, y_true, p, sums, smooth, grad_scale, dx, dw, db, reinterpret_cast<float*>(ws), pixels, C, n_cl,
                  S(stream));
}

int unetb200_adam_step(float* theta, float* m, float* v, const float* g, long long n, const long long* step, float lr0,
                       int const_lr, float lr_fraction, float decay_steps, float beta1, float beta2, float eps,
                       float grad_scale, void* stream) {
  return adam_step(theta, m, v, g, n, step, lr0, const_lr, lr_fraction, decay_steps, beta1, beta2, eps, grad_scale,
                   S(stream));
}
int unetb200_sgd_step(float* theta, const float* g, long long n, const long long* step, float lr0, int const_lr,
                      float lr_fraction, float decay_steps, float grad_scale, void* stream) {
  return sgd_step(theta, g, n, step, lr0, const_lr, lr_fraction, decay_steps, grad_scale, S(stream));
}
int unetb200_step_increment(long long* step, void* stream) { return step_increment(step, S(stream)); }

int unetb200_dropout_mask(unsigned char* keep, long long n, float keep_prob, uint64_t seed, uint32_t stream_id,
                          void* stream) {
  return dropout_mask(keep, n, keep_prob, seed, stream_id, S(stream));
}
int unetb200_tmap_cache_stats(long long* hits, long long* misses) {
  if (hits) *hits = tmap_cache_hits();
  if (misses) *misses = tmap_cache_misses();
  return UB_OK;
}
int unetb200_gather_batch(const void* src, int src_dtype, long long src_rows, const long long* perm, long long first,
                          int rows, long long row_elems, float* dst, void* stream) {
  UB_REQUIRE(src_dtype == UNETB200_F32 || src_dtype == UNETB200_F64, "gather_batch: src_dtype %d is neither F32 nor F64",
             src_dtype);
  return gather_rows(src, src_dtype == UNETB200_F64, src_rows, perm, first, rows, row_elems, dst, S(stream));
}
int unetb200_cast_f32_to_bf16(const float* x, void* y, long long n, void* stream) {
  return cast_f32_to_bf16(x, y, n, S(stream));
}
int unetb200_cast_bf16_to_f32(const void* x, float* y, long long n, void* stream) {
  return cast_bf16_to_f32(x, y, n, S(stream));
}

// ---- the same operators with an explicit element type (UNETB200_BF16 or UNETB200_TF32; include/unetb200.h) ----------
#define UB_DT(dtype, what)                                                                                   \
  UB_REQUIRE((dtype) == UNETB200_BF16 || (dtype) == UNETB200_TF32, what ": dtype %d is neither BF16 nor TF32", \
             (dtype));                                                                                       \
  const int dt = (dtype) == UNETB200_TF32 ? UB_DT_TF32 : UB_DT_BF16

int unetb200_pack_conv3x3_fwd_dt(const float* w, void* out, int Cin, int Cout, int dtype, void* stream) {
  UB_DT(dtype, "pack_conv3x3_fwd");
  return pack_conv_fwd(w, out, 9, Cin, Cout, S(stream), dt);
}
int unetb200_pack_conv3x3_dgrad_dt(const float* w, void* out, int Cin, int Cout, int dtype, void* stream) {
  UB_DT(dtype, "pack_conv3x3_dgrad");
  return pack_conv_dgrad(w, out, 9, Cin, Cout, S(stream), dt);
}
int unetb200_pack_tconv2x2_fwd_dt(const float* w, void* out, int Cin, int Cout, int dtype, void* stream) {
  UB_DT(dtype, "pack_tconv2x2_fwd");
  return pack_tconv_fwd(w, out, Cin, Cout, S(stream), dt);
}
int unetb200_pack_tconv2x2_dgrad_dt(const float* w, void* out, int Cin, int Cout, int dtype, void* stream) {
  UB_DT(dtype, "pack_tconv2x2_dgrad");
  return pack_tconv_dgrad(w, out, Cin, Cout, S(stream), dt);
}
int unetb200_conv3x3_fwd_dt(const void* x0, int C0, const void* x1, int C1, const void* w_packed, const float* bias,
                            void* y, int N, int H, int W, int Cout, int relu, float drop_keep, uint64_t drop_seed,
                            uint32_t drop_stream, int dtype, void* stream) {
  UB_DT(dtype, "conv3x3_fwd");
  ConvSrc s0{x0, C0}, s1{x1, x1 ? C1 : 0};
  DropoutCfg d{drop_keep <= 0.0f ? 1.0f : drop_keep, drop_seed, drop_stream, nullptr};
  return conv3x3_fwd(s0, s1, w_packed, bias, y, N, H, W, Cout, relu, d, S(stream), nullptr, nullptr, dt);
}
int unetb200_conv3x3_dgrad_dt(const void* dz, int Cout, const void* w_packed_d, void* dx0, int C0, const void* mask0,
                              void* dx1, int C1, const void* mask1, float scale, int N, int H, int W, int dtype,
                              void* stream) {
  UB_DT(dtype, "conv3x3_dgrad");
  DgradDst d0{dx0, C0, mask0, scale}, d1{dx1, dx1 ? C1 : 0, mask1, scale};
  return conv3x3_dgrad(dz, Cout, w_packed_d, d0, d1, N, H, W, S(stream), dt);
}
int unetb200_conv3x3_wgrad_dt(const void* dz, int Cout, const void* x0, int C0, const void* x1, int C1, float* dw,
                              void* ws, size_t ws_bytes, int N, int H, int W, int dtype, void* stream) {
  UB_DT(dtype, "conv3x3_wgrad");
  ConvSrc s0{x0, C0}, s1{x1, x1 ? C1 : 0};
  return conv3x3_wgrad(dz, Cout, s0, s1, dw, ws, ws_bytes, N, H, W, S(stream), nullptr, dt);
}
int unetb200_tconv2x2_fwd_dt(const void* x, int Cin, const void* w_packed, const float* bias, void* y, int Cout, int N,
                             int H, int W, int dtype, void* stream) {
  UB_DT(dtype, "tconv2x2_fwd");
  return tconv2x2_fwd(x, Cin, w_packed, bias, y, Cout, N, H, W, S(stream), dt);
}
int unetb200_tconv2x2_dgrad_dt(const void* dy, int Cout, const void* w_packed_d, void* dx, int Cin, const void* mask,
                               float scale, int N, int H, int W, int dtype, void* stream) {
  UB_DT(dtype, "tconv2x2_dgrad");
  return tconv2x2_dgrad(dy, Cout, w_packed_d, dx, Cin, mask, scale, N, H, W, S(stream), nullptr, dt);
}
int unetb200_tconv2x2_wgrad_dt(const void* dy, int Cout, const void* x, int Cin, float* dw, void* ws, size_t ws_bytes,
                               int N, int H, int W, int dtype, void* stream) {
  UB_DT(dtype, "tconv2x2_wgrad");
  return tconv2x2_wgrad(dy, Cout, x, Cin, dw, ws, ws_bytes, N, H, W, S(stream), dt);
}
int unetb200_tconv2x2_wgrad_bias_dt(const void* dy, int Cout, const void* x, int Cin, float* dw, float* db, void* ws,
                                    size_t ws_bytes, int N, int H, int W, int dtype, void* stream) {
  UB_DT(dtype, "tconv2x2_wgrad_bias");
  return tconv2x2_wgrad(dy, Cout, x, Cin, dw, ws, ws_bytes, N, H, W, S(stream), dt, db);
}
int unetb200_conv_first_fwd_dt(const float* x, const float* w, const float* bias, void* y, int N, int H, int W, int Cin,
                               int Cout, int dtype, void* stream) {
  UB_DT(dtype, "conv_first_fwd");
  return conv_first_fwd(x, w, bias, y, N, H, W, Cin, Cout, S(stream), nullptr, dt);
}
int unetb200_conv_first_wgrad_dt(const float* x, const void* dz, float* dw, float* db, void* ws, int N, int H, int W,
                                 int Cin, int Cout, int dtype, void* stream) {
  UB_DT(dtype, "conv_first_wgrad");
  return conv_first_wgrad(x, dz, dw, db, reinterpret_cast<float*>(ws), N, H, W, Cin, Cout, S(stream), dt);
}
int unetb200_bias_grad_dt(const void* dz, float* db, void* ws, long long pixels, int C, int dtype, void* stream) {
  UB_DT(dtype, "bias_grad");
  return colsum_act(dz, db, reinterpret_cast<float*>(ws), pixels, C, S(stream), dt);
}
int unetb200_maxpool2x2_fwd_dt(const void* x, void* y, int N, int H, int W, int C, int dtype, void* stream) {
  UB_DT(dtype, "maxpool2x2_fwd");
  return maxpool2x2_fwd(x, y, N, H, W, C, S(stream), dt);
}
int unetb200_maxpool2x2_bwd_dt(const void* x, const void* dpool, const void* skip_grad, void* dz, int N, int H, int W,
                               int C, int dtype, void* stream) {
  UB_DT(dtype, "maxpool2x2_bwd");
  return maxpool2x2_bwd(x, dpool, skip_grad, dz, N, H, W, C, S(stream), dt);
}
int unetb200_upsample2x2_fwd_dt(const void* x, void* y, int N, int H, int W, int C, int dtype, void* stream) {
  UB_DT(dtype, "upsample2x2_fwd");
  return upsample2x2_fwd(x, y, N, H, W, C, S(stream), dt);
}
int unetb200_upsample2x2_bwd_dt(const void* dy, const void* x, void* dx, int N, int H, int W, int C, int relu_mask,
                                int dtype, void* stream) {
  UB_DT(dtype, "upsample2x2_bwd");
  return upsample2x2_bwd(dy, x, dx, N, H, W, C, relu_mask, S(stream), dt);
}
int unetb200_head_fwd_dt(const void* x, const float* w, const float* b, const float* y_true, float* p, float* sums,
                         void* ws, long long pixels, int C, int n_cl, int dtype, void* stream) {
  UB_DT(dtype, "head_fwd");
  return head_fwd(x, w, b, y_true, p, sums, reinterpret_cast<float*>(ws), pixels, C, n_cl, S(stream), dt);
}
int unetb200_head_bwd_dt(const void* x, const float* w, const float* y_true, const float* p, const float* sums,
                         float smooth, float grad_scale, void* dx, float* dw, float* db, void* ws, long long pixels,
                         int C, int n_cl, int dtype, void* stream) {
  UB_DT(dtype, "head_bwd");
  return head_bwd(x, w, y_true, p, sums, smooth, grad_scale, dx, dw, db, reinterpret_cast<float*>(ws), pixels, C, n_cl,
                  S(stream), dt);
}

}  // extern "C"
