/* libunetb200 -- C ABI of the B200 (sm_100a) UNet training hot path.
 *
 * The reference (dmsuehir/unet_distributed) has no FFI of its own: its hot path is the TensorFlow-1.4 graph that
 * model.py:36-130 builds and test_dist.py:396-398 runs. Each entry point below replaces the TF op(s) one reference
 * line lowers to; the citation names that line. Host code (Python, ctypes) owns every buffer; this library never
 * allocates or frees caller memory.
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the name ends in _host; activations are NHWC bf16 (16-byte aligned) --
 *     or, through the `_dt` entry points with UNETB200_TF32, NHWC fp32 holding tf32-rounded values --
 *     master weights / gradients / optimizer state are fp32 in the TensorFlow variable layouts
 *     (Conv2D kernel [kh][kw][Cin][Cout]; Conv2DTranspose kernel [kh][kw][Cout][Cin]);
 *   - `stream` is a cudaStream_t passed as void*; every call is asynchronous on it and never synchronises;
 *   - return value 0 = OK, negative = error (unetb200_last_error_string() gives the text for the calling thread);
 *   - `ws` arguments are caller-provided scratch of at least the matching *_ws_bytes() size.
 */
#ifndef UNETB200_H_
#define UNETB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Element types named across the ABI. Activations of the tile engines are BF16 (default) or, in the fp32/tf32 precision
 * mode, F32 computed with kind::tf32 tensor-core MMAs (the reference computes in fp32, test_dist.py:196-202). */
enum { UNETB200_BF16 = 0, UNETB200_TF32 = 1, UNETB200_F32 = 2, UNETB200_F64 = 3 };

int unetb200_version(void);
const char* unetb200_last_error_string(void);

/* ---- weight packing: fp32 TF-layout master -> bf16 GEMM operand (run after every optimizer update) ------------- */
/* Conv2D kernel (model.py:47-49, he_uniform HWIO) -> [Cout][9][Cin] for the forward implicit GEMM. */
int unetb200_pack_conv3x3_fwd(const float* w_hwio, void* w_packed, int Cin, int Cout, void* stream);
/* Same kernel, tap-flipped and transposed -> [Cin][9][Cout] for Conv2DBackpropInput. */
int unetb200_pack_conv3x3_dgrad(const float* w_hwio, void* w_packed, int Cin, int Cout, void* stream);
/* Conv2DTranspose kernel (model.py:79-80, [2][2][Cout][Cin]) -> [4*Cout][Cin] (a cast). */
int unetb200_pack_tconv2x2_fwd(const float* w, void* w_packed, int Cin, int Cout, void* stream);
/* -> [Cin][4][Cout] for the transposed conv's input gradient. */
int unetb200_pack_tconv2x2_dgrad(const float* w, void* w_packed, int Cin, int Cout, void* stream);

/* ---- Conv2D 3x3 'same' + bias + ReLU: model.py:51-117 (18 call sites) ------------------------------------------ */
/* Forward. Two sources implement concatenate([x0, x1], -1) (model.py:76-80,...) without materialising it; pass
 * x1=NULL,C1=0 for a single input. Dropout (model.py:60,66; keep<1) is applied after ReLU with scale 1/keep;
 * element i of the output keeps iff the Philox stream (seed, stream_id) says so -- see unetb200_dropout_mask. */
int unetb200_conv3x3_fwd(const void* x0, int C0, const void* x1, int C1, const void* w_packed, const float* bias,
                         void* y, int N, int H, int W, int Cout, int relu, float drop_keep, uint64_t drop_seed,
                         uint32_t drop_stream, void* stream);
/* Conv2DBackpropInput fused with the producer's ReluGrad: dx_k = dgrad(dz)[source k] * (mask_k > 0 ? scale : 0).
 * mask_k is the producer's forward output (NULL = no activation to undo); scale = 1/keep undoes dropout. */
int unetb200_conv3x3_dgrad(const void* dz, int Cout, const void* w_packed_d, void* dx0, int C0, const void* mask0,
                           void* dx1, int C1, const void* mask1, float scale, int N, int H, int W, void* stream);
/* Conv2DBackpropFilter: dw fp32 [3][3][C0+C1][Cout]. */
size_t unetb200_conv3x3_wgrad_ws_bytes(int N, int H, int W, int Cin_total, int Cout);
int unetb200_conv3x3_wgrad(const void* dz, int Cout, const void* x0, int C0, const void* x1, int C1, float* dw,
                           void* ws, size_t ws_bytes, int N, int H, int W, void* stream);
/* First layer (model.py:51, conv1a): fp32 NHWC network input with Cin in 1..8, fp32 HWIO weights, bf16 output. */
int unetb200_conv_first_fwd(const float* x, const float* w_hwio, const float* bias, void* y, int N, int H, int W,
                            int Cin, int Cout, void* stream);
size_t unetb200_conv_first_wgrad_ws_bytes(int Cin, int Cout);
int unetb200_conv_first_wgrad(const float* x, const void* dz, float* dw, float* db, void* ws, int N, int H, int W,
                              int Cin, int Cout, void* stream);
/* BiasAddGrad: db[c] = sum over N*H*W of dz[., c]. */
size_t unetb200_bias_grad_ws_bytes(int C);
int unetb200_bias_grad(const void* dz, float* db, void* ws, long long pixels, int C, void* stream);
/* No counterpart in the reference: the same bias gradient for callers that cannot hand over a workspace because
 * they cannot allocate one -- plug-ins whose host application owns all device memory and forbids cudaMalloc on its
 * worker threads, training loops run under a fixed memory budget with a no-allocation rule, embedders that only
 * ever see the tensors they are given. Inputs and outputs are still the caller's; the partial sums live in a
 * 2,424,832-byte device array built into the library
 * (every GPU that loads the library holds it until the process exits). Limit: C <= 512 (8 * power of two); larger C
 * returns an error that names the limit -- use unetb200_bias_grad then. The first call on each device looks up the
 * array's device address (host-side work): make that call before any stream capture starts. All calls on a device
 * share the array; calls that overlap on the GPU (different streams) race. Calls neither cudaMalloc nor cudaFree
 * nor any allocating routine; nothing in the array needs clearing between calls. */
int unetb200_bias_grad_static(const void* dz, float* db, long long pixels, int C, void* stream);

/* ---- MaxPooling2D 2x2: model.py:53,57,63,69 -------------------------------------------------------------------- */
int unetb200_maxpool2x2_fwd(const void* x, void* y, int N, int H, int W, int C, void* stream);
/* MaxPoolGrad + ReluGrad of the pooled conv + the skip-connection gradient:
 * dz = (skip_grad or 0) + route_to_first_max(dpool) * (x > 0). skip_grad may be NULL or alias dz. */
int unetb200_maxpool2x2_bwd(const void* x, const void* dpool, const void* skip_grad, void* dz, int N, int H, int W,
                            int C, void* stream);

/* ---- decoder up path ------------------------------------------------------------------------------------------- */
/* Conv2DTranspose 2x2 stride 2 (model.py:79-80,90-91,101-102,112-113); (N,H,W) is the INPUT grid. */
int unetb200_tconv2x2_fwd(const void* x, int Cin, const void* w_packed, const float* bias, void* y, int Cout, int N,
                          int H, int W, void* stream);
int unetb200_tconv2x2_dgrad(const void* dy, int Cout, const void* w_packed_d, void* dx, int Cin, const void* mask,
                            float scale, int N, int H, int W, void* stream);
size_t unetb200_tconv2x2_wgrad_ws_bytes(int N, int H, int W, int Cin, int Cout);
int unetb200_tconv2x2_wgrad(const void* dy, int Cout, const void* x, int Cin, float* dw, void* ws, size_t ws_bytes,
                            int N, int H, int W, void* stream);
/* UpSampling2D 2x2 nearest (model.py:76,87,98,109; --use_upsampling); (N,H,W) is the INPUT grid. */
int unetb200_upsample2x2_fwd(const void* x, void* y, int N, int H, int W, int C, void* stream);
int unetb200_upsample2x2_bwd(const void* dy, const void* x, void* dx, int N, int H, int W, int C, int relu_mask,
                             void* stream);

/* ---- head: Conv2D 1x1 + sigmoid (model.py:119-120) fused with dice_coef / dice_coef_loss (model.py:4-21) ------- */
size_t unetb200_head_ws_bytes(long long pixels, int C, int n_cl);
/* p = sigmoid(x.w + b) (fp32 [pixels][n_cl]); sums[3] = {sum y*p, sum y, sum p} over the whole batch tensor. */
int unetb200_head_fwd(const void* x, const float* w, const float* b, const float* y_true, float* p, float* sums,
                      void* ws, long long pixels, int C, int n_cl, void* stream);
/* out4 = {dice_coef_loss, dice_coef, sensitivity, specificity} (model.py:13-21, 4-10, 133-137, 139-143). */
int unetb200_dice_finalize(const float* sums, float smooth, float* out4, void* stream);
/* Stand-alone forms of the same reductions on arbitrary fp32 tensors of n elements, for the eager call pattern
 * `dice_coef_loss(msks, preds)` of test_dist.py:211-215. which: 0 loss, 1 dice, 2 sensitivity, 3 specificity;
 * dp = (*upstream or 1) * d(metric)/d(y_pred). */
size_t unetb200_dice_sums_ws_bytes(void);
int unetb200_dice_sums(const float* y_true, const float* y_pred, long long n, float* sums, void* ws, void* stream);
int unetb200_dice_grad(const float* y_true, const float* sums, float smooth, int which, const float* upstream,
                       float* dp, long long n, void* stream);
/* Gradient of dice_coef_loss through the sigmoid and the 1x1 conv, times grad_scale; dx is masked by x > 0.
 * With sums == NULL, y_true is taken to be an explicit dL/dp instead of the labels (generic autograd). */
int unetb200_head_bwd(const void* x, const float* w, const float* y_true, const float* p, const float* sums,
                      float smooth, float grad_scale, void* dx, float* dw, float* db, void* ws, long long pixels,
                      int C, int n_cl, void* stream);

/* ---- the fp32/tf32 precision mode: every operator above with an explicit element type ----------------------------- */
/* The reference computes in fp32 (test_dist.py:196-202). `_dt` variants take dtype = UNETB200_BF16 (identical to the
 * plain entry point) or UNETB200_TF32: activations, activation gradients, masks and packed weights are then fp32 words
 * (16-byte aligned tensors, same NHWC / packed layouts), rounded to tf32 by whatever produced them; the convolutions
 * run on the tensor cores with tcgen05 kind::tf32 MMAs and fp32 accumulation, the bandwidth kernels in fp32.
 * Every `_dt` output tensor is itself tf32-rounded. Arguments other than dtype are those of the plain entry point. */
int unetb200_pack_conv3x3_fwd_dt(const float* w_hwio, void* w_packed, int Cin, int Cout, int dtype, void* stream);
int unetb200_pack_conv3x3_dgrad_dt(const float* w_hwio, void* w_packed, int Cin, int Cout, int dtype, void* stream);
int unetb200_pack_tconv2x2_fwd_dt(const float* w, void* w_packed, int Cin, int Cout, int dtype, void* stream);
int unetb200_pack_tconv2x2_dgrad_dt(const float* w, void* w_packed, int Cin, int Cout, int dtype, void* stream);
int unetb200_conv3x3_fwd_dt(const void* x0, int C0, const void* x1, int C1, const void* w_packed, const float* bias,
                            void* y, int N, int H, int W, int Cout, int relu, float drop_keep, uint64_t drop_seed,
                            uint32_t drop_stream, int dtype, void* stream);
int unetb200_conv3x3_dgrad_dt(const void* dz, int Cout, const void* w_packed_d, void* dx0, int C0, const void* mask0,
                              void* dx1, int C1, const void* mask1, float scale, int N, int H, int W, int dtype,
                              void* stream);
int unetb200_conv3x3_wgrad_dt(const void* dz, int Cout, const void* x0, int C0, const void* x1, int C1, float* dw,
                              void* ws, size_t ws_bytes, int N, int H, int W, int dtype, void* stream);
int unetb200_tconv2x2_fwd_dt(const void* x, int Cin, const void* w_packed, const float* bias, void* y, int Cout, int N,
                             int H, int W, int dtype, void* stream);
int unetb200_tconv2x2_dgrad_dt(const void* dy, int Cout, const void* w_packed_d, void* dx, int Cin, const void* mask,
                               float scale, int N, int H, int W, int dtype, void* stream);
int unetb200_tconv2x2_wgrad_dt(const void* dy, int Cout, const void* x, int Cin, float* dw, void* ws, size_t ws_bytes,
                               int N, int H, int W, int dtype, void* stream);
/* Weight AND bias gradient of Conv2DTranspose in one pass over dy (db may be NULL); same workspace size as
 * unetb200_tconv2x2_wgrad_ws_bytes. TensorFlow computes the two with separate ops (Conv2DBackpropFilter, BiasAddGrad). */
int unetb200_tconv2x2_wgrad_bias_dt(const void* dy, int Cout, const void* x, int Cin, float* dw, float* db, void* ws,
                                    size_t ws_bytes, int N, int H, int W, int dtype, void* stream);
int unetb200_conv_first_fwd_dt(const float* x, const float* w_hwio, const float* bias, void* y, int N, int H, int W,
                               int Cin, int Cout, int dtype, void* stream);
int unetb200_conv_first_wgrad_dt(const float* x, const void* dz, float* dw, float* db, void* ws, int N, int H, int W,
                                 int Cin, int Cout, int dtype, void* stream);
int unetb200_bias_grad_dt(const void* dz, float* db, void* ws, long long pixels, int C, int dtype, void* stream);
int unetb200_maxpool2x2_fwd_dt(const void* x, void* y, int N, int H, int W, int C, int dtype, void* stream);
int unetb200_maxpool2x2_bwd_dt(const void* x, const void* dpool, const void* skip_grad, void* dz, int N, int H, int W,
                               int C, int dtype, void* stream);
int unetb200_upsample2x2_fwd_dt(const void* x, void* y, int N, int H, int W, int C, int dtype, void* stream);
int unetb200_upsample2x2_bwd_dt(const void* dy, const void* x, void* dx, int N, int H, int W, int C, int relu_mask,
                                int dtype, void* stream);
int unetb200_head_fwd_dt(const void* x, const float* w, const float* b, const float* y_true, float* p, float* sums,
                         void* ws, long long pixels, int C, int n_cl, int dtype, void* stream);
int unetb200_head_bwd_dt(const void* x, const float* w, const float* y_true, const float* p, const float* sums,
                         float smooth, float grad_scale, void* dx, float* dw, float* db, void* ws, long long pixels,
                         int C, int n_cl, int dtype, void* stream);

/* ---- optimizer: tf.train.AdamOptimizer + exponential_decay (test_dist.py:228-248) ------------------------------ */
/* step is a device int64 holding global_step (updates applied so far); the LR uses its value before the increment.
 * grad_scale multiplies g first (1/replicas: SyncReplicasOptimizer averages, test_dist.py:251-258). */
int unetb200_adam_step(float* theta, float* m, float* v, const float* g, long long n, const long long* step, float lr0,
                       int const_lr, float lr_fraction, float decay_steps, float beta1, float beta2, float eps,
                       float grad_scale, void* stream);
/* GradientDescentOptimizer (commented alternative at test_dist.py:239-244). */
int unetb200_sgd_step(float* theta, const float* g, long long n, const long long* step, float lr0, int const_lr,
                      float lr_fraction, float decay_steps, float grad_scale, void* stream);
int unetb200_step_increment(long long* step, void* stream);

/* ---- utilities --------------------------------------------------------------------------------------------------- */
/* keep[i] in {0,1}: the exact dropout decisions the conv epilogue makes for output element i (n multiple of 8). */
int unetb200_dropout_mask(unsigned char* keep, long long n, float keep_prob, uint64_t seed, uint32_t stream_id,
                          void* stream);
/* Measurement hooks (no reference counterpart; bench.py uses them). launch_count: kernels this library has launched
 * in this process. With profiling enabled every tile-engine launch is bracketed by CUDA events on its stream;
 * profile_collect synchronises and returns, per kind {0 igemm (conv/tconv fwd+dgrad), 1 wgrad, 2 other}, the summed
 * milliseconds, algorithmic FLOPs, minimum HBM bytes and launch count (arrays of 3), then clears the records. */
/* Size every grid launched from now on (host-side state, read at launch time, so also at graph capture) for the
 * device's SM count minus n: SMs left free for kernels that run beside this library's persistent one-CTA-per-SM
 * kernels -- the NCCL all-reduces of the gradient buckets (test_dist.py:251-258) during backward. Returns the
 * previous reservation; 0 (default) uses every SM. No reference counterpart. */
int unetb200_set_reserved_sms(int n);
long long unetb200_launch_count(void);
/* Encoded TMA descriptors are cached by (base pointer, element size, dims, strides, box, swizzle), so the eager
 * per-operator path does not re-encode ~150 tensor maps per step; hits / misses of that cache since load. */
int unetb200_tmap_cache_stats(long long* hits, long long* misses);
int unetb200_profile_enable(int on);
int unetb200_profile_collect(double* ms, double* flops, double* bytes, long long* launches);
/* One shuffled batch, or one replica's shard of it, cut out of a training set that stays in HBM: data.py:29-54
 * (get_epoch: shuffle, drop the remainder, reshape into batches) and the shard slice of test_dist.py:390-394.
 * dst[r][0..row_elems) = (float) src[perm[first + r]][...] for r in [0, rows); perm is a DEVICE int64 permutation
 * (NULL = identity). src rows are UNETB200_F32 or UNETB200_F64 (update_channels returns float64 arrays,
 * preprocess.py:301-302, which TF casts at feed time); row_elems (= H*W*C) must be a multiple of 4. */
int unetb200_gather_batch(const void* src, int src_dtype, long long src_rows, const long long* perm, long long first,
                          int rows, long long row_elems, float* dst, void* stream);
int unetb200_cast_f32_to_bf16(const float* x, void* y, long long n, void* stream);
int unetb200_cast_bf16_to_f32(const void* x, float* y, long long n, void* stream);

/* ---- whole-network plan: the graph define_model builds (model.py:36-130) as one launch sequence ---------------- */
/* The plan allocates nothing on the device: query unetb200_plan_arena_bytes(), allocate, then bind the arena and the
 * flat fp32 parameter / gradient buffers. Variables sit in the flat buffers in TF creation order with TF names
 * ("conv1a/kernel", "transConv6/bias", "Mask/kernel", ...) and TF layouts; each starts on a 4-float boundary. */
typedef struct unetb200_plan unetb200_plan;
int unetb200_plan_create(unetb200_plan** out, int N, int H, int W, int in_channels, int n_cl_out, int use_upsampling,
                         float dropout);
/* Same plan in a chosen precision. dtype = UNETB200_BF16 (what unetb200_plan_create builds) or UNETB200_TF32: the
 * fp32/tf32 mode -- activations, activation gradients and packed GEMM operands are fp32 words rounded to tf32 by
 * their producers, every conv / transposed conv runs on the tensor cores with kind::tf32 MMAs (fp32 accumulation),
 * and the bandwidth kernels (pooling, upsampling, head, conv1a, bias gradients) run in fp32. The reference computes in
 * fp32 (test_dist.py:196-202); this mode is the one its 1e-3 tolerance is stated for. Arena tensors then have
 * elem_bytes = 4 (unetb200_plan_tensor). */
int unetb200_plan_create_dt(unetb200_plan** out, int N, int H, int W, int in_channels, int n_cl_out,
                            int use_upsampling, float dropout, int dtype);
void unetb200_plan_destroy(unetb200_plan* plan);
long long unetb200_plan_num_params(const unetb200_plan* plan); /* floats in the flat buffers, padding included */
int unetb200_plan_num_variables(const unetb200_plan* plan);
int unetb200_plan_variable(const unetb200_plan* plan, int i, char* name, int name_cap, long long* offset, int* rank,
                           int* shape4);
size_t unetb200_plan_arena_bytes(const unetb200_plan* plan);
int unetb200_plan_bind(unetb200_plan* plan, void* arena, float* params, float* grads);
/* The arena's sub-buffers (activations, gradients, gate bits, packed operands, workspaces) as [offset, offset + bytes)
 * in allocation order; nothing outside them is ever written. With UNETB200_ARENA_GUARD=<bytes> in the environment at
 * plan creation that many unused bytes follow every sub-buffer: the tests fill the arena with a pattern and check that
 * a training step leaves every byte outside the regions untouched (compute-sanitizer is closed on the GPU pool). */
int unetb200_plan_num_regions(const unetb200_plan* plan);
int unetb200_plan_region(const unetb200_plan* plan, int i, long long* offset, long long* bytes);
/* Arena offset and shape of an activation ("a1".."a9", "c1".."c9", "p1".."p4", "u6".."u9"), its gradient ("g"+name),
 * "preds" (fp32) or "metrics" (fp32 {loss, dice, sensitivity, specificity}). */
int unetb200_plan_tensor(const unetb200_plan* plan, const char* name, long long* offset, int* n, int* h, int* w,
                         int* c, int* elem_bytes);
int unetb200_plan_pack_weights(unetb200_plan* plan, void* stream);
/* Forward (model.py:45-120) and, when msks != NULL, the dice metrics. imgs fp32 [N,H,W,Cin], msks fp32 [N,H,W,n_cl].
 * Dropout stream seed = seed + *seed_add (seed_add: optional device int64, e.g. the global step). */
int unetb200_plan_forward(unetb200_plan* plan, const float* imgs, const float* msks, uint64_t seed,
                          const long long* seed_add, void* stream);
/* Backward in unetb200_plan_num_segments() (= 5) segments, in this order: 0 head + conv9/conv8 levels, 1 conv7/conv6
 * levels, 2 conv5a/b, 3 conv4/conv3, 4 conv2/conv1. Gradients land in the bound flat buffer, multiplied by grad_scale.
 * The loss is dice_coef_loss(msks, preds) of the last forward, unless dpreds (fp32 [N,H,W,n_cl], an explicit
 * dL/dpreds) is given. unetb200_plan_bucket gives the contiguous flat range that is complete after a segment (the
 * buckets tile [0, num_params) exactly): its all-reduce (SyncReplicasOptimizer, test_dist.py:251-258) and optimizer
 * update can start while later segments run; unetb200_plan_pack_range then refreshes the packed GEMM operands of
 * just that range. */
int unetb200_plan_num_segments(const unetb200_plan* plan);
int unetb200_plan_backward(unetb200_plan* plan, const float* imgs, const float* msks, const float* dpreds,
                           float grad_scale, int segment, void* stream);
int unetb200_plan_bucket(const unetb200_plan* plan, int segment, long long* begin, long long* end);
int unetb200_plan_pack_range(unetb200_plan* plan, long long begin, long long end, void* stream);
/* The weight gradients of plan_backward run on a second, plan-owned stream that is forked from and joined back into
 * the caller's stream inside every segment (so the call stays stream-ordered and capturable). enable = 0 keeps all
 * kernels on the caller's stream in program order (used when timing kernels one at a time); returns the old setting.
 * No reference counterpart: TensorFlow's executor schedules independent ops concurrently on its own. */
int unetb200_plan_side_stream(unetb200_plan* plan, int enable);
/* By default every plan_backward segment ends with the caller's stream waiting for the segment's weight gradients.
 * With defer_join enabled it does not: the dgrad chain of the next segment starts without waiting, and whoever
 * consumes the gradients calls unetb200_plan_join(plan, stream) to make ITS stream (e.g. the one that all-reduces and
 * updates a bucket) wait for all weight gradients enqueued so far. defer_join returns the previous setting. */
int unetb200_plan_defer_join(unetb200_plan* plan, int enable);
int unetb200_plan_join(unetb200_plan* plan, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* UNETB200_H_ */
