// UNet step plan: the layer graph of model.py:36-130 as a fixed launch sequence over one caller-owned arena.
// The plan owns no device memory: the caller asks for the arena size, allocates it (PyTorch does) and binds it together
// with the flat fp32 parameter / gradient buffers whose layout (TensorFlow variable order and shapes) the plan reports.
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/unetb200.h"
#include "common.h"
#include "elementwise.cuh"
#include "ops.cuh"

using namespace ub;

namespace {

struct Var {
  std::string name;
  long long offset;
  int rank;
  int shape[4];
  long long count() const {
    long long n = 1;
    for (int i = 0; i < rank; ++i) n *= shape[i];
    return n;
  }
};

struct Layer {
  std::string name;
  int kind;  // 0 conv3x3, 1 tconv, 2 mask
  int cin, cout;
  long long w_off, b_off;     // into the flat parameter buffer (floats)
  size_t pk_fwd, pk_dgrad;    // arena offsets of the packed bf16 operands
};

constexpr size_t kAlign = 1024;
constexpr int kSegments = 5;   // backward segments = gradient buckets
constexpr size_t kNoGate = ~static_cast<size_t>(0);

// NVTX range over one host-side launch sequence (forward, a backward segment): shows up in ncu / Nsight timelines,
// costs nothing without a tool attached (header-only NVTX v3).
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

}  // namespace

struct unetb200_plan {
  int N, H, W, Cin, ncl, ups;
  int dt = UB_DT_BF16;   // element type of activations, gradients and packed operands (UB_DT_TF32: fp32 words)
  int elt = 2;
  float dropout;
  std::vector<Var> vars;
  std::vector<Layer> layers;
  long long nparams = 0;
  size_t arena_bytes = 0;
  // layer indices
  int enc_a[5], enc_b[5], dec_a[4], dec_b[4], tc[4], mask;
  // arena offsets: activations and their gradients (bf16 NHWC)
  size_t a[5], c[5], p[4], u[4], da[4], dc[4];
  size_t g_a[5], g_c[5], g_p[4], g_u[4], g_da[4], g_dc[4];
  // ReLU gates of the activations the backward pass masks by, one bit per element (uint16 per 16 channels); kNoGate
  // where nothing consumes them (the consumer then reads the activation tensor itself)
  size_t gt_a[5], gt_c[5], gt_da[4], gt_dc[4];
  size_t preds, sums, metrics, ws, ws_side;
  size_t drop_bits[5];   // keep bits of the dropout layers (conv3a, conv4a: levels 2, 3), generated once per forward
  std::vector<PackLayer> pack_layers;
  int pack_tiles = 0;
  size_t pack_table = 0;
  size_t ws_bytes = 0;        // main-stream workspace (head, dice)
  size_t ws_side_bytes = 0;   // weight-gradient workspace (side stream)
  // Weight gradients run on a second stream, forked from / joined to the caller's stream with events: they depend
  // only on tensors the main chain (dgrads, pooling) has already produced, so their CTAs fill the SMs that the tail
  // of a persistent dgrad kernel leaves idle, and vice versa. UNETB200_NO_SIDE_STREAM=1 keeps everything in order.
  cudaStream_t side = nullptr;
  cudaEvent_t ev[40];
  int n_ev = 0, ev_next = 0;
  bool use_side = true;
  // deferred join: plan_backward does not make the caller's stream wait for the weight gradients at the end of a
  // segment; whoever consumes the gradients calls unetb200_plan_join(plan, stream) on ITS stream (the trainer's
  // update stream), so the dgrad chain never stalls on the side stream between segments
  bool defer_join = false;
  int ch[5];
  int c4_out;      // channels of c[4] (conv5b)
  int dc_out[4];   // channels of dc[L] (conv{9-L}b)
  bool pool_fused[4];   // upsampling decoder: ResizeNearestNeighborGrad of level L fused into conv{9-L}a's dgrad
  // bound buffers
  uint8_t* arena = nullptr;
  float* params = nullptr;
  float* grads = nullptr;

  // every sub-buffer of the arena, in allocation order (unetb200_plan_region); `guard` bytes are left unused after each
  // one when UNETB200_ARENA_GUARD is set at plan creation (out-of-bounds write detection in the tests)
  std::vector<std::pair<size_t, size_t>> regions;
  size_t guard = 0;
  size_t bump(size_t bytes) {
    size_t off = arena_bytes;
    arena_bytes += (bytes + kAlign - 1) / kAlign * kAlign + guard;
    regions.emplace_back(off, bytes);
    return off;
  }
  int Hl(int L) const { return H >> L; }
  int Wl(int L) const { return W >> L; }
  size_t act_bytes(int L, int C) const { return static_cast<size_t>(N) * Hl(L) * Wl(L) * C * elt; }
  size_t gate_bytes(int L, int C) const { return static_cast<size_t>(N) * Hl(L) * Wl(L) * C / 8; }   // one bit each
  void* at(size_t off) const { return arena + off; }
  unsigned short* gate(size_t off) const {
    return off == kNoGate ? nullptr : reinterpret_cast<unsigned short*>(arena + off);
  }
  const float* W_(int li) const { return params + layers[li].w_off; }
  const float* B_(int li) const { return params + layers[li].b_off; }
  float* dW_(int li) const { return grads + layers[li].w_off; }
  float* dB_(int li) const { return grads + layers[li].b_off; }
  const void* pkf(int li) const { return arena + layers[li].pk_fwd; }
  const void* pkd(int li) const { return arena + layers[li].pk_dgrad; }
};

namespace {

int add_layer(unetb200_plan* P, const std::string& name, int kind, int cin, int cout) {
  Layer l;
  l.name = name;
  l.kind = kind;
  l.cin = cin;
  l.cout = cout;
  Var k, b;
  k.name = name + "/kernel";
  k.rank = 4;
  if (kind == 0) { k.shape[0] = 3; k.shape[1] = 3; k.shape[2] = cin; k.shape[3] = cout; }
  else if (kind == 1) { k.shape[0] = 2; k.shape[1] = 2; k.shape[2] = cout; k.shape[3] = cin; }
  else { k.shape[0] = 1; k.shape[1] = 1; k.shape[2] = cin; k.shape[3] = cout; }
  // every variable starts on a 16-byte boundary of the flat buffer (vectorised Adam, float4 bias loads)
  P->nparams = (P->nparams + 3) / 4 * 4;
  k.offset = P->nparams;
  P->nparams += k.count();
  P->nparams = (P->nparams + 3) / 4 * 4;
  b.name = name + "/bias";
  b.rank = 1;
  b.shape[0] = cout; b.shape[1] = b.shape[2] = b.shape[3] = 1;
  b.offset = P->nparams;
  P->nparams += cout;
  l.w_off = k.offset;
  l.b_off = b.offset;
  l.pk_fwd = l.pk_dgrad = 0;
  P->vars.push_back(k);
  P->vars.push_back(b);
  P->layers.push_back(l);
  return static_cast<int>(P->layers.size()) - 1;
}

size_t max_sz(size_t a, size_t b) { return a > b ? a : b; }

}  // namespace

extern "C" {

int unetb200_plan_create(unetb200_plan** out, int N, int H, int W, int in_channels, int n_cl_out, int use_upsampling,
                         float dropout) {
  return unetb200_plan_create_dt(out, N, H, W, in_channels, n_cl_out, use_upsampling, dropout, UNETB200_BF16);
}

int unetb200_plan_create_dt(unetb200_plan** out, int N, int H, int W, int in_channels, int n_cl_out,
                            int use_upsampling, float dropout, int dtype) {
  UB_REQUIRE(out != nullptr, "plan_create: null output");
  UB_REQUIRE(dtype == UNETB200_BF16 || dtype == UNETB200_TF32, "plan_create: dtype %d is neither BF16 nor TF32", dtype);
  UB_REQUIRE(N >= 1 && H >= 16 && W >= 16 && H % 16 == 0 && W % 16 == 0,
             "plan_create: image %dx%d must be a multiple of 16 (four 2x2 poolings), batch %d >= 1", H, W, N);
  UB_REQUIRE(in_channels >= 1 && in_channels <= 8, "plan_create: in_channels %d outside 1..8", in_channels);
  UB_REQUIRE(n_cl_out >= 1 && n_cl_out <= 4, "plan_create: n_cl_out %d outside 1..4", n_cl_out);
  UB_REQUIRE(dropout >= 0.0f && dropout < 1.0f, "plan_create: dropout %f outside [0,1)", dropout);
  unetb200_plan* P = new unetb200_plan();
  P->N = N; P->H = H; P->W = W; P->Cin = in_channels; P->ncl = n_cl_out; P->ups = use_upsampling ? 1 : 0;
  P->dropout = dropout;
  if (const char* g = getenv("UNETB200_ARENA_GUARD")) {
    const long long gb = atoll(g);
    UB_REQUIRE(gb >= 0 && gb % static_cast<long long>(kAlign) == 0 && gb <= (1ll << 24),
               "plan_create: UNETB200_ARENA_GUARD %lld must be a multiple of %zu (at most 16 MiB)", gb, kAlign);
    P->guard = static_cast<size_t>(gb);
  }
  P->dt = dtype == UNETB200_TF32 ? UB_DT_TF32 : UB_DT_BF16;
  P->elt = dt_bytes(P->dt);
  for (int L = 0; L < 5; ++L) P->ch[L] = 32 << L;
  const int up = P->ups;
  // ---- variables in creation order (model.py:51-120)
  static const char* enc_names[5] = {"conv1", "conv2", "conv3", "conv4", "conv5"};
  for (int L = 0; L < 5; ++L) {
    const int cin = L == 0 ? in_channels : P->ch[L - 1];
    P->enc_a[L] = add_layer(P, std::string(enc_names[L]) + "a", 0, cin, P->ch[L]);
    const int cout_b = (L == 4 && up) ? 256 : P->ch[L];
    P->enc_b[L] = add_layer(P, std::string(enc_names[L]) + "b", 0, P->ch[L], cout_b);
  }
  P->c4_out = up ? 256 : 512;
  for (int L = 3; L >= 0; --L) {
    const std::string idx = std::to_string(9 - L);
    const int below = (L == 3) ? P->c4_out : P->dc_out[L + 1];  // channels of the tensor one level down
    if (!up) P->tc[L] = add_layer(P, "transConv" + idx, 1, below, P->ch[L]);
    else P->tc[L] = -1;
    P->dec_a[L] = add_layer(P, "conv" + idx + "a", 0, 2 * P->ch[L], P->ch[L]);
    P->dc_out[L] = (up && L > 0) ? P->ch[L] / 2 : P->ch[L];
    P->dec_b[L] = add_layer(P, "conv" + idx + "b", 0, P->ch[L], P->dc_out[L]);
  }
  P->mask = add_layer(P, "Mask", 2, 32, n_cl_out);
  P->nparams = (P->nparams + 3) / 4 * 4;

  // ---- arena: packed weights
  for (auto& l : P->layers) {
    if (l.kind == 2 || l.cin < 32) continue;  // Mask and conv1a run from the fp32 master
    const size_t bytes = static_cast<size_t>(l.kind == 0 ? 9 : 4) * l.cin * l.cout * P->elt;
    l.pk_fwd = P->bump(bytes);
    l.pk_dgrad = P->bump(bytes);
  }
  // ---- arena: activations and gradients
  for (int L = 0; L < 5; ++L) {
    P->a[L] = P->bump(P->act_bytes(L, P->ch[L]));
    P->g_a[L] = P->bump(P->act_bytes(L, P->ch[L]));
    const int cb = L == 4 ? P->c4_out : P->ch[L];
    P->c[L] = P->bump(P->act_bytes(L, cb));
    P->g_c[L] = P->bump(P->act_bytes(L, cb));
  }
  for (int L = 0; L < 4; ++L) {
    P->p[L] = P->bump(P->act_bytes(L + 1, P->ch[L]));
    P->g_p[L] = P->bump(P->act_bytes(L + 1, P->ch[L]));
    P->u[L] = P->bump(P->act_bytes(L, P->ch[L]));
    P->g_u[L] = P->bump(P->act_bytes(L, P->ch[L]));
    P->da[L] = P->bump(P->act_bytes(L, P->ch[L]));
    P->g_da[L] = P->bump(P->act_bytes(L, P->ch[L]));
    P->dc[L] = P->bump(P->act_bytes(L, P->dc_out[L]));
    P->g_dc[L] = P->bump(P->act_bytes(L, P->dc_out[L]));
  }
  // ---- arena: ReLU gate bits (1/16 of the activation they describe)
  for (int L = 0; L < 5; ++L) {
    const bool a_gates = L > 0 || (P->dt == UB_DT_BF16 && conv_first_fwd_writes_gates(in_channels, 32));
    P->gt_a[L] = a_gates ? P->bump(P->gate_bytes(L, P->ch[L])) : kNoGate;
    const bool c_gates = L < 4 || !up;   // c[4] is masked by the transposed-conv dgrad only
    P->gt_c[L] = c_gates ? P->bump(P->gate_bytes(L, L == 4 ? P->c4_out : P->ch[L])) : kNoGate;
  }
  for (int L = 0; L < 4; ++L) {
    P->gt_da[L] = P->bump(P->gate_bytes(L, P->ch[L]));
    P->gt_dc[L] = (!up && L > 0) ? P->bump(P->gate_bytes(L, P->dc_out[L])) : kNoGate;
  }
  for (int L = 0; L < 4; ++L) {
    // the coarse tensor below level L is c[4] (L == 3) or dc[L+1]; its gates are needed where the pooling is fused
    P->pool_fused[L] = up && P->dt == UB_DT_BF16 &&
                       conv3x3_dgrad_pool2_ok(N, P->Hl(L), P->Wl(L), P->ch[L], P->ch[L], P->ch[L]);
    if (!P->pool_fused[L]) continue;
    if (L == 3) {
      if (P->gt_c[4] == kNoGate) P->gt_c[4] = P->bump(P->gate_bytes(4, P->c4_out));
    } else if (P->gt_dc[L + 1] == kNoGate) {
      P->gt_dc[L + 1] = P->bump(P->gate_bytes(L + 1, P->dc_out[L + 1]));
    }
  }
  for (int L = 0; L < 5; ++L) {
    P->drop_bits[L] = ((L == 2 || L == 3) && dropout > 0.0f) ? P->bump(P->gate_bytes(L, P->ch[L])) : kNoGate;
  }
  const long long pixels = static_cast<long long>(N) * H * W;
  P->preds = P->bump(static_cast<size_t>(pixels) * n_cl_out * 4);
  P->sums = P->bump(64);
  P->metrics = P->bump(64);
  // ---- workspace: the largest any single op asks for (ops are serialised on one stream)
  // main-stream workspace: the head and conv1a's weight gradient (the last kernel of the dgrad chain; it runs on
  // the main stream so that it overlaps conv1b's weight gradient on the side stream)
  P->ws_bytes = max_sz(unetb200_head_ws_bytes(pixels, 32, n_cl_out), unetb200_conv_first_wgrad_ws_bytes(in_channels, 32));
  P->ws = P->bump(P->ws_bytes);
  size_t ws = 0;
  for (int L = 0; L < 5; ++L) {
    ws = max_sz(ws, unetb200_bias_grad_ws_bytes(P->ch[L]));
    const auto& la = P->layers[P->enc_a[L]];
    const auto& lb = P->layers[P->enc_b[L]];
    if (L > 0) ws = max_sz(ws, conv3x3_wgrad_ws_bytes(N, P->Hl(L), P->Wl(L), la.cin, la.cout, true));
    ws = max_sz(ws, conv3x3_wgrad_ws_bytes(N, P->Hl(L), P->Wl(L), lb.cin, lb.cout, true));
  }
  for (int L = 0; L < 4; ++L) {
    const auto& la = P->layers[P->dec_a[L]];
    const auto& lb = P->layers[P->dec_b[L]];
    ws = max_sz(ws, conv3x3_wgrad_ws_bytes(N, P->Hl(L), P->Wl(L), la.cin, la.cout, true));
    ws = max_sz(ws, conv3x3_wgrad_ws_bytes(N, P->Hl(L), P->Wl(L), lb.cin, lb.cout, true));
    if (!up) {
      const auto& lt = P->layers[P->tc[L]];
      ws = max_sz(ws, tconv2x2_wgrad_ws_bytes(N, P->Hl(L + 1), P->Wl(L + 1), lt.cin, lt.cout));
    }
  }
  P->ws_side_bytes = ws;
  P->ws_side = P->bump(ws);
  // ---- one table for all weight repacks (uploaded at bind time, walked by a single kernel)
  int tile = 0;
  for (const auto& l : P->layers) {
    if (l.kind == 2 || l.cin < 32) continue;
    PackLayer pl;
    pl.src = l.w_off;
    pl.dst_fwd = static_cast<long long>(l.pk_fwd);
    pl.dst_dgrad = static_cast<long long>(l.pk_dgrad);
    pl.kind = l.kind;
    pl.taps = l.kind == 0 ? 9 : 4;
    pl.cin = l.cin; pl.cout = l.cout;
    pl.tile_start = tile;
    P->pack_layers.push_back(pl);
    tile += pl.taps * (l.cin / 32) * (l.cout / 32);
  }
  P->pack_tiles = tile;
  P->pack_table = P->bump(P->pack_layers.size() * sizeof(PackLayer));
  *out = P;
  return UB_OK;
}

void unetb200_plan_destroy(unetb200_plan* P) {
  if (!P) return;
  for (int i = 0; i < P->n_ev; ++i) cudaEventDestroy(P->ev[i]);
  if (P->side) cudaStreamDestroy(P->side);
  delete P;
}

long long unetb200_plan_num_params(const unetb200_plan* P) { return P->nparams; }
int unetb200_plan_num_variables(const unetb200_plan* P) { return static_cast<int>(P->vars.size()); }

int unetb200_plan_variable(const unetb200_plan* P, int i, char* name, int name_cap, long long* offset, int* rank,
                           int* shape4) {
  UB_REQUIRE(i >= 0 && i < static_cast<int>(P->vars.size()), "plan_variable: index %d out of range", i);
  const Var& v = P->vars[i];
  if (name && name_cap > 0) {
    std::strncpy(name, v.name.c_str(), name_cap - 1);
    name[name_cap - 1] = 0;
  }
  if (offset) *offset = v.offset;
  if (rank) *rank = v.rank;
  if (shape4) for (int k = 0; k < 4; ++k) shape4[k] = v.shape[k];
  return UB_OK;
}

size_t unetb200_plan_arena_bytes(const unetb200_plan* P) { return P->arena_bytes; }

int unetb200_plan_num_regions(const unetb200_plan* P) { return static_cast<int>(P->regions.size()); }
int unetb200_plan_region(const unetb200_plan* P, int i, long long* offset, long long* bytes) {
  UB_REQUIRE(i >= 0 && i < static_cast<int>(P->regions.size()), "plan_region: index %d out of range", i);
  *offset = static_cast<long long>(P->regions[i].first);
  *bytes = static_cast<long long>(P->regions[i].second);
  return UB_OK;
}

int unetb200_plan_bind(unetb200_plan* P, void* arena, float* params, float* grads) {
  UB_REQUIRE(arena && params && grads, "plan_bind: null buffer");
  UB_REQUIRE(reinterpret_cast<uintptr_t>(arena) % kAlign == 0, "plan_bind: arena must be %zu-byte aligned", kAlign);
  UB_REQUIRE(reinterpret_cast<uintptr_t>(params) % 16 == 0 && reinterpret_cast<uintptr_t>(grads) % 16 == 0,
             "plan_bind: parameter buffers must be 16-byte aligned");
  P->arena = reinterpret_cast<uint8_t*>(arena);
  P->params = params;
  P->grads = grads;
  if (!P->side) {   // created here, never inside a stream capture
    P->use_side = P->use_side && getenv("UNETB200_NO_SIDE_STREAM") == nullptr;
    // The weight-gradient stream runs at the highest priority: its kernels are what the end-of-step join waits for, so
    // they should never queue behind the dgrad chain (measured with the update stream also high: 2.59 -> 2.57 ms/step,
    // repeatable on one box). UNETB200_SIDE_PRIORITY=low|same overrides.
    int least = 0, greatest = 0;
    UB_CHECK_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    int prio = greatest;
    if (const char* sp = getenv("UNETB200_SIDE_PRIORITY")) prio = sp[0] == 'l' ? least : (sp[0] == 'h' ? greatest : 0);
    UB_CHECK_CUDA(cudaStreamCreateWithPriority(&P->side, cudaStreamNonBlocking, prio));
    for (P->n_ev = 0; P->n_ev < 40; ++P->n_ev) {
      UB_CHECK_CUDA(cudaEventCreateWithFlags(&P->ev[P->n_ev], cudaEventDisableTiming));
    }
  }
  // the repack table lives in the arena; binding is a set-up call, so a synchronous copy is fine here
  UB_CHECK_CUDA(cudaMemcpy(P->at(P->pack_table), P->pack_layers.data(), P->pack_layers.size() * sizeof(PackLayer),
                           cudaMemcpyHostToDevice));
  return UB_OK;
}

// Arena offset of a named tensor (debugging / per-layer parity tests). Names: a1..a5, c1..c5, p1..p4, u6..u9,
// a6..a9, c6..c9 and the same with a leading 'g' for gradients; 'preds', 'metrics', 'sums'.
int unetb200_plan_tensor(const unetb200_plan* P, const char* name, long long* offset, int* n, int* h, int* w, int* c,
                         int* elem_bytes) {
  std::string s(name);
  bool grad = false;
  if (s == "preds") {
    *offset = static_cast<long long>(P->preds); *n = P->N; *h = P->H; *w = P->W; *c = P->ncl; *elem_bytes = 4;
    return UB_OK;
  }
  if (s == "sums") {   // the three dice sums of the last forward: sum(y*p), sum(y), sum(p)
    *offset = static_cast<long long>(P->sums); *n = 1; *h = 1; *w = 1; *c = 3; *elem_bytes = 4;
    return UB_OK;
  }
  if (s == "metrics") {
    *offset = static_cast<long long>(P->metrics); *n = 1; *h = 1; *w = 1; *c = 4; *elem_bytes = 4;
    return UB_OK;
  }
  if (s.size() == 3 && s[0] == 'g') { grad = true; s = s.substr(1); }
  UB_REQUIRE(s.size() == 2 && s[1] >= '1' && s[1] <= '9', "plan_tensor: unknown tensor '%s'", name);
  const int idx = s[1] - '0';
  size_t off = 0;
  int L = 0, C = 0;
  if (idx <= 5) {
    L = idx - 1;
    if (s[0] == 'a') { off = grad ? P->g_a[L] : P->a[L]; C = P->ch[L]; }
    else if (s[0] == 'c') { off = grad ? P->g_c[L] : P->c[L]; C = L == 4 ? P->c4_out : P->ch[L]; }
    else if (s[0] == 'p' && L < 4) { off = grad ? P->g_p[L] : P->p[L]; C = P->ch[L]; L = L + 1; }
    else { UB_REQUIRE(false, "plan_tensor: unknown tensor '%s'", name); }
  } else {
    L = 9 - idx;
    if (s[0] == 'a') { off = grad ? P->g_da[L] : P->da[L]; C = P->ch[L]; }
    else if (s[0] == 'c') { off = grad ? P->g_dc[L] : P->dc[L]; C = P->dc_out[L]; }
    else if (s[0] == 'u') { off = grad ? P->g_u[L] : P->u[L]; C = P->ch[L]; }
    else { UB_REQUIRE(false, "plan_tensor: unknown tensor '%s'", name); }
  }
  *offset = static_cast<long long>(off); *n = P->N; *h = P->Hl(L); *w = P->Wl(L); *c = C; *elem_bytes = P->elt;
  return UB_OK;
}

#define RC(expr)            \
  do {                      \
    int _rc = (expr);       \
    if (_rc) return _rc;    \
  } while (0)

// Refresh the bf16 GEMM operands from the fp32 master weights (after init and after every optimizer update).
int unetb200_plan_pack_weights(unetb200_plan* P, void* stream) {
  UB_REQUIRE(P->arena, "plan_pack_weights: plan is not bound");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  return pack_all(reinterpret_cast<const PackLayer*>(P->at(P->pack_table)), static_cast<int>(P->pack_layers.size()),
                  P->pack_tiles, P->params, P->arena, s, 0, P->dt);
}

// Forward pass (model.py:45-120) + dice metrics (model.py:4-21,133-143) when msks != NULL.
// imgs: fp32 [N,H,W,Cin]; msks: fp32 [N,H,W,n_cl]. Dropout masks come from Philox stream (seed + *seed_add, layer).
int unetb200_plan_forward(unetb200_plan* P, const float* imgs, const float* msks, uint64_t seed,
                          const long long* seed_add, void* stream) {
  UB_REQUIRE(P->arena, "plan_forward: plan is not bound");
  NvtxRange nvtx_range("unetb200.plan_forward");
  PdlScope pdl_scope;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int N = P->N;
  const int dt = P->dt;
  const DropoutCfg no_drop{1.0f, 0, 0, nullptr};
  // dropout keep bits of conv3a / conv4a for this forward: two small kernels instead of Philox in the conv epilogues
  // (where it made conv3a's epilogue slower than its MMAs). They run on the plan's second stream beside conv1a..conv2b
  // (12 us each, Philox-bound); the main stream waits for them right before conv3a.
  static const bool inline_philox = getenv("UNETB200_INLINE_PHILOX") != nullptr;
  const bool bits_on_side = P->use_side && P->side != nullptr;
  cudaEvent_t bits_done = nullptr;
  bool any_bits = false;
  for (int L = 2; L <= 3 && !inline_philox; ++L) any_bits = any_bits || P->drop_bits[L] != kNoGate;
  if (any_bits) {
    cudaStream_t bs = s;
    if (bits_on_side) {
      cudaEvent_t e = P->ev[P->ev_next];
      P->ev_next = (P->ev_next + 1) % P->n_ev;
      UB_CHECK_CUDA(cudaEventRecord(e, s));
      UB_CHECK_CUDA(cudaStreamWaitEvent(P->side, e, 0));   // after whatever last read the bits (previous backward)
      bs = P->side;
    }
    for (int L = 2; L <= 3; ++L) {
      if (P->drop_bits[L] == kNoGate) continue;
      RC(dropout_keep_bits16(P->gate(P->drop_bits[L]), static_cast<long long>(N) * P->Hl(L) * P->Wl(L) * P->ch[L],
                             1.0f - P->dropout, seed, seed_add, static_cast<unsigned>(L), bs));
    }
    if (bits_on_side) {
      bits_done = P->ev[P->ev_next];
      P->ev_next = (P->ev_next + 1) % P->n_ev;
      UB_CHECK_CUDA(cudaEventRecord(bits_done, P->side));
    }
  }
  // encoder
  for (int L = 0; L < 5; ++L) {
    const int H = P->Hl(L), W = P->Wl(L);
    const int la = P->enc_a[L], lb = P->enc_b[L];
    if (L == 0) {
      RC(conv_first_fwd(imgs, P->W_(la), P->B_(la), P->at(P->a[0]), N, H, W, P->Cin, 32, s, P->gate(P->gt_a[0]), dt));
    } else {
      DropoutCfg d = no_drop;
      if (L == 2 && bits_done) UB_CHECK_CUDA(cudaStreamWaitEvent(s, bits_done, 0));   // both layers' bits are ready
      if ((L == 2 || L == 3) && P->dropout > 0.0f) {
        d = DropoutCfg{1.0f - P->dropout, seed, static_cast<unsigned>(L), seed_add,
                       inline_philox ? nullptr : P->gate(P->drop_bits[L])};
      }
      RC(conv3x3_fwd(ConvSrc{P->at(P->p[L - 1]), P->ch[L - 1]}, ConvSrc{nullptr, 0}, P->pkf(la), P->B_(la),
                     P->at(P->a[L]), N, H, W, P->ch[L], 1, d, s, P->gate(P->gt_a[L]), nullptr, dt));
    }
    // MaxPool2D comes out of conv{L+1}b's epilogue where the row-pair kernel runs (128x128 and 64x64 levels)
    const bool fused_pool = L < 4 && dt == UB_DT_BF16 && conv3x3_fwd_maxpool_ok(N, H, W, P->ch[L], 0, P->layers[lb].cout);
    RC(conv3x3_fwd(ConvSrc{P->at(P->a[L]), P->ch[L]}, ConvSrc{nullptr, 0}, P->pkf(lb), P->B_(lb), P->at(P->c[L]), N, H,
                   W, P->layers[lb].cout, 1, no_drop, s, P->gate(P->gt_c[L]), fused_pool ? P->at(P->p[L]) : nullptr, dt));
    if (L < 4 && !fused_pool) RC(maxpool2x2_fwd(P->at(P->c[L]), P->at(P->p[L]), N, H, W, P->ch[L], s, dt));
  }
  // decoder
  for (int L = 3; L >= 0; --L) {
    const int H = P->Hl(L), W = P->Wl(L);
    const size_t below = (L == 3) ? P->c[4] : P->dc[L + 1];
    const int below_C = (L == 3) ? P->c4_out : P->dc_out[L + 1];
    if (P->ups) {
      RC(upsample2x2_fwd(P->at(below), P->at(P->u[L]), N, H / 2, W / 2, below_C, s, dt));
    } else {
      const int lt = P->tc[L];
      RC(tconv2x2_fwd(P->at(below), below_C, P->pkf(lt), P->B_(lt), P->at(P->u[L]), P->ch[L], N, H / 2, W / 2, s, dt));
    }
    const int la = P->dec_a[L], lb = P->dec_b[L];
    RC(conv3x3_fwd(ConvSrc{P->at(P->u[L]), P->ch[L]}, ConvSrc{P->at(P->c[L]), P->ch[L]}, P->pkf(la), P->B_(la),
                   P->at(P->da[L]), N, H, W, P->ch[L], 1, no_drop, s, P->gate(P->gt_da[L]), nullptr, dt));
    RC(conv3x3_fwd(ConvSrc{P->at(P->da[L]), P->ch[L]}, ConvSrc{nullptr, 0}, P->pkf(lb), P->B_(lb), P->at(P->dc[L]), N, H,
                   W, P->dc_out[L], 1, no_drop, s, P->gate(P->gt_dc[L]), nullptr, dt));
  }
  // head
  const long long pixels = static_cast<long long>(N) * P->H * P->W;
  float* preds = reinterpret_cast<float*>(P->at(P->preds));
  float* sums = reinterpret_cast<float*>(P->at(P->sums));
  if (msks) {
    RC(head_fwd(P->at(P->dc[0]), P->W_(P->mask), P->B_(P->mask), msks, preds, sums,
                reinterpret_cast<float*>(P->at(P->ws)), pixels, 32, P->ncl, s, dt));
    RC(dice_finalize(sums, 1.0f, reinterpret_cast<float*>(P->at(P->metrics)), s));
  } else {
    // inference: no labels, the dice sums degenerate to {0, 0, sum p}
    RC(head_fwd(P->at(P->dc[0]), P->W_(P->mask), P->B_(P->mask), nullptr, preds, sums,
                reinterpret_cast<float*>(P->at(P->ws)), pixels, 32, P->ncl, s, dt));
  }
  return UB_OK;
}

// Backward segments (the autodiff graph optimizer.compute_gradients builds at test_dist.py:246-248), split so the
// caller can start the all-reduce (and then the update) of a gradient bucket as soon as the segment that completes it
// has been enqueued. Backward walks the flat buffer from its end to its start, so every bucket is one contiguous range:
//   segment 0: head, conv9b/a, [transConv9], conv8b/a, [transConv8]   (tail of the flat buffer, 0.7 MB)
//   segment 1: conv7b/a, [transConv7], conv6b/a, [transConv6]        (11.5 MB; 7.4 MB with UpSampling2D)
//   segment 2: conv5b, conv5a                                          (the two largest tensors)
//   segment 3: conv4b/a, conv3b/a
//   segment 4: conv2b/a, conv1b/a     (0.26 MB: the only all-reduce that cannot hide behind backward work)
int unetb200_plan_backward(unetb200_plan* P, const float* imgs, const float* msks, const float* dpreds,
                           float grad_scale, int segment, void* stream) {
  UB_REQUIRE(P->arena, "plan_backward: plan is not bound");
  UB_REQUIRE(segment >= 0 && segment < kSegments, "plan_backward: segment %d outside 0..%d", segment, kSegments - 1);
  static const char* const kSegNames[kSegments] = {"unetb200.plan_backward.seg0_head_conv9_conv8",
                                                   "unetb200.plan_backward.seg1_conv7_conv6",
                                                   "unetb200.plan_backward.seg2_conv5",
                                                   "unetb200.plan_backward.seg3_conv4_conv3",
                                                   "unetb200.plan_backward.seg4_conv2_conv1"};
  NvtxRange nvtx_range(kSegNames[segment]);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int N = P->N;
  const int dt = P->dt;
  float* ws = reinterpret_cast<float*>(P->at(P->ws));              // main stream: head
  float* wss = reinterpret_cast<float*>(P->at(P->ws_side));         // weight-gradient stream
  void* wsv = P->at(P->ws_side);
  const size_t wsb = P->ws_side_bytes;
  const float inv_keep = P->dropout > 0.0f ? 1.0f / (1.0f - P->dropout) : 1.0f;
  const bool side_on = P->use_side && P->side != nullptr;
  cudaStream_t w = side_on ? P->side : s;   // where the weight gradients go
  // fork: the side stream waits until the main stream has produced everything enqueued so far
  auto fork = [&]() -> int {
    if (!side_on) return UB_OK;
    cudaEvent_t e = P->ev[P->ev_next];
    P->ev_next = (P->ev_next + 1) % P->n_ev;
    UB_CHECK_CUDA(cudaEventRecord(e, s));
    UB_CHECK_CUDA(cudaStreamWaitEvent(w, e, 0));
    return UB_OK;
  };
  // join: the main stream waits for all weight gradients enqueued so far (end of every segment)
  auto join = [&]() -> int {
    if (!side_on || P->defer_join) return UB_OK;
    cudaEvent_t e = P->ev[P->ev_next];
    P->ev_next = (P->ev_next + 1) % P->n_ev;
    UB_CHECK_CUDA(cudaEventRecord(e, w));
    UB_CHECK_CUDA(cudaStreamWaitEvent(s, e, 0));
    return UB_OK;
  };

  if (segment <= 1) {
    if (segment == 0) {
    const long long pixels = static_cast<long long>(N) * P->H * P->W;
    UB_REQUIRE(msks || dpreds, "plan_backward: need the labels or an explicit dL/dpreds");
    RC(head_bwd(P->at(P->dc[0]), P->W_(P->mask), dpreds ? dpreds : msks, reinterpret_cast<const float*>(P->at(P->preds)),
                dpreds ? nullptr : reinterpret_cast<const float*>(P->at(P->sums)), 1.0f, grad_scale, P->at(P->g_dc[0]), P->dW_(P->mask),
                P->dB_(P->mask), ws, pixels, 32, P->ncl, s, dt));
    }
    for (int L = 2 * segment; L < 2 * segment + 2; ++L) {
      const int H = P->Hl(L), W = P->Wl(L);
      const long long px = static_cast<long long>(N) * H * W;
      const int la = P->dec_a[L], lb = P->dec_b[L];
      // conv{9-L}b
      RC(fork());
      RC(fork());
    RC(conv3x3_wgrad(P->at(P->g_dc[L]), P->dc_out[L], ConvSrc{P->at(P->da[L]), P->ch[L]}, ConvSrc{nullptr, 0},
                       P->dW_(lb), wsv, wsb, N, H, W, w, P->dB_(lb), dt));
      RC(conv3x3_dgrad(P->at(P->g_dc[L]), P->dc_out[L], P->pkd(lb),
                       DgradDst{P->at(P->g_da[L]), P->ch[L], P->at(P->da[L]), 1.0f, P->gate(P->gt_da[L])},
                       DgradDst{nullptr, 0, nullptr, 1.0f},
                       N, H, W, s, dt));
      // conv{9-L}a over [u, skip]
      RC(fork());
      RC(fork());
    RC(conv3x3_wgrad(P->at(P->g_da[L]), P->ch[L], ConvSrc{P->at(P->u[L]), P->ch[L]}, ConvSrc{P->at(P->c[L]), P->ch[L]},
                       P->dW_(la), wsv, wsb, N, H, W, w, P->dB_(la), dt));
      // up path
      const size_t below = (L == 3) ? P->c[4] : P->dc[L + 1];
      const size_t g_below = (L == 3) ? P->g_c[4] : P->g_dc[L + 1];
      const int below_C = (L == 3) ? P->c4_out : P->dc_out[L + 1];
      DgradDst d_up{P->at(P->g_u[L]), P->ch[L], nullptr, 1.0f};
      if (P->pool_fused[L]) {   // 2x2 sums + ReluGrad of the coarse tensor straight from the dgrad epilogue
        d_up = DgradDst{P->at(g_below), P->ch[L], nullptr, 1.0f, P->gate(L == 3 ? P->gt_c[4] : P->gt_dc[L + 1]), 1};
      }
      RC(conv3x3_dgrad(P->at(P->g_da[L]), P->ch[L], P->pkd(la), d_up,
                       DgradDst{P->at(P->g_c[L]), P->ch[L], P->at(P->c[L]), 1.0f, P->gate(P->gt_c[L])}, N, H, W, s, dt));
      if (P->pool_fused[L]) {
        // g_below is complete
      } else if (P->ups) {
        RC(upsample2x2_bwd(P->at(P->g_u[L]), P->at(below), P->at(g_below), N, H / 2, W / 2, below_C, 1, s, dt));
      } else {
        const int lt = P->tc[L];
        RC(fork());
        // the bias gradient (column sums of dY) comes out of the same kernel: it stages every dY tile anyway
        static const bool separate_bias = getenv("UNETB200_TCONV_SEPARATE_BIAS") != nullptr;
        RC(tconv2x2_wgrad(P->at(P->g_u[L]), P->ch[L], P->at(below), below_C, P->dW_(lt), wsv, wsb, N, H / 2, W / 2, w, dt,
                          separate_bias ? nullptr : P->dB_(lt)));
        if (separate_bias) RC(colsum_act(P->at(P->g_u[L]), P->dB_(lt), wss, px, P->ch[L], w, dt));
        RC(tconv2x2_dgrad(P->at(P->g_u[L]), P->ch[L], P->pkd(lt), P->at(g_below), below_C, P->at(below), 1.0f, N, H / 2,
                          W / 2, s, P->gate(L == 3 ? P->gt_c[4] : P->gt_dc[L + 1]), dt));
      }
    }
    RC(join());
    return UB_OK;
  }

  if (segment == 2) {
    const int L = 4;
    const int H = P->Hl(L), W = P->Wl(L);
    const int la = P->enc_a[L], lb = P->enc_b[L];
    RC(fork());
    RC(conv3x3_wgrad(P->at(P->g_c[L]), P->c4_out, ConvSrc{P->at(P->a[L]), P->ch[L]}, ConvSrc{nullptr, 0}, P->dW_(lb), wsv,
                     wsb, N, H, W, w, P->dB_(lb), dt));
    RC(conv3x3_dgrad(P->at(P->g_c[L]), P->c4_out, P->pkd(lb), DgradDst{P->at(P->g_a[L]), P->ch[L], P->at(P->a[L]), 1.0f, P->gate(P->gt_a[L])},
                     DgradDst{nullptr, 0, nullptr, 1.0f}, N, H, W, s, dt));
    RC(fork());
    RC(conv3x3_wgrad(P->at(P->g_a[L]), P->ch[L], ConvSrc{P->at(P->p[L - 1]), P->ch[L - 1]}, ConvSrc{nullptr, 0},
                     P->dW_(la), wsv, wsb, N, H, W, w, P->dB_(la), dt));
    RC(conv3x3_dgrad(P->at(P->g_a[L]), P->ch[L], P->pkd(la), DgradDst{P->at(P->g_p[L - 1]), P->ch[L - 1], nullptr, 1.0f},
                     DgradDst{nullptr, 0, nullptr, 1.0f}, N, H, W, s, dt));
    RC(join());
    return UB_OK;
  }

  for (int L = (segment == 3 ? 3 : 1); L >= (segment == 3 ? 2 : 0); --L) {
    const int H = P->Hl(L), W = P->Wl(L);
    const int la = P->enc_a[L], lb = P->enc_b[L];
    // MaxPoolGrad + skip gradient + ReluGrad of conv{L+1}b, in place in g_c[L]
    RC(maxpool2x2_bwd(P->at(P->c[L]), P->at(P->g_p[L]), P->at(P->g_c[L]), P->at(P->g_c[L]), N, H, W, P->ch[L], s, dt));
    RC(fork());
    RC(conv3x3_wgrad(P->at(P->g_c[L]), P->ch[L], ConvSrc{P->at(P->a[L]), P->ch[L]}, ConvSrc{nullptr, 0}, P->dW_(lb), wsv,
                     wsb, N, H, W, w, P->dB_(lb), dt));
    const bool dropped = (L == 2 || L == 3) && P->dropout > 0.0f;
    RC(conv3x3_dgrad(P->at(P->g_c[L]), P->ch[L], P->pkd(lb),
                     DgradDst{P->at(P->g_a[L]), P->ch[L], P->at(P->a[L]), dropped ? inv_keep : 1.0f, P->gate(P->gt_a[L])},
                     DgradDst{nullptr, 0, nullptr, 1.0f}, N, H, W, s, dt));
    if (L > 0) {
      RC(fork());
      RC(fork());
    RC(conv3x3_wgrad(P->at(P->g_a[L]), P->ch[L], ConvSrc{P->at(P->p[L - 1]), P->ch[L - 1]}, ConvSrc{nullptr, 0},
                       P->dW_(la), wsv, wsb, N, H, W, w, P->dB_(la), dt));
      RC(conv3x3_dgrad(P->at(P->g_a[L]), P->ch[L], P->pkd(la),
                       DgradDst{P->at(P->g_p[L - 1]), P->ch[L - 1], nullptr, 1.0f}, DgradDst{nullptr, 0, nullptr, 1.0f},
                       N, H, W, s, dt));
    } else {
      RC(conv_first_wgrad(imgs, P->at(P->g_a[0]), P->dW_(la), P->dB_(la), ws, N, H, W, P->Cin, 32, s, dt));
    }
  }
  RC(join());
  return UB_OK;
}

// Make `stream` wait for every weight gradient enqueued so far (see defer_join).
int unetb200_plan_join(unetb200_plan* P, void* stream) {
  if (!(P->use_side && P->side != nullptr)) return UB_OK;
  cudaEvent_t e = P->ev[P->ev_next];
  P->ev_next = (P->ev_next + 1) % P->n_ev;
  UB_CHECK_CUDA(cudaEventRecord(e, P->side));
  UB_CHECK_CUDA(cudaStreamWaitEvent(reinterpret_cast<cudaStream_t>(stream), e, 0));
  return UB_OK;
}

int unetb200_plan_defer_join(unetb200_plan* P, int enable) {
  const int old = P->defer_join ? 1 : 0;
  P->defer_join = enable != 0;
  return old;
}

int unetb200_plan_side_stream(unetb200_plan* P, int enable) {
  const int old = P->use_side ? 1 : 0;
  P->use_side = enable != 0 && getenv("UNETB200_NO_SIDE_STREAM") == nullptr;
  return old;
}

// Flat-buffer range [begin, end) (floats) whose gradients are complete once `segment` has run.
int unetb200_plan_bucket(const unetb200_plan* P, int segment, long long* begin, long long* end) {
  UB_REQUIRE(segment >= 0 && segment < kSegments, "plan_bucket: segment %d outside 0..%d", segment, kSegments - 1);
  // first variable of decoder level L (flat order inside a level: [transConv], conv_a, conv_b)
  auto dec_begin = [&](int L) { return P->ups ? P->layers[P->dec_a[L]].w_off : P->layers[P->tc[L]].w_off; };
  const long long cut[kSegments + 1] = {P->nparams, dec_begin(1), dec_begin(3), P->layers[P->enc_a[4]].w_off,
                                        P->layers[P->enc_a[2]].w_off, 0};
  *begin = cut[segment + 1];
  *end = cut[segment];
  return UB_OK;
}

int unetb200_plan_num_segments(const unetb200_plan*) { return kSegments; }

// Refresh the bf16 GEMM operands of the variables inside [begin, end) of the flat buffer only (the update of one
// gradient bucket, applied while the rest of backward still runs).
int unetb200_plan_pack_range(unetb200_plan* P, long long begin, long long end, void* stream) {
  UB_REQUIRE(P->arena, "plan_pack_range: plan is not bound");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int first = -1, last = -1;   // pack_layers is in flat order
  for (int i = 0; i < static_cast<int>(P->pack_layers.size()); ++i) {
    const long long off = P->pack_layers[i].src;
    if (off >= begin && off < end) {
      if (first < 0) first = i;
      last = i;
    }
  }
  if (first < 0) return UB_OK;
  const int tile0 = P->pack_layers[first].tile_start;
  const int tile1 = last + 1 < static_cast<int>(P->pack_layers.size()) ? P->pack_layers[last + 1].tile_start : P->pack_tiles;
  return pack_all(reinterpret_cast<const PackLayer*>(P->at(P->pack_table)) + first, last - first + 1, tile1 - tile0,
                  P->params, P->arena, s, tile0, P->dt);
}

}  // extern "C"
