// Stand-alone numerics check of libunetb200 through its C ABI against plain CPU loops (no Python, fast start-up).
// Used on the GPU box for first-light debugging; the pytest suite is the parity gate proper.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/unetb200.h"

static int g_fail = 0, g_total = 0;

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e = (x);                                                              \
    if (e != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__);  \
      exit(2);                                                                        \
    }                                                                                 \
  } while (0)

static uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  uint32_t r = u + 0x7FFFu + ((u >> 16) & 1u);
  return static_cast<uint16_t>(r >> 16);
}
static float bf2f(uint16_t h) {
  uint32_t u = static_cast<uint32_t>(h) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static float rbf(float f) { return bf2f(f2bf(f)); }

static uint32_t g_rng = 12345u;
static float urand() {
  g_rng = g_rng * 1664525u + 1013904223u;
  return ((g_rng >> 8) & 0xFFFFFF) / 16777216.0f;
}
static float srand1() { return 2.0f * urand() - 1.0f; }

template <class T>
struct DBuf {
  T* p = nullptr;
  size_t n = 0;
  explicit DBuf(size_t n_) : n(n_) { CK(cudaMalloc(&p, (n ? n : 1) * sizeof(T))); CK(cudaMemset(p, 0xCD, (n ? n : 1) * sizeof(T))); }
  ~DBuf() { cudaFree(p); }
  void up(const std::vector<T>& h) { CK(cudaMemcpy(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice)); }
  std::vector<T> down() {
    std::vector<T> h(n);
    CK(cudaMemcpy(h.data(), p, n * sizeof(T), cudaMemcpyDeviceToHost));
    return h;
  }
};

static std::vector<uint16_t> to_bf(const std::vector<float>& v) {
  std::vector<uint16_t> o(v.size());
  for (size_t i = 0; i < v.size(); ++i) o[i] = f2bf(v[i]);
  return o;
}
static std::vector<float> rounded(const std::vector<float>& v) {
  std::vector<float> o(v.size());
  for (size_t i = 0; i < v.size(); ++i) o[i] = rbf(v[i]);
  return o;
}
static std::vector<float> randv(size_t n, float scale, bool relu_like = false) {
  std::vector<float> v(n);
  for (auto& x : v) {
    x = srand1() * scale;
    if (relu_like && x < 0) x = 0;
  }
  return v;
}

static void report(const std::string& name, const std::vector<float>& got, const std::vector<float>& ref, float tol) {
  double maxref = 0, maxerr = 0, se = 0, sr = 0;
  size_t bad = 0, worst = 0;
  for (size_t i = 0; i < ref.size(); ++i) {
    double e = std::fabs(static_cast<double>(got[i]) - ref[i]);
    if (!(e == e)) { e = 1e30; }
    if (e > maxerr) { maxerr = e; worst = i; }
    maxref = std::fmax(maxref, std::fabs(ref[i]));
    se += e * e;
    sr += static_cast<double>(ref[i]) * ref[i];
  }
  const double rel = maxerr / (maxref + 1e-30);
  const double l2 = std::sqrt(se / (sr + 1e-30));
  for (size_t i = 0; i < ref.size(); ++i) {
    if (std::fabs(static_cast<double>(got[i]) - ref[i]) > tol * (maxref + 1e-30)) ++bad;
  }
  const bool ok = rel <= tol;
  ++g_total;
  if (!ok) ++g_fail;
  printf("%-58s %s max|err|/max|ref|=%.3e l2rel=%.3e maxref=%.3e bad=%zu/%zu worst@%zu got=%g ref=%g\n", name.c_str(),
         ok ? "PASS" : "FAIL", rel, l2, maxref, bad, ref.size(), worst, got[worst], ref[worst]);
  fflush(stdout);
}

static void expect_ok(int rc, const char* what) {
  if (rc != 0) {
    printf("%s returned %d: %s\n", what, rc, unetb200_last_error_string());
    ++g_fail;
    ++g_total;
  }
}
static bool sync_ok(const char* what) {
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("%s: kernel failed: %s\n", what, cudaGetErrorString(e));
    exit(3);
  }
  return true;
}

// ----------------------------------------------------------------------------------------------- CPU references
// x: [N,H,W,C] (already concatenated), w: HWIO [3][3][C][Co]
static void ref_conv3x3(const std::vector<float>& x, const std::vector<float>& w, const std::vector<float>& b,
                        std::vector<float>& y, int N, int H, int W, int C, int Co, bool relu) {
  y.assign(static_cast<size_t>(N) * H * W * Co, 0.f);
  std::vector<float> acc(Co);
  for (int n = 0; n < N; ++n)
    for (int h = 0; h < H; ++h)
      for (int ww = 0; ww < W; ++ww) {
        for (int o = 0; o < Co; ++o) acc[o] = b.empty() ? 0.f : b[o];
        for (int r = 0; r < 3; ++r) {
          int hh = h + r - 1;
          if (hh < 0 || hh >= H) continue;
          for (int s = 0; s < 3; ++s) {
            int w2 = ww + s - 1;
            if (w2 < 0 || w2 >= W) continue;
            const float* xp = &x[((static_cast<size_t>(n) * H + hh) * W + w2) * C];
            const float* wp = &w[static_cast<size_t>(r * 3 + s) * C * Co];
            for (int c = 0; c < C; ++c) {
              const float xv = xp[c];
              if (xv == 0.f) continue;
              const float* wr = wp + static_cast<size_t>(c) * Co;
              for (int o = 0; o < Co; ++o) acc[o] += xv * wr[o];
            }
          }
        }
        float* yp = &y[((static_cast<size_t>(n) * H + h) * W + ww) * Co];
        for (int o = 0; o < Co; ++o) yp[o] = relu ? std::fmax(acc[o], 0.f) : acc[o];
      }
}
// dx[n,h,w,c] = sum_{r,s,o} dz[n,h-r+1,w-s+1,o] w[r,s,c,o]
static void ref_conv3x3_dgrad(const std::vector<float>& dz, const std::vector<float>& w, std::vector<float>& dx, int N,
                              int H, int W, int C, int Co) {
  dx.assign(static_cast<size_t>(N) * H * W * C, 0.f);
  for (int n = 0; n < N; ++n)
    for (int h = 0; h < H; ++h)
      for (int ww = 0; ww < W; ++ww) {
        const float* zp = &dz[((static_cast<size_t>(n) * H + h) * W + ww) * Co];
        for (int r = 0; r < 3; ++r) {
          int hh = h + r - 1;
          if (hh < 0 || hh >= H) continue;
          for (int s = 0; s < 3; ++s) {
            int w2 = ww + s - 1;
            if (w2 < 0 || w2 >= W) continue;
            float* xp = &dx[((static_cast<size_t>(n) * H + hh) * W + w2) * C];
            const float* wp = &w[static_cast<size_t>(r * 3 + s) * C * Co];
            for (int c = 0; c < C; ++c) {
              const float* wr = wp + static_cast<size_t>(c) * Co;
              float a = 0.f;
              for (int o = 0; o < Co; ++o) a += zp[o] * wr[o];
              xp[c] += a;
            }
          }
        }
      }
}
static void ref_conv3x3_wgrad(const std::vector<float>& x, const std::vector<float>& dz, std::vector<float>& dw, int N,
                              int H, int W, int C, int Co) {
  dw.assign(static_cast<size_t>(9) * C * Co, 0.f);
  std::vector<double> acc(dw.size(), 0.0);
  for (int n = 0; n < N; ++n)
    for (int h = 0; h < H; ++h)
      for (int ww = 0; ww < W; ++ww) {
        const float* zp = &dz[((static_cast<size_t>(n) * H + h) * W + ww) * Co];
        for (int r = 0; r < 3; ++r) {
          int hh = h + r - 1;
          if (hh < 0 || hh >= H) continue;
          for (int s = 0; s < 3; ++s) {
            int w2 = ww + s - 1;
            if (w2 < 0 || w2 >= W) continue;
            const float* xp = &x[((static_cast<size_t>(n) * H + hh) * W + w2) * C];
            double* ap = &acc[static_cast<size_t>(r * 3 + s) * C * Co];
            for (int c = 0; c < C; ++c) {
              const float xv = xp[c];
              if (xv == 0.f) continue;
              for (int o = 0; o < Co; ++o) ap[static_cast<size_t>(c) * Co + o] += static_cast<double>(xv) * zp[o];
            }
          }
        }
      }
  for (size_t i = 0; i < dw.size(); ++i) dw[i] = static_cast<float>(acc[i]);
}

static std::vector<float> concat_c(const std::vector<float>& a, int Ca, const std::vector<float>& b, int Cb, size_t P) {
  std::vector<float> o(P * (Ca + Cb));
  for (size_t p = 0; p < P; ++p) {
    for (int c = 0; c < Ca; ++c) o[p * (Ca + Cb) + c] = a[p * Ca + c];
    for (int c = 0; c < Cb; ++c) o[p * (Ca + Cb) + Ca + c] = b[p * Cb + c];
  }
  return o;
}

static std::vector<float> bf_down(DBuf<uint16_t>& d) {
  auto h = d.down();
  std::vector<float> o(h.size());
  for (size_t i = 0; i < h.size(); ++i) o[i] = bf2f(h[i]);
  return o;
}

// ----------------------------------------------------------------------------------------------- conv3x3 cases
static void test_conv3x3(int N, int H, int W, int C0, int C1, int Co, bool with_dropout = false) {
  char tag[128];
  snprintf(tag, sizeof(tag), "N%d %dx%d C%d+%d->%d", N, H, W, C0, C1, Co);
  const size_t P = static_cast<size_t>(N) * H * W;
  const int C = C0 + C1;
  auto x0 = rounded(randv(P * C0, 1.0f, true));
  auto x1 = rounded(randv(P * C1, 1.0f, true));
  auto w = randv(static_cast<size_t>(9) * C * Co, 1.0f / std::sqrt(9.0f * C));
  auto wr = rounded(w);
  auto b = randv(Co, 0.1f);
  auto xc = C1 ? concat_c(x0, C0, x1, C1, P) : x0;

  DBuf<uint16_t> dx0(P * C0), dx1(P * (C1 ? C1 : 1)), dy(P * Co), dwp(static_cast<size_t>(9) * C * Co),
      dwpd(static_cast<size_t>(9) * C * Co);
  DBuf<float> dw(static_cast<size_t>(9) * C * Co), db(Co);
  dx0.up(to_bf(x0));
  if (C1) dx1.up(to_bf(x1));
  dw.up(w);
  db.up(b);
  expect_ok(unetb200_pack_conv3x3_fwd(dw.p, dwp.p, C, Co, nullptr), "pack_fwd");
  expect_ok(unetb200_pack_conv3x3_dgrad(dw.p, dwpd.p, C, Co, nullptr), "pack_dgrad");

  // forward
  expect_ok(unetb200_conv3x3_fwd(dx0.p, C0, C1 ? dx1.p : nullptr, C1, dwp.p, db.p, dy.p, N, H, W, Co, 1, 1.0f, 0, 0,
                                 nullptr),
            "conv3x3_fwd");
  sync_ok("conv3x3_fwd");
  std::vector<float> yref;
  ref_conv3x3(xc, wr, b, yref, N, H, W, C, Co, true);
  report(std::string("conv3x3_fwd   ") + tag, bf_down(dy), yref, 1.2e-2f);

  if (with_dropout) {
    DBuf<unsigned char> keep(P * Co);
    expect_ok(unetb200_dropout_mask(keep.p, static_cast<long long>(P * Co), 0.8f, 77, 3, nullptr), "dropout_mask");
    expect_ok(unetb200_conv3x3_fwd(dx0.p, C0, C1 ? dx1.p : nullptr, C1, dwp.p, db.p, dy.p, N, H, W, Co, 1, 0.8f, 77, 3,
                                   nullptr),
              "conv3x3_fwd(drop)");
    sync_ok("conv3x3_fwd(drop)");
    auto k = keep.down();
    std::vector<float> yd(yref.size());
    size_t kept = 0;
    for (size_t i = 0; i < yd.size(); ++i) {
      yd[i] = k[i] ? yref[i] / 0.8f : 0.f;
      kept += k[i];
    }
    printf("   dropout keep fraction %.4f\n", static_cast<double>(kept) / yd.size());
    report(std::string("conv3x3_fwd+drop ") + tag, bf_down(dy), yd, 1.2e-2f);
  }

  // dgrad (mask on source 1 only when two sources, on source 0 when single)
  auto dz = rounded(randv(P * Co, 1.0f));
  DBuf<uint16_t> ddz(P * Co), dgx0(P * C0), dgx1(P * (C1 ? C1 : 1));
  ddz.up(to_bf(dz));
  const float scale = 1.25f;
  if (C1) {
    expect_ok(unetb200_conv3x3_dgrad(ddz.p, Co, dwpd.p, dgx0.p, C0, nullptr, dgx1.p, C1, dx1.p, scale, N, H, W, nullptr),
              "conv3x3_dgrad");
  } else {
    expect_ok(unetb200_conv3x3_dgrad(ddz.p, Co, dwpd.p, dgx0.p, C0, dx0.p, nullptr, 0, nullptr, scale, N, H, W, nullptr),
              "conv3x3_dgrad");
  }
  sync_ok("conv3x3_dgrad");
  std::vector<float> dxref;
  ref_conv3x3_dgrad(dz, wr, dxref, N, H, W, C, Co);
  std::vector<float> r0(P * C0), r1(P * C1);
  for (size_t p = 0; p < P; ++p) {
    for (int c = 0; c < C0; ++c) {
      float v = dxref[p * C + c];
      if (!C1) v = x0[p * C0 + c] > 0 ? v * scale : 0.f;
      r0[p * C0 + c] = v;
    }
    for (int c = 0; c < C1; ++c) {
      float v = dxref[p * C + C0 + c];
      r1[p * C1 + c] = x1[p * C1 + c] > 0 ? v * scale : 0.f;
    }
  }
  report(std::string("conv3x3_dgrad0 ") + tag, bf_down(dgx0), r0, 1.2e-2f);
  if (C1) report(std::string("conv3x3_dgrad1 ") + tag, bf_down(dgx1), r1, 1.2e-2f);

  // wgrad
  const size_t wsb = unetb200_conv3x3_wgrad_ws_bytes(N, H, W, C, Co);
  DBuf<unsigned char> ws(wsb + 16);
  DBuf<float> dgw(static_cast<size_t>(9) * C * Co);
  expect_ok(unetb200_conv3x3_wgrad(ddz.p, Co, dx0.p, C0, C1 ? dx1.p : nullptr, C1, dgw.p, ws.p, wsb, N, H, W, nullptr),
            "conv3x3_wgrad");
  sync_ok("conv3x3_wgrad");
  std::vector<float> dwref;
  ref_conv3x3_wgrad(xc, dz, dwref, N, H, W, C, Co);
  report(std::string("conv3x3_wgrad ") + tag + " ws=" + std::to_string(wsb), dgw.down(), dwref, 2e-3f);

  // bias grad
  DBuf<unsigned char> bws(unetb200_bias_grad_ws_bytes(Co));
  DBuf<float> dgb(Co);
  expect_ok(unetb200_bias_grad(ddz.p, dgb.p, bws.p, static_cast<long long>(P), Co, nullptr), "bias_grad");
  sync_ok("bias_grad");
  std::vector<float> dbref(Co, 0.f);
  {
    std::vector<double> a(Co, 0.0);
    for (size_t p = 0; p < P; ++p)
      for (int o = 0; o < Co; ++o) a[o] += dz[p * Co + o];
    for (int o = 0; o < Co; ++o) dbref[o] = static_cast<float>(a[o]);
  }
  report(std::string("bias_grad     ") + tag, dgb.down(), dbref, 1e-3f);
}

// ----------------------------------------------------------------------------------------------- tconv cases
static void test_tconv(int N, int H, int W, int Ci, int Co) {
  char tag[128];
  snprintf(tag, sizeof(tag), "N%d %dx%d C%d->%d", N, H, W, Ci, Co);
  const size_t P = static_cast<size_t>(N) * H * W;
  auto x = rounded(randv(P * Ci, 1.0f, true));
  auto w = randv(static_cast<size_t>(4) * Co * Ci, 1.0f / std::sqrt(1.0f * Ci));  // [a][b][o][c]
  auto wr = rounded(w);
  auto b = randv(Co, 0.1f);
  DBuf<uint16_t> dx(P * Ci), dy(4 * P * Co), dwp(w.size()), dwpd(w.size());
  DBuf<float> dw(w.size()), db(Co);
  dx.up(to_bf(x));
  dw.up(w);
  db.up(b);
  expect_ok(unetb200_pack_tconv2x2_fwd(dw.p, dwp.p, Ci, Co, nullptr), "pack_tconv_fwd");
  expect_ok(unetb200_pack_tconv2x2_dgrad(dw.p, dwpd.p, Ci, Co, nullptr), "pack_tconv_dgrad");
  expect_ok(unetb200_tconv2x2_fwd(dx.p, Ci, dwp.p, db.p, dy.p, Co, N, H, W, nullptr), "tconv_fwd");
  sync_ok("tconv_fwd");
  std::vector<float> yref(4 * P * Co);
  for (int n = 0; n < N; ++n)
    for (int i = 0; i < H; ++i)
      for (int j = 0; j < W; ++j)
        for (int a = 0; a < 2; ++a)
          for (int bb = 0; bb < 2; ++bb)
            for (int o = 0; o < Co; ++o) {
              float acc = b[o];
              const float* xp = &x[((static_cast<size_t>(n) * H + i) * W + j) * Ci];
              const float* wp = &wr[(static_cast<size_t>(a * 2 + bb) * Co + o) * Ci];
              for (int c = 0; c < Ci; ++c) acc += xp[c] * wp[c];
              yref[((static_cast<size_t>(n) * 2 * H + 2 * i + a) * 2 * W + 2 * j + bb) * Co + o] = acc;
            }
  report(std::string("tconv_fwd     ") + tag, bf_down(dy), yref, 1.2e-2f);

  auto dyv = rounded(randv(4 * P * Co, 1.0f));
  DBuf<uint16_t> ddy(4 * P * Co), dgx(P * Ci);
  ddy.up(to_bf(dyv));
  expect_ok(unetb200_tconv2x2_dgrad(ddy.p, Co, dwpd.p, dgx.p, Ci, dx.p, 1.0f, N, H, W, nullptr), "tconv_dgrad");
  sync_ok("tconv_dgrad");
  std::vector<float> dxref(P * Ci, 0.f);
  std::vector<double> dwacc(w.size(), 0.0);
  for (int n = 0; n < N; ++n)
    for (int i = 0; i < H; ++i)
      for (int j = 0; j < W; ++j) {
        const size_t pin = (static_cast<size_t>(n) * H + i) * W + j;
        for (int a = 0; a < 2; ++a)
          for (int bb = 0; bb < 2; ++bb) {
            const float* gp = &dyv[((static_cast<size_t>(n) * 2 * H + 2 * i + a) * 2 * W + 2 * j + bb) * Co];
            for (int o = 0; o < Co; ++o) {
              const float* wp = &wr[(static_cast<size_t>(a * 2 + bb) * Co + o) * Ci];
              double* ap = &dwacc[(static_cast<size_t>(a * 2 + bb) * Co + o) * Ci];
              for (int c = 0; c < Ci; ++c) {
                dxref[pin * Ci + c] += gp[o] * wp[c];
                ap[c] += static_cast<double>(gp[o]) * x[pin * Ci + c];
              }
            }
          }
      }
  for (size_t i = 0; i < dxref.size(); ++i) dxref[i] = x[i] > 0 ? dxref[i] : 0.f;
  report(std::string("tconv_dgrad   ") + tag, bf_down(dgx), dxref, 1.2e-2f);

  const size_t wsb = unetb200_tconv2x2_wgrad_ws_bytes(N, H, W, Ci, Co);
  DBuf<unsigned char> ws(wsb + 16);
  DBuf<float> dgw(w.size());
  expect_ok(unetb200_tconv2x2_wgrad(ddy.p, Co, dx.p, Ci, dgw.p, ws.p, wsb, N, H, W, nullptr), "tconv_wgrad");
  sync_ok("tconv_wgrad");
  std::vector<float> dwref(w.size());
  for (size_t i = 0; i < w.size(); ++i) dwref[i] = static_cast<float>(dwacc[i]);
  report(std::string("tconv_wgrad   ") + tag + " ws=" + std::to_string(wsb), dgw.down(), dwref, 2e-3f);
}

// ----------------------------------------------------------------------------------------------- bandwidth kernels
static void test_pool_upsample(int N, int H, int W, int C) {
  char tag[128];
  snprintf(tag, sizeof(tag), "N%d %dx%d C%d", N, H, W, C);
  const size_t P = static_cast<size_t>(N) * H * W;
  auto x = rounded(randv(P * C, 1.0f, true));
  // force ties
  for (size_t i = 0; i < x.size(); i += 7) x[i] = 0.5f;
  DBuf<uint16_t> dx(P * C), dyp(P * C / 4);
  dx.up(to_bf(x));
  expect_ok(unetb200_maxpool2x2_fwd(dx.p, dyp.p, N, H, W, C, nullptr), "maxpool_fwd");
  sync_ok("maxpool_fwd");
  std::vector<float> yref(P * C / 4), dzref(P * C, 0.f);
  auto dp = rounded(randv(P * C / 4, 1.0f));
  auto skip = rounded(randv(P * C, 1.0f));
  for (int n = 0; n < N; ++n)
    for (int i = 0; i < H / 2; ++i)
      for (int j = 0; j < W / 2; ++j)
        for (int c = 0; c < C; ++c) {
          size_t idx[4];
          for (int k = 0; k < 4; ++k)
            idx[k] = ((static_cast<size_t>(n) * H + 2 * i + k / 2) * W + 2 * j + k % 2) * C + c;
          int arg = 0;
          for (int k = 1; k < 4; ++k)
            if (x[idx[k]] > x[idx[arg]]) arg = k;
          const size_t o = ((static_cast<size_t>(n) * (H / 2) + i) * (W / 2) + j) * C + c;
          yref[o] = x[idx[arg]];
          for (int k = 0; k < 4; ++k) dzref[idx[k]] = skip[idx[k]];
          if (x[idx[arg]] > 0) dzref[idx[arg]] += dp[o];
        }
  report(std::string("maxpool_fwd   ") + tag, bf_down(dyp), yref, 1e-6f);
  DBuf<uint16_t> ddp(P * C / 4), dsk(P * C);
  ddp.up(to_bf(dp));
  dsk.up(to_bf(skip));
  expect_ok(unetb200_maxpool2x2_bwd(dx.p, ddp.p, dsk.p, dsk.p, N, H, W, C, nullptr), "maxpool_bwd");
  sync_ok("maxpool_bwd");
  report(std::string("maxpool_bwd   ") + tag, bf_down(dsk), dzref, 8e-3f);

  // upsample on the pooled-size tensor
  const int h = H / 2, w = W / 2;
  DBuf<uint16_t> dup(P * C), ddx(P * C / 4);
  expect_ok(unetb200_upsample2x2_fwd(dyp.p, dup.p, N, h, w, C, nullptr), "upsample_fwd");
  sync_ok("upsample_fwd");
  std::vector<float> uref(P * C), ubref(P * C / 4, 0.f);
  for (int n = 0; n < N; ++n)
    for (int i = 0; i < H; ++i)
      for (int j = 0; j < W; ++j)
        for (int c = 0; c < C; ++c) {
          const size_t src = ((static_cast<size_t>(n) * h + i / 2) * w + j / 2) * C + c;
          const size_t dst = ((static_cast<size_t>(n) * H + i) * W + j) * C + c;
          uref[dst] = yref[src];
          ubref[src] += skip[dst];
        }
  for (size_t i = 0; i < ubref.size(); ++i) ubref[i] = yref[i] > 0 ? ubref[i] : 0.f;
  report(std::string("upsample_fwd  ") + tag, bf_down(dup), uref, 1e-6f);
  dsk.up(to_bf(skip));
  expect_ok(unetb200_upsample2x2_bwd(dsk.p, dyp.p, ddx.p, N, h, w, C, 1, nullptr), "upsample_bwd");
  sync_ok("upsample_bwd");
  report(std::string("upsample_bwd  ") + tag, bf_down(ddx), ubref, 8e-3f);
}

static void test_first_and_head(int N, int H, int W, int Cin, int n_cl) {
  char tag[128];
  snprintf(tag, sizeof(tag), "N%d %dx%d Cin%d ncl%d", N, H, W, Cin, n_cl);
  const size_t P = static_cast<size_t>(N) * H * W;
  const int Co = 32;
  auto x = randv(P * Cin, 1.0f);
  auto w = randv(static_cast<size_t>(9) * Cin * Co, 0.5f);
  auto b = randv(Co, 0.1f);
  DBuf<float> dx(P * Cin), dw(w.size()), db(Co);
  DBuf<uint16_t> dy(P * Co);
  dx.up(x); dw.up(w); db.up(b);
  expect_ok(unetb200_conv_first_fwd(dx.p, dw.p, db.p, dy.p, N, H, W, Cin, Co, nullptr), "conv_first_fwd");
  sync_ok("conv_first_fwd");
  std::vector<float> yref;
  ref_conv3x3(x, w, b, yref, N, H, W, Cin, Co, true);
  report(std::string("conv_first_fwd ") + tag, bf_down(dy), yref, 6e-3f);

  auto dz = rounded(randv(P * Co, 1.0f));
  DBuf<uint16_t> ddz(P * Co);
  ddz.up(to_bf(dz));
  DBuf<unsigned char> ws(unetb200_conv_first_wgrad_ws_bytes(Cin, Co));
  DBuf<float> dgw(w.size()), dgb(Co);
  expect_ok(unetb200_conv_first_wgrad(dx.p, ddz.p, dgw.p, dgb.p, ws.p, N, H, W, Cin, Co, nullptr), "conv_first_wgrad");
  sync_ok("conv_first_wgrad");
  std::vector<float> dwref;
  ref_conv3x3_wgrad(x, dz, dwref, N, H, W, Cin, Co);
  report(std::string("conv_first_wgrad ") + tag, dgw.down(), dwref, 1e-3f);
  std::vector<float> dbref(Co, 0.f);
  for (size_t p = 0; p < P; ++p)
    for (int o = 0; o < Co; ++o) dbref[o] += dz[p * Co + o];
  report(std::string("conv_first_bgrad ") + tag, dgb.down(), dbref, 1e-3f);

  // head on a relu-like 32-channel tensor
  auto hx = rounded(randv(P * Co, 1.0f, true));
  auto hw = randv(static_cast<size_t>(Co) * n_cl, 0.3f);
  auto hb = randv(n_cl, 0.1f);
  std::vector<float> yt(P * n_cl);
  for (auto& v : yt) v = urand() < 0.1f ? 1.f : 0.f;
  DBuf<uint16_t> dhx(P * Co), ddx(P * Co);
  DBuf<float> dhw(hw.size()), dhb(n_cl), dyt(yt.size()), dp(yt.size()), dsums(3), dout(4), dgw2(hw.size()), dgb2(n_cl);
  DBuf<unsigned char> hws(unetb200_head_ws_bytes(static_cast<long long>(P), Co, n_cl));
  dhx.up(to_bf(hx)); dhw.up(hw); dhb.up(hb); dyt.up(yt);
  expect_ok(unetb200_head_fwd(dhx.p, dhw.p, dhb.p, dyt.p, dp.p, dsums.p, hws.p, static_cast<long long>(P), Co, n_cl, nullptr),
            "head_fwd");
  expect_ok(unetb200_dice_finalize(dsums.p, 1.0f, dout.p, nullptr), "dice_finalize");
  sync_ok("head_fwd");
  std::vector<float> pref(yt.size());
  double I = 0, Sy = 0, Sp = 0;
  for (size_t p = 0; p < P; ++p)
    for (int k = 0; k < n_cl; ++k) {
      double z = hb[k];
      for (int c = 0; c < Co; ++c) z += static_cast<double>(hx[p * Co + c]) * hw[c * n_cl + k];
      const double pv = 1.0 / (1.0 + std::exp(-z));
      pref[p * n_cl + k] = static_cast<float>(pv);
      I += yt[p * n_cl + k] * pv; Sy += yt[p * n_cl + k]; Sp += pv;
    }
  report(std::string("head_fwd p    ") + tag, dp.down(), pref, 1e-5f);
  std::vector<float> sref = {static_cast<float>(I), static_cast<float>(Sy), static_cast<float>(Sp)};
  report(std::string("head_fwd sums ") + tag, dsums.down(), sref, 1e-5f);
  std::vector<float> oref = {static_cast<float>(-std::log(2 * I + 1) + std::log(Sy + Sp + 1)),
                             static_cast<float>((2 * I + 1) / (Sy + Sp + 1)), static_cast<float>((I + 1) / (Sy + 1)),
                             static_cast<float>((I + 1) / (Sp + 1))};
  report(std::string("dice_finalize ") + tag, dout.down(), oref, 1e-5f);

  expect_ok(unetb200_head_bwd(dhx.p, dhw.p, dyt.p, dp.p, dsums.p, 1.0f, 1.0f, ddx.p, dgw2.p, dgb2.p, hws.p,
                              static_cast<long long>(P), Co, n_cl, nullptr),
            "head_bwd");
  sync_ok("head_bwd");
  std::vector<float> dxref(P * Co, 0.f), dwr(hw.size(), 0.f), dbr(n_cl, 0.f);
  {
    std::vector<double> dwa(hw.size(), 0.0), dba(n_cl, 0.0);
    const double ca = -2.0 / (2 * I + 1), cb = 1.0 / (Sy + Sp + 1);
    for (size_t p = 0; p < P; ++p)
      for (int k = 0; k < n_cl; ++k) {
        const double pv = pref[p * n_cl + k];
        const double dl = (ca * yt[p * n_cl + k] + cb) * pv * (1 - pv);
        dba[k] += dl;
        for (int c = 0; c < Co; ++c) {
          dwa[c * n_cl + k] += hx[p * Co + c] * dl;
          if (hx[p * Co + c] > 0) dxref[p * Co + c] += static_cast<float>(dl * hw[c * n_cl + k]);
        }
      }
    for (size_t i = 0; i < dwr.size(); ++i) dwr[i] = static_cast<float>(dwa[i]);
    for (int k = 0; k < n_cl; ++k) dbr[k] = static_cast<float>(dba[k]);
  }
  report(std::string("head_bwd dx   ") + tag, bf_down(ddx), dxref, 8e-3f);
  report(std::string("head_bwd dw   ") + tag, dgw2.down(), dwr, 1e-4f);
  report(std::string("head_bwd db   ") + tag, dgb2.down(), dbr, 1e-4f);
}

static void test_adam() {
  const long long n = 100003;
  auto th = randv(n, 1.0f), g = randv(n, 0.01f);
  std::vector<float> m(n, 0.f), v(n, 0.f);
  DBuf<float> dth(n + 1), dm(n + 1), dv(n + 1), dg(n + 1);
  DBuf<long long> dstep(1);
  std::vector<long long> st = {0};
  dth.up(th); dm.up(m); dv.up(v); dg.up(g); dstep.up(st);
  std::vector<double> rt(th.begin(), th.end()), rm(n, 0.0), rv(n, 0.0);
  const double lr0 = 5e-4, frac = 0.25, decay = 150, b1 = 0.9, b2 = 0.999, eps = 1e-8, gs = 0.5;
  for (int step = 0; step < 3; ++step) {
    expect_ok(unetb200_adam_step(dth.p, dm.p, dv.p, dg.p, n, dstep.p, 5e-4f, 0, 0.25f, 150.f, 0.9f, 0.999f, 1e-8f, 0.5f,
                                 nullptr),
              "adam_step");
    expect_ok(unetb200_step_increment(dstep.p, nullptr), "step_increment");
    const double lr = lr0 * std::pow(frac, step / decay);
    const double t = step + 1;
    const double lrt = lr * std::sqrt(1 - std::pow(b2, t)) / (1 - std::pow(b1, t));
    for (long long i = 0; i < n; ++i) {
      const double gk = g[i] * gs;
      rm[i] = b1 * rm[i] + (1 - b1) * gk;
      rv[i] = b2 * rv[i] + (1 - b2) * gk * gk;
      rt[i] -= lrt * rm[i] / (std::sqrt(rv[i]) + eps);
    }
  }
  sync_ok("adam");
  auto got = dth.down();
  got.resize(n);
  std::vector<float> ref(n);
  for (long long i = 0; i < n; ++i) ref[i] = static_cast<float>(rt[i]);
  report("adam 3 steps (decaying lr, grad_scale .5)", got, ref, 1e-6f);
}

int main(int argc, char** argv) {
  const bool quick = argc > 1 && std::string(argv[1]) == "quick";
  printf("libunetb200 version %d\n", unetb200_version());
  CK(cudaSetDevice(0));
  // conv3x3: kc=32 path, block_n 32; W=16 single tile
  test_conv3x3(1, 16, 16, 32, 0, 32);
  // kc=64, block_n 64, several tiles, dropout
  test_conv3x3(2, 32, 32, 64, 0, 64, true);
  // two sources (concat-free), kc=32, dual-destination dgrad
  test_conv3x3(2, 16, 16, 32, 32, 32);
  // W=8 tiles (two images per tile), odd N, block_n 256, n_tiles 2
  test_conv3x3(3, 8, 8, 256, 0, 512);
  if (!quick) {
    test_conv3x3(2, 16, 16, 256, 256, 256);
    test_conv3x3(2, 32, 32, 128, 0, 128);
    // more tiles than SMs: persistent loop + TMEM double buffering
    test_conv3x3(5, 64, 64, 32, 0, 64);
    test_conv3x3(1, 128, 128, 64, 0, 32);
  }
  test_tconv(2, 8, 8, 512, 256);
  test_tconv(2, 16, 16, 64, 32);
  if (!quick) {
    test_tconv(1, 32, 32, 128, 64);
    test_tconv(3, 8, 8, 64, 32);
  }
  test_pool_upsample(2, 16, 16, 32);
  test_pool_upsample(1, 8, 8, 512);
  test_first_and_head(2, 32, 32, 1, 1);
  test_first_and_head(1, 16, 16, 4, 2);
  test_adam();
  printf("SUMMARY: %d/%d failed\n", g_fail, g_total);
  return g_fail ? 1 : 0;
}
