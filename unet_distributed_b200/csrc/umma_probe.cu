// Hardware probe for the UMMA shared-memory descriptor and TMA behaviours the conv kernels rely on (or would like to
// rely on). Every experiment loads small-integer bf16 matrices through TMA, issues tcgen05.mma with a descriptor built
// on the host, and compares the TMEM result with an exact host model; mismatches are diagnosed row by row.
// Answers it is after (all printed as PASS/FAIL lines):
//   * K-major SWIZZLE_128B / SWIZZLE_64B tiles with the +32-byte K advance (igemm path);
//   * MN-major SWIZZLE_128B / 64B tiles with LBO/SBO as used by wgrad;
//   * start addresses shifted by whole rows (halo re-use: one smem halo tile, nine tap views), with and without the
//     descriptor's base-offset field, and with a stride-byte-offset that is not a multiple of the swizzle period;
//   * TMA boxes larger than the tensor extent, negative coordinates and zero fill.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "common.h"
#include "ptx.cuh"

using namespace ub;

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e = (x);                                                              \
    if (e != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__);  \
      exit(2);                                                                        \
    }                                                                                 \
  } while (0)

constexpr int kRegion = 64 * 1024;  // bytes per operand region

struct ProbeParams {
  int a_rank, a_nbox;
  int a_coords[4][5];
  int a_off[4];
  int a_bytes[4];
  int b_nbox;
  int b_coords[4][2];
  int b_off[4];
  int b_bytes[4];
  unsigned long long adesc, bdesc;  // start field relative to the operand region
  unsigned int a_kstep, b_kstep;    // added to the descriptor's low word per K step (16-byte units)
  int nk;
  unsigned int idesc;
  int n;
  int run_mma;
};

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ ProbeParams p, const __grid_constant__ CUtensorMap tmA,
             const __grid_constant__ CUtensorMap tmB, float* out, uint4* dump) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
  uint8_t* sa = smem;
  uint8_t* sb = smem + kRegion;
  __shared__ __align__(8) uint64_t bar_load, bar_mma;
  __shared__ uint32_t tmem_slot;
  // poison so that untouched bytes are visible in the dump
  for (int i = threadIdx.x; i < 2 * kRegion / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem)[i] = make_uint4(0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u);
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_init(&bar_load, 1);
    mbar_init(&bar_mma, 1);
    mbar_fence_init();
  }
  if (threadIdx.x < 32) {
    tmem_alloc(&tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    uint32_t tx = 0;
    for (int i = 0; i < p.a_nbox; ++i) tx += p.a_bytes[i];
    for (int i = 0; i < p.b_nbox; ++i) tx += p.b_bytes[i];
    mbar_arrive_expect_tx(&bar_load, tx);
    for (int i = 0; i < p.a_nbox; ++i) {
      const int* c = p.a_coords[i];
      if (p.a_rank == 2) tma_load_2d(sa + p.a_off[i], &tmA, &bar_load, c[0], c[1]);
      else if (p.a_rank == 4) tma_load_4d(sa + p.a_off[i], &tmA, &bar_load, c[0], c[1], c[2], c[3]);
      else tma_load_5d(sa + p.a_off[i], &tmA, &bar_load, c[0], c[1], c[2], c[3], c[4]);
    }
    for (int i = 0; i < p.b_nbox; ++i) tma_load_2d(sb + p.b_off[i], &tmB, &bar_load, p.b_coords[i][0], p.b_coords[i][1]);
  }
  mbar_wait(&bar_load, 0);
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * kRegion / 16; i += blockDim.x) dump[i] = reinterpret_cast<uint4*>(smem)[i];
  if (p.run_mma) {
    if (threadIdx.x == 0) {
      tc_fence_after();
      const uint64_t ad = p.adesc + (smem_u32(sa) >> 4);
      const uint64_t bd = p.bdesc + (smem_u32(sb) >> 4);
      for (int k = 0; k < p.nk; ++k) {
        umma_f16(tmem, ad + static_cast<uint64_t>(k) * p.a_kstep, bd + static_cast<uint64_t>(k) * p.b_kstep, p.idesc,
                 k ? 1u : 0u);
      }
      umma_commit(&bar_mma);
    }
    mbar_wait(&bar_mma, 0);
    tc_fence_after();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = warp * 32 + lane;
    for (int c = 0; c < p.n; c += 16) {
      uint32_t r[16];
      tmem_ld_32x32b_x16(tmem + (static_cast<uint32_t>(warp * 32) << 16) + c, r);
      tmem_ld_wait();
      for (int i = 0; i < 16; ++i) out[row * 256 + c + i] = __uint_as_float(r[i]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------------ host side
static uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return static_cast<uint16_t>((u + 0x7FFFu + ((u >> 16) & 1u)) >> 16);
}
static float bf2f(uint16_t h) {
  uint32_t u = static_cast<uint32_t>(h) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static uint32_t g_rng = 99u;
static int irand(int lo, int hi) {
  g_rng = g_rng * 1664525u + 1013904223u;
  return lo + static_cast<int>((g_rng >> 10) % static_cast<uint32_t>(hi - lo + 1));
}

struct Run {
  std::vector<float> out;        // [128][256]
  std::vector<uint16_t> smemA;   // region image as bf16
  std::vector<uint16_t> smemB;
  bool launched_ok = false;
};

static Run launch(const ProbeParams& p, const CUtensorMap& ma, const CUtensorMap& mb) {
  Run r;
  float* dout;
  uint4* ddump;
  CK(cudaMalloc(&dout, 128 * 256 * 4));
  CK(cudaMalloc(&ddump, 2 * kRegion));
  CK(cudaMemset(dout, 0, 128 * 256 * 4));
  static bool attr = false;
  if (!attr) {
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * kRegion + 1024));
    attr = true;
  }
  probe_kernel<<<1, 128, 2 * kRegion + 1024>>>(p, ma, mb, dout, ddump);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("   kernel error: %s\n", cudaGetErrorString(e));
    exit(4);  // sticky error: the remaining experiments cannot run in this process
  }
  r.launched_ok = true;
  r.out.resize(128 * 256);
  CK(cudaMemcpy(r.out.data(), dout, 128 * 256 * 4, cudaMemcpyDeviceToHost));
  std::vector<uint16_t> img(kRegion);  // 2*kRegion bytes
  CK(cudaMemcpy(img.data(), ddump, 2 * kRegion, cudaMemcpyDeviceToHost));
  r.smemA.assign(img.begin(), img.begin() + kRegion / 2);
  r.smemB.assign(img.begin() + kRegion / 2, img.end());
  cudaFree(dout);
  cudaFree(ddump);
  return r;
}

// Address-bit swizzle model: XOR of 16-byte-chunk bits [4,4+b) with bits [7,7+b).
static uint32_t swz(uint32_t addr, int sw_bytes) {
  const int b = sw_bytes == 128 ? 3 : sw_bytes == 64 ? 2 : sw_bytes == 32 ? 1 : 0;
  const uint32_t mask = (1u << b) - 1u;
  return addr ^ (((addr >> 7) & mask) << 4);
}
static float smem_elem(const std::vector<uint16_t>& img, uint32_t byte_addr) {
  if (byte_addr / 2 >= img.size()) return NAN;
  return bf2f(img[byte_addr / 2]);
}

static int g_fail = 0, g_total = 0;

// Compare D against a host model; on mismatch, say for the first rows which model row they equal (if any).
static void judge(const std::string& name, const Run& r, const std::vector<float>& ref /*[128][n]*/, int n,
                  const std::vector<std::vector<float>>* candidates = nullptr) {
  double maxerr = 0;
  int bad_rows = 0;
  for (int i = 0; i < 128; ++i) {
    bool bad = false;
    for (int j = 0; j < n; ++j) {
      const double e = std::fabs(static_cast<double>(r.out[i * 256 + j]) - ref[i * n + j]);
      if (!(e <= 1e-3)) bad = true;
      if (e == e) maxerr = std::fmax(maxerr, e);
      else maxerr = 1e30;
    }
    bad_rows += bad;
  }
  ++g_total;
  if (bad_rows) ++g_fail;
  printf("%-72s %s bad_rows=%d maxerr=%.3g\n", name.c_str(), bad_rows ? "FAIL" : "PASS", bad_rows, maxerr);
  if (bad_rows && candidates) {
    printf("    row -> matching candidate row: ");
    for (int i = 0; i < 128; i += (i < 24 ? 1 : 8)) {
      int match = -1;
      for (size_t c = 0; c < candidates->size(); ++c) {
        bool ok = true;
        for (int j = 0; j < n && ok; ++j) ok = std::fabs(r.out[i * 256 + j] - (*candidates)[c][j]) <= 1e-3;
        if (ok) { match = static_cast<int>(c); break; }
      }
      printf("%d:%d ", i, match);
    }
    printf("\n");
  }
  fflush(stdout);
}

static CUtensorMap map2d(const void* base, int rows, int inner, int box_rows, int box_inner, int sw) {
  CUtensorMap m;
  const uint64_t dims[2] = {static_cast<uint64_t>(inner), static_cast<uint64_t>(rows)};
  const uint64_t str[1] = {static_cast<uint64_t>(inner) * 2};
  const uint32_t box[2] = {static_cast<uint32_t>(box_inner), static_cast<uint32_t>(box_rows)};
  if (make_tmap_bf16(&m, base, 2, dims, str, box, sw) != 0) {
    printf("   tensor map encode failed: %s\n", last_error());
    memset(&m, 0, sizeof(m));
  }
  return m;
}

struct Mat {
  int rows, cols;
  std::vector<float> v;
  std::vector<uint16_t> h;
  void* d = nullptr;
  Mat(int r, int c) : rows(r), cols(c), v(static_cast<size_t>(r) * c), h(v.size()) {
    for (size_t i = 0; i < v.size(); ++i) {
      v[i] = static_cast<float>(irand(-3, 3));
      h[i] = f2bf(v[i]);
    }
    CK(cudaMalloc(&d, h.size() * 2));
    CK(cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  }
  float at(int r, int c) const { return v[static_cast<size_t>(r) * cols + c]; }
};

// ---- K-major experiment: A rows come from a [256][inner] matrix loaded as one box at region offset 0.
// Descriptor: start = shift_rows*rowbytes, SBO = sbo_bytes, swizzle = inner*2 bytes. Model: D row i uses A row
// shift + (i/8)*(SBO/rowbytes) + i%8.
static void exp_kmajor(const char* label, int inner, int shift_rows, int sbo_bytes, int base_off_mode, int n) {
  const int rowbytes = inner * 2;
  Mat A(256, inner), B(256, inner);
  CUtensorMap ma = map2d(A.d, 256, inner, 256, inner, rowbytes);
  CUtensorMap mb = map2d(B.d, 256, inner, n, inner, rowbytes);
  ProbeParams p;
  memset(&p, 0, sizeof(p));
  p.a_rank = 2; p.a_nbox = 1; p.a_bytes[0] = 256 * rowbytes;
  p.b_nbox = 1; p.b_bytes[0] = n * rowbytes;
  const uint32_t start = shift_rows * rowbytes;
  const uint32_t sw = inner == 64 ? UMMA_SW_128B : UMMA_SW_64B;
  uint32_t bo = 0;
  if (base_off_mode == 1) bo = (start >> 7) & 7;
  p.adesc = umma_smem_desc(start, 0, sbo_bytes, sw, bo);
  p.bdesc = umma_smem_desc(0, 0, 8 * rowbytes, sw, 0);
  p.a_kstep = 2; p.b_kstep = 2;
  p.nk = inner / 16;
  p.idesc = umma_idesc_bf16(128, n, 0, 0);
  p.n = n;
  p.run_mma = 1;
  Run r = launch(p, ma, mb);
  // verify the TMA image model for A first (row r, chunk c at swz(r*rowbytes + c*16))
  int img_bad = 0;
  for (int rr = 0; rr < 256; ++rr)
    for (int k = 0; k < inner; ++k) {
      const uint32_t addr = swz(rr * rowbytes + k * 2, rowbytes);
      if (smem_elem(r.smemA, addr) != A.at(rr, k)) ++img_bad;
    }
  std::vector<float> ref(128 * n);
  const int group_rows = sbo_bytes / rowbytes;
  for (int i = 0; i < 128; ++i) {
    const int ar = shift_rows + (i / 8) * group_rows + i % 8;
    for (int j = 0; j < n; ++j) {
      float acc = 0;
      if (ar < 256)
        for (int k = 0; k < inner; ++k) acc += A.at(ar, k) * B.at(j, k);
      ref[i * n + j] = acc;
    }
  }
  std::vector<std::vector<float>> cand(256, std::vector<float>(n));
  for (int ar = 0; ar < 256; ++ar)
    for (int j = 0; j < n; ++j) {
      float acc = 0;
      for (int k = 0; k < inner; ++k) acc += A.at(ar, k) * B.at(j, k);
      cand[ar][j] = acc;
    }
  char name[160];
  snprintf(name, sizeof(name), "%s K-major inner=%d shift=%d rows SBO=%d baseoff=%s N=%d (tma image bad=%d)", label,
           inner, shift_rows, sbo_bytes, base_off_mode ? "addr" : "0", n, img_bad);
  judge(name, r, ref, n, &cand);
}

// ---- MN-major experiment: A = [kp pixels][am_atoms*atom ch] as am_atoms boxes of [kp][atom]; B likewise with N.
// shift_rows shifts the pixel (K) window of A by whole rows.
static void exp_mnmajor(const char* label, int atom, int n, int shift_rows, int base_off_mode) {
  const int rowbytes = atom * 2;
  const int kp = 64;
  const int m_atoms = 128 / atom;
  const int n_atoms = (n + atom - 1) / atom;
  Mat A(128, 128), B(128, 128);  // [pixels][channels]
  CUtensorMap ma = map2d(A.d, 128, 128, 96, atom, rowbytes);  // 96-row boxes leave room for shifted windows
  CUtensorMap mb = map2d(B.d, 128, 128, kp, atom, rowbytes);
  ProbeParams p;
  memset(&p, 0, sizeof(p));
  p.a_rank = 2; p.a_nbox = m_atoms;
  const int a_atom_bytes = 96 * rowbytes;
  for (int a = 0; a < m_atoms; ++a) {
    p.a_coords[a][0] = a * atom; p.a_coords[a][1] = 0;
    p.a_off[a] = a * a_atom_bytes; p.a_bytes[a] = a_atom_bytes;
  }
  p.b_nbox = n_atoms;
  const int b_atom_bytes = kp * rowbytes;
  for (int a = 0; a < n_atoms; ++a) {
    p.b_coords[a][0] = a * atom; p.b_coords[a][1] = 0;
    p.b_off[a] = a * b_atom_bytes; p.b_bytes[a] = b_atom_bytes;
  }
  const uint32_t sw = atom == 64 ? UMMA_SW_128B : UMMA_SW_64B;
  const uint32_t start = shift_rows * rowbytes;
  uint32_t bo = 0;
  if (base_off_mode == 1) bo = (start >> 7) & 7;
  p.adesc = umma_smem_desc(start, a_atom_bytes, 8 * rowbytes, sw, bo);
  p.bdesc = umma_smem_desc(0, b_atom_bytes, 8 * rowbytes, sw, 0);
  p.a_kstep = (16 * rowbytes) >> 4; p.b_kstep = (16 * rowbytes) >> 4;
  p.nk = kp / 16;
  p.idesc = umma_idesc_bf16(128, n, 1, 1);
  p.n = n;
  p.run_mma = 1;
  Run r = launch(p, ma, mb);
  // D[m][j] = sum_k A[shift+k][m] * B[k][j]
  std::vector<float> ref(128 * n);
  for (int m = 0; m < 128; ++m)
    for (int j = 0; j < n; ++j) {
      float acc = 0;
      for (int k = 0; k < kp; ++k) acc += A.at(shift_rows + k, m) * B.at(k, j);
      ref[m * n + j] = acc;
    }
  // candidates: D row for every (shift', m)
  std::vector<std::vector<float>> cand;
  for (int s = 0; s < 32; ++s)
    for (int m = 0; m < 128; ++m) {
      std::vector<float> c(n);
      for (int j = 0; j < n; ++j) {
        float acc = 0;
        for (int k = 0; k < kp; ++k) acc += A.at(s + k, m) * B.at(k, j);
        c[j] = acc;
      }
      cand.push_back(c);
    }
  char name[160];
  snprintf(name, sizeof(name), "%s MN-major atom=%d N=%d shift=%d rows baseoff=%s (cand idx = shift*128+m)", label, atom,
           n, shift_rows, base_off_mode ? "addr" : "0");
  judge(name, r, ref, n, &cand);
}

// ---- TMA behaviour: 4-D NHWC boxes with negative / out-of-range coordinates and boxes larger than the tensor.
static void exp_tma4d(int N, int H, int W, int C, int box_c, int TW, int TH, int TN, int c0, int w0, int h0, int n0) {
  std::vector<float> v(static_cast<size_t>(N) * H * W * C);
  std::vector<uint16_t> hbuf(v.size());
  for (size_t i = 0; i < v.size(); ++i) {
    v[i] = static_cast<float>(irand(1, 100));
    hbuf[i] = f2bf(v[i]);
  }
  void* d;
  CK(cudaMalloc(&d, hbuf.size() * 2));
  CK(cudaMemcpy(d, hbuf.data(), hbuf.size() * 2, cudaMemcpyHostToDevice));
  CUtensorMap m;
  const uint64_t dims[4] = {static_cast<uint64_t>(C), static_cast<uint64_t>(W), static_cast<uint64_t>(H),
                            static_cast<uint64_t>(N)};
  const uint64_t str[3] = {static_cast<uint64_t>(C) * 2, static_cast<uint64_t>(W) * C * 2,
                           static_cast<uint64_t>(H) * W * C * 2};
  const uint32_t box[4] = {static_cast<uint32_t>(box_c), static_cast<uint32_t>(TW), static_cast<uint32_t>(TH),
                           static_cast<uint32_t>(TN)};
  char name[200];
  snprintf(name, sizeof(name), "TMA 4-D tensor N%d H%d W%d C%d box (c%d,w%d,h%d,n%d) at (%d,%d,%d,%d)", N, H, W, C,
           box_c, TW, TH, TN, c0, w0, h0, n0);
  ++g_total;
  if (make_tmap_bf16(&m, d, 4, dims, str, box, box_c * 2) != 0) {
    printf("%-72s FAIL encode: %s\n", name, last_error());
    ++g_fail;
    cudaFree(d);
    return;
  }
  Mat B(256, 64);
  CUtensorMap mb = map2d(B.d, 256, 64, 8, 64, 128);
  ProbeParams p;
  memset(&p, 0, sizeof(p));
  p.a_rank = 4; p.a_nbox = 1;
  p.a_coords[0][0] = c0; p.a_coords[0][1] = w0; p.a_coords[0][2] = h0; p.a_coords[0][3] = n0;
  p.a_bytes[0] = box_c * TW * TH * TN * 2;
  p.b_nbox = 1; p.b_bytes[0] = 8 * 128;
  p.run_mma = 0;
  Run r = launch(p, m, mb);
  const int rowbytes = box_c * 2;
  int bad = 0;
  for (int tn = 0; tn < TN; ++tn)
    for (int th = 0; th < TH; ++th)
      for (int tw = 0; tw < TW; ++tw)
        for (int c = 0; c < box_c; ++c) {
          const int row = (tn * TH + th) * TW + tw;
          const int n = n0 + tn, h = h0 + th, w = w0 + tw, cc = c0 + c;
          float expect = 0.f;
          if (n >= 0 && n < N && h >= 0 && h < H && w >= 0 && w < W && cc >= 0 && cc < C)
            expect = v[((static_cast<size_t>(n) * H + h) * W + w) * C + cc];
          const float got = smem_elem(r.smemA, swz(row * rowbytes + c * 2, rowbytes));
          if (got != expect) {
            if (bad < 4) printf("    mismatch row %d c %d: got %g expect %g\n", row, c, got, expect);
            ++bad;
          }
        }
  if (bad) ++g_fail;
  printf("%-72s %s bad=%d\n", name, bad ? "FAIL" : "PASS", bad);
  cudaFree(d);
}

int main() {
  CK(cudaSetDevice(0));
  printf("== baseline layouts used by igemm / wgrad\n");
  exp_kmajor("E1", 64, 0, 1024, 0, 64);
  exp_kmajor("E1b", 64, 0, 1024, 0, 256);
  exp_kmajor("E2", 32, 0, 512, 0, 32);
  exp_mnmajor("E7", 64, 64, 0, 0);
  exp_mnmajor("E7b", 64, 128, 0, 0);
  exp_mnmajor("E8", 32, 32, 0, 0);
  exp_mnmajor("E9 partial-atom N", 64, 32, 0, 0);
  printf("== row-shifted start addresses (halo re-use)\n");
  for (int s : {1, 2, 3, 8, 9}) exp_kmajor("E3", 64, s, 1024, 0, 64);
  for (int s : {1, 2, 9}) exp_kmajor("E4", 64, s, 1024, 1, 64);
  for (int s : {0, 1, 11}) exp_kmajor("E5 pitch10", 64, s, 1280, 0, 64);
  for (int s : {1, 11}) exp_kmajor("E5b pitch10", 64, s, 1280, 1, 64);
  for (int s : {0, 1, 19}) exp_kmajor("E5c pitch18", 64, s, 18 * 128, 0, 64);
  for (int s : {1, 2, 5}) exp_kmajor("E6 sw64", 32, s, 512, 0, 32);
  for (int s : {1, 5}) exp_kmajor("E6b sw64", 32, s, 512, 1, 32);
  for (int s : {0, 1, 11}) exp_kmajor("E6c sw64 pitch10", 32, s, 640, 0, 32);
  for (int s : {1, 2, 9, 17}) exp_mnmajor("E10", 64, 64, s, 0);
  for (int s : {1, 9}) exp_mnmajor("E10b", 64, 64, s, 1);
  for (int s : {1, 9}) exp_mnmajor("E10c sw64", 32, 32, s, 0);
  printf("== TMA zero fill / oversized boxes\n");
  exp_tma4d(2, 16, 16, 64, 64, 16, 8, 1, 0, -1, -1, 0);
  exp_tma4d(2, 16, 16, 64, 64, 16, 8, 1, 0, 1, 9, 1);
  exp_tma4d(3, 8, 8, 64, 64, 8, 8, 2, 0, -1, 1, 2);
  exp_tma4d(2, 16, 16, 32, 64, 16, 8, 1, 0, 0, 0, 0);    // channel box larger than C
  exp_tma4d(2, 15, 15, 64, 64, 16, 8, 1, 0, 0, 8, 0);    // W box larger than W
  exp_tma4d(2, 16, 16, 64, 64, 16, 8, 1, 64, 0, 0, 0);   // box entirely out of range
  printf("PROBE SUMMARY: %d/%d failed\n", g_fail, g_total);
  return 0;
}
