#include "common.h"

#include <atomic>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace ub {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// The driver entry point is resolved through the runtime so the library never links libcuda directly.
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<EncodeTiledFn>(p);
    }
  });
  return fn;
}

// Tensor maps are cached per (base pointer, element size, rank, dims, strides, box, swizzle): a training step encodes
// the same ~150 maps over and over, and cuTensorMapEncodeTiled costs ~1 us each on the eager path.
// Unified virtual addressing makes the base pointer unique across devices. The cache is bounded: when full it is cleared.
namespace {
struct TmapKey {
  const void* base;
  int rank, swizzle, elt;
  uint64_t dims[5];
  uint64_t strides[4];
  uint32_t box[5];
  bool operator==(const TmapKey& o) const { return std::memcmp(this, &o, sizeof(TmapKey)) == 0; }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    const unsigned char* p = reinterpret_cast<const unsigned char*>(&k);
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(TmapKey); ++i) h = (h ^ p[i]) * 1099511628211ull;
    return static_cast<size_t>(h);
  }
};
std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tmaps;
std::mutex g_tmaps_mu;
std::atomic<long long> g_tmap_hits{0}, g_tmap_misses{0};
constexpr size_t kTmapCacheMax = 8192;
}  // namespace

long long tmap_cache_hits() { return g_tmap_hits.load(); }
long long tmap_cache_misses() { return g_tmap_misses.load(); }

int make_tmap(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
              const uint32_t* box, int swizzle_bytes, int elt_bytes) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available (no CUDA driver?)");
    return UB_ERR_CUDA;
  }
  TmapKey key;
  std::memset(&key, 0, sizeof(key));   // padding bytes too: the key is compared and hashed as raw bytes
  key.base = base; key.rank = rank; key.swizzle = swizzle_bytes; key.elt = elt_bytes;
  for (int i = 0; i < rank; ++i) { key.dims[i] = dims[i]; key.box[i] = box[i]; }
  for (int i = 0; i + 1 < rank; ++i) key.strides[i] = strides_bytes[i];
  {
    std::lock_guard<std::mutex> lk(g_tmaps_mu);
    auto it = g_tmaps.find(key);
    if (it != g_tmaps.end()) {
      *out = it->second;
      g_tmap_hits.fetch_add(1, std::memory_order_relaxed);
      return UB_OK;
    }
  }
  cuuint64_t gdims[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_NONE;
  if (swizzle_bytes == 128) sw = CU_TENSOR_MAP_SWIZZLE_128B;
  else if (swizzle_bytes == 64) sw = CU_TENSOR_MAP_SWIZZLE_64B;
  else if (swizzle_bytes == 32) sw = CU_TENSOR_MAP_SWIZZLE_32B;
  else if (swizzle_bytes == UB_SWIZZLE_128B_ATOM32) sw = CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B;
  // fp32 tensors hold values already rounded to tf32 by their producers: plain FLOAT32, no conversion on load
  const CUtensorMapDataType dt = elt_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUresult r = fn(out, dt, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdims,
                  gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (CUresult %d): rank=%d elt=%d dims=[%llu,%llu,%llu,%llu,%llu] box=[%u,%u,%u,%u,%u] "
              "stride0=%llu sw=%d base=%p",
              static_cast<int>(r), rank, elt_bytes, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
              (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0),
              (unsigned long long)(rank > 4 ? dims[4] : 0), box[0], rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0,
              rank > 3 ? box[3] : 0, rank > 4 ? box[4] : 0, (unsigned long long)(rank > 1 ? strides_bytes[0] : 0),
              swizzle_bytes, base);
    return UB_ERR_CUDA;
  }
  g_tmap_misses.fetch_add(1, std::memory_order_relaxed);
  {
    std::lock_guard<std::mutex> lk(g_tmaps_mu);
    if (g_tmaps.size() >= kTmapCacheMax) g_tmaps.clear();
    g_tmaps.emplace(key, *out);
  }
  return UB_OK;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box, int swizzle_bytes) {
  return make_tmap(out, base, rank, dims, strides_bytes, box, swizzle_bytes, 2);
}

// ------------------------------------------------------------------------------------------------ launch accounting
namespace {
struct ProfRec {
  cudaEvent_t e0, e1;
  int kind;
  double flops, bytes;
};
std::atomic<long long> g_launches{0};
bool g_prof_on = false;
std::vector<ProfRec> g_prof;
std::mutex g_prof_mu;
}  // namespace

void note_api_call(const char* expr_text) {
  // "cudaGetLastError()" is checked exactly once after every kernel launch in this library
  if (expr_text[4] == 'G' && expr_text[7] == 'L' && std::strcmp(expr_text, "cudaGetLastError()") == 0) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
  }
}
long long launch_count() { return g_launches.load(std::memory_order_relaxed); }
bool profiling_enabled() { return g_prof_on; }

void prof_begin(int kind, double flops, double bytes, cudaStream_t stream) {
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfRec r;
  r.kind = kind;
  r.flops = flops;
  r.bytes = bytes;
  cudaEventCreate(&r.e0);
  cudaEventCreate(&r.e1);
  cudaEventRecord(r.e0, stream);
  g_prof.push_back(r);
}
void prof_end(cudaStream_t stream) {
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (!g_prof.empty()) cudaEventRecord(g_prof.back().e1, stream);
}

extern "C" {
long long unetb200_launch_count(void) { return launch_count(); }
int unetb200_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = on != 0;
  return UB_OK;
}
// Synchronises, then sums per kind: ms[k], flops[k], bytes[k], launches[k] (arrays of 3); clears the records.
int unetb200_profile_collect(double* ms, double* flops, double* bytes, long long* launches) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("profile_collect: %s", cudaGetErrorString(e));
    return UB_ERR_CUDA;
  }
  for (int k = 0; k < UB_NUM_KINDS; ++k) { ms[k] = 0; flops[k] = 0; bytes[k] = 0; launches[k] = 0; }
  FILE* csv = nullptr;
  if (const char* path = getenv("UNETB200_PROFILE_CSV")) csv = fopen(path, "a");
  if (csv) fprintf(csv, "kind,gflop,mbytes,ms,tflops\n");
  for (auto& r : g_prof) {
    float t = 0.f;
    cudaEventElapsedTime(&t, r.e0, r.e1);
    if (csv) fprintf(csv, "%d,%.4f,%.3f,%.5f,%.2f\n", r.kind, r.flops * 1e-9, r.bytes * 1e-6, t, r.flops / (t * 1e-3) * 1e-12);
    ms[r.kind] += t;
    flops[r.kind] += r.flops;
    bytes[r.kind] += r.bytes;
    launches[r.kind] += 1;
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  if (csv) fclose(csv);
  g_prof.clear();
  return UB_OK;
}
}  // extern "C"

namespace {
std::atomic<int> g_reserved_sms{0};
}

namespace {
// UNETB200_PDL=all: every converted launch; =fwd: only inside a PdlScope (the plan's forward pass, where no second
// stream competes for the SMs a draining kernel frees); unset / anything else: never.  Off by default -- measured on one
// B200 inside the captured step (profiles/r02_pdl.txt): 2.534 ms off, 2.560 ms fwd, 2.56-2.61 ms all.  With `all` the
// next dgrad kernel takes the SMs a draining kernel frees before the weight-gradient stream can use them.
int pdl_mode() {
  static const int mode = [] {
    const char* e = getenv("UNETB200_PDL");
    if (!e) return 0;
    return strcmp(e, "all") == 0 ? 2 : (strcmp(e, "fwd") == 0 ? 1 : 0);
  }();
  return mode;
}
thread_local int t_pdl_scope = 0;
}  // namespace

bool pdl_enabled() { return pdl_mode() == 2 || (pdl_mode() == 1 && t_pdl_scope > 0); }
PdlScope::PdlScope() { ++t_pdl_scope; }
PdlScope::~PdlScope() { --t_pdl_scope; }

int set_reserved_sms(int n) {
  if (n < 0) n = 0;
  return g_reserved_sms.exchange(n);
}

int device_sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  const int usable = cached[dev] - g_reserved_sms.load(std::memory_order_relaxed);
  return usable > 8 ? usable : 8;
}

extern "C" int unetb200_set_reserved_sms(int n) { return set_reserved_sms(n); }

}  // namespace ub
