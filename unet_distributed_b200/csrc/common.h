// Host-side shared declarations for libunetb200: error reporting, tensor-map construction, device info.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>

namespace ub {

// Error codes returned across the C ABI (0 = OK, negative = failure; text via unetb200_last_error_string()).
enum : int {
  UB_OK = 0,
  UB_ERR_INVALID = -1,    // bad argument / unsupported shape
  UB_ERR_CUDA = -2,       // CUDA runtime or driver error
  UB_ERR_WORKSPACE = -3,  // workspace too small
  UB_ERR_UNSUPPORTED = -4
};

void set_error(const char* fmt, ...);
const char* last_error();

// Launch accounting and optional per-launch timing (bench.py's roofline numbers come from here).
enum : int { UB_KIND_IGEMM = 0, UB_KIND_WGRAD = 1, UB_KIND_OTHER = 2, UB_NUM_KINDS = 3 };
void note_api_call(const char* expr_text);  // counts kernel launches: every launch site checks cudaGetLastError()
long long launch_count();
bool profiling_enabled();
// Bracket one launch with events on `stream` when profiling is on (no-ops otherwise).
void prof_begin(int kind, double flops, double bytes, cudaStream_t stream);
void prof_end(cudaStream_t stream);

#define UB_CHECK_CUDA(expr)                                                                        \
  do {                                                                                             \
    ::ub::note_api_call(#expr);                                                                    \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) {                                                                       \
      ::ub::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));       \
      return ::ub::UB_ERR_CUDA;                                                                    \
    }                                                                                              \
  } while (0)

#define UB_REQUIRE(cond, ...)                 \
  do {                                        \
    if (!(cond)) {                            \
      ::ub::set_error(__VA_ARGS__);           \
      return ::ub::UB_ERR_INVALID;            \
    }                                         \
  } while (0)

// Encode a tiled bf16 tensor map. dims/box are innermost-first; strides_bytes has rank-1 entries (dims 1..rank-1).
// swizzle_bytes: 128, 64, 32 or 0. Out-of-bounds elements read as zero.
int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box, int swizzle_bytes);
// Same for elements of elt_bytes = 2 (bf16) or 4 (fp32 holding tf32-rounded values). Encoded maps are cached by
// (base, element size, rank, dims, strides, box, swizzle); tmap_cache_hits/misses count look-ups (tests, tuning).
int make_tmap(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
              const uint32_t* box, int swizzle_bytes, int elt_bytes);
long long tmap_cache_hits();
long long tmap_cache_misses();

// Activation / operand element types inside the library (the ABI enum UNETB200_BF16 / UNETB200_TF32 maps onto these).
enum : int { UB_DT_BF16 = 0, UB_DT_TF32 = 1 };
// swizzle_bytes value for make_tmap: 128-byte span swizzled in 32-byte chunks (CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B),
// the only shared-memory layout tcgen05 accepts for MN-major tf32 operands (UMMA layout type SWIZZLE_128B_BASE32B)
enum : int { UB_SWIZZLE_128B_ATOM32 = 1032 };
inline int dt_bytes(int dt) { return dt == UB_DT_TF32 ? 4 : 2; }

// SMs this library sizes its persistent grids and split-K factors for: the device's SM count minus the reservation
// (unetb200_set_reserved_sms). A kernel of one CTA per SM that finds k SMs taken by a concurrent NCCL kernel runs its
// last k CTAs as a second wave and takes twice as long; leaving those SMs out of the grid costs k/148 instead.
int device_sm_count();
int set_reserved_sms(int n);   // returns the previous reservation

// Programmatic dependent launch (UNETB200_PDL=all|fwd|off, see common.cu): the kernel may become resident while the previous
// kernel of the stream drains, runs its prologue (barriers, TMEM allocation, tensor-map prefetch) and then blocks in
// griddepcontrol.wait (pdl_wait() in ptx.cuh) until that kernel has completed and its writes are visible.  ONLY for
// kernels in which every thread executes pdl_wait() before its first global-memory access.  Stream capture keeps the
// edge as a programmatic dependency; an event record or a cross-stream wait between two launches makes it a full one.
bool pdl_enabled();
struct PdlScope {   // launches of this thread inside the scope use the attribute when UNETB200_PDL=fwd
  PdlScope();
  ~PdlScope();
};
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace ub
