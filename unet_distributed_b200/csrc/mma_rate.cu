// Micro-benchmark (development tool, not part of the library): steady-state cost of one tcgen05.mma
// (M=128, K=16, bf16, SS operands) as a function of N, the swizzle mode, the number of accumulators the issue loop
// rotates through and whether the A descriptor changes between MMAs. One CTA per SM, one elected issuing lane.
// Build: make mma_rate; run: build/mma_rate
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ptx.cuh"

using namespace ub;

struct RateParams {
  int a_step;     // A descriptor advances by a_step rows per group (mod 64 rows); 0 = fixed
  int groups;     // groups of NACC * KPT MMAs per CTA
  long long* cycles;
};

__device__ __forceinline__ uint64_t desc_from(uint32_t lo, uint32_t hi) {
  return (static_cast<uint64_t>(hi) << 32) | lo;
}

// N: MMA N; SW128: 128-byte rows (kc = 64, 4 MMAs per tap) or 64-byte rows (kc = 32, 2 MMAs per tap);
// NACC accumulators; the issue loop body is fully unrolled so that every descriptor is base + immediate.
template <int N, bool SW128, int NACC>
__global__ void __launch_bounds__(128, 1) rate_kernel(RateParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t done_bar;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5;
  uint32_t* w = reinterpret_cast<uint32_t*>(smem);
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) w[i] = 0x3c003c00u + (i * 2654435761u >> 28);
  if (threadIdx.x == 0) {
    mbar_init(&done_bar, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(&tmem_base_s, 512);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  constexpr uint32_t kRow = SW128 ? 128 : 64;
  constexpr int KPT = SW128 ? 4 : 2;
  const uint64_t a0 = umma_smem_desc(smem_u32(smem), 0, 8 * kRow, SW128 ? UMMA_SW_128B : UMMA_SW_64B);
  const uint64_t b0 = umma_smem_desc(smem_u32(smem + 32768), 0, 8 * kRow, SW128 ? UMMA_SW_128B : UMMA_SW_64B);
  const uint32_t a_hi = static_cast<uint32_t>(a0 >> 32), b_hi = static_cast<uint32_t>(b0 >> 32);
  const uint32_t a_lo = static_cast<uint32_t>(a0), b_lo = static_cast<uint32_t>(b0);
  const uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
  if (warp == 0) {
    if (elect_one()) {
      const long long t0 = clock64();
      uint32_t arow = 0;
      for (int g = 0; g < p.groups; ++g) {
        const uint32_t a_g = a_lo + ((arow * kRow) >> 4);
        const uint32_t accum = g > 0 ? 1u : 0u;
#pragma unroll
        for (int acc = 0; acc < NACC; ++acc) {
#pragma unroll
          for (int kk = 0; kk < KPT; ++kk) {
            umma_f16(tmem + acc * N, desc_from(a_g + acc * ((8 * kRow) >> 4) + kk * 2, a_hi), desc_from(b_lo + kk * 2, b_hi),
                     idesc, accum | (kk > 0 ? 1u : 0u));
          }
        }
        arow = (arow + p.a_step) & 63;
      }
      umma_commit(&done_bar);
      mbar_wait(&done_bar, 0);
      p.cycles[blockIdx.x] = clock64() - t0;
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

template <int N, bool SW128, int NACC>
static void run(int sms, long long* d_cycles) {
  if (N * NACC > 512) return;
  std::vector<long long> h(sms);
  constexpr int KPT = SW128 ? 4 : 2;
  cudaFuncSetAttribute(rate_kernel<N, SW128, NACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000);
  for (int a_step : {0, 1}) {
    RateParams p;
    p.a_step = a_step;
    p.groups = 4096 / (NACC * KPT);
    p.cycles = d_cycles;
    for (int rep = 0; rep < 2; ++rep) rate_kernel<N, SW128, NACC><<<sms, 128, 70000>>>(p);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("launch failed: %s\n", cudaGetErrorString(e));
      exit(1);
    }
    cudaMemcpy(h.data(), d_cycles, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (long long c : h) mx = c > mx ? c : mx;
    printf("%-5d %-5d %-5d %-6d %-5d %10.1f %10.1f\n", N, SW128 ? 128 : 64, NACC, a_step, KPT,
           static_cast<double>(mx) / 4096.0, 128.0 * N / 256.0);
  }
}

template <int N>
static void run_n(int sms, long long* d) {
  run<N, false, 1>(sms, d);
  run<N, false, 2>(sms, d);
  run<N, false, 4>(sms, d);
  run<N, true, 1>(sms, d);
  run<N, true, 2>(sms, d);
  run<N, true, 4>(sms, d);
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  long long* d_cycles;
  cudaMalloc(&d_cycles, sms * sizeof(long long));
  printf("%-5s %-5s %-5s %-6s %-5s %10s %10s\n", "N", "sw", "nacc", "a_step", "kpt", "cyc/mma", "ideal");
  run_n<32>(sms, d_cycles);
  run_n<64>(sms, d_cycles);
  run_n<96>(sms, d_cycles);
  run_n<128>(sms, d_cycles);
  run_n<256>(sms, d_cycles);
  cudaFree(d_cycles);
  return 0;
}
