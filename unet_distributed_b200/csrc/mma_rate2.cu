// Micro-benchmark (development tool, not part of the library): steady-state cost of one tcgen05.mma.cta_group::2
// (M = 256 across a CTA pair -- 128 accumulator rows per SM --, K = 16, bf16, SS operands) as a function of N, next to
// the cta_group::1 numbers of mma_rate.cu. In a pair every SM reads its own 128 x 16 A tile but only HALF of the
// N x 16 B tile from its shared memory, so if the cta_group::1 cost of 32 + N/4 cycles is operand fetch at
// 128 B/cycle, the pair should cost 32 + N/8 per SM (until the math rate N/2 takes over).
// Both CTAs read their own shared memory only; nothing here touches another GPU.
// Build: make build/mma_rate2; run: build/mma_rate2
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ptx.cuh"

using namespace ub;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish2() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma2_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at the same shared-memory offset in both CTAs of the pair
__device__ __forceinline__ void umma2_commit_both(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3))
      : "memory");
}

__device__ __forceinline__ uint64_t desc_from(uint32_t lo, uint32_t hi) {
  return (static_cast<uint64_t>(hi) << 32) | lo;
}

struct RateParams {
  int groups;          // groups of 4 MMAs (one 64-wide K block) issued by the leader CTA
  long long* cycles;   // per cluster
};

template <int N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) rate2_kernel(RateParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t done_bar;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();
  uint32_t* w = reinterpret_cast<uint32_t*>(smem);
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) w[i] = 0x3c003c00u + (i * 2654435761u >> 28);
  if (threadIdx.x == 0) {
    mbar_init(&done_bar, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc2(&tmem_base_s, 512);
    tmem_relinquish2();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  cluster_sync();            // both CTAs have their operands, barrier and TMEM ready
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  constexpr uint32_t kRow = 128;   // SWIZZLE_128B rows: 64 bf16 of K
  const uint64_t a0 = umma_smem_desc(smem_u32(smem), 0, 8 * kRow, UMMA_SW_128B);             // 128 rows of A per CTA
  const uint64_t b0 = umma_smem_desc(smem_u32(smem + 32768), 0, 8 * kRow, UMMA_SW_128B);     // N/2 rows of B per CTA
  const uint32_t a_hi = static_cast<uint32_t>(a0 >> 32), b_hi = static_cast<uint32_t>(b0 >> 32);
  const uint32_t a_lo = static_cast<uint32_t>(a0), b_lo = static_cast<uint32_t>(b0);
  const uint32_t idesc = umma_idesc_bf16(256, N, 0, 0);
  if (warp == 0) {
    if (rank == 0) {
      if (elect_one()) {
        const long long t0 = clock64();
        for (int g = 0; g < p.groups; ++g) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            umma2_f16(tmem, desc_from(a_lo + kk * 2, a_hi), desc_from(b_lo + kk * 2, b_hi), idesc, (g | kk) ? 1u : 0u);
          }
        }
        umma2_commit_both(&done_bar);
        mbar_wait(&done_bar, 0);
        p.cycles[blockIdx.x >> 1] = clock64() - t0;
      }
    } else {
      if (elect_one()) mbar_wait(&done_bar, 0);   // the peer's tensor core works until the leader's commit fires
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();
  if (warp == 1) tmem_dealloc2(tmem, 512);
}

template <int N>
static void run(int clusters, long long* d_cycles) {
  std::vector<long long> h(clusters);
  cudaFuncSetAttribute(rate2_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000);
  RateParams p;
  p.groups = 1024;
  p.cycles = d_cycles;
  for (int rep = 0; rep < 2; ++rep) rate2_kernel<N><<<2 * clusters, 128, 70000>>>(p);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("N=%d: launch failed: %s\n", N, cudaGetErrorString(e));
    exit(1);
  }
  cudaMemcpy(h.data(), d_cycles, clusters * sizeof(long long), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (long long c : h) mx = c > mx ? c : mx;
  const double cyc = static_cast<double>(mx) / 4096.0;
  printf("%-5d %12.1f %14.1f %14.1f %12.1f\n", N, cyc, 32.0 + N / 8.0, 32.0 + N / 4.0 > N / 2.0 ? 32.0 + N / 4.0 : N / 2.0,
         N / 2.0);
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int clusters = sms / 2;
  long long* d_cycles;
  cudaMalloc(&d_cycles, clusters * sizeof(long long));
  printf("tcgen05.mma.cta_group::2, M=256 (128 rows per SM), K=16, bf16, SWIZZLE_128B, %d CTA pairs\n", clusters);
  printf("%-5s %12s %14s %14s %12s\n", "N", "cyc/mma", "32+N/8 (pred)", "cta_group::1", "math N/2");
  run<32>(clusters, d_cycles);
  run<64>(clusters, d_cycles);
  run<96>(clusters, d_cycles);
  run<128>(clusters, d_cycles);
  run<192>(clusters, d_cycles);
  run<256>(clusters, d_cycles);
  cudaFree(d_cycles);
  return 0;
}
