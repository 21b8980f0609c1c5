// Micro-benchmark (development tool): cost of a dependent-kernel boundary inside a CUDA graph, with and without
// programmatic dependent launch. Chains of K kernels, each one CTA per SM doing ~W cycles of work.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void work_kernel(float* p, int spin, int pdl) {
  if (pdl) asm volatile("griddepcontrol.launch_dependents;");
  __shared__ float s[256];
  s[threadIdx.x] = threadIdx.x;
  __syncthreads();
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  float v = p[blockIdx.x * blockDim.x + threadIdx.x];
  const long long t0 = clock64();
  while (clock64() - t0 < spin) v = v * 1.0001f + s[(threadIdx.x + 1) & 255];
  p[blockIdx.x * blockDim.x + threadIdx.x] = v;
}

static float run_chain(cudaStream_t st, float* d, int sms, int K, int spin, int pdl, int reps) {
  cudaGraph_t g;
  cudaGraphExec_t ge;
  cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
  for (int k = 0; k < K; ++k) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(sms);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, work_kernel, d, spin, pdl);
  }
  cudaStreamEndCapture(st, &g);
  cudaError_t e = cudaGraphInstantiate(&ge, g, 0);
  if (e != cudaSuccess) { printf("instantiate: %s\n", cudaGetErrorString(e)); return -1; }
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaGraphLaunch(ge, st);
  cudaStreamSynchronize(st);
  cudaEventRecord(a, st);
  for (int r = 0; r < reps; ++r) cudaGraphLaunch(ge, st);
  cudaEventRecord(b, st);
  cudaStreamSynchronize(st);
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
  return ms * 1e3f / (reps * K);
}

int main() {
  int sms = 0, khz = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  float* d;
  cudaMalloc(&d, sms * 256 * sizeof(float));
  cudaMemset(d, 0, sms * 256 * sizeof(float));
  cudaStream_t st;
  cudaStreamCreate(&st);
  printf("%-10s %-6s %12s\n", "spin(cyc)", "pdl", "us/kernel");
  for (int spin : {0, 20000, 60000}) {
    for (int pdl = 0; pdl < 2; ++pdl) {
      const float us = run_chain(st, d, sms, 100, spin, pdl, 20);
      printf("%-10d %-6d %12.2f   (work alone ~%.1f us at %d MHz)\n", spin, pdl, us, spin / (khz / 1e3), khz / 1000);
    }
  }
  return 0;
}
