// how much does a kernel boundary cost inside a CUDA graph on B200, with and without programmatic dependent launch?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_plain(float* p, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = p[i] * 1.0001f + 1.f;
}
__global__ void k_pdl(float* p, int n) {
  asm volatile("griddepcontrol.launch_dependents;");
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (i < n) p[i] = p[i] * 1.0001f + 1.f;
}
template <bool PDL>
float run(int n_kernels, int blocks, float* d, int n) {
  cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  cudaGraph_t g; cudaGraphExec_t ge;
  cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
  for (int i = 0; i < n_kernels; ++i) {
    if (!PDL) k_plain<<<blocks, 256, 0, s>>>(d, n);
    else {
      cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(256); cfg.stream = s;
      cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1; cfg.attrs = at; cfg.numAttrs = 1;
      cudaError_t e = cudaLaunchKernelEx(&cfg, k_pdl, d, n);
      if (e != cudaSuccess) { printf("launch err %s\n", cudaGetErrorString(e)); }
    }
  }
  cudaError_t e = cudaStreamEndCapture(s, &g);
  if (e != cudaSuccess) { printf("capture err %s\n", cudaGetErrorString(e)); return -1; }
  e = cudaGraphInstantiate(&ge, g, 0);
  if (e != cudaSuccess) { printf("inst err %s\n", cudaGetErrorString(e)); return -1; }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 5; ++i) cudaGraphLaunch(ge, s);
  cudaStreamSynchronize(s);
  cudaEventRecord(e0, s);
  const int reps = 200;
  for (int i = 0; i < reps; ++i) cudaGraphLaunch(ge, s);
  cudaEventRecord(e1, s); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms * 1e3f / reps;
}
int main() {
  float* d; int n = 148 * 8 * 256; cudaMalloc(&d, n * 4); cudaMemset(d, 0, n * 4);
  for (int blocks : {1, 148, 1184}) {
    float a = run<false>(24, blocks, d, n), b = run<true>(24, blocks, d, n);
    float a1 = run<false>(1, blocks, d, n);
    printf("blocks %4d: 24 plain kernels %.2f us/graph (%.2f per kernel), PDL %.2f us/graph (%.2f per kernel); 1 kernel graph %.2f us\n",
           blocks, a, a / 24, b, b / 24, a1);
  }
  return 0;
}
