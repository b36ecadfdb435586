// micro-benchmark: a chain of N dependent kernels (each reads the previous one's output) replayed from a CUDA graph,
// with and without programmatic dependent launch (griddepcontrol): how much of the kernel-to-kernel gap can PDL hide?
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pdl_bench pdl_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

template <bool PDL>
__global__ void __launch_bounds__(256) k(const float* __restrict__ in, float* __restrict__ out, int n, int work) {
  if (PDL) {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float v = in[i];
    for (int j = 0; j < work; ++j) v = fmaf(v, 1.0001f, 0.001f);
    out[i] = v;
  }
}

template <bool PDL>
float run(float* a, float* b, int n, int work, int chain, int grid, int reps) {
  cudaStream_t s;
  cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  cudaGraph_t g;
  cudaGraphExec_t e;
  cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
  for (int i = 0; i < chain; ++i) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(256);
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = PDL ? 1 : 0;
    const float* in = (i & 1) ? b : a;
    float* out = (i & 1) ? a : b;
    cudaLaunchKernelEx(&cfg, k<PDL>, in, out, n, work);
  }
  cudaStreamEndCapture(s, &g);
  cudaError_t err = cudaGraphInstantiate(&e, g, 0);
  if (err != cudaSuccess) { printf("instantiate: %s\n", cudaGetErrorString(err)); return -1; }
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 5; ++i) cudaGraphLaunch(e, s);
  cudaEventRecord(e0, s);
  for (int i = 0; i < reps; ++i) cudaGraphLaunch(e, s);
  cudaEventRecord(e1, s);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  err = cudaGetLastError();
  if (err != cudaSuccess) printf("error: %s\n", cudaGetErrorString(err));
  return ms * 1e3f / reps;
}

int main() {
  const int n = 1 << 20;
  float *a, *b;
  cudaMalloc(&a, n * 4); cudaMalloc(&b, n * 4);
  cudaMemset(a, 0, n * 4); cudaMemset(b, 0, n * 4);
  for (int work : {0, 200, 1000}) {
    for (int grid : {148, 592, 4096}) {
      float t0 = run<false>(a, b, n, work, 12, grid, 200);
      float t1 = run<true>(a, b, n, work, 12, grid, 200);
      printf("work %4d grid %4d: chain of 12  plain %7.1f us  PDL %7.1f us  saved per boundary %5.2f us\n", work, grid, t0, t1,
             (t0 - t1) / 12);
    }
  }
  // correctness: the chain result must be identical
  return 0;
}
