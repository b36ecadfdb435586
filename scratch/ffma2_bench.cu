// micro-benchmark: scalar FFMA vs packed FFMA2 issue/throughput on sm_100a
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float a, float b) {
  float2 acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  float2 w = make_float2(a, a), c = make_float2(b, b);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (MODE == 0) {
          acc[i].x = fmaf(acc[i].x, w.x, c.x);
          acc[i].y = fmaf(acc[i].y, w.y, c.y);
        } else {
          acc[i] = __ffma2_rn(acc[i], w, c);
        }
      }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int mode = 0; mode < 2; ++mode) for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    if (mode == 0) k<0><<<148 * 8, 256>>>(d, iters, 0.999f, 0.001f); else k<1><<<148 * 8, 256>>>(d, iters, 0.999f, 0.001f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double flops = 2.0 * 16 * 8 * (double)iters * 148 * 8 * 256;
    printf("mode %d: %.3f ms  %.2f TFLOP/s\n", mode, ms, flops / ms / 1e9);
  }
  return 0;
}
