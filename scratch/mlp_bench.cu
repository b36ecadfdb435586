// micro-benchmark: per-thread 16->16->16 residual-ReLU matvec chain (the ARM MLP shape) with the weights coming from
//   0: shared memory, scalar FMA, broadcast LDS.128         1: shared memory, FFMA2 packed over output pairs
//   2: constant memory operands, scalar FMA                  3: constant memory operands, FFMA2 over output pairs
// reports lane-FMA throughput; nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mlp_bench mlp_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#ifndef C
#define C 16
#endif
#ifndef NL
#define NL 2
#endif
__constant__ float cW[NL * C * C];  // [l][i][o]
__constant__ float cB[NL * C];

template <int MODE>
__global__ void __launch_bounds__(128) k(const float* __restrict__ gw, const float* __restrict__ gb, float* out,
                                         int iters) {
  __shared__ __align__(16) float sW[NL * C * C];
  __shared__ __align__(16) float sB[NL * C];
  for (int i = threadIdx.x; i < NL * C * C; i += blockDim.x) sW[i] = gw[i];
  for (int i = threadIdx.x; i < NL * C; i += blockDim.x) sB[i] = gb[i];
  __syncthreads();
  float h[C];
#pragma unroll
  for (int i = 0; i < C; ++i) h[i] = 0.01f * (threadIdx.x % 7) + 0.1f * i;
  for (int it = 0; it < iters; ++it) {
    asm volatile("" ::: "memory");
#pragma unroll
    for (int l = 0; l < NL; ++l) {
      if (MODE == 0 || MODE == 2) {
        float acc[C];
#pragma unroll
        for (int o = 0; o < C; ++o) acc[o] = (MODE == 0 ? sB[l * C + o] : cB[l * C + o]) + h[o];
#pragma unroll
        for (int i = 0; i < C; ++i) {
          if (MODE == 0) {
#pragma unroll
            for (int o4 = 0; o4 < C; o4 += 4) {
              float4 w = *reinterpret_cast<const float4*>(&sW[(l * C + i) * C + o4]);
              acc[o4] = fmaf(h[i], w.x, acc[o4]);
              acc[o4 + 1] = fmaf(h[i], w.y, acc[o4 + 1]);
              acc[o4 + 2] = fmaf(h[i], w.z, acc[o4 + 2]);
              acc[o4 + 3] = fmaf(h[i], w.w, acc[o4 + 3]);
            }
          } else {
#pragma unroll
            for (int o = 0; o < C; ++o) acc[o] = fmaf(h[i], cW[(l * C + i) * C + o], acc[o]);
          }
        }
#pragma unroll
        for (int o = 0; o < C; ++o) h[o] = fmaxf(acc[o], 0.f);
      } else {
        float2 acc[C / 2];
#pragma unroll
        for (int o = 0; o < C / 2; ++o) {
          float2 b = MODE == 1 ? *reinterpret_cast<const float2*>(&sB[l * C + 2 * o])
                               : *reinterpret_cast<const float2*>(&cB[l * C + 2 * o]);
          acc[o] = make_float2(b.x + h[2 * o], b.y + h[2 * o + 1]);
        }
#pragma unroll
        for (int i = 0; i < C; ++i) {
          float2 hh = make_float2(h[i], h[i]);
          if (MODE == 1) {
#pragma unroll
            for (int o4 = 0; o4 < C; o4 += 4) {
              float4 w = *reinterpret_cast<const float4*>(&sW[(l * C + i) * C + o4]);
              acc[o4 / 2] = __ffma2_rn(hh, make_float2(w.x, w.y), acc[o4 / 2]);
              acc[o4 / 2 + 1] = __ffma2_rn(hh, make_float2(w.z, w.w), acc[o4 / 2 + 1]);
            }
          } else {
#pragma unroll
            for (int o = 0; o < C / 2; ++o)
              acc[o] = __ffma2_rn(hh, *reinterpret_cast<const float2*>(&cW[(l * C + i) * C + 2 * o]), acc[o]);
          }
        }
#pragma unroll
        for (int o = 0; o < C / 2; ++o) {
          h[2 * o] = fmaxf(acc[o].x, 0.f);
          h[2 * o + 1] = fmaxf(acc[o].y, 0.f);
        }
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < C; ++i) s += h[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const float* gw, const float* gb, float* d, int ctas_per_sm, int iters) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  int grid = 148 * ctas_per_sm;
  k<MODE><<<grid, 128>>>(gw, gb, d, 10);
  cudaEventRecord(e0);
  k<MODE><<<grid, 128>>>(gw, gb, d, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  double flops = 2.0 * NL * C * C * (double)iters * grid * 128;
  printf("mode %d  warps/SM %2d: %.3f ms  %.2f TFLOP/s  (%s)\n", MODE, ctas_per_sm * 4, ms, flops / ms / 1e9,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  float hw[NL * C * C], hb[NL * C];
  for (int i = 0; i < NL * C * C; ++i) hw[i] = 0.05f * ((i * 37 % 19) - 9) / 9.f;
  for (int i = 0; i < NL * C; ++i) hb[i] = 0.01f * (i % 5);
  float *gw, *gb, *d;
  cudaMalloc(&gw, sizeof(hw));
  cudaMalloc(&gb, sizeof(hb));
  cudaMalloc(&d, 148 * 16 * 128 * 4);
  cudaMemcpy(gw, hw, sizeof(hw), cudaMemcpyHostToDevice);
  cudaMemcpy(gb, hb, sizeof(hb), cudaMemcpyHostToDevice);
  cudaMemcpyToSymbol(cW, hw, sizeof(hw));
  cudaMemcpyToSymbol(cB, hb, sizeof(hb));
  const int iters = 4000;
  for (int c : {2, 4}) {
    run<0>(gw, gb, d, c, iters);
    run<1>(gw, gb, d, c, iters);
    run<2>(gw, gb, d, c, iters);
    run<3>(gw, gb, d, c, iters);
  }
  return 0;
}
