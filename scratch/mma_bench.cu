// micro-benchmark: the ARM MLP shape (16->16->16 residual-ReLU chain per position) on the legacy tensor path
// (mma.sync.m16n8k8 tf32) with error-compensated operands ("3xTF32": a_hi*b_hi + a_lo*b_hi + a_hi*b_lo), against the
// packed-FP32 constant-bank form the shipped kernels use (scratch/mlp_bench.cu mode 3).
//   MODE 0: 3xTF32 (fp32-grade products)      MODE 1: 1xTF32 (upper bound of the tensor path; NOT parity-safe)
// A warp iteration = 32 positions (two 16-row tiles).  The D fragment of layer l is reused as the A fragment of layer
// l+1 by permuting the contraction index (slot t <-> feature 2t, slot t+4 <-> feature 2t+1), so no shuffles are needed.
// Also checks the numerical error of the 3xTF32 chain against a double-precision evaluation.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mma_bench mma_bench.cu
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define C 16
#define NL 2

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void split(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xffffe000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}

// gw: [l][i][o] fp32, gb: [l][o].  x: [n_pos][16] input (only used by the check), out: [n_pos][16]
template <int MODE, bool CHECK>
__global__ void __launch_bounds__(128) k(const float* __restrict__ gw, const float* __restrict__ gb,
                                         const float* __restrict__ x, float* out, int iters) {
  const int lane = threadIdx.x & 31, gid = lane >> 2, tig = lane & 3;
  // B fragments: layer l, k-step ks, n-tile nt: b0 = W[k = slot tig][n = gid], b1 = W[k = slot tig+4][n = gid]
  // slot s of k-step ks <-> feature ks*8 + 2*(s%4) + s/4
  uint32_t bh[NL][2][2][2], bl[NL][2][2][2];
  float bias[NL][2][2];
#pragma unroll
  for (int l = 0; l < NL; ++l)
#pragma unroll
    for (int ks = 0; ks < 2; ++ks)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          int feat = ks * 8 + 2 * tig + r;
          split(gw[(l * C + feat) * C + nt * 8 + gid], bh[l][ks][nt][r], bl[l][ks][nt][r]);
        }
      }
#pragma unroll
  for (int l = 0; l < NL; ++l)
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      bias[l][nt][0] = gb[l * C + nt * 8 + 2 * tig];
      bias[l][nt][1] = gb[l * C + nt * 8 + 2 * tig + 1];
    }
  // activations of two 16-position tiles in "D layout": h[mt][nt][c]: c0,c1 = row gid, features nt*8+2tig+{0,1};
  // c2,c3 = row gid+8
  float h[2][2][4];
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        int row = gid + (c >> 1) * 8, feat = nt * 8 + 2 * tig + (c & 1);
        if (CHECK) h[mt][nt][c] = x[((size_t)warp_global * 32 + mt * 16 + row) * C + feat];
        else h[mt][nt][c] = 0.01f * (lane % 7) + 0.1f * feat;
      }
  for (int it = 0; it < iters; ++it) {
    asm volatile("" ::: "memory");
#pragma unroll
    for (int l = 0; l < NL; ++l) {
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        uint32_t ah[2][4], al[2][4];
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
          // A(ks): a0 = (gid, slot tig) = D c0, a1 = (gid+8, slot tig) = D c2, a2 = (gid, slot tig+4) = D c1, a3 = D c3
          const int map[4] = {0, 2, 1, 3};
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            if (MODE == 0) split(h[mt][ks][map[r]], ah[ks][r], al[ks][r]);
            else ah[ks][r] = __float_as_uint(h[mt][ks][map[r]]);
          }
        }
        float acc[2][4];
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[nt][c] = h[mt][nt][c] + bias[l][nt][c & 1];
#pragma unroll
        for (int ks = 0; ks < 2; ++ks)
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            if (MODE == 0) {
              mma_tf32(acc[nt], al[ks], bh[l][ks][nt]);
              mma_tf32(acc[nt], ah[ks], bl[l][ks][nt]);
            }
            mma_tf32(acc[nt], ah[ks], bh[l][ks][nt]);
          }
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int c = 0; c < 4; ++c) h[mt][nt][c] = fmaxf(acc[nt][c], 0.f);
      }
    }
  }
  if (CHECK) {
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          int row = gid + (c >> 1) * 8, feat = nt * 8 + 2 * tig + (c & 1);
          out[((size_t)warp_global * 32 + mt * 16 + row) * C + feat] = h[mt][nt][c];
        }
  } else {
    float s = 0;
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int c = 0; c < 4; ++c) s += h[mt][nt][c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  }
}

template <int MODE>
void run(const float* gw, const float* gb, float* d, int ctas_per_sm, int iters) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  int grid = 148 * ctas_per_sm;
  k<MODE, false><<<grid, 128>>>(gw, gb, nullptr, d, 10);
  cudaEventRecord(e0);
  k<MODE, false><<<grid, 128>>>(gw, gb, nullptr, d, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  // every thread-iteration covers ONE position's worth of work (a warp covers 32 positions)
  double flops = 2.0 * NL * C * C * (double)iters * grid * 128;
  printf("mma mode %d (%s)  warps/SM %2d: %.3f ms  %.2f effective fp32 TFLOP/s  (%s)\n", MODE,
         MODE == 0 ? "3xTF32" : "1xTF32", ctas_per_sm * 4, ms, flops / ms / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  float hw[NL * C * C], hb[NL * C];
  srand(1);
  for (int i = 0; i < NL * C * C; ++i) hw[i] = 0.5f * (rand() / (float)RAND_MAX - 0.5f);
  for (int i = 0; i < NL * C; ++i) hb[i] = 0.2f * (rand() / (float)RAND_MAX - 0.5f);
  float *gw, *gb, *d, *x;
  const int n_pos = 148 * 4 * 128;
  cudaMalloc(&gw, sizeof(hw));
  cudaMalloc(&gb, sizeof(hb));
  cudaMalloc(&d, (size_t)n_pos * C * 4);
  cudaMalloc(&x, (size_t)n_pos * C * 4);
  cudaMemcpy(gw, hw, sizeof(hw), cudaMemcpyHostToDevice);
  cudaMemcpy(gb, hb, sizeof(hb), cudaMemcpyHostToDevice);
  // numerical check: one pass of the 2-layer chain on random inputs against double precision
  float* hx = (float*)malloc((size_t)n_pos * C * 4);
  float* ho = (float*)malloc((size_t)n_pos * C * 4);
  for (size_t i = 0; i < (size_t)n_pos * C; ++i) hx[i] = 8.f * (rand() / (float)RAND_MAX - 0.5f);
  cudaMemcpy(x, hx, (size_t)n_pos * C * 4, cudaMemcpyHostToDevice);
  for (int mode = 0; mode < 2; ++mode) {
    if (mode == 0) k<0, true><<<148 * 4, 128>>>(gw, gb, x, d, 1);
    else k<1, true><<<148 * 4, 128>>>(gw, gb, x, d, 1);
    cudaMemcpy(ho, d, (size_t)n_pos * C * 4, cudaMemcpyDeviceToHost);
    double worst = 0, worst32 = 0, vmax = 0;
    for (int p = 0; p < 4096; ++p) {
      double h[C];
      float h32[C];
      for (int i = 0; i < C; ++i) h[i] = hx[(size_t)p * C + i], h32[i] = hx[(size_t)p * C + i];
      for (int l = 0; l < NL; ++l) {
        double a[C];
        float a32[C];
        for (int o = 0; o < C; ++o) {
          a[o] = hb[l * C + o] + h[o];
          a32[o] = hb[l * C + o] + h32[o];
          for (int i = 0; i < C; ++i) a[o] += h[i] * hw[(l * C + i) * C + o], a32[o] = fmaf(h32[i], hw[(l * C + i) * C + o], a32[o]);
        }
        for (int o = 0; o < C; ++o) h[o] = a[o] > 0 ? a[o] : 0, h32[o] = a32[o] > 0 ? a32[o] : 0;
      }
      for (int o = 0; o < C; ++o) {
        worst = fmax(worst, fabs(ho[(size_t)p * C + o] - h[o]));
        worst32 = fmax(worst32, fabs(h32[o] - h[o]));
        vmax = fmax(vmax, fabs(h[o]));
      }
    }
    printf("check mode %d: max|mma - f64| = %.3e, max|fp32 fma - f64| = %.3e, max|value| = %.3f  (%s)\n", mode, worst, worst32,
           vmax, cudaGetErrorString(cudaGetLastError()));
  }
  const int iters = 4000;
  for (int c : {2, 4, 8}) {
    run<0>(gw, gb, d, c, iters);
    run<1>(gw, gb, d, c, iters);
  }
  return 0;
}
