// scratch/umma_bench.cu -- prototype: the ARM MLP of the hot path on the 5th-generation tensor cores.
//
// The network is the one k_arm_bwd (csrc/ccb_arm.cu) evaluates per latent position for the HOP / c3x architecture
// (enc/component/core/arm.py:86-101, 227-268): 16 context values -> two residual 16x16 layers with ReLU -> (mu,
// log-scale); the overfitting step needs its forward, the gradient w.r.t. the 16 context values and the weight
// gradients summed over all positions: 3 x 544 = 1632 multiply-adds per position.  The production kernel does all of
// it with packed FP32 FMAs (FFMA2, one thread per position, ~2500 thread instructions per position).  Here the
// contractions are tcgen05.mma.kind::tf32 with accumulators in TMEM:
//
//   * 128 positions = the M rows of one MMA.  A thread (= position) writes its 16 activations into shared memory as
//     the no-swizzle K-major operand image [8-position group][16-byte chunk of 4 features][position], once as the TF32
//     "hi" part (low 13 mantissa bits cleared) and once as the exact remainder "lo".  The weight image stacks
//     [W_hi ; W_lo] as N = 32 rows, so  X_hi [W_hi;W_lo]^T  and  X_lo [W_hi;W_lo]^T  accumulate all four partial
//     products in 32 TMEM columns (the thread adds column j and 16 + j): error-compensated TF32 = fp32-level results;
//   * weight gradients: M = features, K = positions.  scratch/umma_probe.cu shows that kind::tf32 returns zeros for
//     MN-major no-swizzle operands (CUTLASS: "for mn-major tf32 operands, SW128_32B is the only available smem
//     layout", a 4-byte swizzle that cannot alias the K-major image), so every activation is ALSO stored feature-major
//     [feature][position] (K-major again, 144-byte chunk pitch = conflict-free 4-byte stores).  Rows are stacked
//     [x0_hi; x0_lo; x1_hi; x1_lo] (M) against [dpre0_hi; dpre0_lo; dpre1_hi; dpre1_lo] (N = 64): ONE MMA per 8
//     positions accumulates both layers' d W^T (all hi/lo products) in a 64 x 64 TMEM block over ALL tiles of the CTA;
//     it is read once, at the end.  The 2 x 16 output layer and the biases are summed in registers;
//   * a dedicated warp issues the MMAs (one lane), the position warps convert layer outputs (tcgen05.ld -> bias +
//     residual + ReLU -> hi / lo -> st.shared) and hand over through mbarriers; a layer's K-major image is dead once
//     its MMA completed, so ONE transient region serves all five of them.
//
// Checked against a float64 CPU evaluation; timed with CUDA events.  Build / run (B200):
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o scratch/umma_bench scratch/umma_bench.cu
//   timeout 120 scratch/umma_bench [n_positions]
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#ifndef GROUP_LAG_NS
#define GROUP_LAG_NS 400
#endif
constexpr int C = 16;
constexpr int LAYER = C * C + C;            // one hidden layer: weights + biases
constexpr int NP = 2 * LAYER + 2 * C + 2;   // 578 parameters: [w0 b0 w1 b1 wo bo]
constexpr int NPA = (NP + 3) & ~3;
constexpr int GROUP = 128;                  // positions per MMA block (M)
constexpr int NG = 2;                       // position groups per CTA
constexpr int CTA_POS = GROUP * NG;
constexpr int THREADS = CTA_POS + 32 * NG;  // + one MMA warp per group
constexpr int CHUNK = 128;                  // one core matrix: 8 rows x 16 bytes
// K-major transient image [position][feature]: 8-position group = 4 feature chunks
constexpr int KGRP = 4 * CHUNK;                 // 512
constexpr int KIMG = (GROUP / 8) * KGRP;        // 8192 (hi or lo)
// feature-major images [feature][position]: a chunk = 8 features x 4 positions; chunk pitch 144 bytes makes the 32
// lanes' 4-byte stores of one feature hit 32 different banks; 8-feature row group = 32 chunks
constexpr int TLBO = CHUNK + 16;                // 144
constexpr int TSBO = (GROUP / 4) * TLBO;        // 4608
constexpr int TIMG = 8 * TSBO;                  // 8 row groups: [a_hi(2) a_lo(2) b_hi(2) b_lo(2)] = 36864
constexpr int OFF_K = 0;                        // per group: K-major hi | lo
constexpr int OFF_TA = 2 * KIMG;                // x0 / x1 (the M side of the weight-gradient MMA)
constexpr int OFF_TB = OFF_TA + TIMG;           // dpre0 / dpre1 (the N side)
constexpr int GROUP_BYTES = OFF_TB + TIMG;      // 90112
constexpr int WIMG = 2 * 16 * 16 * 4;           // [W_hi ; W_lo]: 32 rows x 16 (K-major)
constexpr int WOIMG = 8 * 16 * 4;               // output layer: rows mu_hi, ls_hi, mu_lo, ls_lo, 4 x zero
constexpr int OFF_W = NG * GROUP_BYTES;         // W0, W1, W0^T, W1^T, Wo
constexpr int OFF_WO = OFF_W + 4 * WIMG;
constexpr int OFF_PAR = OFF_WO + WOIMG;         // plain fp32 copy of the parameters
constexpr int OFF_BAR = OFF_PAR + NPA * 4;      // mbarriers: per group ready, act, done
constexpr int OFF_TMEM = OFF_BAR + 3 * NG * 8;
constexpr int OFF_RED = OFF_TMEM + 16;          // [8 warps][68] register-accumulated sums
constexpr int SMEM_BYTES = OFF_RED + 8 * 68 * 4;
constexpr int TMEM_COLS = 256;                  // per group: 32 (layer result) + 64 (d W^T block), 128-column stride
// M = 128 rows of the weight-gradient MMA = 16 row groups from T_A: groups 8..15 are whatever follows (rows nobody reads)
static_assert(OFF_TA + 16 * TSBO <= SMEM_BYTES - (NG - 1) * GROUP_BYTES, "over-read of the M = 128 operand");

__device__ unsigned g_error;
#ifdef TRACE
__device__ long long g_trace[64];
#define TR(i) do { if (blockIdx.x == 0 && it == 2 && (threadIdx.x & 31) == 0) g_trace[i] = clock64(); } while (0)
#else
#define TR(i) do {} while (0)
#endif  // a bounded wait gave up (the run is invalid, but it ends)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, unsigned parity) {
  for (unsigned spin = 0;; ++spin) {
    unsigned ok;
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
    if ((spin & 1023u) == 1023u && (spin > (1u << 20) || *(volatile unsigned*)&g_error)) {
      g_error = 1;
      return;
    }
  }
}
// One lane of a CONVERGED warp (the operands must be warp-uniform: inside a divergent branch ptxas has to move them to
// uniform registers with a waterfall loop per instruction, ~80 cycles per MMA -- measured with the first version)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n .reg .b32 rx;\n .reg .pred px;\n elect.sync rx|px, 0xffffffff;\n selp.u32 %0, 1, 0, px;\n}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  if (elect_one())
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// shared-memory matrix descriptor, no swizzle (cute/arch/mma_sm100_desc.hpp: SmemDescriptor): start >> 4 in [0,14),
// leading-dimension byte offset >> 4 in [16,30), stride-dimension byte offset >> 4 in [32,46), version 1 in [46,48)
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// instruction descriptor (InstrDescriptor): D = f32, A = B = tf32, majors, N >> 3, M >> 4
constexpr uint32_t make_idesc(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  if (elect_one())
  asm volatile(
      "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(d),
      "l"(da), "l"(db), "r"(idesc), "r"(acc)
      : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t addr, float* v) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, "
      "%18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]),
        "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]),
        "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(addr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld8(uint32_t addr, float* v) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(addr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

// one position's 16 values -> its row of the K-major image (hi, lo) and, with T, rows rg0.. of a feature-major image
template <bool T>
__device__ __forceinline__ void write_act(char* k_hi, char* t_img, int rg0, int row, const float* v) {
  char* p = k_hi + (row >> 3) * KGRP + (row & 7) * 16;
  char* q = t_img + rg0 * TSBO + (row >> 2) * TLBO + (row & 3) * 4;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      h[e] = tf32_hi(v[4 * c + e]);
      l[e] = v[4 * c + e] - h[e];
      if (T) {
        const int i = 4 * c + e;
        *reinterpret_cast<float*>(q + (i >> 3) * TSBO + (i & 7) * 16) = h[e];
        *reinterpret_cast<float*>(q + (2 + (i >> 3)) * TSBO + (i & 7) * 16) = l[e];
      }
    }
    *reinterpret_cast<float4*>(p + c * CHUNK) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(p + c * CHUNK + KIMG) = make_float4(l[0], l[1], l[2], l[3]);
  }
}

constexpr uint32_t ID_N32 = make_idesc(128, 32, 0, 0);
constexpr uint32_t ID_N8 = make_idesc(128, 8, 0, 0);
constexpr uint32_t ID_N64 = make_idesc(128, 64, 0, 0);

// D[128 x N] (+)= (A_hi + A_lo) B^T over K = 16: B stacks its hi and lo rows along N
__device__ __forceinline__ void layer_mma(uint32_t d, uint32_t a_hi, uint32_t b, uint32_t idesc) {
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    const uint64_t db = make_desc(b + s * 2 * CHUNK, CHUNK, 4 * CHUNK);
    umma(d, make_desc(a_hi + s * 2 * CHUNK, CHUNK, KGRP), db, idesc, s);
    umma(d, make_desc(a_hi + KIMG + s * 2 * CHUNK, CHUNK, KGRP), db, idesc, 1u);
  }
}

__global__ void __launch_bounds__(THREADS, 1)
k_arm_mlp_tc(const float* __restrict__ params, const float* __restrict__ x_in, int n_tiles, float* __restrict__ mu_ls,
             float* __restrict__ dx_out, float* __restrict__ part) {
  extern __shared__ __align__(1024) char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s0 = smem_u32(smem);
  float* sp = reinterpret_cast<float*>(smem + OFF_PAR);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + OFF_TMEM);
  auto bar = [&](int g, int which) { return s0 + OFF_BAR + (g * 3 + which) * 8; };  // 0 ready, 1 act, 2 done

  // ---- set-up: zero the images, weight images, parameters, barriers, TMEM
  for (int i = tid; i < OFF_PAR / 16; i += THREADS) reinterpret_cast<float4*>(smem)[i] = make_float4(0, 0, 0, 0);
  for (int i = tid; i < NP; i += THREADS) sp[i] = params[i];
  __syncthreads();
  for (int e = tid; e < 2 * C * C; e += THREADS) {
    const int k = e / (C * C), j = (e / C) % C, i = e % C;
    const float w = sp[k * LAYER + j * C + i], h = tf32_hi(w);
    // forward image: row n = j (+16 for lo), k = i;  transposed image (input gradient): row n = i, k = j
    char* f = smem + OFF_W + k * WIMG + (j >> 3) * 4 * CHUNK + (i >> 2) * CHUNK + (j & 7) * 16 + (i & 3) * 4;
    char* t = smem + OFF_W + (2 + k) * WIMG + (i >> 3) * 4 * CHUNK + (j >> 2) * CHUNK + (i & 7) * 16 + (j & 3) * 4;
    *reinterpret_cast<float*>(f) = h;
    *reinterpret_cast<float*>(f + WIMG / 2) = w - h;
    *reinterpret_cast<float*>(t) = h;
    *reinterpret_cast<float*>(t + WIMG / 2) = w - h;
  }
  for (int e = tid; e < 2 * C; e += THREADS) {  // output layer: rows 0,1 = hi of (mu, ls), rows 2,3 = lo
    const int j = e / C, i = e % C;
    const float w = sp[2 * LAYER + j * C + i], h = tf32_hi(w);
    char* img = smem + OFF_WO + (i >> 2) * CHUNK + j * 16 + (i & 3) * 4;
    *reinterpret_cast<float*>(img) = h;
    *reinterpret_cast<float*>(img + 2 * 16) = w - h;
  }
  if (tid == 0) {
    for (int g = 0; g < NG; ++g) {
      mbar_init(bar(g, 0), GROUP);
      mbar_init(bar(g, 1), 1);
      mbar_init(bar(g, 2), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == NG * 4) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  fence_async_smem();  // the weight images were written through the generic proxy
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *s_tmem;

  int n_my_tiles = 0;
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) ++n_my_tiles;

  if (warp >= NG * 4) {
    // =========================== MMA warps: warp NG*4 + g serves group g, all lanes converged ===========================
    const int g = warp - NG * 4;
    const uint32_t R = s0 + g * GROUP_BYTES, d_act = tmem + g * 128, d_w = d_act + 32;
    unsigned n_ready = 0;
    for (int it = 0; it < n_my_tiles; ++it) {
      for (int ph = 0; ph < 4; ++ph) {
        mbar_wait(bar(g, 0), n_ready++ & 1);
        if (g == 0) TR(32 + ph * 2);
        tc_fence_after();
        // 0, 1: forward through W0, W1; 2, 3: input gradient through W1^T, W0^T
        layer_mma(d_act, R + OFF_K, s0 + OFF_W + (ph == 2 ? 3 : ph == 3 ? 2 : ph) * WIMG, ID_N32);
        tc_commit(bar(g, 1));
        if (g == 0) TR(33 + ph * 2);
        if (ph == 3) {
          // both layers' weight gradients: K = the 128 positions of the tile, 8 (two 4-position chunks) per MMA
#pragma unroll 4
          for (int s = 0; s < GROUP / 8; ++s)
            umma(d_w, make_desc(R + OFF_TA + s * 2 * TLBO, TLBO, TSBO), make_desc(R + OFF_TB + s * 2 * TLBO, TLBO, TSBO),
                 ID_N64, (it > 0 || s > 0) ? 1u : 0u);
          tc_commit(bar(g, 2));
        }
      }
    }
  } else {
    // =========================== position warps ===========================
    const int g = tid / GROUP, row = tid % GROUP;
    char* R = smem + g * GROUP_BYTES;
    const uint32_t t_act = tmem + g * 128 + ((uint32_t)((warp & 3) * 32) << 16);
    unsigned n_act = 0, n_done = 0;
    if (g == 1) __nanosleep(GROUP_LAG_NS);
    float g_wo[2][C], g_b0[C], g_b1[C], g_bo[2] = {0.f, 0.f};  // sums this thread keeps over all its tiles
#pragma unroll
    for (int i = 0; i < C; ++i) g_wo[0][i] = g_wo[1][i] = g_b0[i] = g_b1[i] = 0.f;
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const long long pos = (long long)tile * CTA_POS + tid;
      float x0[C], x1[C], x2[C], a[2 * C];
#pragma unroll
      for (int i = 0; i < C; i += 4) {
        const float4 v = *reinterpret_cast<const float4*>(x_in + pos * C + i);
        x0[i] = v.x; x0[i + 1] = v.y; x0[i + 2] = v.z; x0[i + 3] = v.w;
      }
      if (it > 0) mbar_wait(bar(g, 2), n_done++ & 1);  // the previous tile's weight-gradient MMAs have read T_A / T_B
      // ---- x0 -> layer 0
      write_act<true>(R + OFF_K, R + OFF_TA, 0, row, x0);
      fence_async_smem();
      if (tid == 0) TR(0);
      mbar_arrive(bar(g, 0));
      mbar_wait(bar(g, 1), n_act++ & 1);
      if (tid == 0) TR(1);
      tc_fence_after();
      tmem_ld32(t_act, a);
      if (tid == 0) TR(2);
#pragma unroll
      for (int j = 0; j < C; ++j) x1[j] = fmaxf(a[j] + a[C + j] + sp[C * C + j] + x0[j], 0.f);
      // ---- x1 -> layer 1
      write_act<true>(R + OFF_K, R + OFF_TA, 4, row, x1);
      fence_async_smem();
      tc_fence_before();
      if (tid == 0) TR(4);
      mbar_arrive(bar(g, 0));
      mbar_wait(bar(g, 1), n_act++ & 1);
      if (tid == 0) TR(5);
      tc_fence_after();
      tmem_ld32(t_act, a);
      if (tid == 0) TR(6);
#pragma unroll
      for (int j = 0; j < C; ++j) x2[j] = fmaxf(a[j] + a[C + j] + sp[LAYER + C * C + j] + x1[j], 0.f);
      // ---- x2 -> (mu, log-scale): 2 x 16 multiply-adds per position, cheaper on the CUDA cores than a round trip
      float mu = sp[2 * LAYER + 2 * C], ls = sp[2 * LAYER + 2 * C + 1];
#pragma unroll
      for (int i = 0; i < C; ++i) {
        mu = fmaf(sp[2 * LAYER + i], x2[i], mu);
        ls = fmaf(sp[2 * LAYER + C + i], x2[i], ls);
      }
      *reinterpret_cast<float2*>(mu_ls + pos * 2) = make_float2(mu, ls);
      // stand-in for the Laplace rate backward (smooth, position dependent)
      const float d_mu = 0.01f * mu + 0.003f, d_ls = -0.02f * ls + 0.001f;
      // ---- output layer backward in registers: d Wo, d bo, d x2 -> dpre1
      float dp[C];
      g_bo[0] += d_mu;
      g_bo[1] += d_ls;
#pragma unroll
      for (int i = 0; i < C; ++i) {
        g_wo[0][i] = fmaf(d_mu, x2[i], g_wo[0][i]);
        g_wo[1][i] = fmaf(d_ls, x2[i], g_wo[1][i]);
        const float dxi = sp[2 * LAYER + i] * d_mu + sp[2 * LAYER + C + i] * d_ls;
        dp[i] = x2[i] > 0.f ? dxi : 0.f;
        g_b1[i] += dp[i];
      }
      write_act<true>(R + OFF_K, R + OFF_TB, 4, row, dp);
      fence_async_smem();
      tc_fence_before();
      if (tid == 0) TR(8);
      mbar_arrive(bar(g, 0));
      mbar_wait(bar(g, 1), n_act++ & 1);
      if (tid == 0) TR(9);
      tc_fence_after();
      tmem_ld32(t_act, a);
      if (tid == 0) TR(10);
      // ---- dpre0 = relu'(x1) (W1^T dpre1 + dpre1)
#pragma unroll
      for (int i = 0; i < C; ++i) {
        dp[i] = x1[i] > 0.f ? a[i] + a[C + i] + dp[i] : 0.f;
        g_b0[i] += dp[i];
      }
      write_act<true>(R + OFF_K, R + OFF_TB, 0, row, dp);
      fence_async_smem();
      tc_fence_before();
      if (tid == 0) TR(12);
      mbar_arrive(bar(g, 0));
      mbar_wait(bar(g, 1), n_act++ & 1);
      if (tid == 0) TR(13);
      tc_fence_after();
      tmem_ld32(t_act, a);
      if (tid == 0) TR(14);
#pragma unroll
      for (int i = 0; i < C; i += 4)
        *reinterpret_cast<float4*>(dx_out + pos * C + i) =
            make_float4(a[i] + a[C + i] + dp[i], a[i + 1] + a[C + i + 1] + dp[i + 1], a[i + 2] + a[C + i + 2] + dp[i + 2],
                        a[i + 3] + a[C + i + 3] + dp[i + 3]);
      tc_fence_before();
    }
    // ---- fold: register sums -> one row per warp; the 64 x 64 TMEM block -> shared memory (T_A of this group is free)
    float* red = reinterpret_cast<float*>(smem + OFF_RED) + warp * 68;
#pragma unroll
    for (int i = 0; i < C; ++i) {
      float v0 = g_wo[0][i], v1 = g_wo[1][i], v2 = g_b0[i], v3 = g_b1[i];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        v0 += __shfl_xor_sync(0xffffffffu, v0, o);
        v1 += __shfl_xor_sync(0xffffffffu, v1, o);
        v2 += __shfl_xor_sync(0xffffffffu, v2, o);
        v3 += __shfl_xor_sync(0xffffffffu, v3, o);
      }
      if (lane == 0) {
        red[i] = v0; red[C + i] = v1; red[2 * C + i] = v2; red[3 * C + i] = v3;
      }
    }
    {
      float v0 = g_bo[0], v1 = g_bo[1];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        v0 += __shfl_xor_sync(0xffffffffu, v0, o);
        v1 += __shfl_xor_sync(0xffffffffu, v1, o);
      }
      if (lane == 0) { red[4 * C] = v0; red[4 * C + 1] = v1; }
    }
    if (it > 0) mbar_wait(bar(g, 2), n_done++ & 1);
    tc_fence_after();
    if ((warp & 3) < 2) {  // TMEM lanes 0..63 = rows [x0_hi x0_lo x1_hi x1_lo] of d W^T, 64 columns
      float* blk = reinterpret_cast<float*>(R + OFF_TA);  // [64][65]
      float v[32];
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        tmem_ld32(tmem + g * 128 + 32 + half * 32 + ((uint32_t)((warp & 3) * 32) << 16), v);
#pragma unroll
        for (int j = 0; j < 32; ++j) blk[((warp & 3) * 32 + lane) * 65 + half * 32 + j] = it > 0 ? v[j] : 0.f;
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  {
    float* prow = part + (long long)blockIdx.x * NP;
    const float* red = reinterpret_cast<const float*>(smem + OFF_RED);
    for (int e = tid; e < NP; e += THREADS) {
      float s = 0.f;
      if (e < 2 * LAYER && (e % LAYER) < C * C) {  // d W_k[j][i] = sum of the four hi/lo blocks of rows i, columns j
        const int k = e / LAYER, j = (e % LAYER) / C, i = e % C;
        for (int g = 0; g < NG; ++g) {
          const float* blk = reinterpret_cast<const float*>(smem + g * GROUP_BYTES + OFF_TA);
          for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 2; ++b) s += blk[(k * 32 + a * 16 + i) * 65 + k * 32 + b * 16 + j];
        }
      } else {
        int col;  // column of the per-warp rows: [d wo (32) | d b0 (16) | d b1 (16) | d bo (2)]
        if (e < 2 * LAYER) col = 2 * C + (e / LAYER) * C + (e % LAYER) - C * C;
        else if (e < 2 * LAYER + 2 * C) col = e - 2 * LAYER;
        else col = 4 * C + (e - 2 * LAYER - 2 * C);
        for (int w = 0; w < 8; ++w) s += red[w * 68 + col];
      }
      prow[e] = s;
    }
  }
  tc_fence_after();
  __syncthreads();
  if (warp == NG * 4)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
}

// ------------------------------------------------------------------------------------------------------------
#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      return 1;                                                                        \
    }                                                                                  \
  } while (0)

static float frand(uint32_t& s) {
  s = s * 1664525u + 1013904223u;
  return (float)((s >> 8) & 0xffffff) / 16777216.0f * 2.0f - 1.0f;
}

int main(int argc, char** argv) {
  long long n_pos = argc > 1 ? atoll(argv[1]) : 524288;  // latent positions of a 768x512 image: 524 256
  n_pos = (n_pos + CTA_POS - 1) / CTA_POS * CTA_POS;
  const int n_tiles = (int)(n_pos / CTA_POS);
  int dev = 0, n_sm = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  const int grid = n_tiles < n_sm ? n_tiles : n_sm;
  uint32_t seed = 12345;
  std::vector<float> h_par(NP), h_x((size_t)n_pos * C);
  for (auto& v : h_par) v = 0.25f * frand(seed);
  for (auto& v : h_x) v = 2.0f * frand(seed);
  float *d_par, *d_x, *d_mu, *d_dx, *d_part;
  CK(cudaMalloc(&d_par, NP * 4));
  CK(cudaMalloc(&d_x, (size_t)n_pos * C * 4));
  CK(cudaMalloc(&d_mu, (size_t)n_pos * 2 * 4));
  CK(cudaMalloc(&d_dx, (size_t)n_pos * C * 4));
  CK(cudaMalloc(&d_part, (size_t)grid * NP * 4));
  CK(cudaMemcpy(d_par, h_par.data(), NP * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_x, h_x.data(), (size_t)n_pos * C * 4, cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute(k_arm_mlp_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  printf("positions %lld, tiles %d, grid %d x %d threads, %d bytes of shared memory, %d TMEM columns\n", n_pos, n_tiles, grid,
         THREADS, SMEM_BYTES, TMEM_COLS);
  k_arm_mlp_tc<<<grid, THREADS, SMEM_BYTES>>>(d_par, d_x, n_tiles, d_mu, d_dx, d_part);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  unsigned err = 0;
  CK(cudaMemcpyFromSymbol(&err, g_error, sizeof(err)));
  if (err) printf("!! a bounded mbarrier wait gave up: results invalid\n");
#ifdef TRACE
  {
    long long tr[64];
    CK(cudaMemcpyFromSymbol(tr, g_trace, sizeof(tr)));
    const long long t0 = tr[0];
    printf("trace (cycles since the first arrive of tile iteration 2, CTA 0, group 0):\n");
    for (int ph = 0; ph < 4; ++ph)
      printf("  phase %d: arrive %6lld | mma woke %6lld, issued %6lld | act woke %6lld, ld done %6lld\n", ph, tr[ph * 4] - t0,
             tr[32 + ph * 2] - t0, tr[33 + ph * 2] - t0, tr[ph * 4 + 1] - t0, tr[ph * 4 + 2] - t0);
  }
#endif

  // ---- float64 reference
  std::vector<float> h_mu((size_t)n_pos * 2), h_dx((size_t)n_pos * C), h_part((size_t)grid * NP);
  CK(cudaMemcpy(h_mu.data(), d_mu, h_mu.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(h_dx.data(), d_dx, h_dx.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(h_part.data(), d_part, h_part.size() * 4, cudaMemcpyDeviceToHost));
  std::vector<double> gw(NP, 0.0);
  double e_mu = 0, e_dx = 0, m_mu = 0, m_dx = 0;
  const double* nil = nullptr;
  (void)nil;
  for (long long p = 0; p < n_pos; ++p) {
    double x[3][C], pre;
    for (int i = 0; i < C; ++i) x[0][i] = h_x[p * C + i];
    for (int k = 0; k < 2; ++k)
      for (int j = 0; j < C; ++j) {
        pre = h_par[k * LAYER + C * C + j] + x[k][j];
        for (int i = 0; i < C; ++i) pre += (double)h_par[k * LAYER + j * C + i] * x[k][i];
        x[k + 1][j] = pre > 0 ? pre : 0;
      }
    double out[2];
    for (int o = 0; o < 2; ++o) {
      out[o] = h_par[2 * LAYER + 2 * C + o];
      for (int i = 0; i < C; ++i) out[o] += (double)h_par[2 * LAYER + o * C + i] * x[2][i];
    }
    const double dmu = 0.01 * out[0] + 0.003, dls = -0.02 * out[1] + 0.001;
    double dx[C], dpre[C];
    for (int i = 0; i < C; ++i) {
      dx[i] = h_par[2 * LAYER + i] * dmu + h_par[2 * LAYER + C + i] * dls;
      gw[2 * LAYER + i] += dmu * x[2][i];
      gw[2 * LAYER + C + i] += dls * x[2][i];
    }
    gw[2 * LAYER + 2 * C] += dmu;
    gw[2 * LAYER + 2 * C + 1] += dls;
    for (int k = 1; k >= 0; --k) {
      for (int j = 0; j < C; ++j) dpre[j] = x[k + 1][j] > 0 ? dx[j] : 0;
      for (int j = 0; j < C; ++j) {
        gw[k * LAYER + C * C + j] += dpre[j];
        for (int i = 0; i < C; ++i) gw[k * LAYER + j * C + i] += dpre[j] * x[k][i];
      }
      for (int i = 0; i < C; ++i) {
        double s = dpre[i];
        for (int j = 0; j < C; ++j) s += (double)h_par[k * LAYER + j * C + i] * dpre[j];
        dx[i] = s;
      }
    }
    for (int o = 0; o < 2; ++o) {
      e_mu = fmax(e_mu, fabs(out[o] - h_mu[p * 2 + o]));
      m_mu = fmax(m_mu, fabs(out[o]));
    }
    for (int i = 0; i < C; ++i) {
      e_dx = fmax(e_dx, fabs(dx[i] - h_dx[p * C + i]));
      m_dx = fmax(m_dx, fabs(dx[i]));
    }
  }
  double e_w = 0, m_w = 0;
  for (int e = 0; e < NP; ++e) {
    double s = 0;
    for (int b = 0; b < grid; ++b) s += h_part[(size_t)b * NP + e];
    e_w = fmax(e_w, fabs(s - gw[e]));
    m_w = fmax(m_w, fabs(gw[e]));
  }
  printf("max |err|: (mu, ls) %.3e of %.3e   d x0 %.3e of %.3e   d W %.3e of %.3e\n", e_mu, m_mu, e_dx, m_dx, e_w, m_w);
  const bool ok = !err && e_mu <= 2e-5 * m_mu && e_dx <= 2e-5 * m_dx && e_w <= 2e-5 * m_w;
  printf("%s\n", ok ? "PARITY OK (<= 2e-5 of the largest value)" : "PARITY FAILED");

  // ---- timing
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  const int reps = 20;
  for (int i = 0; i < 3; ++i) k_arm_mlp_tc<<<grid, THREADS, SMEM_BYTES>>>(d_par, d_x, n_tiles, d_mu, d_dx, d_part);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int i = 0; i < reps; ++i) k_arm_mlp_tc<<<grid, THREADS, SMEM_BYTES>>>(d_par, d_x, n_tiles, d_mu, d_dx, d_part);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double us = ms * 1e3 / reps;
  printf("tcgen05 3xTF32: %.1f us per launch, %.2f G positions/s, %.1f effective fp32 TFLOP/s (1632 MAC per position)\n", us,
         n_pos / us * 1e-3, 2.0 * 1632 * n_pos / us * 1e-6);
  printf("for scale: k_arm_bwd (FFMA2, same network + context gather, Laplace rate, halo scatter) takes 68 us with two CTAs\n"
         "per SM and 128 us in the one-CTA-per-SM form the training step uses, for 524 256 positions (profiles/r2_*)\n");
  return ok ? 0 : 2;
}
