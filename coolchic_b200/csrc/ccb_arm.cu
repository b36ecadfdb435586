// G + A + R: causal context gather, ARM MLP, Laplace rate -- forward and backward.
// Restates enc/component/core/arm.py:308-352 (_get_neighbor), :86-101,227-268 (Arm.forward),
// :355-370 (_laplace_cdf) and enc/component/coolchic.py:399-424 (rate), all positions of all
// resolutions in ONE launch: the quantised latents are known at encode time, so the
// decoder's sequential scan becomes an embarrassingly parallel map over M positions.
//
// Layout: one CTA = one TW x TH tile of one (batch, resolution); the tile and its causal
// halo (4 rows above, 4 columns either side, zero outside the grid: arm.py:326-327) are
// staged in shared memory; the MLP weights sit in the constant bank (C <= 16, see c_arm_w) or in
// shared memory (C = 24, 32: read as warp-wide broadcasts); activations live in registers (lane == position).
//
// This file is compiled CCB_N_WEIGHT_SLOTS + 1 times (csrc/build.sh): once without CCB_SLOT -- the public entry
// points and the kernels that read the weights from the tensors (shared memory, no state outside the call) -- and
// once per weight slot with -DCCB_SLOT=k, which yields the constant-bank kernels of C <= 16 bound to that
// translation unit's own __constant__ bank.  A slot is therefore a separate 64 KB constant space with
// compile-time offsets (the only form ptxas keeps as uniform loads feeding FFMA2; a run-time slot offset, even a
// kernel argument, degrades every read to a per-thread LDC as soon as the kernel contains a loop -- checked in SASS).
#include "ccb_internal.cuh"

#ifndef CCB_SLOT
#define CCB_SLOT -1
#endif
#define CCB_CAT2(a, b) a##b
#define CCB_CAT(a, b) CCB_CAT2(a, b)

constexpr int ARM_TH_HOST = 8;  // rows per ARM tile (== ARM_TH below)

#if CCB_SLOT < 0
static int arm_bwd_cap(const ccb_plan* p) { return p->n_sm * (p->arch.dim_arm <= 16 ? 2 : 1); }

// Persistent CTAs of the backward.  Shared weights: one group of CTAs walks the whole tile list.  Per-sample
// weights: one group per batch element (a CTA keeps one weight set and one gradient accumulator).
void ccb_arm_bwd_grid(const ccb_plan* p, bool per_sample, int* gx, int* gy) {
  const int cap = arm_bwd_cap(p);
  if (per_sample) {
    const int B = p->arch.B, per_img = p->n_arm_tiles / B;
    int x = cap / B;
    if (x < 1) x = 1;
    *gx = per_img < x ? per_img : x;
    *gy = B;
  } else {
    *gx = p->n_arm_tiles < cap ? p->n_arm_tiles : cap;
    *gy = 1;
  }
}
int ccb_arm_bwd_max_rows(const ccb_plan* p) {
  int gx, gy, hx, hy;
  ccb_arm_bwd_grid(p, false, &gx, &gy);
  ccb_arm_bwd_grid(p, true, &hx, &hy);
  return gx * gy > hx * hy ? gx * gy : hx * hy;
}
bool ccb_arm_const_ok(const ccb_plan* p) { return p->arch.dim_arm <= 16; }
#endif  // CCB_SLOT < 0

namespace {

constexpr int ARM_TH = 8;        // rows per tile == warps per CTA
constexpr int ARM_THREADS = 256;
constexpr int HALO = 4;

// Causal context offsets (arm.py:373-506; same sets as cpp/cc-frame-decoder.cpp:111-154).
template <int C> struct Ctx;
#define CCB_CTX(CN, ...)                                                        \
  template <> struct Ctx<CN> {                                                  \
    static __host__ __device__ constexpr int dy(int c) {                        \
      constexpr int t[2][CN] = __VA_ARGS__;                                     \
      return t[0][c];                                                           \
    }                                                                           \
    static __host__ __device__ constexpr int dx(int c) {                        \
      constexpr int t[2][CN] = __VA_ARGS__;                                     \
      return t[1][c];                                                           \
    }                                                                           \
  };
CCB_CTX(8, {{-3, -2, -1, -1, -1, 0, 0, 0}, {0, 0, -1, 0, 1, -3, -2, -1}})
CCB_CTX(16, {{-3, -3, -2, -2, -2, -2, -2, -1, -1, -1, -1, -1, -1, 0, 0, 0},
             {0, 1, -2, -1, 0, 1, 2, -3, -2, -1, 0, 1, 2, -3, -2, -1}})
CCB_CTX(24, {{-4, -3, -3, -3, -3, -3, -2, -2, -2, -2, -2, -2, -2, -1, -1, -1, -1, -1, -1, -1, 0, 0, 0, 0},
             {0, -2, -1, 0, 1, 2, -3, -2, -1, 0, 1, 2, 3, -3, -2, -1, 0, 1, 2, 3, -4, -3, -2, -1}})
CCB_CTX(32, {{-4, -4, -4, -4, -3, -3, -3, -3, -3, -3, -3, -2, -2, -2, -2, -2,
              -2, -2, -2, -1, -1, -1, -1, -1, -1, -1, -1, -1, 0,  0,  0,  0},
             {-2, -1, 0, 1, -3, -2, -1, 0, 1, 2, 3, -3, -2, -1, 0, 1,
              2,  3,  4, -4, -3, -2, -1, 0, 1, 2, 3, 4,  -4, -3, -2, -1}})
#undef CCB_CTX

// Weights in the constant bank (C <= 16 only: <= 850 floats).  A warp-uniform weight read from shared memory is
// an LDS.128 per four multiply-adds and caps the matrix-vector chain at ~34 TFLOP/s on B200 (measured,
// scratch/mlp_bench.cu); the same chain with constant-bank operands (LDCU into uniform registers, FFMA/FFMA2 with a
// uniform operand) reaches 52-64 TFLOP/s at 8-32 warps per SM as long as the weights stay within ~2-3 KB.
// The weights change every iteration, so they are copied (device to device, stream ordered, graph capturable)
// into the bank before the kernels that read them.  One bank per translation unit = per weight slot (see the top of
// the file); slots are owned by callers (ccb_weight_slot_acquire), so concurrent encodes never share one, and the
// module path (no slot) reads shared memory.
constexpr int ARM_CONST_FLOATS = 896;
__constant__ __align__(16) float c_arm_w[ARM_CONST_FLOATS];

template <bool CW>
struct ArmW {
  const float* s;  // shared-memory copy (CW == false)
  __device__ __forceinline__ float at(int i) const {
    if constexpr (CW) return c_arm_w[i];
    else return s[i];
  }
  __device__ __forceinline__ float4 at4(int i) const {
    if constexpr (CW) return *reinterpret_cast<const float4*>(&c_arm_w[i]);
    else return *reinterpret_cast<const float4*>(s + i);
  }
};

using ArmArgs = CcbArmArgs;  // ccb_internal.cuh (shared by the per-slot translation units)

// Number of ARM parameters, in [w0,b0,w1,b1,...,wo,bo] order.
template <int C, int NH> __host__ __device__ constexpr int arm_n_params() { return NH * (C * C + C) + 2 * C + 2; }

template <int C, int NH>
__device__ __forceinline__ void load_weights(const ArmArgs& a, float* sw, int b) {
  // sw: [k][C*C + C] for hidden layers, then [2*C + 2]; b: batch element whose weight set is wanted
  const long long wo = (long long)b * a.wbs;
  for (int k = 0; k < NH; ++k) {
    float* dst = sw + k * (C * C + C);
    for (int i = threadIdx.x; i < C * C; i += blockDim.x) dst[i] = a.w[k][wo + i];
    for (int i = threadIdx.x; i < C; i += blockDim.x) dst[C * C + i] = a.b[k][wo + i];
  }
  float* dst = sw + NH * (C * C + C);
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) dst[i] = a.w[NH][wo + i];
  if (threadIdx.x < 2) dst[2 * C + threadIdx.x] = a.b[NH][wo + threadIdx.x];
}

template <int TW>
__device__ __forceinline__ void load_tile(const ArmArgs& a, const ArmTile& t, float* sy) {
  constexpr int SW = TW + 2 * HALO;
  constexpr int SH = ARM_TH + HALO;
  const int h = a.g.h[t.level], w = a.g.w[t.level];
  const float* src = a.yhat + (long long)t.b * a.g.M + a.g.off[t.level];
  stage_tile<SW, false>(sy, SW, src, h, w, t.y0 - HALO, t.x0 - HALO, SH);
  cp_async_wait_all();
}

// dot(w_row, x) with packed fp32 FMAs (FFMA2, sm_100): the weight row comes out of shared memory as float4 =
// two natural pairs, x is read as adjacent pairs, so one issue slot does two multiply-adds; the two lanes of
// the accumulator (even / odd elements) are added at the end.
template <int C, bool CW>
__device__ __forceinline__ float dot_packed(const ArmW<CW>& W, int row, const float* x, float init) {
  float2 acc = make_float2(init, 0.f);
#pragma unroll
  for (int i = 0; i < C; i += 4) {
    const float4 v = W.at4(row + i);
    acc = __ffma2_rn(make_float2(v.x, v.y), make_float2(x[i], x[i + 1]), acc);
    acc = __ffma2_rn(make_float2(v.z, v.w), make_float2(x[i + 2], x[i + 3]), acc);
  }
  return acc.x + acc.y;
}

// Laplace pieces shared by forward and backward.
struct RateOut {
  float rate, scale, diff;  // diff = F(y+.5) - F(y-.5) before the clamp
  float dp, dm;             // y + .5 - mu, y - .5 - mu
  float ep, em;             // exp(-|dp|/s), exp(-|dm|/s)
};

__device__ __forceinline__ float sgn(float v) { return (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f); }

__device__ __forceinline__ float laplace_cdf(float d, float s) {
  // 0.5 - 0.5 * sign(d) * expm1(-|d| / s)   (arm.py:368-370)
  return 0.5f - 0.5f * sgn(d) * expm1f(-fabsf(d) / s);
}

__device__ __forceinline__ float rate_of(float y, float mu, float ls, float& scale_out) {
  float s = expf(fminf(fmaxf(ls - 4.0f, -4.6f), 5.0f));  // arm.py:266
  scale_out = s;
  float p = laplace_cdf(y + 0.5f - mu, s) - laplace_cdf(y - 0.5f - mu, s);
  p = fmaxf(p, 1.52587890625e-05f);  // 2^-16 (coolchic.py:419-423)
  return -log2f(p);
}

// ------------------------------------------------------------------------------------------
// Forward
// ------------------------------------------------------------------------------------------
template <int C, int NH, int P, bool CW>
__global__ void __launch_bounds__(ARM_THREADS) k_arm_fwd(const __grid_constant__ ArmArgs a) {
  constexpr int TW = 32 * P;
  constexpr int SW = TW + 2 * HALO;
  constexpr int SH = ARM_TH + HALO;
  extern __shared__ __align__(128) float smem[];
  float* sy = smem;               // tile first: a TMA destination is 128-byte aligned (SH * SW * 4 is a multiple of 128)
  float* sw = smem + SH * SW;     // weights (unless in the constant bank)
  __shared__ float red[32];
  __shared__ __align__(8) uint64_t s_bar;
  const ArmW<CW> W = {sw};

  const ArmTile t = a.tiles[blockIdx.x];
  if (a.use_tma) {
    // one thread asks the TMA unit for the (x, y, batch) box; outside the grid the unit writes zeros (arm.py:326-327)
    if (threadIdx.x == 0) {
      mbar_init(&s_bar, 1);
      mbar_fence_init();
      mbar_expect_tx(&s_bar, SH * SW * sizeof(float));
      tma_load_3d(sy, &a.tmap[t.level], t.x0 - HALO, t.y0 - HALO, t.b, &s_bar);
    }
    if constexpr (!CW) load_weights<C, NH>(a, sw, t.b);
    __syncthreads();  // the barrier is initialised (and the weights are in place)
    mbar_wait(&s_bar, 0);
  } else {
    if constexpr (!CW) load_weights<C, NH>(a, sw, t.b);
    load_tile<TW>(a, t, sy);
    __syncthreads();
  }

  const int lane = threadIdx.x & 31, wy = threadIdx.x >> 5;
  const int h = a.g.h[t.level], w = a.g.w[t.level];
  const int gy = t.y0 + wy;

  float x[P][C];
#pragma unroll
  for (int p = 0; p < P; ++p) {
    const int lx = lane + 32 * p;
#pragma unroll
    for (int c = 0; c < C; ++c) x[p][c] = sy[(wy + HALO + Ctx<C>::dy(c)) * SW + lx + HALO + Ctx<C>::dx(c)];
  }

#pragma unroll
  for (int k = 0; k < NH; ++k) {
    const int wk = k * (C * C + C);
    float o[P][C];
#pragma unroll
    for (int j = 0; j < C; ++j) {
      const float bj = W.at(wk + C * C + j);
#pragma unroll
      for (int p = 0; p < P; ++p) {
        // residual (arm.py:97-101)
        o[p][j] = fmaxf(dot_packed<C, CW>(W, wk + j * C, x[p], bj + x[p][j]), 0.f);
      }
    }
#pragma unroll
    for (int p = 0; p < P; ++p)
#pragma unroll
      for (int j = 0; j < C; ++j) x[p][j] = o[p][j];
  }

  constexpr int wo = NH * (C * C + C);
  float rsum = 0.f;
#pragma unroll
  for (int p = 0; p < P; ++p) {
    const float mu = dot_packed<C, CW>(W, wo, x[p], W.at(wo + 2 * C));
    const float ls = dot_packed<C, CW>(W, wo + C, x[p], W.at(wo + 2 * C + 1));
    const int gx = t.x0 + lane + 32 * p;
    if (gy < h && gx < w) {
      const float y = sy[(wy + HALO) * SW + lane + 32 * p + HALO];
      float s;
      const float r = rate_of(y, mu, ls, s);
      const long long idx = (long long)t.b * a.g.M + a.g.off[t.level] + (long long)gy * w + gx;
      a.rate[idx] = r;
      if (a.mu) a.mu[idx] = mu;
      if (a.scale) a.scale[idx] = s;
      if (a.log_scale) a.log_scale[idx] = ls;
      rsum += r;
    }
  }
  if (a.rate_part) {
    float tot = block_sum(rsum, red);
    if (threadIdx.x == 0) a.rate_part[blockIdx.x] = tot;
  }
}

// D (16 x 8, fp32) += A (16 x 8) B (8 x 8) on the tensor cores, TF32 operands.  Error-compensated ("3xTF32"): each
// operand is split into its TF32 part (low 13 mantissa bits cleared) and the exact remainder, and the three products
// hi*hi + hi*lo + lo*hi are accumulated in fp32: relative error ~2^-20 per product, the fp32 noise level of the sums.
__device__ __forceinline__ void mma_tf32(float* d, const uint32_t* a, const uint32_t* b) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void split_tf32(float v, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(v) & 0xffffe000u;
  lo = __float_as_uint(v - __uint_as_float(hi));
}

// ------------------------------------------------------------------------------------------
// Backward: recompute the forward per position (cheaper than saving 3*C floats per position
// in HBM), back-propagate to mu / log-scale, the MLP, the C context taps and the weights.
//   * persistent CTAs (2 per SM) walk the tile list; weights are loaded to shared memory once;
//   * d yhat: each position's C context gradients are parked in shared memory, then every
//     position of the tile + its "receiving" halo gathers what it is owed (fixed order) and issues ONE
//     global float atomic.  The halo overlaps the neighbouring tiles, so a receiver can be owed by up to four
//     tiles (its own, left OR right, and the same two of the row band below): three or four float atomics in
//     scheduling order would make the sum order-dependent.  Contributions to the tile's own row band and to
//     the band above therefore go to two different zero-initialised accumulators: each location then receives
//     at most TWO atomic adds, and 0 + a + b == 0 + b + a exactly -- the result is bitwise reproducible; the
//     quantiser backward adds the accumulators in a fixed order;
//   * d W: outer products sum_p dpre[p][j] * x[p][i] are re-laid through a per-warp shared
//     staging buffer so that lanes own (j,i) blocks and accumulate in REGISTERS across all the
//     tiles of the CTA; they are folded once, at the end, without atomics (shared-memory float
//     atomics are CAS loops on this architecture and were the top stall of the first version).
// ------------------------------------------------------------------------------------------
template <int C, int NH, bool CW>
__global__ void __launch_bounds__(ARM_THREADS, (C <= 16 ? 2 : 1)) k_arm_bwd(const __grid_constant__ ArmArgs a,
                                                                            int n_tiles, int n_active) {
  constexpr int TW = 32;
  constexpr int SW = TW + 2 * HALO;
  constexpr int SH = ARM_TH + HALO;
  constexpr int NP = arm_n_params<C, NH>();
  constexpr int NPA = (NP + 3) & ~3;
  constexpr int RS = 2 * C + 4;   // staging row stride (floats)
  constexpr int RJ = C / 8, RI = C / 4;
  constexpr int NPOS = TW * ARM_TH;
  constexpr int NHA = NH > 0 ? NH : 1;
  extern __shared__ __align__(128) float smem[];
  // two y_hat tiles (+halo) first: TMA destinations are 128-byte aligned (SH * SW * 4 = 1920 is a multiple of 128);
  // the TMA path double-buffers (the next tile's box is in flight behind this tile's math), cp.async uses buffer 0
  float* sy_buf[2] = {smem, smem + SH * SW};
  float* sw = smem + 2 * SH * SW;         // [NPA] weights (unused when they sit in the constant bank)
  float* sacc = CW ? sw : sw + NPA;       // [NPA] CTA sums of the weight gradients
  float* sd = sacc + NPA;                 // [(C+1)][NPOS] context gradients (+ self term)
  float* sst = sd + (C + 1) * NPOS;       // [8 warps][32][RS] staging
  __shared__ float srate[ARM_THREADS / 32];  // per-warp rate sums of the current tile
  __shared__ __align__(8) uint64_t s_bar[2];

  const ArmW<CW> W = {sw};
  float* prow = a.part + ((long long)blockIdx.y * gridDim.x + blockIdx.x + a.part_row0) * a.n_arm;
  if (blockIdx.x >= n_active) {  // shared-SM mode (see ccb_arm_backward): this CTA only owes a zero partial row
    for (int i = threadIdx.x; i < NP; i += blockDim.x) prow[i] = 0.f;
    return;
  }
  if constexpr (!CW) load_weights<C, NH>(a, sw, blockIdx.y);
  for (int i = threadIdx.x; i < NPA; i += blockDim.x) sacc[i] = 0.f;

  const int lane = threadIdx.x & 31, wy = threadIdx.x >> 5;
  float* st = sst + wy * 32 * RS;
  const int jg = lane >> 2, ig = lane & 3;
  const int li = lane < C ? lane : 0;

  // weight-gradient accumulators of this lane (whole CTA lifetime)
  float2 aw[NHA][RJ][RI / 2];  // packed pairs along the input index
  float ab[NHA][RJ];
#pragma unroll
  for (int k = 0; k < NHA; ++k)
#pragma unroll
    for (int rj = 0; rj < RJ; ++rj) {
      ab[k][rj] = 0.f;
#pragma unroll
      for (int ri = 0; ri < RI / 2; ++ri) aw[k][rj][ri] = make_float2(0.f, 0.f);
    }
  // C == 16: the hidden layers' weight gradients are 16 x 16 x 32 GEMMs per warp and tile on mma.sync (see below);
  // accumulator fragments: [layer][n block of 8 inputs][c0..c3], bias sums over this lane's fragment rows (j = g, g + 8)
  constexpr bool MMA_DW = (C == 16);
  float dwf[NHA][2][4];
  float dbf[NHA][2];
#pragma unroll
  for (int k = 0; k < NHA; ++k) {
    dbf[k][0] = dbf[k][1] = 0.f;
#pragma unroll
    for (int nb = 0; nb < 2; ++nb)
#pragma unroll
      for (int e = 0; e < 4; ++e) dwf[k][nb][e] = 0.f;
  }
  float ao0 = 0.f, ao1 = 0.f;
  float2 bo = make_float2(0.f, 0.f);  // (d b_mu, d b_ls), packed adds

  (void)n_tiles;
  const int tile_end = a.tile_hi > 0 ? a.tile_hi : (blockIdx.y + 1) * a.tiles_per_group;
  const int tile_first = (a.tile_hi > 0 ? a.tile_lo : blockIdx.y * a.tiles_per_group) + blockIdx.x;
  const bool tma = a.use_tma != 0;
  // TMA: thread 0 posts the (x, y, batch) box of a tile; the unit zero-fills what lies outside the grid (the causal
  // halo of arm.py:326-327) and signals the buffer's mbarrier with the byte count
  auto post = [&](int tile_i, int buf) {
    const ArmTile tt = a.tiles[tile_i];
    mbar_expect_tx(&s_bar[buf], SH * SW * sizeof(float));
    tma_load_3d(sy_buf[buf], &a.tmap[tt.level], tt.x0 - HALO, tt.y0 - HALO, tt.b, &s_bar[buf]);
  };
  if (tma) {
    if (threadIdx.x == 0) {
      mbar_init(&s_bar[0], 1);
      mbar_init(&s_bar[1], 1);
      mbar_fence_init();
      if (tile_first < tile_end) post(tile_first, 0);
    }
    __syncthreads();
  }
  int it = 0;
  for (int tile = tile_first; tile < tile_end; tile += n_active, ++it) {
    const ArmTile t = a.tiles[tile];
    const float* sy = sy_buf[tma ? (it & 1) : 0];
    if (tma) {
      // buffer (it + 1) & 1 was last read before the mid-iteration barrier of iteration it - 1: free to refill
      if (threadIdx.x == 0 && tile + n_active < tile_end) post(tile + n_active, (it + 1) & 1);
      mbar_wait(&s_bar[it & 1], (it >> 1) & 1);
    } else {
      load_tile<TW>(a, t, sy_buf[0]);
    }
    __syncthreads();  // also orders the previous tile's gather (reads sd) before this tile's writes to sd

    const int h = a.g.h[t.level], w = a.g.w[t.level];
    const int gy = t.y0 + wy, gx = t.x0 + lane;
    const bool active = (gy < h) && (gx < w);
    const long long idx = (long long)t.b * a.g.M + a.g.off[t.level] + (long long)gy * w + gx;

    // ---- recompute forward, keeping every layer's input
    float x[NH + 1][C];
#pragma unroll
    for (int c = 0; c < C; ++c) x[0][c] = sy[(wy + HALO + Ctx<C>::dy(c)) * SW + lane + HALO + Ctx<C>::dx(c)];
#pragma unroll
    for (int k = 0; k < NH; ++k) {
#pragma unroll
      for (int j = 0; j < C; ++j) {
        const int row = k * (C * C + C) + j * C;
        const float acc = dot_packed<C, CW>(W, row, x[k], W.at(k * (C * C + C) + C * C + j) + x[k][j]);
        x[k + 1][j] = fmaxf(acc, 0.f);
      }
    }
    constexpr int wo = NH * (C * C + C);
    const float mu = dot_packed<C, CW>(W, wo, x[NH], W.at(wo + 2 * C));
    const float ls = dot_packed<C, CW>(W, wo + C, x[NH], W.at(wo + 2 * C + 1));

    // ---- rate backward
    float d_mu = 0.f, d_ls = 0.f, d_self = 0.f, r_pos = 0.f;
    if (active) {
      const float g = a.d_rate ? a.d_rate[idx] : a.d_rate_uniform;
      const float y = sy[(wy + HALO) * SW + lane + HALO];
      const float lc = ls - 4.0f;
      const float s = expf(fminf(fmaxf(lc, -4.6f), 5.0f));
      const float dp = y + 0.5f - mu, dm = y - 0.5f - mu;
      const float diff = laplace_cdf(dp, s) - laplace_cdf(dm, s);
      r_pos = -log2f(fmaxf(diff, 1.52587890625e-05f));  // the rate itself (coolchic.py:419-424), see rate_part below
      if (diff >= 1.52587890625e-05f) {  // clamp_min passes the gradient where input >= min
        const float gp = -g / (diff * 0.6931471805599453f);  // d(-log2 p)/dp
        const float ep = expf(-fabsf(dp) / s), em = expf(-fabsf(dm) / s);
        const float inv_s = 1.0f / s;
        // dF/dd = 0.5 * sign(d)^2 * exp(-|d|/s) / s ; dF/ds = -0.5 * d * exp(-|d|/s) / s^2
        const float fp = 0.5f * sgn(dp) * sgn(dp) * ep * inv_s;
        const float fm = 0.5f * sgn(dm) * sgn(dm) * em * inv_s;
        d_self = gp * (fp - fm);
        d_mu = -d_self;
        const float d_s = gp * (-0.5f) * (dp * ep - dm * em) * inv_s * inv_s;
        d_ls = (lc >= -4.6f && lc <= 5.0f) ? d_s * s : 0.f;
      }
    }

    // ---- output layer
    float dx[C];
#pragma unroll
    for (int i = 0; i < C; ++i) dx[i] = W.at(wo + i) * d_mu + W.at(wo + C + i) * d_ls;
    {
      // stage x[NH] | (d_mu, d_ls); lane i < C then owns d Wo[0][i], d Wo[1][i]
#pragma unroll
      for (int i = 0; i < C; i += 4)
        *reinterpret_cast<float4*>(st + lane * RS + i) = make_float4(x[NH][i], x[NH][i + 1], x[NH][i + 2], x[NH][i + 3]);
      *reinterpret_cast<float2*>(st + lane * RS + C) = make_float2(d_mu, d_ls);
      __syncwarp();
#pragma unroll 8
      for (int r = 0; r < 32; ++r) {
        const float xv = st[r * RS + li];
        const float2 dv = *reinterpret_cast<const float2*>(st + r * RS + C);
        ao0 = fmaf(dv.x, xv, ao0);
        ao1 = fmaf(dv.y, xv, ao1);
        bo = __fadd2_rn(bo, dv);
      }
      __syncwarp();
    }

    // ---- hidden layers, last to first
#pragma unroll
    for (int k = NH - 1; k >= 0; --k) {
      const int wk = k * (C * C + C);
      float dpre[C];
#pragma unroll
      for (int j = 0; j < C; ++j) dpre[j] = (x[k + 1][j] > 0.f) ? dx[j] : 0.f;
      // weight gradient through the staging buffer
#pragma unroll
      for (int i = 0; i < C; i += 4) {
        *reinterpret_cast<float4*>(st + lane * RS + i) = make_float4(x[k][i], x[k][i + 1], x[k][i + 2], x[k][i + 3]);
        *reinterpret_cast<float4*>(st + lane * RS + C + i) = make_float4(dpre[i], dpre[i + 1], dpre[i + 2], dpre[i + 3]);
      }
      __syncwarp();
      if constexpr (MMA_DW) {
        // d W[j][i] += sum_r dpre[r][j] x[r][i] over the warp's 32 positions: M = j (16), N = i (two blocks of 8),
        // K = r (four steps of 8).  Fragments (m16n8k8, g = lane / 4, t = lane % 4): A[j][r] at rows g, g + 8 and
        // columns t, t + 4; B[r][i] at rows t, t + 4 and column g; D[j][i] at rows g, g + 8 and columns 2t, 2t + 1.
        const int g = lane >> 2, t = lane & 3;
#pragma unroll
        for (int s8 = 0; s8 < 4; ++s8) {
          const float* r0 = st + (8 * s8 + t) * RS;
          const float* r1 = r0 + 4 * RS;
          const float av[4] = {r0[C + g], r0[C + g + 8], r1[C + g], r1[C + g + 8]};
          uint32_t ah[4], al[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) split_tf32(av[e], ah[e], al[e]);
          dbf[k][0] += av[0] + av[2];
          dbf[k][1] += av[1] + av[3];
#pragma unroll
          for (int nb = 0; nb < 2; ++nb) {
            uint32_t bh[2], bl[2];
            split_tf32(r0[nb * 8 + g], bh[0], bl[0]);
            split_tf32(r1[nb * 8 + g], bh[1], bl[1]);
            mma_tf32(dwf[k][nb], ah, bh);
            mma_tf32(dwf[k][nb], ah, bl);
            mma_tf32(dwf[k][nb], al, bh);
          }
        }
      } else {
#pragma unroll 4
      for (int r = 0; r < 32; ++r) {
        float xr[RI], dr[RJ];
        if constexpr (RI % 4 == 0) {
#pragma unroll
          for (int ri = 0; ri < RI; ri += 4) {
            const float4 v = *reinterpret_cast<const float4*>(st + r * RS + ig * RI + ri);
            xr[ri] = v.x; xr[ri + 1] = v.y; xr[ri + 2] = v.z; xr[ri + 3] = v.w;
          }
        } else {
#pragma unroll
          for (int ri = 0; ri < RI; ri += 2) {
            const float2 v = *reinterpret_cast<const float2*>(st + r * RS + ig * RI + ri);
            xr[ri] = v.x; xr[ri + 1] = v.y;
          }
        }
#pragma unroll
        for (int rj = 0; rj < RJ; ++rj) dr[rj] = st[r * RS + C + jg * RJ + rj];
        if constexpr (RJ % 2 == 0) {
#pragma unroll
          for (int rj = 0; rj < RJ; rj += 2) {  // packed adds on the bias-gradient pairs
            const float2 t = __fadd2_rn(make_float2(ab[k][rj], ab[k][rj + 1]), make_float2(dr[rj], dr[rj + 1]));
            ab[k][rj] = t.x;
            ab[k][rj + 1] = t.y;
          }
        }
#pragma unroll
        for (int rj = 0; rj < RJ; ++rj) {
          if constexpr (RJ % 2 != 0) ab[k][rj] += dr[rj];
          const float2 d2 = make_float2(dr[rj], dr[rj]);
#pragma unroll
          for (int ri = 0; ri < RI; ri += 2)
            aw[k][rj][ri / 2] = __ffma2_rn(d2, make_float2(xr[ri], xr[ri + 1]), aw[k][rj][ri / 2]);
        }
      }
      }
      __syncwarp();
      // d input of the layer: residual + W^T dpre
      {
        float2 dx2[C / 2];
#pragma unroll
        for (int i = 0; i < C; i += 2) dx2[i / 2] = make_float2(dpre[i], dpre[i + 1]);
#pragma unroll
        for (int j = 0; j < C; ++j) {
          const float2 dj = make_float2(dpre[j], dpre[j]);
#pragma unroll
          for (int i = 0; i < C; i += 4) {
            const float4 v = W.at4(wk + j * C + i);
            dx2[i / 2] = __ffma2_rn(make_float2(v.x, v.y), dj, dx2[i / 2]);
            dx2[i / 2 + 1] = __ffma2_rn(make_float2(v.z, v.w), dj, dx2[i / 2 + 1]);
          }
        }
#pragma unroll
        for (int i = 0; i < C; i += 2) {
          dx[i] = dx2[i / 2].x;
          dx[i + 1] = dx2[i / 2].y;
        }
      }
    }

    // ---- park the context gradients, then gather per receiving position
    const int lpos = wy * TW + lane;
#pragma unroll
    for (int c = 0; c < C; ++c) sd[c * NPOS + lpos] = dx[c];
    sd[C * NPOS + lpos] = d_self;
    if (a.rate_part) {  // training step: this kernel also emits the per-tile rate sums (no separate ARM forward launch)
      const float rw = warp_sum(r_pos);
      if (lane == 0) srate[wy] = rw;
    }
    __syncthreads();
    if (a.rate_part && threadIdx.x == 0) {
      float tot = 0.f;
#pragma unroll
      for (int i = 0; i < ARM_THREADS / 32; ++i) tot += srate[i];
      a.rate_part[tile] = tot;
    }

    const long long lvl_off = (long long)t.b * a.g.M + a.g.off[t.level];
    for (int e = threadIdx.x; e < SH * SW; e += blockDim.x) {
      const int ey = e / SW, ex = e - ey * SW;
      const int ty = ey - HALO, tx = ex - HALO;  // tile-local coordinates of the receiver
      const int ry = t.y0 + ty, rx = t.x0 + tx;
      if (ry < 0 || ry >= h || rx < 0 || rx >= w) continue;
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const int sy_ = ty - Ctx<C>::dy(c), sx_ = tx - Ctx<C>::dx(c);  // source position (tile-local)
        if (sy_ >= 0 && sy_ < ARM_TH && sx_ >= 0 && sx_ < TW) sum += sd[c * NPOS + sy_ * TW + sx_];
      }
      if (ty >= 0 && tx >= 0 && tx < TW) sum += sd[C * NPOS + ty * TW + tx];
      atomicAdd((ty >= 0 ? a.dyhat : a.dyhat_b) + lvl_off + (long long)ry * w + rx, sum);
    }
  }

  // ---- fold the lane accumulators: one warp at a time adds into the CTA sums (no atomics)
  for (int wsel = 0; wsel < ARM_THREADS / 32; ++wsel) {
    __syncthreads();
    if (wy != wsel) continue;
#pragma unroll
    for (int k = 0; k < NH; ++k) {
      float* acc_k = sacc + k * (C * C + C);
      if constexpr (MMA_DW) {
        const int g = lane >> 2, t = lane & 3;
#pragma unroll
        for (int nb = 0; nb < 2; ++nb) {
          acc_k[g * C + nb * 8 + 2 * t] += dwf[k][nb][0];
          acc_k[g * C + nb * 8 + 2 * t + 1] += dwf[k][nb][1];
          acc_k[(g + 8) * C + nb * 8 + 2 * t] += dwf[k][nb][2];
          acc_k[(g + 8) * C + nb * 8 + 2 * t + 1] += dwf[k][nb][3];
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float v = dbf[k][h];
          v += __shfl_xor_sync(0xffffffffu, v, 1);
          v += __shfl_xor_sync(0xffffffffu, v, 2);
          if (t == 0) acc_k[C * C + g + 8 * h] += v;
        }
      } else {
#pragma unroll
        for (int rj = 0; rj < RJ; ++rj) {
#pragma unroll
          for (int ri = 0; ri < RI; ++ri)
            acc_k[(jg * RJ + rj) * C + ig * RI + ri] += (ri & 1) ? aw[k][rj][ri >> 1].y : aw[k][rj][ri >> 1].x;
          if (ig == 0) acc_k[C * C + jg * RJ + rj] += ab[k][rj];
        }
      }
    }
    float* acc_o = sacc + NH * (C * C + C);
    if (lane < C) {
      acc_o[lane] += ao0;
      acc_o[C + lane] += ao1;
    }
    if (lane == 0) {
      acc_o[2 * C] += bo.x;
      acc_o[2 * C + 1] += bo.y;
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NP; i += blockDim.x) prow[i] = sacc[i];
}

template <int C, int NH, int P, bool CW> size_t fwd_smem() {
  constexpr int NPA = CW ? 0 : (arm_n_params<C, NH>() + 3) & ~3;
  return (size_t)(NPA + (ARM_TH + HALO) * (32 * P + 2 * HALO)) * sizeof(float);
}
template <int C, int NH, bool CW> size_t bwd_smem() {
  constexpr int NPA = (arm_n_params<C, NH>() + 3) & ~3;
  constexpr int SHW = (ARM_TH + HALO) * (32 + 2 * HALO);  // 480 floats = 1920 bytes (a multiple of 128)
  return (size_t)((CW ? 1 : 2) * NPA + 2 * SHW + (C + 1) * 32 * ARM_TH + 8 * 32 * (2 * C + 4)) * sizeof(float);
}

// CW: this translation unit's constant bank (a slot build) or the tensors (the stateless build)
constexpr bool TU_CW = (CCB_SLOT >= 0);

template <int C, int NH>
int launch_fwd(const ccb_plan* p, const ArmArgs& a0, cudaStream_t s) {
  ArmArgs a = a0;
  a.tiles = p->d_arm_tiles_fwd;  // 32 * arm_fwd_p columns per tile
  constexpr int P = C <= 16 ? 2 : 1;
  const size_t sm = fwd_smem<C, NH, P, TU_CW>();
  CCB_CHECK(cudaFuncSetAttribute(k_arm_fwd<C, NH, P, TU_CW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  k_arm_fwd<C, NH, P, TU_CW><<<p->n_arm_tiles_fwd, ARM_THREADS, sm, s>>>(a);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
template <int C, int NH>
int launch_bwd(const ccb_plan* p, const ArmArgs& a0, cudaStream_t s) {
  ArmArgs a = a0;
  const size_t sm = bwd_smem<C, NH, TU_CW>();
  CCB_CHECK(cudaFuncSetAttribute(k_arm_bwd<C, NH, TU_CW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  const int gx = a.grid_x, gy = a.grid_y;
  a.tiles_per_group = p->n_arm_tiles / gy;
  // shared-SM mode: one working CTA per SM (half the register file) so that the kernels of the other stream of the
  // training step keep co-residing; the remaining CTAs of the grid only zero their partial rows
  int n_active = (a.shared_sm && gy == 1 && gx > p->n_sm) ? p->n_sm : gx;
  {
    // CCB_ARM_ACTIVE (tuning switch): working CTAs of the shared-SM form, n_sm .. 2 n_sm
    static const int forced = []() { const char* e = getenv("CCB_ARM_ACTIVE"); return e ? atoi(e) : 0; }();
    if (forced > 0 && a.shared_sm && gy == 1 && gx > p->n_sm) n_active = forced < gx ? forced : gx;
  }
  int launch_x = gx;
  if (a.tile_hi > 0 && a.part_row0 > 0) launch_x = n_active = gx - a.part_row0;  // second part: the rows the first part's idle CTAs zeroed
  k_arm_bwd<C, NH, TU_CW><<<dim3(launch_x, gy), ARM_THREADS, sm, s>>>(a, p->n_arm_tiles, n_active);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}

template <int C>
int dispatch_nh(bool bwd, const ccb_plan* p, const ArmArgs& a, cudaStream_t s) {
  switch (p->arch.n_hidden_arm) {
    case 0: return bwd ? launch_bwd<C, 0>(p, a, s) : launch_fwd<C, 0>(p, a, s);
    case 1: return bwd ? launch_bwd<C, 1>(p, a, s) : launch_fwd<C, 1>(p, a, s);
    case 2: return bwd ? launch_bwd<C, 2>(p, a, s) : launch_fwd<C, 2>(p, a, s);
    case 3: return bwd ? launch_bwd<C, 3>(p, a, s) : launch_fwd<C, 3>(p, a, s);
    default: return CCB_ERR_UNSUPPORTED;
  }
}

int dispatch(bool bwd, const ccb_plan* p, const ArmArgs& a, cudaStream_t s) {
  switch (p->arch.dim_arm) {
    case 8: return dispatch_nh<8>(bwd, p, a, s);
    case 16: return dispatch_nh<16>(bwd, p, a, s);
#if CCB_SLOT < 0
    case 24: return dispatch_nh<24>(bwd, p, a, s);
    case 32: return dispatch_nh<32>(bwd, p, a, s);
#endif
    default: return CCB_ERR_UNSUPPORTED;
  }
}

}  // namespace

#if CCB_SLOT >= 0
// ---- one weight slot: launcher of the constant-bank kernels + the device address of this unit's bank
int CCB_CAT(ccb_arm_slot_launch_, CCB_SLOT)(bool bwd, const ccb_plan* p, const CcbArmArgs& a, cudaStream_t s) {
  return dispatch(bwd, p, a, s);
}
int CCB_CAT(ccb_arm_slot_bank_, CCB_SLOT)(float** out) {
  CCB_CHECK(cudaGetSymbolAddress((void**)out, c_arm_w));  // of the CURRENT device
  return CCB_OK;
}
#else
// ---- public (library-internal) entry points
#define CCB_DECL_SLOT(k)                                                                                  \
  int ccb_arm_slot_launch_##k(bool bwd, const ccb_plan* p, const CcbArmArgs& a, cudaStream_t s);         \
  int ccb_arm_slot_bank_##k(float** out);
CCB_DECL_SLOT(0) CCB_DECL_SLOT(1) CCB_DECL_SLOT(2) CCB_DECL_SLOT(3)
CCB_DECL_SLOT(4) CCB_DECL_SLOT(5) CCB_DECL_SLOT(6) CCB_DECL_SLOT(7)
#undef CCB_DECL_SLOT
typedef int (*arm_slot_launch_fn)(bool, const ccb_plan*, const CcbArmArgs&, cudaStream_t);
typedef int (*arm_slot_bank_fn)(float**);
static const arm_slot_launch_fn g_slot_launch[CCB_N_WEIGHT_SLOTS] = {
    ccb_arm_slot_launch_0, ccb_arm_slot_launch_1, ccb_arm_slot_launch_2, ccb_arm_slot_launch_3,
    ccb_arm_slot_launch_4, ccb_arm_slot_launch_5, ccb_arm_slot_launch_6, ccb_arm_slot_launch_7};
static const arm_slot_bank_fn g_slot_bank[CCB_N_WEIGHT_SLOTS] = {
    ccb_arm_slot_bank_0, ccb_arm_slot_bank_1, ccb_arm_slot_bank_2, ccb_arm_slot_bank_3,
    ccb_arm_slot_bank_4, ccb_arm_slot_bank_5, ccb_arm_slot_bank_6, ccb_arm_slot_bank_7};
static_assert(CCB_N_WEIGHT_SLOTS == 8, "one translation unit per slot: update build.sh and the tables above");

// ---- TMA tensor maps of the y_hat pyramid ------------------------------------------------------------------------
// One 3-D map (x, y, batch) per latent level over the flat [B, M] vector: dims (w, h, B), strides (4, 4 w, 4 M) bytes,
// box (elem_cols, 12, 1), out-of-bounds elements read as zeros.  TMA wants every stride and the base 16-byte aligned:
// all level widths and M multiples of 4 (768x512 and the 256x256 crops qualify; 1080p's 30-wide level does not).
typedef CUresult (*ccb_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static ccb_encode_tiled_fn encode_tiled() {
  static ccb_encode_tiled_fn fn = []() -> ccb_encode_tiled_fn {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<ccb_encode_tiled_fn>(f);
  }();
  return fn;
}

bool ccb_arm_make_tmaps(const ccb_plan* p, const void* yhat, int elem_cols, CUtensorMap* out) {
  if (!p->arm_tma) return false;
  ccb_encode_tiled_fn enc = encode_tiled();
  if (!enc) return false;
  const LevelGeom& g = p->g;
  for (int l = 0; l < g.L; ++l) {
    char* base = (char*)yhat + (size_t)g.off[l] * sizeof(float);
    if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)g.w[l], (cuuint64_t)g.h[l], (cuuint64_t)g.B};
    const cuuint64_t strides[2] = {(cuuint64_t)g.w[l] * sizeof(float), (cuuint64_t)g.M * sizeof(float)};
    const cuuint32_t box[3] = {(cuuint32_t)elem_cols, (cuuint32_t)(ARM_TH_HOST + 4), 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    if (enc(&out[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return false;
  }
  return true;
}

// the constant-bank kernels apply: a slot is given, C <= 16, one weight set for the batch
static bool arm_use_slot(const ccb_plan* p, const ccb_tensors* prm, int slot) {
  return slot >= 0 && slot < CCB_N_WEIGHT_SLOTS && ccb_arm_const_ok(p) && prm->weight_batch_stride == 0;
}

static ArmArgs base_args(const ccb_plan* p, const ccb_tensors* prm, const float* yhat) {
  ArmArgs a = {};
  a.tiles = p->d_arm_tiles;
  a.g = p->g;
  a.yhat = yhat;
  a.wbs = prm->weight_batch_stride;
  for (int k = 0; k <= p->arch.n_hidden_arm; ++k) {
    a.w[k] = prm->arm_w[k];
    a.b[k] = prm->arm_b[k];
  }
  a.n_arm = (int)p->wl.n_arm;
  return a;
}

// Copy the ARM weights of `prm` into the constant bank of `slot` (no-op when the constant-bank kernels do not
// apply).  Stream ordered; one copy when the tensors are contiguous ([w0,b0,w1,b1,...,wo,bo], the arena order).
// The address of a __constant__ symbol differs per device: resolved with the plan's device current.
int ccb_arm_stage_weights(const ccb_plan* p, const ccb_tensors* prm, cudaStream_t s, int slot) {
  const int C = p->arch.dim_arm, NH = p->arch.n_hidden_arm;
  if (!arm_use_slot(p, prm, slot)) return CCB_OK;
  float* dst = nullptr;
  {
    int cur = -1;
    CCB_CHECK(cudaGetDevice(&cur));
    if (cur != p->device) CCB_CHECK(cudaSetDevice(p->device));
    const int rc = g_slot_bank[slot](&dst);
    if (cur != p->device) cudaSetDevice(cur);
    if (rc != CCB_OK) return rc;
  }
  const int n_total = NH * (C * C + C) + 2 * C + 2;
  bool contiguous = true;
  const float* expect = prm->arm_w[0];
  for (int k = 0; k <= NH; ++k) {
    const int nw = k < NH ? C * C : 2 * C, nb = k < NH ? C : 2;
    if (prm->arm_w[k] != expect || prm->arm_b[k] != expect + nw) contiguous = false;
    expect += nw + nb;
  }
  if (contiguous) {
    CCB_CHECK(cudaMemcpyAsync(dst, prm->arm_w[0], sizeof(float) * n_total, cudaMemcpyDeviceToDevice, s));
    return CCB_OK;
  }
  for (int k = 0; k <= NH; ++k) {
    const int nw = k < NH ? C * C : 2 * C, nb = k < NH ? C : 2;
    CCB_CHECK(cudaMemcpyAsync(dst, prm->arm_w[k], sizeof(float) * nw, cudaMemcpyDeviceToDevice, s));
    CCB_CHECK(cudaMemcpyAsync(dst + nw, prm->arm_b[k], sizeof(float) * nb, cudaMemcpyDeviceToDevice, s));
    dst += nw + nb;
  }
  return CCB_OK;
}

int ccb_arm_forward(const ccb_plan* p, const ccb_tensors* prm, const float* yhat, float* rate, float* mu,
                    float* scale, float* log_scale, float* rate_part, cudaStream_t s, int slot, bool weights_staged) {
  const bool use_slot = arm_use_slot(p, prm, slot);
  if (use_slot && !weights_staged) {
    int rc = ccb_arm_stage_weights(p, prm, s, slot);
    if (rc != CCB_OK) return rc;
  }
  ArmArgs a = base_args(p, prm, yhat);
  a.rate = rate;
  a.mu = mu;
  a.scale = scale;
  a.log_scale = log_scale;
  a.rate_part = rate_part;
  a.use_tma = ccb_arm_make_tmaps(p, yhat, 32 * p->arm_fwd_p + 8, a.tmap) ? 1 : 0;
  return use_slot ? g_slot_launch[slot](false, p, a, s) : dispatch(false, p, a, s);
}

int ccb_arm_backward(const ccb_plan* p, const ccb_tensors* prm, const float* yhat, const float* d_rate,
                     float d_rate_uniform, float* dyhat_a, float* dyhat_b, float* part, cudaStream_t s, int slot,
                     bool weights_staged, float* rate_part, bool shared_sm, int tile_lo, int tile_hi, int part_row0) {
  const bool use_slot = arm_use_slot(p, prm, slot);
  if (use_slot && !weights_staged) {
    int rc = ccb_arm_stage_weights(p, prm, s, slot);
    if (rc != CCB_OK) return rc;
  }
  ArmArgs a = base_args(p, prm, yhat);
  a.d_rate = d_rate;
  a.d_rate_uniform = d_rate_uniform;
  a.dyhat = dyhat_a;
  a.dyhat_b = dyhat_b;
  a.part = part;
  a.rate_part = rate_part;  // [n_arm_tiles] per-tile rate sums, or NULL
  a.shared_sm = shared_sm ? 1 : 0;
  ccb_arm_bwd_grid(p, a.wbs != 0, &a.grid_x, &a.grid_y);
  if (tile_hi > 0) {
    if (a.wbs != 0 || a.grid_y != 1 || tile_lo < 0 || tile_hi > p->n_arm_tiles || tile_lo >= tile_hi) return CCB_ERR_INVALID;
    a.tile_lo = tile_lo;
    a.tile_hi = tile_hi;
    a.part_row0 = part_row0;
  }
  a.use_tma = ccb_arm_make_tmaps(p, yhat, 32 + 8, a.tmap) ? 1 : 0;
  return use_slot ? g_slot_launch[slot](true, p, a, s) : dispatch(true, p, a, s);
}
#endif  // CCB_SLOT
