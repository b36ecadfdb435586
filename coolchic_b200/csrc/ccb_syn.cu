// S: synthesis conv stack, forward and backward.
// Restates enc/component/core/synthesis.py:69-84 (SynthesisConv2d: replicate pad (k-1)/2,
// conv2d, + x if residual, ReLU or identity) and :175-277 (Synthesis = sequence of those).
// Supported shape (every decoder config shipped under cfg/: SURVEY.md 8): two 1x1 layers
// (L -> S ReLU, S -> 3) followed by any number of 3x3 3->3 layers.
//   * the two 1x1 layers are fused per pixel: the S-wide hidden activation lives in
//     registers only (at 768x512 it would be 75 MB in HBM);
//   * each 3x3 layer is a shared-memory stencil over a replicate-clamped tile.
//
// Like ccb_arm.cu this file is compiled once without CCB_SLOT (public entry points + the kernels that read the
// weights from the tensors) and once per weight slot with -DCCB_SLOT=k (the constant-bank kernels bound to that
// translation unit's own __constant__ banks: k_syn_head_fwd_c, k_syn_head_bwd<.., SC > 0>, k_syn_tail2_fwd).
#include "ccb_internal.cuh"

#ifndef CCB_SLOT
#define CCB_SLOT -1
#endif
#define CCB_CAT2(a, b) a##b
#define CCB_CAT(a, b) CCB_CAT2(a, b)

namespace {

constexpr int CIN = 7;    // = number of latent resolutions
constexpr int C3 = 3;     // image channels
constexpr int HEAD_P = 4; // pixels per thread in the fused 1x1 pair
constexpr int HW_ROW = 12;  // smem floats per hidden unit: w0[7], b0, w1[3], pad

using HeadArgs = CcbHeadArgs;  // ccb_internal.cuh (shared by the per-slot translation units)

__device__ __forceinline__ void head_load_weights(const HeadArgs& a, float* sw) {
  const long long wo = (long long)blockIdx.y * a.wbs;
  for (int i = threadIdx.x; i < a.S * HW_ROW; i += blockDim.x) {
    int s = i / HW_ROW, c = i - s * HW_ROW;
    float v = 0.f;
    if (c < CIN) v = a.w0[wo + s * CIN + c];
    else if (c == CIN) v = a.b0[wo + s];
    else if (c < CIN + 1 + C3) v = a.w1[wo + (c - CIN - 1) * a.S + s];
    sw[i] = v;
  }
}

// ---- packed-FP32 (FFMA2, sm_100) form of the per-pixel MLP --------------------------------------------
// A thread owns pairs of horizontally adjacent pixels; every value is a float2 = (pixel 2i, pixel 2i+1) and
// the weights sit in shared memory already duplicated as (w, w), so one FFMA2 does the work of two FFMAs in
// one issue slot (the kernels of this file are issue-bound, not FMA-pipe-bound).
constexpr int HW2_ROW = 24;  // floats per hidden unit: (w0[c],w0[c]) x7, (b0,b0), (w1[o],w1[o]) x3, pad 2
constexpr int HEAD2_THREADS = 256;

__device__ __forceinline__ void head_load_weights2(const HeadArgs& a, float* sw) {
  const int SP = (a.S + 1) & ~1;  // rows are consumed two at a time: an odd S gets one all-zero row
  const long long wo = (long long)blockIdx.y * a.wbs;
  for (int i = threadIdx.x; i < SP * HW2_ROW; i += blockDim.x) {
    const int s = i / HW2_ROW, c = (i - s * HW2_ROW) >> 1;
    float v = 0.f;
    if (s >= a.S) v = 0.f;
    else if (c < CIN) v = a.w0[wo + s * CIN + c];
    else if (c == CIN) v = a.b0[wo + s];
    else if (c < CIN + 1 + C3) v = a.w1[wo + (c - CIN - 1) * a.S + s];
    sw[i] = v;
  }
}

// Two horizontally adjacent pixels (pix, pix + 1; pix even) of image blockIdx.y.
struct PixPair {
  long long r0;      // offset of the first pixel inside a [HW] plane
  long long bch;     // blockIdx.y (batch element)
  bool v0, v1, vec;  // validity; vec: both valid and the pair is 8-byte aligned in every plane (even plane size)
};

__device__ __forceinline__ PixPair pix_pair(const HeadArgs& a, long long pix) {
  PixPair q;
  q.v0 = pix < a.HW;
  q.v1 = pix + 1 < a.HW;
  q.r0 = q.v0 ? pix : 0;
  q.bch = blockIdx.y;
  q.vec = q.v1 && ((a.HW & 1) == 0);
  return q;
}

__device__ __forceinline__ float2 load_pair(const float* __restrict__ base, const PixPair& q, int n_ch, int c,
                                            long long HW) {
  const float* src = base + (q.bch * n_ch + c) * HW + q.r0;
  if (q.vec) return *reinterpret_cast<const float2*>(src);
  float2 v = make_float2(0.f, 0.f);
  if (q.v0) v.x = src[0];
  if (q.v1) v.y = src[1];
  return v;
}

__device__ __forceinline__ void store_pair(float* __restrict__ base, const PixPair& q, int n_ch, int c, long long HW,
                                           float2 v) {
  float* dst = base + (q.bch * n_ch + c) * HW + q.r0;
  if (q.vec) {
    *reinterpret_cast<float2*>(dst) = v;
    return;
  }
  if (q.v0) dst[0] = v.x;
  if (q.v1) dst[1] = v.y;
}

template <int NPAIR>
__global__ void __launch_bounds__(HEAD2_THREADS, 3) k_syn_head_fwd(HeadArgs a) {
  extern __shared__ __align__(16) float sw[];
  head_load_weights2(a, sw);
  __syncthreads();
  float2 z[NPAIR][CIN], o[NPAIR][C3];
  PixPair pp[NPAIR];
  const float* b1p = a.b1 + (long long)blockIdx.y * a.wbs;
  const float b1x = b1p[0], b1y = b1p[1], b1z = b1p[2];
#pragma unroll
  for (int q = 0; q < NPAIR; ++q) {
    const long long pair = ((long long)blockIdx.x * NPAIR + q) * HEAD2_THREADS + threadIdx.x;
    pp[q] = pix_pair(a, 2 * pair);
#pragma unroll
    for (int c = 0; c < CIN; ++c) z[q][c] = load_pair(a.z, pp[q], CIN, c, a.HW);
    o[q][0] = make_float2(b1x, b1x);
    o[q][1] = make_float2(b1y, b1y);
    o[q][2] = make_float2(b1z, b1z);
  }
  for (int s = 0; s < a.S; s += 2) {
    float2 w0[2][CIN], bb[2], w1[2][C3];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const float4* wr = reinterpret_cast<const float4*>(sw + (s + u) * HW2_ROW);
      const float4 q0 = wr[0], q1 = wr[1], q2 = wr[2], q3 = wr[3], q4 = wr[4], q5 = wr[5];
      w0[u][0] = make_float2(q0.x, q0.y); w0[u][1] = make_float2(q0.z, q0.w); w0[u][2] = make_float2(q1.x, q1.y);
      w0[u][3] = make_float2(q1.z, q1.w); w0[u][4] = make_float2(q2.x, q2.y); w0[u][5] = make_float2(q2.z, q2.w);
      w0[u][6] = make_float2(q3.x, q3.y);
      bb[u] = make_float2(q3.z, q3.w);
      w1[u][0] = make_float2(q4.x, q4.y); w1[u][1] = make_float2(q4.z, q4.w); w1[u][2] = make_float2(q5.x, q5.y);
    }
#pragma unroll
    for (int q = 0; q < NPAIR; ++q) {
      float2 ha = bb[0], hb = bb[1];
#pragma unroll
      for (int c = 0; c < CIN; ++c) {
        ha = __ffma2_rn(w0[0][c], z[q][c], ha);
        hb = __ffma2_rn(w0[1][c], z[q][c], hb);
      }
      ha.x = fmaxf(ha.x, 0.f); ha.y = fmaxf(ha.y, 0.f);
      hb.x = fmaxf(hb.x, 0.f); hb.y = fmaxf(hb.y, 0.f);
#pragma unroll
      for (int c = 0; c < C3; ++c) o[q][c] = __ffma2_rn(w1[1][c], hb, __ffma2_rn(w1[0][c], ha, o[q][c]));
    }
  }
#pragma unroll
  for (int q = 0; q < NPAIR; ++q) {
#pragma unroll
    for (int c = 0; c < C3; ++c) {
      float2 v = o[q][c];
      if (a.relu1) {
        v.x = fmaxf(v.x, 0.f);
        v.y = fmaxf(v.y, 0.f);
      }
      store_pair(a.out, pp[q], C3, c, a.HW, v);
      if (a.out2) store_pair(a.out2, pp[q], C3, c, a.HW, v);
    }
  }
}

// ---- constant-bank form (S in {8, 16, 40, 48}: the shipped decoder configurations) ---------------------
// The head's weights are copied into c_head_w as [w0^T: 7*S][b0: S][w1: 3*S][b1: 3] (k_head_pack transposes w0 so
// that two adjacent hidden units form an aligned pair everywhere) before the kernels that read them; with
// compile-time offsets every weight becomes a
// uniform-register operand of FFMA / FFMA2 (scalar broadcast `UR.F32`), so the inner loops issue no shared-memory
// loads at all (a broadcast LDS.128 per four multiply-adds capped these loops at ~34 of 72 TFLOP/s, see
// scratch/mlp_bench.cu).  One bank per device: see the note at c_arm_w in ccb_arm.cu.
constexpr int HEAD_CONST_FLOATS = 768;
__constant__ __align__(16) float c_head_w[HEAD_CONST_FLOATS];  // one bank per translation unit = per weight slot
__host__ __device__ constexpr bool head_const_ok(int S) { return S == 8 || S == 16 || S == 40 || S == 48; }

#if CCB_SLOT < 0
__global__ void k_head_pack(const float* __restrict__ w0, const float* __restrict__ b0, const float* __restrict__ w1,
                            const float* __restrict__ b1, int S, float* __restrict__ out) {
  const int n = S * CIN + S + C3 * S + C3;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float v;
    if (i < S * CIN) {
      const int c = i / S, u = i - c * S;
      v = w0[u * CIN + c];
    } else if (i < S * CIN + S) v = b0[i - S * CIN];
    else if (i < S * CIN + S + C3 * S) v = w1[i - S * CIN - S];
    else v = b1[i - S * CIN - S - C3 * S];
    out[i] = v;
  }
}

#endif  // CCB_SLOT < 0

template <int S, int NPAIR>
#ifndef CCB_HEADF_LB
#define CCB_HEADF_LB 4  // 60 registers, no spills: two CTAs fit beside the ARM branch's persistent CTA (66 registers -> one)
#endif
__global__ void __launch_bounds__(HEAD2_THREADS, CCB_HEADF_LB) k_syn_head_fwd_c(HeadArgs a) {
  constexpr int OB0 = S * CIN, OW1 = OB0 + S, OB1 = OW1 + C3 * S;
  const float* const chw = c_head_w;
  float2 z[NPAIR][CIN], o[NPAIR][C3];
  PixPair pp[NPAIR];
#pragma unroll
  for (int q = 0; q < NPAIR; ++q) {
    const long long pair = ((long long)blockIdx.x * NPAIR + q) * HEAD2_THREADS + threadIdx.x;
    pp[q] = pix_pair(a, 2 * pair);
#pragma unroll
    for (int c = 0; c < CIN; ++c) z[q][c] = load_pair(a.z, pp[q], CIN, c, a.HW);
#pragma unroll
    for (int c = 0; c < C3; ++c) o[q][c] = make_float2(chw[OB1 + c], chw[OB1 + c]);
  }
#pragma unroll
  for (int s = 0; s < S; ++s) {
#pragma unroll
    for (int q = 0; q < NPAIR; ++q) {
      float2 h = make_float2(chw[OB0 + s], chw[OB0 + s]);
#pragma unroll
      for (int c = 0; c < CIN; ++c) {
        const float w = chw[c * S + s];
        h = __ffma2_rn(make_float2(w, w), z[q][c], h);
      }
      h.x = fmaxf(h.x, 0.f);
      h.y = fmaxf(h.y, 0.f);
#pragma unroll
      for (int c = 0; c < C3; ++c) {
        const float w = chw[OW1 + c * S + s];
        o[q][c] = __ffma2_rn(make_float2(w, w), h, o[q][c]);
      }
    }
  }
#pragma unroll
  for (int q = 0; q < NPAIR; ++q) {
#pragma unroll
    for (int c = 0; c < C3; ++c) {
      float2 v = o[q][c];
      if (a.relu1) {
        v.x = fmaxf(v.x, 0.f);
        v.y = fmaxf(v.y, 0.f);
      }
      store_pair(a.out, pp[q], C3, c, a.HW, v);
      if (a.out2) store_pair(a.out2, pp[q], C3, c, a.HW, v);
    }
  }
}

// Backward of the fused pair.  Weight-gradient layout inside a partial row:
//   [w0: S*7][b0: S][w1: 3*S][b1: 3]
#if CCB_SLOT < 0
__global__ void __launch_bounds__(256) k_syn_head_bwd_generic(HeadArgs a) {
  extern __shared__ __align__(16) float sw[];
  float* sacc = sw + a.S * HW_ROW;  // [S*7 + S + 3*S + 3]
  const int S = a.S;
  const int n_par = S * CIN + S + C3 * S + C3;
  head_load_weights(a, sw);
  for (int i = threadIdx.x; i < n_par; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long long base = ((long long)blockIdx.x * blockDim.x) * HEAD_P + threadIdx.x;
  float z[HEAD_P][CIN], dz[HEAD_P][CIN], d[HEAD_P][C3];
  long long pb[HEAD_P];
  bool ok[HEAD_P];
  float gb1[C3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int p = 0; p < HEAD_P; ++p) {
    const long long pix = base + (long long)p * blockDim.x;
    ok[p] = pix < a.HW;
    const long long b = blockIdx.y;
    const long long r = ok[p] ? pix : 0;
    pb[p] = b * a.HW * CIN + r;
#pragma unroll
    for (int c = 0; c < CIN; ++c) {
      z[p][c] = ok[p] ? a.z[pb[p] + c * a.HW] : 0.f;
      dz[p][c] = 0.f;
    }
#pragma unroll
    for (int c = 0; c < C3; ++c) {
      float g = ok[p] ? a.d_out[(b * C3 + c) * a.HW + r] : 0.f;
      if (a.relu1 && ok[p]) g = (a.act[(b * C3 + c) * a.HW + r] > 0.f) ? g : 0.f;
      d[p][c] = g;
      gb1[c] += g;
    }
  }
  for (int s = 0; s < S; ++s) {
    const float4 q0 = *reinterpret_cast<const float4*>(sw + s * HW_ROW);
    const float4 q1 = *reinterpret_cast<const float4*>(sw + s * HW_ROW + 4);
    const float4 q2 = *reinterpret_cast<const float4*>(sw + s * HW_ROW + 8);
    const float w0[CIN] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z};
    float g0[CIN], gb0 = 0.f, g1[C3];
#pragma unroll
    for (int c = 0; c < CIN; ++c) g0[c] = 0.f;
#pragma unroll
    for (int c = 0; c < C3; ++c) g1[c] = 0.f;
#pragma unroll
    for (int p = 0; p < HEAD_P; ++p) {
      float h = q1.w;
#pragma unroll
      for (int c = 0; c < CIN; ++c) h = fmaf(w0[c], z[p][c], h);
      const bool on = h > 0.f;
      h = fmaxf(h, 0.f);
      float dh = q2.x * d[p][0];
      dh = fmaf(q2.y, d[p][1], dh);
      dh = fmaf(q2.z, d[p][2], dh);
      dh = on ? dh : 0.f;
#pragma unroll
      for (int c = 0; c < CIN; ++c) {
        dz[p][c] = fmaf(w0[c], dh, dz[p][c]);
        g0[c] = fmaf(dh, z[p][c], g0[c]);
      }
      gb0 += dh;
#pragma unroll
      for (int c = 0; c < C3; ++c) g1[c] = fmaf(d[p][c], h, g1[c]);
    }
#pragma unroll
    for (int c = 0; c < CIN; ++c) {
      float v = warp_sum(g0[c]);
      if (lane == 0) atomicAdd(sacc + s * CIN + c, v);
    }
    {
      float v = warp_sum(gb0);
      if (lane == 0) atomicAdd(sacc + S * CIN + s, v);
    }
#pragma unroll
    for (int c = 0; c < C3; ++c) {
      float v = warp_sum(g1[c]);
      if (lane == 0) atomicAdd(sacc + S * CIN + S + c * S + s, v);
    }
  }
#pragma unroll
  for (int c = 0; c < C3; ++c) {
    float v = warp_sum(gb1[c]);
    if (lane == 0) atomicAdd(sacc + S * CIN + S + C3 * S + c, v);
  }
#pragma unroll
  for (int p = 0; p < HEAD_P; ++p) {
    if (!ok[p]) continue;
#pragma unroll
    for (int c = 0; c < CIN; ++c) a.dz[pb[p] + c * a.HW] = dz[p][c];
  }
  __syncthreads();
  float* prow = a.part + ((long long)blockIdx.y * gridDim.x + blockIdx.x) * n_par;
  for (int i = threadIdx.x; i < n_par; i += blockDim.x) prow[i] = sacc[i];
}

#endif  // CCB_SLOT < 0

// Fast backward of the fused pair (S <= 64).  lane == pixel for the per-pixel math; the weight gradient
//   dW0[s][c] = sum_pix dh[pix][s] * z[pix][c],  dW1[o][s] = sum_pix d[pix][o] * h[pix][s]
// is an outer product summed over pixels: every warp parks, per pixel, the row R = (z0..z6, 1, d0, d1, d2, 0)
// and, per chunk of 16 hidden units, the pairs (dh_s, h_s) in a private shared staging buffer; then lanes
// re-map to (unit pair sg = lane/4, pixel subset ps = lane%4) and accumulate their 2 x 11 entries over
// 8 of the 32 parked pixels in registers.  No shuffles, no atomics inside the loop; persistent CTAs.
constexpr int HB_ROW = 56;  // R (12) + 16 x (dh, h) = 44 used; 56 keeps the MMA fragment loads free of bank conflicts

// D (16 x 8, fp32) += A (16 x 8) B (8 x 8), TF32 operands; error-compensated as in ccb_arm.cu (hi*hi + hi*lo + lo*hi)
__device__ __forceinline__ void hb_mma_tf32(float* d, const uint32_t* a, const uint32_t* b) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void hb_split(float v, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(v) & 0xffffe000u;
  lo = __float_as_uint(v - __uint_as_float(hi));
}

#ifndef CCB_HEAD_LB
#define CCB_HEAD_LB 2  // resident CTAs per SM the head backward is compiled for (and its persistent grid sized to)
#endif
template <int NCHUNK, bool WITH_DZ, int SC>
__global__ void __launch_bounds__(256, CCB_HEAD_LB) k_syn_head_bwd(HeadArgs a, int n_groups) {
  extern __shared__ __align__(16) float sw[];
  constexpr int SP = 16 * NCHUNK;            // padded hidden width
  constexpr bool CW = SC > 0;                // weights in the constant bank (then S == SC at compile time)
  const int S = CW ? SC : a.S;
  const float* const chw = c_head_w;
  const long long wo = (long long)blockIdx.y * a.wbs;
  float* sst = CW ? sw : sw + SP * HW_ROW;  // [8 warps][32][HB_ROW]
  const int n_par = S * CIN + S + C3 * S + C3;
  if constexpr (!CW) {
    for (int i = threadIdx.x; i < SP * HW_ROW; i += blockDim.x) {
      const int ss = i / HW_ROW, c = i - ss * HW_ROW;
      float v = 0.f;
      if (ss < S) {
        if (c < CIN) v = a.w0[wo + ss * CIN + c];
        else if (c == CIN) v = a.b0[wo + ss];
        else if (c < CIN + 1 + C3) v = a.w1[wo + (c - CIN - 1) * S + ss];
      }
      sw[i] = v;
    }
    __syncthreads();
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* st = sst + wid * 32 * HB_ROW;
  // accumulator fragments of the two outer products of a chunk of 16 units (rows s = g, g + 8; columns 2t, 2t + 1):
  //   acc0 = dh^T [z | 1]   (16 x 8: d W0 and d b0),   acc1 = h^T [d0 d1 d2 0 ...]   (16 x 8, columns 0..2: d W1^T)
  const int fg = lane >> 2, ft = lane & 3;
  float acc0[NCHUNK][4], acc1[NCHUNK][4];
#pragma unroll
  for (int ch = 0; ch < NCHUNK; ++ch)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc0[ch][e] = acc1[ch][e] = 0.f;
  float gb1[C3] = {0.f, 0.f, 0.f};

  for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x) {  // n_groups: 256-pixel groups of ONE image
    const long long pix = (long long)grp * 256 + threadIdx.x;
    const bool ok = pix < a.HW;
    const long long b = blockIdx.y;
    const long long r = ok ? pix : 0;
    const long long pb = b * a.HW * CIN + r;
    float z[CIN], dz[CIN], d[C3];
    float2 dz2[CIN];  // constant-bank path: partial sums over even / odd units
#pragma unroll
    for (int c = 0; c < CIN; ++c) {
      z[c] = ok ? a.z[pb + c * a.HW] : 0.f;
      dz[c] = 0.f;
      dz2[c] = make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int c = 0; c < C3; ++c) {
      float g = ok ? a.d_out[(b * C3 + c) * a.HW + r] : 0.f;
      if (a.relu1 && ok) g = (a.act[(b * C3 + c) * a.HW + r] > 0.f) ? g : 0.f;
      d[c] = g;
      gb1[c] += g;
    }
    float* row = st + lane * HB_ROW;
    *reinterpret_cast<float4*>(row) = make_float4(z[0], z[1], z[2], z[3]);
    *reinterpret_cast<float4*>(row + 4) = make_float4(z[4], z[5], z[6], 1.0f);
    *reinterpret_cast<float4*>(row + 8) = make_float4(d[0], d[1], d[2], 0.f);
#pragma unroll
    for (int ch = 0; ch < NCHUNK; ++ch) {
#pragma unroll
      for (int u = 0; u < 16; u += 2) {
        float pr[4];
        if constexpr (CW) {
          // units (su, su + 1) as the two halves of packed FMAs: every weight pair is an aligned pair of the
          // bank ([w0^T: 7][S] | b0 [S] | w1 [3][S]), the pixel's own values are scalar broadcasts
          constexpr int OB0 = SC * CIN, OW1 = OB0 + SC;
          const int su = ch * 16 + u;
          if (su < SC) {
            float2 h2 = *reinterpret_cast<const float2*>(&chw[OB0 + su]);
#pragma unroll
            for (int c = 0; c < CIN; ++c)
              h2 = __ffma2_rn(make_float2(z[c], z[c]), *reinterpret_cast<const float2*>(&chw[c * SC + su]), h2);
            float2 dh2 = make_float2(0.f, 0.f);
#pragma unroll
            for (int c = 0; c < C3; ++c)
              dh2 = __ffma2_rn(make_float2(d[c], d[c]), *reinterpret_cast<const float2*>(&chw[OW1 + c * SC + su]),
                               dh2);
            dh2.x = h2.x > 0.f ? dh2.x : 0.f;
            dh2.y = h2.y > 0.f ? dh2.y : 0.f;
            h2.x = fmaxf(h2.x, 0.f);
            h2.y = fmaxf(h2.y, 0.f);
            if (WITH_DZ) {
#pragma unroll
              for (int c = 0; c < CIN; ++c)
                dz2[c] = __ffma2_rn(dh2, *reinterpret_cast<const float2*>(&chw[c * SC + su]), dz2[c]);
            }
            pr[0] = dh2.x; pr[1] = h2.x; pr[2] = dh2.y; pr[3] = h2.y;
          } else {  // padding units of the last chunk
            pr[0] = pr[1] = pr[2] = pr[3] = 0.f;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int su = ch * 16 + u + j;
            const float* wrow = sw + su * HW_ROW;
            const float4 q0 = *reinterpret_cast<const float4*>(wrow);
            const float4 q1 = *reinterpret_cast<const float4*>(wrow + 4);
            const float4 q2 = *reinterpret_cast<const float4*>(wrow + 8);
            const float w0[CIN] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z};
            float h = q1.w;
#pragma unroll
            for (int c = 0; c < CIN; ++c) h = fmaf(w0[c], z[c], h);
            const bool on = h > 0.f;
            h = fmaxf(h, 0.f);
            float dh = q2.x * d[0];
            dh = fmaf(q2.y, d[1], dh);
            dh = fmaf(q2.z, d[2], dh);
            dh = on ? dh : 0.f;
            if (WITH_DZ) {
#pragma unroll
              for (int c = 0; c < CIN; ++c) dz[c] = fmaf(w0[c], dh, dz[c]);
            }
            pr[2 * j] = dh;
            pr[2 * j + 1] = h;
          }
        }
        *reinterpret_cast<float4*>(row + 12 + 2 * u) = make_float4(pr[0], pr[1], pr[2], pr[3]);
      }
      __syncwarp();
      // the warp's 32 parked pixels = K: four steps of 8.  A fragments from the (dh, h) pairs of units fg, fg + 8 at
      // pixels ft, ft + 4; B fragments from the row R = (z0..z6, 1 | d0, d1, d2, 0, ...) at pixels ft, ft + 4, column fg
#pragma unroll
      for (int s8 = 0; s8 < 4; ++s8) {
        const float* r0 = st + (8 * s8 + ft) * HB_ROW;
        const float* r1 = r0 + 4 * HB_ROW;
        const float2 p00 = *reinterpret_cast<const float2*>(r0 + 12 + 2 * fg);
        const float2 p01 = *reinterpret_cast<const float2*>(r0 + 12 + 2 * fg + 16);
        const float2 p10 = *reinterpret_cast<const float2*>(r1 + 12 + 2 * fg);
        const float2 p11 = *reinterpret_cast<const float2*>(r1 + 12 + 2 * fg + 16);
        uint32_t ah[4], al[4], hh[4], hl[4], bh[2], bl[2];
        hb_split(p00.x, ah[0], al[0]); hb_split(p01.x, ah[1], al[1]); hb_split(p10.x, ah[2], al[2]); hb_split(p11.x, ah[3], al[3]);
        hb_split(p00.y, hh[0], hl[0]); hb_split(p01.y, hh[1], hl[1]); hb_split(p10.y, hh[2], hl[2]); hb_split(p11.y, hh[3], hl[3]);
        hb_split(r0[fg], bh[0], bl[0]);
        hb_split(r1[fg], bh[1], bl[1]);
        hb_mma_tf32(acc0[ch], ah, bh);
        hb_mma_tf32(acc0[ch], ah, bl);
        hb_mma_tf32(acc0[ch], al, bh);
        // columns 4..7 of this block read the first (dh, h) pairs of the row: finite values, columns nobody reads
        hb_split(r0[8 + fg], bh[0], bl[0]);
        hb_split(r1[8 + fg], bh[1], bl[1]);
        hb_mma_tf32(acc1[ch], hh, bh);
        hb_mma_tf32(acc1[ch], hh, bl);
        hb_mma_tf32(acc1[ch], hl, bh);
      }
      __syncwarp();
    }
    if (WITH_DZ && ok) {
#pragma unroll
      for (int c = 0; c < CIN; ++c) a.dz[pb + c * a.HW] = CW ? dz2[c].x + dz2[c].y : dz[c];
    }
  }
  // ---- fold: pixel subsets (lanes ps = 0..3) by shuffles, then the 8 warps through the (now idle) staging buffer:
  // every warp parks its n_par sums in its own slice, the CTA adds the slices (no shared atomics: CAS loops here)
  __syncthreads();
  float* mine = sst + wid * 32 * HB_ROW;  // >= 1408 floats per warp, n_par <= 64 * 11 + 3 = 707
  for (int i = lane; i < n_par; i += 32) mine[i] = 0.f;
  __syncwarp();
#pragma unroll
  for (int ch = 0; ch < NCHUNK; ++ch)
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int ss = ch * 16 + fg + (e >> 1) * 8, col = 2 * ft + (e & 1);  // fragment element -> (unit, column)
      if (ss < S) {
        mine[col < 7 ? ss * CIN + col : S * CIN + ss] = acc0[ch][e];
        if (col < C3) mine[S * CIN + S + col * S + ss] = acc1[ch][e];
      }
    }
#pragma unroll
  for (int c = 0; c < C3; ++c) {
    float v = warp_sum(gb1[c]);
    if (lane == 0) mine[S * CIN + S + C3 * S + c] = v;
  }
  __syncthreads();
  float* prow = a.part + ((long long)blockIdx.y * gridDim.x + blockIdx.x) * n_par;
  for (int i = threadIdx.x; i < n_par; i += blockDim.x) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sst[k * 32 * HB_ROW + i];
    prow[i] = t;
  }
}

// Input gradient of the fused pair only (the critical path of the backward chain: the upsampling backward
// waits for d z, nothing waits for the weight gradients until the reduction at the end of the step, so the
// two-stream schedule runs k_syn_head_bwd<., false> beside the upsampling backward).
template <int NPAIR>
__global__ void __launch_bounds__(HEAD2_THREADS) k_syn_head_dz(HeadArgs a) {
  extern __shared__ __align__(16) float sw[];
  head_load_weights2(a, sw);
  __syncthreads();
  float2 z[NPAIR][CIN], dz[NPAIR][CIN], d[NPAIR][C3];
  PixPair pp[NPAIR];
#pragma unroll
  for (int q = 0; q < NPAIR; ++q) {
    const long long pair = ((long long)blockIdx.x * NPAIR + q) * HEAD2_THREADS + threadIdx.x;
    pp[q] = pix_pair(a, 2 * pair);
#pragma unroll
    for (int c = 0; c < CIN; ++c) {
      z[q][c] = load_pair(a.z, pp[q], CIN, c, a.HW);
      dz[q][c] = make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int c = 0; c < C3; ++c) {
      float2 g = load_pair(a.d_out, pp[q], C3, c, a.HW);
      if (a.relu1) {
        const float2 act = load_pair(a.act, pp[q], C3, c, a.HW);
        g.x = act.x > 0.f ? g.x : 0.f;
        g.y = act.y > 0.f ? g.y : 0.f;
      }
      d[q][c] = g;
    }
  }
  for (int s = 0; s < a.S; s += 2) {
    float2 w0[2][CIN], bb[2], w1[2][C3];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const float4* wr = reinterpret_cast<const float4*>(sw + (s + u) * HW2_ROW);
      const float4 q0 = wr[0], q1 = wr[1], q2 = wr[2], q3 = wr[3], q4 = wr[4], q5 = wr[5];
      w0[u][0] = make_float2(q0.x, q0.y); w0[u][1] = make_float2(q0.z, q0.w); w0[u][2] = make_float2(q1.x, q1.y);
      w0[u][3] = make_float2(q1.z, q1.w); w0[u][4] = make_float2(q2.x, q2.y); w0[u][5] = make_float2(q2.z, q2.w);
      w0[u][6] = make_float2(q3.x, q3.y);
      bb[u] = make_float2(q3.z, q3.w);
      w1[u][0] = make_float2(q4.x, q4.y); w1[u][1] = make_float2(q4.z, q4.w); w1[u][2] = make_float2(q5.x, q5.y);
    }
#pragma unroll
    for (int q = 0; q < NPAIR; ++q) {
      float2 ha = bb[0], hb = bb[1];
#pragma unroll
      for (int c = 0; c < CIN; ++c) {
        ha = __ffma2_rn(w0[0][c], z[q][c], ha);
        hb = __ffma2_rn(w0[1][c], z[q][c], hb);
      }
      float2 da = __fmul2_rn(w1[0][0], d[q][0]), db = __fmul2_rn(w1[1][0], d[q][0]);
      da = __ffma2_rn(w1[0][1], d[q][1], da);
      db = __ffma2_rn(w1[1][1], d[q][1], db);
      da = __ffma2_rn(w1[0][2], d[q][2], da);
      db = __ffma2_rn(w1[1][2], d[q][2], db);
      da.x = ha.x > 0.f ? da.x : 0.f; da.y = ha.y > 0.f ? da.y : 0.f;
      db.x = hb.x > 0.f ? db.x : 0.f; db.y = hb.y > 0.f ? db.y : 0.f;
#pragma unroll
      for (int c = 0; c < CIN; ++c) dz[q][c] = __ffma2_rn(w0[1][c], db, __ffma2_rn(w0[0][c], da, dz[q][c]));
    }
  }
#pragma unroll
  for (int q = 0; q < NPAIR; ++q)
#pragma unroll
    for (int c = 0; c < CIN; ++c) store_pair(a.dz, pp[q], CIN, c, a.HW, dz[q][c]);
}

// ------------------------------------------------------------------------------------------
// 3x3, 3 -> 3 layers
// ------------------------------------------------------------------------------------------
constexpr int CT_W = 32, CT_H = 16, CT_P = 4, CT_THREADS = 128;  // 4 warps x 4-row strips
constexpr int CS_W = CT_W + 2, CS_H = CT_H + 2;

struct ConvArgs {
  const float* x;  // [B,3,H,W] layer input
  float* y;        // [B,3,H,W] layer output
  float* y2;       // optional second copy of the output
  const float* w;  // [3,3,3,3] (cout, cin, ky, kx)
  const float* b;  // [3]
  long long wbs;   // weight batch stride (floats); 0: shared weights
  int B, H, W, residual, relu;
  int tiles_x, tiles_y;
  // backward
  const float* dy;  // gradient w.r.t. the layer output
  float* dx;        // gradient w.r.t. the layer input
  float* part;      // [gridDim.x][84]
};

// A thread owns a vertical strip of CT_P pixels and keeps the 81 + 3 weights in registers: per (c, kx) one
// column of CT_P + 2 inputs is read from shared memory and feeds 9 * CT_P FMAs.
#if CCB_SLOT < 0
__global__ void __launch_bounds__(CT_THREADS) k_syn_conv3_fwd(ConvArgs a) {
  __shared__ float sx[C3 * CS_H * CS_W];
  __shared__ __align__(16) float swt[C3 * C3 * 9 + C3];
  const int tile = blockIdx.x;
  const int b = tile / (a.tiles_x * a.tiles_y);
  const int tr = tile - b * a.tiles_x * a.tiles_y;
  const int y0 = (tr / a.tiles_x) * CT_H, x0 = (tr % a.tiles_x) * CT_W;
  const long long HW = (long long)a.H * a.W;
  if (threadIdx.x < 81) swt[threadIdx.x] = a.w[b * a.wbs + threadIdx.x];
  if (threadIdx.x < 3) swt[81 + threadIdx.x] = a.b[b * a.wbs + threadIdx.x];
  {
    const int lane_ = threadIdx.x & 31, wid_ = threadIdx.x >> 5;
    int gxs[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) gxs[k] = min(max(x0 - 1 + lane_ + 32 * k, 0), a.W - 1);
    const float* xb = a.x + (long long)b * C3 * HW;
    for (int ry = wid_; ry < CS_H; ry += CT_THREADS / 32) {
      const long long ro = (long long)min(max(y0 - 1 + ry, 0), a.H - 1) * a.W;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int rx = lane_ + 32 * k;
        if (rx < CS_W) {
#pragma unroll
          for (int c = 0; c < C3; ++c) cp_async4(sx + c * CS_H * CS_W + ry * CS_W + rx, xb + c * HW + ro + gxs[k], true);
        }
      }
    }
  }
  cp_async_wait_all();
  __syncthreads();
  const int lx = threadIdx.x & 31, ly0 = (threadIdx.x >> 5) * CT_P;
  const int gx = x0 + lx;
  float w[84];
#pragma unroll
  for (int i = 0; i < 84; i += 4) {
    const float4 t = *reinterpret_cast<const float4*>(&swt[i]);
    w[i] = t.x; w[i + 1] = t.y; w[i + 2] = t.z; w[i + 3] = t.w;
  }
  float o[CT_P][C3];
#pragma unroll
  for (int i = 0; i < CT_P; ++i)
#pragma unroll
    for (int oc = 0; oc < C3; ++oc) o[i][oc] = w[81 + oc];
  float ctr[CT_P][C3];  // the centre taps (residual connection)
#pragma unroll
  for (int c = 0; c < C3; ++c)
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      float v[CT_P + 2];
#pragma unroll
      for (int r = 0; r < CT_P + 2; ++r) v[r] = sx[(c * CS_H + ly0 + r) * CS_W + lx + kx];
      if (kx == 1) {
#pragma unroll
        for (int i = 0; i < CT_P; ++i) ctr[i][c] = v[i + 1];
      }
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int oc = 0; oc < C3; ++oc)
#pragma unroll
          for (int i = 0; i < CT_P; ++i) o[i][oc] = fmaf(w[((oc * C3 + c) * 3 + ky) * 3 + kx], v[i + ky], o[i][oc]);
    }
  if (gx >= a.W) return;
#pragma unroll
  for (int i = 0; i < CT_P; ++i) {
    const int gy = y0 + ly0 + i;
    if (gy >= a.H) break;
#pragma unroll
    for (int oc = 0; oc < C3; ++oc) {
      float v = o[i][oc];
      if (a.residual) v += ctr[i][oc];
      if (a.relu) v = fmaxf(v, 0.f);
      a.y[((long long)b * C3 + oc) * HW + (long long)gy * a.W + gx] = v;
      if (a.y2) a.y2[((long long)b * C3 + oc) * HW + (long long)gy * a.W + gx] = v;
    }
  }
}

#endif  // CCB_SLOT < 0

// ------------------------------------------------------------------------------------------
// The two trailing 3x3 layers (synthesis.py:69-84, twice) and, in the training step, the frame clamp, the squared
// error and its gradient (frame.py:183, loss.py:53-87) in ONE launch: a CTA stages a 34 x 34 (replicate-clamped)
// input tile, computes the 32 x 32 intermediate tile it needs in shared memory and the 30 x 30 outputs from it.
// Replicate padding of the SECOND layer means "the intermediate at the clamped position", not "the first layer
// evaluated outside the image": lanes outside the image evaluate the clamped column (same shared-memory reads as
// their neighbour) and out-of-image rows are copied from the clamped row before the second layer reads them.
// Persistent over the tile list (grid <= number of loss partial slots); one squared-error partial per CTA.
// ------------------------------------------------------------------------------------------
constexpr int T2_OUT = 30, T2_MID = 32, T2_IN = 34, T2_THREADS = 256;
// [w(81) b(3)] of the two layers, staged by ccb_syn_stage_weights (uniform-register FFMA operands: no weight
// registers, so four 256-thread CTAs fit per SM and the 468 tiles of a 768x512 image run as ONE wave)
using Tail2Args = CcbTail2Args;  // ccb_internal.cuh

#if CCB_SLOT >= 0
__constant__ __align__(16) float c_tail_w[2 * 84];  // one bank per translation unit = per weight slot
__global__ void __launch_bounds__(T2_THREADS, 4) k_syn_tail2_fwd(Tail2Args a) {
  __shared__ float sx[C3 * T2_IN * T2_IN];
  __shared__ float si[C3 * T2_IN * T2_IN];  // intermediate tile: rows / columns 32, 33 are never written (feed discarded outputs)
  __shared__ float red[32];
  if (a.tail_counter && blockIdx.x == 0 && threadIdx.x == 0) *a.tail_counter = 0u;
  for (int i = threadIdx.x; i < C3 * T2_IN * T2_IN; i += T2_THREADS) si[i] = 0.f;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const long long HW = (long long)a.H * a.W;
  float se = 0.f;
  for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
    const int b = tile / (a.tiles_x * a.tiles_y);
    const int tr = tile - b * a.tiles_x * a.tiles_y;
    const int y0 = (tr / a.tiles_x) * T2_OUT, x0 = (tr % a.tiles_x) * T2_OUT;
    __syncthreads();  // the previous tile's readers are done
#pragma unroll
    for (int c = 0; c < C3; ++c)
      stage_tile<T2_IN, true>(sx + c * T2_IN * T2_IN, T2_IN, a.x + ((long long)b * C3 + c) * HW, a.H, a.W, y0 - 2, x0 - 2,
                              T2_IN);
    cp_async_wait_all();
    __syncthreads();
    // ---- first layer on the 32 x 32 intermediate tile (origin y0-1, x0-1); warp = 4 rows, lane = column
    {
      const float* w = c_tail_w;
      const int ix = x0 - 1 + lane;
      const int lxc = min(max(ix, 0), a.W - 1) - (x0 - 1);  // column actually evaluated (replicate padding of layer 2)
      const int r0 = wid * 4;
      float o[4][C3], ctr[4][C3];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int oc = 0; oc < C3; ++oc) o[i][oc] = w[81 + oc];
#pragma unroll
      for (int c = 0; c < C3; ++c)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          float v[6];
#pragma unroll
          for (int r = 0; r < 6; ++r) v[r] = sx[(c * T2_IN + r0 + r) * T2_IN + lxc + kx];
          if (kx == 1) {
#pragma unroll
            for (int i = 0; i < 4; ++i) ctr[i][c] = v[i + 1];
          }
#pragma unroll
          for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int oc = 0; oc < C3; ++oc)
#pragma unroll
              for (int i = 0; i < 4; ++i) o[i][oc] = fmaf(w[((oc * C3 + c) * 3 + ky) * 3 + kx], v[i + ky], o[i][oc]);
        }
      const bool own_x = lane >= 1 && lane <= T2_OUT && ix < a.W;  // this CTA owns the interior 30 x 30 of the tile
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int iy = y0 - 1 + r0 + i;
#pragma unroll
        for (int oc = 0; oc < C3; ++oc) {
          float v = o[i][oc];
          if (a.res0) v += ctr[i][oc];
          if (a.relu0) v = fmaxf(v, 0.f);
          si[(oc * T2_IN + r0 + i) * T2_IN + lane] = v;
          if (own_x && r0 + i >= 1 && r0 + i <= T2_OUT && iy < a.H)
            a.mid[((long long)b * C3 + oc) * HW + (long long)iy * a.W + ix] = v;
        }
      }
    }
    __syncthreads();
    // ---- out-of-image rows of the intermediate = the clamped row (only row -1 and row H can feed a valid output)
    {
      const int r_top = (y0 == 0) ? 0 : -1;              // local row of image row -1
      const int r_bot = a.H - (y0 - 1);                  // local row of image row H
      if (threadIdx.x < 3 * 32) {
        const int c = threadIdx.x >> 5;
        if (r_top == 0) si[(c * T2_IN + 0) * T2_IN + lane] = si[(c * T2_IN + 1) * T2_IN + lane];
      } else if (threadIdx.x >= 128 && threadIdx.x < 128 + 3 * 32) {
        const int c = (threadIdx.x - 128) >> 5;
        if (r_bot >= 1 && r_bot < T2_MID) si[(c * T2_IN + r_bot) * T2_IN + lane] = si[(c * T2_IN + r_bot - 1) * T2_IN + lane];
      }
    }
    __syncthreads();
    // ---- second layer on the 30 x 30 outputs (+ loss)
    {
      const float* w = c_tail_w + 84;
      const int r0 = wid * 4;
      float o[4][C3], ctr[4][C3];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int oc = 0; oc < C3; ++oc) o[i][oc] = w[81 + oc];
#pragma unroll
      for (int c = 0; c < C3; ++c)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          float v[6];
#pragma unroll
          for (int r = 0; r < 6; ++r) v[r] = si[(c * T2_IN + min(r0 + r, T2_IN - 1)) * T2_IN + lane + kx];
          if (kx == 1) {
#pragma unroll
            for (int i = 0; i < 4; ++i) ctr[i][c] = v[i + 1];
          }
#pragma unroll
          for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int oc = 0; oc < C3; ++oc)
#pragma unroll
              for (int i = 0; i < 4; ++i) o[i][oc] = fmaf(w[((oc * C3 + c) * 3 + ky) * 3 + kx], v[i + ky], o[i][oc]);
        }
      const int gx = x0 + lane;
      if (lane < T2_OUT && gx < a.W) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int gy = y0 + r0 + i;
          if (r0 + i >= T2_OUT || gy >= a.H) break;
#pragma unroll
          for (int oc = 0; oc < C3; ++oc) {
            float v = o[i][oc];
            if (a.res1) v += ctr[i][oc];
            if (a.relu1) v = fmaxf(v, 0.f);
            const long long gi = ((long long)b * C3 + oc) * HW + (long long)gy * a.W + gx;
            if (a.y) a.y[gi] = v;
            if (a.tgt) {
              const float cl = fminf(fmaxf(v, 0.f), 1.f);
              const float e = cl - a.tgt[gi];
              se = fmaf(e, e, se);
              a.d_raw[gi] = (v >= 0.f && v <= 1.f) ? 2.0f * e * a.inv_n_img : 0.f;
            }
          }
        }
      }
    }
  }
  if (a.tgt) {
    __syncthreads();
    const float tot = block_sum(se, red);
    if (threadIdx.x == 0) {
      a.part[2 * blockIdx.x] = tot;
      a.part[2 * blockIdx.x + 1] = 0.f;
    }
    for (int i = gridDim.x + blockIdx.x * T2_THREADS + threadIdx.x; i < a.n_part; i += gridDim.x * T2_THREADS) {
      a.part[2 * i] = 0.f;
      a.part[2 * i + 1] = 0.f;
    }
  }
}

#endif  // CCB_SLOT >= 0

// Backward of a 3x3 layer.  One CTA (8 warps) = one 32 x 32 tile; a thread owns a vertical strip of 4 pixels and keeps
// its 81 + 3 weight-gradient sums in registers (81 independent FMA chains hide the pipe latency); the warp
// then folds them with a butterfly reduce-scatter (~3 instructions per value instead of 10 for a plain
// shuffle reduction) and the CTA writes ONE partial row.  The input gradient uses the same two tiles.
constexpr int BT_W = 32, BT_H = 32, BP = 4, BT_THREADS = 256;
constexpr int BS_W = BT_W + 2, BS_H = BT_H + 2;

#if CCB_SLOT < 0
// After the call lane l holds sum over lanes of v[l] (v is destroyed).
__device__ __forceinline__ float warp_reduce_scatter32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = up ? v[i] : v[i + s];
      const float keep = up ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}

#ifndef CCB_CONVB_LB
#define CCB_CONVB_LB 2
#endif
__global__ void __launch_bounds__(BT_THREADS, CCB_CONVB_LB) k_syn_conv3_bwd(ConvArgs a) {
  extern __shared__ __align__(16) float csm[];
  float* sx = csm;                         // [3][BS_H][BS_W] layer input, replicate-clamped
  float* sg = csm + C3 * BS_H * BS_W;      // [3][BS_H][BS_W] d pre-activation, zero outside the image
  __shared__ float swt[C3 * C3 * 9];
  __shared__ float sacc[BT_THREADS / 32][96];  // one row per warp: no shared atomics (CAS loops on this architecture)
  const long long HW = (long long)a.H * a.W;
  const int tile = blockIdx.x;
  const int b = tile / (a.tiles_x * a.tiles_y);
  if (threadIdx.x < 81) swt[threadIdx.x] = a.w[b * a.wbs + threadIdx.x];
  const int tr = tile - b * a.tiles_x * a.tiles_y;
  const int y0 = (tr / a.tiles_x) * BT_H, x0 = (tr % a.tiles_x) * BT_W;
  // all tile copies in flight at once (LDGSTS); the ReLU mask is applied from the staged output tile afterwards
  float* sy = sg + C3 * BS_H * BS_W;       // [3][BS_H][BS_W] layer output (ReLU layers only)
  {
    // one row loop stages all three arrays of all three channels: row pointer, clamp and validity are computed once
    const int lane_ = threadIdx.x & 31, wid_ = threadIdx.x >> 5;
    int gxs[2];
    bool vxs[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int rx = lane_ + 32 * k, x = x0 - 1 + rx;
      vxs[k] = (x >= 0 && x < a.W);
      gxs[k] = min(max(x, 0), a.W - 1);
    }
    const float* xb = a.x + (long long)b * C3 * HW;
    const float* gb_ = a.dy + (long long)b * C3 * HW;
    const float* yb = a.y + (long long)b * C3 * HW;
    for (int ry = wid_; ry < BS_H; ry += BT_THREADS / 32) {
      const int y = y0 - 1 + ry;
      const bool vy = (y >= 0 && y < a.H);
      const long long ro = (long long)min(max(y, 0), a.H - 1) * a.W;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int rx = lane_ + 32 * k;
        if (rx < BS_W) {
          const long long o = ro + gxs[k];
          const int so = ry * BS_W + rx;
#pragma unroll
          for (int c = 0; c < C3; ++c) {
            cp_async4(sx + c * BS_H * BS_W + so, xb + c * HW + o, true);
            cp_async4(sg + c * BS_H * BS_W + so, gb_ + c * HW + o, vy && vxs[k]);
            if (a.relu) cp_async4(sy + c * BS_H * BS_W + so, yb + c * HW + o, true);
          }
        }
      }
    }
  }
  cp_async_wait_all();
  __syncthreads();
  if (a.relu) {
    for (int i = threadIdx.x; i < C3 * BS_H * BS_W; i += blockDim.x)
      if (!(sy[i] > 0.f)) sg[i] = 0.f;
    __syncthreads();
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int lx = lane, gx = x0 + lx;
  const int ly0 = wid * BP;
  float gw[81], gb[3];
#pragma unroll
  for (int i = 0; i < 81; ++i) gw[i] = 0.f;
  // ---- weight gradients: dW[o][c][ky][kx] += dpre[o](y,x) * xpad[c](y+ky-1, x+kx-1).  sg is zero outside the
  // image, so out-of-range pixels of a partial tile contribute nothing.  Column-wise: one column of BP + 2 inputs
  // per (c, kx) feeds 9 * BP FMAs.
  {
    float gpre[BP][C3];
#pragma unroll
    for (int oc = 0; oc < C3; ++oc) {
      float t = 0.f;
#pragma unroll
      for (int i = 0; i < BP; ++i) {
        gpre[i][oc] = sg[(oc * BS_H + ly0 + i + 1) * BS_W + lx + 1];
        t += gpre[i][oc];
      }
      gb[oc] = t;
    }
#pragma unroll
    for (int c = 0; c < C3; ++c)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        float v[BP + 2];
#pragma unroll
        for (int r = 0; r < BP + 2; ++r) v[r] = sx[(c * BS_H + ly0 + r) * BS_W + lx + kx];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
          for (int oc = 0; oc < C3; ++oc) {
            float t = gw[((oc * C3 + c) * 3 + ky) * 3 + kx];
#pragma unroll
            for (int i = 0; i < BP; ++i) t = fmaf(gpre[i][oc], v[i + ky], t);
            gw[((oc * C3 + c) * 3 + ky) * 3 + kx] = t;
          }
      }
  }
  // ---- fold the 84 sums: warp butterfly, then one shared atomic per warp and value
  {
    float v[32];
#pragma unroll
    for (int grp = 0; grp < 3; ++grp) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int e = grp * 32 + i;
        float t = 0.f;
        if (e < 81) t = gw[e < 81 ? e : 80];
        else if (e < 84) t = gb[(e >= 81 && e < 84) ? e - 81 : 0];
        v[i] = t;
      }
      const float r = warp_reduce_scatter32(v, lane);
      sacc[wid][grp * 32 + lane] = r;
    }
  }
  // ---- input gradient (the weight-gradient registers are free now; the 81 weights take their place):
  //   the padded input position (py,px) receives sum_{o,ky,kx} W[o][c][ky][kx] * dpre[o](py-ky+1, px-kx+1);
  //   replicate padding folds the padded ring onto the border pixels (slow path below).
  {
    float w[81];
#pragma unroll
    for (int i = 0; i < 81; ++i) w[i] = swt[i];
    float dxv[BP][C3];
#pragma unroll
    for (int i = 0; i < BP; ++i)
#pragma unroll
      for (int c = 0; c < C3; ++c) dxv[i][c] = 0.f;
#pragma unroll
    for (int oc = 0; oc < C3; ++oc)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        float gcol[BP + 2];
#pragma unroll
        for (int r = 0; r < BP + 2; ++r) gcol[r] = sg[(oc * BS_H + ly0 + r) * BS_W + lx + 2 - kx];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
          for (int c = 0; c < C3; ++c)
#pragma unroll
            for (int i = 0; i < BP; ++i)
              dxv[i][c] = fmaf(w[((oc * C3 + c) * 3 + ky) * 3 + kx], gcol[i + 2 - ky], dxv[i][c]);
      }
#pragma unroll
    for (int i = 0; i < BP; ++i) {
      const int ly = ly0 + i, gy = y0 + ly;
      if (gy >= a.H || gx >= a.W) continue;
      const bool interior = gy > 0 && gy < a.H - 1 && gx > 0 && gx < a.W - 1;
      if (!interior) {
        float t[C3] = {0.f, 0.f, 0.f};
        const int py_lo = (gy == 0) ? -1 : gy, py_hi = (gy == a.H - 1) ? a.H : gy;
        const int px_lo = (gx == 0) ? -1 : gx, px_hi = (gx == a.W - 1) ? a.W : gx;
        for (int py = py_lo; py <= py_hi; ++py)
          for (int px = px_lo; px <= px_hi; ++px)
            for (int ky = 0; ky < 3; ++ky)
              for (int kx = 0; kx < 3; ++kx) {
                const int ty = py - ky + 1 - y0 + 1, tx = px - kx + 1 - x0 + 1;  // tile-local (+1 halo)
                // outside the halo == outside the image (only reachable from the padded ring): contributes 0
                if (ty < 0 || ty >= BS_H || tx < 0 || tx >= BS_W) continue;
                for (int oc = 0; oc < C3; ++oc) {
                  const float g = sg[(oc * BS_H + ty) * BS_W + tx];
                  for (int c = 0; c < C3; ++c) t[c] = fmaf(swt[((oc * C3 + c) * 3 + ky) * 3 + kx], g, t[c]);
                }
              }
#pragma unroll
        for (int c = 0; c < C3; ++c) dxv[i][c] = t[c];
      }
#pragma unroll
      for (int c = 0; c < C3; ++c) {
        float v = dxv[i][c];
        if (a.residual) v += sg[(c * BS_H + ly + 1) * BS_W + lx + 1];
        a.dx[((long long)b * C3 + c) * HW + (long long)gy * a.W + gx] = v;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 84) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < BT_THREADS / 32; ++k) t += sacc[k][threadIdx.x];
    a.part[(long long)blockIdx.x * 84 + threadIdx.x] = t;
  }
}

#endif  // CCB_SLOT < 0

}  // namespace

constexpr int HEAD_FWD_NPAIR = 2;
constexpr int SYN_TAIL_FLOATS = 2 * 84;

#if CCB_SLOT >= 0
// ---- one weight slot: launchers of the constant-bank kernels + the device addresses of this unit's banks
int CCB_CAT(ccb_syn_slot_head_fwd_, CCB_SLOT)(int S, dim3 grid, const CcbHeadArgs& h, cudaStream_t s) {
  switch (S) {
    case 8: k_syn_head_fwd_c<8, HEAD_FWD_NPAIR><<<grid, HEAD2_THREADS, 0, s>>>(h); break;
    case 16: k_syn_head_fwd_c<16, HEAD_FWD_NPAIR><<<grid, HEAD2_THREADS, 0, s>>>(h); break;
    case 40: k_syn_head_fwd_c<40, HEAD_FWD_NPAIR><<<grid, HEAD2_THREADS, 0, s>>>(h); break;
    case 48: k_syn_head_fwd_c<48, HEAD_FWD_NPAIR><<<grid, HEAD2_THREADS, 0, s>>>(h); break;
    default: return CCB_ERR_UNSUPPORTED;
  }
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
int CCB_CAT(ccb_syn_slot_head_bwd_, CCB_SLOT)(int S, dim3 grid, const CcbHeadArgs& h, int n_groups, cudaStream_t s) {
  const size_t smc = (size_t)(8 * 32 * HB_ROW) * sizeof(float);
#define CCB_HEAD_BWD_C(SV)                                                                                          \
  case SV:                                                                                                          \
    CCB_CHECK(cudaFuncSetAttribute(k_syn_head_bwd<(SV + 15) / 16, true, SV>,                                        \
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smc));                         \
    k_syn_head_bwd<(SV + 15) / 16, true, SV><<<grid, 256, smc, s>>>(h, n_groups);                                   \
    break;
  switch (S) {
    CCB_HEAD_BWD_C(8)
    CCB_HEAD_BWD_C(16)
    CCB_HEAD_BWD_C(40)
    CCB_HEAD_BWD_C(48)
    default: return CCB_ERR_UNSUPPORTED;
  }
#undef CCB_HEAD_BWD_C
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
int CCB_CAT(ccb_syn_slot_tail2_, CCB_SLOT)(int grid, const CcbTail2Args& t, cudaStream_t s) {
  k_syn_tail2_fwd<<<grid, T2_THREADS, 0, s>>>(t);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
int CCB_CAT(ccb_syn_slot_banks_, CCB_SLOT)(float** head, float** tail) {  // of the CURRENT device
  CCB_CHECK(cudaGetSymbolAddress((void**)head, c_head_w));
  CCB_CHECK(cudaGetSymbolAddress((void**)tail, c_tail_w));
  return CCB_OK;
}
#else
// ---- public (library-internal) entry points
#define CCB_DECL_SLOT(k)                                                                                   \
  int ccb_syn_slot_head_fwd_##k(int S, dim3 grid, const CcbHeadArgs& h, cudaStream_t s);                  \
  int ccb_syn_slot_head_bwd_##k(int S, dim3 grid, const CcbHeadArgs& h, int n_groups, cudaStream_t s);    \
  int ccb_syn_slot_tail2_##k(int grid, const CcbTail2Args& t, cudaStream_t s);                            \
  int ccb_syn_slot_banks_##k(float** head, float** tail);
CCB_DECL_SLOT(0) CCB_DECL_SLOT(1) CCB_DECL_SLOT(2) CCB_DECL_SLOT(3)
CCB_DECL_SLOT(4) CCB_DECL_SLOT(5) CCB_DECL_SLOT(6) CCB_DECL_SLOT(7)
#undef CCB_DECL_SLOT
struct SynSlotFns {
  int (*head_fwd)(int, dim3, const CcbHeadArgs&, cudaStream_t);
  int (*head_bwd)(int, dim3, const CcbHeadArgs&, int, cudaStream_t);
  int (*tail2)(int, const CcbTail2Args&, cudaStream_t);
  int (*banks)(float**, float**);
};
#define CCB_SLOT_ROW(k) {ccb_syn_slot_head_fwd_##k, ccb_syn_slot_head_bwd_##k, ccb_syn_slot_tail2_##k, ccb_syn_slot_banks_##k}
static const SynSlotFns g_slot[CCB_N_WEIGHT_SLOTS] = {CCB_SLOT_ROW(0), CCB_SLOT_ROW(1), CCB_SLOT_ROW(2), CCB_SLOT_ROW(3),
                                                      CCB_SLOT_ROW(4), CCB_SLOT_ROW(5), CCB_SLOT_ROW(6), CCB_SLOT_ROW(7)};
#undef CCB_SLOT_ROW
static_assert(CCB_N_WEIGHT_SLOTS == 8, "one translation unit per slot: update build.sh and the table above");

namespace {
HeadArgs head_args(const ccb_plan* p, const ccb_tensors* prm, char* ws) {
  HeadArgs a = {};
  a.z = reinterpret_cast<const float*>(ws + p->off_up[p->g.L - 1]);
  a.w0 = prm->syn_w[0];
  a.b0 = prm->syn_b[0];
  a.w1 = prm->syn_w[1];
  a.b1 = prm->syn_b[1];
  a.S = p->syn_hidden;
  a.relu1 = p->syn[1].relu;
  a.HW = (long long)p->arch.H * p->arch.W;
  a.wbs = prm->weight_batch_stride;
  return a;
}

#if CCB_SLOT < 0
// ------------------------------------------------------------------------------------------------------------------
// Backward of BOTH trailing 3x3 layers in one launch (layers l2 = n - 2 and l3 = n - 1 of a four-layer synthesis).
// A CTA owns a 32x32 tile T of the final input gradient.  It needs d pre-activation of l2 on T grown by one pixel
// (R1, 34x34), i.e. the input gradient of l3 on R1, i.e. d pre-activation of l3 on T grown by two (R2, 36x36): the
// one-pixel ring of l3's input gradient is recomputed by the neighbouring CTAs as well (13 % more of the cheap part)
// instead of travelling through global memory between two launches.  Weight gradients are taken over T only.
// Shared memory: G3 [3][36][36] (zero outside the image), X3 = act[l2] on R1 (replicate-clamped; also l2's ReLU mask),
// X2 = act[l2 - 1] on R1 (clamped), G2 [3][34][34] (zero outside the image).
struct Conv2Args {
  const float* x2;  // [B,3,H,W] input of l2
  const float* x3;  // [B,3,H,W] output of l2 = input of l3
  const float* y3;  // [B,3,H,W] output of l3 (its ReLU mask; unused when !relu3)
  const float* dy;  // [B,3,H,W] gradient w.r.t. the output of l3
  const float* w2;
  const float* w3;
  long long wbs;
  int B, H, W, tiles_x, tiles_y;
  int res2, relu2, res3, relu3;
  float* dx;        // [B,3,H,W] gradient w.r.t. the input of l2
  float* part2;     // [tiles][84]
  float* part3;     // [tiles][84]
};
constexpr int F1 = BT_W + 2;  // 34: R1 edge
constexpr int F2 = BT_W + 4;  // 36: R2 edge
static_assert(BT_W == 32 && BT_H == 32 && BP * (BT_THREADS / 32) == BT_H, "fused 3x3 backward assumes 32x32 tiles");

// 81 + 3 weight-gradient sums of the thread's BP pixels of T: d pre-activation at G[oc][goff + ly][goff + lx] (G is GE
// wide), layer input at X[c][ly + ky][lx + kx] (R1, F1 wide); folded over the warp into row[0..95]
template <int GE>
__device__ __forceinline__ void conv3_dw_tile(const float* G, int goff, const float* X, int lane, int wid, float* row) {
  const int lx = lane, ly0 = wid * BP;
  float gw[81], gb[3];
#pragma unroll
  for (int i = 0; i < 81; ++i) gw[i] = 0.f;
  float gpre[BP][C3];
#pragma unroll
  for (int oc = 0; oc < C3; ++oc) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < BP; ++i) {
      gpre[i][oc] = G[(oc * GE + ly0 + i + goff) * GE + lx + goff];
      t += gpre[i][oc];
    }
    gb[oc] = t;
  }
#pragma unroll
  for (int c = 0; c < C3; ++c)
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      float v[BP + 2];
#pragma unroll
      for (int r = 0; r < BP + 2; ++r) v[r] = X[(c * F1 + ly0 + r) * F1 + lx + kx];
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int oc = 0; oc < C3; ++oc) {
          float t = gw[((oc * C3 + c) * 3 + ky) * 3 + kx];
#pragma unroll
          for (int i = 0; i < BP; ++i) t = fmaf(gpre[i][oc], v[i + ky], t);
          gw[((oc * C3 + c) * 3 + ky) * 3 + kx] = t;
        }
    }
  float v[32];
#pragma unroll
  for (int grp = 0; grp < 3; ++grp) {
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const int e = grp * 32 + i;
      float t = 0.f;
      if (e < 81) t = gw[e < 81 ? e : 80];
      else if (e < 84) t = gb[(e >= 81 && e < 84) ? e - 81 : 0];
      v[i] = t;
    }
    const float r = warp_reduce_scatter32(v, lane);
    row[grp * 32 + lane] = r;
  }
}

// input gradient (without the residual term) of the thread's BP pixels of a 32x32 block whose pixel (0, 0) sits at
// G[.][g0][g0]: sum over (oc, ky, kx) of W[oc][c][ky][kx] * G[oc][g0 + y + 1 - ky][g0 + x + 1 - kx]
template <int GE>
__device__ __forceinline__ void conv3_dx_block(const float* G, int g0, const float* swt, int lane, int wid,
                                               float (&dxv)[BP][C3]) {
  const int lx = lane, ly0 = wid * BP;
  float w[81];
#pragma unroll
  for (int i = 0; i < 81; ++i) w[i] = swt[i];
#pragma unroll
  for (int i = 0; i < BP; ++i)
#pragma unroll
    for (int c = 0; c < C3; ++c) dxv[i][c] = 0.f;
#pragma unroll
  for (int oc = 0; oc < C3; ++oc)
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      float gcol[BP + 2];
#pragma unroll
      for (int r = 0; r < BP + 2; ++r) gcol[r] = G[(oc * GE + g0 + ly0 + r - 1) * GE + g0 + lx + 1 - kx];
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int c = 0; c < C3; ++c)
#pragma unroll
          for (int i = 0; i < BP; ++i)
            dxv[i][c] = fmaf(w[((oc * C3 + c) * 3 + ky) * 3 + kx], gcol[i + 2 - ky], dxv[i][c]);
    }
}

// the same sum for ONE image pixel (gy, gx), generic: every padded position that replicate-clamps onto the pixel is
// included (image borders).  (oy, ox) = image coordinates of G's element (0, 0); G is zero outside the image and the
// sources inside the image all lie within G.
template <int GE>
__device__ __forceinline__ void conv3_dx_pixel(const float* G, int oy, int ox, const float* swt, int gy, int gx, int H,
                                               int W, float (&t)[C3]) {
#pragma unroll
  for (int c = 0; c < C3; ++c) t[c] = 0.f;
  const int py_lo = (gy == 0) ? -1 : gy, py_hi = (gy == H - 1) ? H : gy;
  const int px_lo = (gx == 0) ? -1 : gx, px_hi = (gx == W - 1) ? W : gx;
  for (int py = py_lo; py <= py_hi; ++py)
    for (int px = px_lo; px <= px_hi; ++px)
      for (int ky = 0; ky < 3; ++ky)
        for (int kx = 0; kx < 3; ++kx) {
          const int ly = py - ky + 1 - oy, lx = px - kx + 1 - ox;
          if (ly < 0 || ly >= GE || lx < 0 || lx >= GE) continue;
          for (int oc = 0; oc < C3; ++oc) {
            const float g = G[(oc * GE + ly) * GE + lx];
            for (int c = 0; c < C3; ++c) t[c] = fmaf(swt[((oc * C3 + c) * 3 + ky) * 3 + kx], g, t[c]);
          }
        }
}

__global__ void __launch_bounds__(BT_THREADS, 2) k_syn_conv3x2_bwd(Conv2Args a) {
  extern __shared__ __align__(16) float fsm[];
  float* G3 = fsm;                       // [3][F2][F2] d pre-activation of l3 on R2
  float* X3 = G3 + C3 * F2 * F2;         // [3][F1][F1] act[l2] on R1
  float* X2 = X3 + C3 * F1 * F1;         // [3][F1][F1] act[l2 - 1] on R1
  float* G2 = X2 + C3 * F1 * F1;         // [3][F1][F1] d pre-activation of l2 on R1
  __shared__ float sw3[C3 * C3 * 9], sw2[C3 * C3 * 9];
  __shared__ float sacc3[BT_THREADS / 32][96], sacc2[BT_THREADS / 32][96];
  const long long HW = (long long)a.H * a.W;
  const int tile = blockIdx.x;
  const int b = tile / (a.tiles_x * a.tiles_y);
  if (threadIdx.x < 81) {
    sw3[threadIdx.x] = a.w3[b * a.wbs + threadIdx.x];
    sw2[threadIdx.x] = a.w2[b * a.wbs + threadIdx.x];
  }
  const int tr = tile - b * a.tiles_x * a.tiles_y;
  const int y0 = (tr / a.tiles_x) * BT_H, x0 = (tr % a.tiles_x) * BT_W;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  // ---- staging: X3 / X2 on R1 (clamped), d y of l3 on R2 (zero outside the image)
  {
    const float* x3b = a.x3 + (long long)b * C3 * HW;
    const float* x2b = a.x2 + (long long)b * C3 * HW;
    const float* dyb = a.dy + (long long)b * C3 * HW;
    int gxs[2], gx2[2];
    bool vx2[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int rx = lane + 32 * k;
      gxs[k] = min(max(x0 - 1 + rx, 0), a.W - 1);
      const int x = x0 - 2 + rx;
      vx2[k] = (x >= 0 && x < a.W);
      gx2[k] = min(max(x, 0), a.W - 1);
    }
    for (int ry = wid; ry < F1; ry += BT_THREADS / 32) {
      const long long ro = (long long)min(max(y0 - 1 + ry, 0), a.H - 1) * a.W;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int rx = lane + 32 * k;
        if (rx < F1) {
#pragma unroll
          for (int c = 0; c < C3; ++c) {
            cp_async4(X3 + (c * F1 + ry) * F1 + rx, x3b + c * HW + ro + gxs[k], true);
            cp_async4(X2 + (c * F1 + ry) * F1 + rx, x2b + c * HW + ro + gxs[k], true);
          }
        }
      }
    }
    for (int qy = wid; qy < F2; qy += BT_THREADS / 32) {
      const int y = y0 - 2 + qy;
      const bool vy = (y >= 0 && y < a.H);
      const long long ro = (long long)min(max(y, 0), a.H - 1) * a.W;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int qx = lane + 32 * k;
        if (qx < F2) {
#pragma unroll
          for (int c = 0; c < C3; ++c) cp_async4(G3 + (c * F2 + qy) * F2 + qx, dyb + c * HW + ro + gx2[k], vy && vx2[k]);
        }
      }
    }
  }
  cp_async_wait_all();
  __syncthreads();
  if (a.relu3) {  // l3's own ReLU mask (not the HOP stack: its last layer is linear); plain loads, in-image elements only
    const float* y3b = a.y3 + (long long)b * C3 * HW;
    for (int e = threadIdx.x; e < C3 * F2 * F2; e += blockDim.x) {
      const int c = e / (F2 * F2), r = e - c * F2 * F2, qy = r / F2, qx = r - qy * F2;
      const int y = y0 - 2 + qy, x = x0 - 2 + qx;
      if (y >= 0 && y < a.H && x >= 0 && x < a.W && !(y3b[c * HW + (long long)y * a.W + x] > 0.f)) G3[e] = 0.f;
    }
    __syncthreads();
  }
  // ---- l3: weight gradients over T (T's pixel (0,0) is G3[.][2][2], X3[.][1][1])
  conv3_dw_tile<F2>(G3, 2, X3, lane, wid, sacc3[wid]);
  // ---- l3: input gradient on R1 -> d pre-activation of l2 (G2).  Block part: R1's top-left 32x32
  {
    float dxv[BP][C3];
    conv3_dx_block<F2>(G3, 1, sw3, lane, wid, dxv);
#pragma unroll
    for (int i = 0; i < BP; ++i) {
      const int ry = wid * BP + i, rx = lane;
      const int gy = y0 - 1 + ry, gx = x0 - 1 + rx;
      const bool inside = gy >= 0 && gy < a.H && gx >= 0 && gx < a.W;
      float t[C3] = {dxv[i][0], dxv[i][1], dxv[i][2]};
      if (inside && !(gy > 0 && gy < a.H - 1 && gx > 0 && gx < a.W - 1))
        conv3_dx_pixel<F2>(G3, y0 - 2, x0 - 2, sw3, gy, gx, a.H, a.W, t);
#pragma unroll
      for (int c = 0; c < C3; ++c) {
        float v = t[c];
        if (a.res3) v += G3[(c * F2 + ry + 1) * F2 + rx + 1];
        if (a.relu2 && !(X3[(c * F1 + ry) * F1 + rx] > 0.f)) v = 0.f;
        G2[(c * F1 + ry) * F1 + rx] = inside ? v : 0.f;
      }
    }
  }
  // the L-shaped rest of R1: rows 32, 33 (34 pixels each) and columns 32, 33 of rows 0..31
  for (int e = threadIdx.x; e < 2 * F1 + 2 * BT_H; e += blockDim.x) {
    int ry, rx;
    if (e < 2 * F1) {
      ry = BT_H + e / F1;
      rx = e % F1;
    } else {
      ry = (e - 2 * F1) >> 1;
      rx = BT_W + ((e - 2 * F1) & 1);
    }
    const int gy = y0 - 1 + ry, gx = x0 - 1 + rx;
    const bool inside = gy >= 0 && gy < a.H && gx >= 0 && gx < a.W;
    float t[C3] = {0.f, 0.f, 0.f};
    if (inside) conv3_dx_pixel<F2>(G3, y0 - 2, x0 - 2, sw3, gy, gx, a.H, a.W, t);
    for (int c = 0; c < C3; ++c) {
      float v = t[c];
      if (a.res3) v += G3[(c * F2 + ry + 1) * F2 + rx + 1];
      if (a.relu2 && !(X3[(c * F1 + ry) * F1 + rx] > 0.f)) v = 0.f;
      G2[(c * F1 + ry) * F1 + rx] = inside ? v : 0.f;
    }
  }
  __syncthreads();
  // ---- l2: weight gradients over T (G2[.][1][1], X2[.][1][1]) and the input gradient of T
  conv3_dw_tile<F1>(G2, 1, X2, lane, wid, sacc2[wid]);
  {
    float dxv[BP][C3];
    conv3_dx_block<F1>(G2, 1, sw2, lane, wid, dxv);
#pragma unroll
    for (int i = 0; i < BP; ++i) {
      const int ly = wid * BP + i, gy = y0 + ly, gx = x0 + lane;
      if (gy >= a.H || gx >= a.W) continue;
      float t[C3] = {dxv[i][0], dxv[i][1], dxv[i][2]};
      if (!(gy > 0 && gy < a.H - 1 && gx > 0 && gx < a.W - 1))
        conv3_dx_pixel<F1>(G2, y0 - 1, x0 - 1, sw2, gy, gx, a.H, a.W, t);
#pragma unroll
      for (int c = 0; c < C3; ++c) {
        float v = t[c];
        if (a.res2) v += G2[(c * F1 + ly + 1) * F1 + lane + 1];
        a.dx[((long long)b * C3 + c) * HW + (long long)gy * a.W + gx] = v;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 84) {
    float t3 = 0.f, t2 = 0.f;
#pragma unroll
    for (int k = 0; k < BT_THREADS / 32; ++k) {
      t3 += sacc3[k][threadIdx.x];
      t2 += sacc2[k][threadIdx.x];
    }
    a.part3[(long long)blockIdx.x * 84 + threadIdx.x] = t3;
    a.part2[(long long)blockIdx.x * 84 + threadIdx.x] = t2;
  }
}

#endif  // CCB_SLOT < 0 (fused 3x3 backward)

ConvArgs conv_args(const ccb_plan* p, const ccb_tensors* prm, int layer) {
  ConvArgs a = {};
  a.w = prm->syn_w[layer];
  a.b = prm->syn_b[layer];
  a.wbs = prm->weight_batch_stride;
  a.B = p->arch.B;
  a.H = p->arch.H;
  a.W = p->arch.W;
  a.residual = p->syn[layer].residual;
  a.relu = p->syn[layer].relu;
  a.tiles_x = ceil_div(a.W, CT_W);
  a.tiles_y = ceil_div(a.H, CT_H);
  return a;
}
}  // namespace

constexpr int HEAD_DZ_NPAIR = 2;
// every head kernel runs on a (CTAs per image, B) grid: a CTA never straddles two images, so one CTA needs one
// weight set (per-sample weights) and the pixel index needs no division
static dim3 head2_grid(const ccb_plan* p, int npair) {
  const long long hw = (long long)p->arch.H * p->arch.W;
  const long long per_cta = 2LL * npair * HEAD2_THREADS;
  return dim3((unsigned)((hw + per_cta - 1) / per_cta), (unsigned)p->arch.B);
}
// grid (= partial rows: gx * gy) of the head backward: persistent for S <= 64, one CTA per 1024 pixels otherwise
void ccb_syn_head_grid(const ccb_plan* p, int* gx, int* gy) {
  const long long hw = (long long)p->arch.H * p->arch.W;
  *gy = p->arch.B;
  if (p->syn_hidden > 64) {
    *gx = (int)((hw + 256LL * HEAD_P - 1) / (256LL * HEAD_P));
    return;
  }
  const long long groups = (hw + 255) / 256;
  long long cap = (long long)p->n_sm * CCB_HEAD_LB / p->arch.B;
  if (cap < 1) cap = 1;
  *gx = (int)(groups < cap ? groups : cap);
}
int ccb_syn_head_blocks(const ccb_plan* p) {
  int gx, gy;
  ccb_syn_head_grid(p, &gx, &gy);
  return gx * gy;
}
int ccb_syn_head_params(const ccb_plan* p) { return p->syn_hidden * CIN + p->syn_hidden + C3 * p->syn_hidden + C3; }
int ccb_syn_conv_blocks(const ccb_plan* p) {
  return p->arch.B * ceil_div(p->arch.W, BT_W) * ceil_div(p->arch.H, BT_H);
}

// the constant-bank kernels apply: a slot is given, a shipped head width, one weight set for the batch
static bool syn_const_path(const ccb_plan* p, const ccb_tensors* prm, int slot) {
  return slot >= 0 && slot < CCB_N_WEIGHT_SLOTS && prm->weight_batch_stride == 0 && head_const_ok(p->syn_hidden);
}

// Pack the head's weights (w0 transposed) into a workspace buffer and copy that into the constant bank of `slot`
// (no-op when the constant-bank kernels do not apply).  Stream ordered: one small kernel + device-to-device copies.
// The address of a __constant__ symbol differs per device: resolved with the plan's device current.
int ccb_syn_stage_weights(const ccb_plan* p, const ccb_tensors* prm, char* ws, cudaStream_t s, int slot) {
  const int S = p->syn_hidden;
  if (!syn_const_path(p, prm, slot)) return CCB_OK;
  float *hbase = nullptr, *tbase = nullptr;
  {
    int cur = -1;
    CCB_CHECK(cudaGetDevice(&cur));
    if (cur != p->device) CCB_CHECK(cudaSetDevice(p->device));
    const int rc = g_slot[slot].banks(&hbase, &tbase);
    if (cur != p->device) cudaSetDevice(cur);
    if (rc != CCB_OK) return rc;
  }
  if (p->arch.n_syn_layers == 4) {  // the two trailing 3x3 layers -> c_tail_w (read by k_syn_tail2_fwd)
    const float* w2 = prm->syn_w[2];
    if (prm->syn_b[2] == w2 + 81 && prm->syn_w[3] == w2 + 84 && prm->syn_b[3] == w2 + 165) {
      CCB_CHECK(cudaMemcpyAsync(tbase, w2, sizeof(float) * SYN_TAIL_FLOATS, cudaMemcpyDeviceToDevice, s));
    } else {
      for (int l = 0; l < 2; ++l) {
        CCB_CHECK(cudaMemcpyAsync(tbase + 84 * l, prm->syn_w[2 + l], sizeof(float) * 81, cudaMemcpyDeviceToDevice, s));
        CCB_CHECK(cudaMemcpyAsync(tbase + 84 * l + 81, prm->syn_b[2 + l], sizeof(float) * 3, cudaMemcpyDeviceToDevice, s));
      }
    }
  }
  float* packed = reinterpret_cast<float*>(ws + p->off_head_pack);
  k_head_pack<<<1, 256, 0, s>>>(prm->syn_w[0], prm->syn_b[0], prm->syn_w[1], prm->syn_b[1], S, packed);
  CCB_LAUNCH_CHECK();
  CCB_CHECK(cudaMemcpyAsync(hbase, packed, sizeof(float) * (S * CIN + S + C3 * S + C3), cudaMemcpyDeviceToDevice, s));
  return CCB_OK;
}

// Layer i >= 1 writes its output to ws act[i]; the last layer writes raw_out.
int ccb_syn_forward(const ccb_plan* p, const ccb_tensors* prm, char* ws, float* raw_out, cudaStream_t s, int slot,
                    bool weights_staged, SynLossFuse* loss) {
  const int nl = p->arch.n_syn_layers;
  const bool cpath = syn_const_path(p, prm, slot);
  HeadArgs h = head_args(p, prm, ws);
  h.out = (nl == 2) ? raw_out : reinterpret_cast<float*>(ws + p->off_act[1]);
  h.out2 = (nl == 2 && h.relu1) ? reinterpret_cast<float*>(ws + p->off_act[1]) : nullptr;
  const dim3 hgrid = head2_grid(p, HEAD_FWD_NPAIR);
  if (cpath) {
    if (!weights_staged) {
      int rc = ccb_syn_stage_weights(p, prm, ws, s, slot);
      if (rc != CCB_OK) return rc;
    }
    int rc = g_slot[slot].head_fwd(p->syn_hidden, hgrid, h, s);
    if (rc != CCB_OK) return rc;
  } else {
    size_t sm = (size_t)((p->syn_hidden + 1) & ~1) * HW2_ROW * sizeof(float);
    if (sm > 48 * 1024)
      CCB_CHECK(cudaFuncSetAttribute(k_syn_head_fwd<HEAD_FWD_NPAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    k_syn_head_fwd<HEAD_FWD_NPAIR><<<hgrid, HEAD2_THREADS, sm, s>>>(h);
    CCB_LAUNCH_CHECK();
  }
  if (loss && nl == 4 && cpath) {
    // training step: both 3x3 layers + clamp + squared error + its gradient in one launch
    Tail2Args t = {};
    t.x = reinterpret_cast<const float*>(ws + p->off_act[1]);
    t.mid = reinterpret_cast<float*>(ws + p->off_act[2]);
    t.relu1 = p->syn[3].relu;
    t.y = t.relu1 ? reinterpret_cast<float*>(ws + p->off_act[3]) : nullptr;  // only its ReLU mask is needed later
    t.B = p->arch.B; t.H = p->arch.H; t.W = p->arch.W;
    t.tiles_x = ceil_div(t.W, T2_OUT); t.tiles_y = ceil_div(t.H, T2_OUT);
    t.n_tiles = t.B * t.tiles_x * t.tiles_y;
    t.res0 = p->syn[2].residual; t.relu0 = p->syn[2].relu; t.res1 = p->syn[3].residual;
    t.tgt = loss->target; t.d_raw = loss->d_raw; t.part = loss->part; t.n_part = loss->n_part;
    t.inv_n_img = loss->inv_n_img; t.tail_counter = loss->tail_counter;
    const int grid = t.n_tiles < t.n_part ? t.n_tiles : t.n_part;
    int rc = g_slot[slot].tail2(grid, t, s);
    if (rc != CCB_OK) return rc;
    loss->done = 1;
    return CCB_OK;
  }
  for (int l = 2; l < nl; ++l) {
    ConvArgs c = conv_args(p, prm, l);
    c.x = reinterpret_cast<const float*>(ws + p->off_act[l - 1]);
    c.y = (l == nl - 1) ? raw_out : reinterpret_cast<float*>(ws + p->off_act[l]);
    c.y2 = (l == nl - 1 && c.relu) ? reinterpret_cast<float*>(ws + p->off_act[l]) : nullptr;
    k_syn_conv3_fwd<<<c.B * c.tiles_x * c.tiles_y, CT_THREADS, 0, s>>>(c);
    CCB_LAUNCH_CHECK();
  }
  return CCB_OK;
}

// d_raw_out -> d z (= d up_{L-1}) and the partial weight-gradient rows.
// The last layer's output is needed only for its ReLU mask; when the last layer has a ReLU
// the caller's raw_out is not retained, so forward also keeps a copy in ws act[nl-1].
int ccb_syn_backward(const ccb_plan* p, const ccb_tensors* prm, char* ws, const float* d_raw_out,
                     cudaStream_t s, cudaStream_t dw_stream, int slot, bool weights_staged) {
  const int nl = p->arch.n_syn_layers;
  const float* dcur = d_raw_out;
  static const bool no_fuse = (getenv("CCB_NO_CONV3_FUSE") != nullptr);  // A/B switch: one launch per 3x3 layer
  const bool fuse2 = (nl == 4) && !no_fuse;
  if (fuse2) {
    // both trailing 3x3 layers in one launch (k_syn_conv3x2_bwd)
    Conv2Args c = {};
    c.x2 = reinterpret_cast<const float*>(ws + p->off_act[1]);
    c.x3 = reinterpret_cast<const float*>(ws + p->off_act[2]);
    c.y3 = reinterpret_cast<const float*>(ws + p->off_act[3]);
    c.dy = d_raw_out;
    c.w2 = prm->syn_w[2];
    c.w3 = prm->syn_w[3];
    c.wbs = prm->weight_batch_stride;
    c.B = p->arch.B; c.H = p->arch.H; c.W = p->arch.W;
    c.tiles_x = ceil_div(c.W, BT_W);
    c.tiles_y = ceil_div(c.H, BT_H);
    c.res2 = p->syn[2].residual; c.relu2 = p->syn[2].relu;
    c.res3 = p->syn[3].residual; c.relu3 = p->syn[3].relu;
    c.dx = reinterpret_cast<float*>(ws + p->off_dact[0]);
    c.part2 = reinterpret_cast<float*>(ws + p->off_syn_part[2]);
    c.part3 = reinterpret_cast<float*>(ws + p->off_syn_part[3]);
    const size_t fsm = (size_t)C3 * (F2 * F2 + 3 * F1 * F1) * sizeof(float);
    CCB_CHECK(cudaFuncSetAttribute(k_syn_conv3x2_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsm));
    if (!ccb_debug_skip(8)) k_syn_conv3x2_bwd<<<ccb_syn_conv_blocks(p), BT_THREADS, fsm, s>>>(c);
    CCB_LAUNCH_CHECK();
    dcur = c.dx;
  }
  for (int l = nl - 1; l >= 2 && !fuse2; --l) {
    ConvArgs c = conv_args(p, prm, l);
    c.x = reinterpret_cast<const float*>(ws + p->off_act[l - 1]);
    c.y = reinterpret_cast<float*>(ws + p->off_act[l]);
    c.dy = dcur;
    c.dx = reinterpret_cast<float*>(ws + p->off_dact[(l & 1)]);
    c.part = reinterpret_cast<float*>(ws + p->off_syn_part[l]);
    c.tiles_x = ceil_div(c.W, BT_W);
    c.tiles_y = ceil_div(c.H, BT_H);
    const size_t csm = (size_t)3 * C3 * BS_H * BS_W * sizeof(float);
    CCB_CHECK(cudaFuncSetAttribute(k_syn_conv3_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csm));
    if (!ccb_debug_skip(8)) k_syn_conv3_bwd<<<ccb_syn_conv_blocks(p), BT_THREADS, csm, s>>>(c);
    CCB_LAUNCH_CHECK();
    dcur = c.dx;
  }
  HeadArgs h = head_args(p, prm, ws);
  h.d_out = dcur;
  h.act = reinterpret_cast<const float*>(ws + p->off_act[1]);
  h.dz = reinterpret_cast<float*>(ws + p->off_dup[p->g.L - 1]);
  h.part = reinterpret_cast<float*>(ws + p->off_syn_part[0]);
  const int S = p->syn_hidden;
  int gx, gy;
  ccb_syn_head_grid(p, &gx, &gy);
  const dim3 bgrid(gx, gy);
  if (S > 64) {
    size_t sm = (size_t)(S * HW_ROW + ccb_syn_head_params(p)) * sizeof(float);
    k_syn_head_bwd_generic<<<bgrid, 256, sm, s>>>(h);  // wide heads: combined kernel, one stream
    CCB_LAUNCH_CHECK();
    return CCB_OK;
  }
  const int nchunk = (S + 15) / 16;
  const long long hw = (long long)p->arch.H * p->arch.W;
  const int n_groups = (int)((hw + 255) / 256);  // per image
  if (syn_const_path(p, prm, slot) && !dw_stream) {
    if (!weights_staged) {
      int rc = ccb_syn_stage_weights(p, prm, ws, s, slot);
      if (rc != CCB_OK) return rc;
    }
    if (ccb_debug_skip(16)) return CCB_OK;
    return g_slot[slot].head_bwd(S, bgrid, h, n_groups, s);
  }
  const size_t sm = (size_t)(16 * nchunk * HW_ROW + 8 * 32 * HB_ROW) * sizeof(float);
  if (dw_stream) {
    // d z on the critical stream, the weight gradients beside it
    CCB_CHECK(cudaEventRecord(p->ev_head, s));
    k_syn_head_dz<HEAD_DZ_NPAIR><<<head2_grid(p, HEAD_DZ_NPAIR), HEAD2_THREADS, (size_t)((S + 1) & ~1) * HW2_ROW * sizeof(float), s>>>(h);
    CCB_LAUNCH_CHECK();
    CCB_CHECK(cudaStreamWaitEvent(dw_stream, p->ev_head, 0));
  }
  cudaStream_t ks = dw_stream ? dw_stream : s;
#define CCB_HEAD_BWD(NC)                                                                                              \
  case NC:                                                                                                            \
    if (dw_stream) {                                                                                                  \
      CCB_CHECK(cudaFuncSetAttribute(k_syn_head_bwd<NC, false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); \
      k_syn_head_bwd<NC, false, 0><<<bgrid, 256, sm, ks>>>(h, n_groups);                                               \
    } else {                                                                                                          \
      CCB_CHECK(cudaFuncSetAttribute(k_syn_head_bwd<NC, true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); \
      k_syn_head_bwd<NC, true, 0><<<bgrid, 256, sm, ks>>>(h, n_groups);                                                \
    }                                                                                                                 \
    break;
  switch (nchunk) {
    CCB_HEAD_BWD(1)
    CCB_HEAD_BWD(2)
    CCB_HEAD_BWD(3)
    CCB_HEAD_BWD(4)
    default: return CCB_ERR_UNSUPPORTED;
  }
#undef CCB_HEAD_BWD
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
#endif  // CCB_SLOT
