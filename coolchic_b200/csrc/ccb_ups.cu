// U1 / U2 / U3: learned separable symmetric upsampling pyramid, forward and backward.
// Restates enc/component/core/upsampling.py:
//   :21-89   _Parameterization_Symmetric_1d  (a b c d -> a b c d d c b a / a b c d c b a)
//   :92-209  UpsamplingSeparableSymmetricConv2d        y = conv2d(x, w (x) w, zero pad) + x
//   :212-355 UpsamplingSeparableSymmetricConvTranspose2d  replicate pad k/2, conv_transpose2d
//            stride 2, crop 3k/2-1  ==  the 2-phase polyphase filter
//              y[n] = sum_{t : (n + k/2 - 1 - t) even} w[t] * x[clamp((n + k/2 - 1 - t) / 2)]
//   :476-506 Upsampling.forward  (stage idx: up <- cat(refine(y_{L-2-idx}), crop(upsample(up))))
// One launch per pyramid stage; a CTA produces one 32x32 output tile of one channel with both
// separable passes done through shared memory (no intermediate in HBM).
#include <algorithm>
#include "ccb_internal.cuh"

namespace {

constexpr int UT = 32;         // output tile edge
constexpr int UI = 16;         // input rows/cols under one output tile
constexpr int UH = 2;          // input halo each side (k <= 8)
constexpr int UIT = UI + 2 * UH;   // 20
constexpr int RH = 3;          // refine halo (k <= 7)
constexpr int RIT = UT + 2 * RH;   // 38
constexpr int DYT = 38;        // dY tile edge in the upsample backward; origin 2*r0 - 3
constexpr int UPS_THREADS = 512;  // every stage is latency-bound: more threads = shorter per-CTA critical path
constexpr int UPS_PART = 16;   // floats per partial row: [0..7] transposed-conv taps, [8..14] refine taps

struct UpsArgs {
  // source of the upsample branch: [B, c_in, hs, ws]
  const float* src;
  long long src_bs, src_cs;  // batch / channel strides (floats)
  int c_in, hs, ws;
  // target latent of the refine branch: [B, ht, wt] inside the flat yhat
  const float* tgt;
  long long tgt_bs;
  int ht, wt;
  // destination [B, c_in + 1, ht, wt]
  float* dst;
  int B;
  int K, KP;        // transposed-conv / refine kernel sizes
  const float* w_up;   // [K/2]
  const float* w_pre;  // [(KP+1)/2]
  long long wbs;       // weight batch stride (floats); 0: one filter set for the whole batch
  // backward
  const float* d_dst;  // [B, c_in+1, ht, wt]
  float* d_src;        // [B, c_in, hs, ws] with strides d_src_bs / d_src_cs
  long long d_src_bs, d_src_cs;
  float* d_tgt;        // [B, ht, wt] with batch stride tgt_bs (inside dyhat)
  float* part;         // [n_blocks][UPS_PART]
};

__device__ __forceinline__ void load_filters(const UpsArgs& a, float* wu, float* wp, int b) {
  if (threadIdx.x < 8) {
    int t = threadIdx.x;
    const long long wo = (long long)b * a.wbs;
    wu[t] = (t < a.K) ? a.w_up[wo + (t < a.K / 2 ? t : a.K - 1 - t)] : 0.f;
    int m = a.KP - 1 - t;
    wp[t] = (t < a.KP) ? a.w_pre[wo + (t < m ? t : m)] : 0.f;
  }
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// ------------------------------------------------------------------------------------------
// Forward
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(UPS_THREADS) k_ups_fwd(UpsArgs a) {
  __shared__ float s_in[RIT * RIT];
  __shared__ float s_t[RIT * UT];
  __shared__ float wu[8], wp[8];
  const int ch = blockIdx.z % (a.c_in + 1);
  const int b = blockIdx.z / (a.c_in + 1);
  const int oy0 = blockIdx.y * UT, ox0 = blockIdx.x * UT;
  if (oy0 >= a.ht || ox0 >= a.wt) return;
  load_filters(a, wu, wp, b);
  float* out = a.dst + ((long long)(b * (a.c_in + 1) + ch) * a.ht) * a.wt;

  if (ch == 0) {
    // ---- refine: zero-padded separable odd filter + residual
    const int pad = a.KP / 2;
    const float* src = a.tgt + (long long)b * a.tgt_bs;
    for (int i = threadIdx.x; i < RIT * RIT; i += blockDim.x) {
      int ry = i / RIT, rx = i - ry * RIT;
      int gy = oy0 - RH + ry, gx = ox0 - RH + rx;
      s_in[i] = (gy >= 0 && gy < a.ht && gx >= 0 && gx < a.wt) ? src[(long long)gy * a.wt + gx] : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < RIT * UT; i += blockDim.x) {
      int ry = i / UT, x = i - ry * UT;
      float acc = 0.f;
      for (int t = 0; t < a.KP; ++t) acc = fmaf(wp[t], s_in[ry * RIT + x + RH + t - pad], acc);
      s_t[i] = acc;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < UT * UT; i += blockDim.x) {
      int y = i / UT, x = i - y * UT;
      if (oy0 + y >= a.ht || ox0 + x >= a.wt) continue;
      float acc = 0.f;
      for (int t = 0; t < a.KP; ++t) acc = fmaf(wp[t], s_t[(y + RH + t - pad) * UT + x], acc);
      out[(long long)(oy0 + y) * a.wt + ox0 + x] = acc + s_in[(y + RH) * RIT + x + RH];
    }
    return;
  }

  // ---- upsample x2: replicate-clamped polyphase filter
  const float* src = a.src + (long long)b * a.src_bs + (long long)(ch - 1) * a.src_cs;
  const int iy0 = oy0 / 2 - UH, ix0 = ox0 / 2 - UH;
  for (int i = threadIdx.x; i < UIT * UIT; i += blockDim.x) {
    int ry = i / UIT, rx = i - ry * UIT;
    int gy = clampi(iy0 + ry, 0, a.hs - 1), gx = clampi(ix0 + rx, 0, a.ws - 1);
    s_in[i] = src[(long long)gy * a.ws + gx];
  }
  __syncthreads();
  const int kh = a.K / 2 - 1;
  for (int i = threadIdx.x; i < UIT * UT; i += blockDim.x) {
    int ry = i / UT, x = i - ry * UT;
    float acc = 0.f;
    for (int t = (x + kh) & 1; t < a.K; t += 2) {
      int ax = (ox0 + x + kh - t) >> 1;  // arithmetic shift: floor for negatives (numerator even)
      acc = fmaf(wu[t], s_in[ry * UIT + ax - ix0], acc);
    }
    s_t[i] = acc;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < UT * UT; i += blockDim.x) {
    int y = i / UT, x = i - y * UT;
    if (oy0 + y >= a.ht || ox0 + x >= a.wt) continue;
    float acc = 0.f;
    for (int t = (y + kh) & 1; t < a.K; t += 2) {
      int ay = (oy0 + y + kh - t) >> 1;
      acc = fmaf(wu[t], s_t[(ay - iy0) * UT + x], acc);
    }
    out[(long long)(oy0 + y) * a.wt + ox0 + x] = acc;
  }
}

// ------------------------------------------------------------------------------------------
// Backward
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(UPS_THREADS) k_ups_bwd(UpsArgs a) {
  __shared__ float s_a[DYT * DYT];   // dY tile (zero filled)
  __shared__ float s_b[RIT * RIT];   // refine: x tile ; upsample: input tile (UIT*UIT)
  __shared__ float s_c[RIT * UT];    // refine: t = H(x) [38][32] ; upsample: T = H(in) [20][32]
  __shared__ float s_d[UT * DYT];    // refine: dt = V(dY) [32][38] ; upsample: dT = V^T(dY) [16][38]
  __shared__ float wu[8], wp[8];
  __shared__ float red[32];
  const int ch = blockIdx.z % (a.c_in + 1);
  const int b = blockIdx.z / (a.c_in + 1);
  const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  float* prow = a.part + (long long)bid * UPS_PART;
  load_filters(a, wu, wp, b);
  const float* dY = a.d_dst + ((long long)(b * (a.c_in + 1) + ch) * a.ht) * a.wt;
  float gsum[8];
#pragma unroll
  for (int t = 0; t < 8; ++t) gsum[t] = 0.f;

  if (ch == 0) {
    const int oy0 = blockIdx.y * UT, ox0 = blockIdx.x * UT;
    if (oy0 >= a.ht || ox0 >= a.wt) {
      if (threadIdx.x < UPS_PART) prow[threadIdx.x] = 0.f;
      return;
    }
    const int pad = a.KP / 2;
    const float* xin = a.tgt + (long long)b * a.tgt_bs;
    for (int i = threadIdx.x; i < RIT * RIT; i += blockDim.x) {
      int ry = i / RIT, rx = i - ry * RIT;
      int gy = oy0 - RH + ry, gx = ox0 - RH + rx;
      bool in = (gy >= 0 && gy < a.ht && gx >= 0 && gx < a.wt);
      s_a[i] = in ? dY[(long long)gy * a.wt + gx] : 0.f;
      s_b[i] = in ? xin[(long long)gy * a.wt + gx] : 0.f;
    }
    __syncthreads();
    // t = H(x): rows [oy0-3, oy0+35), tile columns
    for (int i = threadIdx.x; i < RIT * UT; i += blockDim.x) {
      int ry = i / UT, x = i - ry * UT;
      float acc = 0.f;
      for (int t = 0; t < a.KP; ++t) acc = fmaf(wp[t], s_b[ry * RIT + x + RH + t - pad], acc);
      s_c[i] = acc;
    }
    // dt = V(dY): tile rows, columns [ox0-3, ox0+35)
    for (int i = threadIdx.x; i < UT * RIT; i += blockDim.x) {
      int y = i / RIT, rx = i - y * RIT;
      float acc = 0.f;
      for (int t = 0; t < a.KP; ++t) acc = fmaf(wp[t], s_a[(y + RH + t - pad) * RIT + rx], acc);
      s_d[i] = acc;
    }
    __syncthreads();
    float* dT = a.d_tgt + (long long)b * a.tgt_bs;
    for (int i = threadIdx.x; i < UT * UT; i += blockDim.x) {
      int y = i / UT, x = i - y * UT;
      if (oy0 + y >= a.ht || ox0 + x >= a.wt) continue;
      // d x = dY + H(dt)   (the filter is symmetric => self-adjoint under zero padding)
      float acc = 0.f;
      for (int t = 0; t < a.KP; ++t) acc = fmaf(wp[t], s_d[y * RIT + x + RH + t - pad], acc);
      const float dy = s_a[(y + RH) * RIT + x + RH];
      dT[(long long)(oy0 + y) * a.wt + ox0 + x] = acc + dy;
      // filter gradients: vertical taps see t, horizontal taps see dt (x) x
      const float dtv = s_d[y * RIT + x + RH];
      for (int t = 0; t < a.KP; ++t) {
        gsum[t] = fmaf(dy, s_c[(y + RH + t - pad) * UT + x], gsum[t]);
        gsum[t] = fmaf(dtv, s_b[(y + RH) * RIT + x + RH + t - pad], gsum[t]);
      }
    }
    {
      float tot[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) tot[t] = block_sum(gsum[t], red);
      if (threadIdx.x == 0) {
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int m = a.KP - 1 - t;
          prow[t] = 0.f;
          prow[4 + t] = (t > m) ? 0.f : (t == m ? tot[t] : tot[t] + tot[m]);
        }
      }
    }
    return;
  }

  // ---- upsample branch: tile over the SOURCE resolution (16x16 inputs <-> 32x32 outputs)
  const int r0 = blockIdx.y * UI, c0 = blockIdx.x * UI;
  if (r0 >= a.hs || c0 >= a.ws) {
    if (threadIdx.x < UPS_PART) prow[threadIdx.x] = 0.f;
    return;
  }
  const int kh = a.K / 2 - 1;
  const float* src = a.src + (long long)b * a.src_bs + (long long)(ch - 1) * a.src_cs;
  const int iy0 = r0 - UH, ix0 = c0 - UH;
  const int ny0 = 2 * r0 - 3, nx0 = 2 * c0 - 3;
  for (int i = threadIdx.x; i < UIT * UIT; i += blockDim.x) {
    int ry = i / UIT, rx = i - ry * UIT;
    int gy = clampi(iy0 + ry, 0, a.hs - 1), gx = clampi(ix0 + rx, 0, a.ws - 1);
    s_b[i] = src[(long long)gy * a.ws + gx];
  }
  for (int i = threadIdx.x; i < DYT * DYT; i += blockDim.x) {
    int ry = i / DYT, rx = i - ry * DYT;
    int gy = ny0 + ry, gx = nx0 + rx;
    s_a[i] = (gy >= 0 && gy < a.ht && gx >= 0 && gx < a.wt) ? dY[(long long)gy * a.wt + gx] : 0.f;
  }
  __syncthreads();
  // T = H(in): rows of the input tile (+halo), output columns [2c0, 2c0+32)
  for (int i = threadIdx.x; i < UIT * UT; i += blockDim.x) {
    int ry = i / UT, x = i - ry * UT;
    float acc = 0.f;
    for (int t = (x + kh) & 1; t < a.K; t += 2) {
      int ax = clampi((2 * c0 + x + kh - t) >> 1, 0, a.ws - 1);
      // clamped column -> position in the (clamped-loaded) tile: any column that clamps to
      // the same source holds the same value, so index by the unclamped offset
      int ux = ((2 * c0 + x + kh - t) >> 1) - ix0;
      (void)ax;
      acc = fmaf(wu[t], s_b[ry * UIT + ux], acc);
    }
    s_c[i] = acc;
  }
  // dT = V^T(dY): input rows [r0, r0+16), output columns [2c0-3, 2c0+35)
  for (int i = threadIdx.x; i < UI * DYT; i += blockDim.x) {
    int ry = i / DYT, rx = i - ry * DYT;
    const int r = r0 + ry;
    float acc = 0.f;
    if (r < a.hs) {
      const bool edge = (r == 0) || (r == a.hs - 1);
      int n_lo = 2 * r - kh, n_hi = 2 * r + a.K / 2;  // standard support
      if (!edge) {
        for (int n = n_lo; n <= n_hi; ++n) {
          int t = n - 2 * r + kh;  // the single tap with (n + kh - t) / 2 == r
          acc = fmaf(wu[t], s_a[(n - ny0) * DYT + rx], acc);
        }
      } else {
        for (int n = n_lo; n <= n_hi; ++n) {
          if (n < 0 || n >= a.ht) continue;
          for (int t = (n + kh) & 1; t < a.K; t += 2) {
            int ay = clampi((n + kh - t) >> 1, 0, a.hs - 1);
            if (ay == r) acc = fmaf(wu[t], s_a[(n - ny0) * DYT + rx], acc);
          }
        }
      }
    }
    s_d[i] = acc;
  }
  __syncthreads();
  // d in = H^T(dT)
  float* dS = a.d_src + (long long)b * a.d_src_bs + (long long)(ch - 1) * a.d_src_cs;
  for (int i = threadIdx.x; i < UI * UI; i += blockDim.x) {
    int ry = i / UI, rx = i - ry * UI;
    const int r = r0 + ry, c = c0 + rx;
    if (r >= a.hs || c >= a.ws) continue;
    float acc = 0.f;
    const bool edge = (c == 0) || (c == a.ws - 1);
    int n_lo = 2 * c - kh, n_hi = 2 * c + a.K / 2;
    if (!edge) {
      for (int n = n_lo; n <= n_hi; ++n) {
        int t = n - 2 * c + kh;
        acc = fmaf(wu[t], s_d[ry * DYT + n - nx0], acc);
      }
    } else {
      for (int n = n_lo; n <= n_hi; ++n) {
        if (n < 0 || n >= a.wt) continue;
        for (int t = (n + kh) & 1; t < a.K; t += 2) {
          int ax = clampi((n + kh - t) >> 1, 0, a.ws - 1);
          if (ax == c) acc = fmaf(wu[t], s_d[ry * DYT + n - nx0], acc);
        }
      }
    }
    dS[(long long)r * a.ws + c] = acc;
  }
  // filter gradients
  //   vertical taps:   sum_{n,nx in out tile} dY[n][nx] * T[a(n,t)][nx]
  for (int i = threadIdx.x; i < UT * UT; i += blockDim.x) {
    int y = i / UT, x = i - y * UT;
    const int n = 2 * r0 + y, nx = 2 * c0 + x;
    if (n >= a.ht || nx >= a.wt) continue;
    const float dy = s_a[(n - ny0) * DYT + nx - nx0];
    for (int t = (y + kh) & 1; t < a.K; t += 2) {
      int ay = ((n + kh - t) >> 1) - iy0;
      gsum[t] = fmaf(dy, s_c[ay * UT + x], gsum[t]);
    }
  }
  //   horizontal taps: sum_{r in tile, nx in out tile} dT[r][nx] * in[r][a(nx,t)]
  for (int i = threadIdx.x; i < UI * UT; i += blockDim.x) {
    int ry = i / UT, x = i - ry * UT;
    const int r = r0 + ry, nx = 2 * c0 + x;
    if (r >= a.hs || nx >= a.wt) continue;
    const float dt = s_d[ry * DYT + nx - nx0];
    for (int t = (x + kh) & 1; t < a.K; t += 2) {
      int ux = ((nx + kh - t) >> 1) - ix0;
      gsum[t] = fmaf(dt, s_b[(ry + UH) * UIT + ux], gsum[t]);
    }
  }
  {
    float tot[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) tot[t] = block_sum(gsum[t], red);
    if (threadIdx.x == 0) {
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        prow[t] = (t < a.K / 2) ? tot[t] + tot[a.K - 1 - t] : 0.f;
        prow[4 + t] = 0.f;
      }
    }
  }
}

// ==========================================================================================
// Register-tiled kernels for the standard filter sizes (K = 8, KP = 7).
// The generic kernels above spend ~7 warp instructions per output value on index arithmetic and
// shared-memory round trips; here a thread owns a 4x4 output block (2x2 source block), keeps the
// separable intermediate in registers and every tap loop is unrolled at compile time.
//   CTA = 256 threads = 16x16 thread grid = one 64x64 output tile of one channel.
// ==========================================================================================
constexpr int FT = 64;        // output tile edge
constexpr int FXS = 36;       // source tile edge (32 + 2x2 halo), also its row stride
constexpr int FRN = 70;       // refine / dY tile edge (64 + 2x3 halo)
constexpr int FRS = 72;       // its row stride (16-byte aligned rows)

struct Taps8 {
  float w0, w1, w2, w3;       // a b c d of a b c d d c b a
};

// 4 polyphase outputs (2 source positions x 2 phases) from 6 consecutive source samples s[0..5],
// where s[2] is the first of the two source positions.
__device__ __forceinline__ void poly4(const Taps8& w, float s0, float s1, float s2, float s3, float s4, float s5,
                                      float* o) {
  o[0] = fmaf(w.w1, s3, fmaf(w.w3, s2, fmaf(w.w2, s1, w.w0 * s0)));
  o[1] = fmaf(w.w0, s4, fmaf(w.w2, s3, fmaf(w.w3, s2, w.w1 * s1)));
  o[2] = fmaf(w.w1, s4, fmaf(w.w3, s3, fmaf(w.w2, s2, w.w0 * s1)));
  o[3] = fmaf(w.w0, s5, fmaf(w.w2, s4, fmaf(w.w3, s3, w.w1 * s2)));
}

__device__ __forceinline__ void load_row10(const float* row, float* d) {
  const float4 a = *reinterpret_cast<const float4*>(row);
  const float4 b = *reinterpret_cast<const float4*>(row + 4);
  const float2 c = *reinterpret_cast<const float2*>(row + 8);
  d[0] = a.x; d[1] = a.y; d[2] = a.z; d[3] = a.w;
  d[4] = b.x; d[5] = b.y; d[6] = b.z; d[7] = b.w;
  d[8] = c.x; d[9] = c.y;
}

__device__ __forceinline__ void load_row6(const float* row, float* d) {
  const float2 a = *reinterpret_cast<const float2*>(row);
  const float2 b = *reinterpret_cast<const float2*>(row + 2);
  const float2 c = *reinterpret_cast<const float2*>(row + 4);
  d[0] = a.x; d[1] = a.y; d[2] = b.x; d[3] = b.y; d[4] = c.x; d[5] = c.y;
}

// zero-padded [FRN x FRN] tile of a [h, w] plane with origin (y0, x0) -> smem (stride FRS); needs cp_async_wait_all
__device__ __forceinline__ void load_tile_zero(float* sm, const float* __restrict__ g, int h, int w, int y0, int x0) {
  stage_tile<FRN, false>(sm, FRS, g, h, w, y0, x0, FRN);
}

// replicate-clamped [FXS x FXS] tile
__device__ __forceinline__ void load_tile_clamp(float* sm, const float* __restrict__ g, int h, int w, int y0, int x0) {
  stage_tile<FXS, true>(sm, FXS, g, h, w, y0, x0, FXS);
}

__device__ __forceinline__ void store_block4x4(float* out, int h, int w, int y, int x, const float (*v)[4]) {
  const bool vec = ((w & 3) == 0) && (x + 3 < w) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (y + i >= h) break;
    float* o = out + (long long)(y + i) * w + x;
    if (vec) {
      *reinterpret_cast<float4*>(o) = make_float4(v[i][0], v[i][1], v[i][2], v[i][3]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (x + j < w) o[j] = v[i][j];
    }
  }
}

// One 64x64 output tile of one channel (bz = b * (c_in + 1) + ch) of one pyramid stage.
// (oy0, ox0): origin of the tile in the stage's output, multiples of 4 (the chain kernel below places the tiles
// of the early stages where the later stage needs them).
__device__ __forceinline__ void ups_fwd8_tile_at(const UpsArgs& a, float* s_in, int oy0, int ox0, int ch, int b) {
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float* out = a.dst + ((long long)(b * (a.c_in + 1) + ch) * a.ht) * a.wt;
  float res[4][4];

  if (ch == 0) {
    // ---- refine: y = x + (p (x) p) * x, zero padded, p = p0 p1 p2 p3 p2 p1 p0
    const float* wpre = a.w_pre + (long long)b * a.wbs;
    const float p0 = wpre[0], p1 = wpre[1], p2 = wpre[2], p3 = wpre[3];
    const float pt[7] = {p0, p1, p2, p3, p2, p1, p0};
    load_tile_zero(s_in, a.tgt + (long long)b * a.tgt_bs, a.ht, a.wt, oy0 - 3, ox0 - 3);
    cp_async_wait_all();
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) res[i][j] = 0.f;
#pragma unroll
    for (int rr = 0; rr < 10; ++rr) {
      float d[10], hx[4];
      load_row10(&s_in[(4 * ty + rr) * FRS + 4 * tx], d);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float acc = 0.f;
#pragma unroll
        for (int t = 0; t < 7; ++t) acc = fmaf(pt[t], d[j + t], acc);
        hx[j] = acc;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int t = rr - i;  // vertical tap that maps window row rr onto output row i
        if (t >= 0 && t < 7) {
#pragma unroll
          for (int j = 0; j < 4; ++j) res[i][j] = fmaf(pt[t], hx[j], res[i][j]);
        }
        if (rr == i + 3) {
#pragma unroll
          for (int j = 0; j < 4; ++j) res[i][j] += d[j + 3];  // residual
        }
      }
    }
    store_block4x4(out, a.ht, a.wt, oy0 + 4 * ty, ox0 + 4 * tx, res);
    return;
  }

  // ---- upsample x2
  const float* wup = a.w_up + (long long)b * a.wbs;
  Taps8 w = {wup[0], wup[1], wup[2], wup[3]};
  const float* src = a.src + (long long)b * a.src_bs + (long long)(ch - 1) * a.src_cs;
  load_tile_clamp(s_in, src, a.hs, a.ws, oy0 / 2 - 2, ox0 / 2 - 2);
  cp_async_wait_all();
  __syncthreads();
  float h[6][4];
#pragma unroll
  for (int rr = 0; rr < 6; ++rr) {
    float d[6];
    load_row6(&s_in[(2 * ty + rr) * FXS + 2 * tx], d);
    poly4(w, d[0], d[1], d[2], d[3], d[4], d[5], h[rr]);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float o[4];
    poly4(w, h[0][j], h[1][j], h[2][j], h[3][j], h[4][j], h[5][j], o);
    res[0][j] = o[0]; res[1][j] = o[1]; res[2][j] = o[2]; res[3][j] = o[3];
  }
  store_block4x4(out, a.ht, a.wt, oy0 + 4 * ty, ox0 + 4 * tx, res);
}

__device__ __forceinline__ void ups_fwd8_tile(const UpsArgs& a, float* s_in, int bx, int by, int bz) {
  ups_fwd8_tile_at(a, s_in, by * FT, bx * FT, bz % (a.c_in + 1), bz / (a.c_in + 1));
}

__global__ void __launch_bounds__(256) k_ups_fwd8(UpsArgs a) {
  __shared__ __align__(16) float s_in[FRN * FRS];
  ups_fwd8_tile(a, s_in, blockIdx.x, blockIdx.y, blockIdx.z);
}

// Forward of the leading pyramid stages WITHOUT any grid-wide synchronisation.  Every channel of the pyramid is
// its own chain (refine once, then x2 per stage; channels never mix), and a 64x64 tile of stage i+1 needs a 36x36
// window of the same channel of stage i -- which fits in ONE 64x64 tile of stage i placed at (origin/2 - 2).  So
// the CTA that owns a tile of the LAST fused stage recomputes the tiles it depends on, stage by stage, with only
// __syncthreads() in between (it reads back what it wrote itself; other CTAs write the same values to the same
// places, so the canonical stage buffers the backward needs end up complete).  The redundant work is a few tiles
// per CTA on stages that are a handful of tiles in total; what it removes is ~4 us of barrier + tail per stage.
constexpr int UPS_MAX_CHAIN = 5;
struct UpsChain {
  UpsArgs st[UPS_MAX_CHAIN];
  int n;
};

__global__ void __launch_bounds__(256) k_ups_fwd8_chain(UpsChain m) {
  __shared__ __align__(16) float s_in[FRN * FRS];
  const int n = m.n;
  const int nch = m.st[n - 1].c_in + 1;
  const int chf = blockIdx.z % nch, b = blockIdx.z / nch;
  int oy[UPS_MAX_CHAIN], ox[UPS_MAX_CHAIN];
#pragma unroll
  for (int i = UPS_MAX_CHAIN - 1; i >= 0; --i) {
    if (i > n - 1) continue;
    if (i == n - 1) {
      oy[i] = blockIdx.y * FT;
      ox[i] = blockIdx.x * FT;
    } else {
      oy[i] = max(0, oy[i + 1] / 2 - 2) & ~3;  // multiples of 4: the 4x4 register blocks store aligned float4
      ox[i] = max(0, ox[i + 1] / 2 - 2) & ~3;
    }
  }
#pragma unroll
  for (int i = 0; i < UPS_MAX_CHAIN; ++i) {
    if (i >= n) break;
    const int ch = chf - (n - 1 - i);  // channel of this chain at stage i (negative: the chain starts later)
    if (ch < 0) continue;
    ups_fwd8_tile_at(m.st[i], s_in, oy[i], ox[i], ch, b);
    __syncthreads();  // orders this CTA's global writes before its reads of the next stage; frees the tile buffer
  }
}

// Grid-wide barrier for kernels whose whole grid is co-resident (a few dozen CTAs).  `ctr` is zeroed earlier in the
// same step (k_quant<false> / k_loss_partial); k counts the barriers of this launch from 1.
// It is split: a CTA ARRIVES as soon as what the next stage reads is stored (the input gradient of its last tile) and
// WAITS only when it is about to start that stage; its filter-gradient sums and block reductions overlap the other
// CTAs' arrival and the two L2 round trips of the barrier.
__device__ __forceinline__ void grid_arrive(unsigned* ctr) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
  }
}
__device__ __forceinline__ void grid_wait(unsigned* ctr, unsigned k) {
  if (threadIdx.x == 0) {
    const unsigned target = k * gridDim.x;
    unsigned v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
    } while (v < target);
    __threadfence();
  }
  __syncthreads();
}

// The first stages of the pyramid are a handful of tiles each and purely latency-bound as separate launches
// (~6 us per dependent launch for < 0.1 M instructions): one kernel walks them with a grid barrier in between.
constexpr int UPS_MAX_FUSED = 5;
struct UpsMulti {
  UpsArgs st[UPS_MAX_FUSED];
  int gx[UPS_MAX_FUSED], gy[UPS_MAX_FUSED], gz[UPS_MAX_FUSED];
  int n;
  unsigned* barrier;
};

// reduce-scatter of 8 per-lane values over a warp in 9 shuffles; returns the total of value
// index ((lane >> 2) & 7) ... see use below: after the call lanes {0,4,8,...,28} hold totals 0..7.
__device__ __forceinline__ float warp_reduce8(float* g) {
  const int lane = threadIdx.x & 31;
  // 8 -> 4 (xor 16), 4 -> 2 (xor 8), 2 -> 1 (xor 4), then plain sums over xor 2, 1
  float v4[4], v2[2], v1;
  {
    const bool hi = lane & 16;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float send = hi ? g[k] : g[4 + k];
      const float keep = hi ? g[4 + k] : g[k];
      v4[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
  }
  {
    const bool hi = lane & 8;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const float send = hi ? v4[k] : v4[2 + k];
      const float keep = hi ? v4[2 + k] : v4[k];
      v2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
  }
  {
    const bool hi = lane & 4;
    const float send = hi ? v2[0] : v2[1];
    const float keep = hi ? v2[1] : v2[0];
    v1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  v1 += __shfl_xor_sync(0xffffffffu, v1, 2);
  v1 += __shfl_xor_sync(0xffffffffu, v1, 1);
  return v1;  // lane holds the total of index ((lane>>4)&1)*4 + ((lane>>3)&1)*2 + ((lane>>2)&1)
}

// block total of g[0..7], symmetric halves folded (t and last - t; last = 7: 8-tap filter, last = 6: 7-tap filter
// whose middle tap stands alone) -> dst[0..3]; `red` holds 8 warps x 8 floats + 8
__device__ __forceinline__ void block_reduce8_store(float* g, float* red, float* dst, int last) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const float v = warp_reduce8(g);
  if ((lane & 3) == 0) {
    const int idx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    red[wid * 8 + idx] = v;
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k * 8 + threadIdx.x];
    red[64 + threadIdx.x] = t;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    const int m = last - threadIdx.x;
    dst[threadIdx.x] = (threadIdx.x == m) ? red[64 + threadIdx.x] : red[64 + threadIdx.x] + red[64 + m];
  }
}

struct UpsBwdSmem {
  float s_dy[FRN * FRS];
  float s_x[FRN * FRS];
  float wvar[4][8];
  float red[72];
};

__device__ __forceinline__ void ups_bwd8_tile(const UpsArgs& a, UpsBwdSmem& sm, int bx, int by, int bz, int gdx,
                                              int gdy, unsigned* arrive = nullptr) {
  float* s_dy = sm.s_dy;
  float* s_x = sm.s_x;
  float(*wvar)[8] = sm.wvar;
  float* red = sm.red;
  const int ch = bz % (a.c_in + 1);
  const int b = bz / (a.c_in + 1);
  const int bid = (bz * gdy + by) * gdx + bx;
  float* prow = a.part + (long long)bid * UPS_PART;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const float* dY = a.d_dst + ((long long)(b * (a.c_in + 1) + ch) * a.ht) * a.wt;
  float g[8];
#pragma unroll
  for (int t = 0; t < 8; ++t) g[t] = 0.f;

  if (ch == 0) {
    const int oy0 = by * FT, ox0 = bx * FT;
    if (oy0 >= a.ht || ox0 >= a.wt) {
      if (threadIdx.x < UPS_PART) prow[threadIdx.x] = 0.f;
      if (arrive) grid_arrive(arrive);
      return;
    }
    const float* wpre = a.w_pre + (long long)b * a.wbs;
    const float p0 = wpre[0], p1 = wpre[1], p2 = wpre[2], p3 = wpre[3];
    const float pt[7] = {p0, p1, p2, p3, p2, p1, p0};
    load_tile_zero(s_dy, dY, a.ht, a.wt, oy0 - 3, ox0 - 3);
    load_tile_zero(s_x, a.tgt + (long long)b * a.tgt_bs, a.ht, a.wt, oy0 - 3, ox0 - 3);
    cp_async_wait_all();
    __syncthreads();
    float vdy[4][10], hdy[4][4], own[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 10; ++j) vdy[i][j] = 0.f;
#pragma unroll
    for (int rr = 0; rr < 10; ++rr) {
      float d[10];
      load_row10(&s_dy[(4 * ty + rr) * FRS + 4 * tx], d);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int t = rr - i;
        if (t >= 0 && t < 7) {
#pragma unroll
          for (int j = 0; j < 10; ++j) vdy[i][j] = fmaf(pt[t], d[j], vdy[i][j]);
        }
        if (rr == i + 3) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            own[i][j] = d[j + 3];
            float acc = 0.f;
#pragma unroll
            for (int t2 = 0; t2 < 7; ++t2) acc = fmaf(pt[t2], d[j + t2], acc);
            hdy[i][j] = acc;
          }
        }
      }
    }
    // d x = dY + H(V(dY))   (symmetric filter, zero padding => self-adjoint)
    float res[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float acc = own[i][j];
#pragma unroll
        for (int t = 0; t < 7; ++t) acc = fmaf(pt[t], vdy[i][j + t], acc);
        res[i][j] = acc;
      }
    store_block4x4(a.d_tgt + (long long)b * a.tgt_bs, a.ht, a.wt, oy0 + 4 * ty, ox0 + 4 * tx, res);
    if (arrive) grid_arrive(arrive);  // the next stage only needs the input gradient
    // filter gradient: g[t] = sum H(dY)[i][j] x[i+t-3][j] + sum V(dY)[i][j] x[i][j+t-3]
#pragma unroll
    for (int rr = 0; rr < 10; ++rr) {
      float d[10];
      load_row10(&s_x[(4 * ty + rr) * FRS + 4 * tx], d);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int t = rr - i;
        if (t >= 0 && t < 7) {
#pragma unroll
          for (int j = 0; j < 4; ++j) g[t] = fmaf(hdy[i][j], d[j + 3], g[t]);
        }
        if (rr == i + 3) {
#pragma unroll
          for (int t2 = 0; t2 < 7; ++t2)
#pragma unroll
            for (int j = 0; j < 4; ++j) g[t2] = fmaf(vdy[i][j + 3], d[j + t2], g[t2]);
        }
      }
    }
    block_reduce8_store(g, red, prow + 4, 6);
    if (threadIdx.x < 4) prow[threadIdx.x] = 0.f;
    return;
  }

  // ---- upsample branch: CTA = 32x32 source tile <-> 64x64 output tile
  const int r0t = by * 32, c0t = bx * 32;
  if (r0t >= a.hs || c0t >= a.ws) {
    if (threadIdx.x < UPS_PART) prow[threadIdx.x] = 0.f;
    if (arrive) grid_arrive(arrive);
    return;
  }
  const float* wup = a.w_up + (long long)b * a.wbs;
  Taps8 w = {wup[0], wup[1], wup[2], wup[3]};
  if (threadIdx.x < 32) {
    // transposed taps per source row/column class: interior, first (folds the two virtual positions
    // before it: replicate padding), last, both
    const int k = threadIdx.x & 7, v = threadIdx.x >> 3;
    auto wk = [&](int i) -> float {
      if (i < 0 || i > 7) return 0.f;
      const int m = i < 4 ? i : 7 - i;
      return m == 0 ? w.w0 : (m == 1 ? w.w1 : (m == 2 ? w.w2 : w.w3));
    };
    float t = wk(k);
    if (v & 1) t += wk(k + 2) + wk(k + 4);
    if (v & 2) t += wk(k - 2) + wk(k - 4);
    wvar[v][k] = t;
  }
  const float* src = a.src + (long long)b * a.src_bs + (long long)(ch - 1) * a.src_cs;
  load_tile_zero(s_dy, dY, a.ht, a.wt, 2 * r0t - 3, 2 * c0t - 3);
  load_tile_clamp(s_x, src, a.hs, a.ws, r0t - 2, c0t - 2);
  cp_async_wait_all();
  __syncthreads();
  const int r0 = r0t + 2 * ty, c0 = c0t + 2 * tx;
  float own[4][4];
  {
    // d in = (fold o H^T o V^T)(dY)
    float wA[8], wB[8];
    {
      const int vA = (r0 == 0 ? 1 : 0) | (r0 == a.hs - 1 ? 2 : 0);
      const int vB = (r0 + 1 == a.hs - 1 ? 2 : 0);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        wA[k] = wvar[vA][k];
        wB[k] = wvar[vB][k];
      }
    }
    float vt[2][10];
#pragma unroll
    for (int j = 0; j < 10; ++j) vt[0][j] = vt[1][j] = 0.f;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      float d[10];
      load_row10(&s_dy[(4 * ty + i) * FRS + 4 * tx], d);
      if (i < 8) {
#pragma unroll
        for (int j = 0; j < 10; ++j) vt[0][j] = fmaf(wA[i], d[j], vt[0][j]);
      }
      if (i >= 2) {
#pragma unroll
        for (int j = 0; j < 10; ++j) vt[1][j] = fmaf(wB[i - 2], d[j], vt[1][j]);
      }
      if (i >= 3 && i <= 6) {
#pragma unroll
        for (int j = 0; j < 4; ++j) own[i - 3][j] = d[j + 3];
      }
    }
    {
      const int vC = (c0 == 0 ? 1 : 0) | (c0 == a.ws - 1 ? 2 : 0);
      const int vD = (c0 + 1 == a.ws - 1 ? 2 : 0);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        wA[k] = wvar[vC][k];
        wB[k] = wvar[vD][k];
      }
    }
    float* dS = a.d_src + (long long)b * a.d_src_bs + (long long)(ch - 1) * a.d_src_cs;
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      float x0 = 0.f, x1 = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        x0 = fmaf(wA[k], vt[rr][k], x0);
        x1 = fmaf(wB[k], vt[rr][k + 2], x1);
      }
      if (r0 + rr < a.hs) {
        if (c0 < a.ws) dS[(long long)(r0 + rr) * a.ws + c0] = x0;
        if (c0 + 1 < a.ws) dS[(long long)(r0 + rr) * a.ws + c0 + 1] = x1;
      }
    }
  }
  if (arrive) grid_arrive(arrive);  // the next stage only needs the input gradient
  // ---- filter gradient (8 taps, folded onto 4 by k_ups_reduce)
  //   vertical tap t:   sum dY[n][m] * H(x)[a(n,t)][m]      horizontal tap u: sum dY[n][m] * V(x)[n][a(m,u)]
  {
    float h[6][4], vx[4][6];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 6; ++c) vx[i][c] = 0.f;
    // polyphase coefficient of source row rr (0..5) for output row i (0..3), as in poly4
    const float cf[4][6] = {{w.w0, w.w2, w.w3, w.w1, 0.f, 0.f},
                            {0.f, w.w1, w.w3, w.w2, w.w0, 0.f},
                            {0.f, w.w0, w.w2, w.w3, w.w1, 0.f},
                            {0.f, 0.f, w.w1, w.w3, w.w2, w.w0}};
#pragma unroll
    for (int rr = 0; rr < 6; ++rr) {
      float d[6];
      load_row6(&s_x[(2 * ty + rr) * FXS + 2 * tx], d);
      poly4(w, d[0], d[1], d[2], d[3], d[4], d[5], h[rr]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if ((i == 0 && rr <= 3) || (i == 1 && rr >= 1 && rr <= 4) || (i == 2 && rr >= 1 && rr <= 4) ||
            (i == 3 && rr >= 2)) {
#pragma unroll
          for (int c = 0; c < 6; ++c) vx[i][c] = fmaf(cf[i][rr], d[c], vx[i][c]);
        }
      }
    }
    // output row i: phase 0 rows (i = 0, 2) use taps 1,3,5,7 on source rows (hi, hi-1, hi-2, hi-3);
    // phase 1 rows (i = 1, 3) use taps 0,2,4,6.  hi = 3,4,4,5 for i = 0..3.
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int hi = (i == 0) ? 3 : (i == 3 ? 5 : 4);
      const int tb = (i & 1) ? 0 : 1;
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < 4; ++j) g[tb + 2 * q] = fmaf(own[i][j], h[hi - q][j], g[tb + 2 * q]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int hj = (j == 0) ? 3 : (j == 3 ? 5 : 4);
      const int tb = (j & 1) ? 0 : 1;
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int i = 0; i < 4; ++i) g[tb + 2 * q] = fmaf(own[i][j], vx[i][hj - q], g[tb + 2 * q]);
    }
  }
  block_reduce8_store(g, red, prow, 7);
  if (threadIdx.x < 4) prow[4 + threadIdx.x] = 0.f;
}

#ifndef CCB_UPSB_LB
#define CCB_UPSB_LB 3  // 80 registers without spills (the unconstrained build takes 120): three CTAs per SM
#endif
__global__ void __launch_bounds__(256, CCB_UPSB_LB) k_ups_bwd8(UpsArgs a) {
  __shared__ __align__(16) UpsBwdSmem sm;
  ups_bwd8_tile(a, sm, blockIdx.x, blockIdx.y, blockIdx.z, gridDim.x, gridDim.y);
}

__global__ void __launch_bounds__(256) k_ups_bwd8_multi(UpsMulti m) {
  __shared__ __align__(16) UpsBwdSmem sm;
  for (int i = 0; i < m.n; ++i) {
    const int n_tiles = m.gx[i] * m.gy[i] * m.gz[i];
    const bool more = i + 1 < m.n;
    bool arrived = false;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      const int bz = t / (m.gx[i] * m.gy[i]), r = t - bz * m.gx[i] * m.gy[i];
      const bool last = t + (int)gridDim.x >= n_tiles;  // this CTA's last tile of the stage: arrive from inside it
      ups_bwd8_tile(m.st[i], sm, r % m.gx[i], r / m.gx[i], bz, m.gx[i], m.gy[i], (more && last) ? m.barrier : nullptr);
      arrived = last;
      __syncthreads();
    }
    if (more) {
      if (!arrived) grid_arrive(m.barrier);
      grid_wait(m.barrier, (unsigned)(i + 1));
    }
  }
}

UpsArgs stage_args(const ccb_plan* p, const ccb_tensors* prm, char* ws, int idx) {
  const int L = p->g.L;
  const int lvl_t = L - 2 - idx;  // target latent level
  const int lvl_s = lvl_t + 1;
  UpsArgs a = {};
  const float* yhat = reinterpret_cast<const float*>(ws + p->off_yhat);
  a.c_in = idx + 1;
  a.hs = p->g.h[lvl_s];
  a.ws = p->g.w[lvl_s];
  a.ht = p->g.h[lvl_t];
  a.wt = p->g.w[lvl_t];
  if (idx == 0) {
    a.src = yhat + p->g.off[lvl_s];
    a.src_bs = p->g.M;
    a.src_cs = 0;
  } else {
    a.src = reinterpret_cast<const float*>(ws + p->off_up[idx]);
    a.src_cs = (long long)a.hs * a.ws;
    a.src_bs = a.src_cs * a.c_in;
  }
  a.tgt = yhat + p->g.off[lvl_t];
  a.tgt_bs = p->g.M;
  a.dst = reinterpret_cast<float*>(ws + p->off_up[idx + 1]);
  a.B = p->g.B;
  a.K = p->arch.ups_k;
  a.KP = p->arch.ups_pre_k;
  a.w_up = prm->ups_w[idx];
  a.w_pre = prm->pre_w[idx];
  a.wbs = prm->weight_batch_stride;
  return a;
}

}  // namespace

// Number of leading stages (idx 0..n-1) that run inside ONE multi-stage launch: register-tiled kernels only,
// at most UPS_FUSE_TILES tiles per stage (the whole grid is co-resident).  0 or 1 => no fusion.
constexpr int UPS_FUSE_TILES = 64;
static int n_fused_stages(const ccb_plan* p, bool backward) {
  if (!(p->arch.ups_k == 8 && p->arch.ups_pre_k == 7)) return 0;
  const int L = p->g.L;
  int n = 0;
  for (int idx = 0; idx < L - 1 && n < UPS_MAX_FUSED; ++idx) {
    const int lvl_t = L - 2 - idx, lvl_s = lvl_t + 1;
    const int tiles = backward ? ceil_div(p->g.w[lvl_s], 32) * ceil_div(p->g.h[lvl_s], 32) * p->g.B * (idx + 2)
                               : ceil_div(p->g.w[lvl_t], FT) * ceil_div(p->g.h[lvl_t], FT) * p->g.B * (idx + 2);
    if (tiles > UPS_FUSE_TILES) break;
    ++n;
  }
  return n >= 2 ? n : 0;
}

// Leading stages of the FORWARD that run in the chain kernel: the last of them has at most UPS_CHAIN_TILES tiles
// (about one wave of CTAs, each recomputing one tile per earlier stage).
constexpr int UPS_CHAIN_TILES = 160;
static int n_chain_stages(const ccb_plan* p) {
  if (!(p->arch.ups_k == 8 && p->arch.ups_pre_k == 7)) return 0;
  const int L = p->g.L;
  int n = 0;
  for (int idx = 0; idx < L - 1 && n < UPS_MAX_CHAIN; ++idx) {
    const int lvl_t = L - 2 - idx;
    const int tiles = ceil_div(p->g.w[lvl_t], FT) * ceil_div(p->g.h[lvl_t], FT) * p->g.B * (idx + 2);
    if (tiles > UPS_CHAIN_TILES) break;
    ++n;
  }
  return n >= 2 ? n : 0;
}

static unsigned* ups_barrier(const ccb_plan* p, char* ws, bool backward) {
  return reinterpret_cast<unsigned*>(ws + p->off_scalars) + (backward ? 10 : 9);
}

// Finest latent level the forward's chain kernel reads (levels >= it must be quantised before it starts), or 0 when no
// stage runs outside it; finest level whose d yhat the backward's fused launch still writes (levels below it are final
// once the separately launched stages are done), or 0.
int ccb_ups_chain_min_level(const ccb_plan* p) {
  const int nc = n_chain_stages(p);
  return (nc > 0 && nc < p->g.L - 1) ? p->g.L - 1 - nc : 0;
}
int ccb_ups_fused_min_level(const ccb_plan* p) {
  const int nf = n_fused_stages(p, true);
  return (nf > 0 && nf < p->g.L - 1) ? p->g.L - 1 - nf : 0;
}

int ccb_ups_n_launches(const ccb_plan* p, bool backward) {
  const int nf = backward ? n_fused_stages(p, true) : n_chain_stages(p);
  return (p->g.L - 1) - (nf ? nf - 1 : 0);
}

int ccb_ups_forward(const ccb_plan* p, const ccb_tensors* prm, char* ws, cudaStream_t s, cudaEvent_t wait_fine) {
  const int L = p->g.L;
  const int nf = n_chain_stages(p);
  if (nf) {
    UpsChain m = {};
    for (int idx = 0; idx < nf; ++idx) m.st[idx] = stage_args(p, prm, ws, idx);
    m.n = nf;
    const UpsArgs& last = m.st[nf - 1];
    dim3 grid(ceil_div(last.wt, FT), ceil_div(last.ht, FT), p->g.B * (last.c_in + 1));
    k_ups_fwd8_chain<<<grid, 256, 0, s>>>(m);
    CCB_LAUNCH_CHECK();
  }
  // the stages outside the chain kernel read the finer latent levels: quantised by another stream in the split schedule
  if (wait_fine) CCB_CHECK(cudaStreamWaitEvent(s, wait_fine, 0));
  for (int idx = nf; idx < L - 1; ++idx) {
    UpsArgs a = stage_args(p, prm, ws, idx);
    if (a.K == 8 && a.KP == 7) {
      dim3 grid(ceil_div(a.wt, FT), ceil_div(a.ht, FT), p->g.B * (a.c_in + 1));
      k_ups_fwd8<<<grid, 256, 0, s>>>(a);
    } else {
      dim3 grid(ceil_div(a.ws, UI), ceil_div(a.hs, UI), p->g.B * (a.c_in + 1));
      // the refine branch tiles the target (<= 2x source) with 32x32 tiles: covered by the same grid
      k_ups_fwd<<<grid, UPS_THREADS, 0, s>>>(a);
    }
    CCB_LAUNCH_CHECK();
  }
  return CCB_OK;
}

static UpsArgs bwd_stage_args(const ccb_plan* p, const ccb_tensors* prm, char* ws, int idx, float* part) {
  const int L = p->g.L;
  float* dyhat = reinterpret_cast<float*>(ws + p->off_dyhat);
  UpsArgs a = stage_args(p, prm, ws, idx);
  const int lvl_t = L - 2 - idx, lvl_s = lvl_t + 1;
  a.d_dst = reinterpret_cast<const float*>(ws + p->off_dup[idx + 1]);
  if (idx == 0) {
    a.d_src = dyhat + p->g.off[lvl_s];
    a.d_src_bs = p->g.M;
    a.d_src_cs = 0;
  } else {
    a.d_src = reinterpret_cast<float*>(ws + p->off_dup[idx]);
    a.d_src_cs = (long long)a.hs * a.ws;
    a.d_src_bs = a.d_src_cs * a.c_in;
  }
  a.d_tgt = dyhat + p->g.off[lvl_t];
  a.part = part;
  return a;
}

int ccb_ups_backward(const ccb_plan* p, const ccb_tensors* prm, char* ws, const ccb_tensors* grads,
                     cudaStream_t s, UpsPartInfo* info, cudaEvent_t after_first, cudaEvent_t after_big) {
  (void)grads;
  if (!info) return CCB_ERR_INVALID;
  const int L = p->g.L;
  float* part = reinterpret_cast<float*>(ws + p->off_ups_part);
  info->n = L - 1;
  const int nf = n_fused_stages(p, true);
  info->n_fused = nf;
  long long row0 = 0;
  for (int idx = L - 2; idx >= nf; --idx) {
    UpsArgs a = bwd_stage_args(p, prm, ws, idx, part + row0 * UPS_PART);
    const bool fast = (a.K == 8 && a.KP == 7);
    dim3 grid(ceil_div(a.ws, fast ? 32 : UI), ceil_div(a.hs, fast ? 32 : UI), p->g.B * (a.c_in + 1));
    if (fast && ccb_debug_skip(32)) {
    } else if (fast)
      k_ups_bwd8<<<grid, 256, 0, s>>>(a);
    else
      k_ups_bwd<<<grid, UPS_THREADS, 0, s>>>(a);
    CCB_LAUNCH_CHECK();
    if (after_first && idx == L - 2) CCB_CHECK(cudaEventRecord(after_first, s));
    const int n_rows = grid.x * grid.y * grid.z;
    info->part[idx] = a.part;
    info->n_rows[idx] = n_rows;
    row0 += n_rows;
  }
  if (after_big) CCB_CHECK(cudaEventRecord(after_big, s));
  if (nf) {
    UpsMulti m = {};
    int grid = 1;
    for (int k = 0; k < nf; ++k) {
      const int idx = nf - 1 - k;  // descending: the gradient flows from the finer stage to the coarser one
      m.st[k] = bwd_stage_args(p, prm, ws, idx, part + row0 * UPS_PART);
      m.gx[k] = ceil_div(m.st[k].ws, 32);
      m.gy[k] = ceil_div(m.st[k].hs, 32);
      m.gz[k] = p->g.B * (m.st[k].c_in + 1);
      const int n_rows = m.gx[k] * m.gy[k] * m.gz[k];
      grid = std::max(grid, n_rows);
      info->part[idx] = m.st[k].part;
      info->n_rows[idx] = n_rows;
      row0 += n_rows;
    }
    m.n = nf;
    m.barrier = ups_barrier(p, ws, true);
    if (!ccb_debug_skip(2)) k_ups_bwd8_multi<<<grid, 256, 0, s>>>(m);
    CCB_LAUNCH_CHECK();
  }
  if (row0 > p->n_ups_part_rows) return CCB_ERR_INVALID;
  return CCB_OK;
}
