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
  // backward
  const float* d_dst;  // [B, c_in+1, ht, wt]
  float* d_src;        // [B, c_in, hs, ws] with strides d_src_bs / d_src_cs
  long long d_src_bs, d_src_cs;
  float* d_tgt;        // [B, ht, wt] with batch stride tgt_bs (inside dyhat)
  float* part;         // [n_blocks][UPS_PART]
};

__device__ __forceinline__ void load_filters(const UpsArgs& a, float* wu, float* wp) {
  if (threadIdx.x < 8) {
    int t = threadIdx.x;
    wu[t] = (t < a.K) ? a.w_up[t < a.K / 2 ? t : a.K - 1 - t] : 0.f;
    int m = a.KP - 1 - t;
    wp[t] = (t < a.KP) ? a.w_pre[t < m ? t : m] : 0.f;
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
  load_filters(a, wu, wp);
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
  load_filters(a, wu, wp);
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
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      float v = block_sum(gsum[t], red);
      if (threadIdx.x == 0) {
        prow[8 + t] = v;
        prow[t] = 0.f;
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
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    float v = block_sum(gsum[t], red);
    if (threadIdx.x == 0) {
      prow[t] = v;
      prow[8 + t] = 0.f;
    }
  }
}

// Deterministic column sums of the per-CTA partial rows of ALL stages in one launch (CTA == stage),
// folded onto the symmetric halves.
struct UpsReduceArgs {
  const float* part[CCB_MAX_LATENTS];
  int n_rows[CCB_MAX_LATENTS];
  float* g_up[CCB_MAX_LATENTS];
  float* g_pre[CCB_MAX_LATENTS];
  int K, KP;
};

__global__ void __launch_bounds__(256) k_ups_reduce(UpsReduceArgs a) {
  __shared__ float sm[16][17];
  const int st = blockIdx.x;
  const float* part = a.part[st];
  const int n_rows = a.n_rows[st];
  const int col = threadIdx.x & 15, rg = threadIdx.x >> 4;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  int r = rg;
  for (; r + 48 < n_rows; r += 64) {
    s0 += part[(long long)r * UPS_PART + col];
    s1 += part[(long long)(r + 16) * UPS_PART + col];
    s2 += part[(long long)(r + 32) * UPS_PART + col];
    s3 += part[(long long)(r + 48) * UPS_PART + col];
  }
  for (; r < n_rows; r += 16) s0 += part[(long long)r * UPS_PART + col];
  sm[rg][col] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (threadIdx.x < 16) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) v += sm[k][threadIdx.x];
    sm[0][threadIdx.x] = v;
  }
  __syncthreads();
  const float* tot = sm[0];
  const int K = a.K, KP = a.KP;
  if (threadIdx.x < K / 2) a.g_up[st][threadIdx.x] = tot[threadIdx.x] + tot[K - 1 - threadIdx.x];
  if (threadIdx.x < (KP + 1) / 2) {
    int t = threadIdx.x, m = KP - 1 - t;
    a.g_pre[st][t] = (t == m) ? tot[8 + t] : tot[8 + t] + tot[8 + m];
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
  return a;
}

}  // namespace

int ccb_ups_forward(const ccb_plan* p, const ccb_tensors* prm, char* ws, cudaStream_t s) {
  const int L = p->g.L;
  for (int idx = 0; idx < L - 1; ++idx) {
    UpsArgs a = stage_args(p, prm, ws, idx);
    dim3 grid(ceil_div(a.ws, UI), ceil_div(a.hs, UI), p->g.B * (a.c_in + 1));
    // the refine branch tiles the target (<= 2x source) with 32x32 tiles: covered by the same grid
    k_ups_fwd<<<grid, UPS_THREADS, 0, s>>>(a);
    CCB_LAUNCH_CHECK();
  }
  return CCB_OK;
}

int ccb_ups_backward(const ccb_plan* p, const ccb_tensors* prm, char* ws, const ccb_tensors* grads,
                     cudaStream_t s) {
  const int L = p->g.L;
  float* dyhat = reinterpret_cast<float*>(ws + p->off_dyhat);
  float* part = reinterpret_cast<float*>(ws + p->off_ups_part);
  UpsReduceArgs ra = {};
  ra.K = p->arch.ups_k;
  ra.KP = p->arch.ups_pre_k;
  long long row0 = 0;
  for (int idx = L - 2; idx >= 0; --idx) {
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
    a.part = part + row0 * UPS_PART;
    dim3 grid(ceil_div(a.ws, UI), ceil_div(a.hs, UI), p->g.B * (a.c_in + 1));
    k_ups_bwd<<<grid, UPS_THREADS, 0, s>>>(a);
    CCB_LAUNCH_CHECK();
    const int n_rows = grid.x * grid.y * grid.z;
    ra.part[idx] = a.part;
    ra.n_rows[idx] = n_rows;
    ra.g_up[idx] = grads->ups_w[idx];
    ra.g_pre[idx] = grads->pre_w[idx];
    row0 += n_rows;
  }
  if (row0 > p->n_ups_part_rows) return CCB_ERR_INVALID;
  k_ups_reduce<<<L - 1, 256, 0, s>>>(ra);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
