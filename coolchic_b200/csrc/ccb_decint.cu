// Fixed-point upsampling + synthesis of the DECODER, on the encoder's GPU: the exact int32 image the
// reference's C++ decoder will produce, without running it.  Restates
//   cpp/ups_refine_cpu.hpp:11-79    separable refine filter, zero padding, residual
//   cpp/ups_upsample_cpu.hpp:11-90  x2 polyphase upsampling, replicate padding
//   cpp/cc-frame-decoder.cpp:572-680 (run_ups: which filter, which plane) and :767-1042 (run_syn_branch)
//   cpp/synfused_cpu.hpp:16-108     fused 1x1 pair;  cpp/syn_cpu.hpp:21-112 / synlb_cpu.hpp:20-132  1x1 / 3x3 layers
//   cpp/ccdecapi.cpp:59-128         ppm_out rounding to 8 bits
// bit for bit (tests/test_gpu_dec.py against vectors produced by the reference's own C++ on shipped bitstreams).
//
// Integer, HBM/L2-bound work: every kernel is a coalesced map over output pixels, stencils staged in shared memory.
// The pyramid is walked by STAGE, not by latent (the reference walks latent by latent): all planes that reach level t
// use the same filter, so ONE launch per level refines its latent and upsamples every plane below it: 6 launches for 7
// levels, then the fused 1x1 pair and the 3x3 layers (the last one also writes the 8-bit pixels).
#include "ccb_internal.cuh"

namespace {

constexpr int ARM_PREC = 8, UPS_PREC = 12, SYN_PREC = 12;
constexpr int DEC_MAX_K = 8;      // filter taps (ups_k_size <= 8, ups_preconcat_k_size <= 7)
constexpr int DT = 32;            // tile edge (source pixels for the upsampling, output pixels for the refine)

// `sum < 0 ? -(-sum >> n) : sum >> n`: the decoder's rounding toward zero (ups_refine_cpu.hpp:50-53)
__device__ __forceinline__ int trunc_shift(int v, int n) { return v < 0 ? -((-v) >> n) : (v >> n); }

struct DecFilt {
  int k[DEC_MAX_K];
  int n;
};

struct ccb_dec_geom {
  int n, n_sm;
  int h[CCB_MAX_LATENTS], w[CCB_MAX_LATENTS];
  long long off[CCB_MAX_LATENTS];      // latent offsets in the flat finest-first vector
  long long pyr_off[CCB_MAX_LATENTS];  // workspace offsets (ints) of the per-level plane stacks, levels 1 .. n-1
  long long pyr_total;
};

int dec_geom(int H, int W, int n, ccb_dec_geom* g) {
  if (n < 2 || n > CCB_MAX_LATENTS || H < 1 || W < 1) return CCB_ERR_INVALID;
  g->n = n;
  long long off = 0, po = 0;
  for (int t = 0, h = H, w = W; t < n; ++t, h = (h + 1) / 2, w = (w + 1) / 2) {
    g->h[t] = h;
    g->w[t] = w;
    g->off[t] = off;
    off += (long long)h * w;
    g->pyr_off[t] = po;
    if (t >= 1 && t < n - 1) po += (long long)(n - t) * h * w;  // level 0 lives in ups_out, the coarsest is the latent
  }
  g->pyr_total = po;
  int dev = 0;
  g->n_sm = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&g->n_sm, cudaDevAttrMultiProcessorCount, dev);
  return CCB_OK;
}

int dec_max_planes(int n, int n_syn, const ccb_dec_layer* L) {
  int m = n;
  const bool fused = n_syn >= 2 && L[0].ks == 1 && L[1].ks == 1;
  for (int i = fused ? 1 : 0; i < n_syn; ++i) m = L[i].n_out > m ? L[i].n_out : m;
  return m;
}

// ---- refine: dst = trunc((V(trunc(H(src) >> src_prec)) + (src << (12 - src_prec) << 12)) >> 12), zero padding.
// src holds PLAIN integers when src_shift > 0 (the latent, raised by ARM_PRECISION on the fly).
__device__ __forceinline__ void dec_refine_tile(const int* __restrict__ src, int h, int w, const DecFilt& f, int src_shift,
                                                int src_prec, int* __restrict__ dst, int bx, int by) {
  __shared__ int s_in[(DT + 2 * 3) * (DT + 2 * 3 + 1)];
  __shared__ int s_h[(DT + 2 * 3) * (DT + 1)];
  const int pad = f.n / 2;
  const int y0 = by * DT, x0 = bx * DT;
  const int IW = DT + 2 * pad, IS = DT + 2 * 3 + 1;
  for (int i = threadIdx.x; i < IW * IW; i += blockDim.x) {
    const int r = i / IW, c = i - r * IW;
    const int y = y0 - pad + r, x = x0 - pad + c;
    s_in[r * IS + c] = (y >= 0 && y < h && x >= 0 && x < w) ? (src[(long long)y * w + x] << src_shift) : 0;
  }
  __syncthreads();
  // horizontal pass on DT + 2 pad rows; rows outside the image are the vertical ZERO padding of the temporary
  for (int i = threadIdx.x; i < IW * DT; i += blockDim.x) {
    const int r = i / DT, c = i - r * DT;
    const int y = y0 - pad + r;
    int sum = 0;
    if (y >= 0 && y < h) {
      for (int t = 0; t < f.n; ++t) sum += s_in[r * IS + c + t] * f.k[t];
      sum = trunc_shift(sum, src_prec);
    }
    s_h[r * (DT + 1) + c] = sum;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < DT * DT; i += blockDim.x) {
    const int r = i / DT, c = i - r * DT;
    const int y = y0 + r, x = x0 + c;
    if (y >= h || x >= w) continue;
    int sum = 0;
    for (int t = 0; t < f.n; ++t) sum += s_h[(r + t) * (DT + 1) + c] * f.k[t];
    sum += (s_in[(r + pad) * IS + c + pad] << (UPS_PREC - src_prec)) << UPS_PREC;
    dst[(long long)y * w + x] = trunc_shift(sum, UPS_PREC);
  }
}

// ---- x2 upsampling of n_planes planes [h][w] -> [oh][ow] (oh in {2h-1, 2h}), replicate padding.
// Even outputs use the even taps on src[x-p .. x-p+ks-1], odd outputs the odd taps one sample further
// (ups_upsample_cpu.hpp:45-55, 70-81); horizontal results are truncated by the source precision, vertical by 12.
__device__ __forceinline__ void dec_upsample_tile(const int* __restrict__ src, int h, int w, const DecFilt& f, int src_shift,
                                                  int src_prec, int* __restrict__ dst, int oh, int ow, int bx, int by,
                                                  int plane) {
  constexpr int SR = DT + 5;      // source rows / columns a 32 x 32 source tile needs (ks/2 = 4 taps: -2 .. +2)
  __shared__ int s_in[SR * (SR + 1)];
  __shared__ int s_h[SR * (2 * DT + 1)];
  const int ks = f.n / 2, pad = ks / 2;
  const int* sp = src + (long long)plane * h * w;
  int* dp = dst + (long long)plane * oh * ow;
  const int y0 = by * DT, x0 = bx * DT;
  const int NR = DT + ks + 1;  // rows y0-pad .. y0+DT+pad (inclusive of the extra one the odd phase reads)
  for (int i = threadIdx.x; i < NR * NR; i += blockDim.x) {
    const int r = i / NR, c = i - r * NR;
    const int y = min(max(y0 - pad + r, 0), h - 1), x = min(max(x0 - pad + c, 0), w - 1);
    s_in[r * (SR + 1) + c] = sp[(long long)y * w + x] << src_shift;
  }
  __syncthreads();
  // horizontal: NR rows x 2*DT columns
  for (int i = threadIdx.x; i < NR * 2 * DT; i += blockDim.x) {
    const int r = i / (2 * DT), c2 = i - r * 2 * DT;
    const int c = c2 >> 1, odd = c2 & 1;
    int sum = 0;
    for (int t = 0; t < ks; ++t) sum += s_in[r * (SR + 1) + c + t + odd] * f.k[2 * t + odd];
    s_h[r * (2 * DT + 1) + c2] = trunc_shift(sum, src_prec);
  }
  __syncthreads();
  // vertical: 2*DT rows x 2*DT columns
  for (int i = threadIdx.x; i < 4 * DT * DT; i += blockDim.x) {
    const int r2 = i / (2 * DT), c2 = i - r2 * 2 * DT;
    const int oy = 2 * y0 + r2, ox = 2 * x0 + c2;
    if (oy >= oh || ox >= ow) continue;
    const int r = r2 >> 1, odd = r2 & 1;
    int sum = 0;
    for (int t = 0; t < ks; ++t) sum += s_h[(r + t + odd) * (2 * DT + 1) + c2] * f.k[2 * t + odd];
    dp[(long long)oy * ow + ox] = trunc_shift(sum, UPS_PREC);
  }
}

// ---- one pyramid stage = one launch: the refine of this level's own latent (plane 0 of the level's stack) and the
// upsampling of every plane of the coarser stack (planes 1 ..) are independent, so their tiles share a 1-D grid.
struct DecStage {
  const int* lat;   // this level's latent (plain integers), h x w
  const int* src;   // coarser stack: n_up planes of hs x ws (or the coarsest latent itself)
  int* dst;         // this level's stack: 1 + n_up planes of h x w
  int h, w, hs, ws, n_up;
  int src_shift, src_prec;
  int ref_tx, ref_tiles, up_tx, up_tiles;  // tile grid of the refine (over h x w) / of the upsampling (over hs x ws)
  DecFilt fr, fu;
};
__global__ void __launch_bounds__(256) k_dec_stage(DecStage a) {
  int b = blockIdx.x;
  if (b < a.ref_tiles) {
    dec_refine_tile(a.lat, a.h, a.w, a.fr, ARM_PREC, ARM_PREC, a.dst, b % a.ref_tx, b / a.ref_tx);
    return;
  }
  b -= a.ref_tiles;
  const int plane = b / a.up_tiles, t = b - plane * a.up_tiles;
  dec_upsample_tile(a.src, a.hs, a.ws, a.fu, a.src_shift, a.src_prec, a.dst + (long long)a.h * a.w, a.h, a.w, t % a.up_tx,
                    t / a.up_tx, plane);
}

// ---- fused 1x1 pair (synfused_cpu.hpp:16-108): hidden = relu(b0 + W0 x) >> 12, out = trunc((b1 + W1 hidden) >> 12).
// Weights in shared memory as [hidden][n_in] | b0 | [n_out][hidden] | b1 (broadcast reads); thread = pixel.
constexpr int DEC_MAX_IN = 8, DEC_MAX_OUT = 8;
__global__ void __launch_bounds__(256) k_dec_syn_fused(const int* __restrict__ x, long long n_pix, int n_in, int n_hid,
                                                       int n_out, const int* __restrict__ w0, const int* __restrict__ b0,
                                                       const int* __restrict__ w1, const int* __restrict__ b1,
                                                       int* __restrict__ y) {
  extern __shared__ int sw[];
  int* s_w0 = sw;
  int* s_b0 = s_w0 + n_hid * n_in;
  int* s_w1 = s_b0 + n_hid;
  int* s_b1 = s_w1 + n_out * n_hid;
  for (int i = threadIdx.x; i < n_hid * n_in; i += blockDim.x) s_w0[i] = w0[i];
  for (int i = threadIdx.x; i < n_hid; i += blockDim.x) s_b0[i] = b0[i];
  for (int i = threadIdx.x; i < n_out * n_hid; i += blockDim.x) s_w1[i] = w1[i];
  for (int i = threadIdx.x; i < n_out; i += blockDim.x) s_b1[i] = b1[i];
  __syncthreads();
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < n_pix; p += (long long)gridDim.x * blockDim.x) {
    int in[DEC_MAX_IN], acc[DEC_MAX_OUT];
#pragma unroll
    for (int i = 0; i < DEC_MAX_IN; ++i) in[i] = i < n_in ? x[(long long)i * n_pix + p] : 0;
#pragma unroll
    for (int o = 0; o < DEC_MAX_OUT; ++o) acc[o] = o < n_out ? s_b1[o] : 0;
    for (int hdn = 0; hdn < n_hid; ++hdn) {
      int v = s_b0[hdn];
#pragma unroll
      for (int i = 0; i < DEC_MAX_IN; ++i)
        if (i < n_in) v += in[i] * s_w0[hdn * n_in + i];
      v = v < 0 ? 0 : (v >> SYN_PREC);
#pragma unroll
      for (int o = 0; o < DEC_MAX_OUT; ++o)
        if (o < n_out) acc[o] += v * s_w1[o * n_hid + hdn];
    }
#pragma unroll
    for (int o = 0; o < DEC_MAX_OUT; ++o)
      if (o < n_out) y[(long long)o * n_pix + p] = trunc_shift(acc[o], SYN_PREC);
  }
}

// ---- one synthesis layer, ks in {1, 3}, replicate padding (syn_cpu.hpp:21-112, synlb_cpu.hpp:20-132; padding:
// cc-frame-decoder.cpp:838-846).  w [n_out][n_in][ks][ks], b [n_out]; out of place.
__global__ void __launch_bounds__(256) k_dec_syn_layer(const int* __restrict__ x, int H, int W, int n_in, int n_out, int ks,
                                                       const int* __restrict__ w, const int* __restrict__ b, int residual,
                                                       int relu, int* __restrict__ y) {
  extern __shared__ int sw[];
  const int nw = n_out * n_in * ks * ks;
  for (int i = threadIdx.x; i < nw; i += blockDim.x) sw[i] = w[i];
  for (int i = threadIdx.x; i < n_out; i += blockDim.x) sw[nw + i] = b[i];
  __syncthreads();
  const long long HW = (long long)H * W;
  const int pad = ks / 2;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (long long)gridDim.x * blockDim.x) {
    const int py = (int)(p / W), px = (int)(p - (long long)py * W);
    int acc[DEC_MAX_OUT];
#pragma unroll
    for (int o = 0; o < DEC_MAX_OUT; ++o) acc[o] = o < n_out ? sw[nw + o] : 0;
    if (residual) {
#pragma unroll
      for (int o = 0; o < DEC_MAX_OUT; ++o)
        if (o < n_out) acc[o] += x[(long long)o * HW + p] << SYN_PREC;
    }
    for (int i = 0; i < n_in; ++i)
      for (int yy = 0; yy < ks; ++yy) {
        const int sy = min(max(py + yy - pad, 0), H - 1);
        for (int xx = 0; xx < ks; ++xx) {
          const int sx = min(max(px + xx - pad, 0), W - 1);
          const int v = x[(long long)i * HW + (long long)sy * W + sx];
#pragma unroll
          for (int o = 0; o < DEC_MAX_OUT; ++o)
            if (o < n_out) acc[o] += v * sw[((o * n_in + i) * ks + yy) * ks + xx];
        }
      }
#pragma unroll
    for (int o = 0; o < DEC_MAX_OUT; ++o)
      if (o < n_out) {
        const int s = acc[o];
        y[(long long)o * HW + p] = s < 0 ? (relu ? 0 : -((-s) >> SYN_PREC)) : (s >> SYN_PREC);
      }
  }
}

// ---- specialised forms of the two synthesis kernels for the shapes every shipped decoder configuration uses
// (n latents in -> S hidden -> 3; 3 -> 3 3x3 layers).  The generic kernels above issue ~3000 / ~1650 instructions per
// pixel on guarded 8-wide loops and per-tap clamps (ncu: profiles/r1_dec_launches.csv); these run the same integer
// arithmetic in ~700 / ~110.
//   fused pair: one shared-memory row per hidden unit [w0[h][0..N_IN-1], b0[h], w1[0..N_OUT-1][h], pad] read as
//   broadcast LDS.128, two pixels per thread.
template <int N_IN, int N_OUT>
__global__ void __launch_bounds__(256) k_dec_syn_fused_t(const int* __restrict__ x, long long n_pix, int n_hid,
                                                         const int* __restrict__ w0, const int* __restrict__ b0,
                                                         const int* __restrict__ w1, const int* __restrict__ b1,
                                                         int* __restrict__ y) {
  constexpr int ROW = (N_IN + 1 + N_OUT + 3) & ~3;
  extern __shared__ __align__(16) int sw[];
  for (int i = threadIdx.x; i < n_hid * ROW; i += blockDim.x) {
    const int hdn = i / ROW, j = i - hdn * ROW;
    int v = 0;
    if (j < N_IN) v = w0[hdn * N_IN + j];
    else if (j == N_IN) v = b0[hdn];
    else if (j < N_IN + 1 + N_OUT) v = w1[(j - N_IN - 1) * n_hid + hdn];
    sw[i] = v;
  }
  int bias1[N_OUT];
#pragma unroll
  for (int o = 0; o < N_OUT; ++o) bias1[o] = b1[o];
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long p0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; p0 < n_pix; p0 += 2 * stride) {
    const long long p1 = p0 + stride;
    const bool two = p1 < n_pix;
    int in0[N_IN], in1[N_IN], a0[N_OUT], a1[N_OUT];
#pragma unroll
    for (int i = 0; i < N_IN; ++i) {
      in0[i] = x[(long long)i * n_pix + p0];
      in1[i] = two ? x[(long long)i * n_pix + p1] : 0;
    }
#pragma unroll
    for (int o = 0; o < N_OUT; ++o) a0[o] = a1[o] = bias1[o];
    for (int hdn = 0; hdn < n_hid; ++hdn) {
      int r[ROW];
#pragma unroll
      for (int j = 0; j < ROW; j += 4) {
        const int4 t = *reinterpret_cast<const int4*>(&sw[hdn * ROW + j]);
        r[j] = t.x; r[j + 1] = t.y; r[j + 2] = t.z; r[j + 3] = t.w;
      }
      int v0 = r[N_IN], v1 = r[N_IN];
#pragma unroll
      for (int i = 0; i < N_IN; ++i) {
        v0 += in0[i] * r[i];
        v1 += in1[i] * r[i];
      }
      v0 = v0 < 0 ? 0 : (v0 >> SYN_PREC);
      v1 = v1 < 0 ? 0 : (v1 >> SYN_PREC);
#pragma unroll
      for (int o = 0; o < N_OUT; ++o) {
        a0[o] += v0 * r[N_IN + 1 + o];
        a1[o] += v1 * r[N_IN + 1 + o];
      }
    }
#pragma unroll
    for (int o = 0; o < N_OUT; ++o) {
      y[(long long)o * n_pix + p0] = trunc_shift(a0[o], SYN_PREC);
      if (two) y[(long long)o * n_pix + p1] = trunc_shift(a1[o], SYN_PREC);
    }
  }
}

//   3x3, 3 -> 3: CTA = 32 x 16 outputs, replicate-clamped 34 x 18 input tile in shared memory, a thread owns a 4-row
//   strip and keeps the 81 + 3 weights in registers (one column of 6 inputs feeds 36 multiply-adds); the last layer
//   can write the 8-bit pixels (ppm_out) in the same pass.
constexpr int D3_W = 32, D3_H = 16, D3_P = 4, D3_THREADS = 128;
__global__ void __launch_bounds__(D3_THREADS) k_dec_conv3_t(const int* __restrict__ x, int H, int W, const int* __restrict__ w,
                                                            const int* __restrict__ b, int residual, int relu,
                                                            int* __restrict__ y, unsigned char* __restrict__ pix) {
  constexpr int SW_ = D3_W + 2, SH_ = D3_H + 2;
  __shared__ int sx[3 * SH_ * SW_];
  __shared__ __align__(16) int swt[84];
  const long long HW = (long long)H * W;
  const int y0 = blockIdx.y * D3_H, x0 = blockIdx.x * D3_W;
  if (threadIdx.x < 81) swt[threadIdx.x] = w[threadIdx.x];
  else if (threadIdx.x < 84) swt[threadIdx.x] = b[threadIdx.x - 81];
  for (int i = threadIdx.x; i < 3 * SH_ * SW_; i += D3_THREADS) {
    const int c = i / (SH_ * SW_), r = (i - c * SH_ * SW_) / SW_, q = i - c * SH_ * SW_ - r * SW_;
    const int sy = min(max(y0 - 1 + r, 0), H - 1), sxx = min(max(x0 - 1 + q, 0), W - 1);
    sx[i] = x[(long long)c * HW + (long long)sy * W + sxx];
  }
  __syncthreads();
  int wr[84];
#pragma unroll
  for (int i = 0; i < 84; i += 4) {
    const int4 t = *reinterpret_cast<const int4*>(&swt[i]);
    wr[i] = t.x; wr[i + 1] = t.y; wr[i + 2] = t.z; wr[i + 3] = t.w;
  }
  const int lx = threadIdx.x & 31, ly0 = (threadIdx.x >> 5) * D3_P;
  int o[D3_P][3], ctr[D3_P][3];
#pragma unroll
  for (int i = 0; i < D3_P; ++i)
#pragma unroll
    for (int oc = 0; oc < 3; ++oc) o[i][oc] = wr[81 + oc];
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      int v[D3_P + 2];
#pragma unroll
      for (int r = 0; r < D3_P + 2; ++r) v[r] = sx[(c * SH_ + ly0 + r) * SW_ + lx + kx];
      if (kx == 1) {
#pragma unroll
        for (int i = 0; i < D3_P; ++i) ctr[i][c] = v[i + 1];
      }
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int oc = 0; oc < 3; ++oc)
#pragma unroll
          for (int i = 0; i < D3_P; ++i) o[i][oc] += wr[((oc * 3 + c) * 3 + ky) * 3 + kx] * v[i + ky];
    }
  const int gx = x0 + lx;
  if (gx >= W) return;
#pragma unroll
  for (int i = 0; i < D3_P; ++i) {
    const int gy = y0 + ly0 + i;
    if (gy >= H) break;
    const long long p = (long long)gy * W + gx;
#pragma unroll
    for (int oc = 0; oc < 3; ++oc) {
      int sacc = o[i][oc];
      if (residual) sacc += ctr[i][oc] << SYN_PREC;
      const int v = sacc < 0 ? (relu ? 0 : -((-sacc) >> SYN_PREC)) : (sacc >> SYN_PREC);
      y[(long long)oc * HW + p] = v;
      if (pix) pix[3 * p + oc] = (unsigned char)min(max((v * 255 + (1 << (SYN_PREC - 1))) >> SYN_PREC, 0), 255);
    }
  }
}

// ---- ppm_out (ccdecapi.cpp:59-128): [3][H][W] at 12 bits -> interleaved 8-bit RGB
__global__ void __launch_bounds__(256) k_dec_pixels(const int* __restrict__ x, long long HW, unsigned char* __restrict__ pix) {
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (long long)gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      int v = (x[(long long)c * HW + p] * 255 + (1 << (SYN_PREC - 1))) >> SYN_PREC;
      pix[3 * p + c] = (unsigned char)min(max(v, 0), 255);
    }
  }
}

DecFilt make_filt(const int32_t* host_k, int n) {
  DecFilt f = {};
  f.n = n;
  for (int i = 0; i < n; ++i) f.k[i] = host_k[i];
  return f;
}

int grid_1d(const ccb_dec_geom& g, long long n) {
  long long nb = (n + 255) / 256, cap = (long long)g.n_sm * 8;
  return (int)(nb < cap ? (nb < 1 ? 1 : nb) : cap);
}

}  // namespace

extern "C" size_t ccb_dec_int_workspace_bytes(int H, int W, int n_latents, int n_syn, const ccb_dec_layer* layers) {
  ccb_dec_geom g;
  if (dec_geom(H, W, n_latents, &g) != CCB_OK || !layers) return 0;
  const long long HW = (long long)H * W;
  // level stacks 1 .. n-2 | level 0 stack (when the caller does not want it) | two ping-pong images
  return (size_t)(g.pyr_total + (long long)n_latents * HW + 2LL * dec_max_planes(n_latents, n_syn, layers) * HW) * sizeof(int32_t);
}

extern "C" int ccb_dec_int(int H, int W, int n_latents, const int32_t* latents, int ups_k, int n_ups,
                           const int32_t* ups_w_host, int pre_k, int n_pre, const int32_t* pre_w_host, int n_syn,
                           const ccb_dec_layer* layers, void* workspace, size_t workspace_bytes, int32_t* ups_out,
                           int32_t* out, uint8_t* pixels, void* stream) {
  if (!latents || !ups_w_host || !pre_w_host || !layers || !workspace || !out) return CCB_ERR_INVALID;
  // two interleaved filters of EVEN length (the odd phase reads one sample past a half-length padding otherwise:
  // ups_upsample_cpu.hpp:24-25,52), odd refine filter
  if (ups_k < 4 || ups_k > DEC_MAX_K || (ups_k & 3) || pre_k < 1 || pre_k > 7 || !(pre_k & 1)) return CCB_ERR_UNSUPPORTED;
  if (n_ups < 1 || n_pre < 1 || n_syn < 1 || n_syn > CCB_MAX_SYN_LAYERS) return CCB_ERR_INVALID;
  ccb_dec_geom g;
  int rc = dec_geom(H, W, n_latents, &g);
  if (rc) return rc;
  if (workspace_bytes < ccb_dec_int_workspace_bytes(H, W, n_latents, n_syn, layers)) return CCB_ERR_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  const int n = n_latents;
  const long long HW = (long long)H * W;
  int* ws = (int*)workspace;
  int* lvl0 = ups_out ? ups_out : ws + g.pyr_total;
  int* ping = ws + g.pyr_total + (long long)n * HW;
  const int mp = dec_max_planes(n, n_syn, layers);
  int* pong = ping + (long long)mp * HW;
  auto stack = [&](int t) -> int* { return t == 0 ? lvl0 : ws + g.pyr_off[t]; };

  // ---- run_ups, stage by stage (coarse to fine)
  for (int t = n - 2; t >= 0; --t) {
    int* dst = stack(t);
    const int h = g.h[t], w = g.w[t];
    DecStage a = {};
    a.fr = make_filt(pre_w_host + (size_t)((n - 2 - t) % n_pre) * pre_k, pre_k);
    a.fu = make_filt(ups_w_host + (size_t)((n - 2 - t) % n_ups) * ups_k, ups_k);
    const bool raw = (t == n - 2);  // the coarsest level is the latent itself, at ARM precision
    a.lat = latents + g.off[t];
    a.src = raw ? latents + g.off[n - 1] : stack(t + 1);
    a.dst = dst;
    a.h = h; a.w = w; a.hs = g.h[t + 1]; a.ws = g.w[t + 1];
    a.n_up = n - 1 - t;
    a.src_shift = raw ? ARM_PREC : 0;
    a.src_prec = raw ? ARM_PREC : UPS_PREC;
    a.ref_tx = ceil_div(w, DT);
    a.ref_tiles = a.ref_tx * ceil_div(h, DT);
    a.up_tx = ceil_div(a.ws, DT);
    a.up_tiles = a.up_tx * ceil_div(a.hs, DT);
    k_dec_stage<<<a.ref_tiles + a.up_tiles * a.n_up, 256, 0, s>>>(a);
    CCB_LAUNCH_CHECK();
  }

  // ---- run_syn_branch (single branch)
  const int* cur = lvl0;
  int n_in = n, i0 = 0;
  auto target = [&](int layer_idx, const int* in) -> int* {
    if (layer_idx == n_syn - 1) return out;
    return in == ping ? pong : ping;
  };
  if (n_syn >= 2 && layers[0].ks == 1 && layers[1].ks == 1) {
    const int nh = layers[0].n_out, no = layers[1].n_out;
    if (n_in > DEC_MAX_IN || no > DEC_MAX_OUT) return CCB_ERR_UNSUPPORTED;
    int* dst = target(1, cur);
    const size_t sm = (size_t)(nh * n_in + nh + no * nh + no) * sizeof(int);
    if (sm > 48 * 1024) return CCB_ERR_UNSUPPORTED;
    if (n_in == 7 && no == 3 && nh * 12 * sizeof(int) <= 48 * 1024) {
      k_dec_syn_fused_t<7, 3><<<grid_1d(g, (HW + 1) / 2), 256, (size_t)nh * 12 * sizeof(int), s>>>(
          cur, HW, nh, layers[0].w, layers[0].b, layers[1].w, layers[1].b, dst);
    } else {
      k_dec_syn_fused<<<grid_1d(g, HW), 256, sm, s>>>(cur, HW, n_in, nh, no, layers[0].w, layers[0].b, layers[1].w,
                                                       layers[1].b, dst);
    }
    CCB_LAUNCH_CHECK();
    cur = dst;
    n_in = no;
    i0 = 2;
  }
  bool pixels_done = false;
  for (int i = i0; i < n_syn; ++i) {
    const ccb_dec_layer& L = layers[i];
    if ((L.ks != 1 && L.ks != 3) || L.n_out > DEC_MAX_OUT || (L.residual && L.n_out > n_in)) return CCB_ERR_UNSUPPORTED;
    int* dst = target(i, cur);
    const size_t sm = (size_t)(L.n_out * n_in * L.ks * L.ks + L.n_out) * sizeof(int);
    if (sm > 48 * 1024) return CCB_ERR_UNSUPPORTED;
    if (L.ks == 3 && n_in == 3 && L.n_out == 3) {
      const bool last = (i == n_syn - 1);
      k_dec_conv3_t<<<dim3(ceil_div(W, D3_W), ceil_div(H, D3_H)), D3_THREADS, 0, s>>>(cur, H, W, L.w, L.b, L.residual, L.relu,
                                                                                     dst, last ? pixels : nullptr);
      if (last && pixels) pixels_done = true;
    } else {
      k_dec_syn_layer<<<grid_1d(g, HW), 256, sm, s>>>(cur, H, W, n_in, L.n_out, L.ks, L.w, L.b, L.residual, L.relu, dst);
    }
    CCB_LAUNCH_CHECK();
    cur = dst;
    n_in = L.n_out;
  }
  if (cur != out) return CCB_ERR_INVALID;
  if (pixels && !pixels_done) {
    if (n_in != 3) return CCB_ERR_UNSUPPORTED;
    k_dec_pixels<<<grid_1d(g, HW), 256, 0, s>>>(out, HW, pixels);
    CCB_LAUNCH_CHECK();
  }
  return CCB_OK;
}
