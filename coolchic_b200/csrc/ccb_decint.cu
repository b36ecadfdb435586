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
// use the same filter, so one launch upsamples every plane of a level -- 6 refine + 6 upsampling launches for 7 levels.
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
__global__ void __launch_bounds__(256) k_dec_refine(const int* __restrict__ src, int h, int w, DecFilt f, int src_shift,
                                                    int src_prec, int* __restrict__ dst) {
  __shared__ int s_in[(DT + 2 * 3) * (DT + 2 * 3 + 1)];
  __shared__ int s_h[(DT + 2 * 3) * (DT + 1)];
  const int pad = f.n / 2;
  const int y0 = blockIdx.y * DT, x0 = blockIdx.x * DT;
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
__global__ void __launch_bounds__(256) k_dec_upsample(const int* __restrict__ src, int h, int w, DecFilt f, int src_shift,
                                                      int src_prec, int* __restrict__ dst, int oh, int ow) {
  constexpr int SR = DT + 5;      // source rows / columns a 32 x 32 source tile needs (ks/2 = 4 taps: -2 .. +2)
  __shared__ int s_in[SR * (SR + 1)];
  __shared__ int s_h[SR * (2 * DT + 1)];
  const int ks = f.n / 2, pad = ks / 2;
  const int plane = blockIdx.z;
  const int* sp = src + (long long)plane * h * w;
  int* dp = dst + (long long)plane * oh * ow;
  const int y0 = blockIdx.y * DT, x0 = blockIdx.x * DT;
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
    const DecFilt fr = make_filt(pre_w_host + (size_t)((n - 2 - t) % n_pre) * pre_k, pre_k);
    k_dec_refine<<<dim3(ceil_div(w, DT), ceil_div(h, DT)), 256, 0, s>>>(latents + g.off[t], h, w, fr, ARM_PREC, ARM_PREC, dst);
    CCB_LAUNCH_CHECK();
    const DecFilt fu = make_filt(ups_w_host + (size_t)((n - 2 - t) % n_ups) * ups_k, ups_k);
    const int hs = g.h[t + 1], wsrc = g.w[t + 1];
    const bool raw = (t == n - 2);  // the coarsest level is the latent itself, at ARM precision
    const int* src = raw ? latents + g.off[n - 1] : stack(t + 1);
    k_dec_upsample<<<dim3(ceil_div(wsrc, DT), ceil_div(hs, DT), n - 1 - t), 256, 0, s>>>(
        src, hs, wsrc, fu, raw ? ARM_PREC : 0, raw ? ARM_PREC : UPS_PREC, dst + (long long)h * w, h, w);
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
    k_dec_syn_fused<<<grid_1d(g, HW), 256, sm, s>>>(cur, HW, n_in, nh, no, layers[0].w, layers[0].b, layers[1].w,
                                                     layers[1].b, dst);
    CCB_LAUNCH_CHECK();
    cur = dst;
    n_in = no;
    i0 = 2;
  }
  for (int i = i0; i < n_syn; ++i) {
    const ccb_dec_layer& L = layers[i];
    if ((L.ks != 1 && L.ks != 3) || L.n_out > DEC_MAX_OUT || (L.residual && L.n_out > n_in)) return CCB_ERR_UNSUPPORTED;
    int* dst = target(i, cur);
    const size_t sm = (size_t)(L.n_out * n_in * L.ks * L.ks + L.n_out) * sizeof(int);
    if (sm > 48 * 1024) return CCB_ERR_UNSUPPORTED;
    k_dec_syn_layer<<<grid_1d(g, HW), 256, sm, s>>>(cur, H, W, n_in, L.n_out, L.ks, L.w, L.b, L.residual, L.relu, dst);
    CCB_LAUNCH_CHECK();
    cur = dst;
    n_in = L.n_out;
  }
  if (cur != out) return CCB_ERR_INVALID;
  if (pixels) {
    if (n_in != 3) return CCB_ERR_UNSUPPORTED;
    k_dec_pixels<<<grid_1d(g, HW), 256, 0, s>>>(out, HW, pixels);
    CCB_LAUNCH_CHECK();
  }
  return CCB_OK;
}
