// S: synthesis conv stack, forward and backward.
// Restates enc/component/core/synthesis.py:69-84 (SynthesisConv2d: replicate pad (k-1)/2,
// conv2d, + x if residual, ReLU or identity) and :175-277 (Synthesis = sequence of those).
// Supported shape (every decoder config shipped under cfg/: SURVEY.md 8): two 1x1 layers
// (L -> S ReLU, S -> 3) followed by any number of 3x3 3->3 layers.
//   * the two 1x1 layers are fused per pixel: the S-wide hidden activation lives in
//     registers only (at 768x512 it would be 75 MB in HBM);
//   * each 3x3 layer is a shared-memory stencil over a replicate-clamped tile.
#include "ccb_internal.cuh"

namespace {

constexpr int CIN = 7;    // = number of latent resolutions
constexpr int C3 = 3;     // image channels
constexpr int HEAD_P = 4; // pixels per thread in the fused 1x1 pair
constexpr int HW_ROW = 12;  // smem floats per hidden unit: w0[7], b0, w1[3], pad

struct HeadArgs {
  const float* z;   // [B,7,H,W]
  float* out;       // [B,3,H,W]
  float* out2;      // optional second copy (kept for the ReLU mask of the backward)
  const float* w0;  // [S,7]
  const float* b0;  // [S]
  const float* w1;  // [3,S]
  const float* b1;  // [3]
  int S, relu1;
  long long HW;     // H*W
  long long n_pix;  // B*H*W
  // backward
  const float* d_out;  // [B,3,H,W] gradient w.r.t. the layer-1 OUTPUT (post activation)
  const float* act;    // [B,3,H,W] layer-1 output (for its ReLU mask), may be NULL if !relu1
  float* dz;           // [B,7,H,W]
  float* part;         // [gridDim.x][n_head_params]
};

__device__ __forceinline__ void head_load_weights(const HeadArgs& a, float* sw) {
  for (int i = threadIdx.x; i < a.S * HW_ROW; i += blockDim.x) {
    int s = i / HW_ROW, c = i - s * HW_ROW;
    float v = 0.f;
    if (c < CIN) v = a.w0[s * CIN + c];
    else if (c == CIN) v = a.b0[s];
    else if (c < CIN + 1 + C3) v = a.w1[(c - CIN - 1) * a.S + s];
    sw[i] = v;
  }
}

__global__ void __launch_bounds__(256) k_syn_head_fwd(HeadArgs a) {
  extern __shared__ __align__(16) float sw[];
  head_load_weights(a, sw);
  __syncthreads();
  const long long base = ((long long)blockIdx.x * blockDim.x) * HEAD_P + threadIdx.x;
  float z[HEAD_P][CIN], o[HEAD_P][C3];
  long long pb[HEAD_P];
  bool ok[HEAD_P];
#pragma unroll
  for (int p = 0; p < HEAD_P; ++p) {
    const long long pix = base + (long long)p * blockDim.x;
    ok[p] = pix < a.n_pix;
    const long long b = ok[p] ? pix / a.HW : 0;
    const long long r = ok[p] ? pix - b * a.HW : 0;
    pb[p] = b * a.HW * CIN + r;
#pragma unroll
    for (int c = 0; c < CIN; ++c) z[p][c] = ok[p] ? a.z[pb[p] + c * a.HW] : 0.f;
#pragma unroll
    for (int c = 0; c < C3; ++c) o[p][c] = a.b1[c];
  }
  for (int s = 0; s < a.S; ++s) {
    const float4 q0 = *reinterpret_cast<const float4*>(sw + s * HW_ROW);
    const float4 q1 = *reinterpret_cast<const float4*>(sw + s * HW_ROW + 4);
    const float4 q2 = *reinterpret_cast<const float4*>(sw + s * HW_ROW + 8);
#pragma unroll
    for (int p = 0; p < HEAD_P; ++p) {
      float h = q1.w;
      h = fmaf(q0.x, z[p][0], h); h = fmaf(q0.y, z[p][1], h); h = fmaf(q0.z, z[p][2], h);
      h = fmaf(q0.w, z[p][3], h); h = fmaf(q1.x, z[p][4], h); h = fmaf(q1.y, z[p][5], h);
      h = fmaf(q1.z, z[p][6], h);
      h = fmaxf(h, 0.f);
      o[p][0] = fmaf(q2.x, h, o[p][0]);
      o[p][1] = fmaf(q2.y, h, o[p][1]);
      o[p][2] = fmaf(q2.z, h, o[p][2]);
    }
  }
#pragma unroll
  for (int p = 0; p < HEAD_P; ++p) {
    if (!ok[p]) continue;
    const long long pix = base + (long long)p * blockDim.x;
    const long long b = pix / a.HW, r = pix - b * a.HW;
#pragma unroll
    for (int c = 0; c < C3; ++c) {
      float v = o[p][c];
      if (a.relu1) v = fmaxf(v, 0.f);
      a.out[(b * C3 + c) * a.HW + r] = v;
      if (a.out2) a.out2[(b * C3 + c) * a.HW + r] = v;
    }
  }
}

// Backward of the fused pair.  Weight-gradient layout inside a partial row:
//   [w0: S*7][b0: S][w1: 3*S][b1: 3]
__global__ void __launch_bounds__(256) k_syn_head_bwd(HeadArgs a) {
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
    ok[p] = pix < a.n_pix;
    const long long b = ok[p] ? pix / a.HW : 0;
    const long long r = ok[p] ? pix - b * a.HW : 0;
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
  float* prow = a.part + (long long)blockIdx.x * n_par;
  for (int i = threadIdx.x; i < n_par; i += blockDim.x) prow[i] = sacc[i];
}

// ------------------------------------------------------------------------------------------
// 3x3, 3 -> 3 layers
// ------------------------------------------------------------------------------------------
constexpr int CT_W = 32, CT_H = 8;  // tile = 256 pixels, one per thread
constexpr int CS_W = CT_W + 2, CS_H = CT_H + 2;

struct ConvArgs {
  const float* x;  // [B,3,H,W] layer input
  float* y;        // [B,3,H,W] layer output
  float* y2;       // optional second copy of the output
  const float* w;  // [3,3,3,3] (cout, cin, ky, kx)
  const float* b;  // [3]
  int B, H, W, residual, relu;
  int tiles_x, tiles_y;
  // backward
  const float* dy;  // gradient w.r.t. the layer output
  float* dx;        // gradient w.r.t. the layer input
  float* part;      // [gridDim.x][84]
};

__global__ void __launch_bounds__(256) k_syn_conv3_fwd(ConvArgs a) {
  __shared__ float sx[C3 * CS_H * CS_W];
  __shared__ float swt[C3 * C3 * 9 + C3];
  const int tile = blockIdx.x;
  const int b = tile / (a.tiles_x * a.tiles_y);
  const int tr = tile - b * a.tiles_x * a.tiles_y;
  const int y0 = (tr / a.tiles_x) * CT_H, x0 = (tr % a.tiles_x) * CT_W;
  const long long HW = (long long)a.H * a.W;
  if (threadIdx.x < 81) swt[threadIdx.x] = a.w[threadIdx.x];
  if (threadIdx.x < 3) swt[81 + threadIdx.x] = a.b[threadIdx.x];
  for (int i = threadIdx.x; i < C3 * CS_H * CS_W; i += blockDim.x) {
    int c = i / (CS_H * CS_W), r = i - c * CS_H * CS_W;
    int ry = r / CS_W, rx = r - ry * CS_W;
    int gy = min(max(y0 - 1 + ry, 0), a.H - 1), gx = min(max(x0 - 1 + rx, 0), a.W - 1);
    sx[i] = a.x[((long long)b * C3 + c) * HW + (long long)gy * a.W + gx];
  }
  __syncthreads();
  const int ly = threadIdx.x / CT_W, lx = threadIdx.x % CT_W;
  const int gy = y0 + ly, gx = x0 + lx;
  if (gy >= a.H || gx >= a.W) return;
  float o[C3] = {swt[81], swt[82], swt[83]};
#pragma unroll
  for (int c = 0; c < C3; ++c)
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const float v = sx[(c * CS_H + ly + ky) * CS_W + lx + kx];
#pragma unroll
        for (int oc = 0; oc < C3; ++oc) o[oc] = fmaf(swt[((oc * C3 + c) * 3 + ky) * 3 + kx], v, o[oc]);
      }
#pragma unroll
  for (int oc = 0; oc < C3; ++oc) {
    float v = o[oc];
    if (a.residual) v += sx[(oc * CS_H + ly + 1) * CS_W + lx + 1];
    if (a.relu) v = fmaxf(v, 0.f);
    a.y[((long long)b * C3 + oc) * HW + (long long)gy * a.W + gx] = v;
    if (a.y2) a.y2[((long long)b * C3 + oc) * HW + (long long)gy * a.W + gx] = v;
  }
}

// Backward: persistent CTAs loop over tiles; each thread keeps its 81+3 weight-gradient
// sums in registers across tiles and the CTA reduces them once at the end.
__global__ void __launch_bounds__(256) k_syn_conv3_bwd(ConvArgs a) {
  __shared__ float sx[C3 * CS_H * CS_W];   // input, replicate-clamped
  __shared__ float sg[C3 * CS_H * CS_W];   // d pre-activation, zero outside the image
  __shared__ float swt[C3 * C3 * 9];
  __shared__ float red[32];
  const long long HW = (long long)a.H * a.W;
  if (threadIdx.x < 81) swt[threadIdx.x] = a.w[threadIdx.x];
  float gw[81], gb[3];
#pragma unroll
  for (int i = 0; i < 81; ++i) gw[i] = 0.f;
  gb[0] = gb[1] = gb[2] = 0.f;
  const int n_tiles = a.B * a.tiles_x * a.tiles_y;
  const int ly = threadIdx.x / CT_W, lx = threadIdx.x % CT_W;

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int b = tile / (a.tiles_x * a.tiles_y);
    const int tr = tile - b * a.tiles_x * a.tiles_y;
    const int y0 = (tr / a.tiles_x) * CT_H, x0 = (tr % a.tiles_x) * CT_W;
    __syncthreads();
    for (int i = threadIdx.x; i < C3 * CS_H * CS_W; i += blockDim.x) {
      int c = i / (CS_H * CS_W), r = i - c * CS_H * CS_W;
      int ry = r / CS_W, rx = r - ry * CS_W;
      int uy = y0 - 1 + ry, ux = x0 - 1 + rx;
      int gy = min(max(uy, 0), a.H - 1), gx = min(max(ux, 0), a.W - 1);
      const long long o = ((long long)b * C3 + c) * HW + (long long)gy * a.W + gx;
      sx[i] = a.x[o];
      float g = 0.f;
      if (uy >= 0 && uy < a.H && ux >= 0 && ux < a.W) {
        g = a.dy[o];
        if (a.relu && !(a.y[o] > 0.f)) g = 0.f;
      }
      sg[i] = g;
    }
    __syncthreads();
    const int gy = y0 + ly, gx = x0 + lx;
    if (gy < a.H && gx < a.W) {
      // ---- weight gradients: dW[o][c][ky][kx] += dpre[o](y,x) * xpad[c](y+ky-1, x+kx-1)
      float gpre[C3];
#pragma unroll
      for (int oc = 0; oc < C3; ++oc) {
        gpre[oc] = sg[(oc * CS_H + ly + 1) * CS_W + lx + 1];
        gb[oc] += gpre[oc];
      }
#pragma unroll
      for (int c = 0; c < C3; ++c)
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const float v = sx[(c * CS_H + ly + ky) * CS_W + lx + kx];
#pragma unroll
            for (int oc = 0; oc < C3; ++oc)
              gw[((oc * C3 + c) * 3 + ky) * 3 + kx] = fmaf(gpre[oc], v, gw[((oc * C3 + c) * 3 + ky) * 3 + kx]);
          }
      // ---- input gradient.  The padded input position (py,px) receives
      //   sum_{o,ky,kx} W[o][c][ky][kx] * dpre[o](py-ky+1, px-kx+1);  replicate padding folds
      // the padded ring onto the border pixels.
      float dxv[C3] = {0.f, 0.f, 0.f};
      const int py_lo = (gy == 0) ? -1 : gy, py_hi = (gy == a.H - 1) ? a.H : gy;
      const int px_lo = (gx == 0) ? -1 : gx, px_hi = (gx == a.W - 1) ? a.W : gx;
      for (int py = py_lo; py <= py_hi; ++py)
        for (int px = px_lo; px <= px_hi; ++px) {
#pragma unroll
          for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
              const int sy_ = py - ky + 1, sx_ = px - kx + 1;  // output position that read (py,px)
              // tile-local coordinates (+1 halo); positions outside the image hold 0 in sg,
              // positions outside the halo can only be outside the image too
              const int ty = sy_ - y0 + 1, tx = sx_ - x0 + 1;
              if (ty < 0 || ty >= CS_H || tx < 0 || tx >= CS_W) continue;
#pragma unroll
              for (int oc = 0; oc < C3; ++oc) {
                const float g = sg[(oc * CS_H + ty) * CS_W + tx];
#pragma unroll
                for (int c = 0; c < C3; ++c) dxv[c] = fmaf(swt[((oc * C3 + c) * 3 + ky) * 3 + kx], g, dxv[c]);
              }
            }
        }
#pragma unroll
      for (int c = 0; c < C3; ++c) {
        float v = dxv[c];
        if (a.residual) v += sg[(c * CS_H + ly + 1) * CS_W + lx + 1];
        a.dx[((long long)b * C3 + c) * HW + (long long)gy * a.W + gx] = v;
      }
    }
  }
  float* prow = a.part + (long long)blockIdx.x * 84;
#pragma unroll
  for (int i = 0; i < 81; ++i) {
    float v = block_sum(gw[i], red);
    if (threadIdx.x == 0) prow[i] = v;
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float v = block_sum(gb[i], red);
    if (threadIdx.x == 0) prow[81 + i] = v;
  }
}

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
  a.n_pix = a.HW * p->arch.B;
  return a;
}

ConvArgs conv_args(const ccb_plan* p, const ccb_tensors* prm, int layer) {
  ConvArgs a = {};
  a.w = prm->syn_w[layer];
  a.b = prm->syn_b[layer];
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

int ccb_syn_head_blocks(const ccb_plan* p) {
  long long n_pix = (long long)p->arch.B * p->arch.H * p->arch.W;
  return (int)((n_pix + 256LL * HEAD_P - 1) / (256LL * HEAD_P));
}
int ccb_syn_head_params(const ccb_plan* p) { return p->syn_hidden * CIN + p->syn_hidden + C3 * p->syn_hidden + C3; }
int ccb_syn_conv_blocks(const ccb_plan* p) {
  int n_tiles = p->arch.B * ceil_div(p->arch.W, CT_W) * ceil_div(p->arch.H, CT_H);
  int cap = p->n_sm * 4;
  return n_tiles < cap ? n_tiles : cap;
}

// Layer i >= 1 writes its output to ws act[i]; the last layer writes raw_out.
int ccb_syn_forward(const ccb_plan* p, const ccb_tensors* prm, char* ws, float* raw_out, cudaStream_t s) {
  const int nl = p->arch.n_syn_layers;
  HeadArgs h = head_args(p, prm, ws);
  h.out = (nl == 2) ? raw_out : reinterpret_cast<float*>(ws + p->off_act[1]);
  h.out2 = (nl == 2 && h.relu1) ? reinterpret_cast<float*>(ws + p->off_act[1]) : nullptr;
  size_t sm = (size_t)p->syn_hidden * HW_ROW * sizeof(float);
  k_syn_head_fwd<<<ccb_syn_head_blocks(p), 256, sm, s>>>(h);
  CCB_LAUNCH_CHECK();
  for (int l = 2; l < nl; ++l) {
    ConvArgs c = conv_args(p, prm, l);
    c.x = reinterpret_cast<const float*>(ws + p->off_act[l - 1]);
    c.y = (l == nl - 1) ? raw_out : reinterpret_cast<float*>(ws + p->off_act[l]);
    c.y2 = (l == nl - 1 && c.relu) ? reinterpret_cast<float*>(ws + p->off_act[l]) : nullptr;
    k_syn_conv3_fwd<<<c.B * c.tiles_x * c.tiles_y, 256, 0, s>>>(c);
    CCB_LAUNCH_CHECK();
  }
  return CCB_OK;
}

// d_raw_out -> d z (= d up_{L-1}) and the partial weight-gradient rows.
// The last layer's output is needed only for its ReLU mask; when the last layer has a ReLU
// the caller's raw_out is not retained, so forward also keeps a copy in ws act[nl-1].
int ccb_syn_backward(const ccb_plan* p, const ccb_tensors* prm, char* ws, const float* d_raw_out,
                     cudaStream_t s) {
  const int nl = p->arch.n_syn_layers;
  const float* dcur = d_raw_out;
  for (int l = nl - 1; l >= 2; --l) {
    ConvArgs c = conv_args(p, prm, l);
    c.x = reinterpret_cast<const float*>(ws + p->off_act[l - 1]);
    c.y = reinterpret_cast<float*>(ws + p->off_act[l]);
    c.dy = dcur;
    c.dx = reinterpret_cast<float*>(ws + p->off_dact[(l & 1)]);
    c.part = reinterpret_cast<float*>(ws + p->off_syn_part[l]);
    k_syn_conv3_bwd<<<ccb_syn_conv_blocks(p), 256, 0, s>>>(c);
    CCB_LAUNCH_CHECK();
    dcur = c.dx;
  }
  HeadArgs h = head_args(p, prm, ws);
  h.d_out = dcur;
  h.act = reinterpret_cast<const float*>(ws + p->off_act[1]);
  h.dz = reinterpret_cast<float*>(ws + p->off_dup[p->g.L - 1]);
  h.part = reinterpret_cast<float*>(ws + p->off_syn_part[0]);
  size_t sm = (size_t)(p->syn_hidden * HW_ROW + ccb_syn_head_params(p)) * sizeof(float);
  k_syn_head_bwd<<<ccb_syn_head_blocks(p), 256, sm, s>>>(h);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
