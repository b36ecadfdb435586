// I1 + I2: fixed-point ARM on the encoder side, every position in parallel.
// Restates cpp/arm_cpu.cpp:18-106 (custom_conv_11_int32_cpu_X_X_X, == enc/bitstream/armint.py:81-111,
// 244-261) and cpp/cc-contexts.h:20-48 (get_val_mu_indicies).  The decoder runs this loop
// sequentially because each latent is entropy-decoded from the previous ones; the encoder
// knows all quantised latents, so the identical int32 arithmetic (wrap-around, ARM_PRECISION 8)
// runs as one thread per position.
#include "ccb_internal.cuh"

namespace {

constexpr int ARM_PRECISION = 8;
constexpr int ARM_SCALE = 1 << ARM_PRECISION;
constexpr int N_MUQ = 16, N_SIGQ = 50, SIG_LOG_MIN = -1, SIG_LOG_MAX_EXCL = 9;

struct IntArgs {
  LevelGeom g;
  const int32_t* yq;
  const int32_t* w[CCB_MAX_ARM_LAYERS];
  const int32_t* b[CCB_MAX_ARM_LAYERS];
  int nh;
  int32_t *mu, *ls, *mu_r, *mu_i, *sig_i;
  signed char dy[32], dx[32];
};

template <int C>
__global__ void __launch_bounds__(256) k_arm_int(IntArgs a) {
  extern __shared__ int32_t swi[];
  // [k][C*C + C] hidden layers then [2*C + 2]
  for (int k = 0; k < a.nh; ++k) {
    int32_t* d = swi + k * (C * C + C);
    for (int i = threadIdx.x; i < C * C; i += blockDim.x) d[i] = a.w[k][i];
    for (int i = threadIdx.x; i < C; i += blockDim.x) d[C * C + i] = a.b[k][i];
  }
  {
    int32_t* d = swi + a.nh * (C * C + C);
    for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) d[i] = a.w[a.nh][i];
    if (threadIdx.x < 2) d[2 * C + threadIdx.x] = a.b[a.nh][threadIdx.x];
  }
  __syncthreads();
  const long long total = (long long)a.g.B * a.g.M;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int bb = (int)(idx / a.g.M);
    const long long r = idx - (long long)bb * a.g.M;
    int lvl = 0;
#pragma unroll
    for (int l = 1; l < CCB_MAX_LATENTS; ++l)
      if (l < a.g.L && r >= a.g.off[l]) lvl = l;
    const int w = a.g.w[lvl], h = a.g.h[lvl];
    const long long rel = r - a.g.off[lvl];
    const int y = (int)(rel / w), x = (int)(rel - (long long)y * w);
    const int32_t* src = a.yq + (long long)bb * a.g.M + a.g.off[lvl];
    int32_t in[C], out[C];
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int yy = y + a.dy[c], xx = x + a.dx[c];
      int32_t v = 0;
      if (yy >= 0 && yy < h && xx >= 0 && xx < w) v = src[(long long)yy * w + xx];
      in[c] = (int32_t)((uint32_t)v << ARM_PRECISION);  // the decoder stores xx << ARM_PRECISION
    }
    for (int k = 0; k < a.nh; ++k) {
      const int32_t* wk = swi + k * (C * C + C);
#pragma unroll
      for (int j = 0; j < C; ++j) {
        uint32_t acc = (uint32_t)wk[C * C + j] + (uint32_t)in[j] * (uint32_t)ARM_SCALE;  // residual
#pragma unroll
        for (int i = 0; i < C; ++i) acc += (uint32_t)in[i] * (uint32_t)wk[j * C + i];
        const int32_t s = (int32_t)acc;
        out[j] = s < 0 ? 0 : ((s + ARM_SCALE / 2) >> ARM_PRECISION);
      }
#pragma unroll
      for (int j = 0; j < C; ++j) in[j] = out[j];
    }
    const int32_t* wo = swi + a.nh * (C * C + C);
    int32_t res[2];
#pragma unroll
    for (int o = 0; o < 2; ++o) {
      uint32_t acc = (uint32_t)wo[2 * C + o];
#pragma unroll
      for (int i = 0; i < C; ++i) acc += (uint32_t)in[i] * (uint32_t)wo[o * C + i];
      const int32_t s = (int32_t)acc;
      res[o] = s < 0 ? -((-s + ARM_SCALE / 2) >> ARM_PRECISION) : ((s + ARM_SCALE / 2) >> ARM_PRECISION);
    }
    a.mu[idx] = res[0];
    a.ls[idx] = res[1];
    if (a.mu_r || a.mu_i || a.sig_i) {
      // get_val_mu_indicies (cc-contexts.h:20-48)
      const int val_mu = res[0];
      int val_log_sig = res[1];
      int mr = (val_mu >= 0) ? (((val_mu + ARM_SCALE / 2) >> ARM_PRECISION) << ARM_PRECISION)
                             : -((((-val_mu) + ARM_SCALE / 2) >> ARM_PRECISION) << ARM_PRECISION);
      int mi = (val_mu - mr) * N_MUQ;
      mi = mi >= 0 ? ((mi + ARM_SCALE / 2) >> ARM_PRECISION) : -((-mi + ARM_SCALE / 2) >> ARM_PRECISION);
      mi += N_MUQ / 2;
      int si;
      val_log_sig -= SIG_LOG_MIN * ARM_SCALE;
      if (val_log_sig < 0) si = 0;
      else {
        si = val_log_sig * (N_SIGQ / (SIG_LOG_MAX_EXCL - SIG_LOG_MIN)) + ARM_SCALE / 2;
        si >>= ARM_PRECISION;
        if (si >= N_SIGQ) si = N_SIGQ - 1;
      }
      if (a.mu_r) a.mu_r[idx] = mr >> ARM_PRECISION;
      if (a.mu_i) a.mu_i[idx] = mi;
      if (a.sig_i) a.sig_i[idx] = si;
    }
  }
}

const signed char DY8[8] = {-3, -2, -1, -1, -1, 0, 0, 0};
const signed char DX8[8] = {0, 0, -1, 0, 1, -3, -2, -1};
const signed char DY16[16] = {-3, -3, -2, -2, -2, -2, -2, -1, -1, -1, -1, -1, -1, 0, 0, 0};
const signed char DX16[16] = {0, 1, -2, -1, 0, 1, 2, -3, -2, -1, 0, 1, 2, -3, -2, -1};
const signed char DY24[24] = {-4, -3, -3, -3, -3, -3, -2, -2, -2, -2, -2, -2, -2, -1, -1, -1, -1, -1, -1, -1, 0, 0, 0, 0};
const signed char DX24[24] = {0, -2, -1, 0, 1, 2, -3, -2, -1, 0, 1, 2, 3, -3, -2, -1, 0, 1, 2, 3, -4, -3, -2, -1};
const signed char DY32[32] = {-4, -4, -4, -4, -3, -3, -3, -3, -3, -3, -3, -2, -2, -2, -2, -2,
                              -2, -2, -2, -1, -1, -1, -1, -1, -1, -1, -1, -1, 0,  0,  0,  0};
const signed char DX32[32] = {-2, -1, 0, 1, -3, -2, -1, 0, 1, 2, 3, -3, -2, -1, 0, 1,
                              2,  3,  4, -4, -3, -2, -1, 0, 1, 2, 3, 4,  -4, -3, -2, -1};

}  // namespace

extern "C" int ccb_arm_int(const ccb_plan* p, const int32_t* yhat_q, const int32_t* const* w_q,
                           const int32_t* const* b_q, int32_t* mu_q8, int32_t* log_scale_q8,
                           int32_t* mu_rounded, int32_t* mu_index, int32_t* sig_index, void* stream) {
  if (!p || !yhat_q || !w_q || !b_q || !mu_q8 || !log_scale_q8) return CCB_ERR_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  IntArgs a = {};
  a.g = p->g;
  a.yq = yhat_q;
  a.nh = p->arch.n_hidden_arm;
  for (int k = 0; k <= a.nh; ++k) {
    if (!w_q[k] || !b_q[k]) return CCB_ERR_INVALID;
    a.w[k] = w_q[k];
    a.b[k] = b_q[k];
  }
  a.mu = mu_q8;
  a.ls = log_scale_q8;
  a.mu_r = mu_rounded;
  a.mu_i = mu_index;
  a.sig_i = sig_index;
  const int C = p->arch.dim_arm;
  const signed char *dy, *dx;
  switch (C) {
    case 8: dy = DY8; dx = DX8; break;
    case 16: dy = DY16; dx = DX16; break;
    case 24: dy = DY24; dx = DX24; break;
    case 32: dy = DY32; dx = DX32; break;
    default: return CCB_ERR_UNSUPPORTED;
  }
  for (int c = 0; c < C; ++c) {
    a.dy[c] = dy[c];
    a.dx[c] = dx[c];
  }
  const long long total = (long long)p->g.B * p->g.M;
  long long nb = (total + 255) / 256;
  const long long cap = (long long)p->n_sm * 8;
  if (nb > cap) nb = cap;
  if (nb < 1) nb = 1;
  const size_t sm = (size_t)(a.nh * (C * C + C) + 2 * C + 2) * sizeof(int32_t);
  switch (C) {
    case 8: k_arm_int<8><<<(int)nb, 256, sm, s>>>(a); break;
    case 16: k_arm_int<16><<<(int)nb, 256, sm, s>>>(a); break;
    case 24: k_arm_int<24><<<(int)nb, 256, sm, s>>>(a); break;
    case 32: k_arm_int<32><<<(int)nb, 256, sm, s>>>(a); break;
  }
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
