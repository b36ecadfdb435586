// Q: quantiser forward / backward over the flat [B,M] latent vector.
// Restates enc/component/core/quantizer.py:16-41 (softround), :62-107 (kumaraswamy noise),
// :116-232 (quantize) and the gain / AC_MAX_VAL clamp of enc/component/coolchic.py:365-377.
#include "ccb_internal.cuh"

namespace {

struct QuantArgs {
  CPtrTable7 lat;
  PtrTable7 dlat;
  LevelGeom g;
  int q_type, noise_type, ac_max_val;
  float gain;
  float T_host, sigma_host;
  const float* T_dev;
  const float* sigma_dev;
  const float* noise;
  uint64_t seed;
  const uint64_t* offset_dev;
  uint64_t offset_host;
};

__device__ __forceinline__ void locate(const LevelGeom& g, long long r, int& lvl, long long& rel) {
  lvl = 0;
#pragma unroll
  for (int l = 1; l < CCB_MAX_LATENTS; ++l)
    if (l < g.L && r >= g.off[l]) lvl = l;
  rel = r - g.off[lvl];
}

// s(x,T) and, optionally, ds/dx.  Written in the reference's evaluation order
// (tanh(delta/t), division by tanh(1/(2t)); quantizer.py:39-41) -- SURVEY.md H2.
__device__ __forceinline__ float softround_f(float x, float t, float inv_tanh_half_t) {
  float fl = floorf(x);
  float delta = x - fl - 0.5f;
  return fl + 0.5f * tanhf(delta / t) * inv_tanh_half_t + 0.5f;
}
__device__ __forceinline__ float softround_grad(float x, float t, float inv_tanh_half_t) {
  float fl = floorf(x);
  float delta = x - fl - 0.5f;
  float th = tanhf(delta / t);
  return 0.5f * (1.0f - th * th) / t * inv_tanh_half_t;
}

// Raw draw -> additive noise.  raw ~ N(0,1) (gaussian) or U(0,1) (kumaraswamy).
__device__ __forceinline__ float shape_noise(int noise_type, float raw, float sigma, float kum_inv_a,
                                             float kum_inv_b) {
  if (noise_type == CCB_NOISE_GAUSSIAN) return raw * sigma;
  if (noise_type == CCB_NOISE_KUMARASWAMY) {
    // (1 - (1-u)^(1/b))^(1/a) - 0.5   (quantizer.py:100-107)
    float t = 1.0f - powf(1.0f - raw, kum_inv_b);
    return powf(t, kum_inv_a) - 0.5f;
  }
  return 0.f;
}

__device__ __forceinline__ float raw_noise_at(const QuantArgs& a, long long idx, uint64_t offset) {
  if (a.noise) return a.noise[idx];
  // one Philox block per 4 consecutive elements; lanes pick their word
  uint4 r = philox4x32_10((uint64_t)(idx >> 2), offset, a.seed);
  uint32_t w[4] = {r.x, r.y, r.z, r.w};
  int k = (int)(idx & 3);
  if (a.noise_type == CCB_NOISE_GAUSSIAN) {
    // Box-Muller on the pair (k&~1, k|1)
    float u1 = u32_to_unit_open(w[k & 2]);
    float u2 = u32_to_unit_open(w[(k & 2) + 1]);
    float rad = sqrtf(-2.0f * logf(u1));
    float sn, cs;
    sincospif(2.0f * u2, &sn, &cs);
    return (k & 1) ? rad * sn : rad * cs;
  }
  return u32_to_unit_open(w[k]);
}

template <bool BACKWARD>
__global__ void __launch_bounds__(256) k_quant(QuantArgs a, float* __restrict__ yhat,
                                               const float* __restrict__ dyhat, const float* __restrict__ dyhat2) {
  const long long total = (long long)a.g.B * a.g.M;
  const float T = a.T_dev ? *a.T_dev : a.T_host;
  const float sigma = a.sigma_dev ? *a.sigma_dev : a.sigma_host;
  const uint64_t offset = a.offset_dev ? *a.offset_dev : a.offset_host;
  const float inv_th = 1.0f / tanhf(1.0f / (2.0f * T));
  float kum_inv_a = 0.f, kum_inv_b = 0.f;
  if (a.noise_type == CCB_NOISE_KUMARASWAMY) {
    float aa = sigma;
    float bb = (powf(2.0f, aa) * (aa - 1.0f) + 1.0f) / aa;
    kum_inv_a = 1.0f / aa;
    kum_inv_b = 1.0f / bb;
  }
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    int b = (int)(idx / a.g.M);
    long long r = idx - (long long)b * a.g.M;
    int lvl;
    long long rel;
    locate(a.g, r, lvl, rel);
    const long long hw = (long long)a.g.h[lvl] * a.g.w[lvl];
    const long long src = (long long)b * hw + rel;
    const float x = a.lat.p[lvl][src] * a.gain;
    float n = 0.f;
    if (a.noise_type != CCB_NOISE_NONE)
      n = shape_noise(a.noise_type, raw_noise_at(a, idx, offset), sigma, kum_inv_a, kum_inv_b);
    if (!BACKWARD) {
      float y;
      switch (a.q_type) {
        case CCB_Q_NONE: y = x + n; break;
        case CCB_Q_SOFTROUND_ALONE: y = softround_f(x, T, inv_th); break;
        case CCB_Q_SOFTROUND: y = softround_f(softround_f(x, T, inv_th) + n, T, inv_th); break;
        case CCB_Q_STE: {
          // value (s - s) + round(x), evaluated as the reference does (quantizer.py:222-225)
          float sr = softround_f(x, T, inv_th);
          y = (sr - sr) + rintf(x);
        } break;
        default: y = rintf(x); break;  // hardround, true_ste: torch.round == round-half-even
      }
      if (a.ac_max_val != -1) y = fminf(fmaxf(y, -(float)a.ac_max_val), (float)(a.ac_max_val + 1));
      yhat[idx] = y;
    } else {
      float* dl = a.dlat.p[lvl];
      if (dl == nullptr) continue;
      float d = 1.f;
      float y = x;
      switch (a.q_type) {
        case CCB_Q_NONE: y = x + n; break;
        case CCB_Q_SOFTROUND_ALONE:
          d = softround_grad(x, T, inv_th);
          y = softround_f(x, T, inv_th);
          break;
        case CCB_Q_SOFTROUND: {
          float s1 = softround_f(x, T, inv_th) + n;
          d = softround_grad(s1, T, inv_th) * softround_grad(x, T, inv_th);
          y = softround_f(s1, T, inv_th);
        } break;
        case CCB_Q_TRUE_STE: y = rintf(x); break;
        default:  // hardround: zero gradient; ste: the reference returns a detached tensor
          d = 0.f;
          y = rintf(x);
          break;
      }
      if (a.ac_max_val != -1) {
        // torch.clamp backward: pass where min <= y <= max
        if (y < -(float)a.ac_max_val || y > (float)(a.ac_max_val + 1)) d = 0.f;
      }
      const float gy = dyhat2 ? dyhat[idx] + dyhat2[idx] : dyhat[idx];
      dl[src] = gy * d * a.gain;
    }
  }
}

QuantArgs make_args(const ccb_plan* p, const ccb_tensors* prm, const ccb_quant_cfg* q,
                    const ccb_tensors* grads) {
  QuantArgs a;
  for (int i = 0; i < CCB_MAX_LATENTS; ++i) {
    a.lat.p[i] = i < p->g.L ? prm->latent[i] : nullptr;
    a.dlat.p[i] = (grads && i < p->g.L) ? grads->latent[i] : nullptr;
  }
  a.g = p->g;
  a.q_type = q->training ? q->q_type : CCB_Q_HARDROUND;
  a.noise_type = q->training ? q->noise_type : CCB_NOISE_NONE;
  // quantize() only draws noise for the modes that use it
  if (a.q_type != CCB_Q_NONE && a.q_type != CCB_Q_SOFTROUND) a.noise_type = CCB_NOISE_NONE;
  a.ac_max_val = q->ac_max_val;
  a.gain = q->gain;
  a.T_host = q->T_host;
  a.sigma_host = q->sigma_host;
  a.T_dev = q->T_dev;
  a.sigma_dev = q->sigma_dev;
  a.noise = q->noise;
  a.seed = q->seed;
  a.offset_dev = q->offset_dev;
  a.offset_host = q->offset_host;
  return a;
}

int grid_for(const ccb_plan* p) {
  long long total = (long long)p->g.B * p->g.M;
  long long nb = (total + 255) / 256;
  long long cap = (long long)p->n_sm * 8;
  return (int)(nb < cap ? (nb < 1 ? 1 : nb) : cap);
}

}  // namespace

int ccb_quant_forward(const ccb_plan* p, const ccb_tensors* prm, const ccb_quant_cfg* q, float* yhat,
                      cudaStream_t s) {
  QuantArgs a = make_args(p, prm, q, nullptr);
  if (a.q_type == CCB_Q_NONE && a.noise_type == CCB_NOISE_NONE && q->training) return CCB_ERR_INVALID;
  k_quant<false><<<grid_for(p), 256, 0, s>>>(a, yhat, nullptr, nullptr);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}

int ccb_quant_backward(const ccb_plan* p, const ccb_tensors* prm, const ccb_quant_cfg* q, const float* dyhat,
                       const float* dyhat2, const ccb_tensors* grads, cudaStream_t s) {
  QuantArgs a = make_args(p, prm, q, grads);
  bool any = false;
  for (int i = 0; i < p->g.L; ++i) any |= (a.dlat.p[i] != nullptr);
  if (!any) return CCB_OK;
  k_quant<true><<<grid_for(p), 256, 0, s>>>(a, nullptr, dyhat, dyhat2);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
