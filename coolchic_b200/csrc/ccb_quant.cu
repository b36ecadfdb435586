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
  // training step with the latents of one image in one contiguous arena (same finest-first order as y_hat): the
  // forward reads lat_flat[idx] and parks d y_hat / d latent (quantiser derivative x gain) in dq[idx]; the backward is
  // then dlat_flat[idx] = (d y_hat) * dq[idx] -- no noise regeneration, no softround, no level lookup
  const float* lat_flat;
  float* dlat_flat;
  float* dq;
  // this launch covers the groups of four elements [q4_lo, q4_hi) (q4_hi == 0: everything); its sum-of-squares partials
  // start at slot slot0 (the training step may run the fine and the coarse levels as two launches)
  long long q4_lo, q4_hi;
  int slot0;
};

__device__ __forceinline__ void locate(const LevelGeom& g, long long r, int& lvl, long long& rel) {
  lvl = 0;
#pragma unroll
  for (int l = 1; l < CCB_MAX_LATENTS; ++l)
    if (l < g.L && r >= g.off[l]) lvl = l;
  rel = r - g.off[lvl];
}

// tanh through ex2.approx: |error| ~ 1e-7 absolute, which is the fp32 noise floor of the reference's own
// evaluation; the libm tanhf / powf pair made this kernel the most instruction-heavy elementwise pass.
__device__ __forceinline__ float tanh_fast(float x) {
  x = fminf(fmaxf(x, -15.f), 15.f);
  const float e = __expf(2.f * x);
  return 1.f - __fdividef(2.f, e + 1.f);
}

// s(x,T) and ds/dx, written in the reference's evaluation order
// (tanh(delta/t), division by tanh(1/(2t)); quantizer.py:39-41) -- SURVEY.md H2.
__device__ __forceinline__ float softround_val(float x, float inv_t, float inv_tanh_half_t, float& th) {
  const float fl = floorf(x);
  const float delta = x - fl - 0.5f;
  th = tanh_fast(delta * inv_t);
  return fl + 0.5f * th * inv_tanh_half_t + 0.5f;
}
__device__ __forceinline__ float softround_dx(float th, float inv_t, float inv_tanh_half_t) {
  return 0.5f * (1.0f - th * th) * inv_t * inv_tanh_half_t;
}

// Raw draw -> additive noise.  raw ~ N(0,1) (gaussian) or U(0,1) (kumaraswamy).
__device__ __forceinline__ float shape_noise(int noise_type, float raw, float sigma, float kum_inv_a,
                                             float kum_inv_b) {
  if (noise_type == CCB_NOISE_GAUSSIAN) return raw * sigma;
  if (noise_type == CCB_NOISE_KUMARASWAMY) {
    // (1 - (1-u)^(1/b))^(1/a) - 0.5   (quantizer.py:100-107)
    // lg2.approx may return a tiny positive value just below 1: clamp so that t never goes negative (NaN)
    const float t = fmaxf(1.0f - __powf(1.0f - raw, kum_inv_b), 0.f);
    return __powf(t, kum_inv_a) - 0.5f;
  }
  return 0.f;
}

// One thread = 4 consecutive elements of the flat [B*M] vector = one Philox4x32 block (counter = index / 4).
template <bool BACKWARD>
__global__ void __launch_bounds__(256) k_quant(QuantArgs a, float* __restrict__ yhat,
                                               const float* __restrict__ dyhat, const float* __restrict__ dyhat2,
                                               const float* __restrict__ dyhat3, const int has_tail,
                                               const CcbTail tail, unsigned* zero_ctr) {
  __shared__ float red[96];
  // per-step counters (grid barriers of the fused pyramid kernels): the quantiser forward opens every step
  if (!BACKWARD && zero_ctr && blockIdx.x == 0 && threadIdx.x < 2) zero_ctr[threadIdx.x] = 0u;
  float sq = 0.f;  // sum of squares of the latent gradients written by this thread
  if (BACKWARD && has_tail && a.slot0 == 0 && blockIdx.x == 0 && threadIdx.x == 32) ccb_tail_bias(tail);
  const long long total = (long long)a.g.B * a.g.M;
  const float T = a.T_dev ? *a.T_dev : a.T_host;
  const float sigma = a.sigma_dev ? *a.sigma_dev : a.sigma_host;
  const uint64_t offset = a.offset_dev ? *a.offset_dev : a.offset_host;
  const float inv_t = 1.0f / T;
  const float inv_th = 1.0f / tanhf(1.0f / (2.0f * T));
  float kum_inv_a = 0.f, kum_inv_b = 0.f;
  if (a.noise_type == CCB_NOISE_KUMARASWAMY) {
    float aa = sigma;
    float bb = (powf(2.0f, aa) * (aa - 1.0f) + 1.0f) / aa;
    kum_inv_a = 1.0f / aa;
    kum_inv_b = 1.0f / bb;
  }
  const float lo = -(float)a.ac_max_val, hi = (float)(a.ac_max_val + 1);
  const long long n4 = a.q4_hi > 0 ? a.q4_hi : (total + 3) >> 2;
  for (long long q4 = a.q4_lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; q4 < n4;
       q4 += (long long)gridDim.x * blockDim.x) {
    const long long idx0 = q4 << 2;
    if (BACKWARD && a.dq != nullptr) {
      if (idx0 + 3 < total) {
        float4 g4 = *reinterpret_cast<const float4*>(dyhat + idx0);
        if (dyhat2) {  // fixed order: chain + (ARM own-band accumulator + ARM band-below accumulator)
          float4 h4 = *reinterpret_cast<const float4*>(dyhat2 + idx0);
          if (dyhat3) {
            const float4 k4 = *reinterpret_cast<const float4*>(dyhat3 + idx0);
            h4.x += k4.x; h4.y += k4.y; h4.z += k4.z; h4.w += k4.w;
          }
          g4.x += h4.x; g4.y += h4.y; g4.z += h4.z; g4.w += h4.w;
        }
        const float4 d4 = *reinterpret_cast<const float4*>(a.dq + idx0);
        const float4 o4 = make_float4(g4.x * d4.x, g4.y * d4.y, g4.z * d4.z, g4.w * d4.w);
        *reinterpret_cast<float4*>(a.dlat_flat + idx0) = o4;
        sq = fmaf(o4.x, o4.x, sq);
        sq = fmaf(o4.y, o4.y, sq);
        sq = fmaf(o4.z, o4.z, sq);
        sq = fmaf(o4.w, o4.w, sq);
      } else {
        for (int k = 0; k < 4; ++k)
          if (idx0 + k < total) {
            float h = dyhat2 ? dyhat2[idx0 + k] : 0.f;
            if (dyhat2 && dyhat3) h += dyhat3[idx0 + k];
            const float gl = (dyhat2 ? dyhat[idx0 + k] + h : dyhat[idx0 + k]) * a.dq[idx0 + k];
            a.dlat_flat[idx0 + k] = gl;
            sq = fmaf(gl, gl, sq);
          }
      }
      continue;
    }
    float raw[4] = {0.f, 0.f, 0.f, 0.f};
    if (a.noise_type != CCB_NOISE_NONE) {
      if (a.noise) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (idx0 + k < total) raw[k] = a.noise[idx0 + k];
      } else {
        const uint4 r = philox4x32_10((uint64_t)q4, offset, a.seed);
        if (a.noise_type == CCB_NOISE_GAUSSIAN) {
          // Box-Muller on the pairs (0,1) and (2,3)
          float sn, cs;
          float rad = sqrtf(-2.0f * __logf(u32_to_unit_open(r.x)));
          __sincosf(6.283185307179586f * u32_to_unit_open(r.y), &sn, &cs);
          raw[0] = rad * cs;
          raw[1] = rad * sn;
          rad = sqrtf(-2.0f * __logf(u32_to_unit_open(r.z)));
          __sincosf(6.283185307179586f * u32_to_unit_open(r.w), &sn, &cs);
          raw[2] = rad * cs;
          raw[3] = rad * sn;
        } else {
          raw[0] = u32_to_unit_open(r.x);
          raw[1] = u32_to_unit_open(r.y);
          raw[2] = u32_to_unit_open(r.z);
          raw[3] = u32_to_unit_open(r.w);
        }
      }
    }
    float gy[4] = {0.f, 0.f, 0.f, 0.f};
    if (BACKWARD) {
      if (idx0 + 3 < total) {
        const float4 g4 = *reinterpret_cast<const float4*>(dyhat + idx0);
        gy[0] = g4.x; gy[1] = g4.y; gy[2] = g4.z; gy[3] = g4.w;
        // fixed order: chain + (ARM own-band accumulator + ARM band-below accumulator)
        if (dyhat2) {
          float4 h4 = *reinterpret_cast<const float4*>(dyhat2 + idx0);
          if (dyhat3) {
            const float4 k4 = *reinterpret_cast<const float4*>(dyhat3 + idx0);
            h4.x += k4.x; h4.y += k4.y; h4.z += k4.z; h4.w += k4.w;
          }
          gy[0] += h4.x; gy[1] += h4.y; gy[2] += h4.z; gy[3] += h4.w;
        }
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (idx0 + k < total) {
            float h = dyhat2 ? dyhat2[idx0 + k] : 0.f;
            if (dyhat2 && dyhat3) h += dyhat3[idx0 + k];
            gy[k] = dyhat2 ? dyhat[idx0 + k] + h : dyhat[idx0 + k];
          }
      }
    }
    float out[4], dqv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const long long idx = idx0 + k;
      out[k] = 0.f;
      if (idx >= total) continue;
      // 32-bit index arithmetic (the host refuses B*M >= 2^31): a 64-bit division costs ~100 instructions
      int lvl = 0;
      long long src = idx;
      float x;
      if (a.lat_flat != nullptr) {
        x = a.lat_flat[idx] * a.gain;
      } else {
        const unsigned u32 = (unsigned)idx, m32 = (unsigned)a.g.M;
        const int b = (a.g.B == 1) ? 0 : (int)(u32 / m32);
        const long long r = (long long)(u32 - (unsigned)b * m32);
        long long rel;
        locate(a.g, r, lvl, rel);
        const long long hw = (long long)a.g.h[lvl] * a.g.w[lvl];
        src = (long long)b * hw + rel;
        x = a.lat.p[lvl][src] * a.gain;
      }
      const float n = (a.noise_type != CCB_NOISE_NONE) ? shape_noise(a.noise_type, raw[k], sigma, kum_inv_a, kum_inv_b)
                                                      : 0.f;
      float y, d = 1.f, th;
      switch (a.q_type) {
        case CCB_Q_NONE: y = x + n; break;
        case CCB_Q_SOFTROUND_ALONE:
          y = softround_val(x, inv_t, inv_th, th);
          d = softround_dx(th, inv_t, inv_th);
          break;
        case CCB_Q_SOFTROUND: {
          const float s1 = softround_val(x, inv_t, inv_th, th) + n;
          d = softround_dx(th, inv_t, inv_th);
          y = softround_val(s1, inv_t, inv_th, th);
          d *= softround_dx(th, inv_t, inv_th);
        } break;
        case CCB_Q_STE: {
          // value (s - s) + round(x), evaluated as the reference does (quantizer.py:222-225); the
          // reference returns a detached tensor: no latent gradient
          const float sr = softround_val(x, inv_t, inv_th, th);
          y = (sr - sr) + rintf(x);
          d = 0.f;
        } break;
        case CCB_Q_TRUE_STE: y = rintf(x); break;  // identity gradient
        default:  // hardround: torch.round == round-half-even, zero gradient
          y = rintf(x);
          d = 0.f;
          break;
      }
      if (a.ac_max_val != -1) {
        // torch.clamp: value clamped, gradient passes where min <= y <= max
        if (y < lo || y > hi) d = 0.f;
        y = fminf(fmaxf(y, lo), hi);
      }
      if (!BACKWARD) {
        out[k] = y;
        dqv[k] = d * a.gain;
      } else {
        float* dl = a.dlat.p[lvl];
        if (dl != nullptr) {
          const float gl = gy[k] * d * a.gain;
          dl[src] = gl;
          sq = fmaf(gl, gl, sq);
        }
      }
    }
    if (!BACKWARD) {
      if (idx0 + 3 < total) {
        *reinterpret_cast<float4*>(yhat + idx0) = make_float4(out[0], out[1], out[2], out[3]);
        if (a.dq) *reinterpret_cast<float4*>(a.dq + idx0) = make_float4(dqv[0], dqv[1], dqv[2], dqv[3]);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (idx0 + k < total) {
            yhat[idx0 + k] = out[k];
            if (a.dq) a.dq[idx0 + k] = dqv[k];
          }
      }
    }
  }
  if (BACKWARD && has_tail == 1) ccb_tail_finals(tail, sq, red);
  if (BACKWARD && has_tail == 2) {  // partials only: the Adam kernel folds them in its prologue
    const float v = block_sum(sq, red);
    if (threadIdx.x == 0) tail.sq_part[a.slot0 + blockIdx.x] = v;
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
  a.lat_flat = nullptr;
  a.dlat_flat = nullptr;
  a.dq = nullptr;
  a.q4_lo = a.q4_hi = 0;
  a.slot0 = 0;
  return a;
}

// one image whose latent grids sit back to back, finest first: element idx of y_hat is element idx of the arena
bool latents_flat(const ccb_plan* p, const ccb_tensors* t) {
  if (p->g.B != 1 || !t) return false;
  for (int i = 0; i < p->g.L; ++i) {
    if (!t->latent[i]) return false;
    if (i + 1 < p->g.L && t->latent[i + 1] != t->latent[i] + (long long)p->g.h[i] * p->g.w[i]) return false;
  }
  return (reinterpret_cast<uintptr_t>(t->latent[0]) & 15) == 0;
}

int grid_for(const ccb_plan* p) {
  long long total = (long long)p->g.B * p->g.M;
  long long nb = ((total + 3) / 4 + 255) / 256;
  long long cap = (long long)p->n_sm * 8;
  return (int)(nb < cap ? (nb < 1 ? 1 : nb) : cap);
}

// a launch over the elements [e_lo, e_hi) (multiples of four): range fields and grid; `cap` bounds the grid
int range_grid(QuantArgs& a, long long e_lo, long long e_hi, int cap) {
  a.q4_lo = e_lo >> 2;
  a.q4_hi = (e_hi + 3) >> 2;
  long long nb = (a.q4_hi - a.q4_lo + 255) / 256;
  if (nb < 1) nb = 1;
  return (int)(nb < cap ? nb : cap);
}

}  // namespace

bool ccb_quant_flat_ok(const ccb_plan* p, const ccb_tensors* prm, const ccb_tensors* grads) {
  return latents_flat(p, prm) && latents_flat(p, grads);
}

// Element index where latent level `level` starts in the flat y_hat order, if a launch may be cut there (one image,
// a multiple of four, both sides non-empty); else -1.
long long ccb_quant_cut(const ccb_plan* p, int level) {
  if (p->g.B != 1 || level <= 0 || level >= p->g.L) return -1;
  const long long e = p->g.off[level];
  return (e > 0 && e < p->g.M && (e & 3) == 0) ? e : -1;
}

// e_lo / e_hi (elements, multiples of four; both 0: everything): the part of the flat order this launch covers
int ccb_quant_forward(const ccb_plan* p, const ccb_tensors* prm, const ccb_quant_cfg* q, float* yhat,
                      cudaStream_t s, unsigned* zero_ctr, float* dq, long long e_lo, long long e_hi) {
  QuantArgs a = make_args(p, prm, q, nullptr);
  if (latents_flat(p, prm)) a.lat_flat = prm->latent[0];
  a.dq = dq;
  if (e_hi > 0) {
    if ((long long)p->g.B * p->g.M >= (1LL << 31)) return CCB_ERR_UNSUPPORTED;
    const int grid = range_grid(a, e_lo, e_hi, p->n_sm * 8);
    k_quant<false><<<grid, 256, 0, s>>>(a, yhat, nullptr, nullptr, nullptr, 0, CcbTail{}, zero_ctr);
    CCB_LAUNCH_CHECK();
    return CCB_OK;
  }
  if ((long long)p->g.B * p->g.M >= (1LL << 31)) return CCB_ERR_UNSUPPORTED;
  if (a.q_type == CCB_Q_NONE && a.noise_type == CCB_NOISE_NONE && q->training) return CCB_ERR_INVALID;
  k_quant<false><<<grid_for(p), 256, 0, s>>>(a, yhat, nullptr, nullptr, nullptr, 0, CcbTail{}, zero_ctr);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}

int ccb_quant_grid(const ccb_plan* p) { return grid_for(p); }

int ccb_quant_backward(const ccb_plan* p, const ccb_tensors* prm, const ccb_quant_cfg* q, const float* dyhat,
                       const float* dyhat2, const float* dyhat3, const ccb_tensors* grads, cudaStream_t s,
                       CcbTail* tail, int tail_mode, const float* dq, long long e_lo, long long e_hi, int slot0,
                       int* n_slots) {
  QuantArgs a = make_args(p, prm, q, grads);
  if (dq) {
    if (!latents_flat(p, grads)) return CCB_ERR_INVALID;
    a.dq = const_cast<float*>(dq);
    a.dlat_flat = grads->latent[0];
  }
  if (e_hi > 0) {
    // one part of a split backward: parked derivative, partials-only tail, slots [slot0, slot0 + grid)
    if (!dq || !tail || tail_mode != 2 || !n_slots) return CCB_ERR_INVALID;
    a.slot0 = slot0;
    const int grid = range_grid(a, e_lo, e_hi, p->n_sm * 8 - slot0 - (e_lo == 0 ? 64 : 0));
    k_quant<true><<<grid, 256, 0, s>>>(a, nullptr, dyhat, dyhat2, dyhat3, tail_mode, *tail, nullptr);
    CCB_LAUNCH_CHECK();
    *n_slots = grid;
    return CCB_OK;
  }
  bool any = false;
  for (int i = 0; i < p->g.L; ++i) any |= (a.dlat.p[i] != nullptr);
  if (!any) return tail ? CCB_ERR_INVALID : CCB_OK;  // the finals need a carrier: the caller checks beforehand
  if (tail)
    k_quant<true><<<grid_for(p), 256, 0, s>>>(a, nullptr, dyhat, dyhat2, dyhat3, tail_mode, *tail, nullptr);
  else
    k_quant<true><<<grid_for(p), 256, 0, s>>>(a, nullptr, dyhat, dyhat2, dyhat3, 0, CcbTail{}, nullptr);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
