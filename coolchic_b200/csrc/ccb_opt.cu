// LO + OPT: loss (frame clamp + MSE + lambda * bpp), deterministic partial-sum reductions,
// global-norm gradient clipping and multi-tensor Adam.
// Restates enc/component/frame.py:183 (clamp), enc/training/loss.py:135-150 (loss_function),
// enc/training/train.py:265-266 (clip_grad_norm_(., 0.1, 2) ; torch.optim.Adam.step(), defaults).
#include "ccb_internal.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// Column sums of a [n_rows][n_cols] partial area, scattered into gradient tensors.
// ------------------------------------------------------------------------------------------
constexpr int MAX_SEG = 2 * CCB_MAX_ARM_LAYERS + 4;
struct Segments {
  float* dst[MAX_SEG];
  int begin[MAX_SEG];  // first column
  int count[MAX_SEG];
  int n;
};

// One launch serves up to MAX_JOBS partial areas (jobs); CTA = 32 columns of one job, 32 row groups.
constexpr int MAX_JOBS = CCB_MAX_SYN_LAYERS + CCB_MAX_LATENTS;  // synthesis layers + upsampling stages
struct ColsumJobs {
  const float* part[MAX_JOBS];
  int n_rows[MAX_JOBS], n_cols[MAX_JOBS], first_block[MAX_JOBS + 1];
  Segments seg[MAX_JOBS];
  int n;
  float* wsq;  // optional: per-CTA sum of squares of the columns this CTA produced
  int block0;  // first block of this launch (jobs may be spread over two launches; blocks keep their global numbers)
  // per-sample gradients (gridDim.y = B): batch element b owns rows [b * n_rows, (b + 1) * n_rows) of every job and
  // its sums land dst_bstride floats after element b - 1's
  long long dst_bstride;
};

__global__ void __launch_bounds__(1024) k_colsum(ColsumJobs jobs) {
  __shared__ float sm[32][33];
  const int bx = (int)blockIdx.x + jobs.block0;
  int j = 0;
  for (int k = 1; k < jobs.n; ++k)
    if (bx >= jobs.first_block[k]) j = k;
  const int n_rows = jobs.n_rows[j], n_cols = jobs.n_cols[j];
  const float* __restrict__ part = jobs.part[j] + (long long)blockIdx.y * n_rows * n_cols;
  const Segments& seg = jobs.seg[j];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col = (bx - jobs.first_block[j]) * 32 + tx;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (col < n_cols) {
    int r = ty;
    for (; r + 96 < n_rows; r += 128) {  // four independent loads in flight (the kernel is pure L2 latency)
      s0 += part[(long long)r * n_cols + col];
      s1 += part[(long long)(r + 32) * n_cols + col];
      s2 += part[(long long)(r + 64) * n_cols + col];
      s3 += part[(long long)(r + 96) * n_cols + col];
    }
    for (; r < n_rows; r += 32) s0 += part[(long long)r * n_cols + col];
  }
  sm[ty][tx] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (ty == 0) {
    float v = 0.f, sq = 0.f;
    if (col < n_cols) {
#pragma unroll
      for (int k = 0; k < 32; ++k) v += sm[k][tx];
      for (int i = 0; i < seg.n; ++i)
        if (col >= seg.begin[i] && col < seg.begin[i] + seg.count[i]) {
          seg.dst[i][(long long)blockIdx.y * jobs.dst_bstride + col - seg.begin[i]] = v;
          sq = v * v;
        }
    }
    sq = warp_sum(sq);
    if (jobs.wsq && tx == 0) jobs.wsq[bx] = sq;
  }
}

static void add_job(ColsumJobs& jobs, const float* part, int n_rows, int n_cols, const Segments& seg) {
  const int j = jobs.n++;
  jobs.part[j] = part;
  jobs.n_rows[j] = n_rows;
  jobs.n_cols[j] = n_cols;
  jobs.seg[j] = seg;
  jobs.first_block[j + 1] = jobs.first_block[j] + (n_cols + 31) / 32;
}

// ------------------------------------------------------------------------------------------
// Loss
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_loss_partial(const float* __restrict__ raw, const float* __restrict__ tgt,
                                                      long long n_img, const float* __restrict__ rate,
                                                      long long n_rate, float inv_n_img, float* __restrict__ d_raw,
                                                      float* __restrict__ part /* [grid][2] */,
                                                      unsigned* tail_counter) {
  __shared__ float red[32];
  if (tail_counter && blockIdx.x == 0 && threadIdx.x == 0) *tail_counter = 0u;
  float se = 0.f, rs = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long gtid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  auto one = [&](float x, float t, float& d) {
    const float c = fminf(fmaxf(x, 0.f), 1.f);
    const float e = c - t;
    se = fmaf(e, e, se);
    d = (x >= 0.f && x <= 1.f) ? 2.0f * e * inv_n_img : 0.f;
  };
  const bool vec = ((reinterpret_cast<uintptr_t>(raw) | reinterpret_cast<uintptr_t>(tgt) |
                     reinterpret_cast<uintptr_t>(d_raw)) & 15) == 0;
  const long long n4 = vec ? (n_img >> 2) : 0;
  for (long long i = gtid; i < n4; i += stride) {
    const float4 x = reinterpret_cast<const float4*>(raw)[i];
    const float4 t = reinterpret_cast<const float4*>(tgt)[i];
    float4 d;
    one(x.x, t.x, d.x);
    one(x.y, t.y, d.y);
    one(x.z, t.z, d.z);
    one(x.w, t.w, d.w);
    if (d_raw) reinterpret_cast<float4*>(d_raw)[i] = d;
  }
  for (long long i = 4 * n4 + gtid; i < n_img; i += stride) {
    float d;
    one(raw[i], tgt[i], d);
    if (d_raw) d_raw[i] = d;
  }
  if (rate)
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_rate; i += stride) rs += rate[i];
  float a = block_sum(se, red);
  float b = block_sum(rs, red);
  if (threadIdx.x == 0) {
    part[2 * blockIdx.x] = a;
    part[2 * blockIdx.x + 1] = b;
  }
}

__global__ void __launch_bounds__(256) k_loss_final(const float* part, int n_blocks, float inv_n_img,
                                                    float inv_b_npix, float lambda, float* out4,
                                                    const float* rate_part, int n_rate_part) {
  __shared__ float red[32];
  float se = 0.f, rs = 0.f;
  for (int i = threadIdx.x; i < n_blocks; i += blockDim.x) {
    se += part[2 * i];
    rs += part[2 * i + 1];
  }
  if (rate_part)  // per-tile rate sums of the ARM forward replace the rate column
    for (int i = threadIdx.x; i < n_rate_part; i += blockDim.x) rs += rate_part[i];
  float a = block_sum(se, red);
  float b = block_sum(rs, red);
  if (threadIdx.x == 0) {
    const float mse = a * inv_n_img;
    const float bpp = b * inv_b_npix;
    out4[0] = mse + lambda * bpp;
    out4[1] = mse;
    out4[2] = bpp;
  }
}

// ------------------------------------------------------------------------------------------
// clip_grad_norm_ + Adam over a table of tensors
// ------------------------------------------------------------------------------------------
constexpr int MAX_T = CCB_MAX_LATENTS * 3 + CCB_MAX_ARM_LAYERS * 2 + CCB_MAX_SYN_LAYERS * 2;  // 50
struct TensorTable {
  float* p[MAX_T];
  const float* g[MAX_T];
  float* m[MAX_T];
  float* v[MAX_T];
  long long n[MAX_T];
  long long chunk_begin[MAX_T + 1];  // prefix sums of per-tensor chunk counts
  unsigned char update[MAX_T];
  int nt;
};
constexpr int CHUNK = 2048;

__device__ __forceinline__ int find_tensor(const TensorTable& t, long long chunk) {
  int lo = 0;
  for (int i = 1; i < t.nt; ++i)
    if (chunk >= t.chunk_begin[i]) lo = i;
  return lo;
}

__global__ void __launch_bounds__(256) k_sumsq(TensorTable t, float* __restrict__ part) {
  __shared__ float red[32];
  const long long n_chunks = t.chunk_begin[t.nt];
  float s = 0.f;
  for (long long ch = blockIdx.x; ch < n_chunks; ch += gridDim.x) {
    const int ti = find_tensor(t, ch);
    const long long start = (ch - t.chunk_begin[ti]) * CHUNK;
    const long long end = min(start + CHUNK, t.n[ti]);
    const float* g = t.g[ti];
    for (long long i = start + threadIdx.x; i < end; i += blockDim.x) {
      const float x = g[i];
      s = fmaf(x, x, s);
    }
  }
  float v = block_sum(s, red);
  if (threadIdx.x == 0) part[blockIdx.x] = v;
}

// scal[0] = total norm, scal[1] = clip coefficient; advances *step.
__global__ void __launch_bounds__(256) k_norm_final(const float* part, int n_blocks, float max_norm,
                                                    float* scal, int* step, float* norm_out, double beta1,
                                                    double beta2) {
  __shared__ float red[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < n_blocks; i += blockDim.x) s += part[i];
  float v = block_sum(s, red);
  if (threadIdx.x == 0) {
    const float norm = sqrtf(v);
    scal[0] = norm;
    scal[1] = fminf(max_norm / (norm + 1e-6f), 1.0f);  // clip_grad_norm_: clamp(max_norm/(norm+1e-6), max=1)
    if (norm_out) *norm_out = norm;
    const int st = *step + 1;
    *step = st;
    // torch.optim.Adam computes the bias corrections in Python floats (double)
    const double bc1 = 1.0 - pow(beta1, (double)st);
    const double bc2 = 1.0 - pow(beta2, (double)st);
    scal[2] = (float)(1.0 / bc1);
    scal[3] = (float)(1.0 / sqrt(bc2));
  }
}

// Sum of squares over the whole tensor table + the step's finals (CcbTail), for the configurations whose
// quantiser backward does not run (no latent gradient): otherwise k_quant<true> carries the finals.
__global__ void __launch_bounds__(256) k_sumsq_tail(TensorTable t, CcbTail a) {
  __shared__ float red[96];
  if (blockIdx.x == 0 && threadIdx.x == 32) ccb_tail_bias(a);
  const long long n_chunks = t.chunk_begin[t.nt];
  float s = 0.f;
  for (long long ch = blockIdx.x; ch < n_chunks; ch += gridDim.x) {
    const int ti = find_tensor(t, ch);
    const long long start = (ch - t.chunk_begin[ti]) * CHUNK;
    const long long end = min(start + CHUNK, t.n[ti]);
    const float* g = t.g[ti];
    for (long long i = start + threadIdx.x; i < end; i += blockDim.x) {
      const float x = g[i];
      s = fmaf(x, x, s);
    }
  }
  ccb_tail_finals(a, s, red);
}

__global__ void __launch_bounds__(256) k_adam(TensorTable t, const float* __restrict__ scal,
                                              const float* __restrict__ lr_p, float beta2, float omb1, float omb2,
                                              float eps) {
  const long long n_chunks = t.chunk_begin[t.nt];
  const float coef = scal[1];
  const float lr = *lr_p;
  // torch.optim.Adam (single-tensor path): step_size = lr / bc1 ; denom = sqrt(v)/sqrt(bc2) + eps
  const float step_size = lr * scal[2];
  const float inv_sqrt_bc2 = scal[3];
  for (long long ch = blockIdx.x; ch < n_chunks; ch += gridDim.x) {
    const int ti = find_tensor(t, ch);
    if (!t.update[ti]) continue;
    const long long start = (ch - t.chunk_begin[ti]) * CHUNK;
    const long long end = min(start + CHUNK, t.n[ti]);
    float* p = t.p[ti];
    const float* g = t.g[ti];
    float* m = t.m[ti];
    float* v = t.v[ti];
    for (long long i = start + threadIdx.x; i < end; i += blockDim.x) {
      const float gi = g[i] * coef;
      const float mi = m[i] + (gi - m[i]) * omb1;  // lerp_(g, 1-beta1)
      const float vi = fmaf(beta2, v[i], omb2 * gi * gi);
      m[i] = mi;
      v[i] = vi;
      const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
      p[i] = p[i] - step_size * (mi / denom);
    }
  }
}

// Same update over ONE contiguous range: when the caller keeps parameters, gradients and both moments in flat
// arenas (FusedState, ccb_overfit_host) the table walk (a linear tensor search per 2048-element chunk) is pure overhead.
__global__ void __launch_bounds__(256) k_adam_flat(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, long long n,
                                                   const float* __restrict__ scal, const float* __restrict__ lr_p,
                                                   float beta2, float omb1, float omb2, float eps, const int fold,
                                                   const CcbTail tail, const int n_sq) {
  __shared__ float red[96];
  float coef;
  if (fold) {
    // every CTA folds the step's partial arrays itself (a few thousand L2-resident floats, identical result in
    // every CTA): no serial "last CTA" phase between the quantiser backward and the update
    float sq, se, rs;
    ccb_tail_reduce(tail, n_sq, red, sq, se, rs);
    coef = fminf(tail.max_norm / (sqrtf(sq) + 1e-6f), 1.0f);
    if (blockIdx.x == 0 && threadIdx.x == 0) ccb_tail_write(tail, sq, se, rs);
  } else {
    coef = scal[1];
  }
  const float step_size = *lr_p * scal[2];
  const float inv_sqrt_bc2 = scal[3];
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  auto update = [&](float gi, float& mr, float& vr, float& pr) {
    gi *= coef;
    const float mi = mr + (gi - mr) * omb1;  // lerp_(g, 1-beta1)
    const float vi = fmaf(beta2, vr, omb2 * gi * gi);
    mr = mi;
    vr = vi;
    const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    pr = pr - step_size * (mi / denom);
  };
  // 128-bit accesses when the four arenas allow it (they do in FusedState / ccb_overfit_host): four independent loads
  // per thread in flight and, at the benchmark size, a single trip per thread instead of 3-4 dependent ones
  long long done = 0;
  if (((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
        reinterpret_cast<uintptr_t>(v)) & 15) == 0) {
    const long long n4 = n >> 2;
    float4* p4 = reinterpret_cast<float4*>(p);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    for (long long i = tid; i < n4; i += stride) {
      const float4 gq = g4[i];
      float4 mq = m4[i], vq = v4[i], pq = p4[i];
      update(gq.x, mq.x, vq.x, pq.x);
      update(gq.y, mq.y, vq.y, pq.y);
      update(gq.z, mq.z, vq.z, pq.z);
      update(gq.w, mq.w, vq.w, pq.w);
      m4[i] = mq;
      v4[i] = vq;
      p4[i] = pq;
    }
    done = n4 << 2;
  }
  for (long long i = done + tid; i < n; i += stride) {
    float mr = m[i], vr = v[i], pr = p[i];
    update(g[i], mr, vr, pr);
    m[i] = mr;
    v[i] = vr;
    p[i] = pr;
  }
}

void add_tensor(TensorTable& t, float* p, const float* g, float* m, float* v, long long n, bool upd) {
  if (g == nullptr || n == 0) return;
  int i = t.nt++;
  t.p[i] = p;
  t.g[i] = g;
  t.m[i] = m;
  t.v[i] = v;
  t.n[i] = n;
  t.update[i] = upd && p && m && v;
  t.chunk_begin[i + 1] = t.chunk_begin[i] + (n + CHUNK - 1) / CHUNK;
}

TensorTable build_table(const ccb_plan* p, const ccb_tensors* prm, const ccb_tensors* g, const ccb_tensors* m,
                        const ccb_tensors* v, int mask) {
  TensorTable t = {};
  t.chunk_begin[0] = 0;
  const WeightLayout& wl = p->wl;
  for (int i = 0; i < p->g.L; ++i)
    add_tensor(t, prm->latent[i], g->latent[i], m ? m->latent[i] : nullptr, v ? v->latent[i] : nullptr,
               (long long)p->g.B * p->g.h[i] * p->g.w[i], mask & CCB_OPT_LATENT);
  for (int k = 0; k < wl.n_arm_layers; ++k) {
    add_tensor(t, prm->arm_w[k], g->arm_w[k], m ? m->arm_w[k] : nullptr, v ? v->arm_w[k] : nullptr, wl.arm_w_n[k],
               mask & CCB_OPT_ARM);
    add_tensor(t, prm->arm_b[k], g->arm_b[k], m ? m->arm_b[k] : nullptr, v ? v->arm_b[k] : nullptr, wl.arm_b_n[k],
               mask & CCB_OPT_ARM);
  }
  // same order as the flat parameter vector (ccb_plan_n_param_floats): all upsampling filters, then all refine filters
  for (int i = 0; i < p->g.L - 1; ++i)
    add_tensor(t, prm->ups_w[i], g->ups_w[i], m ? m->ups_w[i] : nullptr, v ? v->ups_w[i] : nullptr, wl.ups_n,
               mask & CCB_OPT_UPS);
  for (int i = 0; i < p->g.L - 1; ++i)
    add_tensor(t, prm->pre_w[i], g->pre_w[i], m ? m->pre_w[i] : nullptr, v ? v->pre_w[i] : nullptr, wl.pre_n,
               mask & CCB_OPT_UPS);
  for (int l = 0; l < p->arch.n_syn_layers; ++l) {
    add_tensor(t, prm->syn_w[l], g->syn_w[l], m ? m->syn_w[l] : nullptr, v ? v->syn_w[l] : nullptr, wl.syn_w_n[l],
               mask & CCB_OPT_SYN);
    add_tensor(t, prm->syn_b[l], g->syn_b[l], m ? m->syn_b[l] : nullptr, v ? v->syn_b[l] : nullptr, wl.syn_b_n[l],
               mask & CCB_OPT_SYN);
  }
  return t;
}

}  // namespace

int ccb_syn_head_blocks(const ccb_plan* p);
int ccb_syn_head_params(const ccb_plan* p);
int ccb_syn_conv_blocks(const ccb_plan* p);

// ARM + synthesis partial rows -> gradient tensors (upsampling filters reduce in ccb_ups.cu).
static float* wsq_base(const ccb_plan* p, char* ws) {
  return reinterpret_cast<float*>(ws + p->off_scalars) + 64 + 2 * p->n_loss_blocks + 8 * p->n_sm;
}

int ccb_wsq_layout(const ccb_plan* p, int* arm_base, int* syn_base, int* ups_base) {
  const int n_arm = ((int)p->wl.n_arm + 31) / 32;
  int n_syn = (ccb_syn_head_params(p) + 31) / 32;
  for (int l = 2; l < p->arch.n_syn_layers; ++l) n_syn += 3;  // 84 columns
  *arm_base = 0;
  *syn_base = n_arm;
  *ups_base = n_arm + n_syn;
  return n_arm + n_syn + (p->g.L - 1);  // one column-sum CTA (16 columns) per upsampling stage
}

int ccb_reduce_arm_grads(const ccb_plan* p, char* ws, const ccb_tensors* grads, cudaStream_t s) {
  const WeightLayout& wl = p->wl;
  Segments seg = {};
  for (int k = 0; k < wl.n_arm_layers; ++k) {
    seg.dst[seg.n] = grads->arm_w[k];
    seg.begin[seg.n] = (int)(wl.arm_w_off[k] - wl.arm_base);
    seg.count[seg.n++] = wl.arm_w_n[k];
    seg.dst[seg.n] = grads->arm_b[k];
    seg.begin[seg.n] = (int)(wl.arm_b_off[k] - wl.arm_base);
    seg.count[seg.n++] = wl.arm_b_n[k];
  }
  ColsumJobs jobs = {};
  const bool per_sample = grads->weight_batch_stride != 0;
  int gx, gy;
  ccb_arm_bwd_grid(p, per_sample, &gx, &gy);
  add_job(jobs, reinterpret_cast<const float*>(ws + p->off_arm_part), gx, (int)wl.n_arm, seg);
  jobs.dst_bstride = grads->weight_batch_stride;
  if (!per_sample) {
    int ab, sb, ub;
    ccb_wsq_layout(p, &ab, &sb, &ub);
    jobs.wsq = wsq_base(p, ws) + ab;
  }
  k_colsum<<<dim3(jobs.first_block[jobs.n], gy), 1024, 0, s>>>(jobs);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}

// which (bit mask, contiguous ranges only): 1 = the synthesis partial rows (complete after the synthesis backward),
// 2 = the rows of the pyramid's fused small stages (complete after the whole pyramid backward), 4 = the rows of its
// separately launched large stages (complete before the fused launch starts); 7 = everything in one launch.
// Parts may run as separate launches: blocks keep their global numbers (and their sum-of-squares slots).
int ccb_reduce_syn_grads(const ccb_plan* p, char* ws, const ccb_tensors* grads, cudaStream_t s, const UpsPartInfo* ups,
                         int which) {
  ColsumJobs jobs = {};
  // per-sample gradients: every partial area is ordered by batch element (batch-pure CTAs, b-major), so element b
  // owns the b-th 1/B of its rows
  const bool per_sample = grads->weight_batch_stride != 0;
  const int nb = per_sample ? p->arch.B : 1;
  jobs.dst_bstride = grads->weight_batch_stride;
  {
    const int S = p->syn_hidden;
    Segments seg = {};
    seg.n = 4;
    seg.dst[0] = grads->syn_w[0]; seg.begin[0] = 0; seg.count[0] = S * 7;
    seg.dst[1] = grads->syn_b[0]; seg.begin[1] = S * 7; seg.count[1] = S;
    seg.dst[2] = grads->syn_w[1]; seg.begin[2] = S * 8; seg.count[2] = 3 * S;
    seg.dst[3] = grads->syn_b[1]; seg.begin[3] = S * 11; seg.count[3] = 3;
    add_job(jobs, reinterpret_cast<const float*>(ws + p->off_syn_part[0]), ccb_syn_head_blocks(p) / nb,
            ccb_syn_head_params(p), seg);
  }
  for (int l = 2; l < p->arch.n_syn_layers; ++l) {
    Segments seg = {};
    seg.n = 2;
    seg.dst[0] = grads->syn_w[l]; seg.begin[0] = 0; seg.count[0] = 81;
    seg.dst[1] = grads->syn_b[l]; seg.begin[1] = 81; seg.count[1] = 3;
    add_job(jobs, reinterpret_cast<const float*>(ws + p->off_syn_part[l]), ccb_syn_conv_blocks(p) / nb, 84, seg);
  }
  if (ups) {
    for (int i = 0; i < ups->n; ++i) {
      Segments seg = {};
      seg.n = 2;
      seg.dst[0] = grads->ups_w[i]; seg.begin[0] = 0; seg.count[0] = p->arch.ups_k / 2;
      seg.dst[1] = grads->pre_w[i]; seg.begin[1] = 4; seg.count[1] = (p->arch.ups_pre_k + 1) / 2;
      add_job(jobs, ups->part[i], ups->n_rows[i] / nb, 16, seg);
    }
  }
  if (!per_sample) {
    int ab, sb, ub;
    ccb_wsq_layout(p, &ab, &sb, &ub);
    jobs.wsq = wsq_base(p, ws) + sb;
  }
  const int n_syn_jobs = 1 + (p->arch.n_syn_layers - 2);
  const int j1 = n_syn_jobs < jobs.n ? n_syn_jobs : jobs.n;                   // first pyramid job (stage 0 = coarsest)
  const int j2 = (ups && j1 + ups->n_fused < jobs.n) ? j1 + ups->n_fused : jobs.n;  // first separately launched stage
  const int lo = (which & 1) ? 0 : ((which & 2) ? j1 : j2);
  const int hi = (which & 4) ? jobs.n : ((which & 2) ? j2 : j1);
  if ((which & 5) == 5 && !(which & 2)) return CCB_ERR_INVALID;  // not contiguous
  const int b0 = jobs.first_block[lo], b1 = jobs.first_block[hi];
  if (b1 <= b0) return CCB_OK;
  jobs.block0 = b0;
  k_colsum<<<dim3(b1 - b0, nb), 1024, 0, s>>>(jobs);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}

extern "C" int ccb_loss(const ccb_plan* p, void* workspace, const float* raw_out, const float* rate,
                        const float* target, float lambda, float* d_raw_out, float* out4, void* stream) {
  if (!p || !workspace || !raw_out || !target || !out4) return CCB_ERR_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  float* part = reinterpret_cast<float*>(ws + p->off_scalars) + 64;
  const long long n_img = (long long)p->arch.B * 3 * p->arch.H * p->arch.W;
  const long long n_rate = (long long)p->arch.B * p->g.M;
  const float inv_n_img = 1.0f / (float)n_img;
  const float inv_b_npix = 1.0f / ((float)p->arch.B * (float)p->arch.H * (float)p->arch.W);
  k_loss_partial<<<p->n_loss_blocks, 256, 0, s>>>(raw_out, target, n_img, rate, n_rate, inv_n_img, d_raw_out, part,
                                                  nullptr);
  CCB_LAUNCH_CHECK();
  k_loss_final<<<1, 256, 0, s>>>(part, p->n_loss_blocks, inv_n_img, inv_b_npix, lambda, out4, nullptr, 0);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}

void ccb_loss_fuse_args(const ccb_plan* p, char* ws, const float* target, float* d_raw_out, SynLossFuse* out) {
  const long long n_img = (long long)p->arch.B * 3 * p->arch.H * p->arch.W;
  out->target = target;
  out->d_raw = d_raw_out;
  out->part = reinterpret_cast<float*>(ws + p->off_scalars) + 64;
  out->n_part = p->n_loss_blocks;
  out->inv_n_img = 1.0f / (float)n_img;
  out->tail_counter = reinterpret_cast<unsigned*>(ws + p->off_scalars) + 8;
  out->done = 0;
}

int ccb_loss_image(const ccb_plan* p, char* ws, const float* raw_out, const float* target, float* d_raw_out,
                   cudaStream_t s) {
  float* part = reinterpret_cast<float*>(ws + p->off_scalars) + 64;
  const long long n_img = (long long)p->arch.B * 3 * p->arch.H * p->arch.W;
  k_loss_partial<<<p->n_loss_blocks, 256, 0, s>>>(raw_out, target, n_img, nullptr, 0, 1.0f / (float)n_img, d_raw_out,
                                                  part, reinterpret_cast<unsigned*>(ws + p->off_scalars) + 8);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}

int ccb_loss_finish(const ccb_plan* p, char* ws, float lambda, float* out4, cudaStream_t s) {
  float* part = reinterpret_cast<float*>(ws + p->off_scalars) + 64;
  const long long n_img = (long long)p->arch.B * 3 * p->arch.H * p->arch.W;
  const float inv_b_npix = 1.0f / ((float)p->arch.B * (float)p->arch.H * (float)p->arch.W);
  k_loss_final<<<1, 256, 0, s>>>(part, p->n_loss_blocks, 1.0f / (float)n_img, inv_b_npix, lambda, out4,
                                 reinterpret_cast<const float*>(ws + p->off_rate_part), p->n_arm_tiles);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}

int ccb_tail_prepare(const ccb_plan* p, char* ws, const ccb_tensors* grads, float max_norm, double beta1, double beta2,
                     int32_t* step_dev, float lambda, float* out4, uint64_t* philox_offset_dev, CcbTail* tail) {
  float* scal = reinterpret_cast<float*>(ws + p->off_scalars);
  const int nb = p->n_loss_blocks;
  const long long n_img = (long long)p->arch.B * 3 * p->arch.H * p->arch.W;
  CcbTail& a = *tail;
  a = CcbTail{};
  a.sq_part = scal + 64 + 2 * nb;
  a.counter = reinterpret_cast<unsigned*>(scal) + 8;
  a.loss_part = scal + 64;
  a.n_loss_blocks = nb;
  a.rate_part = reinterpret_cast<const float*>(ws + p->off_rate_part);
  a.n_rate_part = p->n_arm_tiles;  // emitted by k_arm_bwd, one per backward tile
  a.inv_n_img = 1.0f / (float)n_img;
  a.inv_b_npix = 1.0f / ((float)p->arch.B * (float)p->arch.H * (float)p->arch.W);
  a.lambda = lambda;
  a.max_norm = max_norm;
  a.out4 = out4;
  a.scal = scal;
  a.step = step_dev;
  a.beta1 = beta1;
  a.beta2 = beta2;
  a.philox_offset = reinterpret_cast<unsigned long long*>(philox_offset_dev);
  int ab, sb, ub;
  a.n_wsq = ccb_wsq_layout(p, &ab, &sb, &ub);
  a.wsq_part = a.sq_part + 8 * p->n_sm;
  (void)grads;
  return CCB_OK;
}

static bool table_is_flat(const TensorTable& t, long long* total_out) {
  bool flat = t.nt > 0;
  long long total = t.nt > 0 ? t.n[0] : 0;
  for (int i = 0; i < t.nt && flat; ++i) {
    flat = t.update[i];
    if (i > 0) {
      flat = flat && t.p[i] == t.p[i - 1] + t.n[i - 1] && t.g[i] == t.g[i - 1] + t.n[i - 1] &&
             t.m[i] == t.m[i - 1] + t.n[i - 1] && t.v[i] == t.v[i - 1] + t.n[i - 1];
      total += t.n[i];
    }
  }
  *total_out = total;
  return flat;
}

bool ccb_adam_flat_ok(const ccb_plan* p, const ccb_tensors* params, const ccb_tensors* grads,
                      const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, int optimize_mask) {
  if (!optimize_mask) return false;
  TensorTable t = build_table(p, params, grads, exp_avg, exp_avg_sq, optimize_mask);
  long long total;
  return table_is_flat(t, &total);
}

int ccb_adam_only(const ccb_plan* p, char* ws, const ccb_tensors* params, const ccb_tensors* grads,
                  const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, int optimize_mask, const float* lr_dev,
                  double beta1, double beta2, double eps, cudaStream_t s, const CcbTail* tail, int n_sq_parts) {
  if (!optimize_mask) return tail ? CCB_ERR_INVALID : CCB_OK;
  float* scal = reinterpret_cast<float*>(ws + p->off_scalars);
  TensorTable t = build_table(p, params, grads, exp_avg, exp_avg_sq, optimize_mask);
  if (t.nt == 0) return CCB_ERR_INVALID;
  long long total = 0;
  const bool flat = table_is_flat(t, &total);
  if (tail && !flat) return CCB_ERR_INVALID;
  if (flat) {
    const long long nb = (total + 255) / 256;
    const int grid = (int)(nb < (long long)p->n_sm * 4 ? nb : (long long)p->n_sm * 4);
    k_adam_flat<<<grid, 256, 0, s>>>(t.p[0], t.g[0], t.m[0], t.v[0], total, scal, lr_dev, (float)beta2,
                                     (float)(1.0 - beta1), (float)(1.0 - beta2), (float)eps, tail ? 1 : 0,
                                     tail ? *tail : CcbTail{}, n_sq_parts);
  } else {
    k_adam<<<p->n_loss_blocks, 256, 0, s>>>(t, scal, lr_dev, (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2),
                                            (float)eps);
  }
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}

// loss finish + clip + Adam + Philox advance when the quantiser backward does not run: 2 launches.
int ccb_tail_step(const ccb_plan* p, char* ws, const ccb_tensors* params, const ccb_tensors* grads,
                  const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, int optimize_mask, float max_norm,
                  const float* lr_dev, double beta1, double beta2, double eps, int32_t* step_dev, float lambda,
                  float* out4, uint64_t* philox_offset_dev, cudaStream_t s) {
  TensorTable t = build_table(p, params, grads, exp_avg, exp_avg_sq, optimize_mask);
  if (t.nt == 0) return CCB_ERR_INVALID;
  CcbTail a;
  ccb_tail_prepare(p, ws, grads, max_norm, beta1, beta2, step_dev, lambda, out4, philox_offset_dev, &a);
  a.n_wsq = 0;  // the table walk below already covers the weights
  k_sumsq_tail<<<p->n_loss_blocks, 256, 0, s>>>(t, a);
  CCB_LAUNCH_CHECK();
  return ccb_adam_only(p, ws, params, grads, exp_avg, exp_avg_sq, optimize_mask, lr_dev, beta1, beta2, eps, s);
}

extern "C" int ccb_clip_adam_step(const ccb_plan* p, void* workspace, const ccb_tensors* params,
                                  const ccb_tensors* grads, const ccb_tensors* exp_avg,
                                  const ccb_tensors* exp_avg_sq, int optimize_mask, float max_norm,
                                  const float* lr_dev, double beta1, double beta2, double eps,
                                  int32_t* step_dev, float* norm_out, void* stream) {
  if (!p || !workspace || !params || !grads || !lr_dev || !step_dev) return CCB_ERR_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  float* scal = reinterpret_cast<float*>(ws + p->off_scalars);
  float* part = scal + 64 + 2 * p->n_loss_blocks;
  TensorTable t = build_table(p, params, grads, exp_avg, exp_avg_sq, optimize_mask);
  if (t.nt == 0) return CCB_ERR_INVALID;
  const int nb = p->n_loss_blocks;
  k_sumsq<<<nb, 256, 0, s>>>(t, part);
  CCB_LAUNCH_CHECK();
  k_norm_final<<<1, 256, 0, s>>>(part, nb, max_norm, scal, step_dev, norm_out, beta1, beta2);
  CCB_LAUNCH_CHECK();
  if (optimize_mask) {
    k_adam<<<nb, 256, 0, s>>>(t, scal, lr_dev, (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2), (float)eps);
    CCB_LAUNCH_CHECK();
  }
  return CCB_OK;
}
