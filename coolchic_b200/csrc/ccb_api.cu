// C ABI: plans, workspace layout, forward / backward / train-step orchestration.
// The function order inside ccb_forward / ccb_backward is the reference's
// CoolChicEncoder.forward (enc/component/coolchic.py:360-427) and its autograd reverse.
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include <cstdlib>
#include "ccb_internal.cuh"

int ccb_syn_head_blocks(const ccb_plan* p);
int ccb_syn_head_params(const ccb_plan* p);
int ccb_syn_conv_blocks(const ccb_plan* p);

static thread_local char g_cuda_err[512] = "";

void ccb_set_cuda_error(cudaError_t e, const char* file, int line) {
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s (%s) at %s:%d", cudaGetErrorName(e), cudaGetErrorString(e), file, line);
}

extern "C" const char* ccb_last_cuda_error(void) { return g_cuda_err; }

extern "C" const char* ccb_strerror(int status) {
  switch (status) {
    case CCB_OK: return "ok";
    case CCB_ERR_INVALID: return "invalid argument";
    case CCB_ERR_UNSUPPORTED: return "architecture not supported by the sm_100a kernels";
    case CCB_ERR_CUDA: return "CUDA runtime error";
    case CCB_ERR_NOMEM: return "out of memory";
    case CCB_ERR_TOO_LARGE: return "problem size exceeds the static scratch of the allocation-free entry point";
    default: return "unknown status";
  }
}

extern "C" int ccb_version(void) { return 200; }

// ------------------------------------------------------------------------------------------
// Weight slots: who may write which constant bank (one bank set per slot and device, see ccb_arm.cu / ccb_syn.cu)
// ------------------------------------------------------------------------------------------
namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_slot_mutex;
bool g_slot_taken[MAX_DEVICES][CCB_N_WEIGHT_SLOTS];
}  // namespace

extern "C" int ccb_weight_slot_acquire(int device, int* slot) {
  if (!slot || device < 0 || device >= MAX_DEVICES) return CCB_ERR_INVALID;
  std::lock_guard<std::mutex> lock(g_slot_mutex);
  for (int k = 0; k < CCB_N_WEIGHT_SLOTS; ++k)
    if (!g_slot_taken[device][k]) {
      g_slot_taken[device][k] = true;
      *slot = k;
      return CCB_OK;
    }
  *slot = -1;
  return CCB_ERR_NOMEM;
}

extern "C" int ccb_stream_create(int device, void** stream_out) {
  if (!stream_out || device < 0) return CCB_ERR_INVALID;
  *stream_out = nullptr;
  int cur = -1;
  CCB_CHECK(cudaGetDevice(&cur));
  if (cur != device) CCB_CHECK(cudaSetDevice(device));
  cudaStream_t s = nullptr;
  const cudaError_t e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  if (cur != device) cudaSetDevice(cur);
  CCB_CHECK(e);
  *stream_out = s;
  return CCB_OK;
}

extern "C" int ccb_stream_destroy(int device, void* stream) {
  if (!stream || device < 0) return CCB_ERR_INVALID;
  int cur = -1;
  CCB_CHECK(cudaGetDevice(&cur));
  if (cur != device) CCB_CHECK(cudaSetDevice(device));
  const cudaError_t e = cudaStreamDestroy((cudaStream_t)stream);  // deferred by the runtime until its work is done
  if (cur != device) cudaSetDevice(cur);
  CCB_CHECK(e);
  return CCB_OK;
}

extern "C" int ccb_weight_slot_release(int device, int slot) {
  if (device < 0 || device >= MAX_DEVICES || slot < 0 || slot >= CCB_N_WEIGHT_SLOTS) return CCB_ERR_INVALID;
  std::lock_guard<std::mutex> lock(g_slot_mutex);
  if (!g_slot_taken[device][slot]) return CCB_ERR_INVALID;
  g_slot_taken[device][slot] = false;
  return CCB_OK;
}

static size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

extern "C" int ccb_plan_create(ccb_plan** out, const ccb_arch* arch, int device) {
  if (!out || !arch) return CCB_ERR_INVALID;
  *out = nullptr;
  const ccb_arch& A = *arch;
  if (A.B < 1 || A.H < 1 || A.W < 1) return CCB_ERR_INVALID;
  if (A.n_latents != 7) return CCB_ERR_UNSUPPORTED;  // synthesis head is specialised for 7 input features
  if (!(A.dim_arm == 8 || A.dim_arm == 16 || A.dim_arm == 24 || A.dim_arm == 32)) return CCB_ERR_UNSUPPORTED;
  if (A.n_hidden_arm < 0 || A.n_hidden_arm > 3) return CCB_ERR_UNSUPPORTED;
  if (A.ups_k < 2 || A.ups_k > 8 || (A.ups_k & 1)) return CCB_ERR_UNSUPPORTED;
  if (A.ups_pre_k < 1 || A.ups_pre_k > 7 || !(A.ups_pre_k & 1)) return CCB_ERR_UNSUPPORTED;
  if (A.n_syn_layers < 2 || A.n_syn_layers > CCB_MAX_SYN_LAYERS) return CCB_ERR_UNSUPPORTED;
  // [S-1-linear-relu, 3-1-linear-*, (3-3-{linear,residual}-*)*]
  if (A.syn[0].k != 1 || A.syn[0].residual || !A.syn[0].relu || A.syn[0].out_ch < 1 || A.syn[0].out_ch > 256)
    return CCB_ERR_UNSUPPORTED;
  if (A.syn[1].k != 1 || A.syn[1].residual || A.syn[1].out_ch != 3) return CCB_ERR_UNSUPPORTED;
  for (int l = 2; l < A.n_syn_layers; ++l)
    if (A.syn[l].k != 3 || A.syn[l].out_ch != 3) return CCB_ERR_UNSUPPORTED;

  ccb_plan* p = new (std::nothrow) ccb_plan();
  if (!p) return CCB_ERR_NOMEM;
  memset(p, 0, sizeof(*p));
  p->arch = A;
  p->device = device;
  int prev = 0;
  cudaError_t e = cudaGetDevice(&prev);
  struct RestoreDevice {  // every return path below leaves the caller's current device as it was
    int dev;
    bool armed;
    ~RestoreDevice() {
      if (armed) cudaSetDevice(dev);
    }
  } restore{prev, e == cudaSuccess};
  if (e == cudaSuccess) e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&p->n_sm, cudaDevAttrMultiProcessorCount, device);
  if (e != cudaSuccess) {
    ccb_set_cuda_error(e, __FILE__, __LINE__);
    delete p;
    return CCB_ERR_CUDA;
  }

  // ---- latent pyramid (coolchic.py:184-189: ceil(H / 2^i))
  LevelGeom& g = p->g;
  g.L = A.n_latents;
  g.B = A.B;
  long long off = 0;
  for (int i = 0; i < g.L; ++i) {
    g.h[i] = (A.H + (1 << i) - 1) >> i;
    g.w[i] = (A.W + (1 << i) - 1) >> i;
    g.off[i] = off;
    off += (long long)g.h[i] * g.w[i];
  }
  g.M = off;

  // ---- weight layout
  WeightLayout& wl = p->wl;
  const int C = A.dim_arm, NH = A.n_hidden_arm;
  wl.n_arm_layers = NH + 1;
  long long o = 0;
  wl.arm_base = 0;
  for (int k = 0; k <= NH; ++k) {
    wl.arm_w_off[k] = o;
    wl.arm_w_n[k] = (k < NH) ? C * C : 2 * C;
    o += wl.arm_w_n[k];
    wl.arm_b_off[k] = o;
    wl.arm_b_n[k] = (k < NH) ? C : 2;
    o += wl.arm_b_n[k];
  }
  wl.n_arm = o;
  wl.ups_base = o;
  wl.ups_n = A.ups_k / 2;
  wl.pre_n = (A.ups_pre_k + 1) / 2;
  for (int i = 0; i < g.L - 1; ++i) {
    wl.ups_off[i] = o;
    o += wl.ups_n;
  }
  for (int i = 0; i < g.L - 1; ++i) {
    wl.pre_off[i] = o;
    o += wl.pre_n;
  }
  wl.n_ups = o - wl.ups_base;
  wl.syn_base = o;
  int cin = g.L;
  for (int l = 0; l < A.n_syn_layers; ++l) {
    SynLayerGeom& sl = p->syn[l];
    sl.cin = cin;
    sl.cout = A.syn[l].out_ch;
    sl.k = A.syn[l].k;
    sl.residual = A.syn[l].residual;
    sl.relu = A.syn[l].relu;
    wl.syn_w_off[l] = o;
    wl.syn_w_n[l] = sl.cout * sl.cin * sl.k * sl.k;
    o += wl.syn_w_n[l];
    wl.syn_b_off[l] = o;
    wl.syn_b_n[l] = sl.cout;
    o += wl.syn_b_n[l];
    if (sl.residual && sl.cin != sl.cout) {
      delete p;
      return CCB_ERR_INVALID;
    }
    cin = sl.cout;
  }
  wl.n_syn = o - wl.syn_base;
  wl.n_total = o;
  p->syn_hidden = A.syn[0].out_ch;
  p->syn_head_out = A.syn[1].out_ch;
  p->n_tail = A.n_syn_layers - 2;

  // ---- ARM tiles (32 x 8 positions), all (batch, level) pairs in one list
  p->arm_tw = 32;
  p->arm_th = 8;
  std::vector<ArmTile> tiles;
  for (int b = 0; b < A.B; ++b)
    for (int l = 0; l < g.L; ++l)
      for (int y0 = 0; y0 < g.h[l]; y0 += p->arm_th)
        for (int x0 = 0; x0 < g.w[l]; x0 += p->arm_tw) tiles.push_back(ArmTile{b, l, y0, x0});
  p->n_arm_tiles = (int)tiles.size();
  e = cudaMalloc(&p->d_arm_tiles, tiles.size() * sizeof(ArmTile));
  if (e == cudaSuccess)
    e = cudaMemcpy(p->d_arm_tiles, tiles.data(), tiles.size() * sizeof(ArmTile), cudaMemcpyHostToDevice);
  p->arm_fwd_p = (A.dim_arm <= 16) ? 2 : 1;
  {
    // TMA staging of the ARM tiles needs 16-byte aligned row pitches and batch stride (ccb_arm_make_tmaps)
    bool ok = (g.M % 4) == 0;
    for (int i = 0; i < g.L; ++i) ok = ok && (g.w[i] % 4) == 0 && (g.off[i] % 4) == 0;
    const char* env = getenv("CCB_ARM_TMA");  // A/B switch: CCB_ARM_TMA=0 keeps the cp.async staging
    p->arm_tma = (ok && !(env && env[0] == '0')) ? 1 : 0;
  }
  p->d_arm_tiles_fwd = nullptr;
  if (e == cudaSuccess) {
    std::vector<ArmTile> tf;
    for (int b = 0; b < A.B; ++b)
      for (int l = 0; l < g.L; ++l)
        for (int y0 = 0; y0 < g.h[l]; y0 += p->arm_th)
          for (int x0 = 0; x0 < g.w[l]; x0 += p->arm_tw * p->arm_fwd_p) tf.push_back(ArmTile{b, l, y0, x0});
    p->n_arm_tiles_fwd = (int)tf.size();
    e = cudaMalloc(&p->d_arm_tiles_fwd, tf.size() * sizeof(ArmTile));
    if (e == cudaSuccess)
      e = cudaMemcpy(p->d_arm_tiles_fwd, tf.data(), tf.size() * sizeof(ArmTile), cudaMemcpyHostToDevice);
  }
  if (e != cudaSuccess) {
    ccb_set_cuda_error(e, __FILE__, __LINE__);
    if (p->d_arm_tiles) cudaFree(p->d_arm_tiles);
    if (p->d_arm_tiles_fwd) cudaFree(p->d_arm_tiles_fwd);
    delete p;
    return CCB_ERR_CUDA;
  }

  // ---- workspace layout
  size_t w = 0;
  const size_t BM = (size_t)A.B * g.M * sizeof(float);
  const size_t IMG3 = (size_t)A.B * 3 * A.H * A.W * sizeof(float);
  p->off_yhat = w; w = align_up(w + BM);
  p->off_dyhat = w; w = align_up(w + BM);
  p->off_dyhat2 = w; w = align_up(w + BM);
  p->off_dyhat3 = w; w = align_up(w + BM);  // directly after dyhat2: one memset clears both accumulators
  p->off_dq = w; w = align_up(w + BM);
  for (int i = 1; i < g.L; ++i) {
    const int lvl = g.L - 1 - i;  // up_i lives at the resolution of level L-1-i with i+1 channels
    const size_t bytes = (size_t)A.B * (i + 1) * g.h[lvl] * g.w[lvl] * sizeof(float);
    p->off_up[i] = w; w = align_up(w + bytes);
    p->off_dup[i] = w; w = align_up(w + bytes);
  }
  for (int l = 1; l < A.n_syn_layers; ++l) {
    p->off_act[l] = w; w = align_up(w + IMG3);
  }
  p->off_dact[0] = w; w = align_up(w + IMG3);
  p->off_dact[1] = w; w = align_up(w + IMG3);
  p->off_rate_part = w; w = align_up(w + (size_t)p->n_arm_tiles * sizeof(float));
  p->off_arm_part = w; w = align_up(w + (size_t)ccb_arm_bwd_max_rows(p) * wl.n_arm * sizeof(float));
  p->off_syn_part[0] = w;
  w = align_up(w + (size_t)ccb_syn_head_blocks(p) * ccb_syn_head_params(p) * sizeof(float));
  for (int l = 2; l < A.n_syn_layers; ++l) {
    p->off_syn_part[l] = w;
    w = align_up(w + (size_t)ccb_syn_conv_blocks(p) * 84 * sizeof(float));
  }
  {
    // one partial row per CTA of every stage of the upsampling backward
    size_t rows = 0;
    for (int idx = 0; idx < g.L - 1; ++idx) {
      const int lvl_s = g.L - 1 - idx;
      rows += (size_t)ceil_div(g.h[lvl_s], 16) * ceil_div(g.w[lvl_s], 16) * A.B * (idx + 2);
    }
    p->n_ups_part_rows = (int)rows;
    p->off_ups_part = w; w = align_up(w + rows * 16 * sizeof(float));
  }
  p->off_head_pack = w; w = align_up(w + 768 * sizeof(float));
  p->n_loss_blocks = p->n_sm * 4;
  p->off_scalars = w;  // 64 scalars | loss partials [nb][2] | sum-of-squares partials [max(nb, 8 * n_sm)]
  w = align_up(w + (size_t)(64 + 2 * p->n_loss_blocks + 8 * p->n_sm + 512) * sizeof(float));  // + weight-square partials
  p->off_draw = w; w = align_up(w + IMG3);   // d_raw_out of the fused train step
  p->off_raw = w; w = align_up(w + IMG3);    // raw_out of the fused train step
  p->off_rate = w; w = align_up(w + BM);     // rate of the fused train step
  p->ws_bytes = w;
  e = cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_head, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_split, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_big, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_qfine, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_start, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_stage, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_ups, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_join2, cudaEventDisableTiming);
  if (e != cudaSuccess) {
    ccb_set_cuda_error(e, __FILE__, __LINE__);
    cudaFree(p->d_arm_tiles);
    cudaFree(p->d_arm_tiles_fwd);
    delete p;
    return CCB_ERR_CUDA;
  }
  p->step_mutex = new (std::nothrow) std::mutex();
  if (!p->step_mutex) {
    ccb_plan_destroy(p);
    return CCB_ERR_NOMEM;
  }
  *out = p;
  return CCB_OK;
}

extern "C" void ccb_plan_destroy(ccb_plan* p) {
  if (!p) return;
  if (p->step_mutex) delete static_cast<std::mutex*>(p->step_mutex);
  if (p->d_arm_tiles) cudaFree(p->d_arm_tiles);
  if (p->d_arm_tiles_fwd) cudaFree(p->d_arm_tiles_fwd);
  if (p->ev_fork) cudaEventDestroy(p->ev_fork);
  if (p->ev_join) cudaEventDestroy(p->ev_join);
  if (p->ev_head) cudaEventDestroy(p->ev_head);
  if (p->ev_split) cudaEventDestroy(p->ev_split);
  if (p->ev_big) cudaEventDestroy(p->ev_big);
  if (p->ev_qfine) cudaEventDestroy(p->ev_qfine);
  if (p->ev_start) cudaEventDestroy(p->ev_start);
  if (p->ev_stage) cudaEventDestroy(p->ev_stage);
  if (p->ev_ups) cudaEventDestroy(p->ev_ups);
  if (p->ev_join2) cudaEventDestroy(p->ev_join2);
  if (p->host_arena) cudaFree(p->host_arena);
  if (p->host_s) cudaStreamDestroy(p->host_s);
  if (p->host_s2) cudaStreamDestroy(p->host_s2);
  delete p;
}

extern "C" int64_t ccb_plan_n_latent_values(const ccb_plan* p) { return p ? p->g.M : -1; }
extern "C" int ccb_plan_latent_geometry(const ccb_plan* p, int32_t* h, int32_t* w, int64_t* offset) {
  if (!p) return CCB_ERR_INVALID;
  for (int i = 0; i < p->g.L; ++i) {
    if (h) h[i] = p->g.h[i];
    if (w) w[i] = p->g.w[i];
    if (offset) offset[i] = p->g.off[i];
  }
  return CCB_OK;
}
extern "C" size_t ccb_plan_workspace_bytes(const ccb_plan* p) { return p ? p->ws_bytes : 0; }
extern "C" int64_t ccb_plan_n_weight_params(const ccb_plan* p) { return p ? p->wl.n_total : -1; }
extern "C" int64_t ccb_plan_n_param_floats(const ccb_plan* p) {
  return p ? (int64_t)p->g.B * p->g.M + p->wl.n_total : -1;
}

static int check_tensors(const ccb_plan* p, const ccb_tensors* t, bool need_latents) {
  if (!t) return CCB_ERR_INVALID;
  for (int i = 0; i < p->g.L; ++i)
    if (need_latents && !t->latent[i]) return CCB_ERR_INVALID;
  for (int k = 0; k <= p->arch.n_hidden_arm; ++k)
    if (!t->arm_w[k] || !t->arm_b[k]) return CCB_ERR_INVALID;
  for (int i = 0; i < p->g.L - 1; ++i)
    if (!t->ups_w[i] || !t->pre_w[i]) return CCB_ERR_INVALID;
  for (int l = 0; l < p->arch.n_syn_layers; ++l)
    if (!t->syn_w[l] || !t->syn_b[l]) return CCB_ERR_INVALID;
  return CCB_OK;
}

extern "C" int ccb_forward(const ccb_plan* p, const ccb_tensors* params, const ccb_quant_cfg* q, void* workspace,
                           float* raw_out, float* rate, float* mu, float* scale, float* log_scale, float* yhat_out,
                           void* stream) {
  if (!p || !q || !workspace || !raw_out || !rate) return CCB_ERR_INVALID;
  int rc = check_tensors(p, params, true);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  float* yhat = reinterpret_cast<float*>(ws + p->off_yhat);
  if ((rc = ccb_quant_forward(p, params, q, yhat, s, reinterpret_cast<unsigned*>(ws + p->off_scalars) + 9))) return rc;
  if (yhat_out)
    CCB_CHECK(cudaMemcpyAsync(yhat_out, yhat, (size_t)p->g.B * p->g.M * sizeof(float), cudaMemcpyDeviceToDevice, s));
  // module path: every kernel reads the weights from the caller's tensors (no weight slot, no state outside the call)
  if ((rc = ccb_arm_forward(p, params, yhat, rate, mu, scale, log_scale,
                            reinterpret_cast<float*>(ws + p->off_rate_part), s, -1)))
    return rc;
  if ((rc = ccb_ups_forward(p, params, ws, s))) return rc;
  if ((rc = ccb_syn_forward(p, params, ws, raw_out, s, -1))) return rc;
  return CCB_OK;
}

extern "C" int ccb_backward(const ccb_plan* p, const ccb_tensors* params, const ccb_quant_cfg* q, void* workspace,
                            const float* d_raw_out, const float* d_rate, float d_rate_uniform,
                            const ccb_tensors* grads, void* stream) {
  if (!p || !q || !workspace || !d_raw_out) return CCB_ERR_INVALID;
  int rc = check_tensors(p, params, true);
  if (rc) return rc;
  if ((rc = check_tensors(p, grads, false))) return rc;
  if (grads->weight_batch_stride != params->weight_batch_stride) return CCB_ERR_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  const float* yhat = reinterpret_cast<const float*>(ws + p->off_yhat);
  float* dyhat = reinterpret_cast<float*>(ws + p->off_dyhat);
  float* dyhat_a = reinterpret_cast<float*>(ws + p->off_dyhat2);
  float* dyhat_b = reinterpret_cast<float*>(ws + p->off_dyhat3);
  // synthesis -> d z ; upsampling -> d yhat (written) ; the ARM's share goes to its two accumulators (k_arm_bwd)
  if ((rc = ccb_syn_backward(p, params, ws, d_raw_out, s, nullptr, -1))) return rc;
  // the grid-barrier counter of the fused pyramid stages must be zero at every backward (a second backward through
  // the same forward would otherwise find it already advanced)
  CCB_CHECK(cudaMemsetAsync(reinterpret_cast<unsigned*>(ws + p->off_scalars) + 10, 0, sizeof(unsigned), s));
  UpsPartInfo ups_info;
  if ((rc = ccb_ups_backward(p, params, ws, grads, s, &ups_info))) return rc;
  CCB_CHECK(cudaMemsetAsync(dyhat_a, 0, (p->off_dyhat3 - p->off_dyhat2) + (size_t)p->g.B * p->g.M * sizeof(float), s));
  if ((rc = ccb_arm_backward(p, params, yhat, d_rate, d_rate_uniform, dyhat_a, dyhat_b,
                             reinterpret_cast<float*>(ws + p->off_arm_part), s, -1)))
    return rc;
  if ((rc = ccb_reduce_arm_grads(p, ws, grads, s))) return rc;
  if ((rc = ccb_reduce_syn_grads(p, ws, grads, s, &ups_info))) return rc;
  if ((rc = ccb_quant_backward(p, params, q, dyhat, dyhat_a, dyhat_b, grads, s))) return rc;
  return CCB_OK;
}


// Shared body of ccb_train_step / ccb_train_step_timed.
//   * `aux` != NULL: the ARM branch (rate forward, its backward, its weight-gradient reduction) runs on `aux`,
//     forked after the quantiser and joined before the quantiser backward; it only shares yhat (read) with
//     the upsampling -> synthesis -> loss -> backward chain on `s`, and writes its share of d yhat to its
//     own buffer.  Every kernel of the step is latency- rather than throughput-bound at 768x512, so the two
//     chains overlap almost perfectly.
//   * `ev` != NULL (timed variant, serial schedule): one event after every stage, in CCB_N_STAGES order.
bool ccb_debug_skip(int bit) {
  static const int mask = []() {
    const char* e = getenv("CCB_DEBUG_SKIP");
    return e ? atoi(e) : 0;
  }();
  return (mask & bit) != 0;
}

static int train_step_impl(const ccb_plan* p, const ccb_tensors* params, const ccb_tensors* grads,
                           const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, const ccb_quant_cfg* q,
                           void* workspace, const float* target, float lambda, int optimize_mask, float max_norm,
                           const float* lr_dev, double beta1, double beta2, double eps, int32_t* step_dev,
                           uint64_t* philox_offset_dev, float* out4, cudaStream_t s, cudaStream_t aux,
                           cudaEvent_t* ev, int slot) {
  if (!p || !workspace || !target || !out4 || !q) return CCB_ERR_INVALID;
  if (slot >= CCB_N_WEIGHT_SLOTS) return CCB_ERR_INVALID;
  int rc = check_tensors(p, params, true);
  if (rc) return rc;
  if ((rc = check_tensors(p, grads, false))) return rc;
  // the fused iteration trains ONE weight set (loss, clip coefficient and Adam step are per encoder)
  if (params->weight_batch_stride != 0 || grads->weight_batch_stride != 0) return CCB_ERR_UNSUPPORTED;
  // host threads that share this plan (several encoders of one shape trained side by side) enqueue one at a time
  std::lock_guard<std::mutex> step_lock(*static_cast<std::mutex*>(p->step_mutex));
  char* ws = (char*)workspace;
  float* raw = reinterpret_cast<float*>(ws + p->off_raw);
  float* draw = reinterpret_cast<float*>(ws + p->off_draw);
  float* yhat = reinterpret_cast<float*>(ws + p->off_yhat);
  float* dyhat = reinterpret_cast<float*>(ws + p->off_dyhat);
  float* dyhat_arm = reinterpret_cast<float*>(ws + p->off_dyhat2);   // own-band accumulator
  float* dyhat_arm_b = reinterpret_cast<float*>(ws + p->off_dyhat3);  // band-below accumulator (contiguous with it)
  const size_t arm_acc_bytes = (p->off_dyhat3 - p->off_dyhat2) + (size_t)p->g.B * p->g.M * sizeof(float);
  float* rate_part = reinterpret_cast<float*>(ws + p->off_rate_part);
  float* arm_part = reinterpret_cast<float*>(ws + p->off_arm_part);
  const size_t BM = (size_t)p->g.B * p->g.M * sizeof(float);
  ccb_quant_cfg qq = *q;
  qq.training = 1;
  if (philox_offset_dev) qq.offset_dev = philox_offset_dev;
  const float d_rate = lambda / ((float)p->arch.B * (float)p->arch.H * (float)p->arch.W);
  ccb_tensors g = *grads;
  if (qq.q_type == CCB_Q_STE || qq.q_type == CCB_Q_HARDROUND)
    for (int i = 0; i < CCB_MAX_LATENTS; ++i) g.latent[i] = nullptr;  // no latent gradient (quantizer.py:222-225)
  if (ev) aux = nullptr;
  // ARM backward in two launches (two-stream schedule, one weight set): the first part starts with the step and leaves
  // before the heaviest kernels of the chain end, the second runs beside the small pyramid stages of the backward, whose
  // few CTAs leave most SMs idle.  arm_split = first tile of the second part (0: one launch).  CCB_ARM_SPLIT = percent
  // of the tiles in the first part, CCB_ARM_SPLIT_AT = 0 after the synthesis backward, 1 after the largest pyramid stage
  int arm_split = 0;
  {
    static const int pct = []() { const char* e = getenv("CCB_ARM_SPLIT"); return e ? atoi(e) : 100; }();
    int gx, gy;
    ccb_arm_bwd_grid(p, false, &gx, &gy);
    if (aux && pct > 0 && pct < 100 && gx == 2 * p->n_sm && gy == 1 && p->arch.dim_arm <= 16)
      arm_split = (int)((long long)p->n_arm_tiles * pct / 100);
    if (arm_split < p->n_sm || arm_split > p->n_arm_tiles - p->n_sm) arm_split = 0;
  }
  static const int split_at = []() { const char* e = getenv("CCB_ARM_SPLIT_AT"); return e ? atoi(e) : 1; }();
  int stage = 0;
#define STAGE_DONE()                                         \
  do {                                                       \
    if (ev) CCB_CHECK(cudaEventRecord(ev[++stage], s));      \
  } while (0)
  if (ev) CCB_CHECK(cudaEventRecord(ev[0], s));
  // latents of one image in one arena: the forward parks the quantiser derivative, the backward is one multiply
  bool dq_ok = ccb_quant_flat_ok(p, params, &g) && (qq.q_type == CCB_Q_NONE || qq.q_type == CCB_Q_SOFTROUND ||
                                                    qq.q_type == CCB_Q_SOFTROUND_ALONE || qq.q_type == CCB_Q_TRUE_STE);
  static const bool no_dq = (getenv("CCB_NO_QUANT_DQ") != nullptr);  // A/B switch
  if (no_dq) dq_ok = false;
  float* dq = dq_ok ? reinterpret_cast<float*>(ws + p->off_dq) : nullptr;
  // Split quantiser (two-stream schedule, one image, cut points on multiples of four): the coarse levels -- all the
  // pyramid's chain kernel reads -- first on the main stream, the finest level(s) on the second stream beside that
  // kernel; likewise the backward's fine levels are final before the fused small pyramid stages start, and run beside them
  // CCB_QUANT_SPLIT: bit 0 = forward, bit 1 = backward.  Measured on B200 (alternating runs, it/s): none 4 370, forward
  // only 4 245 (the finest level's quantiser on the second stream delays the ARM branch and adds two cross-stream edges),
  // backward only 4 411, both 4 287: the default is the backward split alone.
  static const int qsplit = []() { const char* e = getenv("CCB_QUANT_SPLIT"); return e ? atoi(e) : 2; }();
  const long long cut_f = (aux && !arm_split && (qsplit & 1)) ? ccb_quant_cut(p, ccb_ups_chain_min_level(p)) : -1;
  const long long cut_b = (aux && !arm_split && (qsplit & 2) && dq) ? ccb_quant_cut(p, ccb_ups_fused_min_level(p)) : -1;
  unsigned* zero_ctr = reinterpret_cast<unsigned*>(ws + p->off_scalars) + 9;
  // CCB_EARLY_STAGE=1: the second stream forks at the START of the step (weights -> constant banks and the clearing of
  // the ARM accumulators beside the quantiser forward, so that the ARM backward starts as soon as y_hat exists).
  // Measured on B200, alternating runs: 4 416 it/s against 4 463 for staging after the quantiser with the L2 flushed
  // (4 661 against 4 649 back to back): the earlier ARM start overlaps more of the forward kernels.  Off by default.
  static const bool want_early = (getenv("CCB_EARLY_STAGE") != nullptr);
  const bool early_stage = aux && want_early;
  if (early_stage) {
    CCB_CHECK(cudaEventRecord(p->ev_start, s));
    CCB_CHECK(cudaStreamWaitEvent(aux, p->ev_start, 0));
    if ((rc = ccb_arm_stage_weights(p, params, aux, slot))) return rc;
    if ((rc = ccb_syn_stage_weights(p, params, ws, aux, slot))) return rc;
    CCB_CHECK(cudaEventRecord(p->ev_stage, aux));
    CCB_CHECK(cudaMemsetAsync(dyhat_arm, 0, arm_acc_bytes, aux));
  }
  if (cut_f > 0) {
    if ((rc = ccb_quant_forward(p, params, &qq, yhat, s, zero_ctr, dq, cut_f, p->g.M))) return rc;
  } else {
    if ((rc = ccb_quant_forward(p, params, &qq, yhat, s, zero_ctr, dq))) return rc;
  }
  STAGE_DONE();
  if (aux) {
    // ---- ARM branch on the second stream
    CCB_CHECK(cudaEventRecord(p->ev_fork, s));
    CCB_CHECK(cudaStreamWaitEvent(aux, p->ev_fork, 0));
    if (cut_f > 0) {
      if ((rc = ccb_quant_forward(p, params, &qq, yhat, aux, nullptr, dq, 0, cut_f))) return rc;
      CCB_CHECK(cudaEventRecord(p->ev_qfine, aux));
    }
    if (!early_stage) {
      // weights -> constant banks, beside the pyramid on the main stream (the head forward waits for ev_stage)
      if ((rc = ccb_arm_stage_weights(p, params, aux, slot))) return rc;
      if ((rc = ccb_syn_stage_weights(p, params, ws, aux, slot))) return rc;
      CCB_CHECK(cudaEventRecord(p->ev_stage, aux));
      CCB_CHECK(cudaMemsetAsync(dyhat_arm, 0, arm_acc_bytes, aux));
    }
    // no separate ARM forward: the backward recomputes it anyway and emits the per-tile rate sums itself
    // CCB_DEBUG_SKIP_ARM: timing experiment only (how long is the main chain without the ARM branch beside it?);
    // the iteration's results are wrong with it
    static const bool skip_arm = (getenv("CCB_DEBUG_SKIP_ARM") != nullptr) || ccb_debug_skip(1);
    if (!skip_arm && (rc = ccb_arm_backward(p, params, yhat, nullptr, d_rate, dyhat_arm, dyhat_arm_b, arm_part, aux, slot,
                                            true, rate_part, true, 0, arm_split, 0)))
      return rc;
    if (!arm_split && (rc = ccb_reduce_arm_grads(p, ws, &g, aux))) return rc;
  } else {
    if ((rc = ccb_syn_stage_weights(p, params, ws, s, slot))) return rc;
    if ((rc = ccb_arm_stage_weights(p, params, s, slot))) return rc;
  }
  STAGE_DONE();
  if ((rc = ccb_ups_forward(p, params, ws, s, cut_f > 0 ? p->ev_qfine : nullptr))) return rc;
  STAGE_DONE();
  if (aux) CCB_CHECK(cudaStreamWaitEvent(s, p->ev_stage, 0));
  SynLossFuse lf;
  ccb_loss_fuse_args(p, ws, target, draw, &lf);
  static const bool no_tail_fuse = (getenv("CCB_NO_TAIL_FUSE") != nullptr);
  if ((rc = ccb_syn_forward(p, params, ws, raw, s, slot, true, no_tail_fuse ? nullptr : &lf))) return rc;
  STAGE_DONE();
  if (!lf.done && (rc = ccb_loss_image(p, ws, raw, target, draw, s))) return rc;
  STAGE_DONE();
  static const bool split_head = (getenv("CCB_SPLIT_HEAD") != nullptr);  // measured slower on B200: off by default
  if ((rc = ccb_syn_backward(p, params, ws, draw, s, split_head ? aux : nullptr, slot, true))) return rc;
  if (aux && !arm_split) CCB_CHECK(cudaEventRecord(p->ev_join, aux));
  bool latent_grads = false;
  for (int i = 0; i < p->g.L; ++i) latent_grads |= (g.latent[i] != nullptr);
  const bool fused_tail = latent_grads && ccb_quant_grid(p) <= 8 * p->n_sm;
  // flat arenas: the quantiser backward only leaves its sum-of-squares partials and the Adam kernel folds
  // everything in its prologue; otherwise the last CTA of the quantiser backward does (ticket)
  const bool fold_in_adam = fused_tail && ccb_adam_flat_ok(p, params, &g, exp_avg, exp_avg_sq, optimize_mask);
  // with the fold in the Adam prologue the column sums of the synthesis / pyramid partial rows are independent of
  // the quantiser backward: they run on the second stream -- the synthesis rows as soon as the synthesis backward is
  // done (beside the pyramid backward), the pyramid's few rows beside the quantiser backward; joined before Adam
  const bool reduce_on_aux = aux && fold_in_adam;
  const bool early_syn_reduce = reduce_on_aux && !split_head;
  if (early_syn_reduce) {
    CCB_CHECK(cudaEventRecord(p->ev_head, s));
    CCB_CHECK(cudaStreamWaitEvent(aux, p->ev_head, 0));
    if ((rc = ccb_reduce_syn_grads(p, ws, &g, aux, nullptr, 1))) return rc;
  }
  if (arm_split && split_at == 0) CCB_CHECK(cudaEventRecord(p->ev_split, s));
  STAGE_DONE();
  UpsPartInfo ups_info;
  if ((rc = ccb_ups_backward(p, params, ws, &g, s, &ups_info, (arm_split && split_at != 0) ? p->ev_split : nullptr,
                             early_syn_reduce ? p->ev_big : nullptr)))
    return rc;
  if (arm_split) {
    // second part of the ARM backward, then the column sums of its partial rows
    CCB_CHECK(cudaStreamWaitEvent(aux, p->ev_split, 0));
    if ((rc = ccb_arm_backward(p, params, yhat, nullptr, d_rate, dyhat_arm, dyhat_arm_b, arm_part, aux, slot, true,
                               rate_part, true, arm_split, p->n_arm_tiles, p->n_sm)))
      return rc;
    if ((rc = ccb_reduce_arm_grads(p, ws, &g, aux))) return rc;
    CCB_CHECK(cudaEventRecord(p->ev_join, aux));
  }
  STAGE_DONE();
  CcbTail tail;
  int n_sq_big = 0;  // > 0: the quantiser backward of the fine levels already ran (second stream), with that many partials
  if (fused_tail) ccb_tail_prepare(p, ws, &g, max_norm, beta1, beta2, step_dev, lambda, out4, philox_offset_dev, &tail);
  if (aux) {
    if (reduce_on_aux) {
      CCB_CHECK(cudaEventRecord(p->ev_ups, s));
      if (early_syn_reduce) {
        // rows of the large pyramid stages beside the fused small stages, the latter's few rows after them
        // (enqueued here, but the first launch only waits for the event behind the last large stage)
        CCB_CHECK(cudaStreamWaitEvent(aux, p->ev_big, 0));
        if ((rc = ccb_reduce_syn_grads(p, ws, &g, aux, &ups_info, 4))) return rc;
        // d yhat of the fine levels is final too (the fused small stages only write the coarse ones), and the ARM's
        // share was produced on this stream: their quantiser backward runs here, beside the fused small stages
        if (cut_b > 0 && fold_in_adam && !ccb_debug_skip(4) &&
            (rc = ccb_quant_backward(p, params, &qq, dyhat, dyhat_arm, dyhat_arm_b, &g, aux, &tail, 2, dq, 0, cut_b, 0,
                                     &n_sq_big)))
          return rc;
      }
      CCB_CHECK(cudaStreamWaitEvent(aux, p->ev_ups, 0));
      if ((rc = ccb_reduce_syn_grads(p, ws, &g, aux, &ups_info, early_syn_reduce ? 2 : 7))) return rc;
      CCB_CHECK(cudaEventRecord(p->ev_join2, aux));
    }
    CCB_CHECK(cudaStreamWaitEvent(s, p->ev_join, 0));
  } else {
    // (timed, serial schedule: same shared-SM configuration as the two-stream step, so that the stage time is the
    // launch the step really contains)
    CCB_CHECK(cudaMemsetAsync(dyhat_arm, 0, arm_acc_bytes, s));
    // CCB_TIMED_ARM_ALONE (read per call, timed entry point only): two CTAs per SM, the form the kernel takes when
    // nothing has to co-reside with it (the module path); bench.py reports both
    const bool alone = getenv("CCB_TIMED_ARM_ALONE") != nullptr;
    if ((rc = ccb_arm_backward(p, params, yhat, nullptr, d_rate, dyhat_arm, dyhat_arm_b, arm_part, s, slot, true,
                               rate_part, !alone)))
      return rc;
  }
  STAGE_DONE();
  if (!aux && (rc = ccb_reduce_arm_grads(p, ws, &g, s))) return rc;
  if (!reduce_on_aux && (rc = ccb_reduce_syn_grads(p, ws, &g, s, &ups_info))) return rc;
  STAGE_DONE();
  if (fused_tail) {
    // the quantiser backward also carries the finals (norm, clip, bias corrections, loss, Philox advance)
    int n_sq = ccb_quant_grid(p);
    if (n_sq_big > 0) {
      int n_small = 0;
      if ((rc = ccb_quant_backward(p, params, &qq, dyhat, dyhat_arm, dyhat_arm_b, &g, s, &tail, 2, dq, cut_b, p->g.M,
                                   n_sq_big, &n_small)))
        return rc;
      n_sq = n_sq_big + n_small;
    } else if (!ccb_debug_skip(4) &&
               (rc = ccb_quant_backward(p, params, &qq, dyhat, dyhat_arm, dyhat_arm_b, &g, s, &tail,
                                        fold_in_adam ? 2 : 1, dq))) {
      return rc;
    }
    STAGE_DONE();
    if (reduce_on_aux) CCB_CHECK(cudaStreamWaitEvent(s, p->ev_join2, 0));
    if ((rc = ccb_adam_only(p, ws, params, &g, exp_avg, exp_avg_sq, optimize_mask, lr_dev, beta1, beta2, eps, s,
                            fold_in_adam ? &tail : nullptr, n_sq)))
      return rc;
  } else {
    if ((rc = ccb_quant_backward(p, params, &qq, dyhat, dyhat_arm, dyhat_arm_b, &g, s, nullptr, 1, dq))) return rc;
    STAGE_DONE();
    if ((rc = ccb_tail_step(p, ws, params, &g, exp_avg, exp_avg_sq, optimize_mask, max_norm, lr_dev, beta1, beta2, eps,
                            step_dev, lambda, out4, philox_offset_dev, s)))
      return rc;
  }
  STAGE_DONE();
#undef STAGE_DONE
  return CCB_OK;
}

extern "C" int ccb_plan_train_step_launches(const ccb_plan* p) {
  if (!p) return 0;
  const int n_tail = p->arch.n_syn_layers - 2;
  // quantiser fwd, ARM bwd (with its own forward recompute and rate sums) + column sums, pyramid fwd, head + 3x3 fwd, loss, 3x3 + head bwd, pyramid bwd,
  // column sums (synthesis + pyramid), quantiser bwd (+ finals), Adam; + the head's weight packing when the head
  // reads the constant bank (the two device-to-device copies into the banks are copy nodes, not kernels)
  const int S = p->syn_hidden;
  const int pack = (S == 8 || S == 16 || S == 40 || S == 48) ? 1 : 0;
  const int fwd_tail = (n_tail == 2) ? 1 : n_tail + 1;  // two 3x3 layers + loss fused into one launch
  // column sums: ARM rows (counted with the ARM backward), synthesis rows, rows of the large pyramid stages, rows of the
  // fused small ones (three launches in the two-stream schedule with flat arenas -- FusedState, ccb_overfit_host --
  // which is what this count describes; when every pyramid stage is fused or none is, one of them is empty)
  const int nb = ccb_ups_n_launches(p, true);
  const int ups_colsums = (nb > 1 && nb < p->g.L - 1) ? 2 : 1;
  // the quantiser backward runs as two launches where the flat latent order can be cut at the pyramid's fused level
  const int q_split = ccb_quant_cut(p, ccb_ups_fused_min_level(p)) > 0 ? 1 : 0;
  const int bwd_tail = (n_tail == 2) ? 1 : n_tail;  // two trailing 3x3 layers: one fused backward launch
  return pack + 1 + 2 + ccb_ups_n_launches(p, false) + 1 + fwd_tail + (bwd_tail + 1) + nb + 1 + ups_colsums + 1 + 1 + q_split;
}

extern "C" int ccb_train_step(const ccb_plan* p, const ccb_tensors* params, const ccb_tensors* grads,
                              const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, const ccb_quant_cfg* q,
                              void* workspace, const float* target, float lambda, int optimize_mask,
                              float max_norm, const float* lr_dev, double beta1, double beta2, double eps,
                              int32_t* step_dev, uint64_t* philox_offset_dev, float* out4, void* stream,
                              void* aux_stream, int weight_slot) {
  return train_step_impl(p, params, grads, exp_avg, exp_avg_sq, q, workspace, target, lambda, optimize_mask,
                         max_norm, lr_dev, beta1, beta2, eps, step_dev, philox_offset_dev, out4,
                         (cudaStream_t)stream, (cudaStream_t)aux_stream, nullptr, weight_slot);
}

extern "C" int ccb_train_step_timed(const ccb_plan* p, const ccb_tensors* params, const ccb_tensors* grads,
                                    const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq,
                                    const ccb_quant_cfg* q, void* workspace, const float* target, float lambda,
                                    int optimize_mask, float max_norm, const float* lr_dev, double beta1,
                                    double beta2, double eps, int32_t* step_dev, uint64_t* philox_offset_dev,
                                    float* out4, void* stream, float* stage_ms_host, int weight_slot) {
  if (!stage_ms_host) return CCB_ERR_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  cudaEvent_t ev[CCB_N_STAGES + 1];
  for (int i = 0; i <= CCB_N_STAGES; ++i) CCB_CHECK(cudaEventCreate(&ev[i]));
  int rc = train_step_impl(p, params, grads, exp_avg, exp_avg_sq, q, workspace, target, lambda, optimize_mask,
                           max_norm, lr_dev, beta1, beta2, eps, step_dev, philox_offset_dev, out4, s, nullptr, ev,
                           weight_slot);
  if (rc == CCB_OK) {
    cudaError_t e = cudaEventSynchronize(ev[CCB_N_STAGES]);
    if (e != cudaSuccess) {
      ccb_set_cuda_error(e, __FILE__, __LINE__);
      rc = CCB_ERR_CUDA;
    }
    for (int i = 0; i < CCB_N_STAGES && rc == CCB_OK; ++i) cudaEventElapsedTime(&stage_ms_host[i], ev[i], ev[i + 1]);
  }
  for (int i = 0; i <= CCB_N_STAGES; ++i) cudaEventDestroy(ev[i]);
  return rc;
}

// Carve a flat float buffer into a ccb_tensors, in declaration order.
static void carve(const ccb_plan* p, float* base, ccb_tensors* t) {
  memset(t, 0, sizeof(*t));
  float* c = base;
  for (int i = 0; i < p->g.L; ++i) {
    t->latent[i] = c;
    c += (size_t)p->g.B * p->g.h[i] * p->g.w[i];
  }
  for (int k = 0; k < p->wl.n_arm_layers; ++k) {
    t->arm_w[k] = c; c += p->wl.arm_w_n[k];
    t->arm_b[k] = c; c += p->wl.arm_b_n[k];
  }
  for (int i = 0; i < p->g.L - 1; ++i) { t->ups_w[i] = c; c += p->wl.ups_n; }
  for (int i = 0; i < p->g.L - 1; ++i) { t->pre_w[i] = c; c += p->wl.pre_n; }
  for (int l = 0; l < p->arch.n_syn_layers; ++l) {
    t->syn_w[l] = c; c += p->wl.syn_w_n[l];
    t->syn_b[l] = c; c += p->wl.syn_b_n[l];
  }
}

// Makes `device` current for the scope and puts the caller's device back on every return path.
struct CcbDeviceScope {
  int prev = -1;
  cudaError_t err;
  explicit CcbDeviceScope(int device) {
    err = cudaGetDevice(&prev);
    if (err != cudaSuccess) prev = -1;
    if (err == cudaSuccess && prev != device) err = cudaSetDevice(device);
    else if (err == cudaSuccess) prev = -1;  // already current: nothing to restore
  }
  ~CcbDeviceScope() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

extern "C" int ccb_overfit_host(const ccb_plan* p, float* params_host, const float* target_host,
                                const ccb_quant_cfg* q, float lambda, float lr, int optimize_mask, int n_iters,
                                float* out4_host) {
  if (!p || !params_host || !target_host || !q || !out4_host || n_iters < 1) return CCB_ERR_INVALID;
  CcbDeviceScope dev_scope(p->device);
  CCB_CHECK(dev_scope.err);
  const size_t n_par = (size_t)ccb_plan_n_param_floats(p);
  const size_t n_row = (n_par + 3) & ~size_t(3);  // rows start 16-byte aligned: k_adam_flat takes its 128-bit path
  const size_t n_img = (size_t)p->arch.B * 3 * p->arch.H * p->arch.W;
  float* d_all = nullptr;   // params | grads | m | v, one row of n_row floats each
  float* d_tgt = nullptr;
  char* d_ws = nullptr;
  float* d_small = nullptr;  // out4[4], lr, step(int), offset(u64)
  cudaStream_t s = nullptr, s2 = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  int rc = CCB_OK;
  int slot = -1;  // a weight slot for the duration of the call (none free: the tensor-reading kernels, same results)
  cudaError_t e;
#define OF_CHECK(call)                                 \
  do {                                                 \
    e = (call);                                        \
    if (e != cudaSuccess) {                            \
      ccb_set_cuda_error(e, __FILE__, __LINE__);       \
      rc = CCB_ERR_CUDA;                               \
      goto done;                                       \
    }                                                  \
  } while (0)
  {
    // Device buffers and the stream pair of the host-buffer entry point are owned by the plan and reused by
    // later calls (freed by ccb_plan_destroy): cudaMalloc / cudaFree of ~100 MB cost milliseconds and
    // synchronise the device, which is not part of the encode.
    ccb_plan* mp = const_cast<ccb_plan*>(p);
    if (!mp->host_arena) {
      OF_CHECK(cudaStreamCreateWithFlags(&mp->host_s, cudaStreamNonBlocking));
      OF_CHECK(cudaStreamCreateWithFlags(&mp->host_s2, cudaStreamNonBlocking));
      const size_t a_all = align_up(4 * n_row * sizeof(float)), a_tgt = align_up(n_img * sizeof(float));
      OF_CHECK(cudaMalloc(&mp->host_arena, a_all + a_tgt + align_up(p->ws_bytes) + 256));
      mp->host_off_tgt = a_all;
      mp->host_off_ws = a_all + a_tgt;
      mp->host_off_small = a_all + a_tgt + align_up(p->ws_bytes);
    }
    s = mp->host_s;
    s2 = mp->host_s2;
    d_all = reinterpret_cast<float*>(mp->host_arena);
    d_tgt = reinterpret_cast<float*>(mp->host_arena + mp->host_off_tgt);
    d_ws = mp->host_arena + mp->host_off_ws;
    d_small = reinterpret_cast<float*>(mp->host_arena + mp->host_off_small);
    OF_CHECK(cudaMemsetAsync(d_all + n_row, 0, 3 * n_row * sizeof(float), s));
    OF_CHECK(cudaMemsetAsync(d_small, 0, 64, s));
    OF_CHECK(cudaMemcpyAsync(d_all, params_host, n_par * sizeof(float), cudaMemcpyHostToDevice, s));
    OF_CHECK(cudaMemcpyAsync(d_tgt, target_host, n_img * sizeof(float), cudaMemcpyHostToDevice, s));
    OF_CHECK(cudaMemcpyAsync(d_small + 4, &lr, sizeof(float), cudaMemcpyHostToDevice, s));
    ccb_tensors prm, grd, m1, m2;
    carve(p, d_all, &prm);
    carve(p, d_all + n_row, &grd);
    carve(p, d_all + 2 * n_row, &m1);
    carve(p, d_all + 3 * n_row, &m2);
    ccb_quant_cfg qq = *q;
    qq.noise = nullptr;
    int32_t* step_dev = reinterpret_cast<int32_t*>(d_small + 5);
    uint64_t* off_dev = reinterpret_cast<uint64_t*>(d_small + 6);
    if (ccb_weight_slot_acquire(p->device, &slot) != CCB_OK) slot = -1;
    OF_CHECK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    rc = ccb_train_step(p, &prm, &grd, &m1, &m2, &qq, d_ws, d_tgt, lambda, optimize_mask, 1e-1f, d_small + 4, 0.9,
                        0.999, 1e-8, step_dev, off_dev, d_small, s, s2, slot);
    e = cudaStreamEndCapture(s, &graph);
    if (rc != CCB_OK) goto done;
    OF_CHECK(e);
    OF_CHECK(cudaGraphInstantiate(&exec, graph, 0));
    for (int it = 0; it < n_iters; ++it) OF_CHECK(cudaGraphLaunch(exec, s));
    OF_CHECK(cudaMemcpyAsync(params_host, d_all, n_par * sizeof(float), cudaMemcpyDeviceToHost, s));
    OF_CHECK(cudaMemcpyAsync(out4_host, d_small, 4 * sizeof(float), cudaMemcpyDeviceToHost, s));
    OF_CHECK(cudaStreamSynchronize(s));
  }
done:
#undef OF_CHECK
  if (exec) cudaGraphExecDestroy(exec);
  if (graph) cudaGraphDestroy(graph);
  if (slot >= 0) ccb_weight_slot_release(p->device, slot);
  return rc;
}

// ------------------------------------------------------------------------------------------
// FP32 FMA peak: the roofline denominator of the compute-bound kernels
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
  float x[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = (float)(threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
  if (s == 123.456f) out[0] = s;
}

extern "C" int ccb_fp32_peak(int device, int n_iters, float* tflops_host) {
  if (!tflops_host) return CCB_ERR_INVALID;
  CcbDeviceScope dev_scope(device);
  CCB_CHECK(dev_scope.err);
  int n_sm = 0;
  CCB_CHECK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device));
  float* d = nullptr;
  CCB_CHECK(cudaMalloc(&d, 16));
  cudaEvent_t e0, e1;
  CCB_CHECK(cudaEventCreate(&e0));
  CCB_CHECK(cudaEventCreate(&e1));
  const int blocks = n_sm * 8;
  k_fma_peak<<<blocks, 256>>>(d, 1000, 0.999f, 0.001f);
  float best = 0.f;
  for (int rep = 0; rep < 5; ++rep) {
    CCB_CHECK(cudaEventRecord(e0));
    k_fma_peak<<<blocks, 256>>>(d, n_iters, 0.999f, 0.001f);
    CCB_CHECK(cudaEventRecord(e1));
    CCB_CHECK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CCB_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    const double flops = 2.0 * 16.0 * (double)n_iters * 256.0 * blocks;
    const float tf = (float)(flops / (ms * 1e-3) / 1e12);
    if (tf > best) best = tf;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  *tflops_host = best;
  return CCB_OK;
}
