/* ccb.h -- C ABI of libccb (coolchic_b200/csrc), the B200-native (sm_100a) kernels for
 * Cool-chic's per-image overfitting step.
 *
 * The reference (pepborrell/cool-chic) has no FFI for this path: the hot path is a
 * Python nn.Module API.  The entry points below are what a
 * maintainer's ctypes stub binds underneath those modules; each
 * cites the reference code it replaces (paths relative to /root/reference/coolchic).
 *
 * Conventions
 *   - every function returns 0 on success or a negative ccb_status; it never throws;
 *   - all pointers are DEVICE pointers unless the name ends in _host;
 *   - the library owns no tensor memory: the caller allocates parameters, outputs
 *     and one workspace of ccb_plan_workspace_bytes() bytes per in-flight forward;
 *   - every launch is asynchronous on `stream` (a cudaStream_t passed as void*);
 *   - plans are immutable after creation and can be shared by streams / threads;
 *   - no process-global state on the module path: ccb_forward / ccb_backward / ccb_loss / ccb_clip_adam_step read
 *     every weight from the caller's tensors (any number of calls may be in flight on different streams);
 *   - the fused iteration (ccb_train_step) may additionally keep its weights in one of CCB_N_WEIGHT_SLOTS
 *     constant-bank slots per device (uniform-register operands); a slot is owned by whoever acquired it
 *     (ccb_weight_slot_acquire), so independent encodes on different streams never share one;
 *   - fp32 everywhere except the integer ARM (int32, wrap-around semantics).
 */
#ifndef CCB_H
#define CCB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CCB_MAX_LATENTS 8
#define CCB_MAX_ARM_LAYERS 5 /* hidden layers + output layer */
#define CCB_MAX_SYN_LAYERS 8
#define CCB_N_WEIGHT_SLOTS 8 /* constant-bank weight slots per device (ccb_weight_slot_acquire) */

typedef enum {
  CCB_OK = 0,
  CCB_ERR_INVALID = -1,     /* bad argument */
  CCB_ERR_UNSUPPORTED = -2, /* architecture outside what the kernels cover */
  CCB_ERR_CUDA = -3,        /* a CUDA runtime call failed (see ccb_last_cuda_error) */
  CCB_ERR_NOMEM = -4,
  CCB_ERR_TOO_LARGE = -5    /* problem size beyond the static scratch of an allocation-free entry point */
} ccb_status;

/* quantizer_type (enc/component/core/quantizer.py:116-232) */
typedef enum {
  CCB_Q_NONE = 0,
  CCB_Q_SOFTROUND_ALONE = 1,
  CCB_Q_SOFTROUND = 2,
  CCB_Q_HARDROUND = 3,
  CCB_Q_STE = 4,
  CCB_Q_TRUE_STE = 5
} ccb_qtype;

/* quantizer_noise_type (quantizer.py:62-107,186-199) */
typedef enum { CCB_NOISE_NONE = 0, CCB_NOISE_GAUSSIAN = 1, CCB_NOISE_KUMARASWAMY = 2 } ccb_noise;

/* One synthesis layer "<out>-<k>-<linear|residual>-<none|relu>" (core/synthesis.py:175-260). */
typedef struct {
  int32_t out_ch, k, residual, relu;
} ccb_syn_layer;

/* Architecture = CoolChicEncoderParameter + image size (enc/component/coolchic.py:55-124). */
typedef struct {
  int32_t B, H, W;
  int32_t n_latents;    /* latent_n_grids; one feature per resolution (arm.py:333) */
  int32_t dim_arm;      /* 8 | 16 | 24 | 32 */
  int32_t n_hidden_arm; /* 0..3 */
  int32_t ups_k;        /* even, <= 8 */
  int32_t ups_pre_k;    /* odd,  <= 7 */
  int32_t n_syn_layers;
  ccb_syn_layer syn[CCB_MAX_SYN_LAYERS];
} ccb_arch;

/* Parameter (or gradient, or Adam moment) tensors, named as the reference's state_dict
 * (SURVEY.md 8b): latent_grids.{i}.data [B,1,h_i,w_i]; arm.mlp.{2k}.weight [C,C] / [2,C],
 * .bias; upsampling.conv_transpose2ds.{i}.parametrizations.weight.original [ups_k/2];
 * upsampling.conv2ds.{i}...original [(ups_pre_k+1)/2]; synthesis.layers.{2i}.weight
 * [cout,cin,k,k], .bias [cout].  The never-used upsampling biases are not part of it.
 *
 * weight_batch_stride (floats): 0 = ONE weight set shared by the B batch elements (CoolChicEncoder, NOWholeNet);
 * > 0 = per-sample weights, the way DeltaWholeNet runs the decoder under
 * vmap(functional_call(...)) (hypernet/hypernet.py:1266-1288, hypernet/common.py:227-253): every weight pointer
 * addresses the tensor of batch element 0 and element b's copy lives `b * weight_batch_stride` floats further
 * (one packed [B, n_weights] buffer).  In a gradient set the same stride applies: per-sample gradients.
 * Honoured by ccb_forward / ccb_backward; the fused iteration (ccb_train_step) requires 0. */
typedef struct {
  float* latent[CCB_MAX_LATENTS];
  float* arm_w[CCB_MAX_ARM_LAYERS];
  float* arm_b[CCB_MAX_ARM_LAYERS];
  float* ups_w[CCB_MAX_LATENTS];
  float* pre_w[CCB_MAX_LATENTS];
  float* syn_w[CCB_MAX_SYN_LAYERS];
  float* syn_b[CCB_MAX_SYN_LAYERS];
  int64_t weight_batch_stride;
} ccb_tensors;

/* Arguments of quantize() as CoolChicEncoder.forward passes them (coolchic.py:365-377). */
typedef struct {
  int32_t training; /* 0: eval => noise "none", quantizer "hardround" (coolchic.py:367-368) */
  int32_t q_type;   /* ccb_qtype */
  int32_t noise_type; /* ccb_noise */
  int32_t ac_max_val; /* -1: no clamp; else clamp(y, -v, v+1) (coolchic.py:374-377) */
  float gain;         /* encoder_gain */
  float T_host, sigma_host;          /* used when the _dev pointer is NULL */
  const float* T_dev;                /* soft_round_temperature as a device scalar, or NULL */
  const float* sigma_dev;            /* noise_parameter as a device scalar, or NULL */
  const float* noise;                /* [B,M] raw draws the reference would make: N(0,1) for
                                        gaussian, U(0,1) for kumaraswamy; NULL => in-kernel Philox */
  uint64_t seed;                     /* Philox key when noise == NULL */
  const uint64_t* offset_dev;        /* Philox counter offset (device, advanced by ccb_tick) or NULL */
  uint64_t offset_host;
} ccb_quant_cfg;

typedef struct ccb_plan ccb_plan;

const char* ccb_strerror(int status);
const char* ccb_last_cuda_error(void);
int ccb_version(void);

/* Geometry of CoolChicEncoder.__init__ (coolchic.py:176-193).  `device` is the CUDA ordinal. */
int ccb_plan_create(ccb_plan** out, const ccb_arch* arch, int device);
void ccb_plan_destroy(ccb_plan* plan);
/* M = number of latent values per batch element; sizes/offsets of each latent in the flat [B,M] order. */
int64_t ccb_plan_n_latent_values(const ccb_plan* plan);
int ccb_plan_latent_geometry(const ccb_plan* plan, int32_t* h, int32_t* w, int64_t* offset);
/* Bytes of workspace one forward(+backward) needs. */
size_t ccb_plan_workspace_bytes(const ccb_plan* plan);
/* Number of floats in the weight-gradient partial-sum area (diagnostics). */
int64_t ccb_plan_n_weight_params(const ccb_plan* plan);
/* Number of kernel launches of one ccb_train_step (two-stream schedule, latent gradients on): what bench.py
 * reports as gpu_launches per step. */
int ccb_plan_train_step_launches(const ccb_plan* plan);

/* Constant-bank weight slots of the fused iteration.  The reference keeps one set of weights per nn.Module instance
 * (enc/component/coolchic.py:157-288), so any number of encoders can train side by side (warmup.py:93-129 trains 5);
 * here each concurrently trained encoder owns one slot: its ARM / synthesis weights are copied there (device to
 * device, stream ordered, graph capturable) before the kernels that read them as uniform-register operands.
 * ccb_weight_slot_acquire: *slot = a free slot index of `device`, or CCB_ERR_NOMEM when all are taken (the caller
 * then passes weight_slot = -1: same results from the kernels that read the weights from the tensors). */
int ccb_weight_slot_acquire(int device, int* slot);
int ccb_weight_slot_release(int device, int slot);

/* A private non-blocking CUDA stream on `device` (cudaStreamCreateWithFlags), for callers that train several encoders
 * side by side: frameworks hand out streams from a small shared pool (PyTorch: 32 per device, round robin), so two
 * "different" streams of two host threads can be the same cudaStream_t -- one thread's graph capture then breaks
 * the other's launches.  The reference needs none of this (one encoder at a time on the default stream). */
int ccb_stream_create(int device, void** stream_out);
int ccb_stream_destroy(int device, void* stream);

/* CoolChicEncoder.forward / LatentFreeCoolChicEncoder.forward
 * (coolchic.py:291-479, nlcoolchic.py:151-336): quantize (core/quantizer.py:116-232) -> context
 * gather + ARM (core/arm.py:227-370) -> Laplace rate (coolchic.py:420-445); upsampling
 * (core/upsampling.py:92-506) -> synthesis (core/synthesis.py:69-277).  raw_out [B,3,H,W], rate [B,M]; mu / scale /
 * log_scale / yhat [B,M] are optional (NULL to skip) detailed outputs. */
int ccb_forward(const ccb_plan* plan, const ccb_tensors* params, const ccb_quant_cfg* q, void* workspace,
                float* raw_out, float* rate, float* mu, float* scale, float* log_scale, float* yhat,
                void* stream);

/* Backward of the above (what loss.backward() does through it, train.py:264).
 * d_raw_out [B,3,H,W]; d_rate [B,M] or NULL with d_rate_uniform the same value for all.
 * grads->latent[i] may be NULL (no latent gradient wanted; the reference yields None
 * in "ste" mode, quantizer.py:222-225); all weight gradient pointers must be valid.
 * Gradients are WRITTEN (not accumulated).  `workspace` is the one ccb_forward filled. */
int ccb_backward(const ccb_plan* plan, const ccb_tensors* params, const ccb_quant_cfg* q, void* workspace,
                 const float* d_raw_out, const float* d_rate, float d_rate_uniform,
                 const ccb_tensors* grads, void* stream);

/* FrameEncoder clamp + loss_function, fused (frame.py:183, loss.py:135-150):
 * loss = mean((clamp(raw,0,1)-target)^2) + lambda * sum(rate)/B/(H*W).
 * Writes out[4] = {loss, mse, rate_bpp, 0} (device) and d_raw_out = dloss/draw_out.
 * dloss/drate is the uniform lambda/(B*H*W). */
int ccb_loss(const ccb_plan* plan, void* workspace, const float* raw_out, const float* rate,
             const float* target, float lambda, float* d_raw_out, float* out4, void* stream);

/* torch.nn.utils.clip_grad_norm_(all, max_norm, 2) followed by torch.optim.Adam.step()
 * (train.py:184,265-266; PyTorch defaults: no amsgrad, no weight decay).
 * The norm spans every tensor of `grads` with a non-NULL pointer; the update touches the
 * groups selected by optimize_mask (bit0 latents, bit1 arm, bit2 upsampling, bit3 synthesis;
 * train.py:165-182).  step_dev: device int32 holding the number of steps ALREADY done
 * (incremented by the call); lr_dev: device float.  norm_out (device float, may be NULL)
 * receives the total norm before clipping. */
#define CCB_OPT_LATENT 1
#define CCB_OPT_ARM 2
#define CCB_OPT_UPS 4
#define CCB_OPT_SYN 8
#define CCB_OPT_ALL 15
int ccb_clip_adam_step(const ccb_plan* plan, void* workspace, const ccb_tensors* params,
                       const ccb_tensors* grads, const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq,
                       int optimize_mask, float max_norm, const float* lr_dev, double beta1, double beta2,
                       double eps, int32_t* step_dev, float* norm_out, void* stream);

/* One iteration of train.py:244-268 on device-resident state: forward (train mode) ->
 * loss -> backward -> clip -> Adam.  out4 as in ccb_loss plus out4[3] = grad norm.
 * The Philox offset (q->offset_dev) is advanced so consecutive calls draw fresh noise.
 * aux_stream (may be NULL): a second stream on which the ARM branch runs concurrently with the
 * upsampling/synthesis chain (forked and joined with the plan's events, so the call is capturable in a
 * CUDA graph; host threads sharing one plan are serialised inside the call, while the work they enqueue on
 * different streams runs concurrently on the GPU as long as each caller has its own workspace and weight slot).
 * weight_slot: a slot from ccb_weight_slot_acquire (weights as constant-bank operands), or -1. */
int ccb_train_step(const ccb_plan* plan, const ccb_tensors* params, const ccb_tensors* grads,
                   const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, const ccb_quant_cfg* q,
                   void* workspace, const float* target, float lambda, int optimize_mask, float max_norm,
                   const float* lr_dev, double beta1, double beta2, double eps, int32_t* step_dev,
                   uint64_t* philox_offset_dev, float* out4, void* stream, void* aux_stream, int weight_slot);

/* Same iteration, with a CUDA event recorded on `stream` after every stage; the call synchronises and
 * writes the CCB_N_STAGES per-stage durations (ms) to stage_ms_host, in this order:
 * quantise, ARM+rate fwd, upsampling fwd, synthesis fwd, loss, synthesis bwd, upsampling bwd, ARM bwd,
 * weight-gradient reduction, quantiser bwd, clip+Adam.  Measurement aid for bench.py's roofline line. */
#define CCB_N_STAGES 11
int ccb_train_step_timed(const ccb_plan* plan, const ccb_tensors* params, const ccb_tensors* grads,
                         const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, const ccb_quant_cfg* q,
                         void* workspace, const float* target, float lambda, int optimize_mask, float max_norm,
                         const float* lr_dev, double beta1, double beta2, double eps, int32_t* step_dev,
                         uint64_t* philox_offset_dev, float* out4, void* stream, float* stage_ms_host,
                         int weight_slot);

/* Host-buffer entry point (the "encode one image" call): uploads the target image and the
 * flat parameter vector, runs n_iters iterations of ccb_train_step (captured once in a CUDA
 * graph), downloads parameters and the last out4.  params_host holds every tensor of
 * ccb_tensors back to back in declaration order (latents, arm w/b interleaved per layer,
 * ups, pre, syn w/b interleaved per layer); ccb_plan_n_param_floats() gives its length. */
int64_t ccb_plan_n_param_floats(const ccb_plan* plan);
int ccb_overfit_host(const ccb_plan* plan, float* params_host, const float* target_host,
                     const ccb_quant_cfg* q, float lambda, float lr, int optimize_mask, int n_iters,
                     float* out4_host);

/* Fixed-point ARM on the encoder side, all positions in parallel
 * (cpp/arm_cpu.cpp:18-106 == enc/bitstream/armint.py:81-261; indices: cpp/cc-contexts.h:20-48).
 * yhat_q: int32 [B,M] integer latents (NOT pre-scaled); w_q[k]: hidden [C,C] (out,in) / output
 * [2,C] int32 = round(W*256); b_q[k] = round(b*65536).  Outputs [B,M] int32:
 * mu_q8, log_scale_q8 and (optional, may be NULL) mu_rounded, mu_index, sig_index. */
int ccb_arm_int(const ccb_plan* plan, const int32_t* yhat_q, const int32_t* const* w_q,
                const int32_t* const* b_q, int32_t* mu_q8, int32_t* log_scale_q8, int32_t* mu_rounded,
                int32_t* mu_index, int32_t* sig_index, void* stream);

/* Fixed-point upsampling + synthesis of the decoder, evaluated on the encoder's GPU: the exact int32 planes (and
 * 8-bit pixels) the reference's C++ decoder produces from the same integers
 * (cpp/cc-frame-decoder.cpp:572-680 run_ups, :767-1152 run_syn; cpp/ups_refine_cpu.hpp:11-79;
 * cpp/ups_upsample_cpu.hpp:11-90; cpp/synfused_cpu.hpp:16-108; cpp/syn_cpu.hpp:21-112; cpp/ccdecapi.cpp:59-128).
 * latents: DEVICE int32 [M], finest level first, plain integers (as ccb_arm_int takes them).
 * ups_w_host [n_ups][ups_k], pre_w_host [n_pre][pre_k]: HOST int32 filters at 12 fractional bits, as the decoder holds
 * them after read_ups (cc-frame-decoder.cpp:254-295).  layers: HOST array, w / b are DEVICE int32
 * ([n_out][n_in][ks][ks] at 12 bits, [n_out] at 24 bits: read_syn, cc-frame-decoder.cpp:297-357); one branch.
 * ups_out: DEVICE [n_latents][H][W] (the planes handed to the synthesis) or NULL; out: DEVICE [3][H][W] at 12 bits;
 * pixels: DEVICE uint8 [H][W][3] or NULL.  Asynchronous on `stream`. */
typedef struct ccb_dec_layer {
  int32_t n_out, ks, residual, relu;
  const int32_t* w;
  const int32_t* b;
} ccb_dec_layer;
size_t ccb_dec_int_workspace_bytes(int H, int W, int n_latents, int n_syn, const ccb_dec_layer* layers);
int ccb_dec_int(int H, int W, int n_latents, const int32_t* latents, int ups_k, int n_ups, const int32_t* ups_w_host,
                int pre_k, int n_pre, const int32_t* pre_w_host, int n_syn, const ccb_dec_layer* layers,
                void* workspace, size_t workspace_bytes, int32_t* ups_out, int32_t* out, uint8_t* pixels, void* stream);

/* ccb_loss without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arithmetic as ccb_loss (frame clamp + mse + lambda * bpp, d_raw_out written; d_raw_out may be NULL), but no
 * plan and no workspace: the per-CTA partial sums live in a static device array inside the library
 * (csrc/ccb_noalloc.cu: 32.25 KB per GPU, held until the process exits).  The call itself performs no cudaMalloc /
 * cudaFree and calls no allocating routine; the caller passes inputs and outputs as device pointers.
 *   limit : B*3*H*W <= CCB_NOALLOC_MAX_BLOCKS * 4096 = 16 777 216 elements; larger sizes return CCB_ERR_TOO_LARGE
 *           and launch nothing.
 *   state : every call shares the array and finds the previous call's contents in it; what has to start at zero
 *           (the ticket counter) is cleared at the start of each call, on `stream`.
 *   races : calls that overlap on one GPU (different streams / threads) race on the array: keep them on one stream
 *           or order them with events.
 * n_latent_values = M (rate is [B, M]); out4 = {loss, mse, rate_bpp, 0}.  The first call on a device resolves the
 * array's device address (cudaGetSymbolAddress): make it once outside a graph capture. */
#define CCB_NOALLOC_MAX_BLOCKS 4096
int ccb_loss_noalloc(int device, int B, int H, int W, int64_t n_latent_values, const float* raw_out,
                     const float* rate, const float* target, float lambda, float* d_raw_out, float* out4,
                     void* stream);

/* FP32 FMA micro-benchmark (roofline denominator of the compute-bound kernels):
 * returns measured TFLOP/s in *tflops_host. */
int ccb_fp32_peak(int device, int n_iters, float* tflops_host);

#ifdef __cplusplus
}
#endif
#endif /* CCB_H */
