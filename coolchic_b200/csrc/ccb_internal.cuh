// Internal declarations shared by the libccb translation units (sm_100a only).
#pragma once

#include <cuda.h>  // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint, no -lcuda)
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "ccb.h"

#define CCB_CHECK(call)                                   \
  do {                                                    \
    cudaError_t e__ = (call);                             \
    if (e__ != cudaSuccess) {                             \
      ccb_set_cuda_error(e__, __FILE__, __LINE__);        \
      return CCB_ERR_CUDA;                                \
    }                                                     \
  } while (0)

#define CCB_LAUNCH_CHECK() CCB_CHECK(cudaGetLastError())

void ccb_set_cuda_error(cudaError_t e, const char* file, int line);

// ------------------------------------------------------------------------------------------
// Geometry
// ------------------------------------------------------------------------------------------
struct LevelGeom {
  int h[CCB_MAX_LATENTS];
  int w[CCB_MAX_LATENTS];
  long long off[CCB_MAX_LATENTS];  // offset in the flat [M] order (resolution 0 first)
  long long M;
  int L;
  int B;
};

struct PtrTable7 {
  float* p[CCB_MAX_LATENTS];
};
struct CPtrTable7 {
  const float* p[CCB_MAX_LATENTS];
};

// One ARM tile: a TW x TH block of positions of one (batch, level).
struct ArmTile {
  int b, level, y0, x0;
};

// Arguments of the ARM kernels (ccb_arm.cu; shared by its per-slot translation units).
struct CcbArmArgs {
  const ArmTile* tiles;
  LevelGeom g;
  const float* yhat;
  const float* w[CCB_MAX_ARM_LAYERS];
  const float* b[CCB_MAX_ARM_LAYERS];
  long long wbs;  // weight batch stride (floats); 0: shared weights
  // forward outputs
  float* rate;
  float* mu;
  float* scale;
  float* log_scale;
  float* rate_part;
  // backward
  const float* d_rate;
  float d_rate_uniform;
  float* dyhat;   // accumulator of contributions whose receiver lies in the source tile's own row band
  float* dyhat_b; // accumulator of contributions to the row band above (receiver rows y0-4 .. y0-1)
  float* part;    // [gridDim.y * gridDim.x][n_arm]
  int tiles_per_group;  // tiles walked by one group of CTAs (blockIdx.y): all of them, or one batch element's
  int grid_x, grid_y;   // backward grid (ccb_arm_bwd_grid)
  int n_arm;
  int shared_sm;
  // split launches of the training step (shared weights only): this launch walks tiles [tile_lo, tile_hi) with all its
  // CTAs active and writes its partial rows from row part_row0 on; tile_hi == 0: the whole list
  int tile_lo, tile_hi, part_row0;
  // TMA staging of the tile + causal halo (zero outside the grid == arm.py:326-327 == TMA's out-of-bounds fill):
  // one 3-D tensor map (x, y, batch) per latent level over the flat y_hat vector; use_tma = 0: cp.async staging
  int use_tma;
  alignas(64) CUtensorMap tmap[CCB_MAX_LATENTS];
};

// box of the ARM tile maps: (32 * p + 8) x 12 x 1 floats; false when a level cannot be described (row pitch or base
// not 16-byte aligned): the caller then stages with cp.async
// Timing experiments only: CCB_DEBUG_SKIP=<bit mask> drops launches of the training step (1 ARM backward, 2 the fused
// small pyramid stages of the backward, 4 quantiser backward, 8 3x3 backward, 16 head backward, 32 the large pyramid
// backward stages): the step-time difference is that kernel's share of the critical path.  Results are wrong with it.
bool ccb_debug_skip(int bit);

bool ccb_arm_make_tmaps(const ccb_plan* p, const void* yhat, int elem_cols, CUtensorMap* out);

// Arguments of the synthesis kernels that exist in constant-bank form (ccb_syn.cu and its per-slot translation units).
struct CcbHeadArgs {
  const float* z;   // [B,7,H,W]
  float* out;       // [B,3,H,W]
  float* out2;      // optional second copy (kept for the ReLU mask of the backward)
  const float* w0;  // [S,7]
  const float* b0;  // [S]
  const float* w1;  // [3,S]
  const float* b1;  // [3]
  int S, relu1;
  long long HW;     // H*W; the batch element is blockIdx.y in every head kernel (CTAs never straddle two images)
  long long wbs;    // weight batch stride (floats); 0: shared weights
  // backward
  const float* d_out;  // [B,3,H,W] gradient w.r.t. the layer-1 OUTPUT (post activation)
  const float* act;    // [B,3,H,W] layer-1 output (for its ReLU mask), may be NULL if !relu1
  float* dz;           // [B,7,H,W]
  float* part;         // [gridDim.y * gridDim.x][n_head_params]
};

struct CcbTail2Args {
  const float* x;  // [B,3,H,W] input of the first 3x3 layer
  float* mid;      // [B,3,H,W] output of the first 3x3 layer (kept: the backward needs it)
  float* y;        // [B,3,H,W] output of the second layer, or NULL when only the loss is wanted
  int B, H, W, tiles_x, tiles_y, n_tiles;
  int res0, relu0, res1, relu1;
  const float* tgt;  // fused loss (NULL: none)
  float* d_raw;
  float* part;       // [n_part][2]
  int n_part;
  float inv_n_img;
  unsigned* tail_counter;
};

struct SynLayerGeom {
  int cin, cout, k, residual, relu;
  long long w_off, b_off;  // offsets (floats) inside the weight-parameter area
};

// Weight-parameter area (everything but the latents) as one flat index space:
//   [arm w0,b0,w1,b1,...][ups_w 0..L-2][pre_w 0..L-2][syn w0,b0,...]
struct WeightLayout {
  int n_arm_layers;  // hidden + output
  long long arm_w_off[CCB_MAX_ARM_LAYERS], arm_b_off[CCB_MAX_ARM_LAYERS];
  int arm_w_n[CCB_MAX_ARM_LAYERS], arm_b_n[CCB_MAX_ARM_LAYERS];
  long long ups_off[CCB_MAX_LATENTS], pre_off[CCB_MAX_LATENTS];
  int ups_n, pre_n;  // parameters per filter
  long long syn_w_off[CCB_MAX_SYN_LAYERS], syn_b_off[CCB_MAX_SYN_LAYERS];
  int syn_w_n[CCB_MAX_SYN_LAYERS], syn_b_n[CCB_MAX_SYN_LAYERS];
  long long n_arm, n_ups, n_syn, n_total;  // n_ups covers ups+pre
  long long arm_base, ups_base, syn_base;
};

struct ccb_plan {
  ccb_arch arch;
  int device;
  int n_sm;
  LevelGeom g;
  WeightLayout wl;
  SynLayerGeom syn[CCB_MAX_SYN_LAYERS];
  int syn_hidden;    // S: width of the first (1x1) synthesis layer
  int syn_head_out;  // channels after the second 1x1 layer
  int n_tail;        // k x k layers after the fused 1x1 pair
  // ARM tiling
  int arm_tw, arm_th;
  int n_arm_tiles;
  ArmTile* d_arm_tiles;
  // forward tile list: 32 * arm_fwd_p columns per tile (two latents per thread halve the weight reads per latent)
  int arm_fwd_p, n_arm_tiles_fwd;
  ArmTile* d_arm_tiles_fwd;
  int arm_tma;  // geometry admits TMA tile staging (every level width and M multiples of 4) and CCB_ARM_TMA != 0
  // workspace layout (bytes)
  size_t ws_bytes;
  size_t off_yhat;              // [B,M]
  size_t off_dyhat;             // [B,M]
  size_t off_dyhat2;            // [B,M] ARM share of d yhat: contributions of positions of the receiver's own tile band
  size_t off_dyhat3;            // [B,M] ARM share of d yhat: contributions of the tile band below the receiver
  cudaEvent_t ev_fork, ev_join, ev_head, ev_stage, ev_ups, ev_join2, ev_split, ev_big, ev_qfine, ev_start;
  // std::mutex held while ccb_train_step enqueues (or captures) one iteration: the fork / join events above belong to
  // the plan, so two host threads sharing a plan must not interleave their record / wait pairs
  void* step_mutex;
  // lazily created state of ccb_overfit_host (device arena: params|grads|m|v, target, workspace, scalars)
  char* host_arena = nullptr;
  size_t host_off_tgt = 0, host_off_ws = 0, host_off_small = 0;
  cudaStream_t host_s = nullptr, host_s2 = nullptr;
  // fork/join of the two-stream train step (makes ccb_train_step non-reentrant per plan)
  size_t off_up[CCB_MAX_LATENTS];   // up_i, i = 0..L-1 ; up_0 aliases yhat level L-1 (unused), up_{L-1} = z
  size_t off_dup[CCB_MAX_LATENTS];
  size_t off_act[CCB_MAX_SYN_LAYERS];   // output of synthesis layer i>=1 ([B,c,H,W]); layer 0 is never materialised
  size_t off_dact[2];           // ping-pong d activations [B,4,H,W]
  size_t off_rate_part;         // [n_arm_tiles] per-tile rate sums
  size_t off_dq;                // [B, M] quantiser derivative x gain, parked by the forward of the training step
  size_t off_arm_part;          // [n_arm_tiles][n_arm] partial ARM weight-gradient sums
  size_t off_syn_part[CCB_MAX_SYN_LAYERS];  // [0]: fused 1x1 pair, [l>=2]: 3x3 layer l
  size_t off_head_pack;         // [768] head weights in the constant bank's layout (staging buffer)
  size_t off_scalars;           // small float scratch (norm, loss partials, sum-of-squares partials)
  size_t off_ups_part;          // partial sums for filter gradients
  size_t off_draw, off_raw, off_rate;  // buffers of the fused train step
  int n_ups_part_rows;
  int n_loss_blocks;
};

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// ------------------------------------------------------------------------------------------
// Stage entry points (each returns a ccb_status)
// ------------------------------------------------------------------------------------------
// zero_ctr: two per-step counters (scalars[9], [10]) cleared by the kernel, or NULL
// dq != NULL (training step, ccb_quant_flat_ok): the forward also parks d y_hat / d latent there for the backward
// e_lo / e_hi (elements of the flat y_hat order, see ccb_quant_cut; both 0: everything): the part this launch covers
int ccb_quant_forward(const ccb_plan* p, const ccb_tensors* prm, const ccb_quant_cfg* q, float* yhat,
                      cudaStream_t s, unsigned* zero_ctr, float* dq = nullptr, long long e_lo = 0, long long e_hi = 0);
long long ccb_quant_cut(const ccb_plan* p, int level);
bool ccb_quant_flat_ok(const ccb_plan* p, const ccb_tensors* prm, const ccb_tensors* grads);
struct CcbTail;
// tail != NULL: the kernel also carries the step's finals (see CcbTail)
// d yhat = dyhat + dyhat2 + dyhat3 (the chain's share and the two ARM accumulators; NULL = absent)
int ccb_quant_backward(const ccb_plan* p, const ccb_tensors* prm, const ccb_quant_cfg* q, const float* dyhat,
                       const float* dyhat2, const float* dyhat3, const ccb_tensors* grads, cudaStream_t s,
                       CcbTail* tail = nullptr, int tail_mode = 1, const float* dq = nullptr, long long e_lo = 0,
                       long long e_hi = 0, int slot0 = 0, int* n_slots = nullptr);
int ccb_quant_grid(const ccb_plan* p);
// fills everything of `tail` except sq_part's producer; weight gradients of `grads` are listed in wg[]
int ccb_tail_prepare(const ccb_plan* p, char* ws, const ccb_tensors* grads, float max_norm, double beta1, double beta2,
                     int32_t* step_dev, float lambda, float* out4, uint64_t* philox_offset_dev, CcbTail* tail);
// tail != NULL (flat arenas only, see ccb_adam_flat_ok): the Adam kernel folds the partials itself in its prologue
int ccb_adam_only(const ccb_plan* p, char* ws, const ccb_tensors* params, const ccb_tensors* grads,
                  const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, int optimize_mask, const float* lr_dev,
                  double beta1, double beta2, double eps, cudaStream_t s, const CcbTail* tail = nullptr,
                  int n_sq_parts = 0);
bool ccb_adam_flat_ok(const ccb_plan* p, const ccb_tensors* params, const ccb_tensors* grads,
                      const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, int optimize_mask);
// tail_mode: 0 none, 1 ticket finals inside the kernel, 2 partials only (the Adam kernel folds them)

// Weight slots (CCB_N_WEIGHT_SLOTS per device): `slot` >= 0 selects the constant-bank kernels (C <= 16 / shipped head
// widths; the caller has acquired the slot), slot < 0 the kernels that read the weights from the tensors (no state
// outside the call; these also honour ccb_tensors.weight_batch_stride).
int ccb_arm_forward(const ccb_plan* p, const ccb_tensors* prm, const float* yhat, float* rate, float* mu,
                    float* scale, float* log_scale, float* rate_part, cudaStream_t s, int slot = -1,
                    bool weights_staged = false);
// dyhat_a / dyhat_b: the two zero-initialised [B,M] accumulators of the ARM share of d yhat (see k_arm_bwd)
int ccb_arm_backward(const ccb_plan* p, const ccb_tensors* prm, const float* yhat, const float* d_rate,
                     float d_rate_uniform, float* dyhat_a, float* dyhat_b, float* part, cudaStream_t s, int slot = -1,
                     bool weights_staged = false, float* rate_part = nullptr, bool shared_sm = false, int tile_lo = 0,
                     int tile_hi = 0, int part_row0 = 0);
// weights -> slot `slot` of the constant bank (the ARM kernels of C <= 16 read them there); forward / backward do it
// themselves unless told the bank is current
int ccb_arm_stage_weights(const ccb_plan* p, const ccb_tensors* prm, cudaStream_t s, int slot);
bool ccb_arm_const_ok(const ccb_plan* p);

// grid of the ARM backward: gx CTAs per batch group, gy groups (B with per-sample weights, else 1); every CTA owns
// one partial row [n_arm]
void ccb_arm_bwd_grid(const ccb_plan* p, bool per_sample, int* gx, int* gy);
int ccb_arm_bwd_max_rows(const ccb_plan* p);

// launches of the upsampling forward / backward (fused leading stages count once)
int ccb_ups_n_launches(const ccb_plan* p, bool backward);
int ccb_ups_forward(const ccb_plan* p, const ccb_tensors* prm, char* ws, cudaStream_t s,
                    cudaEvent_t wait_fine = nullptr);
int ccb_ups_chain_min_level(const ccb_plan* p);
int ccb_ups_fused_min_level(const ccb_plan* p);
// per-stage partial rows of the upsampling filter gradients: [n_rows][16], columns 0..3 = folded transposed-conv
// taps, 4..7 = folded refine taps (the symmetric halves are added by the producing CTA); reduced by
// ccb_reduce_syn_grads in the same launch as the synthesis partials
struct UpsPartInfo {
  const float* part[CCB_MAX_LATENTS];
  int n_rows[CCB_MAX_LATENTS];
  int n;
  int n_fused;  // stages 0 .. n_fused - 1 run in the one fused launch at the end, the others are launches of their own
};
// after_first / after_big: events recorded behind the first / the last of the separately launched (large) stages
int ccb_ups_backward(const ccb_plan* p, const ccb_tensors* prm, char* ws, const ccb_tensors* grads,
                     cudaStream_t s, UpsPartInfo* info, cudaEvent_t after_first = nullptr,
                     cudaEvent_t after_big = nullptr);

// Optional fusion of the frame clamp + squared error + its gradient into the last synthesis launch (training step,
// two trailing 3x3 layers): on return `done` tells whether the loss partials were produced (else call ccb_loss_image).
struct SynLossFuse {
  const float* target;
  float* d_raw;
  float* part;  // [n_part][2] (squared error, rate) partial sums
  int n_part;
  float inv_n_img;
  unsigned* tail_counter;
  int done;
};
int ccb_syn_forward(const ccb_plan* p, const ccb_tensors* prm, char* ws, float* raw_out, cudaStream_t s,
                    int slot = -1, bool weights_staged = false, SynLossFuse* loss = nullptr);
void ccb_loss_fuse_args(const ccb_plan* p, char* ws, const float* target, float* d_raw_out, SynLossFuse* out);
// head weights -> the constant bank (read there by the head kernels of the shipped widths)
int ccb_syn_stage_weights(const ccb_plan* p, const ccb_tensors* prm, char* ws, cudaStream_t s, int slot);
// dw_stream != NULL: only d z is computed on `s`; the head's weight gradients run on dw_stream (after an event)
int ccb_syn_backward(const ccb_plan* p, const ccb_tensors* prm, char* ws, const float* d_raw_out,
                     cudaStream_t s, cudaStream_t dw_stream, int slot = -1, bool weights_staged = false);
// grid of the head backward (persistent, batch-pure CTAs): gx CTAs per batch element, gy = B
void ccb_syn_head_grid(const ccb_plan* p, int* gx, int* gy);

// number of CTAs (= weight-square partials) of the three reduction launches, in wsq order: arm | syn | upsampling
int ccb_wsq_layout(const ccb_plan* p, int* arm_base, int* syn_base, int* ups_base);
// grads->weight_batch_stride > 0: the partial rows are summed per batch element (per-sample gradients)
int ccb_reduce_arm_grads(const ccb_plan* p, char* ws, const ccb_tensors* grads, cudaStream_t s);
int ccb_reduce_syn_grads(const ccb_plan* p, char* ws, const ccb_tensors* grads, cudaStream_t s, const UpsPartInfo* ups,
                         int which = 7);
// loss in two halves: image part (squared error partials + d_raw_out), then the final combine with the
// per-tile rate sums the ARM forward left in the workspace
int ccb_loss_image(const ccb_plan* p, char* ws, const float* raw_out, const float* target, float* d_raw_out,
                   cudaStream_t s);
int ccb_tail_step(const ccb_plan* p, char* ws, const ccb_tensors* params, const ccb_tensors* grads,
                  const ccb_tensors* exp_avg, const ccb_tensors* exp_avg_sq, int optimize_mask, float max_norm,
                  const float* lr_dev, double beta1, double beta2, double eps, int32_t* step_dev, float lambda,
                  float* out4, uint64_t* philox_offset_dev, cudaStream_t s);
int ccb_loss_finish(const ccb_plan* p, char* ws, float lambda, float* out4, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// Device helpers
// ------------------------------------------------------------------------------------------
#ifdef __CUDACC__

// Asynchronous 4-byte global -> shared copy (LDGSTS); !valid zero-fills.  All copies of a tile are in
// flight at once: a plain load/store loop serialises on the global latency (19 round trips for a 70x70 tile).
__device__ __forceinline__ void cp_async4(float* sdst, const float* gsrc, bool valid) {
  const unsigned sa = static_cast<unsigned>(__cvta_generic_to_shared(sdst));
  const int sz = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(sa), "l"(gsrc), "r"(sz));
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// Stage a [n_rows x TW] tile of a row-major [H, W] plane into shared memory with cp.async: warps stride over the
// rows, lanes over the columns, so the row pointer / row validity are warp-uniform and the column index, clamp and
// validity of a lane are computed once (a flat `i -> (row, col)` loop spends ~20 instructions per element on
// index arithmetic, more than the stencil itself).  CLAMP: replicate-clamped coordinates; otherwise zero fill
// outside the plane.  The caller issues cp_async_wait_all() + __syncthreads().
template <int TW, bool CLAMP>
__device__ __forceinline__ void stage_tile(float* sm, int sm_stride, const float* __restrict__ g, int H, int W, int y0,
                                           int x0, int n_rows) {
  constexpr int NS = (TW + 31) / 32;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  int gx[NS];
  bool vx[NS];
#pragma unroll
  for (int k = 0; k < NS; ++k) {
    const int rx = lane + 32 * k, x = x0 + rx;
    vx[k] = (rx < TW) && (CLAMP || (x >= 0 && x < W));
    gx[k] = min(max(x, 0), W - 1);
  }
  for (int ry = wid; ry < n_rows; ry += nw) {
    const int y = y0 + ry;
    const bool vy = CLAMP || (y >= 0 && y < H);
    const float* grow = g + (long long)min(max(y, 0), H - 1) * W;
    float* srow = sm + ry * sm_stride;
#pragma unroll
    for (int k = 0; k < NS; ++k) {
      const int rx = lane + 32 * k;
      if (rx < TW) cp_async4(srow + rx, grow + gx[k], vy && vx[k]);
    }
  }
}

// ---- TMA (cp.async.bulk.tensor) + mbarrier helpers --------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// blocks until the phase with parity `parity` of `bar` has completed (try_wait suspends the thread in hardware)
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  const unsigned addr = smem_u32(bar);
  unsigned done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!done);
}
// one box of a 3-D tensor map -> shared memory; out-of-bounds elements arrive as zeros; completes on `bar`
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, int x, int y, int z, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
      : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over the block; result valid in thread 0.  `sm` must hold >= 32 floats.
__device__ __forceinline__ float block_sum(float v, float* sm) {
  v = warp_sum(v);
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sm[wid] = v;
  __syncthreads();
  int nw = (blockDim.x + 31) >> 5;
  float r = 0.f;
  if (wid == 0) {
    r = lane < nw ? sm[lane] : 0.f;
    r = warp_sum(r);
  }
  return r;
}

// Finals of the fused iteration, executed by whichever CTA of the carrying kernel finishes last (atomic ticket):
// total gradient norm / clip coefficient / Adam step and bias corrections, the loss triple and the Philox advance.
// The carrier contributes one sum-of-squares partial per CTA (summed in index order: deterministic); weight
// gradients listed in wg[] are squared here (a few thousand values).
struct CcbTail {
  float* sq_part;             // [gridDim.x of the carrier]
  unsigned* counter;          // zeroed by k_loss_partial earlier in the same step
  const float* loss_part;     // [n_loss_blocks][2]
  int n_loss_blocks;
  const float* rate_part;     // per-tile rate sums of the ARM forward
  int n_rate_part;
  const float* wsq_part;      // sums of squares of the weight gradients, one per CTA of the reduction kernels
  int n_wsq;
  float inv_n_img, inv_b_npix, lambda, max_norm;
  float* out4;                // loss, mse, bpp, norm
  float* scal;                // [0] norm [1] clip coefficient [2] 1/bc1 [3] 1/sqrt(bc2)
  int* step;
  double beta1, beta2;
  unsigned long long* philox_offset;
};

// Adam step counter and bias corrections: ONE thread of the carrier calls this at kernel start (it only depends on
// the previous step count, so it overlaps the kernel body instead of sitting on the serial tail).
// torch.optim.Adam computes the corrections in Python floats (double); beta^st by repeated squaring.
__device__ __forceinline__ void ccb_tail_bias(const CcbTail& a) {
  const int st = *a.step + 1;
  *a.step = st;
  double p1 = 1.0, p2 = 1.0, b1 = a.beta1, b2 = a.beta2;
  for (int e = st; e > 0; e >>= 1) {
    if (e & 1) {
      p1 *= b1;
      p2 *= b2;
    }
    b1 *= b1;
    b2 *= b2;
  }
  a.scal[2] = (float)(1.0 / (1.0 - p1));
  a.scal[3] = (float)(1.0 / sqrt(1.0 - p2));
}

// Block-wide fold of the step's partial arrays (all threads of the CTA call it; every thread gets the totals):
// sum of squared gradients (n_sq carrier partials + weight partials), squared error, rate.  `red` >= 96 floats.
__device__ __forceinline__ void ccb_tail_reduce(const CcbTail& a, int n_sq, float* red, float& sq_o, float& se_o,
                                                float& rs_o) {
  // every partial array is walked in one pass with independent accumulators: the loads overlap
  float sq = 0.f, se = 0.f, rs = 0.f;
  const int n0 = n_sq, n1 = a.n_wsq, n2 = a.n_loss_blocks, n3 = a.n_rate_part;
  const int nmax = max(max(n0, n1), max(n2, n3));
  for (int i0 = threadIdx.x; i0 < nmax; i0 += 4 * blockDim.x) {
    float x[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {  // 16 independent loads in flight per thread: the fold is pure L2 latency
      const int i = i0 + u * blockDim.x;
      x[u][0] = i < n0 ? __ldcg(a.sq_part + i) : 0.f;
      x[u][1] = i < n1 ? __ldcg(a.wsq_part + i) : 0.f;
      x[u][2] = i < n2 ? __ldcg(a.loss_part + 2 * i) : 0.f;
      x[u][3] = i < n3 ? __ldcg(a.rate_part + i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      sq += x[u][0] + x[u][1];
      se += x[u][2];
      rs += x[u][3];
    }
  }
  sq = warp_sum(sq);
  se = warp_sum(se);
  rs = warp_sum(rs);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) {
    red[wid] = sq;
    red[32 + wid] = se;
    red[64 + wid] = rs;
  }
  __syncthreads();
  sq = se = rs = 0.f;
  for (int k = 0; k < nw; ++k) {  // same order in every thread and every CTA: identical results everywhere
    sq += red[k];
    se += red[32 + k];
    rs += red[64 + k];
  }
  sq_o = sq;
  se_o = se;
  rs_o = rs;
}

// norm / clip coefficient / loss triple / Philox advance from the totals (ONE thread of the grid)
__device__ __forceinline__ void ccb_tail_write(const CcbTail& a, float sq, float se, float rs) {
  const float norm = sqrtf(sq);
  a.scal[0] = norm;
  a.scal[1] = fminf(a.max_norm / (norm + 1e-6f), 1.0f);  // clip_grad_norm_: clamp(max_norm/(norm+1e-6), max=1)
  const float mse = se * a.inv_n_img, bpp = rs * a.inv_b_npix;
  a.out4[0] = mse + a.lambda * bpp;
  a.out4[1] = mse;
  a.out4[2] = bpp;
  a.out4[3] = norm;
  if (a.philox_offset) *a.philox_offset += 1ull;
}

// Ticket variant: called by ALL threads of every CTA of the carrier, after its last global write; the CTA that
// finishes last folds and writes.  `red` >= 96 floats of smem.
__device__ __forceinline__ void ccb_tail_finals(const CcbTail& a, float my_sq, float* red) {
  __shared__ bool s_is_last;
  const float v = block_sum(my_sq, red);
  if (threadIdx.x == 0) {
    a.sq_part[blockIdx.x] = v;
    __threadfence();
    s_is_last = (atomicAdd(a.counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_is_last) return;
  __threadfence();
  float sq, se, rs;
  ccb_tail_reduce(a, (int)gridDim.x, red, sq, se, rs);
  if (threadIdx.x == 0) {
    ccb_tail_write(a, sq, se, rs);
    *a.counter = 0u;
  }
}

// Philox4x32-10 (Salmon et al. 2011), counter = (ctr_lo, ctr_hi, 0, 0), key = seed.
__device__ __forceinline__ uint4 philox4x32_10(uint64_t ctr, uint64_t ctr2, uint64_t seed) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = (uint32_t)ctr2, c3 = (uint32_t)(ctr2 >> 32);
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ float u32_to_unit_open(uint32_t x) {
  // (0,1): (x + 0.5) / 2^32
  return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f);
}

#endif  // __CUDACC__
