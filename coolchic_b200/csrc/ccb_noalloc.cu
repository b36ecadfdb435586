// ccb_loss_noalloc: the frame clamp + loss_function of ccb_loss (ccb_opt.cu; enc/component/frame.py:183,
// enc/training/loss.py:135-150) for callers that cannot allocate: code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc.  No counterpart in the reference.
//
// ccb_loss keeps its per-CTA partial sums in the caller's plan workspace.  Here they live in a file-scope
// `static __device__` array of a fixed, modest size (CCB_NOALLOC_MAX_BLOCKS partial pairs + one ticket counter,
// 32.25 KB, held by every GPU that loads the library until the process exits).  The call allocates nothing: no
// cudaMalloc / cudaFree, no plan, no workspace, no allocating library routine.
//
//   * One CTA per 4096 image elements writes ONE partial pair, so the array bounds the problem size:
//     B*3*H*W <= CCB_NOALLOC_MAX_BLOCKS * 4096 = 16 777 216 elements (a 1080x1920 RGB frame is 6.2 M); anything larger
//     is refused with CCB_ERR_TOO_LARGE.
//   * In host code the array's NAME is a host-side stand-in, not a device address.  The device address is taken once
//     per device, with that device current, by cudaGetSymbolAddress (checked), and that pointer is what the kernel gets.
//   * Every call shares the array and finds the previous call's contents in it.  The partial pairs a call reads are
//     all written by that call first; the ticket counter must start at zero, so it is cleared (stream ordered) at the
//     start of every call, and the last CTA leaves it at zero as well.
//   * Calls that overlap on the GPU (two streams, two threads, two processes' worth of work on one device in one
//     process) RACE on the array: issue them on one stream, or order them with events.
#include <mutex>

#include "ccb_internal.cuh"

namespace {

constexpr int NA_THREADS = 256;
constexpr int NA_ITEMS = 16;                              // image elements per thread
constexpr int NA_PER_CTA = NA_THREADS * NA_ITEMS;         // 4096
constexpr int NA_MAX_BLOCKS = CCB_NOALLOC_MAX_BLOCKS;     // partial pairs the scratch holds
constexpr int NA_COUNTER = 2 * NA_MAX_BLOCKS;             // float index of the ticket counter (16-byte aligned)
constexpr int NA_SCRATCH_FLOATS = 2 * NA_MAX_BLOCKS + 64;

static __device__ __align__(256) float g_noalloc_scratch[NA_SCRATCH_FLOATS];

__global__ void __launch_bounds__(NA_THREADS) k_loss_noalloc(const float* __restrict__ raw, const float* __restrict__ tgt,
                                                            long long n_img, const float* __restrict__ rate,
                                                            long long n_rate, float inv_n_img, float inv_b_npix,
                                                            float lambda, float* __restrict__ d_raw,
                                                            float* __restrict__ out4, float* scratch) {
  __shared__ float red[32];
  __shared__ bool s_is_last;
  float se = 0.f, rs = 0.f;
  const long long base = (long long)blockIdx.x * NA_PER_CTA + threadIdx.x;
#pragma unroll
  for (int k = 0; k < NA_ITEMS; ++k) {
    const long long i = base + (long long)k * NA_THREADS;
    if (i < n_img) {
      const float x = raw[i];
      const float c = fminf(fmaxf(x, 0.f), 1.f);  // frame.py:183
      const float e = c - tgt[i];
      se = fmaf(e, e, se);
      if (d_raw) d_raw[i] = (x >= 0.f && x <= 1.f) ? 2.0f * e * inv_n_img : 0.f;
    }
  }
  for (long long i = (long long)blockIdx.x * NA_THREADS + threadIdx.x; i < n_rate; i += (long long)gridDim.x * NA_THREADS)
    rs += rate[i];
  const float a = block_sum(se, red);
  const float b = block_sum(rs, red);
  unsigned* counter = reinterpret_cast<unsigned*>(scratch + NA_COUNTER);
  if (threadIdx.x == 0) {
    scratch[2 * blockIdx.x] = a;
    scratch[2 * blockIdx.x + 1] = b;
    __threadfence();
    s_is_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_is_last) return;
  // the last CTA folds the partial pairs in a fixed order (deterministic) and writes the result
  __threadfence();
  float ts = 0.f, tr = 0.f;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += NA_THREADS) {
    ts += __ldcg(scratch + 2 * i);
    tr += __ldcg(scratch + 2 * i + 1);
  }
  const float tot_se = block_sum(ts, red);
  const float tot_rs = block_sum(tr, red);
  if (threadIdx.x == 0) {
    const float mse = tot_se * inv_n_img, bpp = tot_rs * inv_b_npix;
    out4[0] = mse + lambda * bpp;  // loss.py:135-150
    out4[1] = mse;
    out4[2] = bpp;
    out4[3] = 0.f;
    *counter = 0u;
  }
}

// device address of the scratch array, resolved once per device with that device current
constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mutex;
float* g_na_addr[NA_MAX_DEVICES];

int scratch_address(int device, float** out) {
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_addr[device]) {
    int cur = -1;
    CCB_CHECK(cudaGetDevice(&cur));
    if (cur != device) CCB_CHECK(cudaSetDevice(device));
    float* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (cur != device) cudaSetDevice(cur);
    CCB_CHECK(e);
    if (!ptr) return CCB_ERR_CUDA;
    g_na_addr[device] = ptr;
  }
  *out = g_na_addr[device];
  return CCB_OK;
}

}  // namespace

extern "C" int ccb_loss_noalloc(int device, int B, int H, int W, int64_t n_latent_values, const float* raw_out,
                                const float* rate, const float* target, float lambda, float* d_raw_out, float* out4,
                                void* stream) {
  if (device < 0 || device >= NA_MAX_DEVICES || B < 1 || H < 1 || W < 1 || n_latent_values < 0) return CCB_ERR_INVALID;
  if (!raw_out || !target || !out4 || (n_latent_values > 0 && !rate)) return CCB_ERR_INVALID;
  const long long n_img = (long long)B * 3 * H * W;
  const long long n_blocks = (n_img + NA_PER_CTA - 1) / NA_PER_CTA;
  if (n_blocks > NA_MAX_BLOCKS) return CCB_ERR_TOO_LARGE;  // beyond the static scratch: refused, nothing launched
  float* scratch = nullptr;
  int rc = scratch_address(device, &scratch);
  if (rc != CCB_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  // the ticket counter must start at zero whatever an earlier (possibly failed) call left behind
  CCB_CHECK(cudaMemsetAsync(scratch + NA_COUNTER, 0, sizeof(unsigned), s));
  const long long n_rate = (long long)B * n_latent_values;
  k_loss_noalloc<<<(unsigned)n_blocks, NA_THREADS, 0, s>>>(raw_out, target, n_img, rate, n_rate, 1.0f / (float)n_img,
                                                          1.0f / ((float)B * (float)H * (float)W), lambda, d_raw_out,
                                                          out4, scratch);
  CCB_LAUNCH_CHECK();
  return CCB_OK;
}
