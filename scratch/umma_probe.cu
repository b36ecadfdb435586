// scratch/umma_probe.cu -- which shared-memory float does tcgen05.mma.kind::tf32 read for operand element (mn, k)?
//
// One CTA, one MMA (M = 128, N = 16, K = 8).  The probed operand's image holds image[f] = f + 1 (exact in TF32 up to
// 2048), the other operand is "one-hot in k" (1 where k == k0, through the K-major no-swizzle layout that the forward
// pass of scratch/umma_bench.cu already proved), so D[m][n] = probed(m or n, k0) = 1 + index of the float that was read.
// Eight MMAs (k0 = 0..7) decode the whole address map of a descriptor (major-ness, leading / stride byte offsets).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o scratch/umma_probe scratch/umma_probe.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <vector>

constexpr int IMG_FLOATS = 2048;  // 8 KB per operand image

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}

struct Probe {
  uint32_t idesc, a_lbo, a_sbo, b_lbo, b_sbo;
};

__global__ void __launch_bounds__(128) k_probe(const float* a_img, const float* b_img, Probe pr, float* d_out) {
  __shared__ __align__(1024) float sa[IMG_FLOATS];
  __shared__ __align__(1024) float sb[IMG_FLOATS];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < IMG_FLOATS; i += 128) {
    sa[i] = a_img[i];
    sb[i] = b_img[i];
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(smem_u32(&s_tmem)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;
  if (tid == 0) {
    const uint64_t da = make_desc(smem_u32(sa), pr.a_lbo, pr.a_sbo), db = make_desc(smem_u32(sb), pr.b_lbo, pr.b_sbo);
    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem),
                 "l"(da), "l"(db), "r"(pr.idesc), "r"(0u)
                 : "memory");
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
  }
  for (unsigned spin = 0; spin < (1u << 22); ++spin) {
    unsigned ok;
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok)
                 : "r"(smem_u32(&bar))
                 : "memory");
    if (ok) break;
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(tmem + ((uint32_t)(warp * 32) << 16)));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int i = 0; i < 16; ++i) d_out[tid * 16 + i] = __uint_as_float(r[i]);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tmem) : "memory");
}

constexpr uint32_t make_idesc(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

static float *d_a, *d_b, *d_d;
static std::vector<float> run(const std::vector<float>& a, const std::vector<float>& b, Probe pr) {
  cudaMemcpy(d_a, a.data(), IMG_FLOATS * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_b, b.data(), IMG_FLOATS * 4, cudaMemcpyHostToDevice);
  cudaMemset(d_d, 0, 128 * 16 * 4);
  k_probe<<<1, 128>>>(d_a, d_b, pr, d_d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
  std::vector<float> d(128 * 16);
  cudaMemcpy(d.data(), d_d, d.size() * 4, cudaMemcpyDeviceToHost);
  return d;
}

// the proven K-major no-swizzle image: element (row, k) at (row / 8) * sbo + (k / 4) * lbo + (row % 8) * 16 + (k % 4) * 4
static std::vector<float> one_hot_kmajor(int rows, int k0, int lbo, int sbo) {
  std::vector<float> v(IMG_FLOATS, 0.f);
  for (int r = 0; r < rows; ++r) v[((r / 8) * sbo + (k0 / 4) * lbo + (r % 8) * 16 + (k0 % 4) * 4) / 4] = 1.f;
  return v;
}

int main() {
  cudaMalloc(&d_a, IMG_FLOATS * 4);
  cudaMalloc(&d_b, IMG_FLOATS * 4);
  cudaMalloc(&d_d, 128 * 16 * 4);
  std::vector<float> ramp(IMG_FLOATS);
  for (int i = 0; i < IMG_FLOATS; ++i) ramp[i] = (float)(i + 1);
  const int strides[3][2] = {{128, 256}, {256, 128}, {128, 128}};
  for (int which = 0; which < 2; ++which)  // 0: probe A (rows m), 1: probe B (columns n)
    for (int major = 0; major < 2; ++major)
      for (int sv = 0; sv < 3; ++sv) {
        const int lbo = strides[sv][0], sbo = strides[sv][1];
        printf("\n== operand %c, %s-major, LBO %d, SBO %d: byte offset read for (index, k)\n", which ? 'B' : 'A',
               major ? "MN" : "K", lbo, sbo);
        std::vector<std::vector<float>> res;
        for (int k0 = 0; k0 < 8; ++k0) {
          Probe pr;
          if (which == 0) {
            pr = {make_idesc(128, 16, major, 0), (uint32_t)lbo, (uint32_t)sbo, 128u, 256u};
            res.push_back(run(ramp, one_hot_kmajor(16, k0, 128, 256), pr));
          } else {
            pr = {make_idesc(128, 16, 0, major), 128u, 256u, (uint32_t)lbo, (uint32_t)sbo};
            std::vector<float> a = one_hot_kmajor(128, k0, 128, 256);
            res.push_back(run(a, ramp, pr));
          }
        }
        const int idx[] = {0, 1, 2, 3, 4, 5, 7, 8, 9, 12, 15, 16, 17, 32, 64, 127};
        for (int ii = 0; ii < 16; ++ii) {
          const int i = idx[ii];
          if (which == 1 && i >= 16) continue;
          printf("  %c=%3d:", which ? 'n' : 'm', i);
          for (int k0 = 0; k0 < 8; ++k0) {
            const float v = which ? res[k0][0 * 16 + i] : res[k0][i * 16 + 0];
            printf(" %6d", (int)(v - 1.f) * 4);
          }
          printf("\n");
        }
      }
  return 0;
}
