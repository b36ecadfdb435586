/* TEST INFRASTRUCTURE ONLY -- the product (coolchic_b200) never links, imports or executes this file.
 *
 * CPU restatement of the reference's fixed-point ARM, evaluated at
 * every position of a fully known latent grid (encoder side):
 *
 *   arm_int_grid       follows custom_conv_11_int32_cpu_X_X_X (coolchic/cpp/arm_cpu.cpp:18-106):
 *                      gather (:55-58), residual hidden layers with ReLU and >>8 rounding (:60-82),
 *                      2-output layer with symmetric rounding (:85-99); equivalently
 *                      ArmInt.forward(pure_int) (coolchic/enc/bitstream/armint.py:81-111, 244-261).
 *                      Context offsets follow get_context_indicies (cpp/cc-frame-decoder.cpp:111-153)
 *                      == _get_neighbor's mask order (coolchic/enc/component/core/arm.py:282-370).
 *                      The decoder keeps latents as value << 8 in a zero-padded frame (arm pad 4,
 *                      cc-frame-decoder.cpp:374,489), hence the zero reads outside the grid.
 *   mu_sig_indices     follows get_val_mu_indicies (coolchic/cpp/cc-contexts.h:20-48).
 *
 * Parity pins (tests/test_oracle_int.py): (1) ArmInt(pure_int=True) outputs generated from the imported
 * reference (tests/golden/int_*.npz, script tests/golden/make_int_golden.py); (2) the reference's own C++
 * compiled in place into oracle/_ref (oracle/build.sh): latents decoded from shipped .cool bitstreams are
 * re-encoded by the reference's cc_code_latent_layer_bac with THIS file's (mu, log-scale) and must give the
 * shipped sub-streams byte for byte; (3) ccref_mu_indices == mu_sig_indices on random inputs.
 *
 * All arithmetic is int32 with two's-complement wrap-around (done in uint32 to stay defined in C). */
#include <stdint.h>
#include <stdlib.h>

#define ARM_PRECISION 8
#define ARM_SCALE (1 << ARM_PRECISION)
#define N_MUQ 16
#define N_SIGQ 50
#define SIG_LOG_MIN (-1)
#define SIG_LOG_MAX_EXCL 9

static inline int32_t wadd(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static inline int32_t wmul(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }

/* (dy, dx) of the C causal neighbours, in the reference's order. */
static const signed char CTX8[8][2] = {{-3, 0}, {-2, 0}, {-1, -1}, {-1, 0}, {-1, 1}, {0, -3}, {0, -2}, {0, -1}};
static const signed char CTX16[16][2] = {{-3, 0}, {-3, 1}, {-2, -2}, {-2, -1}, {-2, 0}, {-2, 1}, {-2, 2}, {-1, -3},
                                         {-1, -2}, {-1, -1}, {-1, 0}, {-1, 1}, {-1, 2}, {0, -3}, {0, -2}, {0, -1}};
static const signed char CTX24[24][2] = {{-4, 0}, {-3, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-3, 2}, {-2, -3}, {-2, -2},
                                         {-2, -1}, {-2, 0}, {-2, 1}, {-2, 2}, {-2, 3}, {-1, -3}, {-1, -2}, {-1, -1},
                                         {-1, 0}, {-1, 1}, {-1, 2}, {-1, 3}, {0, -4}, {0, -3}, {0, -2}, {0, -1}};
static const signed char CTX32[32][2] = {
    {-4, -2}, {-4, -1}, {-4, 0}, {-4, 1}, {-3, -3}, {-3, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-3, 2}, {-3, 3},
    {-2, -3}, {-2, -2}, {-2, -1}, {-2, 0}, {-2, 1}, {-2, 2}, {-2, 3}, {-2, 4}, {-1, -4}, {-1, -3}, {-1, -2},
    {-1, -1}, {-1, 0}, {-1, 1}, {-1, 2}, {-1, 3}, {-1, 4}, {0, -4}, {0, -3}, {0, -2}, {0, -1}};

static const signed char (*ctx_table(int C))[2] {
  switch (C) {
    case 8: return CTX8;
    case 16: return CTX16;
    case 24: return CTX24;
    case 32: return CTX32;
  }
  return 0;
}

/* y: [H*W] integer latents; w_hidden: n_hidden x [C][C] in (out, in) order (the ArmIntLinear / state_dict
 * order; the C++ kernel is handed the transpose); b_hidden: n_hidden x [C] at scale 2^16; w_out [2][C];
 * b_out [2].  mu, ls: [H*W] outputs at scale 2^8.  Returns 0, or -1 for an unsupported C. */
int arm_int_grid(const int32_t* y, int H, int W, int C, int n_hidden, const int32_t* w_hidden,
                 const int32_t* b_hidden, const int32_t* w_out, const int32_t* b_out, int32_t* mu, int32_t* ls) {
  const signed char(*ctx)[2] = ctx_table(C);
  if (!ctx || n_hidden < 0) return -1;
  int32_t a[32], b[32];
  for (int py = 0; py < H; ++py)
    for (int px = 0; px < W; ++px) {
      int32_t* in = a;
      int32_t* out = b;
      for (int c = 0; c < C; ++c) {
        int yy = py + ctx[c][0], xx = px + ctx[c][1];
        int32_t v = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? y[(long)yy * W + xx] : 0;
        in[c] = (int32_t)((uint32_t)v << ARM_PRECISION);
      }
      for (int l = 0; l < n_hidden; ++l) {
        const int32_t* wl = w_hidden + (long)l * C * C;
        const int32_t* bl = b_hidden + (long)l * C;
        for (int o = 0; o < C; ++o) {
          int32_t acc = wadd(bl[o], wmul(in[o], ARM_SCALE));
          for (int i = 0; i < C; ++i) acc = wadd(acc, wmul(in[i], wl[o * C + i]));
          out[o] = acc < 0 ? 0 : (wadd(acc, ARM_SCALE / 2) >> ARM_PRECISION);
        }
        int32_t* t = in;
        in = out;
        out = t;
      }
      int32_t res[2];
      for (int o = 0; o < 2; ++o) {
        int32_t acc = b_out[o];
        for (int i = 0; i < C; ++i) acc = wadd(acc, wmul(in[i], w_out[o * C + i]));
        if (acc < 0)
          acc = -(wadd(-acc, ARM_SCALE / 2) >> ARM_PRECISION);
        else
          acc = wadd(acc, ARM_SCALE / 2) >> ARM_PRECISION;
        res[o] = acc;
      }
      mu[(long)py * W + px] = res[0];
      ls[(long)py * W + px] = res[1];
    }
  return 0;
}

void mu_sig_indices(const int32_t* mu, const int32_t* ls, long n, int32_t* mu_rounded, int32_t* mu_index,
                    int32_t* sig_index) {
  for (long i = 0; i < n; ++i) {
    int m = mu[i];
    int mr = m >= 0 ? ((m + ARM_SCALE / 2) >> ARM_PRECISION << ARM_PRECISION)
                    : -((-m + ARM_SCALE / 2) >> ARM_PRECISION << ARM_PRECISION);
    int mi = (m - mr) * N_MUQ;
    mi = mi >= 0 ? ((mi + ARM_SCALE / 2) >> ARM_PRECISION) : -((-mi + ARM_SCALE / 2) >> ARM_PRECISION);
    mi += N_MUQ / 2;
    int s = ls[i] - SIG_LOG_MIN * ARM_SCALE;
    int si;
    if (s < 0)
      si = 0;
    else {
      si = (s * (N_SIGQ / (SIG_LOG_MAX_EXCL - SIG_LOG_MIN)) + ARM_SCALE / 2) >> ARM_PRECISION;
      if (si >= N_SIGQ) si = N_SIGQ - 1;
    }
    mu_rounded[i] = mr >> ARM_PRECISION;
    mu_index[i] = mi;
    sig_index[i] = si;
  }
}
