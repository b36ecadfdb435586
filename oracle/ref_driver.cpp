// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or shipped with coolchic_b200.
//
// Thin C driver around the UNMODIFIED reference decoder sources, compiled where they lie under
// /root/reference/coolchic/cpp by oracle/build.sh into oracle/_ref/libccref.so.  No reference source
// is copied: this file only #includes the reference headers and calls the reference functions.
//
// What it exposes (extern "C", used by tests/ and tests/golden/make_int_golden.py only):
//   * ccref_open / ccref_close      parse a .cool bitstream (cc-bitstream.cpp), decode the integer ARM weights
//                                   (cc_frame_decoder::read_arm, cc-frame-decoder.cpp:195-250) and run the
//                                   reference's sequential ARM + CABAC decode of every latent grid
//                                   (cc_frame_decoder::run_arm, cc-frame-decoder.cpp:455-570, which calls
//                                   custom_conv_11_int32_cpu_X_X_X, arm_cpu.cpp:18-106).
//   * ccref_header / ccref_latent / ccref_arm_weights / ccref_layer_bytes   read the results back.
//   * ccref_decode_layer            run arm_cpu.cpp's kernel on a caller-supplied sub-stream + weights.
//   * ccref_mu_indices              get_val_mu_indicies (cc-contexts.h:20-48) on arrays.
//   * ccref_dec_header / ccref_ups_weights / ccref_syn_weights / ccref_run_ups / ccref_run_syn
//                                   the fixed-point upsampling + synthesis of the opened frame (run_ups / run_syn).
//
// cc_frame_decoder keeps its state private; a test harness has to look at it, so the class is
// included with `private` opened up (layout is unchanged).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "common.h"
#include "TDecBinCoderCABAC.h"
#include "cc-contexts.h"
#include "cc-bac.h"
#include "cc-bitstream.h"
#define private public
#include "cc-frame-decoder.h"
#undef private
#include "arm_cpu.h"

// timing globals the reference decoder expects its application to define (ccdecapi.cpp)
float time_arm_seconds = 0;
float time_ups_seconds = 0;
float time_syn_seconds = 0;
float time_blend_seconds = 0;
float time_warp_seconds = 0;
float time_bpred_seconds = 0;
float time_all_seconds = 0;

int32_t* get_context_indicies(int n_contexts, int stride);  // cc-frame-decoder.cpp:111

namespace {
struct Handle {
  cc_bs bs;
  cc_bs_frame* frame = nullptr;
  cc_frame_decoder* dec = nullptr;
};
}  // namespace

extern "C" {

void* ccref_open(const char* path) {
  Handle* h = new Handle();
  if (!h->bs.open(std::string(path), 0)) {
    delete h;
    return nullptr;
  }
  h->frame = h->bs.decode_frame(0);
  if (h->frame == nullptr) {
    delete h;
    return nullptr;
  }
  h->dec = new cc_frame_decoder(h->bs.m_gop_header, 0, 0, 0);
  h->dec->read_arm(h->frame);
  h->dec->read_ups(h->frame);
  h->dec->read_syn(h->frame);
  h->dec->check_allocations(h->frame);
  h->dec->run_arm(h->frame);
  return h;
}

void ccref_close(void* hv) {
  Handle* h = (Handle*)hv;
  if (!h) return;
  delete h->dec;
  delete h;
}

// out[0..7] = img_h, img_w, n_latents, dim_arm, n_hidden_layers_arm, hls_sig_blksize, ac_max_val_latent, n_ft
int ccref_header(void* hv, int* out) {
  Handle* h = (Handle*)hv;
  auto& fh = h->frame->m_frame_header;
  out[0] = h->bs.m_gop_header.img_h;
  out[1] = h->bs.m_gop_header.img_w;
  out[2] = fh.n_latent_n_resolutions;
  out[3] = fh.dim_arm;
  out[4] = fh.n_hidden_layers_arm;
  out[5] = fh.hls_sig_blksize;
  out[6] = fh.ac_max_val_latent;
  out[7] = fh.latent_n_2d_grid;
  return 0;
}

int ccref_layer_dims(void* hv, int layer, int* hw, int* n_bytes) {
  Handle* h = (Handle*)hv;
  hw[0] = h->dec->m_h_pyramid[layer];
  hw[1] = h->dec->m_w_pyramid[layer];
  *n_bytes = (int)h->frame->m_latents_hevc[layer].size();
  return 0;
}

// decoded latent grid as plain integers (the decoder stores value << ARM_PRECISION)
int ccref_latent(void* hv, int layer, int32_t* out) {
  Handle* h = (Handle*)hv;
  int H = h->dec->m_h_pyramid[layer], W = h->dec->m_w_pyramid[layer];
  if (h->dec->m_zero_layer[layer]) {
    memset(out, 0, sizeof(int32_t) * H * W);
    return 1;
  }
  frame_memory& fm = h->dec->m_plane_pyramid[layer];
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) out[y * W + x] = fm.plane_origin(0)[y * fm.stride + x] >> ARM_PRECISION;
  return 0;
}

int ccref_layer_bytes(void* hv, int layer, unsigned char* out) {
  Handle* h = (Handle*)hv;
  auto& b = h->frame->m_latents_hevc[layer];
  if (!b.empty()) memcpy(out, b.data(), b.size());
  return (int)b.size();
}

// hidden weights are returned in the layout the reference kernel consumes: TRANSPOSED [in][out]
// (cc-frame-decoder.cpp:237-246); output weights [2][C]; biases at 2*ARM_PRECISION.
int ccref_arm_weights(void* hv, int32_t* w_hidden_t, int32_t* b_hidden, int32_t* w_out, int32_t* b_out) {
  Handle* h = (Handle*)hv;
  auto& fh = h->frame->m_frame_header;
  int C = fh.dim_arm;
  for (int l = 0; l < fh.n_hidden_layers_arm; ++l) {
    memcpy(w_hidden_t + (size_t)l * C * C, h->dec->m_mlpw_t[l].data, sizeof(int32_t) * C * C);
    memcpy(b_hidden + (size_t)l * C, h->dec->m_mlpb[l].data, sizeof(int32_t) * C);
  }
  memcpy(w_out, h->dec->m_mlpwOUT.data, sizeof(int32_t) * 2 * C);
  memcpy(b_out, h->dec->m_mlpbOUT.data, sizeof(int32_t) * 2);
  return 0;
}

// Run the reference's sequential integer ARM + CABAC decode (arm_cpu.cpp:18-106) over one latent
// sub-stream with caller-supplied integer weights.  out = decoded integers [H*W].
int ccref_decode_layer(const unsigned char* bytes, int n_bytes, int H, int W, int C, int n_hidden,
                       const int32_t* w_hidden_t, const int32_t* b_hidden, const int32_t* w_out,
                       const int32_t* b_out, int hls_sig_blksize, int32_t* out) {
  const int pad = 4;
  frame_memory fm;
  fm.update_to(H, W, pad, 1);
  memset(fm.raw(), 0, sizeof(int32_t) * fm.plane_stride);
  std::vector<weights_biases> wt(n_hidden > 0 ? n_hidden : 1), bh(n_hidden > 0 ? n_hidden : 1);
  for (int l = 0; l < n_hidden; ++l) {
    wt[l].update_to(C * C);
    memcpy(wt[l].data, w_hidden_t + (size_t)l * C * C, sizeof(int32_t) * C * C);
    bh[l].update_to(C);
    memcpy(bh[l].data, b_hidden + (size_t)l * C, sizeof(int32_t) * C);
  }
  weights_biases wo, bo;
  wo.update_to(2 * C);
  memcpy(wo.data, w_out, sizeof(int32_t) * 2 * C);
  bo.update_to(2);
  memcpy(bo.data, b_out, sizeof(int32_t) * 2);

  InputBitstream bs;
  bs.getFifo().assign(bytes, bytes + n_bytes);
  TDecBinCABAC cabac;
  cabac.init(&bs);
  cabac.start();
  BACContext ctx;
  ctx.set_layer(&cabac, H, W, hls_sig_blksize);
  custom_conv_11_int32_cpu_X_X_X(wt.data(), bh.data(), &wo, &bo, get_context_indicies(C, fm.stride), C, n_hidden,
                                 fm.origin(), H, W, (fm.stride - W) / 2, ctx);
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) out[y * W + x] = fm.plane_origin(0)[y * fm.stride + x] >> ARM_PRECISION;
  return 0;
}

// get_val_mu_indicies (cc-contexts.h:20-48) over arrays
int ccref_mu_indices(const int32_t* mu, const int32_t* log_sig, int n, int32_t* mu_rounded, int32_t* mu_index,
                     int32_t* sig_index) {
  for (int i = 0; i < n; ++i) {
    int a, b, c;
    get_val_mu_indicies(mu[i], log_sig[i], a, b, c);
    mu_rounded[i] = a;
    mu_index[i] = b;
    sig_index[i] = c;
  }
  return 0;
}

// ---- fixed-point upsampling + synthesis: the reference's own run_ups / run_syn on the opened frame
// out[0..5] = ups_k_size, ups_preconcat_k_size, n_ups_kernel, n_ups_preconcat_kernel, n_syn_layers, n_syn_branches
// then per synthesis layer (up to 8): n_out_ft, ks, residual, relu
int ccref_dec_header(void* hv, int* out) {
  Handle* h = (Handle*)hv;
  auto& fh = h->frame->m_frame_header;
  out[0] = fh.ups_k_size;
  out[1] = fh.ups_preconcat_k_size;
  out[2] = fh.n_ups_kernel;
  out[3] = fh.n_ups_preconcat_kernel;
  out[4] = fh.n_syn_layers;
  out[5] = fh.n_syn_branches;
  for (int i = 0; i < fh.n_syn_layers && i < 8; ++i) {
    out[6 + 4 * i] = fh.layers_synthesis[i].n_out_ft;
    out[7 + 4 * i] = fh.layers_synthesis[i].ks;
    out[8 + 4 * i] = fh.layers_synthesis[i].residual ? 1 : 0;
    out[9 + 4 * i] = fh.layers_synthesis[i].relu ? 1 : 0;
  }
  return 0;
}

// integer filters as the decoder holds them (UPS_PRECISION): ups [n_ups][k], preconcat [n_pre][kp]
int ccref_ups_weights(void* hv, int32_t* ups, int32_t* pre) {
  Handle* h = (Handle*)hv;
  auto& fh = h->frame->m_frame_header;
  for (int l = 0; l < h->dec->m_ups_n; ++l)
    memcpy(ups + (size_t)l * fh.ups_k_size, h->dec->m_upsw[l].data, sizeof(int32_t) * fh.ups_k_size);
  for (int l = 0; l < h->dec->m_ups_n_preconcat; ++l)
    memcpy(pre + (size_t)l * fh.ups_preconcat_k_size, h->dec->m_upsw_preconcat[l].data,
           sizeof(int32_t) * fh.ups_preconcat_k_size);
  return 0;
}

// synthesis layer `idx` of branch 0: weights [n_out][n_in][ks][ks] (SYN_WEIGHT_PRECISION), biases [n_out] (2x)
int ccref_syn_weights(void* hv, int idx, int n_w, int32_t* w, int n_b, int32_t* b) {
  Handle* h = (Handle*)hv;
  memcpy(w, h->dec->m_synw[idx].data, sizeof(int32_t) * n_w);
  memcpy(b, h->dec->m_synb[idx].data, sizeof(int32_t) * n_b);
  return 0;
}

// run_ups (cc-frame-decoder.cpp:572-680): ups_out = the n_latents planes handed to the synthesis [n][H][W]
int ccref_run_ups(void* hv, int32_t* ups_out) {
  Handle* h = (Handle*)hv;
  h->dec->run_ups(h->frame);
  const int H = h->bs.m_gop_header.img_h, W = h->bs.m_gop_header.img_w;
  frame_memory& fm = h->dec->m_syn_1;
  const int n = h->frame->m_frame_header.n_latent_n_resolutions;
  for (int p = 0; p < n; ++p)
    for (int y = 0; y < H; ++y)
      memcpy(ups_out + ((size_t)p * H + y) * W, fm.plane_origin(p) + (size_t)y * fm.stride, sizeof(int32_t) * W);
  return 0;
}

// run_syn (cc-frame-decoder.cpp:1044-1152) after ccref_run_ups: out = [3][H][W] at SYN_LAYER_PRECISION
int ccref_run_syn(void* hv, int32_t* out) {
  Handle* h = (Handle*)hv;
  frame_memory* r = h->dec->run_syn(h->frame);
  const int H = h->bs.m_gop_header.img_h, W = h->bs.m_gop_header.img_w;
  for (int p = 0; p < 3; ++p)
    for (int y = 0; y < H; ++y)
      memcpy(out + ((size_t)p * H + y) * W, r->plane_origin(p) + (size_t)y * r->stride, sizeof(int32_t) * W);
  return 0;
}

}  // extern "C"
