// kernels.cuh -- host-side launchers of every kernel in libhatb200.
// All launchers enqueue on `st`, never synchronise, and return HATB200_* codes.
#pragma once
#include "common.cuh"

namespace hb {

// ---- weight packing (pack.cu) ----
// src: PyTorch (O, I, KH, KW) [Linear: (O, I)] of dtype `src_dt`.
// fp32 layout for the CUDA-core kernels: dst[tap][i][n], n pitch = Np (zero padded).
int launch_pack_w_f32(const void* src, int src_dt, int O, int I, int taps, float* dst, int Np, int ncols,
                      cudaStream_t st);
// bf16 K-major layout for tcgen05: dst[tap][n][k], n < Npad, k < Kpad (zero padded).
int launch_pack_w_bf16_shuffle(const void* src, int src_dt, int O, int I, int taps, int r, bf16* dst, int Npad, int Kpad,
                               const float* bias, float* bias_dst, cudaStream_t st);
int launch_pack_w_bf16(const void* src, int src_dt, int O, int I, int taps, bf16* dst, int Npad, int Kpad,
                       int src_i_off, int src_i_total, cudaStream_t st);
// any float dtype vector -> fp32
int launch_to_f32(const void* src, int src_dt, float* dst, size_t n, cudaStream_t st);

// ---- CUDA-core kernels (simt_kernels.cu) ----
int launch_layernorm(const float* x, int M, int C, const float* gamma, const float* beta, void* out, int out_dt,
                     int out_pitch, int zero_to, void* head, int head_c, cudaStream_t st);
// conv_first (hat_arch.py:849-853): y[m, c] = conv3x3((x - mean) * range), x NCHW.
int launch_conv_first(const void* x, int x_dt, int B, int IC, int H, int W, const float* w /*[9][IC][C]*/,
                      const float* bias, int C, const float* mean, float range, float* out, cudaStream_t st);
// conv_last + de-normalisation (hat_arch.py:856-858): in (B*H*W, pitch) tokens of `dt` -> y NCHW of y_dt.
int launch_conv_last(const void* in, int dt, int pitch, int Cin, int B, int H, int W, const float* w /*[9][Cin][OC]*/,
                     const float* bias, int OC, const float* mean, float range, void* y, int y_dt, cudaStream_t st);
// no-upsampler path (hat_arch.py:849-858 with upsampler ''): y = ((x-mean)*range)/range + mean
int launch_affine_roundtrip(const void* x, int x_dt, int B, int IC, int HW, const float* mean, float range, void* y,
                            int y_dt, cudaStream_t st);
// deterministic per-sample column sums: in (B*HW, pitch) channels [0,nc) -> partial (B, nchunk, nc)
int colsum_chunks(int HW);
int launch_colsum(const void* in, int dt, int pitch, int B, int HW, int nc, float* partial, cudaStream_t st);
// ECA gate (hat_arch.py:72-77): mean over HW -> conv1d(k=5, pad 2) across channels -> sigmoid. partial (B,nchunk,C)
int launch_eca_finish(const float* partial, int nchunk, int B, int C, int HW, const float* w5, float* scale,
                      cudaStream_t st);
// ESC dynamic kernel (esc_arch.py:95-100,:121): GAP(pdim) -> 1x1 (pdim->pdim/2) -> GELU -> 1x1 (->pdim*9)
int launch_dyn_finish(const float* partial, int nchunk, int B, int pdim, int HW, const float* w1, const float* b1,
                      const float* w2, const float* b2, float* dyn /*(B,pdim,9)*/, cudaStream_t st);
int launch_hab_finish(const float* eca_partial, int nchunk, int B, int C, int HW, const float* w5, float* scale,
                      const float* dyn_partial, int pdim, const float* w1, const float* b1, const float* w2,
                      const float* b2, float* dyn, cudaStream_t st);
// GatedDconvFFN middle (hat_arch.py:112-116)
int launch_dwgate(const void* in, int dt, int in_pitch, int B, int H, int W, int hid, const float* w /*(2hid,9)*/,
                  const float* b, void* out, int out_pitch, int zero_to, cudaStream_t st);
// sibling variant HATX (hatx_arch.py): spatial-gate FFN middle, key norms, focus / top-k term per (window, key)
int launch_sgate(const void* in, int dt, int in_pitch, int B, int H, int W, int c2, const float* w /*[9][c2]*/,
                 const float* b, void* out, int out_pitch, int zero_to, cudaStream_t st);
int launch_head_copy(const void* in, int dt, int pitch, size_t M, int hc, void* out, cudaStream_t st);
int launch_knorm(const void* qkv, int dt, int pitch, int k_off, int C, size_t M, float* out, cudaStream_t st);
int launch_ocab_focus_prep(const void* sal, int dt, int sal_pitch, const float* knorm, int B, int H, int W, int ws, int ows,
                           int pad, int k_keep, float* kadd, cudaStream_t st);
// out_t[m, c] = a[m, c] + b[m, c] (b optional), fp32 in (M, C); optional fp32 copy to out_f32 (may alias a)
int launch_add_convert(const float* a, const float* b, int M, int C, float* out_f32, void* out_t, int dt, int out_pitch,
                       int zero_to, cudaStream_t st);
// bf16 fast paths (hbm_kernels.cu)
bool dwgate_bf16_ok(int in_pitch, int W, int hid, int out_pitch, int zero_to);
int launch_dwgate_bf16(const bf16* in, int in_pitch, int B, int H, int W, int hid, const float* w, const float* b,
                       bf16* out, int out_pitch, cudaStream_t st);
bool conv_last_bf16_ok(int dt, int pitch, int Cin, int OC);
int launch_conv_last_bf16(const bf16* in, int B, int H, int W, const float* w, const float* bias, int OC,
                          const float* mean, float range, void* y, int y_dt, cudaStream_t st);
int launch_crop_place(const void* src, int dt, int B, int C, int sh, int sw, int cy, int cx, void* dst, int dh, int dw,
                      int oy, int ox, int rh, int rw, cudaStream_t st);
int launch_reflect_pad(const void* src, int dt, int B, int C, int h, int w, void* dst, int ph, int pw, cudaStream_t st);
int launch_nhwc_to_nchw_affine(const float* in, int pitch, int B, int OC, int H, int W, const float* mean, float range,
                               void* y, int y_dt, cudaStream_t st);
int launch_nchw_to_tokens(const void* x, int x_dt, int B, int C, int hw, float* out, cudaStream_t st);
int launch_tokens_to_nchw(const void* in, int dt, int pitch, int B, int C, int hw, void* y, int y_dt, cudaStream_t st);
int launch_tensor2img(const void* x, int dt, int B, int C, int H, int W, float mn, float mx, int rgb2bgr, uint8_t* out,
                      cudaStream_t st);
int launch_psnr_u8(const uint8_t* a, const uint8_t* b, int H, int W, int C, int crop_border, int test_y, double* ws,
                   double* out, cudaStream_t st);

int launch_ssim_u8(const uint8_t* a, const uint8_t* b, int H, int W, int C, int crop_border, int test_y, double* ws,
                   double* out, cudaStream_t st);

// ---- dense conv / GEMM over channels-last tokens ----
struct ConvArgs {
  const void* A; int a_pitch; int a_coff; int Cin;   // input tokens (B*H*W, a_pitch), channels [a_coff, a_coff+Cin)
  int B, H, W, KH, KW;                               // stride 1, zero pad K/2
  int N;                                             // output channels
};
// CUDA-core path, fp32 math, storage dtype `dt` (F32 or BF16); w: fp32 [taps][Cin][Np]
int launch_tapconv_simt(int dt, const ConvArgs& a, const float* w, int Np, const Epilogue& ep, cudaStream_t st);
// tcgen05 path (bf16 storage, fp32 accumulate in TMEM); w: bf16 [taps][Npad][Kpad].
// Output channels are split into n-tiles of tc_bn(O) <= 256 columns; Npad = tc_npad(O) = n_tiles * tc_bn(O).
// n-tiles wider than 32 columns are multiples of 64 so that epilogue stores leave as whole 128-byte rows
inline int tc_bn(int O) { int n16 = round_up(O, 16); int nt = cdiv(n16, 256); return round_up(cdiv(n16, nt), n16 > 32 ? 64 : 16); }
inline int tc_npad(int O) { int n16 = round_up(O, 16); return cdiv(n16, 256) * tc_bn(O); }
int launch_tapconv_tc(const ConvArgs& a, const bf16* w, int Npad, int Kpad, const Epilogue& ep, cudaStream_t st);
bool tapconv_tc_supported(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep);
// persistent weight-stationary tcgen05 GEMM for 1x1 layers (tc_gemm.cu); same packed weights and epilogue
int launch_gemm_tc(const ConvArgs& a, const bf16* w, int Npad, int Kpad, const Epilogue& ep, cudaStream_t st);
bool gemm_tc_supported(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep);
// persistent halo-resident tcgen05 3x3 conv (tc_conv3.cu); same packed weights and epilogue semantics
int launch_conv3_tc(const ConvArgs& a, const bf16* w, int Npad, int Kpad, const Epilogue& ep, cudaStream_t st);
bool conv3_tc_side_ok(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep);
bool conv3_tc_supported(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep);
// picks the tcgen05 kernel for the shape (1x1 -> tc_gemm, 3x3 -> tc_conv3, else / fallback tc_tapconv)
inline bool conv_tc_supported(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep) {
  return gemm_tc_supported(a, Npad, Kpad, ep) || conv3_tc_supported(a, Npad, Kpad, ep) ||
         tapconv_tc_supported(a, Npad, Kpad, ep);
}
inline int launch_conv_tc(const ConvArgs& a, const bf16* w, int Npad, int Kpad, const Epilogue& ep, cudaStream_t st) {
  if (gemm_tc_supported(a, Npad, Kpad, ep)) return launch_gemm_tc(a, w, Npad, Kpad, ep, st);
  if (conv3_tc_supported(a, Npad, Kpad, ep)) return launch_conv3_tc(a, w, Npad, Kpad, ep, st);
  return launch_tapconv_tc(a, w, Npad, Kpad, ep, st);
}

// ---- GatedDconvFFN back half fused (tc_ffn.cu): depthwise 3x3 + gate -> fc2 -> residual + LayerNorm epilogue ----
struct FfnTailArgs {
  const void* h1; int in_pitch;      // fc1 output, bf16 (B*H*W, in_pitch): proj channels [0, hid), gate [hid, 2 hid)
  int B, H, W, hid;
  const float* dw_w;                 // [9][2 hid] tap-major depthwise weights
  const float* dw_b;                 // [2 hid]
  const bf16* w2; int N, Npad, Kpad; // fc2 weights packed [Npad][Kpad] K-major (bias, residual, LN in the Epilogue)
};
bool ffn2_tc_supported(const FfnTailArgs& a, const Epilogue& ep);
int launch_ffn2_tc(const FfnTailArgs& a, const Epilogue& ep, cudaStream_t st);

// ---- ESC 13x13 conv + dynamic depthwise on tensor cores (tc_plk.cu) ----
bool plk_tc_supported(int pdim, int ksize, int H, int W, int out_pitch);
size_t plk_toeplitz_elems();
int launch_pack_plk_toeplitz(const void* src, int dt, bf16* dst, cudaStream_t st);
int launch_plk_pad_copy(const bf16* src, int B, int H, int W, bf16* dst, cudaStream_t st);
int launch_plk_tc(const bf16* xpad, int B, int H, int W, const bf16* wt, const float* dyn, bf16* out, int out_pitch,
                  cudaStream_t st, const SideJob* side = nullptr);

// ---- OCAB attention core (hat_arch.py:354-388) ----
struct AttnArgs {
  const void* qkv; int pitch;        // (M, pitch): q at 0 (pre-scaled), k at C, v at 2C
  int B, H, W, C, heads, ws, ows;
  const float* bias;                 // (heads, (ws+ows-1)^2) compact table, see pack_ocab_bias
  void* out; int out_pitch; int out_zero_to;
  int pad = -1;                      // unfold padding of the key window; -1: (ows - ws) / 2 (HAT)
  const float* kadd = nullptr;       // HATX: (windows, ows^2) additive term per key, -inf = pruned (launch_ocab_focus_prep)
};
int launch_attn_simt(int dt, const AttnArgs& a, cudaStream_t st);
// tcgen05 kernel (tc_attn.cu): head-padded bf16 layout, head_dim padded to 32 channels (pad channels zero):
// q at [0, heads*32) (pre-scaled), k at [heads*32, 2*heads*32), v at [2*heads*32, 3*heads*32); out (M, out_pitch)
// holds heads*32 channels.
struct AttnTcArgs {
  const void* qkv; int pitch;
  int B, H, W, heads, head_dim, ws, ows;
  const float* bias_l2;              // (heads, 2, 39 x 48): compact table * log2(e), rows zero padded, + shifted copy (launch_bias_l2)
  void* out; int out_pitch;
};
int launch_attn_tc(const AttnTcArgs& a, cudaStream_t st);
bool attn_tc_supported(const AttnTcArgs& a);
// bias table in shared memory: 39 rows (dy) at a pitch of 48 floats -- pitch = 16 (mod 32) puts the two query rows a
// warp covers (lanes 0-15 / 16-31, addresses 48 floats apart) on disjoint bank halves, so the per-key bias read is
// conflict free (pitch 39 was 2-way conflicted on 9 of the 32 banks)
// Two copies per head, the second shifted by one float (B[j] = A[j + 1]): queries with odd x read pairs of A, queries
// with even x pairs of B, so every thread's (kx, kx + 1) bias pair is one aligned 8-byte load.
constexpr int ATTN_TC_DP = 32, ATTN_TC_BIAS_PITCH = 48, ATTN_TC_BIAS_TABLE = 39 * ATTN_TC_BIAS_PITCH;
constexpr int ATTN_TC_BIAS_FLOATS = 2 * ATTN_TC_BIAS_TABLE;
// pack-time helpers for the head-padded layout (tc_attn.cu)
int launch_pack_rows_headpad(const void* src, int dt, int C, int heads, int nblk, bf16* dst, int Kpad, int dst_row0,
                             cudaStream_t st);
int launch_pack_vec_headpad(const void* src, int dt, int C, int heads, int nblk, float* dst, int dst0, cudaStream_t st);
int launch_pack_cols_headpad(const void* src, int dt, int O, int C, int heads, bf16* dst, int Npad, int Kpad,
                             cudaStream_t st);
int launch_bias_l2(const float* compact, int heads, int ws, int ows, float* dst, cudaStream_t st);
int launch_headpad_qkv(const bf16* src, int src_pitch, int C, int heads, size_t M, bf16* dst, int dst_pitch,
                       cudaStream_t st);
int launch_headunpad_out(const bf16* src, int src_pitch, int C, int heads, size_t M, bf16* dst, int dst_pitch,
                         cudaStream_t st);
// compact[h][(dy+ws-1)*(ws+ows-1) + dx+ws-1] = table[wrap(rpi[q][k])][h]; also verifies on device that
// the index buffer is translation invariant (err_flag != 0 otherwise).
int launch_pack_ocab_bias(const void* table, int table_dt, const int64_t* rpi, int heads, int ws, int ows,
                          float* compact, int* err_flag, cudaStream_t st);

}  // namespace hb
