// entries.cu -- tile-mode helpers and single-kernel C entry points (unit tests, ncu).
#include <math.h>

#include "model.cuh"

using namespace hb;

namespace {
struct TmpPool {   // test entry points pack weights per call; freed after a sync
  std::vector<void*> ptrs;
  ~TmpPool() { for (void* p : ptrs) cudaFree(p); }
  int alloc(void** out, size_t bytes, cudaStream_t st) {
    HB_CHECK_CUDA(cudaMalloc(out, bytes ? bytes : 16));
    ptrs.push_back(*out);
    HB_CHECK_CUDA(cudaMemsetAsync(*out, 0, bytes ? bytes : 16, st));
    return 0;
  }
};
}  // namespace

extern "C" {

// HATModel.tile_process geometry, hat_model.py:52-108 (integer bookkeeping, bit-exact).
int hatb200_tile_plan(int height, int width, int tile_size, int tile_pad, int scale, int32_t* rects, int cap) {
  HB_REQUIRE(height > 0 && width > 0 && tile_size > 0 && tile_pad >= 0 && scale > 0, HATB200_EINVAL,
             "tile_plan: bad arguments");
  int tiles_x = (width + tile_size - 1) / tile_size;
  int tiles_y = (height + tile_size - 1) / tile_size;
  int n = 0;
  for (int y = 0; y < tiles_y; y++) {
    for (int x = 0; x < tiles_x; x++, n++) {
      if (!rects || n >= cap) continue;
      int x0 = x * tile_size, y0 = y * tile_size;
      int x1 = x0 + tile_size < width ? x0 + tile_size : width;
      int y1 = y0 + tile_size < height ? y0 + tile_size : height;
      int x0p = x0 - tile_pad > 0 ? x0 - tile_pad : 0, x1p = x1 + tile_pad < width ? x1 + tile_pad : width;
      int y0p = y0 - tile_pad > 0 ? y0 - tile_pad : 0, y1p = y1 + tile_pad < height ? y1 + tile_pad : height;
      int32_t* r = rects + (size_t)n * 12;
      r[0] = x0p; r[1] = x1p; r[2] = y0p; r[3] = y1p;
      r[4] = x0 * scale; r[5] = x1 * scale; r[6] = y0 * scale; r[7] = y1 * scale;
      r[8] = (x0 - x0p) * scale; r[9] = r[8] + (x1 - x0) * scale;
      r[10] = (y0 - y0p) * scale; r[11] = r[10] + (y1 - y0) * scale;
    }
  }
  return n;
}

int hatb200_crop_place(const void* src, int dtype, int B, int C, int src_h, int src_w, int crop_y0, int crop_x0,
                       void* dst, int dst_h, int dst_w, int out_y0, int out_x0, int rh, int rw, void* stream) {
  HB_REQUIRE(src && dst, HATB200_EINVAL, "crop_place: null tensor");
  return launch_crop_place(src, dtype, B, C, src_h, src_w, crop_y0, crop_x0, dst, dst_h, dst_w, out_y0, out_x0, rh, rw,
                           (cudaStream_t)stream);
}

int hatb200_reflect_pad(const void* src, int dtype, int B, int C, int h, int w, void* dst, int ph, int pw,
                        void* stream) {
  HB_REQUIRE(src && dst, HATB200_EINVAL, "reflect_pad: null tensor");
  return launch_reflect_pad(src, dtype, B, C, h, w, dst, ph, pw, (cudaStream_t)stream);
}

int hatb200_tensor2img(const void* x, int dtype, int B, int C, int H, int W, float min_v, float max_v, int rgb2bgr,
                       uint8_t* out, void* stream) {
  HB_REQUIRE(x && out && B >= 1 && C >= 1 && H >= 1 && W >= 1, HATB200_EINVAL, "tensor2img: bad arguments");
  return launch_tensor2img(x, dtype, B, C, H, W, min_v, max_v, rgb2bgr, out, (cudaStream_t)stream);
}

int hatb200_psnr_u8(const uint8_t* img, const uint8_t* img2, int H, int W, int C, int crop_border, int test_y_channel,
                    double* workspace, double* out, void* stream) {
  HB_REQUIRE(img && img2 && workspace && out && H >= 1 && W >= 1 && C >= 1, HATB200_EINVAL, "psnr_u8: bad arguments");
  return launch_psnr_u8(img, img2, H, W, C, crop_border, test_y_channel, workspace, out, (cudaStream_t)stream);
}

int hatb200_ssim_u8(const uint8_t* img, const uint8_t* img2, int H, int W, int C, int crop_border, int test_y_channel,
                    double* workspace, double* out, void* stream) {
  HB_REQUIRE(img && img2 && workspace && out && H >= 1 && W >= 1 && C >= 1, HATB200_EINVAL, "ssim_u8: bad arguments");
  return launch_ssim_u8(img, img2, H, W, C, crop_border, test_y_channel, workspace, out, (cudaStream_t)stream);
}

int hatb200_layernorm(const float* x, int M, int C, const float* gamma, const float* beta, void* out, int out_dtype,
                      int out_pitch, void* out_head, int head_c, void* stream) {
  HB_REQUIRE(x && gamma && beta && out && out_pitch >= C, HATB200_EINVAL, "layernorm: bad arguments");
  return launch_layernorm(x, M, C, gamma, beta, out, out_dtype, out_pitch, out_pitch, out_head, head_c,
                          (cudaStream_t)stream);
}

int hatb200_conv_tokens(const void* in, int dtype, int in_pitch, int in_coff, int Cin, int B, int H, int W,
                        const float* w, const float* bias, int N, int KH, int KW, int act, void* out, int out_pitch,
                        int use_tc, void* stream) {
  HB_REQUIRE(in && w && out && N >= 1 && Cin >= 1, HATB200_EINVAL, "conv_tokens: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  TmpPool pool;
  ConvArgs a{in, in_pitch, in_coff, Cin, B, H, W, KH, KW, N};
  Epilogue ep;
  ep.bias = bias; ep.act = act; ep.out_t = out; ep.out_t_pitch = out_pitch;
  int taps = KH * KW, r;
  if (use_tc) {
    HB_REQUIRE(dtype == HATB200_BF16, HATB200_EINVAL, "conv_tokens: the tcgen05 kernel needs bf16 activations");
    int Npad = tc_npad(N), Kpad = Cin <= 16 ? 16 : (Cin <= 32 ? 32 : round_up(Cin, 64));
    bf16* w16;
    HB_TRY(pool.alloc((void**)&w16, (size_t)taps * Npad * Kpad * sizeof(bf16), st));
    HB_TRY(launch_pack_w_bf16(w, HATB200_F32, N, Cin, taps, w16, Npad, Kpad, 0, Cin, st));
    HB_REQUIRE(conv_tc_supported(a, Npad, Kpad, ep), HATB200_EUNSUPPORTED,
               "conv_tokens: shape not supported by the tcgen05 kernel");
    r = launch_conv_tc(a, w16, Npad, Kpad, ep, st);
  } else {
    int Np = round_up(N, 64);
    float* w32;
    HB_TRY(pool.alloc((void**)&w32, (size_t)taps * Cin * Np * sizeof(float), st));
    HB_TRY(launch_pack_w_f32(w, HATB200_F32, N, Cin, taps, w32, Np, Np, st));
    r = launch_tapconv_simt(dtype, a, w32, Np, ep, st);
  }
  cudaError_t e = cudaStreamSynchronize(st);   // test entry: the packed copy dies with this call
  if (r == 0 && e != cudaSuccess) { set_error("conv_tokens: %s", cudaGetErrorString(e)); return HATB200_ECUDA; }
  return r;
}

int hatb200_linear_fused(const void* a, int a_pitch, int Cin, int B, int H, int W, const float* w, const float* bias,
                         int N, int act, const float* resid, const void* u, int u_pitch, const float* eca_scale,
                         float eca_coef, float* out_f32, void* out_t, int out_t_pitch, int out_t_zero_to,
                         const float* gamma, const float* beta, void* ln_out, int ln_pitch, void* ln_head,
                         float* head_colsum, void* stream) {
  HB_REQUIRE(a && w && N >= 1 && Cin >= 1, HATB200_EINVAL, "linear_fused: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  TmpPool pool;
  ConvArgs ca{a, a_pitch, 0, Cin, B, H, W, 1, 1, N};
  Epilogue ep;
  ep.bias = bias; ep.act = act;
  ep.resid = resid; ep.resid_pitch = N;
  ep.eca_u = u; ep.eca_pitch = u_pitch; ep.eca_scale = eca_scale; ep.eca_coef = eca_coef;
  ep.out_f32 = out_f32; ep.out_f32_pitch = N;
  ep.out_t = out_t; ep.out_t_pitch = out_t_pitch; ep.out_t_zero_to = out_t_zero_to;
  if (ln_out) {
    ep.ln_gamma = gamma; ep.ln_beta = beta; ep.ln_out = ln_out; ep.ln_pitch = ln_pitch;
    if (ln_head) { ep.ln_head = ln_head; ep.ln_head_c = 16; ep.ln_head_colsum = head_colsum; }
  }
  const int Npad = tc_npad(N), Kpad = round_up(Cin, 64);
  bf16* w16;
  HB_TRY(pool.alloc((void**)&w16, (size_t)Npad * Kpad * sizeof(bf16), st));
  HB_TRY(launch_pack_w_bf16(w, HATB200_F32, N, Cin, 1, w16, Npad, Kpad, 0, Cin, st));
  HB_REQUIRE(gemm_tc_supported(ca, Npad, Kpad, ep), HATB200_EUNSUPPORTED,
             "linear_fused: shape / epilogue not supported by the persistent tcgen05 GEMM");
  int r = launch_gemm_tc(ca, w16, Npad, Kpad, ep, st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (r == 0 && e != cudaSuccess) { set_error("linear_fused: %s", cudaGetErrorString(e)); return HATB200_ECUDA; }
  return r;
}

int hatb200_conv_fused(const void* a, int a_pitch, int Cin, int B, int H, int W, const float* w, const float* bias, int N,
                       int ksize, int act, const float* resid, float* out_f32, void* out_t, int out_t_pitch,
                       int out_t_zero_to, const float* gamma, const float* beta, void* ln_out, int ln_pitch,
                       void* ln_head, float* head_colsum, float* colsum, void* stream) {
  HB_REQUIRE(a && w && N >= 1 && Cin >= 1 && (ksize == 1 || ksize == 3), HATB200_EINVAL, "conv_fused: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  TmpPool pool;
  ConvArgs ca{a, a_pitch, 0, Cin, B, H, W, ksize, ksize, N};
  Epilogue ep;
  ep.bias = bias; ep.act = act;
  ep.resid = resid; ep.resid_pitch = N;
  ep.out_f32 = out_f32; ep.out_f32_pitch = N;
  ep.out_t = out_t; ep.out_t_pitch = out_t_pitch; ep.out_t_zero_to = out_t_zero_to;
  ep.colsum = colsum;
  if (ln_out) {
    ep.ln_gamma = gamma; ep.ln_beta = beta; ep.ln_out = ln_out; ep.ln_pitch = ln_pitch;
    if (ln_head) { ep.ln_head = ln_head; ep.ln_head_c = 16; ep.ln_head_colsum = head_colsum; }
  }
  const int taps = ksize * ksize, Npad = tc_npad(N), Kpad = round_up(Cin, 64);
  bf16* w16;
  HB_TRY(pool.alloc((void**)&w16, (size_t)taps * Npad * Kpad * sizeof(bf16), st));
  HB_TRY(launch_pack_w_bf16(w, HATB200_F32, N, Cin, taps, w16, Npad, Kpad, 0, Cin, st));
  const bool ok = ksize == 1 ? gemm_tc_supported(ca, Npad, Kpad, ep) : conv3_tc_supported(ca, Npad, Kpad, ep);
  HB_REQUIRE(ok, HATB200_EUNSUPPORTED, "conv_fused: shape / epilogue not supported by the persistent tcgen05 kernels");
  int r = ksize == 1 ? launch_gemm_tc(ca, w16, Npad, Kpad, ep, st) : launch_conv3_tc(ca, w16, Npad, Kpad, ep, st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (r == 0 && e != cudaSuccess) { set_error("conv_fused: %s", cudaGetErrorString(e)); return HATB200_ECUDA; }
  return r;
}

int hatb200_ocab_attention(const void* qkv, int dtype, int pitch, int B, int H, int W, int C, int heads, int ws,
                           int ows, const float* bias_table, void* out, int out_pitch, int use_tc, void* stream) {
  HB_REQUIRE(qkv && bias_table && out, HATB200_EINVAL, "ocab_attention: null tensor");
  AttnArgs a{qkv, pitch, B, H, W, C, heads, ws, ows, bias_table, out, out_pitch, 0};
  if (use_tc) {
    // the tcgen05 kernel works on the head-padded layout the forward's qkv GEMM produces: convert around the call
    cudaStream_t st = (cudaStream_t)stream;
    HB_REQUIRE(dtype == HATB200_BF16 && heads >= 1 && C % heads == 0, HATB200_EUNSUPPORTED,
               "ocab_attention: the tcgen05 kernel needs bf16 and C divisible by heads");
    const int hp = heads * ATTN_TC_DP;
    const size_t M = (size_t)B * H * W;
    TmpPool pool;
    bf16 *qkv_p, *out_p;
    float* bias_l2;
    HB_TRY(pool.alloc((void**)&qkv_p, M * 3 * hp * sizeof(bf16), st));
    HB_TRY(pool.alloc((void**)&out_p, M * hp * sizeof(bf16), st));
    HB_TRY(pool.alloc((void**)&bias_l2, (size_t)heads * ATTN_TC_BIAS_FLOATS * sizeof(float), st));
    AttnTcArgs ta{qkv_p, 3 * hp, B, H, W, heads, C / heads, ws, ows, bias_l2, out_p, hp};
    HB_REQUIRE(attn_tc_supported(ta), HATB200_EUNSUPPORTED, "ocab_attention: shape not supported by the tcgen05 kernel");
    HB_TRY(launch_headpad_qkv((const bf16*)qkv, pitch, C, heads, M, qkv_p, 3 * hp, st));
    HB_TRY(launch_bias_l2(bias_table, heads, ws, ows, bias_l2, st));
    int r = launch_attn_tc(ta, st);
    if (r == 0) r = launch_headunpad_out(out_p, hp, C, heads, M, (bf16*)out, out_pitch, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (r == 0 && e != cudaSuccess) { set_error("ocab_attention: %s", cudaGetErrorString(e)); return HATB200_ECUDA; }
    return r;
  }
  return launch_attn_simt(dtype, a, (cudaStream_t)stream);
}

int hatb200_dwconv_gate(const void* in, int dtype, int in_pitch, int B, int H, int W, int hid, const float* dw_w,
                        const float* dw_b, void* out, int out_pitch, void* stream) {
  HB_REQUIRE(in && dw_w && dw_b && out, HATB200_EINVAL, "dwconv_gate: null tensor");
  cudaStream_t st = (cudaStream_t)stream;
  TmpPool pool;
  float* wt;
  HB_TRY(pool.alloc((void**)&wt, (size_t)9 * 2 * hid * sizeof(float), st));
  HB_TRY(launch_pack_w_f32(dw_w, HATB200_F32, 2 * hid, 1, 9, wt, 2 * hid, 2 * hid, st));
  int r = launch_dwgate(in, dtype, in_pitch, B, H, W, hid, wt, dw_b, out, out_pitch, out_pitch, st);   // pad columns zeroed, as in the forward
  cudaError_t e = cudaStreamSynchronize(st);
  if (r == 0 && e != cudaSuccess) { set_error("dwconv_gate: %s", cudaGetErrorString(e)); return HATB200_ECUDA; }
  return r;
}

}  // extern "C"
