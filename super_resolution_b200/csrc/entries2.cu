// entries2.cu -- more single-kernel C entry points (unit tests, ncu): the kernels that round 1 only covered through
// whole-forward goldens (plk_tc, conv_first, the conv_last tensor-core route, hab_finish, the
// pixel-shuffle TMA store).  Test entries: weights are packed per call and the call synchronises.
#include "model.cuh"

using namespace hb;

namespace {
struct TmpPool2 {
  std::vector<void*> ptrs;
  ~TmpPool2() { for (void* p : ptrs) cudaFree(p); }
  int alloc(void** out, size_t bytes, cudaStream_t st) {
    HB_CHECK_CUDA(cudaMalloc(out, bytes ? bytes : 16));
    ptrs.push_back(*out);
    HB_CHECK_CUDA(cudaMemsetAsync(*out, 0, bytes ? bytes : 16, st));
    return 0;
  }
};
int finish(int r, cudaStream_t st, const char* what) {
  cudaError_t e = cudaStreamSynchronize(st);
  if (r == 0 && e != cudaSuccess) { set_error("%s: %s", what, cudaGetErrorString(e)); return HATB200_ECUDA; }
  return r;
}
}  // namespace

extern "C" {

int hatb200_esc_plk(const void* x16, int B, int H, int W, const float* plk_filter, const float* dyn, void* out,
                    int out_pitch, void* stream) {
  HB_REQUIRE(x16 && plk_filter && out, HATB200_EINVAL, "esc_plk: null tensor");
  HB_REQUIRE(plk_tc_supported(16, 13, H, W, out_pitch), HATB200_EUNSUPPORTED,
             "esc_plk: %dx%d / pitch %d is not supported by the tcgen05 Toeplitz kernel", H, W, out_pitch);
  cudaStream_t st = (cudaStream_t)stream;
  TmpPool2 pool;
  bf16 *wt, *xpad;
  HB_TRY(pool.alloc((void**)&wt, plk_toeplitz_elems() * sizeof(bf16), st));
  HB_TRY(pool.alloc((void**)&xpad, (size_t)B * H * (W + 12) * 16 * sizeof(bf16), st));     // x pads stay zero
  HB_TRY(launch_pack_plk_toeplitz(plk_filter, HATB200_F32, wt, st));
  HB_TRY(launch_plk_pad_copy((const bf16*)x16, B, H, W, xpad, st));
  int r = launch_plk_tc(xpad, B, H, W, wt, dyn, (bf16*)out, out_pitch, st);
  return finish(r, st, "esc_plk");
}

int hatb200_conv_first(const void* x, int x_dtype, int B, int in_chans, int H, int W, const float* w, const float* bias,
                       int C, const float* mean4, float img_range, float* out, void* stream) {
  HB_REQUIRE(x && w && bias && mean4 && out && in_chans >= 1 && in_chans <= 4, HATB200_EINVAL, "conv_first: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  TmpPool2 pool;
  float* wp;
  HB_TRY(pool.alloc((void**)&wp, (size_t)9 * in_chans * C * sizeof(float), st));
  HB_TRY(launch_pack_w_f32(w, HATB200_F32, C, in_chans, 9, wp, C, C, st));
  int r = launch_conv_first(x, x_dtype, B, in_chans, H, W, wp, bias, C, mean4, img_range, out, st);
  return finish(r, st, "conv_first");
}

int hatb200_conv_last(const void* in, int dtype, int B, int H, int W, const float* w, const float* bias, int out_chans,
                      const float* mean4, float img_range, void* y, int y_dtype, int use_tc, void* stream) {
  HB_REQUIRE(in && w && bias && mean4 && y && out_chans >= 1 && out_chans <= 4, HATB200_EINVAL, "conv_last: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  TmpPool2 pool;
  int r;
  if (use_tc) {
    HB_REQUIRE(dtype == HATB200_BF16, HATB200_EINVAL, "conv_last: the tcgen05 route needs bf16 activations");
    const int Npad = tc_npad(out_chans), Kpad = 64;
    bf16* w16;
    float* cl32;
    HB_TRY(pool.alloc((void**)&w16, (size_t)9 * Npad * Kpad * sizeof(bf16), st));
    HB_TRY(pool.alloc((void**)&cl32, (size_t)B * H * W * 4 * sizeof(float), st));
    HB_TRY(launch_pack_w_bf16(w, HATB200_F32, out_chans, 64, 9, w16, Npad, Kpad, 0, 64, st));
    ConvArgs a{in, 64, 0, 64, B, H, W, 3, 3, out_chans};
    Epilogue ep;
    ep.bias = bias;
    if (use_tc == 2) {       // round-1 route: fp32 pixels-last, then the affine transpose kernel
      ep.out_f32 = cl32; ep.out_f32_pitch = 4;
      HB_REQUIRE(y_dtype != HATB200_U8, HATB200_EUNSUPPORTED, "conv_last: U8 output needs the one-pass route");
    } else {                 // one pass: the conv3 epilogue writes the image (or its tensor2img bytes) itself
      ep.nchw_y = y; ep.nchw_dt = y_dtype; ep.nchw_oc = out_chans; ep.nchw_range = img_range; ep.nchw_via_bf16 = 1;
      for (int o = 0; o < 4; o++) ep.nchw_mean[o] = mean4[o];
    }
    HB_REQUIRE(conv3_tc_supported(a, Npad, Kpad, ep), HATB200_EUNSUPPORTED, "conv_last: shape not supported by conv3_tc");
    r = launch_conv3_tc(a, w16, Npad, Kpad, ep, st);
    if (r == 0 && use_tc == 2) r = launch_nhwc_to_nchw_affine(cl32, 4, B, out_chans, H, W, mean4, img_range, y, y_dtype, st);
  } else {
    float* wp;
    HB_TRY(pool.alloc((void**)&wp, (size_t)9 * 64 * out_chans * sizeof(float), st));
    HB_TRY(launch_pack_w_f32(w, HATB200_F32, out_chans, 64, 9, wp, out_chans, out_chans, st));
    r = launch_conv_last(in, dtype, 64, 64, B, H, W, wp, bias, out_chans, mean4, img_range, y, y_dtype, st);
  }
  return finish(r, st, "conv_last");
}

int hatb200_hab_finish(const float* eca_partial, int nchunk, int B, int C, int HW, const float* eca_w5, float* eca_scale,
                       const float* dyn_partial, int pdim, const float* w1, const float* b1, const float* w2,
                       const float* b2, float* dyn, void* stream) {
  HB_REQUIRE(eca_partial && eca_w5 && eca_scale && dyn_partial && w1 && b1 && w2 && b2 && dyn, HATB200_EINVAL,
             "hab_finish: null tensor");
  cudaStream_t st = (cudaStream_t)stream;
  int r = launch_hab_finish(eca_partial, nchunk, B, C, HW, eca_w5, eca_scale, dyn_partial, pdim, w1, b1, w2, b2, dyn, st);
  return finish(r, st, "hab_finish");
}

int hatb200_conv_shuffle(const void* a, int B, int H, int W, const float* w, const float* bias, int r, void* out,
                         void* stream) {
  HB_REQUIRE(a && w && bias && out && (r == 2 || r == 3), HATB200_EINVAL, "conv_shuffle: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  TmpPool2 pool;
  const int O = 64 * r * r, Npad = tc_npad(O), Kpad = 64;
  bf16* w16s;
  float* bias_s;
  HB_TRY(pool.alloc((void**)&w16s, (size_t)9 * Npad * Kpad * sizeof(bf16), st));
  HB_TRY(pool.alloc((void**)&bias_s, (size_t)O * sizeof(float), st));
  HB_TRY(launch_pack_w_bf16_shuffle(w, HATB200_F32, O, 64, 9, r, w16s, Npad, Kpad, bias, bias_s, st));
  ConvArgs ca{a, 64, 0, 64, B, H, W, 3, 3, O};
  Epilogue ep;
  ep.bias = bias_s; ep.out_t = out; ep.out_t_pitch = 64; ep.shuffle_r = r; ep.shuffle_perm = 1;
  HB_REQUIRE(conv3_tc_supported(ca, Npad, Kpad, ep), HATB200_EUNSUPPORTED, "conv_shuffle: shape not supported by conv3_tc");
  int rc = launch_conv3_tc(ca, w16s, Npad, Kpad, ep, st);
  return finish(rc, st, "conv_shuffle");
}

int hatb200_ffn_tail_fused(const void* h1, int in_pitch, int B, int H, int W, int hid, const float* dw_w,
                           const float* dw_b, const float* fc2_w, const float* fc2_b, int N, const float* resid,
                           float* out_f32, const float* gamma, const float* beta, void* ln_out, int ln_pitch,
                           void* stream) {
  HB_REQUIRE(h1 && dw_w && dw_b && fc2_w && fc2_b && resid && out_f32 && gamma && beta && ln_out, HATB200_EINVAL,
             "ffn_tail_fused: null tensor");
  cudaStream_t st = (cudaStream_t)stream;
  TmpPool2 pool;
  const int Npad = tc_npad(N), Kpad = round_up(hid, 64);
  bf16* w16;
  float* dwt;
  HB_TRY(pool.alloc((void**)&w16, (size_t)Npad * Kpad * sizeof(bf16), st));
  HB_TRY(pool.alloc((void**)&dwt, (size_t)9 * 2 * hid * sizeof(float), st));
  HB_TRY(launch_pack_w_bf16(fc2_w, HATB200_F32, N, hid, 1, w16, Npad, Kpad, 0, hid, st));
  HB_TRY(launch_pack_w_f32(dw_w, HATB200_F32, 2 * hid, 1, 9, dwt, 2 * hid, 2 * hid, st));
  FfnTailArgs fa{h1, in_pitch, B, H, W, hid, dwt, dw_b, w16, N, Npad, Kpad};
  Epilogue ep;
  ep.bias = fc2_b; ep.resid = resid; ep.resid_pitch = N; ep.out_f32 = out_f32; ep.out_f32_pitch = N;
  ep.ln_gamma = gamma; ep.ln_beta = beta; ep.ln_out = ln_out; ep.ln_pitch = ln_pitch;
  HB_REQUIRE(ffn2_tc_supported(fa, ep), HATB200_EUNSUPPORTED, "ffn_tail_fused: shape not supported by ffn2_tc");
  int r = launch_ffn2_tc(fa, ep, st);
  return finish(r, st, "ffn_tail_fused");
}

}  // extern "C"
