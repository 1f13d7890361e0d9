/* hatb200.h -- C ABI of the B200-native HAT forward (libhatb200.so).
 *
 * The reference has no native interface for this path: its boundary is the Python
 * class `HAT(nn.Module)` (reference HAT/hat/archs/hat_arch.py:607-859) built by
 * `build_network` (HAT/ESC/basicsr/archs/__init__.py:18-24) and called as
 * `net(img)` (HAT/hat/models/hat_model.py:28-38, :80-88).  This header is the
 * contract a host binds instead; each entry cites the reference
 * lines it replaces.  Plain C: pointers, sizes, a CUDA stream passed as void*.
 *
 * Conventions: every function returns 0 on success and a negative
 * HATB200_E* code otherwise (never aborts); `hatb200_last_error()` gives a
 * thread-local message.  The caller owns all tensors and the workspace; the
 * handle owns only packed weight copies.  All work is enqueued on the given
 * stream, no hidden synchronisation.  A handle is bound to the CUDA device that
 * was current at `hatb200_create`; it is not thread-safe, distinct handles are.
 * There is no CPU path: without a CUDA device every compute entry fails.
 */
#ifndef HATB200_H
#define HATB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HATB200_VERSION 101
#define HATB200_MAX_LAYERS 32

enum { HATB200_F32 = 0, HATB200_BF16 = 1, HATB200_F16 = 2, HATB200_I64 = 3,
       HATB200_U8 = 4   /* only as y_dtype of hatb200_forward: the image as tensor2img bytes, see there */ };

enum {
  HATB200_OK = 0,
  HATB200_EINVAL = -1,   /* bad argument / shape (reference: RuntimeError from .view) */
  HATB200_ECUDA = -2,    /* CUDA runtime error */
  HATB200_EPARAM = -3,   /* missing / mis-shaped state_dict entry */
  HATB200_ENOMEM = -4,   /* workspace too small */
  HATB200_EUNSUPPORTED = -5
};

/* Mirrors the constructor arguments of the reference HAT (hat_arch.py:610-644)
 * after the host has resolved derived sizes. */
typedef struct hatb200_config {
  int32_t struct_size;            /* sizeof(hatb200_config), ABI guard */
  int32_t in_chans;               /* 1..4 */
  int32_t embed_dim;              /* C */
  int32_t num_layers;             /* len(depths) */
  int32_t depths[HATB200_MAX_LAYERS];
  int32_t num_heads[HATB200_MAX_LAYERS];
  int32_t window_size;            /* ws */
  int32_t overlap_win_size;       /* int(ws*overlap_ratio)+ws, hat_arch.py:289 */
  int32_t cab_mid;                /* C // compress_ratio, hat_arch.py:84 */
  int32_t ffn_hidden;             /* int(C*mlp_ratio), hat_arch.py:98 */
  int32_t esc_pdim, esc_kernel;   /* hat_arch.py:638-639 */
  int32_t ocab_esc_enable, ocab_esc_pdim, ocab_esc_kernel; /* hat_arch.py:641-643 */
  int32_t upscale;                /* 2^n or 3, hat_arch.py:593-605 */
  int32_t upsampler;              /* 1 = 'pixelshuffle', 0 = none (hat_arch.py:852) */
  int32_t resi_conv;              /* 1 = '1conv', 0 = 'identity' (hat_arch.py:543-548) */
  int32_t patch_norm;             /* hat_arch.py:683-688 */
  int32_t qkv_bias;
  float conv_scale;               /* hat_arch.py:236 */
  float img_range;                /* hat_arch.py:850 */
  float qk_scale;                 /* <=0: head_dim^-0.5 (hat_arch.py:288) */
  float mean[4];                  /* hat_arch.py:658-662 */
  /* sibling variant HATX (HAT/hat/archs/hatx_arch.py:708-975); all zero for HAT */
  int32_t arch;                   /* 0 = HAT, 1 = HATX: HAB uses the spatial-gate FFN (hatx_arch.py:144-180) */
  int32_t ocab_pad;               /* unfold padding of the OCAB key window: (ows-ws)/2 for HAT (hat_arch.py:296-297),
                                     ceil((ows-ws)/2) for HATX (hatx_arch.py:316-322) */
  int32_t kv_topk_keep;           /* HATX: keys kept per window, max(1, int(kv_topk_ratio * ows^2)); 0 = no pruning */
  int32_t use_focus_bias;         /* HATX: focus head + tanh bias on the logits (hatx_arch.py:405-415) */
} hatb200_config;

typedef struct hatb200_handle hatb200_handle;

int hatb200_version(void);
const char* hatb200_last_error(void);

/* HAT.__init__ (hat_arch.py:610-759): validates the configuration. */
int hatb200_create(const hatb200_config* cfg, hatb200_handle** out);
void hatb200_destroy(hatb200_handle* h);

/* load_state_dict (basicsr base_model.py:289-315): borrow one tensor of the
 * reference state_dict layout by its key.  `dev_ptr` must stay valid until the
 * next hatb200_pack().  Unknown keys -> HATB200_EPARAM. */
int hatb200_set_param(hatb200_handle* h, const char* key, const void* dev_ptr, int dtype,
                      const int64_t* shape, int ndim);
/* Number / names of the state_dict entries this configuration expects. */
int hatb200_num_params(hatb200_handle* h);
const char* hatb200_param_key(hatb200_handle* h, int i);
/* Build the packed device copies (allocates; call outside stream capture). */
int hatb200_pack(hatb200_handle* h, void* stream);

/* Pre-packed weight cache (no reference counterpart: the reference keeps only the fp32 state_dict,
 * base_model.py:207-252).  The packed device copies are a pure function of (configuration, state_dict):
 * hatb200_packed_bytes / hatb200_packed_export copy them into a host blob after hatb200_pack, hatb200_pack_import
 * rebuilds them from such a blob WITHOUT any hatb200_set_param (the caller keys the blob, e.g. by a hash of the
 * state_dict; a blob of another configuration or library version is rejected with HATB200_EPARAM). */
int hatb200_packed_bytes(hatb200_handle* h, size_t* out);
int hatb200_packed_export(hatb200_handle* h, void* host_dst, size_t cap, void* stream);
int hatb200_pack_import(hatb200_handle* h, const void* host_src, size_t bytes, void* stream);

/* Debug/engineering switches: "tensor_cores" (0/1), "sync_debug" (0/1), "op_timing" (0/1), "strict_tc" (0/1: in the
 * bf16 path a shape that no tcgen05 kernel takes fails with HATB200_EUNSUPPORTED instead of running the CUDA-core
 * kernel). */
int hatb200_set_option(hatb200_handle* h, const char* name, int value);

/* Bytes of scratch needed by hatb200_forward for (B,H,W) in the given compute
 * dtype (HATB200_F32 or HATB200_BF16). */
int hatb200_workspace_bytes(hatb200_handle* h, int B, int H, int W, int compute_dtype, size_t* out);

/* HAT.forward (hat_arch.py:848-859).  x: (B,in_chans,H,W) NCHW contiguous,
 * y: (B,in_chans,H*s,W*s) NCHW; x_dtype selects the arithmetic: F32 -> fp32
 * CUDA-core path (parity mode), BF16 -> bf16 tensor-core path (fp32
 * accumulate).  H and W must be multiples of window_size (the reference raises
 * from calculate_mask -> window_partition, hat_arch.py:124-128, :831).
 * y_dtype HATB200_U8 (bf16 path only): y is (B, H*s, W*s, in_chans) bytes = tensor2img(rgb2bgr=True, min_max=(0,1)) of
 * the bf16 image (basicsr/utils/img_util.py:38-95: clamp, * 255, numpy round, channel order reversed), written by the
 * conv_last epilogue itself -- bit-identical to hatb200_tensor2img of the bf16 forward, a quarter of its D2H bytes. */
int hatb200_forward(hatb200_handle* h, const void* x, int x_dtype, int B, int H, int W,
                    void* y, int y_dtype, void* workspace, size_t ws_bytes, void* stream);
/* HAT.forward_features (hat_arch.py:828-846): x is the (B, embed_dim, H, W) output of conv_first, y the
 * (B, embed_dim, H, W) output of the final LayerNorm (patch_embed + norm, the residual groups, norm, patch_unembed).
 * Same arithmetic selection, geometry rules and workspace as hatb200_forward. */
int hatb200_forward_features(hatb200_handle* h, const void* x, int x_dtype, int B, int H, int W,
                             void* y, int y_dtype, void* workspace, size_t ws_bytes, void* stream);
/* Number of kernels launched by the last hatb200_forward on this handle. */
int hatb200_last_launch_count(hatb200_handle* h);
/* Number of ops of the last bf16 hatb200_forward that ran on a CUDA-core kernel because no tcgen05 kernel takes their
 * shape (0 for every published HAT variant at sizes that are multiples of 16). */
int hatb200_last_fallback_count(hatb200_handle* h);

/* Per-op device times (CUDA events) of the last forward run with option "op_timing"=1, as a JSON
 * object {"tag": {"kernel","launches","ms","flops","bytes"}} with ALGORITHMIC flops/bytes per tag (bytes: every
 * stream the launch reads or writes -- activations, weights, fp32 residual in/out, CAB branch, LayerNorm copy --
 * each counted once) and the name of the kernel that served the tag. */
int hatb200_op_profile(hatb200_handle* h, char* json, size_t cap);

/* HATModel.tile_process geometry (hat_model.py:52-108).  Fills up to `cap`
 * rectangles of 12 ints each {in_x0,in_x1,in_y0,in_y1,out_x0,out_x1,out_y0,out_y1,
 * crop_x0,crop_x1,crop_y0,crop_y1} in the reference's tile order; returns the
 * number of tiles (host-only, no CUDA). */
int hatb200_tile_plan(int height, int width, int tile_size, int tile_pad, int scale,
                      int32_t* rects, int cap);
/* dst[:, :, out rect] = src[:, :, crop rect] for NCHW images on device
 * (hat_model.py:106-108). */
int hatb200_crop_place(const void* src, int dtype, int B, int C, int src_h, int src_w,
                       int crop_y0, int crop_x0, void* dst, int dst_h, int dst_w,
                       int out_y0, int out_x0, int rh, int rw, void* stream);
/* F.pad(img, (0,pw,0,ph), 'reflect') of HATModel.pre_process (hat_model.py:16-26). */
int hatb200_reflect_pad(const void* src, int dtype, int B, int C, int h, int w,
                        void* dst, int ph, int pw, void* stream);

/* ---- output post-processing right after the path -------- */

/* basicsr.utils.tensor2img (basicsr/utils/img_util.py:38-95) for a (B,C,H,W) device tensor:
 * clamp to [min_v,max_v], rescale to [0,1], x255, round half to even, uint8, NHWC; channel
 * order reversed when rgb2bgr and C == 3 (cv2.COLOR_RGB2BGR).  out: device (B,H,W,C) bytes --
 * the device-to-host copy shrinks 4x (fp32) / 2x (bf16) and the numpy / cv2 pass disappears. */
int hatb200_tensor2img(const void* x, int dtype, int B, int C, int H, int W, float min_v,
                       float max_v, int rgb2bgr, uint8_t* out, void* stream);
/* basicsr.metrics.calculate_psnr (basicsr/metrics/psnr_ssim.py:12-50) on two device uint8 HWC
 * images (BGR when C == 3): crop_border, optional Y channel (metric_util.py:32-45,
 * color_util.py:60-68).  workspace: HATB200_PSNR_MAX_BLOCKS doubles on device.
 * out (device, 2 doubles): {sum of squared differences, element count};
 * psnr = 10 log10(255^2 * count / sum), +inf when sum == 0. */
#define HATB200_PSNR_MAX_BLOCKS 1024
int hatb200_psnr_u8(const uint8_t* img, const uint8_t* img2, int H, int W, int C, int crop_border,
                    int test_y_channel, double* workspace, double* out, void* stream);

/* basicsr.metrics.calculate_ssim (basicsr/metrics/psnr_ssim.py:85-129, _ssim :170-198) on two device uint8 HWC images
 * (BGR when C == 3): crop_border, optional Y channel, float64 arithmetic; per channel the 11x11 Gaussian window
 * (sigma 1.5) over the valid region, as cv2.filter2D(...)[5:-5, 5:-5].  workspace: HATB200_SSIM_MAX_BLOCKS doubles on
 * device.  out (device, 2 doubles): {sum of the ssim map over all channels, window positions x channels};
 * ssim = out[0] / out[1].  The cropped image must be at least 11x11. */
#define HATB200_SSIM_MAX_BLOCKS 4096
int hatb200_ssim_u8(const uint8_t* img, const uint8_t* img2, int H, int W, int C, int crop_border,
                    int test_y_channel, double* workspace, double* out, void* stream);

/* ---- single-kernel entry points (unit tests, ncu) ------------------------ */

/* nn.LayerNorm(C), eps 1e-5 (hat_arch.py:222,:237,:331,:392,:844).
 * x fp32 (M,C) -> out (M,out_pitch) of out_dtype; optional compact copy of the
 * first `head_c` channels to out_head (M,head_c). */
int hatb200_layernorm(const float* x, int M, int C, const float* gamma, const float* beta,
                      void* out, int out_dtype, int out_pitch, void* out_head, int head_c,
                      void* stream);

/* Dense conv KHxKW (stride 1, zero pad K/2) / Linear (1x1) over channels-last
 * tokens: out[m,n] = sum_tap sum_c in[pix(m)+tap, coff+c] * w[n,c,tap] + bias[n].
 * `w` is the PyTorch OIHW fp32 weight (packed internally per call: test entry).
 * act: 0 none, 1 exact GELU, 2 LeakyReLU(0.01).  use_tc=1 runs the tcgen05
 * kernel (dtype must be BF16). */
int hatb200_conv_tokens(const void* in, int dtype, int in_pitch, int in_coff, int Cin,
                        int B, int H, int W, const float* w, const float* bias, int N,
                        int KH, int KW, int act, void* out, int out_pitch, int use_tc,
                        void* stream);

/* Linear (1x1 conv) with the fused epilogue of the persistent tcgen05 GEMM (bf16 only; csrc/tc_gemm.cu):
 *   v = act(a . w^T + bias) + eca_coef * eca_scale[b,n] * u[m,n] + resid[m,n]
 * (hat_arch.py:117,:236-237,:391-392; esc_arch.py:144).  a: bf16 (M, a_pitch), w: fp32 (N, Cin), resid / out_f32:
 * fp32 (M, N), u: bf16 (M, u_pitch), eca_scale: fp32 (B, N).  Optional outputs: out_f32, out_t bf16 (M,
 * out_t_pitch; columns [N, out_t_zero_to) written as zero), ln_out = LayerNorm(v; gamma, beta, eps 1e-5) bf16 (M,
 * ln_pitch; pad columns zero), ln_head bf16 (M, 16) = first 16 channels of ln_out, head_colsum fp32
 * (M / 128, 16) per-128-token column sums of ln_head.  Any optional pointer may be NULL. */
int hatb200_linear_fused(const void* a, int a_pitch, int Cin, int B, int H, int W, const float* w,
                         const float* bias, int N, int act, const float* resid, const void* u, int u_pitch,
                         const float* eca_scale, float eca_coef, float* out_f32, void* out_t, int out_t_pitch,
                         int out_t_zero_to, const float* gamma, const float* beta, void* ln_out, int ln_pitch,
                         void* ln_head, float* head_colsum, void* stream);

/* Dense conv (ksize 1 or 3, zero pad) with the fused epilogue of the persistent tcgen05 kernels (bf16 only;
 * csrc/tc_gemm.cu, csrc/tc_conv3.cu):  v = act(conv(a, w) + bias) + resid  (hat_arch.py:84-86,:544,:556,:854).
 * w: fp32 (N, Cin, ksize, ksize).  Optional outputs as in hatb200_linear_fused, plus colsum fp32
 * (B * H * W / 128, N): column sums of v over each 128-pixel tile (global-average-pool partials, hat_arch.py:73;
 * tiles are 8 x 16 pixels, numbered row-major inside a sample). */
int hatb200_conv_fused(const void* a, int a_pitch, int Cin, int B, int H, int W, const float* w,
                       const float* bias, int N, int ksize, int act, const float* resid, float* out_f32,
                       void* out_t, int out_t_pitch, int out_t_zero_to, const float* gamma, const float* beta,
                       void* ln_out, int ln_pitch, void* ln_head, float* head_colsum, float* colsum, void* stream);

/* OCAB attention core (hat_arch.py:354-388): qkv (M, pitch) holding q (already
 * scaled) at channel 0, k at C, v at 2C; bias_table (heads, (ws+ows-1)^2)
 * compact relative-position table indexed by (ky-qy+ws-1, kx-qx+ws-1);
 * out (M,out_pitch).  Keys outside the image are zero vectors, not masked.
 * use_tc=1 (BF16, ws 16, ows 24, head_dim <= 32) runs the tcgen05 flash-style kernel;
 * this test entry converts to and from the head-padded layout the forward uses. */
int hatb200_ocab_attention(const void* qkv, int dtype, int pitch, int B, int H, int W, int C,
                           int heads, int ws, int ows, const float* bias_table,
                           void* out, int out_pitch, int use_tc, void* stream);

/* GatedDconvFFN middle (hat_arch.py:112-116): depthwise 3x3 (+bias) over 2*hid
 * channels, then proj * SiLU(gate).  in (M,in_pitch) -> out (M,out_pitch). */
int hatb200_dwconv_gate(const void* in, int dtype, int in_pitch, int B, int H, int W, int hid,
                        const float* dw_w /* (2hid,1,3,3) */, const float* dw_b,
                        void* out, int out_pitch, void* stream);

/* ESC ConvAttn core on the tcgen05 banded-Toeplitz kernel (esc_arch.py:119-123; csrc/tc_plk.cu): x16 bf16 (M, 16) = the
 * first 16 channels of norm1(x); out[:, :16] (bf16, (M, out_pitch)) = conv13x13(x16, plk_filter (16,16,13,13), pad 6)
 * + depthwise3x3(x16, dyn[b] (16,1,3,3), pad 1); dyn (B, 16, 9) fp32 may be NULL.  Other columns of out are untouched. */
int hatb200_esc_plk(const void* x16, int B, int H, int W, const float* plk_filter, const float* dyn, void* out,
                    int out_pitch, void* stream);

/* conv_first + input normalisation (hat_arch.py:849-853): x (B, in_chans, H, W) NCHW -> out fp32 (B*H*W, C) tokens of
 * conv3x3((x - mean) * img_range, w (C, in_chans, 3, 3)) + bias. */
int hatb200_conv_first(const void* x, int x_dtype, int B, int in_chans, int H, int W, const float* w, const float* bias,
                       int C, const float* mean4, float img_range, float* out, void* stream);

/* conv_last + de-normalisation (hat_arch.py:856-858): in (B*H*W, 64) tokens -> y (B, out_chans, H, W) NCHW of
 * conv3x3(in, w (out_chans, 64, 3, 3)) + bias, then / img_range + mean.  use_tc=1 (bf16): the tensor-core route of
 * the forward (N padded to 16 on conv3_tc, whose epilogue applies the affine and writes the NCHW planes -- or, y_dtype
 * HATB200_U8, the (B, H, W, out_chans) tensor2img bytes); 2: the round-1 two-pass form of it (fp32 pixels-last, then a
 * transpose kernel); 0: the CUDA-core kernel. */
int hatb200_conv_last(const void* in, int dtype, int B, int H, int W, const float* w, const float* bias, int out_chans,
                      const float* mean4, float img_range, void* y, int y_dtype, int use_tc, void* stream);

/* The two per-sample "finish" steps of a HAB in one launch: ECA gate (hat_arch.py:72-77) eca_scale (B, C) =
 * sigmoid(conv1d_k5(mean over HW)) from the per-tile partial sums eca_partial (B, nchunk, C), and the ESC dynamic
 * kernel (esc_arch.py:95-100) dyn (B, pdim, 9) = W2 gelu(W1 GAP + b1) + b2 from dyn_partial (B, nchunk, pdim). */
int hatb200_hab_finish(const float* eca_partial, int nchunk, int B, int C, int HW, const float* eca_w5, float* eca_scale,
                       const float* dyn_partial, int pdim, const float* w1, const float* b1, const float* w2,
                       const float* b2, float* dyn, void* stream);

/* Upsample stage (hat_arch.py:593-605): conv3x3 64 -> 64 r^2 with PixelShuffle(r) folded into the TMA store of the
 * persistent conv3 kernel.  a bf16 (B*H*W, 64) tokens, w (64 r^2, 64, 3, 3), out bf16 (B, H r, W r, 64) channels-last:
 * out[b, y r + i, x r + j, c] = conv[b, c r^2 + i r + j, y, x]  (integer address map, bit-exact). */
int hatb200_conv_shuffle(const void* a, int B, int H, int W, const float* w, const float* bias, int r, void* out,
                         void* stream);

/* GatedDconvFFN back half in one persistent tcgen05 kernel (hat_arch.py:112-119, :237; csrc/tc_ffn.cu): h1 bf16
 * (M, in_pitch) = fc1 output (proj channels [0, hid), gate [hid, 2 hid)); G = dw3x3(h1)[:, :hid] * SiLU(dw3x3(h1)[:, hid:])
 * stays in shared memory; v = G . fc2_w^T + fc2_b + resid; out_f32 (M, N) = v; ln_out bf16 (M, ln_pitch) =
 * LayerNorm(v; gamma, beta, eps 1e-5), pad columns zero.  resid / out_f32 fp32 (M, N) may alias. */
int hatb200_ffn_tail_fused(const void* h1, int in_pitch, int B, int H, int W, int hid, const float* dw_w,
                           const float* dw_b, const float* fc2_w, const float* fc2_b, int N, const float* resid,
                           float* out_f32, const float* gamma, const float* beta, void* ln_out, int ln_pitch,
                           void* stream);

/* hatb200_dwconv_gate for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  No counterpart in the reference.
 * Same arguments, arithmetic and results as hatb200_dwconv_gate, but the
 * re-laid-out depthwise weight lives in a file-scope static __device__ array (csrc/noalloc.cu) instead
 * of a per-call allocation: the call makes no cudaMalloc / cudaFree / allocating library call and does
 * not synchronise.  Limit: hid <= HATB200_NOALLOC_MAX_HID (larger -> HATB200_EINVAL with a message);
 * every GPU that loads the library holds 9 * 2 * HATB200_NOALLOC_MAX_HID floats (72 KiB) until exit.
 * All calls share that array (the used part is cleared and rewritten on `stream` by each call):
 * calls on one stream are ordered and safe, calls that overlap on the GPU race. */
#define HATB200_NOALLOC_MAX_HID 1024
int hatb200_dwconv_gate_noalloc(const void* in, int dtype, int in_pitch, int B, int H, int W, int hid,
                                const float* dw_w /* (2hid,1,3,3) */, const float* dw_b,
                                void* out, int out_pitch, void* stream);

#ifdef HATB200_ENGINEERING
/* Engineering probes (csrc/tc_probe.cu), only in libraries built with HATB200_ENGINEERING=1; no reference counterpart.
 * One tcgen05.mma with caller-chosen descriptor fields, and the per-CTA %globaltimer trace of the last persistent
 * GEMM / conv3 launch made with HATB200_GEMM_TRACE set (tools/probe_umma.py, tools/trace_gemm.py, tools/trace_conv3.py). */
int hatb200_probe_umma(const void* a, int R, const void* b, int r0, int sbo_bytes, int base_offset, float* out,
                       void* stream);
int hatb200_debug_gemm_trace(unsigned long long* host, void* stream);
/* tcgen05.cp.128x256b probe (tools/probe_tmem_cp.py): copies a TMA-written 128 x 32 fp32 tile (mode 0: one SWIZZLE_128B
 * box, mode 1: four SWIZZLE_32B boxes) into TMEM and reads it back; the residual preload of csrc/tc_gemm.cu rests on it. */
int hatb200_probe_tmem_cp(const float* src, int mode, float* out, void* stream);
/* Memory-system ceiling of the esc_aggr traffic pattern with no arithmetic (tools/probe_stream.py): o32 = r + a + u,
 * o16 = bf16(o32) over n8 groups of 8 elements, `ctas` CTAs of 1024 threads, optionally walking backwards (rev). */
int hatb200_probe_stream(const void* a, const void* u, const void* r, void* o32, void* o16, size_t n8, int ctas, int rev,
                         void* stream);
#endif

#ifdef __cplusplus
}
#endif
#endif /* HATB200_H */
