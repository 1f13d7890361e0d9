// common.cuh -- shared device/host helpers for libhatb200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/hatb200.h"

typedef __nv_bfloat16 bf16;

namespace hb {

// ---- error plumbing (thread-local message, negative codes; never abort) ----
void set_error(const char* fmt, ...);
extern thread_local int g_launches;   // kernels launched since last reset

#define HB_CHECK_CUDA(expr)                                                        \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) {                                                       \
      hb::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return HATB200_ECUDA;                                                        \
    }                                                                              \
  } while (0)

#define HB_CHECK_LAUNCH()                                                          \
  do {                                                                             \
    hb::g_launches++;                                                              \
    cudaError_t _e = cudaPeekAtLastError();                                        \
    if (_e != cudaSuccess) {                                                       \
      hb::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return HATB200_ECUDA;                                                        \
    }                                                                              \
  } while (0)

#define HB_REQUIRE(cond, code, ...)                                                \
  do {                                                                             \
    if (!(cond)) {                                                                 \
      hb::set_error(__VA_ARGS__);                                                  \
      return (code);                                                               \
    }                                                                              \
  } while (0)

#define HB_TRY(expr)                                                               \
  do {                                                                             \
    int _r = (expr);                                                               \
    if (_r != 0) return _r;                                                        \
  } while (0)

// ---- programmatic dependent launch (PDL) -----------------------------------------------------------------
// The forward is ~290 dependent launches of 10-50 us kernels.  A kernel launched through launch_pdl() may start
// while its predecessor in the stream is still draining: its prologue (barrier init, TMEM allocation, tensor-map
// prefetch, parameter tables, resident weight panels -- nothing a predecessor writes) overlaps that tail.
// Contract for every kernel launched this way: pdl_wait() before the first access to memory another kernel of
// the forward produces and before the first global store; pdl_trigger() right after it, so at most one
// successor is in flight.  Launched without the attribute (HATB200_PDL=0) both instructions are no-ops.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();
// cluster_x > 1 launches thread-block clusters of that many CTAs along x (CTA pairs for tcgen05 cta_group::2)
template <class... KArgs, class... Args>
static inline cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), int cluster_x, dim3 grid, dim3 block, size_t smem,
                                             cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[2];
  int n = 0;
  if (cluster_x > 1) {
    at[n].id = cudaLaunchAttributeClusterDimension;
    at[n].val.clusterDim.x = (unsigned)cluster_x; at[n].val.clusterDim.y = 1; at[n].val.clusterDim.z = 1;
    n++;
  }
  if (pdl_enabled()) {
    at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[n].val.programmaticStreamSerializationAllowed = 1;
    n++;
  }
  cfg.attrs = at; cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
template <class... KArgs, class... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     Args&&... args) {
  return launch_pdl_cluster(kernel, 1, grid, block, smem, st, static_cast<Args&&>(args)...);
}

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline int round_up(int a, int b) { return cdiv(a, b) * b; }

// ---- storage-type helpers: arithmetic is always fp32 ----
__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f<__half>(float v) { return __float2half_rn(v); }

// exact GELU (approximate='none'), SiLU, LeakyReLU(0.01)
__device__ __forceinline__ float gelu_f(float v) { return 0.5f * v * (1.0f + erff(v * 0.70710678118654752440f)); }
__device__ __forceinline__ float silu_f(float v) { return v / (1.0f + expf(-v)); }
__device__ __forceinline__ float lrelu_f(float v) { return v > 0.f ? v : 0.01f * v; }
__device__ __forceinline__ float sigmoid_f(float v) { return 1.0f / (1.0f + expf(-v)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

enum { ACT_NONE = 0, ACT_GELU = 1, ACT_LRELU = 2 };

// A job for EXTRA CTAs of a launch (side_jobs.cuh): kind 0 = none.  Extra CTA j handles sample j (j < B).
struct SideJob {
  int kind = 0;                 // 1: eca_finish, 2: dyn_finish
  int B = 0, nchunk = 0, C = 0; // C: channels (eca) / pdim (dyn)
  float inv_hw = 0.f;
  const float* partial = nullptr;
  const float* w1 = nullptr;    // eca: w5;  dyn: w1
  const float* b1 = nullptr;
  const float* w2 = nullptr;
  const float* b2 = nullptr;
  float* out = nullptr;         // eca: scale (B, C);  dyn: (B, pdim, 9)
};

// Epilogue shared by the CUDA-core and tcgen05 conv/GEMM kernels.
// value order:  v = acc + bias[n]; v += dyn-depthwise term; v *= scale (n < scale_ncols);
//               v = act(v); v += eca_coef * eca_scale[b,n] * U[m,n]; v += resid[m,n]
struct Epilogue {
  const float* bias = nullptr;       // [N]
  int act = ACT_NONE;
  float out_scale = 1.f;             // applied to columns n < scale_ncols
  int scale_ncols = 0;
  const float* resid = nullptr;      // fp32 (M, resid_pitch)
  int resid_pitch = 0;
  const void* eca_u = nullptr;       // T (M, eca_pitch): CAB conv2 output
  int eca_pitch = 0;
  const float* eca_scale = nullptr;  // (B, N) sigmoid gate
  float eca_coef = 0.f;              // conv_scale
  const void* dyn_x = nullptr;       // T (M, dyn_pitch) compact first-pdim channels of the conv input
  int dyn_pitch = 0;
  const float* dyn_k = nullptr;      // (B, pdim, 9) per-sample depthwise 3x3 kernels
  float* out_f32 = nullptr;          // fp32 (M, out_f32_pitch) residual-stream output
  int out_f32_pitch = 0;
  void* out_t = nullptr;             // T (M', out_t_pitch) activation output at channel offset out_t_coff
  int out_t_pitch = 0;
  int out_t_coff = 0;
  int out_t_zero_to = 0;             // columns [N, out_t_zero_to) are written as zero (K padding of the consumer)
  int shuffle_r = 1;                 // PixelShuffle factor folded into the store (hat_arch.py:593-605)
  int shuffle_perm = 0;              // weights / bias rows are pre-permuted to n' = (i*r + j)*64 + c (tc_conv3.cu: each
                                     // 64-column block is then one (i, j) sub-pixel and leaves as one TMA store)
  // fused LayerNorm of the fp32 result row (tcgen05 kernel only; the CUDA-core path runs layernorm separately)
  const float* ln_gamma = nullptr;
  const float* ln_beta = nullptr;
  void* ln_out = nullptr;            // T (M, ln_pitch)
  int ln_pitch = 0;
  void* ln_head = nullptr;           // T (M, ln_head_c) compact copy of the first channels
  int ln_head_c = 0;
  void* ln_head_pad = nullptr;       // optional second copy of the head, x-padded (B, H, W+12, ln_head_c) for tc_plk.cu
  float* ln_head_colsum = nullptr;   // fp32 (num_m_tiles, ln_head_c) per-tile column sums of the LN head
  float* colsum = nullptr;           // fp32 (num_m_tiles, N) per-tile column sums of v (GAP partials)
  int pin_stream = 0;                // tc_gemm.cu: fp32 residual in / out carry the L2 evict_last policy (see forward.cu)
  int walk_back = 0;                 // tc_gemm.cu / tc_conv3.cu: walk the tiles from the last to the first (same result; forward.cu)
  SideJob side;                      // honoured by tc_conv3.cu only (the caller checks conv3_tc_supported first)
  // conv_last in one pass (tc_conv3.cu, FeatLast; hat_arch.py:858): y = (acc + bias) / img_range + mean written straight
  // to the NCHW image (nchw_dt F32 / BF16), or, nchw_dt = HATB200_U8, tensor2img of it (img_util.py:38-95: the value as
  // the compute dtype holds it, clamp to [0, 1], * 255, round half to even) as (B, H, W, oc) bytes with the channel
  // order reversed (RGB -> BGR).  No other output.
  void* nchw_y = nullptr;
  int nchw_dt = 0, nchw_oc = 0, nchw_via_bf16 = 0;
  float nchw_range = 1.f;
  float nchw_mean[4] = {0.f, 0.f, 0.f, 0.f};
};

}  // namespace hb
