// simt_kernels.cu -- CUDA-core kernels of the HAT forward: the HBM-bound ops of the
// bf16 path and every small op of both paths.  fp32 arithmetic, storage type T.
#include "kernels.cuh"
#include "side_jobs.cuh"

namespace hb {

#define HB_DISPATCH_T(dt, T, ...)                                          \
  switch (dt) {                                                            \
    case HATB200_F32: { using T = float; __VA_ARGS__; break; }             \
    case HATB200_BF16: { using T = bf16; __VA_ARGS__; break; }             \
    case HATB200_F16: { using T = __half; __VA_ARGS__; break; }            \
    default: set_error("unsupported dtype %d", dt); return HATB200_EINVAL; \
  }

// ---------------------------------------------------------------------------
// LayerNorm over C, one warp per token (reference: nn.LayerNorm, eps 1e-5;
// hat_arch.py:222,:237,:331,:392,:575,:844).  fp32 statistics, two-pass.
// ---------------------------------------------------------------------------
template <typename T, int PER_LANE>
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, int M, int C,
                                                        const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, T* __restrict__ out,
                                                        int out_pitch, int zero_to, T* __restrict__ head, int head_c) {
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (warp >= M) return;
  const float* row = x + (size_t)warp * C;
  float v[PER_LANE];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < PER_LANE; i++) {
    int c = lane + i * 32;
    v[i] = c < C ? row[c] : 0.f;
    s += v[i];
  }
  float mean = warp_sum(s) / (float)C;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < PER_LANE; i++) {
    int c = lane + i * 32;
    float d = c < C ? v[i] - mean : 0.f;
    q += d * d;
  }
  float rstd = rsqrtf(warp_sum(q) / (float)C + 1e-5f);
  T* orow = out + (size_t)warp * out_pitch;
#pragma unroll
  for (int i = 0; i < PER_LANE; i++) {
    int c = lane + i * 32;
    if (c < C) {
      float y = (v[i] - mean) * rstd * gamma[c] + beta[c];
      orow[c] = from_f<T>(y);
      if (head != nullptr && c < head_c) head[(size_t)warp * head_c + c] = from_f<T>(y);
    } else if (c < zero_to) {
      orow[c] = from_f<T>(0.f);
    }
  }
}

int launch_layernorm(const float* x, int M, int C, const float* gamma, const float* beta, void* out, int out_dt,
                     int out_pitch, int zero_to, void* head, int head_c, cudaStream_t st) {
  HB_REQUIRE(C <= 512 && zero_to <= 512, HATB200_EUNSUPPORTED, "layernorm: C=%d > 512", C);
  int blocks = cdiv(M, 8);
  HB_DISPATCH_T(out_dt, T, {
    if (C <= 256 && zero_to <= 256)
      layernorm_kernel<T, 8><<<blocks, 256, 0, st>>>(x, M, C, gamma, beta, (T*)out, out_pitch, zero_to, (T*)head, head_c);
    else
      layernorm_kernel<T, 16><<<blocks, 256, 0, st>>>(x, M, C, gamma, beta, (T*)out, out_pitch, zero_to, (T*)head, head_c);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------
// conv_first with the input normalisation folded in (hat_arch.py:849-853).
// One thread per (token, channel); the 9*IC inputs are warp-broadcast loads.
// ---------------------------------------------------------------------------
// A warp owns 32 consecutive tokens and one group of CF_G * 4 consecutive channels: the 9*IC normalised inputs of a
// lane's token are loaded once (coalesced along x) and kept in registers; every weight is a float4 at a warp-uniform
// address (one broadcast transaction), reused by the 32 tokens.  0.3 loads per FMA instead of 0.5 with all 27 inputs
// re-read by each of a token's 45 threads (B200: 131 us for 65 536 tokens x 180 channels before).
constexpr int CF_G = 5;                // float4 groups per thread (20 channels)
// Round 2: a BLOCK owns 32 consecutive tokens and all their channels (one warp per 20-channel group), parks the result
// in shared memory and writes the 32 x C floats -- one contiguous piece of the token tensor -- with fully coalesced
// 16-byte stores.  Storing from the accumulators scattered every float4 over 32 different 720-byte rows (78 us per
// forward in the ncu launch list for 47 MB of output).
// ICT / CT: in_chans / embed_dim known at compile time (0 = run time).  With both, the 9 * IC inputs of a token load up
// front and every weight read is ld.shared at an immediate offset with no bounds test: 41 M -> 14 M warp instructions
// (the run-time form spent two thirds of its issue slots on index arithmetic and the `c < C` tests).
template <typename T, int ICT, int CT>
__global__ void __launch_bounds__(512) conv_first_kernel(const T* __restrict__ x, int B, int ICr, int H, int W,
                                                         const float* __restrict__ w, const float* __restrict__ bias,
                                                         int Cr, float4 mean, float range, float* __restrict__ out) {
  const int C = CT > 0 ? CT : Cr;
  constexpr bool FULL = CT > 0 && CT % (CF_G * 4) == 0;     // every warp's 20-channel group is complete
  extern __shared__ __align__(16) float cf_tile[];       // [32 tokens][C], then the weights [9 * IC][C]
  const int IC = ICT > 0 ? ICT : ICr;
  float* cf_w = cf_tile + 32 * C;
  // weights in shared memory: every FMA quad reads a warp-uniform float4, and LDS broadcasts at one per clock where the
  // same read through L1 (round 1) made the kernel load-issue bound (83 us for 318 MFMA)
  for (int i = threadIdx.x; i < 9 * IC * C / 4; i += blockDim.x)
    reinterpret_cast<float4*>(cf_w)[i] = __ldg(reinterpret_cast<const float4*>(w) + i);
  __syncthreads();
  const size_t M = (size_t)B * H * W;
  const int lane = threadIdx.x & 31;
  const int grp = threadIdx.x >> 5;                      // 20-channel group of this warp
  const size_t m0 = (size_t)blockIdx.x * 32;
  const size_t m = m0 + lane;
  const int c0 = grp * CF_G * 4;
  if (m < M) {
    const int px = (int)(m % W);
    const int py = (int)((m / W) % H);
    const int b = (int)(m / ((size_t)W * H));
    const float mean_a[4] = {mean.x, mean.y, mean.z, mean.w};
    float4 acc[CF_G];
#pragma unroll
    for (int q = 0; q < CF_G; q++)
      acc[q] = (bias && (FULL || c0 + 4 * q < C)) ? __ldg(reinterpret_cast<const float4*>(bias + c0 + 4 * q)) : make_float4(0.f, 0.f, 0.f, 0.f);
    const T* xb = x + (size_t)b * IC * H * W;
    auto fma_tap = [&](int t, float v) {                  // t = (ky * 3 + kx) * IC + i
      const float* wr = cf_w + t * C + c0;
#pragma unroll
      for (int q = 0; q < CF_G; q++) {
        if (!FULL && c0 + 4 * q >= C) break;
        const float4 wv = *reinterpret_cast<const float4*>(wr + 4 * q);
        acc[q].x = fmaf(v, wv.x, acc[q].x); acc[q].y = fmaf(v, wv.y, acc[q].y);
        acc[q].z = fmaf(v, wv.z, acc[q].z); acc[q].w = fmaf(v, wv.w, acc[q].w);
      }
    };
    // zero padding applies to the NORMALISED image (hat_arch.py:849-853): a pad pixel contributes nothing
    auto input = [&](int ky, int kx, int i) {
      const int yy = py + ky - 1, xx = px + kx - 1;
      const bool in = yy >= 0 && yy < H && xx >= 0 && xx < W;
      return in ? (to_f(xb[((size_t)i * H + yy) * W + xx]) - mean_a[i]) * range : 0.f;
    };
    if (ICT > 0) {
      float v[9 * (ICT > 0 ? ICT : 1)];                   // all loads in flight before the first FMA needs one
#pragma unroll
      for (int t = 0; t < 9 * ICT; t++) v[t] = input(t / (3 * ICT), (t / ICT) % 3, t % ICT);
#pragma unroll
      for (int t = 0; t < 9 * ICT; t++) fma_tap(t, v[t]);
    } else {
      for (int ky = 0; ky < 3; ky++)
        for (int kx = 0; kx < 3; kx++)
          for (int i = 0; i < IC; i++) fma_tap((ky * 3 + kx) * IC + i, input(ky, kx, i));
    }
#pragma unroll
    for (int q = 0; q < CF_G; q++)
      if (FULL || c0 + 4 * q < C) *reinterpret_cast<float4*>(cf_tile + (size_t)lane * C + c0 + 4 * q) = acc[q];
  }
  __syncthreads();
  const size_t rows = M - m0 < 32 ? M - m0 : 32;
  const size_t n4 = rows * C / 4;                        // the block's rows are one contiguous piece of `out`
  float4* dst = reinterpret_cast<float4*>(out + m0 * C);
  for (size_t i = threadIdx.x; i < n4; i += blockDim.x) dst[i] = reinterpret_cast<const float4*>(cf_tile)[i];
}

// generic channel count (C % 4 != 0): one thread per (token, channel)
template <typename T>
__global__ void __launch_bounds__(256) conv_first_kernel1(const T* __restrict__ x, int B, int IC, int H, int W,
                                                          const float* __restrict__ w, const float* __restrict__ bias,
                                                          int C, float4 mean, float range, float* __restrict__ out) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)B * H * W * C;
  if (idx >= total) return;
  int c = (int)(idx % C);
  size_t m = idx / C;
  int px = (int)(m % W);
  int py = (int)((m / W) % H);
  int b = (int)(m / ((size_t)W * H));
  const float mean_a[4] = {mean.x, mean.y, mean.z, mean.w};
  float acc = bias ? bias[c] : 0.f;
  for (int ky = 0; ky < 3; ky++) {
    int yy = py + ky - 1;
    if (yy < 0 || yy >= H) continue;
    for (int kx = 0; kx < 3; kx++) {
      int xx = px + kx - 1;
      if (xx < 0 || xx >= W) continue;
      for (int i = 0; i < IC; i++) {
        float v = (to_f(x[(((size_t)b * IC + i) * H + yy) * W + xx]) - mean_a[i]) * range;
        acc = fmaf(v, w[((ky * 3 + kx) * IC + i) * C + c], acc);
      }
    }
  }
  out[idx] = acc;
}

int launch_conv_first(const void* x, int x_dt, int B, int IC, int H, int W, const float* w, const float* bias, int C,
                      const float* mean, float range, float* out, cudaStream_t st) {
  HB_REQUIRE(IC >= 1 && IC <= 4, HATB200_EUNSUPPORTED, "conv_first: in_chans=%d not in 1..4", IC);
  float4 m4 = make_float4(mean[0], mean[1], mean[2], mean[3]);
  // vector kernel: one block per 32 tokens, one warp per 20-channel group; scalar kernel: one thread per (token, channel)
  const size_t M = (size_t)B * H * W;
  const int groups = (C + CF_G * 4 - 1) / (CF_G * 4);
  const bool v4 = C % 4 == 0 && groups <= 16 && ((uintptr_t)w & 15) == 0 && ((uintptr_t)out & 15) == 0 &&
                  (!bias || ((uintptr_t)bias & 15) == 0);
  HB_DISPATCH_T(x_dt, T, {
    if (v4 && IC == 3 && C == 180)
      conv_first_kernel<T, 3, 180><<<(unsigned)((M + 31) / 32), groups * 32, (size_t)(32 + 9 * IC) * C * sizeof(float), st>>>(
          (const T*)x, B, IC, H, W, w, bias, C, m4, range, out);
    else if (v4 && IC == 3)
      conv_first_kernel<T, 3, 0><<<(unsigned)((M + 31) / 32), groups * 32, (size_t)(32 + 9 * IC) * C * sizeof(float), st>>>(
          (const T*)x, B, IC, H, W, w, bias, C, m4, range, out);
    else if (v4)
      conv_first_kernel<T, 0, 0><<<(unsigned)((M + 31) / 32), groups * 32, (size_t)(32 + 9 * IC) * C * sizeof(float), st>>>(
          (const T*)x, B, IC, H, W, w, bias, C, m4, range, out);
    else
      conv_first_kernel1<T><<<(unsigned)((M * C + 255) / 256), 256, 0, st>>>((const T*)x, B, IC, H, W, w, bias, C, m4,
                                                                            range, out);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------
// conv_last (Cin -> OC<=4, 3x3) + "/ img_range + mean", NCHW store
// (hat_arch.py:856-858).  HBM/L2-bound: one thread per HR pixel, weights in smem.
// ---------------------------------------------------------------------------
template <typename T, typename TY>
__global__ void __launch_bounds__(256) conv_last_kernel(const T* __restrict__ in, int pitch, int Cin, int B, int H,
                                                        int W, const float* __restrict__ w,
                                                        const float* __restrict__ bias, int OC, float4 mean,
                                                        float range, TY* __restrict__ y) {
  extern __shared__ float sw[];  // [9][Cin][OC]
  for (int i = threadIdx.x; i < 9 * Cin * OC; i += blockDim.x) sw[i] = w[i];
  __syncthreads();
  size_t m = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= (size_t)B * H * W) return;
  int px = (int)(m % W);
  int py = (int)((m / W) % H);
  int b = (int)(m / ((size_t)W * H));
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int ky = 0; ky < 3; ky++) {
    int yy = py + ky - 1;
    if (yy < 0 || yy >= H) continue;
    for (int kx = 0; kx < 3; kx++) {
      int xx = px + kx - 1;
      if (xx < 0 || xx >= W) continue;
      const T* src = in + (((size_t)b * H + yy) * W + xx) * pitch;
      const float* wt = sw + (ky * 3 + kx) * Cin * OC;
      for (int c = 0; c < Cin; c++) {
        float v = to_f(src[c]);
        for (int o = 0; o < OC; o++) acc[o] = fmaf(v, wt[c * OC + o], acc[o]);
      }
    }
  }
  const float mean_a[4] = {mean.x, mean.y, mean.z, mean.w};
  for (int o = 0; o < OC; o++) {
    float v = (acc[o] + (bias ? bias[o] : 0.f)) / range + mean_a[o];
    y[(((size_t)b * OC + o) * H + py) * W + px] = from_f<TY>(v);
  }
}

int launch_conv_last(const void* in, int dt, int pitch, int Cin, int B, int H, int W, const float* w,
                     const float* bias, int OC, const float* mean, float range, void* y, int y_dt, cudaStream_t st) {
  HB_REQUIRE(OC >= 1 && OC <= 4, HATB200_EUNSUPPORTED, "conv_last: out channels=%d not in 1..4", OC);
  if (conv_last_bf16_ok(dt, pitch, Cin, OC))
    return launch_conv_last_bf16((const bf16*)in, B, H, W, w, bias, OC, mean, range, y, y_dt, st);
  size_t total = (size_t)B * H * W;
  float4 m4 = make_float4(mean[0], mean[1], mean[2], mean[3]);
  size_t smem = (size_t)9 * Cin * OC * sizeof(float);
  unsigned blocks = (unsigned)((total + 255) / 256);
  HB_DISPATCH_T(dt, T, {
    HB_DISPATCH_T(y_dt, TY, {
      conv_last_kernel<T, TY><<<blocks, 256, smem, st>>>((const T*)in, pitch, Cin, B, H, W, w, bias, OC, m4,
                                                         range, (TY*)y);
    });
  });
  HB_CHECK_LAUNCH();
  return 0;
}

template <typename T, typename TY>
__global__ void affine_roundtrip_kernel(const T* __restrict__ x, int IC, int HW, size_t total, float4 mean,
                                        float range, TY* __restrict__ y) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const float mean_a[4] = {mean.x, mean.y, mean.z, mean.w};
  float mu = mean_a[(i / HW) % IC];
  y[i] = from_f<TY>(((to_f(x[i]) - mu) * range) / range + mu);
}

int launch_affine_roundtrip(const void* x, int x_dt, int B, int IC, int HW, const float* mean, float range, void* y,
                            int y_dt, cudaStream_t st) {
  size_t total = (size_t)B * IC * HW;
  float4 m4 = make_float4(mean[0], mean[1], mean[2], mean[3]);
  HB_DISPATCH_T(x_dt, T, {
    HB_DISPATCH_T(y_dt, TY, {
      affine_roundtrip_kernel<T, TY><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const T*)x, IC, HW, total, m4,
                                                                                    range, (TY*)y);
    });
  });
  HB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------
// Deterministic per-sample column sums (global average pools of ECA,
// hat_arch.py:73, and of the ESC dynamic-kernel branch, esc_arch.py:96).
// Stage 1: each block sums ROWS_PER_CHUNK consecutive tokens of one sample in a
// fixed order; stage 2 (eca_finish / dyn_finish) adds the chunks in order.
// ---------------------------------------------------------------------------
constexpr int ROWS_PER_CHUNK = 128;
int colsum_chunks(int HW) { return cdiv(HW, ROWS_PER_CHUNK); }

template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ in, int pitch, int HW, int nc, int cpb,
                                                     float* __restrict__ partial) {
  __shared__ float red[256];
  int chunk = blockIdx.x, b = blockIdx.y, nchunk = gridDim.x;
  int groups = 256 / cpb;
  int c = threadIdx.x % cpb, g = threadIdx.x / cpb;
  int r0 = chunk * ROWS_PER_CHUNK;
  int r1 = min(r0 + ROWS_PER_CHUNK, HW);
  for (int cb = 0; cb < nc; cb += cpb) {
    float s = 0.f;
    if (g < groups && cb + c < nc)
      for (int r = r0 + g; r < r1; r += groups) s += to_f(in[((size_t)b * HW + r) * pitch + cb + c]);
    red[threadIdx.x] = s;
    __syncthreads();
    if (g == 0 && cb + c < nc) {
      float t = 0.f;
      for (int k = 0; k < groups; k++) t += red[k * cpb + c];
      partial[((size_t)b * nchunk + chunk) * nc + cb + c] = t;
    }
    __syncthreads();
  }
}

int launch_colsum(const void* in, int dt, int pitch, int B, int HW, int nc, float* partial, cudaStream_t st) {
  int cpb = nc >= 256 ? 256 : (nc > 128 ? 256 : (nc > 64 ? 128 : (nc > 32 ? 64 : (nc > 16 ? 32 : 16))));
  dim3 grid(colsum_chunks(HW), B);
  HB_DISPATCH_T(dt, T, { colsum_kernel<T><<<grid, 256, 0, st>>>((const T*)in, pitch, HW, nc, cpb, partial); });
  HB_CHECK_LAUNCH();
  return 0;
}

__global__ void eca_finish_kernel(const float* __restrict__ partial, int nchunk, int C, float inv_hw,
                                  const float* __restrict__ w5, float* __restrict__ scale) {
  extern __shared__ float mean[];  // C
  pdl_wait();
  pdl_trigger();
  eca_finish_body(blockIdx.x, partial, nchunk, C, inv_hw, w5, scale, mean);
}

int launch_eca_finish(const float* partial, int nchunk, int B, int C, int HW, const float* w5, float* scale,
                      cudaStream_t st) {
  launch_pdl(eca_finish_kernel, dim3(B), dim3(256), C * sizeof(float), st, partial, nchunk, C, 1.0f / (float)HW, w5, scale);
  HB_CHECK_LAUNCH();
  return 0;
}

__global__ void dyn_finish_kernel(const float* __restrict__ partial, int nchunk, int pdim, float inv_hw,
                                  const float* __restrict__ w1, const float* __restrict__ b1,
                                  const float* __restrict__ w2, const float* __restrict__ b2,
                                  float* __restrict__ dyn) {
  __shared__ float mean[64];
  __shared__ float hid[32];
  pdl_wait();
  pdl_trigger();
  dyn_finish_body(blockIdx.x, partial, nchunk, pdim, inv_hw, w1, b1, w2, b2, dyn, mean, hid);
}

// One launch for the two per-sample "finish" steps of a HAB (ECA gate of the CAB branch, dynamic depthwise kernel of
// the ESC branch): blocks [0, B) run the former, [B, 2B) the latter.  They are independent of each other and both
// only need the partial sums written by earlier kernels; a launch of their own each cost ~4 us of an otherwise idle GPU.
struct HabFinishArgs {
  const float* eca_partial; int nchunk, C; const float* w5; float* scale;
  const float* dyn_partial; int pdim; const float* w1; const float* b1; const float* w2; const float* b2; float* dyn;
  float inv_hw; int B;
};
__global__ void hab_finish_kernel(const HabFinishArgs a) {
  extern __shared__ float mean[];  // max(C, 64) + 32
  pdl_wait();
  pdl_trigger();
  if ((int)blockIdx.x < a.B) eca_finish_body(blockIdx.x, a.eca_partial, a.nchunk, a.C, a.inv_hw, a.w5, a.scale, mean);
  else dyn_finish_body(blockIdx.x - a.B, a.dyn_partial, a.nchunk, a.pdim, a.inv_hw, a.w1, a.b1, a.w2, a.b2, a.dyn, mean,
                       mean + (a.C > 64 ? a.C : 64));
}

int launch_hab_finish(const float* eca_partial, int nchunk, int B, int C, int HW, const float* w5, float* scale,
                      const float* dyn_partial, int pdim, const float* w1, const float* b1, const float* w2,
                      const float* b2, float* dyn, cudaStream_t st) {
  HB_REQUIRE(pdim <= 64 && pdim % 2 == 0, HATB200_EUNSUPPORTED, "esc pdim=%d unsupported", pdim);
  HabFinishArgs a{eca_partial, nchunk, C, w5, scale, dyn_partial, pdim, w1, b1, w2, b2, dyn, 1.0f / (float)HW, B};
  launch_pdl(hab_finish_kernel, dim3(2 * B), dim3(256), ((C > 64 ? C : 64) + 32) * sizeof(float), st, a);
  HB_CHECK_LAUNCH();
  return 0;
}

int launch_dyn_finish(const float* partial, int nchunk, int B, int pdim, int HW, const float* w1, const float* b1,
                      const float* w2, const float* b2, float* dyn, cudaStream_t st) {
  HB_REQUIRE(pdim <= 64 && pdim % 2 == 0, HATB200_EUNSUPPORTED, "esc pdim=%d unsupported", pdim);
  launch_pdl(dyn_finish_kernel, dim3(B), dim3(160), 0, st, partial, nchunk, pdim, 1.0f / (float)HW, w1, b1, w2, b2, dyn);
  HB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------
// GatedDconvFFN middle (hat_arch.py:112-116): depthwise 3x3 (+bias, zero pad) on
// 2*hid channels, then out[j] = dw[j] * SiLU(dw[hid + j]).  HBM/L2-bound.
// w: [9][2*hid] tap-major fp32.
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) dwgate_kernel(const T* __restrict__ in, int in_pitch, int B, int H, int W,
                                                     int hid, const float* __restrict__ w,
                                                     const float* __restrict__ bias, T* __restrict__ out,
                                                     int out_pitch, int zero_to) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int cols = max(hid, zero_to);
  size_t total = (size_t)B * H * W * cols;
  if (idx >= total) return;
  int j = (int)(idx % cols);
  size_t m = idx / cols;
  if (j >= hid) { out[m * out_pitch + j] = from_f<T>(0.f); return; }
  int px = (int)(m % W);
  int py = (int)((m / W) % H);
  float p = bias[j], g = bias[hid + j];
#pragma unroll
  for (int ky = 0; ky < 3; ky++) {
    int yy = py + ky - 1;
    if (yy < 0 || yy >= H) continue;
#pragma unroll
    for (int kx = 0; kx < 3; kx++) {
      int xx = px + kx - 1;
      if (xx < 0 || xx >= W) continue;
      const T* src = in + (m + (size_t)((ky - 1) * W + (kx - 1))) * in_pitch;
      const float* wt = w + (ky * 3 + kx) * 2 * hid;
      p = fmaf(to_f(src[j]), wt[j], p);
      g = fmaf(to_f(src[hid + j]), wt[hid + j], g);
    }
  }
  out[m * out_pitch + j] = from_f<T>(p * silu_f(g));
}

int launch_dwgate(const void* in, int dt, int in_pitch, int B, int H, int W, int hid, const float* w, const float* b,
                  void* out, int out_pitch, int zero_to, cudaStream_t st) {
  if (dt == HATB200_BF16 && H % 8 == 0 && W % 16 == 0 && dwgate_bf16_ok(in_pitch, W, hid, out_pitch, zero_to))
    return launch_dwgate_bf16((const bf16*)in, in_pitch, B, H, W, hid, w, b, (bf16*)out, out_pitch, st);
  size_t total = (size_t)B * H * W * (hid > zero_to ? hid : zero_to);
  HB_DISPATCH_T(dt, T, {
    dwgate_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const T*)in, in_pitch, B, H, W, hid, w, b,
                                                                     (T*)out, out_pitch, zero_to);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------
// Tile-mode helpers (hat_model.py:16-26, :106-108)
// ---------------------------------------------------------------------------
template <typename T>
__global__ void crop_place_kernel(const T* __restrict__ src, int sh, int sw, int cy, int cx, T* __restrict__ dst,
                                  int dh, int dw, int oy, int ox, int rh, int rw, size_t total) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int x = (int)(i % rw);
  int y = (int)((i / rw) % rh);
  size_t bc = i / ((size_t)rw * rh);
  dst[(bc * dh + oy + y) * dw + ox + x] = src[(bc * sh + cy + y) * sw + cx + x];
}

int launch_crop_place(const void* src, int dt, int B, int C, int sh, int sw, int cy, int cx, void* dst, int dh, int dw,
                      int oy, int ox, int rh, int rw, cudaStream_t st) {
  HB_REQUIRE(cy >= 0 && cx >= 0 && cy + rh <= sh && cx + rw <= sw && oy >= 0 && ox >= 0 && oy + rh <= dh &&
                 ox + rw <= dw, HATB200_EINVAL, "crop_place: rectangle out of range");
  size_t total = (size_t)B * C * rh * rw;
  if (total == 0) return 0;
  HB_DISPATCH_T(dt, T, {
    crop_place_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const T*)src, sh, sw, cy, cx, (T*)dst, dh,
                                                                         dw, oy, ox, rh, rw, total);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

template <typename T>
__global__ void reflect_pad_kernel(const T* __restrict__ src, int h, int w, T* __restrict__ dst, int H2, int W2,
                                   size_t total) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int x = (int)(i % W2);
  int y = (int)((i / W2) % H2);
  size_t bc = i / ((size_t)W2 * H2);
  int sx = x < w ? x : 2 * (w - 1) - x;   // 'reflect' (edge not repeated), pad on the right/bottom only
  int sy = y < h ? y : 2 * (h - 1) - y;
  dst[i] = src[(bc * h + sy) * w + sx];
}

int launch_reflect_pad(const void* src, int dt, int B, int C, int h, int w, void* dst, int ph, int pw,
                       cudaStream_t st) {
  HB_REQUIRE(ph >= 0 && pw >= 0 && ph < h && pw < w, HATB200_EINVAL,
             "reflect_pad: pad (%d,%d) must be smaller than the image (%d,%d)", ph, pw, h, w);
  size_t total = (size_t)B * C * (h + ph) * (w + pw);
  HB_DISPATCH_T(dt, T, {
    reflect_pad_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const T*)src, h, w, (T*)dst, h + ph,
                                                                          w + pw, total);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

// conv_last tail on the tensor-core route: (B*H*W, pitch) fp32 pixels-last -> (B, OC, H, W) with "/ img_range + mean"
// (hat_arch.py:858).  One thread per pixel: a 16-byte read, OC coalesced plane writes.
template <typename TY>
__global__ void nhwc_to_nchw_affine_kernel(const float* __restrict__ in, int pitch, int OC, size_t hw, float4 mean,
                                           float range, TY* __restrict__ y, size_t total) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const size_t b = i / hw, pix = i - b * hw;
  const float mean_a[4] = {mean.x, mean.y, mean.z, mean.w};
  float v[4];
  if (pitch == 4) {
    const float4 t = *reinterpret_cast<const float4*>(in + i * 4);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  } else {
    for (int o = 0; o < OC; o++) v[o] = in[i * pitch + o];
  }
  for (int o = 0; o < OC; o++) y[(b * OC + o) * hw + pix] = from_f<TY>(v[o] / range + mean_a[o]);
}

int launch_nhwc_to_nchw_affine(const float* in, int pitch, int B, int OC, int H, int W, const float* mean, float range,
                               void* y, int y_dt, cudaStream_t st) {
  HB_REQUIRE(OC >= 1 && OC <= 4 && pitch >= OC, HATB200_EINVAL, "nhwc_to_nchw: OC=%d pitch=%d", OC, pitch);
  const size_t hw = (size_t)H * W, total = (size_t)B * hw;
  const float4 m4 = make_float4(mean[0], mean[1], mean[2], mean[3]);
  const unsigned blocks = (unsigned)((total + 255) / 256);
  if (y_dt == HATB200_F32) nhwc_to_nchw_affine_kernel<float><<<blocks, 256, 0, st>>>(in, pitch, OC, hw, m4, range, (float*)y, total);
  else if (y_dt == HATB200_BF16) nhwc_to_nchw_affine_kernel<bf16><<<blocks, 256, 0, st>>>(in, pitch, OC, hw, m4, range, (bf16*)y, total);
  else { set_error("nhwc_to_nchw: output dtype %d unsupported", y_dt); return HATB200_EINVAL; }
  HB_CHECK_LAUNCH();
  return 0;
}

// tensor2img (basicsr/utils/img_util.py:38-95): clamp to [mn, mx], (v - mn) / (mx - mn), * 255, round half to even
// (numpy .round()), uint8, NCHW -> NHWC with the channel order reversed when rgb2bgr (cv2.COLOR_RGB2BGR).
// One thread per output pixel: reads C planes (coalesced along x), writes C consecutive bytes.
template <typename T>
__global__ void tensor2img_kernel(const T* __restrict__ x, int C, size_t hw, float mn, float mx, int flip,
                                  uint8_t* __restrict__ out, size_t total) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;     // (b, y, x)
  if (i >= total) return;
  size_t b = i / hw, pix = i - b * hw;
  const float span = mx - mn;
  for (int c = 0; c < C; c++) {
    float v = to_f(x[(b * C + (flip ? C - 1 - c : c)) * hw + pix]);
    v = fminf(fmaxf(v, mn), mx);             // Tensor.clamp_: NaN propagates in torch; images are finite
    v = __fdiv_rn(__fsub_rn(v, mn), span);
    v = rintf(__fmul_rn(v, 255.0f));
    out[i * C + c] = (uint8_t)v;
  }
}

int launch_tensor2img(const void* x, int dt, int B, int C, int H, int W, float mn, float mx, int rgb2bgr, uint8_t* out,
                      cudaStream_t st) {
  HB_REQUIRE(mx > mn, HATB200_EINVAL, "tensor2img: min_max (%g, %g)", mn, mx);
  size_t hw = (size_t)H * W, total = (size_t)B * hw;
  if (total == 0) return 0;
  const int flip = rgb2bgr && C == 3;
  HB_DISPATCH_T(dt, T, {
    tensor2img_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const T*)x, C, hw, mn, mx, flip, out, total);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

// calculate_psnr (basicsr/metrics/psnr_ssim.py:12-50) on two uint8 HWC images: crop_border, optional Y channel of
// BGR input (metric_util.py:32-45 -> color_util.py:60-68: float32(u8) / 255, dot with float64 weights + 16, / 255,
// float32, * 255 in float32), squared difference in float64.  Deterministic two-stage sum: per-block partials in a
// fixed tree, then one block adds them in index order.  out[0] = sum of squared differences, out[1] = element count.
__device__ __forceinline__ double y_of_bgr(const uint8_t* p) {
  const float b = __fdiv_rn((float)p[0], 255.f), g = __fdiv_rn((float)p[1], 255.f), r = __fdiv_rn((float)p[2], 255.f);
  double y = __dadd_rn(__dadd_rn(__dmul_rn((double)b, 24.966), __dmul_rn((double)g, 128.553)), __dmul_rn((double)r, 65.481));
  y = __dadd_rn(y, 16.0);
  const float y32 = (float)(y / 255.0);
  return (double)__fmul_rn(y32, 255.f);
}

__global__ void psnr_partial_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, int W, int C, int cb,
                                    int h2, int w2, int test_y, double* __restrict__ partial) {
  __shared__ double red[256];
  const size_t npix = (size_t)h2 * w2;
  double acc = 0.0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
    const int y = (int)(i / w2) + cb, x = (int)(i % w2) + cb;
    const uint8_t* pa = a + ((size_t)y * W + x) * C;
    const uint8_t* pb = b + ((size_t)y * W + x) * C;
    if (test_y && C == 3) {
      const double d = y_of_bgr(pa) - y_of_bgr(pb);
      acc += d * d;
    } else {
      for (int c = 0; c < C; c++) {
        // to_y_channel on a 1-channel image: float32(u8) / 255 * 255 in float32 (metric_util.py:41-45)
        const double va = test_y ? (double)__fmul_rn(__fdiv_rn((float)pa[c], 255.f), 255.f) : (double)pa[c];
        const double vb = test_y ? (double)__fmul_rn(__fdiv_rn((float)pb[c], 255.f), 255.f) : (double)pb[c];
        const double d = va - vb;
        acc += d * d;
      }
    }
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = red[0];
}

__global__ void psnr_finish_kernel(const double* __restrict__ partial, int n, double count, double* __restrict__ out) {
  double s = 0.0;
  for (int i = 0; i < n; i++) s += partial[i];
  out[0] = s;
  out[1] = count;
}

int launch_psnr_u8(const uint8_t* a, const uint8_t* b, int H, int W, int C, int crop_border, int test_y, double* ws,
                   double* out, cudaStream_t st) {
  HB_REQUIRE(crop_border >= 0 && 2 * crop_border < H && 2 * crop_border < W, HATB200_EINVAL,
             "psnr: crop_border %d leaves nothing of a %dx%d image", crop_border, H, W);
  const int h2 = H - 2 * crop_border, w2 = W - 2 * crop_border;
  const size_t npix = (size_t)h2 * w2;
  int blocks = (int)((npix + 1023) / 1024);
  if (blocks > HATB200_PSNR_MAX_BLOCKS) blocks = HATB200_PSNR_MAX_BLOCKS;
  psnr_partial_kernel<<<blocks, 256, 0, st>>>(a, b, W, C, crop_border, h2, w2, test_y, ws);
  HB_CHECK_LAUNCH();
  const double count = (double)npix * ((test_y && C == 3) ? 1 : C);
  psnr_finish_kernel<<<1, 1, 0, st>>>(ws, blocks, count, out);
  HB_CHECK_LAUNCH();
  return 0;
}

// calculate_ssim (basicsr/metrics/psnr_ssim.py:85-129, _ssim :170-198) on two uint8 HWC images: crop_border, optional
// Y channel (as calculate_psnr), float64 arithmetic as the reference (cv2.filter2D on float64): per channel, the 11x11
// Gaussian window (cv2.getGaussianKernel(11, 1.5), outer product) over the VALID region, ssim_map averaged.  A block
// stages a (32+10) x (8+10) tile of both images in shared memory as doubles; one thread = one window position (121
// taps, five running sums).  Blocks walk tiles grid-stride, per-block partial sums are added in index order
// (deterministic).  out[0] = sum of ssim_map over all channels, out[1] = number of window positions x channels.
constexpr int SS_TW = 32, SS_TH = 8, SS_K = 11, SS_HW = SS_TW + SS_K - 1, SS_HH = SS_TH + SS_K - 1;
struct SsimWin { double k[SS_K]; };

__global__ void __launch_bounds__(256) ssim_partial_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b,
                                                           int W, int C, int cb, int h2, int w2, int test_y, int nch,
                                                           int tiles_x, int tiles_y, const SsimWin win,
                                                           double* __restrict__ partial) {
  __shared__ double ta[SS_HH][SS_HW], tb[SS_HH][SS_HW];
  __shared__ double red[256];
  const int vh = h2 - (SS_K - 1), vw = w2 - (SS_K - 1);          // valid window positions
  const double c1 = (0.01 * 255) * (0.01 * 255), c2 = (0.03 * 255) * (0.03 * 255);
  const int lx = threadIdx.x % SS_TW, ly = threadIdx.x / SS_TW;
  const long long ntiles = (long long)tiles_x * tiles_y * nch;
  double acc = 0.0;
  for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int ch = (int)(t / ((long long)tiles_x * tiles_y));
    const int ty = (int)((t / tiles_x) % tiles_y), tx = (int)(t % tiles_x);
    const int x0 = tx * SS_TW, y0 = ty * SS_TH;
    __syncthreads();
    for (int e = threadIdx.x; e < SS_HH * SS_HW; e += 256) {
      const int hy = e / SS_HW, hx = e % SS_HW;
      const int y = y0 + hy, x = x0 + hx;
      double va = 0.0, vb = 0.0;
      if (y < h2 && x < w2) {
        const uint8_t* pa = a + ((size_t)(y + cb) * W + x + cb) * C;
        const uint8_t* pb = b + ((size_t)(y + cb) * W + x + cb) * C;
        if (test_y && C == 3) { va = y_of_bgr(pa); vb = y_of_bgr(pb); }
        else if (test_y) {
          va = (double)__fmul_rn(__fdiv_rn((float)pa[ch], 255.f), 255.f);
          vb = (double)__fmul_rn(__fdiv_rn((float)pb[ch], 255.f), 255.f);
        } else { va = (double)pa[ch]; vb = (double)pb[ch]; }
      }
      ta[hy][hx] = va; tb[hy][hx] = vb;
    }
    __syncthreads();
    if (y0 + ly < vh && x0 + lx < vw) {
      double m1 = 0, m2 = 0, s11 = 0, s22 = 0, s12 = 0;
      for (int i = 0; i < SS_K; i++) {
#pragma unroll
        for (int j = 0; j < SS_K; j++) {
          const double wgt = win.k[i] * win.k[j];
          const double p = ta[ly + i][lx + j], q = tb[ly + i][lx + j];
          m1 += wgt * p; m2 += wgt * q;
          s11 += wgt * (p * p); s22 += wgt * (q * q); s12 += wgt * (p * q);
        }
      }
      const double m11 = m1 * m1, m22 = m2 * m2, m12 = m1 * m2;
      const double v1 = s11 - m11, v2 = s22 - m22, v12 = s12 - m12;
      acc += ((2 * m12 + c1) * (2 * v12 + c2)) / ((m11 + m22 + c1) * (v1 + v2 + c2));
    }
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = red[0];
}

int launch_ssim_u8(const uint8_t* a, const uint8_t* b, int H, int W, int C, int crop_border, int test_y, double* ws,
                   double* out, cudaStream_t st) {
  HB_REQUIRE(crop_border >= 0 && H - 2 * crop_border >= SS_K && W - 2 * crop_border >= SS_K, HATB200_EINVAL,
             "ssim: crop_border %d leaves less than the 11x11 window of a %dx%d image", crop_border, H, W);
  const int h2 = H - 2 * crop_border, w2 = W - 2 * crop_border;
  const int vh = h2 - (SS_K - 1), vw = w2 - (SS_K - 1);
  const int nch = (test_y && C == 3) ? 1 : C;
  const int tiles_x = cdiv(vw, SS_TW), tiles_y = cdiv(vh, SS_TH);
  const long long ntiles = (long long)tiles_x * tiles_y * nch;
  int blocks = ntiles < HATB200_SSIM_MAX_BLOCKS ? (int)ntiles : HATB200_SSIM_MAX_BLOCKS;
  SsimWin win;                        // cv2.getGaussianKernel(11, 1.5): exp(-x^2 / (2 sigma^2)), normalised, float64
  double sum = 0.0;
  for (int i = 0; i < SS_K; i++) {
    const double x = i - (SS_K - 1) * 0.5;
    win.k[i] = exp(-0.5 / (1.5 * 1.5) * x * x);
    sum += win.k[i];
  }
  for (int i = 0; i < SS_K; i++) win.k[i] *= 1.0 / sum;
  ssim_partial_kernel<<<blocks, 256, 0, st>>>(a, b, W, C, crop_border, h2, w2, test_y, nch, tiles_x, tiles_y, win, ws);
  HB_CHECK_LAUNCH();
  psnr_finish_kernel<<<1, 1, 0, st>>>(ws, blocks, (double)vh * vw * nch, out);
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb

namespace hb {
template <typename T>
__global__ void add_convert_kernel(const float* __restrict__ a, const float* __restrict__ b, int C, int cols,
                                   size_t total, float* out_f32, T* __restrict__ out_t, int out_pitch) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  int c = (int)(idx % cols);
  size_t m = idx / cols;
  if (c >= C) { if (out_t) out_t[m * out_pitch + c] = from_f<T>(0.f); return; }
  float v = a[m * C + c] + (b ? b[m * C + c] : 0.f);
  if (out_f32) out_f32[m * C + c] = v;
  if (out_t) out_t[m * out_pitch + c] = from_f<T>(v);
}

int launch_add_convert(const float* a, const float* b, int M, int C, float* out_f32, void* out_t, int dt, int out_pitch,
                       int zero_to, cudaStream_t st) {
  int cols = (out_t && zero_to > C) ? zero_to : C;
  size_t total = (size_t)M * cols;
  HB_DISPATCH_T(dt, T, {
    add_convert_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(a, b, C, cols, total, out_f32, (T*)out_t,
                                                                          out_pitch);
  });
  HB_CHECK_LAUNCH();
  return 0;
}
}  // namespace hb

// ---------------------------------------------------------------------------
// HAT.forward_features boundary (hat_arch.py:828-846): (B, C, H, W) feature maps <-> channels-last tokens.
// 32 x 32 shared-memory tile transpose; both sides coalesced.
// ---------------------------------------------------------------------------
namespace hb {
template <typename TX>
__global__ void nchw_to_tokens_kernel(const TX* __restrict__ x, int C, int hw, float* __restrict__ out) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, c0 = blockIdx.y * 32, p0 = blockIdx.x * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, p = p0 + threadIdx.x;
    tile[i][threadIdx.x] = (c < C && p < hw) ? to_f(x[((size_t)b * C + c) * hw + p]) : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int p = p0 + i, c = c0 + threadIdx.x;
    if (p < hw && c < C) out[((size_t)b * hw + p) * C + c] = tile[threadIdx.x][i];
  }
}

template <typename T, typename TY>
__global__ void tokens_to_nchw_kernel(const T* __restrict__ in, int pitch, int C, int hw, TY* __restrict__ y) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, c0 = blockIdx.y * 32, p0 = blockIdx.x * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int p = p0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (p < hw && c < C) ? to_f(in[((size_t)b * hw + p) * pitch + c]) : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, p = p0 + threadIdx.x;
    if (c < C && p < hw) y[((size_t)b * C + c) * hw + p] = from_f<TY>(tile[threadIdx.x][i]);
  }
}

int launch_nchw_to_tokens(const void* x, int x_dt, int B, int C, int hw, float* out, cudaStream_t st) {
  dim3 grid(cdiv(hw, 32), cdiv(C, 32), B), block(32, 8);
  if (x_dt == HATB200_F32) nchw_to_tokens_kernel<float><<<grid, block, 0, st>>>((const float*)x, C, hw, out);
  else if (x_dt == HATB200_BF16) nchw_to_tokens_kernel<bf16><<<grid, block, 0, st>>>((const bf16*)x, C, hw, out);
  else { set_error("nchw_to_tokens: dtype %d unsupported", x_dt); return HATB200_EINVAL; }
  HB_CHECK_LAUNCH();
  return 0;
}

int launch_tokens_to_nchw(const void* in, int dt, int pitch, int B, int C, int hw, void* y, int y_dt, cudaStream_t st) {
  dim3 grid(cdiv(hw, 32), cdiv(C, 32), B), block(32, 8);
  if (dt == HATB200_F32 && y_dt == HATB200_F32)
    tokens_to_nchw_kernel<float, float><<<grid, block, 0, st>>>((const float*)in, pitch, C, hw, (float*)y);
  else if (dt == HATB200_BF16 && y_dt == HATB200_BF16)
    tokens_to_nchw_kernel<bf16, bf16><<<grid, block, 0, st>>>((const bf16*)in, pitch, C, hw, (bf16*)y);
  else if (dt == HATB200_BF16 && y_dt == HATB200_F32)
    tokens_to_nchw_kernel<bf16, float><<<grid, block, 0, st>>>((const bf16*)in, pitch, C, hw, (float*)y);
  else { set_error("tokens_to_nchw: dtypes %d -> %d unsupported", dt, y_dt); return HATB200_EINVAL; }
  HB_CHECK_LAUNCH();
  return 0;
}
}  // namespace hb

// ---------------------------------------------------------------------------
// Sibling variant HATX (HAT/hat/archs/hatx_arch.py).
// ---------------------------------------------------------------------------
namespace hb {

// SpatialGateDConvFFN middle (hatx_arch.py:166-178): in (M, in_pitch) holds fc1's output (xa = channels [0, c2),
// xb = [c2, 2 c2));  out[:, :c2] = (dw3x3(xa) + bias) * SiLU(xb),  out[:, c2:2 c2] = xb,  columns up to zero_to = 0.
template <typename T>
__global__ void sgate_kernel(const T* __restrict__ in, int in_pitch, int B, int H, int W, int c2,
                             const float* __restrict__ w /*[9][c2]*/, const float* __restrict__ bias,
                             T* __restrict__ out, int out_pitch, int cols) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t total = (size_t)B * H * W * cols;
  if (idx >= total) return;
  const int c = (int)(idx % cols);
  const size_t m = idx / cols;
  float v = 0.f;
  if (c < c2) {
    const int px = (int)(m % W), py = (int)((m / W) % H);
    float a = bias[c];
#pragma unroll
    for (int ky = 0; ky < 3; ky++) {
      const int yy = py + ky - 1;
      if (yy < 0 || yy >= H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; kx++) {
        const int xx = px + kx - 1;
        if (xx < 0 || xx >= W) continue;
        a = fmaf(to_f(in[(m + (size_t)(ky - 1) * W + (kx - 1)) * in_pitch + c]), w[(ky * 3 + kx) * c2 + c], a);
      }
    }
    v = a * silu_f(to_f(in[m * in_pitch + c2 + c]));
  } else if (c < 2 * c2) {
    v = to_f(in[m * in_pitch + c]);
  }
  out[m * out_pitch + c] = from_f<T>(v);
}

template <typename T>
__global__ void head_copy_kernel(const T* __restrict__ in, int pitch, size_t M, int hc, T* __restrict__ out) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= M * hc) return;
  const size_t m = idx / hc;
  const int c = (int)(idx % hc);
  out[idx] = in[m * pitch + c];
}

int launch_head_copy(const void* in, int dt, int pitch, size_t M, int hc, void* out, cudaStream_t st) {
  HB_DISPATCH_T(dt, T, {
    head_copy_kernel<T><<<(unsigned)((M * hc + 255) / 256), 256, 0, st>>>((const T*)in, pitch, M, hc, (T*)out);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

int launch_sgate(const void* in, int dt, int in_pitch, int B, int H, int W, int c2, const float* w, const float* b,
                 void* out, int out_pitch, int zero_to, cudaStream_t st) {
  const int cols = zero_to > 2 * c2 ? zero_to : 2 * c2;
  HB_REQUIRE(in_pitch >= 2 * c2 && out_pitch >= cols, HATB200_EINVAL, "sgate: pitches %d / %d for %d channels", in_pitch,
             out_pitch, 2 * c2);
  const size_t total = (size_t)B * H * W * cols;
  HB_DISPATCH_T(dt, T, {
    sgate_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const T*)in, in_pitch, B, H, W, c2, w, b, (T*)out,
                                                                    out_pitch, cols);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

// ||k||_2 over the C key channels of every token (hatx_arch.py:426): one warp per token, fp32.
template <typename T>
__global__ void knorm_kernel(const T* __restrict__ qkv, int pitch, int k_off, int C, size_t M, float* __restrict__ out) {
  const size_t m = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (m >= M) return;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float v = to_f(qkv[m * pitch + k_off + c]);
    s = fmaf(v, v, s);
  }
  s = warp_sum(s);
  if (lane == 0) out[m] = sqrtf(s);
}

int launch_knorm(const void* qkv, int dt, int pitch, int k_off, int C, size_t M, float* out, cudaStream_t st) {
  HB_DISPATCH_T(dt, T, {
    knorm_kernel<T><<<(unsigned)((M * 32 + 255) / 256), 256, 0, st>>>((const T*)qkv, pitch, k_off, C, M, out);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

// Per (window, key) term of the HATX OCAB (hatx_arch.py:405-441): kadd = tanh(saliency) of the key's token when the
// focus bias is on (0 for keys in the unfold padding), else 0; -inf for keys pruned by the top-k (score = the focus
// term when present, else ||k||_2; zero-padded keys score exactly 0).  The reference masks with -1e4 BEFORE the
// relative-position bias; exp(-1e4 + bias - max) is exactly 0 in fp32, so a pruned key is simply skipped.  Ties at the
// threshold are broken by the lower key index (torch.topk leaves them unspecified).  One block per window.
template <typename T>
__global__ void __launch_bounds__(256) ocab_focus_prep_kernel(const T* __restrict__ sal, int sal_pitch,
                                                              const float* __restrict__ knorm, int H, int W, int ws,
                                                              int ows, int pad, int k_keep, float* __restrict__ kadd) {
  __shared__ float score[1024];
  const int nwx = W / ws, nwy = H / ws, nk = ows * ows;
  const int win = blockIdx.x;
  const int b = win / (nwx * nwy), wr = win % (nwx * nwy), wy = wr / nwx, wx = wr % nwx;
  for (int k = threadIdx.x; k < nk; k += blockDim.x) {
    const int yy = wy * ws - pad + k / ows, xx = wx * ws - pad + k % ows;
    float v = 0.f;
    if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
      const size_t tok = ((size_t)b * H + yy) * W + xx;
      v = sal ? tanhf(to_f(sal[tok * sal_pitch])) : knorm[tok];
    }
    score[k] = v;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < nk; k += blockDim.x) {
    const float sk = score[k];
    bool keep = true;
    if (k_keep > 0 && k_keep < nk) {
      int rank = 0;
      for (int j = 0; j < nk; j++) {
        const float sj = score[j];
        rank += (sj > sk || (sj == sk && j < k)) ? 1 : 0;
      }
      keep = rank < k_keep;
    }
    kadd[(size_t)win * nk + k] = keep ? (sal ? sk : 0.f) : -INFINITY;
  }
}

int launch_ocab_focus_prep(const void* sal, int dt, int sal_pitch, const float* knorm, int B, int H, int W, int ws, int ows,
                           int pad, int k_keep, float* kadd, cudaStream_t st) {
  HB_REQUIRE(ows * ows <= 1024, HATB200_EUNSUPPORTED, "focus_prep: overlap window %d too large", ows);
  HB_REQUIRE(sal || knorm, HATB200_EINVAL, "focus_prep: neither a saliency map nor key norms");
  const int wins = B * (H / ws) * (W / ws);
  HB_DISPATCH_T(dt, T, {
    ocab_focus_prep_kernel<T><<<wins, 256, 0, st>>>((const T*)sal, sal_pitch, knorm, H, W, ws, ows, pad, k_keep, kadd);
  });
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb
