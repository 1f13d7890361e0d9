// pack.cu -- one-time weight re-layout kernels (run at hatb200_pack, not in the forward).
#include "kernels.cuh"

namespace hb {

template <typename T>
__device__ __forceinline__ float ld_as_f(const void* p, size_t i) { return to_f(((const T*)p)[i]); }

__device__ __forceinline__ float ld_dt(const void* p, int dt, size_t i) {
  if (dt == HATB200_F32) return ld_as_f<float>(p, i);
  if (dt == HATB200_BF16) return ld_as_f<bf16>(p, i);
  return ld_as_f<__half>(p, i);
}

// src (O, I, taps) [PyTorch OIHW flattened] -> dst[tap][i][n], n < ncols, row pitch Np, zero for n >= O.
// (dst may point into a wider matrix: concatenation of q_proj / kv_proj along n.)
__global__ void pack_w_f32_kernel(const void* __restrict__ src, int dt, int O, int I, int taps, float* __restrict__ dst,
                                  int Np, int ncols) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)taps * I * ncols;
  if (idx >= total) return;
  int n = (int)(idx % ncols);
  int i = (int)((idx / ncols) % I);
  int tap = (int)(idx / ((size_t)ncols * I));
  dst[((size_t)tap * I + i) * Np + n] = n < O ? ld_dt(src, dt, ((size_t)n * I + i) * taps + tap) : 0.f;
}

int launch_pack_w_f32(const void* src, int src_dt, int O, int I, int taps, float* dst, int Np, int ncols,
                      cudaStream_t st) {
  size_t total = (size_t)taps * I * ncols;
  pack_w_f32_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, src_dt, O, I, taps, dst, Np, ncols);
  HB_CHECK_LAUNCH();
  return 0;
}

// src (O, src_i_total, taps) -> dst[tap][n][k] bf16, n < Npad, k < Kpad; input channel = src_i_off + k (k < I).
__global__ void pack_w_bf16_kernel(const void* __restrict__ src, int dt, int O, int I, int taps,
                                   bf16* __restrict__ dst, int Npad, int Kpad, int src_i_off, int src_i_total) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)taps * Npad * Kpad;
  if (idx >= total) return;
  int k = (int)(idx % Kpad);
  int n = (int)((idx / Kpad) % Npad);
  int tap = (int)(idx / ((size_t)Kpad * Npad));
  float v = 0.f;
  if (n < O && k < I) v = ld_dt(src, dt, ((size_t)n * src_i_total + src_i_off + k) * taps + tap);
  dst[idx] = __float2bfloat16_rn(v);
}

int launch_pack_w_bf16(const void* src, int src_dt, int O, int I, int taps, bf16* dst, int Npad, int Kpad,
                       int src_i_off, int src_i_total, cudaStream_t st) {
  size_t total = (size_t)taps * Npad * Kpad;
  pack_w_bf16_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, src_dt, O, I, taps, dst, Npad, Kpad,
                                                                     src_i_off, src_i_total);
  HB_CHECK_LAUNCH();
  return 0;
}

// Pixel-shuffle layers for tc_conv3.cu: destination row n' = (i*r + j)*64 + c takes source row n = c*r*r + i*r + j
// (hat_arch.py:593-605: out[b, c, h*r + i, w*r + j] = in[b, c*r*r + i*r + j, h, w]).
__global__ void pack_w_bf16_shuffle_kernel(const void* __restrict__ src, int dt, int O, int I, int taps, int r,
                                           bf16* __restrict__ dst, int Npad, int Kpad) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)taps * Npad * Kpad;
  if (idx >= total) return;
  int k = (int)(idx % Kpad);
  int n = (int)((idx / Kpad) % Npad);
  int tap = (int)(idx / ((size_t)Kpad * Npad));
  float v = 0.f;
  if (n < O && k < I) {
    const int ns = (n % 64) * r * r + n / 64;
    v = ld_dt(src, dt, ((size_t)ns * I + k) * taps + tap);
  }
  dst[idx] = __float2bfloat16_rn(v);
}
__global__ void pack_bias_shuffle_kernel(const float* __restrict__ src, int O, int r, float* __restrict__ dst) {
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < O) dst[n] = src[(n % 64) * r * r + n / 64];
}
int launch_pack_w_bf16_shuffle(const void* src, int src_dt, int O, int I, int taps, int r, bf16* dst, int Npad, int Kpad,
                               const float* bias, float* bias_dst, cudaStream_t st) {
  HB_REQUIRE(O == 64 * r * r, HATB200_EINVAL, "pack: pixel-shuffle layer with %d outputs for r=%d", O, r);
  size_t total = (size_t)taps * Npad * Kpad;
  pack_w_bf16_shuffle_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, src_dt, O, I, taps, r, dst, Npad, Kpad);
  HB_CHECK_LAUNCH();
  pack_bias_shuffle_kernel<<<(O + 255) / 256, 256, 0, st>>>(bias, O, r, bias_dst);
  HB_CHECK_LAUNCH();
  return 0;
}

__global__ void to_f32_kernel(const void* __restrict__ src, int dt, float* __restrict__ dst, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = ld_dt(src, dt, i);
}

int launch_to_f32(const void* src, int src_dt, float* dst, size_t n, cudaStream_t st) {
  if (n == 0) return 0;
  to_f32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(src, src_dt, dst, n);
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb
