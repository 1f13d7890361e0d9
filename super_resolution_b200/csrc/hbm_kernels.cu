// hbm_kernels.cu -- bf16 fast paths of the bandwidth-bound ops (16-byte vector access, register / shared
// memory reuse of the 3x3 halo).  The generic (any dtype, any size) versions live in simt_kernels.cu.
#include "kernels.cuh"

namespace hb {

__device__ __forceinline__ void unpack8(const uint4& t, float (&o)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
  for (int j = 0; j < 4; j++) {
    float2 f = __bfloat1622float2(h[j]);
    o[2 * j] = f.x;
    o[2 * j + 1] = f.y;
  }
}

// ---------------------------------------------------------------------------------------------------
// GatedDconvFFN middle (hat_arch.py:112-116), bf16: each thread owns 4 consecutive pixels of one row
// and 8 channel pairs (proj j..j+7, gate hid+j..hid+j+7).  Per kernel row it loads the 6 input columns
// once (sliding-window reuse across the 4 pixels) -> 9 + 9 vector loads per pixel instead of 18 + 36.
// in (M, in_pitch) -> out (M, out_pitch); columns [hid, out_pitch) are written as zero.
// Algorithmic bytes: (2*hid + hid) * 2 B per token.
// ---------------------------------------------------------------------------------------------------
constexpr int DW_PX = 4;

// depthwise 3x3 of 8 consecutive channels starting at `ch` for DW_PX pixels of one row (sliding window)
__device__ __forceinline__ void dw_pass(const bf16* __restrict__ in, int in_pitch, int b, int y, int x0, int H, int W,
                                        int ch, int wpitch, const float* __restrict__ w,
                                        const float* __restrict__ bias, float (&acc)[DW_PX][8]) {
  {
    const float4* b4 = reinterpret_cast<const float4*>(bias + ch);
    const float4 t0 = __ldg(b4), t1 = __ldg(b4 + 1);
#pragma unroll
    for (int p = 0; p < DW_PX; p++) {
      acc[p][0] = t0.x; acc[p][1] = t0.y; acc[p][2] = t0.z; acc[p][3] = t0.w;
      acc[p][4] = t1.x; acc[p][5] = t1.y; acc[p][6] = t1.z; acc[p][7] = t1.w;
    }
  }
#pragma unroll
  for (int ky = 0; ky < 3; ky++) {
    const int yy = y + ky - 1;
    if (yy < 0 || yy >= H) continue;
    float wk[3][8];
#pragma unroll
    for (int kx = 0; kx < 3; kx++) {
      const float4* a4 = reinterpret_cast<const float4*>(w + (size_t)(ky * 3 + kx) * wpitch + ch);
      const float4 t0 = __ldg(a4), t1 = __ldg(a4 + 1);
      wk[kx][0] = t0.x; wk[kx][1] = t0.y; wk[kx][2] = t0.z; wk[kx][3] = t0.w;
      wk[kx][4] = t1.x; wk[kx][5] = t1.y; wk[kx][6] = t1.z; wk[kx][7] = t1.w;
    }
    const bf16* rowp = in + (((size_t)b * H + yy) * W) * in_pitch + ch;
    uint4 raw[DW_PX + 2];
#pragma unroll
    for (int cx = 0; cx < DW_PX + 2; cx++) {          // input columns x0-1 .. x0+DW_PX (all loads in flight together)
      const int xx = x0 + cx - 1;
      raw[cx] = (xx >= 0 && xx < W) ? *reinterpret_cast<const uint4*>(rowp + (size_t)xx * in_pitch)
                                    : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int cx = 0; cx < DW_PX + 2; cx++) {
      float v[8];
      unpack8(raw[cx], v);
#pragma unroll
      for (int kx = 0; kx < 3; kx++) {
        const int p = cx - kx;                         // output pixel that sees this column through tap kx
        if (p < 0 || p >= DW_PX) continue;
#pragma unroll
        for (int c = 0; c < 8; c++) acc[p][c] = fmaf(v[c], wk[kx][c], acc[p][c]);
      }
    }
  }
}

__global__ void __launch_bounds__(128, 4) dwgate_bf16_kernel(const bf16* __restrict__ in, int in_pitch, int B, int H,
                                                             int W, int hid, const float* __restrict__ w,
                                                             const float* __restrict__ bias, bf16* __restrict__ out,
                                                             int out_pitch) {
  const int vec_out = out_pitch / 8;                 // channel vectors per token in the output (incl. zero pad)
  const int wq = W / DW_PX;
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)B * H * wq * vec_out;
  if (idx >= total) return;
  const int cv = (int)(idx % vec_out);
  size_t g = idx / vec_out;
  const int xq = (int)(g % wq);
  const int y = (int)((g / wq) % H);
  const int b = (int)(g / ((size_t)wq * H));
  const int x0 = xq * DW_PX;
  const int j0 = cv * 8;
  const size_t row_tok = ((size_t)b * H + y) * W;
  if (j0 >= hid) {                                    // K padding of fc2's A operand
    const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int p = 0; p < DW_PX; p++) *reinterpret_cast<uint4*>(out + (row_tok + x0 + p) * out_pitch + j0) = z;
    return;
  }
  // gate half first (hat_arch.py:115-116: x_proj * silu(x_gate)), folded to its SiLU, then the proj half
  float s[DW_PX][8];
  dw_pass(in, in_pitch, b, y, x0, H, W, hid + j0, 2 * hid, w, bias, s);
#pragma unroll
  for (int p = 0; p < DW_PX; p++)
#pragma unroll
    for (int c = 0; c < 8; c++) s[p][c] = silu_f(s[p][c]);
  float a[DW_PX][8];
  dw_pass(in, in_pitch, b, y, x0, H, W, j0, 2 * hid, w, bias, a);
#pragma unroll
  for (int p = 0; p < DW_PX; p++) {
    uint4 o;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int c = 0; c < 4; c++)
      h[c] = __floats2bfloat162_rn(a[p][2 * c] * s[p][2 * c], a[p][2 * c + 1] * s[p][2 * c + 1]);
    *reinterpret_cast<uint4*>(out + (row_tok + x0 + p) * out_pitch + j0) = o;
  }
}

bool dwgate_bf16_ok(int in_pitch, int W, int hid, int out_pitch, int zero_to) {
  return hid % 8 == 0 && in_pitch % 8 == 0 && out_pitch % 8 == 0 && W % DW_PX == 0 && zero_to == out_pitch;
}

int launch_dwgate_bf16(const bf16* in, int in_pitch, int B, int H, int W, int hid, const float* w, const float* b,
                       bf16* out, int out_pitch, cudaStream_t st) {
  size_t total = (size_t)B * H * (W / DW_PX) * (out_pitch / 8);
  dwgate_bf16_kernel<<<(unsigned)((total + 127) / 128), 128, 0, st>>>(in, in_pitch, B, H, W, hid, w, b, out, out_pitch);
  HB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// conv_last (64 -> OC<=4, 3x3) + "/ img_range + mean" + NCHW store (hat_arch.py:856-858), bf16 input.
// N=3 is hopeless for MMA tiles; the op is bound by reading 64 ch x 2 B per HR pixel once.  A block
// stages a (32+2) x (8+2) pixel halo tile in shared memory (pixel stride 144 B: conflict-free 16-byte
// reads), then each thread produces one pixel.  Algorithmic bytes: 128 B in + OC*sizeof(out) per pixel.
// ---------------------------------------------------------------------------------------------------
constexpr int CL_TW = 32, CL_TH = 8, CL_C = 64, CL_STRIDE = 144;

template <typename TY, int OC>
__global__ void __launch_bounds__(256) conv_last_bf16_kernel(const bf16* __restrict__ in, int B, int H, int W,
                                                             const float* __restrict__ w,
                                                             const float* __restrict__ bias, float4 mean, float range,
                                                             TY* __restrict__ y) {
  extern __shared__ __align__(16) uint8_t cl_smem[];
  uint8_t* tile = cl_smem;                                                   // (TW+2)*(TH+2) pixels x 144 B
  float* sw = reinterpret_cast<float*>(cl_smem + (CL_TW + 2) * (CL_TH + 2) * CL_STRIDE);   // [9][64][OC]
  const int tx0 = blockIdx.x * CL_TW, ty0 = blockIdx.y * CL_TH, b = blockIdx.z;
  for (int i = threadIdx.x; i < 9 * CL_C * OC; i += 256) sw[i] = w[i];
  // cooperative halo load: (TH+2)*(TW+2) pixels x 8 vectors of 16 B
  for (int e = threadIdx.x; e < (CL_TW + 2) * (CL_TH + 2) * 8; e += 256) {
    const int v = e & 7, pix = e >> 3;
    const int lx = pix % (CL_TW + 2), ly = pix / (CL_TW + 2);
    const int gx = tx0 + lx - 1, gy = ty0 + ly - 1;
    uint4 t = make_uint4(0, 0, 0, 0);
    if (gx >= 0 && gx < W && gy >= 0 && gy < H)
      t = *reinterpret_cast<const uint4*>(in + (((size_t)b * H + gy) * W + gx) * CL_C + v * 8);
    *reinterpret_cast<uint4*>(tile + pix * CL_STRIDE + v * 16) = t;
  }
  __syncthreads();
  const int lx = threadIdx.x % CL_TW, ly = threadIdx.x / CL_TW;
  const int gx = tx0 + lx, gy = ty0 + ly;
  if (gx >= W || gy >= H) return;
  float acc[OC];
#pragma unroll
  for (int o = 0; o < OC; o++) acc[o] = 0.f;
#pragma unroll
  for (int ky = 0; ky < 3; ky++) {
#pragma unroll
    for (int kx = 0; kx < 3; kx++) {
      const uint8_t* src = tile + ((ly + ky) * (CL_TW + 2) + lx + kx) * CL_STRIDE;
      const float* wt = sw + (ky * 3 + kx) * CL_C * OC;
#pragma unroll
      for (int v = 0; v < 8; v++) {
        float f[8];
        unpack8(*reinterpret_cast<const uint4*>(src + v * 16), f);
#pragma unroll
        for (int c = 0; c < 8; c++)
#pragma unroll
          for (int o = 0; o < OC; o++) acc[o] = fmaf(f[c], wt[(v * 8 + c) * OC + o], acc[o]);
      }
    }
  }
  const float mean_a[4] = {mean.x, mean.y, mean.z, mean.w};
#pragma unroll
  for (int o = 0; o < OC; o++)
    y[(((size_t)b * OC + o) * H + gy) * W + gx] = from_f<TY>((acc[o] + bias[o]) / range + mean_a[o]);
}

bool conv_last_bf16_ok(int dt, int pitch, int Cin, int OC) {
  return dt == HATB200_BF16 && pitch == CL_C && Cin == CL_C && OC >= 1 && OC <= 4;
}

template <typename TY>
static int launch_cl(const bf16* in, int B, int H, int W, const float* w, const float* bias, int OC, float4 m4,
                     float range, TY* y, cudaStream_t st) {
  dim3 grid(cdiv(W, CL_TW), cdiv(H, CL_TH), B);
  const int smem = (CL_TW + 2) * (CL_TH + 2) * CL_STRIDE + 9 * CL_C * OC * (int)sizeof(float);
#define HB_CL_LAUNCH(N_)                                                                                        \
  do {                                                                                                          \
    HB_CHECK_CUDA(cudaFuncSetAttribute(conv_last_bf16_kernel<TY, N_>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                       smem));                                                                  \
    conv_last_bf16_kernel<TY, N_><<<grid, 256, smem, st>>>(in, B, H, W, w, bias, m4, range, y);                 \
  } while (0)
  switch (OC) {
    case 1: HB_CL_LAUNCH(1); break;
    case 2: HB_CL_LAUNCH(2); break;
    case 3: HB_CL_LAUNCH(3); break;
    default: HB_CL_LAUNCH(4); break;
  }
#undef HB_CL_LAUNCH
  HB_CHECK_LAUNCH();
  return 0;
}

int launch_conv_last_bf16(const bf16* in, int B, int H, int W, const float* w, const float* bias, int OC,
                          const float* mean, float range, void* y, int y_dt, cudaStream_t st) {
  float4 m4 = make_float4(mean[0], mean[1], mean[2], mean[3]);
  if (y_dt == HATB200_F32) return launch_cl<float>(in, B, H, W, w, bias, OC, m4, range, (float*)y, st);
  if (y_dt == HATB200_BF16) return launch_cl<bf16>(in, B, H, W, w, bias, OC, m4, range, (bf16*)y, st);
  set_error("conv_last: output dtype %d unsupported", y_dt);
  return HATB200_EINVAL;
}

}  // namespace hb
