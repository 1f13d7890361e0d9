// hbm_kernels.cu -- bf16 fast paths of the bandwidth-bound ops (16-byte vector access, register / shared
// memory reuse of the 3x3 halo).  The generic (any dtype, any size) versions live in simt_kernels.cu.
#include "kernels.cuh"
#include "tc_common.cuh"

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
// GatedDconvFFN middle (hat_arch.py:112-116), bf16: depthwise 3x3 (+bias, zero pad) on 2*hid channels, then
// out[j] = dw[j] * SiLU(dw[hid + j]).  in (M, in_pitch) -> out (M, out_pitch); columns [hid, out_pitch) are
// written as zero (K padding of fc2).  Algorithmic bytes: (2*hid + hid) * 2 B per token.
//
// History (ncu, B200): v1 (thread = 8 channels x 4 px of one row, 16-byte loads) issued 89 instructions per output
// value at 1.4 TB/s; v2 (thread = channel pair x 4 px x 8-row strip, every input row loaded once) was latency
// bound at 1.7 TB/s: 12 dependent global loads per row and 18 warps per SM.  v3 below:
//   * a CTA owns a 16 x 8 pixel tile and 64 proj + 64 gate channels; ONE thread issues two 4-D TMA boxes
//     {64 ch, 18 x, 10 y} (halo included; TMA's out-of-bounds zero fill IS the conv padding) into shared memory,
//     four CTAs per SM keep ~180 KB of loads in flight with no registers or address arithmetic spent on them;
//   * a thread owns one channel pair (36 weights in registers), 4 pixels and the 8-row strip: each halo row is
//     read once from shared memory (conflict-free 4-byte reads) and scattered into three rolling accumulator
//     rows with packed FFMA2 (both channels of the pair per instruction);
//   * SiLU through tanh.approx: x * sigmoid(x) = h + h * tanh(h), h = x / 2  (one MUFU instead of EX2 + RCP).
// ---------------------------------------------------------------------------------------------------
constexpr int DW_TW = 16, DW_TH = 8, DW_PX = 4, DW_CC = 64;        // tile, pixels per thread, channels per CTA
constexpr int DW_HW = DW_TW + 2, DW_HH = DW_TH + 2;               // halo
constexpr int DW_HALF_BYTES = DW_HH * DW_HW * DW_CC * 2;          // 23040
constexpr int DW_THREADS = (DW_CC / 2) * (DW_TW / DW_PX);         // 128
constexpr int DW_SMEM = 2 * DW_HALF_BYTES + 128;

__device__ __forceinline__ unsigned long long dw_pk(float a, float b) {
  return (unsigned long long)__float_as_uint(a) | ((unsigned long long)__float_as_uint(b) << 32);
}
// bf16 pair -> packed f32x2 (low half first)
__device__ __forceinline__ unsigned long long dw_unpack(uint32_t v) {
  return (unsigned long long)(v << 16) | ((unsigned long long)(v & 0xffff0000u) << 32);
}
__device__ __forceinline__ unsigned long long dw_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint32_t dw_lds(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ float dw_silu(float x) {
  const float h = 0.5f * x;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  return fmaf(h, t, h);
}

__global__ void __launch_bounds__(DW_THREADS, 4)
dwgate_bf16_kernel(const __grid_constant__ CUtensorMap tmIn, int H, int W, int hid, const float* __restrict__ w,
                   const float* __restrict__ bias, bf16* __restrict__ out, int out_pitch) {
  extern __shared__ __align__(128) uint8_t dw_smem[];          // TMA destination (no swizzle): 128-byte aligned
  uint8_t* sp = dw_smem;
  uint8_t* sg = sp + DW_HALF_BYTES;
  __shared__ __align__(8) uint64_t bar;
  // Blocks are dispatched in blockIdx order, z slowest; fc1 (the producer of `in`) wrote sample 0 first and sample B - 1
  // last, and 94 MB of H1 do not all stay in L2.  Walking the samples (and tiles) BACKWARDS starts with what fc1 wrote
  // last -- still resident -- instead of evicting it while re-fetching the oldest part from HBM.
  const int tiles_x = W / DW_TW;
  const int bxr = (int)(gridDim.x - 1 - blockIdx.x);
  const int tx = bxr % tiles_x, ty = bxr / tiles_x;
  const int x0 = tx * DW_TW, y0 = ty * DW_TH;
  const int c0 = blockIdx.y * DW_CC;
  const int b = (int)(gridDim.z - 1 - blockIdx.z);
  if (threadIdx.x == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_barrier_init();
    pdl_wait();                     // fc1's output; the weight registers below load meanwhile
    pdl_trigger();
    tc::mbar_arrive_expect_tx(&bar, 2 * DW_HALF_BYTES);
    tc::tma_load_4d(sp, &tmIn, &bar, c0, x0 - 1, y0 - 1, b);
    tc::tma_load_4d(sg, &tmIn, &bar, hid + c0, x0 - 1, y0 - 1, b);
  }
  const int pair = threadIdx.x & 31, xq = threadIdx.x >> 5;
  const int c = c0 + 2 * pair;                                    // first channel of this thread's pair
  const int xs = xq * DW_PX;                                      // first pixel column inside the tile
  const size_t img = (size_t)b * H * W;
  bf16* orow = out + (img + (size_t)y0 * W + x0 + xs) * out_pitch + c;
  const bool active = c < hid;            // lanes past hid only write the K padding of fc2's A operand
  const int cl = active ? c : 0;          // keep the (unused) loads of inactive lanes in bounds
  unsigned long long wp[9], wg[9];
#pragma unroll
  for (int t = 0; t < 9; t++) {
    const float2 a = __ldg(reinterpret_cast<const float2*>(w + (size_t)t * 2 * hid + cl));
    const float2 g = __ldg(reinterpret_cast<const float2*>(w + (size_t)t * 2 * hid + hid + cl));
    wp[t] = dw_pk(a.x, a.y);
    wg[t] = dw_pk(g.x, g.y);
  }
  const float2 bpf = __ldg(reinterpret_cast<const float2*>(bias + cl));
  const float2 bgf = __ldg(reinterpret_cast<const float2*>(bias + hid + cl));
  const unsigned long long bp = dw_pk(bpf.x, bpf.y), bg = dw_pk(bgf.x, bgf.y);
  // rolling accumulators: slot 0 = output row r-1, 1 = row r, 2 = row r+1 (r = current halo row)
  unsigned long long ap[3][DW_PX], ag[3][DW_PX];
#pragma unroll
  for (int s = 0; s < 3; s++)
#pragma unroll
    for (int p = 0; p < DW_PX; p++) { ap[s][p] = bp; ag[s][p] = bg; }
  __syncthreads();                                                // barrier initialised and armed (block-uniform point)
  pdl_wait();                                                     // every thread stores: the previous kernel must be done
  tc::mbar_wait(&bar, 0);
  if (!active) {
    if (c < out_pitch)
      for (int y = 0; y < DW_TH; y++)
#pragma unroll
        for (int p = 0; p < DW_PX; p++)
          *reinterpret_cast<uint32_t*>(orow + ((size_t)y * W + p) * out_pitch) = 0u;
    return;
  }
  const uint32_t tp = tc::smem_u32(sp) + (xs * DW_CC + 2 * pair) * 2;   // halo column xs of row 0, this pair
  const uint32_t tg = tp + DW_HALF_BYTES;
#pragma unroll
  for (int i = 0; i < DW_HH; i++) {                               // halo row i = image row y0 - 1 + i
    uint32_t rp[DW_PX + 2], rg[DW_PX + 2];
#pragma unroll
    for (int cx = 0; cx < DW_PX + 2; cx++) {
      rp[cx] = dw_lds(tp + (i * DW_HW + cx) * (DW_CC * 2));
      rg[cx] = dw_lds(tg + (i * DW_HW + cx) * (DW_CC * 2));
    }
#pragma unroll
    for (int cx = 0; cx < DW_PX + 2; cx++) {
      const unsigned long long vp = dw_unpack(rp[cx]), vg = dw_unpack(rg[cx]);
#pragma unroll
      for (int kx = 0; kx < 3; kx++) {
        const int p = cx - kx;                                    // output pixel that sees this column through tap kx
        if (p < 0 || p >= DW_PX) continue;
#pragma unroll
        for (int ky = 0; ky < 3; ky++) {                          // output row (i - 1) + 1 - ky lives in slot 2 - ky
          const int s = 2 - ky, t = ky * 3 + kx;
          if (i - ky < 0 || i - ky >= DW_TH) continue;            // that output row is outside the strip
          ap[s][p] = dw_fma2(vp, wp[t], ap[s][p]);
          ag[s][p] = dw_fma2(vg, wg[t], ag[s][p]);
        }
      }
    }
    if (i >= 2) {                                                 // output row i - 2 of the strip is complete
#pragma unroll
      for (int p = 0; p < DW_PX; p++) {
        const float px_ = __uint_as_float((uint32_t)ap[0][p]), py_ = __uint_as_float((uint32_t)(ap[0][p] >> 32));
        const float gx_ = __uint_as_float((uint32_t)ag[0][p]), gy_ = __uint_as_float((uint32_t)(ag[0][p] >> 32));
        const __nv_bfloat162 h = __floats2bfloat162_rn(px_ * dw_silu(gx_), py_ * dw_silu(gy_));
        *reinterpret_cast<__nv_bfloat162*>(orow + ((size_t)(i - 2) * W + p) * out_pitch) = h;
      }
    }
#pragma unroll
    for (int p = 0; p < DW_PX; p++) {                             // roll (register renaming after full unroll)
      ap[0][p] = ap[1][p]; ap[1][p] = ap[2][p]; ap[2][p] = bp;
      ag[0][p] = ag[1][p]; ag[1][p] = ag[2][p]; ag[2][p] = bg;
    }
  }
}

bool dwgate_bf16_ok(int in_pitch, int W, int hid, int out_pitch, int zero_to) {
  return hid % 8 == 0 && in_pitch % 8 == 0 && out_pitch % 2 == 0 && W % DW_TW == 0 && zero_to == out_pitch &&
         in_pitch >= 2 * hid && tc::encode_tiled_fn() != nullptr;
}

int launch_dwgate_bf16(const bf16* in, int in_pitch, int B, int H, int W, int hid, const float* w, const float* b,
                       bf16* out, int out_pitch, cudaStream_t st) {
  HB_REQUIRE(H % DW_TH == 0 && W % DW_TW == 0, HATB200_EUNSUPPORTED, "dwgate: %dx%d is not a multiple of the 16x8 tile", W, H);
  CUtensorMap tm;
  {
    uint64_t dims[4] = {(uint64_t)in_pitch, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    uint64_t str[3] = {(uint64_t)in_pitch * 2, (uint64_t)W * in_pitch * 2, (uint64_t)H * W * in_pitch * 2};
    uint32_t box[4] = {DW_CC, DW_HW, DW_HH, 1};
    HB_TRY(tc::make_tmap_bf16(&tm, in, 4, dims, str, box, 0));     // inner_bytes 0: no swizzle
  }
  static bool attr = false;
  if (!attr) {
    HB_CHECK_CUDA(cudaFuncSetAttribute(dwgate_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DW_SMEM));
    attr = true;
  }
  dim3 grid((W / DW_TW) * (H / DW_TH), cdiv(out_pitch, DW_CC), B);
  launch_pdl(dwgate_bf16_kernel, grid, dim3(DW_THREADS), DW_SMEM, st, tm, H, W, hid, w, b, out, out_pitch);
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
