// tc_plk.cu -- ESC ConvAttn core (esc_arch.py:119-123) on the tensor cores: the 13x13 dense conv over the first
// 16 channels plus the per-sample dynamic depthwise 3x3, written back into channels [0,16) of the token tensor.
//
// A direct tap-GEMM has N = 16 output channels and K = 16 per tap (169 taps): hopeless tiles.  Instead, FOUR
// neighbouring output pixels share one accumulator row (N = 4 px x 16 oc = 64) and one kernel ROW becomes one
// contraction of K = 16 px x 16 ic = 256 against a banded (Toeplitz) weight matrix
//     Wt[dy][(p, oc)][(j, ic)] = w[oc][ic][dy][j - p]   (0 <= j - p <= 12, else 0),
// 13/16 of the MACs are useful.  The A operand needs no im2col: input pixels live in a zero-padded (6 px each
// side) channels-last buffer, a "row" of A is just 512 contiguous bytes starting at pixel 4*gx, so consecutive A
// rows OVERLAP by 384 bytes.  The CTA loads the (16+12) x (32+12) pixel halo once by TMA (28 x 11 rows of 128 B,
// SWIZZLE_128B) and every (dy, k-block) MMA addresses it in place with a shifted start address and an
// 8-row-group stride of one halo row (1408 B) -- tcgen05 swizzles on absolute address bits, so unaligned starts
// are legal (measured with tc_probe.cu).  Only the 32 KB weight slab per dy streams through a TMA ring.
// Tile: 32 x 16 output pixels = 128 accumulator rows (8 groups x 16 image rows), 64 TMEM columns.
#include <stdlib.h>

#include "kernels.cuh"
#include "tc_common.cuh"
#include "side_jobs.cuh"

namespace hb {

using namespace tc;

namespace {
constexpr int PK = 13, PH = 6, PD = 16;              // kernel size, half, channels
constexpr int TGX = 8, TROWS = 16;                   // tile: 8 pixel groups (32 px) x 16 rows
constexpr int HGX = TGX + 3, HROWS = TROWS + 2 * PH; // halo: 11 groups x 28 rows
constexpr int HROW_BYTES = HGX * 128;                // 1408
constexpr int HALO_BYTES = HROWS * HROW_BYTES;       // 39424
constexpr int BST_BYTES = 4 * 64 * 128;              // one dy slab: 4 k-blocks x 64 rows x 128 B
constexpr int B_STAGES = 4;
constexpr int PLK_SMEM = B_STAGES * BST_BYTES + ((HALO_BYTES + 1023) / 1024) * 1024 + 1024;
constexpr int PLK_THREADS = 192;

struct PlkParams {
  int B, H, W, tiles_x, tiles_y;
  const bf16* xpad;          // (B, H, W + 12, 16) zero padded in x
  const float* dyn;          // (B, 16, 9) per-sample depthwise kernels (may be null)
  bf16* out;                 // token tensor, channels [0,16) of each token are overwritten
  int out_pitch;
  int main_ctas;             // blockIdx.x >= main_ctas: extra CTA running `side` (side_jobs.cuh) for sample blockIdx.x - main_ctas
  SideJob side;
};
}  // namespace

__global__ void __launch_bounds__(PLK_THREADS, 1)
plk_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, const PlkParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t halo_bar, full_bar[B_STAGES], empty_bar[B_STAGES], done_bar;
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float sdk[9 * PD];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  if ((int)blockIdx.x >= p.main_ctas) {
    // extra CTAs: a per-sample job of earlier kernels' results (the ECA gate the NEXT kernel needs), on SMs this
    // launch leaves idle (128 tiles on 148 SMs at the bench shape)
    pdl_wait();
    pdl_trigger();
    run_side_job(p.side, (int)blockIdx.x - p.main_ctas, reinterpret_cast<float*>(smem));
    return;
  }
  uint8_t* sm_b = smem;
  uint8_t* sm_h = smem + B_STAGES * BST_BYTES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  const int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, b = tile / (p.tiles_x * p.tiles_y);
  const int gx0 = tx * TGX, y0 = ty * TROWS;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmX);
    prefetch_tmap(&tmW);
    mbar_init(&halo_bar, 1);
    mbar_init(&done_bar, 1);
    for (int s = 0; s < B_STAGES; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      // the first ring fill is weights only (parameters): it may run ahead of the previous kernel's completion
      auto issue_w = [&](int dy) {
        const int s = dy % B_STAGES;
        mbar_wait(&empty_bar[s], (((uint32_t)(dy / B_STAGES)) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&full_bar[s], BST_BYTES);
        for (int kb = 0; kb < 4; kb++) tma_load_2d(sm_b + s * BST_BYTES + kb * 8192, &tmW, &full_bar[s], kb * 64, dy * 64);
      };
      const int pre = B_STAGES < PK ? B_STAGES : PK;
      for (int dy = 0; dy < pre; dy++) issue_w(dy);
      pdl_wait();
      pdl_trigger();
      mbar_arrive_expect_tx(&halo_bar, HALO_BYTES);
      tma_load_4d(sm_h, &tmX, &halo_bar, 0, gx0, y0 - PH, b);
      for (int dy = pre; dy < PK; dy++) issue_w(dy);
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(128, 64);
      mbar_wait(&halo_bar, 0);
      tc_fence_after();
      const uint32_t h_addr = smem_u32(sm_h);
      for (int dy = 0; dy < PK; dy++) {
        const int s = dy % B_STAGES;
        mbar_wait(&full_bar[s], ((uint32_t)(dy / B_STAGES)) & 1u);
        tc_fence_after();
        const uint32_t b_addr = smem_u32(sm_b + s * BST_BYTES);
#pragma unroll
        for (int kb = 0; kb < 4; kb++) {
#pragma unroll
          for (int k = 0; k < 4; k++) {
            // A: accumulator row (y, gx) reads halo row y + dy, group gx + kb; 8-row groups are one halo row apart
            uint64_t ad = make_smem_desc(h_addr + dy * HROW_BYTES + kb * 128 + k * 32, 128);
            ad &= ~(0x3FFFull << 32);
            ad |= (uint64_t)(HROW_BYTES >> 4) << 32;
            const uint64_t bd = make_smem_desc(b_addr + kb * 8192 + k * 32, 128);
            umma_bf16(tmem_base, ad, bd, idesc, (dy | kb | k) != 0 ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[s]);
      }
      umma_commit(&done_bar);
    }
  } else {
    // ===== epilogue: one accumulator row (4 pixels x 16 channels) per thread =====
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int y = y0 + row / TGX, x = (gx0 + row % TGX) * 4;
    const int WP = p.W + 2 * PH;
    pdl_wait();          // reads the per-sample dynamic kernels and the padded head, writes the output
    // this sample's dynamic 3x3 kernels, tap-major, staged once: the loop below reads them as broadcast float4s
    // (144 scalar read-only loads per pixel before)
    if (p.dyn) {
      for (int j = threadIdx.x - 64; j < PD * 9; j += 128) sdk[(j % 9) * PD + j / 9] = p.dyn[(size_t)b * PD * 9 + j];
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }
    mbar_wait(&halo_bar, 0);      // the TMA-written halo is read with plain shared loads below: observe its barrier
    mbar_wait(&done_bar, 0);
    tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t h_sm = smem_u32(sm_h);
    {
      const bool dk = p.dyn != nullptr;
      const bool live = x < p.W;          // half-empty tiles at the right image edge (W % 32 == 16)
#pragma unroll
      for (int px = 0; px < 4; px++) {
        uint32_t r[16];
        tmem_ld16(trow + px * 16, r);     // warp-collective (.sync.aligned): every lane executes it
        tmem_ld_wait();
        if (!live) continue;
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = __uint_as_float(r[j]);
        if (dk) {      // + conv2d(x, dyn, padding 1, groups 16)  (esc_arch.py:123)
#pragma unroll
          // The 3x3 neighbourhood comes from the halo the MMAs just read (shared memory, SWIZZLE_128B lines of 4 px x 16 ch;
          // rows outside the image are TMA zero fill, columns outside are the buffer's own zero padding) -- the same
          // pixels through global loads cost 6.4 of the launch's 24.7 us (18 dependent 32-byte loads per pixel).
          for (int ky = 0; ky < 3; ky++) {
            const int hr = row / TGX + PH + ky - 1;                   // halo row
#pragma unroll
            for (int kx = 0; kx < 3; kx++) {
              const int xr = (row % TGX) * 4 + px + kx + PH - 1;      // pixel inside the halo row
              const int L = hr * HGX + (xr >> 2), q = xr & 3;         // 128-byte line, pixel inside it
              const uint32_t la = h_sm + (uint32_t)L * 128u;
              uint4 t0, t1;
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(t0.x), "=r"(t0.y), "=r"(t0.z), "=r"(t0.w)
                           : "r"(la + (uint32_t)(((2 * q) ^ (L & 7)) << 4)));
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(t1.x), "=r"(t1.y), "=r"(t1.z), "=r"(t1.w)
                           : "r"(la + (uint32_t)(((2 * q + 1) ^ (L & 7)) << 4)));
              const __nv_bfloat162* h0 = reinterpret_cast<const __nv_bfloat162*>(&t0);
              const __nv_bfloat162* h1 = reinterpret_cast<const __nv_bfloat162*>(&t1);
              const float4* wk = reinterpret_cast<const float4*>(sdk + (ky * 3 + kx) * PD);
              const float4 w0 = wk[0], w1 = wk[1], w2 = wk[2], w3 = wk[3];
              const float wv[16] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w,
                                    w2.x, w2.y, w2.z, w2.w, w3.x, w3.y, w3.z, w3.w};
#pragma unroll
              for (int j = 0; j < 4; j++) {
                const float2 f0 = __bfloat1622float2(h0[j]), f1 = __bfloat1622float2(h1[j]);
                v[2 * j] = fmaf(f0.x, wv[2 * j], v[2 * j]);
                v[2 * j + 1] = fmaf(f0.y, wv[2 * j + 1], v[2 * j + 1]);
                v[8 + 2 * j] = fmaf(f1.x, wv[8 + 2 * j], v[8 + 2 * j]);
                v[9 + 2 * j] = fmaf(f1.y, wv[9 + 2 * j], v[9 + 2 * j]);
              }
            }
          }
        }
        uint4 o[2];
        __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(o);
#pragma unroll
        for (int j = 0; j < 8; j++) oh[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
        bf16* dst = p.out + (((size_t)b * p.H + y) * p.W + x + px) * p.out_pitch;
        reinterpret_cast<uint4*>(dst)[0] = o[0];
        reinterpret_cast<uint4*>(dst)[1] = o[1];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

// ---- pack: (16, 16, 13, 13) -> banded Wt[dy][(p, oc)][(j, ic)] bf16, 13 x 64 x 256 ----
__global__ void pack_plk_toeplitz_kernel(const void* __restrict__ src, int dt, bf16* __restrict__ dst) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= PK * 64 * 256) return;
  const int k = idx % 256, n = (idx / 256) % 64, dy = idx / (256 * 64);
  const int j = k / PD, ic = k % PD, pp = n / PD, oc = n % PD;
  const int dx = j - pp;
  float v = 0.f;
  if (dx >= 0 && dx < PK) {
    const size_t s = (((size_t)oc * PD + ic) * PK + dy) * PK + dx;
    v = dt == HATB200_F32 ? ((const float*)src)[s] : (dt == HATB200_BF16 ? to_f(((const bf16*)src)[s]) : to_f(((const __half*)src)[s]));
  }
  dst[idx] = __float2bfloat16_rn(v);
}
int launch_pack_plk_toeplitz(const void* src, int dt, bf16* dst, cudaStream_t st) {
  pack_plk_toeplitz_kernel<<<(PK * 64 * 256 + 255) / 256, 256, 0, st>>>(src, dt, dst);
  HB_CHECK_LAUNCH();
  return 0;
}
size_t plk_toeplitz_elems() { return (size_t)PK * 64 * 256; }

// compact (M, 16) -> x-padded (B, H, W + 12, 16); pad columns must already be zero
__global__ void plk_pad_copy_kernel(const bf16* __restrict__ src, int B, int H, int W, bf16* __restrict__ dst) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;       // one 16-byte half token per thread
  size_t total = (size_t)B * H * W * 2;
  if (idx >= total) return;
  size_t tok = idx >> 1;
  int half = (int)(idx & 1);
  int x = (int)(tok % W);
  size_t by = tok / W;
  reinterpret_cast<uint4*>(dst + (by * (W + 2 * PH) + x + PH) * PD)[half] = reinterpret_cast<const uint4*>(src + tok * PD)[half];
}
int launch_plk_pad_copy(const bf16* src, int B, int H, int W, bf16* dst, cudaStream_t st) {
  size_t total = (size_t)B * H * W * 2;
  plk_pad_copy_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, B, H, W, dst);
  HB_CHECK_LAUNCH();
  return 0;
}

bool plk_tc_supported(int pdim, int ksize, int H, int W, int out_pitch) {
  if (pdim != PD || ksize != PK) return false;
  if (H % TROWS || W % 4 || (out_pitch * 2) % 16) return false;
  if (tc::encode_tiled_fn() == nullptr) return false;
  const char* v = getenv("HATB200_PLK_V2");
  if (v && *v == '0') return false;
  return true;
}

int launch_plk_tc(const bf16* xpad, int B, int H, int W, const bf16* wt, const float* dyn, bf16* out, int out_pitch,
                  cudaStream_t st, const SideJob* side) {
  CUtensorMap tmX, tmW;
  const int G = (W + 2 * PH) / 4;
  {
    uint64_t dims[4] = {64, (uint64_t)G, (uint64_t)H, (uint64_t)B};
    uint64_t str[3] = {128, (uint64_t)G * 128, (uint64_t)H * G * 128};
    uint32_t box[4] = {64, HGX, HROWS, 1};
    HB_TRY(tc::make_tmap_bf16(&tmX, xpad, 4, dims, str, box, 128));
  }
  {
    uint64_t dims[2] = {256, (uint64_t)PK * 64};
    uint64_t str[1] = {512};
    uint32_t box[2] = {64, 64};
    HB_TRY(tc::make_tmap_bf16(&tmW, wt, 2, dims, str, box, 128));
  }
  PlkParams p;
  p.B = B; p.H = H; p.W = W;
  p.tiles_x = cdiv(W, TGX * 4); p.tiles_y = H / TROWS;
  p.xpad = xpad; p.dyn = dyn; p.out = out; p.out_pitch = out_pitch;
  static bool attr_set = false;
  if (!attr_set) {
    HB_CHECK_CUDA(cudaFuncSetAttribute(plk_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PLK_SMEM));
    attr_set = true;
  }
  p.main_ctas = B * p.tiles_x * p.tiles_y;
  if (side) p.side = *side;
  launch_pdl(plk_tc_kernel, dim3(p.main_ctas + (side ? side->B : 0)), dim3(PLK_THREADS), PLK_SMEM, st, tmX, tmW, p);
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb
