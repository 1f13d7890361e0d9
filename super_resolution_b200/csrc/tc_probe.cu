// tc_probe.cu -- engineering probe (tests only): which shared-memory rows does a tcgen05 K-major SWIZZLE_128B
// A descriptor read when its start address is NOT 1024-byte aligned and its 8-row groups are SBO bytes apart?
// A (R rows x 64 bf16, one 128-byte row each) is TMA-loaded into shared memory as one SWIZZLE_128B box;
// B (64 x 64) is whatever the caller passes (identity -> D reproduces the rows the MMA actually read).
// D = A_desc * B^T with M = 128, N = 64, K = 64; A_desc starts at row r0, group stride sbo, base-offset field bo.
// The answer decides whether conv halos / Toeplitz windows can be addressed in place (tc_plk.cu).
#include "kernels.cuh"
#include "tc_common.cuh"

namespace hb {
using namespace tc;

__global__ void __launch_bounds__(128, 1)
umma_probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int a_bytes, int r0,
                  int sbo, int bo, float* __restrict__ out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar, done;
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sb = smem;               // 64 x 128 B
  uint8_t* sa = smem + 8192;        // R x 128 B
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_init(&done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bar, (uint32_t)(8192 + a_bytes));
    tma_load_2d(sb, &tmB, &bar, 0, 0);
    for (int r = 0; r * 128 < a_bytes; r += 128) tma_load_2d(sa + r * 128, &tmA, &bar, 0, r);
    mbar_wait(&bar, 0);
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(128, 64);
    for (int k = 0; k < 4; k++) {
      uint64_t ad = make_smem_desc(smem_u32(sa) + r0 * 128 + k * 32, 128);
      ad &= ~(0x3FFFull << 32);
      ad |= (uint64_t)((uint32_t)sbo >> 4) << 32;
      ad |= (uint64_t)(bo & 7) << 49;
      const uint64_t bd = make_smem_desc(smem_u32(sb) + k * 32, 128);
      umma_bf16(tmem_base, ad, bd, idesc, k);
    }
    umma_commit(&done);
  }
  mbar_wait(&done, 0);
  tc_fence_after();
  const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16);
  for (int c = 0; c < 4; c++) {
    uint32_t r[16];
    tmem_ld16(trow + c * 16, r);
    tmem_ld_wait();
    for (int j = 0; j < 16; j++) out[(size_t)(warp * 32 + lane) * 64 + c * 16 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}
}  // namespace hb

using namespace hb;

// a: (R, 64) bf16 device, R multiple of 128 and <= 1024; b: (64, 64) bf16 device; out: (128, 64) fp32 device
extern "C" int hatb200_probe_umma(const void* a, int R, const void* b, int r0, int sbo_bytes, int base_offset,
                                  float* out, void* stream) {
  HB_REQUIRE(a && b && out && R % 128 == 0 && R >= 128 && R <= 1024, HATB200_EINVAL, "probe_umma: bad arguments");
  CUtensorMap tmA, tmB;
  {
    uint64_t dims[2] = {64, (uint64_t)R};
    uint64_t str[1] = {128};
    uint32_t box[2] = {64, 128};
    HB_TRY(tc::make_tmap_bf16(&tmA, a, 2, dims, str, box, 128));
  }
  {
    uint64_t dims[2] = {64, 64};
    uint64_t str[1] = {128};
    uint32_t box[2] = {64, 64};
    HB_TRY(tc::make_tmap_bf16(&tmB, b, 2, dims, str, box, 128));
  }
  size_t smem = 8192 + (size_t)R * 128 + 1024;
  HB_CHECK_CUDA(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  umma_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(tmA, tmB, R * 128, r0, sbo_bytes, base_offset, out);
  HB_CHECK_LAUNCH();
  return 0;
}

// ---- probe 2: tcgen05.cp.128x256b (shared memory -> TMEM, issued by one thread).  Does a 128-row x 32-fp32 tile that
// TMA wrote with SWIZZLE_128B (mode 0) / as four 8-column SWIZZLE_32B boxes (mode 1) land as "row r -> lane r,
// column c -> TMEM column c" when copied 8 columns (256 bits) at a time with the matching K-major descriptor?
// If yes, the fp32 residual can be PRELOADED into the accumulator and the MMA accumulates on top of it.
namespace hb {
__global__ void __launch_bounds__(128, 1)
tmem_cp_probe_kernel(const __grid_constant__ CUtensorMap tm, int mode, float* __restrict__ out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar, done;
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_init(&done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&tmem_slot, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bar, 128 * 128);
    if (mode == 0) tma_load_2d(smem, &tm, &bar, 0, 0);
    else for (int i = 0; i < 4; i++) tma_load_2d(smem + i * 4096, &tm, &bar, 8 * i, 0);
    mbar_wait(&bar, 0);
    tc_fence_after();
    for (int i = 0; i < 4; i++) {
      const uint64_t d = mode == 0 ? make_smem_desc(smem_u32(smem) + 32 * i, 128) : make_smem_desc(smem_u32(smem) + 4096 * i, 32);
      asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(tmem_base + 8 * i), "l"(d) : "memory");
    }
    umma_commit(&done);
  }
  mbar_wait(&done, 0);
  tc_fence_after();
  const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16);
  uint32_t r[32];
  tmem_ld32(trow, r);
  tmem_ld_wait();
  for (int j = 0; j < 32; j++) out[(size_t)(warp * 32 + lane) * 32 + j] = __uint_as_float(r[j]);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 32);
  }
}
}  // namespace hb

// src, out: (128, 32) fp32 device
extern "C" int hatb200_probe_tmem_cp(const float* src, int mode, float* out, void* stream) {
  CUtensorMap tm;
  uint64_t dims[2] = {32, 128};
  uint64_t str[1] = {128};
  uint32_t box[2] = {mode == 0 ? 32u : 8u, 128};
  HB_TRY(tc::make_tmap_f32(&tm, src, 2, dims, str, box, mode == 0 ? 128 : 32));
  const size_t smem = 16384 + 1024;
  HB_CHECK_CUDA(cudaFuncSetAttribute(tmem_cp_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tmem_cp_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(tm, mode, out);
  HB_CHECK_LAUNCH();
  return 0;
}

// ---- probe 3: what does the memory system give a kernel with the esc_aggr traffic pattern and no arithmetic?
// o32 = r + a + u (fp32), o16 = bf16(o32): reads two bf16 streams + one fp32 stream, writes one fp32 + one bf16 stream,
// 16-byte accesses, grid-stride, `ctas` CTAs of 1024 threads.  n8 = number of 8-element groups.
namespace hb {
__global__ void __launch_bounds__(1024) stream_probe_kernel(const uint4* __restrict__ a, const uint4* __restrict__ u,
                                                            const float4* __restrict__ r, float4* __restrict__ o32,
                                                            uint4* __restrict__ o16, size_t n8, int rev) {
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n8; i0 += (size_t)gridDim.x * blockDim.x) {
    const size_t i = rev ? n8 - 1 - i0 : i0;
    const uint4 av = a[i], uv = u[i];
    const float4 r0 = r[2 * i], r1 = r[2 * i + 1];
    const uint32_t aw[4] = {av.x, av.y, av.z, av.w}, uw[4] = {uv.x, uv.y, uv.z, uv.w};
    float v[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      v[2 * j] += __uint_as_float(aw[j] << 16) + __uint_as_float(uw[j] << 16);
      v[2 * j + 1] += __uint_as_float(aw[j] & 0xffff0000u) + __uint_as_float(uw[j] & 0xffff0000u);
      __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
      o[j] = *reinterpret_cast<uint32_t*>(&h);
    }
    o32[2 * i] = make_float4(v[0], v[1], v[2], v[3]);
    o32[2 * i + 1] = make_float4(v[4], v[5], v[6], v[7]);
    o16[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}
}  // namespace hb
extern "C" int hatb200_probe_stream(const void* a, const void* u, const void* r, void* o32, void* o16, size_t n8, int ctas,
                                    int rev, void* stream) {
  stream_probe_kernel<<<ctas, 1024, 0, (cudaStream_t)stream>>>((const uint4*)a, (const uint4*)u, (const float4*)r,
                                                               (float4*)o32, (uint4*)o16, n8, rev);
  HB_CHECK_LAUNCH();
  return 0;
}

// engineering: copy the last traced GEMM launch's per-CTA event timestamps (ns) to the host: (1024, 64) uint64
namespace hb { unsigned long long* gemm_trace_buffer(); }
extern "C" int hatb200_debug_gemm_trace(unsigned long long* host, void* stream) {
  unsigned long long* d = hb::gemm_trace_buffer();
  HB_REQUIRE(d && host, HATB200_EINVAL, "gemm trace not enabled (HATB200_GEMM_TRACE=1)");
  HB_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  HB_CHECK_CUDA(cudaMemcpy(host, d, 1024 * 64 * 8, cudaMemcpyDeviceToHost));
  return 0;
}
