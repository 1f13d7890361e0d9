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

// engineering: copy the last traced GEMM launch's per-CTA event timestamps (ns) to the host: (1024, 64) uint64
namespace hb { unsigned long long* gemm_trace_buffer(); }
extern "C" int hatb200_debug_gemm_trace(unsigned long long* host, void* stream) {
  unsigned long long* d = hb::gemm_trace_buffer();
  HB_REQUIRE(d && host, HATB200_EINVAL, "gemm trace not enabled (HATB200_GEMM_TRACE=1)");
  HB_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  HB_CHECK_CUDA(cudaMemcpy(host, d, 1024 * 64 * 8, cudaMemcpyDeviceToHost));
  return 0;
}
