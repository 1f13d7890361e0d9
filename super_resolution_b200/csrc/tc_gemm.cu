// tc_gemm.cu -- Linear / 1x1 conv over channels-last tokens as a PERSISTENT, weight-stationary tcgen05 GEMM.
//
//   out[m, n] = sum_k A[m, k] * W[n, k]          (bf16 x bf16 -> fp32 in TMEM), fused epilogue (tc_epilogue.cuh)
//
// Why a second kernel next to tc_tapconv.cu: a 1x1 layer has K <= 384, so a 128-row tile does only 3-6
// k-blocks of MMA work while the one-tile-per-CTA kernel re-streams the whole weight panel (up to
// 147 KB) from L2 for every tile and pays TMEM allocation + barrier setup + an un-overlapped epilogue
// per tile -- measured 35-110 TFLOP/s, L2->SM bandwidth bound.  Here:
//   * one CTA per SM (grid <= #SMs), each CTA owns one n-tile (<= 256 output channels) and walks the
//     128-token tiles m = blockIdx.x, blockIdx.x + gridDim.x, ...;
//   * the weight panel (BN x Kpad bf16) is loaded ONCE per CTA by TMA and stays in shared memory;
//   * only the activation tile (128 x 64 bf16 per k-block, SWIZZLE_128B) streams through a TMA ring;
//   * two TMEM accumulator stages: while epilogue warpgroup g drains tile i, the MMA warp already
//     accumulates tile i+1 into the other stage (warpgroup 0 takes even tiles, warpgroup 1 odd tiles,
//     so a token row still lives in ONE thread and the fused LayerNorm needs no cross-thread reduce).
// Warp roles (320 threads): warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2-5 / 6-9 the
// two epilogue warpgroups (TMEM lane quarter = warp % 4).
// Replaces F.linear / 1x1 F.conv2d at hat_arch.py:111,:117,:347-351,:391-392 and esc_arch.py:144.
#include <stdlib.h>

#include "kernels.cuh"
#include "tc_epilogue2.cuh"

namespace hb {

using namespace tc;

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define GM_TRACE(slot) do { if (p.trace) p.trace[((size_t)(blockIdx.y * gridDim.x + blockIdx.x)) * 64 + (slot)] = gtime_ns(); } while (0)

constexpr int GM_THREADS = 320;
constexpr int GM_MAX_STAGES = 8;
constexpr int GM_TILE_M = 128;
constexpr int GM_BK = 64;
constexpr int GM_DYN_SMEM = 227 * 1024 - 18 * 1024;   // static: 16 KB store staging + barriers   // static shared memory (barriers, 8 KB reduction scratch) stays below 10 KB

struct GemmParams {
  int M, H, W;              // tokens; image size (pixel coordinates / per-sample chunks for the epilogue)
  int a_coff, kblocks;      // first input channel, Kpad / 64
  int N, BN, Npad;
  int m_tiles;              // M / 128
  int stages;               // activation ring depth
  int b_bytes;              // BN * Kpad * 2: resident weight panel
  uint32_t acc_stride;      // TMEM columns between the two accumulator stages
  uint32_t tmem_cols;       // allocation (power of two)
  int op_resid, op_u;       // epilogue operands fed by TMA (fp32 residual stream, bf16 CAB branch)
  int op_stage_bytes;       // bytes of one operand-ring stage (resid chunk + U chunk), 1024-aligned
  int op_u_off;
  int dbg;                     // engineering bit mask (HATB200_GEMM_DBG): 1 = no out_t store, 2 = no bias
  unsigned long long* trace;   // engineering: per-CTA event timestamps (HATB200_GEMM_TRACE=1), else null
};

template <int CW>
__global__ void __launch_bounds__(GM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmU, const GemmParams p,
               const Epilogue ep) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[GM_MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[GM_MAX_STAGES];
  __shared__ __align__(8) uint64_t b_bar;
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ __align__(8) uint64_t op_full[2][2];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) uint8_t wstage[8][2048];   // per epilogue warp: store staging (32 rows x 64 B)
  __shared__ float hsum[2][4][16];                    // per warpgroup: column sums of the LN head

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_b = smem;                                   // [kblocks][BN rows x 128 B], SWIZZLE_128B
  uint8_t* smem_a = smem + ((p.b_bytes + 1023) & ~1023);    // [stages][128 rows x 128 B]
  constexpr int A_BYTES = GM_TILE_M * GM_BK * 2;
  uint8_t* smem_op = smem_a + (size_t)p.stages * A_BYTES;   // [2 warpgroups][2 stages][op_stage_bytes]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.y * p.BN;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < p.stages; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&b_bar, 1);
    for (int g = 0; g < 2; g++) {
      mbar_init(&tmem_full_bar[g], 1);
      mbar_init(&tmem_empty_bar[g], 128);
      mbar_init(&op_full[g][0], 1);
      mbar_init(&op_full[g][1], 1);
    }
    if (p.op_resid) prefetch_tmap(&tmR);
    if (p.op_u) prefetch_tmap(&tmU);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      mbar_arrive_expect_tx(&b_bar, (uint32_t)p.b_bytes);
      for (int kb = 0; kb < p.kblocks; kb++)
        tma_load_2d(smem_b + (size_t)kb * p.BN * 128, &tmB, &b_bar, kb * GM_BK, n0);
      int it = 0;
      for (int t = blockIdx.x; t < p.m_tiles; t += gridDim.x) {
        for (int kb = 0; kb < p.kblocks; kb++, it++) {
          const int s = it % p.stages;
          const uint32_t par = (uint32_t)(it / p.stages) & 1u;
          mbar_wait(&empty_bar[s], par ^ 1u);
          mbar_arrive_expect_tx(&full_bar[s], A_BYTES);
          tma_load_2d(smem_a + (size_t)s * A_BYTES, &tmA, &full_bar[s], p.a_coff + kb * GM_BK, t * GM_TILE_M);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(GM_TILE_M, p.BN);
      GM_TRACE(0);
      mbar_wait(&b_bar, 0);
      GM_TRACE(1);
      int it = 0, i = 0;
      for (int t = blockIdx.x; t < p.m_tiles; t += gridDim.x, i++) {
        const int g = i & 1;
        mbar_wait(&tmem_empty_bar[g], (((uint32_t)i >> 1) & 1u) ^ 1u);     // epilogue has drained this stage
        tc_fence_after();
        if (i < 6) GM_TRACE(8 + 2 * i);
        const uint32_t acc = tmem_base + (uint32_t)g * p.acc_stride;
        for (int kb = 0; kb < p.kblocks; kb++, it++) {
          const int s = it % p.stages;
          const uint32_t par = (uint32_t)(it / p.stages) & 1u;
          mbar_wait(&full_bar[s], par);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem_a + (size_t)s * A_BYTES);
          const uint32_t sb = smem_u32(smem_b + (size_t)kb * p.BN * 128);
#pragma unroll
          for (int k = 0; k < GM_BK / 16; k++) {
            const uint64_t ad = make_smem_desc(sa + k * 32, 128);
            const uint64_t bd = make_smem_desc(sb + k * 32, 128);
            umma_bf16(acc, ad, bd, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);       // frees the activation slot once these MMAs have read it
        }
        umma_commit(&tmem_full_bar[g]);     // accumulator of tile i complete
        if (i < 6) GM_TRACE(9 + 2 * i);
      }
    }
  } else {
    // ===== epilogue warpgroups: warps 2-5 take even local tiles, warps 6-9 odd ones =====
    const int g = (warp - 2) >> 2;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int hw = p.H * p.W;
    EpiCtx c;
    c.H = p.H; c.W = p.W; c.N = p.N; c.n0 = n0; c.BN = p.BN;
    c.last_ntile = blockIdx.y == gridDim.y - 1;
    c.nchunk = hw / GM_TILE_M;
    c.quarter = quarter; c.lane = lane; c.gtid = (threadIdx.x - 64) & 127; c.bar_id = 1 + g;
    c.sred = nullptr;
    c.wstage = wstage[warp - 2];
    c.hsum = hsum[g];
    c.trow = tmem_base + (uint32_t)g * p.acc_stride + ((uint32_t)(quarter * 32) << 16);
    // operand ring of this warpgroup: residual / CAB-branch chunks arrive by TMA two chunks ahead
    OpRing ring;
    OpRing* rg = nullptr;
    if (p.op_resid || p.op_u) {
      rg = &ring;
      ring.stage[0] = smem_op + (size_t)(g * 2) * p.op_stage_bytes;
      ring.stage[1] = ring.stage[0] + p.op_stage_bytes;
      ring.full[0] = &op_full[g][0];
      ring.full[1] = &op_full[g][1];
      ring.tm_resid = &tmR; ring.tm_u = &tmU;
      ring.has_resid = p.op_resid != 0; ring.has_u = p.op_u != 0;
      ring.bytes = (uint32_t)((p.op_resid ? GM_TILE_M * CW * 4 : 0) + (p.op_u ? GM_TILE_M * CW * 2 : 0));
      ring.u_off = p.op_u_off;
      ring.nch = p.BN / CW;
      ring.tile_stride = 2 * gridDim.x;
      ring.m_tiles = p.m_tiles;
      ring.k = 0;
      if (c.gtid == 0) {                 // prologue: the first two chunks of this warpgroup's first tile(s)
        int tn = blockIdx.x + g * gridDim.x, cn = 0;
        for (uint32_t kk = 0; kk < 2; kk++) {
          if (tn < p.m_tiles) ring_issue<CW>(ring, kk, tn, cn, n0);
          if (++cn >= ring.nch) { cn = 0; tn += ring.tile_stride; }
        }
      }
    }
    int i = 0;
    for (int t = blockIdx.x; t < p.m_tiles; t += gridDim.x, i++) {
      if ((i & 1) != g) continue;
      const int m = t * GM_TILE_M + row;
      c.m = m;
      c.b = m / hw;
      const int rem = m - c.b * hw;
      c.py = rem / p.W; c.px = rem - c.py * p.W;
      c.chunk = (t * GM_TILE_M - c.b * hw) / GM_TILE_M;
      const bool tr = c.gtid == 0 && (i >> 1) < 4;
      if (tr) GM_TRACE(24 + g * 16 + 4 * (i >> 1));
      mbar_wait(&tmem_full_bar[g], ((uint32_t)i >> 1) & 1u);
      tc_fence_after();
      if (tr) GM_TRACE(25 + g * 16 + 4 * (i >> 1));
      run_epilogue2<CW>(ep, c, rg, row, t);
      tc_fence_before();
      mbar_arrive(&tmem_empty_bar[g]);
      if (tr) GM_TRACE(26 + g * 16 + 4 * (i >> 1));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ---- host ------------------------------------------------------------------------------------------
static unsigned long long* g_gemm_trace = nullptr;
unsigned long long* gemm_trace_buffer() { return g_gemm_trace; }

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        n <= 0)
      n = 148;
  }
  return n;
}

static bool gemm_geometry(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep, GemmParams& p, int& cw) {
  if (a.KH != 1 || a.KW != 1) return false;
  const long long M = (long long)a.B * a.H * a.W;
  if (M % GM_TILE_M || (a.H * a.W) % GM_TILE_M) return false;
  if (Kpad % GM_BK || Kpad < a.Cin) return false;
  if ((a.a_pitch * 2) % 16 || (a.a_coff * 2) % 16) return false;
  p.BN = tc_bn(a.N);
  if (Npad != tc_npad(a.N) || p.BN < 16 || p.BN > 256 || p.BN % 16) return false;
  // epilogue features / alignment the second-generation epilogue relies on (else: tc_tapconv.cu handles the layer)
  if (ep.colsum || ep.dyn_k || ep.shuffle_r != 1) return false;
  if (a.N % 4) return false;
  if (ep.out_f32 && (ep.out_f32_pitch % 4 || ((uintptr_t)ep.out_f32 & 15))) return false;
  if (ep.out_t && (ep.out_t_pitch % 8 || ep.out_t_coff % 8 || ((uintptr_t)ep.out_t & 15))) return false;
  if (ep.out_t && (ep.out_t_zero_to % 8 || (a.N % 8 && ep.out_t_zero_to < round_up(a.N, 8)))) return false;   // 16-byte store granules
  if (ep.ln_out && ep.ln_pitch < round_up(a.N, 8)) return false;
  if (ep.ln_out && (ep.ln_pitch % 8 || ((uintptr_t)ep.ln_out & 15))) return false;
  if (ep.resid && (ep.resid_pitch % 4 || ((uintptr_t)ep.resid & 15))) return false;
  if (ep.eca_u && (ep.eca_pitch % 8 || ((uintptr_t)ep.eca_u & 15) || ep.eca_pitch < round_up(a.N, 32))) return false;
  p.M = (int)M; p.H = a.H; p.W = a.W;
  p.a_coff = a.a_coff; p.kblocks = Kpad / GM_BK;
  p.N = a.N; p.Npad = Npad;
  p.m_tiles = (int)(M / GM_TILE_M);
  p.b_bytes = p.BN * Kpad * 2;
  p.acc_stride = (uint32_t)round_up(p.BN, 32);
  uint32_t cols = 32;
  while (cols < 2 * p.acc_stride) cols <<= 1;
  p.tmem_cols = cols;
  p.op_resid = ep.resid != nullptr; p.op_u = ep.eca_u != nullptr;
  const int budget = GM_DYN_SMEM - 1024 - round_up(p.b_bytes, 1024);
  for (cw = (p.BN % 32 == 0) ? 32 : 16; cw >= 16; cw -= 16) {
    const int rb = p.op_resid ? GM_TILE_M * cw * 4 : 0, ub = p.op_u ? GM_TILE_M * cw * 2 : 0;
    p.op_u_off = rb;
    p.op_stage_bytes = round_up(rb + ub, 1024);
    int st = (budget - 4 * p.op_stage_bytes) / (GM_TILE_M * GM_BK * 2);
    if (st >= 2) {
      p.stages = st > GM_MAX_STAGES ? GM_MAX_STAGES : st;
      return true;
    }
  }
  return false;
}

bool gemm_tc_supported(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep) {
  GemmParams p;
  int cw;
  if (!gemm_geometry(a, Npad, Kpad, ep, p, cw)) return false;
  if (ep.ln_out && (Npad != p.BN || ep.ln_pitch > 512)) return false;     // LN needs the whole row in one CTA
  if (ep.ln_head && ep.ln_head_c > 16) return false;
  if (tc::encode_tiled_fn() == nullptr) return false;
  const char* v = getenv("HATB200_GEMM_V2");
  if (v && *v == '0') return false;
  return true;
}

int launch_gemm_tc(const ConvArgs& a, const bf16* w, int Npad, int Kpad, const Epilogue& ep, cudaStream_t st) {
  GemmParams p;
  int cw;
  HB_REQUIRE(gemm_geometry(a, Npad, Kpad, ep, p, cw), HATB200_EUNSUPPORTED, "gemm_tc: unsupported shape");
  CUtensorMap tmA, tmB, tmR, tmU;
  {
    uint64_t dims[2] = {(uint64_t)a.a_pitch, (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)a.a_pitch * 2};
    uint32_t box[2] = {GM_BK, GM_TILE_M};
    HB_TRY(tc::make_tmap_bf16(&tmA, a.A, 2, dims, str, box, GM_BK * 2));
  }
  {
    uint64_t dims[2] = {(uint64_t)Kpad, (uint64_t)Npad};
    uint64_t str[1] = {(uint64_t)Kpad * 2};
    uint32_t box[2] = {GM_BK, (uint32_t)p.BN};
    HB_TRY(tc::make_tmap_bf16(&tmB, w, 2, dims, str, box, GM_BK * 2));
  }
  tmR = tmA; tmU = tmA;     // placeholders when the operand is absent (never dereferenced)
  if (p.op_resid) {
    uint64_t dims[2] = {(uint64_t)p.N, (uint64_t)p.M};            // columns >= N read as zero
    uint64_t str[1] = {(uint64_t)ep.resid_pitch * 4};
    uint32_t box[2] = {(uint32_t)cw, GM_TILE_M};
    HB_TRY(tc::make_tmap_f32(&tmR, ep.resid, 2, dims, str, box, cw * 4));
  }
  if (p.op_u) {
    uint64_t dims[2] = {(uint64_t)ep.eca_pitch, (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)ep.eca_pitch * 2};
    uint32_t box[2] = {(uint32_t)cw, GM_TILE_M};
    HB_TRY(tc::make_tmap_bf16(&tmU, ep.eca_u, 2, dims, str, box, cw * 2));
  }
  p.trace = nullptr;
  { const char* d = getenv("HATB200_GEMM_DBG"); p.dbg = d ? atoi(d) : 0; }
  if (getenv("HATB200_GEMM_TRACE")) {
    if (!g_gemm_trace) {
      HB_CHECK_CUDA(cudaMalloc(&g_gemm_trace, 1024 * 64 * 8));
    }
    HB_CHECK_CUDA(cudaMemsetAsync(g_gemm_trace, 0, 1024 * 64 * 8, st));
    p.trace = g_gemm_trace;
  }
  const size_t smem = (size_t)round_up(p.b_bytes, 1024) + (size_t)p.stages * GM_TILE_M * GM_BK * 2 +
                      4 * (size_t)p.op_stage_bytes + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    HB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, GM_DYN_SMEM));
    HB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, GM_DYN_SMEM));
    attr_set = true;
  }
  const int n_tiles = Npad / p.BN;
  int per_n = sm_count() / n_tiles;
  if (per_n < 1) per_n = 1;
  if (per_n > p.m_tiles) per_n = p.m_tiles;
  dim3 grid(per_n, n_tiles);
  if (cw == 32) gemm_tc_kernel<32><<<grid, GM_THREADS, smem, st>>>(tmA, tmB, tmR, tmU, p, ep);
  else gemm_tc_kernel<16><<<grid, GM_THREADS, smem, st>>>(tmA, tmB, tmR, tmU, p, ep);
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb
