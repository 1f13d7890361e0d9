// tc_tapconv.cu -- dense conv / Linear over channels-last tokens on the 5th-gen tensor cores.
//
//   out[m, n] = sum_tap sum_k A[pix(m) + tap, k] * W[tap][n][k]        (bf16 x bf16 -> fp32 in TMEM)
//
// A conv KHxKW is KH*KW shifted GEMMs: one CTA owns a 16x8-pixel tile (128 rows = UMMA M) and a
// BN-wide slice of output channels.  For every (tap, 64-channel block) the producer warp issues ONE
// 4-D TMA box {64 ch, 16 x, 8 y, 1 b} at the shifted pixel coordinates -- TMA's out-of-bounds
// zero fill IS the conv's zero padding, and the box lands in shared memory already in the
// canonical K-major SWIZZLE_128B operand layout -- plus a 2-D box of the packed weights.  One thread
// issues tcgen05.mma; completion is tracked with tcgen05.commit -> mbarrier.  Four epilogue warps
// read the accumulator with tcgen05.ld (one token row per thread) and apply the fused epilogue:
// bias, dynamic depthwise term, q-scale, GELU/LeakyReLU, ECA-rescaled CAB branch, fp32 residual,
// fp32 stream store, bf16 store (optionally pixel-shuffled), per-tile column sums (global average
// pool partials) and a full LayerNorm of the row (the row lives in ONE thread: no reduction).
// Replaces F.conv2d / F.linear at hat_arch.py:84-86,:111,:117,:347-351,:391-392,:544,:854-856 and
// esc_arch.py:122-123,:144.
#include <stdlib.h>

#include "kernels.cuh"
#include "tc_epilogue.cuh"

namespace hb {

using namespace tc;

constexpr int TC_THREADS = 192;     // warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-5: epilogue
constexpr int TC_MAX_STAGES = 8;
constexpr int TILE_W = 16, TILE_H = 8, TILE_M = 128;

struct TcParams {
  int B, H, W, KH, KW;
  int a_coff, kblocks, BK;
  int N, BN, Npad;
  int tiles_x, tiles_y;
  int stages, stage_bytes, a_bytes;
  uint32_t tmem_cols;
};

// ---- the kernel --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TC_THREADS)
tapconv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcParams p,
                  const Epilogue ep) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[TC_MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[TC_MAX_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float sred[4][256];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // tile coordinates: 16 x 8 pixels of one sample
  const int tile = blockIdx.x;
  const int tx = tile % p.tiles_x;
  const int ty = (tile / p.tiles_x) % p.tiles_y;
  const int b = tile / (p.tiles_x * p.tiles_y);
  const int x0 = tx * TILE_W, y0 = ty * TILE_H;
  const int n0 = blockIdx.y * p.BN;
  const int total_it = p.KH * p.KW * p.kblocks;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < p.stages; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  pdl_trigger();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one()) {
      const int ph = p.KH / 2, pw = p.KW / 2;
      const uint32_t tx_bytes = (uint32_t)(p.a_bytes + p.BN * p.BK * 2);
      RingPos rp;
      int tap = 0;
      for (int ky = 0; ky < p.KH; ky++)
       for (int kx = 0; kx < p.KW; kx++, tap++) {
        const int dy = ky - ph, dx = kx - pw;
        for (int kb = 0; kb < p.kblocks; kb++, rp.next(p.stages)) {
          const int s = rp.s;
          mbar_wait(&empty_bar[s], rp.ph ^ 1u);
          uint8_t* sa = smem + (size_t)s * p.stage_bytes;
          mbar_arrive_expect_tx(&full_bar[s], tx_bytes);
          tma_load_4d(sa, &tmA, &full_bar[s], p.a_coff + kb * p.BK, x0 + dx, y0 + dy, b);
          tma_load_2d(sa + p.a_bytes, &tmB, &full_bar[s], kb * p.BK, tap * p.Npad + n0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(TILE_M, p.BN);
      const uint32_t row_bytes = (uint32_t)p.BK * 2;
      RingPos rp;
      for (int it = 0; it < total_it; it++, rp.next(p.stages)) {
        const int s = rp.s;
        mbar_wait(&full_bar[s], rp.ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + (size_t)s * p.stage_bytes);
        const uint32_t sb = sa + (uint32_t)p.a_bytes;
        for (int k = 0; k < p.BK / 16; k++) {
          const uint64_t ad = make_smem_desc(sa + k * 32, row_bytes);
          const uint64_t bd = make_smem_desc(sb + k * 32, row_bytes);
          umma_bf16(tmem_base, ad, bd, idesc, (it | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);     // frees the smem slot once these MMAs have read it
      }
      umma_commit(&tmem_full_bar);      // accumulator complete
    }
  } else {
    // ===== epilogue: warps 2..5, TMEM lane quarter = warp % 4, one token row per thread =====
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    EpiCtx c;
    c.py = y0 + row / TILE_W; c.px = x0 + row % TILE_W;
    c.b = b; c.H = p.H; c.W = p.W;
    c.m = (b * p.H + c.py) * p.W + c.px;
    c.trow = tmem_base + ((uint32_t)(quarter * 32) << 16);
    c.N = p.N; c.n0 = n0; c.BN = p.BN;
    c.last_ntile = blockIdx.y == gridDim.y - 1;
    c.chunk = ty * p.tiles_x + tx;                 // tile index inside the sample
    c.nchunk = p.tiles_x * p.tiles_y;
    c.quarter = quarter; c.lane = lane; c.gtid = threadIdx.x - 64; c.bar_id = 1;
    c.sred = sred;
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    run_epilogue(ep, c);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ---- host ------------------------------------------------------------------------------------------
namespace tc {

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

static int make_tmap_any(CUtensorMap* out, CUtensorMapDataType dt, const void* base, int rank, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box, int inner_bytes) {
  EncodeTiledFn fn = encode_tiled_fn();
  HB_REQUIRE(fn != nullptr, HATB200_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t gdim[5], gstr[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; i++) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; i++) gstr[i] = strides_bytes[i];
  CUtensorMapSwizzle sw = inner_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                        : inner_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                        : inner_bytes == 0 ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_32B;
  CUresult r = fn(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  HB_REQUIRE(r == CUDA_SUCCESS, HATB200_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d)", (int)r, rank);
  return 0;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box, int inner_bytes) {
  return make_tmap_any(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, base, rank, dims, strides_bytes, box, inner_bytes);
}
int make_tmap_f32(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                  const uint32_t* box, int inner_bytes) {
  return make_tmap_any(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, base, rank, dims, strides_bytes, box, inner_bytes);
}

}  // namespace tc

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

static bool tc_geometry(const ConvArgs& a, int Npad, int Kpad, TcParams& p) {
  if (a.W % TILE_W || a.H % TILE_H) return false;
  if (!(a.KH & 1) || !(a.KW & 1)) return false;
  p.BK = Kpad == 16 ? 16 : (Kpad == 32 ? 32 : 64);
  if (Kpad % p.BK || Kpad < a.Cin) return false;
  if ((a.a_pitch * 2) % 16 || (a.a_coff * 2) % 16) return false;
  p.BN = tc_bn(a.N);
  if (Npad != tc_npad(a.N) || p.BN < 16 || p.BN > 256 || p.BN % 16) return false;
  p.B = a.B; p.H = a.H; p.W = a.W; p.KH = a.KH; p.KW = a.KW;
  p.a_coff = a.a_coff; p.kblocks = Kpad / p.BK; p.N = a.N; p.Npad = Npad;
  p.tiles_x = a.W / TILE_W; p.tiles_y = a.H / TILE_H;
  p.a_bytes = TILE_M * p.BK * 2;
  p.stage_bytes = round_up(p.a_bytes + p.BN * p.BK * 2, 1024);
  uint32_t cols = 32;
  while ((int)cols < p.BN) cols <<= 1;
  p.tmem_cols = cols;
  // pipeline depth: long main loops (3x3 convs) take the whole SM's shared memory; short ones (1x1) leave room
  // for a second resident CTA whose epilogue overlaps this one's main loop
  int iters = a.KH * a.KW * p.kblocks;
  int budget = iters >= 12 ? 200 * 1024 : 100 * 1024;
  int st = budget / p.stage_bytes;
  st = st > 6 ? 6 : st;
  st = st > iters ? iters : st;
  st = st < 2 ? (iters < 2 ? 1 : 2) : st;
  int forced = env_int("HATB200_TC_STAGES", 0);
  if (forced > 0) st = forced > TC_MAX_STAGES ? TC_MAX_STAGES : forced;
  p.stages = st;
  return true;
}

bool tapconv_tc_supported(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep) {
  TcParams p;
  if (!tc_geometry(a, Npad, Kpad, p)) return false;
  if (ep.ln_out && (Npad != p.BN || ep.ln_pitch > 512)) return false;     // LN needs the whole row in one CTA
  if (ep.ln_head && ep.ln_head_c > 16) return false;
  if (ep.dyn_k && a.N > 16) return false;
  if (ep.shuffle_perm) return false;                 // pre-permuted rows: tc_conv3.cu only
  if (tc::encode_tiled_fn() == nullptr) return false;
  return true;
}

int launch_tapconv_tc(const ConvArgs& a, const bf16* w, int Npad, int Kpad, const Epilogue& ep, cudaStream_t st) {
  TcParams p;
  HB_REQUIRE(tc_geometry(a, Npad, Kpad, p), HATB200_EUNSUPPORTED, "tapconv_tc: unsupported shape");
  CUtensorMap tmA, tmB;
  {
    uint64_t dims[4] = {(uint64_t)a.a_pitch, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.B};
    uint64_t str[3] = {(uint64_t)a.a_pitch * 2, (uint64_t)a.W * a.a_pitch * 2, (uint64_t)a.H * a.W * a.a_pitch * 2};
    uint32_t box[4] = {(uint32_t)p.BK, TILE_W, TILE_H, 1};
    HB_TRY(tc::make_tmap_bf16(&tmA, a.A, 4, dims, str, box, p.BK * 2));
  }
  {
    uint64_t dims[2] = {(uint64_t)Kpad, (uint64_t)a.KH * a.KW * Npad};
    uint64_t str[1] = {(uint64_t)Kpad * 2};
    uint32_t box[2] = {(uint32_t)p.BK, (uint32_t)p.BN};
    HB_TRY(tc::make_tmap_bf16(&tmB, w, 2, dims, str, box, p.BK * 2));
  }
  size_t smem = (size_t)p.stages * p.stage_bytes + 1024;
  static size_t smem_attr = 0;
  if (smem > smem_attr) {
    HB_CHECK_CUDA(cudaFuncSetAttribute(tapconv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 6144));
    smem_attr = 227 * 1024;
  }
  dim3 grid(a.B * p.tiles_x * p.tiles_y, Npad / p.BN);
  launch_pdl(tapconv_tc_kernel, grid, dim3(TC_THREADS), smem, st, tmA, tmB, p, ep);
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb
