// tc_ffn.cu -- the back half of GatedDconvFFN (hat_arch.py:112-119) as ONE persistent tcgen05 kernel:
//
//   G[p, c]  = dw3x3(H1)[p, c] * SiLU(dw3x3(H1)[p, hid + c])              (depthwise 3x3 + bias, zero pad; c < hid)
//   v[p, n]  = sum_c G[p, c] * W2[n, c] + b2[n] + resid[p, n]             (fc2 + the HAB residual, hat_arch.py:237)
//   outputs: fp32 residual stream, LayerNorm(v) in bf16 (+ 16-channel head): the FeatResLn epilogue of tc_epilogue3.cuh
//
// Why: op by op the gated middle G (M x 384 bf16 = 50 MB at configs[1]) was written by dwgate_bf16_kernel and read
// back by gemm_tc_kernel 36 times per forward (the chain fc1 -> dwgate -> fc2 moves 11.9 of the 19 GB of
// a step).  Here G never exists in HBM: a CTA walks 8 (x) x 16 (y) pixel tiles; per 32-channel CHUNK the producer
// TMA-loads the (8+2) x (16+2) halo of the 32 proj and the 32 gate channels of H1 (out-of-bounds zero fill = the conv
// padding), eight "dw" warps turn it into a 128-pixel x 32-channel bf16 A tile in shared memory (K-major,
// SWIZZLE_64B), and the MMA thread accumulates that K = 32 slice of fc2 into TMEM (fp32) while the next chunk is
// being convolved.  After the last chunk the accumulator goes through the shared fused epilogue (fp32 residual via a
// TMA ring, fp32 store, row LayerNorm with the row parked in TMEM, bf16 store) on a dedicated warpgroup while the
// next tile's chunks run into the other accumulator stage.
//
// Warp roles (896 threads): 0 TMA producer (H1 halos, fc2 weight slices), 1 MMA issuer + TMEM owner, 2 / 3 residual-ring
// producers, 4-7 / 8-11 the two epilogue warpgroups (alternate tiles = alternate accumulator stages; TMEM lane quarter =
// warp % 4), 12-27 depthwise / gate warps.  B200 trace: the epilogue costs ~15 us per tile and warpgroup, the depthwise
// stage 18-20 us per tile on eight warps at ~40 % issue utilisation (2 per scheduler: fixed-latency dependencies, 563
// instructions per chunk), so both are doubled; 896 threads start at 72 registers and re-balance with setmaxnreg
// (producers 40, dw 56, epilogue 120: the pool is what the CTA was launched with, 896 x 72).
// dw thread = one channel pair x one pixel column x a 4-row strip, proj and gate in two passes with three rolling
// accumulator rows (packed FFMA2); conflict-free 4-byte reads of the plain 64-byte halo rows at immediate offsets.
#include <stdlib.h>

#include "tc_epilogue3.cuh"

namespace hb {

using namespace tc;
using namespace epi3;

unsigned long long* gemm_trace_arm(cudaStream_t st);

namespace {

__device__ __forceinline__ unsigned long long ff_time() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define FF_TRACE(slot) do { if (p.trace) p.trace[(size_t)blockIdx.x * 64 + (slot)] = ff_time(); } while (0)

constexpr int FF_THREADS = 896;
constexpr int FF_DW_T0 = 384;                // first depthwise thread (warp 12)
constexpr int FF_TW = 8, FF_TH = 16, FF_HW = FF_TW + 2, FF_HH = FF_TH + 2;
constexpr int FF_CK = 32;                    // gated channels per chunk = K of one fc2 slice
constexpr int FF_HALF = 12288;               // one staged half (proj or gate): 180 halo pixels x 64 B, rounded up
constexpr int FF_STAGE = 2 * FF_HALF;
constexpr int FF_MAX_ST = 4, FF_MAX_WST = 4, FF_GST = 2;
constexpr int FF_GT = 128 * FF_CK * 2;       // G A-tile: 128 rows x 64 B
constexpr int FF_DW_WARPS = 16, FF_DW_THREADS = FF_DW_WARPS * 32;
constexpr int FF_DW_ROWS = 4;                // output rows per depthwise thread (strip); 16 rows = 4 strips
constexpr int FF_SMEM_LIMIT = 227 * 1024 - 2048;

struct FfnParams {
  int B, H, W, tiles_x, tiles_y, tiles;
  int hid, nchunks;
  int N, BN;
  int st, wst, w_stage_bytes;
  int ops, op_stage_bytes;
  uint32_t acc_stride, tmem_cols;
  int off_stg, off_g, off_w, off_wtab, off_op, off_stg32, off_tab, smem_bytes;
  int stg_bufs;
  unsigned long long* trace;   // engineering: per-CTA timestamps (HATB200_GEMM_TRACE)
  int dbg;                // engineering (HATB200_FFN_DBG): 1 dw warps skip the arithmetic, 2 epilogue skips run_tile, 4 no H1 loads
};

__device__ __forceinline__ void ff_tile_coords(const FfnParams& p, int t, int& x0, int& y0, int& b) {
  const int tx = t % p.tiles_x, ty = (t / p.tiles_x) % p.tiles_y;
  b = t / (p.tiles_x * p.tiles_y);
  x0 = tx * FF_TW; y0 = ty * FF_TH;
}

__device__ __forceinline__ unsigned long long ff_unpack(uint32_t v) {      // bf16 pair -> packed f32x2 (low half first)
  return (unsigned long long)(v << 16) | ((unsigned long long)(v & 0xffff0000u) << 32);
}
__device__ __forceinline__ unsigned long long ff_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint32_t ff_lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ unsigned long long ff_lds64(uint32_t addr) {
  unsigned long long v;
  asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void ff_sts32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
template <int N>
__device__ __forceinline__ void ff_reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void ff_reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
__device__ __forceinline__ float ff_silu(float x) {      // x * sigmoid(x) = h + h * tanh(h), h = x / 2 (one MUFU)
  const float h = 0.5f * x;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  return fmaf(h, t, h);
}

}  // namespace

template <bool GROUPED>
__global__ void __launch_bounds__(FF_THREADS, 1)
ffn2_tc_kernel(const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmW,
               const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmOF,
               const __grid_constant__ CUtensorMap tmLN, const FfnParams p, const Epilogue ep,
               const float* __restrict__ dw_w, const float* __restrict__ dw_b) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t stg_full[FF_MAX_ST];
  __shared__ __align__(8) uint64_t stg_empty[FF_MAX_ST];
  __shared__ __align__(8) uint64_t g_full[FF_GST];
  __shared__ __align__(8) uint64_t g_empty[FF_GST];
  __shared__ __align__(8) uint64_t w_full[FF_MAX_WST];
  __shared__ __align__(8) uint64_t w_empty[FF_MAX_WST];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ __align__(8) uint64_t op_full[2][MAX_OPS];
  __shared__ __align__(8) uint64_t op_empty[2][MAX_OPS];
  __shared__ uint32_t tmem_slot;
  __shared__ float hsum[8][16];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* tab_bias = reinterpret_cast<float*>(smem + p.off_tab);
  float* tab_gamma = tab_bias + p.BN;
  float* tab_beta = tab_gamma + p.BN;
  float* tab_coef = tab_beta + p.BN;
  float* wtab = reinterpret_cast<float*>(smem + p.off_wtab);      // [chunk][10][64]: 9 taps + bias, proj 32 | gate 32

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmH);
    prefetch_tmap(&tmW);
    for (int s = 0; s < p.st; s++) { mbar_init(&stg_full[s], 1); mbar_init(&stg_empty[s], FF_DW_WARPS); }
    for (int s = 0; s < FF_GST; s++) { mbar_init(&g_full[s], FF_DW_WARPS); mbar_init(&g_empty[s], 1); }
    for (int s = 0; s < p.wst; s++) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int g = 0; g < 2; g++) { mbar_init(&tmem_full_bar[g], 1); mbar_init(&tmem_empty_bar[g], 128); }
    for (int g = 0; g < 2; g++)
      for (int s = 0; s < MAX_OPS; s++) { mbar_init(&op_full[g][s], 1); mbar_init(&op_empty[g][s], 4); }
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < p.BN; i += FF_THREADS) {
    const bool ok = i < p.N;
    tab_bias[i] = (ok && ep.bias) ? __ldg(ep.bias + i) : 0.f;
    tab_gamma[i] = (ok && ep.ln_gamma) ? __ldg(ep.ln_gamma + i) : 0.f;
    tab_beta[i] = (ok && ep.ln_beta) ? __ldg(ep.ln_beta + i) : 0.f;
  }
  // depthwise weights (parameters: loaded ahead of the programmatic dependency), zero beyond hid
  for (int i = threadIdx.x; i < p.nchunks * 640; i += FF_THREADS) {
    const int j = i / 640, r = i - j * 640, tap = r >> 6, cc = r & 63;
    const int c = j * FF_CK + (cc & 31);
    float v = 0.f;
    if (c < p.hid) {
      const int ch = (cc >> 5) * p.hid + c;
      v = tap < 9 ? __ldg(dw_w + (size_t)tap * 2 * p.hid + ch) : __ldg(dw_b + ch);
    }
    wtab[i] = v;
  }
  if (warp == 1) tmem_alloc(&tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int t_first = (int)blockIdx.x, t_step = (int)gridDim.x;

  // register re-balancing by warpgroup: every role branch starts with the setmaxnreg of its warpgroup, so that the
  // code it dominates is compiled against that budget
  if (warp < 4) {
  ff_reg_dec<40>();
  if (warp == 0) {
    // ===== TMA producer: per chunk one fc2 weight slice (BN rows x 32 k) and the proj / gate halo boxes of H1 =====
    if (elect_one()) {
      pdl_wait();          // H1 is the previous kernel's output
      pdl_trigger();
      RingPos sp, wp;
      const uint32_t stg_bytes = (uint32_t)(2 * FF_HH * FF_HW * FF_CK * 2), w_bytes = (uint32_t)(p.BN * FF_CK * 2);
      for (int t = t_first; t < p.tiles; t += t_step) {
        int x0, y0, b;
        ff_tile_coords(p, t, x0, y0, b);
        for (int j = 0; j < p.nchunks; j++) {
          {
            const int s = wp.s;
            if (p.dbg & 8) mbar_wait(&w_empty[s], wp.ph ^ 1u); else mbar_wait_relaxed(&w_empty[s], wp.ph ^ 1u);
            mbar_arrive_expect_tx(&w_full[s], w_bytes);
            tma_load_2d(smem + p.off_w + (size_t)s * p.w_stage_bytes, &tmW, &w_full[s], j * FF_CK, 0);
            wp.next(p.wst);
          }
          const int s = sp.s;
          if (p.dbg & 8) mbar_wait(&stg_empty[s], sp.ph ^ 1u); else mbar_wait_relaxed(&stg_empty[s], sp.ph ^ 1u);
          uint8_t* dst = smem + p.off_stg + (size_t)s * FF_STAGE;
          if (p.dbg & 4) { mbar_arrive(&stg_full[s]); sp.next(p.st); continue; }
          mbar_arrive_expect_tx(&stg_full[s], stg_bytes);
          tma_load_4d(dst, &tmH, &stg_full[s], j * FF_CK, x0 - 1, y0 - 1, b);
          tma_load_4d(dst + FF_HALF, &tmH, &stg_full[s], p.hid + j * FF_CK, x0 - 1, y0 - 1, b);
          sp.next(p.st);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: per chunk two K = 16 MMAs (A = the chunk's G tile, B = the fc2 slice), fp32 in TMEM =====
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(128, p.BN);
      RingPos gp, wp;
      int i = 0;
      for (int t = t_first; t < p.tiles; t += t_step, i++) {
        const int g = i & 1;
        mbar_wait(&tmem_empty_bar[g], (((uint32_t)i >> 1) & 1u) ^ 1u);     // epilogue has drained this stage
        tc_fence_after();
        const uint32_t acc = tmem_base + (uint32_t)g * p.acc_stride;
        for (int j = 0; j < p.nchunks; j++) {
          if (p.dbg & 8) mbar_wait(&g_full[gp.s], gp.ph); else mbar_wait_relaxed(&g_full[gp.s], gp.ph, 32);
          mbar_wait(&w_full[wp.s], wp.ph);
          tc_fence_after();
          const uint64_t ad = make_smem_desc(smem_u32(smem + p.off_g + (size_t)gp.s * FF_GT), 64);
          const uint64_t bd = make_smem_desc(smem_u32(smem + p.off_w + (size_t)wp.s * p.w_stage_bytes), 64);
#pragma unroll
          for (int k = 0; k < FF_CK / 16; k++) umma_bf16(acc, ad + 2 * k, bd + 2 * k, idesc, (uint32_t)(j | k));
          umma_commit(&g_empty[gp.s]);        // the G tile and the weight slice are free once these MMAs have read them
          umma_commit(&w_empty[wp.s]);
          gp.next(FF_GST);
          wp.next(p.wst);
        }
        umma_commit(&tmem_full_bar[g]);
      }
    }
  } else {
    // ===== residual-ring producer of epilogue warpgroup g =====
    const int g = warp - 2;
    if (elect_one()) {
      prefetch_tmap(&tmR);
      pdl_wait();
      uint8_t* ring = smem + p.off_op + (size_t)g * p.ops * p.op_stage_bytes;
      RingPos k;
      int i = 0;
      for (int t = t_first; t < p.tiles; t += t_step, i++) {
        if ((i & 1) != g) continue;
        Tile4D ts;
        ff_tile_coords(p, t, ts.x0, ts.y0, ts.b);
        produce_tile(ts, &tmR, &tmR, ring, op_full[g], op_empty[g], p.ops, p.op_stage_bytes, 0, 1, 0, 0, p.BN / CW, k);
      }
    }
  }
  } else if (warp < 12) {
    // ===== epilogue warpgroups: FeatResLn, warpgroup g takes the tiles of accumulator stage g =====
    ff_reg_inc<120>();
    const int g = (warp - 4) >> 2;
    WarpState<GROUPED> ws;
    ws.g = g; ws.ew = warp - 4; ws.quarter = warp & 3; ws.lane = lane;
    ws.row = ws.quarter * 32 + lane;
    ws.wg_tid = ((warp - 4) & 3) * 32 + lane;
    ws.BN = p.BN; ws.N = p.N; ws.n0 = 0; ws.nch = p.BN / CW; ws.H = p.H; ws.W = p.W; ws.hw = p.H * p.W;
    ws.last_ntile = true;
    ws.tab_bias = tab_bias; ws.tab_gamma = tab_gamma; ws.tab_beta = tab_beta; ws.coef = tab_coef + g * p.BN;
    ws.ring = smem + p.off_op + (size_t)g * p.ops * p.op_stage_bytes;
    ws.op_full = op_full[g]; ws.op_empty = op_empty[g];
    ws.ops = p.ops; ws.op_stage_bytes = p.op_stage_bytes; ws.op_u_off = 0;
    ws.has_resid = 1; ws.has_u = 0; ws.share_stg = 1;
    ws.hsum = hsum; ws.sred = nullptr;
    ws.st32.nbuf = ws.st16.nbuf = p.stg_bufs;
    ws.st32.base = smem + p.off_stg32 + (size_t)ws.ew * (p.stg_bufs * decltype(ws.st32)::BUF);
    ws.st16.base = ws.st32.base;             // fp32 staging is idle during the LayerNorm pass and vice versa
    if (lane == 0) { prefetch_tmap(&tmOF); prefetch_tmap(&tmLN); }
    pdl_wait();
    int i = 0;
    for (int t = t_first; t < p.tiles; t += t_step, i++) {
      if ((i & 1) != g) continue;
      Tile4D ts;
      ff_tile_coords(p, t, ts.x0, ts.y0, ts.b);
      RowInfo ri;
      ri.b = ts.b;
      ri.py = ts.y0 + ws.row / FF_TW; ri.px = ts.x0 + ws.row % FF_TW;
      ri.m = (ts.b * p.H + ri.py) * p.W + ri.px;
      ri.nchunk = p.tiles_x * p.tiles_y;
      ri.chunk = t - ts.b * ri.nchunk;
      ri.trow = tmem_base + (uint32_t)g * p.acc_stride + ((uint32_t)(ws.quarter * 32) << 16);
      const bool tr = ws.wg_tid == 0 && i < 4;
      if (tr) FF_TRACE(16 + 4 * i);
      if (p.dbg & 8) mbar_wait(&tmem_full_bar[g], ((uint32_t)i >> 1) & 1u);
      else mbar_wait_relaxed(&tmem_full_bar[g], ((uint32_t)i >> 1) & 1u, 128);
      tc_fence_after();
      if (tr) FF_TRACE(17 + 4 * i);
      if (!(p.dbg & 2)) run_tile<FeatResLn, GROUPED>(ep, ws, ri, ts, &tmOF, &tmOF, &tmLN);
      tc_fence_before();
      mbar_arrive(&tmem_empty_bar[g]);
      if (tr) FF_TRACE(18 + 4 * i);
    }
    if (elect_one()) bulk_wait_all();
  } else {
    // ===== depthwise 3x3 + gate: staged halo chunk -> 128 x 32 bf16 A tile (K-major, SWIZZLE_64B) =====
    ff_reg_dec<56>();
    const int dt = (int)threadIdx.x - FF_DW_T0;
    const int cp = dt & 15, x = (dt >> 4) & 7, strip = dt >> 7;             // channel pair, pixel column, 4-row strip
    const uint32_t cp_gran = (uint32_t)(cp >> 2), cp_off = (uint32_t)(cp & 3) * 4u;
    RingPos sp, gp;
    int ti = 0;
    long long w_stg = 0, w_g = 0;
    for (int t = t_first; t < p.tiles; t += t_step, ti++) {
      if (dt == 0 && ti < 4) FF_TRACE(4 * ti);
      for (int j = 0; j < p.nchunks; j++) {
        const uint32_t wt = smem_u32(wtab + (size_t)j * 640) + (uint32_t)cp * 8u;
        const bool live = j * FF_CK + 2 * cp < p.hid;            // channels past hid: K padding of fc2, written as zero
        long long c0 = 0;
        if (p.trace) c0 = clock64();
        mbar_wait(&stg_full[sp.s], sp.ph);
        if (p.trace) { const long long c1 = clock64(); w_stg += c1 - c0; c0 = c1; }
        mbar_wait(&g_empty[gp.s], gp.ph ^ 1u);
        if (p.trace) w_g += clock64() - c0;
        if (p.dbg & 1) {
          __syncwarp();
          if (lane == 0) { mbar_arrive(&g_full[gp.s]); mbar_arrive(&stg_empty[sp.s]); }
          sp.next(p.st); gp.next(FF_GST);
          continue;
        }
        // staged halves are plain (unswizzled) 64-byte rows, halo row hr = hy * 10 + hx: a warp reads two adjacent rows
        // = 128 contiguous bytes (conflict-free), and every address is this thread's base plus an immediate
        const uint32_t sp_base = smem_u32(smem + p.off_stg + (size_t)sp.s * FF_STAGE) +
                                 (uint32_t)((strip * FF_DW_ROWS * FF_HW + x) * 64 + cp * 4);
        const uint32_t g_base = smem_u32(smem + p.off_g + (size_t)gp.s * FF_GT);
        // Two passes (proj, then gate) over the strip's FF_DW_ROWS + 2 halo rows with three rolling accumulator rows each:
        // nine weight pairs live at a time, so that sixteen of these warps fit the register file next to the epilogue.
        unsigned long long kept[FF_DW_ROWS];
#pragma unroll
        for (int half = 0; half < 2; half++) {
          unsigned long long wv[9];
#pragma unroll
          for (int tp = 0; tp < 9; tp++) wv[tp] = ff_lds64(wt + tp * 256 + half * 128);
          const unsigned long long bv = ff_lds64(wt + 9 * 256 + half * 128);
          unsigned long long acc[3];
          acc[0] = acc[1] = acc[2] = bv;
#pragma unroll
          for (int i = 0; i < FF_DW_ROWS + 2; i++) {              // halo row strip * ROWS + i = image row y0 - 1 + strip * ROWS + i
            uint32_t rv[3];
#pragma unroll
            for (int cx = 0; cx < 3; cx++) rv[cx] = ff_lds32(sp_base + (uint32_t)((i * FF_HW + cx) * 64 + half * FF_HALF));
#pragma unroll
            for (int cx = 0; cx < 3; cx++) {
              const unsigned long long v = ff_unpack(rv[cx]);
#pragma unroll
              for (int ky = 0; ky < 3; ky++) {                    // output row i - ky of the strip lives in slot 2 - ky
                if (i - ky < 0 || i - ky >= FF_DW_ROWS) continue;
                acc[2 - ky] = ff_fma2(v, wv[ky * 3 + cx], acc[2 - ky]);
              }
            }
            if (i >= 2) {                                         // output row strip * ROWS + i - 2 is complete
              if (half == 0) {
                kept[i - 2] = acc[0];
              } else {
                const unsigned long long pv = kept[i - 2];
                const float px_ = __uint_as_float((uint32_t)pv), py_ = __uint_as_float((uint32_t)(pv >> 32));
                const float gx_ = __uint_as_float((uint32_t)acc[0]), gy_ = __uint_as_float((uint32_t)(acc[0] >> 32));
                const __nv_bfloat162 h = __floats2bfloat162_rn(px_ * ff_silu(gx_), py_ * ff_silu(gy_));
                const uint32_t r = (uint32_t)((strip * FF_DW_ROWS + i - 2) * FF_TW + x);
                ff_sts32(g_base + r * 64u + ((cp_gran ^ ((r >> 1) & 3u)) << 4) + cp_off,
                         live ? *reinterpret_cast<const uint32_t*>(&h) : 0u);
              }
            }
            acc[0] = acc[1]; acc[1] = acc[2]; acc[2] = bv;
          }
        }
        fence_proxy_async();               // the A tile is read by the tensor core through the async proxy
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&g_full[gp.s]);
          mbar_arrive(&stg_empty[sp.s]);
        }
        sp.next(p.st);
        gp.next(FF_GST);
      }
      if (dt == 0 && ti < 4) FF_TRACE(4 * ti + 1);
    }
    if (dt == 0 && p.trace) {
      p.trace[(size_t)blockIdx.x * 64 + 32] = (unsigned long long)w_stg;
      p.trace[(size_t)blockIdx.x * 64 + 33] = (unsigned long long)w_g;
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
static int ff_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        n <= 0)
      n = 148;
  }
  return n;
}

static bool ffn2_geometry(const FfnTailArgs& a, const Epilogue& ep, FfnParams& p, bool& grouped) {
  if (a.W % FF_TW || a.H % FF_TH) return false;
  if (a.hid < 8 || a.hid % 8 || (a.in_pitch * 2) % 16 || a.in_pitch < 2 * a.hid) return false;
  p.nchunks = cdiv(a.hid, FF_CK);
  if (p.nchunks > 16 || a.Kpad < p.nchunks * FF_CK || a.Kpad % 8) return false;
  p.BN = tc_bn(a.N);
  if (a.Npad != tc_npad(a.N) || a.Npad != p.BN || p.BN < 16 || p.BN > 256 || p.BN % 16) return false;   // one n-tile (LN)
  if (feat_select(ep) != FEAT_RESLN) return false;
  if ((ep.out_f32_pitch * 4) % 16 || ((uintptr_t)ep.out_f32 & 15)) return false;
  if ((ep.ln_pitch * 2) % 16 || ((uintptr_t)ep.ln_out & 15) || ep.ln_pitch < a.N) return false;
  if ((ep.resid_pitch * 4) % 16 || ((uintptr_t)ep.resid & 15)) return false;
  if (ep.ln_head && ep.ln_head_c > 16) return false;
  p.B = a.B; p.H = a.H; p.W = a.W;
  p.tiles_x = a.W / FF_TW; p.tiles_y = a.H / FF_TH; p.tiles = a.B * p.tiles_x * p.tiles_y;
  p.hid = a.hid; p.N = a.N;
  p.acc_stride = (uint32_t)round_up(p.BN, 32);
  uint32_t cols = 32;
  while (cols < 2 * p.acc_stride) cols <<= 1;
  if (cols > 512) return false;
  p.tmem_cols = cols;
  grouped = p.BN % 64 == 0;
  p.w_stage_bytes = round_up(p.BN * FF_CK * 2, 1024);
  p.op_stage_bytes = 128 * CW * 4;
  const int tab = round_up(5 * p.BN * 4, 1024);
  const int wtab = round_up(p.nchunks * 640 * 4, 1024);
  // staging depth first (it hides the L2 latency of the halo boxes), then the residual ring, then the weight ring
  for (int pass = 0; pass < 16; pass++) {
    // candidates in order of preference: staging depth 3 / 2, double / single epilogue staging, ring depth, weight ring
    const int st = (pass & 8) ? 2 : 3, bufs = (pass & 4) ? 1 : 2, ops = (pass & 2) ? 2 : 3, wst = (pass & 1) ? 2 : 3;
    const int stg32 = 8 * bufs * 32 * (grouped ? 128 : 64);
    int off = 0;
    p.off_stg = off; off += st * FF_STAGE;
    p.off_g = off; off += FF_GST * FF_GT;
    p.off_w = off; off += wst * p.w_stage_bytes;
    p.off_wtab = off; off += wtab;
    p.off_op = off; off += 2 * ops * p.op_stage_bytes;
    p.off_stg32 = off; off += stg32;
    p.off_tab = off; off += tab;
    if (off > FF_SMEM_LIMIT) continue;
    p.st = st; p.ops = ops; p.wst = wst; p.stg_bufs = bufs; p.smem_bytes = off;
    return true;
  }
  return false;
}

bool ffn2_tc_supported(const FfnTailArgs& a, const Epilogue& ep) {
  FfnParams p;
  bool grouped;
  if (!ffn2_geometry(a, ep, p, grouped)) return false;
  if (tc::encode_tiled_fn() == nullptr) return false;
  return true;
}

int launch_ffn2_tc(const FfnTailArgs& a, const Epilogue& ep, cudaStream_t st) {
  FfnParams p;
  bool grouped;
  HB_REQUIRE(ffn2_geometry(a, ep, p, grouped), HATB200_EUNSUPPORTED, "ffn2_tc: unsupported shape");
  CUtensorMap tmH, tmW, tmR, tmOF, tmLN;
  {   // H1 (B, H, W, 2 hid of in_pitch): channels >= 2 hid and pixels outside the image read as zero
    uint64_t dims[4] = {(uint64_t)(2 * a.hid), (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.B};
    uint64_t str[3] = {(uint64_t)a.in_pitch * 2, (uint64_t)a.W * a.in_pitch * 2, (uint64_t)a.H * a.W * a.in_pitch * 2};
    uint32_t box[4] = {FF_CK, FF_HW, FF_HH, 1};
    HB_TRY(tc::make_tmap_bf16(&tmH, a.h1, 4, dims, str, box, 0));        // plain 64-byte rows (no swizzle)
  }
  {   // fc2 weights [Npad][Kpad] K-major: one slice = BN rows x 32 k
    uint64_t dims[2] = {(uint64_t)a.Kpad, (uint64_t)a.Npad};
    uint64_t str[1] = {(uint64_t)a.Kpad * 2};
    uint32_t box[2] = {FF_CK, (uint32_t)p.BN};
    HB_TRY(tc::make_tmap_bf16(&tmW, a.w2, 2, dims, str, box, FF_CK * 2));
  }
  const uint32_t w32 = grouped ? 32 : 16, w16 = grouped ? 64 : 16;
  auto map4 = [&](CUtensorMap* tm, bool f32, const void* base, int width, int pitch, uint32_t bw, uint32_t bx, uint32_t by) -> int {
    const int es = f32 ? 4 : 2;
    uint64_t dims[4] = {(uint64_t)width, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.B};
    uint64_t str[3] = {(uint64_t)pitch * es, (uint64_t)a.W * pitch * es, (uint64_t)a.H * a.W * pitch * es};
    uint32_t box[4] = {bw, bx, by, 1};
    return f32 ? tc::make_tmap_f32(tm, base, 4, dims, str, box, (int)bw * es) : tc::make_tmap_bf16(tm, base, 4, dims, str, box, (int)bw * es);
  };
  HB_TRY(map4(&tmR, true, ep.resid, p.N, ep.resid_pitch, CW, FF_TW, FF_TH));
  HB_TRY(map4(&tmOF, true, ep.out_f32, p.N, ep.out_f32_pitch, w32, FF_TW, 4));
  HB_TRY(map4(&tmLN, false, ep.ln_out, ep.ln_pitch, ep.ln_pitch, w16, FF_TW, 4));
  p.dbg = getenv("HATB200_FFN_DBG") ? atoi(getenv("HATB200_FFN_DBG")) : 0;
  p.trace = gemm_trace_arm(st);
  const size_t smem = (size_t)p.smem_bytes + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    const int lim = FF_SMEM_LIMIT + 1024;
    HB_CHECK_CUDA(cudaFuncSetAttribute(ffn2_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
    HB_CHECK_CUDA(cudaFuncSetAttribute(ffn2_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
    attr_set = true;
  }
  int grid = ff_sm_count();
  if (grid > p.tiles) grid = p.tiles;
  if (grouped) launch_pdl(ffn2_tc_kernel<true>, dim3(grid), dim3(FF_THREADS), smem, st, tmH, tmW, tmR, tmOF, tmLN, p, ep, a.dw_w, a.dw_b);
  else launch_pdl(ffn2_tc_kernel<false>, dim3(grid), dim3(FF_THREADS), smem, st, tmH, tmW, tmR, tmOF, tmLN, p, ep, a.dw_w, a.dw_b);
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb
