// tc_gemm.cu -- Linear / 1x1 conv over channels-last tokens as a PERSISTENT tcgen05 GEMM with a warp-independent,
// TMA-fed and TMA-drained fused epilogue (third generation).
//
//   out[m, n] = sum_k A[m, k] * W[n, k]          (bf16 x bf16 -> fp32 in TMEM)
//   v = acc + bias; v *= q-scale; v = act(v); v += conv_scale * eca_scale[b,n] * U[m,n]; v += resid[m,n]
//   outputs: fp32 residual stream, bf16 activation, LayerNorm(v) in bf16 (+ compact 16-channel head, its x-padded
//   copy and its per-tile column sums)                                   (value order of common.cuh, struct Epilogue)
//
// Why a third generation (B200 measurements of the second one, profiles/r01_ncu_full_gemm_v2_summary.csv and the
// globaltimer trace tools/trace_gemm.py): a 128 x 240 tile took ~9 us to drain even for a bias-only epilogue, the
// tensor pipe was 1-3 % busy, 38 thread-instructions per output element, stalls = long scoreboard (per-thread
// global operands), the warpgroup barrier that guarded the 2-stage operand ring, and instruction fetch (117 KB of
// SASS).  Here:
//   * one CTA per SM walks 128-token tiles; the weight panel stays resident in shared memory when it fits
//     (else it streams with the activations); two TMEM accumulator stages, the two epilogue warpgroups take
//     alternate tiles so a token row lives in ONE thread (row-local LayerNorm, fp32 row parked in TMEM);
//   * every epilogue WARP is independent: no warpgroup barrier inside a tile.  The global operands (fp32 residual,
//     bf16 CAB branch) are TMA-loaded by a dedicated producer warp per warpgroup into a ring of 128-row x CW-column
//     chunks, released as soon as the four warps have pulled their rows into registers;
//   * every output leaves through TMA: a warp parks its 32 rows x CW columns in a private double-buffered staging
//     tile (swizzled, conflict-free) and one lane issues cp.async.bulk.tensor stores.  The tensor maps clip the
//     ragged edge (N = 180 of 192), TMA loads zero-fill it, weights / bias / gamma / beta are zero padded: the
//     epilogue has no per-column predicates at all;
//   * bias, gamma, beta and the per-sample ECA coefficients live in shared-memory tables; fp32 pair math uses
//     the packed FFMA2 / FADD2 / FMUL2 forms.
// Warp roles (384 threads): 0 TMA producer (A, W), 1 MMA issuer + TMEM owner, 2-5 / 6-9 epilogue warpgroups
// (TMEM lane quarter = warp % 4), 10 / 11 operand-ring producers of warpgroup 0 / 1.
// Replaces F.linear / 1x1 F.conv2d at hat_arch.py:111,:117,:347-351,:391-392 and esc_arch.py:144.
#include <stdlib.h>

#include "tc_epilogue3.cuh"

namespace hb {

using namespace tc;
using namespace epi3;

namespace {

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define G3_TRACE(slot) do { if (p.trace) p.trace[((size_t)(blockIdx.y * gridDim.x + blockIdx.x)) * 64 + (slot)] = gtime_ns(); } while (0)

constexpr int G3_THREADS = 384;
constexpr int G3_MAX_STAGES = 8;       // activation ring
constexpr int G3_MAX_OPS = MAX_OPS;    // operand ring stages per warpgroup
constexpr int G3_TILE_M = 128;
constexpr int G3_BK = 64;
constexpr int G3_SMEM_LIMIT = 227 * 1024 - 2048;   // static shared memory (barriers) stays below 2 KB
constexpr int G3_PRE_MAX = 6;          // preload ring stages (one stage = 128 rows x 32 fp32 columns, SWIZZLE_128B)
constexpr int G3_PRE_BYTES = G3_TILE_M * 32 * 4;

struct G3Params {
  int M, H, W, hw;
  int a_coff, kblocks;
  int N, BN, Npad;
  int m_tiles;
  int b_resident;           // weight panel resident in shared memory (else streamed per k-block with A)
  int stages, stage_bytes;  // activation ring (stage = A tile [+ W k-block])
  int ops, op_stage_bytes, op_u_off;   // operand ring per warpgroup
  int has_resid, has_u;
  int pre, pre_stages, off_pre;        // fp32 residual preloaded into the accumulator (tcgen05.cp), its TMA ring
  int rev;                             // walk the token tiles from the last to the first (Epilogue::walk_back)
  int pin;                             // fp32 residual stream in / out with the L2 evict_last policy (Epilogue::pin_stream)
  int split;                           // N-split CTA pairs: BN is HALF of the LayerNorm row, the cluster peer owns the rest
  uint32_t acc_stride, tmem_cols;
  // shared memory offsets (bytes from the 1024-aligned base)
  int off_b, off_a, off_op, off_stg32, off_stg16, off_tab, smem_bytes;
  int stg_bufs;             // staging buffers per warp and element type (2, or 1 when both types stage in pass 1)
  int share_stg;            // fp32 and bf16 staging alias (pass 1 writes only fp32, pass 2 only bf16)
  int nwg;                  // epilogue warpgroups: 2 (one per accumulator stage) or 4 (two per stage, each half of the columns)
  int prof;                      // engineering: epilogue wait-cycle counters on top of the timestamps
  int l2_ahead;                  // L2-prefetch the warpgroup's next tile's ring operands (HATB200_G3_L2AHEAD=1; measured slower: 10.82 vs 10.60 ms/step)
  unsigned long long* trace;
};

}  // namespace

// ---- the kernel ----------------------------------------------------------------------------------------
// NWG = 4: the epilogue warps are latency bound at two per scheduler (ncu: 18-33 % issue-slot utilisation), and a warp
// can only read its own 32 TMEM lanes, so the extra parallelism comes from the columns: two warpgroups per accumulator
// stage, each running the unchanged per-tile epilogue on one half of the n-tile (plain bf16 layers only: no row
// statistics, no operand ring; 576 threads, no ring-producer warps).
//
// PRE (round 2): the fp32 residual never passes through the epilogue threads.  A dedicated producer warp TMA-loads it
// in 128-row x 32-column boxes (SWIZZLE_128B), the MMA-issuing thread copies each box into the accumulator with four
// tcgen05.cp.128x256b (shared memory -> TMEM; tools/probe_tmem_cp.py: row r -> lane r, column c -> column c), and
// the tile's MMAs then ACCUMULATE on top of it (the tensor pipe runs cp and mma in issue order).  `acc` arrives in the
// epilogue as resid + A W^T: no ring stage, no shared-memory read and no add for the largest operand of the layer.
//
// SPLIT (round 2): the two CTAs of a cluster share every 128-token tile -- rank r owns columns [r * BN, (r + 1) * BN) of
// the row (BN = 96 of N = 180): half of the weight panel (resident again for K = 384), half of the residual boxes, half
// of the epilogue per tile, 512 tiles on 74 pairs = 6.92 -> 7 rounds instead of 3.46 -> 4 on 148 CTAs.  Plain
// cta_group::1 MMAs (M128 x N96); only the LayerNorm statistics cross the pair: (mean, M2) of each half row go to the
// peer's shared memory (st.shared::cluster + remote mbarrier arrive) and are merged (tc_epilogue3.cuh).  fp32 / LN stores
// use per-rank tensor maps (tmOF / tmOF1, tmLN / tmLN1) based at the rank's first column so that boxes clip at its edge.
template <class F, bool GROUPED, int NWG = 2, bool PRE = false, bool SPLIT = false>
__global__ void __launch_bounds__(NWG == 4 ? 64 + 128 * 4 : (PRE ? G3_THREADS + 32 : G3_THREADS), 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmU,
               const __grid_constant__ CUtensorMap tmOF, const __grid_constant__ CUtensorMap tmOT,
               const __grid_constant__ CUtensorMap tmLN, const __grid_constant__ CUtensorMap tmOF1,
               const __grid_constant__ CUtensorMap tmLN1, const G3Params p, const Epilogue ep) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[G3_MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[G3_MAX_STAGES];
  __shared__ __align__(8) uint64_t b_bar;
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ __align__(8) uint64_t op_full[2][G3_MAX_OPS];
  __shared__ __align__(8) uint64_t op_empty[2][G3_MAX_OPS];
  __shared__ __align__(8) uint64_t pre_full[G3_PRE_MAX], pre_empty[G3_PRE_MAX];
  __shared__ uint32_t tmem_slot;
  __shared__ float hsum[8][16];                       // per epilogue warp: column sums of the LN head
  __shared__ __align__(8) float2 xch[SPLIT ? 2 * 2 * 128 : 1];   // SPLIT: (mean, M2) of the peer's half rows
  __shared__ __align__(8) uint64_t xch_bar[2];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int crank = SPLIT ? (int)cluster_ctarank() : 0;                     // rank in the N-split pair
  const int walker = SPLIT ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;      // which tiles this CTA walks ...
  const int wstep = SPLIT ? (int)(gridDim.x >> 1) : (int)gridDim.x;         // ... and in what stride
  const int n0 = SPLIT ? crank * p.BN : blockIdx.y * p.BN;
  // first token of the t-th tile this launch walks: backwards when p.rev (the launch then STARTS with the rows its
  // producer wrote last, which L2 still holds -- see launch_gemm_tc)
  auto row0 = [&](int t) { return (p.rev ? p.m_tiles - 1 - t : t) * G3_TILE_M; };
  constexpr int A_BYTES = G3_TILE_M * G3_BK * 2;
  float* tab_bias = reinterpret_cast<float*>(smem + p.off_tab);   // [BN]
  float* tab_gamma = tab_bias + p.BN;                             // [BN]
  float* tab_beta = tab_gamma + p.BN;                             // [BN]
  float* tab_coef = tab_beta + p.BN;                              // [2][BN]: per-warpgroup ECA coefficients of its tile

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < p.stages; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&b_bar, 1);
    for (int s = 0; s < G3_PRE_MAX; s++) {
      mbar_init(&pre_full[s], 1);
      mbar_init(&pre_empty[s], 1);
    }
    for (int g = 0; g < 2; g++) {
      mbar_init(&xch_bar[g], 128);
      mbar_init(&tmem_full_bar[g], 1);
      mbar_init(&tmem_empty_bar[g], 64 * NWG);
      for (int s = 0; s < G3_MAX_OPS; s++) {
        mbar_init(&op_full[g][s], 1);
        mbar_init(&op_empty[g][s], 4);
      }
    }
    fence_barrier_init();
  }
  // tables: zero padded beyond N, so pad columns come out as exact zeros with no predicate
  for (int i = threadIdx.x; i < p.BN; i += (int)blockDim.x) {
    const int n = n0 + i;
    const bool ok = n < p.N;
    tab_bias[i] = (ok && ep.bias) ? __ldg(ep.bias + n) : 0.f;
    tab_gamma[i] = (ok && ep.ln_gamma) ? __ldg(ep.ln_gamma + n) : 0.f;
    tab_beta[i] = (ok && ep.ln_beta) ? __ldg(ep.ln_beta + n) : 0.f;
  }
  if (warp == 1) tmem_alloc(&tmem_slot, p.tmem_cols);
  tc_fence_before();
  if (SPLIT) cluster_sync_all();      // the peer's exchange barriers are initialised before anything remote arrives on them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ===== TMA producer: weights once (resident) or per k-block (streamed), activations through the ring =====
    // (single issuing threads are picked with elect.sync, not `lane == 0`: ptxas then knows the region is
    //  one-thread and emits UTMALDG / UTCHMMA from uniform registers without a per-lane waterfall loop)
    if (elect_one()) {
      if (p.b_resident) {
        mbar_arrive_expect_tx(&b_bar, (uint32_t)(p.kblocks * p.BN * 128));
        for (int kb = 0; kb < p.kblocks; kb++)
          tma_load_2d(smem + p.off_b + (size_t)kb * p.BN * 128, &tmB, &b_bar, kb * G3_BK, n0);
      }
      pdl_wait();        // weights are parameters; activations come from the previous kernel
      pdl_trigger();
      const uint32_t tx = (uint32_t)(A_BYTES + (p.b_resident ? 0 : p.BN * 128));
      RingPos rp;
      for (int t = walker; t < p.m_tiles; t += wstep) {
        const int tn = t + wstep;
        for (int kb = 0; kb < p.kblocks; kb++, rp.next(p.stages)) {
          const int s = rp.s;
          if (PRE && p.l2_ahead && tn < p.m_tiles) tma_prefetch_2d(&tmA, p.a_coff + kb * G3_BK, row0(tn));
          mbar_wait(&empty_bar[s], rp.ph ^ 1u);
          uint8_t* sa = smem + p.off_a + (size_t)s * p.stage_bytes;
          mbar_arrive_expect_tx(&full_bar[s], tx);
          tma_load_2d(sa, &tmA, &full_bar[s], p.a_coff + kb * G3_BK, row0(t));
          if (!p.b_resident) tma_load_2d(sa + A_BYTES, &tmB, &full_bar[s], kb * G3_BK, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one elected thread) =====
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(G3_TILE_M, p.BN);
      pdl_wait();
      G3_TRACE(0);
      if (p.b_resident) mbar_wait(&b_bar, 0);
      G3_TRACE(1);
      RingPos rp, pp;
      int i = 0;
      for (int t = walker; t < p.m_tiles; t += wstep, i++) {
        const int g = i & 1;
        mbar_wait(&tmem_empty_bar[g], (((uint32_t)i >> 1) & 1u) ^ 1u);     // epilogue has drained this stage
        tc_fence_after();
        if (i < 6) G3_TRACE(8 + 2 * i);
        const uint32_t acc = tmem_base + (uint32_t)g * p.acc_stride;
        if (PRE) {
          for (int c = 0; c < p.BN / 32; c++, pp.next(p.pre_stages)) {
            mbar_wait(&pre_full[pp.s], pp.ph);
            tc_fence_after();
            const uint64_t rd = make_smem_desc(smem_u32(smem + p.off_pre + (size_t)pp.s * G3_PRE_BYTES), 128);
#pragma unroll
            for (int q = 0; q < 4; q++)       // 8 fp32 columns (256 bits) per copy; +32 bytes inside the swizzle atom
              asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(acc + (uint32_t)(32 * c + 8 * q)), "l"(rd + 2 * q) : "memory");
            umma_commit(&pre_empty[pp.s]);    // frees the stage once the copies have read it
          }
        }
        for (int kb = 0; kb < p.kblocks; kb++, rp.next(p.stages)) {
          const int s = rp.s;
          mbar_wait(&full_bar[s], rp.ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + p.off_a + (size_t)s * p.stage_bytes);
          const uint32_t sb = p.b_resident ? smem_u32(smem + p.off_b + (size_t)kb * p.BN * 128) : sa + A_BYTES;
          const uint64_t ad = make_smem_desc(sa, 128), bd = make_smem_desc(sb, 128);
#pragma unroll
          for (int k = 0; k < G3_BK / 16; k++) umma_bf16(acc, ad + 2 * k, bd + 2 * k, idesc, PRE ? 1u : (uint32_t)(kb | k));
          umma_commit(&empty_bar[s]);       // frees the ring slot once these MMAs have read it
        }
        umma_commit(&tmem_full_bar[g]);     // accumulator of tile i complete
        if (i < 6) G3_TRACE(9 + 2 * i);
      }
    }
  } else if (PRE && warp == 12) {
    // ===== residual preload producer: 128 x 32 fp32 boxes of every tile of this CTA, as far ahead as the ring allows =====
    if (elect_one()) {
      prefetch_tmap(&tmR);
      pdl_wait();
      RingPos pp;
      const uint64_t pol = p.pin ? l2_policy_evict_last() : 0;
      for (int t = walker; t < p.m_tiles; t += wstep) {
        const int tn = t + wstep;                      // this CTA's next tile: start it towards L2 now (p.l2_ahead)
        for (int c = 0; c < p.BN / 32; c++, pp.next(p.pre_stages)) {
          if (p.l2_ahead && tn < p.m_tiles) tma_prefetch_2d(&tmR, n0 + 32 * c, row0(tn));
          mbar_wait(&pre_empty[pp.s], pp.ph ^ 1u);
          mbar_arrive_expect_tx(&pre_full[pp.s], (uint32_t)G3_PRE_BYTES);
          if (p.pin) tma_load_2d_hint(smem + p.off_pre + (size_t)pp.s * G3_PRE_BYTES, &tmR, &pre_full[pp.s], n0 + 32 * c, row0(t), pol);
          else tma_load_2d(smem + p.off_pre + (size_t)pp.s * G3_PRE_BYTES, &tmR, &pre_full[pp.s], n0 + 32 * c, row0(t));
        }
      }
    }
  } else if (warp >= 2 + 4 * NWG) {
    // ===== operand-ring producer of warpgroup g: residual / CAB-branch chunks, as far ahead as the ring allows =====
    const int g = warp - 10;
    const int ring_resid = PRE ? 0 : p.has_resid;
    if (feat_ring<F>() && (ring_resid || p.has_u) && elect_one()) {
      if (ring_resid) prefetch_tmap(&tmR);
      if (p.has_u) prefetch_tmap(&tmU);
      pdl_wait();
      uint8_t* ring = smem + p.off_op + (size_t)g * p.ops * p.op_stage_bytes;
      RingPos k;
      int i = 0;
      for (int t = walker; t < p.m_tiles; t += wstep, i++) {
        if ((i & 1) != g) continue;
        const Tile2D ts{row0(t)};
        const int tn = t + 2 * wstep;                   // this warpgroup's next tile
        const Tile2D nx{row0(tn)};
        produce_tile(ts, &tmR, &tmU, ring, op_full[g], op_empty[g], p.ops, p.op_stage_bytes, p.op_u_off, ring_resid,
                     p.has_u, n0, p.BN / CW, k, (p.l2_ahead && tn < p.m_tiles) ? &nx : nullptr,
                     p.pin ? l2_policy_evict_last() : 0);
      }
    }
  } else {
    // ===== epilogue warps: warpgroup g = (warp - 2) / 4 takes local tiles i with (i & 1) == g =====
    const int wgi = (warp - 2) >> 2;
    const int g = wgi & 1, half = NWG == 4 ? wgi >> 1 : 0;      // accumulator stage, column half of the n-tile
    const int HB = NWG == 4 ? p.BN / 2 : p.BN, hoff = half * HB;
    WarpState<GROUPED> ws;
    ws.g = g; ws.ew = warp - 2; ws.quarter = warp & 3; ws.lane = lane;
    ws.row = ws.quarter * 32 + lane;                     // row of the tile owned by this thread (TMEM lane)
    ws.wg_tid = ((warp - 2) & 3) * 32 + lane;            // 0..127 inside the warpgroup
    ws.BN = HB; ws.N = p.N; ws.n0 = n0 + hoff; ws.nch = HB / CW; ws.H = p.H; ws.W = p.W; ws.hw = p.hw;
    ws.last_ntile = blockIdx.y == gridDim.y - 1 && half == NWG / 2 - 1;
    ws.tab_bias = tab_bias + hoff; ws.tab_gamma = tab_gamma + hoff; ws.tab_beta = tab_beta + hoff; ws.coef = tab_coef + g * p.BN;
    ws.ring = smem + p.off_op + (size_t)g * p.ops * p.op_stage_bytes;
    ws.op_full = op_full[g]; ws.op_empty = op_empty[g];
    ws.ops = p.ops; ws.op_stage_bytes = p.op_stage_bytes; ws.op_u_off = p.op_u_off;
    ws.has_resid = PRE ? 0 : p.has_resid; ws.has_u = p.has_u; ws.share_stg = p.share_stg;
    ws.prof = p.trace != nullptr && p.prof;
    ws.hsum = hsum; ws.sred = nullptr;                   // no tile column sums in the 1x1 layers
    ws.so_f32 = 0; ws.xch = xch; ws.xch_bar = xch_bar; ws.peer = (uint32_t)(crank ^ 1);
    const CUtensorMap* pOF = (SPLIT && crank) ? &tmOF1 : &tmOF;
    const CUtensorMap* pLN = (SPLIT && crank) ? &tmLN1 : &tmLN;
    ws.st32.nbuf = ws.st16.nbuf = p.stg_bufs;
    if (p.pin) ws.st32.policy = l2_policy_evict_last();
    ws.st32.base = smem + p.off_stg32 + (size_t)ws.ew * (p.stg_bufs * decltype(ws.st32)::BUF);
    ws.st16.base = smem + p.off_stg16 + (size_t)ws.ew * (p.stg_bufs * decltype(ws.st16)::BUF);
    if (lane == 0) {
      if (ep.out_f32) prefetch_tmap(pOF);
      if (ep.out_t) prefetch_tmap(&tmOT);
      if (ep.ln_out) prefetch_tmap(pLN);
    }
    pdl_wait();          // the epilogue reads the ECA gate / ring operands and writes the outputs
    int i = 0;
    for (int t = walker; t < p.m_tiles; t += wstep, i++) {
      if ((i & 1) != g) continue;
      const int m0 = row0(t);
      RowInfo ri;
      ri.m = m0 + ws.row;
      ri.b = m0 / p.hw;                                  // a tile never straddles two samples (hw % 128 == 0)
      const int rem = ri.m - ri.b * p.hw;
      ri.py = rem / p.W; ri.px = rem - ri.py * p.W;
      ri.chunk = (m0 - ri.b * p.hw) / G3_TILE_M; ri.nchunk = p.hw / G3_TILE_M;
      ri.trow = tmem_base + (uint32_t)g * p.acc_stride + (uint32_t)hoff + ((uint32_t)(ws.quarter * 32) << 16);
      const bool tr = ws.wg_tid == 0 && half == 0 && (i >> 1) < 4;
      if (tr) G3_TRACE(24 + g * 16 + 4 * (i >> 1));
      mbar_wait(&tmem_full_bar[g], ((uint32_t)i >> 1) & 1u);
      tc_fence_after();
      if (tr) G3_TRACE(25 + g * 16 + 4 * (i >> 1));
      const Tile2D ts{m0};
      run_tile<F, GROUPED, Tile2D, PRE, SPLIT>(ep, ws, ri, ts, pOF, &tmOT, pLN);
      tc_fence_before();
      mbar_arrive_relaxed(&tmem_empty_bar[g]);      // guards the TMEM stage only (tc_common.cuh)
      if (tr) G3_TRACE(26 + g * 16 + 4 * (i >> 1));
    }
    if (elect_one()) bulk_wait_all();       // staging buffers must outlive the last stores
    if (ws.prof && ws.wg_tid == 0 && half == 0) {
      unsigned long long* tr = p.trace + (size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 64 + 56 + 4 * g;
      tr[0] = (unsigned long long)ws.t_ring; tr[1] = (unsigned long long)ws.t_stg;
      tr[2] = (unsigned long long)ws.t_p1; tr[3] = (unsigned long long)ws.t_p2;
    }
  }
  tc_fence_before();
  if (SPLIT) cluster_sync_all();      // no CTA leaves while its peer may still write into its shared memory
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ---- host ------------------------------------------------------------------------------------------
static unsigned long long* g_gemm_trace = nullptr;
unsigned long long* gemm_trace_buffer() { return g_gemm_trace; }
// engineering: (re)arm the per-CTA timestamp buffer when HATB200_GEMM_TRACE is set; null otherwise
unsigned long long* gemm_trace_arm(cudaStream_t st) {
  if (!getenv("HATB200_GEMM_TRACE")) return nullptr;
  if (!g_gemm_trace && cudaMalloc(&g_gemm_trace, 1024 * 64 * 8) != cudaSuccess) return nullptr;
  cudaMemsetAsync(g_gemm_trace, 0, 1024 * 64 * 8, st);
  return g_gemm_trace;
}

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

static bool gemm_geometry(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep, G3Params& p, bool& grouped) {
  if (a.KH != 1 || a.KW != 1) return false;
  const long long M = (long long)a.B * a.H * a.W;
  if (M % G3_TILE_M || (a.H * a.W) % G3_TILE_M) return false;
  if (Kpad % G3_BK || Kpad < a.Cin) return false;
  if ((a.a_pitch * 2) % 16 || (a.a_coff * 2) % 16) return false;
  p.BN = tc_bn(a.N);
  if (Npad != tc_npad(a.N) || p.BN < 16 || p.BN > 256 || p.BN % 16) return false;
  // N-split CTA pairs for the residual + LayerNorm layers whose row is one n-tile (see the kernel)
  // MEASURED SLOWER on B200 and therefore opt-in (HATB200_G3_SPLIT=1; parity-green, tests/test_gpu_kernels.py): esc_aggr
  // 46.4 -> 51.8 us, fc2 42.2 -> 48.3 us.  Pass 1 of a half row halves (5.5 -> 2.8 us per tile) but every tile is a
  // rendezvous of the two CTAs -- the statistics exchange waits 2-4 us per tile for the slower peer (trace_gemm.py:
  // pass 2 2.3 -> 4.4 us), which costs more than 7 half tiles instead of 4 whole ones saves.
  const char* split_env = getenv("HATB200_G3_SPLIT");
  const bool no_split = !(split_env && split_env[0] == '1');
  static const bool no_pre0 = getenv("HATB200_G3_PRE") && getenv("HATB200_G3_PRE")[0] == '0';
  const int feat0 = feat_select(ep);
  p.split = (!no_split && ep.ln_out && Npad == p.BN && p.BN % 64 == 0 && M / G3_TILE_M >= 2 &&
             ((feat0 == FEAT_RESLN && !no_pre0) || feat0 == FEAT_RESULN)) ? 1 : 0;
  if (p.split) p.BN /= 2;
  // features this kernel does not implement (tc_tapconv.cu handles those layers)
  if (ep.colsum || ep.dyn_k || ep.shuffle_r != 1) return false;
  if (ep.scale_ncols % CW) return false;
  // TMA alignment: 16-byte row pitches and bases
  if (ep.out_f32 && ((ep.out_f32_pitch * 4) % 16 || ((uintptr_t)ep.out_f32 & 15))) return false;
  if (ep.out_t && ((ep.out_t_pitch * 2) % 16 || (ep.out_t_coff * 2) % 16 || ((uintptr_t)ep.out_t & 15))) return false;
  if (ep.out_t && ep.out_t_pitch - ep.out_t_coff < (ep.out_t_zero_to > a.N ? ep.out_t_zero_to : a.N)) return false;
  if (ep.ln_out && ((ep.ln_pitch * 2) % 16 || ((uintptr_t)ep.ln_out & 15) || ep.ln_pitch < a.N)) return false;
  if (ep.resid && ((ep.resid_pitch * 4) % 16 || ((uintptr_t)ep.resid & 15))) return false;
  if (ep.eca_u && ((ep.eca_pitch * 2) % 16 || ((uintptr_t)ep.eca_u & 15) || ep.eca_pitch < a.N)) return false;
  if (ep.ln_out && Npad != p.BN * (p.split ? 2 : 1)) return false;     // LayerNorm needs the whole row in one CTA (pair)
  if (ep.ln_head && (ep.ln_head_c > 16 || !ep.ln_out)) return false;
  p.M = (int)M; p.H = a.H; p.W = a.W; p.hw = a.H * a.W;
  p.a_coff = a.a_coff; p.kblocks = Kpad / G3_BK;
  p.N = a.N; p.Npad = Npad;
  p.m_tiles = (int)(M / G3_TILE_M);
  p.acc_stride = (uint32_t)round_up(p.BN, 32);
  uint32_t cols = 32;
  while (cols < 2 * p.acc_stride) cols <<= 1;
  if (cols > 512) return false;
  p.tmem_cols = cols;
  p.has_resid = ep.resid != nullptr; p.has_u = ep.eca_u != nullptr;
  static const bool no_pre = getenv("HATB200_G3_PRE") && getenv("HATB200_G3_PRE")[0] == '0';       // engineering: A/B
  // not with the CAB branch in the same epilogue (esc_aggr): its operand ring keeps the epilogue the long pole, and the
  // preload then only lengthens the issuer's serial chain (B200: 46.5 us without, 47.8 us with)
  p.pre = (!no_pre && p.has_resid && !p.has_u && p.BN % 32 == 0 && feat_select(ep) != FEAT_PLAIN) ? 1 : 0;
  p.pre_stages = 0; p.off_pre = 0;
  const bool ring = (p.has_resid && !p.pre) || p.has_u;
  // stores leave as 128-byte rows (2 fp32 / 4 bf16 chunks per TMA store); a split row ends its last bf16 group early and
  // the per-rank tensor map clips the box there
  grouped = p.BN % 64 == 0 || (p.split && p.BN % 32 == 0);
  const bool f32o = ep.out_f32 != nullptr, bf16o = ep.out_t != nullptr || ep.ln_out != nullptr;
  // fp32 staging is idle during the LayerNorm pass and vice versa: share it unless pass 1 also writes bf16
  p.share_stg = f32o && bf16o && ep.out_t == nullptr;
  const int b_bytes = p.BN * Kpad * 2;
  const int tab = round_up(5 * p.BN * 4, 1024);
  // preload kernels: the serial chain is the issuer's (residual boxes, then the k-blocks), so shared memory goes to
  // activation / preload stages before a second staging buffer (B200 sweep: fc2 43.0 -> 41.6 us, OCAB mlp2 59.7 -> 50.9)
  p.stg_bufs = ((f32o && ep.out_t != nullptr) || p.pre) ? 1 : 2;
  static const bool nwg2 = getenv("HATB200_G3_NWG") && getenv("HATB200_G3_NWG")[0] == '2';
  p.nwg = (!nwg2 && grouped && p.BN % 128 == 0 && feat_select(ep) == FEAT_PLAIN) ? 4 : 2;
  const int ewarps = 4 * p.nwg;
  const int stg32 = f32o ? ewarps * p.stg_bufs * 32 * (grouped ? 128 : 64) : 0;
  const int stg16 = bf16o && !p.share_stg ? ewarps * p.stg_bufs * 32 * (grouped ? 128 : 32) : 0;
  const int rb = (p.has_resid && !p.pre) ? G3_TILE_M * CW * 4 : 0, ub = p.has_u ? G3_TILE_M * CW * 2 : 0;
  p.op_u_off = rb;
  p.op_stage_bytes = round_up(rb + ub, 1024);
  // Shared-memory plan.  Candidates: weights resident or streamed, staging double or single buffered, operand
  // ring depth; each leaves some number of activation stages.  The epilogue -- not the MMA -- is the long pole of
  // the ring kernels (one tile's MMAs take 1-3 us, its epilogue 8-10 us), and what the ring waits for is one
  // memory latency per `ops` chunks, so ring depth is bought before activation stages beyond 2.
  static const int force_ops = getenv("HATB200_G3_OPS") ? atoi(getenv("HATB200_G3_OPS")) : 0;          // engineering
  static const int force_bufs = getenv("HATB200_G3_STGBUFS") ? atoi(getenv("HATB200_G3_STGBUFS")) : 0;
  static const int force_res = getenv("HATB200_G3_RESIDENT") ? atoi(getenv("HATB200_G3_RESIDENT")) : -1;
  static const int force_pre = getenv("HATB200_G3_PRESTAGES") ? atoi(getenv("HATB200_G3_PRESTAGES")) : 0;
  const int smem_limit = G3_SMEM_LIMIT - (p.split ? 4096 : 0);     // the pair's exchange buffer is static shared memory
  int best_score = -1;
  G3Params best = p;
  for (int resident = 1; resident >= 0; resident--) {
    if (force_res >= 0 && resident != force_res) continue;
    for (int bufs = p.stg_bufs; bufs >= 1; bufs--) {
      if (force_bufs && bufs != force_bufs) continue;
      const int s32 = stg32 / p.stg_bufs * bufs, s16 = stg16 / p.stg_bufs * bufs;
      const int stage = G3_TILE_M * G3_BK * 2 + (resident ? 0 : p.BN * 128);
      const int fixed = tab + s32 + s16 + (resident ? round_up(b_bytes, 1024) : 0);
      for (int ops = ring ? G3_MAX_OPS : 0; ops >= (ring ? 2 : 0); ops--)
      for (int pre = p.pre ? 4 : 0; pre >= (p.pre ? 2 : 0); pre--) {
        if (force_ops && ring && ops != force_ops) continue;
        if (force_pre && p.pre && pre != force_pre) continue;
        const int avail = smem_limit - fixed - 2 * ops * p.op_stage_bytes - pre * G3_PRE_BYTES;
        int st = avail > 0 ? avail / stage : 0;
        if (st > G3_MAX_STAGES) st = G3_MAX_STAGES;
        if (st > p.kblocks * 2 && st > 4) st = p.kblocks * 2 > 4 ? p.kblocks * 2 : 4;
        if (st < 2) continue;
        // score: ring depth first (x4), then activation stages (capped at 4), resident weights (+3 when K is
        // short enough that the panel load is not a long prologue), double-buffered staging (+2)
        const int score = (ring ? 4 * ops : 0) + (st > 4 ? 4 : st) + (resident ? (p.kblocks <= 3 ? 3 : 1) : 0) + 2 * (bufs - 1) +
                          3 * (pre > 4 ? 4 : pre);
        if (score > best_score) {
          best_score = score;
          best = p;
          best.b_resident = resident; best.stage_bytes = stage; best.stages = st; best.ops = ops; best.stg_bufs = bufs;
          best.pre_stages = pre;
        }
      }
    }
  }
  if (best_score < 0) return false;
  {
    const int bufs0 = p.stg_bufs;
    p = best;
    const int s32 = stg32 / bufs0 * p.stg_bufs, s16 = stg16 / bufs0 * p.stg_bufs;
    const int resident = p.b_resident, stage = p.stage_bytes;
    if (getenv("HATB200_TRACE"))
      fprintf(stderr, "[gemm_tc] N=%d K=%d BN=%d resident=%d stages=%d ops=%d stg_bufs=%d share=%d nwg=%d pre=%d x %d\n", p.N, Kpad,
              p.BN, resident, p.stages, p.ops, p.stg_bufs, p.share_stg, p.nwg, p.pre, p.pre_stages);
    int off = 0;
    p.off_b = off; off += resident ? round_up(b_bytes, 1024) : 0;
    p.off_a = off; off += p.stages * stage;
    p.off_op = off; off += 2 * p.ops * p.op_stage_bytes;
    off = round_up(off, 1024);
    p.off_pre = off; off += p.pre_stages * G3_PRE_BYTES;
    p.off_stg32 = off; off += s32;
    p.off_stg16 = p.share_stg ? p.off_stg32 : off; off += s16;
    p.off_tab = off; off += tab;
    p.smem_bytes = off;
    return off <= smem_limit;
  }
}

bool gemm_tc_supported(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep) {
  G3Params p;
  bool grouped;
  if (!gemm_geometry(a, Npad, Kpad, ep, p, grouped)) return false;
  if (tc::encode_tiled_fn() == nullptr) return false;
  const char* v = getenv("HATB200_GEMM_V3");
  if (v && *v == '0') return false;
  return true;
}

int launch_gemm_tc(const ConvArgs& a, const bf16* w, int Npad, int Kpad, const Epilogue& ep, cudaStream_t st) {
  G3Params p;
  bool grouped;
  HB_REQUIRE(gemm_geometry(a, Npad, Kpad, ep, p, grouped), HATB200_EUNSUPPORTED, "gemm_tc: unsupported shape");
  CUtensorMap tmA, tmB, tmR, tmU, tmOF, tmOT, tmLN;
  {
    uint64_t dims[2] = {(uint64_t)a.a_pitch, (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)a.a_pitch * 2};
    uint32_t box[2] = {G3_BK, G3_TILE_M};
    HB_TRY(tc::make_tmap_bf16(&tmA, a.A, 2, dims, str, box, G3_BK * 2));
  }
  {
    uint64_t dims[2] = {(uint64_t)Kpad, (uint64_t)Npad};
    uint64_t str[1] = {(uint64_t)Kpad * 2};
    uint32_t box[2] = {G3_BK, (uint32_t)p.BN};
    HB_TRY(tc::make_tmap_bf16(&tmB, w, 2, dims, str, box, G3_BK * 2));
  }
  tmR = tmA; tmU = tmA; tmOF = tmA; tmOT = tmA; tmLN = tmA;     // placeholders for absent operands (never dereferenced)
  const uint32_t w32 = grouped ? 32 : 16, w16 = grouped ? 64 : 16;   // store box widths (columns)
  if (p.has_resid) {     // columns >= N read as zero
    uint64_t dims[2] = {(uint64_t)p.N, (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)ep.resid_pitch * 4};
    uint32_t box[2] = {p.pre ? 32u : (uint32_t)CW, G3_TILE_M};
    HB_TRY(tc::make_tmap_f32(&tmR, ep.resid, 2, dims, str, box, p.pre ? 128 : CW * 4));
  }
  if (p.has_u) {         // columns >= N read as zero (U's own pad columns are never touched)
    uint64_t dims[2] = {(uint64_t)p.N, (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)ep.eca_pitch * 2};
    uint32_t box[2] = {(uint32_t)CW, G3_TILE_M};
    HB_TRY(tc::make_tmap_bf16(&tmU, ep.eca_u, 2, dims, str, box, CW * 2));
  }
  CUtensorMap tmOF1, tmLN1;
  tmOF1 = tmA; tmLN1 = tmA;
  if (ep.out_f32) {      // stores clip at N
    uint64_t dims[2] = {(uint64_t)(p.split ? p.BN : p.N), (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)ep.out_f32_pitch * 4};
    uint32_t box[2] = {w32, 32};
    HB_TRY(tc::make_tmap_f32(&tmOF, ep.out_f32, 2, dims, str, box, w32 * 4));
    if (p.split) {       // rank 1: columns [BN, N), map based at column BN
      dims[0] = (uint64_t)(p.N - p.BN);
      HB_TRY(tc::make_tmap_f32(&tmOF1, ep.out_f32 + p.BN, 2, dims, str, box, w32 * 4));
    }
  }
  if (ep.out_t) {        // stores clip at max(N, zero_to): the pad columns in between receive computed zeros
    const int width = ep.out_t_zero_to > p.N ? ep.out_t_zero_to : p.N;
    uint64_t dims[2] = {(uint64_t)width, (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)ep.out_t_pitch * 2};
    uint32_t box[2] = {w16, 32};
    HB_TRY(tc::make_tmap_bf16(&tmOT, (const bf16*)ep.out_t + ep.out_t_coff, 2, dims, str, box, w16 * 2));
  }
  if (ep.ln_out) {       // LayerNorm output: pad columns up to ln_pitch are written as zero
    uint64_t dims[2] = {(uint64_t)(p.split ? p.BN : ep.ln_pitch), (uint64_t)p.M};
    uint64_t str[1] = {(uint64_t)ep.ln_pitch * 2};
    uint32_t box[2] = {w16, 32};
    HB_TRY(tc::make_tmap_bf16(&tmLN, ep.ln_out, 2, dims, str, box, w16 * 2));
    if (p.split) {       // rank 1: columns [BN, ln_pitch) (its pad columns come out as computed zeros)
      HB_REQUIRE(ep.ln_pitch <= 2 * p.BN && ep.ln_pitch > p.BN, HATB200_EUNSUPPORTED, "gemm_tc: split row wider than the pair");
      dims[0] = (uint64_t)(ep.ln_pitch - p.BN);
      HB_TRY(tc::make_tmap_bf16(&tmLN1, (const bf16*)ep.ln_out + p.BN, 2, dims, str, box, w16 * 2));
    }
  }
  p.rev = ep.walk_back ? 1 : 0;
  p.pin = ep.pin_stream ? 1 : 0;
  p.trace = nullptr; p.prof = 0;
  p.l2_ahead = getenv("HATB200_G3_L2AHEAD") ? atoi(getenv("HATB200_G3_L2AHEAD")) : 0;
  if (getenv("HATB200_GEMM_TRACE")) {
    if (!g_gemm_trace) HB_CHECK_CUDA(cudaMalloc(&g_gemm_trace, 1024 * 64 * 8));
    HB_CHECK_CUDA(cudaMemsetAsync(g_gemm_trace, 0, 1024 * 64 * 8, st));
    p.trace = g_gemm_trace;
    p.prof = getenv("HATB200_GEMM_PROF") != nullptr;
  }
  const size_t smem = (size_t)p.smem_bytes + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    const int lim = G3_SMEM_LIMIT + 1024;
#define G3_ATTR(F_) \
    HB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<F_, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim)); \
    HB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<F_, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim))
    G3_ATTR(FeatAll); G3_ATTR(FeatPlain); G3_ATTR(FeatResLn); G3_ATTR(FeatResULn);
#undef G3_ATTR
    HB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<FeatPlain, true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
#define G3_ATTR_PRE(F_) \
    HB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<F_, false, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim)); \
    HB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<F_, true, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim))
    G3_ATTR_PRE(FeatAll); G3_ATTR_PRE(FeatResLn); G3_ATTR_PRE(FeatResULn);
#undef G3_ATTR_PRE
    HB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<FeatResLn, true, 2, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim - 4096));
    HB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<FeatResULn, true, 2, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim - 4096));
    attr_set = true;
  }
  const int n_tiles = p.split ? 1 : Npad / p.BN;
  int per_n = sm_count() / n_tiles;
  if (per_n < 1) per_n = 1;
  if (per_n > p.m_tiles) per_n = p.m_tiles;
  dim3 grid(per_n, n_tiles);
  if (p.split) {         // clusters of two along x: 74 pairs walk the tiles, rank = column half
    HB_REQUIRE(grouped, HATB200_EUNSUPPORTED, "gemm_tc: split needs grouped stores");
    int pairs = sm_count() / 2;
    if (pairs > p.m_tiles) pairs = p.m_tiles;
    grid = dim3(2 * pairs, 1);
    if (feat_select(ep) == FEAT_RESLN && p.pre)
      launch_pdl_cluster(gemm_tc_kernel<FeatResLn, true, 2, true, true>, 2, grid, dim3(G3_THREADS + 32), smem, st, tmA, tmB, tmR, tmU,
                         tmOF, tmOT, tmLN, tmOF1, tmLN1, p, ep);
    else if (feat_select(ep) == FEAT_RESULN && !p.pre)
      launch_pdl_cluster(gemm_tc_kernel<FeatResULn, true, 2, false, true>, 2, grid, dim3(G3_THREADS), smem, st, tmA, tmB, tmR, tmU,
                         tmOF, tmOT, tmLN, tmOF1, tmLN1, p, ep);
    else {
      set_error("gemm_tc: split planned for an epilogue without a split kernel");
      return HATB200_EUNSUPPORTED;
    }
    HB_CHECK_LAUNCH();
    return 0;
  }
  const bool ring = p.has_resid || p.has_u;
  // epilogue specialisation: the first compile-time feature set that covers this layer exactly, else the generic one
#define G3_LAUNCH(F_)                                                                                                \
  do {                                                                                                               \
    if (grouped) launch_pdl(gemm_tc_kernel<F_, true>, grid, dim3(G3_THREADS), smem, st, tmA, tmB, tmR, tmU, tmOF, tmOT, tmLN, tmOF1, tmLN1, p, ep); \
    else launch_pdl(gemm_tc_kernel<F_, false>, grid, dim3(G3_THREADS), smem, st, tmA, tmB, tmR, tmU, tmOF, tmOT, tmLN, tmOF1, tmLN1, p, ep);       \
  } while (0)
#define G3_LAUNCH_R(F_)                                                                                              \
  do {                                                                                                               \
    if (!p.pre) G3_LAUNCH(F_);                                                                                       \
    else if (grouped) launch_pdl(gemm_tc_kernel<F_, true, 2, true>, grid, dim3(G3_THREADS + 32), smem, st, tmA, tmB, tmR, tmU, tmOF, tmOT, tmLN, tmOF1, tmLN1, p, ep); \
    else launch_pdl(gemm_tc_kernel<F_, false, 2, true>, grid, dim3(G3_THREADS + 32), smem, st, tmA, tmB, tmR, tmU, tmOF, tmOT, tmLN, tmOF1, tmLN1, p, ep);            \
  } while (0)
  static const bool generic_only = getenv("HATB200_EPI_GENERIC") != nullptr;     // engineering: A/B the specialisation
  switch (generic_only ? FEAT_ALL : feat_select(ep)) {
    case FEAT_PLAIN:
      if (p.nwg == 4)
        launch_pdl(gemm_tc_kernel<FeatPlain, true, 4>, grid, dim3(64 + 128 * 4), smem, st, tmA, tmB, tmR, tmU, tmOF, tmOT, tmLN, tmOF1, tmLN1, p, ep);
      else G3_LAUNCH(FeatPlain);
      break;
    case FEAT_RESLN: G3_LAUNCH_R(FeatResLn); break;
    case FEAT_RESULN: G3_LAUNCH_R(FeatResULn); break;
    default: G3_LAUNCH_R(FeatAll); break;
  }
#undef G3_LAUNCH_R
#undef G3_LAUNCH
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb
