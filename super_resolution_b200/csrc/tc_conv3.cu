// tc_conv3.cu -- dense 3x3 conv (stride 1, zero pad 1) over channels-last tokens as a PERSISTENT, halo-resident
// tcgen05 implicit GEMM with the warp-independent TMA epilogue of tc_epilogue3.cuh.
//
//   out[pix, n] = sum_{ky,kx} sum_k A[pix + (ky-1, kx-1), k] * W[ky*3+kx][n][k]      (bf16 x bf16 -> fp32 in TMEM)
//
// Why (B200 measurements of the one-tile-per-CTA kernel tc_tapconv.cu, profiles/r01_*): CAB conv1 / conv2 ran at
// 91 / 60 us per launch with the tensor pipe 2-3 % busy.  Every (tap, 64-channel block) re-fetched its 16 KB
// activation box from L2 (9x the tile), a CTA lived for ONE tile (barrier / TMEM set-up and an un-overlapped
// epilogue per tile, 1-2 CTAs per SM), and the epilogue stored one 16-byte granule per thread and line.  Here:
//   * one CTA per SM walks 8 (x) x 16 (y) pixel tiles.  The (8+2) x (16+2) halo of a tile is TMA-loaded ONCE per
//     64-channel block ({64 ch, 10 x, 18 y} box, out-of-bounds zero fill = the conv padding) and all nine taps
//     address it IN PLACE: the A descriptor of tap (ky, kx) starts at halo row ky*10 + kx and its 8-row groups
//     (8 consecutive x of one image row) are one halo row (1280 B) apart -- tcgen05 swizzles on absolute address
//     bits, so the unaligned start is legal (measured with tc_probe.cu, used by tc_plk.cu);
//   * only the weight slices (BN x 64 per tap and block) stream through a TMA ring; halos are double buffered when
//     shared memory allows, so the next tile's halo lands while this tile's 9 x kblocks x 4 MMAs run;
//   * two TMEM accumulator stages; the two epilogue warpgroups take alternate tiles and run the shared epilogue
//     (bias, GELU / LeakyReLU, fp32 residual through a TMA ring, fp32 / bf16 TMA stores as 4-D boxes
//     {cols, 8 x, 4 y}, pool partials, row-local LayerNorm + 16-channel head).
// Warp roles (384 threads): 0 TMA producer (halos, weights), 1 MMA issuer + TMEM owner, 2-5 / 6-9 epilogue
// warpgroups, 10 / 11 operand-ring producers.
//
// CTA pairs (template TWO, the default when the tile count is even): the one-CTA main loop is bound by shared-memory
// bandwidth, not by the tensor pipe -- an M=128 x N MMA re-reads 4 KB of A and N*32 B of B per k-step while TMA writes
// the next weight slice into the same banks (N=192: 107 + 64 B/clk of 128; N=64: 192 B/clk), measured as 330-370 ns
// per 24 KB slice against 195 ns of tensor time.  With tcgen05 cta_group::2 the two SMs of a TPC run one M=256 MMA
// over two tiles: each CTA keeps its own halo and accumulator rows but only HALF of every weight slice, so the
// weight bytes written to and read from each SM's shared memory halve.  Rank 0 issues all MMAs; both producers load
// into their own shared memory with the bytes completing on rank 0's full barriers; commits multicast to the
// empty / accumulator-full barriers of both CTAs; both epilogues arrive on rank 0's accumulator-empty barrier.
// Replaces F.conv2d at hat_arch.py:84-86 (CAB), :544 (RHAG conv), :854 (conv_after_body), :855 (conv_before_upsample).
#include <stdlib.h>

#include "side_jobs.cuh"
#include "tc_epilogue3.cuh"

namespace hb {

using namespace tc;
using namespace epi3;

unsigned long long* gemm_trace_arm(cudaStream_t st);

namespace {

__device__ __forceinline__ unsigned long long c3_time() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define C3_TRACE(slot) do { if (p.trace) { p.trace[(size_t)blockIdx.x * 64 + (slot)] = c3_time(); p.trace[(size_t)blockIdx.x * 64 + 32 + (slot)] = (unsigned long long)clock64(); } } while (0)

constexpr int C3_THREADS = 384;
constexpr int C3_TW = 8, C3_TH = 16;                   // tile: 128 pixels = UMMA M
constexpr int C3_HW = C3_TW + 2, C3_HH = C3_TH + 2;    // halo
constexpr int C3_HROW = C3_HW * 128;                   // bytes between two halo image rows of one 64-channel block
constexpr int C3_HALO_KB = ((C3_HH * C3_HW * 128 + 1023) / 1024) * 1024;   // 23552: one 64-channel block of a halo
constexpr int C3_MAX_HST = 4, C3_MAX_WST = 16;
constexpr int C3_SMEM_LIMIT = 227 * 1024 - 2048;

struct C3Params {
  int B, H, W, hw, tiles_x, tiles_y, tiles;
  int a_coff, kblocks;
  int N, BN, Npad;
  int hst, halo_stage_bytes;     // halo stages, kblocks * C3_HALO_KB
  int wst, w_stage_bytes, grp;   // weight ring: stages of grp slices of BN (BN / 2 for a CTA pair) x 128 bytes
  int ops, op_stage_bytes, op_u_off, has_resid, has_u;
  uint32_t acc_stride, tmem_cols;
  int off_halo, off_w, off_op, off_stg32, off_stg16, off_tab, off_sred, smem_bytes;
  int stg_bufs, share_stg;
  unsigned long long* trace;     // engineering: per-CTA timestamps (HATB200_GEMM_TRACE)
  int rot;                       // start tap rotated per CTA (HATB200_C3_ROT=0 disables)
  int shuf_r;                    // > 1: pixel-shuffle store (columns pre-permuted to (i, j, c))
  int lean;                      // MMA issue through the unrolled c3_issue_tiles (grp == 3, kblocks 1 or 3, no tap rotation)
  int rev;                       // walk the tiles from the last to the first (Epilogue::walk_back)
  int main_ctas;                 // CTAs that walk tiles; blockIdx.x >= main_ctas runs ep.side (side_jobs.cuh) and exits
  int probe;                     // engineering (HATB200_C3_PROBE=1, with tracing): time 64 back-to-back MMAs before the tile loop
};

// Everything the unrolled MMA issuer needs, as shared-window addresses (descriptor fields are address >> 4).
struct C3Issue {
  int tiles, t_first, t_step, hst, wst;
  uint32_t idesc, tmem_base, acc_stride;
  uint32_t halo16, halo_stage16;         // halo stage 0, stage stride
  uint32_t w16, wslice16;                // weight ring stage 0, one slice
  uint32_t w_full, w_empty, halo_full, halo_empty, tmem_full, tmem_empty;   // barrier arrays (8 bytes per entry)
  unsigned long long* trace;
};

// One thread issues every tile's 36 * KB MMAs as straight-line code: all 9 taps x KB k-blocks unrolled, each operand
// descriptor the tile's (A) or ring stage's (B) base plus an immediate, so that an MMA costs one UIADD3.64 + UTCHMMA
// instead of the looped form's ~70 instructions per group of four (ring cursor, tap stepping, R2UR moves, elect /
// reconvergence).  A ring stage holds 3 slices (producer: grp == 3).
template <int KB, bool TWO>
__device__ __noinline__ void c3_issue_tiles(const C3Issue a) {
  const uint64_t a_tmpl = (make_smem_desc(0, 128) & ~(0x3FFFull << 32)) | ((uint64_t)(C3_HROW >> 4) << 32);
  const uint64_t b_tmpl = make_smem_desc(0, 128);
  uint32_t full_a = a.w_full, empty_a = a.w_empty, wb = a.w16, wph = 0;
  int ws_i = 0;
  uint32_t hs = 0, hph = 0;
  int i = 0;
  unsigned long long* tr = a.trace ? a.trace + (size_t)blockIdx.x * 64 : nullptr;
  for (int t = a.t_first; t < a.tiles; t += a.t_step, i++) {
    const uint32_t g = (uint32_t)i & 1u;
    if (tr && i < 4) tr[4 * i] = c3_time();
    mbar_wait_a(a.tmem_empty + 8 * g, (((uint32_t)i >> 1) & 1u) ^ 1u);     // epilogue has drained this stage
    if (tr && i < 4) tr[4 * i + 1] = c3_time();
    mbar_wait_a(a.halo_full + 8 * hs, hph);
    tc_fence_after();
    if (tr && i < 4) { tr[4 * i + 2] = c3_time(); tr[32 + 4 * i + 2] = (unsigned long long)clock64(); }
    const uint32_t acc = a.tmem_base + g * a.acc_stride;
    const uint64_t ad0 = a_tmpl | (uint64_t)(a.halo16 + hs * a.halo_stage16);
#pragma unroll
    for (int st = 0; st < 3 * KB; st++) {
      mbar_wait_a(full_a, wph);
      tc_fence_after();
      const uint64_t bd0 = b_tmpl | (uint64_t)wb;
#pragma unroll
      for (int q = 0; q < 3; q++) {
        const int n = st * 3 + q, tap = n / KB, kb = n % KB, ky = tap / 3, kx = tap % 3;
        // accumulator row (y, x) reads halo row (y + ky) * 10 + x + kx of k-block kb
        const uint64_t ad = ad0 + (uint64_t)((ky * C3_HW + kx) * (128 >> 4) + kb * (C3_HALO_KB >> 4));
        const uint64_t bd = bd0 + (uint64_t)((uint32_t)q * a.wslice16);
#pragma unroll
        for (int k = 0; k < 4; k++) {
          if (TWO) umma2_bf16(acc, ad + 2 * k, bd + 2 * k, a.idesc, (n | k) ? 1u : 0u);
          else umma_bf16(acc, ad + 2 * k, bd + 2 * k, a.idesc, (n | k) ? 1u : 0u);
        }
      }
      if (TWO) umma2_commit_a(empty_a);          // frees the ring stage once its MMAs have read it
      else umma_commit_a(empty_a);
      full_a += 8; empty_a += 8; wb += 3 * a.wslice16;
      if (++ws_i == a.wst) { ws_i = 0; full_a = a.w_full; empty_a = a.w_empty; wb = a.w16; wph ^= 1u; }
    }
    if (tr && i < 4) { tr[4 * i + 3] = c3_time(); tr[32 + 4 * i + 3] = (unsigned long long)clock64(); }
    if (TWO) {
      umma2_commit_a(a.halo_empty + 8 * hs);
      umma2_commit_a(a.tmem_full + 8 * g);
    } else {
      umma_commit_a(a.halo_empty + 8 * hs);      // halo may be overwritten
      umma_commit_a(a.tmem_full + 8 * g);        // accumulator of tile i complete
    }
    if (++hs == (uint32_t)a.hst) { hs = 0; hph ^= 1u; }
  }
}

// the t-th tile of the walk (backwards when p.rev: Epilogue::walk_back); returns the tile's index in image order
__device__ __forceinline__ int tile_coords(const C3Params& p, int t, int& x0, int& y0, int& b) {
  if (p.rev) t = p.tiles - 1 - t;
  const int tx = t % p.tiles_x, ty = (t / p.tiles_x) % p.tiles_y;
  b = t / (p.tiles_x * p.tiles_y);
  x0 = tx * C3_TW; y0 = ty * C3_TH;
  return t;
}

}  // namespace

template <class F, bool GROUPED, bool TWO>
__global__ void __launch_bounds__(C3_THREADS, 1)
conv3_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmOF,
                const __grid_constant__ CUtensorMap tmOT, const __grid_constant__ CUtensorMap tmLN, const C3Params p,
                const Epilogue ep) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t halo_full[C3_MAX_HST];
  __shared__ __align__(8) uint64_t halo_empty[C3_MAX_HST];
  __shared__ __align__(8) uint64_t w_full[C3_MAX_WST];
  __shared__ __align__(8) uint64_t w_empty[C3_MAX_WST];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ __align__(8) uint64_t op_full[2][MAX_OPS];
  __shared__ __align__(8) uint64_t op_empty[2][MAX_OPS];
  __shared__ uint32_t tmem_slot;
  __shared__ float hsum[8][16];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  if ((int)blockIdx.x >= p.main_ctas) {
    // extra CTAs (whole clusters): a small per-sample job that only needs results of EARLIER kernels.  They are
    // dispatched when the first tile-walking CTAs retire, i.e. inside this launch's tail, and cost no launch.
    pdl_wait();
    pdl_trigger();
    run_side_job(ep.side, (int)blockIdx.x - p.main_ctas, reinterpret_cast<float*>(smem));
    return;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.y * p.BN;
  // CTA pair: rank r of cluster q takes tile 2 * pair + r of every pair the cluster walks
  const uint32_t rank = TWO ? cluster_ctarank() : 0u;
  const int t_first = TWO ? (int)(blockIdx.x & ~1u) + (int)rank : (int)blockIdx.x;
  const int t_step = p.main_ctas;
  float* tab_bias = reinterpret_cast<float*>(smem + p.off_tab);
  float* tab_gamma = tab_bias + p.BN;
  float* tab_beta = tab_gamma + p.BN;
  float* tab_coef = tab_beta + p.BN;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < p.hst; s++) { mbar_init(&halo_full[s], 1); mbar_init(&halo_empty[s], 1); }
    for (int s = 0; s < p.wst; s++) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int g = 0; g < 2; g++) {
      mbar_init(&tmem_full_bar[g], 1);
      mbar_init(&tmem_empty_bar[g], TWO ? 256 : 128);
      for (int s = 0; s < MAX_OPS; s++) { mbar_init(&op_full[g][s], 1); mbar_init(&op_empty[g][s], 4); }
    }
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < p.BN; i += C3_THREADS) {
    const int n = n0 + i;
    const bool ok = n < p.N;
    tab_bias[i] = (ok && ep.bias) ? __ldg(ep.bias + n) : 0.f;
    tab_gamma[i] = (ok && ep.ln_gamma) ? __ldg(ep.ln_gamma + n) : 0.f;
    tab_beta[i] = (ok && ep.ln_beta) ? __ldg(ep.ln_beta + n) : 0.f;
  }
  if (warp == 1) {
    if (TWO) tmem_alloc2(&tmem_slot, p.tmem_cols);
    else tmem_alloc(&tmem_slot, p.tmem_cols);
  }
  tc_fence_before();
  if (TWO) cluster_sync_all();        // the peer's barriers are initialised before anything remote touches them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  // All CTAs stream the same 9 * kblocks weight slices per tile; each starts at a different tap (and wraps) so that
  // 148 SMs do not ask the same L2 lines for the same slice at the same moment.  The accumulation order of a tile
  // then depends on its CTA: results stay deterministic, sums differ in the last fp32 bit between tiles.
  const int nsl = 9 * p.kblocks;
  const int tap0 = p.rot ? (int)((TWO ? blockIdx.x >> 1 : blockIdx.x) % 9u) : 0;
  if (warp == 0) {
    // ===== TMA producer: one halo per tile (kblocks boxes), 9 * kblocks weight slices per tile =====
    // (the single issuing thread is picked with elect.sync, not `lane == 0`: ptxas then knows the region is
    //  one-thread and emits UTMALDG / UTCHMMA straight from uniform registers instead of a per-lane waterfall loop)
    if (elect_one()) {
      pdl_wait();        // halos are the previous kernel's output
      pdl_trigger();
      RingPos hp, wp;
      // Halos never block the weight stream: the next halo goes out as soon as its stage is free, polled between
      // weight slices and inside the wait for a free ring stage.  (Issued at a fixed point of the slice sequence, a
      // halo that had to wait for the previous tile's MMAs held back every weight slice behind it, and each tile then
      // paid an L2 round trip for its last slices: B200 trace 0.24 us per slice whatever N, 2.2 us per K = 64 tile
      // whose MMAs take 0.5 us.)
      int halo_t = t_first;
      auto try_halo = [&](bool block) {
        if (halo_t >= p.tiles) return;
        const int hs = hp.s;
        if (block) mbar_wait(&halo_empty[hs], hp.ph ^ 1u);
        else if (!mbar_try_wait(&halo_empty[hs], hp.ph ^ 1u)) return;
        int x0, y0, b;
        tile_coords(p, halo_t, x0, y0, b);
        uint8_t* hb_ = smem + p.off_halo + (size_t)hs * p.halo_stage_bytes;
        const uint32_t hbytes = (uint32_t)(p.kblocks * C3_HH * C3_HW * 128);
        if (TWO) {       // both halos of the pair complete on rank 0's barrier
          if (rank == 0) mbar_arrive_expect_tx(&halo_full[hs], 2 * hbytes);
          const uint32_t bar0 = mapa_u32(&halo_full[hs], 0);
          for (int kb = 0; kb < p.kblocks; kb++)
            tma2_load_4d(hb_ + (size_t)kb * C3_HALO_KB, &tmA, bar0, p.a_coff + kb * 64, x0 - 1, y0 - 1, b);
        } else {
          mbar_arrive_expect_tx(&halo_full[hs], hbytes);
          for (int kb = 0; kb < p.kblocks; kb++)
            tma_load_4d(hb_ + (size_t)kb * C3_HALO_KB, &tmA, &halo_full[hs], p.a_coff + kb * 64, x0 - 1, y0 - 1, b);
        }
        hp.next(p.hst);
        halo_t += t_step;
      };
      // A ring stage carries `grp` consecutive weight slices (one full barrier, one release): see the MMA issuer.
      const uint32_t slice_bytes = (uint32_t)(p.w_stage_bytes / p.grp);
      try_halo(true);
      if (p.lean) {
        // One 4-D box per ring stage ([tap][k-block][row][64] view of the weights: 3 k-blocks of a tap, or 3 taps of
        // a one-block layer): this thread runs alone on its scheduler slot at ~10 cycles per dependent instruction,
        // and a per-slice loop (~45 instructions + 1 TMA each) fed the ring slower than the MMAs drain it.
        const uint32_t full0 = smem_u32(&w_full[0]), empty0 = smem_u32(&w_empty[0]);
        const uint32_t full0c = TWO ? mapa_u32(&w_full[0], 0) : full0;
        const uint32_t dst0 = smem_u32(smem + p.off_w), stage_b = (uint32_t)p.w_stage_bytes;
        const uint32_t tx = (TWO ? 2u : 1u) * stage_b;
        const int row0 = n0 + (int)rank * (p.BN / 2) * (TWO ? 1 : 0);
        const int nst = 3 * p.kblocks, dtap = p.kblocks == 3 ? 1 : 3;
        uint32_t full_a = full0, full_c = full0c, empty_a = empty0, dst = dst0, ph = 1u;
        int ws_i = 0;
        for (int t = t_first; t < p.tiles; t += t_step) {
          for (int st = 0, tap = 0; st < nst; st++, tap += dtap) {
            try_halo(false);
            if (!mbar_try_wait_a(empty_a, ph)) {
              const long long t0 = clock64();
              while (!mbar_try_wait_a(empty_a, ph)) {
                try_halo(false);
                if (clock64() - t0 > 4000000000ll) {
                  printf("libhatb200: conv3 weight ring wait timed out (block %d,%d)\n", blockIdx.x, blockIdx.y);
                  __trap();
                }
              }
            }
            if (!TWO || rank == 0) mbar_arrive_expect_tx_a(full_a, tx);
            if (TWO) tma2_load_4d_a(dst, &tmB, full_c, 0, row0, 0, tap);
            else tma_load_4d_a(dst, &tmB, full_a, 0, row0, 0, tap);
            full_a += 8; full_c += 8; empty_a += 8; dst += stage_b;
            if (++ws_i == p.wst) { ws_i = 0; full_a = full0; full_c = full0c; empty_a = empty0; dst = dst0; ph ^= 1u; }
          }
        }
      } else
      for (int t = t_first; t < p.tiles; t += t_step) {
        int tap = tap0, q = 0;
        for (int j = 0; j < 9; j++) {
          for (int kb = 0; kb < p.kblocks; kb++) {
            try_halo(false);
            const int s = wp.s;
            uint8_t* dst = smem + p.off_w + (size_t)s * p.w_stage_bytes + (size_t)q * slice_bytes;
            if (q == 0) {
              if (!mbar_try_wait(&w_empty[s], wp.ph ^ 1u)) {
                const long long t0 = clock64();
                while (!mbar_try_wait(&w_empty[s], wp.ph ^ 1u)) {
                  try_halo(false);
                  if (clock64() - t0 > 4000000000ll) {
                    printf("libhatb200: conv3 weight ring wait timed out (block %d,%d)\n", blockIdx.x, blockIdx.y);
                    __trap();
                  }
                }
              }
              if (!TWO) mbar_arrive_expect_tx(&w_full[s], (uint32_t)p.w_stage_bytes);
              else if (rank == 0) mbar_arrive_expect_tx(&w_full[s], 2u * (uint32_t)p.w_stage_bytes);
            }
            if (TWO)            // this CTA's half of the slice (rows [rank * BN/2, +BN/2)); both halves complete on rank 0
              tma2_load_2d(dst, &tmB, mapa_u32(&w_full[s], 0), kb * 64, tap * p.Npad + n0 + (int)rank * (p.BN / 2));
            else
              tma_load_2d(dst, &tmB, &w_full[s], kb * 64, tap * p.Npad + n0);
            if (++q == p.grp) { q = 0; wp.next(p.wst); }
          }
          if (++tap == 9) tap = 0;
        }
      }
      while (halo_t < p.tiles) try_halo(true);
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the loop so that ring cursors, descriptors and barrier addresses stay
    // warp-uniform (uniform registers, no R2UR per operand); one elected lane issues each tcgen05 instruction =====
    if (rank == 0) {
      const uint32_t idesc = make_idesc_bf16(TWO ? 256 : 128, p.BN);
      // The instructions this warp executes between two MMA groups are NOT hidden by the tensor pipe: tcgen05.mma
      // issue blocks on a queue about two MMAs deep, and a per-slice iteration (barrier poll, fence, descriptor
      // assembly, R2UR moves, commit, ring step) costs ~200 cycles against 4 x 48 (N = 64) / 4 x 96 (N = 192) cycles
      // of MMAs -- B200 trace: 225 / 330 ns per slice, 95 / 110 cycles per MMA even with the ring removed.  So a ring
      // stage carries `grp` slices (12 MMAs for grp = 3) behind ONE barrier poll and ONE commit, and everything else
      // is stepped incrementally out of registers.
      const uint32_t full0 = smem_u32(&w_full[0]), empty0 = smem_u32(&w_empty[0]);
      const uint32_t wb0 = (smem_u32(smem + p.off_w) & 0x3FFFF) >> 4;
      const int wst = p.wst, kblocks = p.kblocks, grp = p.grp;
      const uint32_t wslice = (uint32_t)(p.w_stage_bytes / grp) >> 4;
      // A descriptor template: 128-byte swizzled rows, 8-row groups one halo image row (C3_HROW bytes) apart
      const uint64_t a_tmpl = (make_smem_desc(0, 128) & ~(0x3FFFull << 32)) | ((uint64_t)(C3_HROW >> 4) << 32);
      const uint64_t b_tmpl = make_smem_desc(0, 128);
      uint32_t full_a = full0, empty_a = empty0, wb = wb0, wph = 0;
      int ws_i = 0;
      RingPos hp;
      int i = 0;
      if ((p.probe & 1) && p.trace) {    // MMA issue rate in this kernel's environment, nothing else running on the pipe
        const uint32_t ha = (smem_u32(smem + p.off_halo) & 0x3FFFF) >> 4;
        if (elect_one()) {
          const long long c0 = clock64();
          for (int n = 0; n < 16; n++) {
            const uint64_t ad = a_tmpl | (uint64_t)(ha + (uint32_t)((n % 9) * 8)), bd = b_tmpl | (uint64_t)wb0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
              if (TWO) umma2_bf16(tmem_base, ad + 2 * k, bd + 2 * k, idesc, 1u);
              else umma_bf16(tmem_base, ad + 2 * k, bd + 2 * k, idesc, 1u);
            }
          }
          p.trace[(size_t)blockIdx.x * 64 + 31] = (unsigned long long)(clock64() - c0);
        }
        __syncwarp();
      }
      if (p.lean) {
        if (elect_one()) {
          C3Issue ia;
          ia.tiles = p.tiles; ia.t_first = t_first; ia.t_step = t_step; ia.hst = p.hst; ia.wst = wst;
          ia.idesc = idesc; ia.tmem_base = tmem_base; ia.acc_stride = p.acc_stride;
          ia.halo16 = (smem_u32(smem + p.off_halo) & 0x3FFFF) >> 4; ia.halo_stage16 = (uint32_t)p.halo_stage_bytes >> 4;
          ia.w16 = wb0; ia.wslice16 = wslice;
          ia.w_full = full0; ia.w_empty = empty0;
          ia.halo_full = smem_u32(&halo_full[0]); ia.halo_empty = smem_u32(&halo_empty[0]);
          ia.tmem_full = smem_u32(&tmem_full_bar[0]); ia.tmem_empty = smem_u32(&tmem_empty_bar[0]);
          ia.trace = p.trace;
          if (p.kblocks == 3) c3_issue_tiles<3, TWO>(ia);
          else c3_issue_tiles<1, TWO>(ia);
        }
        __syncwarp();
      } else
      for (int t = t_first; t < p.tiles; t += t_step, i++, hp.next(p.hst)) {
        const int g = i & 1;
        if (i < 4 && lane == 0) C3_TRACE(4 * i);
        mbar_wait(&tmem_empty_bar[g], (((uint32_t)i >> 1) & 1u) ^ 1u);     // epilogue has drained this stage
        const int hs = hp.s;
        if (i < 4 && lane == 0) C3_TRACE(4 * i + 1);
        mbar_wait(&halo_full[hs], hp.ph);
        tc_fence_after();
        if (i < 4 && lane == 0) C3_TRACE(4 * i + 2);
        const uint32_t acc = tmem_base + (uint32_t)g * p.acc_stride;
        const uint32_t ha = (smem_u32(smem + p.off_halo + (size_t)hs * p.halo_stage_bytes) & 0x3FFFF) >> 4;
        uint32_t first = 0;
        int ky = tap0 / 3, kx = tap0 - ky * 3, q = 0;
        for (int j = 0; j < 9; j++) {
          // accumulator row (y, x) reads halo row (y + ky) * 10 + x + kx
          uint32_t a_lo = ha + (uint32_t)((ky * C3_HW + kx) * (128 >> 4));
          for (int kb = 0; kb < kblocks; kb++) {
            if (q == 0) {
              mbar_wait_a(full_a, wph);
              tc_fence_after();
            }
            const uint64_t ad = a_tmpl | (uint64_t)a_lo, bd = b_tmpl | (uint64_t)wb;
            const bool last = q == grp - 1;
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < 4; k++) {
                if (TWO) umma2_bf16(acc, ad + 2 * k, bd + 2 * k, idesc, first | (uint32_t)k);
                else umma_bf16(acc, ad + 2 * k, bd + 2 * k, idesc, first | (uint32_t)k);
              }
              if (last) {                              // frees the ring stage once its MMAs have read it
                if (TWO) umma2_commit_a(empty_a);
                else umma_commit_a(empty_a);
              }
            }
            __syncwarp();
            first = 1;
            a_lo += (uint32_t)(C3_HALO_KB >> 4);
            wb += wslice;
            if (last) {
              q = 0;
              full_a += 8; empty_a += 8;
              if (++ws_i == wst) { ws_i = 0; full_a = full0; empty_a = empty0; wb = wb0; wph ^= 1u; }
            } else {
              q++;
            }
          }
          if (++kx == 3) { kx = 0; if (++ky == 3) ky = 0; }
        }
        if (i < 4 && lane == 0) C3_TRACE(4 * i + 3);
        if (elect_one()) {
          if (TWO) {
            umma2_commit(&halo_empty[hs]);
            umma2_commit(&tmem_full_bar[g]);
          } else {
            umma_commit(&halo_empty[hs]);     // halo may be overwritten
            umma_commit(&tmem_full_bar[g]);   // accumulator of tile i complete
          }
        }
        __syncwarp();
      }
    }
  } else if (warp >= 10) {
    // ===== operand-ring producer of warpgroup g (fp32 residual chunks) =====
    const int g = warp - 10;
    if (feat_ring<F>() && p.has_resid && elect_one()) {
      prefetch_tmap(&tmR);
      pdl_wait();
      uint8_t* ring = smem + p.off_op + (size_t)g * p.ops * p.op_stage_bytes;
      RingPos k;
      int i = 0;
      for (int t = t_first; t < p.tiles; t += t_step, i++) {
        if ((i & 1) != g) continue;
        Tile4D ts;
        tile_coords(p, t, ts.x0, ts.y0, ts.b);
        produce_tile(ts, &tmR, &tmR, ring, op_full[g], op_empty[g], p.ops, p.op_stage_bytes, p.op_u_off, p.has_resid, 0,
                     n0, p.BN / CW, k);
      }
    }
  } else {
    // ===== epilogue warps =====
    const int g = (warp - 2) >> 2;
    WarpState<GROUPED> ws;
    ws.g = g; ws.ew = warp - 2; ws.quarter = warp & 3; ws.lane = lane;
    ws.row = ws.quarter * 32 + lane;
    ws.wg_tid = ((warp - 2) & 3) * 32 + lane;
    ws.BN = p.BN; ws.N = p.N; ws.n0 = n0; ws.nch = p.BN / CW; ws.H = p.H; ws.W = p.W; ws.hw = p.hw;
    ws.last_ntile = blockIdx.y == gridDim.y - 1;
    ws.tab_bias = tab_bias; ws.tab_gamma = tab_gamma; ws.tab_beta = tab_beta; ws.coef = tab_coef + g * p.BN;
    ws.ring = smem + p.off_op + (size_t)g * p.ops * p.op_stage_bytes;
    ws.op_full = op_full[g]; ws.op_empty = op_empty[g];
    ws.ops = p.ops; ws.op_stage_bytes = p.op_stage_bytes; ws.op_u_off = p.op_u_off;
    ws.has_resid = p.has_resid; ws.has_u = 0; ws.share_stg = p.share_stg;
    ws.hsum = hsum;
    ws.sred = reinterpret_cast<float (*)[256]>(smem + p.off_sred + (size_t)g * 4096);
    ws.st32.nbuf = ws.st16.nbuf = p.stg_bufs;
    ws.st32.base = smem + p.off_stg32 + (size_t)ws.ew * (p.stg_bufs * decltype(ws.st32)::BUF);
    ws.st16.base = smem + p.off_stg16 + (size_t)ws.ew * (p.stg_bufs * decltype(ws.st16)::BUF);
    if (lane == 0) {
      if (ep.out_f32) prefetch_tmap(&tmOF);
      if (ep.out_t) prefetch_tmap(&tmOT);
      if (ep.ln_out) prefetch_tmap(&tmLN);
    }
    pdl_wait();          // the epilogue writes the outputs (and reads the residual ring)
    const uint32_t empty0[2] = {TWO ? mapa_u32(&tmem_empty_bar[0], 0) : 0u, TWO ? mapa_u32(&tmem_empty_bar[1], 0) : 0u};
    int i = 0;
    for (int t = t_first; t < p.tiles; t += t_step, i++) {
      if ((i & 1) != g) continue;
      Tile4D ts;
      const int te = tile_coords(p, t, ts.x0, ts.y0, ts.b);
      ts.r = p.shuf_r; ts.H = p.H;
      RowInfo ri;
      ri.b = ts.b;
      ri.py = ts.y0 + ws.row / C3_TW; ri.px = ts.x0 + ws.row % C3_TW;
      ri.m = (ts.b * p.H + ri.py) * p.W + ri.px;
      ri.nchunk = p.tiles_x * p.tiles_y;
      ri.chunk = te - ts.b * ri.nchunk;
      ri.trow = tmem_base + (uint32_t)g * p.acc_stride + ((uint32_t)(ws.quarter * 32) << 16);
      const bool tr = ws.wg_tid == 0 && i < 4;
      if (tr) C3_TRACE(16 + 4 * i);
      mbar_wait(&tmem_full_bar[g], ((uint32_t)i >> 1) & 1u);
      tc_fence_after();
      if (tr) C3_TRACE(17 + 4 * i);
      if (F::LAST) {
        // conv_last in one pass (hat_arch.py:858): the accumulator's first columns are the image channels.  Four
        // columns per row out of TMEM, the stage handed back at once, then (acc + bias) / img_range + mean straight
        // into the NCHW planes (8-pixel runs per image row), or the tensor2img bytes of it (Epilogue::nchw_*).
        uint32_t a4[4];
        tmem_ld4(ri.trow, a4);
        tmem_ld_wait();
        tc_fence_before();
        if (TWO) mbar_arrive_relaxed_cluster(empty0[g]);       // relaxed: see tc_common.cuh (image stores are in flight)
        else mbar_arrive_relaxed(&tmem_empty_bar[g]);
        const size_t hw = (size_t)p.H * p.W, pix = (size_t)ri.py * p.W + ri.px;
#pragma unroll
        for (int o = 0; o < 4; o++) {
          if (o >= ep.nchw_oc || (p.probe & 2)) break;
          float v = __fadd_rn(__fdiv_rn(__uint_as_float(a4[o]) + tab_bias[o], ep.nchw_range), ep.nchw_mean[o]);
          if (ep.nchw_dt == HATB200_F32) {
            reinterpret_cast<float*>(ep.nchw_y)[((size_t)ts.b * ep.nchw_oc + o) * hw + pix] = v;
          } else if (ep.nchw_dt == HATB200_BF16) {
            reinterpret_cast<bf16*>(ep.nchw_y)[((size_t)ts.b * ep.nchw_oc + o) * hw + pix] = __float2bfloat16_rn(v);
          } else {                        // HATB200_U8: tensor2img of the value as the compute dtype holds it
            if (ep.nchw_via_bf16) v = __bfloat162float(__float2bfloat16_rn(v));
            v = fminf(fmaxf(v, 0.f), 1.f);
            reinterpret_cast<uint8_t*>(ep.nchw_y)[((size_t)ts.b * hw + pix) * ep.nchw_oc + (ep.nchw_oc - 1 - o)] =
                (uint8_t)rintf(__fmul_rn(v, 255.0f));
          }
        }
      } else {
        run_tile<F, GROUPED>(ep, ws, ri, ts, &tmOF, &tmOT, &tmLN);
        tc_fence_before();
        // rank 0 issues the next MMAs into both CTAs' stage g.  Relaxed (tc_common.cuh): the barrier only guards the TMEM
        // stage, whose loads / stores have completed (wait::ld / wait::st); a release would also wait for the generic
        // stores of the tile (pool partials, LayerNorm head) at GPU scope
        if (TWO) mbar_arrive_relaxed_cluster(empty0[g]);
        else mbar_arrive_relaxed(&tmem_empty_bar[g]);
      }
      if (tr) C3_TRACE(18 + 4 * i);
    }
    if (elect_one()) bulk_wait_all();
  }
  tc_fence_before();
  if (TWO) cluster_sync_all();        // the peer may still read this CTA's operands / arrive on its barriers
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (TWO) tmem_dealloc2(tmem_base, p.tmem_cols);
    else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ---- host ------------------------------------------------------------------------------------------
// set when a cluster launch was refused (no co-schedulable SM pairs, e.g. a partitioned GPU): one-CTA kernels from then on
static bool g_c3_pairs_failed = false;

static int c3_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        n <= 0)
      n = 148;
  }
  return n;
}

static bool conv3_geometry(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep, C3Params& p, bool& grouped,
                           bool& two) {
  if (a.KH != 3 || a.KW != 3) return false;
  if (a.W % C3_TW || a.H % C3_TH) return false;
  if (Kpad % 64 || Kpad < a.Cin) return false;
  if ((a.a_pitch * 2) % 16 || (a.a_coff * 2) % 16) return false;
  p.BN = tc_bn(a.N);
  if (Npad != tc_npad(a.N) || p.BN < 16 || p.BN > 256 || p.BN % 16) return false;
  if (ep.dyn_k || ep.eca_u || ep.scale_ncols) return false;      // tc_tapconv.cu keeps those layers
  if (ep.nchw_y && (ep.nchw_oc < 1 || ep.nchw_oc > 4 || a.N != ep.nchw_oc || Npad != p.BN || ep.out_f32 || ep.out_t || ep.ln_out ||
                    ep.colsum || ep.resid || ep.act != ACT_NONE || ep.shuffle_r != 1))
    return false;
  if (ep.shuffle_r != 1) {      // conv + PixelShuffle: only with (i, j, c)-ordered rows, plain bf16 output of 64 channels
    const int r = ep.shuffle_r;
    if (!ep.shuffle_perm || (r != 2 && r != 3) || a.N != 64 * r * r || p.BN % 64) return false;
    if (!ep.out_t || ep.out_f32 || ep.ln_out || ep.colsum || ep.resid || ep.out_t_pitch != 64 || ep.out_t_coff ||
        ep.out_t_zero_to > a.N)
      return false;
  } else if (ep.shuffle_perm) {
    return false;
  }
  p.shuf_r = ep.shuffle_r;
  if (ep.out_f32 && ((ep.out_f32_pitch * 4) % 16 || ((uintptr_t)ep.out_f32 & 15))) return false;
  if (ep.out_t && ((ep.out_t_pitch * 2) % 16 || (ep.out_t_coff * 2) % 16 || ((uintptr_t)ep.out_t & 15))) return false;
  if (ep.out_t && ep.shuffle_r == 1 && ep.out_t_pitch - ep.out_t_coff < (ep.out_t_zero_to > a.N ? ep.out_t_zero_to : a.N)) return false;
  if (ep.ln_out && ((ep.ln_pitch * 2) % 16 || ((uintptr_t)ep.ln_out & 15) || ep.ln_pitch < a.N)) return false;
  if (ep.resid && ((ep.resid_pitch * 4) % 16 || ((uintptr_t)ep.resid & 15))) return false;
  if (ep.ln_out && Npad != p.BN) return false;
  if (ep.ln_head && (ep.ln_head_c > 16 || !ep.ln_out)) return false;
  if (ep.colsum && p.BN > 256) return false;
  p.B = a.B; p.H = a.H; p.W = a.W; p.hw = a.H * a.W;
  p.tiles_x = a.W / C3_TW; p.tiles_y = a.H / C3_TH; p.tiles = a.B * p.tiles_x * p.tiles_y;
  p.a_coff = a.a_coff; p.kblocks = Kpad / 64;
  p.N = a.N; p.Npad = Npad;
  // CTA pairs need an even tile count (both ranks walk the same number of tiles) and an even number of CTAs per n-tile
  static const bool two_off = getenv("HATB200_C3_TWO") && getenv("HATB200_C3_TWO")[0] == '0';
  two = !two_off && !g_c3_pairs_failed && p.tiles % 2 == 0 && p.tiles >= 2 && c3_sm_count() / (Npad / p.BN) >= 2;
  p.acc_stride = (uint32_t)round_up(p.BN, 32);
  uint32_t cols = 32;
  while (cols < 2 * p.acc_stride) cols <<= 1;
  if (cols > 512) return false;
  p.tmem_cols = cols;
  p.has_resid = ep.resid != nullptr; p.has_u = 0;
  const bool ring = p.has_resid;
  grouped = p.BN % 64 == 0;
  const bool f32o = ep.out_f32 != nullptr, bf16o = ep.out_t != nullptr || ep.ln_out != nullptr;
  p.share_stg = f32o && bf16o && ep.out_t == nullptr;
  const int tab = round_up(5 * p.BN * 4, 1024);
  const int sred = ep.colsum ? 2 * 4096 : 0;
  p.op_u_off = 128 * CW * 4;
  p.op_stage_bytes = ring ? 128 * CW * 4 : 0;
  p.halo_stage_bytes = p.kblocks * C3_HALO_KB;
  const int slice_bytes = (two ? p.BN / 2 : p.BN) * 128;     // a CTA of a pair holds half of every weight slice
  p.ops = ring ? 2 : 0;
  // Prefer ring stages of 3 slices (>= 2 of them: one barrier round trip per 12 MMAs), double-buffered halos and
  // double-buffered staging; fall back to single-slice stages (>= 3) before giving up the second halo buffer.
  static const int force_grp = getenv("HATB200_C3_GRP") ? atoi(getenv("HATB200_C3_GRP")) : 0;
  for (int pass = 0; pass < 16; pass++) {
    static const int force_hst = getenv("HATB200_C3_HST") ? atoi(getenv("HATB200_C3_HST")) : 0;
    static const int force_bufs = getenv("HATB200_C3_BUFS") ? atoi(getenv("HATB200_C3_BUFS")) : 0;
    // pass order: (grp 3, hst 2, bufs 2/1), (grp 1, hst 2, bufs 2/1), (grp 3, hst 1, ..), (grp 1, hst 1, ..), then want - 1
    int grp = (pass & 2) ? 1 : 3;
    int hst = (pass & 4) ? 1 : 2, bufs = (pass & 1) || (f32o && ep.out_t) ? 1 : 2;
    if (force_grp) grp = force_grp;
    const int want = (grp == 3 ? 2 : 3) - (pass >= 8 ? 1 : 0);
    // conv_last (64 -> 3 at HR size): 0.5 us of MMAs per tile and an epilogue of a few instructions, so the halo of tile
    // t + 2 (issued when tile t's MMAs retire) is late with two stages; a third one took 78.3 -> 65.9 us
    if (ep.nchw_y && !(pass & 4)) hst = 3;
    if (force_hst) hst = force_hst;
    if (force_bufs) bufs = force_bufs;
    p.grp = grp;
    static const bool lean_off = getenv("HATB200_C3_LEAN") && getenv("HATB200_C3_LEAN")[0] == '0';
    p.lean = !lean_off && grp == 3 && (p.kblocks == 1 || p.kblocks == 3);
    p.w_stage_bytes = grp * slice_bytes;
    const int stg32 = f32o ? 8 * bufs * 32 * (grouped ? 128 : 64) : 0;
    const int stg16 = bf16o && !p.share_stg ? 8 * bufs * 32 * (grouped ? 128 : 32) : 0;
    const int fixed = tab + sred + stg32 + stg16 + 2 * p.ops * p.op_stage_bytes + hst * p.halo_stage_bytes;
    const int avail = C3_SMEM_LIMIT - fixed;
    const int wst = avail > 0 ? avail / p.w_stage_bytes : 0;
    if (wst < want || wst < 1) continue;
    p.hst = hst; p.stg_bufs = bufs;
    p.wst = wst > C3_MAX_WST ? C3_MAX_WST : wst;
    if (ring) {                      // spend what is left on a deeper operand ring
      int left = avail - p.wst * p.w_stage_bytes;
      while (p.ops < MAX_OPS && left >= 2 * p.op_stage_bytes) { p.ops++; left -= 2 * p.op_stage_bytes; }
    }
    int off = 0;
    p.off_halo = off; off += hst * p.halo_stage_bytes;
    p.off_w = off; off += p.wst * p.w_stage_bytes;
    p.off_op = off; off += 2 * p.ops * p.op_stage_bytes;
    p.off_stg32 = off; off += stg32;
    p.off_stg16 = p.share_stg ? p.off_stg32 : off; off += stg16;
    p.off_tab = off; off += tab;
    p.off_sred = off; off += sred;
    p.smem_bytes = off;
    if (getenv("HATB200_TRACE"))
      fprintf(stderr, "[conv3_tc] N=%d K=%d BN=%d two=%d grp=%d wst=%d hst=%d stg_bufs=%d ops=%d smem=%d\n", p.N, Kpad, p.BN,
              (int)two, p.grp, p.wst, p.hst, p.stg_bufs, p.ops, p.smem_bytes);
    return off <= C3_SMEM_LIMIT;
  }
  return false;
}

bool conv3_tc_supported(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep) {
  C3Params p;
  bool grouped, two;
  if (!conv3_geometry(a, Npad, Kpad, ep, p, grouped, two)) return false;
  if (tc::encode_tiled_fn() == nullptr) return false;
  const char* v = getenv("HATB200_CONV3");
  if (v && *v == '0') return false;
  return true;
}

// can this launch carry ep.side as extra CTAs?  (the tcgen05 kernel takes the shape and it is one n-tile)
bool conv3_tc_side_ok(const ConvArgs& a, int Npad, int Kpad, const Epilogue& ep) {
  return conv3_tc_supported(a, Npad, Kpad, ep) && Npad == tc_bn(a.N);
}

int launch_conv3_tc(const ConvArgs& a, const bf16* w, int Npad, int Kpad, const Epilogue& ep, cudaStream_t st) {
  C3Params p;
  bool grouped, two;
  HB_REQUIRE(conv3_geometry(a, Npad, Kpad, ep, p, grouped, two), HATB200_EUNSUPPORTED, "conv3_tc: unsupported shape");
  CUtensorMap tmA, tmB, tmR, tmOF, tmOT, tmLN;
  {
    uint64_t dims[4] = {(uint64_t)a.a_pitch, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.B};
    uint64_t str[3] = {(uint64_t)a.a_pitch * 2, (uint64_t)a.W * a.a_pitch * 2, (uint64_t)a.H * a.W * a.a_pitch * 2};
    uint32_t box[4] = {64, C3_HW, C3_HH, 1};
    HB_TRY(tc::make_tmap_bf16(&tmA, a.A, 4, dims, str, box, 128));
  }
  if (p.lean) {      // [tap][k-block][row][64]: a ring stage (3 k-blocks of one tap, or 3 taps of a one-block layer) is one box
    uint64_t dims[4] = {64, (uint64_t)Npad, (uint64_t)p.kblocks, 9};
    uint64_t str[3] = {(uint64_t)Kpad * 2, 128, (uint64_t)Npad * Kpad * 2};
    uint32_t box[4] = {64, (uint32_t)(two ? p.BN / 2 : p.BN), p.kblocks == 3 ? 3u : 1u, p.kblocks == 3 ? 1u : 3u};
    HB_TRY(tc::make_tmap_bf16(&tmB, w, 4, dims, str, box, 128));
  } else {
    uint64_t dims[2] = {(uint64_t)Kpad, (uint64_t)9 * Npad};
    uint64_t str[1] = {(uint64_t)Kpad * 2};
    uint32_t box[2] = {64, (uint32_t)(two ? p.BN / 2 : p.BN)};
    HB_TRY(tc::make_tmap_bf16(&tmB, w, 2, dims, str, box, 128));
  }
  tmR = tmA; tmOF = tmA; tmOT = tmA; tmLN = tmA;      // placeholders for absent operands (never dereferenced)
  const uint32_t w32 = grouped ? 32 : 16, w16 = grouped ? 64 : 16;
  auto map4 = [&](CUtensorMap* tm, bool f32, const void* base, int width, int pitch, uint32_t bw, uint32_t bx, uint32_t by) -> int {
    const int es = f32 ? 4 : 2;
    uint64_t dims[4] = {(uint64_t)width, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.B};
    uint64_t str[3] = {(uint64_t)pitch * es, (uint64_t)a.W * pitch * es, (uint64_t)a.H * a.W * pitch * es};
    uint32_t box[4] = {bw, bx, by, 1};
    return f32 ? tc::make_tmap_f32(tm, base, 4, dims, str, box, (int)bw * es) : tc::make_tmap_bf16(tm, base, 4, dims, str, box, (int)bw * es);
  };
  if (p.has_resid) HB_TRY(map4(&tmR, true, ep.resid, p.N, ep.resid_pitch, CW, C3_TW, C3_TH));
  if (ep.out_f32) HB_TRY(map4(&tmOF, true, ep.out_f32, p.N, ep.out_f32_pitch, w32, C3_TW, 4));
  if (ep.out_t && p.shuf_r > 1) {     // (B, H*r, W*r, 64) output as {c, j, x, i, b*H + y}: one (i, j) sub-pixel per store
    const uint64_t r = (uint64_t)p.shuf_r, W = (uint64_t)a.W;
    uint64_t dims[5] = {64, r, W, r, (uint64_t)a.B * a.H};
    uint64_t str[4] = {64 * 2, r * 64 * 2, W * r * 64 * 2, r * W * r * 64 * 2};
    uint32_t box[5] = {64, 1, C3_TW, 1, 4};
    HB_TRY(tc::make_tmap_bf16(&tmOT, ep.out_t, 5, dims, str, box, 128));
  } else if (ep.out_t) {
    const int width = ep.out_t_zero_to > p.N ? ep.out_t_zero_to : p.N;
    HB_TRY(map4(&tmOT, false, (const bf16*)ep.out_t + ep.out_t_coff, width, ep.out_t_pitch, w16, C3_TW, 4));
  }
  if (ep.ln_out) HB_TRY(map4(&tmLN, false, ep.ln_out, ep.ln_pitch, ep.ln_pitch, w16, C3_TW, 4));
  p.rev = ep.walk_back ? 1 : 0;
  p.trace = gemm_trace_arm(st);
  p.rot = getenv("HATB200_C3_ROT") ? atoi(getenv("HATB200_C3_ROT")) : 1;
  if (p.lean) p.rot = 0;         // the unrolled issuer walks the taps in order
  p.probe = getenv("HATB200_C3_PROBE") ? atoi(getenv("HATB200_C3_PROBE")) : 0;
  const size_t smem = (size_t)p.smem_bytes + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    const int lim = C3_SMEM_LIMIT + 1024;
#define C3_ATTR1(F_, G_, T_) \
    HB_CHECK_CUDA(cudaFuncSetAttribute(conv3_tc_kernel<F_, G_, T_>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim))
#define C3_ATTR(F_) \
    C3_ATTR1(F_, false, false); C3_ATTR1(F_, true, false); C3_ATTR1(F_, false, true); C3_ATTR1(F_, true, true)
    C3_ATTR(FeatAll); C3_ATTR(FeatPlain); C3_ATTR(FeatColsum); C3_ATTR(FeatResLn); C3_ATTR(FeatLast);
#undef C3_ATTR
#undef C3_ATTR1
    attr_set = true;
  }
  const int n_tiles = Npad / p.BN;
  int per_n = c3_sm_count() / n_tiles;
  if (per_n < 1) per_n = 1;
  if (per_n > p.tiles) per_n = p.tiles;
  if (two) per_n &= ~1;
  dim3 grid(per_n, n_tiles);
  p.main_ctas = per_n;
  if (ep.side.kind) {        // extra CTAs (whole clusters) for the side job: one per sample
    HB_REQUIRE(n_tiles == 1, HATB200_EUNSUPPORTED, "conv3_tc: a side job needs a single n-tile");
    grid.x = (unsigned)(per_n + round_up(ep.side.B, 2));
  }
#define C3_LAUNCH(F_)                                                                                              \
  do {                                                                                                             \
    if (two) {                                                                                                     \
      if (grouped) launch_pdl_cluster(conv3_tc_kernel<F_, true, true>, 2, grid, dim3(C3_THREADS), smem, st, tmA, tmB, tmR, tmOF, tmOT, tmLN, p, ep);  \
      else launch_pdl_cluster(conv3_tc_kernel<F_, false, true>, 2, grid, dim3(C3_THREADS), smem, st, tmA, tmB, tmR, tmOF, tmOT, tmLN, p, ep);        \
    } else {                                                                                                       \
      if (grouped) launch_pdl(conv3_tc_kernel<F_, true, false>, grid, dim3(C3_THREADS), smem, st, tmA, tmB, tmR, tmOF, tmOT, tmLN, p, ep);  \
      else launch_pdl(conv3_tc_kernel<F_, false, false>, grid, dim3(C3_THREADS), smem, st, tmA, tmB, tmR, tmOF, tmOT, tmLN, p, ep);        \
    }                                                                                                              \
  } while (0)
  static const bool generic_only = getenv("HATB200_EPI_GENERIC") != nullptr;
  switch (generic_only ? FEAT_ALL : feat_select(ep)) {
    case FEAT_PLAIN: C3_LAUNCH(FeatPlain); break;
    case FEAT_COLSUM: C3_LAUNCH(FeatColsum); break;
    case FEAT_RESLN: C3_LAUNCH(FeatResLn); break;
    case FEAT_LAST: C3_LAUNCH(FeatLast); break;
    default: C3_LAUNCH(FeatAll); break;
  }
#undef C3_LAUNCH
  if (two && cudaPeekAtLastError() != cudaSuccess) {     // pairs cannot be scheduled here: retry as single CTAs
    cudaGetLastError();
    g_c3_pairs_failed = true;
    return launch_conv3_tc(a, w, Npad, Kpad, ep, st);
  }
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb
