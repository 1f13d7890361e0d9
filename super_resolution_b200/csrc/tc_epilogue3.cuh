// tc_epilogue3.cuh -- the warp-independent, TMA-fed and TMA-drained fused epilogue shared by the persistent tcgen05
// kernels (tc_gemm.cu: Linear / 1x1; tc_conv3.cu: halo-resident 3x3 conv).
//
//   v = acc + bias; v *= q-scale; v = act(v); v += conv_scale * eca_scale[b,n] * U[m,n]; v += resid[m,n]
//   outputs: fp32 residual stream, bf16 activation, per-tile column sums (global-average-pool partials),
//   LayerNorm(v) in bf16 (+ compact 16-channel head, its x-padded copy and its per-tile column sums)
//
// One epilogue thread owns ONE row of the 128-row accumulator tile (TMEM lane = row) and walks its columns in
// chunks of CW = 16 (tcgen05.ld 32x32b.x16, software pipelined).  Every epilogue WARP is independent: global
// operands (fp32 residual, bf16 CAB branch) arrive through a TMA ring of 128-row x CW-column chunks filled by a
// dedicated producer warp and are released as soon as the four warps of the warpgroup have pulled their rows into
// registers; every output leaves through TMA from a private, double-buffered, swizzled staging tile (Stager).
// Tensor maps clip the ragged edge (N = 180 of 192), TMA loads zero-fill it, and bias / gamma / beta / ECA tables
// are zero padded, so the epilogue has no per-column predicates.  fp32 pair math uses FFMA2 / FADD2 / FMUL2.
// How a tile's rows map to memory is a policy (Tile2D: 128 consecutive tokens; Tile4D: 8 x 16 pixels of a sample).
#pragma once
#include "kernels.cuh"
#include "tc_common.cuh"

namespace hb {
namespace epi3 {

using namespace tc;

constexpr int CW = 16;                 // epilogue column chunk (one tcgen05.ld 32x32b.x16 per thread)
constexpr int MAX_OPS = 4;             // operand ring stages per warpgroup

__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, const void* src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(m),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* m, const void* src, int c0, int c1, int c2, int c3, int c4) {
  asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::"l"(m),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
               : "memory");
}
// ---- small PTX helpers -------------------------------------------------------------------------------
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(m),
               "r"(smem_u32(src)), "r"(c0), "r"(c1)
               : "memory");
}
// L2 eviction-priority policy for the fp32 residual stream (47 MB at the bench shape, read and rewritten in place by two
// launches of every block): evict_last keeps it resident across the ~600 MB of other traffic per block.
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void tma_store_2d_hint(const CUtensorMap* m, const void* src, int c0, int c1, uint64_t pol) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group.L2::cache_hint [%0, {%2, %3}], [%1], %4;" ::"l"(m),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], "
      "[%2], %5;" ::"r"(smem_u32(dst)),
      "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long pk(float a, float b) {
  return (unsigned long long)__float_as_uint(a) | ((unsigned long long)__float_as_uint(b) << 32);
}
__device__ __forceinline__ void upk(unsigned long long v, float& a, float& b) {
  a = __uint_as_float((uint32_t)v); b = __uint_as_float((uint32_t)(v >> 32));
}
__device__ __forceinline__ void fma2(float& d0, float& d1, float a0, float a1, float b0, float b1, float c0, float c1) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pk(a0, a1)), "l"(pk(b0, b1)), "l"(pk(c0, c1)));
  upk(r, d0, d1);
}
__device__ __forceinline__ void add2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a0, a1)), "l"(pk(b0, b1)));
  upk(r, d0, d1);
}
__device__ __forceinline__ void mul2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  unsigned long long r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a0, a1)), "l"(pk(b0, b1)));
  upk(r, d0, d1);
}


// Exact (erf) GELU of two values, nn.GELU() of the reference (hat_arch.py:85, :114), for the bf16 tensor-core path:
//   gelu(v) = 0.5 v (1 + erf(v / sqrt 2)) = max(v, 0) - 0.5 |v| P(t) exp(-v^2 / 2),   t = 1 / (1 + p |v| / sqrt 2)
// with Abramowitz-Stegun 7.1.26 for erfc (|error of erf| <= 1.5e-7; of this gelu <= 2.2e-7 over [-12, 12] in float64, and the
// result is rounded to bf16 next).  19 issue slots per PAIR (packed FFMA2 / FMUL2, 4 MUFU) against ~80 for two erff calls
// whose two argument ranges diverge inside a warp: CAB conv1 and the OCAB MLP were bound by this one line
// (6.46 M warp instructions per conv1 launch, 80 % of them erff).
__device__ __forceinline__ void gelu2(float& v0, float& v1) {
  const float a0 = fabsf(v0), a1 = fabsf(v1);
  float d0, d1, q0, q1, s0, s1, w0, w1;
  fma2(d0, d1, a0, a1, 0.23164189f, 0.23164189f, 1.f, 1.f);
  float t0, t1;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(d0));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(d1));
  fma2(q0, q1, t0, t1, 1.061405429f, 1.061405429f, -1.453152027f, -1.453152027f);
  fma2(q0, q1, q0, q1, t0, t1, 1.421413741f, 1.421413741f);
  fma2(q0, q1, q0, q1, t0, t1, -0.284496736f, -0.284496736f);
  fma2(q0, q1, q0, q1, t0, t1, 0.254829592f, 0.254829592f);
  mul2(q0, q1, q0, q1, t0, t1);
  mul2(s0, s1, v0, v1, v0, v1);
  mul2(s0, s1, s0, s1, -0.72134752f, -0.72134752f);            // -0.5 log2(e)
  mul2(w0, w1, a0, a1, ex2_approx(s0), ex2_approx(s1));
  mul2(w0, w1, w0, w1, q0, q1);
  fma2(v0, v1, w0, w1, -0.5f, -0.5f, fmaxf(v0, 0.f), fmaxf(v1, 0.f));
}

// swizzle of 16-byte granule index for a tile whose rows are RB bytes wide (TMA SWIZZLE_{32,64,128}B, tile base
// aligned to the swizzle period)
template <int RB>
__device__ __forceinline__ int sw_gran(int row) {
  return RB == 128 ? (row & 7) : (RB == 64 ? ((row >> 1) & 3) : ((row >> 2) & 1));
}
// this thread's row (CW fp32) of a swizzled tile
__device__ __forceinline__ void row_read_f32(const uint8_t* tile, int row, float (&o)[CW]) {
  constexpr int RB = CW * 4;
  const int sw = sw_gran<RB>(row);
  const uint32_t base = smem_u32(tile) + row * RB;
#pragma unroll
  for (int c = 0; c < RB / 16; c++) {
    const uint4 t = lds128(base + ((c ^ sw) << 4));
    o[4 * c] = __uint_as_float(t.x); o[4 * c + 1] = __uint_as_float(t.y);
    o[4 * c + 2] = __uint_as_float(t.z); o[4 * c + 3] = __uint_as_float(t.w);
  }
}
__device__ __forceinline__ void row_read_bf16(const uint8_t* tile, int row, float (&o)[CW]) {
  constexpr int RB = CW * 2;
  const int sw = sw_gran<RB>(row);
  const uint32_t base = smem_u32(tile) + row * RB;
#pragma unroll
  for (int c = 0; c < RB / 16; c++) {
    const uint4 t = lds128(base + ((c ^ sw) << 4));
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {
      o[8 * c + 2 * j] = __uint_as_float(w[j] << 16);
      o[8 * c + 2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
    }
  }
}
// CW consecutive floats of a shared-memory table (same address for the whole warp: broadcast)
__device__ __forceinline__ void tab_read(const float* tab, float (&o)[CW]) {
#pragma unroll
  for (int c = 0; c < CW / 4; c++) {
    const uint4 t = lds128(smem_u32(tab) + 16 * c);
    o[4 * c] = __uint_as_float(t.x); o[4 * c + 1] = __uint_as_float(t.y);
    o[4 * c + 2] = __uint_as_float(t.z); o[4 * c + 3] = __uint_as_float(t.w);
  }
}

// Per-warp output staging.  GW consecutive chunks (CW columns each, EB bytes per element) of the warp's 32 rows are
// collected in one swizzled tile whose rows are RB = CW * EB * GW bytes (128 B when grouped) and leave with ONE TMA
// store; two buffers alternate so that the store of group q overlaps the arithmetic of group q + 1.
template <int EB, int GW>
struct Stager {
  static constexpr int RB = CW * EB * GW;
  static constexpr int BUF = 32 * RB;
  uint8_t* base;
  uint32_t n = 0;                    // stores issued by this warp
  uint64_t policy = 0;               // != 0: L2 cache-hint policy of the stores (Tile2D only)
  int nbuf = 2;                      // 2: double buffered; 1: single buffer (shared memory is short)
  __device__ __forceinline__ uint8_t* buf() const { return base + (nbuf == 2 ? (n & 1) : 0) * BUF; }
  // before the first write of a group: the store that last read this buffer must be done reading it
  __device__ __forceinline__ void begin(int lane) const {
    if (elect_one()) { if (nbuf == 2) bulk_wait_read<1>(); else bulk_wait_read<0>(); }
    __syncwarp();
  }
  __device__ __forceinline__ void put_f32(int lane, int slot, const float (&v)[CW]) const {
    const uint32_t row = smem_u32(buf()) + lane * RB;
    const int sw = sw_gran<RB>(lane);
#pragma unroll
    for (int c = 0; c < 4; c++)
      sts128(row + (((slot * 4 + c) ^ sw) << 4), __float_as_uint(v[4 * c]), __float_as_uint(v[4 * c + 1]),
             __float_as_uint(v[4 * c + 2]), __float_as_uint(v[4 * c + 3]));
  }
  __device__ __forceinline__ void put_bf16(int lane, int slot, const float (&v)[CW]) const {
    const uint32_t row = smem_u32(buf()) + lane * RB;
    const int sw = sw_gran<RB>(lane);
#pragma unroll
    for (int c = 0; c < 2; c++) {
      uint4 t;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
      for (int j = 0; j < 4; j++) h[j] = __floats2bfloat162_rn(v[8 * c + 2 * j], v[8 * c + 2 * j + 1]);
      sts128(row + (((slot * 2 + c) ^ sw) << 4), t.x, t.y, t.z, t.w);
    }
  }
  __device__ __forceinline__ void put_zero(int lane) const {
    const uint32_t row = smem_u32(buf()) + lane * RB;
#pragma unroll
    for (int c = 0; c < RB / 16; c++) sts128(row + 16 * c, 0u, 0u, 0u, 0u);
  }
  // after the last write of a group: hand the tile to the async proxy; columns past the map's width are clipped
  template <class TS>
  __device__ __forceinline__ void flush(const CUtensorMap* tm, int lane, int c0, const TS& ts, int quarter) {
    fence_proxy_async();
    __syncwarp();
    if (elect_one()) {               // same leader every time (full mask): bulk groups are per-thread state
      ts.store(tm, buf(), c0, quarter, policy);
      bulk_commit();
    }
    n++;
  }
};


// ---- tile policies: how the 128 rows of an accumulator tile map to memory --------------------------------
// Tile2D: rows are 128 consecutive tokens m0 .. m0+127; store maps are 2-D (col, token), box {w, 32}.
struct Tile2D {
  int m0;
  __device__ __forceinline__ void store(const CUtensorMap* tm, const void* src, int col, int quarter, uint64_t pol = 0) const {
    if (pol) tma_store_2d_hint(tm, src, col, m0 + quarter * 32, pol);
    else tma_store_2d(tm, src, col, m0 + quarter * 32);
  }
  __device__ __forceinline__ void load(void* dst, const CUtensorMap* tm, uint64_t* bar, int col, uint64_t pol = 0) const {
    if (pol) tma_load_2d_hint(dst, tm, bar, col, m0, pol);
    else tma_load_2d(dst, tm, bar, col, m0);
  }
  __device__ __forceinline__ void prefetch(const CUtensorMap* tm, int col) const { tma_prefetch_2d(tm, col, m0); }
};
// Tile4D: rows are the 8 (x) x 16 (y) pixels of one sample starting at (x0, y0), row r = (y0 + r / 8, x0 + r % 8);
// maps are 4-D (col, x, y, b): store box {w, 8, 4, 1} (one warp = 4 image rows), load box {CW, 8, 16, 1}.
struct Tile4D {
  int x0, y0, b;
  int r = 1, H = 0;      // r > 1: conv + PixelShuffle(r) with (i, j, c)-ordered columns: the 64-column block starting at
                         // `col` is sub-pixel (i, j) = (col/64 / r, col/64 % r) of every pixel of the tile; the store map
                         // is 5-D {c, j, x, i, b*H + y} over the (B, H*r, W*r, 64) output
  __device__ __forceinline__ void store(const CUtensorMap* tm, const void* src, int col, int quarter, uint64_t = 0) const {
    if (r > 1) {
      const int blk = col >> 6, i = blk / r, j = blk - i * r;
      tma_store_5d(tm, src, 0, j, x0, i, b * H + y0 + quarter * 4);
    } else {
      tma_store_4d(tm, src, col, x0, y0 + quarter * 4, b);
    }
  }
  __device__ __forceinline__ void load(void* dst, const CUtensorMap* tm, uint64_t* bar, int col, uint64_t = 0) const {
    tma_load_4d(dst, tm, bar, col, x0, y0, b);
  }
  __device__ __forceinline__ void prefetch(const CUtensorMap* tm, int col) const { tma_prefetch_4d(tm, col, x0, y0, b); }
};

// Column sums of a 32-row x 16-column register block: after five butterfly rounds lane l (l even) holds the sum
// over the warp's 32 rows of column ((l>>4)&1)*8 + ((l>>3)&1)*4 + ((l>>2)&1)*2 + ((l>>1)&1).  16 shuffles
// instead of 80 for sixteen independent warp reductions; the order of additions is fixed (deterministic).
__device__ __forceinline__ float colsum16(const float (&v)[CW], int lane) {
  float a[8], b4[4], c2[2];
  const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const float send = h16 ? v[j] : v[j + 8];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 16);
    a[j] = (h16 ? v[j + 8] : v[j]) + recv;
  }
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const float send = h8 ? a[j] : a[j + 4];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 8);
    b4[j] = (h8 ? a[j + 4] : a[j]) + recv;
  }
#pragma unroll
  for (int j = 0; j < 2; j++) {
    const float send = h4 ? b4[j] : b4[j + 2];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 4);
    c2[j] = (h4 ? b4[j + 2] : b4[j]) + recv;
  }
  const float send = h2 ? c2[0] : c2[1];
  const float recv = __shfl_xor_sync(0xffffffffu, send, 2);
  float d = (h2 ? c2[1] : c2[0]) + recv;
  d += __shfl_xor_sync(0xffffffffu, d, 1);
  return d;
}
__device__ __forceinline__ int colsum16_col(int lane) {
  return ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
}

// Everything an epilogue warp keeps across tiles.
template <bool GROUPED>
struct WarpState {
  static constexpr int GW32 = GROUPED ? 2 : 1, GW16 = GROUPED ? 4 : 1;   // chunks per store: 128-byte rows when grouped
  Stager<4, GW32> st32;
  Stager<2, GW16> st16;
  const float *tab_bias, *tab_gamma, *tab_beta;   // shared-memory tables [BN], zero beyond N
  float* coef;                                    // this warpgroup's ECA coefficient table [BN]
  const uint8_t* ring;                            // this warpgroup's operand ring
  uint64_t *op_full, *op_empty;                   // [ops] barriers of that ring
  int ops, op_stage_bytes, op_u_off, has_resid, has_u, share_stg;
  float (*hsum)[16];                              // [8][16]: per epilogue warp column sums of the LN head
  float (*sred)[256];                             // [4][256]: this warpgroup's scratch for tile column sums
  RingPos k;                                      // operand-ring cursor of this warpgroup (consumer side)
  int g, ew, quarter, lane, wg_tid, row;
  int BN, N, n0, nch, H, W, hw;
  // N-split CTA pair (tc_gemm.cu, SPLIT): this CTA owns columns [n0, n0 + BN) of a LayerNorm row, its cluster peer the
  // rest; the fp32 store map is based at column n0 (so_f32 = 0 then), row statistics are exchanged through xch
  int so_f32 = 0;                                 // column origin of fp32 stores in tmOF coordinates (n0 unless split)
  float2* xch = nullptr;                          // [2 parities][2 stages][128 rows]: (mean, M2) written by the PEER
  uint64_t* xch_bar = nullptr;                    // [2 stages]: 128 remote arrivals per tile
  uint32_t peer = 0;                              // cluster rank of the peer
  uint32_t xch_n = 0;                             // exchanges done by this warp (parity / buffer)
  bool last_ntile;
  bool prof = false;                              // engineering: accumulate wait cycles (HATB200_GEMM_TRACE)
  long long t_ring = 0, t_stg = 0, t_p1 = 0, t_p2 = 0;
};

// Per-tile facts of the calling thread's row.
struct RowInfo {
  uint32_t trow;       // TMEM address: this warp's lane quarter, column 0 of the tile's accumulator stage
  int b;               // sample
  int m;               // token index of this thread's row
  int py, px;          // pixel of this thread's row
  int chunk, nchunk;   // index of this tile inside its sample / tiles per sample (partial-sum slots)
};

// Compile-time feature sets.  A kernel instantiated on F contains code only for the features F enables; with
// EXACT the structural ones (U, RESID, OUT32, OUT16, LN, COLSUM) are known PRESENT as well, so the chunk loop has
// no feature branches at all (the generic body costs 266 instructions per 16-column chunk, 40 of them arithmetic,
// and the epilogue warps -- two per scheduler -- are issue/latency bound: profiles/r01_ncu_fc2_*).  ACT, SCALE and
// HEAD stay run-time inside a set that enables them.  The host picks the first matching set (feat_match).
struct FeatAll    { static constexpr bool EXACT = 0, ACT = 1, SCALE = 1, U = 1, RESID = 1, OUT32 = 1, OUT16 = 1, LN = 1, HEAD = 1, COLSUM = 1, LAST = 0; };
struct FeatPlain  { static constexpr bool EXACT = 1, ACT = 1, SCALE = 1, U = 0, RESID = 0, OUT32 = 0, OUT16 = 1, LN = 0, HEAD = 0, COLSUM = 0, LAST = 0; };
struct FeatColsum { static constexpr bool EXACT = 1, ACT = 0, SCALE = 0, U = 0, RESID = 0, OUT32 = 0, OUT16 = 1, LN = 0, HEAD = 0, COLSUM = 1, LAST = 0; };
struct FeatResLn  { static constexpr bool EXACT = 1, ACT = 0, SCALE = 0, U = 0, RESID = 1, OUT32 = 1, OUT16 = 0, LN = 1, HEAD = 1, COLSUM = 0, LAST = 0; };
struct FeatResULn { static constexpr bool EXACT = 1, ACT = 0, SCALE = 0, U = 1, RESID = 1, OUT32 = 1, OUT16 = 0, LN = 1, HEAD = 0, COLSUM = 0, LAST = 0; };
// conv_last: no generic output at all, the kernel's own few lines write the image (Epilogue::nchw_y)
struct FeatLast   { static constexpr bool EXACT = 1, ACT = 0, SCALE = 0, U = 0, RESID = 0, OUT32 = 0, OUT16 = 0, LN = 0, HEAD = 0, COLSUM = 0, LAST = 1; };
template <class F> constexpr bool feat_ring() { return F::U || F::RESID; }

// does the feature set F cover this epilogue exactly (EXACT) / at least (generic)?
template <class F>
inline bool feat_match(const Epilogue& ep) {
  const bool u = ep.eca_u != nullptr, r = ep.resid != nullptr, o32 = ep.out_f32 != nullptr, o16 = ep.out_t != nullptr;
  const bool ln = ep.ln_out != nullptr, cs = ep.colsum != nullptr;
  if (ep.act != ACT_NONE && !F::ACT) return false;
  if (ep.scale_ncols > 0 && !F::SCALE) return false;
  if (ep.ln_head && !F::HEAD) return false;
  if (F::EXACT) return u == F::U && r == F::RESID && o32 == F::OUT32 && o16 == F::OUT16 && ln == F::LN && cs == F::COLSUM;
  return (!u || F::U) && (!r || F::RESID) && (!o32 || F::OUT32) && (!o16 || F::OUT16) && (!ln || F::LN) && (!cs || F::COLSUM);
}
enum { FEAT_ALL = 0, FEAT_PLAIN, FEAT_COLSUM, FEAT_RESLN, FEAT_RESULN, FEAT_LAST };
inline int feat_select(const Epilogue& ep) {
  if (ep.nchw_y) return FEAT_LAST;
  if (feat_match<FeatPlain>(ep)) return FEAT_PLAIN;
  if (feat_match<FeatColsum>(ep)) return FEAT_COLSUM;
  if (feat_match<FeatResLn>(ep)) return FEAT_RESLN;
  if (feat_match<FeatResULn>(ep)) return FEAT_RESULN;
  return FEAT_ALL;
}

// The whole epilogue of one tile for the calling warp.  F: feature set; TS: Tile2D / Tile4D.
// PRE: the accumulator already contains the fp32 residual (tc_gemm.cu preloads it with tcgen05.cp), the epilogue skips it.
// SPLIT: the row is shared with the cluster peer (see WarpState); LayerNorm statistics are merged across the pair.
template <class F, bool GROUPED, class TS, bool PRE = false, bool SPLIT = false>
__device__ __forceinline__ void run_tile(const Epilogue& ep, WarpState<GROUPED>& ws, const RowInfo& ri, const TS& ts,
                                         const CUtensorMap* tmOF, const CUtensorMap* tmOT, const CUtensorMap* tmLN) {
  constexpr int GW32 = WarpState<GROUPED>::GW32, GW16 = WarpState<GROUPED>::GW16;
  constexpr bool RING = feat_ring<F>();
  const int lane = ws.lane, quarter = ws.quarter, row = ws.row, g = ws.g, n0 = ws.n0, nch = ws.nch;
  // feature present?  compile-time for an EXACT set, a uniform run-time test otherwise
  const bool has_u = F::U && (F::EXACT || ws.has_u), has_resid = !PRE && F::RESID && (F::EXACT || ws.has_resid);
  const bool do_ln = F::LN && (F::EXACT || ep.ln_out != nullptr);
  const bool out32 = F::OUT32 && (F::EXACT || ep.out_f32 != nullptr), out16 = F::OUT16 && (F::EXACT || ep.out_t != nullptr);
  const bool do_cs = F::COLSUM && (F::EXACT || ep.colsum != nullptr);
  const int act = F::ACT ? ep.act : ACT_NONE;
  const int scale_ncols = F::SCALE ? ep.scale_ncols : 0;
  const float out_scale = ep.out_scale;
  const uint32_t trow = ri.trow;
  if (has_u) {
    // per-sample ECA coefficients of this tile (hat_arch.py:236: conv_scale * sigmoid gate), zero beyond N
    asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");          // previous tile's readers are done
    for (int j = ws.wg_tid; j < ws.BN; j += 128) {
      const int n = n0 + j;
      ws.coef[j] = n < ws.N ? ep.eca_coef * ep.eca_scale[(size_t)ri.b * ws.N + n] : 0.f;
    }
    asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
  }
  float ln_k = 0.f, s1a = 0.f, s1b = 0.f, s2a = 0.f, s2b = 0.f;
  if (ws.share_stg) { if (elect_one()) bulk_wait_read<0>(); __syncwarp(); }     // previous tile's bf16 stores have left it
  long long tp = 0;
  if (ws.prof) tp = clock64();
  // ---------------- pass 1: value, fp32 / bf16 outputs, column sums, LayerNorm statistics ----------------
  // Software pipeline over two register sets: while chunk ci is processed out of one, chunk ci + 1 loads into
  // the other (no register copies between iterations).
  auto pass1 = [&](const int ci, uint32_t (&rc)[CW], uint32_t (&rn)[CW]) __attribute__((always_inline)) {
    const int nb = ci * CW;                       // column inside the n-tile
    float v[CW], tb[CW];
    tab_read(ws.tab_bias + nb, tb);
    float u[CW], rs[CW], cf[CW];
    int rel_s = -1;
    if (RING && (has_u || has_resid)) {
      const int s = ws.k.s;
      long long tw = 0;
      if (ws.prof) tw = clock64();
      mbar_wait(&ws.op_full[s], ws.k.ph);
      if (ws.prof) ws.t_ring += clock64() - tw;
      const uint8_t* st = ws.ring + (size_t)s * ws.op_stage_bytes;
      if (has_u) { row_read_bf16(st + ws.op_u_off, row, u); tab_read(ws.coef + nb, cf); }
      if (has_resid) row_read_f32(st, row, rs);
      rel_s = s;
      ws.k.next(ws.ops);
    }
    tmem_ld_wait();
    if (ci + 1 < nch) tmem_ld16(trow + (uint32_t)(nb + CW), rn);
#pragma unroll
    for (int j = 0; j < CW; j += 2)
      add2(v[j], v[j + 1], __uint_as_float(rc[j]), __uint_as_float(rc[j + 1]), tb[j], tb[j + 1]);
    if (F::SCALE && n0 + nb < scale_ncols) {      // q pre-scaling (hat_arch.py:375); scale_ncols % CW == 0
#pragma unroll
      for (int j = 0; j < CW; j += 2) mul2(v[j], v[j + 1], v[j], v[j + 1], out_scale, out_scale);
    }
    if (F::ACT && act == ACT_GELU) {
#pragma unroll
      for (int j = 0; j < CW; j += 2) gelu2(v[j], v[j + 1]);
    } else if (F::ACT && act == ACT_LRELU) {
#pragma unroll
      for (int j = 0; j < CW; j++) v[j] = lrelu_f(v[j]);
    }
    if (has_u) {
#pragma unroll
      for (int j = 0; j < CW; j += 2) fma2(v[j], v[j + 1], cf[j], cf[j + 1], u[j], u[j + 1], v[j], v[j + 1]);
    }
    if (has_resid) {
#pragma unroll
      for (int j = 0; j < CW; j += 2) add2(v[j], v[j + 1], v[j], v[j + 1], rs[j], rs[j + 1]);
    }
    if (RING && rel_s >= 0) {
      // Release the ring stage only AFTER its values have been consumed: the shared-memory loads above are
      // asynchronous, and an arrive issued right behind them let the producer's next TMA box land in the stage while
      // loads of the slower warps were still outstanding (B200, M = 65536: ~0.3 % of the rows, always TMEM lanes
      // 0-63 = the two warps that share their schedulers with the TMA / MMA warps, picked up a later chunk's
      // residual -- errors of +-5 in the fp32 stream that the small-M parity tests never reached).
      __syncwarp();
      if (lane == 0) mbar_arrive(&ws.op_empty[rel_s]);
    }
    if (do_ln) {
      if (ci == 0) ln_k = v[0];
#pragma unroll
      for (int j = 0; j < CW; j += 2) {
        float d0, d1;
        add2(d0, d1, v[j], v[j + 1], -ln_k, -ln_k);
        add2(s1a, s1b, s1a, s1b, d0, d1);
        fma2(s2a, s2b, d0, d1, d0, d1, s2a, s2b);
      }
      uint32_t w[CW];
#pragma unroll
      for (int j = 0; j < CW; j++) w[j] = __float_as_uint(v[j]);
      tmem_st16(trow + (uint32_t)nb, w);          // fp32 row parked in TMEM for the normalise pass
    }
    if (do_cs) {                                  // per-tile column sums of v (global-average-pool partials)
      const float cs = colsum16(v, lane);
      if ((lane & 1) == 0) ws.sred[quarter][nb + colsum16_col(lane)] = cs;
    }
    if (out32) {
      const int slot = ci % GW32;
      long long tw = 0;
      if (ws.prof) tw = clock64();
      if (slot == 0) ws.st32.begin(lane);
      if (ws.prof) ws.t_stg += clock64() - tw;
      ws.st32.put_f32(lane, slot, v);
      if (slot == GW32 - 1 || ci == nch - 1) ws.st32.flush(tmOF, lane, (SPLIT ? ws.so_f32 : n0) + (ci - slot) * CW, ts, quarter);
    }
    if (out16) {
      const int slot = ci % GW16;
      if (slot == 0) ws.st16.begin(lane);
      ws.st16.put_bf16(lane, slot, v);
      if (slot == GW16 - 1 || ci == nch - 1) ws.st16.flush(tmOT, lane, n0 + (ci - slot) * CW, ts, quarter);
    }
  };
  uint32_t ra[CW], rb[CW];
  tmem_ld16(trow, ra);
  for (int ci = 0; ci < nch; ci += 2) {
    pass1(ci, ra, rb);
    if (ci + 1 < nch) pass1(ci + 1, rb, ra);
  }
  // K-padding columns of the consumer that no n-tile covers: [Npad, out_t_zero_to), written by the last n-tile
  if (out16 && ws.last_ntile) {
    for (int nb = ws.BN; n0 + nb < ep.out_t_zero_to; nb += CW * GW16) {
      ws.st16.begin(lane);
      ws.st16.put_zero(lane);
      ws.st16.flush(tmOT, lane, n0 + nb, ts, quarter);
    }
  }
  if (do_cs) {
    asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
    for (int n = ws.wg_tid; n < ws.BN; n += 128)
      if (n0 + n < ws.N)
        ep.colsum[((size_t)ri.b * ri.nchunk + ri.chunk) * ws.N + n0 + n] =
            ws.sred[0][n] + ws.sred[1][n] + ws.sred[2][n] + ws.sred[3][n];
    asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");           // sred is reused by the next tile
  }
  if (ws.prof) { const long long t = clock64(); ws.t_p1 += t - tp; tp = t; }
  // ---------------- pass 2: LayerNorm of the row (eps 1e-5), bf16 output, 16-channel head ----------------
  if (do_ln) {
    if (ws.share_stg) { if (elect_one()) bulk_wait_read<0>(); __syncwarp(); }   // fp32 stores have left the shared staging
    tmem_st_wait();
    // pad columns hold exact zeros: remove their contribution to the shifted sums
    const int nmine = SPLIT ? max(0, min(ws.BN, ws.N - n0)) : ws.N;       // valid columns of this CTA
    const float npad = (float)(ws.BN - nmine);
    const float s1 = s1a + s1b + npad * ln_k;
    const float s2 = s2a + s2b - npad * ln_k * ln_k;
    float rstd, nmr;
    if (SPLIT) {
      // (mean, M2) of my columns -> the peer's shared memory, its barrier; mine arrive the same way; merge (Chan et al.)
      const float fm = (float)nmine, fp = (float)(ws.N - nmine);
      const float mean_m = ln_k + s1 / fm;
      const float m2_m = fmaxf(s2 - s1 * s1 / fm, 0.f);
      const uint32_t slot = ((ws.xch_n & 1u) * 2u + (uint32_t)g) * 128u + (uint32_t)row;
      const uint32_t remote = mapa_u32(ws.xch + slot, ws.peer);
      asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(remote), "f"(mean_m), "f"(m2_m) : "memory");
      mbar_arrive_cluster(mapa_u32(&ws.xch_bar[g], ws.peer));
      mbar_wait_cluster(&ws.xch_bar[g], ws.xch_n & 1u);
      const float2 pr = ws.xch[slot];
      ws.xch_n++;
      const float inv_n = 1.f / (float)ws.N;
      const float delta = pr.x - mean_m;
      const float mean = (fm * mean_m + fp * pr.x) * inv_n;
      const float m2 = m2_m + pr.y + delta * delta * fm * fp * inv_n;
      rstd = rsqrtf(m2 * inv_n + 1e-5f);
      nmr = -mean * rstd;
    } else {
      const float inv_n = 1.f / (float)ws.N;
      const float mean_d = s1 * inv_n;
      const float var = fmaxf(s2 * inv_n - mean_d * mean_d, 0.f);
      rstd = rsqrtf(var + 1e-5f);
      nmr = -(ln_k + mean_d) * rstd;    // (x - mean) * rstd = x * rstd + nmr
    }
    const bool head = F::HEAD && ep.ln_head != nullptr && (!SPLIT || n0 == 0);
    auto pass2 = [&](const int ci, uint32_t (&rc)[CW], uint32_t (&rn)[CW]) __attribute__((always_inline)) {
      const int nb = ci * CW;
      float ga[CW], be[CW], v[CW];
      tab_read(ws.tab_gamma + nb, ga);
      tab_read(ws.tab_beta + nb, be);
      tmem_ld_wait();
      if (ci + 1 < nch) tmem_ld16(trow + (uint32_t)(nb + CW), rn);
#pragma unroll
      for (int j = 0; j < CW; j += 2) {
        float t0, t1;
        fma2(t0, t1, __uint_as_float(rc[j]), __uint_as_float(rc[j + 1]), rstd, rstd, nmr, nmr);
        fma2(v[j], v[j + 1], t0, t1, ga[j], ga[j + 1], be[j], be[j + 1]);
      }
      const int slot = ci % GW16;
      if (slot == 0) ws.st16.begin(lane);
      ws.st16.put_bf16(lane, slot, v);
      if (slot == GW16 - 1 || ci == nch - 1) ws.st16.flush(tmLN, lane, (ci - slot) * CW, ts, quarter);
      if (ci == 0 && head) {          // compact / x-padded copies of the first channels + their column sums
        uint4 q[2];
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(q);
#pragma unroll
        for (int j = 0; j < 8; j++) h[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
        bf16* hd = (bf16*)ep.ln_head + (size_t)ri.m * ep.ln_head_c;
        if (ep.ln_head_c == 16) {
          reinterpret_cast<uint4*>(hd)[0] = q[0];
          reinterpret_cast<uint4*>(hd)[1] = q[1];
        } else {
#pragma unroll
          for (int j = 0; j < 16; j++)
            if (j < ep.ln_head_c) hd[j] = __float2bfloat16_rn(v[j]);
        }
        if (ep.ln_head_pad && ep.ln_head_c == 16) {
          bf16* hp = (bf16*)ep.ln_head_pad + (((size_t)ri.b * ws.H + ri.py) * (ws.W + 12) + ri.px + 6) * 16;
          reinterpret_cast<uint4*>(hp)[0] = q[0];
          reinterpret_cast<uint4*>(hp)[1] = q[1];
        }
        if (ep.ln_head_colsum) {      // sums of the bf16-rounded head values, fixed order
#pragma unroll
          for (int j = 0; j < 16; j++) {
            const float sj = warp_sum(__bfloat162float(__float2bfloat16_rn(v[j])));
            if (lane == 0) ws.hsum[ws.ew][j] = sj;
          }
        }
      }
    };
    tmem_ld16(trow, ra);
    pass2(0, ra, rb);                 // chunk 0 carries the head: kept out of the steady-state loop
    for (int ci = 1; ci < nch; ci += 2) {
      pass2(ci, rb, ra);
      if (ci + 1 < nch) pass2(ci + 1, ra, rb);
    }
    for (int nb = ws.BN; !SPLIT && nb < ep.ln_pitch; nb += CW * GW16) {   // pad columns beyond the (single) n-tile
      ws.st16.begin(lane);
      ws.st16.put_zero(lane);
      ws.st16.flush(tmLN, lane, nb, ts, quarter);
    }
    if (head && ep.ln_head_colsum) {
      asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
      if (ws.wg_tid < ep.ln_head_c && ws.wg_tid < 16) {
        // warps of this warpgroup in row order: quarter q is warp ((q + 2) & 3) of the group
        const int e0 = g * 4;
        ep.ln_head_colsum[((size_t)ri.b * ri.nchunk + ri.chunk) * ep.ln_head_c + ws.wg_tid] =
            ws.hsum[e0 + 2][ws.wg_tid] + ws.hsum[e0 + 3][ws.wg_tid] + ws.hsum[e0 + 0][ws.wg_tid] + ws.hsum[e0 + 1][ws.wg_tid];
      }
      asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");       // hsum is reused by the next tile
    }
  }
  if (ws.prof) ws.t_p2 += clock64() - tp;
}

// Operand-ring producer of one warpgroup for one tile (one elected thread): residual / CAB-branch chunks.
// (Measured and dropped: L2-prefetching the warpgroup's next tile, and everything of its first tile the ring
//  cannot hold yet -- 27 MB of prefetch at kernel start delays the first A/B stages by 3-4 us and ring depth
//  2 -> 4 moves a launch by < 5 %; the epilogue warps wait for accumulators, not for the ring.)
template <class TS>
__device__ __forceinline__ void produce_tile(const TS& ts, const CUtensorMap* tmR,
                                             const CUtensorMap* tmU, uint8_t* ring,
                                             uint64_t* op_full, uint64_t* op_empty, int ops, int op_stage_bytes,
                                             int op_u_off, int has_resid, int has_u, int n0, int nch, RingPos& k,
                                             const TS* next = nullptr, uint64_t resid_policy = 0) {
  const uint32_t bytes = (uint32_t)((has_resid ? 128 * CW * 4 : 0) + (has_u ? 128 * CW * 2 : 0));
  for (int ci = 0; ci < nch; ci++, k.next(ops)) {
    const int s = k.s;
    // one tile ahead, chunk by chunk: the warpgroup's NEXT tile's operands start towards L2 while this tile's chunks
    // go through the ring (a ring stage then waits for an L2 hit, not for HBM)
    if (next) {
      if (has_resid) next->prefetch(tmR, n0 + ci * CW);
      if (has_u) next->prefetch(tmU, n0 + ci * CW);
    }
    mbar_wait(&op_empty[s], k.ph ^ 1u);
    uint8_t* st = ring + (size_t)s * op_stage_bytes;
    mbar_arrive_expect_tx(&op_full[s], bytes);
    if (has_resid) ts.load(st, tmR, &op_full[s], n0 + ci * CW, resid_policy);
    if (has_u) ts.load(st + op_u_off, tmU, &op_full[s], n0 + ci * CW);
  }
}

}  // namespace epi3
}  // namespace hb
