// tc_epilogue2.cuh -- epilogue of the persistent GEMM (tc_gemm.cu), second generation.
//
// Same arithmetic and value order as tc_epilogue.cuh (one token row per thread, row-local LayerNorm), but
//   * the global operands of the epilogue -- the fp32 residual stream and the bf16 CAB branch U -- no longer
//     come from per-thread global loads (ncu: long-scoreboard stalls on a 128-thread warpgroup, 0.8 TB/s):
//     one elected thread of the warpgroup TMA-loads them two column chunks ahead into a swizzled 2-stage
//     shared-memory ring (chunk = 128 rows x CW columns), prefetching across tile boundaries, and every thread
//     reads its own row with conflict-free 16-byte shared loads;
//   * columns are walked CW = 32 (or 16) at a time with a predicate-free fast path for full chunks and row
//     base pointers hoisted out of the loop (v1 issued ~700 instructions per 16 columns).
//   * stores are coalesced: measured on B200, one thread per row storing 16 B at a 1.5 KB stride costs ~90 cycles
//     per warp instruction (32 different lines) and made a 128 x 240 tile take 5.5 us to drain; so each warp
//     parks its 32 rows x CW columns in shared memory (in place in the operand slab for the fp32 stream, a 2 KB
//     per-warp scratch for bf16) and writes them out with whole rows per quarter-warp (3x faster).
#pragma once
#include "tc_epilogue.cuh"

namespace hb {

struct OpRing {                 // per epilogue warpgroup
  uint8_t* stage[2];            // [resid chunk | U chunk]
  uint64_t* full[2];
  const CUtensorMap* tm_resid;  // fp32 (M, N): box {CW, 128}, swizzle = CW * 4 bytes
  const CUtensorMap* tm_u;      // bf16 (M, pitch): box {CW, 128}, swizzle = CW * 2 bytes
  bool has_resid, has_u;
  uint32_t bytes;               // per stage transaction bytes
  int u_off;                    // byte offset of the U chunk inside a stage
  int nch;                      // chunks per tile
  int tile_stride;              // tiles between two consecutive tiles of this warpgroup (2 * gridDim.x)
  int m_tiles;
  uint32_t k;                   // chunks consumed so far by this warpgroup
};

template <int CW>
__device__ __forceinline__ void ring_issue(const OpRing& rg, uint32_t kk, int tile, int ci, int n0) {
  const int s = kk & 1;
  mbar_arrive_expect_tx(rg.full[s], rg.bytes);
  if (rg.has_resid) tma_load_2d(rg.stage[s], rg.tm_resid, rg.full[s], n0 + ci * CW, tile * 128);
  if (rg.has_u) tma_load_2d(rg.stage[s] + rg.u_off, rg.tm_u, rg.full[s], n0 + ci * CW, tile * 128);
}

// CW fp32 values of this thread's row from a swizzled chunk (row bytes = CW * 4: SWIZZLE_128B / SWIZZLE_64B)
template <int CW>
__device__ __forceinline__ void ring_read_f32(const uint8_t* st, int row, float (&o)[CW]) {
  constexpr int RB = CW * 4, NV = RB / 16;
  const int sw = RB == 128 ? (row & 7) : ((row >> 1) & 3);
  const uint8_t* base = st + row * RB;
#pragma unroll
  for (int c = 0; c < NV; c++) {
    const float4 t = *reinterpret_cast<const float4*>(base + ((c ^ sw) << 4));
    o[4 * c] = t.x; o[4 * c + 1] = t.y; o[4 * c + 2] = t.z; o[4 * c + 3] = t.w;
  }
}
// CW bf16 values (row bytes = CW * 2: SWIZZLE_64B / SWIZZLE_32B)
template <int CW>
__device__ __forceinline__ void ring_read_bf16(const uint8_t* st, int row, float (&o)[CW]) {
  constexpr int RB = CW * 2, NV = RB / 16;
  const int sw = RB == 64 ? ((row >> 1) & 3) : ((row >> 2) & 1);
  const uint8_t* base = st + row * RB;
#pragma unroll
  for (int c = 0; c < NV; c++) {
    const uint4 t = *reinterpret_cast<const uint4*>(base + ((c ^ sw) << 4));
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const float2 f = __bfloat1622float2(h[j]);
      o[8 * c + 2 * j] = f.x; o[8 * c + 2 * j + 1] = f.y;
    }
  }
}

template <int CW>
__device__ __forceinline__ void tmem_ld_cw(uint32_t addr, uint32_t (&r)[CW]) {
  if constexpr (CW == 32) tmem_ld32(addr, r); else tmem_ld16(addr, r);
}
template <int CW>
__device__ __forceinline__ void tmem_st_cw(uint32_t addr, const uint32_t (&r)[CW]) {
  if constexpr (CW == 32) tmem_st32(addr, r); else tmem_st16(addr, r);
}

// vector helpers over CW columns, FULL = all CW columns are real (no predicates)
template <int CW, bool FULL>
__device__ __forceinline__ void ldv_f32(const float* p, int valid, float (&o)[CW]) {
  if (FULL) {
    const float4* q = reinterpret_cast<const float4*>(p);
#pragma unroll
    for (int i = 0; i < CW / 4; i++) {
      const float4 t = __ldg(q + i);
      o[4 * i] = t.x; o[4 * i + 1] = t.y; o[4 * i + 2] = t.z; o[4 * i + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < CW; i++) o[i] = i < valid ? __ldg(p + i) : 0.f;
  }
}
template <int CW, bool FULL>
__device__ __forceinline__ void stv_f32(float* p, int valid, const float (&v)[CW]) {
  if (FULL) {
    float4* q = reinterpret_cast<float4*>(p);
#pragma unroll
    for (int i = 0; i < CW / 4; i++) q[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  } else {
#pragma unroll
    for (int i = 0; i < CW; i++)
      if (i < valid) p[i] = v[i];
  }
}
// bf16 store; columns [valid, zero_valid) are written as zero (K padding of the consumer)
template <int CW, bool FULL>
__device__ __forceinline__ void stv_bf16(bf16* p, int valid, int zero_valid, const float (&v)[CW]) {
  if (FULL) {
    uint4* q = reinterpret_cast<uint4*>(p);
#pragma unroll
    for (int i = 0; i < CW / 8; i++) {
      uint4 t;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
      for (int j = 0; j < 4; j++) h[j] = __floats2bfloat162_rn(v[8 * i + 2 * j], v[8 * i + 2 * j + 1]);
      q[i] = t;
    }
  } else {
#pragma unroll
    for (int i = 0; i < CW; i++) {
      if (i < valid) p[i] = __float2bfloat16_rn(v[i]);
      else if (i < zero_valid) p[i] = __float2bfloat16_rn(0.f);
    }
  }
}


__device__ __forceinline__ int sw_of(int row_bytes, int row) {
  return row_bytes == 128 ? (row & 7) : (row_bytes == 64 ? ((row >> 1) & 3) : ((row >> 2) & 1));
}
// thread parks CW values of its own row (row index `r` inside `slab`, rows of RB bytes, swizzled)
template <int CW>
__device__ __forceinline__ void park_f32(uint8_t* slab, int r, const float (&v)[CW]) {
  constexpr int RB = CW * 4;
  const int sw = sw_of(RB, r);
#pragma unroll
  for (int c = 0; c < RB / 16; c++)
    *reinterpret_cast<float4*>(slab + r * RB + ((c ^ sw) << 4)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
}
template <int CW>
__device__ __forceinline__ void park_bf16(uint8_t* slab, int r, const float (&v)[CW]) {
  constexpr int RB = CW * 2;
  const int sw = sw_of(RB, r);
#pragma unroll
  for (int c = 0; c < RB / 16; c++) {
    uint4 t;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
    for (int j = 0; j < 4; j++) h[j] = __floats2bfloat162_rn(v[8 * c + 2 * j], v[8 * c + 2 * j + 1]);
    *reinterpret_cast<uint4*>(slab + r * RB + ((c ^ sw) << 4)) = t;
  }
}
// warp writes 32 parked rows (slab rows r0 .. r0+31, RB bytes each) to global rows g0 + i * pitch_bytes;
// 16-byte granules at byte offset >= limit_bytes are skipped
template <int RB>
__device__ __forceinline__ void flush_rows(const uint8_t* slab, int r0, int lane, uint8_t* g0, size_t pitch_bytes,
                                           int limit_bytes) {
  constexpr int LPR = RB / 16, RPI = 32 / LPR;
  const int ch = lane % LPR, rl = lane / LPR;
  if ((ch << 4) >= limit_bytes) return;
#pragma unroll
  for (int it = 0; it < LPR; it++) {
    const int rr = it * RPI + rl, r = r0 + rr;
    const uint4 t = *reinterpret_cast<const uint4*>(slab + r * RB + ((ch ^ sw_of(RB, r)) << 4));
    *reinterpret_cast<uint4*>(g0 + (size_t)rr * pitch_bytes + (ch << 4)) = t;
  }
}

// One column chunk of the main pass.  Returns with v[] = final values of the chunk (zero beyond `valid`).
template <int CW, bool FULL>
__device__ __forceinline__ void epi2_chunk(const Epilogue& ep, const EpiCtx& c, const OpRing* rg, int ci, int row,
                                           float* of32, bf16* ot, bool do_ln, float& ln_k, float& ln_s1,
                                           float& ln_s2) {
  const int nb = c.n0 + ci * CW;
  const int valid = c.N - nb;
  uint32_t r[CW];
  tmem_ld_cw<CW>(c.trow + (uint32_t)(ci * CW), r);
  float bv[CW], sc[CW];
  if (ep.bias) ldv_f32<CW, FULL>(ep.bias + nb, valid, bv);
  if (ep.eca_u) ldv_f32<CW, FULL>(ep.eca_scale + (size_t)c.b * c.N + nb, valid, sc);
  tmem_ld_wait();
  float v[CW];
#pragma unroll
  for (int j = 0; j < CW; j++) v[j] = __uint_as_float(r[j]);
  if (ep.bias) {
#pragma unroll
    for (int j = 0; j < CW; j++) v[j] += bv[j];
  }
  if (ep.scale_ncols > 0) {
    if (nb + CW <= ep.scale_ncols) {
#pragma unroll
      for (int j = 0; j < CW; j++) v[j] *= ep.out_scale;
    } else if (nb < ep.scale_ncols) {
#pragma unroll
      for (int j = 0; j < CW; j++) if (nb + j < ep.scale_ncols) v[j] *= ep.out_scale;
    }
  }
  if (ep.act == ACT_GELU) {
#pragma unroll
    for (int j = 0; j < CW; j++) v[j] = gelu_f(v[j]);
  } else if (ep.act == ACT_LRELU) {
#pragma unroll
    for (int j = 0; j < CW; j++) v[j] = lrelu_f(v[j]);
  }
  if (rg) {
    const uint8_t* st = rg->stage[rg->k & 1];
    if (rg->has_u) {
      float u[CW];
      ring_read_bf16<CW>(st + rg->u_off, row, u);
#pragma unroll
      for (int j = 0; j < CW; j++) v[j] = fmaf(ep.eca_coef * sc[j], u[j], v[j]);
    }
    if (rg->has_resid) {
      float rs[CW];
      ring_read_f32<CW>(st, row, rs);
#pragma unroll
      for (int j = 0; j < CW; j++) v[j] += rs[j];
    }
  }
  if (!FULL) {
#pragma unroll
    for (int j = 0; j < CW; j++) if (j >= valid) v[j] = 0.f;
  }
  const int lrow = row & 31;                       // row inside this warp's 32-row slice
  if (of32) {
    if (rg && rg->has_resid) {                     // park in place in the residual slab (own row already read)
      uint8_t* slab = rg->stage[rg->k & 1];
      park_f32<CW>(slab, row, v);
      __syncwarp();
      flush_rows<CW * 4>(slab, row - lrow, c.lane, reinterpret_cast<uint8_t*>(of32 - (size_t)lrow * ep.out_f32_pitch + nb),
                         (size_t)ep.out_f32_pitch * 4, (FULL ? CW : valid) * 4);
    } else {
      stv_f32<CW, FULL>(of32 + nb, valid, v);
    }
  }
  if (ot) {
    const int zv = ep.out_t_zero_to - nb;
    const int lim = FULL ? CW : (valid > zv ? valid : (zv < CW ? zv : CW));
    if (lim > 0) {
      park_bf16<CW>(c.wstage, lrow, v);
      __syncwarp();
      flush_rows<CW * 2>(c.wstage, 0, c.lane, reinterpret_cast<uint8_t*>(ot - (size_t)lrow * ep.out_t_pitch + nb),
                         (size_t)ep.out_t_pitch * 2, lim * 2);
      __syncwarp();
    }
  }
  if (do_ln) {
    if (ci == 0) ln_k = v[0];
#pragma unroll
    for (int j = 0; j < CW; j++) {
      if (FULL || j < valid) {
        const float dlt = v[j] - ln_k;
        ln_s1 += dlt;
        ln_s2 = fmaf(dlt, dlt, ln_s2);
      }
    }
    uint32_t w[CW];
#pragma unroll
    for (int j = 0; j < CW; j++) w[j] = __float_as_uint(v[j]);
    tmem_st_cw<CW>(c.trow + (uint32_t)(ci * CW), w);     // fp32 row parked in TMEM for the normalise pass
  }
}

// LayerNorm normalise pass over one chunk
template <int CW, bool FULL>
__device__ __forceinline__ void epi2_ln_chunk(const Epilogue& ep, const EpiCtx& c, int ci, bf16* lo, float mean,
                                              float rstd) {
  const int nb = c.n0 + ci * CW;
  const int valid = c.N - nb;
  uint32_t r[CW];
  tmem_ld_cw<CW>(c.trow + (uint32_t)(ci * CW), r);
  float g[CW], be[CW];
  ldv_f32<CW, FULL>(ep.ln_gamma + nb, valid, g);
  ldv_f32<CW, FULL>(ep.ln_beta + nb, valid, be);
  tmem_ld_wait();
  float v[CW];
#pragma unroll
  for (int j = 0; j < CW; j++) {
    const float xv = __uint_as_float(r[j]);
    v[j] = (FULL || j < valid) ? (xv - mean) * rstd * g[j] + be[j] : 0.f;
  }
  {
    const int lrow = c.m & 31;                     // tiles start at multiples of 128 rows
    const int zv = ep.ln_pitch - nb;
    const int lim = FULL ? CW : (valid > zv ? valid : (zv < CW ? zv : CW));
    if (lim > 0) {
      park_bf16<CW>(c.wstage, lrow, v);
      __syncwarp();
      flush_rows<CW * 2>(c.wstage, 0, c.lane, reinterpret_cast<uint8_t*>(lo - (size_t)lrow * ep.ln_pitch + nb),
                         (size_t)ep.ln_pitch * 2, lim * 2);
      __syncwarp();
    }
  }
  if (ci == 0 && ep.ln_head) {      // compact / x-padded copies of the first 16 channels + their column sums
    float h16[16];
#pragma unroll
    for (int j = 0; j < 16; j++) h16[j] = v[j];
    bf16* hd = (bf16*)ep.ln_head + (size_t)c.m * ep.ln_head_c;
    if (ep.ln_head_c == 16) {
      st16_bf16(hd, h16, 16, 16);
    } else {
#pragma unroll
      for (int j = 0; j < 16; j++)
        if (j < ep.ln_head_c) hd[j] = __float2bfloat16_rn(h16[j]);
    }
    if (ep.ln_head_pad && ep.ln_head_c == 16) {
      bf16* hp = (bf16*)ep.ln_head_pad + (((size_t)c.b * c.H + c.py) * (c.W + 12) + c.px + 6) * 16;
      st16_bf16(hp, h16, 16, 16);
    }
    if (ep.ln_head_colsum) {
#pragma unroll
      for (int j = 0; j < 16; j++) {
        float s = warp_sum(__bfloat162float(__float2bfloat16_rn(h16[j])));
        if (c.lane == 0) c.hsum[c.quarter][j] = s;
      }
    }
  }
}

// Whole-tile epilogue.  rg == nullptr: no TMA-fed operands (bias / activation only).
template <int CW>
__device__ __forceinline__ void run_epilogue2(const Epilogue& ep, const EpiCtx& c, OpRing* rg, int row, int tile) {
  const int nch = c.BN / CW;
  const bool do_ln = ep.ln_out != nullptr;
  float* of32 = ep.out_f32 ? ep.out_f32 + (size_t)c.m * ep.out_f32_pitch : nullptr;
  bf16* ot = ep.out_t ? (bf16*)ep.out_t + (size_t)c.m * ep.out_t_pitch + ep.out_t_coff : nullptr;
  float ln_k = 0.f, ln_s1 = 0.f, ln_s2 = 0.f;
  for (int ci = 0; ci < nch; ci++) {
    const int valid = c.N - (c.n0 + ci * CW);
    if (rg) mbar_wait(rg->full[rg->k & 1], (rg->k >> 1) & 1u);
    if (valid >= CW) epi2_chunk<CW, true>(ep, c, rg, ci, row, of32, ot, do_ln, ln_k, ln_s1, ln_s2);
    else epi2_chunk<CW, false>(ep, c, rg, ci, row, of32, ot, do_ln, ln_k, ln_s1, ln_s2);
    if (rg) {
      epi_bar(c.bar_id);                       // every thread of the warpgroup has read this stage
      if (c.gtid == 0) {                       // refill it with the chunk two ahead (possibly of the next tile)
        int cn = ci + 2, tn = tile;
        if (cn >= nch) { cn -= nch; tn += rg->tile_stride; }
        if (tn < rg->m_tiles) ring_issue<CW>(*rg, rg->k, tn, cn, c.n0);
      }
      rg->k++;
    }
  }
  if (ot && c.last_ntile) {                    // trailing K-padding columns that no n-tile covers
    for (int n = c.n0 + c.BN; n < ep.out_t_zero_to; n++) ot[n] = __float2bfloat16_rn(0.f);
  }
  if (do_ln) {
    tmem_st_wait();
    const float inv_n = 1.f / (float)c.N;
    const float mean_d = ln_s1 * inv_n;
    const float var = fmaxf(ln_s2 * inv_n - mean_d * mean_d, 0.f);
    const float rstd = rsqrtf(var + 1e-5f);
    const float mean = ln_k + mean_d;
    bf16* lo = (bf16*)ep.ln_out + (size_t)c.m * ep.ln_pitch;
    for (int ci = 0; ci < nch; ci++) {
      const int valid = c.N - (c.n0 + ci * CW);
      if (valid >= CW) epi2_ln_chunk<CW, true>(ep, c, ci, lo, mean, rstd);
      else epi2_ln_chunk<CW, false>(ep, c, ci, lo, mean, rstd);
    }
    for (int n = c.n0 + c.BN; n < ep.ln_pitch; n++) lo[n] = __float2bfloat16_rn(0.f);
    if (ep.ln_head_colsum) {
      epi_bar(c.bar_id);
      if (c.gtid < ep.ln_head_c && c.gtid < 16)
        ep.ln_head_colsum[((size_t)c.b * c.nchunk + c.chunk) * ep.ln_head_c + c.gtid] =
            c.hsum[0][c.gtid] + c.hsum[1][c.gtid] + c.hsum[2][c.gtid] + c.hsum[3][c.gtid];
      epi_bar(c.bar_id);                       // hsum is reused by this warpgroup's next tile
    }
  }
}

}  // namespace hb
