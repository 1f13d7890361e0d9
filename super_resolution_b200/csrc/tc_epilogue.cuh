// tc_epilogue.cuh -- the fused epilogue shared by the tcgen05 kernels (tc_tapconv.cu, tc_gemm.cu).
//
// One epilogue thread owns ONE token row of the 128-row accumulator tile (TMEM lane = row) and walks
// its BN columns in chunks of 16 (tcgen05.ld 32x32b.x16).  Value order (common.cuh, struct Epilogue):
//   v = acc + bias; v += dynamic depthwise 3x3 (esc_arch.py:123); v *= q-scale; v = act(v);
//   v += conv_scale * eca_scale[b,n] * U[m,n] (hat_arch.py:236); v += resid[m,n]
// Outputs: fp32 residual stream, bf16 activation (optionally pixel-shuffled, hat_arch.py:593-605),
// per-tile column sums (global-average-pool partials in a fixed order -> deterministic), and a full
// LayerNorm of the row (eps 1e-5; the row lives in one thread, so the statistics need no reduction;
// the fp32 row is parked in TMEM between the two passes).
// Global operands of chunk c+1 (residual, CAB branch) are prefetched while chunk c is processed.
#pragma once
#include "tc_common.cuh"

namespace hb {

using namespace tc;

struct EpiCtx {
  uint32_t trow;      // TMEM address of this thread's lane quarter, column 0 of the accumulator
  int m;              // token row (global)
  int b, py, px;      // sample, pixel coordinates of the token
  int H, W;           // image size of this launch (tokens are (b, y, x) row-major)
  int N, n0, BN;      // real output channels, first column of this CTA's n-tile, n-tile width
  bool last_ntile;    // this CTA owns the last n-tile (writes trailing K-padding zeros)
  int chunk;          // index of this 128-row tile inside its sample (column-sum partial slot)
  int nchunk;         // tiles per sample
  int quarter, lane;  // TMEM lane quarter (warp % 4), lane
  int gtid;           // 0..127 inside the epilogue warpgroup
  int bar_id;         // named barrier of this epilogue warpgroup (128 threads)
  float (*sred)[256]; // [4][256] shared scratch of this warpgroup
  uint8_t* wstage = nullptr;      // tc_epilogue2: 2 KB store-staging scratch of this WARP
  float (*hsum)[16] = nullptr;    // tc_epilogue2: [4][16] per-warp column sums of the LN head
};

__device__ __forceinline__ void epi_bar(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

__device__ __forceinline__ bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// 16 fp32 from global; `valid` of them exist
__device__ __forceinline__ void ld16_f32(const float* p, int valid, float (&o)[16]) {
  if (valid >= 16 && al16(p)) {
    const float4* q = reinterpret_cast<const float4*>(p);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float4 t = __ldg(q + i);
      o[4 * i] = t.x; o[4 * i + 1] = t.y; o[4 * i + 2] = t.z; o[4 * i + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; i++) o[i] = i < valid ? __ldg(p + i) : 0.f;
  }
}
// same, but for data written earlier in this launch sequence by other kernels (plain loads)
__device__ __forceinline__ void ld16_f32_plain(const float* p, int valid, float (&o)[16]) {
  if (valid >= 16 && al16(p)) {
    const float4* q = reinterpret_cast<const float4*>(p);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float4 t = q[i];
      o[4 * i] = t.x; o[4 * i + 1] = t.y; o[4 * i + 2] = t.z; o[4 * i + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; i++) o[i] = i < valid ? p[i] : 0.f;
  }
}
__device__ __forceinline__ void ld16_bf16(const bf16* p, int valid, float (&o)[16]) {
  if (valid >= 16 && al16(p)) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
#pragma unroll
    for (int i = 0; i < 2; i++) {
      uint4 t = q[i];
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float2 f = __bfloat1622float2(h[j]);
        o[8 * i + 2 * j] = f.x; o[8 * i + 2 * j + 1] = f.y;
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; i++) o[i] = i < valid ? __bfloat162float(p[i]) : 0.f;
  }
}
// store 16 values as bf16; `valid` of them are real, columns [valid, zero_valid) are written as zero
__device__ __forceinline__ void st16_bf16(bf16* p, const float (&v)[16], int valid, int zero_valid) {
  if (valid >= 16 && al16(p)) {
    uint4 t[2];
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(t);
#pragma unroll
    for (int i = 0; i < 8; i++) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = t[0];
    q[1] = t[1];
  } else {
#pragma unroll
    for (int i = 0; i < 16; i++) {
      if (i < valid) p[i] = __float2bfloat16_rn(v[i]);
      else if (i < zero_valid) p[i] = __float2bfloat16_rn(0.f);
    }
  }
}
__device__ __forceinline__ void st16_f32(float* p, const float (&v)[16], int valid) {
  if (valid >= 16 && al16(p)) {
    float4* q = reinterpret_cast<float4*>(p);
#pragma unroll
    for (int i = 0; i < 4; i++) q[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  } else {
#pragma unroll
    for (int i = 0; i < 16; i++)
      if (i < valid) p[i] = v[i];
  }
}

__device__ __forceinline__ void run_epilogue(const Epilogue& ep, const EpiCtx& c) {
  const int N = c.N, n0 = c.n0, m = c.m, b = c.b, lane = c.lane, quarter = c.quarter;
  const bool do_ln = ep.ln_out != nullptr;
  const int nch = c.BN / 16;

  float ln_k = 0.f, ln_s1 = 0.f, ln_s2 = 0.f;
  // prefetched global operands of the current chunk
  float rs[16], u[16];
  {
    const int valid = N - n0;
    if (ep.resid && valid > 0) ld16_f32_plain(ep.resid + (size_t)m * ep.resid_pitch + n0, valid, rs);
    if (ep.eca_u && valid > 0) ld16_bf16((const bf16*)ep.eca_u + (size_t)m * ep.eca_pitch + n0, valid, u);
  }
  for (int ci = 0; ci < nch; ci++) {
    const int nb = n0 + ci * 16;
    const int valid = N - nb;                                   // real columns in this chunk (may be <= 0)
    uint32_t r[16];
    tmem_ld16(c.trow + (uint32_t)(ci * 16), r);
    // prefetch the next chunk's residual / CAB-branch operands while the TMEM load is in flight
    float rs_n[16], u_n[16];
    {
      const int nbn = nb + 16, vn = N - nbn;
      if (ci + 1 < nch && vn > 0) {
        if (ep.resid) ld16_f32_plain(ep.resid + (size_t)m * ep.resid_pitch + nbn, vn, rs_n);
        if (ep.eca_u) ld16_bf16((const bf16*)ep.eca_u + (size_t)m * ep.eca_pitch + nbn, vn, u_n);
      }
    }
    tmem_ld_wait();
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; j++) v[j] = __uint_as_float(r[j]);
    if (valid > 0) {
      if (ep.bias) {
        float bv[16];
        ld16_f32(ep.bias + nb, valid, bv);
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] += bv[j];
      }
      if (ep.dyn_k) {       // per-sample dynamic depthwise 3x3 on the compact channels (esc_arch.py:123)
        const bf16* xs = (const bf16*)ep.dyn_x;
        for (int ky = 0; ky < 3; ky++) {
          const int yy = c.py + ky - 1;
          if (yy < 0 || yy >= c.H) continue;
          for (int kx = 0; kx < 3; kx++) {
            const int xx = c.px + kx - 1;
            if (xx < 0 || xx >= c.W) continue;
            float xv[16];
            ld16_bf16(xs + (size_t)(m + (ky - 1) * c.W + (kx - 1)) * ep.dyn_pitch + nb, valid, xv);
#pragma unroll
            for (int j = 0; j < 16; j++)
              if (j < valid) v[j] = fmaf(xv[j], __ldg(ep.dyn_k + ((size_t)b * N + nb + j) * 9 + ky * 3 + kx), v[j]);
          }
        }
      }
      if (ep.scale_ncols > 0) {
#pragma unroll
        for (int j = 0; j < 16; j++) if (nb + j < ep.scale_ncols) v[j] *= ep.out_scale;
      }
      if (ep.act == ACT_GELU) {
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = gelu_f(v[j]);
      } else if (ep.act == ACT_LRELU) {
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = lrelu_f(v[j]);
      }
      if (ep.eca_u) {
        float sc[16];
        ld16_f32_plain(ep.eca_scale + (size_t)b * N + nb, valid, sc);
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = fmaf(ep.eca_coef * sc[j], u[j], v[j]);
      }
      if (ep.resid) {
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] += rs[j];
      }
#pragma unroll
      for (int j = 0; j < 16; j++) if (j >= valid) v[j] = 0.f;
    } else {
#pragma unroll
      for (int j = 0; j < 16; j++) v[j] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 16; j++) { rs[j] = rs_n[j]; u[j] = u_n[j]; }
    // ---- outputs ----
    if (ep.out_f32 && valid > 0) st16_f32(ep.out_f32 + (size_t)m * ep.out_f32_pitch + nb, v, valid);
    if (ep.out_t) {
      bf16* ot = (bf16*)ep.out_t;
      if (ep.shuffle_r == 1) {
        const int zv = ep.out_t_zero_to - nb;
        if (valid > 0 || zv > 0) st16_bf16(ot + (size_t)m * ep.out_t_pitch + ep.out_t_coff + nb, v, valid, zv);
      } else if (valid > 0) {
        // PixelShuffle(r) folded into the store: out[b, c, y*r+i, x*r+j] = in[b, c*r*r + i*r + j, y, x]
        const int r_ = ep.shuffle_r, rr = r_ * r_;
        const size_t orow_w = (size_t)c.W * r_;
        if (r_ == 2 && valid >= 16) {
          const int c0 = nb >> 2;   // 4 channels, 4 sub-pixels
#pragma unroll
          for (int ij = 0; ij < 4; ij++) {
            const size_t otk =
                ((size_t)b * c.H * 2 + (size_t)c.py * 2 + (ij >> 1)) * orow_w + (size_t)c.px * 2 + (ij & 1);
            __nv_bfloat162 h0 = __floats2bfloat162_rn(v[ij], v[4 + ij]);
            __nv_bfloat162 h1 = __floats2bfloat162_rn(v[8 + ij], v[12 + ij]);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&h0);
            pk.y = *reinterpret_cast<uint32_t*>(&h1);
            *reinterpret_cast<uint2*>(ot + otk * ep.out_t_pitch + ep.out_t_coff + c0) = pk;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; j++) {       // fully unrolled: v[] must stay in registers
            if (j < valid) {
              const int n = nb + j;
              const int cc = n / rr, ij = n - cc * rr, i = ij / r_, jj = ij - i * r_;
              const size_t otk = ((size_t)b * c.H * r_ + (size_t)c.py * r_ + i) * orow_w + (size_t)c.px * r_ + jj;
              ot[otk * ep.out_t_pitch + ep.out_t_coff + cc] = __float2bfloat16_rn(v[j]);
            }
          }
        }
      }
    }
    if (ep.colsum) {          // per-tile column sums of v (global-average-pool partials, fixed order)
#pragma unroll
      for (int j = 0; j < 16; j++) {
        float s = warp_sum(v[j]);
        if (lane == 0) c.sred[quarter][ci * 16 + j] = s;
      }
    }
    if (do_ln) {
      if (ci == 0) ln_k = v[0];
#pragma unroll
      for (int j = 0; j < 16; j++) {
        if (j < valid) {
          const float dlt = v[j] - ln_k;
          ln_s1 += dlt;
          ln_s2 = fmaf(dlt, dlt, ln_s2);
        }
      }
      uint32_t w[16];
#pragma unroll
      for (int j = 0; j < 16; j++) w[j] = __float_as_uint(v[j]);
      tmem_st16(c.trow + (uint32_t)(ci * 16), w);     // keep the fp32 row in TMEM for the normalise pass
    }
  }
  // trailing K-padding columns of the consumer that no n-tile covers
  if (ep.out_t && ep.shuffle_r == 1 && c.last_ntile) {
    bf16* ot = (bf16*)ep.out_t + (size_t)m * ep.out_t_pitch + ep.out_t_coff;
    for (int n = n0 + c.BN; n < ep.out_t_zero_to; n++) ot[n] = __float2bfloat16_rn(0.f);
  }
  if (ep.colsum) {
    epi_bar(c.bar_id);
    for (int n = c.gtid; n < c.BN; n += 128)
      if (n0 + n < N)
        ep.colsum[((size_t)b * c.nchunk + c.chunk) * N + n0 + n] =
            c.sred[0][n] + c.sred[1][n] + c.sred[2][n] + c.sred[3][n];
  }
  if (do_ln) {
    // fused LayerNorm of the row (eps 1e-5), shifted single-pass statistics in fp32
    tmem_st_wait();
    const float inv_n = 1.f / (float)N;
    const float mean_d = ln_s1 * inv_n;
    const float var = fmaxf(ln_s2 * inv_n - mean_d * mean_d, 0.f);
    const float rstd = rsqrtf(var + 1e-5f);
    const float mean = ln_k + mean_d;
    bf16* lo = (bf16*)ep.ln_out + (size_t)m * ep.ln_pitch;
    if (ep.ln_head_colsum && ep.colsum) epi_bar(c.bar_id);   // sred reuse after the colsum readers
    for (int ci = 0; ci < nch; ci++) {
      const int nb = n0 + ci * 16;
      const int valid = N - nb;
      uint32_t r[16];
      tmem_ld16(c.trow + (uint32_t)(ci * 16), r);
      float g[16], be[16];
      if (valid > 0) {
        ld16_f32(ep.ln_gamma + nb, valid, g);
        ld16_f32(ep.ln_beta + nb, valid, be);
      }
      tmem_ld_wait();
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; j++) {
        const float xv = __uint_as_float(r[j]);
        v[j] = j < valid ? (xv - mean) * rstd * g[j] + be[j] : 0.f;
      }
      st16_bf16(lo + nb, v, valid, ep.ln_pitch - nb);
      if (ci == 0 && ep.ln_head) {
        bf16* hd = (bf16*)ep.ln_head + (size_t)m * ep.ln_head_c;
        if (ep.ln_head_c == 16) {
          st16_bf16(hd, v, 16, 16);
        } else {
#pragma unroll
          for (int j = 0; j < 16; j++)
            if (j < ep.ln_head_c) hd[j] = __float2bfloat16_rn(v[j]);
        }
        if (ep.ln_head_pad && ep.ln_head_c == 16) {
          bf16* hp = (bf16*)ep.ln_head_pad + (((size_t)b * c.H + c.py) * (c.W + 12) + c.px + 6) * 16;
          st16_bf16(hp, v, 16, 16);
        }
        if (ep.ln_head_colsum) {
          // sums of the bf16-rounded head values (what the separate colsum kernel would read)
#pragma unroll
          for (int j = 0; j < 16; j++) {
            float s = warp_sum(__bfloat162float(__float2bfloat16_rn(v[j])));
            if (lane == 0) c.sred[quarter][j] = s;
          }
        }
      }
    }
    for (int n = n0 + c.BN; n < ep.ln_pitch; n++) lo[n] = __float2bfloat16_rn(0.f);
    if (ep.ln_head_colsum) {
      epi_bar(c.bar_id);
      if (c.gtid < ep.ln_head_c && c.gtid < 16)
        ep.ln_head_colsum[((size_t)b * c.nchunk + c.chunk) * ep.ln_head_c + c.gtid] =
            c.sred[0][c.gtid] + c.sred[1][c.gtid] + c.sred[2][c.gtid] + c.sred[3][c.gtid];
    }
  }
  // persistent kernels reuse sred for the next tile of this warpgroup: readers must be done first
  if (ep.colsum || ep.ln_head_colsum) epi_bar(c.bar_id);
}

}  // namespace hb
