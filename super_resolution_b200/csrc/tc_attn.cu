// tc_attn.cu -- OCAB overlapping-window attention core on the 5th-gen tensor cores (flash style).
//
// Reference: OCAB.forward, hat_arch.py:354-388 (window_size 16, overlap window 24, 6 heads):
//   per (window, head):  S = (q*scale) k^T + rel-pos-bias   (256 x 576),  softmax over the 576 keys,  O = P v
// Keys/values of a window are the 24x24 pixels around it; positions outside the image are ZERO vectors
// that still take softmax mass (nn.Unfold pads after the projection, hat_arch.py:296-297,:359-360).
//
// B200 mapping.  The qkv GEMM writes a head-padded layout (head_dim 30 -> 32: q | k | v, each heads x 32
// channels, pad channels exactly zero), so every operand tile is a TMA box of the token tensor:
//   Q   {32 ch, 8 x of one parity, 16 y} -> 128 rows x 64 B  (K-major A operand, SWIZZLE_64B), twice: warpgroup 0
//                                            takes the even query columns of the window, warpgroup 1 the odd ones
//   K,V {32 ch, 24 x, 4 y} per key chunk ->  96 rows x 64 B  (K: K-major B operand; V: MN-major B operand)
// with TMA out-of-bounds zero fill standing in for the unfold padding -- no unfold tensor, no S matrix in HBM
// (reference: 906 MB of fp32 logits at the bench config).
// A persistent CTA walks (window, head) units.  Warp roles (320 threads):
//   warp 0   TMA producer: Q + bias tables (2 stages), K/V chunks (5-stage ring, prefetches across units)
//   warp 1   MMA issuer:   S[g] = Q[g] K_c^T  (M128 N96 K32)  and  O[g] = P[g] V_c  (M128 N32 K96)
//   warps 2-5 / 6-9  softmax warpgroups g = 0/1, one query row per thread (rows 0-127 / 128-255):
//            one pass per 96-key chunk: S*log2e + bias, exp2 against a fixed per-row reference exponent (the maximum
//            of the row's first 32 logits), row sum, P as bf16 into shared memory in the UMMA K-major layout (double
//            buffered).  O accumulates over the six chunks in TMEM (no per-chunk rescale: a fixed reference needs none);
//            a chunk that would overflow takes the exact `rescale_chunk` path (see the kernel).
// TMEM: S double-buffered per warpgroup (4 x 96 columns) + O per warpgroup, alternating per unit (4 x 32) = 512.
#include "kernels.cuh"
#include "tc_common.cuh"

namespace hb {

using namespace tc;
unsigned long long* gemm_trace_arm(cudaStream_t st);

namespace {
constexpr int AT_THREADS = 320;
constexpr int WS = 16, OWS = 24, PADW = 4, DP = 32, TN = 39;     // window, overlap window, ring, padded head dim, 16+24-1
constexpr int BP = ATTN_TC_BIAS_PITCH;                           // bias table row pitch in shared memory (see kernels.cuh)
constexpr int KC = 96, NCH = 6, KROWS = 4;                       // keys per chunk (4 key rows), chunks per unit
constexpr int KV_STAGES = 5;
constexpr int Q_BYTES = 256 * 64, BIAS_FLOATS = ATTN_TC_BIAS_FLOATS, BIAS_BYTES = BIAS_FLOATS * 4;
constexpr int QST_BYTES = (Q_BYTES + BIAS_BYTES + 1023) / 1024 * 1024;   // 31744: Q tiles stay swizzle aligned
constexpr int K_BYTES = KC * 64, KVST_BYTES = 2 * K_BYTES;       // 6144, 12288
constexpr int P_BYTES = 3 * 128 * 64;                            // 24576 per warpgroup and buffer
constexpr int SMEM_BYTES = 2 * QST_BYTES + KV_STAGES * KVST_BYTES + 4 * P_BYTES + 1024;      // 224256
static_assert(BIAS_BYTES % 16 == 0 && QST_BYTES % 128 == 0, "bulk copy / TMA destination alignment");
constexpr uint32_t S_COLS = 96, O_COL0 = 384;
constexpr float LOG2E = 1.4426950408889634f;

struct AttnParams {
  int B, H, W, heads, nwx, nwy, units;
  int k_off, v_off;          // first k / v channel
  const float* bias_l2;      // (heads, 2, 39 x 48): compact rel-pos bias * log2(e), rows padded to BP, + copy shifted by one
  bf16* out;
  int out_pitch;
  unsigned long long* trace;   // engineering (HATB200_GEMM_TRACE): per-CTA cycle counters of warp 2, else null
};

// MN-major (B = V[key][d]) shared-memory descriptor: 64-byte rows, SWIZZLE_64B, 8-key groups 512 B apart
__device__ __forceinline__ uint64_t make_desc_mn64(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(512 >> 4) << 16;      // LBO (next 32-wide MN atom; unused for N = 32)
  d |= (uint64_t)(512 >> 4) << 32;      // SBO: next group of 8 K rows
  d |= 1ull << 46;
  d |= 4ull << 61;                      // SWIZZLE_64B
  return d;
}
// packed fp32 pairs: one issue slot for two lanes of work (the softmax warps are issue bound, not pipe bound)
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return *reinterpret_cast<float2*>(&r);
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&r);
}
__device__ __forceinline__ float2 lds_f32x2(uint32_t a) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ float lds_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
// byte offset of key l (0..95) of a chunk inside the bias table: 4 key rows of 24
__device__ __forceinline__ constexpr uint32_t bias_off(int l) { return (uint32_t)((l / OWS) * BP + (l % OWS)) * 4u; }
}  // namespace

template <int DBG>
__global__ void __launch_bounds__(AT_THREADS, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t q_full[2], q_empty[2];
  __shared__ __align__(8) uint64_t kv_full[KV_STAGES], kv_empty[KV_STAGES];
  __shared__ __align__(8) uint64_t s_full[2][2], s_empty[2][2];
  __shared__ __align__(8) uint64_t p_full[2][2], pv_done[2][2];
  __shared__ uint32_t tmem_slot;
  __shared__ unsigned long long uoff[2];     // per Q stage: element offset of the unit's (window origin, head) in `out`

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sm_q = smem;                                  // [2][Q 16 KB | bias 6 KB]
  uint8_t* sm_kv = sm_q + 2 * QST_BYTES;                 // [KV_STAGES][K 6 KB | V 6 KB]
  uint8_t* sm_p = sm_kv + KV_STAGES * KVST_BYTES;        // [2 warpgroups][2 buffers][3 blocks of 128 rows x 64 B]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmKV);
    for (int i = 0; i < 2; i++) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1 + 256);
      for (int j = 0; j < 2; j++) {
        mbar_init(&p_full[i][j], 128);
        mbar_init(&s_full[i][j], 1);
        mbar_init(&s_empty[i][j], 128);
        mbar_init(&pv_done[i][j], 1);
      }
    }
    for (int i = 0; i < KV_STAGES; i++) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  pdl_trigger();
  const uint32_t tmem_base = tmem_slot;
  const int wins = p.nwx * p.nwy;

  if (warp == 0) {
    // ===== TMA producer =====  (runs a whole unit ahead: relaxed waits keep its polling off the issue port)
    if (elect_one()) {
      int ui = 0, kvit = 0;
      for (int u = blockIdx.x; u < p.units; u += gridDim.x, ui++) {
        const int head = u % p.heads, win = u / p.heads;
        const int b = win / wins, wr = win % wins, wy = wr / p.nwx, wx = wr % p.nwx;
        const int qb = ui & 1;
        mbar_wait_relaxed(&q_empty[qb], (((uint32_t)ui >> 1) & 1u) ^ 1u);
        uint8_t* qs = sm_q + qb * QST_BYTES;
        // the softmax warps need the unit only for their output address: one offset through shared memory instead of
        // four integer divisions per thread and unit (published by the arrive below)
        uoff[qb] = (((unsigned long long)b * p.H + wy * WS) * p.W + wx * WS) * (unsigned long long)p.out_pitch + head * DP;
        mbar_arrive_expect_tx(&q_full[qb], Q_BYTES + BIAS_BYTES);
        tma_load_4d(qs, &tmQ, &q_full[qb], head * DP, 0, wx * (WS / 2), b * p.H + wy * WS);          // even columns
        tma_load_4d(qs + Q_BYTES / 2, &tmQ, &q_full[qb], head * DP, 1, wx * (WS / 2), b * p.H + wy * WS);   // odd columns
        bulk_load_1d(qs + Q_BYTES, p.bias_l2 + (size_t)head * BIAS_FLOATS, BIAS_BYTES, &q_full[qb]);
        for (int kc = 0; kc < NCH; kc++, kvit++) {
          const int st = kvit % KV_STAGES;
          mbar_wait_relaxed(&kv_empty[st], (((uint32_t)(kvit / KV_STAGES)) & 1u) ^ 1u);
          uint8_t* ks = sm_kv + st * KVST_BYTES;
          mbar_arrive_expect_tx(&kv_full[st], KVST_BYTES);
          const int x0 = wx * WS - PADW, y0 = wy * WS - PADW + kc * KROWS;
          tma_load_4d(ks, &tmKV, &kv_full[st], p.k_off + head * DP, x0, y0, b);
          tma_load_4d(ks + K_BYTES, &tmKV, &kv_full[st], p.v_off + head * DP, x0, y0, b);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    // One flat software pipeline over all chunks of all units of this CTA:  S(c) is issued before PV(c - 1), also
    // across a unit boundary, so the first logits of the next unit are in TMEM while the softmax warpgroups still
    // finish the previous one.
    if (elect_one()) {
      const uint32_t idesc_s = make_idesc_bf16(128, KC);
      const uint32_t idesc_o = make_idesc_bf16(128, DP) | (1u << 16);     // B operand MN-major
      int ui = 0, kvit = 0;
      int pv_slot = -1, pv_kc = 0, pv_ui = 0;                             // the PV still to be issued
      auto issue_pv = [&]() {             // O(unit) += P(kc) V(kc): accumulates over the six chunks of a unit in TMEM
        const uint32_t v_addr = smem_u32(sm_kv + pv_slot * KVST_BYTES + K_BYTES);
        const int pb = pv_kc & 1;
        for (int g = 0; g < 2; g++) {
          mbar_wait(&p_full[g][pb], (uint32_t)(pv_ui * 3 + (pv_kc >> 1)) & 1u);
          tc_fence_after();
          const uint32_t p_addr = smem_u32(sm_p + (g * 2 + pb) * P_BYTES);
          const uint32_t d_o = tmem_base + O_COL0 + (uint32_t)(g * 2 + (pv_ui & 1)) * DP;
#pragma unroll
          for (int j = 0; j < KC / 16; j++) {
            const uint64_t ad = make_smem_desc(p_addr + (j >> 1) * 8192 + (j & 1) * 32, 64);
            const uint64_t bd = make_desc_mn64(v_addr + j * 1024);
            umma_bf16(d_o, ad, bd, idesc_o, (j != 0 || pv_kc != 0) ? 1u : 0u);
          }
          umma_commit(&pv_done[g][pb]);
        }
        umma_commit(&kv_empty[pv_slot]);
      };
      for (int u = blockIdx.x; u < p.units; u += gridDim.x, ui++) {
        const int qb = ui & 1;
        const uint32_t q_addr = smem_u32(sm_q + qb * QST_BYTES);
        mbar_wait(&q_full[qb], ((uint32_t)ui >> 1) & 1u);
        tc_fence_after();
        for (int kc = 0; kc < NCH; kc++, kvit++) {
          const int st = kvit % KV_STAGES;
          mbar_wait(&kv_full[st], ((uint32_t)(kvit / KV_STAGES)) & 1u);
          tc_fence_after();
          const uint32_t k_addr = smem_u32(sm_kv + st * KVST_BYTES);
          const int sb = kc & 1;
          for (int g = 0; g < 2; g++) {
            mbar_wait(&s_empty[g][sb], ((uint32_t)(ui * 3 + (kc >> 1)) & 1u) ^ 1u);
            tc_fence_after();
            const uint32_t d_s = tmem_base + (uint32_t)(g * 2 + sb) * S_COLS;
#pragma unroll
            for (int k = 0; k < 2; k++) {
              const uint64_t ad = make_smem_desc(q_addr + g * 8192 + k * 32, 64);
              const uint64_t bd = make_smem_desc(k_addr + k * 32, 64);
              umma_bf16(d_s, ad, bd, idesc_s, k);
            }
            umma_commit(&s_full[g][sb]);
          }
          if (kc == NCH - 1) umma_commit(&q_empty[qb]);      // Q tile consumed once these MMAs complete
          if (pv_slot >= 0) issue_pv();
          pv_slot = st; pv_kc = kc; pv_ui = ui;
        }
      }
      if (pv_slot >= 0) issue_pv();
    }
  } else {
    // ===== softmax warpgroups =====
    const int g = (warp - 2) >> 2;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;            // query row inside this warpgroup's half
    const int qy = row >> 3, qx = 2 * (row & 7) + g;  // warpgroup g holds the window's columns of parity g (see tmQ)
    const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
    const int psw = (row >> 1) & 3;                 // SWIZZLE_64B: 16-byte chunk index ^= (row / 2) % 4
    const float2 L2 = make_float2(LOG2E, LOG2E);
    // Deferred unit epilogue: the last PV of a unit is still in flight when its softmax is done, so the normalise +
    // store of unit i runs after chunk 0 of unit i + 1 has been handed to the tensor core (O alternates per unit).
    bool pend = false;
    int pend_ui = 0;
    float pend_l = 1.f;
    bf16* pend_row = nullptr;
    auto flush = [&]() {
      mbar_wait(&pv_done[g][1], (uint32_t)(pend_ui * 3 + 2) & 1u);
      tc_fence_after();
      uint32_t ro[32];
      tmem_ld32(tmem_base + lane_sel + O_COL0 + (uint32_t)(g * 2 + (pend_ui & 1)) * DP, ro);
      tmem_ld_wait();
      const float inv = 1.f / pend_l;
#pragma unroll
      for (int c = 0; c < 4; c++) {       // 32 (padded) head channels of one token
        uint4 t;
        __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
        for (int j = 0; j < 4; j++)
          h2[j] = __floats2bfloat162_rn(__uint_as_float(ro[8 * c + 2 * j]) * inv, __uint_as_float(ro[8 * c + 2 * j + 1]) * inv);
        reinterpret_cast<uint4*>(pend_row)[c] = t;
      }
      pend = false;
    };
    long long tc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, c0 = 0;
    const bool tr = DBG == 5 && p.trace != nullptr;
#define AT_T(i) do { if (tr) { const long long c1 = clock64(); tc[i] += c1 - c0; c0 = c1; } } while (0)
    if (tr) c0 = clock64();
    int ui = 0;
    const size_t my_off = ((size_t)qy * p.W + qx) * p.out_pitch;     // this thread's query inside the window
    for (int u = blockIdx.x; u < p.units; u += gridDim.x, ui++) {
      const int qb = ui & 1;
      mbar_wait(&q_full[qb], ((uint32_t)ui >> 1) & 1u);
      const unsigned long long unit_off = uoff[qb];
      AT_T(0);
      // bias(q, k) = table[(ky - qy + 15) * BP + kx - qx + 15]: this thread's (qy, qx) goes into the base address.
      // 15 - qx is even for warpgroup 1 (aligned pairs of table A) and odd for warpgroup 0, which reads the copy
      // shifted by one float at index - 1.
      const uint32_t bias_a = smem_u32(sm_q + qb * QST_BYTES + Q_BYTES) +
                              (uint32_t)((1 - g) * ATTN_TC_BIAS_TABLE + (WS - 1 - qy) * BP + (WS - 1 - qx) - (1 - g)) * 4u;
      const uint32_t o_addr = tmem_base + lane_sel + O_COL0 + (uint32_t)(g * 2 + (ui & 1)) * DP;
      float m_ref = 0.f;
      float2 l2 = make_float2(0.f, 0.f);            // row sum, even / odd keys

      // ONE pass per 96-key chunk against a FIXED reference exponent per row (round 1: two passes per chunk with the
      // logits parked back in TMEM, and an online-softmax rescale of O per chunk).  Softmax is shift invariant, so
      // the reference does not have to be the row maximum: m_ref = max of the row's first 32 logits, p = 2^(v - m_ref)
      // for all 576 keys, P in bf16 (the exponent range of fp32), row sum and O in fp32 -- any logit up to ~60
      // octaves above the reference is exact.  With a fixed reference there is no rescale, so the tensor core
      // accumulates O over the six chunks in TMEM and the softmax warps never touch it until the unit is done.
      // Guard: a chunk whose row sum reaches 2^60 (or is inf / NaN) sends the warp through `rescale_chunk` -- exact
      // maximum of the chunk, P recomputed, O and the row sum rescaled in place.  Not taken on image-like data.
      // TMEM loads are software pipelined (sub-block s + 1 in flight while s is processed), the bias comes through
      // ld.shared at immediate offsets, the arithmetic is packed f32x2 (FFMA2 / FADD2: half the issue slots), and P is
      // double buffered so PV(kc) only has to finish before chunk kc + 2 is written.
      for (int kc = 0; kc < NCH; kc++) {
        const int sb = kc & 1;
        const uint32_t s_addr = tmem_base + lane_sel + (uint32_t)(g * 2 + sb) * S_COLS;
        mbar_wait(&s_full[g][sb], (uint32_t)(ui * 3 + (kc >> 1)) & 1u);
        tc_fence_after();
        AT_T(1);
        const uint32_t bchunk = bias_a + (uint32_t)(kc * KROWS * BP) * 4u;
        const uint32_t prow = smem_u32(sm_p) + (uint32_t)((g * 2 + sb) * P_BYTES + row * 64);
        uint32_t ra[32], rb[32];
        tmem_ld32(s_addr, ra);
        tmem_ld_wait();
        AT_T(2);
        tmem_ld32(s_addr + 32, rb);                              // in flight while sub-block 0 is processed
        if (kc == 0) {
          // logits of sub-block 0 in the log2 domain; their maximum is this row's reference
          float smax = -INFINITY;
#pragma unroll
          for (int j = 0; j < 16; j++) {
            const float2 v = ffma2(make_float2(__uint_as_float(ra[2 * j]), __uint_as_float(ra[2 * j + 1])), L2,
                                   lds_f32x2(bchunk + bias_off(2 * j)));
            smax = fmaxf(smax, fmaxf(v.x, v.y));
            ra[2 * j] = __float_as_uint(v.x); ra[2 * j + 1] = __float_as_uint(v.y);
          }
          m_ref = smax;
        }
        const float2 nm = make_float2(-m_ref, -m_ref);
        float2 cs = make_float2(0.f, 0.f);
        // P buffer `sb` was last read by PV(kc - 2) (of the previous unit for kc < 2)
        const int pvn = ui * 3 + (kc >> 1);                      // completions of pv_done[g][sb] needed so far
        if (pvn > 0) mbar_wait(&pv_done[g][sb], (uint32_t)(pvn - 1) & 1u);
        AT_T(3);
        auto emit = [&](const int sub, uint32_t (&r)[32], const bool logits_ready) __attribute__((always_inline)) {
          uint32_t pk[16];
#pragma unroll
          for (int j = 0; j < 16; j++) {
            float2 v = make_float2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
            if (!logits_ready)
              v = ffma2(v, L2, lds_f32x2(bchunk + bias_off(sub * 32 + 2 * j)));
            v = fadd2(v, nm);
            const float2 pp = make_float2(ex2_approx(v.x), ex2_approx(v.y));
            cs = fadd2(cs, pp);
            __nv_bfloat162 h2 = __floats2bfloat162_rn(pp.x, pp.y);
            pk[j] = *reinterpret_cast<uint32_t*>(&h2);
          }
#pragma unroll
          for (int c = 0; c < 4; c++)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(prow + (uint32_t)(sub * 8192) + (uint32_t)((c ^ psw) << 4)),
                         "r"(pk[4 * c]), "r"(pk[4 * c + 1]), "r"(pk[4 * c + 2]), "r"(pk[4 * c + 3]) : "memory");
        };
        if (kc == 0) emit(0, ra, true); else emit(0, ra, false);
        AT_T(4);
        tmem_ld_wait();
        AT_T(5);
        tmem_ld32(s_addr + 64, ra);                              // sub-block 2 in flight while sub-block 1 is processed
        emit(1, rb, false);
        AT_T(4);
        tmem_ld_wait();
        AT_T(5);
        emit(2, ra, false);
        AT_T(4);
        if (__any_sync(0xffffffffu, !(cs.x + cs.y < 1.1529215e18f))) {
          // ---- rescale_chunk (rare): exact maximum, P again, O and the row sum moved to the new reference ----
          float mx = m_ref;
#pragma unroll 1
          for (int sub = 0; sub < 3; sub++) {
            tmem_ld32(s_addr + sub * 32, ra);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j++) mx = fmaxf(mx, fmaf(__uint_as_float(ra[j]), LOG2E, lds_f32(bchunk + bias_off(sub * 32 + j))));
          }
          const float alpha = ex2_approx(m_ref - mx);            // 1 for the rows of the warp that did not overflow
          m_ref = mx;
          cs = make_float2(0.f, 0.f);
#pragma unroll 1
          for (int sub = 0; sub < 3; sub++) {
            tmem_ld32(s_addr + sub * 32, ra);
            tmem_ld_wait();
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; j++) {
              const float p0 = ex2_approx(fmaf(__uint_as_float(ra[2 * j]), LOG2E, lds_f32(bchunk + bias_off(sub * 32 + 2 * j))) - mx);
              const float p1 = ex2_approx(fmaf(__uint_as_float(ra[2 * j + 1]), LOG2E, lds_f32(bchunk + bias_off(sub * 32 + 2 * j + 1))) - mx);
              cs.x += p0; cs.y += p1;
              __nv_bfloat162 h2 = __floats2bfloat162_rn(p0, p1);
              pk[j] = *reinterpret_cast<uint32_t*>(&h2);
            }
#pragma unroll
            for (int c = 0; c < 4; c++)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(prow + (uint32_t)(sub * 8192) + (uint32_t)((c ^ psw) << 4)),
                           "r"(pk[4 * c]), "r"(pk[4 * c + 1]), "r"(pk[4 * c + 2]), "r"(pk[4 * c + 3]) : "memory");
          }
          l2.x *= alpha; l2.y *= alpha;
          if (kc > 0) {                  // O holds chunks 0 .. kc - 1 at the old reference; PV(kc - 1) must have landed
            mbar_wait(&pv_done[g][sb ^ 1], (uint32_t)(ui * 3 + ((kc - 1) >> 1)) & 1u);
            tc_fence_after();
            tmem_ld32(o_addr, ra);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j++) ra[j] = __float_as_uint(__uint_as_float(ra[j]) * alpha);
            tmem_st32(o_addr, ra);
            tmem_st_wait();
          }
        }
        l2 = fadd2(l2, cs);
        tc_fence_before();
        mbar_arrive(&s_empty[g][sb]);                            // the logits are consumed: S(kc + 2) may overwrite them
        fence_proxy_async();            // P (generic-proxy stores) -> visible to the tensor core (async proxy)
        mbar_arrive(&p_full[g][sb]);
        AT_T(6);
        if (kc == 0 && pend) flush();
        AT_T(7);
      }
      mbar_arrive(&q_empty[qb]);        // bias tables of this stage consumed
      // O of this unit is complete once PV(5) has landed; it is picked up after chunk 0 of the next unit (`flush`)
      pend = true; pend_ui = ui; pend_l = l2.x + l2.y;
      pend_row = p.out + unit_off + my_off;
    }
    if (pend) flush();
    AT_T(7);
    if (tr && lane == 0 && (warp == 2 || warp == 6))
      for (int i = 0; i < 10; i++) p.trace[(size_t)blockIdx.x * 64 + (warp == 6 ? 16 : 0) + i] = (unsigned long long)tc[i];
#undef AT_T
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ---- host ------------------------------------------------------------------------------------------
bool attn_tc_supported(const AttnTcArgs& a) {
  if (a.ws != WS || a.ows != OWS || a.head_dim > DP || a.head_dim < 1) return false;
  if (a.H % WS || a.W % WS) return false;
  if ((a.pitch * 2) % 16 || (a.out_pitch * 2) % 16) return false;
  if (a.pitch < 3 * a.heads * DP || a.out_pitch < a.heads * DP) return false;
  if (tc::encode_tiled_fn() == nullptr) return false;
  return true;
}

int launch_attn_tc(const AttnTcArgs& a, cudaStream_t st) {
  HB_REQUIRE(attn_tc_supported(a), HATB200_EUNSUPPORTED, "attn_tc: unsupported shape");
  CUtensorMap tmQ, tmKV;
  uint64_t dims[4] = {(uint64_t)a.pitch, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.B};
  uint64_t str[3] = {(uint64_t)a.pitch * 2, (uint64_t)a.W * a.pitch * 2, (uint64_t)a.H * a.W * a.pitch * 2};
  {   // Q: the x axis split into (x / 2, parity) so that one box is the 8 columns of one parity of a window
    uint64_t qdims[4] = {(uint64_t)a.pitch, 2, (uint64_t)a.W / 2, (uint64_t)a.H * a.B};
    uint64_t qstr[3] = {(uint64_t)a.pitch * 2, (uint64_t)a.pitch * 4, (uint64_t)a.W * a.pitch * 2};
    uint32_t box[4] = {DP, 1, WS / 2, WS};
    HB_TRY(tc::make_tmap_bf16(&tmQ, a.qkv, 4, qdims, qstr, box, DP * 2));
  }
  {
    uint32_t box[4] = {DP, OWS, KROWS, 1};
    HB_TRY(tc::make_tmap_bf16(&tmKV, a.qkv, 4, dims, str, box, DP * 2));
  }
  AttnParams p;
  p.B = a.B; p.H = a.H; p.W = a.W; p.heads = a.heads;
  p.nwx = a.W / WS; p.nwy = a.H / WS;
  p.units = a.B * p.nwx * p.nwy * a.heads;
  p.k_off = a.heads * DP; p.v_off = 2 * a.heads * DP;
  p.bias_l2 = a.bias_l2;
  p.out = (bf16*)a.out; p.out_pitch = a.out_pitch;
  p.trace = nullptr;
#ifdef HATB200_ENGINEERING
  if (getenv("HATB200_ATTN_TRACE")) {       // cycle budget of the softmax warps (tools/trace_attn.py)
    p.trace = gemm_trace_arm(st);
    cudaFuncSetAttribute(attn_tc_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    launch_pdl(attn_tc_kernel<5>, dim3(p.units < 148 ? p.units : 148), dim3(AT_THREADS), SMEM_BYTES, st, tmQ, tmKV, p);
    HB_CHECK_LAUNCH();
    return 0;
  }
#endif
  static bool attr_set = false;
  if (!attr_set) {
    HB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    attr_set = true;
  }
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int grid = p.units < sms ? p.units : sms;
  launch_pdl(attn_tc_kernel<0>, dim3(grid), dim3(AT_THREADS), SMEM_BYTES, st, tmQ, tmKV, p);
  HB_CHECK_LAUNCH();
  return 0;
}

// ---- head-padded packing helpers (pack time / test entry) ---------------------------------------------
// dst[(blk*heads + h)*32 + i][k] = src[blk*C + h*d + i][k] (i < d, k < C), zero elsewhere; dst is (nblk*heads*32, Kpad)
__global__ void pack_rows_headpad_kernel(const void* __restrict__ src, int dt, int C, int heads, int d, int nblk,
                                         bf16* __restrict__ dst, int Kpad, int dst_row0) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)nblk * heads * DP * Kpad;
  if (idx >= total) return;
  int k = (int)(idx % Kpad);
  int r = (int)(idx / Kpad);
  int i = r % DP, h = (r / DP) % heads, blk = r / (DP * heads);
  float v = 0.f;
  if (i < d && k < C) {
    size_t s = ((size_t)blk * C + h * d + i) * C + k;
    v = dt == HATB200_F32 ? ((const float*)src)[s] : (dt == HATB200_BF16 ? to_f(((const bf16*)src)[s]) : to_f(((const __half*)src)[s]));
  }
  dst[((size_t)dst_row0 + r) * Kpad + k] = __float2bfloat16_rn(v);
}
__global__ void pack_vec_headpad_kernel(const void* __restrict__ src, int dt, int C, int heads, int d, int nblk,
                                        float* __restrict__ dst, int dst0) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nblk * heads * DP) return;
  int i = idx % DP, h = (idx / DP) % heads, blk = idx / (DP * heads);
  float v = 0.f;
  if (i < d) {
    size_t s = (size_t)blk * C + h * d + i;
    v = dt == HATB200_F32 ? ((const float*)src)[s] : (dt == HATB200_BF16 ? to_f(((const bf16*)src)[s]) : to_f(((const __half*)src)[s]));
  }
  dst[dst0 + idx] = v;
}
// dst[n][h*32 + i] = src[n][h*d + i] (n < O), zero elsewhere; dst is (Npad, Kpad)
__global__ void pack_cols_headpad_kernel(const void* __restrict__ src, int dt, int O, int C, int heads, int d,
                                         bf16* __restrict__ dst, int Npad, int Kpad) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)Npad * Kpad;
  if (idx >= total) return;
  int k = (int)(idx % Kpad), n = (int)(idx / Kpad);
  int h = k / DP, i = k % DP;
  float v = 0.f;
  if (n < O && h < heads && i < d) {
    size_t s = (size_t)n * C + h * d + i;
    v = dt == HATB200_F32 ? ((const float*)src)[s] : (dt == HATB200_BF16 ? to_f(((const bf16*)src)[s]) : to_f(((const __half*)src)[s]));
  }
  dst[idx] = __float2bfloat16_rn(v);
}
__global__ void bias_l2_kernel(const float* __restrict__ compact, int heads, int rows, float* __restrict__ dst) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= heads * BIAS_FLOATS) return;
  int h = idx / BIAS_FLOATS, i = idx % BIAS_FLOATS;
  int copy = i / ATTN_TC_BIAS_TABLE, j = i % ATTN_TC_BIAS_TABLE + copy;      // copy 1 is shifted by one float
  int r = j / BP, c = j % BP;
  dst[idx] = (r < TN && c < TN) ? compact[(size_t)h * rows + r * TN + c] * LOG2E : 0.f;
}

int launch_pack_rows_headpad(const void* src, int dt, int C, int heads, int nblk, bf16* dst, int Kpad, int dst_row0,
                             cudaStream_t st) {
  size_t total = (size_t)nblk * heads * DP * Kpad;
  pack_rows_headpad_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, dt, C, heads, C / heads, nblk, dst, Kpad,
                                                                           dst_row0);
  HB_CHECK_LAUNCH();
  return 0;
}
int launch_pack_vec_headpad(const void* src, int dt, int C, int heads, int nblk, float* dst, int dst0, cudaStream_t st) {
  int total = nblk * heads * DP;
  pack_vec_headpad_kernel<<<(total + 255) / 256, 256, 0, st>>>(src, dt, C, heads, C / heads, nblk, dst, dst0);
  HB_CHECK_LAUNCH();
  return 0;
}
int launch_pack_cols_headpad(const void* src, int dt, int O, int C, int heads, bf16* dst, int Npad, int Kpad,
                             cudaStream_t st) {
  size_t total = (size_t)Npad * Kpad;
  pack_cols_headpad_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, dt, O, C, heads, C / heads, dst, Npad, Kpad);
  HB_CHECK_LAUNCH();
  return 0;
}
int launch_bias_l2(const float* compact, int heads, int ws, int ows, float* dst, cudaStream_t st) {
  int rows = (ws + ows - 1) * (ws + ows - 1);
  HB_REQUIRE(rows == TN * TN, HATB200_EUNSUPPORTED, "bias_l2: window 16 / overlap window 24 only");
  bias_l2_kernel<<<(heads * BIAS_FLOATS + 255) / 256, 256, 0, st>>>(compact, heads, rows, dst);
  HB_CHECK_LAUNCH();
  return 0;
}

// ---- layout converters for the single-kernel test entry (compact <-> head-padded) ----
__global__ void headpad_qkv_kernel(const bf16* __restrict__ src, int src_pitch, int C, int heads, int d, size_t M,
                                   bf16* __restrict__ dst, int dst_pitch) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t per = (size_t)3 * heads * DP;
  if (idx >= M * per) return;
  size_t m = idx / per;
  int c = (int)(idx % per);
  int i = c % DP, h = (c / DP) % heads, blk = c / (DP * heads);
  dst[m * dst_pitch + c] = i < d ? src[m * src_pitch + blk * C + h * d + i] : __float2bfloat16_rn(0.f);
}
__global__ void headunpad_out_kernel(const bf16* __restrict__ src, int src_pitch, int C, int heads, int d, size_t M,
                                     bf16* __restrict__ dst, int dst_pitch) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= M * C) return;
  size_t m = idx / C;
  int c = (int)(idx % C);
  dst[m * dst_pitch + c] = src[m * src_pitch + (c / d) * DP + c % d];
}
int launch_headpad_qkv(const bf16* src, int src_pitch, int C, int heads, size_t M, bf16* dst, int dst_pitch,
                       cudaStream_t st) {
  size_t total = M * 3 * heads * DP;
  headpad_qkv_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, src_pitch, C, heads, C / heads, M, dst, dst_pitch);
  HB_CHECK_LAUNCH();
  return 0;
}
int launch_headunpad_out(const bf16* src, int src_pitch, int C, int heads, size_t M, bf16* dst, int dst_pitch,
                         cudaStream_t st) {
  size_t total = M * C;
  headunpad_out_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, src_pitch, C, heads, C / heads, M, dst, dst_pitch);
  HB_CHECK_LAUNCH();
  return 0;
}

}  // namespace hb
