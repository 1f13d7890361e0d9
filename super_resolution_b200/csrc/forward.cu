// forward.cu -- hatb200_forward: the HAT forward (hat_arch.py:848-859) as one stream of
// kernel launches over a single channels-last token layout.  No layout copies
// (reference: 691 copy_ calls per forward), no attention mask
// (dead in the reference, hat_arch.py:831), no unfold / softmax materialisation.
//
// Data layout in HBM (M = B*H*W tokens in (b, y, x) order):
//   residual stream  fp32 (M, C)      -- S (conv_first out, long skip), XA (group input), XB (running)
//   activations      T    (M, pitch)  -- T = bf16 (tensor-core path) or fp32 (parity path); pitch = C rounded
//                                        up to 64, pad columns are written as zero (K padding of the consumer)
#include <stdlib.h>

#include "model.cuh"

namespace hb {

struct Dims {
  int B, H, W, M, C, Cp, mid, midp, hid, hp2, hp4, qkvp, h1p, pdim, opdim, hpad;
  int f1, f1p;   // fc1 output channels of a HAB and their pitch: 2 hid (HAT gated FFN) or hid (HATX spatial-gate FFN)
  int dt;      // storage dtype of activations
  size_t es;   // element size of T
};

static Dims make_dims(const hatb200_config& c, int B, int H, int W, int dt) {
  Dims d;
  d.B = B; d.H = H; d.W = W; d.M = B * H * W;
  d.C = c.embed_dim; d.Cp = round_up(d.C, 64);
  d.mid = c.cab_mid; d.midp = round_up(d.mid, 64);
  d.hid = c.ffn_hidden; d.hp2 = round_up(d.hid, 64); d.hp4 = round_up(2 * d.hid, 64);
  d.f1 = c.arch == 1 ? d.hid : 2 * d.hid; d.f1p = round_up(d.f1, 64);
  if (c.arch == 1 && c.use_focus_bias && round_up(c.embed_dim / 4, 64) > d.midp) d.midp = round_up(c.embed_dim / 4, 64);
  d.qkvp = round_up(3 * d.C, 64);
  d.hpad = 0;                                   // tcgen05 attention: heads * 32 channels per q / k / v
  for (int L = 0; L < c.num_layers; L++)
    if (c.num_heads[L] * ATTN_TC_DP > d.hpad) d.hpad = c.num_heads[L] * ATTN_TC_DP;
  if (3 * d.hpad > d.qkvp) d.qkvp = round_up(3 * d.hpad, 64);
  d.h1p = d.hp4 > d.qkvp ? d.hp4 : d.qkvp;
  d.pdim = c.esc_pdim; d.opdim = c.ocab_esc_enable ? c.ocab_esc_pdim : 0;
  d.dt = dt; d.es = dt == HATB200_F32 ? 4 : 2;
  return d;
}

static void plan(const hatb200_config& c, const Dims& d, char* base, Workspace& w) {
  size_t off = 0;
  auto take = [&](size_t bytes) -> void* {
    void* p = base ? base + off : nullptr;
    off += (bytes + 255) & ~(size_t)255;
    return p;
  };
  size_t M = (size_t)d.M;
  w.S = (float*)take(M * d.C * 4);
  w.XA = (float*)take(M * d.C * 4);
  w.XB = (float*)take(M * d.C * 4);
  w.XN = take(M * d.Cp * d.es);
  w.XN2 = take(M * d.Cp * d.es);
  w.U = take(M * d.Cp * d.es);
  int pd = d.pdim > d.opdim ? d.pdim : d.opdim;
  w.X16 = take(M * pd * d.es);
  w.X16P = take((size_t)d.B * d.H * (d.W + 12) * 16 * 2);
  w.T1 = take(M * d.midp * d.es);
  w.H1 = take(M * d.h1p * d.es);
  size_t gp = d.hp2 > d.Cp ? d.hp2 : d.Cp;
  if ((size_t)round_up(d.hpad, 64) > gp) gp = round_up(d.hpad, 64);
  w.G = take(M * gp * d.es);
  if (c.upsampler == 1) {
    int s = c.upscale;
    w.F64 = take(M * 64 * d.es);
    if (s == 3) w.UP1 = take(M * 9 * 64 * d.es);
    if (s == 2 || s == 4) w.UP1 = take(M * 4 * 64 * d.es);
    if (s == 4) w.UP2 = take(M * 16 * 64 * d.es);
    w.CL32 = d.dt == HATB200_BF16 ? (float*)take(M * s * s * 4 * sizeof(float)) : nullptr;   // conv_last, pixels-last fp32
  }
  int nchunk = colsum_chunks(d.H * d.W);
  w.part = (float*)take((size_t)d.B * nchunk * d.C * 4);
  w.part16 = (float*)take((size_t)d.B * nchunk * 64 * 4);
  w.eca_s = (float*)take((size_t)d.B * d.C * 4);
  w.dyn = (float*)take((size_t)d.B * 64 * 9 * 4);
  if (c.arch == 1 && (c.use_focus_bias || c.kv_topk_keep > 0)) {
    const int ows = c.overlap_win_size, ws = c.window_size;
    w.SAL = take(M * 16 * d.es);
    w.KN = (float*)take(M * 4);
    w.KADD = (float*)take((size_t)d.B * (d.H / ws) * (d.W / ws) * ows * ows * 4);
  }
  w.bytes = off;
}

struct Fwd {
  hatb200_handle* h;
  Dims d;
  Workspace w;
  cudaStream_t st;

  bool tc() const { return h->use_tc && d.dt == HATB200_BF16; }

  // op timing (bench roofline breakdown): CUDA events around each op on the launching stream
  int t_begin(const char* tag, double flops, double bytes, const char* kernel = nullptr) {
    if (!h->op_timing) return -1;
    while (h->ev_pool.size() < h->ev_used + 2) {
      cudaEvent_t e;
      if (cudaEventCreate(&e) != cudaSuccess) return -1;
      h->ev_pool.push_back(e);
    }
    OpRec r{tag, kernel, h->ev_pool[h->ev_used], h->ev_pool[h->ev_used + 1], flops, bytes};
    h->ev_used += 2;
    cudaEventRecord(r.e0, st);
    h->recs.push_back(r);
    return (int)h->recs.size() - 1;
  }
  void t_end(int id) { if (id >= 0) cudaEventRecord(h->recs[id].e1, st); }

  // the bf16 path met a shape no tcgen05 kernel takes: counted (hatb200_last_fallback_count), an error under
  // the "strict_tc" option -- never silent (the CUDA-core kernels are 10-20x slower)
  int fallback(const char* tag, const char* kernel) {
    h->last_fallbacks++;
    static const bool trace = getenv("HATB200_TRACE") != nullptr;
    if (trace) fprintf(stderr, "[hatb200] %s: no tcgen05 kernel takes this shape -> %s\n", tag, kernel);
    HB_REQUIRE(!h->strict_tc, HATB200_EUNSUPPORTED, "strict_tc: '%s' would run on the CUDA-core kernel %s", tag, kernel);
    return 0;
  }

  // ALGORITHMIC bytes of one fused conv / GEMM launch: every stream the epilogue reads or writes, each once
  double conv_bytes(const ConvW& cw, double Mt, const Epilogue& ep) const {
    const double fo = cw.algoO ? cw.algoO : cw.O, fi = cw.algoI ? cw.algoI : cw.I;
    double b = Mt * fi * d.es + (double)cw.taps() * fo * fi * d.es;          // activations in, weights
    if (ep.out_t) b += Mt * fo * d.es * (ep.shuffle_r > 1 ? 1.0 : 1.0);
    if (ep.out_f32) b += Mt * fo * 4;
    if (ep.resid) b += Mt * fo * 4;
    if (ep.eca_u) b += Mt * fo * d.es;
    if (ep.ln_out) b += Mt * fo * d.es;
    if (ep.ln_head) b += Mt * ep.ln_head_c * d.es * (ep.ln_head_pad ? 2 : 1);
    return b;
  }
  const char* conv_kernel(const ConvW& cw, const ConvArgs& a, const Epilogue& ep, const char* tag) const {
    if (tc() && !tc_disabled(tag)) {
      if (gemm_tc_supported(a, cw.Npad, cw.Kpad, ep)) return "gemm_tc";
      if (conv3_tc_supported(a, cw.Npad, cw.Kpad, ep)) return "conv3_tc";
      if (tapconv_tc_supported(a, cw.Npad, cw.Kpad, ep)) return "tapconv_tc";
    }
    return "tapconv_simt";
  }

  int dbg(const char* what) {
    if (!h->sync_debug) return 0;
    static const bool trace = getenv("HATB200_TRACE") != nullptr;
    if (trace) fprintf(stderr, "[hatb200] sync after %s\n", what);
    cudaError_t e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
      set_error("forward: '%s' failed: %s", what, cudaGetErrorString(e));
      return HATB200_ECUDA;
    }
    return 0;
  }

  // dense conv / linear over tokens; tensor cores when available for this shape
  int conv(const char* tag, const ConvW& cw, const void* A, int a_pitch, int a_coff, int H, int W, Epilogue ep) {
    double Mt = (double)d.B * H * W;
    const double fo = cw.algoO ? cw.algoO : cw.O, fi = cw.algoI ? cw.algoI : cw.I;
    int tid = -1;
    if (h->op_timing) {
      ConvArgs a{A, a_pitch, a_coff, cw.I, d.B, H, W, cw.KH, cw.KW, cw.O};
      Epilogue e2 = ep;
      if (!e2.bias) e2.bias = cw.bias;
      tid = t_begin(tag, 2.0 * Mt * fo * fi * cw.taps(), conv_bytes(cw, Mt, ep), conv_kernel(cw, a, e2, tag));
    }
    int r = conv_impl(tag, cw, A, a_pitch, a_coff, H, W, ep);
    t_end(tid);
    return r;
  }

  // conv + PixelShuffle (hat_arch.py:593-605): the persistent conv3 kernel with (i, j, c)-ordered weight rows when it
  // takes the shape (each 64-column block = one sub-pixel = one TMA store), else the generic route
  int conv_shuffle(const char* tag, const ConvW& cw, const void* A, int H, int W, Epilogue ep) {
    if (tc() && cw.w16s && cw.shuffle_r == ep.shuffle_r && !tc_disabled(tag)) {
      ConvArgs a{A, 64, 0, cw.I, d.B, H, W, cw.KH, cw.KW, cw.O};
      Epilogue e2 = ep;
      e2.shuffle_perm = 1; e2.bias = cw.bias_s;
      if (conv3_tc_supported(a, cw.Npad, cw.Kpad, e2)) {
        double Mt = (double)d.B * H * W;
        int tid = t_begin(tag, 2.0 * Mt * cw.O * cw.I * cw.taps(), conv_bytes(cw, Mt, ep), "conv3_tc");
        int r = launch_conv3_tc(a, cw.w16s, cw.Npad, cw.Kpad, e2, st);
        t_end(tid);
        HB_TRY(r);
        return dbg("conv3_tc shuffle");
      }
    }
    return conv(tag, cw, A, 64, 0, H, W, ep);
  }

  // The fp32 residual stream of the blocks (XB: M x C x 4 bytes, read and rewritten in place by esc_aggr and fc2 of every
  // block) is kept in L2 with the evict_last policy when it is comfortably smaller than L2 (47 MB of 126 MB at the
  // bench shape); larger streams would only thrash among themselves.
  int pin_stream() const {
    static const int mode = getenv("HATB200_L2_PIN") ? atoi(getenv("HATB200_L2_PIN")) : 1;
    return mode && (size_t)d.M * d.C * 4 <= ((size_t)56 << 20);
  }

  // will conv(tag, ...) run on the persistent 3x3 kernel in a way that can carry a side job (side_jobs.cuh)?
  bool side_ok(const char* tag, const ConvW& cw, const void* A, int a_pitch, int H, int W, Epilogue ep) const {
    static const bool off = getenv("HATB200_SIDE_JOBS") && getenv("HATB200_SIDE_JOBS")[0] == '0';     // engineering: A/B
    if (off || !tc() || tc_disabled(tag)) return false;
    ConvArgs a{A, a_pitch, 0, cw.I, d.B, H, W, cw.KH, cw.KW, cw.O};
    if (!ep.bias) ep.bias = cw.bias;
    return !gemm_tc_supported(a, cw.Npad, cw.Kpad, ep) && conv3_tc_side_ok(a, cw.Npad, cw.Kpad, ep);
  }

  // engineering switch: HATB200_TC_DISABLE=tag1,tag2 keeps the named ops on the CUDA-core kernel (bisecting)
  static bool tc_disabled(const char* tag) {
    const char* v = getenv("HATB200_TC_DISABLE");
    if (!v || !*v) return false;
    std::string list = std::string(",") + v + ",";
    return list.find(std::string(",") + tag + ",") != std::string::npos;
  }

  int conv_impl(const char* tag, const ConvW& cw, const void* A, int a_pitch, int a_coff, int H, int W, Epilogue ep) {
    ConvArgs a{A, a_pitch, a_coff, cw.I, d.B, H, W, cw.KH, cw.KW, cw.O};
    if (!ep.bias) ep.bias = cw.bias;
    static const bool trace = getenv("HATB200_TRACE") != nullptr;
    if (trace)
      fprintf(stderr, "[hatb200] %s: N=%d Cin=%d k=%d Npad=%d Kpad=%d gemm=%d tap=%d\n", tag, cw.O, cw.I, cw.KH, cw.Npad,
              cw.Kpad, (int)(tc() && gemm_tc_supported(a, cw.Npad, cw.Kpad, ep)),
              (int)(tc() && tapconv_tc_supported(a, cw.Npad, cw.Kpad, ep)));
    if (tc() && !tc_disabled(tag) && conv_tc_supported(a, cw.Npad, cw.Kpad, ep)) {
      HB_TRY(launch_conv_tc(a, cw.w16, cw.Npad, cw.Kpad, ep, st));
      return dbg("tapconv_tc");
    }
    // CUDA-core path: LayerNorm / column sums are separate kernels
    if (tc()) HB_TRY(fallback(tag, "tapconv_simt"));
    Epilogue e2 = ep;
    e2.ln_gamma = e2.ln_beta = nullptr; e2.ln_out = e2.ln_head = nullptr;
    e2.ln_head_colsum = nullptr; e2.colsum = nullptr;
    HB_TRY(launch_tapconv_simt(d.dt, a, cw.w32, cw.Np, e2, st));
    HB_TRY(dbg("tapconv_simt"));
    int hw = H * W;
    if (ep.ln_out) {
      HB_REQUIRE(ep.out_f32 != nullptr, HATB200_EINVAL, "conv: fused LN needs the fp32 stream output");
      HB_TRY(launch_layernorm(ep.out_f32, d.B * hw, cw.O, ep.ln_gamma, ep.ln_beta, ep.ln_out, d.dt, ep.ln_pitch,
                              ep.ln_pitch, ep.ln_head, ep.ln_head_c, st));
      if (ep.ln_head_colsum)
        HB_TRY(launch_colsum(ep.ln_head, d.dt, ep.ln_head_c, d.B, hw, ep.ln_head_c, ep.ln_head_colsum, st));
      if (ep.ln_head_pad)     // the x-padded copy the Toeplitz 13x13 kernel reads (written by the fused epilogue otherwise)
        HB_TRY(launch_plk_pad_copy((const bf16*)ep.ln_head, d.B, H, W, (bf16*)ep.ln_head_pad, st));
    }
    if (ep.colsum) HB_TRY(launch_colsum(ep.out_t, d.dt, ep.out_t_pitch, d.B, hw, cw.O, ep.colsum, st));
    return dbg("conv tail");
  }

  // standalone LayerNorm of the fp32 stream (+ compact head and its column sums)
  int ln(const float* x, const LNW& n, void* out, int head_c) {
    int tid = t_begin("layernorm", 8.0 * d.M * d.C, (double)d.M * d.C * (4 + d.es), "layernorm");
    int r = ln_impl(x, n, out, head_c);
    t_end(tid);
    return r;
  }
  bool plk_tc(int pdim) const {
    return tc() && !tc_disabled("esc_plk13") && plk_tc_supported(pdim, h->cfg.esc_kernel, d.H, d.W, d.Cp);
  }
  int ln_impl(const float* x, const LNW& n, void* out, int head_c) {
    HB_TRY(launch_layernorm(x, d.M, d.C, n.g, n.b, out, d.dt, d.Cp, d.Cp, head_c ? w.X16 : nullptr, head_c, st));
    if (head_c) HB_TRY(launch_colsum(w.X16, d.dt, head_c, d.B, d.H * d.W, head_c, w.part16, st));
    if (head_c && plk_tc(head_c)) HB_TRY(launch_plk_pad_copy((const bf16*)w.X16, d.B, d.H, d.W, (bf16*)w.X16P, st));
    return dbg("layernorm");
  }

  void set_ln(Epilogue& ep, const LNW* n, void* out, int head_c) {
    if (!n) return;
    ep.ln_gamma = n->g; ep.ln_beta = n->b; ep.ln_out = out; ep.ln_pitch = d.Cp;
    if (head_c && !(tc() && head_c > 16)) {      // wider heads (HATX configs: esc_pdim 24 / 32) are extracted by head_after()
      ep.ln_head = w.X16; ep.ln_head_c = head_c; ep.ln_head_colsum = w.part16;
      if (plk_tc(head_c)) ep.ln_head_pad = w.X16P;
    }
  }

  // The fused epilogues carry a compact head of at most 16 channels; for a wider ESC head the LayerNorm output is
  // copied out and pooled by two small kernels right after the fused layer.
  int head_after(int head_c) {
    if (!(tc() && head_c > 16)) return 0;
    HB_TRY(launch_head_copy(w.XN, d.dt, d.Cp, (size_t)d.M, head_c, w.X16, st));
    HB_TRY(launch_colsum(w.X16, d.dt, head_c, d.B, d.H * d.W, head_c, w.part16, st));
    return dbg("head_after");
  }

  // ESC ConvAttn core on xn (esc_arch.py:119-123): writes the mixed first-pdim channels into
  // dst[:, :pdim] (dst already holds xn elsewhere).  Needs X16 + part16 from the preceding LN.
  // `side`: a per-sample job for extra CTAs of the 13x13 kernel (only passed when that kernel runs, see hab())
  int esc_mix(const EscW& e, void* dst, bool dyn_done = false, const SideJob* side = nullptr) {
    int nchunk = colsum_chunks(d.H * d.W);
    if (!dyn_done) HB_TRY(launch_dyn_finish(w.part16, nchunk, d.B, e.pdim, d.H * d.W, e.w1, e.b1, e.w2, e.b2, w.dyn, st));
    if (e.plk_wt && plk_tc(e.pdim)) {
      double Mt = (double)d.M;
      int tid = t_begin("esc_plk13", 2.0 * Mt * e.pdim * e.pdim * (e.plk.taps() + 9.0 / e.pdim), Mt * 2 * e.pdim * d.es, "plk_tc");
      int r = launch_plk_tc((const bf16*)w.X16P, d.B, d.H, d.W, e.plk_wt, w.dyn, (bf16*)dst, d.Cp, st, side);
      t_end(tid);
      HB_TRY(r);
      return dbg("plk_tc");
    }
    Epilogue ep;
    ep.dyn_x = w.X16; ep.dyn_pitch = e.pdim; ep.dyn_k = w.dyn;
    ep.out_t = dst; ep.out_t_pitch = d.Cp; ep.out_t_coff = 0;
    return conv("esc_plk13", e.plk, w.X16, e.pdim, 0, d.H, d.W, ep);
  }

  // HAB.forward (hat_arch.py:217-238).  Entry: XN = norm1(src) (+X16, part16).  Exit: dst = block output
  // and XN = next_ln(dst) (+X16) when next_ln is given.
  int hab(const HabW& b, const float* src, float* dst, const LNW* next_ln, int next_head) {
    const hatb200_config& c = h->cfg;
    int nchunk = colsum_chunks(d.H * d.W);
    // The two per-sample "finish" steps (ESC dynamic kernel from the LN head's pool partials, ECA gate from conv2's)
    // ride as extra CTAs of kernels that run anyway: dyn_finish in cab_conv1's tail (its inputs exist since the previous
    // block), eca_finish on the SMs the 13x13 kernel leaves idle (the gate is only needed by esc_aggr after it).
    // Whatever cannot ride (CUDA-core routes, other shapes) keeps its own launch.
    const float inv_hw = 1.0f / (float)(d.H * d.W);
    SideJob dynj; dynj.kind = 2; dynj.B = d.B; dynj.nchunk = nchunk; dynj.C = b.esc.pdim; dynj.inv_hw = inv_hw;
    dynj.partial = w.part16; dynj.w1 = b.esc.w1; dynj.b1 = b.esc.b1; dynj.w2 = b.esc.w2; dynj.b2 = b.esc.b2; dynj.out = w.dyn;
    SideJob ecaj; ecaj.kind = 1; ecaj.B = d.B; ecaj.nchunk = nchunk; ecaj.C = d.C; ecaj.inv_hw = inv_hw;
    ecaj.partial = w.part; ecaj.w1 = b.eca_w5; ecaj.out = w.eca_s;
    bool dyn_rides = false;
    {  // CAB conv1 + GELU
      Epilogue ep; ep.act = ACT_GELU; ep.out_t = w.T1; ep.out_t_pitch = d.midp; ep.out_t_zero_to = d.midp;
      dyn_rides = b.esc.pdim <= 64 && b.esc.pdim % 2 == 0 && side_ok("cab_conv1", b.cab0, w.XN, d.Cp, d.H, d.W, ep);
      if (dyn_rides) ep.side = dynj;
      HB_TRY(conv("cab_conv1", b.cab0, w.XN, d.Cp, 0, d.H, d.W, ep));
    }
    static const bool side_off = getenv("HATB200_SIDE_JOBS") && getenv("HATB200_SIDE_JOBS")[0] == '0';
    const bool eca_rides = !side_off && tc() && b.esc.plk_wt && plk_tc(b.esc.pdim);
    {  // CAB conv2 -> U, with GAP partials for ECA
      Epilogue ep; ep.out_t = w.U; ep.out_t_pitch = d.Cp; ep.colsum = w.part;
      HB_TRY(conv("cab_conv2", b.cab2, w.T1, d.midp, 0, d.H, d.W, ep));
      if (!dyn_rides && !eca_rides)      // both in one launch of their own
        HB_TRY(launch_hab_finish(w.part, nchunk, d.B, d.C, d.H * d.W, b.eca_w5, w.eca_s, w.part16, b.esc.pdim, b.esc.w1,
                                 b.esc.b1, b.esc.w2, b.esc.b2, w.dyn, st));
      else if (!dyn_rides)
        HB_TRY(launch_dyn_finish(w.part16, nchunk, d.B, b.esc.pdim, d.H * d.W, b.esc.w1, b.esc.b1, b.esc.w2, b.esc.b2, w.dyn, st));
      else if (!eca_rides)
        HB_TRY(launch_eca_finish(w.part, nchunk, d.B, d.C, d.H * d.W, b.eca_w5, w.eca_s, st));
    }
    HB_TRY(esc_mix(b.esc, w.XN, true, eca_rides ? &ecaj : nullptr));   // XN[:, :pdim] <- plk conv + dynamic dw (CAB has consumed XN)
    {  // aggr 1x1 + HAB combine (hat_arch.py:236) [+ norm2]
      Epilogue ep;
      ep.eca_u = w.U; ep.eca_pitch = d.Cp; ep.eca_scale = w.eca_s; ep.eca_coef = c.conv_scale;
      ep.resid = src; ep.resid_pitch = d.C; ep.out_f32 = dst; ep.out_f32_pitch = d.C;
      set_ln(ep, &b.n2, w.XN2, 0);
      // backwards: the fp32 stream `src` was written front to back by the previous block three launches ago, so its
      // tail is what L2 still holds; starting there saves HBM reads that a front-to-back walk would evict first
      static const bool fwd_only = getenv("HATB200_WALK_BACK") && getenv("HATB200_WALK_BACK")[0] == '0';
      ep.walk_back = fwd_only ? 0 : 1;
      ep.pin_stream = pin_stream();
      HB_TRY(conv("esc_aggr", b.esc.aggr, w.XN, d.Cp, 0, d.H, d.W, ep));
    }
    if (c.arch == 1) {  // HATX SpatialGateDConvFFN (hatx_arch.py:160-180): fc1 -> [dw3x3(xa) * SiLU(xb) | xb] -> fc2
      Epilogue ep; ep.out_t = w.H1; ep.out_t_pitch = d.f1p;
      HB_TRY(conv("ffn_fc1", b.fc1, w.XN2, d.Cp, 0, d.H, d.W, ep));
      int tid = t_begin("ffn_spatial_gate", 2.0 * d.M * (d.hid / 2) * 9, (double)d.M * 2 * d.hid * d.es, "sgate");
      HB_TRY(launch_sgate(w.H1, d.dt, d.f1p, d.B, d.H, d.W, d.hid / 2, b.dw_w, b.dw_b, w.G, d.hp2, d.hp2, st));
      t_end(tid);
      HB_TRY(dbg("sgate"));
      Epilogue e2; e2.resid = dst; e2.resid_pitch = d.C; e2.out_f32 = dst; e2.out_f32_pitch = d.C;
      set_ln(e2, next_ln, w.XN, next_head);
      HB_TRY(conv("ffn_fc2", b.fc2, w.G, d.hp2, 0, d.H, d.W, e2));
      HB_TRY(head_after(next_ln ? next_head : 0));
    } else {  // GatedDconvFFN (hat_arch.py:107-119)
      Epilogue ep; ep.out_t = w.H1; ep.out_t_pitch = d.hp4;
      HB_TRY(conv("ffn_fc1", b.fc1, w.XN2, d.Cp, 0, d.H, d.W, ep));
      Epilogue e2; e2.resid = dst; e2.resid_pitch = d.C; e2.out_f32 = dst; e2.out_f32_pitch = d.C;
      set_ln(e2, next_ln, w.XN, next_head);
      e2.pin_stream = pin_stream();
      // depthwise 3x3 + gate + fc2 + residual + LayerNorm in one persistent tcgen05 kernel: the gated middle G never
      // reaches HBM (tc_ffn.cu)
      FfnTailArgs fa{w.H1, d.hp4, d.B, d.H, d.W, d.hid, b.dw_w, b.dw_b, b.fc2.w16, b.fc2.O, b.fc2.Npad, b.fc2.Kpad};
      Epilogue e3 = e2; e3.bias = b.fc2.bias;
      // Measured on B200 (profiles/r02_ffn_fused_trace.txt): 88-94 us per launch against 39 + 43 us for dwgate + fc2 --
      // the depthwise stage is FMA-pipe bound (17-20 us per 128-pixel tile whether 8 or 16 warps run it) and shares its
      // issue slots with the epilogue warps, so the fused kernel is opt-in (HATB200_FFN_FUSED=1), not the default.
      static const bool ffn_fused = getenv("HATB200_FFN_FUSED") && getenv("HATB200_FFN_FUSED")[0] == '1';
      if (ffn_fused && tc() && !tc_disabled("ffn_dw_fc2") && ffn2_tc_supported(fa, e3)) {
        const double Mt = (double)d.M;
        int tid = t_begin("ffn_dw_fc2", 2.0 * Mt * (2.0 * d.hid * 9 + (double)d.hid * d.C),
                          Mt * (2.0 * d.hid * d.es) + conv_bytes(b.fc2, Mt, e2) - Mt * d.hid * d.es, "ffn_tc");
        int r = launch_ffn2_tc(fa, e3, st);
        t_end(tid);
        HB_TRY(r);
        HB_TRY(head_after(next_ln ? next_head : 0));
        return dbg("ffn2_tc");
      }
      int tid = t_begin("ffn_dwconv_gate", 2.0 * d.M * 2 * d.hid * 9, (double)d.M * 3 * d.hid * d.es, "dwgate");
      HB_TRY(launch_dwgate(w.H1, d.dt, d.hp4, d.B, d.H, d.W, d.hid, b.dw_w, b.dw_b, w.G, d.hp2, d.hp2, st));
      t_end(tid);
      HB_TRY(dbg("dwgate"));
      HB_TRY(conv("ffn_fc2", b.fc2, w.G, d.hp2, 0, d.H, d.W, e2));
      HB_TRY(head_after(next_ln ? next_head : 0));
    }
    return 0;
  }

  // OCAB.forward (hat_arch.py:326-393).  Entry: XN = norm1(x) (+X16 when esc).  x updated in place.
  // Exit: XN2 = T copy of x (input of the RHAG conv; XN receives that conv's fused LN) when want_copy.
  int ocab(const OcabW& o, float* x, bool want_copy) {
    const hatb200_config& c = h->cfg;
    int dh = d.C / o.heads;
    float scale = c.qk_scale > 0.f ? c.qk_scale : 1.0f / sqrtf((float)dh);
    // tcgen05 attention (bf16 path, default geometry): head-padded qkv / proj weights
    const int hp = o.heads * ATTN_TC_DP;
    AttnTcArgs ta{w.H1, 3 * hp, d.B, d.H, d.W, o.heads, dh, c.window_size, c.overlap_win_size, o.bias_l2, w.G, hp};
    // HATX focus bias / top-k pruning (hatx_arch.py:405-441) live in the CUDA-core attention kernel
    const bool hatx_opts = c.arch == 1 && (c.use_focus_bias || c.kv_topk_keep > 0);
    const bool use_attn_tc = tc() && !hatx_opts && o.bias_l2 != nullptr && !tc_disabled("ocab_attention") &&
                             attn_tc_supported(ta);
    if (c.ocab_esc_enable) {   // K/V from y = ConvAttnWrapper(xn) (hat_arch.py:336-344)
      HB_CHECK_CUDA(cudaMemcpyAsync(w.XN2, w.XN, (size_t)d.M * d.Cp * d.es, cudaMemcpyDeviceToDevice, st));
      HB_TRY(esc_mix(o.esc, w.XN2));
      Epilogue ey; ey.out_t = w.U; ey.out_t_pitch = d.Cp; ey.out_t_zero_to = d.Cp;
      HB_TRY(conv("esc_aggr", o.esc.aggr, w.XN2, d.Cp, 0, d.H, d.W, ey));
      if (use_attn_tc) {       // head-padded layout of the tcgen05 attention: q at [0, hp), k / v at [hp, 3 hp)
        Epilogue eq; eq.out_scale = scale; eq.scale_ncols = hp; eq.out_t = w.H1; eq.out_t_pitch = 3 * hp;
        HB_TRY(conv("ocab_qkv", o.q_tc, w.XN, d.Cp, 0, d.H, d.W, eq));
        Epilogue ek; ek.out_t = w.H1; ek.out_t_pitch = 3 * hp; ek.out_t_coff = hp;
        HB_TRY(conv("ocab_qkv", o.kv_tc, w.U, d.Cp, 0, d.H, d.W, ek));
      } else {
        Epilogue eq; eq.out_scale = scale; eq.scale_ncols = d.C; eq.out_t = w.H1; eq.out_t_pitch = d.qkvp;
        HB_TRY(conv("ocab_qkv", o.q, w.XN, d.Cp, 0, d.H, d.W, eq));
        Epilogue ek; ek.out_t = w.H1; ek.out_t_pitch = d.qkvp; ek.out_t_coff = d.C;
        HB_TRY(conv("ocab_qkv", o.kv, w.U, d.Cp, 0, d.H, d.W, ek));
      }
    } else if (use_attn_tc) {
      Epilogue ep; ep.out_scale = scale; ep.scale_ncols = hp; ep.out_t = w.H1; ep.out_t_pitch = 3 * hp;
      HB_TRY(conv("ocab_qkv", o.qkv_tc, w.XN, d.Cp, 0, d.H, d.W, ep));
    } else {
      Epilogue ep; ep.out_scale = scale; ep.scale_ncols = d.C; ep.out_t = w.H1; ep.out_t_pitch = d.qkvp;
      HB_TRY(conv("ocab_qkv", o.qkv, w.XN, d.Cp, 0, d.H, d.W, ep));
    }
    {
      double nk = (double)c.overlap_win_size * c.overlap_win_size;
      int tid = t_begin("ocab_attention", 4.0 * d.M * nk * d.C, (double)d.M * 4 * d.C * d.es,
                        use_attn_tc ? "attn_tc" : "attn_simt");
      if (tc() && !use_attn_tc) HB_TRY(fallback("ocab_attention", "attn_simt"));
      if (use_attn_tc) {
        HB_TRY(launch_attn_tc(ta, st));
      } else {
        AttnArgs a{w.H1, d.qkvp, d.B, d.H, d.W, d.C, o.heads, c.window_size, c.overlap_win_size, o.bias_compact,
                   w.G, d.Cp, d.Cp};
        a.pad = c.ocab_pad;
        if (hatx_opts) {
          const void* ysrc = c.ocab_esc_enable ? w.U : w.XN;          // y: the tensor K / V were projected from
          const void* sal = nullptr;
          if (c.use_focus_bias) {      // focus_head(y): 1x1 C -> C/4, GELU, 1x1 -> 1 (hatx_arch.py:355-361, :408)
            Epilogue e0; e0.act = ACT_GELU; e0.out_t = w.T1; e0.out_t_pitch = d.midp; e0.out_t_zero_to = d.midp;
            HB_TRY(conv_impl("ocab_focus", o.focus0, ysrc, d.Cp, 0, d.H, d.W, e0));
            Epilogue e1; e1.out_t = w.SAL; e1.out_t_pitch = 16;
            HB_TRY(conv_impl("ocab_focus", o.focus1, w.T1, d.midp, 0, d.H, d.W, e1));
            sal = w.SAL;
          } else {                     // score = ||k||_2 over all C key channels (hatx_arch.py:426)
            HB_TRY(launch_knorm(w.H1, d.dt, d.qkvp, d.C, d.C, (size_t)d.M, w.KN, st));
          }
          HB_TRY(launch_ocab_focus_prep(sal, d.dt, 16, sal ? nullptr : w.KN, d.B, d.H, d.W, c.window_size,
                                        c.overlap_win_size, c.ocab_pad, c.kv_topk_keep, w.KADD, st));
          a.kadd = w.KADD;
        }
        HB_TRY(launch_attn_simt(d.dt, a, st));
      }
      t_end(tid);
    }
    HB_TRY(dbg("attention"));
    {  // proj + shortcut, norm2
      Epilogue ep; ep.resid = x; ep.resid_pitch = d.C; ep.out_f32 = x; ep.out_f32_pitch = d.C;
      ep.pin_stream = pin_stream();
      set_ln(ep, &o.n2, w.XN2, 0);
      if (use_attn_tc) HB_TRY(conv("ocab_proj", o.proj_tc, w.G, hp, 0, d.H, d.W, ep));
      else HB_TRY(conv("ocab_proj", o.proj, w.G, d.Cp, 0, d.H, d.W, ep));
    }
    {  // MLP (hat_arch.py:309-313)
      Epilogue ep; ep.act = ACT_GELU; ep.out_t = w.H1; ep.out_t_pitch = d.hp2; ep.out_t_zero_to = d.hp2;
      HB_TRY(conv("ocab_mlp0", o.mlp0, w.XN2, d.Cp, 0, d.H, d.W, ep));
      Epilogue e2; e2.resid = x; e2.resid_pitch = d.C; e2.out_f32 = x; e2.out_f32_pitch = d.C;
      e2.pin_stream = pin_stream();
      if (want_copy) { e2.out_t = w.XN2; e2.out_t_pitch = d.Cp; e2.out_t_zero_to = d.Cp; }
      HB_TRY(conv("ocab_mlp2", o.mlp2, w.H1, d.hp2, 0, d.H, d.W, e2));
    }
    return 0;
  }

  // features_only: HAT.forward_features (hat_arch.py:828-846) -- x is the (B, C, H, W) output of conv_first, y the
  // (B, C, H, W) output of the final LayerNorm; everything between is the same launch stream as the full forward
  int run(const void* x, int x_dt, void* y, int y_dt, bool features_only = false) {
    const hatb200_config& c = h->cfg;
    if (c.upsampler != 1 && !features_only)   // hat_arch.py:852: no tail -> only the (de)normalisation remains
      return launch_affine_roundtrip(x, x_dt, d.B, c.in_chans, d.H * d.W, c.mean, c.img_range, y, y_dt, st);
    if (tc()) HB_CHECK_CUDA(cudaMemsetAsync(w.X16P, 0, (size_t)d.B * d.H * (d.W + 12) * 16 * 2, st));   // x pads stay zero
    if (features_only) {
      HB_TRY(launch_nchw_to_tokens(x, x_dt, d.B, d.C, d.H * d.W, w.S, st));
    } else {
      HB_TRY(launch_conv_first(x, x_dt, d.B, c.in_chans, d.H, d.W, h->conv_first_w, h->conv_first_b, d.C, c.mean,
                               c.img_range, w.S, st));
    }
    HB_TRY(dbg("conv_first"));
    // patch_embed (+norm) -> XA: the running stream at the first group input (hat_arch.py:836)
    if (c.patch_norm) {
      HB_TRY(launch_layernorm(w.S, d.M, d.C, h->pe_norm.g, h->pe_norm.b, w.XA, HATB200_F32, d.C, d.C, nullptr, 0, st));
    } else {
      HB_CHECK_CUDA(cudaMemcpyAsync(w.XA, w.S, (size_t)d.M * d.C * 4, cudaMemcpyDeviceToDevice, st));
    }
    // first LN of the first block
    auto first_ln_of = [&](int L) -> const LNW* {
      const LayerW& lw = h->layers[L];
      return lw.habs.empty() ? &lw.ocab.n1 : &lw.habs[0].n1;
    };
    auto first_head_of = [&](int L) -> int {
      const LayerW& lw = h->layers[L];
      return lw.habs.empty() ? d.opdim : d.pdim;
    };
    HB_TRY(ln(w.XA, *first_ln_of(0), w.XN, first_head_of(0)));
    for (int L = 0; L < c.num_layers; L++) {
      const LayerW& lw = h->layers[L];
      int nb = (int)lw.habs.size();
      const float* src = w.XA;       // RHAG input stays in XA; blocks run in XB
      for (int i = 0; i < nb; i++) {
        const LNW* nl = i + 1 < nb ? &lw.habs[i + 1].n1 : &lw.ocab.n1;
        int nh = i + 1 < nb ? d.pdim : d.opdim;
        HB_TRY(hab(lw.habs[i], src, w.XB, nl, nh));
        src = w.XB;
      }
      if (nb == 0) HB_CHECK_CUDA(cudaMemcpyAsync(w.XB, w.XA, (size_t)d.M * d.C * 4, cudaMemcpyDeviceToDevice, st));
      HB_TRY(ocab(lw.ocab, w.XB, c.resi_conv != 0));
      // RHAG tail (hat_arch.py:555-556): XA = XA + conv(XB)  [+ next LN]
      bool last = L + 1 == c.num_layers;
      const LNW* nl = last ? &h->final_norm : first_ln_of(L + 1);
      int nh = last ? 0 : first_head_of(L + 1);
      if (c.resi_conv) {
        Epilogue ep; ep.resid = w.XA; ep.resid_pitch = d.C; ep.out_f32 = w.XA; ep.out_f32_pitch = d.C;
        set_ln(ep, nl, w.XN, nh);
        HB_TRY(conv("rhag_conv", lw.conv, w.XN2, d.Cp, 0, d.H, d.W, ep));
        HB_TRY(head_after(nl ? nh : 0));
      } else {
        HB_TRY(launch_add_convert(w.XA, w.XB, d.M, d.C, w.XA, nullptr, d.dt, 0, 0, st));
        HB_TRY(ln(w.XA, *nl, w.XN, nh));
      }
    }
    if (features_only) {      // XN = norm(x) (hat_arch.py:844) -> patch_unembed: (B, C, H, W)
      // (both the '1conv' and the 'identity' loop tails leave the final LayerNorm in XN)
      HB_TRY(launch_tokens_to_nchw(w.XN, d.dt, d.Cp, d.B, d.C, d.H * d.W, y, y_dt, st));
      return dbg("forward_features");
    }
    // XN = norm(x) (hat_arch.py:844); conv_after_body + long skip (hat_arch.py:854) -> XN2 (T)
    if (c.resi_conv) {
      Epilogue ep; ep.resid = w.S; ep.resid_pitch = d.C; ep.out_t = w.XN2; ep.out_t_pitch = d.Cp; ep.out_t_zero_to = d.Cp;
      HB_TRY(conv("conv_after_body", h->conv_after_body, w.XN, d.Cp, 0, d.H, d.W, ep));
    } else {
      HB_TRY(launch_layernorm(w.XA, d.M, d.C, h->final_norm.g, h->final_norm.b, w.XB, HATB200_F32, d.C, d.C, nullptr,
                              0, st));
      HB_TRY(launch_add_convert(w.XB, w.S, d.M, d.C, nullptr, w.XN2, d.dt, d.Cp, d.Cp, st));
    }
    {  // conv_before_upsample + LeakyReLU (hat_arch.py:855)
      Epilogue ep; ep.act = ACT_LRELU; ep.out_t = w.F64; ep.out_t_pitch = 64;
      HB_TRY(conv("conv_before_upsample", h->conv_before_up, w.XN2, d.Cp, 0, d.H, d.W, ep));
    }
    // Upsample: conv + PixelShuffle folded into the store (hat_arch.py:593-605)
    const void* hr = w.F64;
    int hh = d.H, hw = d.W;
    if (c.upscale == 3) {
      Epilogue ep; ep.out_t = w.UP1; ep.out_t_pitch = 64; ep.shuffle_r = 3;
      HB_TRY(conv_shuffle("upsample_conv", h->up0, w.F64, hh, hw, ep));
      hr = w.UP1; hh *= 3; hw *= 3;
    } else {
      if (c.upscale >= 2) {
        Epilogue ep; ep.out_t = w.UP1; ep.out_t_pitch = 64; ep.shuffle_r = 2;
        HB_TRY(conv_shuffle("upsample_conv", h->up0, w.F64, hh, hw, ep));
        hr = w.UP1; hh *= 2; hw *= 2;
      }
      if (c.upscale >= 4) {
        Epilogue ep; ep.out_t = w.UP2; ep.out_t_pitch = 64; ep.shuffle_r = 2;
        HB_TRY(conv_shuffle("upsample_conv", h->up1, w.UP1, hh, hw, ep));
        hr = w.UP2; hh *= 2; hw *= 2;
      }
    }
    int tid = t_begin("conv_last", 2.0 * d.B * hh * hw * 64 * 9 * c.in_chans,
                      (double)d.B * hh * hw * (64 * d.es + c.in_chans * (y_dt == HATB200_F32 ? 4 : 2)),
                      tc() ? "conv3_tc+nhwc_to_nchw_affine" : "conv_last");
    // tensor-core route: N = 3 padded to a 16-column MMA tile on the persistent conv3 kernel (fp32 pixels-last out),
    // then one pass that applies "/ img_range + mean" and writes the NCHW planes.  (The CUDA-core kernel below spends
    // one shared-memory weight load per FMA: 172 us for 16 x 256 x 256 pixels.)
    if (tc() && h->conv_last_tc.w16 && c.in_chans <= 4 && !tc_disabled("conv_last")) {
      // one pass (round 2): the conv3 epilogue itself applies "/ img_range + mean" and writes the NCHW planes -- or,
      // y_dtype U8, the tensor2img bytes (clamp, * 255, round half to even, HWC, BGR) of the bf16 image
      const ConvW& cw = h->conv_last_tc;
      ConvArgs a{hr, 64, 0, 64, d.B, hh, hw, 3, 3, cw.O};
      Epilogue ep; ep.bias = cw.bias;
      ep.nchw_y = y; ep.nchw_dt = y_dt; ep.nchw_oc = c.in_chans; ep.nchw_range = c.img_range; ep.nchw_via_bf16 = 1;
      for (int o = 0; o < 4; o++) ep.nchw_mean[o] = o < c.in_chans ? c.mean[o] : 0.f;
      static const bool two_pass = getenv("HATB200_CONV_LAST_2PASS") != nullptr;       // engineering: A/B
      if (!two_pass && conv3_tc_supported(a, cw.Npad, cw.Kpad, ep)) {
        if (tid >= 0) h->recs[tid].kernel = "conv3_tc";
        HB_TRY(launch_conv3_tc(a, cw.w16, cw.Npad, cw.Kpad, ep, st));
        t_end(tid);
        return dbg("conv_last one pass");
      }
    }
    HB_REQUIRE(y_dt != HATB200_U8, HATB200_EUNSUPPORTED,
               "y_dtype U8 needs the one-pass tcgen05 conv_last (bf16 compute, tensor cores on, <= 4 image channels)");
    if (tc() && w.CL32 && h->conv_last_tc.w16 && c.in_chans <= 4 && !tc_disabled("conv_last")) {
      const ConvW& cw = h->conv_last_tc;
      ConvArgs a{hr, 64, 0, 64, d.B, hh, hw, 3, 3, cw.O};
      Epilogue ep; ep.bias = cw.bias; ep.out_f32 = w.CL32; ep.out_f32_pitch = 4;
      if (conv3_tc_supported(a, cw.Npad, cw.Kpad, ep)) {
        HB_TRY(launch_conv3_tc(a, cw.w16, cw.Npad, cw.Kpad, ep, st));
        HB_TRY(launch_nhwc_to_nchw_affine(w.CL32, 4, d.B, c.in_chans, hh, hw, c.mean, c.img_range, y, y_dt, st));
        t_end(tid);
        return dbg("conv_last tc");
      }
    }
    HB_TRY(launch_conv_last(hr, d.dt, 64, 64, d.B, hh, hw, h->conv_last_w, h->conv_last_b, c.in_chans, c.mean,
                            c.img_range, y, y_dt, st));
    t_end(tid);
    return dbg("conv_last");
  }
};

static int check_geometry(hatb200_handle* h, int B, int H, int W) {
  HB_REQUIRE(h, HATB200_EINVAL, "null handle");
  int ws = h->cfg.window_size;
  HB_REQUIRE(B >= 1 && H >= ws && W >= ws && H % ws == 0 && W % ws == 0, HATB200_EINVAL,
             "input %dx%d (batch %d) must be a positive multiple of window_size %d in both dimensions", H, W, B, ws);
  HB_REQUIRE((long long)B * H * W * 16 * 64 < (1ll << 31), HATB200_EUNSUPPORTED,
             "input too large for 32-bit token indexing: use tile mode");
  return 0;
}

}  // namespace hb

using namespace hb;

extern "C" int hatb200_workspace_bytes(hatb200_handle* h, int B, int H, int W, int compute_dtype, size_t* out) {
  HB_TRY(check_geometry(h, B, H, W));
  HB_REQUIRE(out, HATB200_EINVAL, "workspace_bytes: null out");
  HB_REQUIRE(compute_dtype == HATB200_F32 || compute_dtype == HATB200_BF16, HATB200_EUNSUPPORTED,
             "compute dtype %d: only F32 and BF16 are implemented", compute_dtype);
  Dims d = make_dims(h->cfg, B, H, W, compute_dtype);
  Workspace w;
  plan(h->cfg, d, nullptr, w);
  *out = w.bytes + 256;
  return 0;
}

static int forward_common(hatb200_handle* h, const void* x, int x_dtype, int B, int H, int W, void* y, int y_dtype,
                          void* workspace, size_t ws_bytes, void* stream, bool features_only);

extern "C" int hatb200_forward(hatb200_handle* h, const void* x, int x_dtype, int B, int H, int W, void* y,
                               int y_dtype, void* workspace, size_t ws_bytes, void* stream) {
  return forward_common(h, x, x_dtype, B, H, W, y, y_dtype, workspace, ws_bytes, stream, false);
}

extern "C" int hatb200_forward_features(hatb200_handle* h, const void* x, int x_dtype, int B, int H, int W, void* y,
                                        int y_dtype, void* workspace, size_t ws_bytes, void* stream) {
  return forward_common(h, x, x_dtype, B, H, W, y, y_dtype, workspace, ws_bytes, stream, true);
}

static int forward_common(hatb200_handle* h, const void* x, int x_dtype, int B, int H, int W, void* y, int y_dtype,
                          void* workspace, size_t ws_bytes, void* stream, bool features_only) {
  HB_TRY(check_geometry(h, B, H, W));
  HB_REQUIRE(x && y, HATB200_EINVAL, "forward: null tensor");
  HB_REQUIRE(x_dtype == HATB200_F32 || x_dtype == HATB200_BF16, HATB200_EUNSUPPORTED,
             "forward: input dtype %d (only F32 and BF16 are implemented)", x_dtype);
  HB_REQUIRE(y_dtype == HATB200_F32 || y_dtype == HATB200_BF16 || (y_dtype == HATB200_U8 && !features_only),
             HATB200_EUNSUPPORTED, "forward: output dtype %d", y_dtype);
  HB_REQUIRE(h->packed, HATB200_EPARAM, "forward: weights are not packed (hatb200_set_param + hatb200_pack first)");
  Fwd f;
  f.h = h;
  f.st = (cudaStream_t)stream;
  f.d = make_dims(h->cfg, B, H, W, x_dtype);
  plan(h->cfg, f.d, nullptr, f.w);
  HB_REQUIRE(workspace != nullptr && ws_bytes >= f.w.bytes + 256, HATB200_ENOMEM,
             "forward: workspace of %zu bytes is smaller than the %zu required", ws_bytes, f.w.bytes + 256);
  char* base = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  plan(h->cfg, f.d, base, f.w);
  g_launches = 0;
  h->last_fallbacks = 0;
  h->recs.clear();
  h->ev_used = 0;
  int r = f.run(x, x_dtype, y, y_dtype, features_only);
  h->last_launches = g_launches;
  return r;
}
