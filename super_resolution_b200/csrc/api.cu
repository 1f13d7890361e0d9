// api.cu -- C ABI of libhatb200: handle lifetime, state_dict registry, weight packing.
// (forward orchestration lives in forward.cu.)
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "model.cuh"

namespace hb {

static thread_local char g_err[1024] = "";
thread_local int g_launches = 0;

// HATB200_PDL=0 launches every kernel fully serialised (bisecting aid); default on
bool pdl_enabled() {
  static const bool on = !(getenv("HATB200_PDL") && getenv("HATB200_PDL")[0] == '0');
  return on;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static void add_param(hatb200_handle* h, const std::string& key, std::vector<int64_t> shape, bool is_int = false) {
  ParamRef p;
  p.key = key;
  p.shape = std::move(shape);
  p.is_int = is_int;
  h->index[key] = (int)h->params.size();
  h->params.push_back(std::move(p));
}

static void add_wb(hatb200_handle* h, const std::string& prefix, std::vector<int64_t> wshape, bool bias = true) {
  int64_t o = wshape[0];
  add_param(h, prefix + ".weight", std::move(wshape));
  if (bias) add_param(h, prefix + ".bias", {o});
}

static void add_esc(hatb200_handle* h, const std::string& core, int C, int pdim) {
  add_wb(h, core + ".plk.dwc_proj.1", {pdim / 2, pdim, 1, 1});
  add_wb(h, core + ".plk.dwc_proj.3", {pdim * 9, pdim / 2, 1, 1});
  add_wb(h, core + ".aggr", {C, C, 1, 1});
}

// The reference state_dict layout (hat_arch.py:669-757), in the
// reference's own key order.
static void build_registry(hatb200_handle* h) {
  const hatb200_config& c = h->cfg;
  const int C = c.embed_dim, ws = c.window_size, ows = c.overlap_win_size;
  add_param(h, "relative_position_index_SA", {ws * ws, ws * ws}, true);
  add_param(h, "relative_position_index_OCA", {ws * ws, ows * ows}, true);
  add_wb(h, "conv_first", {C, c.in_chans, 3, 3});
  if (c.patch_norm) add_wb(h, "patch_embed.norm", {C});
  for (int L = 0; L < c.num_layers; L++) {
    std::string lp = "layers." + std::to_string(L);
    for (int i = 0; i < c.depths[L]; i++) {
      std::string bp = lp + ".residual_group.blocks." + std::to_string(i);
      add_wb(h, bp + ".norm1", {C});
      add_param(h, bp + ".esc_attn.plk_filter", {c.esc_pdim, c.esc_pdim, c.esc_kernel, c.esc_kernel});
      add_esc(h, bp + ".esc_attn.core", C, c.esc_pdim);
      add_wb(h, bp + ".conv_block.cab.0", {c.cab_mid, C, 3, 3});
      add_wb(h, bp + ".conv_block.cab.2", {C, c.cab_mid, 3, 3});
      add_param(h, bp + ".conv_block.cab.3.conv.weight", {1, 1, 5});
      add_wb(h, bp + ".norm2", {C});
      if (c.arch == 1) {      // HATX spatial-gate FFN (hatx_arch.py:144-158): fc1 C -> hid, dw on hid / 2, fc2 hid -> C
        add_wb(h, bp + ".mlp.fc1", {c.ffn_hidden, C});
        add_wb(h, bp + ".mlp.dw", {c.ffn_hidden / 2, 1, 3, 3});
        add_wb(h, bp + ".mlp.fc2", {C, c.ffn_hidden});
      } else {
        add_wb(h, bp + ".mlp.fc1", {2 * c.ffn_hidden, C});
        add_wb(h, bp + ".mlp.dw", {2 * c.ffn_hidden, 1, 3, 3});
        add_wb(h, bp + ".mlp.fc2", {C, c.ffn_hidden});
      }
    }
    std::string op = lp + ".residual_group.overlap_attn";
    int T = ws + ows - 1;
    add_param(h, op + ".relative_position_bias_table", {T * T, c.num_heads[L]});
    if (c.ocab_esc_enable)
      add_param(h, op + ".esc_plk", {c.ocab_esc_pdim, c.ocab_esc_pdim, c.ocab_esc_kernel, c.ocab_esc_kernel});
    add_wb(h, op + ".norm1", {C});
    add_wb(h, op + ".q_proj", {C, C}, c.qkv_bias != 0);
    add_wb(h, op + ".kv_proj", {2 * C, C}, c.qkv_bias != 0);
    add_wb(h, op + ".proj", {C, C});
    add_wb(h, op + ".norm2", {C});
    add_wb(h, op + ".mlp.0", {c.ffn_hidden, C});
    add_wb(h, op + ".mlp.2", {C, c.ffn_hidden});
    if (c.ocab_esc_enable) add_esc(h, op + ".esc_core", C, c.ocab_esc_pdim);
    if (c.arch == 1 && c.use_focus_bias) {      // hatx_arch.py:355-361
      add_wb(h, op + ".focus_head.0", {C / 4, C, 1, 1});
      add_wb(h, op + ".focus_head.2", {1, C / 4, 1, 1});
    }
    if (c.resi_conv) add_wb(h, lp + ".conv", {C, C, 3, 3});
  }
  add_wb(h, "norm", {C});
  if (c.resi_conv) add_wb(h, "conv_after_body", {C, C, 3, 3});
  if (c.upsampler == 1) {
    add_wb(h, "conv_before_upsample.0", {64, C, 3, 3});
    if (c.upscale == 3) {
      add_wb(h, "upsample.0", {9 * 64, 64, 3, 3});
    } else {
      int n = 0;
      for (int s = c.upscale; s > 1; s >>= 1) n++;
      for (int i = 0; i < n; i++) add_wb(h, "upsample." + std::to_string(2 * i), {4 * 64, 64, 3, 3});
    }
    add_wb(h, "conv_last", {c.in_chans, 64, 3, 3});
  }
}

}  // namespace hb

using namespace hb;

extern "C" {

int hatb200_version(void) { return HATB200_VERSION; }
const char* hatb200_last_error(void) { return hb::g_err; }

int hatb200_create(const hatb200_config* cfg, hatb200_handle** out) {
  HB_REQUIRE(cfg && out, HATB200_EINVAL, "create: null argument");
  HB_REQUIRE(cfg->struct_size == (int)sizeof(hatb200_config), HATB200_EINVAL,
             "create: config struct_size %d != %d (ABI mismatch)", cfg->struct_size, (int)sizeof(hatb200_config));
  const hatb200_config& c = *cfg;
  HB_REQUIRE(c.in_chans >= 1 && c.in_chans <= 4, HATB200_EUNSUPPORTED, "in_chans=%d not in 1..4", c.in_chans);
  HB_REQUIRE(c.num_layers >= 1 && c.num_layers <= HATB200_MAX_LAYERS, HATB200_EINVAL, "num_layers=%d", c.num_layers);
  HB_REQUIRE(c.embed_dim >= 8 && c.embed_dim <= 256, HATB200_EUNSUPPORTED, "embed_dim=%d not in 8..256", c.embed_dim);
  HB_REQUIRE(c.window_size >= 2 && c.window_size <= 16, HATB200_EUNSUPPORTED, "window_size=%d not in 2..16",
             c.window_size);
  HB_REQUIRE(c.arch == 0 || c.arch == 1, HATB200_EINVAL, "arch=%d (0 = HAT, 1 = HATX)", c.arch);
  HB_REQUIRE(c.overlap_win_size >= c.window_size && c.overlap_win_size <= 32, HATB200_EUNSUPPORTED,
             "overlap window %d must be in [window %d, 32]", c.overlap_win_size, c.window_size);
  if (c.arch == 0)
    HB_REQUIRE(((c.overlap_win_size - c.window_size) & 1) == 0 && c.ocab_pad == (c.overlap_win_size - c.window_size) / 2 &&
                   c.kv_topk_keep == 0 && c.use_focus_bias == 0, HATB200_EUNSUPPORTED,
               "HAT: overlap window %d must exceed window %d by an even amount (pad %d) and the HATX options must be off",
               c.overlap_win_size, c.window_size, c.ocab_pad);
  // nn.Unfold(ows, stride ws, pad) must yield exactly one key window per query window (hat_arch.py:296-297)
  HB_REQUIRE(2 * c.ocab_pad >= c.overlap_win_size - c.window_size && 2 * c.ocab_pad < c.overlap_win_size, HATB200_EUNSUPPORTED,
             "ocab_pad=%d does not tile windows of %d with overlap windows of %d", c.ocab_pad, c.window_size,
             c.overlap_win_size);
  HB_REQUIRE(c.kv_topk_keep >= 0 && c.kv_topk_keep <= c.overlap_win_size * c.overlap_win_size, HATB200_EINVAL,
             "kv_topk_keep=%d", c.kv_topk_keep);
  if (c.arch == 1) HB_REQUIRE(c.ffn_hidden % 2 == 0, HATB200_EINVAL, "Hidden(%d) must be even for spatial gate split.", c.ffn_hidden);
  HB_REQUIRE(c.cab_mid >= 1 && c.ffn_hidden >= 1, HATB200_EINVAL, "cab_mid=%d ffn_hidden=%d", c.cab_mid, c.ffn_hidden);
  HB_REQUIRE(c.esc_pdim >= 2 && c.esc_pdim <= 64 && c.esc_pdim <= c.embed_dim && (c.esc_kernel & 1),
             HATB200_EUNSUPPORTED, "esc_pdim=%d esc_kernel=%d", c.esc_pdim, c.esc_kernel);
  if (c.ocab_esc_enable)
    HB_REQUIRE(c.ocab_esc_pdim >= 2 && c.ocab_esc_pdim <= 64 && c.ocab_esc_pdim <= c.embed_dim &&
                   (c.ocab_esc_kernel & 1), HATB200_EUNSUPPORTED, "ocab_esc_pdim=%d", c.ocab_esc_pdim);
  for (int L = 0; L < c.num_layers; L++) {
    HB_REQUIRE(c.depths[L] >= 0 && c.num_heads[L] >= 1 && c.embed_dim % c.num_heads[L] == 0 &&
                   c.embed_dim / c.num_heads[L] <= 64, HATB200_EUNSUPPORTED,
               "layer %d: depth=%d heads=%d embed_dim=%d", L, c.depths[L], c.num_heads[L], c.embed_dim);
  }
  if (c.upsampler == 1)   // hat_arch.py:603-604
    HB_REQUIRE(c.upscale == 3 || (c.upscale >= 1 && (c.upscale & (c.upscale - 1)) == 0), HATB200_EINVAL,
               "scale %d is not supported. Supported scales: 2^n and 3.", c.upscale);
  hatb200_handle* h = new hatb200_handle();
  h->cfg = c;
  cudaError_t e = cudaGetDevice(&h->device);
  if (e != cudaSuccess) {
    set_error("create: no CUDA device (%s); libhatb200 has no CPU path", cudaGetErrorString(e));
    delete h;
    return HATB200_ECUDA;
  }
  build_registry(h);
  *out = h;
  return 0;
}

static void free_packed(hatb200_handle* h) {
  for (void* p : h->allocs) cudaFree(p);
  h->allocs.clear();
  h->layers.clear();
  h->packed = false;
}

void hatb200_destroy(hatb200_handle* h) {
  if (!h) return;
  free_packed(h);
  for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
  delete h;
}

int hatb200_num_params(hatb200_handle* h) { return h ? (int)h->params.size() : 0; }

const char* hatb200_param_key(hatb200_handle* h, int i) {
  if (!h || i < 0 || i >= (int)h->params.size()) return nullptr;
  return h->params[i].key.c_str();
}

int hatb200_set_param(hatb200_handle* h, const char* key, const void* dev_ptr, int dtype, const int64_t* shape,
                      int ndim) {
  HB_REQUIRE(h && key && dev_ptr && shape, HATB200_EINVAL, "set_param: null argument");
  auto it = h->index.find(key);
  HB_REQUIRE(it != h->index.end(), HATB200_EPARAM, "set_param: unexpected key '%s'", key);
  ParamRef& p = h->params[it->second];
  bool ok = ndim == (int)p.shape.size();
  for (int i = 0; ok && i < ndim; i++) ok = shape[i] == p.shape[i];
  if (!ok) {
    std::string want, got;
    for (auto s : p.shape) want += std::to_string(s) + ",";
    for (int i = 0; i < ndim; i++) got += std::to_string(shape[i]) + ",";
    set_error("set_param: size mismatch for %s: expected (%s) got (%s)", key, want.c_str(), got.c_str());
    return HATB200_EPARAM;
  }
  if (p.is_int)
    HB_REQUIRE(dtype == HATB200_I64, HATB200_EPARAM, "set_param: %s must be int64", key);
  else
    HB_REQUIRE(dtype == HATB200_F32 || dtype == HATB200_BF16 || dtype == HATB200_F16, HATB200_EPARAM,
               "set_param: %s has unsupported dtype %d", key, dtype);
  p.ptr = dev_ptr;
  p.dtype = dtype;
  h->packed = false;
  return 0;
}

int hatb200_set_option(hatb200_handle* h, const char* name, int value) {
  HB_REQUIRE(h && name, HATB200_EINVAL, "set_option: null argument");
  if (!strcmp(name, "tensor_cores")) h->use_tc = value != 0;
  else if (!strcmp(name, "sync_debug")) h->sync_debug = value != 0;
  else if (!strcmp(name, "op_timing")) h->op_timing = value != 0;
  else if (!strcmp(name, "strict_tc")) h->strict_tc = value != 0;
  else { set_error("set_option: unknown option '%s'", name); return HATB200_EINVAL; }
  return 0;
}

int hatb200_last_launch_count(hatb200_handle* h) { return h ? h->last_launches : 0; }
int hatb200_last_fallback_count(hatb200_handle* h) { return h ? h->last_fallbacks : 0; }

// Per-op device times of the last forward run with option "op_timing" = 1, as JSON:
// {"tag": {"kernel": name, "launches": n, "ms": t, "flops": f, "bytes": b}, ...}.  Synchronises.
int hatb200_op_profile(hatb200_handle* h, char* json, size_t cap) {
  HB_REQUIRE(h && json && cap > 2, HATB200_EINVAL, "op_profile: bad arguments");
  struct Agg { int n = 0; double ms = 0, flops = 0, bytes = 0; std::string kernel; };
  std::map<std::string, Agg> agg;
  std::vector<std::string> order;
  for (const OpRec& r : h->recs) {
    HB_CHECK_CUDA(cudaEventSynchronize(r.e1));
    float ms = 0.f;
    HB_CHECK_CUDA(cudaEventElapsedTime(&ms, r.e0, r.e1));
    if (!agg.count(r.tag)) order.push_back(r.tag);
    Agg& a = agg[r.tag];
    a.n++; a.ms += ms; a.flops += r.flops; a.bytes += r.bytes;
    const std::string k = r.kernel ? r.kernel : "?";
    if (a.kernel.empty()) a.kernel = k;
    else if (a.kernel.find(k) == std::string::npos) a.kernel += "+" + k;     // a tag served by more than one kernel
  }
  std::string out = "{";
  for (size_t i = 0; i < order.size(); i++) {
    const Agg& a = agg[order[i]];
    char line[384];
    snprintf(line, sizeof(line),
             "%s\"%s\": {\"kernel\": \"%s\", \"launches\": %d, \"ms\": %.6f, \"flops\": %.6e, \"bytes\": %.6e}",
             i ? ", " : "", order[i].c_str(), a.kernel.c_str(), a.n, a.ms, a.flops, a.bytes);
    out += line;
  }
  out += "}";
  HB_REQUIRE(out.size() + 1 <= cap, HATB200_ENOMEM, "op_profile: buffer of %zu bytes too small (%zu needed)", cap,
             out.size() + 1);
  memcpy(json, out.c_str(), out.size() + 1);
  return 0;
}

}  // extern "C"
