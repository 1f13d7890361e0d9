// model.cuh -- handle, state_dict registry and packed-weight structures of libhatb200.
#pragma once
#include <map>
#include <string>
#include <vector>

#include "kernels.cuh"

namespace hb {

struct ParamRef {            // one state_dict entry (borrowed device pointer)
  std::string key;
  std::vector<int64_t> shape;
  bool is_int = false;       // int64 index buffer
  const void* ptr = nullptr;
  int dtype = -1;
  size_t numel() const { size_t n = 1; for (auto s : shape) n *= (size_t)s; return n; }
};

struct ConvW {               // dense conv / Linear weight in both kernel layouts
  int O = 0, I = 0, KH = 1, KW = 1;
  float* w32 = nullptr; int Np = 0;            // fp32 [taps][I][Np]        (CUDA-core kernels)
  bf16* w16 = nullptr; int Npad = 0, Kpad = 0;  // bf16 [taps][Npad][Kpad]   (tcgen05 kernels)
  float* bias = nullptr;                        // fp32 [O] (zeros when the layer has no bias)
  bf16* w16s = nullptr; float* bias_s = nullptr; int shuffle_r = 0;   // pixel-shuffle layers: rows permuted to (i, j, c)
  int algoO = 0, algoI = 0;                     // un-padded sizes for FLOP accounting when O / I include layout padding
  int taps() const { return KH * KW; }
};

struct LNW { float* g = nullptr; float* b = nullptr; };

struct EscW {                // ESC ConvAttn (esc_arch.py:89-145): plk_filter, dwc_proj, aggr
  int pdim = 0;
  ConvW plk;                 // (pdim, pdim, k, k), no bias
  bf16* plk_wt = nullptr;    // banded Toeplitz form [k][4*pdim][16*pdim] for tc_plk.cu (pdim 16, k 13 only)
  float *w1 = nullptr, *b1 = nullptr, *w2 = nullptr, *b2 = nullptr;   // dwc_proj.1 / dwc_proj.3
  ConvW aggr;                // 1x1 C->C
};

struct HabW {
  LNW n1, n2;
  EscW esc;
  ConvW cab0, cab2;
  float* eca_w5 = nullptr;
  ConvW fc1, fc2;
  float *dw_w = nullptr, *dw_b = nullptr;      // [9][2*hid] tap-major, [2*hid]
};

struct OcabW {
  LNW n1, n2;
  ConvW qkv;                 // q_proj rows [0,C) then kv_proj rows [C,3C)
  ConvW q, kv;               // separate copies (ocab_esc_enable: K/V come from another input)
  float* bias_compact = nullptr;   // (heads, (ws+ows-1)^2)
  // tcgen05 attention path: head-padded (head_dim -> 32) qkv / proj weights and the log2e-scaled bias table
  ConvW qkv_tc, proj_tc;
  ConvW focus0, focus1;            // HATX focus head: 1x1 C -> C/4 (GELU) -> 1 (hatx_arch.py:355-361)
  ConvW q_tc, kv_tc;               // ocab_esc_enable: separate head-padded q / kv GEMMs (different inputs)
  float* bias_l2 = nullptr;        // (heads, ATTN_TC_BIAS_FLOATS)
  ConvW proj, mlp0, mlp2;
  EscW esc;
  int heads = 0;
};

struct LayerW {
  std::vector<HabW> habs;
  OcabW ocab;
  ConvW conv;
};

struct Workspace {           // carved from the caller's scratch buffer
  float *S = nullptr, *XA = nullptr, *XB = nullptr;      // fp32 (M, C): shallow, group input, running stream
  void *XN = nullptr, *XN2 = nullptr, *U = nullptr;      // T (M, Cp)
  void *X16 = nullptr;                                   // T (M, pdim)
  void *X16P = nullptr;                                  // bf16 (B, H, W+12, 16): x-padded copy for tc_plk.cu
  void *T1 = nullptr;                                    // T (M, midp)
  void *H1 = nullptr;                                    // T (M, max(hp4, qkvp))
  void *G = nullptr;                                     // T (M, hp2)
  void *F64 = nullptr, *UP1 = nullptr, *UP2 = nullptr;   // tail
  float* CL32 = nullptr;                                 // conv_last output, pixels-last fp32 (B*Hr*Wr, 4), bf16 path
  float *part = nullptr, *part16 = nullptr, *eca_s = nullptr, *dyn = nullptr;
  void* SAL = nullptr;                                   // HATX: T (M, 16) focus-head output (column 0)
  float *KN = nullptr, *KADD = nullptr;                  // HATX: per-token ||k||_2; per (window, key) additive term / -inf
  size_t bytes = 0;
};

}  // namespace hb

namespace hb {
struct OpRec { const char* tag; const char* kernel; cudaEvent_t e0, e1; double flops, bytes; };
}

struct hatb200_handle {
  hatb200_config cfg;
  bool op_timing = false;            // record CUDA events around every op (bench roofline breakdown)
  std::vector<hb::OpRec> recs;
  std::vector<cudaEvent_t> ev_pool;
  size_t ev_used = 0;
  int device = 0;
  std::vector<hb::ParamRef> params;
  std::map<std::string, int> index;
  bool packed = false;
  bool use_tc = true;
  bool sync_debug = false;
  bool strict_tc = false;            // bf16 path: a shape no tcgen05 kernel takes is an error instead of a CUDA-core launch
  int last_launches = 0;
  int last_fallbacks = 0;            // ops of the last bf16 forward that ran on a CUDA-core kernel instead of tcgen05
  std::vector<void*> allocs;
  std::vector<size_t> alloc_bytes;   // sizes of `allocs` (packed-weight export / import)
  int* err_flag = nullptr;
  // packed weights
  float *conv_first_w = nullptr, *conv_first_b = nullptr;   // [9][IC][C]
  hb::LNW pe_norm, final_norm;
  std::vector<hb::LayerW> layers;
  hb::ConvW conv_after_body, conv_before_up, up0, up1;
  float *conv_last_w = nullptr, *conv_last_b = nullptr;     // [9][64][OC]
  hb::ConvW conv_last_tc;                                        // the same layer for tc_conv3.cu (N padded to 16)
  float* mean_dev = nullptr;
};
