// noalloc.cu -- hatb200_dwconv_gate_noalloc: the GatedDconvFFN middle (hat_arch.py:112-116) for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// hatb200_dwconv_gate (entries.cu) re-lays the PyTorch (2*hid, 1, 3, 3) depthwise weight as [tap][channel] in
// a cudaMalloc'ed temporary.  Here that scratch is a file-scope device array of a fixed, modest size: every
// GPU that loads libhatb200.so holds HATB200_NOALLOC_MAX_HID * 2 * 9 floats (72 KiB) until the process exits.
// All calls share the array and it keeps the last call's contents: the part a call uses is cleared and
// rewritten on the caller's stream first, so consecutive calls on ONE stream are safe; calls that overlap on
// the GPU (different streams, different threads) race on it.
//
// In host code `g_dwgate_scratch` is only a host-side stand-in for the symbol, never a device address:
// the address is looked up once per device with cudaGetSymbolAddress (that device current) and cached.
#include <mutex>

#include "model.cuh"

using namespace hb;

namespace {

constexpr int kMaxDevices = 64;
constexpr size_t kScratchFloats = (size_t)9 * 2 * HATB200_NOALLOC_MAX_HID;

static __device__ __align__(256) float g_dwgate_scratch[kScratchFloats];

std::mutex g_scratch_mu;
float* g_scratch_dev[kMaxDevices];   // device address of g_dwgate_scratch on each device, null until looked up

int scratch_address(float** out) {
  int dev = -1;
  HB_CHECK_CUDA(cudaGetDevice(&dev));
  HB_REQUIRE(dev >= 0 && dev < kMaxDevices, HATB200_EUNSUPPORTED, "dwconv_gate_noalloc: device ordinal %d", dev);
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (!g_scratch_dev[dev]) {
    float* p = nullptr;
    HB_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_dwgate_scratch));
    HB_REQUIRE(p != nullptr, HATB200_ECUDA, "dwconv_gate_noalloc: cudaGetSymbolAddress returned a null address");
    g_scratch_dev[dev] = p;
  }
  *out = g_scratch_dev[dev];
  return 0;
}

}  // namespace

extern "C" int hatb200_dwconv_gate_noalloc(const void* in, int dtype, int in_pitch, int B, int H, int W, int hid,
                                           const float* dw_w, const float* dw_b, void* out, int out_pitch,
                                           void* stream) {
  HB_REQUIRE(in && dw_w && dw_b && out, HATB200_EINVAL, "dwconv_gate_noalloc: null tensor");
  HB_REQUIRE(hid >= 1 && hid <= HATB200_NOALLOC_MAX_HID, HATB200_EINVAL,
             "dwconv_gate_noalloc: hid %d is outside [1, %d], the size of the static weight scratch "
             "(use hatb200_dwconv_gate, which allocates)", hid, HATB200_NOALLOC_MAX_HID);
  cudaStream_t st = (cudaStream_t)stream;
  float* wt = nullptr;
  HB_TRY(scratch_address(&wt));
  const size_t used = (size_t)9 * 2 * hid;
  HB_CHECK_CUDA(cudaMemsetAsync(wt, 0, used * sizeof(float), st));   // last call's contents out
  HB_TRY(launch_pack_w_f32(dw_w, HATB200_F32, 2 * hid, 1, 9, wt, 2 * hid, 2 * hid, st));
  return launch_dwgate(in, dtype, in_pitch, B, H, W, hid, wt, dw_b, out, out_pitch, out_pitch, st);
}
