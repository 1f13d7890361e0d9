// tc_stub.cu -- placeholder until the tcgen05 kernels land: reports "unsupported" so every op
// takes the CUDA-core path.  (Replaced by tc_tapconv.cu / tc_attn.cu.)
#include "kernels.cuh"
namespace hb {
bool tapconv_tc_supported(const ConvArgs&, int, int, const Epilogue&) { return false; }
int launch_tapconv_tc(const ConvArgs&, const bf16*, int, int, const Epilogue&, cudaStream_t) {
  set_error("tcgen05 conv kernel not built");
  return HATB200_EUNSUPPORTED;
}
bool attn_tc_supported(const AttnArgs&) { return false; }
int launch_attn_tc(const AttnArgs&, cudaStream_t) {
  set_error("tcgen05 attention kernel not built");
  return HATB200_EUNSUPPORTED;
}
}  // namespace hb
