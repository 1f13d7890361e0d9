// tc_stub.cu -- placeholder until the tcgen05 attention kernel lands: reports "unsupported" so the
// attention core takes the CUDA-core kernel.  (To be replaced by tc_attn.cu.)
#include "kernels.cuh"
namespace hb {
bool attn_tc_supported(const AttnArgs&) { return false; }
int launch_attn_tc(const AttnArgs&, cudaStream_t) {
  set_error("tcgen05 attention kernel not built");
  return HATB200_EUNSUPPORTED;
}
}  // namespace hb
