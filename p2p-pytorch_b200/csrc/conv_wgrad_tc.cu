// placeholder until the tcgen05 weight-gradient kernel lands
#include "tc_common.cuh"
namespace p2p {
bool conv_wgrad_tc_supported(const p2p_conv*) { return false; }
int conv_wgrad_tc(const p2p_conv*, const void*, const void*, float*, cudaStream_t) {
  set_error("tcgen05 wgrad not built");
  return -1;
}
}  // namespace p2p
