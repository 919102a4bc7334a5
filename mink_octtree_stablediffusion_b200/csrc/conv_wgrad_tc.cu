// Tensor-core weight gradient (placeholder until the tcgen05 kernel lands):
// reports "not eligible" so mk_conv_wgrad uses the SIMT split-row kernel.
#include "common.cuh"
namespace mk {
bool conv_wgrad_tc_eligible(int, int64_t, int, int, int) { return false; }
int64_t conv_wgrad_tc_workspace(int64_t, int, int, int) { return 0; }
int conv_wgrad_tc(int, const void *, const void *, const int32_t *, const uint32_t *, int64_t, int64_t, int64_t, int, int,
                  int, void *, void *, cudaStream_t) {
  set_error("conv_wgrad_tc: not built");
  return MK_EUNSUPPORTED;
}
}  // namespace mk
