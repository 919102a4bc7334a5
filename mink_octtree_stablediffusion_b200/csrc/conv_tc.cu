// placeholder, replaced below
#include "common.cuh"
namespace mk {
bool conv_fwd_tc_eligible(int, int64_t, int, int, int) { return false; }
int64_t conv_fwd_tc_workspace(int, int, int) { return 0; }
int conv_fwd_tc(int, const void *, const void *, int, int, const int32_t *, int64_t, int64_t, int, int, int, void *, void *, cudaStream_t) { return MK_EUNSUPPORTED; }
bool conv_wgrad_tc_eligible(int, int64_t, int, int, int) { return false; }
int64_t conv_wgrad_tc_workspace(int, int, int) { return 0; }
int conv_wgrad_tc(int, const void *, const void *, const int32_t *, int64_t, int64_t, int, int, int, void *, void *, cudaStream_t) { return MK_EUNSUPPORTED; }
}
extern "C" int mk_has_tcgen05(void) { return 0; }
