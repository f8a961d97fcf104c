// tcgen05 / TMA Gram-tile path (f32 dense).  Placeholder translation unit: the path reports "unavailable" so the
// dispatcher in cg.cu selects the SIMT kernel.
#include "internal.cuh"

namespace b200 {

void tc_prepare_split(ctx &) {}
bool tc_available(const ctx &) { return false; }
void tc_kernel_matvec(ctx &, const float *, float *) { fail(B200SVM_ERR_INVALID, "tcgen05 matvec path is not available in this build"); }

}  // namespace b200
