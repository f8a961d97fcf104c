// Multi-GPU exchange (NCCL).  Placeholder: single-rank only.
#include "internal.cuh"

namespace b200 {
void comm_destroy(ctx &) {}
void comm_allreduce_sum(ctx &c, void *, int64_t) {
    if (c.world > 1) { fail(B200SVM_ERR_NCCL, "NCCL communicator not initialised"); }
}
}  // namespace b200

extern "C" {
int b200svm_comm_unique_id(void *) { return B200SVM_ERR_NCCL; }
int b200svm_comm_init(b200svm_handle, const void *, int, int world) { return world == 1 ? B200SVM_OK : B200SVM_ERR_NCCL; }
}
