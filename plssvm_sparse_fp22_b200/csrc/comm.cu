// Multi-GPU exchange: one process per GPU, NCCL over NVLink 5 / NVSwitch on the handle's stream.
//
// Replaces the reference's host-staged gpu_csvm::device_reduction (src/plssvm/backends/gpu_csvm.cpp:365-386: D2H from every
// device, add on the host, H2D to every device).  Here each rank computes the partial K d of its tiles / rows and ONE
// ncclAllReduce(sum) of m reals per matvec makes the full vector resident on every rank; the O(m) CG algebra is then replicated
// bit-identically on all ranks, so no scalar collectives and no host round trips are needed.
//
// NCCL is bound at run time (dlopen): under Python the copy torch already loaded is reused (RTLD_NOLOAD), a stand-alone C++
// host picks up the system libnccl.so.2.  Only the six stable entry points below are used.
#include "internal.cuh"

#include <dlfcn.h>

#include <cstring>

namespace b200 {

namespace {

// minimal NCCL ABI (nccl.h: ncclUniqueId is 128 opaque bytes; ncclDataType_t: float = 7, double = 8; ncclRedOp_t: sum = 0)
struct nccl_unique_id {
    char internal[128];
};
using nccl_comm_t = void *;
constexpr int NCCL_FLOAT32 = 7, NCCL_FLOAT64 = 8, NCCL_SUM = 0, NCCL_SUCCESS = 0;

struct nccl_api {
    int (*GetUniqueId)(nccl_unique_id *){ nullptr };
    int (*CommInitRank)(nccl_comm_t *, int, nccl_unique_id, int){ nullptr };
    int (*CommDestroy)(nccl_comm_t){ nullptr };
    int (*AllReduce)(const void *, void *, size_t, int, int, nccl_comm_t, cudaStream_t){ nullptr };
    int (*AllGather)(const void *, void *, size_t, int, nccl_comm_t, cudaStream_t){ nullptr };
    const char *(*GetErrorString)(int){ nullptr };
    int (*GetVersion)(int *){ nullptr };
    void *lib{ nullptr };
    std::string origin;
};

nccl_api &api() {
    static nccl_api a;
    if (a.lib != nullptr) { return a; }
    const char *env = std::getenv("B200SVM_NCCL_LIB");
    struct cand {
        const char *name;
        int flags;
    };
    const cand cands[] = { { env, RTLD_NOW | RTLD_GLOBAL },
                           { "libnccl.so.2", RTLD_NOW | RTLD_NOLOAD },  // the copy the process (torch) already mapped
                           { "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL },
                           { "libnccl.so", RTLD_NOW | RTLD_GLOBAL } };
    for (const cand &c : cands) {
        if (c.name == nullptr) { continue; }
        void *h = dlopen(c.name, c.flags);
        if (h != nullptr) {
            a.lib = h;
            a.origin = c.name;
            break;
        }
    }
    if (a.lib == nullptr) { fail(B200SVM_ERR_NCCL, "libnccl.so.2 could not be loaded (set B200SVM_NCCL_LIB to its path)"); }
#define B200_NCCL_SYM(field, sym)                                                                  \
    a.field = reinterpret_cast<decltype(a.field)>(dlsym(a.lib, sym));                              \
    if (a.field == nullptr) { fail(B200SVM_ERR_NCCL, std::string("NCCL symbol missing: ") + sym); }
    B200_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    B200_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    B200_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    B200_NCCL_SYM(AllReduce, "ncclAllReduce")
    B200_NCCL_SYM(AllGather, "ncclAllGather")
    B200_NCCL_SYM(GetErrorString, "ncclGetErrorString")
    B200_NCCL_SYM(GetVersion, "ncclGetVersion")
#undef B200_NCCL_SYM
    return a;
}

void nccl_check(int rc, const char *what) {
    if (rc != NCCL_SUCCESS) { fail(B200SVM_ERR_NCCL, std::string("NCCL error '") + api().GetErrorString(rc) + "' in " + what); }
}

}  // namespace

struct comm_state {
    nccl_comm_t comm{ nullptr };
};

void comm_destroy(ctx &c) {
    if (c.comm != nullptr) {
        if (c.comm->comm != nullptr) { api().CommDestroy(c.comm->comm); }
        delete c.comm;
        c.comm = nullptr;
    }
    c.rank = 0;
    c.world = 1;
}

// in-place sum over ranks of `count` reals (type of the handle) on the handle's stream
void comm_allreduce_sum(ctx &c, void *buf, int64_t count) {
    if (c.world <= 1 || count == 0) { return; }
    if (c.comm == nullptr || c.comm->comm == nullptr) { fail(B200SVM_ERR_NCCL, "NCCL communicator not initialised"); }
    nccl_check(api().AllReduce(buf, buf, (size_t) count, c.dtype == B200SVM_F32 ? NCCL_FLOAT32 : NCCL_FLOAT64, NCCL_SUM, c.comm->comm, c.stream),
               "ncclAllReduce");
    ++c.stats.collectives;
    c.stats.collective_bytes += count * (int64_t) c.esize();
}

// Row range [lo, hi) of the m rows of the implicit matrix owned by `rank` (full-row paths: equal row counts)
void partition_rows(int64_t m, int rank, int world, int64_t *lo, int64_t *hi) {
    if (world <= 1) {
        *lo = 0;
        *hi = m;
        return;
    }
    *lo = m * rank / world;
    *hi = m * (rank + 1) / world;
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200svm_comm_unique_id(void *id128) {
    try {
        if (id128 == nullptr) { return B200SVM_ERR_INVALID; }
        nccl_unique_id id{};
        nccl_check(api().GetUniqueId(&id), "ncclGetUniqueId");
        std::memcpy(id128, &id, sizeof(id));
        return B200SVM_OK;
    } catch (const error &e) { return e.code; }
}

int b200svm_comm_init(b200svm_handle h, const void *id128, int rank, int world) {
    ctx *c = reinterpret_cast<ctx *>(h);
    if (c == nullptr) { return B200SVM_ERR_INVALID; }
    try {
        B200_CUDA_CHECK(cudaSetDevice(c->device));
        B200_REQUIRE(world >= 1 && rank >= 0 && rank < world, "rank / world out of range");
        B200_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        comm_destroy(*c);
        if (world > 1) {
            B200_REQUIRE(id128 != nullptr, "null NCCL unique id");
            nccl_unique_id id{};
            std::memcpy(&id, id128, sizeof(id));
            c->comm = new comm_state();
            nccl_check(api().CommInitRank(&c->comm->comm, world, id, rank), "ncclCommInitRank");
        }
        c->rank = rank;
        c->world = world;
        // tile / row ownership changed: device-side work lists are rebuilt lazily
        if (c->has_data && c->has_split) { tc_prepare_split(*c); }
        return B200SVM_OK;
    } catch (const error &e) {
        c->last_error = e.msg;
        return e.code;
    }
}

int b200svm_partition_rows(int64_t m, int rank, int world, int64_t *lo, int64_t *hi) {
    if (lo == nullptr || hi == nullptr || world < 1 || rank < 0 || rank >= world || m < 0) { return B200SVM_ERR_INVALID; }
    partition_rows(m, rank, world, lo, hi);
    return B200SVM_OK;
}

}  // extern "C"
