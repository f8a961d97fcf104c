// Multi-GPU exchange: one process per GPU, NCCL over NVLink 5 / NVSwitch on the handle's stream.
//
// Replaces the reference's host-staged gpu_csvm::device_reduction (src/plssvm/backends/gpu_csvm.cpp:365-386: D2H from every
// device, add on the host, H2D to every device).  Here each rank computes the partial K d of its tiles / rows and stores it
// straight into every peer's exchange buffer over NVLink (CUDA-IPC mapped peer memory, xchg.cuh); the consumer is fused into the
// O(m) CG kernel that follows, so the full vector becomes resident on every rank without a collective call and the CG algebra is
// replicated bit-identically on all ranks: no scalar collectives, no host round trips.  NCCL carries the set-up
// handshake, the barrier in front of a buffer release, and the fallback (one ncclAllReduce per matvec) when the peers cannot be
// mapped or B200SVM_OPT_PEER_EXCHANGE = 1.
//
// NCCL is bound at run time (dlopen): under Python the copy torch already loaded is reused (RTLD_NOLOAD), a stand-alone C++
// host picks up the system libnccl.so.2.  Only the six stable entry points below are used.
#include "internal.cuh"
#include "xchg.cuh"

#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <vector>

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

// ---------------------------------------------------------------------------------------------------------------------
// Peer-memory exchange (protocol in xchg.cuh): the sum (or gather) of the matvec result over the ranks without
// NCCL on the critical path.
//   sum mode    (tile-partitioned paths): rank r stores its full-length partial vector into slot r of EVERY peer's buffer; the
//               consumer adds the `world` slots in rank order — the same order on every rank, so the replicated CG state stays
//               bit-identical.  The tcgen05 path does the stores from its own partial-vector reduction kernel (gram_tc.cu).
//   gather mode (row-partitioned paths): rank r stores only its rows [lo, hi) into slot 0 of every peer — 1 / world of the bytes.
// ---------------------------------------------------------------------------------------------------------------------
// generic producer: P[lo, hi) -> slot `slot` of every peer (sum mode: lo = 0, hi = m, slot = rank; gather mode: slot = 0)
template <typename T>
__global__ void __launch_bounds__(256) k_xchg_push(const T *__restrict__ P, int64_t lo, int64_t hi, xchg_push x, int slot, const T *__restrict__ done) {
    __shared__ bool is_last;
    if (done != nullptr && *done != T(0)) { return; }  // CG has converged: this queued exchange is skipped on every rank (xchg.cuh)
    const unsigned long long seq = xchg_next_seq(x);
    for (int r = 0; r < x.world; ++r) {
        T *dst = xchg_slot<T>(x, r, seq, slot);
        for (int64_t i = lo + (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (int64_t) gridDim.x * blockDim.x) { dst[i] = P[i]; }
    }
    xchg_publish(x, seq, &is_last);
}

struct comm_state {
    nccl_comm_t comm{ nullptr };
    // peer exchange
    bool xchg_tried{ false }, xchg_ok{ false };
    int64_t ld{ 0 };            // reals per slot
    size_t flags_offset{ 0 };   // byte offset of the flags inside a buffer
    void *local{ nullptr };
    void *peers[MAX_PEERS]{};
    unsigned long long *dseq{ nullptr };  // device: exchanges performed (advanced by the producers)
    unsigned int *ticket{ nullptr };
    dbuf bar;                   // one float for the barrier allreduce
    int world{ 1 }, rank{ 0 };
};

namespace {

// every rank has reached this point and drained its stream: no peer kernel can still be storing into this rank's buffer
void comm_barrier(ctx &c) {
    comm_state &cs = *c.comm;
    if (cs.comm == nullptr || c.world <= 1) { return; }
    cs.bar.reserve(sizeof(float));
    if (cudaMemsetAsync(cs.bar.p, 0, sizeof(float), c.stream) != cudaSuccess) { return; }
    if (api().AllReduce(cs.bar.p, cs.bar.p, 1, NCCL_FLOAT32, NCCL_SUM, cs.comm, c.stream) != NCCL_SUCCESS) { return; }
    cudaStreamSynchronize(c.stream);
}

// collective: the exported buffer is unmapped by the peers and freed only after a cross-rank barrier (a peer's queued producer
// may still be storing into it otherwise)
void xchg_release(ctx &c) {
    comm_state &cs = *c.comm;
    if (cs.xchg_ok) {
        cudaStreamSynchronize(c.stream);
        comm_barrier(c);
        for (int r = 0; r < cs.world; ++r) {
            if (r != cs.rank && cs.peers[r] != nullptr) { cudaIpcCloseMemHandle(cs.peers[r]); }
        }
        comm_barrier(c);  // everybody has unmapped before anybody frees
    }
    if (cs.local != nullptr) { cudaFree(cs.local); }
    if (cs.ticket != nullptr) { cudaFree(cs.ticket); }
    if (cs.dseq != nullptr) { cudaFree(cs.dseq); }
    cs.local = nullptr;
    cs.ticket = nullptr;
    cs.dseq = nullptr;
    cs.xchg_ok = false;
}

// collective: every rank calls this at the same point (first exchange of at least `count` reals)
void xchg_setup(ctx &c, int64_t count) {
    comm_state &cs = *c.comm;
    cs.xchg_tried = true;
    cs.world = c.world;
    cs.rank = c.rank;
    const char *env = std::getenv("B200SVM_PEER_EXCHANGE");
    bool want = c.world <= MAX_PEERS && !(env != nullptr && env[0] == '0') && c.opt_peer_exchange != 1;
    cs.ld = (count + 255) / 256 * 256;
    const size_t data_bytes = (size_t) 2 * (size_t) c.world * (size_t) cs.ld * c.esize();
    cs.flags_offset = (data_bytes + 255) / 256 * 256;
    const size_t bytes = cs.flags_offset + (size_t) 2 * c.world * sizeof(unsigned long long);
    cudaIpcMemHandle_t mine{};
    int ok = want ? 1 : 0;
    if (ok) {
        ok = cudaMalloc(&cs.local, bytes) == cudaSuccess && cudaMalloc(&cs.ticket, sizeof(unsigned int)) == cudaSuccess &&
             cudaMalloc(reinterpret_cast<void **>(&cs.dseq), sizeof(unsigned long long)) == cudaSuccess;
        if (ok) {
            ok = cudaMemsetAsync(cs.local, 0, bytes, c.stream) == cudaSuccess && cudaMemsetAsync(cs.ticket, 0, sizeof(unsigned int), c.stream) == cudaSuccess &&
                 cudaMemsetAsync(cs.dseq, 0, sizeof(unsigned long long), c.stream) == cudaSuccess && cudaIpcGetMemHandle(&mine, cs.local) == cudaSuccess;
        }
    }
    // all-gather (handle, ok) through NCCL — setup only
    struct item {
        cudaIpcMemHandle_t h;
        int ok;
        int pad[15];
    };
    static_assert(sizeof(item) == 128, "item layout");
    item send{};
    send.h = mine;
    send.ok = ok;
    dbuf d_send, d_recv;
    d_send.reserve(sizeof(item));
    d_recv.reserve(sizeof(item) * (size_t) c.world);
    B200_CUDA_CHECK(cudaMemcpyAsync(d_send.p, &send, sizeof(item), cudaMemcpyHostToDevice, c.stream));
    nccl_check(api().AllGather(d_send.p, d_recv.p, sizeof(item), /* ncclUint8 */ 1, cs.comm, c.stream), "ncclAllGather");
    std::vector<item> all((size_t) c.world);
    B200_CUDA_CHECK(cudaMemcpyAsync(all.data(), d_recv.p, sizeof(item) * (size_t) c.world, cudaMemcpyDeviceToHost, c.stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
    bool everyone = true;
    for (const item &it : all) { everyone = everyone && it.ok != 0; }
    int opened = everyone ? 1 : 0;
    if (everyone) {
        for (int r = 0; r < c.world; ++r) {
            cs.peers[r] = nullptr;
            if (r == c.rank) {
                cs.peers[r] = cs.local;
            } else if (cudaIpcOpenMemHandle(&cs.peers[r], all[(size_t) r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                (void) cudaGetLastError();
                cs.peers[r] = nullptr;
                opened = 0;
            }
        }
    }
    // second round: did every rank map every peer?  (sum of the flags == world)
    float flag = (float) opened, total = 0.f;
    B200_CUDA_CHECK(cudaMemcpyAsync(d_send.p, &flag, sizeof(float), cudaMemcpyHostToDevice, c.stream));
    nccl_check(api().AllReduce(d_send.p, d_send.p, 1, NCCL_FLOAT32, NCCL_SUM, cs.comm, c.stream), "ncclAllReduce");
    B200_CUDA_CHECK(cudaMemcpyAsync(&total, d_send.p, sizeof(float), cudaMemcpyDeviceToHost, c.stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
    cs.xchg_ok = everyone && (int) (total + 0.5f) == c.world;
    if (!cs.xchg_ok) {
        if (everyone) {
            for (int r = 0; r < c.world; ++r) {
                if (r != c.rank && cs.peers[r] != nullptr) { cudaIpcCloseMemHandle(cs.peers[r]); }
            }
        }
        if (cs.local != nullptr) { cudaFree(cs.local); }
        if (cs.ticket != nullptr) { cudaFree(cs.ticket); }
        if (cs.dseq != nullptr) { cudaFree(cs.dseq); }
        cs.local = nullptr;
        cs.ticket = nullptr;
        cs.dseq = nullptr;
        B200_REQUIRE(c.opt_peer_exchange != 2, "peer-memory exchange was required but CUDA IPC mapping of the peers failed");
    }
}

void ensure_xchg(ctx &c, int64_t count) {
    if (c.comm == nullptr || c.comm->comm == nullptr) { fail(B200SVM_ERR_NCCL, "NCCL communicator not initialised"); }
    comm_state &cs = *c.comm;
    if (!cs.xchg_tried || (cs.xchg_ok && count > cs.ld)) {
        if (cs.xchg_ok) { xchg_release(c); }
        xchg_setup(c, count);
    }
}

}  // namespace

// collective when the handle owns a communicator (barrier before the exchange buffers are released)
void comm_destroy(ctx &c) {
    if (c.comm != nullptr) {
        cudaStreamSynchronize(c.stream);
        xchg_release(c);
        if (c.comm->comm != nullptr) { api().CommDestroy(c.comm->comm); }
        delete c.comm;
        c.comm = nullptr;
    }
    c.rank = 0;
    c.world = 1;
}

// Prepare one exchange of `count` reals per slot.  plan.peer == true: the caller's producer kernel stores through plan.push (or
// comm_push does) and the O(m) epilogue consumes plan.view; false: no peer mapping — the caller sums with comm_allreduce_sum.
// Collective on first use (and when `count` outgrows the buffers).
xchg_plan comm_exchange_begin(ctx &c, int64_t count, int mode) {
    xchg_plan plan{};
    if (c.world <= 1 || count == 0) { return plan; }
    ensure_xchg(c, count);
    comm_state &cs = *c.comm;
    if (!cs.xchg_ok) { return plan; }
    plan.peer = true;
    for (int r = 0; r < c.world; ++r) { plan.push.peer[r] = cs.peers[r]; }
    plan.push.world = c.world;
    plan.push.rank = c.rank;
    plan.push.ld = cs.ld;
    plan.push.flags_offset = cs.flags_offset;
    plan.push.dseq = cs.dseq;
    plan.push.ticket = cs.ticket;
    plan.view.base = cs.local;
    plan.view.ld = cs.ld;
    plan.view.slots = mode == XCHG_GATHER ? 1 : c.world;
    plan.view.world = c.world;
    plan.view.parity_stride = (int64_t) c.world * cs.ld;
    plan.view.flags = reinterpret_cast<const unsigned long long *>(static_cast<char *>(cs.local) + cs.flags_offset);
    plan.view.dseq = cs.dseq;
    ++c.stats.collectives;
    ++c.stats.peer_exchanges;
    return plan;
}

bool comm_peer_mode(ctx &c, int64_t count) {
    if (c.world <= 1 || count == 0) { return false; }
    ensure_xchg(c, count);
    return c.comm->xchg_ok;
}

// generic producer of an exchange prepared by comm_exchange_begin: P[lo, hi) of this rank -> every peer
void comm_push(ctx &c, const xchg_plan &plan, const void *P, int64_t lo, int64_t hi, int mode, const void *done) {
    const int64_t count = hi - lo;
    const int grid = (int) std::max<int64_t>(1, std::min<int64_t>((count + 2047) / 2048, 2 * NUM_SMS));
    const int slot = mode == XCHG_GATHER ? 0 : c.rank;
    if (c.dtype == B200SVM_F32) {
        k_xchg_push<float><<<grid, 256, 0, c.stream>>>(static_cast<const float *>(P), lo, hi, plan.push, slot, static_cast<const float *>(done));
    } else {
        k_xchg_push<double><<<grid, 256, 0, c.stream>>>(static_cast<const double *>(P), lo, hi, plan.push, slot, static_cast<const double *>(done));
    }
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
    c.stats.collective_bytes += count * (int64_t) c.esize() * (int64_t) (c.world - 1);
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

// in-place all-gather over NVLink: rank r's `bytes_per_rank` bytes already sit at buf + r * bytes_per_rank; afterwards every rank holds all
// shares (set_data: every rank uploads 1 / world of the training matrix over PCIe and fetches the rest from its peers)
void comm_allgather_inplace(ctx &c, void *buf, size_t bytes_per_rank) {
    if (c.world <= 1 || bytes_per_rank == 0) { return; }
    if (c.comm == nullptr || c.comm->comm == nullptr) { fail(B200SVM_ERR_NCCL, "NCCL communicator not initialised"); }
    nccl_check(api().AllGather(static_cast<char *>(buf) + (size_t) c.rank * bytes_per_rank, buf, bytes_per_rank, /* ncclUint8 */ 1, c.comm->comm, c.stream),
               "ncclAllGather");
    ++c.stats.collectives;
    c.stats.collective_bytes += (int64_t) bytes_per_rank * (int64_t) (c.world - 1);
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
