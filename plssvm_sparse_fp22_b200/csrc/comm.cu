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
// Peer-memory exchange: the allreduce of the matvec result without NCCL on the critical path.
// Every rank owns a buffer  [2 parities][world slots][ld reals] + [2][world] sequence flags, mapped into all peers through
// CUDA IPC.  Producer (k_xchg_push): stores this rank's partial vector into slot `rank` of EVERY peer's buffer over NVLink,
// then the last block publishes the sequence number in every peer's flag (system-scope fence before it).  Consumer
// (k_abar_finish): one thread per source rank spins on the local flags until they reach the sequence number, then the
// `world` slots are summed in rank order — the same order on every rank, so the replicated CG state stays bit-identical.
// Two parities suffice: the producer of exchange k + 2 on any rank runs after that rank's consumer of k + 1, which waited
// for every peer's producer of k + 1, which ran after that peer's consumer of k.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int MAX_PEERS = 16;

struct peer_table {
    void *base[MAX_PEERS];
};

template <typename T>
__global__ void __launch_bounds__(256) k_xchg_push(const T *__restrict__ P, int64_t m, peer_table peers, int world, int rank, int64_t ld, int parity,
                                                   unsigned long long seq, size_t flags_offset, unsigned int *ticket) {
    __shared__ bool is_last;
    const size_t slot = ((size_t) parity * (size_t) world + (size_t) rank) * (size_t) ld;
    for (int r = 0; r < world; ++r) {
        T *dst = static_cast<T *>(peers.base[r]) + slot;
        for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) { dst[i] = P[i]; }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last && threadIdx.x < world) {
        __threadfence_system();  // every block's stores (ordered before its ticket) are visible system-wide before the flag
        volatile unsigned long long *flag =
            reinterpret_cast<volatile unsigned long long *>(static_cast<char *>(peers.base[threadIdx.x]) + flags_offset) + (size_t) parity * world + rank;
        *flag = seq;
        if (threadIdx.x == 0) { *ticket = 0u; }
    }
}

struct comm_state {
    nccl_comm_t comm{ nullptr };
    // peer exchange
    bool xchg_tried{ false }, xchg_ok{ false };
    int64_t ld{ 0 };            // reals per slot
    size_t flags_offset{ 0 };   // byte offset of the flags inside a buffer
    void *local{ nullptr };
    peer_table peers{};
    unsigned long long seq{ 0 };
    unsigned int *ticket{ nullptr };
    int world{ 1 }, rank{ 0 };
};

namespace {

void xchg_release(comm_state &cs) {
    if (cs.xchg_ok) {
        for (int r = 0; r < cs.world; ++r) {
            if (r != cs.rank && cs.peers.base[r] != nullptr) { cudaIpcCloseMemHandle(cs.peers.base[r]); }
        }
    }
    if (cs.local != nullptr) { cudaFree(cs.local); }
    if (cs.ticket != nullptr) { cudaFree(cs.ticket); }
    cs.local = nullptr;
    cs.ticket = nullptr;
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
        ok = cudaMalloc(&cs.local, bytes) == cudaSuccess && cudaMalloc(&cs.ticket, sizeof(unsigned int)) == cudaSuccess;
        if (ok) {
            ok = cudaMemsetAsync(cs.local, 0, bytes, c.stream) == cudaSuccess && cudaMemsetAsync(cs.ticket, 0, sizeof(unsigned int), c.stream) == cudaSuccess &&
                 cudaIpcGetMemHandle(&mine, cs.local) == cudaSuccess;
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
            cs.peers.base[r] = nullptr;
            if (r == c.rank) {
                cs.peers.base[r] = cs.local;
            } else if (cudaIpcOpenMemHandle(&cs.peers.base[r], all[(size_t) r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                (void) cudaGetLastError();
                cs.peers.base[r] = nullptr;
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
                if (r != c.rank && cs.peers.base[r] != nullptr) { cudaIpcCloseMemHandle(cs.peers.base[r]); }
            }
        }
        if (cs.local != nullptr) { cudaFree(cs.local); }
        if (cs.ticket != nullptr) { cudaFree(cs.ticket); }
        cs.local = nullptr;
        cs.ticket = nullptr;
        B200_REQUIRE(c.opt_peer_exchange != 2, "peer-memory exchange was required but CUDA IPC mapping of the peers failed");
    }
    cs.seq = 0;
}

}  // namespace

void comm_destroy(ctx &c) {
    if (c.comm != nullptr) {
        cudaStreamSynchronize(c.stream);
        xchg_release(*c.comm);
        if (c.comm->comm != nullptr) { api().CommDestroy(c.comm->comm); }
        delete c.comm;
        c.comm = nullptr;
    }
    c.rank = 0;
    c.world = 1;
}

// Sum over ranks of the partial vector P (count reals of the handle's type).  Peer exchange: P is pushed to every rank and the
// caller's O(m) epilogue sums `world` slots (returned in *x); NCCL fallback: in-place allreduce, x->base == nullptr.
void comm_exchange(ctx &c, void *P, int64_t count, xchg_view *x) {
    x->base = nullptr;
    if (c.world <= 1 || count == 0) { return; }
    if (c.comm == nullptr || c.comm->comm == nullptr) { fail(B200SVM_ERR_NCCL, "NCCL communicator not initialised"); }
    comm_state &cs = *c.comm;
    if (!cs.xchg_tried || (cs.xchg_ok && count > cs.ld)) {
        if (cs.xchg_ok) {
            B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
            xchg_release(cs);
        }
        xchg_setup(c, count);
    }
    if (!cs.xchg_ok) {
        comm_allreduce_sum(c, P, count);
        return;
    }
    ++cs.seq;
    const int parity = (int) (cs.seq & 1ull);
    const int grid = (int) std::max<int64_t>(1, std::min<int64_t>((count + 2047) / 2048, 2 * NUM_SMS));
    if (c.dtype == B200SVM_F32) {
        k_xchg_push<float><<<grid, 256, 0, c.stream>>>(static_cast<const float *>(P), count, cs.peers, c.world, c.rank, cs.ld, parity, cs.seq, cs.flags_offset,
                                                       cs.ticket);
    } else {
        k_xchg_push<double><<<grid, 256, 0, c.stream>>>(static_cast<const double *>(P), count, cs.peers, c.world, c.rank, cs.ld, parity, cs.seq,
                                                        cs.flags_offset, cs.ticket);
    }
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
    ++c.stats.collectives;
    ++c.stats.peer_exchanges;
    c.stats.collective_bytes += count * (int64_t) c.esize() * (int64_t) (c.world - 1);
    x->base = static_cast<char *>(cs.local) + (size_t) parity * (size_t) c.world * (size_t) cs.ld * c.esize();
    x->ld = cs.ld;
    x->slots = c.world;
    x->flags = reinterpret_cast<const unsigned long long *>(static_cast<char *>(cs.local) + cs.flags_offset) + (size_t) parity * c.world;
    x->seq = cs.seq;
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
