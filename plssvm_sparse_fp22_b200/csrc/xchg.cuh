// Peer-memory exchange over NVLink: device-side producer / consumer helpers shared by the kernels that fuse the
// exchange with their own work (gram_tc.cu: partial-vector reduction + push; cg.cu: wait + sum inside the O(m) CG kernels).
//
// Every rank owns one buffer  [2 parities][world slots][ld reals] | [2 parities][world] sequence flags  that all peers map
// through CUDA IPC.  An exchange is numbered by a DEVICE-resident counter `dseq` that only advances when the exchange is really
// performed: a producer that finds the CG convergence flag set returns without pushing and without advancing it, and the
// matching consumer returns without waiting.  The flag is part of the replicated CG state (bit-identical on every rank), so all
// ranks skip the same exchanges and their counters stay equal.  Performed exchanges therefore carry consecutive numbers, which is
// what makes two parities sufficient: the producer of exchange k + 2 on any rank runs after that rank's consumer of k + 1, which
// waited for every peer's producer of k + 1, each of which ran after that peer's consumer of k.
#pragma once

#include <cstddef>
#include <cstdint>

namespace b200 {

constexpr int MAX_PEERS = 16;

// producer side (kernel argument, by value)
struct xchg_push {
    void *peer[MAX_PEERS];        // base of every rank's exchange buffer as mapped into this process (peer[rank] = the local one)
    int world{ 1 }, rank{ 0 };
    int64_t ld{ 0 };              // reals per slot
    size_t flags_offset{ 0 };     // byte offset of the flags inside a buffer
    unsigned long long *dseq{ nullptr };  // exchanges performed so far (local device memory)
    unsigned int *ticket{ nullptr };      // last-block ticket of the producer grid
};

// consumer side
//   sum mode   : `slots` = world partial vectors of stride ld; the consumer adds them in rank order
//   gather mode: `slots` = 1; every rank stored its own rows [lo, hi) of the one vector
struct xchg_view {
    void *base{ nullptr };        // parity 0 data of the local buffer (nullptr: no peer exchange — the vector was all-reduced in place)
    int64_t ld{ 0 };
    int slots{ 0 };
    int world{ 1 };
    int64_t parity_stride{ 0 };   // reals between the two parities
    const unsigned long long *flags{ nullptr };  // [2][world]
    const unsigned long long *dseq{ nullptr };
};

#ifdef __CUDACC__

// number of the exchange this producer grid performs (every block reads the counter before it takes its ticket, the counter is
// advanced by the last block only)
__device__ __forceinline__ unsigned long long xchg_next_seq(const xchg_push &x) { return *reinterpret_cast<volatile unsigned long long *>(x.dseq) + 1ull; }

template <typename T>
__device__ __forceinline__ T *xchg_slot(const xchg_push &x, int peer, unsigned long long seq, int slot) {
    return static_cast<T *>(x.peer[peer]) + ((size_t) (seq & 1ull) * (size_t) x.world + (size_t) slot) * (size_t) x.ld;
}

// Call by ALL threads of EVERY block after the block's stores: the last block to arrive publishes `seq` in every peer's flag of
// this rank and advances the local counter.  `is_last` is a shared-memory scratch flag.
__device__ __forceinline__ void xchg_publish(const xchg_push &x, unsigned long long seq, bool *is_last) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(x.ticket, 1u);
        *is_last = (t == gridDim.x * gridDim.y - 1);
    }
    __syncthreads();
    if (*is_last) {
        __threadfence_system();  // every block's stores (ordered before its ticket) are visible system-wide before the flags
        if ((int) threadIdx.x < x.world) {
            volatile unsigned long long *flag = reinterpret_cast<volatile unsigned long long *>(static_cast<char *>(x.peer[threadIdx.x]) + x.flags_offset) +
                                                (size_t) (seq & 1ull) * x.world + x.rank;
            *flag = seq;
        }
        if (threadIdx.x == 0) {
            *x.ticket = 0u;
            *reinterpret_cast<volatile unsigned long long *>(x.dseq) = seq;
        }
    }
}

// Consumer: block until every source has published the current exchange; returns the parity-selected data base.
// Call by all threads of a block (contains __syncthreads).
template <typename T>
__device__ __forceinline__ const T *xchg_wait(const xchg_view &v) {
    const unsigned long long seq = *reinterpret_cast<const volatile unsigned long long *>(v.dseq);  // advanced by the producer grid that ran before us
    const int parity = (int) (seq & 1ull);
    if ((int) threadIdx.x < v.world) {
        const volatile unsigned long long *f = v.flags + (size_t) parity * v.world + threadIdx.x;
        while (*f < seq) {}
        __threadfence_system();
    }
    __syncthreads();
    return static_cast<const T *>(v.base) + (size_t) parity * (size_t) v.parity_stride;
}

#endif  // __CUDACC__

}  // namespace b200
