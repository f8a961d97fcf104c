// Device-resident Conjugate Gradient (restates src/plssvm/backends/OpenMP/csvm.cpp:82-170; the reference's GPU variant keeps
// the vector algebra on the host with 4-5 PCIe round trips per iteration, src/plssvm/backends/gpu_csvm.cpp:259-309).
//
// All vectors and scalars live in HBM.  Per iteration: [sums of d] -> [K d] -> [Abar epilogue + d.Ad] -> [x, r update + r.r]
// -> [d update + sums of d]; every reduction is a fused two-stage deterministic grid sum (device_utils.cuh: grid_sum).
// The host only reads one scalar (delta) per iteration for the stop test `delta <= eps^2 delta0` (csvm.cpp:145).
#include "device_utils.cuh"

#include <algorithm>

namespace b200 {

constexpr int VEC_THREADS = 256;

static int vec_grid(int64_t m) {
    int64_t g = (m + VEC_THREADS * 4 - 1) / (VEC_THREADS * 4);
    if (g > MAX_PARTIAL_BLOCKS) { g = MAX_PARTIAL_BLOCKS; }
    if (g < 1) { g = 1; }
    return (int) g;
}

// scal[idx_sum] = sum v ; scal[idx_qv] = q . v
template <typename T>
__global__ void __launch_bounds__(VEC_THREADS) k_sums2(const T *__restrict__ v, const T *__restrict__ q, int64_t m, T *__restrict__ scal,
                                                       int idx_sum, int idx_qv, T *partials, unsigned int *ticket) {
    __shared__ T red[32];
    T acc[2] = { T(0), T(0) };
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) {
        const T vi = v[i];
        acc[0] += vi;
        acc[1] = fma(q[i], vi, acc[1]);
    }
    T out[2];
    if (grid_sum<T, 2>(acc, partials, ticket, red, out)) {
        scal[idx_sum] = out[0];
        scal[idx_qv] = out[1];
    }
}

// Final O(m) pass of the implicit matrix.  The O(m^2) kernels deliver, per entry, the reference's centred value
// k(x_i,x_j) + QA_cost - q_i - q_j (svm_kernel.cpp:32) times d_j, summed into partial vectors P[s][i]; here the diagonal
// term of svm_kernel.cpp:33-34 is added:  (Abar d)_i = sum_s P[s][i] + d_i / C  (+ sum(d) / C for the linear Z Z^T form).
// MODE 0: out = Abar d,            scal[S_DAD] = d . out                                  (CG:  Ad, d.Ad)
// MODE 1: out += add * Abar d,     delta_old = delta, delta = out.out, beta = delta/old   (CG:  residual refresh, csvm.cpp:130-135)
// MODE 2: out += add * Abar d,     delta = delta0 = out.out                               (CG:  initial residual, csvm.cpp:94-98)
// MODE 3: out += add * Abar d                                                             (run_device_kernel test hook)
// CENTER 0: P already holds the per-entry centred sums (dense tile kernels)
//        1: dense linear Z Z^T form: + sum(d) / C
//        2: P holds the un-centred K d (sparse paths): + (QA_cost - q_i) sum(d) - q.d   (the rank-1 part of svm_kernel.cpp:32)
//        3: P holds the raw S v of the sparse dense-tile path: K d = u_i (sum v + S v) (rbf) or c0^deg sum v + S v (polynomial) is
//           formed here, AFTER the partial vectors of the ranks have been combined, then as 2
// Several ranks: P / splits describe the slots of the peer exchange (xchg.cuh); the kernel waits for every source first.
template <typename T>
struct finish_args {
    const T *P;        // local partial vectors (x.base == nullptr) ...
    int64_t ldP;
    int splits;
    int center;
    xchg_view x;       // ... or the exchange buffer of this rank
    const T *d;
    const T *q;
    int64_t m;
    T inv_cost, qa, add, target;
    T *out;
    T *scal;
    T *partials;
    unsigned int *ticket;
    T *mailbox;        // pinned host slot that receives delta (MODE 1), or nullptr
    const T *u;        // CENTER 3: u_i = exp(-gamma |x_i|^2) (rbf)
    T c0pow;           // CENTER 3: coef0^degree (polynomial)
    int rbf;
};

// CENTER 2 / 3: the un-centred K d of element i -> + the rank-1 centring
template <typename T>
__device__ __forceinline__ T centre_sparse(const finish_args<T> &a, T kd, int64_t i, T sum_d, T q_d, T sum_v) {
    if (a.center == 3) { kd = a.rbf ? a.u[i] * (sum_v + kd) : fma(a.c0pow, sum_v, kd); }
    return kd + fma(a.qa - a.q[i], sum_d, -q_d);
}

// (K d)_i from the partial vectors
template <typename T>
__device__ __forceinline__ T gather_kd(const T *__restrict__ P, int64_t ldP, int splits, bool remote, int64_t i) {
    T kd = T(0);
    if (remote) {
        // the ranks' slots, read past L1 (peer-written), combined in a fixed order that is identical on every rank: the balanced tree
        // of group_tree (gram_tc.cu) for 2 / 4 / 8 ranks — with the tile paths' canonical groups the sum is then bit-identical to the
        // one a single GPU forms — else rank order
        if (splits == 2 || splits == 4 || splits == 8) {
            T v[8];
#pragma unroll
            for (int s = 0; s < 8; ++s) { v[s] = s < splits ? __ldcg(&P[(int64_t) s * ldP + i]) : T(0); }
            kd = splits == 8 ? ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7])) : (splits == 4 ? (v[0] + v[1]) + (v[2] + v[3]) : v[0] + v[1]);
        } else {
            for (int s = 0; s < splits; ++s) { kd += __ldcg(&P[(int64_t) s * ldP + i]); }
        }
    } else {
        for (int s = 0; s < splits; ++s) { kd += P[(int64_t) s * ldP + i]; }
    }
    return kd;
}

template <typename T, int MODE>
__global__ void __launch_bounds__(VEC_THREADS) k_abar_finish(const finish_args<T> a) {
    __shared__ T red[32];
    if constexpr (MODE == 0 || MODE == 1) {
        if (a.scal[S_DONE] != T(0)) { return; }  // CG has converged: this queued iteration (and its exchange, xchg.cuh) is skipped on every rank
    }
    const bool remote = a.x.base != nullptr;
    const T *P = remote ? xchg_wait<T>(a.x) : a.P;
    const int64_t ldP = remote ? a.x.ld : a.ldP;
    const int splits = remote ? a.x.slots : a.splits;
    const T sum_d = a.center != 0 ? a.scal[S_SUM_D] : T(0);
    const T q_d = a.center >= 2 ? a.scal[S_Q_D] : T(0);
    const T sum_v = a.center == 3 ? a.scal[S_TMP0] : T(0);
    const T rank1 = a.center == 1 ? sum_d * a.inv_cost : T(0);
    T acc[1] = { T(0) };
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < a.m; i += (int64_t) gridDim.x * blockDim.x) {
        T kd = gather_kd<T>(P, ldP, splits, remote, i);
        const T di = a.d[i];
        if (a.center >= 2) { kd = centre_sparse(a, kd, i, sum_d, q_d, sum_v); }
        const T val = kd + rank1 + di * a.inv_cost;
        if constexpr (MODE == 0) {
            a.out[i] = val;
            acc[0] = fma(di, val, acc[0]);
        } else {
            const T r = fma(a.add, val, a.out[i]);
            a.out[i] = r;
            acc[0] = fma(r, r, acc[0]);
        }
    }
    if constexpr (MODE != 3) {
        T res[1];
        if (grid_sum<T, 1>(acc, a.partials, a.ticket, red, res)) {
            if constexpr (MODE == 0) {
                a.scal[S_DAD] = res[0];
            } else if constexpr (MODE == 1) {
                const T old = a.scal[S_DELTA];
                a.scal[S_DELTA_OLD] = old;
                a.scal[S_DELTA] = res[0];
                a.scal[S_BETA] = res[0] / old;
                if (res[0] <= a.target) { a.scal[S_DONE] = T(1); }  // csvm.cpp:145
                if (a.mailbox != nullptr) { *reinterpret_cast<volatile T *>(a.mailbox) = res[0]; }
            } else {
                a.scal[S_DELTA] = res[0];
                a.scal[S_DELTA0] = res[0];
            }
        }
    }
}

// alpha_cd = delta / d.Ad ; x += alpha_cd d ; (REFRESH == 0:) r -= alpha_cd Ad, delta_old = delta, delta = r.r, beta = delta / delta_old
template <typename T, int REFRESH>
__global__ void __launch_bounds__(VEC_THREADS) k_cg_update(T *__restrict__ x, T *__restrict__ r, const T *__restrict__ d,
                                                           const T *__restrict__ Ad, int64_t m, T *__restrict__ scal, T *partials,
                                                           unsigned int *ticket, T target, T *mailbox) {
    __shared__ T red[32];
    if (scal[S_DONE] != T(0)) { return; }  // converged at an earlier iteration: alpha stays frozen
    const T delta = scal[S_DELTA];
    const T alpha_cd = delta / scal[S_DAD];  // csvm.cpp:125
    T acc[1] = { T(0) };
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) {
        x[i] = fma(alpha_cd, d[i], x[i]);  // csvm.cpp:128
        if constexpr (REFRESH == 0) {
            const T ri = fma(-alpha_cd, Ad[i], r[i]);  // csvm.cpp:138
            r[i] = ri;
            acc[0] = fma(ri, ri, acc[0]);
        }
    }
    if constexpr (REFRESH == 0) {
        T res[1];
        if (grid_sum<T, 1>(acc, partials, ticket, red, res)) {
            scal[S_ALPHA_CD] = alpha_cd;
            scal[S_DELTA_OLD] = delta;         // csvm.cpp:142
            scal[S_DELTA] = res[0];            // csvm.cpp:143
            scal[S_BETA] = res[0] / delta;     // csvm.cpp:153
            if (res[0] <= target) { scal[S_DONE] = T(1); }  // csvm.cpp:145 (set by the last block: no block of this grid reads it afterwards)
            if (mailbox != nullptr) { *reinterpret_cast<volatile T *>(mailbox) = res[0]; }
        }
    }
}

// d = beta d + r (csvm.cpp:155) ; scal[S_SUM_D] = sum d ; scal[S_Q_D] = q.d (for the next matvec's rank-1 epilogue)
// pk != nullptr (tcgen05 path): the packed per-index operand {c_i, q_i, d_i, 0} of the tile kernels gets the new d_i
template <typename T>
__global__ void __launch_bounds__(VEC_THREADS) k_cg_direction(T *__restrict__ d, const T *__restrict__ r, const T *__restrict__ q, int64_t m,
                                                              T *__restrict__ scal, T *partials, unsigned int *ticket, float4 *__restrict__ pk) {
    __shared__ T red[32];
    if (scal[S_DONE] != T(0)) { return; }  // (the direction of a converged solve is never used again)
    const T beta = scal[S_BETA];
    T acc[2] = { T(0), T(0) };
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) {
        const T di = fma(beta, d[i], r[i]);
        d[i] = di;
        if (pk != nullptr) { pk[i].z = (float) di; }
        acc[0] += di;
        acc[1] = fma(q[i], di, acc[1]);
    }
    T out[2];
    if (grid_sum<T, 2>(acc, partials, ticket, red, out)) {
        scal[S_SUM_D] = out[0];
        scal[S_Q_D] = out[1];
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// One ordinary CG iteration after the O(m^2) kernel as ONE kernel (csvm.cpp:121-155 without the refresh branch):
//   A  Ad = sum of the partial vectors (+ centring, + d / C), d.Ad          [several ranks: wait for the peers' slots first]
//   B  alpha_cd = delta / d.Ad ; x += alpha_cd d ; r -= alpha_cd Ad ; delta' = r.r ; stop flag
//   C  beta = delta' / delta ; d = beta d + r (+ the packed operand of the tile kernels) ; sum d, q.d
// Phases are separated by grid-wide barriers; every scalar is a two-stage sum in a fixed order (block partials in global memory,
// every block adds them in the same order), so all blocks — and all ranks — hold bit-identical values.  delta' goes straight to
// the host's pinned mailbox: no copy-engine operation between two iterations.
// ---------------------------------------------------------------------------------------------------------------------
template <typename T>
struct step_args {
    finish_args<T> f;   // P / exchange view, d, q, m, inv_cost, qa, target, out = Ad, scal, partials ([4][MAX_PARTIAL_BLOCKS]), mailbox
    T *x, *r, *dvec;
    float4 *pk;
    unsigned int *bar;  // grid barrier state (two words)
};

// Grid-wide barrier of a grid whose blocks are all co-resident (at most one per SM, launched alone on the stream): sense-reversing
// and self-resetting, bar[0] = arrivals, bar[1] = generation.  A plain launch with this barrier measured faster than a
// cooperative launch with grid.sync() (the cooperative launch itself costs ~30 us per iteration: profiles/r2_summary.md).
__device__ __forceinline__ void grid_barrier(unsigned int *bar) {
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile unsigned int *vbar = bar;
        const unsigned int gen = vbar[1];
        __threadfence();
        if (atomicAdd(bar, 1u) == gridDim.x - 1) {
            vbar[0] = 0u;
            __threadfence();
            atomicAdd(bar + 1, 1u);
        } else {
            while (vbar[1] == gen) {}
        }
        __threadfence();
    }
    __syncthreads();
}

// sum of `count` block partials in a fixed order; result in every thread of the block
template <typename T>
__device__ __forceinline__ T total_of(const T *partials, int count, T *red) {
    T s = T(0);
    if (threadIdx.x < 32) {
        for (int b = threadIdx.x; b < count; b += 32) { s += __ldcg(&partials[b]); }
        s = warp_sum(s);
        if (threadIdx.x == 0) { red[0] = s; }
    }
    __syncthreads();
    s = red[0];
    __syncthreads();
    return s;
}

template <typename T>
__global__ void __launch_bounds__(VEC_THREADS) k_cg_step(const step_args<T> a) {
    __shared__ T red[32];
    const finish_args<T> &f = a.f;
    if (f.scal[S_DONE] != T(0)) { return; }  // uniform over the grid: the flag is only written behind the last barrier of such a grid
    const bool remote = f.x.base != nullptr;
    const T *P = remote ? xchg_wait<T>(f.x) : f.P;
    const int64_t ldP = remote ? f.x.ld : f.ldP;
    const int splits = remote ? f.x.slots : f.splits;
    const T delta = f.scal[S_DELTA];
    const T sum_d = f.center != 0 ? f.scal[S_SUM_D] : T(0);
    const T q_d = f.center >= 2 ? f.scal[S_Q_D] : T(0);
    const T sum_v = f.center == 3 ? f.scal[S_TMP0] : T(0);
    const T rank1 = f.center == 1 ? sum_d * f.inv_cost : T(0);
    const int64_t first = (int64_t) blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t) gridDim.x * blockDim.x;
    const int nblk = (int) gridDim.x;
    T *Ad = f.out;
    // ---- A
    T acc = T(0);
    for (int64_t i = first; i < f.m; i += stride) {
        T kd = gather_kd<T>(P, ldP, splits, remote, i);
        const T di = a.dvec[i];
        if (f.center >= 2) { kd = centre_sparse(f, kd, i, sum_d, q_d, sum_v); }
        const T val = kd + rank1 + di * f.inv_cost;
        Ad[i] = val;
        acc = fma(di, val, acc);
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) { f.partials[blockIdx.x] = acc; }
    grid_barrier(a.bar);
    const T dAd = total_of(f.partials, nblk, red);
    // ---- B
    const T alpha_cd = delta / dAd;  // csvm.cpp:125
    acc = T(0);
    for (int64_t i = first; i < f.m; i += stride) {
        a.x[i] = fma(alpha_cd, a.dvec[i], a.x[i]);       // csvm.cpp:128
        const T ri = fma(-alpha_cd, Ad[i], a.r[i]);      // csvm.cpp:138
        a.r[i] = ri;
        acc = fma(ri, ri, acc);
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) { f.partials[MAX_PARTIAL_BLOCKS + blockIdx.x] = acc; }
    grid_barrier(a.bar);
    const T delta_new = total_of(f.partials + MAX_PARTIAL_BLOCKS, nblk, red);
    // ---- C
    const T beta = delta_new / delta;  // csvm.cpp:153
    T s0 = T(0), s1 = T(0);
    for (int64_t i = first; i < f.m; i += stride) {
        const T di = fma(beta, a.dvec[i], a.r[i]);  // csvm.cpp:155
        a.dvec[i] = di;
        if (a.pk != nullptr) { a.pk[i].z = (float) di; }
        s0 += di;
        s1 = fma(f.q[i], di, s1);
    }
    if (f.center != 0) {  // uniform: only the rank-1 epilogues of the linear / sparse paths need the sums of d
        s0 = block_sum(s0, red);
        if (threadIdx.x == 0) { f.partials[2 * MAX_PARTIAL_BLOCKS + blockIdx.x] = s0; }
        s1 = block_sum(s1, red);
        if (threadIdx.x == 0) { f.partials[3 * MAX_PARTIAL_BLOCKS + blockIdx.x] = s1; }
        grid_barrier(a.bar);
        if (blockIdx.x == 0) {
            s0 = total_of(f.partials + 2 * MAX_PARTIAL_BLOCKS, nblk, red);
            s1 = total_of(f.partials + 3 * MAX_PARTIAL_BLOCKS, nblk, red);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        f.scal[S_DAD] = dAd;
        f.scal[S_ALPHA_CD] = alpha_cd;
        f.scal[S_DELTA_OLD] = delta;      // csvm.cpp:142
        f.scal[S_DELTA] = delta_new;      // csvm.cpp:143
        f.scal[S_BETA] = beta;
        if (f.center != 0) {
            f.scal[S_SUM_D] = s0;
            f.scal[S_Q_D] = s1;
        }
        if (delta_new <= f.target) { f.scal[S_DONE] = T(1); }  // csvm.cpp:145
        if (f.mailbox != nullptr) { *reinterpret_cast<volatile T *>(f.mailbox) = delta_new; }
    }
}

// Multi-GPU: fold this rank's column-split partials of its rows [lo, hi) into one vector, P[0][i] = sum_s P[s][i]; the rows are
// then pushed to the peers (gather).  zero_rest (NCCL fallback: the vectors are summed with one ncclAllReduce): P[0][i] = 0 elsewhere
template <typename T>
__global__ void __launch_bounds__(VEC_THREADS) k_fold_splits(T *__restrict__ P, int64_t ldP, int splits, int64_t m, int64_t lo, int64_t hi, int zero_rest) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) {
        if (i >= lo && i < hi) {
            T s = T(0);
            for (int k = 0; k < splits; ++k) { s += P[(int64_t) k * ldP + i]; }
            P[i] = s;
        } else if (zero_rest) {
            P[i] = T(0);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(VEC_THREADS) k_fill(T *__restrict__ v, int64_t m, T value) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) { v[i] = value; }
}

// ---------------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------------
// account the (start, stop) pairs [ev_tail, upto): the caller knows that these matvecs have completed
void drain_matvec_events(ctx &c, int64_t upto) {
    for (; c.ev_tail < upto; ++c.ev_tail) {
        const int k = (int) (c.ev_tail % ctx::EV_RING);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, c.ev[2 * k], c.ev[2 * k + 1]) == cudaSuccess) {
            c.stats.matvec_ms += ms;
            ++c.stats.matvecs_timed;
        }
    }
}

// account the (start, stop) pairs around the vector kernels of the iterations [sv_tail, upto): known to be complete
static void drain_vector_events(ctx &c, int64_t upto) {
    for (; c.sv_tail < upto; ++c.sv_tail) {
        const int k = (int) (c.sv_tail % ctx::EV_RING);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, c.sv[2 * k], c.sv[2 * k + 1]) == cudaSuccess) {
            c.stats.cg_vector_ms += ms;
            ++c.stats.cg_vector_timed;
        }
    }
}

void sync_stream(ctx &c) {
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
    drain_matvec_events(c, c.ev_head);
    drain_vector_events(c, c.sv_head);
    if (c.cb_pending) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, c.cb_ev[0], c.cb_ev[1]) == cudaSuccess) { c.stats.cache_build_ms += ms; }
        c.cb_pending = false;
    }
}

template <typename T>
void sums2(ctx &c, const T *v, int idx_sum, int idx_qv) {
    const int64_t m = c.m();
    k_sums2<T><<<vec_grid(m), VEC_THREADS, 0, c.stream>>>(v, c.q.as<T>(), m, c.scal.as<T>(), idx_sum, idx_qv, c.partials.as<T>(),
                                                          c.counters.as<unsigned int>());
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}
template void sums2<float>(ctx &, const float *, int, int);
template void sums2<double>(ctx &, const double *, int, int);

namespace {

struct kd_result {
    int64_t ldP;
    int splits;
    int center;  // see k_abar_finish
    const void *u{ nullptr };  // CENTER 3
    double c0pow{ 0 };
    int rbf{ 0 };
    xchg_view x; // several ranks, peer exchange: where the partial vectors of all ranks arrive (base == nullptr: c.P holds the sum)
};

template <typename T>
bool tc_active(const ctx &c) {
    if constexpr (sizeof(T) == 4) {
        return !c.is_csr && c.kernel != B200SVM_LINEAR && ((c.opt_matvec_path == 2) || (c.opt_matvec_path == 0 && tc_available(c)));
    } else {
        return false;
    }
}

// K d for the reduced system (rows/cols 0..m-1) into c.P as `splits` partial vectors (or into the peers' exchange buffers).
//   done      : CG convergence flag checked by every kernel of the matvec (nullptr outside the CG loop)
//   have_sums : scal[S_SUM_D] / scal[S_Q_D] already describe d (written by the direction update)
//   packed    : tcgen05 path, the packed operand already holds d
template <typename T>
kd_result kernel_times_vector(ctx &c, const T *d, const T *done, bool have_sums = false, bool packed = false) {
    const int64_t m = c.m();
    const int64_t ldP = (m + 255) / 256 * 256;
    constexpr int MAXS = 64;
    kd_result res{};
    res.ldP = ldP;
    res.splits = 1;
    res.center = 0;
    const bool timed = c.ev_head - c.ev_tail < ctx::EV_RING;
    const int ev_slot = (int) (c.ev_head % ctx::EV_RING);
    if (timed) { B200_CUDA_CHECK(cudaEventRecord(c.ev[2 * ev_slot], c.stream)); }
    // row-partitioned paths over several ranks: every rank pushes its rows [lo, hi) to all peers (gather); without peer
    // mapping the rows of the others are zeroed and the vectors are summed by ncclAllReduce
    int64_t lo = 0, hi = m;
    partition_rows(m, c.rank, c.world, &lo, &hi);
    const auto exchange = [&](int mode) {
        if (c.world <= 1) { return; }
        const xchg_plan plan = comm_exchange_begin(c, m, mode);
        if (plan.peer) {
            comm_push(c, plan, c.P.p, mode == XCHG_GATHER ? lo : 0, mode == XCHG_GATHER ? hi : m, mode, done);
            res.x = plan.view;
        } else {
            comm_allreduce_sum(c, c.P.p, m);
        }
    };
    if (c.is_csr) {
        // sparse rows: un-centred K d of this rank's rows, rank-1 centring in the O(m) epilogue
        c.P.reserve((size_t) ldP * sizeof(T));
        if (!have_sums) { sums2<T>(c, d, S_SUM_D, S_Q_D); }
        const bool zero_rest = c.world > 1 && !comm_peer_mode(c, m);
        // (false: dense-tile S cache — tile-partitioned RAW partial vectors S v of full length, summed over the ranks; K d is formed
        //  from the sum by the O(m) epilogue, centre mode 3)
        const bool rows = sp_kernel_matvec<T>(c, d, c.P.as<T>(), zero_rest, done);
        exchange(rows ? XCHG_GATHER : XCHG_SUM);
        res.center = rows ? 2 : 3;
        if (!rows) { sp_finish_params(c, &res.u, &res.c0pow, &res.rbf); }
        c.stats.matvec_path = 4;
    } else if (c.kernel == B200SVM_LINEAR) {
        // centred linear kernel: sum_j (K_ij + K_nn - K_in - K_nj) d_j = Z (Z^T d), Z = X - 1 x_n^T : two GEMVs (HBM bound);
        // replicated on every rank (O(m F) per matvec: an exchange would cost as much as the GEMV pair at these sizes)
        if (!c.has_Z) {
            c.Z.reserve((size_t) m * c.ldx * sizeof(T));
            center_rows<T>(c, c.X.as<T>(), m, c.ldx, c.X.as<T>() + m * c.ldx, c.Z.as<T>());
            c.has_Z = true;
        }
        c.P.reserve((size_t) ldP * sizeof(T));
        c.wvec.reserve((size_t) c.ldx * sizeof(T));
        if (!have_sums) { sums2<T>(c, d, S_SUM_D, S_Q_D); }
        xt_v<T>(c, c.Z.as<T>(), m, c.ldx, d, c.wvec.as<T>());
        x_w<T>(c, c.Z.as<T>(), m, c.ldx, c.wvec.as<T>(), T(0), c.P.as<T>());
        res.center = 1;
        c.stats.matvec_path = 3;
    } else if (tc_active<T>(c)) {
        if constexpr (sizeof(T) == 4) {
            c.P.reserve((size_t) ldP * sizeof(T));
            // partial over this rank's tiles; several ranks: the partial-vector reduction stores straight into the peers
            xchg_plan plan{};
            if (c.world > 1) { plan = comm_exchange_begin(c, m, XCHG_SUM); }
            tc_kernel_matvec(c, reinterpret_cast<const float *>(d), c.P.as<float>(), packed, plan.peer ? &plan.push : nullptr,
                             reinterpret_cast<const float *>(done));
            if (plan.peer) {
                res.x = plan.view;
                c.stats.collective_bytes += m * (int64_t) sizeof(float) * (int64_t) (c.world - 1);
            } else if (c.world > 1) {
                comm_allreduce_sum(c, c.P.p, m);
            }
        }
        c.stats.matvec_path = 2;
    } else if (sizeof(T) == 8 && dense_f64_cache_ready(c)) {
        // dense double polynomial / rbf with the kernel matrix cached as lower-triangle tiles: symmetric HBM-bound sweep over this
        // rank's tiles; several ranks: the partial-vector reduction stores straight into the peers
        if constexpr (sizeof(T) == 8) {
            c.P.reserve((size_t) ldP * sizeof(T));
            xchg_plan plan{};
            if (c.world > 1) { plan = comm_exchange_begin(c, m, XCHG_SUM); }
            tile_store_matvec_f64(c, c.dts, c.q.as<double>(), d, c.QA_cost, m, c.P.as<double>(), plan.peer ? &plan.push : nullptr, done);
            if (plan.peer) {
                res.x = plan.view;
                c.stats.collective_bytes += m * (int64_t) sizeof(double) * (int64_t) (c.world - 1);
            } else if (c.world > 1) {
                comm_allreduce_sum(c, c.P.p, m);
            }
        }
        c.stats.matvec_path = 1;
    } else {
        // full rows [lo, hi) of the implicit matrix on this rank (all rows when world == 1)
        c.P.reserve((size_t) (MAXS * ldP) * sizeof(T));
        res.splits = 0;
        if (hi > lo) {
            res.splits = kmv_simt_launch<T>(c, c.X.as<T>() + lo * c.ldx, c.norms.as<T>() + lo, hi - lo, c.ldx, c.X.as<T>(), c.norms.as<T>(), m,
                                            c.ldx, d, c.P.as<T>() + lo, ldP, MAXS, c.q.as<T>() + lo, c.q.as<T>(), (T) c.QA_cost);
        }
        if (c.world > 1) {
            const bool zero_rest = !comm_peer_mode(c, m);
            k_fold_splits<T><<<vec_grid(m), VEC_THREADS, 0, c.stream>>>(c.P.as<T>(), ldP, res.splits, m, lo, hi, zero_rest ? 1 : 0);
            count_launch(c);
            B200_CUDA_CHECK(cudaGetLastError());
            res.splits = 1;
            exchange(XCHG_GATHER);
        }
        c.stats.matvec_path = 1;
    }
    if (timed) {
        B200_CUDA_CHECK(cudaEventRecord(c.ev[2 * ev_slot + 1], c.stream));
        ++c.ev_head;
    }
    ++c.stats.matvecs;
    return res;
}

template <typename T>
finish_args<T> make_finish_args(ctx &c, const kd_result &kd, const T *d, T add, T *out, T target, T *mailbox) {
    finish_args<T> a{};
    a.P = c.P.as<T>();
    a.ldP = kd.ldP;
    a.splits = kd.splits;
    a.center = kd.center;
    a.x = kd.x;
    a.d = d;
    a.q = c.q.as<T>();
    a.m = c.m();
    a.inv_cost = T(1) / (T) c.cost;
    a.qa = (T) c.QA_cost;
    a.add = add;
    a.target = target;
    a.out = out;
    a.scal = c.scal.as<T>();
    a.partials = c.partials.as<T>();
    a.ticket = c.counters.as<unsigned int>();
    a.mailbox = mailbox;
    a.u = static_cast<const T *>(kd.u);
    a.c0pow = (T) kd.c0pow;
    a.rbf = kd.rbf;
    return a;
}

template <typename T, int MODE>
void abar_finish(ctx &c, const kd_result &kd, const T *d, T add, T *out, T target = T(-1), T *mailbox = nullptr) {
    const finish_args<T> a = make_finish_args<T>(c, kd, d, add, out, target, mailbox);
    k_abar_finish<T, MODE><<<vec_grid(a.m), VEC_THREADS, 0, c.stream>>>(a);
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}

// blocks of the CG step kernel: all of them must be co-resident for its grid barrier (at most one small block per SM always is)
int step_grid(const ctx &c, int64_t m) {
    static int sms[64] = {};
    int &n = sms[c.device & 63];
    if (n == 0 && cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, c.device) != cudaSuccess) { n = NUM_SMS; }
    const int64_t g = (m + VEC_THREADS * 4 - 1) / (VEC_THREADS * 4);
    return (int) std::max<int64_t>(1, std::min<int64_t>(g, std::min(n, NUM_SMS)));
}

template <typename T>
void cg_step(ctx &c, const kd_result &kd, T *x, T *r, T *d, T *Ad, T target, T *mailbox, float4 *pk) {
    step_args<T> a{};
    a.f = make_finish_args<T>(c, kd, d, T(1), Ad, target, mailbox);
    a.x = x;
    a.r = r;
    a.dvec = d;
    a.pk = pk;
    a.bar = c.counters.as<unsigned int>() + 2;
    k_cg_step<T><<<step_grid(c, a.f.m), VEC_THREADS, 0, c.stream>>>(a);
    count_launch(c);
}

}  // namespace

// ret += add * Abar d (test hook for run_device_kernel; d and ret are device vectors of m reals)
template <typename T>
void apply_abar(ctx &c, const T *d, T add, T *ret) {
    // a solve that failed half-way may have left the convergence flag set: the kernels below would skip their work
    B200_CUDA_CHECK(cudaMemsetAsync(c.scal.as<T>() + S_DONE, 0, sizeof(T), c.stream));
    const kd_result kd = kernel_times_vector<T>(c, d, static_cast<const T *>(nullptr));
    abar_finish<T, 3>(c, kd, d, add, ret);
}
template void apply_abar<float>(ctx &, const float *, float, float *);
template void apply_abar<double>(ctx &, const double *, double, double *);

template <typename T>
int64_t cg_solve(ctx &c, const T *b, int64_t imax, T eps, T *x, double *delta_out) {
    const int64_t m = c.m();
    const size_t vb = (size_t) m * sizeof(T);
    c.v_r.reserve(vb);
    c.v_d.reserve(vb);
    c.v_Ad.reserve(vb);
    T *r = c.v_r.as<T>(), *d = c.v_d.as<T>(), *Ad = c.v_Ad.as<T>();
    T *scal = c.scal.as<T>();
    T *mailbox = static_cast<T *>(c.pinned);  // pinned host memory, written by the CG kernels directly (mapped, UVA)
    const int g = vec_grid(m);
    unsigned int *ticket = c.counters.as<unsigned int>();
    T *partials = c.partials.as<T>();
    const T *done = scal + S_DONE;
    static const bool unfused = [] {
        const char *e = std::getenv("B200SVM_CG_UNFUSED");  // A/B switch: three vector kernels per iteration instead of the fused step
        return e != nullptr && e[0] == '1';
    }();

    // alpha = 1 (csvm.cpp:85), r = b (:91), r -= Abar alpha (:94), delta0 = delta = r.r (:97-98), d = r (:101)
    B200_CUDA_CHECK(cudaMemsetAsync(scal + S_DONE, 0, sizeof(T), c.stream));
    k_fill<T><<<g, VEC_THREADS, 0, c.stream>>>(x, m, T(1));
    count_launch(c);
    B200_CUDA_CHECK(cudaMemcpyAsync(r, b, vb, cudaMemcpyDeviceToDevice, c.stream));
    {
        const kd_result kd = kernel_times_vector<T>(c, x, done);
        abar_finish<T, 2>(c, kd, x, T(-1), r);
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(d, r, vb, cudaMemcpyDeviceToDevice, c.stream));
    B200_CUDA_CHECK(cudaMemcpyAsync(mailbox, scal + S_DELTA0, sizeof(T), cudaMemcpyDeviceToHost, c.stream));
    sync_stream(c);
    const T delta0 = mailbox[0];
    T delta = delta0;
    const T target = eps * eps * delta0;
    float4 *pk = tc_active<T>(c) ? static_cast<float4 *>(tc_pack_buffer(c)) : nullptr;
    bool d_described = false;  // scal[S_SUM_D] / scal[S_Q_D] and the packed operand describe d (true after the first direction update)

    // The host runs `ahead` iterations in front of the one whose delta it has read, so the stream never drains between
    // iterations (and ranks never wait for each other's launch latency).  The stop test of csvm.cpp:145 is taken on the device:
    // the kernel that produces delta raises scal[S_DONE], which turns the already queued iterations — their O(m^2) kernels, their
    // exchanges and their vector updates — into no-ops: alpha and delta are exactly those of the converged iteration whatever
    // `ahead` is.
    const int64_t ahead = std::min<int64_t>(c.opt_sync_every, ctx::CG_RING - 2);
    int64_t marks[ctx::CG_RING] = {}, vmarks[ctx::CG_RING] = {};
    int64_t converged = -1, polled = 0;
    const auto poll = [&](int64_t it) {
        const int slot = (int) (it % ctx::CG_RING);
        B200_CUDA_CHECK(cudaEventSynchronize(c.cg_ev[slot]));
        drain_matvec_events(c, marks[slot]);
        drain_vector_events(c, vmarks[slot]);
        delta = *reinterpret_cast<volatile T *>(mailbox + 1 + slot);
        return delta <= target;
    };
    int64_t run = 0;
    try {
        for (; run < imax; ++run) {
            const int slot = (int) (run % ctx::CG_RING);
            // Ad = Abar d (:121-122) and d.Ad
            const kd_result kd = kernel_times_vector<T>(c, d, done, d_described, d_described && pk != nullptr);
            const bool vtimed = c.sv_head - c.sv_tail < ctx::EV_RING && run % 50 != 49;
            const int vslot = (int) (c.sv_head % ctx::EV_RING);
            if (vtimed) { B200_CUDA_CHECK(cudaEventRecord(c.sv[2 * vslot], c.stream)); }
            if (run % 50 == 49) {
                // x += alpha_cd d ; r = b - Abar x (:130-135)
                abar_finish<T, 0>(c, kd, d, T(1), Ad);
                k_cg_update<T, 1><<<g, VEC_THREADS, 0, c.stream>>>(x, r, d, Ad, m, scal, partials, ticket, target, static_cast<T *>(nullptr));
                count_launch(c);
                B200_CUDA_CHECK(cudaMemcpyAsync(r, b, vb, cudaMemcpyDeviceToDevice, c.stream));
                const kd_result kx = kernel_times_vector<T>(c, x, done);
                abar_finish<T, 1>(c, kx, x, T(-1), r, target, mailbox + 1 + slot);
                k_cg_direction<T><<<g, VEC_THREADS, 0, c.stream>>>(d, r, c.q.as<T>(), m, scal, partials, ticket, pk);
                count_launch(c);
            } else if (unfused) {
                abar_finish<T, 0>(c, kd, d, T(1), Ad);
                k_cg_update<T, 0><<<g, VEC_THREADS, 0, c.stream>>>(x, r, d, Ad, m, scal, partials, ticket, target, mailbox + 1 + slot);
                k_cg_direction<T><<<g, VEC_THREADS, 0, c.stream>>>(d, r, c.q.as<T>(), m, scal, partials, ticket, pk);
                count_launch(c, 2);
            } else {
                cg_step<T>(c, kd, x, r, d, Ad, target, mailbox + 1 + slot, pk);
            }
            B200_CUDA_CHECK(cudaGetLastError());
            if (vtimed) {
                B200_CUDA_CHECK(cudaEventRecord(c.sv[2 * vslot + 1], c.stream));
                ++c.sv_head;
            }
            d_described = true;
            B200_CUDA_CHECK(cudaEventRecord(c.cg_ev[slot], c.stream));  // delta of this iteration has reached the mailbox
            marks[slot] = c.ev_head;
            vmarks[slot] = c.sv_head;
            while (converged < 0 && polled + ahead <= run) {
                if (poll(polled)) { converged = polled; }
                ++polled;
            }
            if (converged >= 0) { break; }
        }
        const int64_t queued = std::min(run + 1, imax);
        while (converged < 0 && polled < queued) {
            if (poll(polled)) { converged = polled; }
            ++polled;
        }
    } catch (...) {
        cudaMemsetAsync(scal + S_DONE, 0, sizeof(T), c.stream);  // a later matvec on this handle must not be skipped
        throw;
    }
    B200_CUDA_CHECK(cudaMemsetAsync(scal + S_DONE, 0, sizeof(T), c.stream));
    sync_stream(c);
    if (delta_out != nullptr) { *delta_out = (double) delta; }
    const int64_t last = converged >= 0 ? converged : imax - 1;
    return last + 1 < imax ? last + 1 : imax;  // std::min(run + 1, imax), csvm.cpp:162-167
}
template int64_t cg_solve<float>(ctx &, const float *, int64_t, float, float *, double *);
template int64_t cg_solve<double>(ctx &, const double *, int64_t, double, double *, double *);

}  // namespace b200
