// Device-resident Conjugate Gradient (restates src/plssvm/backends/OpenMP/csvm.cpp:82-170; the reference's GPU variant keeps
// the vector algebra on the host with 4-5 PCIe round trips per iteration, src/plssvm/backends/gpu_csvm.cpp:259-309).
//
// All vectors and scalars live in HBM.  Per iteration: [sums of d] -> [K d] -> [Abar epilogue + d.Ad] -> [x, r update + r.r]
// -> [d update + sums of d]; every reduction is a fused two-stage deterministic grid sum (device_utils.cuh: grid_sum).
// The host only reads one scalar (delta) per iteration for the stop test `delta <= eps^2 delta0` (csvm.cpp:145).
#include "device_utils.cuh"

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
template <typename T, int MODE>
__global__ void __launch_bounds__(VEC_THREADS) k_abar_finish(const T *__restrict__ P, int64_t ldP, int splits, const T *__restrict__ d,
                                                             int64_t m, T inv_cost, int center, const T *__restrict__ q, T qa, T add,
                                                             T *__restrict__ out, T *__restrict__ scal, T *partials, unsigned int *ticket,
                                                             T target, const unsigned long long *wait_flags, unsigned long long wait_seq) {
    __shared__ T red[32];
    if constexpr (MODE == 0 || MODE == 1) {
        if (scal[S_DONE] != T(0)) { return; }  // CG has converged: this queued iteration must not touch r / delta
    }
    // multi-GPU peer exchange: the `splits` partial vectors are the slots the ranks pushed into this rank's buffer; wait until
    // every source has published this exchange's sequence number (one spinning thread per source), then read past L1
    const bool remote = wait_flags != nullptr;
    if (remote) {
        if (threadIdx.x < splits) {
            const volatile unsigned long long *f = wait_flags + threadIdx.x;
            while (*f < wait_seq) {}
            __threadfence_system();
        }
        __syncthreads();
    }
    const T sum_d = center != 0 ? scal[S_SUM_D] : T(0);
    const T q_d = center == 2 ? scal[S_Q_D] : T(0);
    const T rank1 = center == 1 ? sum_d * inv_cost : T(0);
    T acc[1] = { T(0) };
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) {
        T kd = T(0);
        if (remote) {
            for (int s = 0; s < splits; ++s) { kd += __ldcg(&P[(int64_t) s * ldP + i]); }  // rank order: identical on every rank
        } else {
            for (int s = 0; s < splits; ++s) { kd += P[(int64_t) s * ldP + i]; }
        }
        const T di = d[i];
        if (center == 2) { kd += fma(qa - q[i], sum_d, -q_d); }
        const T val = kd + rank1 + di * inv_cost;
        if constexpr (MODE == 0) {
            out[i] = val;
            acc[0] = fma(di, val, acc[0]);
        } else {
            const T r = fma(add, val, out[i]);
            out[i] = r;
            acc[0] = fma(r, r, acc[0]);
        }
    }
    if constexpr (MODE != 3) {
        T res[1];
        if (grid_sum<T, 1>(acc, partials, ticket, red, res)) {
            if constexpr (MODE == 0) {
                scal[S_DAD] = res[0];
            } else if constexpr (MODE == 1) {
                const T old = scal[S_DELTA];
                scal[S_DELTA_OLD] = old;
                scal[S_DELTA] = res[0];
                scal[S_BETA] = res[0] / old;
                if (res[0] <= target) { scal[S_DONE] = T(1); }  // csvm.cpp:145
            } else {
                scal[S_DELTA] = res[0];
                scal[S_DELTA0] = res[0];
            }
        }
    }
}

// alpha_cd = delta / d.Ad ; x += alpha_cd d ; (REFRESH == 0:) r -= alpha_cd Ad, delta_old = delta, delta = r.r, beta = delta / delta_old
template <typename T, int REFRESH>
__global__ void __launch_bounds__(VEC_THREADS) k_cg_update(T *__restrict__ x, T *__restrict__ r, const T *__restrict__ d,
                                                           const T *__restrict__ Ad, int64_t m, T *__restrict__ scal, T *partials,
                                                           unsigned int *ticket, T target) {
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
        }
    }
}

// d = beta d + r (csvm.cpp:155) ; scal[S_SUM_D] = sum d ; scal[S_Q_D] = q.d (for the next matvec's rank-1 epilogue)
template <typename T>
__global__ void __launch_bounds__(VEC_THREADS) k_cg_direction(T *__restrict__ d, const T *__restrict__ r, const T *__restrict__ q, int64_t m,
                                                              T *__restrict__ scal, T *partials, unsigned int *ticket) {
    __shared__ T red[32];
    const T beta = scal[S_BETA];
    T acc[2] = { T(0), T(0) };
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) {
        const T di = fma(beta, d[i], r[i]);
        d[i] = di;
        acc[0] += di;
        acc[1] = fma(q[i], di, acc[1]);
    }
    T out[2];
    if (grid_sum<T, 2>(acc, partials, ticket, red, out)) {
        scal[S_SUM_D] = out[0];
        scal[S_Q_D] = out[1];
    }
}

// Multi-GPU: fold this rank's column-split partials into one full-length vector, zero outside its row range [lo, hi)
// (the vector is then summed over ranks with one ncclAllReduce):  P[0][i] = lo <= i < hi ? sum_s P[s][i] : 0
template <typename T>
__global__ void __launch_bounds__(VEC_THREADS) k_fold_splits(T *__restrict__ P, int64_t ldP, int splits, int64_t m, int64_t lo, int64_t hi) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) {
        T s = T(0);
        if (i >= lo && i < hi) {
            for (int k = 0; k < splits; ++k) { s += P[(int64_t) k * ldP + i]; }
        }
        P[i] = s;
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

void sync_stream(ctx &c) {
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
    drain_matvec_events(c, c.ev_head);
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
    xchg_view x; // multi-GPU peer exchange: where the partial vectors of all ranks arrive (base == nullptr: c.P holds the sum)
};

// sum the rank's partial vector (c.P, m reals) over the ranks: peer-memory exchange (the sum happens in k_abar_finish) or NCCL
template <typename T>
void reduce_over_ranks(ctx &c, int64_t m, kd_result &res) {
    if (c.world <= 1) { return; }
    comm_exchange(c, c.P.p, m, &res.x);
    if (res.x.base != nullptr) {
        res.ldP = res.x.ld;
        res.splits = res.x.slots;
    }
}

// K d for the reduced system (rows/cols 0..m-1) into c.P as `splits` partial vectors.
template <typename T>
kd_result kernel_times_vector(ctx &c, const T *d) {
    const int64_t m = c.m();
    const int64_t ldP = (m + 255) / 256 * 256;
    constexpr int MAXS = 64;
    kd_result res{ ldP, 1, 0, xchg_view{} };
    const bool timed = c.ev_head - c.ev_tail < ctx::EV_RING;
    const int ev_slot = (int) (c.ev_head % ctx::EV_RING);
    if (timed) { B200_CUDA_CHECK(cudaEventRecord(c.ev[2 * ev_slot], c.stream)); }
    if (c.is_csr) {
        // sparse rows: un-centred K d of this rank's rows, rank-1 centring in the O(m) epilogue
        c.P.reserve((size_t) ldP * sizeof(T));
        sums2<T>(c, d, S_SUM_D, S_Q_D);
        sp_kernel_matvec<T>(c, d, c.P.as<T>());
        reduce_over_ranks<T>(c, m, res);
        res.center = 2;
        c.stats.matvec_path = 4;
    } else if (c.kernel == B200SVM_LINEAR) {
        // centred linear kernel: sum_j (K_ij + K_nn - K_in - K_nj) d_j = Z (Z^T d), Z = X - 1 x_n^T : two GEMVs (HBM bound)
        if (!c.has_Z) {
            c.Z.reserve((size_t) m * c.ldx * sizeof(T));
            center_rows<T>(c, c.X.as<T>(), m, c.ldx, c.X.as<T>() + m * c.ldx, c.Z.as<T>());
            c.has_Z = true;
        }
        c.P.reserve((size_t) ldP * sizeof(T));
        c.wvec.reserve((size_t) c.ldx * sizeof(T));
        sums2<T>(c, d, S_SUM_D, S_Q_D);
        xt_v<T>(c, c.Z.as<T>(), m, c.ldx, d, c.wvec.as<T>());
        x_w<T>(c, c.Z.as<T>(), m, c.ldx, c.wvec.as<T>(), T(0), c.P.as<T>());
        res.center = 1;
        c.stats.matvec_path = 3;
    } else {
        bool use_tc = false;
        if constexpr (sizeof(T) == 4) { use_tc = (c.opt_matvec_path == 2) || (c.opt_matvec_path == 0 && tc_available(c)); }
        if (use_tc) {
            if constexpr (sizeof(T) == 4) {
                c.P.reserve((size_t) ldP * sizeof(T));
                tc_kernel_matvec(c, reinterpret_cast<const float *>(d), c.P.as<float>());  // partial over this rank's tiles
            }
            reduce_over_ranks<T>(c, m, res);
            c.stats.matvec_path = 2;
        } else {
            // full rows [lo, hi) of the implicit matrix on this rank (all rows when world == 1)
            int64_t lo = 0, hi = m;
            partition_rows(m, c.rank, c.world, &lo, &hi);
            c.P.reserve((size_t) (MAXS * ldP) * sizeof(T));
            res.splits = 0;
            if (hi > lo) {
                res.splits = kmv_simt_launch<T>(c, c.X.as<T>() + lo * c.ldx, c.norms.as<T>() + lo, hi - lo, c.ldx, c.X.as<T>(), c.norms.as<T>(), m,
                                                c.ldx, d, c.P.as<T>() + lo, ldP, MAXS, c.q.as<T>() + lo, c.q.as<T>(), (T) c.QA_cost);
            }
            if (c.world > 1) {
                k_fold_splits<T><<<vec_grid(m), VEC_THREADS, 0, c.stream>>>(c.P.as<T>(), ldP, res.splits, m, lo, hi);
                count_launch(c);
                B200_CUDA_CHECK(cudaGetLastError());
                res.splits = 1;
                reduce_over_ranks<T>(c, m, res);
            }
            c.stats.matvec_path = 1;
        }
    }
    if (timed) {
        B200_CUDA_CHECK(cudaEventRecord(c.ev[2 * ev_slot + 1], c.stream));
        ++c.ev_head;
    }
    ++c.stats.matvecs;
    return res;
}

template <typename T, int MODE>
void abar_finish(ctx &c, const kd_result &kd, const T *d, T add, T *out, T target = T(-1)) {
    const int64_t m = c.m();
    const T *P = kd.x.base != nullptr ? static_cast<const T *>(kd.x.base) : c.P.as<T>();
    k_abar_finish<T, MODE><<<vec_grid(m), VEC_THREADS, 0, c.stream>>>(P, kd.ldP, kd.splits, d, m, T(1) / (T) c.cost, kd.center, c.q.as<T>(),
                                                                       (T) c.QA_cost, add, out, c.scal.as<T>(), c.partials.as<T>(),
                                                                       c.counters.as<unsigned int>(), target, kd.x.flags, kd.x.seq);
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}

}  // namespace

// ret += add * Abar d (test hook for run_device_kernel; d and ret are device vectors of m reals)
template <typename T>
void apply_abar(ctx &c, const T *d, T add, T *ret) {
    const kd_result kd = kernel_times_vector<T>(c, d);
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
    T *mailbox = static_cast<T *>(c.pinned);
    const int g = vec_grid(m);
    unsigned int *ticket = c.counters.as<unsigned int>();
    T *partials = c.partials.as<T>();

    // alpha = 1 (csvm.cpp:85), r = b (:91), r -= Abar alpha (:94), delta0 = delta = r.r (:97-98), d = r (:101)
    B200_CUDA_CHECK(cudaMemsetAsync(scal + S_DONE, 0, sizeof(T), c.stream));
    k_fill<T><<<g, VEC_THREADS, 0, c.stream>>>(x, m, T(1));
    count_launch(c);
    B200_CUDA_CHECK(cudaMemcpyAsync(r, b, vb, cudaMemcpyDeviceToDevice, c.stream));
    {
        const kd_result kd = kernel_times_vector<T>(c, x);
        abar_finish<T, 2>(c, kd, x, T(-1), r);
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(d, r, vb, cudaMemcpyDeviceToDevice, c.stream));
    B200_CUDA_CHECK(cudaMemcpyAsync(mailbox, scal + S_DELTA0, sizeof(T), cudaMemcpyDeviceToHost, c.stream));
    sync_stream(c);
    const T delta0 = mailbox[0];
    T delta = delta0;
    const T target = eps * eps * delta0;

    // The host runs `ahead` iterations in front of the one whose delta it has read, so the stream never drains between
    // iterations (and ranks never wait for each other's launch latency inside the allreduce).  The stop test of csvm.cpp:145 is
    // taken on the device: the kernel that produces delta raises scal[S_DONE], which turns the already queued iterations into
    // no-ops — alpha and delta are exactly those of the converged iteration whatever `ahead` is.
    const int64_t ahead = std::min<int64_t>(c.opt_sync_every, ctx::CG_RING - 2);
    int64_t marks[ctx::CG_RING] = {};
    int64_t converged = -1, polled = 0;
    const auto poll = [&](int64_t it) {
        const int slot = (int) (it % ctx::CG_RING);
        B200_CUDA_CHECK(cudaEventSynchronize(c.cg_ev[slot]));
        drain_matvec_events(c, marks[slot]);
        delta = mailbox[1 + slot];
        return delta <= target;
    };
    int64_t run = 0;
    for (; run < imax; ++run) {
        // Ad = Abar d (:121-122) and d.Ad
        const kd_result kd = kernel_times_vector<T>(c, d);
        abar_finish<T, 0>(c, kd, d, T(1), Ad);
        if (run % 50 == 49) {
            // x += alpha_cd d ; r = b - Abar x (:130-135)
            k_cg_update<T, 1><<<g, VEC_THREADS, 0, c.stream>>>(x, r, d, Ad, m, scal, partials, ticket, target);
            count_launch(c);
            B200_CUDA_CHECK(cudaMemcpyAsync(r, b, vb, cudaMemcpyDeviceToDevice, c.stream));
            const kd_result kx = kernel_times_vector<T>(c, x);
            abar_finish<T, 1>(c, kx, x, T(-1), r, target);
        } else {
            k_cg_update<T, 0><<<g, VEC_THREADS, 0, c.stream>>>(x, r, d, Ad, m, scal, partials, ticket, target);
            count_launch(c);
        }
        B200_CUDA_CHECK(cudaGetLastError());
        // delta of this iteration -> pinned mailbox; the direction update is queued right behind it
        const int slot = (int) (run % ctx::CG_RING);
        B200_CUDA_CHECK(cudaMemcpyAsync(mailbox + 1 + slot, scal + S_DELTA, sizeof(T), cudaMemcpyDeviceToHost, c.stream));
        B200_CUDA_CHECK(cudaEventRecord(c.cg_ev[slot], c.stream));
        marks[slot] = c.ev_head;
        k_cg_direction<T><<<g, VEC_THREADS, 0, c.stream>>>(d, r, c.q.as<T>(), m, scal, partials, ticket);
        count_launch(c);
        B200_CUDA_CHECK(cudaGetLastError());
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
    B200_CUDA_CHECK(cudaMemsetAsync(scal + S_DONE, 0, sizeof(T), c.stream));
    sync_stream(c);
    if (delta_out != nullptr) { *delta_out = (double) delta; }
    const int64_t last = converged >= 0 ? converged : imax - 1;
    return last + 1 < imax ? last + 1 : imax;  // std::min(run + 1, imax), csvm.cpp:162-167
}
template int64_t cg_solve<float>(ctx &, const float *, int64_t, float, float *, double *);
template int64_t cg_solve<double>(ctx &, const double *, int64_t, double, double *, double *);

}  // namespace b200
