// SIMT kernels (sm_100a): squared norms, q-vector, the generic tiled "kernel-matrix x vector" product used by the
// O(m^2) matvec (double, and float when the tcgen05 path is not selected) and by predict, and the two-GEMV linear path.
//
// Reference behaviour restated (not ported): src/plssvm/backends/OpenMP/svm_kernel.cpp:21-47 (matvec),
// q_kernel.cpp:19-50 (q), csvm.cpp:173-244 (w, predict); CUDA analogues src/plssvm/backends/CUDA/{svm,q,predict}_kernel.cu.
// Differences by design: Gram form for all kernels (rbf via |a|^2 + |b|^2 - 2<a,b>), full rows without atomics
// (deterministic), rank-1 terms (QA_cost, q) applied in an O(m) epilogue, linear kernel as X (X^T d).
#include "device_utils.cuh"

#include <cstdarg>

namespace b200 {

// ---------------------------------------------------------------------------------------------------------------------
// squared norms: one warp per row
// ---------------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_row_norms(const T *__restrict__ X, int64_t rows, int64_t ld, T *__restrict__ out) {
    const int64_t row = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) { return; }
    const int lane = threadIdx.x & 31;
    const T *x = X + row * ld;
    T s = 0;
    for (int64_t k = lane; k < ld; k += 32) { s = fma(x[k], x[k], s); }
    s = warp_sum(s);
    if (lane == 0) { out[row] = s; }
}

template <typename T>
void compute_norms(ctx &c, const T *X, int64_t rows, int64_t ld, int64_t, T *out) {
    if (rows == 0) { return; }
    k_row_norms<T><<<(unsigned) ((rows + 7) / 8), 256, 0, c.stream>>>(X, rows, ld, out);
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}
template void compute_norms<float>(ctx &, const float *, int64_t, int64_t, int64_t, float *);
template void compute_norms<double>(ctx &, const double *, int64_t, int64_t, int64_t, double *);

// ---------------------------------------------------------------------------------------------------------------------
// q[i] = k(x_i, x_last), i < m ; also k(x_last, x_last) -> scal[S_KXX]       (one warp per row)
// ---------------------------------------------------------------------------------------------------------------------
template <typename T, int KERNEL>
__global__ void __launch_bounds__(256) k_generate_q(const T *__restrict__ X, const T *__restrict__ norms, int64_t n, int64_t ld,
                                                    int degree, T gamma, T coef0, T *__restrict__ q, T *__restrict__ scal) {
    const int64_t row = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= n) { return; }
    const int lane = threadIdx.x & 31;
    const T *x = X + row * ld;
    const T *xl = X + (n - 1) * ld;
    T s = 0;
    for (int64_t k = lane; k < ld; k += 32) { s = fma(x[k], xl[k], s); }
    s = warp_sum(s);
    if (lane == 0) {
        const T v = kernel_value<T, KERNEL>(s, norms[row], norms[n - 1], degree, gamma, coef0);
        if (row < n - 1) {
            q[row] = v;
        } else {
            scal[S_KXX] = KERNEL == B200SVM_RBF ? T(1) : v;  // exp(-gamma * 0) is exactly 1 in the reference
        }
    }
}

template <typename T>
void generate_q(ctx &c) {
    const unsigned grid = (unsigned) ((c.n + 7) / 8);
    const T *X = c.X.as<T>();
    const T *nr = c.norms.as<T>();
    T *q = c.q.as<T>();
    T *sc = c.scal.as<T>();
    switch (c.kernel) {
        case B200SVM_LINEAR: k_generate_q<T, B200SVM_LINEAR><<<grid, 256, 0, c.stream>>>(X, nr, c.n, c.ldx, c.degree, (T) c.gamma, (T) c.coef0, q, sc); break;
        case B200SVM_POLYNOMIAL: k_generate_q<T, B200SVM_POLYNOMIAL><<<grid, 256, 0, c.stream>>>(X, nr, c.n, c.ldx, c.degree, (T) c.gamma, (T) c.coef0, q, sc); break;
        default: k_generate_q<T, B200SVM_RBF><<<grid, 256, 0, c.stream>>>(X, nr, c.n, c.ldx, c.degree, (T) c.gamma, (T) c.coef0, q, sc); break;
    }
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}
template void generate_q<float>(ctx &);
template void generate_q<double>(ctx &);

// ---------------------------------------------------------------------------------------------------------------------
// Generic tiled kernel-matrix x vector:  P[s][i] = sum_{j in column split s} k(A_i, B_j) v_j
//   (CENTER: k(A_i, B_j) + QA_cost - q_i - q_j, the entry of the reduced system without its 1/C diagonal)
//   A: rowsA x ldA, B: rowsB x ldB (row-major, zero padded to a multiple of 16 features), squared norms precomputed.
//   256 threads = 16 x 16, each owning a TM x TN register micro-tile of the BM x BN Gram tile (BM = 16 TM, BN = 16 TN);
//   K is streamed in BK = 16 slices through double-buffered shared memory (register prefetch of the next slice).
//   The Gram tile never leaves registers: kernel function + multiply by v_j + row reduction happen in the epilogue.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int BK = 16;

template <typename T>
struct vec16;
template <>
struct vec16<float> {
    using type = float4;
    static constexpr int N = 4;
};
template <>
struct vec16<double> {
    using type = double2;
    static constexpr int N = 2;
};

// STORE (double): instead of reducing against v, the kernel values k(A_i, B_j) of the lower block triangle go to the
// symmetric tile store (128 x 256 tiles, pair-column-major double2; tile (I, J) at tile_cum(I) - tile0 + J): P = the tiles, ldP = tile0,
// qa = global index of A's first row; a CTA covers the 64-column tiles below the end of its row block's diagonal strip only.
template <typename T, int KERNEL, int TM, int TN, bool CENTER, bool STORE = false>
__global__ void __launch_bounds__(256)
    k_kmv_tile(const T *__restrict__ A, const T *__restrict__ normA, int64_t rowsA, int64_t ldA, const T *__restrict__ B,
               const T *__restrict__ normB, int64_t rowsB, int64_t ldB, int64_t K, const T *__restrict__ v, int degree, T gamma, T coef0,
               const T *__restrict__ qrow, const T *__restrict__ qcol, T qa, T *__restrict__ P, int64_t ldP, int col_tiles,
               const T *__restrict__ done) {
    constexpr int BM = 16 * TM, BN = 16 * TN;
    static_assert(!STORE || (sizeof(T) == 8 && TM == 4 && TN == 4 && !CENTER), "the tile-store epilogue is written for the 64 x 64 double tile");
    if (done != nullptr && *done != T(0)) { return; }  // CG convergence flag (scal[S_DONE]); uniform over the grid
    using V = typename vec16<T>::type;
    constexpr int VN = vec16<T>::N;
    constexpr int EA = BM * BK / 256, TPRA = BK / EA;  // elements per thread / threads per row when loading the A slice
    constexpr int EB = BN * BK / 256, TPRB = BK / EB;
    static_assert(EA % VN == 0 && EB % VN == 0, "slice loads are 16-byte vectors");

    __shared__ __align__(16) T As[2][BK][BM];
    __shared__ __align__(16) T Bs[2][BK][BN];

    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    const int64_t row0 = (int64_t) blockIdx.x * BM;
    int64_t grow0 = 0;  // STORE: global index of the CTA's first row
    if constexpr (STORE) {
        grow0 = (int64_t) qa + row0;
        col_tiles = (int) (256 * ((grow0 >> 8) + 1) / BN);  // columns below the end of the diagonal strip of row block grow0 / 128
    }
    const int ct0 = (int) ((int64_t) blockIdx.y * col_tiles / gridDim.y);
    const int ct1 = (int) ((int64_t) (blockIdx.y + 1) * col_tiles / gridDim.y);

    const int la_row = t / TPRA, la_k = (t % TPRA) * EA;
    const int lb_row = t / TPRB, lb_k = (t % TPRB) * EB;
    const int64_t a_row = min(row0 + la_row, rowsA - 1);  // out-of-range rows re-read the last row; masked in the epilogue
    const T *a_ptr = A + a_row * ldA + la_k;

    T rowsum[TM];
    T na[TM];
    T ai[TM];  // CENTER: QA_cost - q_i (the reference's per-entry centring, svm_kernel.cpp:32)
#pragma unroll
    for (int r = 0; r < TM; ++r) {
        rowsum[r] = T(0);
        const int64_t row = row0 + (r >> 2) * 64 + ty * 4 + (r & 3);
        na[r] = row < rowsA ? normA[row] : T(0);
        ai[r] = T(0);
        if constexpr (CENTER) { ai[r] = row < rowsA ? qa - qrow[row] : T(0); }
    }

    const int nk = (int) (K / BK);
    for (int ct = ct0; ct < ct1; ++ct) {
        const int64_t col0 = (int64_t) ct * BN;
        const int64_t b_row = min(col0 + lb_row, rowsB - 1);
        const T *b_ptr = B + b_row * ldB + lb_k;

        T acc[TM][TN];
#pragma unroll
        for (int r = 0; r < TM; ++r) {
#pragma unroll
            for (int cc = 0; cc < TN; ++cc) { acc[r][cc] = T(0); }
        }

        V ra[EA / VN], rb[EB / VN];
#pragma unroll
        for (int e = 0; e < EA / VN; ++e) { ra[e] = __ldg(reinterpret_cast<const V *>(a_ptr) + e); }
#pragma unroll
        for (int e = 0; e < EB / VN; ++e) { rb[e] = __ldg(reinterpret_cast<const V *>(b_ptr) + e); }
#pragma unroll
        for (int e = 0; e < EA; ++e) { As[0][la_k + e][la_row] = reinterpret_cast<const T *>(ra)[e]; }
#pragma unroll
        for (int e = 0; e < EB; ++e) { Bs[0][lb_k + e][lb_row] = reinterpret_cast<const T *>(rb)[e]; }
        __syncthreads();

        for (int kt = 0; kt < nk; ++kt) {
            const int cur = kt & 1;
            if (kt + 1 < nk) {
#pragma unroll
                for (int e = 0; e < EA / VN; ++e) { ra[e] = __ldg(reinterpret_cast<const V *>(a_ptr + (int64_t) (kt + 1) * BK) + e); }
#pragma unroll
                for (int e = 0; e < EB / VN; ++e) { rb[e] = __ldg(reinterpret_cast<const V *>(b_ptr + (int64_t) (kt + 1) * BK) + e); }
            }
#pragma unroll
            for (int k = 0; k < BK; ++k) {
                T a[TM], b[TN];
#pragma unroll
                for (int r4 = 0; r4 < TM / 4; ++r4) {
#pragma unroll
                    for (int e = 0; e < 4 / VN; ++e) {
                        const V val = *reinterpret_cast<const V *>(&As[cur][k][r4 * 64 + ty * 4 + e * VN]);
#pragma unroll
                        for (int u = 0; u < VN; ++u) { a[r4 * 4 + e * VN + u] = reinterpret_cast<const T *>(&val)[u]; }
                    }
                }
#pragma unroll
                for (int c4 = 0; c4 < TN / 4; ++c4) {
#pragma unroll
                    for (int e = 0; e < 4 / VN; ++e) {
                        const V val = *reinterpret_cast<const V *>(&Bs[cur][k][c4 * 64 + tx * 4 + e * VN]);
#pragma unroll
                        for (int u = 0; u < VN; ++u) { b[c4 * 4 + e * VN + u] = reinterpret_cast<const T *>(&val)[u]; }
                    }
                }
#pragma unroll
                for (int r = 0; r < TM; ++r) {
#pragma unroll
                    for (int cc = 0; cc < TN; ++cc) { acc[r][cc] = fma(a[r], b[cc], acc[r][cc]); }
                }
            }
            if (kt + 1 < nk) {
#pragma unroll
                for (int e = 0; e < EA; ++e) { As[cur ^ 1][la_k + e][la_row] = reinterpret_cast<const T *>(ra)[e]; }
#pragma unroll
                for (int e = 0; e < EB; ++e) { Bs[cur ^ 1][lb_k + e][lb_row] = reinterpret_cast<const T *>(rb)[e]; }
            }
            __syncthreads();
        }

        if constexpr (STORE) {
            // kernel values of this thread's 4 x 4 block -> two column pairs x four consecutive rows = two 64-byte runs of the tile
            const int64_t grow = grow0 + ty * 4, col = col0 + tx * 4;
            const int64_t Ib = grow >> 7, kk2 = Ib >> 1;
            const int64_t cum = (Ib & 1) ? (kk2 + 1) * (kk2 + 1) : kk2 * (kk2 + 1);  // = tile_store_cum(Ib)
            double2 *tb = reinterpret_cast<double2 *>(P) + (cum - ldP + (col >> 8)) * (int64_t) (128 * 256 / 2) + ((col & 255) >> 1) * 128 + (grow & 127);
#pragma unroll
            for (int r = 0; r < TM; ++r) {
                if (row0 + ty * 4 + r < rowsA) {
                    T kv[TN];
#pragma unroll
                    for (int cc = 0; cc < TN; ++cc) {
                        const bool valid = col + cc < rowsB;
                        kv[cc] = valid ? kernel_value<T, KERNEL>(acc[r][cc], na[r], normB[valid ? col + cc : 0], degree, gamma, coef0) : T(0);
                    }
                    tb[r] = make_double2(kv[0], kv[1]);
                    tb[128 + r] = make_double2(kv[2], kv[3]);
                }
            }
        } else {
        // epilogue of this Gram tile: kernel function, times v_j, row reduction in registers
#pragma unroll
        for (int cc = 0; cc < TN; ++cc) {
            const int64_t col = col0 + (cc >> 2) * 64 + tx * 4 + (cc & 3);
            const bool valid = col < rowsB;
            const T vj = valid ? v[col] : T(0);
            const T nb = valid ? normB[col] : T(0);
            T qj = T(0);
            if constexpr (CENTER) { qj = valid ? qcol[col] : T(0); }
#pragma unroll
            for (int r = 0; r < TM; ++r) {
                T kv = kernel_value<T, KERNEL>(acc[r][cc], na[r], nb, degree, gamma, coef0);
                if constexpr (CENTER) { kv += ai[r] - qj; }
                rowsum[r] = valid ? fma(kv, vj, rowsum[r]) : rowsum[r];
            }
        }
        }
    }

    if constexpr (!STORE) {
        // reduce over the 16 threads (tx) sharing a row; lanes of one row group are contiguous within a half-warp
#pragma unroll
        for (int r = 0; r < TM; ++r) {
            T s = rowsum[r];
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); }
            const int64_t row = row0 + (r >> 2) * 64 + ty * 4 + (r & 3);
            if (tx == 0 && row < rowsA) { P[(int64_t) blockIdx.y * ldP + row] = s; }
        }
    }
}

template <typename T>
struct simt_tile {
    static constexpr int TM = sizeof(T) == 4 ? 8 : 4;
    static constexpr int TN = sizeof(T) == 4 ? 8 : 4;
};

// Launch: returns the number of column splits written to P ([splits][ldP]).
template <typename T>
int kmv_simt_launch(ctx &c, const T *A, const T *normA, int64_t rowsA, int64_t ldA, const T *B, const T *normB, int64_t rowsB, int64_t ldB,
                    const T *v, T *P, int64_t ldP, int max_splits, const T *qrow, const T *qcol, T qa) {
    constexpr int TM = simt_tile<T>::TM, TN = simt_tile<T>::TN;
    constexpr int BM = 16 * TM, BN = 16 * TN;
    const int64_t row_tiles = (rowsA + BM - 1) / BM;
    const int col_tiles = (int) ((rowsB + BN - 1) / BN);
    // at least ~3 waves of CTAs when the problem allows it
    int64_t splits = (3 * 2 * NUM_SMS + row_tiles - 1) / row_tiles;
    if (splits > col_tiles) { splits = col_tiles; }
    if (splits > max_splits) { splits = max_splits; }
    if (splits < 1) { splits = 1; }
    const int64_t K = ldA < ldB ? ldA : ldB;
    const dim3 grid((unsigned) row_tiles, (unsigned) splits);
#define B200_LAUNCH_KMV(KN)                                                                                                              \
    do {                                                                                                                                 \
        if (qrow != nullptr) {                                                                                                           \
            k_kmv_tile<T, KN, TM, TN, true><<<grid, 256, 0, c.stream>>>(A, normA, rowsA, ldA, B, normB, rowsB, ldB, K, v, c.degree,      \
                                                                         (T) c.gamma, (T) c.coef0, qrow, qcol, qa, P, ldP, col_tiles,    \
                                                                         c.scal.as<T>() + S_DONE);                                       \
        } else {                                                                                                                         \
            k_kmv_tile<T, KN, TM, TN, false><<<grid, 256, 0, c.stream>>>(A, normA, rowsA, ldA, B, normB, rowsB, ldB, K, v, c.degree,     \
                                                                          (T) c.gamma, (T) c.coef0, qrow, qcol, qa, P, ldP, col_tiles,   \
                                                                          c.scal.as<T>() + S_DONE);                                      \
        }                                                                                                                                \
    } while (0)
    switch (c.kernel) {
        case B200SVM_LINEAR: B200_LAUNCH_KMV(B200SVM_LINEAR); break;
        case B200SVM_POLYNOMIAL: B200_LAUNCH_KMV(B200SVM_POLYNOMIAL); break;
        default: B200_LAUNCH_KMV(B200SVM_RBF); break;
    }
#undef B200_LAUNCH_KMV
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
    return (int) splits;
}
template int kmv_simt_launch<float>(ctx &, const float *, const float *, int64_t, int64_t, const float *, const float *, int64_t, int64_t,
                                    const float *, float *, int64_t, int, const float *, const float *, float);
template int kmv_simt_launch<double>(ctx &, const double *, const double *, int64_t, int64_t, const double *, const double *, int64_t, int64_t,
                                     const double *, double *, int64_t, int, const double *, const double *, double);

// ---------------------------------------------------------------------------------------------------------------------
// Dense double polynomial / rbf: cached kernel matrix.  The reference keeps the matrix implicit because its devices
// cannot hold it; 65536 points in double are 17 GB of lower-triangle tiles on a B200.  First matvec after set_data: one pass of the
// SIMT Gram kernel in STORE mode (FP64 FMA pipe bound, the same O(m^2 F) work as one implicit matvec, lower triangle only); every
// later matvec: the HBM-bound symmetric sweep k_tile_sweep_f64.  Row blocks are dealt to the ranks by equal tile count.
// ---------------------------------------------------------------------------------------------------------------------
bool dense_f64_cache_ready(ctx &c) {
    if (c.dtype != B200SVM_F64 || c.is_csr || c.kernel == B200SVM_LINEAR || c.opt_dense_cache == 1 || c.m() < 1) { return false; }
    const int64_t m = c.m();
    if (c.dts_valid && c.dts_rank == c.rank && c.dts_world == c.world && c.dts_m == m) { return true; }
    if (c.dts_refused) { return false; }
    c.dts_valid = false;
    if (c.dts == nullptr || c.dts_rank != c.rank || c.dts_world != c.world || c.dts_m != m) {
        tile_store_destroy(c.dts);
        c.dts = nullptr;
        c.dts = tile_store_create(c, m, sizeof(double));
        c.dts_rank = c.rank;
        c.dts_world = c.world;
        c.dts_m = m;
    }
    const size_t need = tile_store_bytes(c.dts);
    size_t free_b = 0, total_b = 0;
    B200_CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
    free_b += tile_store_allocated(c.dts);  // the buffer of a previous assembly is reused
    const double budget = c.opt_sparse_cache_mb > 0 ? (double) c.opt_sparse_cache_mb * 1048576.0 : 0.85 * (double) free_b;
    if ((double) need > budget) {
        B200_REQUIRE(c.opt_dense_cache != 2, "the cached kernel-matrix tiles do not fit the device memory budget");
        c.dts_refused = true;
        return false;
    }
    double *Kc = static_cast<double *>(tile_store_alloc(c.dts));
    int64_t row_lo = 0, row_hi = 0, tile0 = 0;
    tile_store_rows(c.dts, &row_lo, &row_hi, &tile0);
    row_hi = std::min(row_hi, m);
    if (c.cb_pending) { sync_stream(c); }
    B200_CUDA_CHECK(cudaEventRecord(c.cb_ev[0], c.stream));
    B200_CUDA_CHECK(cudaMemsetAsync(Kc, 0, need, c.stream));  // rows / columns >= m of the padded tiles
    if (row_hi > row_lo) {
        const int64_t rows = row_hi - row_lo;
        const dim3 grid((unsigned) ((rows + 63) / 64), 8);
        const double *A = c.X.as<double>() + row_lo * c.ldx;
        const double *nA = c.norms.as<double>() + row_lo;
        if (c.kernel == B200SVM_RBF) {
            k_kmv_tile<double, B200SVM_RBF, 4, 4, false, true><<<grid, 256, 0, c.stream>>>(A, nA, rows, c.ldx, c.X.as<double>(), c.norms.as<double>(), m, c.ldx, c.ldx, nullptr,
                                                                                       c.degree, c.gamma, c.coef0, nullptr, nullptr, (double) row_lo, Kc, tile0, 0, nullptr);
        } else {
            k_kmv_tile<double, B200SVM_POLYNOMIAL, 4, 4, false, true><<<grid, 256, 0, c.stream>>>(A, nA, rows, c.ldx, c.X.as<double>(), c.norms.as<double>(), m, c.ldx, c.ldx,
                                                                                              nullptr, c.degree, c.gamma, c.coef0, nullptr, nullptr, (double) row_lo, Kc, tile0,
                                                                                              0, nullptr);
        }
        count_launch(c);
        B200_CUDA_CHECK(cudaGetLastError());
    }
    B200_CUDA_CHECK(cudaEventRecord(c.cb_ev[1], c.stream));
    c.cb_pending = true;
    c.dts_valid = true;
    c.stats.dense_cache_bytes = (int64_t) need;
    return true;
}

void dense_f64_cache_invalidate(ctx &c) {
    c.dts_valid = false;
    c.dts_refused = false;
    if (c.dtype == B200SVM_F64 && !c.is_csr) { c.stats.dense_cache_bytes = 0; }
}

// ---------------------------------------------------------------------------------------------------------------------
// Linear kernel: K d = X (X^T d) — two HBM-bound GEMVs instead of an O(m^2 F) pairwise pass.
//   k_xt_v : wpart[chunk][f] = sum_{i in row chunk} v_i X[i][f]     (thread per feature: coalesced across f)
//   k_sum_chunks : w[f] = sum_chunk wpart[chunk][f]                  (fixed order => deterministic)
//   k_x_w  : out[i] = X[i] . w                                       (warp per row)
// The same kernels implement update_w (w = X^T alpha over all n rows; reference src/plssvm/backends/OpenMP/csvm.cpp:173-188).
// ---------------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_xt_v(const T *__restrict__ X, int64_t rows, int64_t ld, const T *__restrict__ v,
                                              T *__restrict__ wpart, int64_t rows_per_chunk) {
    const int64_t f = (int64_t) blockIdx.x * 256 + threadIdx.x;
    if (f >= ld) { return; }
    const int64_t r0 = (int64_t) blockIdx.y * rows_per_chunk;
    const int64_t r1 = min(rows, r0 + rows_per_chunk);
    T s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    int64_t i = r0;
    for (; i + 3 < r1; i += 4) {
        s0 = fma(v[i], X[i * ld + f], s0);
        s1 = fma(v[i + 1], X[(i + 1) * ld + f], s1);
        s2 = fma(v[i + 2], X[(i + 2) * ld + f], s2);
        s3 = fma(v[i + 3], X[(i + 3) * ld + f], s3);
    }
    for (; i < r1; ++i) { s0 = fma(v[i], X[i * ld + f], s0); }
    wpart[(int64_t) blockIdx.y * ld + f] = (s0 + s1) + (s2 + s3);
}

template <typename T>
__global__ void __launch_bounds__(256) k_sum_chunks(const T *__restrict__ wpart, int64_t ld, int chunks, T *__restrict__ w) {
    const int64_t f = (int64_t) blockIdx.x * 256 + threadIdx.x;
    if (f >= ld) { return; }
    T s = 0;
    for (int ch = 0; ch < chunks; ++ch) { s += wpart[(int64_t) ch * ld + f]; }
    w[f] = s;
}

template <typename T>
__global__ void __launch_bounds__(256) k_x_w(const T *__restrict__ X, int64_t rows, int64_t ld, const T *__restrict__ w, T bias,
                                             T *__restrict__ out) {
    const int64_t row = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) { return; }
    const int lane = threadIdx.x & 31;
    const T *x = X + row * ld;
    T s = 0;
    for (int64_t k = lane; k < ld; k += 32) { s = fma(x[k], w[k], s); }
    s = warp_sum(s);
    if (lane == 0) { out[row] = bias + s; }
}

// w = X[0:rows]^T v   (w: ld reals)
template <typename T>
void xt_v(ctx &c, const T *X, int64_t rows, int64_t ld, const T *v, T *w) {
    const int64_t fblocks = (ld + 255) / 256;
    int64_t chunks = (4 * NUM_SMS + fblocks - 1) / fblocks;
    if (chunks > (rows + 63) / 64) { chunks = (rows + 63) / 64; }
    if (chunks < 1) { chunks = 1; }
    const int64_t rpc = (rows + chunks - 1) / chunks;
    chunks = (rows + rpc - 1) / rpc;
    c.wpart.reserve((size_t) (chunks * ld) * sizeof(T));
    k_xt_v<T><<<dim3((unsigned) fblocks, (unsigned) chunks), 256, 0, c.stream>>>(X, rows, ld, v, c.wpart.as<T>(), rpc);
    k_sum_chunks<T><<<(unsigned) fblocks, 256, 0, c.stream>>>(c.wpart.as<T>(), ld, (int) chunks, w);
    count_launch(c, 2);
    B200_CUDA_CHECK(cudaGetLastError());
}
// out[i] = bias + X[i] . w
template <typename T>
void x_w(ctx &c, const T *X, int64_t rows, int64_t ld, const T *w, T bias, T *out) {
    if (rows == 0) { return; }
    k_x_w<T><<<(unsigned) ((rows + 7) / 8), 256, 0, c.stream>>>(X, rows, ld, w, bias, out);
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}
template void xt_v<float>(ctx &, const float *, int64_t, int64_t, const float *, float *);
template void xt_v<double>(ctx &, const double *, int64_t, int64_t, const double *, double *);
template void x_w<float>(ctx &, const float *, int64_t, int64_t, const float *, float, float *);
template void x_w<double>(ctx &, const double *, int64_t, int64_t, const double *, double, double *);

// Z[i] = X[i] - X[last]  (linear kernel: K_ij + K_nn - K_in - K_nj = <x_i - x_n, x_j - x_n>, so the reduced system is
// Z Z^T + 1/C (1 1^T + I) — no cancellation of large terms)
template <typename T>
__global__ void __launch_bounds__(256) k_center_rows(const T *__restrict__ X, int64_t rows, int64_t ld, const T *__restrict__ xlast,
                                                     T *__restrict__ Z) {
    const int64_t total = rows * ld;
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t) gridDim.x * blockDim.x) {
        Z[e] = X[e] - xlast[e % ld];
    }
}
template <typename T>
void center_rows(ctx &c, const T *X, int64_t rows, int64_t ld, const T *xlast, T *Z) {
    k_center_rows<T><<<8 * NUM_SMS, 256, 0, c.stream>>>(X, rows, ld, xlast, Z);
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}
template void center_rows<float>(ctx &, const float *, int64_t, int64_t, const float *, float *);
template void center_rows<double>(ctx &, const double *, int64_t, int64_t, const double *, double *);

template <typename T>
void update_w(ctx &c, const T *alpha_dev, T *w_dev) {
    xt_v<T>(c, c.X.as<T>(), c.n, c.ldx, alpha_dev, w_dev);
}
template void update_w<float>(ctx &, const float *, float *);
template void update_w<double>(ctx &, const double *, double *);

// out[p] = bias + sum_s P[s][p]
template <typename T>
__global__ void __launch_bounds__(256) k_sum_splits_bias(const T *__restrict__ P, int64_t ldP, int splits, int64_t rows, T bias, T *__restrict__ out) {
    const int64_t i = (int64_t) blockIdx.x * 256 + threadIdx.x;
    if (i >= rows) { return; }
    T s = 0;
    for (int sp = 0; sp < splits; ++sp) { s += P[(int64_t) sp * ldP + i]; }
    out[i] = bias + s;
}

// predict (reference src/plssvm/backends/OpenMP/csvm.cpp:191-244): decision values for dense points already on the device
template <typename T>
void predict_dense(ctx &c, const T *alpha_dev, T bias, const T *pts_dev, int64_t P, int64_t ld, T *out_dev) {
    if (P == 0) { return; }
    if (c.kernel == B200SVM_LINEAR) {
        c.wvec.reserve((size_t) c.ldx * sizeof(T));
        update_w<T>(c, alpha_dev, c.wvec.as<T>());
        x_w<T>(c, pts_dev, P, ld, c.wvec.as<T>(), bias, out_dev);
        return;
    }
    c.pts_norms.reserve((size_t) P * sizeof(T));
    compute_norms<T>(c, pts_dev, P, ld, c.F, c.pts_norms.as<T>());
    constexpr int MAXS = 64;
    const int64_t ldP = (P + 63) / 64 * 64;
    c.P.reserve((size_t) (MAXS * ldP) * sizeof(T));
    const int splits = kmv_simt_launch<T>(c, pts_dev, c.pts_norms.as<T>(), P, ld, c.X.as<T>(), c.norms.as<T>(), c.n, c.ldx, alpha_dev,
                                          c.P.as<T>(), ldP, MAXS, static_cast<const T *>(nullptr), static_cast<const T *>(nullptr), T(0));
    k_sum_splits_bias<T><<<(unsigned) ((P + 255) / 256), 256, 0, c.stream>>>(c.P.as<T>(), ldP, splits, P, bias, out_dev);
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}
template void predict_dense<float>(ctx &, const float *, float, const float *, int64_t, int64_t, float *);
template void predict_dense<double>(ctx &, const double *, double, const double *, int64_t, int64_t, double *);

// ---------------------------------------------------------------------------------------------------------------------
// FMA-pipe probe (measurement hook, b200svm_probe_fma_tflops): the FP32 / FP64 FMA peak of this device, measured — the denominator
// of the roofline of the SIMT Gram-tile kernel ("must be measured on the box").  16 independent FMA chains per thread.
// ---------------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_fma_probe(T *__restrict__ out, int iters, T a, T b) {
    T acc[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { acc[k] = (T) (threadIdx.x + k); }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 16; ++k) { acc[k] = fma(acc[k], a, b); }
    }
    T s = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) { s += acc[k]; }
    if (s == (T) -1.2345) { out[0] = s; }  // never true: keeps the chains alive
}

template <typename T>
double probe_fma_tflops(ctx &c) {
    const int iters = sizeof(T) == 4 ? 8192 : 1024, blocks = NUM_SMS * 8;
    dbuf sink;
    sink.reserve(sizeof(T));
    cudaEvent_t e0, e1;
    B200_CUDA_CHECK(cudaEventCreate(&e0));
    B200_CUDA_CHECK(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        B200_CUDA_CHECK(cudaEventRecord(e0, c.stream));
        k_fma_probe<T><<<blocks, 256, 0, c.stream>>>(sink.as<T>(), iters, (T) 0.999, (T) 1e-3);
        B200_CUDA_CHECK(cudaEventRecord(e1, c.stream));
        B200_CUDA_CHECK(cudaEventSynchronize(e1));
        float ms = 0.f;
        B200_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        const double tf = 2.0 * 16.0 * iters * 256.0 * blocks / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) { best = tf; }
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    count_launch(c, 4);
    return best;
}
// Shared-memory integer atomic probe (b200svm_probe_smem_atomic_gops): fire-and-forget 32-bit ATOMS.ADD to pseudo-random slots of an
// 8192-entry accumulator — the instruction the on-the-fly sparse kernel (k_sp_kmv) issues once per Gram product (float) or twice
// (double).  Its measured rate is the roofline of that kernel.
__global__ void __launch_bounds__(256) k_atoms_probe(int *__restrict__ out, int iters) {
    __shared__ int acc[8192];
    for (int k = threadIdx.x; k < 8192; k += 256) { acc[k] = 0; }
    __syncthreads();
    unsigned int x = threadIdx.x * 2654435761u + blockIdx.x * 40503u + 12345u;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x = x * 1664525u + 1013904223u;
            atomicAdd(&acc[(x >> 12) & 8191u], (int) (x & 7u));
        }
    }
    __syncthreads();
    if (acc[threadIdx.x] == -123456789) { out[0] = 1; }
}

double probe_smem_atomic_gops(ctx &c) {
    const int iters = 2048, blocks = NUM_SMS * 16;
    dbuf sink;
    sink.reserve(sizeof(int));
    cudaEvent_t e0, e1;
    B200_CUDA_CHECK(cudaEventCreate(&e0));
    B200_CUDA_CHECK(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        B200_CUDA_CHECK(cudaEventRecord(e0, c.stream));
        k_atoms_probe<<<blocks, 256, 0, c.stream>>>(sink.as<int>(), iters);
        B200_CUDA_CHECK(cudaEventRecord(e1, c.stream));
        B200_CUDA_CHECK(cudaEventSynchronize(e1));
        float ms = 0.f;
        B200_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        const double g = 8.0 * iters * 256.0 * blocks / (ms * 1e-3) / 1e9;
        if (rep > 0 && g > best) { best = g; }
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    count_launch(c, 4);
    return best;
}

template double probe_fma_tflops<float>(ctx &);
template double probe_fma_tflops<double>(ctx &);

}  // namespace b200

// ---------------------------------------------------------------------------------------------------------------------
// b200svm_update_w_noalloc: w = X^T alpha with the chunk partials of k_xt_v in a file-scope static device array instead of a
// cudaMalloc'ed buffer (include/b200svm.h).  No counterpart in the reference; added for callers
// that cannot allocate (CUDA graph capture, serving loops).  Nothing here allocates: no handle, no dbuf, no std::string.
// ---------------------------------------------------------------------------------------------------------------------
namespace b200 {
namespace {

// every GPU that loads libb200svm.so holds these B200SVM_NOALLOC_SCRATCH_BYTES until the process exits
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200SVM_NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
void *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};  // device address of g_noalloc_scratch per device (the symbol itself is a host stand-in)

// w[f] = sum_chunk wpart[chunk][f] for the F real features only (k_sum_chunks writes all ld columns of a handle-owned buffer)
template <typename T>
__global__ void __launch_bounds__(256) k_sum_chunks_noalloc(const T *__restrict__ wpart, int64_t ld, int64_t F, int chunks, T *__restrict__ w) {
    const int64_t f = (int64_t) blockIdx.x * 256 + threadIdx.x;
    if (f >= F) { return; }
    T s = 0;
    for (int ch = 0; ch < chunks; ++ch) { s += wpart[(int64_t) ch * ld + f]; }
    w[f] = s;
}

int noalloc_fail(char *err, size_t err_len, int code, const char *fmt, ...) {
    if (err != nullptr && err_len > 0) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

// device that owns a device pointer; -1 if it is not device memory
int noalloc_device_of(const void *p) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return -1;
    }
    return (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) ? at.device : -1;
}

template <typename T>
int update_w_noalloc(const T *X, int64_t n, int64_t F, int64_t ld, const T *alpha, T *w, cudaStream_t stream, void *scratch, char *err, size_t err_len) {
    // same chunking as xt_v above, so the result equals b200svm_update_w's when the leading dimensions agree
    const int64_t fblocks = (ld + 255) / 256;
    int64_t chunks = (4 * NUM_SMS + fblocks - 1) / fblocks;
    if (chunks > (n + 63) / 64) { chunks = (n + 63) / 64; }
    if (chunks < 1) { chunks = 1; }
    const int64_t rpc = (n + chunks - 1) / chunks;
    chunks = (n + rpc - 1) / rpc;
    const unsigned long long need = (unsigned long long) chunks * (unsigned long long) ld * sizeof(T);
    if (need > (unsigned long long) B200SVM_NOALLOC_SCRATCH_BYTES) {
        return noalloc_fail(err, err_len, B200SVM_ERR_INVALID,
                            "b200svm_update_w_noalloc: %lld chunks x ld %lld x %zu bytes = %llu bytes of scratch exceed the static array of %llu bytes "
                            "(B200SVM_NOALLOC_SCRATCH_BYTES); use b200svm_update_w for this size",
                            (long long) chunks, (long long) ld, sizeof(T), need, (unsigned long long) B200SVM_NOALLOC_SCRATCH_BYTES);
    }
    if (fblocks > 2147483647LL || chunks > 65535) { return noalloc_fail(err, err_len, B200SVM_ERR_INVALID, "b200svm_update_w_noalloc: grid too large"); }
    // every (chunk, f < ld) element that k_sum_chunks_noalloc reads is written by k_xt_v first: nothing of the previous call's
    // contents survives into the result, so nothing has to be cleared
    T *wpart = static_cast<T *>(scratch);
    k_xt_v<T><<<dim3((unsigned) fblocks, (unsigned) chunks), 256, 0, stream>>>(X, n, ld, alpha, wpart, rpc);
    k_sum_chunks_noalloc<T><<<(unsigned) ((F + 255) / 256), 256, 0, stream>>>(wpart, ld, F, (int) chunks, w);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { return noalloc_fail(err, err_len, B200SVM_ERR_CUDA, "b200svm_update_w_noalloc: kernel launch failed: %s", cudaGetErrorString(e)); }
    return B200SVM_OK;
}

}  // namespace
}  // namespace b200

extern "C" int b200svm_update_w_noalloc(int dtype, const void *X_dev, int64_t n, int64_t F, int64_t ld, const void *alpha_dev, void *w_out_dev,
                                        void *cuda_stream, char *err, size_t err_len) {
    using namespace b200;
    if (err != nullptr && err_len > 0) { err[0] = '\0'; }
    if (dtype != B200SVM_F32 && dtype != B200SVM_F64) { return noalloc_fail(err, err_len, B200SVM_ERR_INVALID, "b200svm_update_w_noalloc: dtype must be B200SVM_F32 or B200SVM_F64"); }
    if (X_dev == nullptr || alpha_dev == nullptr || w_out_dev == nullptr) { return noalloc_fail(err, err_len, B200SVM_ERR_INVALID, "b200svm_update_w_noalloc: null pointer"); }
    if (n < 1 || F < 1 || ld < F) { return noalloc_fail(err, err_len, B200SVM_ERR_INVALID, "b200svm_update_w_noalloc: need n >= 1, F >= 1 and ld >= F (got n %lld, F %lld, ld %lld)", (long long) n, (long long) F, (long long) ld); }
    const int dev = noalloc_device_of(X_dev);
    if (dev < 0 || noalloc_device_of(alpha_dev) != dev || noalloc_device_of(w_out_dev) != dev) {
        return noalloc_fail(err, err_len, B200SVM_ERR_INVALID, "b200svm_update_w_noalloc: X, alpha and w_out must be device pointers of one GPU (there is no CPU fallback and no staging copy)");
    }
    if (dev >= NOALLOC_MAX_DEVICES) { return noalloc_fail(err, err_len, B200SVM_ERR_INVALID, "b200svm_update_w_noalloc: device ordinal %d >= %d", dev, NOALLOC_MAX_DEVICES); }
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != dev) { e = cudaSetDevice(dev); }
    if (e != cudaSuccess) { return noalloc_fail(err, err_len, B200SVM_ERR_CUDA, "b200svm_update_w_noalloc: cannot select device %d: %s", dev, cudaGetErrorString(e)); }
    int rc = B200SVM_OK;
    // once per device, with that device current: the device address of the static array (racing first calls store the same value)
    void *scratch = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
    if (scratch == nullptr) {
        e = cudaGetSymbolAddress(&scratch, g_noalloc_scratch);
        if (e != cudaSuccess || scratch == nullptr) {
            rc = noalloc_fail(err, err_len, B200SVM_ERR_CUDA, "b200svm_update_w_noalloc: cudaGetSymbolAddress of the static scratch array failed on device %d: %s", dev, cudaGetErrorString(e));
        } else {
            __atomic_store_n(&g_noalloc_ptr[dev], scratch, __ATOMIC_RELEASE);
        }
    }
    if (rc == B200SVM_OK) {
        cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
        rc = dtype == B200SVM_F32 ? update_w_noalloc<float>(static_cast<const float *>(X_dev), n, F, ld, static_cast<const float *>(alpha_dev), static_cast<float *>(w_out_dev), stream, scratch, err, err_len)
                                  : update_w_noalloc<double>(static_cast<const double *>(X_dev), n, F, ld, static_cast<const double *>(alpha_dev), static_cast<double *>(w_out_dev), stream, scratch, err, err_len);
    }
    if (prev >= 0 && prev != dev) { cudaSetDevice(prev); }
    return rc;
}
