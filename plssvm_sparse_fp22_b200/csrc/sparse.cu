// Sparse (CSR / FP22-packed) hot path: q-vector, implicit kernel-matrix x vector, w-vector and predict over sparse rows.
//
// The reference has no sparse in-memory representation: it densifies LIBSVM files and runs
// src/plssvm/backends/OpenMP/svm_kernel.cpp:21-47 over dense rows.  The semantics restated here are those of that kernel and of
// q_kernel.cpp:19-50 / csvm.cpp:173-244 applied to rows whose zeros are not stored.
//
// Algorithm.  With G = X X^T:
//   rbf         K_ij = u_i u_j (1 + expm1(2 gamma G_ij)),  u_i = exp(-gamma |x_i|^2)
//   polynomial  K_ij = c0^deg + [(gamma G_ij + c0)^deg - c0^deg]
// The bracket S_ij vanishes unless rows i and j share a feature, so K v = rank-1 part + S v and only sum_f c_f^2 products are formed
// instead of m^2 F.  G is never stored: for every (row i, block b of W columns) a warp accumulates x_if * x_jf over the features of
// row i into a warp-private shared-memory accumulator, driven by a column-blocked transpose of X (per block, per feature: the list
// of (j_local, x_jf)), compacts the touched slots from a bitmap and evaluates the kernel function only on those.
//   linear      K v = X (X^T v): two O(nnz) passes over the same structures.
// FP22 values (1|8|13 = top 22 bits of binary32, packed with a 10-bit column delta / local column) are decoded in registers.
#include "device_utils.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace b200 {

namespace sp {

// ---------------------------------------------------------------------------------------------------------------------
// device views
// ---------------------------------------------------------------------------------------------------------------------
struct csr_view {
    const int64_t *rowptr;   // [rows + 1] (FP22: counts packed entries, padding included)
    const int32_t *col;      // [nnz]   (unused for FP22)
    const void *val;         // T[nnz]  or uint32 packed (fp22 << 10 | column_delta)
    int64_t rows;
};

// Transpose of X (CSC, rows ascending inside a column) cut into column blocks of the implicit matrix:
//   block b  = rows [b W, (b+1) W) of X  -> one pass of the warp-private accumulator
//   group g  = G consecutive blocks      -> one kernel launch (its slice of the transpose stays L2-resident)
// A lane that owns feature f walks column f sequentially across the blocks of a group, so its loads are contiguous; the
// per-(feature, block) end pointers give every segment a known trip count (loads issue back to back, no pointer chasing).
struct csc_view {
    const uint32_t *ptrT;    // [F][nb + 1] index of the first entry of column f inside block b; ptrT[f][nb] = end of the column
    const uint16_t *j;       // [entries] row of X relative to its block (plain formats)
    const void *val;         // T[entries], or uint32 packed (fp22 << 10 | row_in_block), W <= 1024
    int64_t F;
    int W;                   // rows of X per block (power of two, 128..4096; <= 1024 for FP22)
    int nb;                  // blocks
    int G;                   // blocks per group
    int ng;                  // groups
};

__device__ __forceinline__ float fp22_value(uint32_t w) { return __uint_as_float(w & 0xFFFFFC00u); }

__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) { v += t; }
    }
    return v;
}

// One chunk of up to 32 stored entries of a row: lane -> (feature f, value x).  FP22: the column is the running sum of the
// 10-bit deltas (col_0 = delta_0), carried across chunks in `colrun`; padding entries decode to x = +0.
template <typename T, bool FP22>
__device__ __forceinline__ bool load_entry(const csr_view &A, int64_t p, int64_t p1, int lane, int &colrun, int &f, T &x) {
    const bool valid = p + lane < p1;
    if constexpr (FP22) {
        const uint32_t w = valid ? __ldg(static_cast<const uint32_t *>(A.val) + p + lane) : 0u;
        const int incl = warp_incl_scan((int) (w & 1023u), lane);
        f = colrun + incl;
        colrun += __shfl_sync(0xffffffffu, incl, 31);
        x = (T) fp22_value(w);
    } else {
        f = valid ? __ldg(A.col + p + lane) : 0;
        x = valid ? __ldg(static_cast<const T *>(A.val) + p + lane) : T(0);
    }
    return valid;
}

// ---------------------------------------------------------------------------------------------------------------------
// O(nnz) kernels: one warp per row
// ---------------------------------------------------------------------------------------------------------------------
// norms[i] = |x_i|^2 ; u[i] = exp(-gamma |x_i|^2) (rbf only, else 1)
template <typename T, bool FP22>
__global__ void __launch_bounds__(256) k_sp_norms(csr_view A, T gamma, int rbf, T *__restrict__ norms, T *__restrict__ u) {
    const int64_t row = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= A.rows) { return; }
    const int lane = threadIdx.x & 31;
    const int64_t p0 = A.rowptr[row], p1 = A.rowptr[row + 1];
    T s = 0;
    int colrun = 0;
    for (int64_t p = p0; p < p1; p += 32) {
        int f;
        T x;
        load_entry<T, FP22>(A, p, p1, lane, colrun, f, x);
        s = fma(x, x, s);
    }
    s = warp_sum(s);
    if (lane == 0) {
        norms[row] = s;
        if (u != nullptr) { u[row] = rbf ? (T) exp(-(double) gamma * (double) s) : T(1); }
    }
}

// dense[f] = x_row[f] (dense must be zero before)
template <typename T, bool FP22>
__global__ void __launch_bounds__(32) k_sp_densify_row(csr_view A, int64_t row, T *__restrict__ dense) {
    const int lane = threadIdx.x;
    const int64_t p0 = A.rowptr[row], p1 = A.rowptr[row + 1];
    int colrun = 0;
    for (int64_t p = p0; p < p1; p += 32) {
        int f;
        T x;
        if (load_entry<T, FP22>(A, p, p1, lane, colrun, f, x) && x != T(0)) { dense[f] = x; }
    }
}

// out[i] = x_i . w  for rows [0, A.rows)       MODE 0: out = bias + dot
//                                              MODE 1: out = kernel_value(dot, norms[i], norm_last)  (q-vector; row == last -> scal[S_KXX])
template <typename T, bool FP22, int MODE, int KERNEL>
__global__ void __launch_bounds__(256) k_sp_row_dot(csr_view A, const T *__restrict__ w, T bias, const T *__restrict__ norms, int64_t last, int degree,
                                                    T gamma, T coef0, T *__restrict__ out, T *__restrict__ scal) {
    const int64_t row = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= A.rows) { return; }
    const int lane = threadIdx.x & 31;
    const int64_t p0 = A.rowptr[row], p1 = A.rowptr[row + 1];
    T s = 0;
    int colrun = 0;
    for (int64_t p = p0; p < p1; p += 32) {
        int f;
        T x;
        if (load_entry<T, FP22>(A, p, p1, lane, colrun, f, x)) { s = fma(x, w[f], s); }
    }
    s = warp_sum(s);
    if (lane == 0) {
        if constexpr (MODE == 0) {
            out[row] = bias + s;
        } else {
            const T v = kernel_value<T, KERNEL>(s, norms[row], norms[last], degree, gamma, coef0);
            if (row < last) {
                out[row] = v;
            } else if (row == last) {
                scal[S_KXX] = KERNEL == B200SVM_RBF ? T(1) : v;  // exp(-gamma * 0) is exactly 1 in the reference
            }
        }
    }
}

// w[f] = sum_j x_jf v_j  via the transpose (thread per feature, fixed order => deterministic); v has nb * W entries
template <typename T, bool FP22>
__global__ void __launch_bounds__(256) k_sp_xt_v(csc_view B, const T *__restrict__ v, T *__restrict__ w) {
    const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= B.F) { return; }
    const uint32_t *pt = B.ptrT + f * (B.nb + 1);
    T s = 0;
    for (int b = 0; b < B.nb; ++b) {
        const T *vb = v + (int64_t) b * B.W;
        for (uint32_t e = pt[b], e1 = pt[b + 1]; e < e1; ++e) {
            if constexpr (FP22) {
                const uint32_t wd = __ldg(static_cast<const uint32_t *>(B.val) + e);
                s = fma((T) fp22_value(wd), vb[wd & 1023u], s);
            } else {
                s = fma(__ldg(static_cast<const T *>(B.val) + e), vb[__ldg(B.j + e)], s);
            }
        }
    }
    w[f] = s;
}

// ---------------------------------------------------------------------------------------------------------------------
// S(g): the part of the kernel value that vanishes when two rows share no feature
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double expm1_fast(double z) {
    if (fabs(z) < 0x1p-7) {  // Taylor to z^7: truncation z^7 / 5040 -> relative 2^-42 / 5040 < 1e-16
        double p = fma(z, 1.0 / 7.0, 1.0);
        p = fma(z * (1.0 / 6.0), p, 1.0);
        p = fma(z * (1.0 / 5.0), p, 1.0);
        p = fma(z * (1.0 / 4.0), p, 1.0);
        p = fma(z * (1.0 / 3.0), p, 1.0);
        p = fma(z * (1.0 / 2.0), p, 1.0);
        return z * p;
    }
    return expm1(z);
}
__device__ __forceinline__ float expm1_fast(float z) {
    if (fabsf(z) < 0x1p-4f) {  // Taylor to z^5: truncation relative z^5 / 720 < 2e-9
        float p = fmaf(z, 1.f / 5.f, 1.f);
        p = fmaf(z * (1.f / 4.f), p, 1.f);
        p = fmaf(z * (1.f / 3.f), p, 1.f);
        p = fmaf(z * (1.f / 2.f), p, 1.f);
        return z * p;
    }
    return expm1f(z);
}

struct kfun {
    int degree;
    double a;       // rbf: 2 gamma ; polynomial: gamma
    double coef0;
    double c0pow;   // coef0^degree
    double fx_scale, fx_inv;  // 2^s and 2^-s of the fixed-point Gram accumulation
};

template <typename T, int KERNEL>
__device__ __forceinline__ T s_value(T g, const kfun &k) {
    if constexpr (KERNEL == B200SVM_RBF) {
        return expm1_fast((T) k.a * g);
    } else {
        const T base = fma((T) k.a, g, (T) k.coef0);
        T r;
        switch (k.degree) {  // warp-uniform
            case 1: r = base; break;
            case 2: r = base * base; break;
            case 3: r = base * base * base; break;
            case 4: r = (base * base) * (base * base); break;
            default: r = ipow(base, k.degree); break;
        }
        return r - (T) k.c0pow;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// out[row] += sum_{j in group g} S(G_row,j) v_j      rows [row_lo, row_hi) of A, grid-strided over warps
//   per-warp shared memory: acc T[W] | touched bitmap uint32[W/32] | touched list uint16[W]
//   A lane keeps up to K features of the row in registers (value, feature, cursor into the feature's column) and, block after
//   block, consumes the entries whose row falls into the block (segment end from ptrT: known trip count, loads batched by 4).
//   COMPACT: touched slots are recorded in a bitmap and compacted, the kernel function runs on those only (sparse fill);
//   otherwise all W slots are evaluated (S(0) = 0), which is cheaper when most slots are hit.
// ---------------------------------------------------------------------------------------------------------------------
// Fixed-point accumulation of the Gram entries in 32-bit limbs: shared-memory floating-point atomics are CAS loops on sm_100,
// 32-bit integer atomics are single fire-and-forget instructions.  A product p is scaled by 2^s (s chosen from the largest
// squared row norm, which bounds every partial sum by Cauchy-Schwarz), rounded to an integer q and split into limbs with 10 bits
// of head-room each, so up to 1024 products can be added per slot without a carry; limbs are recombined in the epilogue.
// The sum is exact in the fixed-point grid (2^-42 / 2^-61 of the bound for float / double) and independent of the order of the
// additions => bit-reproducible.
template <typename T>
struct fixp;
template <>
struct fixp<float> {
    static constexpr int LIMBS = 2;
    static constexpr int BITS = 42;  // |q| <= 2^42
    static __device__ __forceinline__ void add(uint32_t *acc, int W, int jl, float p, float scale) {
        const long long q = __float2ll_rn(p * scale);
        atomicAdd(&acc[jl], (uint32_t) (q & 0x3FFFFFll));
        atomicAdd(reinterpret_cast<int *>(&acc[W + jl]), (int) (q >> 22));
    }
    static __device__ __forceinline__ float take(uint32_t *acc, int W, int jl, double inv_scale) {
        const uint32_t lo = acc[jl];
        const int hi = (int) acc[W + jl];
        acc[jl] = 0u;
        acc[W + jl] = 0u;
        const long long q = ((long long) hi << 22) + (long long) lo;
        return (float) q * (float) inv_scale;  // 2^-s: exact scaling, one rounding of q to 24 bits
    }
};
template <>
struct fixp<double> {
    static constexpr int LIMBS = 3;
    static constexpr int BITS = 61;  // |q| <= 2^61
    static __device__ __forceinline__ void add(uint32_t *acc, int W, int jl, double p, double scale) {
        const long long q = __double2ll_rn(p * scale);
        atomicAdd(&acc[jl], (uint32_t) (q & 0x1FFFFFll));
        atomicAdd(&acc[W + jl], (uint32_t) ((q >> 21) & 0x1FFFFFll));
        atomicAdd(reinterpret_cast<int *>(&acc[2 * W + jl]), (int) (q >> 42));
    }
    static __device__ __forceinline__ double take(uint32_t *acc, int W, int jl, double inv_scale) {
        const uint32_t lo = acc[jl], mid = acc[W + jl];
        const int hi = (int) acc[2 * W + jl];
        acc[jl] = 0u;
        acc[W + jl] = 0u;
        acc[2 * W + jl] = 0u;
        const long long q = ((long long) hi << 42) + ((long long) mid << 21) + (long long) lo;
        return (double) q * inv_scale;
    }
};

template <typename T, bool FP22>
__device__ __forceinline__ void load_b_entry(const csc_view &B, uint32_t e, int &jl, T &xv) {
    if constexpr (FP22) {
        const uint32_t wd = __ldg(static_cast<const uint32_t *>(B.val) + e);
        jl = (int) (wd & 1023u);
        xv = (T) fp22_value(wd);
    } else {
        jl = (int) __ldg(B.j + e);
        xv = __ldg(static_cast<const T *>(B.val) + e);
    }
}

template <typename T, int KERNEL, bool FP22, int K, int MAXT, bool COMPACT, int GS>
__global__ void __launch_bounds__(MAXT, 1)
    k_sp_kmv(csr_view A, int64_t row_lo, int64_t row_hi, csc_view B, int g, const T *__restrict__ v, kfun kf, T *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char sp_smem[];
    using FX = fixp<T>;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int W = B.W, words = W >> 5;
    const size_t per_warp = (size_t) W * 4 * FX::LIMBS + (size_t) words * 4 + (size_t) W * 2;
    unsigned char *mine = sp_smem + (size_t) warp * per_warp;
    uint32_t *acc = reinterpret_cast<uint32_t *>(mine);
    uint32_t *bits = acc + (size_t) W * FX::LIMBS;
    uint16_t *list = reinterpret_cast<uint16_t *>(bits + words);
    for (int k = lane; k < W * FX::LIMBS + words; k += 32) { acc[k] = 0u; }
    const T fx_scale = (T) kf.fx_scale;
    const double fx_inv = kf.fx_inv;
    __syncwarp();

    constexpr int UNR = K <= 8 ? K : 1;  // the 32-features-per-lane fallback keeps its state in local memory
    const int b0 = g * B.G, b1 = min(B.nb, b0 + B.G);
    const int64_t gwarp = (int64_t) blockIdx.x * nwarps + warp, twarps = (int64_t) gridDim.x * nwarps;
    for (int64_t row = row_lo + gwarp; row < row_hi; row += twarps) {
        const int64_t p0 = A.rowptr[row], p1 = A.rowptr[row + 1];
        // ---- this lane's features: value, end-pointer row of ptrT, cursor
        T x[K];
        const uint32_t *pt[K];
        uint32_t cur[K];
        int colrun = 0;
#pragma unroll UNR
        for (int k = 0; k < K; ++k) {
            int f = 0;
            x[k] = T(0);
            pt[k] = nullptr;
            cur[k] = 0u;
            if (p0 + (int64_t) k * 32 < p1) {  // warp-uniform
                const bool valid = load_entry<T, FP22>(A, p0 + (int64_t) k * 32, p1, lane, colrun, f, x[k]);
                if (valid && x[k] != T(0)) {
                    pt[k] = B.ptrT + (int64_t) f * (B.nb + 1) + b0;
                    cur[k] = __ldg(pt[k]);
                }
            }
        }
        T rowsum = T(0);
        for (int b = b0; b < b1; ++b) {
            // ---- segment ends of this block (independent loads), then the products
            uint32_t e1[K];
#pragma unroll UNR
            for (int k = 0; k < K; ++k) { e1[k] = pt[k] != nullptr ? __ldg(pt[k] + (b - b0) + 1) : cur[k]; }
            // GS lanes share one feature's segment (GS = 1: a lane walks its own features; larger GS for longer segments)
            constexpr int NG = 32 / GS;
#pragma unroll UNR
            for (int k = 0; k < K; ++k) {
                if (p0 + (int64_t) k * 32 >= p1) { break; }  // warp-uniform: no more stored entries in this row
                for (int l0 = 0; l0 < 32; l0 += NG) {
                    if (p0 + (int64_t) k * 32 + l0 >= p1) { break; }
                    uint32_t e, ee;
                    T xs;
                    if constexpr (GS == 1) {
                        e = cur[k];
                        ee = e1[k];
                        xs = x[k];
                    } else {
                        const int src = l0 + lane / GS;
                        e = __shfl_sync(0xffffffffu, cur[k], src) + (uint32_t) (lane % GS);
                        ee = __shfl_sync(0xffffffffu, e1[k], src);
                        xs = __shfl_sync(0xffffffffu, x[k], src);
                    }
                    for (; e < ee; e += GS) {
                        int jl;
                        T xv;
                        load_b_entry<T, FP22>(B, e, jl, xv);
                        FX::add(acc, W, jl, xs * xv, fx_scale);  // lanes of one warp may meet on a slot
                        if constexpr (COMPACT) { atomicOr(&bits[jl >> 5], 1u << (jl & 31)); }
                    }
                }
                cur[k] = e1[k];
            }
            __syncwarp();
            const T *vb = v + (int64_t) b * W;
            if constexpr (COMPACT) {
                // ---- compact the touched slots (bitmap -> list), fixed order
                int cnt = 0;
                for (int w0 = 0; w0 < words; w0 += 32) {
                    uint32_t word = (w0 + lane < words) ? bits[w0 + lane] : 0u;
                    const int c = __popc(word);
                    const int incl = warp_incl_scan(c, lane);
                    int pos = cnt + incl - c;
                    if (word != 0u) { bits[w0 + lane] = 0u; }
                    while (word != 0u) {
                        const int bit = __ffs((int) word) - 1;
                        list[pos++] = (uint16_t) (((w0 + lane) << 5) + bit);
                        word &= word - 1u;
                    }
                    cnt += __shfl_sync(0xffffffffu, incl, 31);
                }
                __syncwarp();
                // ---- kernel function on the touched slots only, times v_j
                for (int t = lane; t < cnt; t += 32) {
                    const int jl = list[t];
                    const T gv = FX::take(acc, W, jl, fx_inv);
                    rowsum = fma(s_value<T, KERNEL>(gv, kf), __ldg(vb + jl), rowsum);
                }
            } else {
                for (int jl = lane; jl < W; jl += 32) {
                    const T gv = FX::take(acc, W, jl, fx_inv);
                    rowsum = fma(s_value<T, KERNEL>(gv, kf), __ldg(vb + jl), rowsum);
                }
            }
            __syncwarp();
        }
        rowsum = warp_sum(rowsum);
        if (lane == 0) { out[row] += rowsum; }
    }
}

// weights of the columns: v_j = d_j u_j (rbf) or d_j, zero-extended to nbW ; also sums[0] = sum v (two-stage deterministic)
template <typename T>
__global__ void __launch_bounds__(256) k_sp_weights(const T *__restrict__ d, const T *__restrict__ u, int64_t m, int64_t nbW, T *__restrict__ v,
                                                    T *__restrict__ scal, int idx_sumv, T *partials, unsigned int *ticket) {
    __shared__ T red[32];
    T acc[1] = { T(0) };
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nbW; i += (int64_t) gridDim.x * blockDim.x) {
        T val = T(0);
        if (i < m) { val = u != nullptr ? d[i] * u[i] : d[i]; }
        v[i] = val;
        acc[0] += val;
    }
    T res[1];
    if (grid_sum<T, 1>(acc, partials, ticket, red, res)) { scal[idx_sumv] = res[0]; }
}

// K v from the S-part:  rbf  out_i = u_i (sum_v + S_i) ; polynomial  out_i = c0^deg sum_v + S_i ;  (+ bias)
// rows [lo, hi) only; the rest of [0, rows) is zeroed when ZERO_REST (multi-GPU partial vectors)
template <typename T>
__global__ void __launch_bounds__(256) k_sp_finish(T *__restrict__ S, const T *__restrict__ u, int64_t rows, int64_t lo, int64_t hi, int rbf, T c0pow,
                                                   const T *__restrict__ scal, int idx_sumv, T bias, int zero_rest) {
    const T sumv = scal[idx_sumv];
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += (int64_t) gridDim.x * blockDim.x) {
        if (i >= lo && i < hi) {
            S[i] = bias + (rbf ? u[i] * (sumv + S[i]) : fma(c0pow, sumv, S[i]));
        } else if (zero_rest) {
            S[i] = T(0);
        }
    }
}

}  // namespace sp

// ---------------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------------
namespace {

struct sparse_state {
    bool fp22{ false };
    int64_t n{ 0 }, F{ 0 };
    // rows (A side)
    dbuf a_rowptr, a_col, a_val;
    int64_t a_entries{ 0 };
    int64_t a_max_row{ 0 };  // longest row (stored entries)
    double max_norm2{ 0 };   // largest squared row norm (bounds every Gram entry)
    // transpose (B side): CSC + per-group column pointers
    dbuf b_ptrT, b_j, b_val;
    int W{ 0 }, nb{ 0 }, G{ 1 }, ng{ 1 };
    int64_t b_entries{ 0 };
    double avg_seg{ 0 };     // mean stored entries per (feature, block) segment
    double fill{ 0 };  // expected fraction of accumulator slots hit per (row, block) visit
    // per-row data and work vectors
    dbuf norms, u, v, xl, wvec;
    // launch geometry of k_sp_kmv
    int warps{ 0 };
    size_t smem{ 0 };
    // predict staging
    dbuf p_rowptr, p_col, p_val, p_norms, p_u, p_out;
};

sparse_state &state_of(ctx &c) {
    if (c.sp_opaque == nullptr) {
        c.sp_opaque = new sparse_state();
        c.sp_opaque_free = [](void *p) { delete static_cast<sparse_state *>(p); };
    }
    return *static_cast<sparse_state *>(c.sp_opaque);
}

sp::csr_view view_rows(const sparse_state &st, int64_t lo, int64_t hi) {
    return sp::csr_view{ st.a_rowptr.as<int64_t>() + lo, st.a_col.as<int32_t>(), st.a_val.p, hi - lo };
}
sp::csc_view view_blocks(const sparse_state &st) {
    return sp::csc_view{ st.b_ptrT.as<uint32_t>(), st.b_j.as<uint16_t>(), st.b_val.p, st.F, st.W, st.nb, st.G, st.ng };
}

void upload(ctx &c, dbuf &dst, const void *src, size_t bytes) {
    dst.reserve(std::max<size_t>(bytes, 16));
    if (bytes != 0) {
        B200_CUDA_CHECK(cudaMemcpyAsync(dst.p, src, bytes, cudaMemcpyHostToDevice, c.stream));
        c.stats.h2d_bytes += (int64_t) bytes;
    }
}

// Host copy of the CSR input in the handle's real type (FP22: decoded values, absolute columns, padding entries dropped).
template <typename T>
struct host_csr {
    std::vector<int64_t> rowptr;
    std::vector<int32_t> col;
    std::vector<T> val;
    std::vector<uint32_t> bits;  // FP22: the 22-bit value pattern (already shifted: word & ~1023)
};

template <typename T>
host_csr<T> decode_input(const int64_t *rowptr, const int32_t *colidx, const void *values, int64_t n, int64_t F, int fmt) {
    host_csr<T> h;
    h.rowptr.resize((size_t) n + 1);
    const int64_t entries = rowptr[n];
    h.col.reserve((size_t) entries);
    h.val.reserve((size_t) entries);
    h.rowptr[0] = 0;
    for (int64_t i = 0; i < n; ++i) {
        B200_REQUIRE(rowptr[i + 1] >= rowptr[i], "CSR row pointers must be non-decreasing");
        int64_t colrun = 0;
        int32_t prev = -1;
        for (int64_t p = rowptr[i]; p < rowptr[i + 1]; ++p) {
            if (fmt == B200SVM_VAL_FP22) {
                const uint32_t w = static_cast<const uint32_t *>(values)[p];
                colrun += (int64_t) (w & 1023u);
                const uint32_t vb = w & 0xFFFFFC00u;
                if ((vb & 0x7FFFFFFFu) == 0u) { continue; }  // padding entry / stored zero
                B200_REQUIRE(colrun < F, "FP22 column delta runs past the number of features");
                float x;
                std::memcpy(&x, &vb, 4);
                h.col.push_back((int32_t) colrun);
                h.val.push_back((T) x);
                h.bits.push_back(vb);
            } else {
                const int32_t cidx = colidx[p];
                B200_REQUIRE(cidx >= 0 && cidx < F, "CSR column index out of range");
                B200_REQUIRE(cidx > prev, "CSR column indices must be sorted and unique within a row");
                prev = cidx;
                const T x = fmt == B200SVM_VAL_F32 ? (T) static_cast<const float *>(values)[p] : (T) static_cast<const double *>(values)[p];
                h.col.push_back(cidx);
                h.val.push_back(x);
            }
        }
        h.rowptr[(size_t) i + 1] = (int64_t) h.col.size();
    }
    return h;
}

// Transpose (CSC, rows ascending) + per-(feature, block) pointers.  FP22: entries are re-packed as (fp22 << 10 | row_in_block).
template <typename T>
void build_transpose(ctx &c, sparse_state &st, const host_csr<T> &h) {
    const int64_t n = st.n, F = st.F;
    const int64_t nnz = (int64_t) h.col.size();
    const int nb = st.nb, W = st.W;
    B200_REQUIRE(nnz < (int64_t) 0xFFFFFFFFll, "more than 2^32 - 1 stored entries are not supported by the sparse path");
    std::vector<int64_t> colptr((size_t) F + 1, 0);
    for (int64_t p = 0; p < nnz; ++p) { ++colptr[(size_t) h.col[(size_t) p] + 1]; }
    double sumsq = 0;
    for (int64_t f = 0; f < F; ++f) {
        sumsq += (double) colptr[(size_t) f + 1] * (double) colptr[(size_t) f + 1];
        colptr[(size_t) f + 1] += colptr[(size_t) f];
    }
    st.fill = n > 0 ? std::min(1.0, sumsq / ((double) n * (double) n)) : 0.0;  // products per visit / W (upper bound of the hit fraction)
    std::vector<uint16_t> tj;
    std::vector<T> tval;
    std::vector<uint32_t> pack;
    std::vector<int32_t> trow((size_t) nnz);  // global row of every transposed entry (for the block pointers)
    if (st.fp22) {
        pack.resize((size_t) nnz);
    } else {
        tj.resize((size_t) nnz);
        tval.resize((size_t) nnz);
    }
    {
        std::vector<int64_t> cursor(colptr.begin(), colptr.end() - 1);
        for (int64_t j = 0; j < n; ++j) {
            const uint32_t jl = (uint32_t) (j % W);
            for (int64_t p = h.rowptr[(size_t) j]; p < h.rowptr[(size_t) j + 1]; ++p) {
                const size_t e = (size_t) cursor[(size_t) h.col[(size_t) p]]++;
                trow[e] = (int32_t) j;
                if (st.fp22) {
                    pack[e] = h.bits[(size_t) p] | jl;
                } else {
                    tj[e] = (uint16_t) jl;
                    tval[e] = h.val[(size_t) p];
                }
            }
        }
    }
    std::vector<uint32_t> ptrT((size_t) F * (size_t) (nb + 1));
#pragma omp parallel for schedule(static)
    for (int64_t f = 0; f < F; ++f) {
        uint32_t *pt = ptrT.data() + (size_t) f * (size_t) (nb + 1);
        int64_t e = colptr[(size_t) f];
        const int64_t e1 = colptr[(size_t) f + 1];
        for (int b = 0; b < nb; ++b) {
            pt[b] = (uint32_t) e;
            const int64_t limit = (int64_t) (b + 1) * W;
            while (e < e1 && trow[(size_t) e] < limit) { ++e; }
        }
        pt[nb] = (uint32_t) e1;
    }
    st.b_entries = nnz;
    st.avg_seg = (double) nnz / std::max(1.0, (double) F * (double) nb);
    if (st.fp22) {
        st.b_j.reserve(16);
        upload(c, st.b_val, pack.data(), pack.size() * sizeof(uint32_t));
    } else {
        upload(c, st.b_j, tj.data(), tj.size() * sizeof(uint16_t));
        upload(c, st.b_val, tval.data(), tval.size() * sizeof(T));
    }
    upload(c, st.b_ptrT, ptrT.data(), ptrT.size() * sizeof(uint32_t));
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));  // host vectors go out of scope
}

// k_sp_kmv variants: K features per lane in registers -> rows of up to 32 K stored entries
template <typename T, int KERNEL, bool FP22>
void kmv_set_smem(int smem) {
#define B200_SP_ATTR(KK, MT, CP, GG) \
    B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, KERNEL, FP22, KK, MT, CP, GG>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem))
#define B200_SP_ATTR_G(KK, MT, CP) \
    B200_SP_ATTR(KK, MT, CP, 1);   \
    B200_SP_ATTR(KK, MT, CP, 8)
    B200_SP_ATTR_G(4, 1024, true);
    B200_SP_ATTR_G(8, 512, true);
    B200_SP_ATTR_G(32, 256, true);
    B200_SP_ATTR_G(4, 1024, false);
    B200_SP_ATTR_G(8, 512, false);
    B200_SP_ATTR_G(32, 256, false);
#undef B200_SP_ATTR_G
#undef B200_SP_ATTR
}

template <typename T>
void set_kmv_attributes(size_t smem) {
    static size_t configured = 0;
    if (smem <= configured) { return; }
    kmv_set_smem<T, B200SVM_RBF, false>((int) smem);
    kmv_set_smem<T, B200SVM_POLYNOMIAL, false>((int) smem);
    if constexpr (sizeof(T) == 4) {
        kmv_set_smem<T, B200SVM_RBF, true>((int) smem);
        kmv_set_smem<T, B200SVM_POLYNOMIAL, true>((int) smem);
    }
    configured = smem;
}

template <typename T>
void norms_and_u(ctx &c, const sparse_state &st, const sp::csr_view &A, T *norms, T *u) {
    if (A.rows == 0) { return; }
    const unsigned grid = (unsigned) ((A.rows + 7) / 8);
    const int rbf = c.kernel == B200SVM_RBF ? 1 : 0;
    if (st.fp22) {
        if constexpr (sizeof(T) == 4) { sp::k_sp_norms<T, true><<<grid, 256, 0, c.stream>>>(A, (T) c.gamma, rbf, norms, u); }
    } else {
        sp::k_sp_norms<T, false><<<grid, 256, 0, c.stream>>>(A, (T) c.gamma, rbf, norms, u);
    }
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}

template <typename T>
void set_data_typed(ctx &c, const int64_t *rowptr, const int32_t *colidx, const void *values, int64_t n, int64_t F, int fmt) {
    sparse_state &st = state_of(c);
    st.fp22 = fmt == B200SVM_VAL_FP22;
    st.n = n;
    st.F = F;
    const host_csr<T> h = decode_input<T>(rowptr, colidx, values, n, F, fmt);
    // rows: FP22 keeps the caller's packed stream (decoded in registers by the kernels); otherwise (col, val) in the handle's type
    if (st.fp22) {
        st.a_entries = rowptr[n];
        upload(c, st.a_rowptr, rowptr, ((size_t) n + 1) * sizeof(int64_t));
        upload(c, st.a_val, values, (size_t) st.a_entries * sizeof(uint32_t));
        st.a_col.reserve(16);
    } else {
        st.a_entries = (int64_t) h.col.size();
        upload(c, st.a_rowptr, h.rowptr.data(), h.rowptr.size() * sizeof(int64_t));
        upload(c, st.a_col, h.col.data(), h.col.size() * sizeof(int32_t));
        upload(c, st.a_val, h.val.data(), h.val.size() * sizeof(T));
    }
    st.max_norm2 = 0;
    for (int64_t i = 0; i < n; ++i) {
        double nrm = 0;
        for (int64_t p = h.rowptr[(size_t) i]; p < h.rowptr[(size_t) i + 1]; ++p) { nrm += (double) h.val[(size_t) p] * (double) h.val[(size_t) p]; }
        st.max_norm2 = std::max(st.max_norm2, nrm);
    }
    st.a_max_row = 0;
    {
        const int64_t *rpx = st.fp22 ? rowptr : h.rowptr.data();
        for (int64_t i = 0; i < n; ++i) { st.a_max_row = std::max(st.a_max_row, rpx[i + 1] - rpx[i]); }
    }
    // block width (rows of X per accumulator pass): option or default; shared memory per warp = W (sizeof(T) + 2) + W / 8
    const int W = c.opt_sparse_block > 0 ? (int) c.opt_sparse_block : 1024;
    B200_REQUIRE(W >= 128 && W <= 4096 && (W & (W - 1)) == 0, "sparse block width must be a power of two in [128, 4096]");
    B200_REQUIRE(!st.fp22 || W <= 1024, "FP22 storage packs a 10-bit row-in-block index: sparse block width must be <= 1024");
    st.W = W;
    st.nb = (int) ((n + W - 1) / W);
    // blocks per launch group: keep the group's slice of the transpose L2-resident (~48 MB)
    const double bytes_per_block = (double) h.col.size() / std::max(1, st.nb) * (st.fp22 ? 4.0 : 2.0 + sizeof(T)) + 4.0 * (double) F;
    st.G = (int) std::max(1.0, std::min((double) st.nb, 48.0e6 / std::max(1.0, bytes_per_block)));
    if (c.opt_sparse_group > 0) { st.G = (int) std::min<int64_t>(c.opt_sparse_group, st.nb); }
    st.ng = (st.nb + st.G - 1) / st.G;
    build_transpose<T>(c, st, h);
    const size_t per_warp = (size_t) W * 4 * (sizeof(T) == 4 ? 2 : 3) + (size_t) W / 8 + (size_t) W * 2;
    st.warps = (int) std::min<size_t>(32, (size_t) (220 * 1024) / per_warp);
    st.smem = per_warp * (size_t) st.warps;
    set_kmv_attributes<T>(st.smem);
    // per-row norms, u = exp(-gamma |x|^2), work vectors
    st.norms.reserve((size_t) n * sizeof(T));
    st.u.reserve((size_t) n * sizeof(T));
    st.v.reserve((size_t) st.nb * W * sizeof(T));
    st.xl.reserve((size_t) F * sizeof(T));
    st.wvec.reserve((size_t) F * sizeof(T));
    norms_and_u<T>(c, st, view_rows(st, 0, n), st.norms.as<T>(), st.u.as<T>());
    c.n = n;
    c.F = F;
    c.ldx = F;
    c.nnz = st.a_entries;
    c.value_format = fmt;
}

}  // namespace

void sp_set_data(ctx &c, const int64_t *rowptr, const int32_t *colidx, const void *values, int64_t n, int64_t F, int64_t nnz, int fmt) {
    B200_REQUIRE(rowptr != nullptr && rowptr[0] == 0 && rowptr[n] == nnz, "CSR row pointers inconsistent with nnz");
    B200_REQUIRE(values != nullptr || nnz == 0, "No data points provided!");
    B200_REQUIRE(fmt == B200SVM_VAL_FP22 || colidx != nullptr || nnz == 0, "null column index array");
    B200_REQUIRE(fmt >= B200SVM_VAL_F32 && fmt <= B200SVM_VAL_FP22, "unknown CSR value format");
    B200_REQUIRE(fmt != B200SVM_VAL_FP22 || c.dtype == B200SVM_F32, "FP22 value storage requires a float handle");
    if (c.dtype == B200SVM_F32) {
        set_data_typed<float>(c, rowptr, colidx, values, n, F, fmt);
    } else {
        set_data_typed<double>(c, rowptr, colidx, values, n, F, fmt);
    }
}

namespace {

#define B200_SP_DISPATCH_FP22(st, T, CALL_TRUE, CALL_FALSE) \
    do {                                                    \
        if ((st).fp22) {                                    \
            if constexpr (sizeof(T) == 4) { CALL_TRUE; }    \
        } else {                                            \
            CALL_FALSE;                                     \
        }                                                   \
    } while (0)

// max_norm2 bounds |G_ij| and every partial sum of its products (Cauchy-Schwarz): scale = 2^(BITS - ceil(log2 bound))
sp::kfun make_kfun(const ctx &c, double max_norm2 = 1.0) {
    sp::kfun k{};
    {
        const int bits = c.dtype == B200SVM_F32 ? 42 : 61;
        int e = 0;
        if (max_norm2 > 0 && std::isfinite(max_norm2)) { std::frexp(max_norm2, &e); }  // max_norm2 < 2^e
        k.fx_scale = std::ldexp(1.0, bits - e);
        k.fx_inv = std::ldexp(1.0, e - bits);
    }
    k.degree = c.degree;
    k.coef0 = c.coef0;
    if (c.kernel == B200SVM_RBF) {
        k.a = 2.0 * c.gamma;
        k.c0pow = 0.0;
    } else {
        k.a = c.gamma;
        k.c0pow = std::pow(c.coef0, (double) c.degree);
    }
    return k;
}

// out[row] (rows [lo, hi) of A, zero on entry) += sum_j S(G_row,j) v_j over all column groups; max_row = longest row of A
template <typename T, int KERNEL, bool FP22>
void launch_kmv_variant(ctx &c, const sparse_state &st, const sp::csr_view &A, int64_t lo, int64_t hi, int64_t max_row, double max_norm2,
                        const T *v, T *out) {
    const sp::csc_view B = view_blocks(st);
    const sp::kfun kf = make_kfun(c, max_norm2);
    const int64_t rows = hi - lo;
    B200_REQUIRE(max_row <= 1024, "rows with more than 1024 stored entries are not supported by the sparse path: pass the data dense");
    const int K = max_row <= 128 ? 4 : (max_row <= 256 ? 8 : 32);
    const int warps = std::min(st.warps, K == 4 ? 32 : (K == 8 ? 16 : 8));
    const int grid = (int) std::min<int64_t>(NUM_SMS, (rows + warps - 1) / warps);
    const size_t smem = st.smem / (size_t) st.warps * (size_t) warps;
    const bool compact = st.fill < 0.5;
    // lanes per (feature, block) segment: one lane walks its own features unless the segments are long (measured on B200:
    // 8 lanes per segment only pays beyond ~64 entries per segment)
    const int gs = c.opt_sparse_lanes > 0 ? (int) c.opt_sparse_lanes : (st.avg_seg < 64.0 ? 1 : 8);
    for (int g = 0; g < st.ng; ++g) {
#define B200_SP_LAUNCH_G(KK, MT, CP)                                                                                               \
    do {                                                                                                                           \
        if (gs == 1) {                                                                                                             \
            sp::k_sp_kmv<T, KERNEL, FP22, KK, MT, CP, 1><<<grid, warps * 32, smem, c.stream>>>(A, lo, hi, B, g, v, kf, out);       \
        } else {                                                                                                                   \
            sp::k_sp_kmv<T, KERNEL, FP22, KK, MT, CP, 8><<<grid, warps * 32, smem, c.stream>>>(A, lo, hi, B, g, v, kf, out);       \
        }                                                                                                                          \
    } while (0)
#define B200_SP_LAUNCH(KK, MT)                \
    do {                                      \
        if (compact) {                        \
            B200_SP_LAUNCH_G(KK, MT, true);   \
        } else {                              \
            B200_SP_LAUNCH_G(KK, MT, false);  \
        }                                     \
    } while (0)
        if (K == 4) {
            B200_SP_LAUNCH(4, 1024);
        } else if (K == 8) {
            B200_SP_LAUNCH(8, 512);
        } else {
            B200_SP_LAUNCH(32, 256);
        }
#undef B200_SP_LAUNCH_G
#undef B200_SP_LAUNCH
        count_launch(c);
        B200_CUDA_CHECK(cudaGetLastError());
    }
}

template <typename T>
void launch_kmv(ctx &c, const sparse_state &st, const sp::csr_view &A, int64_t lo, int64_t hi, int64_t max_row, double max_norm2, const T *v, T *out) {
    if (hi <= lo) { return; }
    if (c.kernel == B200SVM_RBF) {
        B200_SP_DISPATCH_FP22(st, T, (launch_kmv_variant<T, B200SVM_RBF, true>(c, st, A, lo, hi, max_row, max_norm2, v, out)),
                              (launch_kmv_variant<T, B200SVM_RBF, false>(c, st, A, lo, hi, max_row, max_norm2, v, out)));
    } else {
        B200_SP_DISPATCH_FP22(st, T, (launch_kmv_variant<T, B200SVM_POLYNOMIAL, true>(c, st, A, lo, hi, max_row, max_norm2, v, out)),
                              (launch_kmv_variant<T, B200SVM_POLYNOMIAL, false>(c, st, A, lo, hi, max_row, max_norm2, v, out)));
    }
}

template <typename T>
void weights(ctx &c, const T *d, const T *u, int64_t count, int64_t nbW, T *v) {
    int64_t g = (nbW + 1023) / 1024;
    g = std::max<int64_t>(1, std::min<int64_t>(g, MAX_PARTIAL_BLOCKS));
    sp::k_sp_weights<T><<<(unsigned) g, 256, 0, c.stream>>>(d, u, count, nbW, v, c.scal.as<T>(), S_TMP0, c.partials.as<T>(), c.counters.as<unsigned int>());
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}

template <typename T>
void row_dot(ctx &c, const sparse_state &st, const sp::csr_view &A, const T *w, T bias, T *out) {
    if (A.rows == 0) { return; }
    const unsigned grid = (unsigned) ((A.rows + 7) / 8);
    B200_SP_DISPATCH_FP22(st, T,
                          (sp::k_sp_row_dot<T, true, 0, B200SVM_LINEAR><<<grid, 256, 0, c.stream>>>(A, w, bias, nullptr, 0, 0, T(0), T(0), out, nullptr)),
                          (sp::k_sp_row_dot<T, false, 0, B200SVM_LINEAR><<<grid, 256, 0, c.stream>>>(A, w, bias, nullptr, 0, 0, T(0), T(0), out, nullptr)));
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}

template <typename T>
void xt_v_blocks(ctx &c, const sparse_state &st, const T *v, T *w) {
    const sp::csc_view B = view_blocks(st);
    const unsigned grid = (unsigned) ((st.F + 255) / 256);
    B200_SP_DISPATCH_FP22(st, T, (sp::k_sp_xt_v<T, true><<<grid, 256, 0, c.stream>>>(B, v, w)), (sp::k_sp_xt_v<T, false><<<grid, 256, 0, c.stream>>>(B, v, w)));
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}

}  // namespace

// q[i] = k(x_i, x_n), i < m ; k(x_n, x_n) -> scal[S_KXX]     (src/plssvm/backends/OpenMP/q_kernel.cpp:19-50)
template <typename T>
void sp_generate_q(ctx &c) {
    sparse_state &st = state_of(c);
    const int64_t n = st.n;
    B200_CUDA_CHECK(cudaMemsetAsync(st.xl.p, 0, (size_t) st.F * sizeof(T), c.stream));
    const sp::csr_view A = view_rows(st, 0, n);
    B200_SP_DISPATCH_FP22(st, T, (sp::k_sp_densify_row<T, true><<<1, 32, 0, c.stream>>>(A, n - 1, st.xl.as<T>())),
                          (sp::k_sp_densify_row<T, false><<<1, 32, 0, c.stream>>>(A, n - 1, st.xl.as<T>())));
    const unsigned grid = (unsigned) ((n + 7) / 8);
    const T *xl = st.xl.as<T>();
    const T *nr = st.norms.as<T>();
    T *q = c.q.as<T>();
    T *sc = c.scal.as<T>();
#define B200_SP_Q(KN)                                                                                                                                  \
    B200_SP_DISPATCH_FP22(st, T, (sp::k_sp_row_dot<T, true, 1, KN><<<grid, 256, 0, c.stream>>>(A, xl, T(0), nr, n - 1, c.degree, (T) c.gamma, (T) c.coef0, q, sc)), \
                          (sp::k_sp_row_dot<T, false, 1, KN><<<grid, 256, 0, c.stream>>>(A, xl, T(0), nr, n - 1, c.degree, (T) c.gamma, (T) c.coef0, q, sc)))
    switch (c.kernel) {
        case B200SVM_LINEAR: B200_SP_Q(B200SVM_LINEAR); break;
        case B200SVM_POLYNOMIAL: B200_SP_Q(B200SVM_POLYNOMIAL); break;
        default: B200_SP_Q(B200SVM_RBF); break;
    }
#undef B200_SP_Q
    count_launch(c, 2);
    B200_CUDA_CHECK(cudaGetLastError());
}
template void sp_generate_q<float>(ctx &);
template void sp_generate_q<double>(ctx &);

// out[0..m) = K d over the reduced system for the rows owned by this rank (zero elsewhere when world > 1); un-centred:
// the caller adds (QA_cost - q_i) sum(d) - q.d + d_i / C (k_abar_finish, centre mode 2).
template <typename T>
void sp_kernel_matvec(ctx &c, const T *d, T *out) {
    sparse_state &st = state_of(c);
    const int64_t m = c.m();
    int64_t lo = 0, hi = m;
    partition_rows(m, c.rank, c.world, &lo, &hi);
    const int64_t nbW = (int64_t) st.nb * st.W;
    T *v = st.v.as<T>();
    if (c.kernel == B200SVM_LINEAR) {
        weights<T>(c, d, static_cast<const T *>(nullptr), m, nbW, v);
        xt_v_blocks<T>(c, st, v, st.wvec.as<T>());
        if (c.world > 1) { B200_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t) m * sizeof(T), c.stream)); }
        row_dot<T>(c, st, view_rows(st, lo, hi), st.wvec.as<T>(), T(0), out + lo);
        return;
    }
    const bool rbf = c.kernel == B200SVM_RBF;
    weights<T>(c, d, rbf ? st.u.as<T>() : static_cast<const T *>(nullptr), m, nbW, v);
    B200_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t) m * sizeof(T), c.stream));
    launch_kmv<T>(c, st, view_rows(st, 0, st.n), lo, hi, st.a_max_row, st.max_norm2, v, out);
    const sp::kfun kf = make_kfun(c);
    int64_t g = std::max<int64_t>(1, std::min<int64_t>((m + 1023) / 1024, 8 * NUM_SMS));
    sp::k_sp_finish<T><<<(unsigned) g, 256, 0, c.stream>>>(out, st.u.as<T>(), m, lo, hi, rbf ? 1 : 0, (T) kf.c0pow, c.scal.as<T>(), S_TMP0, T(0),
                                                           c.world > 1 ? 1 : 0);
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}
template void sp_kernel_matvec<float>(ctx &, const float *, float *);
template void sp_kernel_matvec<double>(ctx &, const double *, double *);

// w[f] = sum_i alpha_i x_i[f] over all n rows   (src/plssvm/backends/OpenMP/csvm.cpp:173-188)
template <typename T>
void sp_update_w(ctx &c, const T *alpha_dev, T *w_dev) {
    sparse_state &st = state_of(c);
    weights<T>(c, alpha_dev, static_cast<const T *>(nullptr), st.n, (int64_t) st.nb * st.W, st.v.as<T>());
    xt_v_blocks<T>(c, st, st.v.as<T>(), w_dev);
}
template void sp_update_w<float>(ctx &, const float *, float *);
template void sp_update_w<double>(ctx &, const double *, double *);

// decision values for CSR points against the CSR support vectors of this handle (src/plssvm/backends/OpenMP/csvm.cpp:191-244)
template <typename T>
void sp_predict(ctx &c, const T *alpha_dev, T bias, const int64_t *rowptr, const int32_t *colidx, const void *values, int64_t P, int64_t nnz, int fmt,
                T *out_host) {
    sparse_state &st = state_of(c);
    B200_REQUIRE(rowptr != nullptr && rowptr[0] == 0 && rowptr[P] == nnz, "CSR row pointers of the points inconsistent with nnz");
    B200_REQUIRE((fmt == B200SVM_VAL_FP22) == st.fp22, "points must use the same value storage class (FP22 or plain) as the support vectors");
    const int64_t nbW = (int64_t) st.nb * st.W;
    const bool linear = c.kernel == B200SVM_LINEAR, rbf = c.kernel == B200SVM_RBF;
    if (linear) {
        sp_update_w<T>(c, alpha_dev, st.wvec.as<T>());
    } else {
        weights<T>(c, alpha_dev, rbf ? st.u.as<T>() : static_cast<const T *>(nullptr), st.n, nbW, st.v.as<T>());  // sum v -> scal[S_TMP0]
    }
    // stage the points in chunks of at most ~2^27 stored entries
    const int64_t max_entries = int64_t(1) << 27;
    std::vector<int64_t> rp;
    std::vector<int32_t> ci;
    std::vector<T> vals;
    for (int64_t r0 = 0; r0 < P;) {
        int64_t r1 = r0 + 1;
        while (r1 < P && rowptr[r1 + 1] - rowptr[r0] <= max_entries) { ++r1; }
        const int64_t rows = r1 - r0, e0 = rowptr[r0], entries = rowptr[r1] - e0;
        rp.resize((size_t) rows + 1);
        for (int64_t i = 0; i <= rows; ++i) { rp[(size_t) i] = rowptr[r0 + i] - e0; }
        upload(c, st.p_rowptr, rp.data(), rp.size() * sizeof(int64_t));
        if (st.fp22) {
            upload(c, st.p_val, static_cast<const uint32_t *>(values) + e0, (size_t) entries * sizeof(uint32_t));
            st.p_col.reserve(16);
        } else {
            B200_REQUIRE(colidx != nullptr || entries == 0, "null column index array");
            ci.assign(colidx + e0, colidx + e0 + entries);
            for (int32_t f : ci) { B200_REQUIRE(f >= 0 && f < st.F, "Number of features per data point must match the number of features per predict point!"); }
            vals.resize((size_t) entries);
            for (int64_t p = 0; p < entries; ++p) {
                vals[(size_t) p] = fmt == B200SVM_VAL_F32 ? (T) static_cast<const float *>(values)[e0 + p] : (T) static_cast<const double *>(values)[e0 + p];
            }
            upload(c, st.p_col, ci.data(), ci.size() * sizeof(int32_t));
            upload(c, st.p_val, vals.data(), vals.size() * sizeof(T));
        }
        st.p_out.reserve((size_t) rows * sizeof(T));
        const sp::csr_view A{ st.p_rowptr.as<int64_t>(), st.p_col.as<int32_t>(), st.p_val.p, rows };
        if (linear) {
            row_dot<T>(c, st, A, st.wvec.as<T>(), bias, st.p_out.as<T>());
        } else {
            st.p_norms.reserve((size_t) rows * sizeof(T));
            st.p_u.reserve((size_t) rows * sizeof(T));
            norms_and_u<T>(c, st, A, st.p_norms.as<T>(), st.p_u.as<T>());
            B200_CUDA_CHECK(cudaMemsetAsync(st.p_out.p, 0, (size_t) rows * sizeof(T), c.stream));
            int64_t max_row = 0;
            double max_norm2 = st.max_norm2;
            for (int64_t i = 0; i < rows; ++i) {
                max_row = std::max(max_row, rp[(size_t) i + 1] - rp[(size_t) i]);
                double nrm = 0;
                for (int64_t p = rowptr[r0 + i]; p < rowptr[r0 + i + 1]; ++p) {
                    double x;
                    if (st.fp22) {
                        const uint32_t vb = static_cast<const uint32_t *>(values)[p] & 0xFFFFFC00u;
                        float xf;
                        std::memcpy(&xf, &vb, 4);
                        x = xf;
                    } else {
                        x = fmt == B200SVM_VAL_F32 ? (double) static_cast<const float *>(values)[p] : static_cast<const double *>(values)[p];
                    }
                    nrm += x * x;
                }
                max_norm2 = std::max(max_norm2, nrm);
            }
            launch_kmv<T>(c, st, A, 0, rows, max_row, max_norm2, st.v.as<T>(), st.p_out.as<T>());
            const sp::kfun kf = make_kfun(c);
            int64_t g = std::max<int64_t>(1, std::min<int64_t>((rows + 1023) / 1024, 8 * NUM_SMS));
            sp::k_sp_finish<T><<<(unsigned) g, 256, 0, c.stream>>>(st.p_out.as<T>(), st.p_u.as<T>(), rows, 0, rows, rbf ? 1 : 0, (T) kf.c0pow,
                                                                   c.scal.as<T>(), S_TMP0, bias, 0);
            count_launch(c);
            B200_CUDA_CHECK(cudaGetLastError());
        }
        B200_CUDA_CHECK(cudaMemcpyAsync(out_host + r0, st.p_out.p, (size_t) rows * sizeof(T), cudaMemcpyDeviceToHost, c.stream));
        c.stats.d2h_bytes += (int64_t) ((size_t) rows * sizeof(T));
        sync_stream(c);  // staging vectors are reused by the next chunk
        r0 = r1;
    }
}
template void sp_predict<float>(ctx &, const float *, float, const int64_t *, const int32_t *, const void *, int64_t, int64_t, int, float *);
template void sp_predict<double>(ctx &, const double *, double, const int64_t *, const int32_t *, const void *, int64_t, int64_t, int, double *);

}  // namespace b200
