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
// instead of m^2 F.  G is never stored: for every (row i, block b of W columns) a CTA accumulates x_if * x_jf over the features of
// row i into a shared-memory accumulator, driven by a column-blocked transpose of X (per block, per feature: the list of
// (j_local, x_jf)), compacts the touched slots from a bitmap and evaluates the kernel function only on those.
//   linear      K v = X (X^T v): two O(nnz) passes over the same structures.
// FP22 rows (1|8|13 = top 22 bits of binary32, packed with a 10-bit column delta) are decoded in registers.
#include "device_utils.cuh"

#include <cub/device/device_radix_sort.cuh>

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
//   block b  = rows [b W, (b+1) W) of X  -> one pass of a CTA's shared-memory accumulator
//   group g  = G consecutive blocks      -> one kernel launch (its slice of the transpose stays L2-resident)
// The per-(feature, block) pointers give every segment a known extent: a group of lanes reads it as one contiguous run.
struct csc_view {
    const uint32_t *ptrT;    // [F][nb + 1] index of the first entry of column f inside block b; ptrT[f][nb] = end of the column
    const uint16_t *j;       // [entries] row of X relative to its block
    const void *val;         // T[entries] (FP22 storage: the FP22-exact binary32 value; the row index needs more than 10 bits)
    int64_t F;
    int W;                   // rows of X per block (power of two, 128..32768)
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
// stats (optional, predict points): [0] = bits of the largest squared norm as a double (non-negative => ordered like the integer),
//                                   [1] = non-zero if a column index lies outside [0, F) or is not ascending within its row
template <typename T, bool FP22>
__global__ void __launch_bounds__(256) k_sp_norms(csr_view A, T gamma, int rbf, T *__restrict__ norms, T *__restrict__ u, int64_t F,
                                                  unsigned long long *__restrict__ stats) {
    const int64_t row = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= A.rows) { return; }
    const int lane = threadIdx.x & 31;
    const int64_t p0 = A.rowptr[row], p1 = A.rowptr[row + 1];
    T s = 0;
    int colrun = 0;
    unsigned int bad = 0u;
    int prev_last = -1;
    for (int64_t p = p0; p < p1; p += 32) {
        int f;
        T x;
        const bool valid = load_entry<T, FP22>(A, p, p1, lane, colrun, f, x);
        s = fma(x, x, s);
        if (stats != nullptr) {
            // ascending and in range (FP22: the running sum of the deltas is ascending by construction; only the range can fail)
            const int before = __shfl_up_sync(0xffffffffu, f, 1);
            const int prev = lane == 0 ? prev_last : before;
            if (valid && (f < 0 || f >= F)) { bad |= 1u; }
            if (valid && !FP22 && f <= prev) { bad |= 2u; }
            prev_last = __shfl_sync(0xffffffffu, f, 31);
        }
    }
    s = warp_sum(s);
    if (stats != nullptr) {
        bad = __reduce_or_sync(0xffffffffu, bad);
        if (lane == 0) {
            atomicMax(&stats[0], (unsigned long long) __double_as_longlong((double) s));
            if (bad != 0u) { atomicOr(&stats[1], (unsigned long long) bad); }  // bit 0: out of range, bit 1: not ascending
        }
    }
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
template <typename T>
__global__ void __launch_bounds__(256) k_sp_xt_v(csc_view B, const T *__restrict__ v, T *__restrict__ w) {
    const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= B.F) { return; }
    const uint32_t *pt = B.ptrT + f * (B.nb + 1);
    T s = 0;
    for (int b = 0; b < B.nb; ++b) {
        const T *vb = v + (int64_t) b * B.W;
        for (uint32_t e = pt[b], e1 = pt[b + 1]; e < e1; ++e) {
            s = fma(__ldg(static_cast<const T *>(B.val) + e), vb[__ldg(B.j + e)], s);
        }
    }
    w[f] = s;
}

// ---------------------------------------------------------------------------------------------------------------------
// S(g): the part of the kernel value that vanishes when two rows share no feature
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double expm1_taylor(double z) {  // |z| < 2^-7, to z^7: truncation z^7 / 5040 -> relative 2^-42 / 5040 < 1e-16
    double p = fma(z, 1.0 / 5040.0, 1.0 / 720.0);
    p = fma(p, z, 1.0 / 120.0);
    p = fma(p, z, 1.0 / 24.0);
    p = fma(p, z, 1.0 / 6.0);
    p = fma(p, z, 0.5);
    p = fma(p, z, 1.0);
    return z * p;
}
__device__ __forceinline__ float expm1_taylor(float z) {  // |z| < 2^-4, to z^5: truncation relative z^5 / 720 < 2e-9
    float p = fmaf(z, 1.f / 120.f, 1.f / 24.f);
    p = fmaf(p, z, 1.f / 6.f);
    p = fmaf(p, z, 0.5f);
    p = fmaf(p, z, 1.f);
    return z * p;
}
__device__ __forceinline__ bool expm1_small(double z) { return fabs(z) < 0x1p-7; }
__device__ __forceinline__ bool expm1_small(float z) { return fabsf(z) < 0x1p-4f; }
__device__ __forceinline__ double expm1_fast(double z) { return expm1_small(z) ? expm1_taylor(z) : expm1(z); }
__device__ __forceinline__ float expm1_fast(float z) { return expm1_small(z) ? expm1_taylor(z) : expm1f(z); }

struct kfun {
    int degree;
    double a;       // rbf: 2 gamma ; polynomial: gamma
    double coef0;
    double c0pow;   // coef0^degree
    double fx_scale, fx_inv;  // 2^s and 2^-s of the fixed-point Gram accumulation
    int lb;                   // double: bit at which q is split into limbs
};

// kernel constants in the handle's type, hoisted out of the slot loops.  za = a 2^-s folds the fixed-point scale into the factor of
// the Gram value: a (q 2^-s) and (a 2^-s) q are the same real number and 2^-s is a power of two, so both round identically.
template <typename T>
struct kconst {
    T za, coef0, c0pow;
    int degree;
};
template <typename T>
__device__ __forceinline__ kconst<T> make_kconst(const kfun &k) {
    return kconst<T>{ (T) k.a * (T) k.fx_inv, (T) k.coef0, (T) k.c0pow, k.degree };
}

// S of the fixed-point Gram sum q (converted to T): rbf expm1(2 gamma G), polynomial (gamma G + c0)^deg - c0^deg
template <typename T, int KERNEL>
__device__ __forceinline__ T s_of_q(T q, const kconst<T> &k) {
    if constexpr (KERNEL == B200SVM_RBF) {
        return expm1_fast(k.za * q);
    } else {
        const T base = fma(k.za, q, k.coef0);
        T r;
        switch (k.degree) {  // warp-uniform
            case 1: r = base; break;
            case 2: r = base * base; break;
            case 3: r = base * base * base; break;
            case 4: r = (base * base) * (base * base); break;
            default: r = ipow(base, k.degree); break;
        }
        return r - k.c0pow;
    }
}
// four slots at once: one branch for the whole quad on the rbf fast path (every |z| small, which is the rule: z = 2 gamma G)
template <typename T, int KERNEL>
__device__ __forceinline__ void s_of_q4(const T (&q)[4], const kconst<T> &k, T (&sv)[4]) {
    if constexpr (KERNEL == B200SVM_RBF) {
        T z[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { z[u] = k.za * q[u]; }
        if (expm1_small(fmax(fmax(fabs(z[0]), fabs(z[1])), fmax(fabs(z[2]), fabs(z[3]))))) {
#pragma unroll
            for (int u = 0; u < 4; ++u) { sv[u] = expm1_taylor(z[u]); }
        } else {
#pragma unroll
            for (int u = 0; u < 4; ++u) { sv[u] = expm1_fast(z[u]); }
        }
    } else {
#pragma unroll
        for (int u = 0; u < 4; ++u) { sv[u] = s_of_q<T, KERNEL>(q[u], k); }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// out[row] += sum_{j in group g} S(G_row,j) v_j      rows [row_lo, row_hi) of A, one CTA per row (grid-strided)
//
// Shared memory of a CTA: the accumulator of one block of W columns (fixed-point limbs), a touched bitmap, the compacted list of
// touched slots, and the current row: (feature, value) of its stored entries plus the [start, end) of each feature's segment of
// the transpose inside the current block.  Per block:
//   1. owner threads prefetch the end pointer of the NEXT block's segments (latency hidden behind the products);
//   2. groups of GS lanes walk one segment each (contiguous (row_in_block, value) pairs => coalesced), one integer atomic per limb
//      into the shared accumulator and one atomicOr into the bitmap; segments longer than `thr` are taken by the whole CTA;
//   3. COMPACT: the bitmap is compacted (block scan, fixed order) and the kernel function runs on the touched slots only;
//      otherwise all W slots are evaluated (S(0) = 0), cheaper when most slots are hit.
// Every reduction has a fixed order and the accumulation is integer => bit-reproducible.
// ---------------------------------------------------------------------------------------------------------------------
// Fixed-point accumulation of the Gram entries: shared-memory floating-point (and 64-bit integer) atomics are CAS loops on
// sm_100, 32-bit integer atomics are single fire-and-forget instructions.  A product p is scaled by 2^s, s chosen from the largest
// squared row norm (which bounds every Gram entry and every partial sum by Cauchy-Schwarz), and rounded to an integer q.
//   float : one signed 32-bit limb, |sum q| < 2^30 (resolution 2^-30 of the bound; binary32 carries 2^-24 of the value)
//   double: q is split at bit `lb` into an unsigned low limb (32 - lb bits of head-room: up to 2^(32-lb) additions per slot
//           without a carry) and a signed high limb; lb + 30 significant bits (52 for rows of up to 1024 stored entries)
// The sum is exact on the fixed-point grid and independent of the order of the additions.
// The accumulator is addressed through its 32-bit shared-memory address (computed once per kernel) with explicit ld / st / red
// instructions: the address arithmetic of a generic pointer into dynamic shared memory is otherwise redone for every access.
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void red_s32(uint32_t a, int q) { asm volatile("red.shared.add.s32 [%0], %1;" ::"r"(a), "r"(q) : "memory"); }
__device__ __forceinline__ void red_u32(uint32_t a, uint32_t q) { asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(q) : "memory"); }
__device__ __forceinline__ void red_s32_if(bool p, uint32_t a, int q) {
    asm volatile("{\n\t.reg .pred pp;\n\tsetp.ne.b32 pp, %2, 0;\n\t@pp red.shared.add.s32 [%0], %1;\n\t}" ::"r"(a), "r"(q), "r"((int) p) : "memory");
}
__device__ __forceinline__ void red_u32_if(bool p, uint32_t a, uint32_t q) {
    asm volatile("{\n\t.reg .pred pp;\n\tsetp.ne.b32 pp, %2, 0;\n\t@pp red.shared.add.u32 [%0], %1;\n\t}" ::"r"(a), "r"(q), "r"((int) p) : "memory");
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint4 lds_v4(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_v4_zero(uint32_t a) { asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(a), "r"(0u) : "memory"); }

__device__ __forceinline__ void load4(const float *p, float (&v)[4]) {
    const float4 w = __ldg(reinterpret_cast<const float4 *>(p));
    v[0] = w.x, v[1] = w.y, v[2] = w.z, v[3] = w.w;
}
__device__ __forceinline__ void load4(const double *p, double (&v)[4]) {
    const double2 a = __ldg(reinterpret_cast<const double2 *>(p)), b = __ldg(reinterpret_cast<const double2 *>(p) + 1);
    v[0] = a.x, v[1] = a.y, v[2] = b.x, v[3] = b.y;
}

// add(): p is the product ALREADY scaled by 2^s (the row value is scaled once when the row is loaded: a power of two commutes with
// the rounding of the product).  take_q(): the accumulated integer converted to T (NOT scaled back: 2^-s is folded into the kernel
// constant, kconst::za) with the slot cleared for the next visit; take_q4(): four consecutive slots (jl a multiple of 4).
template <typename T>
struct fixp;
template <>
struct fixp<float> {
    static constexpr int LIMBS = 1;
    static __device__ __forceinline__ void add(uint32_t acc_s, int, int jl, float p, int) { red_s32(acc_s + 4u * (uint32_t) jl, __float2int_rn(p)); }
    static __device__ __forceinline__ void add_if(bool on, uint32_t acc_s, int, int jl, float p, int) {
        red_s32_if(on, acc_s + 4u * (uint32_t) jl, __float2int_rn(p));
    }
    static __device__ __forceinline__ float take_q(uint32_t acc_s, int, int jl, int) {
        const uint32_t a = acc_s + 4u * (uint32_t) jl;
        const int q = (int) lds_u32(a);
        sts_u32(a, 0u);
        return (float) q;  // one rounding of q to 24 bits
    }
    static __device__ __forceinline__ void take_q4(uint32_t acc_s, int, int jl, int, float (&q)[4]) {
        const uint32_t a = acc_s + 4u * (uint32_t) jl;
        const uint4 v = lds_v4(a);
        sts_v4_zero(a);
        q[0] = (float) (int) v.x;
        q[1] = (float) (int) v.y;
        q[2] = (float) (int) v.z;
        q[3] = (float) (int) v.w;
    }
};
template <>
struct fixp<double> {
    static constexpr int LIMBS = 2;
    static __device__ __forceinline__ void add(uint32_t acc_s, int W, int jl, double p, int lb) {
        const long long q = __double2ll_rn(p);
        const uint32_t a = acc_s + 4u * (uint32_t) jl;
        red_u32(a, (uint32_t) (q & ((1ll << lb) - 1ll)));
        red_s32(a + 4u * (uint32_t) W, (int) (q >> lb));
    }
    static __device__ __forceinline__ void add_if(bool on, uint32_t acc_s, int W, int jl, double p, int lb) {
        const long long q = __double2ll_rn(p);
        const uint32_t a = acc_s + 4u * (uint32_t) jl;
        red_u32_if(on, a, (uint32_t) (q & ((1ll << lb) - 1ll)));
        red_s32_if(on, a + 4u * (uint32_t) W, (int) (q >> lb));
    }
    static __device__ __forceinline__ double join(uint32_t lo, uint32_t hi, int lb) { return (double) (((long long) (int) hi << lb) + (long long) lo); }
    static __device__ __forceinline__ double take_q(uint32_t acc_s, int W, int jl, int lb) {
        const uint32_t a = acc_s + 4u * (uint32_t) jl, ah = a + 4u * (uint32_t) W;
        const uint32_t lo = lds_u32(a), hi = lds_u32(ah);
        sts_u32(a, 0u);
        sts_u32(ah, 0u);
        return join(lo, hi, lb);
    }
    static __device__ __forceinline__ void take_q4(uint32_t acc_s, int W, int jl, int lb, double (&q)[4]) {
        const uint32_t a = acc_s + 4u * (uint32_t) jl, ah = a + 4u * (uint32_t) W;
        const uint4 lo = lds_v4(a), hi = lds_v4(ah);
        sts_v4_zero(a);
        sts_v4_zero(ah);
        q[0] = join(lo.x, hi.x, lb);
        q[1] = join(lo.y, hi.y, lb);
        q[2] = join(lo.z, hi.z, lb);
        q[3] = join(lo.w, hi.w, lb);
    }
};

// N rounds of 32 consecutive entries of one (feature, block) segment of the transpose: lane l takes entries l, l + 32, ...; x = row
// value 2^s.  GUARD: round u only if 32 u < left (left = entries from this lane on).
template <typename T>
constexpr int SEG_ROUNDS = sizeof(T) == 4 ? 6 : 2;  // registers: a double round costs three (index, value pair)
template <typename T, int N, bool COMPACT, bool GUARD>
__device__ __forceinline__ void seg_rounds(const uint16_t *__restrict__ pj, const T *__restrict__ pv, int left, T x, uint32_t acc_s, int W, int lb,
                                           uint32_t *bits) {
    int ja[N];
    T va[N];
#pragma unroll
    for (int u = 0; u < N; ++u) {
        if (!GUARD || 32 * u < left) {
            ja[u] = (int) __ldg(pj + 32 * u);
            va[u] = __ldg(pv + 32 * u);
        }
    }
#pragma unroll
    for (int u = 0; u < N; ++u) {
        if (!GUARD || 32 * u < left) {
            fixp<T>::add(acc_s, W, ja[u], x * va[u], lb);
            if constexpr (COMPACT) { atomicOr(&bits[ja[u] >> 5], 1u << (ja[u] & 31)); }
        }
    }
}

constexpr int SP_MAX_OWN = 4;  // stored entries of a row per thread: rows of up to 4 * blockDim.x entries

// MODE 0: out[row] += sum_j S_ij v_j (S never stored)
// MODE 1: count the touched slots of every (block, row) visit            -> cache.ptr / cache.cnt [b * rows + (row - row_lo)]
// MODE 2: store (row_in_block, S value) of every touched slot, ascending -> cache.j / cache.val at cache.ptr[...]   
// MODE 3: one-pass assembly: the visit reserves its (padded) entries with one atomicAdd on a cursor (capacity = an upper bound
//         known from the column counts) and stores them at once; ptr / cnt describe where.  Which visit lands where depends on
//         the schedule, what it stores does not: the SpMV result is unchanged.
// MODE 4: dense-tile assembly (float, S densely filled): row i stores S(G_ij) of ALL columns j below the end of its
//         diagonal 256-column strip into the 128 x 256 tiles (I = i / 128, J = j / 256) of the lower block triangle, in the
//         quad-column-major layout the tile sweep reads (kc_quad_index of gram_tc.cu); blocks above the diagonal strip are skipped.
struct cache_view {
    int64_t *ptr;   // [nb * rows + 1] first entry of a visit
    int32_t *cnt;   // [nb * rows] entries of a visit (multiple of 4)
    uint16_t *j;    // [entries]
    void *val;      // T[entries]
    int64_t rows;   // rows of the implicit matrix held by this rank
    unsigned long long *cursor;  // MODE 3: next free entry; cursor[1] != 0: capacity exceeded
    int64_t capacity;
    float *Kc;      // MODE 4: tiles of this rank, storage index of tile (I, J) = cum(I) - tile0 + J
    int64_t tile0;
};

__device__ __forceinline__ int64_t tile_cum(int64_t I) {  // = tile_store_cum (gram_tc.cu)
    const int64_t k = I >> 1;
    return (I & 1) ? (k + 1) * (k + 1) : k * (k + 1);
}

template <typename T, int KERNEL, bool FP22, bool COMPACT, int MODE>
__global__ void __launch_bounds__(1024)
    k_sp_kmv(csr_view A, int64_t row_lo, int64_t row_hi, csc_view B, int g, int rmax, int gs_log2, int thr, const T *__restrict__ v, kfun kf,
             T *__restrict__ out, cache_view cache, const T *__restrict__ done) {
    static_assert(MODE == 0 || MODE == 4 || COMPACT, "the indexed cache passes work from the compacted slot list");
    static_assert(MODE != 4 || (!COMPACT && sizeof(T) == 4), "dense tiles: float, every slot evaluated");
    if (*done != T(0)) { return; }  // CG convergence flag (scal[S_DONE]); uniform over the grid
    extern __shared__ __align__(16) unsigned char sp_smem[];
    __shared__ int wtot[32];
    __shared__ T red[32];
    __shared__ int nlong_s;
    __shared__ long long base_s;
    using FX = fixp<T>;
    const int t = threadIdx.x, TT = blockDim.x, lane = t & 31, warp = t >> 5, nwarps = TT >> 5;
    const int W = B.W, words = W >> 5;
    uint32_t *acc = reinterpret_cast<uint32_t *>(sp_smem);
    uint32_t *bits = acc + (size_t) W * FX::LIMBS;
    T *xs = reinterpret_cast<T *>(bits + words);
    int32_t *feat = reinterpret_cast<int32_t *>(xs + rmax);
    uint32_t *s0 = reinterpret_cast<uint32_t *>(feat + rmax);
    uint32_t *s1 = s0 + rmax;
    uint16_t *longk = reinterpret_cast<uint16_t *>(s1 + rmax);
    uint16_t *list = longk + rmax;
    for (int k = t; k < W * FX::LIMBS + words; k += TT) { acc[k] = 0u; }
    if (t == 0) { nlong_s = 0; }
    const uint32_t acc_s = smem_u32(acc);
    const T fx_scale = (T) kf.fx_scale;
    const kconst<T> kc = make_kconst<T>(kf);
    const int lb = kf.lb;
    const int GS = 1 << gs_log2, NG = TT >> gs_log2, gid = t >> gs_log2, gl = t & (GS - 1);
    const int b0 = g * B.G, b1 = min(B.nb, b0 + B.G);
    const int64_t ldp = (int64_t) B.nb + 1;
    const uint16_t *__restrict__ Bj = B.j;
    const T *__restrict__ Bv = static_cast<const T *>(B.val);
    __syncthreads();

    for (int64_t row = row_lo + blockIdx.x; row < row_hi; row += gridDim.x) {
        const int64_t p0 = A.rowptr[row];
        const int R = (int) (A.rowptr[row + 1] - p0);
        // MODE 4: columns below the end of the row's diagonal strip only; an empty row still has to store its zeros
        const int64_t jmax = MODE == 4 ? 256 * ((row >> 8) + 1) : 0;
        const int b1r = MODE == 4 ? (int) min((int64_t) b1, (jmax + W - 1) / W) : b1;
        if (MODE == 4 ? b0 >= b1r : R == 0) { continue; }  // block-uniform
        // ---- the row: (feature, value 2^s) of every stored entry; entry k is owned by thread k % TT
        if constexpr (FP22) {
            int carry = 0;
            for (int base = 0; base < R; base += TT) {
                const int k = base + t;
                const uint32_t w = k < R ? __ldg(static_cast<const uint32_t *>(A.val) + p0 + k) : 0u;
                const int incl = warp_incl_scan((int) (w & 1023u), lane);
                if (lane == 31) { wtot[warp] = incl; }
                __syncthreads();
                int off = 0, tot = 0;
                for (int w2 = 0; w2 < nwarps; ++w2) {
                    const int c = wtot[w2];
                    off += w2 < warp ? c : 0;
                    tot += c;
                }
                if (k < R) {
                    const T x = (T) fp22_value(w);
                    xs[k] = x * fx_scale;
                    feat[k] = x != T(0) ? carry + off + incl : -1;
                }
                carry += tot;
                __syncthreads();
            }
        } else {
            for (int k = t; k < R; k += TT) {
                const T x = __ldg(static_cast<const T *>(A.val) + p0 + k);
                xs[k] = x * fx_scale;  // the fixed-point scale 2^s rides on the row value
                feat[k] = x != T(0) ? __ldg(A.col + p0 + k) : -1;
            }
        }
        // ---- segments of the first block of the group
#pragma unroll
        for (int q = 0; q < SP_MAX_OWN; ++q) {
            const int k = t + q * TT;
            if (k < R) {
                const int f = feat[k];
                uint32_t a = 0u, e = 0u;
                if (f >= 0) {
                    const uint32_t *pt = B.ptrT + (int64_t) f * ldp + b0;
                    a = __ldg(pt);
                    e = __ldg(pt + 1);
                }
                s0[k] = a;
                s1[k] = e;
            }
        }
        __syncthreads();

        T rowsum = T(0);
        for (int b = b0; b < b1r; ++b) {
            // ---- 1. prefetch the next block's segment ends; note the long segments of this block
            uint32_t nxt[SP_MAX_OWN];
#pragma unroll
            for (int q = 0; q < SP_MAX_OWN; ++q) {
                const int k = t + q * TT;
                nxt[q] = 0u;
                if (k < R) {
                    const int f = feat[k];
                    const uint32_t e1 = s1[k];
                    if ((int) (e1 - s0[k]) > thr) { longk[atomicAdd(&nlong_s, 1)] = (uint16_t) k; }
                    nxt[q] = e1;
                    if (f >= 0 && b + 1 < b1r) { nxt[q] = __ldg(B.ptrT + (int64_t) f * ldp + b + 2); }
                }
            }
            // ---- 2. products
            if (gs_log2 == 5) {
                // one warp per segment, in chunks of SEG_ROUNDS rounds of 32 entries: every load of a chunk is in flight before its
                // first product, so a typical segment exposes the L2 latency once; full chunks (warp-uniform test) run unpredicated
                for (int k = warp; k < R; k += nwarps) {
                    const uint32_t a0 = s0[k];
                    const int len = (int) (s1[k] - a0);
                    if (len <= 0 || len > thr) { continue; }
                    const T x = xs[k];
                    const uint16_t *pj = Bj + a0 + lane;
                    const T *pv = Bv + a0 + lane;
                    int rest = len;
                    for (; rest >= 32 * SEG_ROUNDS<T>; rest -= 32 * SEG_ROUNDS<T>, pj += 32 * SEG_ROUNDS<T>, pv += 32 * SEG_ROUNDS<T>) {
                        seg_rounds<T, SEG_ROUNDS<T>, COMPACT, false>(pj, pv, 0, x, acc_s, W, lb, bits);
                    }
                    if (rest > 0) { seg_rounds<T, SEG_ROUNDS<T>, COMPACT, true>(pj, pv, rest - lane, x, acc_s, W, lb, bits); }
                }
            } else {
                // GS < 32 lanes per segment (short segments: the visit is bound by its instruction count, not by the load latency):
                // a group of GS lanes walks two segments at once, one entry of each per lane and round
                for (int k0 = gid; k0 < R; k0 += 2 * NG) {
                    uint32_t e[2], e1[2];
                    T x[2];
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int k = k0 + u * NG;
                        e[u] = 0u;
                        e1[u] = 0u;
                        x[u] = T(0);
                        if (k < R) {
                            const uint32_t a0 = s0[k], a1 = s1[k];
                            if ((int) (a1 - a0) <= thr) {
                                e[u] = a0 + (uint32_t) gl;
                                e1[u] = a1;
                                x[u] = xs[k];
                            }
                        }
                    }
                    while ((e[0] < e1[0]) | (e[1] < e1[1])) {
                        int ja[2];
                        T va[2];
#pragma unroll
                        for (int u = 0; u < 2; ++u) {
                            if (e[u] < e1[u]) {
                                ja[u] = (int) __ldg(Bj + e[u]);
                                va[u] = __ldg(Bv + e[u]);
                            }
                        }
#pragma unroll
                        for (int u = 0; u < 2; ++u) {
                            if (e[u] < e1[u]) {
                                FX::add(acc_s, W, ja[u], x[u] * va[u], lb);
                                if constexpr (COMPACT) { atomicOr(&bits[ja[u] >> 5], 1u << (ja[u] & 31)); }
                                e[u] += (uint32_t) GS;
                            }
                        }
                    }
                }
            }
            __syncthreads();
            const int nl = nlong_s;
            if (nl > 0) {  // block-uniform
                for (int i = 0; i < nl; ++i) {
                    const int k = longk[i];
                    const T x = xs[k];
                    for (uint32_t e = s0[k] + (uint32_t) t, e1 = s1[k]; e < e1; e += (uint32_t) TT) {
                        const int ja = (int) __ldg(Bj + e);
                        FX::add(acc_s, W, ja, x * __ldg(Bv + e), lb);
                        if constexpr (COMPACT) { atomicOr(&bits[ja >> 5], 1u << (ja & 31)); }
                    }
                }
                __syncthreads();
                if (t == 0) { nlong_s = 0; }
            }
            // ---- rotate the segment pointers (owner threads; everybody is done reading them)
#pragma unroll
            for (int q = 0; q < SP_MAX_OWN; ++q) {
                const int k = t + q * TT;
                if (k < R) {
                    s0[k] = s1[k];
                    s1[k] = nxt[q];
                }
            }
            // ---- 3. kernel function on the touched slots, times v_j
            const T *vb = v + (int64_t) b * W;
            if constexpr (COMPACT) {
                int cnt = 0;
                for (int w0 = 0; w0 < words; w0 += TT) {
                    const int idx = w0 + t;
                    uint32_t word = idx < words ? bits[idx] : 0u;
                    if (word != 0u) { bits[idx] = 0u; }
                    const int c = __popc(word);
                    const int incl = warp_incl_scan(c, lane);
                    if (lane == 31) { wtot[warp] = incl; }
                    __syncthreads();
                    const int cw = lane < nwarps ? wtot[lane] : 0;  // every warp scans the (at most 32) warp totals
                    const int incl2 = warp_incl_scan(cw, lane);
                    const int off = __shfl_sync(0xffffffffu, incl2 - cw, warp);
                    const int tot = __shfl_sync(0xffffffffu, incl2, 31);
                    int pos = cnt + off + incl - c;
                    while (word != 0u) {
                        const int bit = __ffs((int) word) - 1;
                        list[pos++] = (uint16_t) ((idx << 5) + bit);
                        word &= word - 1u;
                    }
                    cnt += tot;
                    __syncthreads();
                }
                if constexpr (MODE == 0) {
                    for (int i = t; i < cnt; i += TT) {
                        const int jl = list[i];
                        rowsum = fma(s_of_q<T, KERNEL>(FX::take_q(acc_s, W, jl, lb), kc), __ldg(vb + jl), rowsum);
                    }
                } else {
                    const int64_t visit = (int64_t) b * cache.rows + (row - row_lo);
                    const int cnt4 = (cnt + 3) & ~3;  // visits are padded to whole quads (vector loads)
                    if constexpr (MODE == 1) {
                        if (t == 0) {
                            cache.ptr[visit] = cnt4;
                            cache.cnt[visit] = cnt4;
                        }
                        for (int i = t; i < cnt; i += TT) { (void) FX::take_q(acc_s, W, list[i], lb); }
                    } else {
                        int64_t base;
                        if constexpr (MODE == 2) {
                            base = cache.ptr[visit];
                        } else {
                            if (t == 0) {
                                long long b64 = (long long) atomicAdd(cache.cursor, (unsigned long long) cnt4);
                                if (b64 + cnt4 > cache.capacity) {
                                    atomicExch(cache.cursor + 1, 1ull);
                                    b64 = -1;
                                }
                                cache.ptr[visit] = b64 < 0 ? 0 : b64;
                                cache.cnt[visit] = b64 < 0 ? 0 : cnt4;
                                base_s = b64;
                            }
                            __syncthreads();
                            base = base_s;
                            if (base < 0) {  // block-uniform: out of capacity, the host falls back to the two-pass assembly
                                for (int i = t; i < cnt; i += TT) { (void) FX::take_q(acc_s, W, list[i], lb); }
                                __syncthreads();
                                continue;
                            }
                        }
                        T *sval = static_cast<T *>(cache.val);
                        for (int i = t; i < cnt; i += TT) {
                            const int jl = list[i];
                            cache.j[base + i] = (uint16_t) jl;
                            sval[base + i] = s_of_q<T, KERNEL>(FX::take_q(acc_s, W, jl, lb), kc);
                        }
                        for (int i = cnt + t; i < ((cnt + 3) & ~3); i += TT) {
                            cache.j[base + i] = (uint16_t) 0;
                            sval[base + i] = T(0);
                        }
                    }
                }
            } else if constexpr (MODE == 4) {
                const int64_t col_base = (int64_t) b * W;
                const int ncols = (int) min((int64_t) W, jmax - col_base);  // multiple of 256 (W and jmax are)
                const int rr = (int) (row & 127);
                float4 *tiles = reinterpret_cast<float4 *>(cache.Kc) + (tile_cum(row >> 7) - cache.tile0) * (int64_t) (128 * 256 / 4);
                for (int q4 = t; q4 < (ncols >> 2); q4 += TT) {
                    const int jl = q4 << 2;
                    T qv[4], sv[4];
                    FX::take_q4(acc_s, W, jl, lb, qv);
                    s_of_q4<T, KERNEL>(qv, kc, sv);
                    const int64_t col = col_base + jl;
                    // consecutive rows (other CTAs, same time) fill the neighbouring 16 bytes: the sectors merge in L2
                    __stcg(tiles + (col >> 8) * (int64_t) (128 * 256 / 4) + (int64_t) (((int) (col & 255)) >> 2) * 128 + rr, make_float4(sv[0], sv[1], sv[2], sv[3]));
                }
                for (int jl = ncols + 4 * t; jl < W; jl += 4 * TT) {  // slots above the diagonal strip
                    T qv[4];
                    FX::take_q4(acc_s, W, jl, lb, qv);
                }
            } else {
                // two quads of consecutive slots per thread and round; the weights (L1 / L2) of both are requested first
                constexpr int QUADS = sizeof(T) == 4 ? 2 : 1;  // registers
                for (int jl = 4 * t; jl < W; jl += 4 * QUADS * TT) {
                    const int jl2 = jl + 4 * TT;
                    const bool two = QUADS == 2 && jl2 < W;
                    T qv[4], sv[4], va[4], vb4[4];
                    load4(vb + jl, va);
                    if (two) { load4(vb + jl2, vb4); }
                    FX::take_q4(acc_s, W, jl, lb, qv);
                    s_of_q4<T, KERNEL>(qv, kc, sv);
#pragma unroll
                    for (int u = 0; u < 4; ++u) { rowsum = fma(sv[u], va[u], rowsum); }
                    if (two) {
                        FX::take_q4(acc_s, W, jl2, lb, qv);
                        s_of_q4<T, KERNEL>(qv, kc, sv);
#pragma unroll
                        for (int u = 0; u < 4; ++u) { rowsum = fma(sv[u], vb4[u], rowsum); }
                    }
                }
            }
            __syncthreads();
        }
        // ---- row sum in a fixed order
        if constexpr (MODE == 0) {
            rowsum = warp_sum(rowsum);
            if (lane == 0) { red[warp] = rowsum; }
            __syncthreads();
            if (t == 0) {
                T s = red[0];
                for (int w2 = 1; w2 < nwarps; ++w2) { s += red[w2]; }
                out[row] += s;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Cached S: P[b][r] = sum over the stored entries of visit (b, r) of S * v[b W + j]
// One CTA per (block b, chunk of rows): the W weights of the block are staged in shared memory once, then every warp streams the
// entries of its rows (coalesced 2-byte index + value loads; the gather hits shared memory).  HBM-bound: sizeof(T) + 2 bytes per entry.
// ---------------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_sp_cached_mv(cache_view S, int W, int rows_per_cta, const T *__restrict__ v, T *__restrict__ P,
                                                      const T *__restrict__ done) {
    extern __shared__ __align__(16) unsigned char sp_smem[];
    if (*done != T(0)) { return; }  // CG convergence flag (scal[S_DONE]); uniform over the grid
    T *vs = reinterpret_cast<T *>(sp_smem);
    const int b = blockIdx.y;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, nwarps = blockDim.x >> 5;
    const T *vb = v + (int64_t) b * W;
    for (int k = t; k < W; k += blockDim.x) { vs[k] = vb[k]; }
    __syncthreads();
    const int64_t r0 = (int64_t) blockIdx.x * rows_per_cta, r1 = min(S.rows, r0 + rows_per_cta);
    const T *__restrict__ sval = static_cast<const T *>(S.val);
    const uint16_t *__restrict__ sj = S.j;
    for (int64_t r = r0 + warp; r < r1; r += nwarps) {
        const int64_t visit = (int64_t) b * S.rows + r;
        const int64_t q0 = S.ptr[visit] >> 2, q1 = q0 + (S.cnt[visit] >> 2);  // quads of entries (visits are padded to multiples of 4)
        T s0 = T(0), s1 = T(0), s2 = T(0), s3 = T(0);
        const uint2 *__restrict__ jq = reinterpret_cast<const uint2 *>(sj);
        auto quad = [&](const uint2 jj, const T a0, const T a1, const T a2, const T a3) {
            s0 = fma(a0, vs[jj.x & 0xFFFFu], s0);
            s1 = fma(a1, vs[jj.x >> 16], s1);
            s2 = fma(a2, vs[jj.y & 0xFFFFu], s2);
            s3 = fma(a3, vs[jj.y >> 16], s3);
        };
        // four quads in flight per lane (predicated: no tail loop), streamed past L1/L2 residency with ld.global.cs
        for (int64_t q = q0 + lane; q < q1; q += 128) {
            uint2 jj[4];
            T a[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t qq = q + 32 * u;
                if (qq < q1) {
                    jj[u] = __ldcs(jq + qq);
                    if constexpr (sizeof(T) == 4) {
                        const float4 f = __ldcs(reinterpret_cast<const float4 *>(sval) + qq);
                        a[u][0] = f.x;
                        a[u][1] = f.y;
                        a[u][2] = f.z;
                        a[u][3] = f.w;
                    } else {
                        const double2 d0 = __ldcs(reinterpret_cast<const double2 *>(sval) + 2 * qq);
                        const double2 d1 = __ldcs(reinterpret_cast<const double2 *>(sval) + 2 * qq + 1);
                        a[u][0] = d0.x;
                        a[u][1] = d0.y;
                        a[u][2] = d1.x;
                        a[u][3] = d1.y;
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (q + 32 * u < q1) { quad(jj[u], a[u][0], a[u][1], a[u][2], a[u][3]); }
            }
        }
        const T sum = warp_sum((s0 + s1) + (s2 + s3));
        if (lane == 0) { P[visit] = sum; }
    }
}

// out[r] = sum_b P[b][r] in block order (fixed => deterministic), rows [0, rows) -> out[lo + r]
template <typename T>
__global__ void __launch_bounds__(256) k_sp_fold_blocks(const T *__restrict__ P, int nb, int64_t rows, T *__restrict__ out) {
    for (int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (int64_t) gridDim.x * blockDim.x) {
        T s = T(0);
        for (int b = 0; b < nb; ++b) { s += P[(int64_t) b * rows + r]; }
        out[r] = s;
    }
}

// exclusive prefix sum of int64 counts, in place, three passes (4096 elements per CTA)
constexpr int SCAN_CHUNK = 4096;
__global__ void __launch_bounds__(1024) k_scan_chunks(int64_t *x, int64_t n, int64_t *chunk_sum) {
    __shared__ int64_t wsum[32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int64_t base = (int64_t) blockIdx.x * SCAN_CHUNK + (int64_t) t * 4;
    int64_t a[4], tot = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        a[k] = base + k < n ? x[base + k] : 0;
        tot += a[k];
    }
    int64_t incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int64_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) { incl += u; }
    }
    if (lane == 31) { wsum[warp] = incl; }
    __syncthreads();
    int64_t off = 0, all = 0;
    for (int w = 0; w < 32; ++w) {
        off += w < warp ? wsum[w] : 0;
        all += wsum[w];
    }
    int64_t run = off + incl - tot;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (base + k < n) { x[base + k] = run; }
        run += a[k];
    }
    if (t == 0) { chunk_sum[blockIdx.x] = all; }
}
__global__ void __launch_bounds__(32) k_scan_top(int64_t *chunk_sum, int64_t chunks) {
    if (threadIdx.x == 0) {
        int64_t run = 0;
        for (int64_t k = 0; k < chunks; ++k) {
            const int64_t c = chunk_sum[k];
            chunk_sum[k] = run;
            run += c;
        }
    }
}
__global__ void __launch_bounds__(1024) k_scan_add(int64_t *x, int64_t n, const int64_t *chunk_sum) {
    const int64_t base = (int64_t) blockIdx.x * SCAN_CHUNK + (int64_t) threadIdx.x * 4;
    const int64_t off = chunk_sum[blockIdx.x];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (base + k < n) { x[base + k] += off; }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Transpose of the rows on the device (set_data_csr): emit (column key, row) per stored entry, stable radix sort by key (CSR
// order is row-major, so a stable sort by column leaves the rows ascending inside a column), gather, and binary searches for
// the column and (column, block) pointers.  Stored zeros / FP22 padding get the key F and fall off the end.
// ---------------------------------------------------------------------------------------------------------------------
template <typename T, bool FP22>
__global__ void __launch_bounds__(256) k_sp_emit(csr_view A, int64_t F, uint32_t *__restrict__ key, uint32_t *__restrict__ idx,
                                                 int32_t *__restrict__ erow, T *__restrict__ eval) {
    const int64_t row = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= A.rows) { return; }
    const int lane = threadIdx.x & 31;
    const int64_t p0 = A.rowptr[row], p1 = A.rowptr[row + 1];
    int colrun = 0;
    for (int64_t p = p0; p < p1; p += 32) {
        int f;
        T x;
        if (load_entry<T, FP22>(A, p, p1, lane, colrun, f, x)) {
            const int64_t e = p + lane;
            key[e] = (x != T(0) && f >= 0 && f < F) ? (uint32_t) f : (uint32_t) F;
            idx[e] = (uint32_t) e;
            erow[e] = (int32_t) row;
            if (eval != nullptr) { eval[e] = x; }
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_sp_gather(const uint32_t *__restrict__ sidx, const int32_t *__restrict__ erow, const T *__restrict__ eval,
                                                   int64_t count, int W, uint16_t *__restrict__ bj, int32_t *__restrict__ srow, T *__restrict__ bval) {
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < count; e += (int64_t) gridDim.x * blockDim.x) {
        const uint32_t p = sidx[e];
        const int32_t r = erow[p];
        srow[e] = r;
        bj[e] = (uint16_t) (r & (W - 1));
        bval[e] = eval[p];
    }
}

// colptr[f] = first sorted entry whose key is >= f, f = 0..F  (colptr[F] = number of entries kept)
__global__ void __launch_bounds__(256) k_sp_colptr(const uint32_t *__restrict__ skey, int64_t count, int64_t F, uint32_t *__restrict__ colptr) {
    const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (f > F) { return; }
    int64_t lo = 0, hi = count;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if ((int64_t) skey[mid] < f) { lo = mid + 1; } else { hi = mid; }
    }
    colptr[f] = (uint32_t) lo;
}

// ptrT[f][b] = first entry of column f whose row is >= b W, b = 0..nb
__global__ void __launch_bounds__(256) k_sp_block_ptr(const uint32_t *__restrict__ colptr, const int32_t *__restrict__ srow, int64_t F, int nb, int W,
                                                      uint32_t *__restrict__ ptrT) {
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= F * (nb + 1)) { return; }
    const int64_t f = t / (nb + 1);
    const int b = (int) (t - f * (nb + 1));
    int64_t lo = colptr[f], hi = colptr[f + 1];
    const int64_t limit = (int64_t) b * W;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if ((int64_t) srow[mid] < limit) { lo = mid + 1; } else { hi = mid; }
    }
    ptrT[t] = (uint32_t) lo;
}

template <typename S, typename T>
__global__ void __launch_bounds__(256) k_sp_convert(const S *__restrict__ src, int64_t count, T *__restrict__ dst) {
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < count; e += (int64_t) gridDim.x * blockDim.x) { dst[e] = (T) src[e]; }
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

// K v from the S-part:  rbf  out_i = u_i (sum_v + S_i) ; polynomial  out_i = c0^deg sum_v + S_i ;  (+ bias)   [linear in S_i]
// rows [lo, hi) only; the rest of [0, rows) is zeroed when ZERO_REST (multi-GPU partial vectors)
template <typename T>
__global__ void __launch_bounds__(256) k_sp_finish(T *__restrict__ S, const T *__restrict__ u, int64_t rows, int64_t lo, int64_t hi, int rbf, T c0pow,
                                                   const T *__restrict__ scal, int idx_sumv, T bias, int zero_rest, int add_sumv) {
    // add_sumv == 0: a tile-partitioned partial vector of another rank than 0 — the rank-1 term is added once, by rank 0
    const T sumv = add_sumv ? scal[idx_sumv] : T(0);
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
    int threads{ 256 };
    // cached S of the rows [c_lo, c_hi) owned by this rank: visit-major (block, row) CSR with 16-bit row-in-block indices
    dbuf c_ptr, c_cnt, c_cursor, c_j, c_val, c_P, c_chunks;
    double sum_cf2{ 0 };  // sum over the features of (rows containing it)^2 = Gram products of a full matvec
    bool c_valid{ false }, c_refused{ false };
    int64_t c_lo{ 0 }, c_hi{ 0 }, c_entries{ 0 };
    // ... or, when S is densely filled (float): dense lower-triangle tiles swept like the dense kernel-matrix cache
    tile_store *ts{ nullptr };
    bool t_valid{ false }, t_refused{ false };
    int t_rank{ -1 }, t_world{ 0 };
    int64_t t_m{ -1 };
    ~sparse_state() {
        tile_store_destroy(ts);
        if (p_stream != nullptr) { cudaStreamDestroy(p_stream); }
    }
    // predict staging: two sets (upload + norms of chunk i + 1 on p_stream while chunk i computes); p_stats is also the set-up scratch
    struct pstage {
        dbuf rowptr, col, val, norms, u, out, stats;
    };
    pstage ps[2];
    dbuf p_stats;
    cudaStream_t p_stream{ nullptr };
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

#define B200_SP_DISPATCH_FP22(st, T, CALL_TRUE, CALL_FALSE) \
    do {                                                    \
        if ((st).fp22) {                                    \
            if constexpr (sizeof(T) == 4) { CALL_TRUE; }    \
        } else {                                            \
            CALL_FALSE;                                     \
        }                                                   \
    } while (0)

void upload(ctx &c, dbuf &dst, const void *src, size_t bytes) {
    dst.reserve(std::max<size_t>(bytes, 16));
    if (bytes != 0) {
        B200_CUDA_CHECK(cudaMemcpyAsync(dst.p, src, bytes, cudaMemcpyHostToDevice, c.stream));
        c.stats.h2d_bytes += (int64_t) bytes;
    }
}

// shared memory of one k_sp_kmv CTA: accumulator limbs | bitmap | row (value, feature, segment start / end, long list) | slot list
template <typename T>
size_t kmv_smem_bytes(int W, int rmax, bool compact = true) {  // the slot list (last region) is used by the compacting variants only
    return (size_t) W * 4 * sp::fixp<T>::LIMBS + (size_t) W / 8 + (size_t) rmax * (sizeof(T) + 4 + 4 + 4 + 2) + (compact ? (size_t) W * 2 : 0);
}

template <typename T, int KERNEL, bool FP22>
void kmv_set_smem(int smem) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, KERNEL, FP22, true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, KERNEL, FP22, false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, KERNEL, FP22, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, KERNEL, FP22, true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, KERNEL, FP22, true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if constexpr (sizeof(T) == 4) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, KERNEL, FP22, false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    }
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
void norms_and_u(ctx &c, const sparse_state &st, const sp::csr_view &A, T *norms, T *u, unsigned long long *stats = nullptr) {
    if (A.rows == 0) { return; }
    const unsigned grid = (unsigned) ((A.rows + 7) / 8);
    const int rbf = c.kernel == B200SVM_RBF ? 1 : 0;
    if (st.fp22) {
        if constexpr (sizeof(T) == 4) { sp::k_sp_norms<T, true><<<grid, 256, 0, c.stream>>>(A, (T) c.gamma, rbf, norms, u, st.F, stats); }
    } else {
        sp::k_sp_norms<T, false><<<grid, 256, 0, c.stream>>>(A, (T) c.gamma, rbf, norms, u, st.F, stats);
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
    st.c_valid = false;
    st.c_refused = false;
    st.t_valid = false;
    st.t_refused = false;  // (the tile store itself is kept: a data set of the same size reuses its plan and its buffers)
    c.stats.sparse_cache_entries = 0;
    c.stats.dense_cache_bytes = 0;
    const int64_t entries = rowptr[n];
    B200_REQUIRE(entries < (int64_t) 0xFFFF0000ll, "more than 2^32 - 2^16 stored entries are not supported by the sparse path");
    st.a_max_row = 0;
    for (int64_t i = 0; i < n; ++i) {
        B200_REQUIRE(rowptr[i + 1] >= rowptr[i], "CSR row pointers must be non-decreasing");
        st.a_max_row = std::max(st.a_max_row, rowptr[i + 1] - rowptr[i]);
    }
    // rows (A side): the caller's arrays go to the device as they are — FP22 keeps the packed stream (decoded in registers by the
    // kernels); values of the other real type are converted on the device
    st.a_entries = entries;
    upload(c, st.a_rowptr, rowptr, ((size_t) n + 1) * sizeof(int64_t));
    const int64_t cgrid = std::max<int64_t>(1, std::min<int64_t>((entries + 255) / 256, 16 * NUM_SMS));
    dbuf raw;
    if (st.fp22) {
        upload(c, st.a_val, values, (size_t) entries * sizeof(uint32_t));
        st.a_col.reserve(16);
    } else {
        upload(c, st.a_col, colidx, (size_t) entries * sizeof(int32_t));
        const bool same_type = fmt == B200SVM_VAL_F32 ? sizeof(T) == 4 : sizeof(T) == 8;
        if (same_type) {
            upload(c, st.a_val, values, (size_t) entries * sizeof(T));
        } else {
            st.a_val.reserve(std::max<size_t>((size_t) entries * sizeof(T), 16));
            if (fmt == B200SVM_VAL_F32) {
                upload(c, raw, values, (size_t) entries * sizeof(float));
                if (entries > 0) { sp::k_sp_convert<float, T><<<(unsigned) cgrid, 256, 0, c.stream>>>(raw.as<float>(), entries, st.a_val.as<T>()); }
            } else {
                upload(c, raw, values, (size_t) entries * sizeof(double));
                if (entries > 0) { sp::k_sp_convert<double, T><<<(unsigned) cgrid, 256, 0, c.stream>>>(raw.as<double>(), entries, st.a_val.as<T>()); }
            }
            count_launch(c);
        }
    }
    // block width (rows of X per accumulator pass): option or default (8192: 50 KB float / 84 KB double per CTA => 4 / 2 CTAs per SM)
    // Float rows whose S will be densely filled (two rows share nnz^2 / (n^2 F) features at least — Cauchy-Schwarz on the column counts):
    // blocks of 16384 with 512 threads — half the (row, block) visits; measured on the configs[4] geometry (predict, 65536 points x
    // 262144 support vectors, one B200): 820 k points/s against 699 k with 8192 / 256 (profiles/r2_summary.md)
    const double shared_lb = (double) entries * (double) entries / std::max(1.0, (double) n * (double) n * (double) F);
    const bool wide = sizeof(T) == 4 && shared_lb >= 0.5 && n >= 32768;
    const int W = c.opt_sparse_block > 0 ? (int) c.opt_sparse_block : (wide ? 16384 : 8192);
    B200_REQUIRE(W >= 128 && W <= 32768 && (W & (W - 1)) == 0, "sparse block width must be a power of two in [128, 32768]");
    B200_REQUIRE(kmv_smem_bytes<T>(W, 1024) <= (size_t) 200 * 1024, "sparse block width too large for the shared memory of one SM");
    st.W = W;
    st.nb = (int) ((n + W - 1) / W);
    // per-row norms, u = exp(-gamma |x|^2); the same pass validates the columns and finds the largest norm
    st.norms.reserve((size_t) n * sizeof(T));
    st.u.reserve((size_t) n * sizeof(T));
    st.p_stats.reserve(2 * sizeof(unsigned long long));
    B200_CUDA_CHECK(cudaMemsetAsync(st.p_stats.p, 0, 2 * sizeof(unsigned long long), c.stream));
    const sp::csr_view A = view_rows(st, 0, n);
    norms_and_u<T>(c, st, A, st.norms.as<T>(), st.u.as<T>(), st.p_stats.as<unsigned long long>());
    // ---- transpose on the device
    const size_t cap = std::max<size_t>((size_t) entries, 1);
    dbuf key_in, key_out, idx_in, idx_out, erow, srow, eval, temp, colptr;
    key_in.reserve(cap * 4);
    key_out.reserve(cap * 4);
    idx_in.reserve(cap * 4);
    idx_out.reserve(cap * 4);
    erow.reserve(cap * 4);
    srow.reserve(cap * 4);
    const T *values_T = st.a_val.as<T>();
    if (st.fp22) {
        eval.reserve(cap * sizeof(T));
        values_T = eval.as<T>();
    }
    if (n > 0) {
        const unsigned grid = (unsigned) ((n + 7) / 8);
        B200_SP_DISPATCH_FP22(st, T,
                              (sp::k_sp_emit<T, true><<<grid, 256, 0, c.stream>>>(A, F, key_in.as<uint32_t>(), idx_in.as<uint32_t>(), erow.as<int32_t>(), eval.as<T>())),
                              (sp::k_sp_emit<T, false><<<grid, 256, 0, c.stream>>>(A, F, key_in.as<uint32_t>(), idx_in.as<uint32_t>(), erow.as<int32_t>(), nullptr)));
        count_launch(c);
    }
    int key_bits = 1;
    while ((int64_t(1) << key_bits) <= F) { ++key_bits; }
    size_t temp_bytes = 0;
    B200_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, key_in.as<uint32_t>(), key_out.as<uint32_t>(), idx_in.as<uint32_t>(),
                                                    idx_out.as<uint32_t>(), (int64_t) entries, 0, key_bits, c.stream));
    temp.reserve(std::max<size_t>(temp_bytes, 16));
    if (entries > 0) {
        B200_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(temp.p, temp_bytes, key_in.as<uint32_t>(), key_out.as<uint32_t>(), idx_in.as<uint32_t>(),
                                                        idx_out.as<uint32_t>(), (int64_t) entries, 0, key_bits, c.stream));
    }
    colptr.reserve(((size_t) F + 1) * 4);
    sp::k_sp_colptr<<<(unsigned) ((F + 1 + 255) / 256), 256, 0, c.stream>>>(key_out.as<uint32_t>(), entries, F, colptr.as<uint32_t>());
    st.b_j.reserve(cap * sizeof(uint16_t));
    st.b_val.reserve(cap * sizeof(T));
    if (entries > 0) {
        sp::k_sp_gather<T><<<(unsigned) cgrid, 256, 0, c.stream>>>(idx_out.as<uint32_t>(), erow.as<int32_t>(), values_T, entries, W, st.b_j.as<uint16_t>(),
                                                                   srow.as<int32_t>(), st.b_val.as<T>());
    }
    const int64_t np = F * (int64_t) (st.nb + 1);
    st.b_ptrT.reserve(std::max<size_t>((size_t) np * 4, 16));
    sp::k_sp_block_ptr<<<(unsigned) ((np + 255) / 256), 256, 0, c.stream>>>(colptr.as<uint32_t>(), srow.as<int32_t>(), F, st.nb, W, st.b_ptrT.as<uint32_t>());
    count_launch(c, 4);
    B200_CUDA_CHECK(cudaGetLastError());
    // ---- back to the host: validation flags, largest norm, column counts (for the fill estimate and the launch geometry)
    std::vector<uint32_t> cp((size_t) F + 1);
    unsigned long long stats_h[2] = { 0ull, 0ull };
    B200_CUDA_CHECK(cudaMemcpyAsync(cp.data(), colptr.p, cp.size() * 4, cudaMemcpyDeviceToHost, c.stream));
    B200_CUDA_CHECK(cudaMemcpyAsync(stats_h, st.p_stats.p, sizeof(stats_h), cudaMemcpyDeviceToHost, c.stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
    B200_REQUIRE((stats_h[1] & 1ull) == 0ull, st.fp22 ? "FP22 column delta runs past the number of features" : "CSR column index out of range");
    B200_REQUIRE((stats_h[1] & 2ull) == 0ull, "CSR column indices must be sorted and unique within a row");
    std::memcpy(&st.max_norm2, &stats_h[0], sizeof(double));
    double sumsq = 0;
    for (int64_t f = 0; f < F; ++f) {
        const double cf = (double) (cp[(size_t) f + 1] - cp[(size_t) f]);
        sumsq += cf * cf;
    }
    st.b_entries = (int64_t) cp[(size_t) F];
    st.sum_cf2 = sumsq;
    st.fill = n > 0 ? std::min(1.0, sumsq / ((double) n * (double) n)) : 0.0;  // products per visit / W (upper bound of the hit fraction)
    st.avg_seg = (double) st.b_entries / std::max(1.0, (double) F * (double) st.nb);
    // blocks per launch group: keep the group's slice of the transpose L2-resident (~48 MB)
    const double bytes_per_block = (double) st.b_entries / std::max(1, st.nb) * (2.0 + sizeof(T)) + 4.0 * (double) F;
    st.G = (int) std::max(1.0, std::min((double) st.nb, 48.0e6 / std::max(1.0, bytes_per_block)));
    if (c.opt_sparse_group > 0) { st.G = (int) std::min<int64_t>(c.opt_sparse_group, st.nb); }
    st.ng = (st.nb + st.G - 1) / st.G;
    st.threads = sizeof(T) == 4 && W < 16384 ? 256 : 512;
    st.v.reserve((size_t) st.nb * W * sizeof(T));
    st.xl.reserve((size_t) F * sizeof(T));
    st.wvec.reserve((size_t) F * sizeof(T));
    c.n = n;
    c.F = F;
    c.ldx = F;
    c.nnz = st.a_entries;
    c.value_format = fmt;
}

}  // namespace

void sp_invalidate_cache(ctx &c) {
    if (c.sp_opaque != nullptr) {
        sparse_state &st = state_of(c);
        st.c_valid = false;
        st.c_refused = false;
        st.t_valid = false;
        st.t_refused = false;
        c.stats.sparse_cache_entries = 0;
        c.stats.dense_cache_bytes = 0;
    }
}

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


template <typename T>
T pow_like_device(T b, int degree) {
    switch (degree) {
        case 1: return b;
        case 2: return b * b;
        case 3: return b * b * b;
        case 4: return (b * b) * (b * b);
        default: break;
    }
    const bool neg = degree < 0;
    unsigned int u = neg ? (unsigned int) (-degree) : (unsigned int) degree;
    T r = T(1), x = b;
    while (u != 0u) {  // ipow() of device_utils.cuh
        if (u & 1u) { r *= x; }
        x *= x;
        u >>= 1;
    }
    return neg ? T(1) / r : r;
}

// max_norm2 bounds |G_ij| and every partial sum of its products (Cauchy-Schwarz): scale = 2^(bits - ceil(log2 bound));
// max_row bounds the number of additions per slot (head-room of the low limb of the double accumulator)
sp::kfun make_kfun(const ctx &c, double max_norm2 = 1.0, int64_t max_row = 1024) {
    sp::kfun k{};
    {
        int headroom = 1;
        while ((int64_t(1) << headroom) < max_row) { ++headroom; }
        k.lb = 32 - headroom;
        const int bits = c.dtype == B200SVM_F32 ? 30 : k.lb + 30;
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
        // coef0^degree in the handle's type with the multiplication order of s_value(): S(0) is then exactly zero, so evaluating
        // untouched accumulator slots (dense epilogue) adds nothing
        k.c0pow = c.dtype == B200SVM_F32 ? (double) pow_like_device<float>((float) c.coef0, c.degree) : pow_like_device<double>(c.coef0, c.degree);
    }
    return k;
}

// out[row] (rows [lo, hi) of A, zero on entry) += sum_j S(G_row,j) v_j over all column groups; max_row = longest row of A
// mode 0: on-the-fly matvec; 1 / 2: count / fill pass of the S cache, 3: one-pass assembly (`cache` describes the destination)
template <typename T, int KERNEL, bool FP22>
void launch_kmv_variant(ctx &c, const sparse_state &st, const sp::csr_view &A, int64_t lo, int64_t hi, int64_t max_row, double max_norm2,
                        const T *v, T *out, int mode, const sp::cache_view &cache) {
    const sp::csc_view B = view_blocks(st);
    const sp::kfun kf = make_kfun(c, max_norm2, max_row);
    const int64_t rows = hi - lo;
    B200_REQUIRE(max_row <= 4096, "rows with more than 4096 stored entries are not supported by the sparse path: pass the data dense");
    int threads = c.opt_sparse_threads > 0 ? (int) c.opt_sparse_threads : st.threads;
    while ((int64_t) threads * sp::SP_MAX_OWN < max_row) { threads *= 2; }
    const int rmax = (int) std::max<int64_t>(32, (max_row + 31) / 32 * 32);
    const bool compact = (mode != 0 && mode != 4) || (mode == 0 && st.fill < 0.5);
    const size_t smem = kmv_smem_bytes<T>(st.W, rmax, compact);
    B200_REQUIRE(smem <= (size_t) 227 * 1024, "sparse block width and row length exceed the shared memory of one SM");
    set_kmv_attributes<T>(kmv_smem_bytes<T>(st.W, rmax));
    const int per_sm = (int) std::max<size_t>(1, std::min<size_t>((size_t) 2048 / (size_t) threads, (size_t) (227 * 1024) / (smem + 1024)));
    const int grid = (int) std::min<int64_t>((int64_t) NUM_SMS * per_sm, rows);
    // lanes per (feature, block) segment: the power of two next to the mean segment length (option: fixed)
    int gs_log2 = 0;
    if (c.opt_sparse_lanes > 0) {
        while ((2 << gs_log2) <= (int) c.opt_sparse_lanes && gs_log2 < 5) { ++gs_log2; }
    } else {
        while ((double) (1 << gs_log2) < st.avg_seg && gs_log2 < 5) { ++gs_log2; }
    }
    const int thr = std::max(64, 16 << gs_log2);
    const T *done = c.scal.as<T>() + S_DONE;
    for (int g = 0; g < st.ng; ++g) {
        if (mode == 1) {
            sp::k_sp_kmv<T, KERNEL, FP22, true, 1><<<grid, threads, smem, c.stream>>>(A, lo, hi, B, g, rmax, gs_log2, thr, v, kf, out, cache, done);
        } else if (mode == 2) {
            sp::k_sp_kmv<T, KERNEL, FP22, true, 2><<<grid, threads, smem, c.stream>>>(A, lo, hi, B, g, rmax, gs_log2, thr, v, kf, out, cache, done);
        } else if (mode == 3) {
            sp::k_sp_kmv<T, KERNEL, FP22, true, 3><<<grid, threads, smem, c.stream>>>(A, lo, hi, B, g, rmax, gs_log2, thr, v, kf, out, cache, done);
        } else if (mode == 4) {
            if constexpr (sizeof(T) == 4) {
                sp::k_sp_kmv<T, KERNEL, FP22, false, 4><<<grid, threads, smem, c.stream>>>(A, lo, hi, B, g, rmax, gs_log2, thr, v, kf, out, cache, done);
            }
        } else if (compact) {
            sp::k_sp_kmv<T, KERNEL, FP22, true, 0><<<grid, threads, smem, c.stream>>>(A, lo, hi, B, g, rmax, gs_log2, thr, v, kf, out, cache, done);
        } else {
            sp::k_sp_kmv<T, KERNEL, FP22, false, 0><<<grid, threads, smem, c.stream>>>(A, lo, hi, B, g, rmax, gs_log2, thr, v, kf, out, cache, done);
        }
        count_launch(c);
        B200_CUDA_CHECK(cudaGetLastError());
    }
}

template <typename T>
void launch_kmv(ctx &c, const sparse_state &st, const sp::csr_view &A, int64_t lo, int64_t hi, int64_t max_row, double max_norm2, const T *v, T *out,
                int mode = 0, const sp::cache_view &cache = sp::cache_view{}) {
    if (hi <= lo) { return; }
    if (c.kernel == B200SVM_RBF) {
        B200_SP_DISPATCH_FP22(st, T, (launch_kmv_variant<T, B200SVM_RBF, true>(c, st, A, lo, hi, max_row, max_norm2, v, out, mode, cache)),
                              (launch_kmv_variant<T, B200SVM_RBF, false>(c, st, A, lo, hi, max_row, max_norm2, v, out, mode, cache)));
    } else {
        B200_SP_DISPATCH_FP22(st, T, (launch_kmv_variant<T, B200SVM_POLYNOMIAL, true>(c, st, A, lo, hi, max_row, max_norm2, v, out, mode, cache)),
                              (launch_kmv_variant<T, B200SVM_POLYNOMIAL, false>(c, st, A, lo, hi, max_row, max_norm2, v, out, mode, cache)));
    }
}

// Assemble the S cache of rows [lo, hi) (two passes of the on-the-fly kernel: count, prefix sum, fill).  Returns false (and
// remembers the refusal) when the entries do not fit the memory budget; the caller then keeps the on-the-fly path.
template <typename T>
bool build_cache(ctx &c, sparse_state &st, int64_t lo, int64_t hi) {
    const int64_t rows = hi - lo;
    const int64_t visits = (int64_t) st.nb * rows;
    st.c_valid = false;
    if (rows <= 0) { return false; }
    st.c_ptr.reserve((size_t) (visits + 1) * sizeof(int64_t));
    st.c_cnt.reserve((size_t) visits * sizeof(int32_t));
    st.c_cursor.reserve(2 * sizeof(unsigned long long));
    if (c.cb_pending) { sync_stream(c); }
    B200_CUDA_CHECK(cudaEventRecord(c.cb_ev[0], c.stream));
    const sp::csr_view A = view_rows(st, 0, st.n);
    size_t free_b = 0, total_b = 0;
    B200_CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
    free_b += st.c_j.bytes + st.c_val.bytes + st.c_P.bytes;  // buffers of a previous cache are reused
    const double budget = c.opt_sparse_cache_mb > 0 ? (double) c.opt_sparse_cache_mb * 1048576.0 : 0.85 * (double) free_b;
    const auto bytes_for = [&](double entries) { return entries * (double) (sizeof(T) + 2) + (double) visits * sizeof(T) + (double) (64 << 20); };
    sp::cache_view cv{ st.c_ptr.as<int64_t>(), st.c_cnt.as<int32_t>(), nullptr, nullptr, rows, st.c_cursor.as<unsigned long long>(), 0 };
    int64_t entries = -1;

    // ---- one pass: every product touches one slot, so the products of this rank's rows (estimated from the column counts,
    //      15 % slack for the row split) plus the padding bound the entries; a visit never holds more than W
    const char *two_pass_env = std::getenv("B200SVM_CACHE_TWO_PASS");
    const bool allow_one_pass = !(two_pass_env != nullptr && two_pass_env[0] == '1');
    const double bound = std::min(st.sum_cf2 * ((double) rows / (double) std::max<int64_t>(st.n, 1)) * 1.15, (double) visits * (double) st.W) + 4.0 * (double) visits + 1024.0;
    if (allow_one_pass && bytes_for(bound) <= budget && bound < 9.0e18) {
        const int64_t capacity = (int64_t) bound;
        st.c_j.reserve((size_t) capacity * 2);
        st.c_val.reserve((size_t) capacity * sizeof(T));
        cv.j = st.c_j.as<uint16_t>();
        cv.val = st.c_val.p;
        cv.capacity = capacity;
        B200_CUDA_CHECK(cudaMemsetAsync(st.c_ptr.p, 0, (size_t) (visits + 1) * sizeof(int64_t), c.stream));
        B200_CUDA_CHECK(cudaMemsetAsync(st.c_cnt.p, 0, (size_t) visits * sizeof(int32_t), c.stream));
        B200_CUDA_CHECK(cudaMemsetAsync(st.c_cursor.p, 0, 2 * sizeof(unsigned long long), c.stream));
        launch_kmv<T>(c, st, A, lo, hi, st.a_max_row, st.max_norm2, static_cast<const T *>(nullptr), static_cast<T *>(nullptr), 3, cv);
        unsigned long long cur[2] = { 0ull, 0ull };
        B200_CUDA_CHECK(cudaMemcpyAsync(cur, st.c_cursor.p, sizeof(cur), cudaMemcpyDeviceToHost, c.stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
        if (cur[1] == 0ull) { entries = (int64_t) cur[0]; }  // else: the estimate was too small for this row range -> exact sizing below
    }

    // ---- two passes: count, prefix sum, fill (exact allocation)
    if (entries < 0) {
        B200_CUDA_CHECK(cudaMemsetAsync(st.c_ptr.p, 0, (size_t) (visits + 1) * sizeof(int64_t), c.stream));
        B200_CUDA_CHECK(cudaMemsetAsync(st.c_cnt.p, 0, (size_t) visits * sizeof(int32_t), c.stream));
        launch_kmv<T>(c, st, A, lo, hi, st.a_max_row, st.max_norm2, static_cast<const T *>(nullptr), static_cast<T *>(nullptr), 1, cv);
        const int64_t n = visits + 1, chunks = (n + sp::SCAN_CHUNK - 1) / sp::SCAN_CHUNK;
        st.c_chunks.reserve((size_t) chunks * sizeof(int64_t));
        sp::k_scan_chunks<<<(unsigned) chunks, 1024, 0, c.stream>>>(st.c_ptr.as<int64_t>(), n, st.c_chunks.as<int64_t>());
        sp::k_scan_top<<<1, 32, 0, c.stream>>>(st.c_chunks.as<int64_t>(), chunks);
        sp::k_scan_add<<<(unsigned) chunks, 1024, 0, c.stream>>>(st.c_ptr.as<int64_t>(), n, st.c_chunks.as<int64_t>());
        count_launch(c, 3);
        B200_CUDA_CHECK(cudaGetLastError());
        B200_CUDA_CHECK(cudaMemcpyAsync(&entries, st.c_ptr.as<int64_t>() + visits, sizeof(int64_t), cudaMemcpyDeviceToHost, c.stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
        if (bytes_for((double) entries) > budget) {
            B200_REQUIRE(c.opt_sparse_cache != 2, "the cached kernel-matrix entries do not fit the device memory budget");
            st.c_refused = true;
            return false;
        }
        st.c_j.reserve(std::max<size_t>((size_t) entries * 2, 16));
        st.c_val.reserve(std::max<size_t>((size_t) entries * sizeof(T), 16));
        cv.j = st.c_j.as<uint16_t>();
        cv.val = st.c_val.p;
        launch_kmv<T>(c, st, A, lo, hi, st.a_max_row, st.max_norm2, static_cast<const T *>(nullptr), static_cast<T *>(nullptr), 2, cv);
    }
    st.c_P.reserve((size_t) visits * sizeof(T));
    B200_CUDA_CHECK(cudaEventRecord(c.cb_ev[1], c.stream));
    c.cb_pending = true;
    st.c_lo = lo;
    st.c_hi = hi;
    st.c_entries = entries;
    st.c_valid = true;
    c.stats.sparse_cache_entries = entries;
    return true;
}

// Dense-tile store of S (float, densely filled).  Returns false when it does not apply or does not fit.
template <typename T>
bool build_tile_store(ctx &c, sparse_state &st) {
    if constexpr (sizeof(T) != 4) {
        return false;
    } else {
        const int64_t m = c.m();
        // expected fill of S: sum_f c_f^2 / n^2 products per pair, hit with probability 1 - exp(-that); dense tiles cost 2 bytes per
        // ordered pair, the indexed layout 6 bytes per hit: dense wins above a fill of 1 / 3
        const double per_pair = st.n > 0 ? st.sum_cf2 / ((double) st.n * (double) st.n) : 0.0;
        const double fill_est = 1.0 - std::exp(-per_pair);
        const bool want = c.opt_sparse_tiles == 2 || (c.opt_sparse_tiles == 0 && fill_est >= 0.4);
        if (!want || st.W < 256 || m < 1) { return false; }
        if (st.ts == nullptr || st.t_rank != c.rank || st.t_world != c.world || st.t_m != m) {
            tile_store_destroy(st.ts);
            st.ts = nullptr;
            st.ts = tile_store_create(c, m);
            st.t_rank = c.rank;
            st.t_world = c.world;
            st.t_m = m;
        }
        const size_t need = tile_store_bytes(st.ts);
        size_t free_b = 0, total_b = 0;
        B200_CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
        free_b += tile_store_allocated(st.ts);  // the buffer of a previous assembly is reused
        const double budget = c.opt_sparse_cache_mb > 0 ? (double) c.opt_sparse_cache_mb * 1048576.0 : 0.85 * (double) free_b;
        if ((double) need > budget) {
            st.t_refused = true;
            return false;
        }
        float *Kc = static_cast<float *>(tile_store_alloc(st.ts));
        int64_t row_lo = 0, row_hi = 0, tile0 = 0;
        tile_store_rows(st.ts, &row_lo, &row_hi, &tile0);
        if (c.cb_pending) { sync_stream(c); }
        B200_CUDA_CHECK(cudaEventRecord(c.cb_ev[0], c.stream));
        // rows >= m of the padded row blocks are never written: zero the tiles of those row blocks (they end this rank's storage)
        const int64_t I0 = std::max(row_lo / 128, m / 128);
        if (I0 * 128 < row_hi) {
            const int64_t off = (tile_store_cum(I0) - tile0) * (int64_t) (128 * 256);
            B200_CUDA_CHECK(cudaMemsetAsync(Kc + off, 0, need - (size_t) off * sizeof(float), c.stream));
        }
        sp::cache_view cv{};
        cv.rows = 0;
        cv.Kc = Kc;
        cv.tile0 = tile0;
        const sp::csr_view A = view_rows(st, 0, st.n);
        launch_kmv<T>(c, st, A, row_lo, std::min(row_hi, m), st.a_max_row, st.max_norm2, static_cast<const T *>(nullptr), static_cast<T *>(nullptr), 4, cv);
        B200_CUDA_CHECK(cudaEventRecord(c.cb_ev[1], c.stream));
        c.cb_pending = true;
        st.t_valid = true;
        c.stats.dense_cache_bytes = (int64_t) need;
        c.stats.sparse_cache_entries = tile_store_tiles(st.ts) * (int64_t) (128 * 256);
        return true;
    }
}

// out[lo..hi) = S v from the cache
template <typename T>
void cached_matvec(ctx &c, sparse_state &st, const T *v, T *out) {
    const int64_t rows = st.c_hi - st.c_lo;
    const sp::cache_view cv{ st.c_ptr.as<int64_t>(), st.c_cnt.as<int32_t>(), st.c_j.as<uint16_t>(), st.c_val.p, rows, nullptr, 0 };
    const int rows_per_cta = 512;
    static size_t attr_smem[2] = { 0, 0 };
    const size_t smem = (size_t) st.W * sizeof(T);
    if (smem > attr_smem[sizeof(T) == 8]) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_cached_mv<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        attr_smem[sizeof(T) == 8] = smem;
    }
    const dim3 grid((unsigned) ((rows + rows_per_cta - 1) / rows_per_cta), (unsigned) st.nb);
    sp::k_sp_cached_mv<T><<<grid, 256, smem, c.stream>>>(cv, st.W, rows_per_cta, v, st.c_P.as<T>(), c.scal.as<T>() + S_DONE);
    const int64_t g = std::max<int64_t>(1, std::min<int64_t>((rows + 255) / 256, 8 * NUM_SMS));
    sp::k_sp_fold_blocks<T><<<(unsigned) g, 256, 0, c.stream>>>(st.c_P.as<T>(), st.nb, rows, out + st.c_lo);
    count_launch(c, 2);
    B200_CUDA_CHECK(cudaGetLastError());
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
    sp::k_sp_xt_v<T><<<grid, 256, 0, c.stream>>>(B, v, w);
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

// out[lo..hi) = K d over the reduced system for the rows owned by this rank (zero elsewhere when zero_rest); un-centred:
// the caller adds (QA_cost - q_i) sum(d) - q.d + d_i / C (k_abar_finish, centre mode 2).
void sp_finish_params(ctx &c, const void **u, double *c0pow, int *rbf) {
    sparse_state &st = state_of(c);
    *u = st.u.p;
    *c0pow = make_kfun(c).c0pow;
    *rbf = c.kernel == B200SVM_RBF ? 1 : 0;
}

template <typename T>
bool sp_kernel_matvec(ctx &c, const T *d, T *out, bool zero_rest, const T *done) {
    sparse_state &st = state_of(c);
    // (the sparse kernels read scal[S_DONE] themselves; it is cleared by the caller outside the CG loop)
    const float *done_flag = nullptr;
    if constexpr (sizeof(T) == 4) { done_flag = done != nullptr ? done : c.scal.as<float>() + S_DONE; }
    (void) done_flag;
    const int64_t m = c.m();
    int64_t lo = 0, hi = m;
    partition_rows(m, c.rank, c.world, &lo, &hi);
    const int64_t nbW = (int64_t) st.nb * st.W;
    T *v = st.v.as<T>();
    if (c.kernel == B200SVM_LINEAR) {
        weights<T>(c, d, static_cast<const T *>(nullptr), m, nbW, v);
        xt_v_blocks<T>(c, st, v, st.wvec.as<T>());
        if (zero_rest) { B200_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t) m * sizeof(T), c.stream)); }
        row_dot<T>(c, st, view_rows(st, lo, hi), st.wvec.as<T>(), T(0), out + lo);
        return true;
    }
    const bool rbf = c.kernel == B200SVM_RBF;
    weights<T>(c, d, rbf ? st.u.as<T>() : static_cast<const T *>(nullptr), m, nbW, v);
    B200_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t) m * sizeof(T), c.stream));
    // a densely filled S (float) is kept as dense lower-triangle tiles and swept like the dense kernel-matrix cache: the tiles
    // (not the rows) are dealt to the ranks, the result is a full-length partial vector
    if (c.opt_sparse_cache != 1) {
        if (st.t_valid && (st.t_rank != c.rank || st.t_world != c.world)) { st.t_valid = false; }
        if (!st.t_valid && !st.t_refused && !st.c_valid) { build_tile_store<T>(c, st); }
    } else {
        st.t_valid = false;
    }
    if (st.t_valid) {
        if constexpr (sizeof(T) == 4) { tile_store_matvec(c, st.ts, v, m, out, done_flag); }
        return false;  // raw S v of this rank's tiles: K d is formed after the ranks' partial vectors have been combined (cg.cu, centre mode 3)
    }
    // S is cached on first use when it fits the memory budget (then the matvec is an HBM-bound SpMV), else recomputed on the fly
    if (c.opt_sparse_cache != 1 && hi > lo) {
        if (st.c_valid && (st.c_lo != lo || st.c_hi != hi)) { st.c_valid = false; }
        if (!st.c_valid && !st.c_refused) { build_cache<T>(c, st, lo, hi); }
    } else {
        st.c_valid = false;
    }
    if (st.c_valid) {
        cached_matvec<T>(c, st, v, out);
    } else {
        launch_kmv<T>(c, st, view_rows(st, 0, st.n), lo, hi, st.a_max_row, st.max_norm2, v, out);
    }
    const sp::kfun kf = make_kfun(c);
    int64_t g = std::max<int64_t>(1, std::min<int64_t>((m + 1023) / 1024, 8 * NUM_SMS));
    sp::k_sp_finish<T><<<(unsigned) g, 256, 0, c.stream>>>(out, st.u.as<T>(), m, lo, hi, rbf ? 1 : 0, (T) kf.c0pow, c.scal.as<T>(), S_TMP0, T(0),
                                                           zero_rest ? 1 : 0, 1);
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
    return true;
}
template bool sp_kernel_matvec<float>(ctx &, const float *, float *, bool, const float *);
template bool sp_kernel_matvec<double>(ctx &, const double *, double *, bool, const double *);

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
    // The points are staged in chunks of at most ~2^24 stored entries through two sets of buffers: while the kernels of chunk i run
    // on the handle's stream, the host uploads chunk i + 1 on a second stream (values already in the handle's type, and FP22 words,
    // straight from the caller's arrays) and one kernel there computes its norms, u, the largest norm (bound of the fixed-point
    // scale, needed on the host before the launch) and validates the column indices.
    int64_t max_entries = int64_t(1) << 24;
    if (const char *e = std::getenv("B200SVM_PREDICT_CHUNK")) {  // test hook: stored entries per chunk
        const long long v = std::atoll(e);
        if (v > 0) { max_entries = v; }
    }
    const bool same_type = st.fp22 || (fmt == B200SVM_VAL_F32 ? sizeof(T) == 4 : sizeof(T) == 8);
    if (st.p_stream == nullptr) { B200_CUDA_CHECK(cudaStreamCreateWithFlags(&st.p_stream, cudaStreamNonBlocking)); }
    std::vector<int64_t> rp;
    std::vector<T> vals;
    unsigned long long *mailbox = static_cast<unsigned long long *>(c.pinned);
    struct chunk {
        int64_t r0{ 0 }, rows{ 0 }, max_row{ 0 };
        double norm2{ 0 };
    };
    const auto chunk_end = [&](int64_t r0) {
        int64_t r1 = r0 + 1;
        while (r1 < P && rowptr[r1 + 1] - rowptr[r0] <= max_entries) { ++r1; }
        return r1;
    };
    // upload + norms of the chunk starting at r0 into set `s` on the staging stream; returns after both are complete
    const auto prepare = [&](int64_t r0, int s) {
        sparse_state::pstage &ps = st.ps[s];
        chunk ch;
        ch.r0 = r0;
        const int64_t r1 = chunk_end(r0);
        ch.rows = r1 - r0;
        const int64_t rows = ch.rows, e0 = rowptr[r0], entries = rowptr[r1] - e0;
        rp.resize((size_t) rows + 1);
        for (int64_t i = 0; i <= rows; ++i) {
            rp[(size_t) i] = rowptr[r0 + i] - e0;
            if (i > 0) {
                B200_REQUIRE(rp[(size_t) i] >= rp[(size_t) i - 1], "CSR row pointers must be non-decreasing");
                ch.max_row = std::max(ch.max_row, rp[(size_t) i] - rp[(size_t) i - 1]);
            }
        }
        cudaStream_t keep = c.stream;
        c.stream = st.p_stream;  // upload() and norms_and_u() launch on the handle's current stream
        try {
            upload(c, ps.rowptr, rp.data(), rp.size() * sizeof(int64_t));
            if (st.fp22) {
                upload(c, ps.val, static_cast<const uint32_t *>(values) + e0, (size_t) entries * sizeof(uint32_t));
                ps.col.reserve(16);
            } else {
                B200_REQUIRE(colidx != nullptr || entries == 0, "null column index array");
                upload(c, ps.col, colidx + e0, (size_t) entries * sizeof(int32_t));
                if (same_type) {
                    upload(c, ps.val, static_cast<const T *>(values) + e0, (size_t) entries * sizeof(T));
                } else {
                    vals.resize((size_t) entries);
                    for (int64_t q = 0; q < entries; ++q) {
                        vals[(size_t) q] = fmt == B200SVM_VAL_F32 ? (T) static_cast<const float *>(values)[e0 + q] : (T) static_cast<const double *>(values)[e0 + q];
                    }
                    upload(c, ps.val, vals.data(), vals.size() * sizeof(T));
                }
            }
            ps.out.reserve((size_t) rows * sizeof(T));
            ps.norms.reserve((size_t) rows * sizeof(T));
            ps.u.reserve((size_t) rows * sizeof(T));
            ps.stats.reserve(2 * sizeof(unsigned long long));
            const sp::csr_view A{ ps.rowptr.as<int64_t>(), ps.col.as<int32_t>(), ps.val.p, rows };
            B200_CUDA_CHECK(cudaMemsetAsync(ps.stats.p, 0, 2 * sizeof(unsigned long long), c.stream));
            norms_and_u<T>(c, st, A, ps.norms.as<T>(), ps.u.as<T>(), ps.stats.as<unsigned long long>());
            B200_CUDA_CHECK(cudaMemcpyAsync(mailbox + 2 * s, ps.stats.p, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
            B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
        } catch (...) {
            c.stream = keep;
            throw;
        }
        c.stream = keep;
        B200_REQUIRE(mailbox[2 * s + 1] == 0ull, "Number of features per data point must match the number of features per predict point!");
        std::memcpy(&ch.norm2, &mailbox[2 * s], sizeof(double));
        return ch;
    };
    sync_stream(c);  // the weights are complete and the staging sets are free before the second stream touches anything
    chunk cur = prepare(0, 0);
    for (int i = 0;; ++i) {
        sparse_state::pstage &ps = st.ps[i & 1];
        const int64_t rows = cur.rows;
        const sp::csr_view A{ ps.rowptr.as<int64_t>(), ps.col.as<int32_t>(), ps.val.p, rows };
        if (linear) {
            row_dot<T>(c, st, A, st.wvec.as<T>(), bias, ps.out.as<T>());
        } else {
            B200_CUDA_CHECK(cudaMemsetAsync(ps.out.p, 0, (size_t) rows * sizeof(T), c.stream));
            launch_kmv<T>(c, st, A, 0, rows, cur.max_row, std::max(st.max_norm2, cur.norm2), st.v.as<T>(), ps.out.as<T>());
            const sp::kfun kf = make_kfun(c);
            int64_t g = std::max<int64_t>(1, std::min<int64_t>((rows + 1023) / 1024, 8 * NUM_SMS));
            sp::k_sp_finish<T><<<(unsigned) g, 256, 0, c.stream>>>(ps.out.as<T>(), ps.u.as<T>(), rows, 0, rows, rbf ? 1 : 0, (T) kf.c0pow,
                                                                   c.scal.as<T>(), S_TMP0, bias, 0, 1);
            count_launch(c);
            B200_CUDA_CHECK(cudaGetLastError());
        }
        const int64_t next = cur.r0 + rows;
        chunk nxt;
        if (next < P) { nxt = prepare(next, (i + 1) & 1); }  // overlaps the kernels queued above; the other set was drained by the copy below
        B200_CUDA_CHECK(cudaMemcpyAsync(out_host + cur.r0, ps.out.p, (size_t) rows * sizeof(T), cudaMemcpyDeviceToHost, c.stream));
        c.stats.d2h_bytes += (int64_t) ((size_t) rows * sizeof(T));
        sync_stream(c);
        if (next >= P) { break; }
        cur = nxt;
    }
}
template void sp_predict<float>(ctx &, const float *, float, const int64_t *, const int32_t *, const void *, int64_t, int64_t, int, float *);
template void sp_predict<double>(ctx &, const double *, double, const int64_t *, const int32_t *, const void *, int64_t, int64_t, int, double *);

}  // namespace b200
