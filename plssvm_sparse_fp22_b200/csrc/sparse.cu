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

struct blk_view {
    const uint32_t *ptr;     // [nb][F + 1] entry offsets relative to blk[b]
    const int64_t *blk;      // [nb + 1] first entry of each column block
    const uint16_t *j;       // [nnzB] local column (row of X) inside its block   (unused for FP22)
    const void *val;         // T[nnzB] or uint32 packed (fp22 << 10 | j_local), W = 1024
    int64_t F;
    int W;                   // columns of the implicit matrix per block (power of two, <= 4096; 1024 for FP22)
    int nb;
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

// w[f] = sum_j x_jf v_j  via the blocked transpose (thread per feature, fixed order => deterministic); v is zero-extended to nb * W
template <typename T, bool FP22>
__global__ void __launch_bounds__(256) k_sp_xt_v(blk_view B, const T *__restrict__ v, T *__restrict__ w) {
    const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= B.F) { return; }
    T s = 0;
    for (int b = 0; b < B.nb; ++b) {
        const uint32_t *ptr = B.ptr + (int64_t) b * (B.F + 1);
        const int64_t base = B.blk[b];
        const T *vb = v + (int64_t) b * B.W;
        for (int64_t e = base + ptr[f], e1 = base + ptr[f + 1]; e < e1; ++e) {
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
};

template <typename T, int KERNEL>
__device__ __forceinline__ T s_value(T g, const kfun &k) {
    if constexpr (KERNEL == B200SVM_RBF) {
        return expm1_fast((T) k.a * g);
    } else {
        return ipow(fma((T) k.a, g, (T) k.coef0), k.degree) - (T) k.c0pow;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// out[row] += sum_{j in blocks [b0, b1)} S(G_row,j) v_j      rows [row_lo, row_hi) of A, grid-strided over warps
//   per-warp shared memory: acc T[W] | touched bitmap uint32[W/32] | touched list uint16[W]
// ---------------------------------------------------------------------------------------------------------------------
template <typename T, int KERNEL, bool FP22>
__global__ void __launch_bounds__(1024, 1)
    k_sp_kmv(csr_view A, int64_t row_lo, int64_t row_hi, blk_view B, int b0, int b1, const T *__restrict__ v, kfun kf, T *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char sp_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int W = B.W, words = W >> 5;
    const size_t per_warp = (size_t) W * sizeof(T) + (size_t) words * 4 + (size_t) W * 2;
    unsigned char *mine = sp_smem + (size_t) warp * per_warp;
    T *acc = reinterpret_cast<T *>(mine);
    uint32_t *bits = reinterpret_cast<uint32_t *>(mine + (size_t) W * sizeof(T));
    uint16_t *list = reinterpret_cast<uint16_t *>(mine + (size_t) W * sizeof(T) + (size_t) words * 4);
    for (int k = lane; k < W; k += 32) { acc[k] = T(0); }
    for (int k = lane; k < words; k += 32) { bits[k] = 0u; }
    __syncwarp();

    const int64_t gwarp = (int64_t) blockIdx.x * nwarps + warp, twarps = (int64_t) gridDim.x * nwarps;
    for (int64_t row = row_lo + gwarp; row < row_hi; row += twarps) {
        const int64_t p0 = A.rowptr[row], p1 = A.rowptr[row + 1];
        T rowsum = T(0);
        for (int b = b0; b < b1; ++b) {
            const uint32_t *ptr = B.ptr + (int64_t) b * (B.F + 1);
            const int64_t base = B.blk[b];
            // ---- scatter: lane <-> feature of the row, each lane walks that feature's segment of block b
            int colrun = 0;
            for (int64_t p = p0; p < p1; p += 32) {
                int f;
                T x;
                const bool valid = load_entry<T, FP22>(A, p, p1, lane, colrun, f, x);
                if (valid && x != T(0)) {
                    const int64_t e0 = base + __ldg(ptr + f), e1 = base + __ldg(ptr + f + 1);
                    for (int64_t e = e0; e < e1; ++e) {
                        int jl;
                        T xv;
                        if constexpr (FP22) {
                            const uint32_t wd = __ldg(static_cast<const uint32_t *>(B.val) + e);
                            jl = (int) (wd & 1023u);
                            xv = (T) fp22_value(wd);
                        } else {
                            jl = (int) __ldg(B.j + e);
                            xv = __ldg(static_cast<const T *>(B.val) + e);
                        }
                        atomicAdd(&acc[jl], x * xv);                 // lanes of one warp may meet on a slot
                        atomicOr(&bits[jl >> 5], 1u << (jl & 31));
                    }
                }
            }
            __syncwarp();
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
            const T *vb = v + (int64_t) b * W;
            for (int t = lane; t < cnt; t += 32) {
                const int jl = list[t];
                const T g = acc[jl];
                acc[jl] = T(0);
                rowsum = fma(s_value<T, KERNEL>(g, kf), __ldg(vb + jl), rowsum);
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
    // column-blocked transpose (B side)
    dbuf b_ptr, b_blk, b_j, b_val;
    int W{ 0 }, nb{ 0 };
    int64_t b_entries{ 0 };
    // per-row data and work vectors
    dbuf norms, u, v, xl, wvec;
    // launch geometry of k_sp_kmv
    int warps{ 0 };
    size_t smem{ 0 };
    int blocks_per_launch{ 1 };
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
sp::blk_view view_blocks(const sparse_state &st) {
    return sp::blk_view{ st.b_ptr.as<uint32_t>(), st.b_blk.as<int64_t>(), st.b_j.as<uint16_t>(), st.b_val.p, st.F, st.W, st.nb };
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

// Column-blocked transpose: for block b (rows [bW, (b+1)W) of X) and feature f the entries (j_local, x_jf), j ascending.
template <typename T>
void build_blocks(ctx &c, sparse_state &st, const host_csr<T> &h) {
    const int64_t n = st.n, F = st.F;
    const int W = st.W, nb = st.nb;
    std::vector<uint32_t> ptr((size_t) nb * (size_t) (F + 1), 0u);
    std::vector<int64_t> blk((size_t) nb + 1, 0);
    for (int b = 0; b < nb; ++b) {
        const int64_t r0 = (int64_t) b * W, r1 = std::min<int64_t>(n, r0 + W);
        blk[(size_t) b + 1] = blk[(size_t) b] + (h.rowptr[(size_t) r1] - h.rowptr[(size_t) r0]);
    }
    const int64_t total = blk[(size_t) nb];
    std::vector<uint16_t> bj;
    std::vector<T> bval;
    std::vector<uint32_t> bpack;
    if (st.fp22) {
        bpack.resize((size_t) total);
    } else {
        bj.resize((size_t) total);
        bval.resize((size_t) total);
    }
#pragma omp parallel
    {
        std::vector<uint32_t> cursor((size_t) F);
#pragma omp for schedule(dynamic, 1)
        for (int b = 0; b < nb; ++b) {
            uint32_t *pb = ptr.data() + (size_t) b * (size_t) (F + 1);
            const int64_t r0 = (int64_t) b * W, r1 = std::min<int64_t>(n, r0 + W);
            for (int64_t p = h.rowptr[(size_t) r0]; p < h.rowptr[(size_t) r1]; ++p) { ++pb[(size_t) h.col[(size_t) p] + 1]; }
            for (int64_t f = 0; f < F; ++f) { pb[f + 1] += pb[f]; }
            std::copy(pb, pb + F, cursor.begin());
            const int64_t base = blk[(size_t) b];
            for (int64_t j = r0; j < r1; ++j) {
                for (int64_t p = h.rowptr[(size_t) j]; p < h.rowptr[(size_t) j + 1]; ++p) {
                    const size_t e = (size_t) (base + cursor[(size_t) h.col[(size_t) p]]++);
                    if (st.fp22) {
                        bpack[e] = h.bits[(size_t) p] | (uint32_t) (j - r0);
                    } else {
                        bj[e] = (uint16_t) (j - r0);
                        bval[e] = h.val[(size_t) p];
                    }
                }
            }
        }
    }
    st.b_entries = total;
    upload(c, st.b_ptr, ptr.data(), ptr.size() * sizeof(uint32_t));
    upload(c, st.b_blk, blk.data(), blk.size() * sizeof(int64_t));
    if (st.fp22) {
        upload(c, st.b_val, bpack.data(), bpack.size() * sizeof(uint32_t));
    } else {
        upload(c, st.b_j, bj.data(), bj.size() * sizeof(uint16_t));
        upload(c, st.b_val, bval.data(), bval.size() * sizeof(T));
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));  // host vectors go out of scope
}

template <typename T>
void set_kmv_attributes(size_t smem) {
    static size_t configured = 0;
    if (smem <= configured) { return; }
    B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, B200SVM_RBF, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, B200SVM_POLYNOMIAL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    if constexpr (sizeof(T) == 4) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, B200SVM_RBF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        B200_CUDA_CHECK(cudaFuncSetAttribute(sp::k_sp_kmv<T, B200SVM_POLYNOMIAL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
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
    // block width: FP22 packs a 10-bit local column; otherwise the option or a default sized for >= 11 resident warps per SM
    int W = st.fp22 ? 1024 : (c.opt_sparse_block > 0 ? (int) c.opt_sparse_block : 2048);
    B200_REQUIRE(W >= 128 && W <= 4096 && (W & (W - 1)) == 0, "sparse block width must be a power of two in [128, 4096]");
    st.W = W;
    st.nb = (int) ((n + W - 1) / W);
    build_blocks<T>(c, st, h);
    const size_t per_warp = (size_t) W * sizeof(T) + (size_t) W / 8 + (size_t) W * 2;
    st.warps = (int) std::min<size_t>(32, (size_t) (220 * 1024) / per_warp);
    st.smem = per_warp * (size_t) st.warps;
    set_kmv_attributes<T>(st.smem);
    // keep the B slices of one launch L2-resident (~64 MB)
    const double bytes_per_block = (double) st.b_entries / std::max(1, st.nb) * (st.fp22 ? 4.0 : 2.0 + sizeof(T)) + 4.0 * (double) (F + 1);
    st.blocks_per_launch = (int) std::max(1.0, std::min((double) st.nb, 64.0e6 / std::max(1.0, bytes_per_block)));
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

sp::kfun make_kfun(const ctx &c) {
    sp::kfun k{};
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

// out[row] (rows [lo, hi) of A, zero on entry) += sum_j S(G_row,j) v_j over all column blocks
template <typename T>
void launch_kmv(ctx &c, const sparse_state &st, const sp::csr_view &A, int64_t lo, int64_t hi, const T *v, T *out) {
    if (hi <= lo) { return; }
    const sp::blk_view B = view_blocks(st);
    const sp::kfun kf = make_kfun(c);
    const int64_t rows = hi - lo;
    const int grid = (int) std::min<int64_t>(NUM_SMS, (rows + st.warps - 1) / st.warps);
    const int threads = st.warps * 32;
    for (int b0 = 0; b0 < st.nb; b0 += st.blocks_per_launch) {
        const int b1 = std::min(st.nb, b0 + st.blocks_per_launch);
        if (c.kernel == B200SVM_RBF) {
            B200_SP_DISPATCH_FP22(st, T, (sp::k_sp_kmv<T, B200SVM_RBF, true><<<grid, threads, st.smem, c.stream>>>(A, lo, hi, B, b0, b1, v, kf, out)),
                                  (sp::k_sp_kmv<T, B200SVM_RBF, false><<<grid, threads, st.smem, c.stream>>>(A, lo, hi, B, b0, b1, v, kf, out)));
        } else {
            B200_SP_DISPATCH_FP22(st, T,
                                  (sp::k_sp_kmv<T, B200SVM_POLYNOMIAL, true><<<grid, threads, st.smem, c.stream>>>(A, lo, hi, B, b0, b1, v, kf, out)),
                                  (sp::k_sp_kmv<T, B200SVM_POLYNOMIAL, false><<<grid, threads, st.smem, c.stream>>>(A, lo, hi, B, b0, b1, v, kf, out)));
        }
        count_launch(c);
        B200_CUDA_CHECK(cudaGetLastError());
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
    const sp::blk_view B = view_blocks(st);
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
    launch_kmv<T>(c, st, view_rows(st, 0, st.n), lo, hi, v, out);
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
            launch_kmv<T>(c, st, A, 0, rows, st.v.as<T>(), st.p_out.as<T>());
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
