/* TEST INFRASTRUCTURE — CPU restatement ("port") of the reference's LS-SVM hot path.  NOT product code.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 *
 * Parity pin: validated against oracle/_ref (the unmodified reference OpenMP backend compiled from
 * /root/reference) and against the reference's golden predict fixture (tests/golden/, see tests/golden/make_golden.py).
 *
 * Every function cites the reference file:line it restates (paths relative to /root/reference).
 * Compiled strict IEEE (-ffp-contract=off, no fast-math); fma() calls below are the reference's explicit std::fma.
 *
 * The file is instantiated twice (float: suffix _f32, double: suffix _f64) through the REAL/SFX macros.
 */
#ifndef LSSVM_ORACLE_BODY

#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)

#define LSSVM_ORACLE_BODY
#define REAL float
#define SFX _f32
#define FMA fmaf
#define POW powf
#define EXP expf
#include "lssvm_oracle.c"
#undef REAL
#undef SFX
#undef FMA
#undef POW
#undef EXP

#define REAL double
#define SFX _f64
#define FMA fma
#define POW pow
#define EXP exp
#include "lssvm_oracle.c"
#undef REAL
#undef SFX
#undef FMA
#undef POW
#undef EXP

/* ------------------------------------------------------------------------------------------------
 * FP22 value codec.  The reference tree contains no FP22 code: the format is DEFINED
 * by this build and pinned by its own known-answer tests (tests/test_fp22.py) — "parity unpinned" by the
 * reference for this item.
 *   FP22 = IEEE-754 binary32 with the mantissa rounded (round-to-nearest-even) to 13 bits:
 *          1 sign | 8 exponent | 13 mantissa, i.e. the top 22 bits of the binary32 pattern.
 *   Packed sparse entry (32 bit) = (fp22 << 10) | column_delta, column_delta in [0, 1023].
 *   Within a row: col_0 = delta_0, col_k = col_{k-1} + delta_k.  Gaps > 1023 are bridged by padding
 *   entries (fp22 = +0, delta = 1023).
 * ------------------------------------------------------------------------------------------------ */
uint32_t lssvm_fp22_encode(float x) {
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7f800000u) == 0x7f800000u) { /* inf / nan: truncate, keep nan-ness */
        uint32_t t = u >> 10;
        if ((u & 0x007fffffu) != 0 && (t & 0x1fffu) == 0) { t |= 1u; }
        return t;
    }
    /* round to nearest even on the 10 dropped bits (carry may ripple into the exponent: correct behaviour) */
    const uint32_t lsb = (u >> 10) & 1u;
    u += 0x1ffu + lsb;
    return u >> 10;
}
float lssvm_fp22_decode(uint32_t v22) {
    const uint32_t u = v22 << 10;
    float x;
    memcpy(&x, &u, 4);
    return x;
}
float lssvm_fp22_round(float x) { return lssvm_fp22_decode(lssvm_fp22_encode(x)); }

#else /* ======================= templated body ======================= */

#define F(name) CAT(name, SFX)

/* include/plssvm/detail/operators.hpp:113-121 — dot product as a sequential std::fma chain */
REAL F(lssvm_dot)(const REAL *a, const REAL *b, long n) {
    REAL val = 0;
    for (long i = 0; i < n; ++i) { val = FMA(a[i], b[i], val); }
    return val;
}
/* include/plssvm/detail/operators.hpp:139-147 — plain sum */
REAL F(lssvm_sum)(const REAL *a, long n) {
    REAL val = 0;
    for (long i = 0; i < n; ++i) { val += a[i]; }
    return val;
}
/* include/plssvm/detail/operators.hpp:157-167 — squared Euclidean distance, fma chain on the differences */
REAL F(lssvm_sqdist)(const REAL *a, const REAL *b, long n) {
    REAL val = 0;
    for (long i = 0; i < n; ++i) {
        const REAL diff = a[i] - b[i];
        val = FMA(diff, diff, val);
    }
    return val;
}
/* include/plssvm/kernel_types.hpp:63-85 — kernel: 0 linear, 1 polynomial, 2 rbf (enum order kernel_types.hpp:27-34) */
REAL F(lssvm_kernel)(int kernel, const REAL *xi, const REAL *xj, long nf, int degree, REAL gamma, REAL coef0) {
    switch (kernel) {
        case 0: return F(lssvm_dot)(xi, xj, nf);
        case 1: return POW(FMA(gamma, F(lssvm_dot)(xi, xj, nf), coef0), (REAL) degree);
        default: return EXP(-gamma * F(lssvm_sqdist)(xi, xj, nf));
    }
}
/* src/plssvm/backends/OpenMP/q_kernel.cpp:19-50 — q[i] = k(x_i, x_last), i < n-1 */
void F(lssvm_generate_q)(int kernel, const REAL *X, long n, long nf, int degree, REAL gamma, REAL coef0, REAL *q) {
    #pragma omp parallel for
    for (long i = 0; i < n - 1; ++i) { q[i] = F(lssvm_kernel)(kernel, X + i * nf, X + (n - 1) * nf, nf, degree, gamma, coef0); }
}
/* src/plssvm/backends/OpenMP/svm_kernel.cpp:21-47 — ret += add * Abar * d, lower triangle mirrored.
 * Same 64x64 blocking and loop order as the reference (OPENMP_BLOCK_SIZE = 64, include/plssvm/constants.hpp);
 * executed serially, which is what the reference does with OMP_NUM_THREADS=1.  `cost` here is 1/C
 * (the caller passes 1 / cost_, src/plssvm/backends/OpenMP/csvm.cpp:70). */
void F(lssvm_matvec)(int kernel, const REAL *X, long n, long nf, int degree, REAL gamma, REAL coef0,
                     const REAL *q, const REAL *d, REAL QA_cost, REAL cost, REAL add, REAL *ret) {
    const long dept = n - 1;
    const long B = 64;
    for (long i = 0; i < dept; i += B) {
        for (long j = 0; j < dept; j += B) {
            for (long ii = 0; ii < B && ii + i < dept; ++ii) {
                REAL ret_iii = 0;
                for (long jj = 0; jj < B && jj + j < dept; ++jj) {
                    if (ii + i >= jj + j) {
                        const REAL temp = (F(lssvm_kernel)(kernel, X + (ii + i) * nf, X + (jj + j) * nf, nf, degree, gamma, coef0) + QA_cost - q[ii + i] - q[jj + j]) * add;
                        if (ii + i == jj + j) {
                            ret_iii += (temp + cost * add) * d[ii + i];
                        } else {
                            ret_iii += temp * d[jj + j];
                            ret[jj + j] += temp * d[ii + i];
                        }
                    }
                }
                ret[ii + i] += ret_iii;
            }
        }
    }
}
/* Row-range form of the same operator (full rows, no mirroring): ret[i] += add * sum_j Abar_ij d_j for
 * row_lo <= i < row_hi.  Same Abar definition (svm_kernel.cpp:32-34); used (a) OpenMP-parallel for quick parity
 * checks at sizes where the serial mirrored loop is slow and (b) as the bounded sample of bench.py's cpu_baseline. */
void F(lssvm_matvec_rows)(int kernel, const REAL *X, long n, long nf, int degree, REAL gamma, REAL coef0,
                          const REAL *q, const REAL *d, REAL QA_cost, REAL cost, REAL add, long row_lo, long row_hi, REAL *ret) {
    const long dept = n - 1;
    #pragma omp parallel for schedule(dynamic, 4)
    for (long i = row_lo; i < row_hi; ++i) {
        REAL acc = 0;
        for (long j = 0; j < dept; ++j) {
            const REAL temp = (F(lssvm_kernel)(kernel, X + i * nf, X + j * nf, nf, degree, gamma, coef0) + QA_cost - q[i] - q[j]) * add;
            acc += (i == j) ? (temp + cost * add) * d[i] : temp * d[j];
        }
        ret[i] += acc;
    }
}
/* src/plssvm/backends/OpenMP/csvm.cpp:82-170 — CG driver: alpha0 = 1, refresh every 50th iteration (run % 50 == 49),
 * stop when delta <= eps^2 delta0, at most imax iterations.  Returns the number of iterations executed
 * (= std::min(run + 1, imax) of csvm.cpp:162-167).  use_rows != 0 swaps in the parallel row form of the operator. */
long F(lssvm_solver_cg)(int kernel, const REAL *X, long n, long nf, int degree, REAL gamma, REAL coef0, REAL cost_C,
                        const REAL *b, const REAL *q, REAL QA_cost, long imax, REAL eps, int use_rows, REAL *alpha, REAL *delta_out) {
    const long dept = n - 1;
    const REAL inv_cost = 1 / cost_C;
    REAL *r = (REAL *) malloc(sizeof(REAL) * dept);
    REAL *Ad = (REAL *) malloc(sizeof(REAL) * dept);
    REAL *d = (REAL *) malloc(sizeof(REAL) * dept);
#define MATVEC(vec, addv, out)                                                                                                              \
    do {                                                                                                                                    \
        if (use_rows) F(lssvm_matvec_rows)(kernel, X, n, nf, degree, gamma, coef0, q, vec, QA_cost, inv_cost, addv, 0, dept, out);          \
        else F(lssvm_matvec)(kernel, X, n, nf, degree, gamma, coef0, q, vec, QA_cost, inv_cost, addv, out);                                 \
    } while (0)
    for (long i = 0; i < dept; ++i) { alpha[i] = 1; r[i] = b[i]; }           /* :85, :91 */
    MATVEC(alpha, (REAL) -1, r);                                             /* :94 */
    REAL delta = F(lssvm_dot)(r, r, dept);                                   /* :97 */
    const REAL delta0 = delta;                                               /* :98 */
    memcpy(d, r, sizeof(REAL) * dept);                                       /* :101 */
    long run = 0;
    for (; run < imax; ++run) {                                              /* :114 */
        for (long i = 0; i < dept; ++i) { Ad[i] = 0; }                       /* :121 */
        MATVEC(d, (REAL) 1, Ad);                                             /* :122 */
        const REAL alpha_cd = delta / F(lssvm_dot)(d, Ad, dept);             /* :125 */
        for (long i = 0; i < dept; ++i) { alpha[i] += alpha_cd * d[i]; }     /* :128 */
        if (run % 50 == 49) {                                                /* :130 */
            memcpy(r, b, sizeof(REAL) * dept);
            MATVEC(alpha, (REAL) -1, r);                                     /* :135 */
        } else {
            for (long i = 0; i < dept; ++i) { r[i] -= alpha_cd * Ad[i]; }    /* :138 */
        }
        const REAL delta_old = delta;                                        /* :142 */
        delta = F(lssvm_dot)(r, r, dept);                                    /* :143 */
        if (delta <= eps * eps * delta0) { break; }                          /* :145 */
        const REAL beta = delta / delta_old;                                 /* :153 */
        for (long i = 0; i < dept; ++i) { d[i] = beta * d[i] + r[i]; }       /* :155 */
    }
#undef MATVEC
    if (delta_out) { *delta_out = delta; }
    free(r); free(Ad); free(d);
    return run + 1 < imax ? run + 1 : imax;
}
/* src/plssvm/csvm.cpp:207-267 — learn(): q, b = y - y_n, QA_cost = k(x_n,x_n) + 1/C, CG, bias, alpha_n = -sum(alpha).
 * Writes n alphas, returns iterations; *rho_out = -bias (model files store rho, src/plssvm/csvm.cpp:123). */
long F(lssvm_learn)(int kernel, const REAL *X, const REAL *y, long n, long nf, int degree, REAL gamma, REAL coef0, REAL cost_C, REAL eps,
                    int use_rows, REAL *alpha, REAL *rho_out) {
    const long dept = n - 1;
    REAL *q = (REAL *) malloc(sizeof(REAL) * dept);
    REAL *b = (REAL *) malloc(sizeof(REAL) * dept);
    F(lssvm_generate_q)(kernel, X, n, nf, degree, gamma, coef0, q);                                   /* :232-235 */
    for (long i = 0; i < dept; ++i) { b[i] = y[i] - y[n - 1]; }                                      /* :238-239 */
    const REAL QA_cost = F(lssvm_kernel)(kernel, X + (n - 1) * nf, X + (n - 1) * nf, nf, degree, gamma, coef0) + 1 / cost_C; /* :243 */
    const long it = F(lssvm_solver_cg)(kernel, X, n, nf, degree, gamma, coef0, cost_C, b, q, QA_cost, nf, eps, use_rows, alpha, 0); /* :256 imax = num_features */
    const REAL s = F(lssvm_sum)(alpha, dept);
    const REAL bias = y[n - 1] + QA_cost * s - F(lssvm_dot)(q, alpha, dept);                         /* :257 */
    alpha[dept] = -s;                                                                                /* :258 */
    *rho_out = -bias;
    free(q); free(b);
    return it;
}
/* src/plssvm/backends/OpenMP/csvm.cpp:173-188 — w[f] = sum_i alpha_i x_i[f] */
void F(lssvm_update_w)(const REAL *X, const REAL *alpha, long n, long nf, REAL *w) {
    #pragma omp parallel for
    for (long f = 0; f < nf; ++f) {
        REAL t = 0;
        for (long i = 0; i < n; ++i) { t += alpha[i] * X[i * nf + f]; }
        w[f] = t;
    }
}
/* src/plssvm/backends/OpenMP/csvm.cpp:191-244 — decision values: bias + w.p (linear) or bias + sum_i alpha_i k(x_i,p); bias = -rho */
void F(lssvm_predict)(int kernel, const REAL *X, const REAL *alpha, long n, long nf, int degree, REAL gamma, REAL coef0, REAL rho,
                      const REAL *points, long np, REAL *out) {
    const REAL bias = -rho;
    if (kernel == 0) {
        REAL *w = (REAL *) malloc(sizeof(REAL) * nf);
        F(lssvm_update_w)(X, alpha, n, nf, w);
        #pragma omp parallel for
        for (long p = 0; p < np; ++p) { out[p] = bias + F(lssvm_dot)(w, points + p * nf, nf); }      /* :226 */
        free(w);
    } else {
        #pragma omp parallel for
        for (long p = 0; p < np; ++p) {
            REAL t = 0;
            for (long i = 0; i < n; ++i) { t += alpha[i] * F(lssvm_kernel)(kernel, X + i * nf, points + p * nf, nf, degree, gamma, coef0); } /* :231-233 */
            out[p] = bias + t;
        }
    }
}
/* include/plssvm/detail/operators.hpp:174-177 — sign: x > 0 ? +1 : -1  (0 -> -1) */
void F(lssvm_sign)(const REAL *v, long n, REAL *out) {
    for (long i = 0; i < n; ++i) { out[i] = v[i] > (REAL) 0 ? (REAL) 1 : (REAL) -1; }
}

/* ---- sparse (CSR) restatement: same kernel definitions evaluated on sorted CSR rows.  The reference stores data
 * dense; these are cross-checked against the dense functions above on densified inputs. ---- */
static REAL F(csr_dot)(const int32_t *ca, const REAL *va, long na, const int32_t *cb, const REAL *vb, long nb) {
    REAL val = 0;
    long a = 0, b = 0;
    while (a < na && b < nb) {
        if (ca[a] == cb[b]) { val = FMA(va[a], vb[b], val); ++a; ++b; }
        else if (ca[a] < cb[b]) { ++a; }
        else { ++b; }
    }
    return val;
}
static REAL F(csr_sqdist)(const int32_t *ca, const REAL *va, long na, const int32_t *cb, const REAL *vb, long nb) {
    REAL val = 0;
    long a = 0, b = 0;
    while (a < na || b < nb) {
        REAL diff;
        if (b >= nb || (a < na && ca[a] < cb[b])) { diff = va[a]; ++a; }
        else if (a >= na || cb[b] < ca[a]) { diff = -vb[b]; ++b; }
        else { diff = va[a] - vb[b]; ++a; ++b; }
        val = FMA(diff, diff, val);
    }
    return val;
}
REAL F(lssvm_csr_kernel)(int kernel, const int64_t *rp_a, const int32_t *ci_a, const REAL *v_a, long i,
                         const int64_t *rp_b, const int32_t *ci_b, const REAL *v_b, long j, int degree, REAL gamma, REAL coef0) {
    const long sa = rp_a[i], na = rp_a[i + 1] - sa, sb = rp_b[j], nb = rp_b[j + 1] - sb;
    switch (kernel) {
        case 0: return F(csr_dot)(ci_a + sa, v_a + sa, na, ci_b + sb, v_b + sb, nb);
        case 1: return POW(FMA(gamma, F(csr_dot)(ci_a + sa, v_a + sa, na, ci_b + sb, v_b + sb, nb), coef0), (REAL) degree);
        default: return EXP(-gamma * F(csr_sqdist)(ci_a + sa, v_a + sa, na, ci_b + sb, v_b + sb, nb));
    }
}
void F(lssvm_csr_generate_q)(int kernel, const int64_t *rp, const int32_t *ci, const REAL *v, long n, int degree, REAL gamma, REAL coef0, REAL *q) {
    #pragma omp parallel for schedule(dynamic, 16)
    for (long i = 0; i < n - 1; ++i) { q[i] = F(lssvm_csr_kernel)(kernel, rp, ci, v, i, rp, ci, v, n - 1, degree, gamma, coef0); }
}
void F(lssvm_csr_matvec_rows)(int kernel, const int64_t *rp, const int32_t *ci, const REAL *v, long n, int degree, REAL gamma, REAL coef0,
                              const REAL *q, const REAL *d, REAL QA_cost, REAL cost, REAL add, long row_lo, long row_hi, REAL *ret) {
    const long dept = n - 1;
    #pragma omp parallel for schedule(dynamic, 4)
    for (long i = row_lo; i < row_hi; ++i) {
        REAL acc = 0;
        for (long j = 0; j < dept; ++j) {
            const REAL temp = (F(lssvm_csr_kernel)(kernel, rp, ci, v, i, rp, ci, v, j, degree, gamma, coef0) + QA_cost - q[i] - q[j]) * add;
            acc += (i == j) ? (temp + cost * add) * d[i] : temp * d[j];
        }
        ret[i] += acc;
    }
}
void F(lssvm_csr_predict)(int kernel, const int64_t *rp, const int32_t *ci, const REAL *v, const REAL *alpha, long n, int degree, REAL gamma, REAL coef0,
                          REAL rho, const int64_t *prp, const int32_t *pci, const REAL *pv, long np, REAL *out) {
    #pragma omp parallel for schedule(dynamic, 4)
    for (long p = 0; p < np; ++p) {
        REAL t = 0;
        for (long i = 0; i < n; ++i) { t += alpha[i] * F(lssvm_csr_kernel)(kernel, rp, ci, v, i, prp, pci, pv, p, degree, gamma, coef0); }
        out[p] = -rho + t;
    }
}

#undef F
#endif
