// Internal declarations shared by the translation units of libb200svm.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/b200svm.h"
#include "xchg.cuh"

namespace b200 {

// ---------------------------------------------------------------------------------------------------------------------
// error handling: C-ABI functions catch `error` and turn it into a status code + message
// ---------------------------------------------------------------------------------------------------------------------
struct error {
    int code;
    std::string msg;
};

[[noreturn]] inline void fail(int code, const std::string &msg) { throw error{ code, msg }; }

#define B200_CUDA_CHECK(expr)                                                                                                \
    do {                                                                                                                     \
        const cudaError_t err__ = (expr);                                                                                    \
        if (err__ != cudaSuccess) {                                                                                          \
            ::b200::fail(B200SVM_ERR_CUDA, std::string("CUDA error '") + cudaGetErrorString(err__) + "' in " #expr " (" +    \
                                               __FILE__ + ":" + std::to_string(__LINE__) + ")");                             \
        }                                                                                                                    \
    } while (0)

#define B200_REQUIRE(cond, text)                                            \
    do {                                                                    \
        if (!(cond)) { ::b200::fail(B200SVM_ERR_INVALID, std::string(text)); } \
    } while (0)

// ---------------------------------------------------------------------------------------------------------------------
// RAII device buffer (replaces the reference's detail::gpu_device_ptr, include/plssvm/backends/gpu_device_ptr.hpp)
// ---------------------------------------------------------------------------------------------------------------------
struct dbuf {
    void *p{ nullptr };
    size_t bytes{ 0 };
    dbuf() = default;
    dbuf(const dbuf &) = delete;
    dbuf &operator=(const dbuf &) = delete;
    ~dbuf() { release(); }
    void release() {
        if (p != nullptr) { cudaFree(p); }
        p = nullptr;
        bytes = 0;
    }
    // (re)allocate if too small; contents are not preserved
    void reserve(size_t nbytes) {
        if (nbytes <= bytes) { return; }
        release();
        const cudaError_t e = cudaMalloc(&p, nbytes);
        if (e != cudaSuccess) {
            p = nullptr;
            fail(B200SVM_ERR_ALLOC, std::string("cudaMalloc of ") + std::to_string(nbytes) + " bytes failed: " + cudaGetErrorString(e));
        }
        bytes = nbytes;
    }
    template <typename T>
    T *as() const { return static_cast<T *>(p); }
};

constexpr int NUM_SMS = 148;          // B200
constexpr int MAX_PARTIAL_BLOCKS = 1024;  // upper bound of blocks in a fused vector-op + reduction launch

// device scalars of the CG loop (index into ctx.scal, element type = real type of the handle)
enum scal_index : int {
    S_DELTA = 0,
    S_DELTA_OLD,
    S_DELTA0,
    S_DAD,
    S_SUM_D,
    S_Q_D,
    S_ALPHA_CD,
    S_BETA,
    S_SUM_X,  // sum(alpha) for the bias
    S_Q_X,    // q . alpha for the bias
    S_KXX,    // k(x_n, x_n)
    S_TMP0,
    S_TMP1,
    S_DONE,   // CG: non-zero once delta <= eps^2 delta0; later iterations already queued by the host become no-ops
    S_COUNT
};

struct comm_state;  // NCCL communicator wrapper (comm.cu)
struct tile_store;  // symmetric tile store (gram_tc.cu)

struct ctx {
    // configuration
    int dtype{ B200SVM_F64 };
    int kernel{ B200SVM_LINEAR };
    int degree{ 3 };
    double gamma{ 0 }, coef0{ 0 }, cost{ 1 };
    int device{ 0 };
    cudaStream_t own_stream{ nullptr };
    cudaStream_t stream{ nullptr };
    std::string last_error;

    // options
    int64_t opt_matvec_path{ 0 };
    int64_t opt_sync_every{ 1 };
    int64_t opt_tc_mode{ 0 };       // tcgen05 cross terms: 0 / 2 = fp8 chain, 1 = three f16 MMAs
    int64_t opt_sparse_block{ 0 };  // 0 = default
    int64_t opt_sparse_group{ 0 };  // 0 = auto
    int64_t opt_sparse_lanes{ 0 };  // 0 = auto
    int64_t opt_sparse_threads{ 0 };  // 0 = auto
    int64_t opt_sparse_cache{ 0 };    // 0 = auto, 1 = off, 2 = required
    int64_t opt_sparse_cache_mb{ 0 }; // 0 = 85 % of the free memory (both caches)
    int64_t opt_peer_exchange{ 0 };   // multi-GPU: 0 = peer-memory exchange if the peers can be mapped, 1 = NCCL allreduce, 2 = required
    int64_t opt_dense_cache{ 0 };     // tcgen05 path: 0 = auto, 1 = off, 2 = required
    int64_t opt_sparse_tiles{ 0 };    // CSR paths, float: dense-tile store of a densely filled S: 0 = auto (expected fill >= 0.4), 1 = never, 2 = always

    // data
    bool has_data{ false };
    bool is_csr{ false };
    int64_t n{ 0 }, F{ 0 }, ldx{ 0 };  // ldx in elements
    dbuf X;                            // dense n x ldx, row-major, real type
    dbuf norms;                        // n squared norms (real type)
    dbuf Z;                            // linear kernel: rows 0..m-1 of X minus the last row
    bool has_Z{ false };
    // CSR
    int64_t nnz{ 0 };
    int value_format{ 0 };
    dbuf csr_rowptr, csr_colidx, csr_vals;
    dbuf csc_colptr, csc_rowidx, csc_vals;  // transpose (for X^T d)
    // fp16 hi/lo split of X for the tcgen05 path (f32 dense only)
    dbuf Xh_hi, Xh_lo;
    int64_t ldh{ 0 };     // leading dimension (elements) of the split operands
    int64_t n_pad{ 0 };   // rows padded to the tile size
    float split_scale{ 1.f };
    bool has_split{ false };

    // hot-path state
    dbuf q;        // m
    bool has_q{ false };
    double QA_cost{ 0 };
    dbuf scal;     // S_COUNT reals
    dbuf partials; // reduction partials: 4 * MAX_PARTIAL_BLOCKS reals
    dbuf counters; // unsigned int[8] "last block" tickets
    dbuf P;        // matvec partial buffer [splits][m_pad]
    dbuf wvec;     // F reals (linear path: X^T d) + split partials
    dbuf wpart;
    dbuf v_x, v_r, v_d, v_Ad, v_b, v_tmp;  // CG vectors (m reals)
    dbuf pts, pts_norms, pred_out, alpha_dev;  // predict staging
    void *pinned{ nullptr };                   // small pinned host mailbox (scalars / flags)

    struct tile_store *dts{ nullptr };          // dense double polynomial / rbf: cached kernel-matrix tiles (kernels_simt.cu)
    bool dts_valid{ false }, dts_refused{ false };
    int dts_rank{ -1 }, dts_world{ 0 };
    int64_t dts_m{ -1 };
    void *tc_opaque{ nullptr };                 // tcgen05 path state (gram_tc.cu)
    void (*tc_opaque_free)(void *){ nullptr };
    void *sp_opaque{ nullptr };                 // sparse path state (sparse.cu)
    void (*sp_opaque_free)(void *){ nullptr };

    // multi-GPU
    comm_state *comm{ nullptr };
    int rank{ 0 }, world{ 1 };

    // stats
    b200svm_stats stats{};
    static constexpr int EV_RING = 16;
    cudaEvent_t ev[2 * EV_RING]{};  // ring of (start, stop) pairs around the O(m^2) kernel, drained when known to be complete
    int64_t ev_head{ 0 }, ev_tail{ 0 };
    cudaEvent_t sv[2 * EV_RING]{};  // ring of (start, stop) pairs around the O(m) vector kernels of a CG iteration
    int64_t sv_head{ 0 }, sv_tail{ 0 };
    cudaEvent_t cb_ev[2]{};         // around the assembly of a kernel-matrix cache (dense tiles / sparse S)
    bool cb_pending{ false };
    static constexpr int CG_RING = 16;
    cudaEvent_t cg_ev[CG_RING]{};   // one per queued CG iteration (its delta has reached the pinned mailbox)

    size_t esize() const { return dtype == B200SVM_F32 ? 4 : 8; }
    int64_t m() const { return n - 1; }
};

// ---------------------------------------------------------------------------------------------------------------------
// entry points implemented per translation unit (T = float / double)
// ---------------------------------------------------------------------------------------------------------------------
template <typename T> void compute_norms(ctx &c, const T *X, int64_t rows, int64_t ld, int64_t F, T *out);
template <typename T> void generate_q(ctx &c);                                              // q + k(x_n,x_n) -> c.q, scal[S_KXX]
template <typename T> void kernel_matvec(ctx &c, const T *d, T *out_Kd);                    // out = K d (m), poly/rbf O(m^2) or linear X(X^T d)
template <typename T> void apply_abar(ctx &c, const T *d, T add, T *ret);                   // ret += add * Abar d  (device vectors)
template <typename T> int64_t cg_solve(ctx &c, const T *b_dev, int64_t imax, T eps, T *alpha_dev, double *delta_out);
template <typename T> void update_w(ctx &c, const T *alpha_dev, T *w_dev);
// qrow / qcol: q indexed like the rows of A / rows of B (both nullptr = no centring)
template <typename T> int kmv_simt_launch(ctx &c, const T *A, const T *normA, int64_t rowsA, int64_t ldA, const T *B, const T *normB, int64_t rowsB,
                                          int64_t ldB, const T *v, T *P, int64_t ldP, int max_splits, const T *qrow, const T *qcol, T qa);
template <typename T> void center_rows(ctx &c, const T *X, int64_t rows, int64_t ld, const T *xlast, T *Z);
template <typename T> void xt_v(ctx &c, const T *X, int64_t rows, int64_t ld, const T *v, T *w);
template <typename T> void x_w(ctx &c, const T *X, int64_t rows, int64_t ld, const T *w, T bias, T *out);
double probe_smem_atomic_gops(ctx &c);                   // measured shared-memory 32-bit atomic-add rate of the device (1e9 / s)
template <typename T> double probe_fma_tflops(ctx &c);  // measured FMA peak of the device (TFLOP/s, 2 flops per FMA)
template <typename T> void predict_dense(ctx &c, const T *alpha_dev, T bias, const T *pts_dev, int64_t P, int64_t ld, T *out_dev);

// tcgen05 path (gram_tc.cu), f32 dense only
void tc_prepare_split(ctx &c);                                   // build fp16 hi/lo operands + TMA descriptors
bool tc_available(const ctx &c);
void tc_invalidate_cache(ctx &c);  // drop the cached kernel-matrix tiles (option changed); rebuilt on the next matvec
// out = centred K d over this rank's tiles (symmetric tcgen05 tiles or the sweep over the stored tiles).  packed: pk already holds d
// (written by the CG direction kernel).  push != nullptr: the partial-vector reduction stores into the peers instead of `out`.
void tc_kernel_matvec(ctx &c, const float *d, float *out_Kd, bool packed, const xchg_push *push, const float *done);
void *tc_pack_buffer(ctx &c);                                     // float4 pk[n_pad] of the tcgen05 path (nullptr when not prepared)

// symmetric tile store (gram_tc.cu): a densely filled S of a CSR data set (float) or the kernel matrix of dense double data as dense
// lower-triangle tiles, swept like the dense float cache
bool dense_f64_cache_ready(ctx &c);        // dense double polynomial / rbf: tiles assembled (on first use) and valid?
void dense_f64_cache_invalidate(ctx &c);
tile_store *tile_store_create(ctx &c, int64_t m, size_t esize = 4);  // plan for c.rank / c.world (contiguous row-block ranges of equal tile count); esize 4 / 8
void tile_store_destroy(tile_store *ts);
int64_t tile_store_tiles(const tile_store *ts);
size_t tile_store_bytes(const tile_store *ts);             // bytes of the stored tiles of this rank
size_t tile_store_allocated(const tile_store *ts);         // bytes already allocated for them (a re-assembly reuses the buffer)
void tile_store_rows(const tile_store *ts, int64_t *row_lo, int64_t *row_hi, int64_t *first_tile);  // rows of this rank (multiples of 128), storage offset
int64_t tile_store_cum(int64_t I);                         // tiles of the row blocks < I; tile (I, J) is stored at cum(I) - first_tile + J
void *tile_store_alloc(tile_store *ts);                    // float: [tiles][64 column quads][128 rows] float4 (kc_quad_index); double: [tiles][128 column pairs][128 rows] double2
void tile_store_matvec(ctx &c, tile_store *ts, const float *w, int64_t m, float *out, const float *done);
void tile_store_matvec_f64(ctx &c, tile_store *ts, const double *q, const double *w, double qa, int64_t m, double *out, const xchg_push *push, const double *done);

// sparse path (sparse.cu)
void sp_invalidate_cache(ctx &c);  // drop the cached S (options changed); rebuilt on the next matvec
void sp_set_data(ctx &c, const int64_t *rowptr, const int32_t *colidx, const void *values, int64_t n, int64_t F, int64_t nnz, int fmt);
template <typename T> void sp_generate_q(ctx &c);
// out = K d (un-centred).  Returns true: rows [lo, hi) of this rank are in out[lo, hi) (the rest is zeroed when zero_rest);
// false: out holds the RAW partial vector S v over this rank's TILES of a densely filled S (to be summed over the ranks, then turned
// into K d by the O(m) epilogue: sp_finish_params)
template <typename T> bool sp_kernel_matvec(ctx &c, const T *d, T *out, bool zero_rest, const T *done);
void sp_finish_params(ctx &c, const void **u, double *c0pow, int *rbf);  // what the O(m) epilogue needs to turn a raw S v into K d (centre mode 3)
template <typename T> void sp_update_w(ctx &c, const T *alpha_dev, T *w_dev);
template <typename T> void sp_predict(ctx &c, const T *alpha_dev, T bias, const int64_t *rowptr, const int32_t *colidx, const void *values, int64_t P,
                                      int64_t nnz, int fmt, T *out_host);

// multi-GPU (comm.cu)
void comm_destroy(ctx &c);
void comm_allreduce_sum(ctx &c, void *buf, int64_t count);  // in place, real type of the handle
void comm_allgather_inplace(ctx &c, void *buf, size_t bytes_per_rank);  // rank r's share sits at buf + r * bytes_per_rank
// peer-memory exchange (protocol and device helpers: xchg.cuh)
enum xchg_mode : int { XCHG_SUM = 0, XCHG_GATHER = 1 };
struct xchg_plan {
    bool peer{ false };  // false: no peer mapping (NCCL fallback) — sum with comm_allreduce_sum
    xchg_push push;      // producer side
    xchg_view view;      // consumer side
};
xchg_plan comm_exchange_begin(ctx &c, int64_t count, int mode);  // one exchange of `count` reals per slot; collective on first use
bool comm_peer_mode(ctx &c, int64_t count);                      // will exchanges of this size use peer memory?  (collective on first use)
void comm_push(ctx &c, const xchg_plan &plan, const void *P, int64_t lo, int64_t hi, int mode, const void *done);  // generic producer: P[lo, hi) -> every peer
void partition_rows(int64_t m, int rank, int world, int64_t *lo, int64_t *hi);

inline void count_launch(ctx &c, int k = 1) { c.stats.kernel_launches += k; }
void drain_matvec_events(ctx &c, int64_t upto);   // account the matvec event pairs [ev_tail, upto) known to be complete
void sync_stream(ctx &c);           // cudaStreamSynchronize + drain
template <typename T> void sums2(ctx &c, const T *v, int idx_sum, int idx_qv);

}  // namespace b200
