/*
 * b200svm.h — C ABI of the B200-native LS-SVM hot path (libb200svm.so).
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types, no exceptions.  Every entry point returns
 * B200SVM_OK (0) or an error code; the message is available through b200svm_last_error().  The entry points are what a
 * PLSSVM backend (a subclass of plssvm::csvm<T>, reference include/plssvm/csvm.hpp:33-278) calls from its five protected
 * pure virtuals; each declaration cites the reference interface it replaces (paths relative to the reference tree).
 * There is no CPU fallback: every compute entry point fails with B200SVM_ERR_CUDA when no sm_100 device is usable.
 *
 * Pointers to matrices and vectors of the DENSE entry points (set_data_dense, set_q, matvec, cg_solve, learn, update_w, predict_dense,
 * generate_q) may be host pointers or device pointers of the handle's GPU — unified addressing tells them apart, device-resident
 * inputs / outputs cost a device-to-device copy only (plssvm_sparse_fp22_b200/torch_api.py drives the library from torch CUDA tensors
 * this way).  CSR arrays are host pointers.  `n` = number of data points, `F` = number of features,
 * m = n - 1 = size of the reduced system ("dept" in the reference).  Real type is selected per handle (dtype).
 */
#ifndef B200SVM_H
#define B200SVM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200SVM_OK 0
#define B200SVM_ERR_INVALID 1  /* bad argument / call order                           */
#define B200SVM_ERR_CUDA 2     /* CUDA runtime / driver error, or no usable device    */
#define B200SVM_ERR_NCCL 3     /* NCCL error                                          */
#define B200SVM_ERR_ALLOC 4    /* host or device allocation failure                   */

/* real type of the handle: plssvm::csvm<float> / plssvm::csvm<double> (include/plssvm/csvm.hpp:33-39) */
#define B200SVM_F32 0
#define B200SVM_F64 1
/* plssvm::kernel_type (include/plssvm/kernel_types.hpp:27-34) */
#define B200SVM_LINEAR 0
#define B200SVM_POLYNOMIAL 1
#define B200SVM_RBF 2
/* value storage of CSR data (the reference stores dense vector<vector<real>> only) */
#define B200SVM_VAL_F32 0
#define B200SVM_VAL_F64 1
#define B200SVM_VAL_FP22 2 /* packed 32-bit entries: (fp22 << 10) | column_delta, see DESIGN.md "FP22" */

typedef struct b200svm_ctx *b200svm_handle;

/* Error text of the last failing call on this handle (or of the last failing b200svm_create when h == NULL). */
const char *b200svm_last_error(b200svm_handle h);

/* Library/ABI version (major * 1000 + minor). */
int b200svm_version(void);

/* Replaces the backend constructor: plssvm::csvm<T>::csvm(const parameter<T>&) (src/plssvm/csvm.cpp:39-57) +
 * plssvm::cuda::csvm ctor device selection (src/plssvm/backends/CUDA/csvm.cu:38-80).  `device` = CUDA ordinal.
 * gamma must be > 0 for polynomial / rbf (the caller resolves the 1/F default, src/plssvm/parameter.cpp:150-152). */
int b200svm_create(b200svm_handle *h, int dtype, int kernel, int degree, double gamma, double coef0, double cost, int device);
int b200svm_destroy(b200svm_handle h);

/* Launch all work of this handle on an existing CUDA stream (cudaStream_t passed as void*); NULL = the handle's own
 * stream.  Lets a harness time the kernels with events recorded on the launching stream. */
int b200svm_set_stream(b200svm_handle h, void *cuda_stream);

/* Replaces setup_data_on_device() (include/plssvm/csvm.hpp:188; GPU analogue src/plssvm/backends/gpu_csvm.cpp:129-157):
 * copies the n x F training matrix (row-major, leading dimension ld elements) to HBM and prepares device-side layouts.
 * The last row is the reference's "last point" x_n; rows 0..n-2 span the reduced system. */
int b200svm_set_data_dense(b200svm_handle h, const void *X, int64_t n, int64_t F, int64_t ld);
/* Same for CSR rows (sorted, unique column indices per row).  rowptr int64[n+1], colidx int32[nnz];
 * values: float[nnz] / double[nnz] for VAL_F32 / VAL_F64, or uint32 packed entries for VAL_FP22 (colidx ignored). */
int b200svm_set_data_csr(b200svm_handle h, const int64_t *rowptr, const int32_t *colidx, const void *values, int64_t n, int64_t F,
                         int64_t nnz, int value_format);

/* Replaces generate_q() (include/plssvm/csvm.hpp:193; src/plssvm/backends/OpenMP/q_kernel.cpp:19-50):
 * q[i] = k(x_i, x_n), i < m.  Keeps q resident on the device; q_out (m reals) may be NULL.
 * Also evaluates k(x_n, x_n) (src/plssvm/csvm.cpp:243) — see b200svm_get_kxx_last. */
int b200svm_generate_q(b200svm_handle h, void *q_out);
int b200svm_get_kxx_last(b200svm_handle h, double *kxx);
/* Test hook: overwrite the device-resident q and QA_cost (mirrors mock_csvm::set_QA_cost, tests/backends/OpenMP/mock_openmp_csvm.hpp:38). */
int b200svm_set_q(b200svm_handle h, const void *q, double QA_cost);

/* Replaces run_device_kernel (src/plssvm/backends/OpenMP/csvm.cpp:67-79 -> svm_kernel.cpp:21-47):
 * ret += add * Abar * d, Abar_ij = k(x_i,x_j) + QA_cost - q_i - q_j + delta_ij / C.  d, ret: m reals. */
int b200svm_matvec(b200svm_handle h, const void *d, double add, void *ret_inout);

/* Replaces solver_CG (include/plssvm/csvm.hpp:204; src/plssvm/backends/OpenMP/csvm.cpp:82-170): alpha_0 = 1, residual
 * refresh every 50th iteration, stop when delta <= eps^2 delta_0, at most imax iterations; all vectors stay on the device.
 * b: m reals; q: m reals or NULL (use the device-resident q); alpha_out: m reals.  iters = iterations executed
 * (the value the reference prints, csvm.cpp:162-167); delta = final residual r.r (may be NULL). */
int b200svm_cg_solve(b200svm_handle h, const void *b, const void *q, double QA_cost, int64_t imax, double eps, void *alpha_out,
                     int64_t *iters, double *delta);

/* Whole csvm::learn() numeric path (src/plssvm/csvm.cpp:228-258) on the device: q, b = y - y_n, QA_cost, CG (imax = F),
 * bias, alpha_n = -sum(alpha).  y: n reals (+-1).  alpha_out: n reals; rho_out = -bias. */
int b200svm_learn(b200svm_handle h, const void *y, double eps, void *alpha_out, double *rho_out, int64_t *iters);

/* Replaces update_w() (include/plssvm/csvm.hpp:208; src/plssvm/backends/OpenMP/csvm.cpp:173-188): w[f] = sum_i alpha_i x_i[f]
 * over the data set by set_data_*; alpha: n reals; w_out: F reals. */
int b200svm_update_w(b200svm_handle h, const void *alpha, void *w_out);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The same computation as b200svm_update_w — w[f] = sum_i alpha_i
 * X[i][f], i < n, f < F — without a handle and without any allocation: it calls no cudaMalloc / cudaFree / allocating library routine.
 * The chunk partials of the first kernel live in a file-scope static __device__ array of B200SVM_NOALLOC_SCRATCH_BYTES that every GPU
 * loading libb200svm.so holds until the process exits.  X_dev (n rows, leading dimension ld >= F reals), alpha_dev (n reals) and
 * w_out_dev (F reals) are DEVICE pointers of one GPU; dtype = B200SVM_F32 / B200SVM_F64; cuda_stream = cudaStream_t of that GPU (NULL =
 * the default stream); the call only enqueues two kernels and does not synchronise.  Limit: chunks * ld * sizeof(real) <=
 * B200SVM_NOALLOC_SCRATCH_BYTES with chunks = min(ceil(592 / ceil(ld / 256)), ceil(n / 64)) — which holds for every n when ld <= 524288
 * doubles or ld <= 1048576 floats; larger sizes are rejected with B200SVM_ERR_INVALID (use b200svm_update_w).  All calls share the one array per GPU and it keeps
 * the previous call's contents (every element read is written first, so nothing needs clearing): calls that overlap on the same GPU —
 * different streams, or threads — race; serialise them.  On failure the message is copied into err (err_len bytes incl. the NUL; err may
 * be NULL). */
#define B200SVM_NOALLOC_SCRATCH_BYTES (4u << 20)
int b200svm_update_w_noalloc(int dtype, const void *X_dev, int64_t n, int64_t F, int64_t ld, const void *alpha_dev, void *w_out_dev,
                             void *cuda_stream, char *err, size_t err_len);

/* Replaces predict(points) (include/plssvm/csvm.hpp:214; src/plssvm/backends/OpenMP/csvm.cpp:191-244): decision values
 * out[p] = -rho + sum_i alpha_i k(x_i, p) over the support vectors set by set_data_* (alpha: n reals). */
int b200svm_predict_dense(b200svm_handle h, const void *alpha, double rho, const void *points, int64_t P, int64_t ld, void *out);
int b200svm_predict_csr(b200svm_handle h, const void *alpha, double rho, const int64_t *rowptr, const int32_t *colidx, const void *values,
                        int64_t P, int64_t nnz, int value_format, void *out);

/* Multi-GPU (one process per GPU): rows/tiles of the implicit matrix are partitioned over `world` ranks; the CG loop
 * exchanges with NCCL on the handle's stream.  The unique id (128 bytes) is created on rank 0 and distributed by the
 * caller (e.g. torch.distributed broadcast).  Replaces the reference's host-staged device_reduction
 * (src/plssvm/backends/gpu_csvm.cpp:365-386).  With a communicator the following calls are COLLECTIVE (every rank makes them in the same
 * order, with the same data): b200svm_set_data_dense (each rank uploads 1 / world of the rows and all-gathers the rest over NVLink),
 * b200svm_matvec / b200svm_cg_solve / b200svm_learn (one exchange of the matvec result per application of the implicit matrix through
 * CUDA-IPC mapped peer buffers), b200svm_comm_init and b200svm_destroy (barrier before the exchange buffers are released).  Results are
 * bit-identical on every rank and, on the tile-partitioned paths, bit-identical for 1, 2, 4 and 8 ranks. */
int b200svm_comm_unique_id(void *id128);
int b200svm_comm_init(b200svm_handle h, const void *id128, int rank, int world);
/* Host-only helpers (no device needed): the ownership rules the ranks use, exposed so that the sharding can be tested on
 * CPU-only machines.  Rows [lo, hi) of the m rows of the implicit matrix (full-row paths: SIMT, sparse), and the slice
 * [first, first + count) of the `total` lower-triangle 128 x 256 tiles of an n-point problem (tcgen05 path). */
int b200svm_partition_rows(int64_t m, int rank, int world, int64_t *lo, int64_t *hi);
int b200svm_partition_tiles(int64_t n, int rank, int world, int64_t *first, int64_t *count, int64_t *total);
/* Row-block ranges of the symmetric tile stores (double dense tiles, dense-tile S of CSR data): rows [row_lo, row_hi) (multiples of 128) of
 * rank `rank`, contiguous ranges of equal tile count.  b200svm_canonical_groups(world) = number of canonical tile groups (8 when world
 * divides 8): a rank owns groups / world consecutive ones, so the group boundaries are the same for 1, 2, 4 and 8 ranks — the basis of
 * the bit-identical results across GPU counts. */
int b200svm_partition_tile_rows(int64_t m, int rank, int world, int64_t *row_lo, int64_t *row_hi);
int b200svm_canonical_groups(int world);

/* Tuning / introspection. */
#define B200SVM_OPT_MATVEC_PATH 1 /* 0 = auto, 1 = SIMT full-square, 2 = tcgen05 symmetric tiles (f32 dense only) */
#define B200SVM_OPT_SYNC_EVERY 2  /* CG look-ahead: the host queues up to k iterations beyond the one whose delta it has read (default 1; 0 = synchronous).
                                     A device-side flag freezes alpha / delta at the converged iteration, so results do not depend on k. */
#define B200SVM_OPT_SPARSE_GROUP 4 /* CSR paths: blocks per launch group (0 = sized so that a group's slice of the transpose fits L2) */
#define B200SVM_OPT_SPARSE_LANES 5 /* CSR paths: lanes sharing one (feature, block) segment: 0 auto, or a power of two up to 32 */
#define B200SVM_OPT_SPARSE_THREADS 7 /* CSR paths: threads of the CTA that owns one row (0 auto: 256 float / 512 double; power of two 64..1024) */
#define B200SVM_OPT_TC_MODE 6 /* tcgen05 path, cross terms of the fp16x2 split: 0 = default (fp8 chain), 1 = three f16 MMAs, 2 = fp8 chain; set before set_data */
#define B200SVM_OPT_SPARSE_CACHE 8 /* CSR paths, polynomial / rbf: 0 = cache S = f(X X^T) on first use if it fits (then the matvec is an HBM-bound SpMV), 1 = never (recompute on the fly), 2 = required (error if it does not fit) */
#define B200SVM_OPT_SPARSE_CACHE_MB 9 /* memory budget of the S cache / the dense kernel-matrix cache in MiB (0 = 85 % of the free device memory) */
#define B200SVM_OPT_PEER_EXCHANGE 11 /* multi-GPU sum of the matvec partials: 0 = stores into CUDA-IPC mapped peer buffers over NVLink, summed in rank order by the O(m) epilogue (falls back to NCCL if the peers cannot be mapped), 1 = ncclAllReduce, 2 = peer exchange required; set before comm_init */
#define B200SVM_OPT_DENSE_CACHE 10 /* dense float polynomial / rbf (tcgen05 path): 0 = store the kernel-matrix tiles on first use if they fit (later matvecs are HBM-bound sweeps), 1 = never (recompute every matvec on the tensor cores), 2 = required */
#define B200SVM_OPT_SPARSE_TILES 12 /* CSR paths, float handles: keep a densely filled S as dense 128 x 256 lower-triangle tiles (4 bytes per stored entry, half the matrix, no indices) swept like the dense kernel-matrix cache: 0 = when the expected fill is >= 0.4, 1 = never, 2 = always */
#define B200SVM_OPT_SPARSE_BLOCK 3 /* CSR paths: columns of the implicit matrix per block (power of two, 128..32768, default 8192; set before set_data_csr) */
int b200svm_set_option(b200svm_handle h, int option, int64_t value);

typedef struct b200svm_stats {
    int64_t kernel_launches;  /* CUDA kernels launched by this handle since creation / last reset */
    int64_t matvecs;          /* applications of Abar                                              */
    double matvec_ms;         /* accumulated device time of the O(m^2) kernel (CUDA events)       */
    int64_t matvec_path;      /* path used by the last matvec: 1 SIMT, 2 tcgen05, 3 dense GEMV pair, 4 sparse */
    int64_t h2d_bytes;        /* host->device bytes copied                                        */
    int64_t d2h_bytes;        /* device->host bytes copied                                        */
    int64_t collectives;      /* NCCL collectives enqueued                                        */
    int64_t collective_bytes; /* payload bytes of those collectives (per rank)                    */
    int64_t peer_exchanges;   /* of those collectives: done by peer-memory stores instead of NCCL  */
    int64_t matvecs_timed;    /* matvecs whose duration is included in matvec_ms (all of them unless the event ring overflowed) */
    double cache_build_ms;        /* device time spent assembling a kernel-matrix cache (dense tiles or sparse S); also inside matvec_ms */
    int64_t dense_cache_bytes;    /* dense tcgen05 path: bytes of kernel-matrix tiles cached on this rank (0 = recomputed every matvec) */
    int64_t sparse_cache_entries; /* CSR paths: stored entries of the cached S matrix on this rank (0 = recomputed on the fly) */
    double cg_vector_ms;          /* accumulated device time of the O(m) vector kernels of the CG iterations (CUDA events; several ranks: includes waiting for the peers' slots) */
    int64_t cg_vector_timed;      /* iterations included in cg_vector_ms */
} b200svm_stats;
int b200svm_get_stats(b200svm_handle h, b200svm_stats *out);
/* Measurement hook: the FP32 / FP64 FMA throughput of the handle's device in TFLOP/s (2 flops per FMA), measured with a register-only
 * kernel — the roofline denominator of the SIMT Gram-tile kernels (double precision dense polynomial / rbf). */
int b200svm_probe_fma_tflops(b200svm_handle h, int dtype, double *tflops);
/* Measurement hook: rate (1e9 per second) of 32-bit shared-memory atomic adds to spread addresses on the handle's device — the
 * instruction the on-the-fly sparse kernel issues per Gram product; roofline denominator of b200svm_predict_csr. */
int b200svm_probe_smem_atomic_gops(b200svm_handle h, double *gops);
int b200svm_reset_stats(b200svm_handle h);

#ifdef __cplusplus
}
#endif
#endif /* B200SVM_H */
