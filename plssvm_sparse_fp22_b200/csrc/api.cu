// C ABI of libb200svm.so (see include/b200svm.h for the contract and the reference interfaces each entry replaces).
#include "internal.cuh"

#include <algorithm>
#include <cstring>
#include <exception>
#include <new>

using namespace b200;

namespace {

thread_local std::string g_create_error;

template <typename Fn>
int guarded(ctx *c, Fn &&fn, bool need_handle = true) {
    if (c == nullptr && need_handle) {
        g_create_error = "null b200svm_handle";
        return B200SVM_ERR_INVALID;
    }
    try {
        if (c != nullptr) { B200_CUDA_CHECK(cudaSetDevice(c->device)); }  // callers may hop threads (src/plssvm/csvm.cpp:230-235)
        fn();
        return B200SVM_OK;
    } catch (const error &e) {
        (c != nullptr ? c->last_error : g_create_error) = e.msg;
        return e.code;
    } catch (const std::bad_alloc &) {
        (c != nullptr ? c->last_error : g_create_error) = "host allocation failed";
        return B200SVM_ERR_ALLOC;
    } catch (const std::exception &e) {
        (c != nullptr ? c->last_error : g_create_error) = e.what();
        return B200SVM_ERR_INVALID;
    }
}

#define DISPATCH_T(c, expr_f32, expr_f64)   \
    do {                                    \
        if ((c)->dtype == B200SVM_F32) {    \
            using T = float;                \
            expr_f32;                       \
        } else {                            \
            using T = double;               \
            expr_f64;                       \
        }                                   \
    } while (0)

// The dense entry points take HOST pointers or DEVICE pointers (unified addressing tells them apart): a caller that already
// holds its data in HBM — e.g. torch CUDA tensors, plssvm_sparse_fp22_b200/torch_api.py — pays a device-to-device copy only.
bool on_device(const void *p) {
    cudaPointerAttributes a{};
    if (p == nullptr || cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        (void) cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}
void h2d(ctx &c, void *dst, const void *src, size_t bytes) {
    if (bytes == 0) { return; }
    B200_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, c.stream));
    if (!on_device(src)) { c.stats.h2d_bytes += (int64_t) bytes; }
}
void d2h(ctx &c, void *dst, const void *src, size_t bytes) {
    if (bytes == 0) { return; }
    B200_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, c.stream));
    if (!on_device(dst)) { c.stats.d2h_bytes += (int64_t) bytes; }
}

// copy a host row-major matrix (rows x F, leading dimension ld) into a zero-padded device matrix (rows x ldd)
void stage_matrix(ctx &c, void *dst, int64_t ldd, const void *src, int64_t rows, int64_t F, int64_t ld) {
    const size_t es = c.esize();
    B200_CUDA_CHECK(cudaMemsetAsync(dst, 0, (size_t) rows * ldd * es, c.stream));
    B200_CUDA_CHECK(cudaMemcpy2DAsync(dst, (size_t) ldd * es, src, (size_t) ld * es, (size_t) F * es, (size_t) rows, cudaMemcpyDefault, c.stream));
    if (!on_device(src)) { c.stats.h2d_bytes += (int64_t) ((size_t) rows * F * es); }
}

void require_data(const ctx &c) { B200_REQUIRE(c.has_data, "no data set: call b200svm_set_data_dense / b200svm_set_data_csr first"); }
void require_q(const ctx &c) { B200_REQUIRE(c.has_q, "q vector missing: call b200svm_generate_q or b200svm_set_q first"); }

template <typename T>
void do_generate_q(ctx &c, void *q_out) {
    require_data(c);
    B200_REQUIRE(c.n >= 2, "at least two data points are required");
    c.q.reserve((size_t) c.m() * sizeof(T));
    if (c.is_csr) {
        sp_generate_q<T>(c);
    } else {
        generate_q<T>(c);
    }
    c.has_q = true;
    if (q_out != nullptr) { d2h(c, q_out, c.q.p, (size_t) c.m() * sizeof(T)); }
    sync_stream(c);
}

template <typename T>
T read_scalar(ctx &c, int idx) {
    T *mb = static_cast<T *>(c.pinned);
    B200_CUDA_CHECK(cudaMemcpyAsync(mb, c.scal.as<T>() + idx, sizeof(T), cudaMemcpyDeviceToHost, c.stream));
    sync_stream(c);
    return mb[0];
}

}  // namespace

extern "C" {

int b200svm_version(void) { return 1000; }

const char *b200svm_last_error(b200svm_handle h) {
    return h != nullptr ? reinterpret_cast<ctx *>(h)->last_error.c_str() : g_create_error.c_str();
}

int b200svm_create(b200svm_handle *h, int dtype, int kernel, int degree, double gamma, double coef0, double cost, int device) {
    if (h == nullptr) {
        g_create_error = "null handle pointer";
        return B200SVM_ERR_INVALID;
    }
    *h = nullptr;
    ctx *c = nullptr;
    const auto create_body = [&]() {
        B200_REQUIRE(dtype == B200SVM_F32 || dtype == B200SVM_F64, "dtype must be B200SVM_F32 or B200SVM_F64");
        B200_REQUIRE(kernel >= B200SVM_LINEAR && kernel <= B200SVM_RBF, "unknown kernel type");
        B200_REQUIRE(cost != 0.0, "cost must not be 0");
        int count = 0;
        const cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count == 0) {
            fail(B200SVM_ERR_CUDA, std::string("no CUDA device available (") + cudaGetErrorString(e) + "): libb200svm has no CPU fallback");
        }
        B200_REQUIRE(device >= 0 && device < count, "device ordinal out of range");
        cudaDeviceProp prop{};
        B200_CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
        if (prop.major != 10) {
            fail(B200SVM_ERR_CUDA, std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major) + std::to_string(prop.minor) +
                                       "; libb200svm is built for sm_100a (B200) only");
        }
        B200_CUDA_CHECK(cudaSetDevice(device));
        c = new ctx();
        c->dtype = dtype;
        c->kernel = kernel;
        c->degree = degree;
        c->gamma = gamma;
        c->coef0 = coef0;
        c->cost = cost;
        c->device = device;
        B200_CUDA_CHECK(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
        c->stream = c->own_stream;
        for (auto &ev : c->ev) { B200_CUDA_CHECK(cudaEventCreate(&ev)); }
        for (auto &ev : c->sv) { B200_CUDA_CHECK(cudaEventCreate(&ev)); }
        for (auto &ev : c->cg_ev) { B200_CUDA_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)); }
        for (auto &ev : c->cb_ev) { B200_CUDA_CHECK(cudaEventCreate(&ev)); }
        B200_CUDA_CHECK(cudaMallocHost(&c->pinned, 512));
        c->scal.reserve(S_COUNT * sizeof(double));
        c->partials.reserve(4 * MAX_PARTIAL_BLOCKS * sizeof(double));
        c->counters.reserve(8 * sizeof(unsigned int));
        B200_CUDA_CHECK(cudaMemsetAsync(c->scal.p, 0, c->scal.bytes, c->stream));
        B200_CUDA_CHECK(cudaMemsetAsync(c->counters.p, 0, c->counters.bytes, c->stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    };
    const int rc = guarded(nullptr, create_body, false);
    if (rc != B200SVM_OK) {
        delete c;
        return rc;
    }
    *h = reinterpret_cast<b200svm_handle>(c);
    return B200SVM_OK;
}

int b200svm_destroy(b200svm_handle h) {
    if (h == nullptr) { return B200SVM_OK; }
    ctx *c = reinterpret_cast<ctx *>(h);
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    comm_destroy(*c);
    tile_store_destroy(c->dts);
    if (c->tc_opaque != nullptr && c->tc_opaque_free != nullptr) { c->tc_opaque_free(c->tc_opaque); }
    if (c->sp_opaque != nullptr && c->sp_opaque_free != nullptr) { c->sp_opaque_free(c->sp_opaque); }
    for (auto &ev : c->ev) {
        if (ev != nullptr) { cudaEventDestroy(ev); }
    }
    for (auto &ev : c->cg_ev) {
        if (ev != nullptr) { cudaEventDestroy(ev); }
    }
    for (auto &ev : c->sv) {
        if (ev != nullptr) { cudaEventDestroy(ev); }
    }
    for (auto &ev : c->cb_ev) {
        if (ev != nullptr) { cudaEventDestroy(ev); }
    }
    if (c->pinned != nullptr) { cudaFreeHost(c->pinned); }
    if (c->own_stream != nullptr) { cudaStreamDestroy(c->own_stream); }
    delete c;
    return B200SVM_OK;
}

int b200svm_set_stream(b200svm_handle h, void *cuda_stream) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        B200_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        c->stream = cuda_stream != nullptr ? static_cast<cudaStream_t>(cuda_stream) : c->own_stream;
    });
}

int b200svm_set_option(b200svm_handle h, int option, int64_t value) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        switch (option) {
            case B200SVM_OPT_MATVEC_PATH:
                B200_REQUIRE(value >= 0 && value <= 2, "matvec path must be 0 (auto), 1 (SIMT) or 2 (tcgen05)");
                c->opt_matvec_path = value;
                break;
            case B200SVM_OPT_SYNC_EVERY:
                B200_REQUIRE(value >= 0 && value <= 8, "CG look-ahead must be in [0, 8] iterations");
                c->opt_sync_every = value;
                break;
            case B200SVM_OPT_SPARSE_BLOCK:
                B200_REQUIRE(value == 0 || (value >= 128 && value <= 32768 && (value & (value - 1)) == 0), "sparse block width must be 0 or a power of two in [128, 32768]");
                c->opt_sparse_block = value;
                break;
            case B200SVM_OPT_TC_MODE:
                B200_REQUIRE(value >= 0 && value <= 2, "tcgen05 mode must be 0 (default), 1 (f16 x 3) or 2 (fp8 cross terms)");
                c->opt_tc_mode = value;
                if (c->has_data && c->has_split) { tc_prepare_split(*c); }
                break;
            case B200SVM_OPT_SPARSE_LANES:
                B200_REQUIRE(value >= 0 && value <= 32 && (value & (value - 1)) == 0, "lanes per segment must be 0 (auto) or a power of two up to 32");
                c->opt_sparse_lanes = value;
                break;
            case B200SVM_OPT_SPARSE_THREADS:
                B200_REQUIRE(value == 0 || (value >= 64 && value <= 1024 && (value & (value - 1)) == 0), "threads per row must be 0 (auto) or a power of two in [64, 1024]");
                c->opt_sparse_threads = value;
                break;
            case B200SVM_OPT_SPARSE_CACHE:
                B200_REQUIRE(value >= 0 && value <= 2, "sparse cache mode must be 0 (auto), 1 (off) or 2 (required)");
                c->opt_sparse_cache = value;
                sp_invalidate_cache(*c);
                break;
            case B200SVM_OPT_SPARSE_CACHE_MB:
                B200_REQUIRE(value >= 0, "cache budget must be >= 0 MiB");
                c->opt_sparse_cache_mb = value;
                sp_invalidate_cache(*c);
                tc_invalidate_cache(*c);
                dense_f64_cache_invalidate(*c);
                break;
            case B200SVM_OPT_PEER_EXCHANGE:
                B200_REQUIRE(value >= 0 && value <= 2, "peer exchange mode must be 0 (auto), 1 (NCCL allreduce) or 2 (required)");
                B200_REQUIRE(c->comm == nullptr, "set the peer exchange mode before b200svm_comm_init");
                c->opt_peer_exchange = value;
                break;
            case B200SVM_OPT_DENSE_CACHE:
                B200_REQUIRE(value >= 0 && value <= 2, "dense cache mode must be 0 (auto), 1 (off) or 2 (required)");
                c->opt_dense_cache = value;
                tc_invalidate_cache(*c);
                dense_f64_cache_invalidate(*c);
                break;
            case B200SVM_OPT_SPARSE_TILES:
                B200_REQUIRE(value >= 0 && value <= 2, "sparse tile store mode must be 0 (auto), 1 (never) or 2 (always)");
                c->opt_sparse_tiles = value;
                sp_invalidate_cache(*c);
                break;
            case B200SVM_OPT_SPARSE_GROUP:
                B200_REQUIRE(value >= 0, "blocks per group must be >= 0");
                c->opt_sparse_group = value;
                break;
            default: fail(B200SVM_ERR_INVALID, "unknown option");
        }
    });
}

int b200svm_probe_fma_tflops(b200svm_handle h, int dtype, double *tflops) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        B200_REQUIRE(tflops != nullptr && (dtype == B200SVM_F32 || dtype == B200SVM_F64), "bad arguments");
        *tflops = dtype == B200SVM_F32 ? probe_fma_tflops<float>(*c) : probe_fma_tflops<double>(*c);
    });
}

int b200svm_probe_smem_atomic_gops(b200svm_handle h, double *gops) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        B200_REQUIRE(gops != nullptr, "null output pointer");
        *gops = probe_smem_atomic_gops(*c);
    });
}

int b200svm_get_stats(b200svm_handle h, b200svm_stats *out) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        B200_REQUIRE(out != nullptr, "null stats pointer");
        *out = c->stats;
    });
}

int b200svm_reset_stats(b200svm_handle h) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        const int64_t cached = c->stats.sparse_cache_entries, dense = c->stats.dense_cache_bytes;  // properties of the resident data
        sync_stream(*c);                                       // account the events still in flight before zeroing
        c->stats = b200svm_stats{};
        c->stats.sparse_cache_entries = cached;
        c->stats.dense_cache_bytes = dense;
    });
}

int b200svm_set_data_dense(b200svm_handle h, const void *X, int64_t n, int64_t F, int64_t ld) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        B200_REQUIRE(X != nullptr, "No data points provided!");               // src/plssvm/csvm.cpp:42
        B200_REQUIRE(n > 0, "Data set is empty!");                             // src/plssvm/csvm.cpp:44
        B200_REQUIRE(F > 0, "No features provided for the data points!");      // src/plssvm/csvm.cpp:48
        B200_REQUIRE(ld >= F, "leading dimension smaller than the number of features");
        const size_t es = c->esize();
        c->is_csr = false;
        c->has_data = false;
        c->has_q = false;
        c->has_split = false;
        c->has_Z = false;
        dense_f64_cache_invalidate(*c);
        c->n = n;
        c->F = F;
        c->ldx = (F + 15) / 16 * 16;
        c->norms.reserve((size_t) n * es);
        static const bool full_upload = [] {
            const char *e = std::getenv("B200SVM_FULL_UPLOAD");  // A/B switch: every rank uploads the whole matrix over PCIe
            return e != nullptr && e[0] == '1';
        }();
        if (c->world > 1 && c->comm != nullptr && !full_upload && !on_device(X)) {
            // several ranks (every rank is handed the same matrix): upload 1 / world of the rows over PCIe, fetch the rest from the
            // peers over NVLink (in-place ncclAllGather) — the host-to-device copy is the largest part of a multi-GPU set_data
            const int64_t per = (n + c->world - 1) / c->world;
            c->X.reserve((size_t) per * c->world * c->ldx * es);
            const int64_t lo = std::min<int64_t>(n, per * c->rank), hi = std::min<int64_t>(n, per * (c->rank + 1));
            char *share = static_cast<char *>(c->X.p) + (size_t) per * c->rank * c->ldx * es;
            B200_CUDA_CHECK(cudaMemsetAsync(share, 0, (size_t) per * c->ldx * es, c->stream));
            if (hi > lo) { stage_matrix(*c, share, c->ldx, static_cast<const char *>(X) + (size_t) lo * ld * es, hi - lo, F, ld); }
            comm_allgather_inplace(*c, c->X.p, (size_t) per * c->ldx * es);
        } else {
            c->X.reserve((size_t) n * c->ldx * es);
            stage_matrix(*c, c->X.p, c->ldx, X, n, F, ld);
        }
        DISPATCH_T(c, compute_norms<T>(*c, c->X.as<T>(), n, c->ldx, F, c->norms.as<T>()), compute_norms<T>(*c, c->X.as<T>(), n, c->ldx, F, c->norms.as<T>()));
        c->has_data = true;
        if (c->dtype == B200SVM_F32 && c->kernel != B200SVM_LINEAR && c->opt_matvec_path != 1) { tc_prepare_split(*c); }
        sync_stream(*c);
    });
}

int b200svm_generate_q(b200svm_handle h, void *q_out) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() { DISPATCH_T(c, do_generate_q<T>(*c, q_out), do_generate_q<T>(*c, q_out)); });
}

int b200svm_get_kxx_last(b200svm_handle h, double *kxx) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        require_q(*c);
        B200_REQUIRE(kxx != nullptr, "null output pointer");
        DISPATCH_T(c, *kxx = (double) read_scalar<T>(*c, S_KXX), *kxx = (double) read_scalar<T>(*c, S_KXX));
    });
}

int b200svm_set_q(b200svm_handle h, const void *q, double QA_cost) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        require_data(*c);
        B200_REQUIRE(c->n >= 2, "at least two data points are required");
        if (q != nullptr) {
            c->q.reserve((size_t) c->m() * c->esize());
            h2d(*c, c->q.p, q, (size_t) c->m() * c->esize());
            c->has_q = true;
        }
        c->QA_cost = QA_cost;
        sync_stream(*c);
    });
}

int b200svm_matvec(b200svm_handle h, const void *d, double add, void *ret_inout) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        require_data(*c);
        require_q(*c);
        B200_REQUIRE(d != nullptr && ret_inout != nullptr, "null vector");
        const size_t vb = (size_t) c->m() * c->esize();
        c->v_tmp.reserve(vb);
        c->v_b.reserve(vb);
        h2d(*c, c->v_tmp.p, d, vb);
        h2d(*c, c->v_b.p, ret_inout, vb);
        DISPATCH_T(c, apply_abar<T>(*c, c->v_tmp.as<T>(), (T) add, c->v_b.as<T>()), apply_abar<T>(*c, c->v_tmp.as<T>(), (T) add, c->v_b.as<T>()));
        d2h(*c, ret_inout, c->v_b.p, vb);
        sync_stream(*c);
    });
}

int b200svm_cg_solve(b200svm_handle h, const void *b, const void *q, double QA_cost, int64_t imax, double eps, void *alpha_out, int64_t *iters,
                     double *delta) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        require_data(*c);
        B200_REQUIRE(b != nullptr && alpha_out != nullptr, "null vector");
        B200_REQUIRE(c->n >= 2, "at least two data points are required");
        const size_t vb = (size_t) c->m() * c->esize();
        if (q != nullptr) {
            c->q.reserve(vb);
            h2d(*c, c->q.p, q, vb);
            c->has_q = true;
        }
        require_q(*c);
        c->QA_cost = QA_cost;
        c->v_b.reserve(vb);
        c->v_x.reserve(vb);
        h2d(*c, c->v_b.p, b, vb);
        int64_t it = 0;
        DISPATCH_T(c, it = cg_solve<T>(*c, c->v_b.as<T>(), imax, (T) eps, c->v_x.as<T>(), delta),
                   it = cg_solve<T>(*c, c->v_b.as<T>(), imax, (T) eps, c->v_x.as<T>(), delta));
        d2h(*c, alpha_out, c->v_x.p, vb);
        sync_stream(*c);
        if (iters != nullptr) { *iters = it; }
    });
}

}  // extern "C"

namespace {

// csvm::learn() numeric path (src/plssvm/csvm.cpp:228-258)
template <typename T>
void do_learn(ctx &c, const void *y_host, double eps, void *alpha_out, double *rho_out, int64_t *iters) {
    require_data(c);
    B200_REQUIRE(y_host != nullptr, "No labels given for training! Maybe the data is only usable for prediction?");  // csvm.cpp:214
    B200_REQUIRE(alpha_out != nullptr && rho_out != nullptr, "null output pointer");
    B200_REQUIRE(c.n >= 2, "at least two data points are required");
    const int64_t n = c.n, m = c.m();
    std::vector<T> y_copy;
    if (on_device(y_host)) {  // device-resident labels: n reals come to the host for b = y - y_n and the bias
        y_copy.resize((size_t) n);
        B200_CUDA_CHECK(cudaMemcpyAsync(y_copy.data(), y_host, (size_t) n * sizeof(T), cudaMemcpyDefault, c.stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
        y_host = y_copy.data();
    }
    const T *y = static_cast<const T *>(y_host);
    // q = generate_q() ; QA_cost = k(x_n, x_n) + 1 / C   (csvm.cpp:232-243)
    c.q.reserve((size_t) m * sizeof(T));
    if (c.is_csr) {
        sp_generate_q<T>(c);
    } else {
        generate_q<T>(c);
    }
    c.has_q = true;
    // b = y[0..m) - y_n  (csvm.cpp:238-239)
    std::vector<T> b((size_t) m);
    for (int64_t i = 0; i < m; ++i) { b[(size_t) i] = y[i] - y[n - 1]; }
    c.v_b.reserve((size_t) m * sizeof(T));
    c.v_x.reserve((size_t) n * sizeof(T));
    h2d(c, c.v_b.p, b.data(), (size_t) m * sizeof(T));
    const T kxx = read_scalar<T>(c, S_KXX);  // syncs: b may go out of scope afterwards
    const T QA_cost = kxx + T(1) / (T) c.cost;
    c.QA_cost = (double) QA_cost;
    // alpha = solver_CG(b, num_features, epsilon, q)   (csvm.cpp:256)
    const int64_t it = cg_solve<T>(c, c.v_b.as<T>(), c.F, (T) eps, c.v_x.as<T>(), nullptr);
    // bias = y_n + QA_cost * sum(alpha) - q.alpha ; alpha_n = -sum(alpha)   (csvm.cpp:257-258)
    sums2<T>(c, c.v_x.as<T>(), S_SUM_X, S_Q_X);
    T *mb = static_cast<T *>(c.pinned);
    B200_CUDA_CHECK(cudaMemcpyAsync(mb, c.scal.as<T>() + S_SUM_X, 2 * sizeof(T), cudaMemcpyDeviceToHost, c.stream));
    d2h(c, alpha_out, c.v_x.p, (size_t) m * sizeof(T));
    sync_stream(c);
    const T sum_alpha = mb[0], q_alpha = mb[1];
    const T bias = y[n - 1] + QA_cost * sum_alpha - q_alpha;
    const T alpha_n = -sum_alpha;
    if (on_device(alpha_out)) {
        B200_CUDA_CHECK(cudaMemcpyAsync(static_cast<T *>(alpha_out) + m, &alpha_n, sizeof(T), cudaMemcpyDefault, c.stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
    } else {
        static_cast<T *>(alpha_out)[m] = alpha_n;
    }
    *rho_out = -(double) bias;
    if (iters != nullptr) { *iters = it; }
}

template <typename T>
void do_update_w(ctx &c, const void *alpha, void *w_out) {
    require_data(c);
    B200_REQUIRE(alpha != nullptr && w_out != nullptr, "No alphas provided for prediction!");
    c.alpha_dev.reserve((size_t) c.n * sizeof(T));
    c.wvec.reserve((size_t) c.ldx * sizeof(T));
    h2d(c, c.alpha_dev.p, alpha, (size_t) c.n * sizeof(T));
    if (c.is_csr) {
        sp_update_w<T>(c, c.alpha_dev.as<T>(), c.wvec.as<T>());
    } else {
        update_w<T>(c, c.alpha_dev.as<T>(), c.wvec.as<T>());
    }
    d2h(c, w_out, c.wvec.p, (size_t) c.F * sizeof(T));
    sync_stream(c);
}

template <typename T>
void do_predict_dense(ctx &c, const void *alpha, double rho, const void *points, int64_t P, int64_t ld, void *out) {
    require_data(c);
    B200_REQUIRE(!c.is_csr, "dense points against CSR support vectors: pass the points as CSR (b200svm_predict_csr)");
    B200_REQUIRE(alpha != nullptr, "No alphas provided for prediction!");  // src/plssvm/backends/OpenMP/csvm.cpp:207
    if (P == 0) { return; }                                                // csvm.cpp:198-200
    B200_REQUIRE(points != nullptr && out != nullptr, "null pointer");
    B200_REQUIRE(ld >= c.F, "Number of features per data point must match the number of features per predict point!");
    c.alpha_dev.reserve((size_t) c.n * sizeof(T));
    h2d(c, c.alpha_dev.p, alpha, (size_t) c.n * sizeof(T));
    // stage the points in chunks of at most ~1 GiB
    int64_t chunk = (int64_t) ((size_t(1) << 30) / ((size_t) c.ldx * sizeof(T)));
    if (chunk < 1024) { chunk = 1024; }
    if (chunk > P) { chunk = P; }
    c.pts.reserve((size_t) chunk * c.ldx * sizeof(T));
    c.pred_out.reserve((size_t) chunk * sizeof(T));
    const T bias = (T) (-rho);  // src/plssvm/csvm.cpp:42
    for (int64_t p0 = 0; p0 < P; p0 += chunk) {
        const int64_t rows = P - p0 < chunk ? P - p0 : chunk;
        stage_matrix(c, c.pts.p, c.ldx, static_cast<const T *>(points) + p0 * ld, rows, c.F, ld);
        predict_dense<T>(c, c.alpha_dev.as<T>(), bias, c.pts.as<T>(), rows, c.ldx, c.pred_out.as<T>());
        d2h(c, static_cast<T *>(out) + p0, c.pred_out.p, (size_t) rows * sizeof(T));
        sync_stream(c);
    }
}

}  // namespace

extern "C" {

int b200svm_learn(b200svm_handle h, const void *y, double eps, void *alpha_out, double *rho_out, int64_t *iters) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() { DISPATCH_T(c, do_learn<T>(*c, y, eps, alpha_out, rho_out, iters), do_learn<T>(*c, y, eps, alpha_out, rho_out, iters)); });
}

int b200svm_update_w(b200svm_handle h, const void *alpha, void *w_out) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() { DISPATCH_T(c, do_update_w<T>(*c, alpha, w_out), do_update_w<T>(*c, alpha, w_out)); });
}

int b200svm_predict_dense(b200svm_handle h, const void *alpha, double rho, const void *points, int64_t P, int64_t ld, void *out) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        DISPATCH_T(c, do_predict_dense<T>(*c, alpha, rho, points, P, ld, out), do_predict_dense<T>(*c, alpha, rho, points, P, ld, out));
    });
}

int b200svm_set_data_csr(b200svm_handle h, const int64_t *rowptr, const int32_t *colidx, const void *values, int64_t n, int64_t F, int64_t nnz,
                         int value_format) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        B200_REQUIRE(rowptr != nullptr, "No data points provided!");        // src/plssvm/csvm.cpp:42
        B200_REQUIRE(n > 0, "Data set is empty!");                           // src/plssvm/csvm.cpp:44
        B200_REQUIRE(F > 0, "No features provided for the data points!");    // src/plssvm/csvm.cpp:48
        c->has_data = false;
        c->has_q = false;
        c->has_split = false;
        c->has_Z = false;
        sp_set_data(*c, rowptr, colidx, values, n, F, nnz, value_format);
        c->is_csr = true;
        c->has_data = true;
        sync_stream(*c);
    });
}

int b200svm_predict_csr(b200svm_handle h, const void *alpha, double rho, const int64_t *rowptr, const int32_t *colidx, const void *values, int64_t P,
                        int64_t nnz, int value_format, void *out) {
    ctx *c = reinterpret_cast<ctx *>(h);
    return guarded(c, [&]() {
        require_data(*c);
        B200_REQUIRE(c->is_csr, "CSR predict needs CSR support vectors (b200svm_set_data_csr)");
        B200_REQUIRE(alpha != nullptr, "No alphas provided for prediction!");  // src/plssvm/backends/OpenMP/csvm.cpp:207
        if (P == 0) { return; }                                                // csvm.cpp:198-200
        B200_REQUIRE(rowptr != nullptr && out != nullptr, "null pointer");
        c->alpha_dev.reserve((size_t) c->n * c->esize());
        h2d(*c, c->alpha_dev.p, alpha, (size_t) c->n * c->esize());
        DISPATCH_T(c, sp_predict<T>(*c, c->alpha_dev.as<T>(), (T) (-rho), rowptr, colidx, values, P, nnz, value_format, static_cast<T *>(out)),
                   sp_predict<T>(*c, c->alpha_dev.as<T>(), (T) (-rho), rowptr, colidx, values, P, nnz, value_format, static_cast<T *>(out)));
        sync_stream(*c);
    });
}

}  // extern "C"
