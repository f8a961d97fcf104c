// Test hooks (extern "C") over the host's parsers / writer so that tests/test_host_formats.py can compare them with the
// reference's own parsers compiled into oracle/_ref (oracle/ref_driver.cpp oref_parse_*).  Not needed by the executables.
#include "plssvm/core.hpp"

#include <omp.h>

#include <cstring>
#include <string>

namespace {
thread_local std::string g_err;

template <typename T, typename F>
int guarded(F &&f) {
    try {
        g_err.clear();
        f();
        return 0;
    } catch (const plssvm::invalid_file_format_exception &e) {
        g_err = e.what();
        return 2;
    } catch (const plssvm::file_not_found_exception &e) {
        g_err = e.what();
        return 3;
    } catch (const std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

template <typename T>
void copy_dense(const plssvm::csr_matrix<T> &m, void *X_out) {
    if (X_out == nullptr) { return; }
    T *x = static_cast<T *>(X_out);
    std::memset(x, 0, m.rows * m.cols * sizeof(T));
    for (std::size_t i = 0; i < m.rows; ++i) {
        for (std::int64_t p = m.rowptr[i]; p < m.rowptr[i + 1]; ++p) { x[i * m.cols + (std::size_t) m.col[(std::size_t) p]] = m.val[(std::size_t) p]; }
    }
}

template <typename T>
int parse_data(const char *path, long *n, long *F, long *nnz, int *has_labels, void *X_out, void *y_out, double *gamma_out) {
    return guarded<T>([&]() {
        plssvm::parameter<T> p;
        p.print_info = false;
        const std::string file{ path };
        if (file.size() >= 5 && file.substr(file.size() - 5) == ".arff") {
            p.parse_arff_file(file);
        } else {
            p.parse_libsvm_file(file);
        }
        *n = (long) p.sparse_data_ptr->rows;
        *F = (long) p.sparse_data_ptr->cols;
        *nnz = (long) p.sparse_data_ptr->nnz();
        *has_labels = p.value_ptr != nullptr;
        if (gamma_out != nullptr) { *gamma_out = (double) p.gamma; }
        copy_dense(*p.sparse_data_ptr, X_out);
        if (y_out != nullptr && p.value_ptr != nullptr) { std::memcpy(y_out, p.value_ptr->data(), p.value_ptr->size() * sizeof(T)); }
    });
}

template <typename T>
int parse_model(const char *path, int *kernel, int *degree, double *gamma, double *coef0, double *rho, long *n, long *F, void *X_out, void *alpha_out,
                void *y_out) {
    return guarded<T>([&]() {
        plssvm::parameter<T> p;
        p.print_info = false;
        p.parse_model_file(path);
        *kernel = (int) p.kernel;
        *degree = p.degree;
        *gamma = (double) p.gamma;
        *coef0 = (double) p.coef0;
        *rho = (double) p.rho;
        *n = (long) p.sparse_data_ptr->rows;
        *F = (long) p.sparse_data_ptr->cols;
        copy_dense(*p.sparse_data_ptr, X_out);
        if (alpha_out != nullptr) { std::memcpy(alpha_out, p.alpha_ptr->data(), p.alpha_ptr->size() * sizeof(T)); }
        if (y_out != nullptr && p.value_ptr != nullptr) { std::memcpy(y_out, p.value_ptr->data(), p.value_ptr->size() * sizeof(T)); }
    });
}

// test hook: the model text writer on host data (dense rows, or the same rows as CSR when `as_csr`), no device involved
template <typename T>
int write_model(const int kernel, const int degree, const double gamma, const double coef0, const void *X, const void *y, const void *alpha, const long n,
                const long F, const double rho, const char *path, const int as_csr) {
    return guarded<T>([&]() {
        const T *Xp = static_cast<const T *>(X);
        std::vector<std::vector<T>> rows((std::size_t) n);
        for (long i = 0; i < n; ++i) { rows[(std::size_t) i].assign(Xp + i * F, Xp + (i + 1) * F); }
        const std::vector<T> labels(static_cast<const T *>(y), static_cast<const T *>(y) + n);
        const std::vector<T> alphas(static_cast<const T *>(alpha), static_cast<const T *>(alpha) + n);
        const plssvm::csr_matrix<T> csr = plssvm::csr_matrix<T>::from_dense(rows);
        plssvm::detail::write_model_file<T>(path, static_cast<plssvm::kernel_type>(kernel), degree, (T) gamma, (T) coef0, (T) rho, labels, alphas,
                                            as_csr != 0 ? &csr : nullptr, &rows, (std::size_t) F, false);
    });
}
}  // namespace

extern "C" {
const char *b200host_last_error() { return g_err.c_str(); }
void b200host_set_threads(int threads) { omp_set_num_threads(threads > 0 ? threads : 1); }  // parser team size (tests)
int b200host_parse_data_file(int dtype, const char *path, long *n, long *F, long *nnz, int *has_labels, void *X_out, void *y_out, double *gamma_out) {
    return dtype == 0 ? parse_data<float>(path, n, F, nnz, has_labels, X_out, y_out, gamma_out) : parse_data<double>(path, n, F, nnz, has_labels, X_out, y_out, gamma_out);
}
int b200host_parse_model_file(int dtype, const char *path, int *kernel, int *degree, double *gamma, double *coef0, double *rho, long *n, long *F, void *X_out,
                              void *alpha_out, void *y_out) {
    return dtype == 0 ? parse_model<float>(path, kernel, degree, gamma, coef0, rho, n, F, X_out, alpha_out, y_out)
                      : parse_model<double>(path, kernel, degree, gamma, coef0, rho, n, F, X_out, alpha_out, y_out);
}
int b200host_format_shortest(int dtype, double v, char *out, int cap) {
    const std::string s = dtype == 0 ? plssvm::detail::format_shortest((float) v) : plssvm::detail::format_shortest(v);
    if ((int) s.size() + 1 > cap) { return 1; }
    std::memcpy(out, s.c_str(), s.size() + 1);
    return 0;
}
int b200host_write_model(int dtype, int kernel, int degree, double gamma, double coef0, const void *X, const void *y, const void *alpha, long n, long F,
                         double rho, const char *path, int as_csr) {
    return dtype == 0 ? write_model<float>(kernel, degree, gamma, coef0, X, y, alpha, n, F, rho, path, as_csr)
                      : write_model<double>(kernel, degree, gamma, coef0, X, y, alpha, n, F, rho, path, as_csr);
}
unsigned b200host_fp22_encode(float x) { return plssvm::detail::fp22_encode(x); }
}  // extern "C"
