// TEST INFRASTRUCTURE — not product code.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load the library built from this file.
//
// C-ABI driver around the UNMODIFIED reference OpenMP backend (compiled from /root/reference by
// oracle/Makefile into oracle/_ref/).  It exposes protected members the same way the reference's own
// test mock does (tests/backends/OpenMP/mock_openmp_csvm.hpp:23-46):
//   generate_q          -> plssvm::openmp::csvm<T>::generate_q          (src/plssvm/backends/OpenMP/csvm.cpp:50-64)
//   run_device_kernel   -> plssvm::openmp::csvm<T>::run_device_kernel   (src/plssvm/backends/OpenMP/csvm.cpp:67-79)
//   solver_CG           -> plssvm::openmp::csvm<T>::solver_CG           (src/plssvm/backends/OpenMP/csvm.cpp:82-170)
//   learn               -> plssvm::csvm<T>::learn                       (src/plssvm/csvm.cpp:207-267)
//   predict             -> plssvm::openmp::csvm<T>::predict             (src/plssvm/backends/OpenMP/csvm.cpp:191-244)
// The CG iteration count is not returned by the reference; it is parsed from the solver's own
// "Finished after N iterations" line (src/plssvm/backends/OpenMP/csvm.cpp:162-167) captured from stdout.

#include "plssvm/backends/OpenMP/csvm.hpp"
#include "plssvm/csvm.hpp"
#include "plssvm/kernel_types.hpp"
#include "plssvm/parameter.hpp"

#include <omp.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace {

template <typename T>
class exposed_csvm : public plssvm::openmp::csvm<T> {
    using base_type = plssvm::openmp::csvm<T>;

  public:
    explicit exposed_csvm(const plssvm::parameter<T> &params) : base_type{ params } {}
    using base_type::generate_q;
    using base_type::run_device_kernel;
    using base_type::solver_CG;
    using base_type::setup_data_on_device;
    using base_type::predict;
    void set_QA_cost(const T v) { base_type::QA_cost_ = v; }
    T get_QA_cost() const { return base_type::QA_cost_; }
    T get_bias() const { return base_type::bias_; }
    const std::vector<T> &get_alpha() const { return *base_type::alpha_ptr_; }
    const std::vector<std::vector<T>> &get_data() const { return *base_type::data_ptr_; }
    T kfun(const std::vector<T> &a, const std::vector<T> &b) { return base_type::kernel_function(a, b); }
};

template <typename T>
std::shared_ptr<const std::vector<std::vector<T>>> make_data(const void *X, long n, long F) {
    const T *x = static_cast<const T *>(X);
    auto v = std::make_shared<std::vector<std::vector<T>>>(static_cast<std::size_t>(n));
    for (long i = 0; i < n; ++i) {
        (*v)[i].assign(x + i * F, x + (i + 1) * F);
    }
    return v;
}

template <typename T>
plssvm::parameter<T> make_params(int kernel, int degree, double gamma, double coef0, double cost, double eps, bool print) {
    plssvm::parameter<T> p;
    p.kernel = static_cast<plssvm::kernel_type>(kernel);
    p.degree = degree;
    p.gamma = static_cast<T>(gamma);
    p.coef0 = static_cast<T>(coef0);
    p.cost = static_cast<T>(cost);
    p.epsilon = static_cast<T>(eps);
    p.print_info = print;
    p.backend = plssvm::backend_type::openmp;
    p.target = plssvm::target_platform::cpu;
    return p;
}

// RAII stdout capture (fd level, so fmt::print's FILE* output is included).
struct stdout_capture {
    int saved_fd{ -1 };
    char path[64]{};
    stdout_capture() {
        std::fflush(stdout);
        std::snprintf(path, sizeof(path), "/tmp/oref_capture_%d_XXXXXX", static_cast<int>(getpid()));
        const int fd = mkstemp(path);
        saved_fd = dup(1);
        dup2(fd, 1);
        close(fd);
    }
    std::string finish() {
        std::fflush(stdout);
        dup2(saved_fd, 1);
        close(saved_fd);
        saved_fd = -1;
        std::string out;
        if (FILE *f = std::fopen(path, "rb")) {
            char buf[4096];
            std::size_t k = 0;
            while ((k = std::fread(buf, 1, sizeof(buf), f)) > 0) { out.append(buf, k); }
            std::fclose(f);
        }
        std::remove(path);
        return out;
    }
    ~stdout_capture() {
        if (saved_fd >= 0) { finish(); }
    }
};

long parse_iterations(const std::string &log) {
    const char *key = "Finished after ";
    const auto pos = log.rfind(key);
    if (pos == std::string::npos) { return -1; }
    return std::strtol(log.c_str() + pos + std::strlen(key), nullptr, 10);
}

template <typename T>
int do_generate_q(int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F, void *q_out) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, cost, 1e-3, false);
    p.data_ptr = make_data<T>(X, n, F);
    exposed_csvm<T> svm{ p };
    const std::vector<T> q = svm.generate_q();
    std::memcpy(q_out, q.data(), q.size() * sizeof(T));
    return 0;
}

template <typename T>
int do_matvec(int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F,
              const void *q_in, const void *d_in, double QA_cost, double add, void *ret_inout, double *seconds) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, cost, 1e-3, false);
    p.data_ptr = make_data<T>(X, n, F);
    exposed_csvm<T> svm{ p };
    svm.set_QA_cost(static_cast<T>(QA_cost));
    const long m = n - 1;
    const std::vector<T> q(static_cast<const T *>(q_in), static_cast<const T *>(q_in) + m);
    const std::vector<T> d(static_cast<const T *>(d_in), static_cast<const T *>(d_in) + m);
    std::vector<T> ret(static_cast<T *>(ret_inout), static_cast<T *>(ret_inout) + m);
    const auto t0 = std::chrono::steady_clock::now();
    svm.run_device_kernel(q, ret, d, svm.get_data(), static_cast<T>(add));
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds != nullptr) { *seconds = std::chrono::duration<double>(t1 - t0).count(); }
    std::memcpy(ret_inout, ret.data(), ret.size() * sizeof(T));
    return 0;
}

template <typename T>
int do_solver_cg(int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F,
                 const void *b_in, const void *q_in, double QA_cost, long imax, double eps, void *alpha_out, long *iters, double *seconds) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, cost, eps, true);
    p.data_ptr = make_data<T>(X, n, F);
    stdout_capture cap;
    exposed_csvm<T> svm{ p };
    svm.set_QA_cost(static_cast<T>(QA_cost));
    const long m = n - 1;
    const std::vector<T> q(static_cast<const T *>(q_in), static_cast<const T *>(q_in) + m);
    const std::vector<T> b(static_cast<const T *>(b_in), static_cast<const T *>(b_in) + m);
    const auto t0 = std::chrono::steady_clock::now();
    const std::vector<T> alpha = svm.solver_CG(b, static_cast<std::size_t>(imax), static_cast<T>(eps), q);
    const auto t1 = std::chrono::steady_clock::now();
    const std::string log = cap.finish();
    if (iters != nullptr) { *iters = parse_iterations(log); }
    if (seconds != nullptr) { *seconds = std::chrono::duration<double>(t1 - t0).count(); }
    std::memcpy(alpha_out, alpha.data(), alpha.size() * sizeof(T));
    return 0;
}

template <typename T>
int do_learn(int kernel, int degree, double gamma, double coef0, double cost, double eps, const void *X, const void *y, long n, long F,
             void *alpha_out, double *rho_out, long *iters, double *seconds) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, cost, eps, true);
    p.data_ptr = make_data<T>(X, n, F);
    p.value_ptr = std::make_shared<const std::vector<T>>(static_cast<const T *>(y), static_cast<const T *>(y) + n);
    stdout_capture cap;
    exposed_csvm<T> svm{ p };
    const auto t0 = std::chrono::steady_clock::now();
    svm.learn();
    const auto t1 = std::chrono::steady_clock::now();
    const std::string log = cap.finish();
    if (iters != nullptr) { *iters = parse_iterations(log); }
    if (seconds != nullptr) { *seconds = std::chrono::duration<double>(t1 - t0).count(); }
    std::memcpy(alpha_out, svm.get_alpha().data(), static_cast<std::size_t>(n) * sizeof(T));
    *rho_out = -static_cast<double>(svm.get_bias());  // model file stores rho = -bias (src/plssvm/csvm.cpp:123)
    return 0;
}

template <typename T>
int do_predict(int kernel, int degree, double gamma, double coef0, const void *X, const void *alpha, long n, long F, double rho,
               const void *points, long P, void *out) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, 1.0, 1e-3, false);
    p.data_ptr = make_data<T>(X, n, F);
    p.alpha_ptr = std::make_shared<const std::vector<T>>(static_cast<const T *>(alpha), static_cast<const T *>(alpha) + n);
    p.rho = static_cast<T>(rho);
    exposed_csvm<T> svm{ p };
    const auto pts = make_data<T>(points, P, F);
    const std::vector<T> res = svm.predict(*pts);
    std::memcpy(out, res.data(), res.size() * sizeof(T));
    return 0;
}

template <typename T>
int do_kernel_function(int kernel, int degree, double gamma, double coef0, const void *a, const void *b, long F, double *out) {
    const std::vector<T> va(static_cast<const T *>(a), static_cast<const T *>(a) + F);
    const std::vector<T> vb(static_cast<const T *>(b), static_cast<const T *>(b) + F);
    switch (kernel) {
        case 0: *out = plssvm::kernel_function<plssvm::kernel_type::linear>(va, vb); break;
        case 1: *out = plssvm::kernel_function<plssvm::kernel_type::polynomial>(va, vb, degree, static_cast<T>(gamma), static_cast<T>(coef0)); break;
        case 2: *out = plssvm::kernel_function<plssvm::kernel_type::rbf>(va, vb, static_cast<T>(gamma)); break;
        default: return 1;
    }
    return 0;
}


// ---- file formats (reference parsers / writer; used to pin the N1 row: src/plssvm/parameter.cpp:122-520, src/plssvm/csvm.cpp:60-204) ----
template <typename T>
int do_parse_data_file(const char *path, long *n, long *F, int *has_labels, void *X_out, void *y_out) {
    plssvm::parameter<T> p;
    p.print_info = false;
    p.parse_file(path, p.data_ptr);
    *n = static_cast<long>(p.data_ptr->size());
    *F = *n > 0 ? static_cast<long>((*p.data_ptr)[0].size()) : 0;
    *has_labels = p.value_ptr != nullptr ? 1 : 0;
    if (X_out != nullptr) {
        T *x = static_cast<T *>(X_out);
        for (long i = 0; i < *n; ++i) { std::memcpy(x + i * *F, (*p.data_ptr)[i].data(), sizeof(T) * *F); }
    }
    if (y_out != nullptr && p.value_ptr != nullptr) { std::memcpy(y_out, p.value_ptr->data(), sizeof(T) * *n); }
    return 0;
}

template <typename T>
int do_parse_model_file(const char *path, int *kernel, int *degree, double *gamma, double *coef0, double *rho, long *n, long *F,
                        void *X_out, void *alpha_out, void *y_out) {
    plssvm::parameter<T> p;
    p.print_info = false;
    p.parse_model_file(path);
    *kernel = static_cast<int>(p.kernel);
    *degree = p.degree;
    *gamma = p.gamma;
    *coef0 = p.coef0;
    *rho = p.rho;
    *n = static_cast<long>(p.data_ptr->size());
    *F = *n > 0 ? static_cast<long>((*p.data_ptr)[0].size()) : 0;
    if (X_out != nullptr) {
        T *x = static_cast<T *>(X_out);
        for (long i = 0; i < *n; ++i) { std::memcpy(x + i * *F, (*p.data_ptr)[i].data(), sizeof(T) * *F); }
    }
    if (alpha_out != nullptr && p.alpha_ptr != nullptr) { std::memcpy(alpha_out, p.alpha_ptr->data(), sizeof(T) * *n); }
    if (y_out != nullptr && p.value_ptr != nullptr) { std::memcpy(y_out, p.value_ptr->data(), sizeof(T) * *n); }
    return 0;
}

template <typename T>
int do_write_model(int kernel, int degree, double gamma, double coef0, const void *X, const void *y, const void *alpha, long n, long F, double rho, const char *path) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, 1.0, 1e-3, false);
    p.data_ptr = make_data<T>(X, n, F);
    p.value_ptr = std::make_shared<const std::vector<T>>(static_cast<const T *>(y), static_cast<const T *>(y) + n);
    p.alpha_ptr = std::make_shared<const std::vector<T>>(static_cast<const T *>(alpha), static_cast<const T *>(alpha) + n);
    p.rho = static_cast<T>(rho);
    exposed_csvm<T> svm{ p };
    svm.write_model(path);
    return 0;
}

template <typename F>
int guarded(F &&f) {
    try {
        return f();
    } catch (const std::exception &e) {
        std::fprintf(stderr, "[oracle/_ref] exception: %s\n", e.what());
        return 2;
    }
}

}  // namespace

#define DISPATCH(dtype, fn, ...) guarded([&]() { return (dtype) == 0 ? fn<float>(__VA_ARGS__) : fn<double>(__VA_ARGS__); })

extern "C" {

int oref_num_threads() { return omp_get_max_threads(); }
void oref_set_num_threads(int t) { omp_set_num_threads(t); }

int oref_generate_q(int dtype, int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F, void *q_out) {
    return DISPATCH(dtype, do_generate_q, kernel, degree, gamma, coef0, cost, X, n, F, q_out);
}
int oref_matvec(int dtype, int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F,
                const void *q, const void *d, double QA_cost, double add, void *ret_inout, double *seconds) {
    return DISPATCH(dtype, do_matvec, kernel, degree, gamma, coef0, cost, X, n, F, q, d, QA_cost, add, ret_inout, seconds);
}
int oref_solver_cg(int dtype, int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F,
                   const void *b, const void *q, double QA_cost, long imax, double eps, void *alpha_out, long *iters, double *seconds) {
    return DISPATCH(dtype, do_solver_cg, kernel, degree, gamma, coef0, cost, X, n, F, b, q, QA_cost, imax, eps, alpha_out, iters, seconds);
}
int oref_learn(int dtype, int kernel, int degree, double gamma, double coef0, double cost, double eps, const void *X, const void *y, long n, long F,
               void *alpha_out, double *rho_out, long *iters, double *seconds) {
    return DISPATCH(dtype, do_learn, kernel, degree, gamma, coef0, cost, eps, X, y, n, F, alpha_out, rho_out, iters, seconds);
}
int oref_predict(int dtype, int kernel, int degree, double gamma, double coef0, const void *X, const void *alpha, long n, long F, double rho,
                 const void *points, long P, void *out) {
    return DISPATCH(dtype, do_predict, kernel, degree, gamma, coef0, X, alpha, n, F, rho, points, P, out);
}
int oref_kernel_function(int dtype, int kernel, int degree, double gamma, double coef0, const void *a, const void *b, long F, double *out) {
    return DISPATCH(dtype, do_kernel_function, kernel, degree, gamma, coef0, a, b, F, out);
}

int oref_parse_data_file(int dtype, const char *path, long *n, long *F, int *has_labels, void *X_out, void *y_out) {
    return DISPATCH(dtype, do_parse_data_file, path, n, F, has_labels, X_out, y_out);
}
int oref_parse_model_file(int dtype, const char *path, int *kernel, int *degree, double *gamma, double *coef0, double *rho, long *n, long *F,
                          void *X_out, void *alpha_out, void *y_out) {
    return DISPATCH(dtype, do_parse_model_file, path, kernel, degree, gamma, coef0, rho, n, F, X_out, alpha_out, y_out);
}
int oref_write_model(int dtype, int kernel, int degree, double gamma, double coef0, const void *X, const void *y, const void *alpha, long n, long F, double rho, const char *path) {
    return DISPATCH(dtype, do_write_model, kernel, degree, gamma, coef0, X, y, alpha, n, F, rho, path);
}

}  // extern "C"
