// TEST INFRASTRUCTURE — C-ABI driver that runs the REFERENCE's own csvm::learn() / predict_label() / accuracy() control flow
// (src/plssvm/csvm.cpp:207-366, compiled unmodified from /root/reference) on top of the B200 backend subclass
// (integration/src/plssvm/backends/B200/csvm.cpp).  It proves the drop-in claim: the reference's base class drives
// libb200svm.so through the five virtuals exactly as it drives its own OpenMP / CUDA backends.
// Built by `make -C oracle backend` into oracle/_ref/libplssvm_b200_backend.so (needs /root/reference; the .so travels).
#include "plssvm/backends/B200/csvm.hpp"
#include "plssvm/csvm.hpp"
#include "plssvm/parameter.hpp"

#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace {

thread_local std::string g_error;

template <typename T>
class exposed : public plssvm::b200::csvm<T> {
    using base = plssvm::b200::csvm<T>;

  public:
    explicit exposed(const plssvm::parameter<T> &p) : base{ p } {}
    T get_bias() const { return base::bias_; }
    const std::vector<T> &get_alpha() const { return *base::alpha_ptr_; }
};

template <typename T>
std::shared_ptr<const std::vector<std::vector<T>>> make_data(const void *X, long n, long F) {
    const T *x = static_cast<const T *>(X);
    auto v = std::make_shared<std::vector<std::vector<T>>>(static_cast<std::size_t>(n));
    for (long i = 0; i < n; ++i) { (*v)[i].assign(x + i * F, x + (i + 1) * F); }
    return v;
}

template <typename T>
plssvm::parameter<T> make_params(int kernel, int degree, double gamma, double coef0, double cost, double eps, int target) {
    plssvm::parameter<T> p;
    p.kernel = static_cast<plssvm::kernel_type>(kernel);
    p.degree = degree;
    p.gamma = static_cast<T>(gamma);
    p.coef0 = static_cast<T>(coef0);
    p.cost = static_cast<T>(cost);
    p.epsilon = static_cast<T>(eps);
    p.print_info = false;
    p.backend = plssvm::backend_type::cuda;  // the reference's enum has no B200 entry; make_csvm is bypassed
    p.target = static_cast<plssvm::target_platform>(target);
    return p;
}

template <typename T>
int do_learn(int kernel, int degree, double gamma, double coef0, double cost, double eps, int target, const void *X, const void *y, long n, long F,
             void *alpha_out, double *rho_out, long *iters, int *sparse_rows, double *accuracy) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, cost, eps, target);
    p.data_ptr = make_data<T>(X, n, F);
    p.value_ptr = std::make_shared<const std::vector<T>>(static_cast<const T *>(y), static_cast<const T *>(y) + n);
    exposed<T> svm{ p };
    svm.learn();  // reference control flow: setup_data_on_device, generate_q || b, QA_cost, solver_CG, bias, alpha_n
    std::memcpy(alpha_out, svm.get_alpha().data(), static_cast<std::size_t>(n) * sizeof(T));
    *rho_out = -static_cast<double>(svm.get_bias());
    if (iters != nullptr) { *iters = static_cast<long>(svm.last_cg_iterations()); }
    if (sparse_rows != nullptr) { *sparse_rows = svm.uses_sparse_rows() ? 1 : 0; }
    if (accuracy != nullptr) { *accuracy = static_cast<double>(svm.accuracy()); }
    return 0;
}

template <typename T>
int do_predict(int kernel, int degree, double gamma, double coef0, const void *X, const void *alpha, long n, long F, double rho, const void *points,
               long P, long F_points, void *decision_out, void *label_out) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, 1.0, 1e-3, 0);
    p.data_ptr = make_data<T>(X, n, F);
    p.alpha_ptr = std::make_shared<const std::vector<T>>(static_cast<const T *>(alpha), static_cast<const T *>(alpha) + n);
    p.rho = static_cast<T>(rho);
    exposed<T> svm{ p };
    const auto pts = make_data<T>(points, P, F_points);
    if (decision_out != nullptr) {
        const std::vector<T> res = svm.predict(*pts);
        std::memcpy(decision_out, res.data(), res.size() * sizeof(T));
    }
    if (label_out != nullptr) {
        const std::vector<T> lab = svm.predict_label(*pts);
        std::memcpy(label_out, lab.data(), lab.size() * sizeof(T));
    }
    return 0;
}

template <typename Fn>
int guarded(Fn &&f) {
    try {
        g_error.clear();
        return f();
    } catch (const plssvm::b200::backend_exception &e) {
        g_error = std::string("b200::backend_exception: ") + e.what();
        return 3;
    } catch (const plssvm::exception &e) {
        g_error = std::string("plssvm::exception: ") + e.what();
        return 2;
    } catch (const std::exception &e) {
        g_error = std::string("std::exception: ") + e.what();
        return 1;
    }
}

}  // namespace

#define DISPATCH(dtype, fn, ...) guarded([&]() { return (dtype) == 0 ? fn<float>(__VA_ARGS__) : fn<double>(__VA_ARGS__); })

extern "C" {

const char *b200int_last_error() { return g_error.c_str(); }

int b200int_learn(int dtype, int kernel, int degree, double gamma, double coef0, double cost, double eps, int target, const void *X, const void *y, long n,
                  long F, void *alpha_out, double *rho_out, long *iters, int *sparse_rows, double *accuracy) {
    return DISPATCH(dtype, do_learn, kernel, degree, gamma, coef0, cost, eps, target, X, y, n, F, alpha_out, rho_out, iters, sparse_rows, accuracy);
}

int b200int_predict(int dtype, int kernel, int degree, double gamma, double coef0, const void *X, const void *alpha, long n, long F, double rho,
                    const void *points, long P, long F_points, void *decision_out, void *label_out) {
    return DISPATCH(dtype, do_predict, kernel, degree, gamma, coef0, X, alpha, n, F, rho, points, P, F_points, decision_out, label_out);
}

}  // extern "C"
