/**
 * B200 backend for PLSSVM — implementation of plssvm::b200::csvm<T> on top of the C ABI of libb200svm.so (include/b200svm.h).
 * Each virtual is a thin translation: std::vector <-> plain pointers, status code -> plssvm::b200::backend_exception.
 * Reference semantics followed: src/plssvm/backends/OpenMP/csvm.cpp (the oracle backend), error texts as asserted by the
 * reference's tests (tests/backends/generic_tests.hpp, tests/backends/CUDA/test.cpp:55).
 */
#include "plssvm/backends/B200/csvm.hpp"

#include "b200svm.h"

#include "plssvm/detail/assert.hpp"    // PLSSVM_ASSERT
#include "plssvm/kernel_types.hpp"     // plssvm::kernel_type
#include "plssvm/target_platforms.hpp"  // plssvm::target_platform

#include "fmt/chrono.h"   // format std::chrono
#include "fmt/core.h"     // fmt::print, fmt::format
#include "fmt/ostream.h"  // operator<< overloads of the enums

#include <algorithm>    // std::all_of
#include <chrono>       // std::chrono
#include <cstdint>      // std::int64_t
#include <type_traits>  // std::is_same_v
#include <vector>       // std::vector

namespace plssvm::b200 {

backend_exception::backend_exception(const std::string &msg, source_location loc) :
    ::plssvm::exception{ msg, "b200::backend_exception", loc } {}

namespace {

b200svm_handle as_handle(void *p) { return static_cast<b200svm_handle>(p); }

// rows are stored sparsely on the device when at most this fraction of the entries is non-zero (and no row is too long for the
// sparse kernels); the reference itself always densifies (src/plssvm/parameter.cpp:41-116)
constexpr double sparse_density_threshold = 0.05;
constexpr std::size_t sparse_max_row = 4096;  // the sparse kernels stage one row in shared memory: longer rows go dense (libb200svm rejects them)

}  // namespace

template <typename T>
void csvm<T>::check(const int status) const {
    if (status != B200SVM_OK) {
        throw backend_exception{ fmt::format("B200 backend error {}: {}", status, b200svm_last_error(as_handle(handle_))) };
    }
}

template <typename T>
csvm<T>::csvm(const parameter<T> &params) :
    base_type{ params } {
    // same target check as the CUDA backend (src/plssvm/backends/CUDA/csvm.cu:41-48)
    if (target_ != target_platform::automatic && target_ != target_platform::gpu_nvidia) {
        throw backend_exception{ fmt::format("Invalid target platform '{}' for the B200 backend!", target_) };
    }
    if (print_info_) {
        fmt::print("Using B200 (sm_100a, libb200svm {}) as backend.\n\n", b200svm_version());
    }
    b200svm_handle h = nullptr;
    const int dtype = std::is_same_v<T, float> ? B200SVM_F32 : B200SVM_F64;
    // gamma_ == 0 means "1 / num_features", resolved once the data is known (src/plssvm/csvm.cpp:222-224); the handle wants the value
    const double gamma = gamma_ != T{ 0 } ? static_cast<double>(gamma_) : 1.0 / static_cast<double>(std::max<std::size_t>(num_features_, 1));
    const int status = b200svm_create(&h, dtype, static_cast<int>(kernel_), degree_, gamma, static_cast<double>(coef0_), static_cast<double>(cost_), 0);
    if (status != B200SVM_OK) {
        throw backend_exception{ fmt::format("B200 backend error {}: {}", status, b200svm_last_error(nullptr)) };
    }
    handle_ = h;
}

template <typename T>
csvm<T>::~csvm() {
    // must not throw (src/plssvm/backends/CUDA/csvm.cu:82-92)
    if (handle_ != nullptr) { b200svm_destroy(as_handle(handle_)); }
}

template <typename T>
void csvm<T>::setup_data_on_device() {
    PLSSVM_ASSERT(data_ptr_ != nullptr, "No data is provided!");
    const std::vector<std::vector<real_type>> &data = *data_ptr_;
    const std::size_t n = data.size(), F = num_features_;
    std::size_t nnz = 0, longest = 0;
    for (const std::vector<real_type> &row : data) {
        std::size_t c = 0;
        for (const real_type v : row) { c += v != real_type{ 0 }; }
        nnz += c;
        longest = std::max(longest, c);
    }
    sparse_rows_ = static_cast<double>(nnz) <= sparse_density_threshold * static_cast<double>(n) * static_cast<double>(F) && longest <= sparse_max_row;
    if (print_info_) {
        fmt::print("B200 backend: {} x {} data set, {:.3g} % non-zero, longest row {} entries -> stored {} on the device.\n", n, F,
                   100.0 * static_cast<double>(nnz) / std::max(1.0, static_cast<double>(n) * static_cast<double>(F)), longest, sparse_rows_ ? "as CSR rows" : "dense");
    }
    if (sparse_rows_) {
        std::vector<std::int64_t> rowptr(n + 1, 0);
        std::vector<std::int32_t> colidx;
        std::vector<real_type> values;
        colidx.reserve(nnz);
        values.reserve(nnz);
        for (std::size_t i = 0; i < n; ++i) {
            for (std::size_t f = 0; f < F; ++f) {
                if (data[i][f] != real_type{ 0 }) {
                    colidx.push_back(static_cast<std::int32_t>(f));
                    values.push_back(data[i][f]);
                }
            }
            rowptr[i + 1] = static_cast<std::int64_t>(colidx.size());
        }
        check(b200svm_set_data_csr(as_handle(handle_), rowptr.data(), colidx.data(), values.data(), static_cast<std::int64_t>(n), static_cast<std::int64_t>(F),
                                   static_cast<std::int64_t>(nnz), std::is_same_v<T, float> ? B200SVM_VAL_F32 : B200SVM_VAL_F64));
    } else {
        std::vector<real_type> flat(n * F);
        for (std::size_t i = 0; i < n; ++i) { std::copy(data[i].begin(), data[i].end(), flat.begin() + static_cast<std::ptrdiff_t>(i * F)); }
        check(b200svm_set_data_dense(as_handle(handle_), flat.data(), static_cast<std::int64_t>(n), static_cast<std::int64_t>(F), static_cast<std::int64_t>(F)));
    }
    data_on_device_ = true;
}

template <typename T>
void csvm<T>::ensure_data() {
    if (!data_on_device_) { setup_data_on_device(); }
}

template <typename T>
auto csvm<T>::generate_q() -> std::vector<real_type> {
    ensure_data();
    std::vector<real_type> q(num_data_points_ - 1);
    check(b200svm_generate_q(as_handle(handle_), q.data()));
    return q;
}

template <typename T>
auto csvm<T>::solver_CG(const std::vector<real_type> &b, const std::size_t imax, const real_type eps, const std::vector<real_type> &q) -> std::vector<real_type> {
    ensure_data();
    std::vector<real_type> alpha(b.size());
    std::int64_t iterations = 0;
    double delta = 0.0;
    const auto start_time = std::chrono::steady_clock::now();
    check(b200svm_cg_solve(as_handle(handle_), b.data(), q.data(), static_cast<double>(QA_cost_), static_cast<std::int64_t>(imax), static_cast<double>(eps),
                           alpha.data(), &iterations, &delta));
    last_iterations_ = static_cast<std::size_t>(iterations);
    if (print_info_) {
        const auto dur = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - start_time);
        fmt::print("Finished after {} iterations with a residuum of {} and an average iteration time of {}.\n", iterations, delta,
                   dur / std::max<std::int64_t>(iterations, 1));
    }
    return alpha;
}

template <typename T>
void csvm<T>::update_w() {
    ensure_data();
    PLSSVM_ASSERT(alpha_ptr_ != nullptr, "No alphas provided!");
    w_.resize(num_features_);
    check(b200svm_update_w(as_handle(handle_), alpha_ptr_->data(), w_.data()));
}

template <typename T>
auto csvm<T>::predict(const std::vector<std::vector<real_type>> &points) -> std::vector<real_type> {
    PLSSVM_ASSERT(data_ptr_ != nullptr, "No data is provided!");
    PLSSVM_ASSERT(!data_ptr_->empty(), "Data set is empty!");
    if (points.empty()) { return std::vector<real_type>{}; }
    // same checks and texts as src/plssvm/backends/OpenMP/csvm.cpp:203-209
    if (!std::all_of(points.begin(), points.end(), [&](const std::vector<real_type> &point) { return point.size() == points.front().size(); })) {
        throw exception{ "All points in the prediction point vector must have the same number of features!" };
    } else if (points.front().size() != data_ptr_->front().size()) {
        throw exception{ fmt::format("Number of features per data point ({}) must match the number of features per predict point ({})!",
                                     data_ptr_->front().size(), points.front().size()) };
    } else if (alpha_ptr_ == nullptr) {
        throw exception{ "No alphas provided for prediction!" };
    }
    ensure_data();
    const auto start_time = std::chrono::steady_clock::now();
    const std::size_t P = points.size(), F = points.front().size();
    std::vector<real_type> out(P);
    const double rho = -static_cast<double>(bias_);  // model files carry rho = -bias (src/plssvm/csvm.cpp:42,123)
    if (sparse_rows_) {
        std::vector<std::int64_t> rowptr(P + 1, 0);
        std::vector<std::int32_t> colidx;
        std::vector<real_type> values;
        for (std::size_t p = 0; p < P; ++p) {
            for (std::size_t f = 0; f < F; ++f) {
                if (points[p][f] != real_type{ 0 }) {
                    colidx.push_back(static_cast<std::int32_t>(f));
                    values.push_back(points[p][f]);
                }
            }
            rowptr[p + 1] = static_cast<std::int64_t>(colidx.size());
        }
        check(b200svm_predict_csr(as_handle(handle_), alpha_ptr_->data(), rho, rowptr.data(), colidx.data(), values.data(), static_cast<std::int64_t>(P),
                                  rowptr[P], std::is_same_v<T, float> ? B200SVM_VAL_F32 : B200SVM_VAL_F64, out.data()));
    } else {
        std::vector<real_type> flat(P * F);
        for (std::size_t p = 0; p < P; ++p) { std::copy(points[p].begin(), points[p].end(), flat.begin() + static_cast<std::ptrdiff_t>(p * F)); }
        check(b200svm_predict_dense(as_handle(handle_), alpha_ptr_->data(), rho, flat.data(), static_cast<std::int64_t>(P), static_cast<std::int64_t>(F), out.data()));
    }
    if (print_info_) {
        fmt::print("Predicted {} data points in {}.\n", points.size(),
                   std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - start_time));
    }
    return out;
}

template class csvm<float>;
template class csvm<double>;

}  // namespace plssvm::b200
