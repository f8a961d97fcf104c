// plssvm::csvm<T> of the stand-alone host: the reference's public interface (include/plssvm/csvm.hpp:33-278) on top of the C ABI.
// learn() keeps the whole numeric path of src/plssvm/csvm.cpp:228-258 on the device (b200svm_learn); write_model() reproduces the
// LIBSVM model layout of src/plssvm/csvm.cpp:60-204 (header order, "{alpha} " then "{index}:{value:e} " skipping zeros, all +1 rows
// before the -1 rows).
#include "plssvm/core.hpp"

#include "b200svm.h"

#include <omp.h>

#include <algorithm>
#include <charconv>
#include <chrono>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>
#include <type_traits>

namespace plssvm {

namespace {

b200svm_handle as_handle(void *p) { return static_cast<b200svm_handle>(p); }
constexpr double sparse_density_threshold = 0.05;
constexpr std::size_t sparse_max_row = 1024;

template <typename T>
T sign(const T x) { return x > T{ 0 } ? T{ +1 } : T{ -1 }; }

template <typename E>
std::string to_text(const E e) {
    std::ostringstream ss;
    ss << e;
    return ss.str();
}

// FP22-packed CSR of float rows: (fp22 << 10) | column_delta, padding entries (value +0, delta 1023) bridge gaps > 1023
template <typename T>
void pack_fp22(const csr_matrix<T> &m, std::vector<std::int64_t> &rowptr, std::vector<std::uint32_t> &packed) {
    rowptr.assign(m.rows + 1, 0);
    packed.clear();
    packed.reserve(m.nnz());
    for (std::size_t i = 0; i < m.rows; ++i) {
        std::int64_t prev = 0;
        for (std::int64_t p = m.rowptr[i]; p < m.rowptr[i + 1]; ++p) {
            std::int64_t delta = m.col[(std::size_t) p] - prev;
            while (delta > 1023) {
                packed.push_back(1023u);
                delta -= 1023;
            }
            packed.push_back((detail::fp22_encode((float) m.val[(std::size_t) p]) << 10) | (std::uint32_t) delta);
            prev = m.col[(std::size_t) p];
        }
        rowptr[i + 1] = (std::int64_t) packed.size();
    }
}

}  // namespace

template <typename T>
void csvm<T>::check(const int status) const {
    if (status != B200SVM_OK) {
        throw b200::backend_exception{ "B200 backend error " + std::to_string(status) + ": " + b200svm_last_error(as_handle(handle_)) };
    }
}

template <typename T>
csvm<T>::csvm(const parameter<T> &params) :
    target_{ params.target }, kernel_{ params.kernel }, degree_{ params.degree }, gamma_{ params.gamma }, coef0_{ params.coef0 }, cost_{ params.cost },
    epsilon_{ params.epsilon }, print_info_{ params.print_info }, storage_{ params.storage }, data_ptr_{ params.data_ptr },
    sparse_ptr_{ params.sparse_data_ptr }, value_ptr_{ params.value_ptr }, alpha_ptr_{ params.alpha_ptr }, bias_{ -params.rho } {
    // same checks and texts as the reference constructor (src/plssvm/csvm.cpp:39-57)
    if (data_ptr_ == nullptr && sparse_ptr_ == nullptr) { throw exception{ "No data points provided!" }; }
    if (sparse_ptr_ == nullptr) {
        if (data_ptr_->empty()) { throw exception{ "Data set is empty!" }; }
        if (!std::all_of(data_ptr_->begin(), data_ptr_->end(), [&](const std::vector<T> &p) { return p.size() == data_ptr_->front().size(); })) {
            throw exception{ "All points in the data vector must have the same number of features!" };
        }
        if (data_ptr_->front().empty()) { throw exception{ "No features provided for the data points!" }; }
        num_data_points_ = data_ptr_->size();
        num_features_ = data_ptr_->front().size();
    } else {
        if (sparse_ptr_->rows == 0) { throw exception{ "Data set is empty!" }; }
        if (sparse_ptr_->cols == 0) { throw exception{ "No features provided for the data points!" }; }
        num_data_points_ = sparse_ptr_->rows;
        num_features_ = sparse_ptr_->cols;
    }
    if (alpha_ptr_ != nullptr && alpha_ptr_->size() != num_data_points_) {
        throw exception{ "Number of weights (" + std::to_string(alpha_ptr_->size()) + ") must match the number of data points (" +
                         std::to_string(num_data_points_) + ")!" };
    }
    // (no label-count check here: plssvm-predict hands over the TEST labels next to the model's support vectors, src/main_predict.cpp:92-105)
    if (target_ != target_platform::automatic && target_ != target_platform::gpu_nvidia) {
        throw b200::backend_exception{ "Invalid target platform '" + to_text(target_) + "' for the B200 backend!" };
    }
    if (gamma_ == T{ 0 }) { gamma_ = T{ 1 } / static_cast<T>(num_features_); }  // src/plssvm/parameter.cpp:150-152
    if (storage_ == value_storage::fp22 && !std::is_same_v<T, float>) { throw exception{ "FP22 value storage needs single precision (float)!" }; }
    if (print_info_) { std::cout << "Using B200 (sm_100a, libb200svm " << b200svm_version() << ") as backend.\n\n"; }
    b200svm_handle h = nullptr;
    const int status = b200svm_create(&h, std::is_same_v<T, float> ? B200SVM_F32 : B200SVM_F64, static_cast<int>(kernel_), degree_, (double) gamma_,
                                      (double) coef0_, (double) cost_, 0);
    if (status != B200SVM_OK) { throw b200::backend_exception{ "B200 backend error " + std::to_string(status) + ": " + b200svm_last_error(nullptr) }; }
    handle_ = h;
}

template <typename T>
csvm<T>::~csvm() {
    if (handle_ != nullptr) { b200svm_destroy(as_handle(handle_)); }
}

template <typename T>
void csvm<T>::setup_data_on_device() {
    if (data_on_device_) { return; }
    // sparse rows when few entries are non-zero (the reference only knows dense rows) — or always for FP22 storage
    std::shared_ptr<const csr_matrix<T>> sp = sparse_ptr_;
    if (sp == nullptr) { sp = std::make_shared<const csr_matrix<T>>(csr_matrix<T>::from_dense(*data_ptr_)); }
    const double density = (double) sp->nnz() / ((double) sp->rows * (double) sp->cols);
    const bool fits = sp->longest_row() <= sparse_max_row;
    sparse_rows_ = fits && (storage_ == value_storage::fp22 || density <= sparse_density_threshold);
    if (storage_ == value_storage::fp22 && !fits) { throw exception{ "FP22 value storage needs rows of at most 1024 stored entries!" }; }
    if (sparse_rows_) {
        sparse_ptr_ = sp;
        if (storage_ == value_storage::fp22) {
            std::vector<std::int64_t> rowptr;
            std::vector<std::uint32_t> packed;
            pack_fp22(*sp, rowptr, packed);
            check(b200svm_set_data_csr(as_handle(handle_), rowptr.data(), nullptr, packed.data(), (std::int64_t) sp->rows, (std::int64_t) sp->cols,
                                       rowptr.back(), B200SVM_VAL_FP22));
        } else {
            check(b200svm_set_data_csr(as_handle(handle_), sp->rowptr.data(), sp->col.data(), sp->val.data(), (std::int64_t) sp->rows, (std::int64_t) sp->cols,
                                       (std::int64_t) sp->nnz(), std::is_same_v<T, float> ? B200SVM_VAL_F32 : B200SVM_VAL_F64));
        }
    } else {
        std::vector<T> flat(num_data_points_ * num_features_, T{ 0 });
        for (std::size_t i = 0; i < sp->rows; ++i) {
            for (std::int64_t p = sp->rowptr[i]; p < sp->rowptr[i + 1]; ++p) { flat[i * num_features_ + (std::size_t) sp->col[(std::size_t) p]] = sp->val[(std::size_t) p]; }
        }
        check(b200svm_set_data_dense(as_handle(handle_), flat.data(), (std::int64_t) num_data_points_, (std::int64_t) num_features_, (std::int64_t) num_features_));
    }
    data_on_device_ = true;
}

template <typename T>
void csvm<T>::learn() {
    if (value_ptr_ == nullptr) { throw exception{ "No labels given for training! Maybe the data is only usable for prediction?" }; }  // csvm.cpp:214
    if (value_ptr_->size() != num_data_points_) {
        throw exception{ "Number of labels (" + std::to_string(value_ptr_->size()) + ") must match the number of data points (" +
                         std::to_string(num_data_points_) + ")!" };
    }
    const auto t0 = std::chrono::steady_clock::now();
    setup_data_on_device();
    const auto t1 = std::chrono::steady_clock::now();
    if (print_info_) {
        std::cout << "Setup for solving the optimization problem done in " << std::chrono::duration_cast<std::chrono::milliseconds>(t1 - t0).count() << "ms.\n";
    }
    std::vector<T> alpha(num_data_points_);
    double rho = 0.0;
    std::int64_t iterations = 0;
    check(b200svm_learn(as_handle(handle_), value_ptr_->data(), (double) epsilon_, alpha.data(), &rho, &iterations));
    iterations_ = (std::size_t) iterations;
    bias_ = static_cast<T>(-rho);
    alpha_ptr_ = std::make_shared<const std::vector<T>>(std::move(alpha));
    if (print_info_) {
        std::cout << "Finished after " << iterations << " iterations.\nSolved minimization problem (r = b - Ax) using CG in "
                  << std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t1).count() << "ms.\n";
    }
}

template <typename T>
void csvm<T>::check_points(const std::size_t features) const {
    if (features != num_features_) {
        throw exception{ "Number of features per data point (" + std::to_string(num_features_) + ") must match the number of features per predict point (" +
                         std::to_string(features) + ")!" };  // src/plssvm/backends/OpenMP/csvm.cpp:205-206
    } else if (alpha_ptr_ == nullptr) {
        throw exception{ "No alphas provided for prediction!" };  // :207-208
    }
}

template <typename T>
auto csvm<T>::predict(const csr_matrix<real_type> &points) -> std::vector<real_type> {
    if (points.rows == 0) { return {}; }
    check_points(points.cols);
    setup_data_on_device();
    std::vector<T> out(points.rows);
    const double rho = -(double) bias_;
    if (sparse_rows_) {
        if (storage_ == value_storage::fp22) {
            std::vector<std::int64_t> rowptr;
            std::vector<std::uint32_t> packed;
            pack_fp22(points, rowptr, packed);
            check(b200svm_predict_csr(as_handle(handle_), alpha_ptr_->data(), rho, rowptr.data(), nullptr, packed.data(), (std::int64_t) points.rows, rowptr.back(),
                                      B200SVM_VAL_FP22, out.data()));
        } else {
            check(b200svm_predict_csr(as_handle(handle_), alpha_ptr_->data(), rho, points.rowptr.data(), points.col.data(), points.val.data(),
                                      (std::int64_t) points.rows, (std::int64_t) points.nnz(), std::is_same_v<T, float> ? B200SVM_VAL_F32 : B200SVM_VAL_F64,
                                      out.data()));
        }
    } else {
        // dense support vectors: stage the points densely in chunks of ~256 MiB
        const std::size_t chunk = std::max<std::size_t>(1, (std::size_t{ 1 } << 28) / (num_features_ * sizeof(T)));
        std::vector<T> flat;
        for (std::size_t p0 = 0; p0 < points.rows; p0 += chunk) {
            const std::size_t rows = std::min(chunk, points.rows - p0);
            flat.assign(rows * num_features_, T{ 0 });
            for (std::size_t i = 0; i < rows; ++i) {
                for (std::int64_t p = points.rowptr[p0 + i]; p < points.rowptr[p0 + i + 1]; ++p) {
                    flat[i * num_features_ + (std::size_t) points.col[(std::size_t) p]] = points.val[(std::size_t) p];
                }
            }
            check(b200svm_predict_dense(as_handle(handle_), alpha_ptr_->data(), rho, flat.data(), (std::int64_t) rows, (std::int64_t) num_features_, out.data() + p0));
        }
    }
    return out;
}

template <typename T>
auto csvm<T>::predict(const std::vector<std::vector<real_type>> &points) -> std::vector<real_type> {
    if (points.empty()) { return {}; }
    if (!std::all_of(points.begin(), points.end(), [&](const std::vector<T> &p) { return p.size() == points.front().size(); })) {
        throw exception{ "All points in the prediction point vector must have the same number of features!" };  // OpenMP/csvm.cpp:203-204
    }
    check_points(points.front().size());
    setup_data_on_device();
    if (sparse_rows_) { return predict(csr_matrix<T>::from_dense(points)); }
    const std::size_t P = points.size(), F = num_features_;
    std::vector<T> flat(P * F), out(P);
    for (std::size_t p = 0; p < P; ++p) { std::copy(points[p].begin(), points[p].end(), flat.begin() + (std::ptrdiff_t) (p * F)); }
    check(b200svm_predict_dense(as_handle(handle_), alpha_ptr_->data(), -(double) bias_, flat.data(), (std::int64_t) P, (std::int64_t) F, out.data()));
    return out;
}

template <typename T>
auto csvm<T>::predict(const std::vector<real_type> &point) -> real_type {
    if (point.size() != num_features_) {
        throw exception{ "Number of features per data point (" + std::to_string(num_features_) + ") must match the number of features of the predict point (" +
                         std::to_string(point.size()) + ")!" };  // src/plssvm/csvm.cpp:322-324
    }
    return predict(std::vector<std::vector<T>>(1, point))[0];
}

template <typename T>
auto csvm<T>::predict_label(const std::vector<real_type> &point) -> real_type { return sign(predict(point)); }

template <typename T>
auto csvm<T>::predict_label(const std::vector<std::vector<real_type>> &points) -> std::vector<real_type> {
    std::vector<T> classes = predict(points);
    for (T &c : classes) { c = sign(c); }  // operators::sign: x > 0 ? +1 : -1
    return classes;
}

template <typename T>
auto csvm<T>::predict_label(const csr_matrix<real_type> &points) -> std::vector<real_type> {
    std::vector<T> classes = predict(points);
    for (T &c : classes) { c = sign(c); }
    return classes;
}

template <typename T>
auto csvm<T>::accuracy(const std::vector<std::vector<real_type>> &points, const std::vector<real_type> &correct_labels) -> real_type {
    if (points.size() != correct_labels.size()) {
        throw exception{ "Number of data points (" + std::to_string(points.size()) + ") must match number of correct labels (" +
                         std::to_string(correct_labels.size()) + ")!" };  // src/plssvm/csvm.cpp:291-293
    }
    if (points.empty()) { return T{ 0 }; }
    const std::vector<T> predictions = predict(points);
    unsigned long long correct = 0;
    for (std::size_t i = 0; i < predictions.size(); ++i) { correct += predictions[i] * correct_labels[i] > T{ 0 }; }
    return static_cast<T>(correct) / static_cast<T>(points.size());
}

template <typename T>
auto csvm<T>::accuracy() -> real_type {
    if (value_ptr_ == nullptr) { throw exception{ "No labels given! Maybe the data is only usable for prediction?" }; }
    const std::vector<T> predictions = sparse_ptr_ != nullptr ? predict(*sparse_ptr_) : predict(*data_ptr_);
    unsigned long long correct = 0;
    for (std::size_t i = 0; i < predictions.size(); ++i) { correct += predictions[i] * (*value_ptr_)[i] > T{ 0 }; }
    return static_cast<T>(correct) / static_cast<T>(predictions.size());
}

template <typename T>
void csvm<T>::write_model(const std::string &model_name) {
    if (alpha_ptr_ == nullptr) {
        throw exception{ "No alphas given! Maybe a call to 'learn()' is missing?" };
    } else if (value_ptr_ == nullptr) {
        throw exception{ "No labels given! Maybe the data is only usable for prediction?" };
    } else if (value_ptr_->size() != num_data_points_) {
        throw exception{ "Number of labels (" + std::to_string(value_ptr_->size()) + ") must match the number of data points (" +
                         std::to_string(num_data_points_) + ")!" };  // src/plssvm/csvm.cpp:69-71
    }
    detail::write_model_file<T>(model_name, kernel_, degree_, gamma_, coef0_, static_cast<T>(-bias_), *value_ptr_, *alpha_ptr_, sparse_ptr_.get(),
                                data_ptr_.get(), num_features_, print_info_);
}

namespace detail {

// The text of a LIBSVM model file as src/plssvm/csvm.cpp:60-204 writes it.
template <typename T>
void write_model_file(const std::string &model_name, const kernel_type kernel, const int degree, const T gamma, const T coef0, const T rho,
                      const std::vector<T> &labels, const std::vector<T> &alphas, const csr_matrix<T> *sparse,
                      const std::vector<std::vector<T>> *dense, const std::size_t num_features, const bool print_info) {
    unsigned long long count_pos = 0, count_neg = 0;
    for (const T v : labels) {
        count_pos += v > 0;
        count_neg += v < 0;
    }
    std::ostringstream header;
    header << "svm_type c_svc\nkernel_type " << kernel << "\n";
    switch (kernel) {
        case kernel_type::linear: break;
        case kernel_type::polynomial:
            header << "degree " << degree << "\ngamma " << detail::format_shortest(gamma) << "\ncoef0 " << detail::format_shortest(coef0) << "\n";
            break;
        case kernel_type::rbf: header << "gamma " << detail::format_shortest(gamma) << "\n"; break;
    }
    header << "nr_class 2\ntotal_sv " << count_pos + count_neg << "\nrho " << detail::format_shortest(rho) << "\nlabel 1 -1\nnr_sv " << count_pos
           << " " << count_neg << "\nSV\n";
    if (print_info) { std::cout << "\nOptimization finished\n" << header.str() << "\n"; }
    std::FILE *out = std::fopen(model_name.c_str(), "wb");
    if (out == nullptr) { throw exception{ "Can't open model file '" + model_name + "' for writing!" }; }
    std::fputs(header.str().c_str(), out);
    // All support vectors of class 1 first (src/plssvm/csvm.cpp:160-172), then class -1 (:175-195), in file order inside a class.
    // Batches of rows are formatted by an OpenMP team in contiguous chunks and written in chunk order (the reference formats inside
    // `omp parallel` as well); "{index}:{value:e} " is produced with std::to_chars(scientific, 6), which prints what "%e" prints.
    std::vector<std::size_t> order;
    order.reserve(labels.size());
    for (const bool positive : { true, false }) {
        for (std::size_t i = 0; i < labels.size(); ++i) {
            const T y = labels[i];
            if (positive ? y > 0 : y < 0) { order.push_back(i); }
        }
    }
    const auto format_row = [&](const std::size_t i, std::string &dst) {
        char buf[64];
        dst += detail::format_shortest(alphas[i]);
        dst.push_back(' ');
        const auto emit = [&](const std::size_t f, const T v) {
            if (v != T{ 0 }) {
                char *p = std::to_chars(buf, buf + sizeof(buf), f).ptr;
                *p++ = ':';
                p = std::to_chars(p, buf + sizeof(buf) - 1, (double) v, std::chars_format::scientific, 6).ptr;
                *p++ = ' ';
                dst.append(buf, p);
            }
        };
        if (sparse != nullptr) {
            for (std::int64_t p = sparse->rowptr[i]; p < sparse->rowptr[i + 1]; ++p) {
                emit((std::size_t) sparse->col[(std::size_t) p], sparse->val[(std::size_t) p]);
            }
        } else {
            for (std::size_t f = 0; f < num_features; ++f) { emit(f, (*dense)[i][f]); }
        }
        dst.push_back('\n');
    };
    constexpr std::size_t batch = std::size_t{ 1 } << 15;
    std::vector<std::string> parts;
    for (std::size_t b0 = 0; b0 < order.size(); b0 += batch) {
        const std::size_t nb = std::min(batch, order.size() - b0);
        const int teams = (int) std::max<std::size_t>(1, std::min<std::size_t>((std::size_t) omp_get_max_threads(), nb / 64 + 1));
        parts.assign((std::size_t) teams, std::string{});
#pragma omp parallel num_threads(teams)
        {
            const std::size_t t = (std::size_t) omp_get_thread_num(), nt = (std::size_t) omp_get_num_threads();
            std::string &dst = parts[t];
            for (std::size_t k = b0 + nb * t / nt; k < b0 + nb * (t + 1) / nt; ++k) { format_row(order[k], dst); }
        }
        for (const std::string &part : parts) {
            if (!part.empty() && std::fwrite(part.data(), 1, part.size(), out) != part.size()) {
                std::fclose(out);
                throw exception{ "Can't write to model file '" + model_name + "'!" };
            }
        }
    }
    std::fclose(out);
    if (print_info) { std::cout << "Wrote model file ('" << model_name << "') with " << count_pos + count_neg << " support vectors.\n"; }
}

template void write_model_file<float>(const std::string &, kernel_type, int, float, float, float, const std::vector<float> &, const std::vector<float> &,
                                      const csr_matrix<float> *, const std::vector<std::vector<float>> *, std::size_t, bool);
template void write_model_file<double>(const std::string &, kernel_type, int, double, double, double, const std::vector<double> &,
                                       const std::vector<double> &, const csr_matrix<double> *, const std::vector<std::vector<double>> *, std::size_t, bool);

}  // namespace detail

template <typename T>
std::unique_ptr<csvm<T>> make_csvm(const parameter<T> &params) {
    switch (params.backend) {
        case backend_type::automatic:
        case backend_type::b200:
        case backend_type::cuda:  // `-b cuda` of existing scripts lands on the one NVIDIA backend this build has
            return std::make_unique<csvm<T>>(params);
        default: break;
    }
    throw unsupported_backend_exception{ "No " + to_text(params.backend) + " backend available!" };  // include/plssvm/csvm_factory.hpp:60-108
}

template class csvm<float>;
template class csvm<double>;
template std::unique_ptr<csvm<float>> make_csvm(const parameter<float> &);
template std::unique_ptr<csvm<double>> make_csvm(const parameter<double> &);

}  // namespace plssvm
