/**
 * Stand-alone C++ host for libb200svm.so with PLSSVM's API surface (reference: include/plssvm/core.hpp:14-29 and the headers it
 * pulls in): plssvm::parameter / parameter_train / parameter_predict, plssvm::csvm, plssvm::make_csvm, kernel / backend / target
 * enums, the exception types, the LIBSVM / ARFF data formats and the LIBSVM model format.  Needs no reference sources.
 *
 * Differences by design: data parsed from files is kept as CSR (the reference densifies, src/plssvm/parameter.cpp:41-116);
 * the dense `data_ptr` view of the reference API is materialised as well while it is small.  There is exactly one backend (b200):
 * backend_type values other than automatic / b200 / cuda are rejected — no multi-backend dispatch, no CPU fallback.
 */
#ifndef PLSSVM_B200_HOST_CORE_HPP_
#define PLSSVM_B200_HOST_CORE_HPP_

#include <cstddef>
#include <cstdint>
#include <iosfwd>
#include <memory>
#include <stdexcept>
#include <string>
#include <string_view>
#include <vector>

namespace plssvm {

// ------------------------------------------------------------------------------------------------------------------
// exceptions (reference: include/plssvm/exceptions/exceptions.hpp:26-118) — same names, what() texts asserted by the reference's tests
// ------------------------------------------------------------------------------------------------------------------
class exception : public std::runtime_error {
  public:
    explicit exception(const std::string &msg, std::string_view class_name = "exception") :
        std::runtime_error{ msg }, class_name_{ class_name } {}
    [[nodiscard]] std::string what_with_loc() const { return std::string{ what() } + "\nException type: " + class_name_; }

  private:
    std::string class_name_;
};
#define PLSSVM_B200_EXCEPTION(name)                                                    \
    class name : public exception {                                                    \
      public:                                                                          \
        explicit name(const std::string &msg) : exception{ msg, #name } {}            \
    }
PLSSVM_B200_EXCEPTION(invalid_file_format_exception);
PLSSVM_B200_EXCEPTION(file_not_found_exception);
PLSSVM_B200_EXCEPTION(unsupported_backend_exception);
PLSSVM_B200_EXCEPTION(unsupported_kernel_type_exception);
namespace b200 {
PLSSVM_B200_EXCEPTION(backend_exception);
}
#undef PLSSVM_B200_EXCEPTION

// ------------------------------------------------------------------------------------------------------------------
// enums (reference: include/plssvm/kernel_types.hpp:27-34, backend_types.hpp:22-35, target_platforms.hpp:22-33)
// ------------------------------------------------------------------------------------------------------------------
enum class kernel_type { linear = 0, polynomial = 1, rbf = 2 };
enum class backend_type { automatic, openmp, cuda, hip, opencl, sycl, b200 };
enum class target_platform { automatic, cpu, gpu_nvidia, gpu_amd, gpu_intel };

std::ostream &operator<<(std::ostream &out, kernel_type kernel);
std::istream &operator>>(std::istream &in, kernel_type &kernel);  // "linear"/"0", "polynomial"/"poly"/"1", "rbf"/"2" (case-insensitive)
std::ostream &operator<<(std::ostream &out, backend_type backend);
std::istream &operator>>(std::istream &in, backend_type &backend);
std::ostream &operator<<(std::ostream &out, target_platform target);
std::istream &operator>>(std::istream &in, target_platform &target);

/// value storage of sparse rows on the device
enum class value_storage { native, fp22 };

// ------------------------------------------------------------------------------------------------------------------
// sparse rows
// ------------------------------------------------------------------------------------------------------------------
template <typename T>
struct csr_matrix {
    std::size_t rows{ 0 }, cols{ 0 };
    std::vector<std::int64_t> rowptr{ 0 };
    std::vector<std::int32_t> col;
    std::vector<T> val;

    [[nodiscard]] std::size_t nnz() const noexcept { return col.size(); }
    [[nodiscard]] std::size_t longest_row() const noexcept;
    [[nodiscard]] std::vector<std::vector<T>> to_dense() const;
    static csr_matrix from_dense(const std::vector<std::vector<T>> &data);
};

// ------------------------------------------------------------------------------------------------------------------
// parameter (reference: include/plssvm/parameter.hpp:36-235) — public data members with the reference's names and defaults
// ------------------------------------------------------------------------------------------------------------------
template <typename T>
class parameter {
  public:
    using real_type = T;
    virtual ~parameter() = default;

    void parse_libsvm_file(const std::string &filename, bool test = false);
    void parse_arff_file(const std::string &filename, bool test = false);
    void parse_model_file(const std::string &filename);
    void parse_train_file(const std::string &filename);  // .arff -> ARFF else LIBSVM; labels required
    void parse_test_file(const std::string &filename);

    kernel_type kernel = kernel_type::linear;
    int degree = 3;
    real_type gamma = real_type{ 0.0 };  // 0 -> 1 / num_features once the data is known (src/plssvm/parameter.cpp:150-152)
    real_type coef0 = real_type{ 0.0 };
    real_type cost = real_type{ 1.0 };
    real_type epsilon = static_cast<real_type>(0.001);
    bool print_info = true;
    backend_type backend = backend_type::automatic;
    target_platform target = target_platform::automatic;
    value_storage storage = value_storage::native;  // extension: --fp22

    std::string input_filename{};
    std::string model_filename{};
    std::string predict_filename{};

    std::shared_ptr<const std::vector<std::vector<real_type>>> data_ptr{};       // dense view (reference API); may be empty for large sparse files
    std::shared_ptr<const std::vector<std::vector<real_type>>> test_data_ptr{};
    std::shared_ptr<const csr_matrix<real_type>> sparse_data_ptr{};              // what the parsers produce
    std::shared_ptr<const csr_matrix<real_type>> sparse_test_data_ptr{};
    std::shared_ptr<const std::vector<real_type>> value_ptr{};
    std::shared_ptr<const std::vector<real_type>> alpha_ptr{};
    real_type rho = real_type{ 0.0 };

  protected:
    [[nodiscard]] std::string model_name_from_input() const;
    [[nodiscard]] std::string predict_name_from_input() const;
};

template <typename T>
std::ostream &operator<<(std::ostream &out, const parameter<T> &params);

/// plssvm-train command line (reference: src/plssvm/parameter_train.cpp:38-142)
template <typename T>
class parameter_train : public parameter<T> {
  public:
    explicit parameter_train(std::string input_filename);
    parameter_train(int argc, char **argv);
};
/// plssvm-predict command line (reference: src/plssvm/parameter_predict.cpp:38-115)
template <typename T>
class parameter_predict : public parameter<T> {
  public:
    parameter_predict(std::string input_filename, std::string model_filename);
    parameter_predict(int argc, char **argv);
};

// ------------------------------------------------------------------------------------------------------------------
// csvm (reference: include/plssvm/csvm.hpp:33-278) — public interface kept; the five backend virtuals are the C-ABI calls
// ------------------------------------------------------------------------------------------------------------------
template <typename T>
class csvm {
  public:
    using real_type = T;
    explicit csvm(const parameter<T> &params);
    ~csvm();
    csvm(const csvm &) = delete;
    csvm &operator=(const csvm &) = delete;

    void write_model(const std::string &filename);
    void learn();
    [[nodiscard]] real_type accuracy();
    [[nodiscard]] real_type accuracy(const std::vector<std::vector<real_type>> &points, const std::vector<real_type> &correct_labels);
    [[nodiscard]] real_type predict(const std::vector<real_type> &point);
    [[nodiscard]] std::vector<real_type> predict(const std::vector<std::vector<real_type>> &points);
    [[nodiscard]] std::vector<real_type> predict(const csr_matrix<real_type> &points);
    [[nodiscard]] real_type predict_label(const std::vector<real_type> &point);
    [[nodiscard]] std::vector<real_type> predict_label(const std::vector<std::vector<real_type>> &points);
    [[nodiscard]] std::vector<real_type> predict_label(const csr_matrix<real_type> &points);

    [[nodiscard]] std::size_t last_cg_iterations() const noexcept { return iterations_; }
    [[nodiscard]] real_type bias() const noexcept { return bias_; }
    [[nodiscard]] const std::vector<real_type> &alpha() const { return *alpha_ptr_; }
    [[nodiscard]] bool uses_sparse_rows() const noexcept { return sparse_rows_; }

  private:
    void check(int status) const;
    void setup_data_on_device();
    void check_points(std::size_t features) const;

    const target_platform target_;
    const kernel_type kernel_;
    const int degree_;
    real_type gamma_;
    const real_type coef0_;
    real_type cost_;
    const real_type epsilon_;
    const bool print_info_;
    const value_storage storage_;
    std::shared_ptr<const std::vector<std::vector<real_type>>> data_ptr_;
    std::shared_ptr<const csr_matrix<real_type>> sparse_ptr_;
    std::shared_ptr<const std::vector<real_type>> value_ptr_;
    std::shared_ptr<const std::vector<real_type>> alpha_ptr_;
    std::size_t num_data_points_{ 0 }, num_features_{ 0 };
    real_type bias_{ 0 };
    void *handle_{ nullptr };
    bool data_on_device_{ false }, sparse_rows_{ false };
    std::size_t iterations_{ 0 };
};

/// reference: include/plssvm/csvm_factory.hpp:53-111 — one backend, so this only validates the request
template <typename T>
[[nodiscard]] std::unique_ptr<csvm<T>> make_csvm(const parameter<T> &params);

namespace detail {
/// text of a floating point value as fmt's "{}" prints it (shortest round-trip digits; exponent form below 1e-4 / from 1e16)
std::string format_shortest(double v);
std::string format_shortest(float v);
/// LIBSVM model text (src/plssvm/csvm.cpp:60-204): header, then "{alpha} {index}:{value:e} ..." per support vector, zeros skipped, all rows of
/// class +1 before those of class -1.  Exactly one of `sparse` / `dense` is used (`sparse` when not null).
template <typename T>
void write_model_file(const std::string &model_name, kernel_type kernel, int degree, T gamma, T coef0, T rho, const std::vector<T> &labels,
                      const std::vector<T> &alphas, const csr_matrix<T> *sparse, const std::vector<std::vector<T>> *dense, std::size_t num_features,
                      bool print_info);
/// FP22: binary32 rounded to nearest-even at 13 mantissa bits
std::uint32_t fp22_encode(float x);
float fp22_decode(std::uint32_t v22);
}  // namespace detail

}  // namespace plssvm

#endif  // PLSSVM_B200_HOST_CORE_HPP_
