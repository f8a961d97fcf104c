/**
 * B200 backend for PLSSVM: a subclass of the reference's own plssvm::csvm<T> (include/plssvm/csvm.hpp:33-278) whose five
 * protected virtuals call libb200svm.so through its C ABI (include/b200svm.h).  This header and src/plssvm/backends/B200/csvm.cpp
 * are the ONLY files a maintainer adds to the reference tree; they compile against the reference's unmodified
 * headers.  Layout mirrors the existing backends (include/plssvm/backends/OpenMP/csvm.hpp, include/plssvm/backends/CUDA/csvm.hpp).
 */
#ifndef PLSSVM_BACKENDS_B200_CSVM_HPP_
#define PLSSVM_BACKENDS_B200_CSVM_HPP_

#include "plssvm/csvm.hpp"                   // plssvm::csvm
#include "plssvm/exceptions/exceptions.hpp"  // plssvm::exception
#include "plssvm/exceptions/source_location.hpp"
#include "plssvm/parameter.hpp"  // plssvm::parameter

#include <cstddef>  // std::size_t
#include <string>   // std::string
#include <vector>   // std::vector

namespace plssvm::b200 {

/// Exception type thrown if a problem with the B200 backend occurs (same shape as plssvm::cuda::backend_exception,
/// include/plssvm/backends/CUDA/exceptions.hpp).
class backend_exception : public exception {
  public:
    explicit backend_exception(const std::string &msg, source_location loc = source_location::current());
};

/**
 * @brief A C-SVM implementation for NVIDIA B200 (sm_100a): every numeric step runs in libb200svm.so, nothing falls back to the CPU.
 * @tparam T the type of the data, `float` or `double`
 */
template <typename T>
class csvm : public ::plssvm::csvm<T> {
  protected:
    using base_type = ::plssvm::csvm<T>;
    using base_type::alpha_ptr_;
    using base_type::bias_;
    using base_type::coef0_;
    using base_type::cost_;
    using base_type::data_ptr_;
    using base_type::degree_;
    using base_type::gamma_;
    using base_type::kernel_;
    using base_type::num_data_points_;
    using base_type::num_features_;
    using base_type::print_info_;
    using base_type::QA_cost_;
    using base_type::target_;
    using base_type::w_;

  public:
    using base_type::predict;
    using real_type = typename base_type::real_type;

    /**
     * @throws plssvm::csvm::csvm() exceptions
     * @throws plssvm::b200::backend_exception if the target platform isn't automatic or gpu_nvidia, or no sm_100 device is usable
     */
    explicit csvm(const parameter<T> &params);
    ~csvm() override;

    csvm(const csvm &) = delete;
    csvm &operator=(const csvm &) = delete;

    /// @copydoc plssvm::csvm::predict(const std::vector<std::vector<real_type>>&)
    [[nodiscard]] std::vector<real_type> predict(const std::vector<std::vector<real_type>> &points) override;

    /// CG iterations executed by the last solver_CG call (the value the reference only prints, OpenMP/csvm.cpp:162-167).
    [[nodiscard]] std::size_t last_cg_iterations() const noexcept { return last_iterations_; }
    /// Rows stored sparsely on the device (CSR) instead of dense?  Chosen in setup_data_on_device from the zero fraction.
    [[nodiscard]] bool uses_sparse_rows() const noexcept { return sparse_rows_; }

  protected:
    void setup_data_on_device() override;
    [[nodiscard]] std::vector<real_type> generate_q() override;
    std::vector<real_type> solver_CG(const std::vector<real_type> &b, std::size_t imax, real_type eps, const std::vector<real_type> &q) override;
    void update_w() override;

  private:
    void check(int status) const;  // status code of the C ABI -> backend_exception with the library's message
    void ensure_data();            // predict-only objects (model files) never see setup_data_on_device()

    void *handle_{ nullptr };  // b200svm_handle
    bool data_on_device_{ false };
    bool sparse_rows_{ false };
    std::size_t last_iterations_{ 0 };
};

extern template class csvm<float>;
extern template class csvm<double>;

}  // namespace plssvm::b200

#endif  // PLSSVM_BACKENDS_B200_CSVM_HPP_
