// TEST / BASELINE INFRASTRUCTURE — not product code.  Only tests/ and bench.py's baseline legs may load the library built from
// this file.
//
// C-ABI driver around the UNMODIFIED reference CUDA backend ("the kernels to beat"), compiled by
// `make -C oracle ref_cuda` with nvcc for sm_100 from the sources where they lie under /root/reference:
//   src/plssvm/backends/CUDA/{csvm,svm_kernel,q_kernel,predict_kernel}.cu, detail/{device_ptr,utility}.cu, backends/gpu_csvm.cpp
// Protected members are exposed the way the reference's own test mock does (tests/backends/CUDA/mock_cuda_csvm.hpp).
//   orefcuda_solver_cg : plssvm::detail::gpu_csvm::solver_CG  (src/plssvm/backends/gpu_csvm.cpp:185-324: kernel on the device,
//                        CG vector algebra on the host, 4-5 PCIe copies per iteration)
//   orefcuda_matvec_ms : one run_svm_kernel launch (device_kernel_{linear,poly,radial}, src/plssvm/backends/CUDA/svm_kernel.cu)
//                        timed with CUDA events, data resident
#include "plssvm/backends/CUDA/csvm.hpp"
#include "plssvm/backends/CUDA/detail/device_ptr.cuh"
#include "plssvm/csvm.hpp"
#include "plssvm/detail/execution_range.hpp"
#include "plssvm/kernel_types.hpp"
#include "plssvm/parameter.hpp"

#include <cuda_runtime.h>
#include <unistd.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace {

template <typename T>
class exposed_cuda_csvm : public plssvm::cuda::csvm<T> {
    using base_type = plssvm::cuda::csvm<T>;

  public:
    using device_ptr_type = plssvm::cuda::detail::device_ptr<T>;
    explicit exposed_cuda_csvm(const plssvm::parameter<T> &params) : base_type{ params } {}
    using base_type::generate_q;
    using base_type::run_device_kernel;
    using base_type::setup_data_on_device;
    using base_type::solver_CG;
    void set_QA_cost(const T v) { base_type::QA_cost_ = v; }
    std::size_t num_devices() const { return base_type::devices_.size(); }
    std::size_t dept() const { return base_type::dept_; }
    std::size_t boundary() const { return base_type::boundary_size_; }
};

template <typename T>
std::shared_ptr<const std::vector<std::vector<T>>> make_data(const void *X, long n, long F) {
    const T *x = static_cast<const T *>(X);
    auto v = std::make_shared<std::vector<std::vector<T>>>(static_cast<std::size_t>(n));
    for (long i = 0; i < n; ++i) { (*v)[i].assign(x + i * F, x + (i + 1) * F); }
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
    p.backend = plssvm::backend_type::cuda;
    p.target = plssvm::target_platform::gpu_nvidia;
    return p;
}

struct stdout_capture {
    int saved_fd{ -1 };
    char path[64]{};
    stdout_capture() {
        std::fflush(stdout);
        std::snprintf(path, sizeof(path), "/tmp/orefcuda_capture_%d_XXXXXX", static_cast<int>(getpid()));
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
int do_solver_cg(int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F, const void *b_in, const void *q_in,
                 double QA_cost, long imax, double eps, void *alpha_out, long *iters, double *seconds, double *setup_seconds) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, cost, eps, true);
    p.data_ptr = make_data<T>(X, n, F);
    const long m = n - 1;
    const std::vector<T> b(static_cast<const T *>(b_in), static_cast<const T *>(b_in) + m);
    const std::vector<T> q(static_cast<const T *>(q_in), static_cast<const T *>(q_in) + m);
    stdout_capture cap;
    std::vector<T> alpha;
    try {
        exposed_cuda_csvm<T> svm{ p };
        const auto s0 = std::chrono::steady_clock::now();
        svm.setup_data_on_device();
        cudaDeviceSynchronize();
        const auto s1 = std::chrono::steady_clock::now();
        svm.set_QA_cost(static_cast<T>(QA_cost));
        const auto t0 = std::chrono::steady_clock::now();
        alpha = svm.solver_CG(b, static_cast<std::size_t>(imax), static_cast<T>(eps), q);
        cudaDeviceSynchronize();
        const auto t1 = std::chrono::steady_clock::now();
        if (seconds != nullptr) { *seconds = std::chrono::duration<double>(t1 - t0).count(); }
        if (setup_seconds != nullptr) { *setup_seconds = std::chrono::duration<double>(s1 - s0).count(); }
    } catch (const std::exception &e) {
        const std::string log = cap.finish();
        std::fprintf(stderr, "orefcuda_solver_cg: %s\n", e.what());
        return 2;
    }
    const std::string log = cap.finish();
    if (iters != nullptr) { *iters = parse_iterations(log); }
    std::memcpy(alpha_out, alpha.data(), alpha.size() * sizeof(T));
    return 0;
}

template <typename T>
int do_matvec_ms(int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F, const void *q_in, const void *d_in,
                 double QA_cost, double add, void *ret_inout, int reps, double *ms_out) {
    auto p = make_params<T>(kernel, degree, gamma, coef0, cost, 1e-3, false);
    p.data_ptr = make_data<T>(X, n, F);
    const long m = n - 1;
    try {
        exposed_cuda_csvm<T> svm{ p };
        svm.setup_data_on_device();
        svm.set_QA_cost(static_cast<T>(QA_cost));
        using dptr = typename exposed_cuda_csvm<T>::device_ptr_type;
        const std::size_t len = svm.dept() + svm.boundary();
        // the linear kernel splits the FEATURES over all visible devices (src/plssvm/backends/gpu_csvm.cpp:129-157); every device adds
        // its share into its own copy of ret, the copies are summed on the host like gpu_csvm::device_reduction (:365-386)
        const std::size_t nd = svm.num_devices();
        std::vector<dptr> q_d(nd), r_d(nd), x_d(nd);
        for (std::size_t dev = 0; dev < nd; ++dev) {
            q_d[dev] = dptr{ len, static_cast<int>(dev) };
            r_d[dev] = dptr{ len, static_cast<int>(dev) };
            x_d[dev] = dptr{ len, static_cast<int>(dev) };
            q_d[dev].memset(0);
            r_d[dev].memset(0);
            x_d[dev].memset(0);
            q_d[dev].memcpy_to_device(static_cast<const T *>(q_in), 0, m);
            x_d[dev].memcpy_to_device(static_cast<const T *>(d_in), 0, m);
        }
        r_d[0].memcpy_to_device(static_cast<const T *>(ret_inout), 0, m);
        for (std::size_t dev = 0; dev < nd; ++dev) { svm.run_device_kernel(dev, q_d[dev], r_d[dev], x_d[dev], static_cast<T>(add)); }  // the call whose result is returned
        std::vector<T> acc(static_cast<std::size_t>(m), T(0)), part(static_cast<std::size_t>(m));
        for (std::size_t dev = 0; dev < nd; ++dev) {
            cudaSetDevice(static_cast<int>(dev));
            cudaDeviceSynchronize();
            r_d[dev].memcpy_to_host(part.data(), 0, m);
            for (long i = 0; i < m; ++i) { acc[static_cast<std::size_t>(i)] += part[static_cast<std::size_t>(i)]; }
        }
        std::memcpy(ret_inout, acc.data(), static_cast<std::size_t>(m) * sizeof(T));
        cudaSetDevice(0);
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        double best = 1e30, sum = 0;
        for (int k = 0; k < reps; ++k) {  // timing: device 0's share (the whole kernel for polynomial / rbf, which the reference runs on one device)
            cudaEventRecord(e0, 0);
            svm.run_device_kernel(0, q_d[0], r_d[0], x_d[0], static_cast<T>(add));
            cudaEventRecord(e1, 0);
            cudaEventSynchronize(e1);
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            best = ms < best ? ms : best;
            sum += ms;
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        if (ms_out != nullptr) {
            ms_out[0] = reps > 0 ? sum / reps : 0.0;
            ms_out[1] = reps > 0 ? best : 0.0;
        }
    } catch (const std::exception &e) {
        std::fprintf(stderr, "orefcuda_matvec_ms: %s\n", e.what());
        return 2;
    }
    return 0;
}

}  // namespace

extern "C" {

int orefcuda_solver_cg(int dtype, int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F, const void *b, const void *q,
                       double QA_cost, long imax, double eps, void *alpha_out, long *iters, double *seconds, double *setup_seconds) {
    return dtype == 0 ? do_solver_cg<float>(kernel, degree, gamma, coef0, cost, X, n, F, b, q, QA_cost, imax, eps, alpha_out, iters, seconds, setup_seconds)
                      : do_solver_cg<double>(kernel, degree, gamma, coef0, cost, X, n, F, b, q, QA_cost, imax, eps, alpha_out, iters, seconds, setup_seconds);
}

int orefcuda_matvec_ms(int dtype, int kernel, int degree, double gamma, double coef0, double cost, const void *X, long n, long F, const void *q, const void *d,
                       double QA_cost, double add, void *ret_inout, int reps, double *ms_out) {
    return dtype == 0 ? do_matvec_ms<float>(kernel, degree, gamma, coef0, cost, X, n, F, q, d, QA_cost, add, ret_inout, reps, ms_out)
                      : do_matvec_ms<double>(kernel, degree, gamma, coef0, cost, X, n, F, q, d, QA_cost, add, ret_inout, reps, ms_out);
}

}  // extern "C"
