// plssvm-predict (reference: src/main_predict.cpp:31-115): parse test + model files, predict labels, write them joined by '\n'
// (no trailing newline, src/main_predict.cpp:80-81), print the accuracy line when the test file carries labels (:92-105).
#include "plssvm/core.hpp"

#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>

template <typename real_type>
int run(int argc, char **argv) {
    plssvm::parameter_predict<real_type> params{ argc, argv };
    if (params.print_info) {
        std::cout << "\ntask: prediction\nkernel type: " << params.kernel << " -> ";
        switch (params.kernel) {
            case plssvm::kernel_type::linear: std::cout << "u'*v\n"; break;
            case plssvm::kernel_type::polynomial:
                std::cout << "(gamma*u'*v + coef0)^degree\ngamma: " << plssvm::detail::format_shortest(params.gamma)
                          << "\ncoef0: " << plssvm::detail::format_shortest(params.coef0) << "\ndegree: " << params.degree << "\n";
                break;
            case plssvm::kernel_type::rbf: std::cout << "exp(-gamma*|u-v|^2)\ngamma: " << plssvm::detail::format_shortest(params.gamma) << "\n"; break;
        }
        std::cout << "rho: " << plssvm::detail::format_shortest(params.rho) << "\ninput file (data set): '" << params.input_filename
                  << "'\ninput file (model): '" << params.model_filename << "'\noutput file (prediction): '" << params.predict_filename << "'\n\n";
    }
    // the test labels (if any) were parsed last into value_ptr; the model's own labels are not needed for prediction
    const auto test_labels = params.value_ptr;
    auto svm = plssvm::make_csvm(params);
    const std::vector<real_type> labels = svm->predict_label(*params.sparse_test_data_ptr);
    {
        std::ofstream out{ params.predict_filename };
        for (std::size_t i = 0; i < labels.size(); ++i) { out << (i ? "\n" : "") << plssvm::detail::format_shortest(labels[i]); }
    }
    if (params.print_info) { std::cout << "Wrote prediction file ('" << params.predict_filename << "') with " << labels.size() << " labels.\n"; }
    if (test_labels != nullptr) {
        unsigned long long correct = 0;
        for (std::size_t i = 0; i < labels.size(); ++i) { correct += (*test_labels)[i] * labels[i] > real_type{ 0 }; }
        std::cout << "Accuracy = " << plssvm::detail::format_shortest(static_cast<real_type>(correct) / static_cast<real_type>(labels.size()) * real_type{ 100 })
                  << "% (" << correct << "/" << labels.size() << ") (classification)\n";
    }
    return EXIT_SUCCESS;
}

int main(int argc, char *argv[]) {
    try {
        bool single = false;
#if defined(PLSSVM_EXECUTABLES_USE_SINGLE_PRECISION)
        single = true;
#endif
        const char *env = std::getenv("PLSSVM_SINGLE_PRECISION");
        if (env != nullptr && std::strcmp(env, "0") != 0) { single = true; }
        for (int i = 1; i < argc; ++i) { single = single || std::strcmp(argv[i], "--fp22") == 0; }
        return single ? run<float>(argc, argv) : run<double>(argc, argv);
    } catch (const plssvm::exception &e) {
        std::cerr << e.what_with_loc() << std::endl;
        return EXIT_FAILURE;
    } catch (const std::exception &e) {
        std::cerr << e.what() << std::endl;
        return EXIT_FAILURE;
    }
}
