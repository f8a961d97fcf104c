// plssvm-train (reference: src/plssvm/main_train.cpp:27-85): parse the command line, print the task summary, learn, write the model.
// Precision: double unless built with -DPLSSVM_EXECUTABLES_USE_SINGLE_PRECISION (same switch as the reference, src/main_train.cpp:21-25)
// or started with --fp22 / env PLSSVM_SINGLE_PRECISION=1.
#include "plssvm/core.hpp"

#include <cstdlib>
#include <cstring>
#include <iostream>

template <typename real_type>
int run(int argc, char **argv) {
    plssvm::parameter_train<real_type> params{ argc, argv };
    if (params.print_info) {
        std::cout << "\ntask: training\nkernel type: " << params.kernel << " -> ";
        switch (params.kernel) {
            case plssvm::kernel_type::linear: std::cout << "u'*v\n"; break;
            case plssvm::kernel_type::polynomial:
                std::cout << "(gamma*u'*v + coef0)^degree\ngamma: " << plssvm::detail::format_shortest(params.gamma)
                          << "\ncoef0: " << plssvm::detail::format_shortest(params.coef0) << "\ndegree: " << params.degree << "\n";
                break;
            case plssvm::kernel_type::rbf: std::cout << "exp(-gamma*|u-v|^2)\ngamma: " << plssvm::detail::format_shortest(params.gamma) << "\n"; break;
        }
        std::cout << "cost: " << plssvm::detail::format_shortest(params.cost) << "\nepsilon: " << plssvm::detail::format_shortest(params.epsilon)
                  << "\ninput file (data set): '" << params.input_filename << "'\noutput file (model): '" << params.model_filename << "'\n\n";
    }
    auto svm = plssvm::make_csvm(params);
    svm->learn();
    svm->write_model(params.model_filename);
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
