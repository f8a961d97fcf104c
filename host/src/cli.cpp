// Command lines of plssvm-train / plssvm-predict (reference: src/plssvm/parameter_train.cpp:38-142, parameter_predict.cpp:38-115).
// cxxopts is not available offline, so the same short / long flags and positionals are parsed by hand:
//   train:   -t/--kernel_type -d/--degree -g/--gamma -r/--coef0 -c/--cost -e/--epsilon -b/--backend -p/--target_platform -q/--quiet -h/--help
//            training_set_file [model_file]
//   predict: -b/--backend -p/--target_platform -q/--quiet -h/--help   test_file model_file [output_file]
// cxxopts also accepts the positionals by name (--input/--model, --test/--model/--output; parameter_train.cpp:60-61,
// parameter_predict.cpp:54-56); a named value wins over the positional slot it fills, the remaining positionals fill the free slots in order.
// Extension of this build: --fp22 (store sparse feature values as FP22 on the device; single precision only).
#include "plssvm/core.hpp"

#include <cstdlib>
#include <iostream>
#include <map>
#include <sstream>

namespace plssvm {

namespace {

struct option_spec {
    char short_name;
    const char *long_name;
    bool takes_value;
};

struct parsed_args {
    std::map<std::string, std::string> values;
    std::vector<std::string> positionals;
};

[[noreturn]] void fail_usage(const std::string &msg, const std::string &help) {
    std::cout << msg << "\n" << help;
    std::exit(EXIT_FAILURE);
}

parsed_args parse_args(int argc, char **argv, const std::vector<option_spec> &specs, const std::string &help) {
    parsed_args res;
    const auto find_long = [&](const std::string &name) -> const option_spec * {
        for (const option_spec &s : specs) {
            if (name == s.long_name) { return &s; }
        }
        return nullptr;
    };
    const auto find_short = [&](const char c) -> const option_spec * {
        for (const option_spec &s : specs) {
            if (s.short_name != 0 && c == s.short_name) { return &s; }
        }
        return nullptr;
    };
    bool only_positionals = false;
    for (int i = 1; i < argc; ++i) {
        const std::string arg = argv[i];
        if (only_positionals || arg.size() < 2 || arg[0] != '-' || (arg[1] != '-' && (std::isdigit((unsigned char) arg[1]) || arg[1] == '.'))) {
            res.positionals.push_back(arg);
            continue;
        }
        if (arg == "--") {
            only_positionals = true;
            continue;
        }
        const option_spec *spec = nullptr;
        std::string value;
        bool has_value = false;
        if (arg[1] == '-') {
            const auto eq = arg.find('=');
            const std::string name = arg.substr(2, eq == std::string::npos ? std::string::npos : eq - 2);
            spec = find_long(name);
            if (spec == nullptr) { fail_usage("Option '" + name + "' does not exist", help); }
            if (eq != std::string::npos) {
                value = arg.substr(eq + 1);
                has_value = true;
            }
        } else {
            spec = find_short(arg[1]);
            if (spec == nullptr) { fail_usage(std::string{ "Option '" } + arg[1] + "' does not exist", help); }
            if (arg.size() > 2) {
                value = arg.substr(2);
                has_value = true;
            }
        }
        if (spec->takes_value) {
            if (!has_value) {
                if (i + 1 >= argc) { fail_usage(std::string{ "Option '" } + spec->long_name + "' is missing an argument", help); }
                value = argv[++i];
            }
            res.values[spec->long_name] = value;
        } else {
            res.values[spec->long_name] = "true";
        }
    }
    return res;
}

// Slots of the positional list: a slot given by name is taken from the option, the others consume the bare arguments in order.
std::vector<std::string> fill_slots(const parsed_args &args, const std::vector<const char *> &names) {
    std::vector<std::string> slots(names.size());
    std::size_t next = 0;
    for (std::size_t k = 0; k < names.size(); ++k) {
        const auto it = args.values.find(names[k]);
        if (it != args.values.end()) {
            slots[k] = it->second;
        } else if (next < args.positionals.size()) {
            slots[k] = args.positionals[next++];
        }
    }
    return slots;
}

template <typename V>
V value_as(const parsed_args &args, const char *name, const V fallback, const std::string &help) {
    const auto it = args.values.find(name);
    if (it == args.values.end()) { return fallback; }
    std::istringstream iss{ it->second };
    V v{};
    iss >> v;
    if (iss.fail() || !(iss >> std::ws).eof()) { fail_usage(std::string{ "Argument '" } + it->second + "' failed to parse for option '" + name + "'", help); }
    return v;
}

const char *train_help =
    "LS-SVM on NVIDIA B200 (libb200svm)\nUsage:\n  plssvm-train [OPTION...] training_set_file [model_file]\n\n"
    "  -t, --kernel_type arg      set type of kernel function.\n"
    "                             \t 0 -- linear: u'*v\n"
    "                             \t 1 -- polynomial: (gamma*u'*v + coef0)^degree\n"
    "                             \t 2 -- radial basis function: exp(-gamma*|u-v|^2) (default: 0)\n"
    "  -d, --degree arg           set degree in kernel function (default: 3)\n"
    "  -g, --gamma arg            set gamma in kernel function (default: 1 / num_features)\n"
    "  -r, --coef0 arg            set coef0 in kernel function (default: 0)\n"
    "  -c, --cost arg             set the parameter C (default: 1)\n"
    "  -e, --epsilon arg          set the tolerance of termination criterion (default: 0.001)\n"
    "  -b, --backend arg          choose the backend: automatic|b200 (default: automatic)\n"
    "  -p, --target_platform arg  choose the target platform: automatic|gpu_nvidia (default: automatic)\n"
    "      --fp22                 store sparse feature values as FP22 on the device (single precision only)\n"
    "  -q, --quiet                quiet mode (no outputs)\n"
    "  -h, --help                 print this helper message\n";

const char *predict_help =
    "LS-SVM on NVIDIA B200 (libb200svm)\nUsage:\n  plssvm-predict [OPTION...] test_file model_file [output_file]\n\n"
    "  -b, --backend arg          choose the backend: automatic|b200 (default: automatic)\n"
    "  -p, --target_platform arg  choose the target platform: automatic|gpu_nvidia (default: automatic)\n"
    "      --fp22                 store sparse feature values as FP22 on the device (single precision only)\n"
    "  -q, --quiet                quiet mode (no outputs)\n"
    "  -h, --help                 print this helper message\n";

}  // namespace

template <typename T>
parameter_train<T>::parameter_train(std::string input_filename) {
    this->parse_train_file(input_filename);
}

template <typename T>
parameter_train<T>::parameter_train(int argc, char **argv) {
    const std::vector<option_spec> specs = { { 't', "kernel_type", true }, { 'd', "degree", true }, { 'g', "gamma", true }, { 'r', "coef0", true },
                                             { 'c', "cost", true }, { 'e', "epsilon", true }, { 'b', "backend", true }, { 'p', "target_platform", true },
                                             { 0, "fp22", false }, { 'q', "quiet", false }, { 'h', "help", false }, { 0, "input", true }, { 0, "model", true } };
    const parsed_args args = parse_args(argc, argv, specs, train_help);
    if (args.values.count("help")) {
        std::cout << train_help;
        std::exit(EXIT_SUCCESS);
    }
    this->kernel = value_as(args, "kernel_type", this->kernel, train_help);
    this->degree = value_as(args, "degree", this->degree, train_help);
    if (args.values.count("gamma")) {
        this->gamma = value_as(args, "gamma", this->gamma, train_help);
        if (this->gamma == T{ 0 }) {
            std::cerr << "gamma = 0.0 is not allowed, it doesnt make any sense!\n";  // parameter_train.cpp:89-93
            std::cout << train_help;
            std::exit(EXIT_FAILURE);
        }
    }
    this->coef0 = value_as(args, "coef0", this->coef0, train_help);
    this->cost = value_as(args, "cost", this->cost, train_help);
    this->epsilon = value_as(args, "epsilon", this->epsilon, train_help);
    this->backend = value_as(args, "backend", this->backend, train_help);
    this->target = value_as(args, "target_platform", this->target, train_help);
    this->storage = args.values.count("fp22") ? value_storage::fp22 : value_storage::native;
    this->print_info = args.values.count("quiet") == 0;
    const std::vector<std::string> slots = fill_slots(args, { "input", "model" });
    if (slots[0].empty()) {
        std::cerr << "Error missing input file!";  // parameter_train.cpp:127-131
        std::cout << train_help;
        std::exit(EXIT_FAILURE);
    }
    this->input_filename = slots[0];
    this->model_filename = !slots[1].empty() ? slots[1] : this->model_name_from_input();
    this->parse_train_file(this->input_filename);
}

template <typename T>
parameter_predict<T>::parameter_predict(std::string input_filename, std::string model_filename) {
    this->parse_model_file(model_filename);
    this->parse_test_file(input_filename);
    this->predict_filename = this->predict_name_from_input();
}

template <typename T>
parameter_predict<T>::parameter_predict(int argc, char **argv) {
    const std::vector<option_spec> specs = { { 'b', "backend", true }, { 'p', "target_platform", true }, { 0, "fp22", false }, { 'q', "quiet", false },
                                             { 'h', "help", false }, { 0, "test", true }, { 0, "model", true }, { 0, "output", true } };
    const parsed_args args = parse_args(argc, argv, specs, predict_help);
    if (args.values.count("help")) {
        std::cout << predict_help;
        std::exit(EXIT_SUCCESS);
    }
    this->backend = value_as(args, "backend", this->backend, predict_help);
    this->target = value_as(args, "target_platform", this->target, predict_help);
    this->storage = args.values.count("fp22") ? value_storage::fp22 : value_storage::native;
    this->print_info = args.values.count("quiet") == 0;
    const std::vector<std::string> slots = fill_slots(args, { "test", "model", "output" });
    if (slots[0].empty()) {
        std::cerr << "Error missing test file!";  // parameter_predict.cpp:88-92
        std::cout << predict_help;
        std::exit(EXIT_FAILURE);
    }
    if (slots[1].empty()) {
        std::cerr << "Error missing model file!";  // parameter_predict.cpp:96-100
        std::cout << predict_help;
        std::exit(EXIT_FAILURE);
    }
    this->parse_model_file(slots[1]);
    this->parse_test_file(slots[0]);
    this->predict_filename = !slots[2].empty() ? slots[2] : this->predict_name_from_input();
}

template class parameter_train<float>;
template class parameter_train<double>;
template class parameter_predict<float>;
template class parameter_predict<double>;

}  // namespace plssvm
