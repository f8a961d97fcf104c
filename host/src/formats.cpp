// Enums, sparse rows and the file formats of the stand-alone host: LIBSVM / ARFF data files and LIBSVM model files.
// Behaviour follows the reference parsers (src/plssvm/parameter.cpp:41-116 LIBSVM lines, :180-363 ARFF, :366-520 model header,
// src/plssvm/detail/file_reader.cpp:140-151 comment / blank-line handling) including their quirks: feature indices are used as
// written (0-based), the feature count is the largest index + 1, labels go through sign(), a line whose first token contains ':'
// has no label.  Unlike the reference, rows are parsed straight into CSR.
#include "plssvm/core.hpp"

#include <omp.h>

#include <exception>
#include <algorithm>
#include <cctype>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <limits>
#include <sstream>

namespace plssvm {

namespace {

std::string lower(std::string s) {
    std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char) std::tolower(c); });
    return s;
}
std::string upper(std::string s) {
    std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char) std::toupper(c); });
    return s;
}
std::string_view trim_left(std::string_view s) {
    const auto p = s.find_first_not_of(" \t\v\r\f");
    return p == std::string_view::npos ? std::string_view{} : s.substr(p);
}
std::string_view trim(std::string_view s) {
    s = trim_left(s);
    const auto p = s.find_last_not_of(" \t\v\r\f\n");
    return p == std::string_view::npos ? std::string_view{} : s.substr(0, p + 1);
}
bool starts_with(std::string_view s, std::string_view p) { return s.substr(0, p.size()) == p; }
bool ends_with(std::string_view s, std::string_view p) { return s.size() >= p.size() && s.substr(s.size() - p.size()) == p; }

// the reference's detail::convert_to (include/plssvm/detail/string_conversion.hpp:39-63): leading blanks skipped, the longest
// numeric prefix is converted (trailing characters are ignored, as from_chars does), message quotes the left-trimmed text
template <typename V, typename E = std::runtime_error>
V convert_to(std::string_view token) {
    const std::string_view t = trim_left(token);
    V v{};
    const auto res = std::from_chars(t.data(), t.data() + t.size(), v);
    if (res.ec != std::errc{}) {
        throw E{ std::string{ "Can't convert '" } + std::string{ t } + "' to a value of type " +
                 (std::is_floating_point_v<V> ? (sizeof(V) == 4 ? "float" : "double") : "unsigned long") + "!" };
    }
    return v;
}

template <typename T>
T sign(const T x) { return x > T{ 0 } ? T{ +1 } : T{ -1 }; }  // include/plssvm/detail/operators.hpp:174-177

// lines of a text file without blank lines and lines starting with the comment character (file_reader.cpp:140-151)
struct text_file {
    std::string content;
    std::vector<std::string_view> lines;
    text_file(const std::string &filename, const char comment) {
        std::ifstream in{ filename, std::ios::binary };
        if (!in) { throw file_not_found_exception{ "Couldn't find file: '" + filename + "'!" }; }
        in.seekg(0, std::ios::end);
        const std::streamoff size = in.tellg();
        in.seekg(0, std::ios::beg);
        content.resize(size > 0 ? (std::size_t) size : 0);
        if (size > 0) { in.read(content.data(), size); }
        std::string_view rest{ content };
        while (!rest.empty()) {
            const auto nl = rest.find('\n');
            std::string_view line = trim(rest.substr(0, nl));
            if (!line.empty() && line.front() != comment) { lines.push_back(line); }
            if (nl == std::string_view::npos) { break; }
            rest.remove_prefix(nl + 1);
        }
    }
};

// One LIBSVM line -> (label / alpha, sorted unique non-zero entries appended to col / val).  Returns the row width (largest index + 1).
template <typename T>
std::size_t parse_libsvm_line(std::string_view line, T &first, bool &has_first, std::vector<std::pair<std::int32_t, T>> &row,
                              std::vector<std::int32_t> &col, std::vector<T> &val) {
    std::string_view::size_type pos = line.find_first_of(" \n");
    const std::string_view::size_type first_colon = line.find_first_of(":\n");
    if (first_colon >= pos) {
        first = convert_to<T, invalid_file_format_exception>(line.substr(0, pos));
    } else {
        has_first = false;  // any unlabeled line marks the file as unlabeled (the reference sets values[0], parameter.cpp:62)
        pos = 0;
    }
    row.clear();
    std::size_t row_size = 0;
    while (pos != std::string_view::npos) {
        const auto colon = line.find_first_of(':', pos);
        if (colon == std::string_view::npos) { break; }
        const auto index = convert_to<unsigned long, invalid_file_format_exception>(line.substr(pos, colon - pos));
        pos = colon + 1;
        const auto next = line.find_first_of(' ', pos);
        const T value = convert_to<T, invalid_file_format_exception>(line.substr(pos, next == std::string_view::npos ? next : next - pos));
        if (index > (unsigned long) std::numeric_limits<std::int32_t>::max() - 1) { throw invalid_file_format_exception{ "Feature index too large!" }; }
        row_size = std::max<std::size_t>(row_size, index + 1);
        row.emplace_back((std::int32_t) index, value);
        pos = next;
    }
    // later duplicates overwrite earlier ones (dense semantics of the reference: vline[index] = value); zeros are not stored
    std::stable_sort(row.begin(), row.end(), [](const auto &a, const auto &b) { return a.first < b.first; });
    for (std::size_t k = 0; k < row.size(); ++k) {
        if (k + 1 < row.size() && row[k + 1].first == row[k].first) { continue; }
        if (row[k].second != T{ 0 }) {
            col.push_back(row[k].first);
            val.push_back(row[k].second);
        }
    }
    return row_size;
}

// Parses lines [start, end) of LIBSVM text into CSR; first[i] = label / alpha of line i.  Returns false if the lines carry no label.
// The lines are parsed by an OpenMP team in contiguous chunks (the reference parses with `omp parallel for`, parameter.cpp:55-99) and
// stitched in order; of several malformed lines the first one in file order is reported, as a sequential pass would.
template <typename T>
bool parse_libsvm_lines(const text_file &f, const std::size_t start, csr_matrix<T> &m, std::vector<T> &first) {
    const std::size_t n = f.lines.size() - start;
    m = csr_matrix<T>{};
    m.rows = n;
    m.rowptr.assign(n + 1, 0);
    first.assign(n, T{ 0 });
    const int teams = (int) std::max<std::size_t>(1, std::min<std::size_t>((std::size_t) omp_get_max_threads(), n / 256 + 1));
    struct chunk {
        std::vector<std::int32_t> col;
        std::vector<T> val;
        std::size_t max_size{ 0 };
        bool has_first{ true };
        std::exception_ptr error{};
    };
    std::vector<chunk> chunks((std::size_t) teams);
#pragma omp parallel num_threads(teams)
    {
        const std::size_t t = (std::size_t) omp_get_thread_num();
        const std::size_t nt = (std::size_t) omp_get_num_threads();
        if (t < chunks.size()) {
            // contiguous line ranges; with fewer threads than chunks a thread takes several
            for (std::size_t ci = t; ci < chunks.size(); ci += nt) {
                chunk &ck = chunks[ci];
                const std::size_t lo = n * ci / chunks.size(), hi = n * (ci + 1) / chunks.size();
                std::vector<std::pair<std::int32_t, T>> row;
                if (hi > lo) {  // one allocation per chunk: an "index:value " token is at least 4 characters long
                    const std::string_view a = f.lines[lo + start], b = f.lines[hi - 1 + start];
                    const std::size_t bytes = (std::size_t) ((b.data() + b.size()) - a.data());
                    ck.col.reserve(bytes / 4);
                    ck.val.reserve(bytes / 4);
                }
                try {
                    for (std::size_t i = lo; i < hi; ++i) {
                        const std::size_t before = ck.col.size();
                        ck.max_size = std::max(ck.max_size, parse_libsvm_line<T>(f.lines[i + start], first[i], ck.has_first, row, ck.col, ck.val));
                        m.rowptr[i + 1] = (std::int64_t) (ck.col.size() - before);  // entries of the line; turned into offsets below
                    }
                } catch (...) { ck.error = std::current_exception(); }
            }
        }
    }
    for (const chunk &ck : chunks) {  // file order: the first malformed line wins
        if (ck.error) { std::rethrow_exception(ck.error); }
    }
    bool has_first = true;
    std::size_t max_size = 0, total = 0;
    std::vector<std::size_t> offset(chunks.size());
    for (std::size_t ci = 0; ci < chunks.size(); ++ci) {
        has_first = has_first && chunks[ci].has_first;
        max_size = std::max(max_size, chunks[ci].max_size);
        offset[ci] = total;
        total += chunks[ci].col.size();
    }
    if (max_size == 0) { throw invalid_file_format_exception{ "Can't parse file: no data points are given!" }; }
    for (std::size_t i = 0; i < n; ++i) { m.rowptr[i + 1] += m.rowptr[i]; }
    m.col.resize(total);
    m.val.resize(total);
#pragma omp parallel for num_threads(teams) schedule(static, 1)
    for (std::size_t ci = 0; ci < chunks.size(); ++ci) {
        std::copy(chunks[ci].col.begin(), chunks[ci].col.end(), m.col.begin() + (std::ptrdiff_t) offset[ci]);
        std::copy(chunks[ci].val.begin(), chunks[ci].val.end(), m.val.begin() + (std::ptrdiff_t) offset[ci]);
    }
    m.cols = max_size;
    return has_first;
}

constexpr std::size_t dense_view_limit_bytes = std::size_t{ 64 } << 20;  // the vector-of-vectors view of the reference API costs more than the parse itself

template <typename T>
std::shared_ptr<const std::vector<std::vector<T>>> dense_view(const csr_matrix<T> &m) {
    if (m.rows * m.cols * sizeof(T) > dense_view_limit_bytes) { return nullptr; }
    return std::make_shared<const std::vector<std::vector<T>>>(m.to_dense());
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// enums
// ---------------------------------------------------------------------------------------------------------------------
std::ostream &operator<<(std::ostream &out, const kernel_type kernel) {
    switch (kernel) {
        case kernel_type::linear: return out << "linear";
        case kernel_type::polynomial: return out << "polynomial";
        case kernel_type::rbf: return out << "rbf";
    }
    return out << "unknown";
}
std::istream &operator>>(std::istream &in, kernel_type &kernel) {
    std::string s;
    in >> s;
    s = lower(s);
    if (s == "linear" || s == "0") {
        kernel = kernel_type::linear;
    } else if (s == "polynomial" || s == "poly" || s == "1") {
        kernel = kernel_type::polynomial;
    } else if (s == "rbf" || s == "2") {
        kernel = kernel_type::rbf;
    } else {
        in.setstate(std::ios::failbit);
    }
    return in;
}
std::ostream &operator<<(std::ostream &out, const backend_type backend) {
    static const char *names[] = { "automatic", "openmp", "cuda", "hip", "opencl", "sycl", "b200" };
    return out << names[(int) backend];
}
std::istream &operator>>(std::istream &in, backend_type &backend) {
    std::string s;
    in >> s;
    s = lower(s);
    static const char *names[] = { "automatic", "openmp", "cuda", "hip", "opencl", "sycl", "b200" };
    for (int k = 0; k < 7; ++k) {
        if (s == names[k]) {
            backend = (backend_type) k;
            return in;
        }
    }
    in.setstate(std::ios::failbit);
    return in;
}
std::ostream &operator<<(std::ostream &out, const target_platform target) {
    static const char *names[] = { "automatic", "cpu", "gpu_nvidia", "gpu_amd", "gpu_intel" };
    return out << names[(int) target];
}
std::istream &operator>>(std::istream &in, target_platform &target) {
    std::string s;
    in >> s;
    s = lower(s);
    static const char *names[] = { "automatic", "cpu", "gpu_nvidia", "gpu_amd", "gpu_intel" };
    for (int k = 0; k < 5; ++k) {
        if (s == names[k]) {
            target = (target_platform) k;
            return in;
        }
    }
    in.setstate(std::ios::failbit);
    return in;
}

// ---------------------------------------------------------------------------------------------------------------------
// csr_matrix
// ---------------------------------------------------------------------------------------------------------------------
template <typename T>
std::size_t csr_matrix<T>::longest_row() const noexcept {
    std::int64_t m = 0;
    for (std::size_t i = 0; i < rows; ++i) { m = std::max(m, rowptr[i + 1] - rowptr[i]); }
    return (std::size_t) m;
}
template <typename T>
std::vector<std::vector<T>> csr_matrix<T>::to_dense() const {
    std::vector<std::vector<T>> d(rows, std::vector<T>(cols, T{ 0 }));
    for (std::size_t i = 0; i < rows; ++i) {
        for (std::int64_t p = rowptr[i]; p < rowptr[i + 1]; ++p) { d[i][(std::size_t) col[(std::size_t) p]] = val[(std::size_t) p]; }
    }
    return d;
}
template <typename T>
csr_matrix<T> csr_matrix<T>::from_dense(const std::vector<std::vector<T>> &data) {
    csr_matrix<T> m;
    m.rows = data.size();
    m.cols = data.empty() ? 0 : data.front().size();
    m.rowptr.assign(m.rows + 1, 0);
    for (std::size_t i = 0; i < m.rows; ++i) {
        for (std::size_t f = 0; f < data[i].size(); ++f) {
            if (data[i][f] != T{ 0 }) {
                m.col.push_back((std::int32_t) f);
                m.val.push_back(data[i][f]);
            }
        }
        m.rowptr[i + 1] = (std::int64_t) m.col.size();
    }
    return m;
}
template struct csr_matrix<float>;
template struct csr_matrix<double>;

// ---------------------------------------------------------------------------------------------------------------------
// parameter: file names and parsers
// ---------------------------------------------------------------------------------------------------------------------
namespace {
std::string basename_of(const std::string &path) {
    const auto pos = path.find_last_of("/\\");
    return pos == std::string::npos ? path : path.substr(pos + 1);
}
}  // namespace

template <typename T>
std::string parameter<T>::model_name_from_input() const { return basename_of(input_filename) + ".model"; }  // include/plssvm/parameter.hpp:217-221
template <typename T>
std::string parameter<T>::predict_name_from_input() const { return basename_of(input_filename) + ".predict"; }

template <typename T>
void parameter<T>::parse_libsvm_file(const std::string &filename, const bool test) {
    input_filename = filename;
    if (model_filename.empty()) { model_filename = model_name_from_input(); }
    const text_file f{ filename, '#' };
    auto m = std::make_shared<csr_matrix<T>>();
    std::vector<T> labels;
    const bool has_labels = parse_libsvm_lines<T>(f, 0, *m, labels);
    if (gamma == T{ 0 }) { gamma = T{ 1 } / static_cast<T>(m->cols); }  // parameter.cpp:150-152
    if (has_labels) {
        for (T &v : labels) { v = sign(v); }  // parameter.cpp:160-163
        value_ptr = std::make_shared<const std::vector<T>>(std::move(labels));
    } else {
        value_ptr = nullptr;
    }
    if (print_info) {
        std::cout << "Read " << m->rows << " data points with " << m->cols << " features (" << m->nnz() << " stored entries) using the libsvm parser from file '"
                  << filename << "'.\n";
    }
    (test ? test_data_ptr : data_ptr) = dense_view(*m);
    (test ? sparse_test_data_ptr : sparse_data_ptr) = std::move(m);
}

template <typename T>
void parameter<T>::parse_arff_file(const std::string &filename, const bool test) {
    input_filename = filename;
    if (model_filename.empty()) { model_filename = model_name_from_input(); }
    const text_file f{ filename, '%' };
    std::size_t max_size = 0, header = 0;
    bool has_label = false;
    for (; header < f.lines.size(); ++header) {
        const std::string line = upper(std::string{ f.lines[header] });
        if (starts_with(line, "@RELATION")) {
            continue;
        } else if (starts_with(line, "@ATTRIBUTE")) {
            if (line.find("NUMERIC") == std::string::npos) {
                throw invalid_file_format_exception{ "Can only use NUMERIC features, but '" + std::string{ f.lines[header] } + "' was given!" };
            }
            if (has_label) {
                throw invalid_file_format_exception{ "Only the last ATTRIBUTE may be CLASS!" };
            } else if (line.find("CLASS") != std::string::npos) {
                has_label = true;
            }
            ++max_size;
        } else if (starts_with(line, "@DATA")) {
            break;
        }
    }
    if (max_size == 0) {
        throw invalid_file_format_exception{ "Can't parse file: no ATTRIBUTES are defined!" };
    } else if (header + 1 >= f.lines.size()) {
        throw invalid_file_format_exception{ "Can't parse file: no data points are given or @DATA is missing!" };
    }
    const std::size_t n = f.lines.size() - (header + 1);
    const std::size_t num_features = has_label ? max_size - 1 : max_size;
    auto m = std::make_shared<csr_matrix<T>>();
    m->rows = n;
    m->cols = num_features;
    m->rowptr.assign(n + 1, 0);
    std::vector<T> labels(n, T{ 0 });
    // One data line -> its non-zero (index, value) entries in ascending index order (what the reference's dense row holds,
    // parameter.cpp:238-330: a later entry of a sparse line overwrites an earlier one with the same index), appended to col / val.
    const auto parse_line = [&](const std::size_t i, std::vector<std::pair<std::int32_t, T>> &row, std::vector<std::int32_t> &col, std::vector<T> &val) {
        std::string_view line = f.lines[i + header + 1];
        if (starts_with(line, "@")) { throw invalid_file_format_exception{ "Read @ inside data section!: '" + std::string{ line } + "'" }; }
        if (starts_with(line, "{")) {
            if (!ends_with(line, "}")) {
                throw invalid_file_format_exception{ "Missing closing '}' for sparse data point " + std::to_string(i) + " description!" };
            }
            row.clear();
            bool is_class_set = false;
            std::string_view::size_type pos = 1;
            while (true) {
                auto next_pos = line.find_first_of(' ', pos);
                if (next_pos == std::string_view::npos) { break; }
                const auto index = convert_to<unsigned long, invalid_file_format_exception>(line.substr(pos, next_pos - pos));
                if (index >= max_size) {
                    throw invalid_file_format_exception{ "Too many features given! Trying to add feature at position " + std::to_string(index) +
                                                         " but max position is " + std::to_string(num_features - 1) + "!" };
                }
                pos = next_pos + 1;
                next_pos = line.find_first_of(",}", pos);
                if (index == max_size - 1 && has_label) {
                    is_class_set = true;
                    std::string_view rest = line.substr(pos);
                    if (!rest.empty() && rest.back() == '}') { rest.remove_suffix(1); }
                    labels[i] = sign(convert_to<T, invalid_file_format_exception>(rest));
                } else {
                    row.emplace_back((std::int32_t) index, convert_to<T, invalid_file_format_exception>(line.substr(pos, next_pos - pos)));
                }
                line.remove_prefix(next_pos + 1);
                line = trim_left(line);
                pos = 0;
            }
            if (!is_class_set && has_label) { throw invalid_file_format_exception{ "Missing label for data point " + std::to_string(i) + "!" }; }
            std::stable_sort(row.begin(), row.end(), [](const auto &x, const auto &y) { return x.first < y.first; });
            for (std::size_t k = 0; k < row.size(); ++k) {
                if (k + 1 < row.size() && row[k + 1].first == row[k].first) { continue; }  // overwritten by a later entry
                if (row[k].second != T{ 0 }) {
                    col.push_back(row[k].first);
                    val.push_back(row[k].second);
                }
            }
        } else {
            const auto keep = [&](const std::size_t j, const T v) {
                if (v != T{ 0 }) {
                    col.push_back((std::int32_t) j);
                    val.push_back(v);
                }
            };
            std::string_view::size_type pos = 0, next_pos = 0;
            for (std::size_t j = 0; j < max_size - 1; ++j) {
                next_pos = line.find_first_of(',', pos);
                if (next_pos == std::string_view::npos) {
                    throw invalid_file_format_exception{ "Invalid number of features/labels! Found " + std::to_string(j) + " but should be " +
                                                         std::to_string(max_size - 1) + "!" };
                }
                keep(j, convert_to<T, invalid_file_format_exception>(line.substr(pos, next_pos - pos)));
                pos = next_pos + 1;
            }
            next_pos = line.find_first_of(',', pos);
            if (next_pos != std::string_view::npos) {
                throw invalid_file_format_exception{ "Too many features! Superfluous '" + std::string{ line.substr(next_pos) } + "' for data point " +
                                                     std::to_string(i) + "!" };
            }
            if (has_label) {
                labels[i] = sign(convert_to<T, invalid_file_format_exception>(line.substr(pos)));
            } else {
                keep(num_features - 1, convert_to<T, invalid_file_format_exception>(line.substr(pos)));
            }
        }
    };
    // contiguous chunks of lines per team member, stitched in file order; the first malformed line in file order is reported
    // (same scheme as parse_libsvm_lines; the reference parses the data section inside `omp parallel for`, parameter.cpp:236)
    struct chunk {
        std::vector<std::int32_t> col;
        std::vector<T> val;
        std::exception_ptr error{};
    };
    const int teams = (int) std::max<std::size_t>(1, std::min<std::size_t>((std::size_t) omp_get_max_threads(), n / 256 + 1));
    std::vector<chunk> chunks((std::size_t) teams);
#pragma omp parallel num_threads(teams)
    {
        const std::size_t t = (std::size_t) omp_get_thread_num(), nt = (std::size_t) omp_get_num_threads();
        std::vector<std::pair<std::int32_t, T>> row;
        for (std::size_t ci = t; ci < chunks.size(); ci += nt) {
            chunk &ck = chunks[ci];
            try {
                for (std::size_t i = n * ci / chunks.size(); i < n * (ci + 1) / chunks.size(); ++i) {
                    const std::size_t before = ck.col.size();
                    parse_line(i, row, ck.col, ck.val);
                    m->rowptr[i + 1] = (std::int64_t) (ck.col.size() - before);  // entries of the line; turned into offsets below
                }
            } catch (...) { ck.error = std::current_exception(); }
        }
    }
    for (const chunk &ck : chunks) {
        if (ck.error) { std::rethrow_exception(ck.error); }
    }
    for (std::size_t i = 0; i < n; ++i) { m->rowptr[i + 1] += m->rowptr[i]; }
    m->col.reserve((std::size_t) m->rowptr[n]);
    m->val.reserve((std::size_t) m->rowptr[n]);
    for (const chunk &ck : chunks) {
        m->col.insert(m->col.end(), ck.col.begin(), ck.col.end());
        m->val.insert(m->val.end(), ck.val.begin(), ck.val.end());
    }
    if (gamma == T{ 0 }) { gamma = T{ 1 } / static_cast<T>(num_features); }
    value_ptr = has_label ? std::make_shared<const std::vector<T>>(std::move(labels)) : nullptr;
    if (print_info) {
        std::cout << "Read " << n << " data points with " << num_features << " features using the arff parser from file '" << filename << "'.\n";
    }
    (test ? test_data_ptr : data_ptr) = dense_view(*m);
    (test ? sparse_test_data_ptr : sparse_data_ptr) = std::move(m);
}

template <typename T>
void parameter<T>::parse_train_file(const std::string &filename) {
    if (ends_with(filename, ".arff")) {
        parse_arff_file(filename, false);
    } else {
        parse_libsvm_file(filename, false);
    }
    if (value_ptr == nullptr) {
        throw invalid_file_format_exception{ "Missing labels for train file!" };  // src/plssvm/parameter.cpp:524-529
    }
}

template <typename T>
void parameter<T>::parse_test_file(const std::string &filename) {
    const std::string model = model_filename;
    if (ends_with(filename, ".arff")) {
        parse_arff_file(filename, true);
    } else {
        parse_libsvm_file(filename, true);
    }
    model_filename = model;
}

// LIBSVM model file (reference: src/plssvm/parameter.cpp:366-520; written by csvm::write_model, src/plssvm/csvm.cpp:60-204)
template <typename T>
void parameter<T>::parse_model_file(const std::string &filename) {
    model_filename = filename;
    const text_file f{ filename, '#' };
    value_ptr = nullptr;
    unsigned long long num_sv = 0;
    T label_first{ 0 }, label_second{ 0 };
    bool rho_set = false;
    std::size_t header = 0;
    for (; header < f.lines.size(); ++header) {
        const std::string line = lower(std::string{ trim(f.lines[header]) });
        std::string_view value{ line };
        value.remove_prefix(std::min(value.find_first_of(' ') + 1, value.size()));
        value = trim_left(value);
        if (starts_with(line, "svm_type")) {
            if (value != "c_svc") { throw invalid_file_format_exception{ "Can only use c_svc as svm_type, but '" + std::string{ value } + "' was given!" }; }
        } else if (starts_with(line, "kernel_type")) {
            std::istringstream iss{ std::string{ value } };
            iss >> kernel;
            if (iss.fail()) { throw invalid_file_format_exception{ "Unrecognized kernel type '" + std::string{ value } + "'!" }; }
        } else if (starts_with(line, "gamma")) {
            gamma = convert_to<T>(value);
        } else if (starts_with(line, "degree")) {
            degree = (int) convert_to<long>(value);
        } else if (starts_with(line, "coef0")) {
            coef0 = convert_to<T>(value);
        } else if (starts_with(line, "nr_class")) {
            const auto nr_class = convert_to<unsigned long>(value);
            if (nr_class != 2) { throw invalid_file_format_exception{ "Can only use 2 classes, but " + std::to_string(nr_class) + " were given!" }; }
        } else if (starts_with(line, "total_sv")) {
            num_sv = convert_to<unsigned long long>(value);
            if (num_sv == 0) { throw invalid_file_format_exception{ "The number of support vectors must be greater than 0, but is 0!" }; }
        } else if (starts_with(line, "rho")) {
            rho = convert_to<T>(value);
            rho_set = true;
        } else if (starts_with(line, "label")) {
            const std::string_view first_label = value.substr(0, value.find_first_of(' '));
            label_first = convert_to<T>(first_label);
            value.remove_prefix(std::min(first_label.size() + 1, value.size()));
            const std::string_view second_label = value.substr(0, value.find_first_of(" \n"));
            label_second = convert_to<T, invalid_file_format_exception>(second_label);
            value.remove_prefix(std::min(second_label.size() + 1, value.size()));
            value = trim_left(value);
            if (!value.empty() || (label_first != 1 && label_first != -1) || (label_second != 1 && label_second != -1)) {
                throw invalid_file_format_exception{ "Only the labels 1 and -1 are allowed, but '" + line + "' were given!" };
            }
        } else if (starts_with(line, "nr_sv")) {
            const std::string_view first_num = value.substr(0, value.find_first_of(' '));
            const auto num_first = convert_to<unsigned long long>(first_num);
            value.remove_prefix(std::min(first_num.size() + 1, value.size()));
            const std::string_view second_num = value.substr(0, value.find_first_of(" \n"));
            const auto num_second = convert_to<unsigned long long>(second_num);
            value.remove_prefix(std::min(second_num.size() + 1, value.size()));
            value = trim_left(value);
            if (!value.empty()) {
                throw invalid_file_format_exception{ "Only two numbers are allowed, but more were given '" + line + "'!" };
            } else if (num_first + num_second != num_sv) {
                throw invalid_file_format_exception{ "The number of positive and negative support vectors doesn't add up to the total number: " +
                                                     std::to_string(num_first) + " + " + std::to_string(num_second) + " != " + std::to_string(num_sv) + "!" };
            }
            std::vector<T> values(num_sv);
            std::fill(values.begin(), values.begin() + (std::ptrdiff_t) num_first, label_first);
            std::fill(values.begin() + (std::ptrdiff_t) num_first, values.end(), label_second);
            value_ptr = std::make_shared<const std::vector<T>>(std::move(values));
        } else if (line == "sv") {
            break;
        } else {
            throw invalid_file_format_exception{ "Unrecognized header entry '" + std::string{ f.lines[header] } + "'! Maybe SV is missing?" };
        }
    }
    if (num_sv == 0) {
        throw invalid_file_format_exception{ "Missing total number of support vectors!" };
    } else if (label_first == 0 || label_second == 0) {
        throw invalid_file_format_exception{ "Missing labels!" };
    } else if (value_ptr == nullptr) {
        throw invalid_file_format_exception{ "Missing number of support vectors per class!" };
    } else if (!rho_set) {
        throw invalid_file_format_exception{ "Missing rho value!" };
    } else if (header + 1 >= f.lines.size()) {
        throw invalid_file_format_exception{ "Can't parse file: no support vectors are given or SV is missing!" };
    }
    auto m = std::make_shared<csr_matrix<T>>();
    std::vector<T> alphas;
    parse_libsvm_lines<T>(f, header + 1, *m, alphas);
    if (m->rows != num_sv) {
        throw invalid_file_format_exception{ "Found " + std::to_string(m->rows) + " support vectors but total_sv is " + std::to_string(num_sv) + "!" };
    }
    alpha_ptr = std::make_shared<const std::vector<T>>(std::move(alphas));
    if (print_info) {
        std::cout << "Read " << m->rows << " support vectors with " << m->cols << " features using the libsvm model parser from file '" << filename << "'.\n";
    }
    data_ptr = dense_view(*m);
    sparse_data_ptr = std::move(m);
}

template <typename T>
std::ostream &operator<<(std::ostream &out, const parameter<T> &params) {
    out << "kernel_type       " << params.kernel << "\ndegree            " << params.degree << "\ngamma             " << detail::format_shortest(params.gamma)
        << "\ncoef0             " << detail::format_shortest(params.coef0) << "\ncost              " << detail::format_shortest(params.cost)
        << "\nepsilon           " << detail::format_shortest(params.epsilon) << "\nprint_info        " << (params.print_info ? "true" : "false")
        << "\nbackend           " << params.backend << "\ntarget platform   " << params.target << "\ninput_filename    '" << params.input_filename
        << "'\nmodel_filename    '" << params.model_filename << "'\npredict_filename  '" << params.predict_filename << "'\nrho               "
        << detail::format_shortest(params.rho) << "\nreal_type         " << (sizeof(T) == 4 ? "float" : "double") << "\n";
    return out;
}

template class parameter<float>;
template class parameter<double>;
template std::ostream &operator<<(std::ostream &, const parameter<float> &);
template std::ostream &operator<<(std::ostream &, const parameter<double> &);

// ---------------------------------------------------------------------------------------------------------------------
// number formatting / FP22
// ---------------------------------------------------------------------------------------------------------------------
namespace detail {

namespace {
template <typename V>
std::string shortest(const V v) {
    if (v != v) { return "nan"; }
    if (v == std::numeric_limits<V>::infinity()) { return "inf"; }
    if (v == -std::numeric_limits<V>::infinity()) { return "-inf"; }
    if (v == V{ 0 }) { return std::signbit(v) ? "-0" : "0"; }
    // shortest round-trip digits in scientific form: d[.ddd]e[+-]XX
    char buf[64];
    const auto res = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);
    std::string_view sci{ buf, (std::size_t) (res.ptr - buf) };
    const bool neg = sci.front() == '-';
    if (neg) { sci.remove_prefix(1); }
    const auto epos = sci.find('e');
    std::string digits;
    for (const char ch : sci.substr(0, epos)) {
        if (ch != '.') { digits.push_back(ch); }
    }
    const int exp10 = std::stoi(std::string{ sci.substr(epos + 1) });
    std::string out = neg ? "-" : "";
    if (exp10 < -4 || exp10 >= 16) {  // fmt's switch to exponent notation for the default float format
        out += digits.substr(0, 1);
        if (digits.size() > 1) { out += "." + digits.substr(1); }
        char e[16];
        std::snprintf(e, sizeof(e), "e%c%02d", exp10 < 0 ? '-' : '+', exp10 < 0 ? -exp10 : exp10);
        out += e;
    } else if (exp10 < 0) {
        out += "0." + std::string((std::size_t) (-exp10 - 1), '0') + digits;
    } else if ((std::size_t) exp10 + 1 >= digits.size()) {
        out += digits + std::string((std::size_t) exp10 + 1 - digits.size(), '0');
    } else {
        out += digits.substr(0, (std::size_t) exp10 + 1) + "." + digits.substr((std::size_t) exp10 + 1);
    }
    return out;
}
}  // namespace

std::string format_shortest(const double v) { return shortest(v); }
std::string format_shortest(const float v) { return shortest(v); }

std::uint32_t fp22_encode(const float x) {
    std::uint32_t u;
    std::memcpy(&u, &x, 4);
    if ((u & 0x7f800000u) == 0x7f800000u) {  // inf / nan: truncate, keep nan-ness
        std::uint32_t t = u >> 10;
        if ((u & 0x007fffffu) != 0 && (t & 0x1fffu) == 0) { t |= 1u; }
        return t;
    }
    const std::uint32_t lsb = (u >> 10) & 1u;  // round to nearest even on the 10 dropped bits
    u += 0x1ffu + lsb;
    return u >> 10;
}
float fp22_decode(const std::uint32_t v22) {
    const std::uint32_t u = v22 << 10;
    float x;
    std::memcpy(&x, &u, 4);
    return x;
}

}  // namespace detail

}  // namespace plssvm
