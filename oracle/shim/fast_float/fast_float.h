// TEST INFRASTRUCTURE (oracle build shim) — not product code.
// The reference parses numbers with fast_float v3.4.0 (include/plssvm/detail/string_conversion.hpp:45),
// which is fetched from the network at configure time and is absent here. Both fast_float::from_chars and
// std::from_chars are correctly rounded, so forwarding is value-identical.
#pragma once
#include <charconv>
namespace fast_float {
using from_chars_result = std::from_chars_result;
template <typename T>
inline from_chars_result from_chars(const char *first, const char *last, T &value) noexcept {
    return std::from_chars(first, last, value);
}
}  // namespace fast_float
