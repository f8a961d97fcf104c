// TEST INFRASTRUCTURE (oracle build shim) — not product code.
// Force-included (-include) when compiling the reference's sources with fmt >= 9 (torch ships fmt 12 headers):
// fmt no longer formats operator<< types implicitly, which the reference (written against fmt 8.1.1) relies on
// (e.g. src/plssvm/csvm.cpp:94-96, src/plssvm/backends/OpenMP/csvm.cpp:37).
#pragma once
#include "fmt/format.h"
#include "fmt/ostream.h"
#include "fmt/ranges.h"

#include "plssvm/backend_types.hpp"
#include "plssvm/backends/SYCL/implementation_type.hpp"
#include "plssvm/backends/SYCL/kernel_invocation_type.hpp"
#include "plssvm/kernel_types.hpp"
#include "plssvm/target_platforms.hpp"

template <> struct fmt::formatter<plssvm::kernel_type> : fmt::ostream_formatter {};
template <> struct fmt::formatter<plssvm::target_platform> : fmt::ostream_formatter {};
template <> struct fmt::formatter<plssvm::backend_type> : fmt::ostream_formatter {};
template <> struct fmt::formatter<plssvm::sycl_generic::implementation_type> : fmt::ostream_formatter {};
template <> struct fmt::formatter<plssvm::sycl_generic::kernel_invocation_type> : fmt::ostream_formatter {};
