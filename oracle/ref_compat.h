// TEST INFRASTRUCTURE ONLY (oracle/): force-included (-include) when compiling the
// UNMODIFIED reference sources under /root/reference/src for oracle/_ref.
// The reference calls AT_DISPATCH_*(tensor.type(), ...) (src/correlation_kernels.cu:145,176,
// src/altcorr_kernel.cu:309); torch >= 2.x dropped the DeprecatedTypeProperties overload of
// ::detail::scalar_type that the macro needs.  Supplying that one overload lets the sources
// compile where they lie, without patching them.
#pragma once
#include <torch/extension.h>
namespace detail {
inline at::ScalarType scalar_type(const at::DeprecatedTypeProperties& t) { return t.scalarType(); }
}  // namespace detail
