// Force-included (-include) when compiling the UNMODIFIED reference sources
// /root/reference/models/csrc/*.cu against torch >= 2.1, where
// AT_DISPATCH_*(x.type(), ...) no longer accepts at::DeprecatedTypeProperties.
// The reference passes x.type() at 13 sites (e.g. raymarching.cu:79,110,152,309,431);
// this overload restores the conversion so that the sources compile where they lie.
// TEST INFRASTRUCTURE ONLY (builds oracle/_ref/vren_ref*.so, the GPU-side oracle).
#pragma once
#include <torch/extension.h>
namespace detail {
inline at::ScalarType scalar_type(const at::DeprecatedTypeProperties& t) {
  return t.scalarType();
}
}  // namespace detail
