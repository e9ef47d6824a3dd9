// Force-included (-include) when building the UNMODIFIED reference extension
// sources from /root/reference against torch 2.11 (see oracle/build_ref.py).
// Arithmetic-neutral shims only; no reference source is copied or edited.
//   1. matching_kernels.cu:103 passes the deprecated `Tensor::type()` to
//      AT_DISPATCH_FLOATING_TYPES_AND_HALF; torch 2.11's dispatch macro calls
//      ::detail::scalar_type(), which lost its DeprecatedTypeProperties overload.
//   2. gn_kernels.cu:802,1219,1629 call torch::linalg::linalg_norm, whose header
//      (torch/linalg.h) no longer exists; forward to at::linalg_norm.
#pragma once
#include <torch/extension.h>
#include <ATen/core/DeprecatedTypeProperties.h>

namespace detail {
inline at::ScalarType scalar_type(const at::DeprecatedTypeProperties& t) {
  return t.scalarType();
}
}  // namespace detail

namespace torch {
namespace linalg {
inline at::Tensor linalg_norm(const at::Tensor& self,
                              const std::optional<at::Scalar>& ord,
                              at::OptionalIntArrayRef dim, bool keepdim,
                              std::optional<at::ScalarType> dtype) {
  return at::linalg_norm(self, ord, dim, keepdim, dtype);
}
}  // namespace linalg
}  // namespace torch
