// TEST INFRASTRUCTURE ONLY -- stand-in for eigenmath/scalar_utils.h (un-vendored).
#ifndef ORACLE_SHIM_EIGENMATH_SCALAR_UTILS_H_
#define ORACLE_SHIM_EIGENMATH_SCALAR_UTILS_H_

namespace eigenmath {
template <typename T>
constexpr T Square(T x) {
  return x * x;
}
}  // namespace eigenmath

#endif  // ORACLE_SHIM_EIGENMATH_SCALAR_UTILS_H_
