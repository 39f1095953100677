// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/algorithm/container.h.
#ifndef ORACLE_SHIM_ABSL_ALGORITHM_CONTAINER_H_
#define ORACLE_SHIM_ABSL_ALGORITHM_CONTAINER_H_

#include <algorithm>

namespace absl {
template <typename C, typename OutIt>
OutIt c_copy(const C& c, OutIt out) {
  return std::copy(c.begin(), c.end(), out);
}
template <typename C>
void c_reverse(C& c) {
  std::reverse(c.begin(), c.end());
}
template <typename C, typename T>
auto c_upper_bound(C& c, const T& value) -> decltype(c.begin()) {
  return std::upper_bound(c.begin(), c.end(), value);
}
template <typename C, typename T, typename Compare>
auto c_upper_bound(C& c, const T& value, Compare comp) -> decltype(c.begin()) {
  return std::upper_bound(c.begin(), c.end(), value, comp);
}
template <typename C, typename T>
auto c_lower_bound(C& c, const T& value) -> decltype(c.begin()) {
  return std::lower_bound(c.begin(), c.end(), value);
}
}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_ALGORITHM_CONTAINER_H_
