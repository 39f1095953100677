// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/algorithm/container.h.
#ifndef ORACLE_SHIM_ABSL_ALGORITHM_CONTAINER_H_
#define ORACLE_SHIM_ABSL_ALGORITHM_CONTAINER_H_

#include <algorithm>

namespace absl {
template <typename C>
void c_reverse(C& c) {
  std::reverse(c.begin(), c.end());
}
}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_ALGORITHM_CONTAINER_H_
