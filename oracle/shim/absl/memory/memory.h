// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/memory/memory.h.
#ifndef ORACLE_SHIM_ABSL_MEMORY_MEMORY_H_
#define ORACLE_SHIM_ABSL_MEMORY_MEMORY_H_

#include <memory>

namespace absl {
template <typename T>
std::unique_ptr<T> WrapUnique(T* p) {
  return std::unique_ptr<T>(p);
}
}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_MEMORY_MEMORY_H_
