// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/strings/substitute.h
// (the formatted text only ever reaches CHECK's discarded stream).
#ifndef ORACLE_SHIM_ABSL_STRINGS_SUBSTITUTE_H_
#define ORACLE_SHIM_ABSL_STRINGS_SUBSTITUTE_H_

#include <string>

namespace absl {
template <typename... Args>
std::string Substitute(const char* format, const Args&...) {
  return std::string(format);
}
}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_STRINGS_SUBSTITUTE_H_
