// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/strings/str_format.h
// (messages only; no test compares them).
#ifndef ORACLE_SHIM_ABSL_STRINGS_STR_FORMAT_H_
#define ORACLE_SHIM_ABSL_STRINGS_STR_FORMAT_H_

#include <cstdio>
#include <string>

namespace absl {
template <typename... Args>
std::string StrFormat(const char* fmt, const Args&... args) {
  char buf[256];
  std::snprintf(buf, sizeof(buf), fmt, args...);
  return std::string(buf);
}
}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_STRINGS_STR_FORMAT_H_
