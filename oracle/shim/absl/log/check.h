// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/log/check.h: CHECK(cond)
// aborts with a message; anything streamed into it is discarded.
#ifndef ORACLE_SHIM_ABSL_LOG_CHECK_H_
#define ORACLE_SHIM_ABSL_LOG_CHECK_H_

#include <cstdio>
#include <cstdlib>

namespace oracle_shim {
struct CheckSink {
  template <typename T>
  CheckSink& operator<<(const T&) { return *this; }
};
inline CheckSink CheckFail(bool ok, const char* what) {
  if (!ok) {
    std::fprintf(stderr, "CHECK failed: %s\n", what);
    std::abort();
  }
  return CheckSink();
}
}  // namespace oracle_shim

#define CHECK(cond) ::oracle_shim::CheckFail(static_cast<bool>(cond), #cond)
#define CHECK_OK(status) CHECK((status).ok())
#define CHECK_EQ(a, b) CHECK((a) == (b))
#define CHECK_NE(a, b) CHECK((a) != (b))
#define CHECK_LE(a, b) CHECK((a) <= (b))
#define CHECK_LT(a, b) CHECK((a) < (b))
#define CHECK_GE(a, b) CHECK((a) >= (b))
#define CHECK_GT(a, b) CHECK((a) > (b))

#endif  // ORACLE_SHIM_ABSL_LOG_CHECK_H_
