// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/log/absl_check.h.
#ifndef ORACLE_SHIM_ABSL_LOG_ABSL_CHECK_H_
#define ORACLE_SHIM_ABSL_LOG_ABSL_CHECK_H_

#include "absl/log/check.h"

#define ABSL_CHECK(cond) CHECK(cond)
#define ABSL_CHECK_LT(a, b) CHECK_LT(a, b)
#define ABSL_CHECK_LE(a, b) CHECK_LE(a, b)
#define ABSL_CHECK_EQ(a, b) CHECK_EQ(a, b)

#endif  // ORACLE_SHIM_ABSL_LOG_ABSL_CHECK_H_
