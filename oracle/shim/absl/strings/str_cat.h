// TEST INFRASTRUCTURE ONLY -- stand-in for absl/strings/str_cat.h (StrCat lives in
// the status.h stand-in, which the reference's sources reach it through).
#ifndef ORACLE_SHIM_ABSL_STRINGS_STR_CAT_H_
#define ORACLE_SHIM_ABSL_STRINGS_STR_CAT_H_

#include "absl/status/status.h"

#endif  // ORACLE_SHIM_ABSL_STRINGS_STR_CAT_H_
