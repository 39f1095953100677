// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/log/log.h: LOG(severity)
// discards what is streamed into it (messages only; nothing compares them).
#ifndef ORACLE_SHIM_ABSL_LOG_LOG_H_
#define ORACLE_SHIM_ABSL_LOG_LOG_H_

#include "absl/log/check.h"

#define LOG(severity) ::oracle_shim::CheckSink()

#endif  // ORACLE_SHIM_ABSL_LOG_LOG_H_
