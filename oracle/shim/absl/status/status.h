// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// Minimal stand-in for absl/status/status.h (abseil-cpp is not installed in
// this image), sufficient for the reference's time_optimal_path_timing.{h,cc}.
// StatusCode numeric values are abseil's.
#ifndef ORACLE_SHIM_ABSL_STATUS_STATUS_H_
#define ORACLE_SHIM_ABSL_STATUS_STATUS_H_

#include <cstdio>
#include <string>
#include <utility>

// abseil's status.h brings Span along; bspline_base.h relies on that
#include "absl/types/span.h"

namespace absl {

enum class StatusCode : int {
  kOk = 0,
  kCancelled = 1,
  kUnknown = 2,
  kInvalidArgument = 3,
  kDeadlineExceeded = 4,
  kNotFound = 5,
  kAlreadyExists = 6,
  kPermissionDenied = 7,
  kResourceExhausted = 8,
  kFailedPrecondition = 9,
  kAborted = 10,
  kOutOfRange = 11,
  kUnimplemented = 12,
  kInternal = 13,
  kUnavailable = 14,
  kDataLoss = 15,
  kUnauthenticated = 16,
};

class Status {
 public:
  Status() = default;
  Status(StatusCode code, std::string msg)
      : code_(code), msg_(std::move(msg)) {}
  bool ok() const { return code_ == StatusCode::kOk; }
  void IgnoreError() const {}
  std::string ToString() const { return ok() ? std::string("OK") : msg_; }
  StatusCode code() const { return code_; }
  const std::string& message() const { return msg_; }

 private:
  StatusCode code_ = StatusCode::kOk;
  std::string msg_;
};

inline Status OkStatus() { return Status(); }
inline Status FailedPreconditionError(std::string m) {
  return Status(StatusCode::kFailedPrecondition, std::move(m));
}
inline Status NotFoundError(std::string m) {
  return Status(StatusCode::kNotFound, std::move(m));
}
inline Status InvalidArgumentError(std::string m) {
  return Status(StatusCode::kInvalidArgument, std::move(m));
}
inline Status InternalError(std::string m) {
  return Status(StatusCode::kInternal, std::move(m));
}
inline Status OutOfRangeError(std::string m) {
  return Status(StatusCode::kOutOfRange, std::move(m));
}
inline Status UnimplementedError(std::string m) {
  return Status(StatusCode::kUnimplemented, std::move(m));
}
inline Status AbortedError(std::string m) {
  return Status(StatusCode::kAborted, std::move(m));
}

// The reference .cc calls absl::StrCat without including str_cat.h directly.
// (No iostreams here: this library is dlopen()ed with a static libstdc++.)
namespace shim_detail {
inline void Append(std::string* out, const char* v) { out->append(v); }
inline void Append(std::string* out, const std::string& v) { out->append(v); }
inline void Append(std::string* out, int v) { out->append(std::to_string(v)); }
inline void Append(std::string* out, long v) { out->append(std::to_string(v)); }
inline void Append(std::string* out, unsigned long v) { out->append(std::to_string(v)); }
inline void Append(std::string* out, long long v) { out->append(std::to_string(v)); }
inline void Append(std::string* out, unsigned v) { out->append(std::to_string(v)); }
inline void Append(std::string* out, double v) {
  char buf[40];
  std::snprintf(buf, sizeof(buf), "%g", v);
  out->append(buf);
}
}  // namespace shim_detail
template <typename... Args>
std::string StrCat(const Args&... args) {
  std::string out;
  (shim_detail::Append(&out, args), ...);
  return out;
}

}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_STATUS_STATUS_H_
