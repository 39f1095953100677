// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/status/statusor.h.
#ifndef ORACLE_SHIM_ABSL_STATUS_STATUSOR_H_
#define ORACLE_SHIM_ABSL_STATUS_STATUSOR_H_

#include <type_traits>
#include <utility>

#include "absl/status/status.h"

namespace absl {

template <typename T>
class StatusOr {
 public:
  StatusOr(const Status& s) : status_(s) {}              // NOLINT
  StatusOr(T v) : value_(std::move(v)) {}                // NOLINT
  template <typename U, typename = typename std::enable_if<
                            std::is_convertible<U, T>::value &&
                            !std::is_same<typename std::decay<U>::type, T>::value &&
                            !std::is_same<typename std::decay<U>::type, Status>::value>::type>
  StatusOr(U&& v) : value_(std::forward<U>(v)) {}        // NOLINT
  bool ok() const { return status_.ok(); }
  const Status& status() const { return status_; }
  T& operator*() { return value_; }
  const T& operator*() const { return value_; }
  T* operator->() { return &value_; }
  const T* operator->() const { return &value_; }

 private:
  Status status_;
  T value_{};
};

}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_STATUS_STATUSOR_H_
