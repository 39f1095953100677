// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/time/time.h, sufficient
// for the reference's trajectory_buffer.{h,cc} and time.h: absl::Time as int64
// nanoseconds since the Unix epoch (absl::Time() IS the epoch), absl::Duration
// as int64 nanoseconds, Duration / Duration as the floating-point quotient
// (absl::FDivDuration).  Times in the tests are far inside int64 ns, where
// abseil's own representation (seconds + quarter-nanosecond ticks) holds the
// same values exactly.
#ifndef ORACLE_SHIM_ABSL_TIME_TIME_H_
#define ORACLE_SHIM_ABSL_TIME_TIME_H_

#include <cstdint>

namespace absl {

class Duration {
 public:
  constexpr Duration() = default;
  constexpr explicit Duration(int64_t ns) : ns_(ns) {}
  constexpr int64_t ns() const { return ns_; }

 private:
  int64_t ns_ = 0;
};
constexpr Duration Nanoseconds(int64_t n) { return Duration(n); }
constexpr Duration Seconds(int64_t n) { return Duration(n * 1000000000); }
inline double operator/(Duration a, Duration b) {
  return static_cast<double>(a.ns()) / static_cast<double>(b.ns());
}

class Time {
 public:
  constexpr Time() = default;
  constexpr explicit Time(int64_t ns) : ns_(ns) {}
  constexpr int64_t ns() const { return ns_; }
  friend bool operator==(Time a, Time b) { return a.ns_ == b.ns_; }
  friend bool operator!=(Time a, Time b) { return a.ns_ != b.ns_; }
  friend bool operator<(Time a, Time b) { return a.ns_ < b.ns_; }

 private:
  int64_t ns_ = 0;
};
constexpr Time FromUnixNanos(int64_t ns) { return Time(ns); }
constexpr int64_t ToUnixNanos(Time t) { return t.ns(); }

}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_TIME_TIME_H_
