// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/time/time.h, sufficient
// for the reference's trajectory_buffer.{h,cc}, trajectory_planner.h,
// path_timing_trajectory.cc and time.h: absl::Time as int64 nanoseconds since the
// Unix epoch (absl::Time() IS the epoch), absl::Duration as int64 nanoseconds,
// Duration / Duration as the floating-point quotient (absl::FDivDuration).  Times
// in the tests are far inside int64 ns, where abseil's own representation
// (seconds + quarter-nanosecond ticks) holds the same values exactly; Seconds(double)
// rounds to the nearest nanosecond (abseil: to the nearest quarter nanosecond -- the
// planner only passes its time step, a whole number of nanoseconds in every test).
#ifndef ORACLE_SHIM_ABSL_TIME_TIME_H_
#define ORACLE_SHIM_ABSL_TIME_TIME_H_

#include <cmath>
#include <cstdint>
#include <type_traits>

namespace absl {

class Duration {
 public:
  constexpr Duration() = default;
  constexpr explicit Duration(int64_t ns) : ns_(ns) {}
  constexpr int64_t ns() const { return ns_; }
  friend constexpr bool operator==(Duration a, Duration b) { return a.ns_ == b.ns_; }
  friend constexpr bool operator!=(Duration a, Duration b) { return a.ns_ != b.ns_; }
  friend constexpr bool operator<(Duration a, Duration b) { return a.ns_ < b.ns_; }
  friend constexpr bool operator<=(Duration a, Duration b) { return a.ns_ <= b.ns_; }
  friend constexpr bool operator>(Duration a, Duration b) { return a.ns_ > b.ns_; }
  friend constexpr bool operator>=(Duration a, Duration b) { return a.ns_ >= b.ns_; }
  friend constexpr Duration operator+(Duration a, Duration b) { return Duration(a.ns_ + b.ns_); }
  friend constexpr Duration operator-(Duration a, Duration b) { return Duration(a.ns_ - b.ns_); }

 private:
  int64_t ns_ = 0;
};
constexpr Duration Nanoseconds(int64_t n) { return Duration(n); }
constexpr Duration Milliseconds(int64_t n) { return Duration(n * 1000000); }
template <typename T, typename std::enable_if<std::is_integral<T>::value, int>::type = 0>
constexpr Duration Seconds(T n) {
  return Duration(static_cast<int64_t>(n) * 1000000000);
}
template <typename T, typename std::enable_if<std::is_floating_point<T>::value, int>::type = 0>
inline Duration Seconds(T n) {
  return Duration(static_cast<int64_t>(std::llround(static_cast<double>(n) * 1e9)));
}
inline double operator/(Duration a, Duration b) {
  return static_cast<double>(a.ns()) / static_cast<double>(b.ns());
}

class Time {
 public:
  constexpr Time() = default;
  constexpr explicit Time(int64_t ns) : ns_(ns) {}
  constexpr int64_t ns() const { return ns_; }
  friend constexpr bool operator==(Time a, Time b) { return a.ns_ == b.ns_; }
  friend constexpr bool operator!=(Time a, Time b) { return a.ns_ != b.ns_; }
  friend constexpr bool operator<(Time a, Time b) { return a.ns_ < b.ns_; }
  friend constexpr bool operator<=(Time a, Time b) { return a.ns_ <= b.ns_; }
  friend constexpr bool operator>(Time a, Time b) { return a.ns_ > b.ns_; }
  friend constexpr bool operator>=(Time a, Time b) { return a.ns_ >= b.ns_; }
  friend constexpr Time operator+(Time a, Duration d) { return Time(a.ns_ + d.ns()); }
  friend constexpr Time operator-(Time a, Duration d) { return Time(a.ns_ - d.ns()); }
  friend constexpr Duration operator-(Time a, Time b) { return Duration(a.ns_ - b.ns_); }

 private:
  int64_t ns_ = 0;
};
constexpr Time FromUnixNanos(int64_t ns) { return Time(ns); }
constexpr Time FromUnixSeconds(int64_t s) { return Time(s * 1000000000); }
constexpr int64_t ToUnixNanos(Time t) { return t.ns(); }
namespace time_internal {
constexpr Time FromUnixDuration(Duration d) { return Time(d.ns()); }
}  // namespace time_internal

}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_TIME_TIME_H_
