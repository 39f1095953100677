// Minimal stand-ins for the vocabulary types the reference's public API uses
// (absl::Status / StatusOr / Span / Time / Duration, eigenmath::VectorXd), which
// are not available in this image.  Numeric StatusCode values are abseil's, so
// adapters to the real types are one-liners where those libraries exist.
#ifndef TP_BASE_H_
#define TP_BASE_H_

#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <initializer_list>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

namespace tp {

enum class StatusCode : int {
  kOk = 0,
  kInvalidArgument = 3,
  kDeadlineExceeded = 4,
  kNotFound = 5,
  kResourceExhausted = 8,
  kFailedPrecondition = 9,
  kAborted = 10,
  kOutOfRange = 11,
  kUnimplemented = 12,
  kInternal = 13,
};

class Status {
 public:
  Status() = default;
  Status(StatusCode code, std::string msg) : code_(code), msg_(std::move(msg)) {}
  bool ok() const { return code_ == StatusCode::kOk; }
  StatusCode code() const { return code_; }
  const std::string& message() const { return msg_; }

 private:
  StatusCode code_ = StatusCode::kOk;
  std::string msg_;
};
inline Status OkStatus() { return Status(); }
inline Status InvalidArgumentError(std::string m) { return {StatusCode::kInvalidArgument, std::move(m)}; }
inline Status FailedPreconditionError(std::string m) { return {StatusCode::kFailedPrecondition, std::move(m)}; }
inline Status OutOfRangeError(std::string m) { return {StatusCode::kOutOfRange, std::move(m)}; }
inline Status InternalError(std::string m) { return {StatusCode::kInternal, std::move(m)}; }
inline Status NotFoundError(std::string m) { return {StatusCode::kNotFound, std::move(m)}; }
inline Status UnimplementedError(std::string m) { return {StatusCode::kUnimplemented, std::move(m)}; }
// Status from a C-ABI return code (include/tpb.h uses the same numbering).
inline Status FromCode(int32_t code, const char* msg) {
  return code == 0 ? Status() : Status(static_cast<StatusCode>(code), msg ? msg : "");
}

template <typename T>
class StatusOr {
 public:
  StatusOr(const Status& s) : status_(s) { assert(!s.ok()); }  // NOLINT
  StatusOr(const T& v) : value_(v) {}                          // NOLINT
  StatusOr(T&& v) : value_(std::move(v)) {}                    // NOLINT
  bool ok() const { return status_.ok(); }
  const Status& status() const { return status_; }
  const T& operator*() const { return value_; }
  T& operator*() { return value_; }
  const T* operator->() const { return &value_; }
  const T& value() const { return value_; }

 private:
  Status status_;
  T value_{};
};

template <typename T>
class Span {
 public:
  Span() = default;
  Span(T* p, size_t n) : p_(p), n_(n) {}
  template <typename V, typename = decltype(std::declval<V&>().data())>
  Span(V& v) : p_(v.data()), n_(v.size()) {}  // NOLINT
  template <typename V, typename = decltype(std::declval<const V&>().data()),
            typename U = T, typename = typename std::enable_if<std::is_const<U>::value>::type>
  Span(const V& v) : p_(v.data()), n_(v.size()) {}  // NOLINT
  Span(std::initializer_list<typename std::remove_const<T>::type> l)
      : p_(l.begin()), n_(l.size()) {}
  T* data() const { return p_; }
  size_t size() const { return n_; }
  bool empty() const { return n_ == 0; }
  T& operator[](size_t i) const { return p_[i]; }
  T* begin() const { return p_; }
  T* end() const { return p_ + n_; }
  T& front() const { return p_[0]; }
  T& back() const { return p_[n_ - 1]; }

 private:
  T* p_ = nullptr;
  size_t n_ = 0;
};

// Dense dynamic column vector of doubles (the subset of eigenmath::VectorXd the
// hot-path API touches).
class VectorXd {
 public:
  VectorXd() = default;
  explicit VectorXd(size_t n) : d_(n, 0.0) {}
  VectorXd(std::initializer_list<double> l) : d_(l) {}
  VectorXd(const double* p, size_t n) : d_(p, p + n) {}
  static VectorXd Zero(size_t n) { return VectorXd(n); }
  static VectorXd Constant(size_t n, double v) {
    VectorXd r(n);
    for (auto& x : r.d_) x = v;
    return r;
  }
  void resize(size_t n) { d_.resize(n, 0.0); }
  void setZero() { for (auto& x : d_) x = 0.0; }
  size_t size() const { return d_.size(); }
  size_t rows() const { return d_.size(); }
  double* data() { return d_.data(); }
  const double* data() const { return d_.data(); }
  double& operator[](size_t i) { return d_[i]; }
  const double& operator[](size_t i) const { return d_[i]; }
  double& operator()(size_t i) { return d_[i]; }
  const double& operator()(size_t i) const { return d_[i]; }
  const double* begin() const { return d_.data(); }
  const double* end() const { return d_.data() + d_.size(); }
  bool operator==(const VectorXd& o) const { return d_ == o.d_; }

 private:
  std::vector<double> d_;
};

// Minimal stand-ins for eigenmath::Vector3d / Quaterniond / Pose3d / Matrix6Xd:
// plain storage in Eigen's coefficient order (x, y, z, w), no arithmetic (the
// Cartesian path's arithmetic runs on the device).
struct Vector3d {
  double v[3] = {0.0, 0.0, 0.0};
  Vector3d() = default;
  Vector3d(double x, double y, double z) : v{x, y, z} {}
  double& operator[](size_t i) { return v[i]; }
  const double& operator[](size_t i) const { return v[i]; }
  double& operator()(size_t i) { return v[i]; }
  const double& operator()(size_t i) const { return v[i]; }
};
struct Quaterniond {
  double c[4] = {0.0, 0.0, 0.0, 1.0};  // x, y, z, w
  Quaterniond() = default;
  Quaterniond(double w, double x, double y, double z) : c{x, y, z, w} {}  // Eigen's ctor order
  static Quaterniond Identity() { return Quaterniond(); }
  double x() const { return c[0]; }
  double y() const { return c[1]; }
  double z() const { return c[2]; }
  double w() const { return c[3]; }
  const double* coeffs() const { return c; }
  double* coeffs() { return c; }
};
class Pose3d {
 public:
  Pose3d() = default;
  explicit Pose3d(const Vector3d& t) : t_(t) {}
  explicit Pose3d(const Quaterniond& q) : q_(q) {}
  Pose3d(const Quaterniond& q, const Vector3d& t) : t_(t), q_(q) {}
  const Vector3d& translation() const { return t_; }
  Vector3d& translation() { return t_; }
  const Quaterniond& quaternion() const { return q_; }
  void setQuaternion(const Quaterniond& q) { q_ = q; }

 private:
  Vector3d t_;
  Quaterniond q_;
};
// 6 x n, row-major storage, (row, col) access like Eigen's operator().
class Matrix6Xd {
 public:
  Matrix6Xd() = default;
  Matrix6Xd(size_t rows, size_t cols) : cols_(cols), d_(6 * cols, 0.0) { (void)rows; }
  void resize(size_t rows, size_t cols) {
    (void)rows;
    cols_ = cols;
    d_.assign(6 * cols, 0.0);
  }
  void setZero() { for (auto& x : d_) x = 0.0; }
  size_t rows() const { return 6; }
  size_t cols() const { return cols_; }
  double& operator()(size_t r, size_t c) { return d_[r * cols_ + c]; }
  const double& operator()(size_t r, size_t c) const { return d_[r * cols_ + c]; }
  const double* data() const { return d_.data(); }

 private:
  size_t cols_ = 0;
  std::vector<double> d_;
};

// absl::Time / absl::Duration only matter through int64-nanosecond conversions
// (trajectory_planning/time.h:23-30).
struct Duration {
  int64_t ns = 0;
};
struct Time {
  int64_t ns = 0;
};
inline Duration Nanoseconds(int64_t n) { return {n}; }
inline Duration Milliseconds(double m) { return {static_cast<int64_t>(m * 1e6)}; }
inline Duration Seconds(double s) { return {static_cast<int64_t>(s * 1e9)}; }
inline Time FromUnixNanos(int64_t n) { return {n}; }
inline int64_t ToUnixNanos(Time t) { return t.ns; }
inline Time operator+(Time t, Duration d) { return {t.ns + d.ns}; }
inline Duration operator-(Time a, Time b) { return {a.ns - b.ns}; }
inline bool operator<(Time a, Time b) { return a.ns < b.ns; }
inline bool operator>(Time a, Time b) { return a.ns > b.ns; }
inline bool operator>=(Time a, Time b) { return a.ns >= b.ns; }
inline bool operator<=(Time a, Time b) { return a.ns <= b.ns; }
inline bool operator==(Time a, Time b) { return a.ns == b.ns; }

}  // namespace tp

namespace trajectory_planning {
// time.h:23-30
inline tp::Time TimeFromSec(double seconds) {
  return tp::FromUnixNanos(static_cast<int64_t>(seconds * static_cast<double>(1000000000LL)));
}
inline double TimeToSec(const tp::Time& time) {
  return static_cast<double>(tp::ToUnixNanos(time)) / static_cast<double>(1000000000LL);
}
}  // namespace trajectory_planning

#endif  // TP_BASE_H_
