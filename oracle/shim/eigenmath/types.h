// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// Minimal stand-in for eigenmath/types.h (x_edr_eigenmath v1.0.0 is an
// un-vendored Bazel module; Eigen is not installed in this image), sufficient
// for the reference's rescale_to_stop.cc and sampled_trajectory.cc.
//
// eigenmath::VectorXd is Eigen's dynamic column vector of doubles.  The
// stand-in evaluates every operator eagerly, coefficient by coefficient, in
// the order Eigen's lazy expression templates evaluate the same expression per
// coefficient (a + (0.5 * v) * d etc.), so results are bit-identical for the
// expressions those two files contain.
#ifndef ORACLE_SHIM_EIGENMATH_TYPES_H_
#define ORACLE_SHIM_EIGENMATH_TYPES_H_

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <vector>

namespace Eigen {
constexpr int Infinity = -1;
}  // namespace Eigen

namespace eigenmath {

class VectorXd {
 public:
  VectorXd() = default;
  explicit VectorXd(int n) : d_(static_cast<size_t>(n), 0.0) {}
  static VectorXd Zero(int n) { return VectorXd(n); }
  static VectorXd Constant(int n, double v) {
    VectorXd r(n);
    std::fill(r.d_.begin(), r.d_.end(), v);
    return r;
  }
  int size() const { return static_cast<int>(d_.size()); }
  double& operator[](int i) { return d_[static_cast<size_t>(i)]; }
  const double& operator[](int i) const { return d_[static_cast<size_t>(i)]; }
  double& operator()(int i) { return d_[static_cast<size_t>(i)]; }
  const double& operator()(int i) const { return d_[static_cast<size_t>(i)]; }
  double* data() { return d_.data(); }
  const double* data() const { return d_.data(); }
  void setZero() { std::fill(d_.begin(), d_.end(), 0.0); }
  void setZero(int n) { d_.assign(static_cast<size_t>(n), 0.0); }
  void setConstant(int n, double v) { d_.assign(static_cast<size_t>(n), v); }

  template <int P>
  double lpNorm() const {
    static_assert(P == Eigen::Infinity, "only the infinity norm is provided");
    double m = 0.0;
    for (double x : d_) m = std::max(m, std::abs(x));
    return m;
  }
  double minCoeff() const { return *std::min_element(d_.begin(), d_.end()); }
  double maxCoeff() const { return *std::max_element(d_.begin(), d_.end()); }

 private:
  std::vector<double> d_;
};

inline VectorXd operator+(const VectorXd& a, const VectorXd& b) {
  VectorXd r(a.size());
  for (int i = 0; i < a.size(); ++i) r[i] = a[i] + b[i];
  return r;
}
inline VectorXd operator-(const VectorXd& a, const VectorXd& b) {
  VectorXd r(a.size());
  for (int i = 0; i < a.size(); ++i) r[i] = a[i] - b[i];
  return r;
}
inline VectorXd operator-(const VectorXd& a) {
  VectorXd r(a.size());
  for (int i = 0; i < a.size(); ++i) r[i] = -a[i];
  return r;
}
inline VectorXd operator*(const VectorXd& a, double s) {
  VectorXd r(a.size());
  for (int i = 0; i < a.size(); ++i) r[i] = a[i] * s;
  return r;
}
inline VectorXd operator*(double s, const VectorXd& a) {
  VectorXd r(a.size());
  for (int i = 0; i < a.size(); ++i) r[i] = s * a[i];
  return r;
}

}  // namespace eigenmath

#endif  // ORACLE_SHIM_EIGENMATH_TYPES_H_
