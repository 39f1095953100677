// TEST INFRASTRUCTURE ONLY -- stand-in for eigenmath/manifolds.h (x_edr_eigenmath
// v1.0.0, un-vendored).  UNPINNED: ExpSO3 / LogSO3 are the textbook exponential and
// logarithm maps with series near zero, the same restatement as orc_exp_so3 /
// orc_log_so3 in oracle/tp_oracle_cart.c.  Reached only from BSplineQ's degree > 2
// knot-insertion solver (see so3.h).
#ifndef ORACLE_SHIM_EIGENMATH_MANIFOLDS_H_
#define ORACLE_SHIM_EIGENMATH_MANIFOLDS_H_

#include <cmath>

#include "eigenmath/so3.h"
#include "eigenmath/types.h"

namespace eigenmath {

inline SO3d ExpSO3(const Vector3d& v) {
  const double th2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  const double th = std::sqrt(th2);
  const double k = (th < 1e-8) ? 0.5 - th2 / 48.0 : std::sin(0.5 * th) / th;
  return SO3d(Quaterniond(std::cos(0.5 * th), k * v[0], k * v[1], k * v[2]));
}

inline Vector3d LogSO3(const SO3d& r) {
  const Quaterniond& q = r.quaternion();
  const double n = std::sqrt(q.x() * q.x() + q.y() * q.y() + q.z() * q.z());
  const double k = (n < 1e-10) ? 2.0 / q.w() : 2.0 * std::atan2(n, q.w()) / n;
  return Vector3d(k * q.x(), k * q.y(), k * q.z());
}

}  // namespace eigenmath

#endif  // ORACLE_SHIM_EIGENMATH_MANIFOLDS_H_
