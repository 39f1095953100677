// TEST INFRASTRUCTURE ONLY -- stand-in for eigenmath/so3.h (x_edr_eigenmath v1.0.0,
// un-vendored).  UNPINNED: SO3d is restated as a unit quaternion with the group
// product of Eigen's scalar quaternion formulas.  Reached only from BSplineQ's
// knot-insertion solver for degree > 2 (splines/bsplineq.cc:433-494), which the
// degree-2 Cartesian path (kSplineOrder) never runs.
#ifndef ORACLE_SHIM_EIGENMATH_SO3_H_
#define ORACLE_SHIM_EIGENMATH_SO3_H_

#include "eigenmath/types.h"

namespace eigenmath {

class SO3d {
 public:
  SO3d() = default;
  explicit SO3d(const Quaterniond& q) : q_(q) {}
  const Quaterniond& quaternion() const { return q_; }
  SO3d operator*(const SO3d& b) const { return SO3d(q_ * b.q_); }
  SO3d inverse() const { return SO3d(q_.conjugate()); }

 private:
  Quaterniond q_;
};

}  // namespace eigenmath

#endif  // ORACLE_SHIM_EIGENMATH_SO3_H_
