// TEST INFRASTRUCTURE ONLY -- stand-in for eigenmath/pose3.h (x_edr_eigenmath
// v1.0.0, un-vendored).  UNPINNED: Pose3d is restated as (translation, unit
// quaternion) with a * b = (ta + qa * tb, qa * qb) and inverse() = (-(q^* t), q^*),
// Eigen's scalar quaternion formulas (the same choices as oracle/tp_oracle_cart.c).
#ifndef ORACLE_SHIM_EIGENMATH_POSE3_H_
#define ORACLE_SHIM_EIGENMATH_POSE3_H_

#include "eigenmath/types.h"

namespace eigenmath {

class Pose3d {
 public:
  Pose3d() = default;
  Pose3d(const Quaterniond& q, const Vector3d& t) : t_(t), q_(q) {}
  Vector3d& translation() { return t_; }
  const Vector3d& translation() const { return t_; }
  const Quaterniond& quaternion() const { return q_; }
  void setQuaternion(const Quaterniond& q) { q_ = q; }
  Pose3d inverse() const {
    const Quaterniond qc = q_.conjugate();
    return Pose3d(qc, -(qc * t_));
  }
  Pose3d operator*(const Pose3d& b) const { return Pose3d(q_ * b.q_, t_ + q_ * b.t_); }

 private:
  Vector3d t_;
  Quaterniond q_;
};

}  // namespace eigenmath

#endif  // ORACLE_SHIM_EIGENMATH_POSE3_H_
