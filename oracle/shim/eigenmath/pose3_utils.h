// TEST INFRASTRUCTURE ONLY -- stand-in for eigenmath/pose3_utils.h (x_edr_eigenmath
// v1.0.0, un-vendored).  UNPINNED: PoseErrorBetween is restated as the Euclidean
// distance of the translations and the angular distance of the rotations.  Only
// .translation reaches a result (timeable_path_cartesian_spline.cc:425-437 sums
// translation + translation).
#ifndef ORACLE_SHIM_EIGENMATH_POSE3_UTILS_H_
#define ORACLE_SHIM_EIGENMATH_POSE3_UTILS_H_

#include "eigenmath/pose3.h"
#include "eigenmath/types.h"

namespace eigenmath {

struct PoseError {
  double translation;
  double rotation;
};

inline PoseError PoseErrorBetween(const Pose3d& a, const Pose3d& b) {
  return PoseError{(a.translation() - b.translation()).norm(),
                   a.quaternion().angularDistance(b.quaternion())};
}

}  // namespace eigenmath

#endif  // ORACLE_SHIM_EIGENMATH_POSE3_UTILS_H_
