// TEST INFRASTRUCTURE ONLY -- stand-in for eigenmath/utils.h (x_edr_eigenmath
// v1.0.0, un-vendored).  UNPINNED: DistanceFromLineSegment is restated as the
// distance to the SEGMENT with the line parameter clamped to [0, 1]
// (the same choice as oracle/tp_oracle_switch.c and csrc/tpb_switch.cu; the
// reference's tests, path_tools_test.cc:41-115, hold for it).
#ifndef ORACLE_SHIM_EIGENMATH_UTILS_H_
#define ORACLE_SHIM_EIGENMATH_UTILS_H_

#include <cmath>

#include "eigenmath/types.h"

namespace eigenmath {

template <typename V, typename Scalar>
void DistanceFromLineSegment(const V& a, const V& b, const V& p, Scalar* distance,
                             Scalar* line_parameter) {
  Scalar ab2 = 0, dot = 0;
  for (int d = 0; d < a.size(); ++d) {
    ab2 += (b[d] - a[d]) * (b[d] - a[d]);
    dot += (p[d] - a[d]) * (b[d] - a[d]);
  }
  Scalar t = (ab2 > 0) ? dot / ab2 : 0;
  t = (t < 0) ? 0 : ((t > 1) ? 1 : t);
  Scalar s = 0;
  for (int d = 0; d < a.size(); ++d) {
    const Scalar e = a[d] + t * (b[d] - a[d]) - p[d];
    s += e * e;
  }
  *distance = std::sqrt(s);
  *line_parameter = t;
}

// Declared only: path_tools.cc (ComputeStoppingPoint) is not compiled.
VectorXd ScaleDownToLimits(const VectorXd& value, const VectorXd& limits);

}  // namespace eigenmath

#endif  // ORACLE_SHIM_EIGENMATH_UTILS_H_
