// TEST INFRASTRUCTURE ONLY -- stand-in for eigenmath/interpolation.h
// (x_edr_eigenmath v1.0.0, un-vendored).  UNPINNED: the five-argument
// InterpolateLinear(x, x0, x1, y0, y1) used by trajectory_buffer.cc is restated
// as y0 + ((x - x0) / (x1 - x0)) * (y1 - y0), returning y0 when x1 == x0 (the
// reference's own test reads the last sample through a (last, last) bracket and
// expects that sample: trajectory_buffer_test.cc:503-528).
#ifndef ORACLE_SHIM_EIGENMATH_INTERPOLATION_H_
#define ORACLE_SHIM_EIGENMATH_INTERPOLATION_H_

#include "eigenmath/types.h"

namespace eigenmath {
inline VectorXd InterpolateLinear(double x, double x0, double x1, const VectorXd& y0,
                                  const VectorXd& y1) {
  if (x1 == x0) return y0;
  const double t = (x - x0) / (x1 - x0);
  VectorXd r(y0.size());
  for (int i = 0; i < y0.size(); ++i) r[i] = y0[i] + t * (y1[i] - y0[i]);
  return r;
}
}  // namespace eigenmath

#endif  // ORACLE_SHIM_EIGENMATH_INTERPOLATION_H_
