// TEST INFRASTRUCTURE ONLY -- stand-in for eigenmath/interpolation.h
// (x_edr_eigenmath v1.0.0, un-vendored).  UNPINNED (stated in DESIGN.md):
//  * InterpolateLinear(t, a, b) is restated as a + t * (b - a)
//    (the oracle port and the kernels make the same choice);
//  * the five-argument InterpolateLinear(x, x0, x1, y0, y1) used by
//    trajectory_buffer.cc is restated as y0 + ((x - x0) / (x1 - x0)) * (y1 - y0),
//    returning y0 when x1 == x0 (the reference's own test reads the last sample
//    through a (last, last) bracket and expects that sample:
//    trajectory_buffer_test.cc:503-528).
#ifndef ORACLE_SHIM_EIGENMATH_INTERPOLATION_H_
#define ORACLE_SHIM_EIGENMATH_INTERPOLATION_H_

#include "eigenmath/types.h"

namespace eigenmath {
inline double InterpolateLinear(double t, double a, double b) { return a + t * (b - a); }
inline VectorXd InterpolateLinear(double t, const VectorXd& a, const VectorXd& b) {
  VectorXd r(a.size());
  for (int i = 0; i < a.size(); ++i) r[i] = a[i] + t * (b[i] - a[i]);
  return r;
}
template <int N>
Eigen::VectorNf<N> InterpolateLinear(double t, const Eigen::VectorNf<N>& a,
                                     const Eigen::VectorNf<N>& b) {
  Eigen::VectorNf<N> r;
  for (int i = 0; i < N; ++i) r[i] = a[i] + t * (b[i] - a[i]);
  return r;
}
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
