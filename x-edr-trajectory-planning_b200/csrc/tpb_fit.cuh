// Control polygon and knot vector of the degree-2 path splines, shared by the
// joint-space and the Cartesian fit kernels (one thread per path).
#pragma once

#include "tpb_device.cuh"

namespace tpb {

// PolyLineToBspline3WaypointsCornerOffset, splines/spline_utils.cc:25-45.
__device__ inline void corner_offset(const double* delta, int D, double radius, double* off) {
  constexpr double kMinNorm = 1e-6;
  constexpr double kMinWaypointSpacingFactor = 4.0;  // spline_utils.h:46
  double s = 0.0;
  for (int d = 0; d < D; ++d) s += delta[d] * delta[d];
  const double norm = sqrt(s);
  if (norm > kMinNorm) {
    for (int d = 0; d < D; ++d) off[d] = delta[d] / norm;
  } else {
    for (int d = 0; d < D; ++d) off[d] = 0.0;
  }
  if (norm > kMinWaypointSpacingFactor * radius) {
    for (int d = 0; d < D; ++d) off[d] = off[d] * radius;
  } else {
    const double f = 1.0 / kMinWaypointSpacingFactor;
    for (int d = 0; d < D; ++d) off[d] = (off[d] * f) * norm;
  }
}

// PolyLineToBspline3Waypoints (VectorXd), splines/spline_utils.cc:47-102:
// W corners -> 3W-2 control points (4 copies of the point if W == 1).
__device__ inline int polyline_to_bspline3(const double* wp, int W, int D, double rounding,
                                           double* c) {
  double delta[kMaxDofs], off[kMaxDofs];
  if (W == 1) {
    for (int i = 0; i < 4; ++i)
      for (int d = 0; d < D; ++d) c[i * D + d] = wp[d];
    return 4;
  }
  const int n = 3 * W - 2;
  for (int i = 0; i < W; ++i)
    for (int d = 0; d < D; ++d) c[3 * i * D + d] = wp[i * D + d];
  for (int i = 1; i < W - 1; ++i) {
    const int k = 3 * i, knext = 3 * (i + 1), klast = 3 * (i - 1);
    for (int d = 0; d < D; ++d) delta[d] = c[knext * D + d] - c[k * D + d];
    corner_offset(delta, D, rounding, off);
    for (int d = 0; d < D; ++d) c[(k + 1) * D + d] = c[k * D + d] + off[d];
    for (int d = 0; d < D; ++d) delta[d] = c[klast * D + d] - c[k * D + d];
    corner_offset(delta, D, rounding, off);
    for (int d = 0; d < D; ++d) c[(k - 1) * D + d] = c[k * D + d] + off[d];
  }
  for (int d = 0; d < D; ++d) delta[d] = c[3 * D + d] - c[d];
  corner_offset(delta, D, rounding, off);
  for (int d = 0; d < D; ++d) c[D + d] = c[d] + off[d];
  for (int d = 0; d < D; ++d) delta[d] = c[(n - 4) * D + d] - c[(n - 1) * D + d];
  corner_offset(delta, D, rounding, off);
  for (int d = 0; d < D; ++d) c[(n - 2) * D + d] = c[(n - 1) * D + d] + off[d];
  return n;
}

// MakeUniformKnotVector on [0,1] (bspline_base.cc:350-381) for degree 2
// (kSplineOrder, timeable_path_joint_spline.h:89).  Returns the knot count.
__device__ inline int uniform_knots_degree2(int n, double* kn) {
  constexpr int p = 2;
  const int nk = n + p + 1;
  for (int i = 0; i <= p; ++i) kn[i] = 0.0;
  const double denom = nk - 2.0 * (p + 1.0) + 1.0;
  const double spacing = (1.0 / denom) * (1.0 - 0.0);
  for (int i = p + 1; i < nk - p - 1; ++i) kn[i] = kn[i - 1] + spacing;
  for (int i = nk - p - 1; i < nk; ++i) kn[i] = 1.0;
  return nk;
}

// BSplineBase::UpdateBasis, splines/bspline_base.cc:249-265 (NURBS A2.2).
template <int P>
__device__ __forceinline__ void basis_funs(const double* __restrict__ knots, int span, double u,
                                           double (&basis)[P + 1]) {
  double left[P + 1], right[P + 1];
  basis[0] = 1.0;
#pragma unroll
  for (int j = 1; j <= P; ++j) {
    left[j] = u - knots[span + 1 - j];
    right[j] = knots[span + j] - u;
  }
#pragma unroll
  for (int j = 1; j <= P; ++j) {
    double saved = 0.0;
#pragma unroll
    for (int r = 0; r < j; ++r) {
      const double tmp = basis[r] / (right[r + 1] + left[j - r]);
      basis[r] = saved + right[r + 1] * tmp;
      saved = left[j - r] * tmp;
    }
    basis[j] = saved;
  }
}

}  // namespace tpb
