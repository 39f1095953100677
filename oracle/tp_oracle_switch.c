/* TEST INFRASTRUCTURE ONLY -- not part of the product.
 *
 * CPU restatement of the online re-planning geometry of a joint path:
 *   BSplineT::InsertKnotAndUpdateControlPointsRef   splines/bspline.h:244-278
 *   BSplineBase::CanInsertKnot / InsertKnotIntoKnotVector  bspline_base.cc:166-214
 *   BSplineT::TruncateSplineAt                      splines/bspline.h:402-427
 *   BSplineT::ExtendWithControlPoints               splines/bspline.h:429-511
 *   ProjectPointOnPath                              path_tools.h:59-104
 *   TimeableJointSplinePath::SwitchToWaypointPath   timeable_path_joint_spline.cc:209-250
 *
 * PARITY UNPINNED in two places, both because of un-vendored dependencies:
 *  - TruncateSplineAt calls InsertKnotAndUpdateControlPoints, which solves a
 *    small linear system with Eigen's colPivHouseholderQr (bspline.h:281-399;
 *    Eigen is not in /root/reference).  The restatement uses the reference's
 *    own ...Ref variant (NURBS Alg. 5.1), which the reference's tests hold to
 *    the same curve (bspline_test.cc "InsertKnot..." compare both); results
 *    agree to rounding, not bit for bit.
 *  - eigenmath::DistanceFromLineSegment is not in /root/reference; restated as
 *    the distance to the SEGMENT (line parameter clamped to [0, 1]).
 * Pinned through the properties the reference's tests state: the curve is
 * unchanged on [umin, u_end] by truncation and by extension, knot insertion
 * leaves the curve unchanged (bspline_test.cc), ProjectPointOnPath's four cases
 * (path_tools_test.cc:41-115), and the switching scenario of
 * path_timing_trajectory_test.cc:299-436.
 */
#include <float.h>
#include <math.h>
#include <string.h>

#include "tp_oracle.h"

#ifndef ORC_MAX_DEGREE
#define ORC_MAX_DEGREE 8
#endif

enum { kOk = 0, kInvalidArgument = 3, kFailedPrecondition = 9, kUnimplemented = 12 };

/* bspline_base.cc:166-195 */
static int can_insert_knot(const orc_spline* s, double knot, int multiplicity) {
  if (multiplicity > s->degree + 1) return kInvalidArgument;
  if (s->num_knots + multiplicity > s->knot_capacity) return kFailedPrecondition;
  if (s->num_knots < 2) return kFailedPrecondition;
  if (knot <= s->knots[0] || knot >= s->knots[s->num_knots - 1]) return kInvalidArgument;
  return kOk;
}

/* bspline.h:244-278 */
static void insert_knot_ref(orc_spline* s, double knot) {
  const int p = s->degree, D = s->dim;
  double scratch[ORC_MAX_DEGREE][64];
  const int span = orc_knot_span(s->knots, s->num_knots, p, knot);
  /* shift the unchanged points behind the affected region */
  memmove(s->points + (size_t)(span + 1) * D, s->points + (size_t)span * D,
          sizeof(double) * (size_t)(s->num_points - span) * D);
  for (int i = 0; i < p; ++i) {
    const int k = span + i - p + 1;
    const double alpha = (knot - s->knots[k]) / (s->knots[k + p] - s->knots[k]);
    for (int d = 0; d < D; ++d) {
      scratch[i][d] =
          alpha * s->points[(size_t)k * D + d] + (1.0 - alpha) * s->points[(size_t)(k - 1) * D + d];
    }
  }
  for (int i = 0; i < p; ++i) {
    memcpy(s->points + (size_t)(span - p + 1 + i) * D, scratch[i], sizeof(double) * D);
  }
  /* InsertKnotIntoKnotVector (bspline_base.cc:197-214) */
  memmove(s->knots + span + 2, s->knots + span + 1,
          sizeof(double) * (size_t)(s->num_knots - span - 1));
  s->knots[span + 1] = knot;
  s->num_knots += 1;
  s->num_points += 1;
}

int orc_spline_insert_knot(orc_spline* s, double knot, int multiplicity) {
  const int st = can_insert_knot(s, knot, multiplicity);
  if (st != kOk) return st;
  for (int i = 0; i < multiplicity; ++i) insert_knot_ref(s, knot);
  return kOk;
}

/* bspline.h:402-427 */
int orc_spline_truncate(orc_spline* s, double u_end) {
  if (s->num_knots == 0) return kOk;
  const double umin = s->knots[0], umax = s->knots[s->num_knots - 1];
  if (u_end >= umax) return kOk;
  if (u_end <= umin) {
    s->num_knots = 0;
    s->num_points = 0;
    return kOk;
  }
  const int st = orc_spline_insert_knot(s, u_end, s->degree + 1);
  if (st != kOk) return st;
  const int span = orc_knot_span(s->knots, s->num_knots, s->degree, u_end);
  s->num_knots = span + 1;
  s->num_points = s->num_knots - s->degree - 1;
  return kOk;
}

/* bspline.h:429-511 (degree 2 only) */
int orc_spline_extend(orc_spline* s, const double* pts, int n) {
  const int p = s->degree, D = s->dim;
  if (p != 2) return kUnimplemented;
  const int new_num_points = s->num_points + n;
  const int added_knots = (n + 1 + p + 1) - 2 * p; /* NumKnots(n + 1, p) - 2 p */
  const int new_num_knots = s->num_knots + added_knots;
  if (s->num_knots < 2 * p + 2) return kFailedPrecondition;
  if (new_num_points > s->knot_capacity - p - 1) return kFailedPrecondition;
  if (new_num_knots > s->knot_capacity) return kFailedPrecondition;
  if (n < 2) return kUnimplemented;
  const double u_join = s->knots[s->num_knots - 1];
  const double old_knot_range = s->knots[s->num_knots - 1] - s->knots[0];
  const int old_inner = s->num_knots - 2 * p - 1;
  const int new_inner = new_num_knots - 2 * p - 1;
  const double new_knot_range = (old_knot_range * new_inner) / old_inner;
  const int upper = new_num_knots - p;
  const int start = s->num_knots - p - 1;
  const int size = upper - start;
  { /* Eigen 3.4 setLinSpaced(size, low, high), |high| >= |low| branch */
    const double low = old_knot_range, high = new_knot_range;
    const int size1 = (size == 1) ? 1 : size - 1;
    const double step = (size == 1) ? 0.0 : (high - low) / (double)(size - 1);
    for (int i = 0; i < size; ++i) {
      s->knots[start + i] = (i == size1) ? high : low + (double)i * step;
    }
  }
  for (int i = 0; i < p + 1; ++i) s->knots[new_num_knots - p - 1 + i] = s->knots[0] + new_knot_range;
  s->num_knots = new_num_knots;
  /* KnotSpan uses num_points_ of the OLD spline for u == last knot only; u_join is
   * now an inner knot */
  const int join_span = orc_knot_span(s->knots, s->num_knots, p, u_join);
  const int mod = s->num_points - 1;
  double basis[ORC_MAX_DEGREE + 1];
  orc_basis(s->knots, join_span, p, u_join, basis);
  if (!(basis[1] > 0)) return 13; /* CHECK in the reference */
  for (int d = 0; d < D; ++d) {
    s->points[(size_t)mod * D + d] =
        1.0 / (basis[1]) *
        (s->points[(size_t)mod * D + d] - basis[0] * s->points[(size_t)(mod - 1) * D + d]);
  }
  memcpy(s->points + (size_t)s->num_points * D, pts, sizeof(double) * (size_t)n * D);
  s->num_points = new_num_points;
  return kOk;
}

/* eigenmath::DistanceFromLineSegment: UNPINNED restatement (see the header) */
static void distance_from_line_segment(const double* a, const double* b, const double* p, int D,
                                       double* distance, double* line_parameter) {
  double ab2 = 0.0, dot = 0.0;
  for (int d = 0; d < D; ++d) {
    ab2 += (b[d] - a[d]) * (b[d] - a[d]);
    dot += (p[d] - a[d]) * (b[d] - a[d]);
  }
  double t = (ab2 > 0.0) ? dot / ab2 : 0.0;
  t = (t < 0.0) ? 0.0 : ((t > 1.0) ? 1.0 : t);
  double s = 0.0;
  for (int d = 0; d < D; ++d) {
    const double e = a[d] + t * (b[d] - a[d]) - p[d];
    s += e * e;
  }
  *distance = sqrt(s);
  *line_parameter = t;
}

/* path_tools.h:59-104 */
int orc_project_point_on_path(const double* waypoints, int W, int D, const double* point,
                              int* waypoint_index, double* distance_to_path,
                              double* line_parameter, double* projected_point) {
  if (W == 0) return kInvalidArgument;
  if (W == 1) {
    double s = 0.0;
    for (int d = 0; d < D; ++d) s += (waypoints[d] - point[d]) * (waypoints[d] - point[d]);
    *waypoint_index = 0;
    *distance_to_path = sqrt(s);
    *line_parameter = 0.0;
    memcpy(projected_point, waypoints, sizeof(double) * D);
    return kOk;
  }
  *distance_to_path = DBL_MAX;
  *line_parameter = 0.0;
  *waypoint_index = 0;
  for (int i = 0; i < W - 1; ++i) {
    double dist = 0.0, t = 0.0;
    distance_from_line_segment(waypoints + (size_t)i * D, waypoints + (size_t)(i + 1) * D, point, D,
                               &dist, &t);
    if (dist < *distance_to_path) {
      *distance_to_path = dist;
      *line_parameter = t;
      *waypoint_index = i;
    }
  }
  const double* a = waypoints + (size_t)(*waypoint_index) * D;
  const double* b = a + D;
  for (int d = 0; d < D; ++d) projected_point[d] = a[d] + *line_parameter * (b[d] - a[d]);
  return kOk;
}

/* timeable_path_joint_spline.cc:209-250.  new_ctrl receives the control polygon of
 * the new waypoints (control_points_ after the call), capacity 3 (W + 1) - 2 rows. */
int orc_joint_switch(orc_spline* s, double keep_path_until, const double* waypoints, int W,
                     double rounding, double* new_ctrl, int* num_new_ctrl) {
  const int D = s->dim;
  double switch_position[64], projected[64];
  double wp[(ORC_MAX_SWITCH_WAYPOINTS + 1) * 64];
  int st = orc_spline_truncate(s, keep_path_until);
  if (st != kOk) return st;
  if (s->num_knots == 0) return 11; /* EvalCurve on a cleared curve: kOutOfRange */
  st = orc_eval_curve(s->knots, s->num_knots, s->points, D, s->degree, keep_path_until,
                      switch_position);
  if (st != kOk) return st;
  if (W > ORC_MAX_SWITCH_WAYPOINTS) return kInvalidArgument;
  int index = 0;
  double dist = 0.0, t = 0.0;
  st = orc_project_point_on_path(waypoints, W, D, switch_position, &index, &dist, &t, projected);
  if (st != kOk) return st;
  int n = 0;
  double linf = 0.0;
  for (int d = 0; d < D; ++d) {
    const double e = fabs(switch_position[d] - projected[d]);
    if (e > linf) linf = e;
  }
  if (linf > 1e-3) { /* kEpsilon */
    memcpy(wp, projected, sizeof(double) * D);
    n = 1;
  }
  const int first = (t >= 0) ? index + 1 : index;
  for (int i = first; i < W; ++i) {
    memcpy(wp + (size_t)n * D, waypoints + (size_t)i * D, sizeof(double) * D);
    ++n;
  }
  if (n == 0) return kInvalidArgument; /* the reference would index an empty vector */
  const int nc = orc_polyline_to_bspline3(wp, n, D, rounding, new_ctrl);
  *num_new_ctrl = nc;
  return orc_spline_extend(s, new_ctrl, nc);
}
