/* TEST INFRASTRUCTURE ONLY -- the CPU oracle, Cartesian-path part (see
 * tp_oracle.h).  Never linked into or called from the product.
 *
 * Plain-C restatement of the reference's Cartesian timeable path (BASELINE
 * config 5).  Citations are into /root/reference/trajectory_planning/.
 *
 *   quaternion B-spline      splines/bsplineq.cc:99-146, 223-244, 309-317
 *   pose control polygon     splines/spline_utils.cc:104-203
 *   fit / sample / rows      timeable_path_cartesian_spline.cc:39-68, 396-595
 *   fake kinematics          path_timing_trajectory_test.cc:548-587
 *
 * PARITY PIN
 *  - QuatExp against the Mathematica table of splines/bsplineq_test.cc:106-168
 *    (tests/golden/quat_exp_mathematica.json); degree-1 quaternion spline ==
 *    piecewise slerp (bsplineq_test.cc:309-344); pose control points against
 *    the values of splines/spline_utils_test.cc:43-137.
 *  - UNPINNED at the ulp level (arithmetic lives in un-vendored Eigen 3.4.0 /
 *    x_edr_eigenmath 1.0.0): quaternion product / squaredNorm summation order
 *    (Eigen's scalar formulas here), Pose3d composition (t1 + q1*t2 with Eigen's
 *    quaternion-vector formula), PoseErrorBetween (Euclidean distance of the
 *    translations), ExpSO3 / LogSO3 / LogSO3DerivativeManifold of the test's
 *    fake kinematics (standard closed forms; no reference test pins values).
 *
 * Poses are 7 doubles (tx, ty, tz, qx, qy, qz, qw); quaternions 4 doubles in
 * Eigen's coefficient order (x, y, z, w).
 */
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "tp_oracle.h"

/* Eigen::NumTraits<double>::dummy_precision() */
static const double kDummyPrecision = 1e-12;

/* ---------------------------------------------------------------- Eigen bits */
static double sq_norm4(const double* q) {
  return q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
}
/* Eigen stableNorm() of a 3-vector: scale * sqrt(sum((v/scale)^2)) */
static double stable_norm3(const double* v) {
  double scale = fabs(v[0]);
  if (fabs(v[1]) > scale) scale = fabs(v[1]);
  if (fabs(v[2]) > scale) scale = fabs(v[2]);
  if (!(scale > 0)) return 0.0;
  const double inv = 1.0 / scale;
  const double a = v[0] * inv, b = v[1] * inv, c = v[2] * inv;
  return scale * sqrt(a * a + b * b + c * c);
}
/* Eigen stableNormalized(): v / (sqrt(sum((v/w)^2)) * w), w = max|v_i| */
static void stable_normalized3(const double* v, double* out) {
  double w = fabs(v[0]);
  if (fabs(v[1]) > w) w = fabs(v[1]);
  if (fabs(v[2]) > w) w = fabs(v[2]);
  double z = 0.0;
  if (w > 0) {
    const double a = v[0] / w, b = v[1] / w, c = v[2] / w;
    z = a * a + b * b + c * c;
  }
  if (z > 0) {
    const double d = sqrt(z) * w;
    out[0] = v[0] / d;
    out[1] = v[1] / d;
    out[2] = v[2] / d;
  } else {
    out[0] = v[0];
    out[1] = v[1];
    out[2] = v[2];
  }
}
/* Eigen quat_product (scalar path) */
void orc_quat_mul(const double* a, const double* b, double* out) {
  const double x = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
  const double y = a[3] * b[1] + a[1] * b[3] + a[2] * b[0] - a[0] * b[2];
  const double z = a[3] * b[2] + a[2] * b[3] + a[0] * b[1] - a[1] * b[0];
  const double w = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
  out[0] = x;
  out[1] = y;
  out[2] = z;
  out[3] = w;
}
/* QuaternionBase::inverse(): conjugate / squaredNorm (zero if the norm is 0) */
static void quat_inverse(const double* q, double* out) {
  const double n2 = sq_norm4(q);
  if (n2 > 0) {
    out[0] = -q[0] / n2;
    out[1] = -q[1] / n2;
    out[2] = -q[2] / n2;
    out[3] = q[3] / n2;
  } else {
    out[0] = out[1] = out[2] = out[3] = 0.0;
  }
}
/* QuaternionBase::_transformVector */
static void quat_rotate(const double* q, const double* v, double* out) {
  double uv[3] = {q[1] * v[2] - q[2] * v[1], q[2] * v[0] - q[0] * v[2],
                  q[0] * v[1] - q[1] * v[0]};
  uv[0] += uv[0];
  uv[1] += uv[1];
  uv[2] += uv[2];
  const double c[3] = {q[1] * uv[2] - q[2] * uv[1], q[2] * uv[0] - q[0] * uv[2],
                       q[0] * uv[1] - q[1] * uv[0]};
  out[0] = v[0] + q[3] * uv[0] + c[0];
  out[1] = v[1] + q[3] * uv[1] + c[1];
  out[2] = v[2] + q[3] * uv[2] + c[2];
}

/* splines/bsplineq.cc:99-109 */
static void normalize_if_necessary_positive_real(double* q) {
  if (q[3] < 0) {
    q[0] *= -1.0;
    q[1] *= -1.0;
    q[2] *= -1.0;
    q[3] *= -1.0;
  }
  if (fabs(sq_norm4(q) - 1.0) > kDummyPrecision) {
    const double n = sqrt(sq_norm4(q));
    q[0] /= n;
    q[1] /= n;
    q[2] /= n;
    q[3] /= n;
  }
}

/* QuatLog, splines/bsplineq.cc:113-126 */
void orc_quat_log(const double* q, double* out) {
  const double squared_norm_q = sq_norm4(q);
  const double norm_v = stable_norm3(q);
  out[3] = 0.5 * log(squared_norm_q);
  if (norm_v > kDummyPrecision) {
    double n[3];
    stable_normalized3(q, n);
    const double ang = atan2(norm_v, q[3]);
    out[0] = n[0] * ang;
    out[1] = n[1] * ang;
    out[2] = n[2] * ang;
  } else {
    out[0] = q[0];
    out[1] = q[1];
    out[2] = q[2];
  }
}

/* QuatExp, splines/bsplineq.cc:128-135 */
void orc_quat_exp(const double* q, double* out) {
  const double norm_v = stable_norm3(q);
  double n[3];
  stable_normalized3(q, n);
  const double s = sin(norm_v);
  const double e = exp(q[3]);
  out[3] = cos(norm_v);
  out[0] = n[0] * s;
  out[1] = n[1] * s;
  out[2] = n[2] * s;
  out[0] *= e;
  out[1] *= e;
  out[2] *= e;
  out[3] *= e;
}

/* QuatPower, splines/bsplineq.cc:137-146 */
void orc_quat_power(const double* q, double power, double* out) {
  double in[4] = {q[0], q[1], q[2], q[3]};
  normalize_if_necessary_positive_real(in);
  double lg[4];
  orc_quat_log(in, lg);
  lg[0] *= power;
  lg[1] *= power;
  lg[2] *= power;
  lg[3] *= power;
  orc_quat_exp(lg, out);
}

/* BSplineQ::EvalCurve, splines/bsplineq.cc:223-244 with UpdateCumulativeBasis
 * (:309-317).  Returns 0, or 11 (kOutOfRange). */
int orc_quat_eval_curve(const double* knots, int num_knots, const double* quats,
                        int degree, double u, double* out) {
  if (u < knots[0] || u > knots[num_knots - 1]) return 11;
  const int span = orc_knot_span(knots, num_knots, degree, u);
  double basis[16], cum[16];
  orc_basis(knots, span, degree, u, basis);
  cum[degree - 1] = basis[degree];
  for (int i = degree - 2; i >= 0; --i) cum[i] = cum[i + 1] + basis[i + 1];
  double q[4];
  memcpy(q, quats + 4 * (span - degree), sizeof(q));
  for (int i = 0; i < degree; ++i) {
    double inv[4], rel[4], pw[4], t[4];
    quat_inverse(quats + 4 * (span - degree + i), inv);
    orc_quat_mul(inv, quats + 4 * (span - degree + i + 1), rel);
    orc_quat_power(rel, cum[i], pw);
    orc_quat_mul(q, pw, t);
    memcpy(q, t, sizeof(q));
  }
  normalize_if_necessary_positive_real(q);
  memcpy(out, q, sizeof(q));
  return 0;
}

/* QuatPowerDerivative, splines/bsplineq.cc:148-160: exp(p log q) * log q */
void orc_quat_power_derivative(const double* q, double power, double* out) {
  double in[4] = {q[0], q[1], q[2], q[3]};
  normalize_if_necessary_positive_real(in);
  double lg[4], sc[4], ex[4];
  orc_quat_log(in, lg);
  for (int i = 0; i < 4; ++i) sc[i] = lg[i] * power;
  orc_quat_exp(sc, ex);
  orc_quat_mul(ex, lg, out);
}

/* BSplineQ::EvalCurveAndDerivative, splines/bsplineq.cc:246-307 with
 * UpdateCumulativeBasisAndDerivative (:319-333).  The value is
 * points[span - p] * (prod_k QuatPower(rel_k, Bc_k)) WITHOUT the final normalisation of
 * EvalCurve (only the sign flip + normalize when w < 0).  Returns 0 or 11. */
int orc_quat_eval_curve_derivative(const double* knots, int num_knots, const double* quats,
                                   int degree, double u, double* out, double* dout) {
  if (u < knots[0] || u > knots[num_knots - 1]) return 11;
  const int span = orc_knot_span(knots, num_knots, degree, u);
  double ders[2 * 16], cum[16], dcum[16];
  orc_basis_ders(knots, span, degree, 1, u, ders);
  const double* b0 = ders;
  const double* b1 = ders + (degree + 1);
  cum[degree - 1] = b0[degree];
  dcum[degree - 1] = b1[degree];
  for (int i = degree - 2; i >= 0; --i) {
    cum[i] = cum[i + 1] + b0[i + 1];
    dcum[i] = dcum[i + 1] + b1[i + 1];
  }
  double pw[17][4]; /* quat_powers_: [0] = identity */
  double rel[16][4];
  pw[0][0] = pw[0][1] = pw[0][2] = 0.0;
  pw[0][3] = 1.0;
  for (int k = 0; k < degree; ++k) {
    double inv[4], f[4];
    quat_inverse(quats + 4 * (span - degree + k), inv);
    orc_quat_mul(inv, quats + 4 * (span - degree + k + 1), rel[k]);
    orc_quat_power(rel[k], cum[k], f);
    orc_quat_mul(pw[k], f, pw[k + 1]);
  }
  double q[4], d[4] = {0.0, 0.0, 0.0, 0.0};
  orc_quat_mul(quats + 4 * (span - degree), pw[degree], q);
  for (int k = 0; k < degree; ++k) {
    double dk[4], inv[4], t1[4], t2[4], t3[4];
    orc_quat_power_derivative(rel[k], cum[k], dk);
    for (int i = 0; i < 4; ++i) dk[i] *= dcum[k];
    orc_quat_mul(pw[k], dk, t1);
    quat_inverse(pw[k + 1], inv);
    orc_quat_mul(t1, inv, t2);
    orc_quat_mul(t2, pw[degree], t3);
    for (int i = 0; i < 4; ++i) d[i] += t3[i];
  }
  double dd[4];
  orc_quat_mul(quats + 4 * (span - degree), d, dd);
  if (q[3] < 0) {
    for (int i = 0; i < 4; ++i) q[i] *= -1.0;
    const double z = sq_norm4(q);
    if (z > 0) {
      const double n = sqrt(z);
      for (int i = 0; i < 4; ++i) q[i] /= n;
    }
    for (int i = 0; i < 4; ++i) dd[i] *= -1.0;
  }
  memcpy(out, q, sizeof(q));
  memcpy(dout, dd, sizeof(dd));
  return 0;
}

/* ---------------------------------------------------------------- poses */
static void pose_identity(double* p) {
  p[0] = p[1] = p[2] = 0.0;
  p[3] = p[4] = p[5] = 0.0;
  p[6] = 1.0;
}
/* a * b: (qa*qb, ta + qa*tb) */
static void pose_mul(const double* a, const double* b, double* out) {
  double r[3], q[4];
  quat_rotate(a + 3, b, r);
  orc_quat_mul(a + 3, b + 3, q);
  out[0] = a[0] + r[0];
  out[1] = a[1] + r[1];
  out[2] = a[2] + r[2];
  memcpy(out + 3, q, sizeof(q));
}
/* inverse: (conj(q), -(conj(q) * t)) for unit q */
static void pose_inverse(const double* a, double* out) {
  double qc[4] = {-a[3], -a[4], -a[5], a[6]};
  double r[3];
  quat_rotate(qc, a, r);
  out[0] = -r[0];
  out[1] = -r[1];
  out[2] = -r[2];
  memcpy(out + 3, qc, sizeof(qc));
}

/* CornerOffset (Pose3d), splines/spline_utils.cc:105-146, with Eigen's
 * AngleAxis <-> Quaternion conversions. */
static void pose_corner_offset(const double* delta, double translation_radius,
                               double rotation_radius, double* offset) {
  static const double kMinRadius = 1e-6;
  static const double kMinWaypointSpacingFactor = 4.0; /* spline_utils.h:46 */
  pose_identity(offset);
  if (translation_radius < kMinRadius || rotation_radius < kMinRadius) return;
  const double translation_norm =
      sqrt(delta[0] * delta[0] + delta[1] * delta[1] + delta[2] * delta[2]);
  /* AngleAxis(quaternion) */
  const double* q = delta + 3;
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  if (n < DBL_EPSILON) n = stable_norm3(q);
  double angle, axis[3];
  if (n != 0.0) {
    angle = 2.0 * atan2(n, fabs(q[3]));
    if (q[3] < 0.0) n = -n;
    axis[0] = q[0] / n;
    axis[1] = q[1] / n;
    axis[2] = q[2] / n;
  } else {
    angle = 0.0;
    axis[0] = 1.0;
    axis[1] = 0.0;
    axis[2] = 0.0;
  }
  const double pct_trans = (translation_norm == 0.0) ? INFINITY : translation_radius / translation_norm;
  const double pct_rot = (angle == 0.0) ? INFINITY : rotation_radius / angle;
  double pct = (pct_rot < pct_trans) ? pct_rot : pct_trans; /* std::min(trans, rot) */
  if (pct > (1.0 / kMinWaypointSpacingFactor)) pct = 1.0 / kMinWaypointSpacingFactor;
  offset[0] = delta[0] * pct;
  offset[1] = delta[1] * pct;
  offset[2] = delta[2] * pct;
  angle *= pct;
  /* Quaternion(AngleAxis) */
  const double ha = 0.5 * angle;
  const double s = sin(ha);
  offset[6] = cos(ha);
  offset[3] = s * axis[0];
  offset[4] = s * axis[1];
  offset[5] = s * axis[2];
}

/* PolyLineToBspline3Waypoints (Pose3d), splines/spline_utils.cc:148-203 */
int orc_pose_polyline_to_bspline3(const double* corners, int W, double translation_radius,
                                  double rotational_radius, double* out) {
  if (W == 1) {
    for (int i = 0; i < 4; ++i) memcpy(out + 7 * i, corners, 7 * sizeof(double));
    return 4;
  }
  const int n = 3 * W - 2;
  double inv[7], rel[7], off[7];
  for (int i = 0; i < W; ++i) memcpy(out + 7 * 3 * i, corners + 7 * i, 7 * sizeof(double));
  for (int idx = 1; idx < W - 1; ++idx) {
    const int k = 3 * idx, knext = 3 * (idx + 1), klast = 3 * (idx - 1);
    pose_inverse(out + 7 * k, inv);
    pose_mul(inv, out + 7 * knext, rel);
    pose_corner_offset(rel, translation_radius, rotational_radius, off);
    pose_mul(out + 7 * k, off, out + 7 * (k + 1));
    pose_mul(inv, out + 7 * klast, rel);
    pose_corner_offset(rel, translation_radius, rotational_radius, off);
    pose_mul(out + 7 * k, off, out + 7 * (k - 1));
  }
  pose_inverse(out, inv);
  pose_mul(inv, out + 7 * 3, rel);
  pose_corner_offset(rel, translation_radius, rotational_radius, off);
  pose_mul(out, off, out + 7 * 1);
  pose_inverse(out + 7 * (n - 1), inv);
  pose_mul(inv, out + 7 * (n - 4), rel);
  pose_corner_offset(rel, translation_radius, rotational_radius, off);
  pose_mul(out + 7 * (n - 1), off, out + 7 * (n - 2));
  return n;
}

/* FitSplineToWaypoints, timeable_path_cartesian_spline.cc:396-462.  Degree 2
 * (kSplineOrder).  Returns the number of control points. */
int orc_cart_fit(const double* pose_wp, const double* joint_wp, int W, int dim,
                 double translation_rounding, double rotation_rounding, double* pose_ctrl,
                 double* joint_ctrl, double* knots) {
  const int n = orc_polyline_to_bspline3(joint_wp, W, dim, rotation_rounding, joint_ctrl);
  orc_pose_polyline_to_bspline3(pose_wp, W, translation_rounding, rotation_rounding,
                                pose_ctrl);
  orc_make_uniform_knots(n, 2, 0.0, 1.0, knots);
  double len_t = 0.0;
  for (int i = 0; i < n - 1; ++i) {
    const double* a = pose_ctrl + 7 * i;
    const double* b = pose_ctrl + 7 * (i + 1);
    const double dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    len_t += sqrt(dx * dx + dy * dy + dz * dz); /* PoseErrorBetween(...).translation */
  }
  /* sic: translation + translation (:432-434), rotation is not counted */
  const double kMinimumFinalKnotValue = 0.1;
  const double kPathParameterPerPolygonLength = 10.0;
  double weighted = len_t + len_t;
  if (weighted < kMinimumFinalKnotValue) weighted = kMinimumFinalKnotValue;
  const double scale = weighted * kPathParameterPerPolygonLength;
  for (int i = 0; i < n + 3; ++i) knots[i] *= scale;
  return n;
}

/* The spline part of SamplePath for a fresh path at path_start = 0
 * (timeable_path_cartesian_spline.cc:464-500): IK indices -1..N-1, i.e. N+1
 * target samples of which the first two are both parameter 0. */
void orc_cart_sample_targets(const double* knots, int num_knots, const double* pose_ctrl,
                             const double* joint_ctrl, int dim, int N, double delta,
                             double* pose_targets, double* joint_targets) {
  orc_cart_sample_targets_from(knots, num_knots, pose_ctrl, joint_ctrl, dim, N, delta, 0,
                               pose_targets, joint_targets);
}

/* The same for a later planning window: IK indices start_index-1 ..
 * start_index+N-1 with start_index = PathIkIndex(path_start) (:671-678).  The
 * reference keeps the IK solutions of earlier windows in path_ik_positions_ and
 * only evaluates the indices it has not seen (:469-483); the targets themselves
 * depend on the index alone, which is what this returns. */
void orc_cart_sample_targets_from(const double* knots, int num_knots, const double* pose_ctrl,
                                  const double* joint_ctrl, int dim, int N, double delta,
                                  int start_index, double* pose_targets,
                                  double* joint_targets) {
  const int n = num_knots - 3;
  double* trans = (double*)malloc(sizeof(double) * 3 * (size_t)n);
  double* quats = (double*)malloc(sizeof(double) * 4 * (size_t)n);
  for (int i = 0; i < n; ++i) {
    memcpy(trans + 3 * i, pose_ctrl + 7 * i, 3 * sizeof(double));
    memcpy(quats + 4 * i, pose_ctrl + 7 * i + 3, 4 * sizeof(double));
  }
  const double kend = knots[num_knots - 1];
  for (int m = 0; m <= N; ++m) {
    const int i = start_index + m - 1;
    const double parameter = (i < 0) ? 0.0 : i * delta;
    double* pt = pose_targets + 7 * (size_t)m;
    double* jt = joint_targets + (size_t)dim * m;
    if (parameter < kend - delta) {
      orc_eval_curve(knots, num_knots, joint_ctrl, dim, 2, parameter, jt);
      orc_eval_curve(knots, num_knots, trans, 3, 2, parameter, pt);
      orc_quat_eval_curve(knots, num_knots, quats, 2, parameter, pt + 3);
    } else {
      memcpy(pt, pose_ctrl + 7 * (size_t)(n - 1), 7 * sizeof(double));
      memcpy(jt, joint_ctrl + (size_t)dim * (n - 1), sizeof(double) * dim);
    }
  }
  free(trans);
  free(quats);
}

/* ComputePathDerivatives, timeable_path_cartesian_spline.cc:39-68 */
void orc_cart_derivatives(const double* pos, int N, int dim, double delta, double* d1,
                          double* d2) {
  const double inv = 1.0 / delta;
  for (int i = 0; i < N - 1; ++i)
    for (int d = 0; d < dim; ++d)
      d1[(size_t)i * dim + d] = inv * (pos[(size_t)(i + 1) * dim + d] - pos[(size_t)i * dim + d]);
  for (int d = 0; d < dim; ++d) d1[(size_t)(N - 1) * dim + d] = 0.0;
  for (int i = 1; i < N - 1; ++i)
    for (int d = 0; d < dim; ++d)
      d2[(size_t)i * dim + d] = inv * (d1[(size_t)(i + 1) * dim + d] - d1[(size_t)i * dim + d]);
  for (int d = 0; d < dim; ++d) {
    d2[d] = 0.0;
    d2[(size_t)(N - 1) * dim + d] = 0.0;
  }
}

/* ConstraintSetup, timeable_path_cartesian_spline.cc:551-595.  jac: [N][6][dim]
 * row-major (rows 0-2 linear, 3-5 angular). */
void orc_cart_constraints(const double* d1, const double* d2, const double* jac, int N,
                          int dim, const double* vmax, const double* amax, double safety,
                          double max_trans_vel, double max_rot_vel, double* cons) {
  const int C = 2 * dim + 2;
  for (int idx = 0; idx < N; ++idx) {
    double* c = cons + (size_t)idx * C * 4;
    const double* a1 = d1 + (size_t)idx * dim;
    const double* a2 = d2 + (size_t)idx * dim;
    for (int d = 0; d < dim; ++d) {
      c[d * 4 + 0] = a1[d];
      c[d * 4 + 1] = a2[d];
      c[d * 4 + 3] = amax[d] * safety;
      c[d * 4 + 2] = -amax[d] * safety;
      c[(dim + d) * 4 + 0] = 0.0;
      c[(dim + d) * 4 + 1] = pow(a1[d], 2);
      c[(dim + d) * 4 + 3] = pow(vmax[d] * safety, 2);
      c[(dim + d) * 4 + 2] = 0.0;
    }
    const double* J = jac + (size_t)idx * 6 * dim;
    double v[6];
    for (int r = 0; r < 6; ++r) {
      double s = 0.0;
      for (int d = 0; d < dim; ++d) s += J[r * dim + d] * a1[d];
      v[r] = s;
    }
    c[(2 * dim) * 4 + 0] = 0.0;
    c[(2 * dim) * 4 + 1] = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    c[(2 * dim) * 4 + 3] = pow(max_trans_vel, 2);
    c[(2 * dim) * 4 + 2] = -c[(2 * dim) * 4 + 3];
    c[(2 * dim + 1) * 4 + 0] = 0.0;
    c[(2 * dim + 1) * 4 + 1] = v[3] * v[3] + v[4] * v[4] + v[5] * v[5];
    c[(2 * dim + 1) * 4 + 3] = pow(max_rot_vel, 2);
    c[(2 * dim + 1) * 4 + 2] = -c[(2 * dim + 1) * 4 + 3];
  }
}

/* ---------------------------------------------------------------- fake kinematics
 * path_timing_trajectory_test.cc:548-587: translation = q[0:3], rotation =
 * ExpSO3(q[3:6]); IK copies q[6] from the joint target. */
void orc_exp_so3(const double* v, double* q) {
  const double th2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  const double th = sqrt(th2);
  double k; /* sin(th/2)/th */
  if (th < 1e-8) {
    k = 0.5 - th2 / 48.0;
  } else {
    k = sin(0.5 * th) / th;
  }
  q[0] = k * v[0];
  q[1] = k * v[1];
  q[2] = k * v[2];
  q[3] = cos(0.5 * th);
}
void orc_log_so3(const double* q, double* v) {
  const double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  double k; /* theta / n */
  if (n < 1e-10) {
    k = 2.0 / q[3];
  } else {
    k = 2.0 * atan2(n, q[3]) / n;
  }
  v[0] = k * q[0];
  v[1] = k * q[1];
  v[2] = k * q[2];
}
/* d LogSO3(q) / d (qx, qy, qz, qw), 3x4 row-major */
void orc_log_so3_derivative(const double* q, double* out) {
  const double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2];
  const double n = sqrt(n2);
  const double w = q[3];
  const double den = n2 + w * w;
  double k, c; /* k = theta/n; c = (2w/den - k) / n^2 */
  if (n < 1e-6) {
    k = 2.0 / w;
    c = (-4.0 / 3.0) / (w * w * w);
  } else {
    k = 2.0 * atan2(n, w) / n;
    c = (2.0 * w / den - k) / n2;
  }
  for (int r = 0; r < 3; ++r) {
    for (int col = 0; col < 3; ++col)
      out[r * 4 + col] = ((r == col) ? k : 0.0) + c * q[r] * q[col];
    out[r * 4 + 3] = -2.0 * q[r] / den;
  }
}
/* FakePathIK: ik [M][7] from pose targets [M][7] and joint targets [M][7] */
void orc_fake_ik(const double* pose_targets, const double* joint_targets, int M, double* ik) {
  for (int i = 0; i < M; ++i) {
    const double* p = pose_targets + 7 * (size_t)i;
    double* o = ik + 7 * (size_t)i;
    o[0] = p[0];
    o[1] = p[1];
    o[2] = p[2];
    orc_log_so3(p + 3, o + 3);
    o[6] = joint_targets[7 * (size_t)i + 6];
  }
}
/* FakeJacobian: 6x7 row-major */
void orc_fake_jacobian(const double* joints, double* J) {
  double q[4], dl[12];
  orc_exp_so3(joints + 3, q);
  orc_log_so3_derivative(q, dl);
  memset(J, 0, sizeof(double) * 42);
  J[0 * 7 + 0] = 1.0;
  J[1 * 7 + 1] = 1.0;
  J[2 * 7 + 2] = 1.0;
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c) J[(3 + r) * 7 + 3 + c] = dl[r * 4 + c];
}

/* Whole pipeline for one fresh Cartesian path (7 dofs, fake kinematics), single
 * window at path_start = 0: SetWaypoints, SamplePath, ConstraintSetup,
 * OptimizePathParameter, assembly, uniform resampling. */
int orc_plan_cart_fake(const orc_cart_cfg* cfg, const double* pose_wp, const double* joint_wp,
                       const double* vmax, const double* amax, orc_plan_out* out,
                       double* pose_ctrl_out, double* knots_out, double* pose_targets_out,
                       double* cons_out) {
  const int N = cfg->N, D = 7, W = cfg->W;
  const int nmax = (W == 1) ? 4 : 3 * W - 2;
  const int C = 2 * D + 2;
  double* pose_ctrl = (double*)malloc(sizeof(double) * 7 * (size_t)nmax);
  double* joint_ctrl = (double*)malloc(sizeof(double) * D * (size_t)nmax);
  double* knots = (double*)malloc(sizeof(double) * (size_t)(nmax + 3));
  double* pt = (double*)malloc(sizeof(double) * 7 * (size_t)(N + 1));
  double* jt = (double*)malloc(sizeof(double) * D * (size_t)(N + 1));
  double* ik = (double*)malloc(sizeof(double) * D * (size_t)(N + 1));
  double* jac = (double*)malloc(sizeof(double) * 42 * (size_t)N);
  double* cons = (double*)malloc(sizeof(double) * (size_t)N * C * 4);
  const int n = orc_cart_fit(pose_wp, joint_wp, W, D, cfg->translation_rounding,
                             cfg->rotation_rounding, pose_ctrl, joint_ctrl, knots);
  const int nk = n + 3;
  const double kend = knots[nk - 1];
  const double delta = (cfg->delta > 0) ? cfg->delta : kend / (N - 1);
  out->knot_end = kend;
  out->delta_used = delta;
  orc_cart_sample_targets(knots, nk, pose_ctrl, joint_ctrl, D, N, delta, pt, jt);
  orc_fake_ik(pt, jt, N + 1, ik);
  /* new_ik_path[1:] are the path positions (:512-513) */
  memcpy(out->q, ik + D, sizeof(double) * D * (size_t)N);
  orc_cart_derivatives(out->q, N, D, delta, out->dq, out->ddq);
  for (int i = 0; i < N; ++i) orc_fake_jacobian(out->q + (size_t)D * i, jac + 42 * (size_t)i);
  orc_cart_constraints(out->dq, out->ddq, jac, N, D, vmax, amax, cfg->safety,
                       cfg->max_translational_velocity, cfg->max_rotational_velocity, cons);
  if (pose_ctrl_out) memcpy(pose_ctrl_out, pose_ctrl, sizeof(double) * 7 * (size_t)n);
  if (knots_out) memcpy(knots_out, knots, sizeof(double) * (size_t)nk);
  if (pose_targets_out) memcpy(pose_targets_out, pt, sizeof(double) * 7 * (size_t)(N + 1));
  if (cons_out) memcpy(cons_out, cons, sizeof(double) * (size_t)N * C * 4);
  const double horizon = 0.0 + delta * (N - 1);
  const int max_loops = (100 > 10 * N) ? 100 : 10 * N;
  orc_solution sol;
  memset(&sol, 0, sizeof(sol));
  sol.time = out->time;
  sol.s = out->s;
  sol.sd = out->sd;
  sol.sdd = out->sdd;
  const int st = orc_solve(N, C, cons, 0.0, horizon, 0.0, 0.0, cfg->start_sec, max_loops, &sol);
  out->last_extremal_index = sol.last_extremal_index;
  out->M = 0;
  out->duration = NAN;
  if (st == ORC_OK) {
    out->duration = out->time[N - 1] - out->time[0];
    orc_assemble(out->dq, out->ddq, out->sd, out->sdd, amax, N, D, out->vel, out->acc);
    if (cfg->dt > 0 && out->r_t) {
      int M = orc_resample_count(out->time[N - 1], cfg->start_sec, cfg->dt);
      if (M > out->M_cap) M = out->M_cap;
      out->M = M;
      orc_resample_uniform(N, D, out->time, out->s, out->sd, out->sdd, out->q, out->vel,
                           out->acc, amax, cfg->start_sec, cfg->dt, M, out->r_t, out->r_pos,
                           out->r_vel, out->r_acc, out->r_s, out->r_sd, out->r_sdd);
    }
  }
  free(pose_ctrl);
  free(joint_ctrl);
  free(knots);
  free(pt);
  free(jt);
  free(ik);
  free(jac);
  free(cons);
  return st;
}

/* ---------------------------------------------------------------- online re-planning
 * TimeableCartesianSplinePath::SwitchToWaypointPath (timeable_path_cartesian_spline.cc:
 * 183-296) over the joint, translation and rotation splines:
 *   BSplineQ::InsertKnotAndUpdateControlPoints (degree <= 2: slerp of neighbouring control
 *     quaternions, splines/bsplineq.cc:355-431), TruncateSplineAt (:497-519),
 *     ExtendWithControlPoints (:521-587)
 *   BSplineT<Vector3d / VectorXd>::TruncateSplineAt / ExtendWithControlPoints: orc_spline_*
 *     (tp_oracle_switch.c; closed-form knot insertion where the reference solves a small
 *     system by QR -- same curve, control points to rounding, see that file's header)
 * The three splines keep knot vectors of their own in the reference; they stay equal, so one
 * is kept here (the joint spline's) and the other two work on copies of it. */
static void quat_slerp(const double* a, double t, const double* b, double* out) {
  const double one = 1.0 - DBL_EPSILON;
  const double d = a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3];
  const double abs_d = fabs(d);
  double scale0, scale1;
  if (abs_d >= one) {
    scale0 = 1.0 - t;
    scale1 = t;
  } else {
    const double theta = acos(abs_d);
    const double sin_theta = sin(theta);
    scale0 = sin((1.0 - t) * theta) / sin_theta;
    scale1 = sin(t * theta) / sin_theta;
  }
  if (d < 0.0) scale1 = -scale1;
  for (int i = 0; i < 4; ++i) out[i] = scale0 * a[i] + scale1 * b[i];
}

typedef struct {
  int knot_capacity, num_knots, num_points;
  double* knots; /* [knot_capacity] */
  double* quats; /* [knot_capacity - 3][4] */
} qspline;

/* one insertion, degree 2 (bsplineq.cc:355-431) */
static void qspline_insert_knot(qspline* s, double knot) {
  const int p = 2;
  double scratch[2][4];
  const int span = orc_knot_span(s->knots, s->num_knots, p, knot);
  for (int i = 0; i < p; ++i) {
    const int k = span + i - p + 1;
    const double alpha = (knot - s->knots[k]) / (s->knots[k + p] - s->knots[k]);
    quat_slerp(s->quats + 4 * (k - 1), alpha, s->quats + 4 * k, scratch[i]);
  }
  memmove(s->knots + span + 2, s->knots + span + 1,
          sizeof(double) * (size_t)(s->num_knots - span - 1));
  s->knots[span + 1] = knot;
  s->num_knots += 1;
  /* copy_backward(points + span, points + old_count, points + old_count + 1) */
  memmove(s->quats + 4 * (span + 1), s->quats + 4 * span,
          sizeof(double) * 4 * (size_t)(s->num_points - span));
  s->num_points += 1;
  memcpy(s->quats + 4 * (span - p + 1), scratch, sizeof(scratch));
}

static int qspline_truncate(qspline* s, double u_end) {
  const int p = 2;
  if (s->num_knots == 0) return 0;
  if (u_end >= s->knots[s->num_knots - 1]) return 0;
  if (u_end <= s->knots[0]) {
    s->num_knots = 0;
    s->num_points = 0;
    return 0;
  }
  /* CanInsertKnot (bspline_base.cc:166-195) */
  if (s->num_knots + p + 1 > s->knot_capacity) return 9;
  for (int i = 0; i < p + 1; ++i) qspline_insert_knot(s, u_end);
  const int span = orc_knot_span(s->knots, s->num_knots, p, u_end);
  s->num_knots = span + 1;
  s->num_points = s->num_knots - p - 1;
  return 0;
}

static int qspline_extend(qspline* s, const double* quats, int n) {
  const int p = 2;
  const int new_num_points = s->num_points + n;
  const int added_knots = (n + 1 + p + 1) - 2 * p;
  const int new_num_knots = s->num_knots + added_knots;
  if (new_num_points > s->knot_capacity - p - 1) return 9;
  if (new_num_knots > s->knot_capacity) return 9;
  if (n < 2) return 12;
  const double u_join = s->knots[s->num_knots - 1];
  double quat_join[4] = {0, 0, 0, 1};
  orc_quat_eval_curve(s->knots, s->num_knots, s->quats, p, u_join, quat_join); /* IgnoreError */
  const double old_knot_range = s->knots[s->num_knots - 1] - s->knots[0];
  const int old_inner = s->num_knots - 2 * p - 1;
  const int new_inner = new_num_knots - 2 * p - 1;
  const double new_knot_range = (old_knot_range * new_inner) / old_inner;
  const int start = s->num_knots - p - 1;
  const int size = (new_num_knots - p) - start;
  {
    const double low = old_knot_range, high = new_knot_range;
    const int size1 = (size == 1) ? 1 : size - 1;
    const double step = (size == 1) ? 0.0 : (high - low) / (double)(size - 1);
    for (int i = 0; i < size; ++i) s->knots[start + i] = (i == size1) ? high : low + (double)i * step;
  }
  for (int i = 0; i < p + 1; ++i) s->knots[new_num_knots - p - 1 + i] = s->knots[0] + new_knot_range;
  s->num_knots = new_num_knots;
  const int imod = s->num_points - 1;
  const int span = orc_knot_span(s->knots, s->num_knots, p, u_join);
  double basis[3];
  orc_basis(s->knots, span, p, u_join, basis);
  const double cum0 = basis[2] + basis[1]; /* cumulative_basis_[0] (:309-317) */
  if (!(cum0 > 0.0)) return 13;            /* CHECK_GT in the reference */
  double inv[4], rel[4], lg[4], ex[4], np_[4];
  quat_inverse(s->quats + 4 * (imod - 1), inv);
  orc_quat_mul(inv, quat_join, rel);
  orc_quat_log(rel, lg);
  for (int i = 0; i < 4; ++i) lg[i] /= cum0;
  if (lg[3] < 0) {
    for (int i = 0; i < 4; ++i) lg[i] *= -1.0;
  }
  orc_quat_exp(lg, ex);
  orc_quat_mul(s->quats + 4 * (imod - 1), ex, np_);
  memcpy(s->quats + 4 * imod, np_, sizeof(np_));
  memcpy(s->quats + 4 * s->num_points, quats, sizeof(double) * 4 * (size_t)n);
  s->num_points = new_num_points;
  return 0;
}

/* joint: the joint spline (knots + joint control points); pose_ctrl [knot_capacity - 3][7]:
 * control points of the translation / rotation splines.  Returns an absl::StatusCode number.
 * Scratch is sized for ORC_MAX_SWITCH_WAYPOINTS new waypoints and joint->knot_capacity. */
int orc_cart_switch(orc_spline* joint, double* pose_ctrl, double keep_path_until,
                    const double* pose_waypoints, const double* joint_waypoints, int W,
                    double translation_rounding, double rotation_rounding) {
  const int D = joint->dim, cap = joint->knot_capacity, pcap = cap - 3;
  if (W > ORC_MAX_SWITCH_WAYPOINTS) return 3;
  double* tk = (double*)malloc(sizeof(double) * (size_t)cap * 2);
  double* rk = tk + cap;
  double* tp = (double*)malloc(sizeof(double) * (size_t)pcap * 7);
  double* rq = tp + (size_t)pcap * 3;
  memcpy(tk, joint->knots, sizeof(double) * (size_t)cap);
  memcpy(rk, joint->knots, sizeof(double) * (size_t)cap);
  for (int i = 0; i < joint->num_points; ++i) {
    memcpy(tp + 3 * i, pose_ctrl + 7 * i, sizeof(double) * 3);
    memcpy(rq + 4 * i, pose_ctrl + 7 * i + 3, sizeof(double) * 4);
  }
  orc_spline ts = {2, 3, cap, joint->num_knots, joint->num_points, tk, tp};
  qspline rs = {cap, joint->num_knots, joint->num_points, rk, rq};
  int st = orc_spline_truncate(joint, keep_path_until);
  if (st == 0) st = qspline_truncate(&rs, keep_path_until);
  if (st == 0) st = orc_spline_truncate(&ts, keep_path_until);
  double jsw[64], qsw[4] = {0, 0, 0, 1}, tsw[3] = {0, 0, 0};
  if (st == 0 && joint->num_knots == 0) st = 11; /* EvalCurve on a cleared curve */
  if (st == 0)
    st = orc_eval_curve(joint->knots, joint->num_knots, joint->points, D, 2, keep_path_until, jsw);
  if (st == 0) st = orc_quat_eval_curve(rs.knots, rs.num_knots, rs.quats, 2, keep_path_until, qsw);
  if (st == 0) st = orc_eval_curve(ts.knots, ts.num_knots, ts.points, 3, 2, keep_path_until, tsw);
  double jproj[64], tproj[3], jdist = 0, jt = 0, tdist = 0, tt = 0;
  int jidx = 0, tidx = 0;
  if (st == 0)
    st = orc_project_point_on_path(joint_waypoints, W, D, jsw, &jidx, &jdist, &jt, jproj);
  double twp[ORC_MAX_SWITCH_WAYPOINTS * 3];
  for (int i = 0; i < W; ++i) memcpy(twp + 3 * i, pose_waypoints + 7 * i, sizeof(double) * 3);
  if (st == 0) st = orc_project_point_on_path(twp, W, 3, tsw, &tidx, &tdist, &tt, tproj);
  double njw[(ORC_MAX_SWITCH_WAYPOINTS + 1) * 64], npw[(ORC_MAX_SWITCH_WAYPOINTS + 1) * 7];
  int n = 0;
  if (st == 0) {
    double linf = 0.0;
    for (int d = 0; d < 3; ++d) {
      const double e = fabs(tsw[d] - tproj[d]);
      if (e > linf) linf = e;
    }
    if (linf > 1e-3) { /* kEpsilon */
      if (tidx + 1 >= W) {
        st = 3; /* the reference would read pose_waypoints[W] */
      } else {
        memcpy(njw, jproj, sizeof(double) * (size_t)D);
        memcpy(npw, tproj, sizeof(double) * 3);
        quat_slerp(pose_waypoints + 7 * tidx + 3, tt, pose_waypoints + 7 * (tidx + 1) + 3, npw + 3);
        n = 1;
      }
    }
  }
  if (st == 0) {
    const int first = (tt >= 0) ? tidx + 1 : tidx;
    for (int i = first; i < W; ++i) {
      memcpy(njw + (size_t)n * D, joint_waypoints + (size_t)i * D, sizeof(double) * (size_t)D);
      memcpy(npw + (size_t)n * 7, pose_waypoints + (size_t)i * 7, sizeof(double) * 7);
      ++n;
    }
    if (n == 0) st = 3; /* the reference would index an empty vector */
  }
  if (st == 0) {
    const int ncmax = (n == 1) ? 4 : 3 * n - 2;
    double* jc = (double*)malloc(sizeof(double) * (size_t)ncmax * (D + 7 + 3 + 4));
    double* pc = jc + (size_t)ncmax * D;
    double* tc = pc + (size_t)ncmax * 7;
    double* qc = tc + (size_t)ncmax * 3;
    const int nc = orc_polyline_to_bspline3(njw, n, D, rotation_rounding, jc);
    orc_pose_polyline_to_bspline3(npw, n, translation_rounding, rotation_rounding, pc);
    for (int i = 0; i < nc; ++i) {
      memcpy(tc + 3 * i, pc + 7 * i, sizeof(double) * 3);
      memcpy(qc + 4 * i, pc + 7 * i + 3, sizeof(double) * 4);
    }
    st = orc_spline_extend(joint, jc, nc);
    if (st == 0) st = orc_spline_extend(&ts, tc, nc);
    if (st == 0) st = qspline_extend(&rs, qc, nc);
    free(jc);
  }
  /* the splines' control points as they stand, also after a failure */
  const int np_out = (ts.num_points > rs.num_points) ? ts.num_points : rs.num_points;
  for (int i = 0; i < np_out && i < pcap; ++i) {
    if (i < ts.num_points) memcpy(pose_ctrl + 7 * i, tp + 3 * i, sizeof(double) * 3);
    if (i < rs.num_points) memcpy(pose_ctrl + 7 * i + 3, rq + 4 * i, sizeof(double) * 4);
  }
  free(tk);
  free(tp);
  return st;
}
