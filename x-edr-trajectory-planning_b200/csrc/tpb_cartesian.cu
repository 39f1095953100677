// Cartesian timeable path (BASELINE config 5) for sm_100a, compiled with
// -fmad=false like the rest of the path-timing kernels.
//
//  k_cart_fit             one thread per path: pose + joint waypoints -> control
//                         polygons and knots (spline_utils.cc:47-203,
//                         timeable_path_cartesian_spline.cc:396-462)
//  k_cart_sample_targets  one thread per (path, IK index): joint / translation /
//                         quaternion spline EvalCurve (bspline.h:514-536,
//                         bsplineq.cc:113-146, 223-244, 309-317) -> the pose and
//                         joint targets SamplePath hands to the path IK
//                         (timeable_path_cartesian_spline.cc:464-500)
//  k_cart_ik_builtin      built-in device kinematics: the IK of the reference
//                         test's fake robot (path_timing_trajectory_test.cc:548-570)
//  k_cart_constraints     one thread per (path, sample): forward-difference path
//                         derivatives (:39-68) and the 2D+2 constraint rows of
//                         ConstraintSetup (:551-595), Jacobian either uploaded by
//                         the host (user JacobianFunc) or of the built-in model
//
// A user-supplied PathIKFunc / JacobianFunc is host code with sequential
// warm-start semantics: the host layer downloads the targets, runs the user's
// callbacks and uploads joint positions / Jacobians (tpb_cart_set_ik_positions,
// tpb_cart_setup_constraints); everything else stays on the device.
#include <cfloat>
#include <cstdlib>

#include "tpb_engine.h"
#include "tpb_fit.cuh"
#include "tpb_joint_limit.cuh"
#include "tpb_quat.cuh"

namespace tpb {
namespace {

// (quaternion arithmetic of splines/bsplineq.cc: tpb_quat.cuh)
__device__ __forceinline__ void store_quat(double* p, const Quat& q) {
  p[0] = q.x;
  p[1] = q.y;
  p[2] = q.z;
  p[3] = q.w;
}

// poses: 7 doubles (tx, ty, tz, qx, qy, qz, qw); a * b = (qa*qb, ta + qa*tb)
__device__ inline void pose_mul(const double* a, const double* b, double* out) {
  const Quat qa = load_quat(a + 3);
  double r[3];
  qrot(qa, b, r);
  const Quat q = qmul(qa, load_quat(b + 3));
  out[0] = a[0] + r[0];
  out[1] = a[1] + r[1];
  out[2] = a[2] + r[2];
  store_quat(out + 3, q);
}
__device__ inline void pose_inverse(const double* a, double* out) {
  const Quat qc{-a[3], -a[4], -a[5], a[6]};
  double r[3];
  qrot(qc, a, r);
  out[0] = -r[0];
  out[1] = -r[1];
  out[2] = -r[2];
  store_quat(out + 3, qc);
}

// CornerOffset (Pose3d), splines/spline_utils.cc:105-146, with Eigen's
// AngleAxis <-> Quaternion conversions.
__device__ inline void pose_corner_offset(const double* delta, double translation_radius,
                                          double rotation_radius, double* offset) {
  constexpr double kMinRadius = 1e-6;
  constexpr double kMinWaypointSpacingFactor = 4.0;
  offset[0] = offset[1] = offset[2] = 0.0;
  offset[3] = offset[4] = offset[5] = 0.0;
  offset[6] = 1.0;
  if (translation_radius < kMinRadius || rotation_radius < kMinRadius) return;
  const double translation_norm =
      sqrt(delta[0] * delta[0] + delta[1] * delta[1] + delta[2] * delta[2]);
  const Quat q = load_quat(delta + 3);
  double n = sqrt(q.x * q.x + q.y * q.y + q.z * q.z);
  if (n < DBL_EPSILON) n = stable_norm3(q);
  double angle, ax, ay, az;
  if (n != 0.0) {
    angle = 2.0 * atan2(n, fabs(q.w));
    if (q.w < 0.0) n = -n;
    ax = q.x / n;
    ay = q.y / n;
    az = q.z / n;
  } else {
    angle = 0.0;
    ax = 1.0;
    ay = 0.0;
    az = 0.0;
  }
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  const double pct_trans = (translation_norm == 0.0) ? inf : translation_radius / translation_norm;
  const double pct_rot = (angle == 0.0) ? inf : rotation_radius / angle;
  double pct = (pct_rot < pct_trans) ? pct_rot : pct_trans;
  if (pct > (1.0 / kMinWaypointSpacingFactor)) pct = 1.0 / kMinWaypointSpacingFactor;
  offset[0] = delta[0] * pct;
  offset[1] = delta[1] * pct;
  offset[2] = delta[2] * pct;
  angle *= pct;
  const double ha = 0.5 * angle;
  const double s = sin(ha);
  offset[6] = cos(ha);
  offset[3] = s * ax;
  offset[4] = s * ay;
  offset[5] = s * az;
}

// PolyLineToBspline3Waypoints (Pose3d), splines/spline_utils.cc:148-203
__device__ inline int pose_polyline_to_bspline3(const double* corners, int W, double tr,
                                                double rr, double* out) {
  if (W == 1) {
    for (int i = 0; i < 4; ++i)
      for (int k = 0; k < 7; ++k) out[7 * i + k] = corners[k];
    return 4;
  }
  const int n = 3 * W - 2;
  double inv[7], rel[7], off[7];
  for (int i = 0; i < W; ++i)
    for (int k = 0; k < 7; ++k) out[7 * 3 * i + k] = corners[7 * i + k];
  for (int idx = 1; idx < W - 1; ++idx) {
    const int k = 3 * idx, knext = 3 * (idx + 1), klast = 3 * (idx - 1);
    pose_inverse(out + 7 * k, inv);
    pose_mul(inv, out + 7 * knext, rel);
    pose_corner_offset(rel, tr, rr, off);
    pose_mul(out + 7 * k, off, out + 7 * (k + 1));
    pose_mul(inv, out + 7 * klast, rel);
    pose_corner_offset(rel, tr, rr, off);
    pose_mul(out + 7 * k, off, out + 7 * (k - 1));
  }
  pose_inverse(out, inv);
  pose_mul(inv, out + 7 * 3, rel);
  pose_corner_offset(rel, tr, rr, off);
  pose_mul(out, off, out + 7);
  pose_inverse(out + 7 * (n - 1), inv);
  pose_mul(inv, out + 7 * (n - 4), rel);
  pose_corner_offset(rel, tr, rr, off);
  pose_mul(out + 7 * (n - 1), off, out + 7 * (n - 2));
  return n;
}

__global__ void k_cart_fit(Dims dm, DeviceArrays a, double rotation_rounding,
                           double translation_rounding, bool auto_delta) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B) return;
  const int D = dm.D;
  const int W = min(max(a.num_wp[b], 1), dm.Wmax);  // see k_joint_fit
  const double* jwp = a.waypoints + static_cast<size_t>(b) * dm.Wmax * D;
  const double* pwp = a.pose_wp + static_cast<size_t>(b) * dm.Wmax * 7;
  double* jc = a.ctrl + static_cast<size_t>(b) * dm.max_points * D;
  double* pc = a.pose_ctrl + static_cast<size_t>(b) * dm.max_points * 7;
  double* kn = a.knots + static_cast<size_t>(b) * (dm.max_points + 3);
  // FitSplineToWaypoints, timeable_path_cartesian_spline.cc:396-462
  const int n = polyline_to_bspline3(jwp, W, D, rotation_rounding, jc);
  pose_polyline_to_bspline3(pwp, W, translation_rounding, rotation_rounding, pc);
  const int nk = uniform_knots_degree2(n, kn);
  double len_t = 0.0;
  for (int i = 0; i < n - 1; ++i) {
    const double* p0 = pc + 7 * i;
    const double* p1 = pc + 7 * (i + 1);
    const double dx = p0[0] - p1[0], dy = p0[1] - p1[1], dz = p0[2] - p1[2];
    len_t += sqrt(dx * dx + dy * dy + dz * dz);
  }
  // sic: translation + translation (:432-434); rotation is not counted
  constexpr double kMinimumFinalKnotValue = 0.1;
  constexpr double kPathParameterPerPolygonLength = 10.0;
  double weighted = len_t + len_t;
  if (weighted < kMinimumFinalKnotValue) weighted = kMinimumFinalKnotValue;
  const double scale = weighted * kPathParameterPerPolygonLength;
  for (int i = 0; i < nk; ++i) kn[i] *= scale;
  a.num_points[b] = n;
  a.knot_cap[b] = max(2 * nk, 100);  // timeable_path_cartesian_spline.cc:404-413
  const double kend = kn[nk - 1];
  a.knot_end[b] = kend;
  a.delta[b] = auto_delta ? kend / (dm.N - 1) : a.delta_in[b];
}

// ---------------------------------------------------------------------------
// TimeableCartesianSplinePath::SwitchToWaypointPath (timeable_path_cartesian_spline.cc:
// 183-296), one thread per path, in place on the path's knots[] / ctrl[] / pose_ctrl[]
// (which hold the reference's knot capacity by now).  The joint, translation and rotation
// splines keep equal knot vectors in the reference, so one insertion step serves all
// three: lerp for the two vector splines (bspline.h:244-278, the closed form, see
// tpb_switch.cu), slerp for the quaternions (bsplineq.cc:355-431).
// ---------------------------------------------------------------------------
enum : int32_t { kSwOk = 0, kSwInvalidArgument = 3, kSwFailedPrecondition = 9, kSwOutOfRange = 11,
                 kSwInternal = 13 };

__device__ int span2(const double* knots, int num_knots, double u) {
  if (u == knots[num_knots - 1]) return num_knots - 4;  // num_points - 1
  int lo = 2, hi = num_knots - 2;
  while (lo < hi) {
    const int mid = lo + ((hi - lo) >> 1);
    if (knots[mid] <= u) {
      lo = mid + 1;
    } else {
      hi = mid;
    }
  }
  return lo - 1;
}

__device__ void cart_insert_knot(double* knots, double* pts, double* pose, int D, int& num_knots,
                                 int& num_points, double knot) {
  double sj[2][kMaxDofs], sp[2][7];
  const int span = span2(knots, num_knots, knot);
  for (int i = 0; i < 2; ++i) {
    const int k = span + i - 1;
    const double alpha = (knot - knots[k]) / (knots[k + 2] - knots[k]);
    for (int d = 0; d < D; ++d) sj[i][d] = alpha * pts[k * D + d] + (1.0 - alpha) * pts[(k - 1) * D + d];
    for (int d = 0; d < 3; ++d) sp[i][d] = alpha * pose[k * 7 + d] + (1.0 - alpha) * pose[(k - 1) * 7 + d];
    store_quat(sp[i] + 3, qslerp(load_quat(pose + (k - 1) * 7 + 3), alpha, load_quat(pose + k * 7 + 3)));
  }
  for (int i = num_points - 1; i >= span; --i) {  // copy_backward
    for (int d = 0; d < D; ++d) pts[(i + 1) * D + d] = pts[i * D + d];
    for (int d = 0; d < 7; ++d) pose[(i + 1) * 7 + d] = pose[i * 7 + d];
  }
  for (int i = 0; i < 2; ++i) {
    for (int d = 0; d < D; ++d) pts[(span - 1 + i) * D + d] = sj[i][d];
    for (int d = 0; d < 7; ++d) pose[(span - 1 + i) * 7 + d] = sp[i][d];
  }
  for (int i = num_knots - 1; i >= span + 1; --i) knots[i + 1] = knots[i];
  knots[span + 1] = knot;
  num_knots += 1;
  num_points += 1;
}

// ProjectPointOnPath (path_tools.h:59-104) over rows `stride` doubles apart
__device__ void project_on_polyline(const double* wps, int stride, int W, int dim, const double* point,
                                    int* index, double* line_parameter, double* projected) {
  *index = 0;
  *line_parameter = 0.0;
  if (W == 1) {
    for (int d = 0; d < dim; ++d) projected[d] = wps[d];
    return;
  }
  double best = DBL_MAX;
  for (int i = 0; i < W - 1; ++i) {
    const double* p0 = wps + i * stride;
    const double* p1 = p0 + stride;
    double ab2 = 0.0, dot = 0.0;
    for (int d = 0; d < dim; ++d) {
      ab2 += (p1[d] - p0[d]) * (p1[d] - p0[d]);
      dot += (point[d] - p0[d]) * (p1[d] - p0[d]);
    }
    double t = (ab2 > 0.0) ? dot / ab2 : 0.0;
    t = (t < 0.0) ? 0.0 : ((t > 1.0) ? 1.0 : t);
    double sum = 0.0;
    for (int d = 0; d < dim; ++d) {
      const double e = p0[d] + t * (p1[d] - p0[d]) - point[d];
      sum += e * e;
    }
    const double dist = sqrt(sum);
    if (dist < best) {
      best = dist;
      *line_parameter = t;
      *index = i;
    }
  }
  const double* p0 = wps + *index * stride;
  for (int d = 0; d < dim; ++d) projected[d] = p0[d] + *line_parameter * (p0[stride + d] - p0[d]);
}

__global__ void k_cart_switch(Dims dm, DeviceArrays a, SwitchArgs s, double rotation_rounding,
                              double translation_rounding) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B) return;
  const int D = dm.D;
  const int W = s.num_waypoints ? s.num_waypoints[b] : s.max_waypoints;
  if (W <= 0) {  // this path keeps its splines
    s.status[b] = kSwOk;
    s.modified[b] = 0;
    return;
  }
  s.modified[b] = 1;  // kModifiedPath, also when the switch fails (:194)
  double* knots = a.knots + static_cast<size_t>(b) * (dm.max_points + 3);
  double* pts = a.ctrl + static_cast<size_t>(b) * dm.max_points * D;
  double* pose = a.pose_ctrl + static_cast<size_t>(b) * dm.max_points * 7;
  const double* jw = s.waypoints + static_cast<size_t>(b) * s.max_waypoints * D;
  const double* pw = s.pose_waypoints + static_cast<size_t>(b) * s.max_waypoints * 7;
  double* jlist = a.waypoints + static_cast<size_t>(b) * dm.Wmax * D;  // new waypoint lists
  double* plist = a.pose_wp + static_cast<size_t>(b) * dm.Wmax * 7;
  const int cap = a.knot_cap[b];
  int num_points = a.num_points[b];
  int num_knots = (num_points > 0) ? num_points + 3 : 0;
  const double keep = s.keep_path_until[b];
  int32_t st = kSwOk;

  // ---- TruncateSplineAt of the three splines (bspline.h:402-427, bsplineq.cc:497-519)
  if (num_knots > 0 && !(keep >= knots[num_knots - 1])) {
    if (keep <= knots[0]) {
      num_knots = 0;
      num_points = 0;
    } else if (num_knots + 3 > cap) {
      st = kSwFailedPrecondition;  // CanInsertKnot (bspline_base.cc:166-195)
    } else {
      for (int i = 0; i < 3; ++i) cart_insert_knot(knots, pts, pose, D, num_knots, num_points, keep);
      const int span = span2(knots, num_knots, keep);
      num_knots = span + 1;
      num_points = num_knots - 3;
    }
  }
  // ---- EvalCurve(keep_path_until) of the joint and the translation spline (:203-212; the
  // rotation's value is not used afterwards and its range check is the same)
  double jsw[kMaxDofs], tsw[3];
  if (st == kSwOk) {
    if (num_knots == 0 || keep < knots[0] || keep > knots[num_knots - 1]) {
      st = kSwOutOfRange;
    } else {
      const int span = span2(knots, num_knots, keep);
      double basis[3];
      basis_funs<2>(knots, span, keep, basis);
      for (int d = 0; d < D; ++d) jsw[d] = 0.0;
      for (int d = 0; d < 3; ++d) tsw[d] = 0.0;
      for (int i = 0; i <= 2; ++i) {
        for (int d = 0; d < D; ++d) jsw[d] += basis[i] * pts[(span - 2 + i) * D + d];
        for (int d = 0; d < 3; ++d) tsw[d] += basis[i] * pose[(span - 2 + i) * 7 + d];
      }
    }
  }
  if (st == kSwOk) {
    int jidx, tidx;
    double jt, tt, jproj[kMaxDofs], tproj[3];
    project_on_polyline(jw, D, W, D, jsw, &jidx, &jt, jproj);
    project_on_polyline(pw, 7, W, 3, tsw, &tidx, &tt, tproj);
    // ---- the new waypoint lists (:238-276)
    int n = 0;
    double linf = 0.0;
    for (int d = 0; d < 3; ++d) linf = fmax(linf, fabs(tsw[d] - tproj[d]));
    if (linf > 1e-3) {
      if (tidx + 1 >= W) {
        st = kSwInvalidArgument;  // the reference would read pose_waypoints[W]
      } else {
        for (int d = 0; d < D; ++d) jlist[d] = jproj[d];
        for (int d = 0; d < 3; ++d) plist[d] = tproj[d];
        store_quat(plist + 3, qslerp(load_quat(pw + tidx * 7 + 3), tt, load_quat(pw + (tidx + 1) * 7 + 3)));
        n = 1;
      }
    }
    if (st == kSwOk) {
      const int first = (tt >= 0) ? tidx + 1 : tidx;
      for (int i = first; i < W; ++i) {
        for (int d = 0; d < D; ++d) jlist[n * D + d] = jw[i * D + d];
        for (int d = 0; d < 7; ++d) plist[n * 7 + d] = pw[i * 7 + d];
        ++n;
      }
      // ---- control polygons + ExtendWithControlPoints of the three splines
      // (bspline.h:429-511, bsplineq.cc:521-587)
      const int nc = (n == 1) ? 4 : 3 * n - 2;
      const int new_num_points = num_points + nc;
      const int new_num_knots = num_knots + (nc + 4) - 4;
      if (n == 0) {
        st = kSwInvalidArgument;  // the reference would index an empty vector
      } else if (num_knots < 6 || new_num_points > cap - 3 || new_num_knots > cap) {
        st = kSwFailedPrecondition;
      } else {
        polyline_to_bspline3(jlist, n, D, rotation_rounding, pts + num_points * D);
        pose_polyline_to_bspline3(plist, n, translation_rounding, rotation_rounding,
                                  pose + num_points * 7);
        const double u_join = knots[num_knots - 1];
        // the rotation curve's value at the join, on the truncated spline (:548-549)
        Quat q_join;
        {
          double basis[3];
          const int span = span2(knots, num_knots, u_join);
          basis_funs<2>(knots, span, u_join, basis);
          const double c1 = basis[2], c0 = c1 + basis[1];
          const Quat p0 = load_quat(pose + (span - 2) * 7 + 3);
          const Quat p1 = load_quat(pose + (span - 1) * 7 + 3);
          const Quat p2 = load_quat(pose + span * 7 + 3);
          q_join = qmul(qmul(p0, qpow(qmul(qinv(p0), p1), c0)), qpow(qmul(qinv(p1), p2), c1));
          normalize_positive_real(q_join);
        }
        const double old_knot_range = knots[num_knots - 1] - knots[0];
        const int old_inner = num_knots - 5;
        const int new_inner = new_num_knots - 5;
        const double new_knot_range = (old_knot_range * new_inner) / old_inner;
        const int start = num_knots - 3;
        const int size = (new_num_knots - 2) - start;
        {  // Eigen's setLinSpaced(size, low, high) for |high| >= |low|
          const double low = old_knot_range, high = new_knot_range;
          const int size1 = (size == 1) ? 1 : size - 1;
          const double step = (size == 1) ? 0.0 : (high - low) / static_cast<double>(size - 1);
          for (int i = 0; i < size; ++i) {
            knots[start + i] = (i == size1) ? high : low + static_cast<double>(i) * step;
          }
        }
        for (int i = 0; i < 3; ++i) knots[new_num_knots - 3 + i] = knots[0] + new_knot_range;
        // the old end points move so that the curves keep their values at the join
        const int join_span = span2(knots, new_num_knots, u_join);
        const int mod = num_points - 1;
        double basis[3];
        basis_funs<2>(knots, join_span, u_join, basis);
        const double cum0 = basis[2] + basis[1];
        if (!(basis[1] > 0) || !(cum0 > 0.0)) {
          st = kSwInternal;
        } else {
          for (int d = 0; d < D; ++d) {
            pts[mod * D + d] = 1.0 / (basis[1]) * (pts[mod * D + d] - basis[0] * pts[(mod - 1) * D + d]);
          }
          for (int d = 0; d < 3; ++d) {
            pose[mod * 7 + d] = 1.0 / (basis[1]) * (pose[mod * 7 + d] - basis[0] * pose[(mod - 1) * 7 + d]);
          }
          const Quat pm = load_quat(pose + (mod - 1) * 7 + 3);
          Quat lg = qlog(qmul(qinv(pm), q_join));
          lg.x /= cum0;
          lg.y /= cum0;
          lg.z /= cum0;
          lg.w /= cum0;
          if (lg.w < 0) {  // (sic, :575-577: the whole logarithm is negated)
            lg.x *= -1.0;
            lg.y *= -1.0;
            lg.z *= -1.0;
            lg.w *= -1.0;
          }
          store_quat(pose + mod * 7 + 3, qmul(pm, qexp(lg)));
        }
        num_knots = new_num_knots;
        num_points = new_num_points;
      }
    }
  }
  a.num_points[b] = num_points;
  if (num_knots > 0) a.knot_end[b] = knots[num_knots - 1];
  s.status[b] = st;
}

// [B][old_rows][width] -> [B][new_rows][width]
__global__ void k_restride_rows(int width, int old_rows, int new_rows, const double* __restrict__ in,
                                double* __restrict__ out) {
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < old_rows * width; i += blockDim.x) {
    out[static_cast<size_t>(b) * new_rows * width + i] = in[static_cast<size_t>(b) * old_rows * width + i];
  }
}

constexpr int kCartThreads = 128;

// IK index i = PathIkIndex(path_start) + m - 1 for m in [0, N] (:469-483, 671-678);
// for a fresh path parameter 0 is evaluated twice (:478-482).  The reference keeps the
// IK solutions of earlier windows and evaluates only the indices it has not seen; the
// targets depend on the index alone.
__global__ void __launch_bounds__(kCartThreads)
k_cart_sample_targets(Dims dm, DeviceArrays a) {
  extern __shared__ double smem[];
  const int b = path_of_block();
  if (b >= dm.B) return;  // uniform per CTA (grid folded over y and z)
  if (a.active != nullptr && a.active[b] == 0) return;  // uniform per CTA
  const int D = dm.D;
  const int n = a.num_points[b];
  const int nk = n + 3;
  double* s_knots = smem;                        // [max_points + 3]
  double* s_jc = s_knots + dm.max_points + 3;    // [max_points][D]
  double* s_pc = s_jc + dm.max_points * D;       // [max_points][7]
  {
    const double* gk = a.knots + static_cast<size_t>(b) * (dm.max_points + 3);
    const double* gj = a.ctrl + static_cast<size_t>(b) * dm.max_points * D;
    const double* gp = a.pose_ctrl + static_cast<size_t>(b) * dm.max_points * 7;
    for (int i = threadIdx.x; i < nk; i += blockDim.x) s_knots[i] = gk[i];
    for (int i = threadIdx.x; i < n * D; i += blockDim.x) s_jc[i] = gj[i];
    for (int i = threadIdx.x; i < n * 7; i += blockDim.x) s_pc[i] = gp[i];
  }
  __syncthreads();
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m > dm.N) return;
  const double delta = a.delta[b];
  const int start_index = static_cast<int>(round(a.path_start[b] / delta));  // PathIkIndex
  const int i = start_index + m - 1;
  const double parameter = (i < 0) ? 0.0 : i * delta;
  const double kend = s_knots[nk - 1];
  double* pt = a.pose_targets + (static_cast<size_t>(b) * (dm.N + 1) + m) * 7;
  double* jt = a.joint_targets + (static_cast<size_t>(b) * (dm.N + 1) + m) * D;
  if (parameter < kend - delta) {
    // EvalCurve range check (bspline.h:519-522) cannot fail: 0 <= parameter < kend
    const int span = knot_span(s_knots, nk, 2, parameter);
    double basis[3];
    basis_funs<2>(s_knots, span, parameter, basis);
    for (int d = 0; d < D; ++d) {
      double v = 0.0;
#pragma unroll
      for (int k = 0; k <= 2; ++k) v += basis[k] * s_jc[(span - 2 + k) * D + d];
      jt[d] = v;
    }
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      double v = 0.0;
#pragma unroll
      for (int k = 0; k <= 2; ++k) v += basis[k] * s_pc[(span - 2 + k) * 7 + d];
      pt[d] = v;
    }
    // BSplineQ::EvalCurve (bsplineq.cc:223-244), cumulative basis (:309-317)
    double cum[2];
    cum[1] = basis[2];
    cum[0] = cum[1] + basis[1];
    Quat q = load_quat(s_pc + (span - 2) * 7 + 3);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const Quat p0 = load_quat(s_pc + (span - 2 + k) * 7 + 3);
      const Quat p1 = load_quat(s_pc + (span - 2 + k + 1) * 7 + 3);
      q = qmul(q, qpow(qmul(qinv(p0), p1), cum[k]));
    }
    normalize_positive_real(q);
    store_quat(pt + 3, q);
  } else {
    for (int k = 0; k < 7; ++k) pt[k] = s_pc[(n - 1) * 7 + k];
    for (int d = 0; d < D; ++d) jt[d] = s_jc[(n - 1) * D + d];
  }
}

// ---- built-in kinematics model TPB_KIN_FAKE7 (path_timing_trajectory_test.cc:548-587)
__device__ __forceinline__ Quat exp_so3(const double* v) {
  const double th2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  const double th = sqrt(th2);
  double k;
  if (th < 1e-8) {
    k = 0.5 - th2 / 48.0;
  } else {
    k = sin(0.5 * th) / th;
  }
  return Quat{k * v[0], k * v[1], k * v[2], cos(0.5 * th)};
}
__device__ __forceinline__ void log_so3(const Quat& q, double* v) {
  const double n = sqrt(q.x * q.x + q.y * q.y + q.z * q.z);
  double k;
  if (n < 1e-10) {
    k = 2.0 / q.w;
  } else {
    k = 2.0 * atan2(n, q.w) / n;
  }
  v[0] = k * q.x;
  v[1] = k * q.y;
  v[2] = k * q.z;
}
// d LogSO3(q) / d (qx, qy, qz, qw), 3x4 row-major
__device__ __forceinline__ void log_so3_derivative(const Quat& q, double* out) {
  const double qv[3] = {q.x, q.y, q.z};
  const double n2 = q.x * q.x + q.y * q.y + q.z * q.z;
  const double n = sqrt(n2);
  const double w = q.w;
  const double den = n2 + w * w;
  double k, c;
  if (n < 1e-6) {
    k = 2.0 / w;
    c = (-4.0 / 3.0) / (w * w * w);
  } else {
    k = 2.0 * atan2(n, w) / n;
    c = (2.0 * w / den - k) / n2;
  }
  for (int r = 0; r < 3; ++r) {
    for (int col = 0; col < 3; ++col)
      out[r * 4 + col] = ((r == col) ? k : 0.0) + c * qv[r] * qv[col];
    out[r * 4 + 3] = -2.0 * qv[r] / den;
  }
}

// FakePathIK: path position i = IK of target i + 1 (new_ik_path[1:], :512-513)
__global__ void k_cart_ik_fake7(Dims dm, DeviceArrays a) {
  const int b = path_of_block();
  if (b >= dm.B) return;  // uniform per CTA (grid folded over y and z)
  if (a.active != nullptr && a.active[b] == 0) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= dm.N) return;
  const double* pt = a.pose_targets + (static_cast<size_t>(b) * (dm.N + 1) + i + 1) * 7;
  const double* jt = a.joint_targets + (static_cast<size_t>(b) * (dm.N + 1) + i + 1) * 7;
  double* o = a.q + (static_cast<size_t>(b) * dm.N + i) * 7;
  o[0] = pt[0];
  o[1] = pt[1];
  o[2] = pt[2];
  log_so3(load_quat(pt + 3), o + 3);
  o[6] = jt[6];
}

// ComputePathDerivatives (:39-68) + ConstraintSetup (:551-595)
template <bool kFake7>
__global__ void __launch_bounds__(kCartThreads)
k_cart_constraints(Dims dm, DeviceArrays a, double safety, double max_trans_vel,
                   double max_rot_vel) {
  const int b = path_of_block();
  if (b >= dm.B) return;  // uniform per CTA (grid folded over y and z)
  if (a.active != nullptr && a.active[b] == 0) return;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int N = dm.N, D = dm.D, C = dm.C;
  if (idx >= N) return;
  const size_t row = (static_cast<size_t>(b) * N + idx) * D;
  const double delta = a.delta[b];
  const double inv = 1.0 / delta;
  double d1[kMaxDofs], d2[kMaxDofs], p0[kMaxDofs];
  for (int d = 0; d < D; ++d) {
    p0[d] = a.q[row + d];
    double f1 = 0.0, f2 = 0.0;
    if (idx < N - 1) {
      const double p1 = a.q[row + D + d];
      f1 = inv * (p1 - p0[d]);
      if (idx >= 1) {
        // first derivative of the next sample (0 at the last one)
        double g1 = 0.0;
        if (idx + 1 < N - 1) g1 = inv * (a.q[row + 2 * D + d] - p1);
        f2 = inv * (g1 - f1);
      }
    }
    d1[d] = f1;
    d2[d] = f2;
    a.dq[row + d] = f1;
    a.ddq[row + d] = f2;
  }
  double* c = a.cons + (static_cast<size_t>(b) * N + idx) * C * 4;
  for (int d = 0; d < D; ++d) {
    const double am = a.amax[static_cast<size_t>(b) * D + d];
    const double vm = a.vmax[static_cast<size_t>(b) * D + d];
    c[d * 4 + 0] = d1[d];
    c[d * 4 + 1] = d2[d];
    c[d * 4 + 2] = -am * safety;
    c[d * 4 + 3] = am * safety;
    const double vs = vm * safety;
    c[(D + d) * 4 + 0] = 0.0;
    c[(D + d) * 4 + 1] = d1[d] * d1[d];
    c[(D + d) * 4 + 2] = 0.0;
    c[(D + d) * 4 + 3] = vs * vs;
  }
  double v[6];
  if (kFake7) {
    // FakeJacobian: rows 0-2 = identity on q[0:3]; rows 3-5, columns 3-6 =
    // d LogSO3 / d quaternion coefficients at ExpSO3(q[3:6])
    double dl[12];
    log_so3_derivative(exp_so3(p0 + 3), dl);
    for (int r = 0; r < 3; ++r) {
      // full row sums like the dense product jacobian * q' (zeros included)
      double s = 0.0;
      for (int d = 0; d < 7; ++d) s += ((d == r) ? 1.0 : 0.0) * d1[d];
      v[r] = s;
      double t = 0.0;
      for (int d = 0; d < 7; ++d) t += ((d >= 3) ? dl[r * 4 + d - 3] : 0.0) * d1[d];
      v[3 + r] = t;
    }
  } else {
    const double* J = a.jac + (static_cast<size_t>(b) * N + idx) * 6 * D;
    for (int r = 0; r < 6; ++r) {
      double s = 0.0;
      for (int d = 0; d < D; ++d) s += J[r * D + d] * d1[d];
      v[r] = s;
    }
  }
  const double vt2 = max_trans_vel * max_trans_vel;
  const double vr2 = max_rot_vel * max_rot_vel;
  c[(2 * D) * 4 + 0] = 0.0;
  c[(2 * D) * 4 + 1] = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  c[(2 * D) * 4 + 2] = -vt2;
  c[(2 * D) * 4 + 3] = vt2;
  c[(2 * D + 1) * 4 + 0] = 0.0;
  c[(2 * D + 1) * 4 + 1] = v[3] * v[3] + v[4] * v[4] + v[5] * v[5];
  c[(2 * D + 1) * 4 + 2] = -vr2;
  c[(2 * D + 1) * 4 + 3] = vr2;
  if (idx == 0) {
    // SetupProblem: s in [parameter_start_, parameter_end_] (:546-547)
    const double path_start = a.path_start[b];
    a.s_start[b] = path_start;
    a.s_end[b] = path_start + delta * (N - 1);
  }
}

// k_cart_constraints + the per-sample limit curve in one pass for kD <= 7 dofs
// (2*kD + 2 <= 16 rows): the rows are still written for the sweeps, which read
// them per sample, but the LP and FindSddMin/Max run on the registers that built
// them (JointLimit<kD, 2>, bit-identical to lp_max_sd2 / find_sdd_min_max on the
// stored rows), so k_generic_limit's second pass over `cons` disappears.
template <bool kFake7, int kD>
__global__ void __launch_bounds__(kCartThreads)
k_cart_constraints_limit(Dims dm, DeviceArrays a, double safety, double max_trans_vel,
                         double max_rot_vel) {
  __shared__ double s_aS[kD], s_vS2[kD];
  const int b = path_of_block();
  if (b >= dm.B) return;  // uniform per CTA (grid folded over y and z)
  if (a.active != nullptr && a.active[b] == 0) return;  // uniform per CTA
  if (threadIdx.x < kD) {
    const double am = a.amax[static_cast<size_t>(b) * kD + threadIdx.x];
    const double vm = a.vmax[static_cast<size_t>(b) * kD + threadIdx.x];
    s_aS[threadIdx.x] = am * safety;
    const double vs = vm * safety;
    s_vS2[threadIdx.x] = vs * vs;
  }
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int N = dm.N;
  constexpr int C = 2 * kD + 2;
  if (idx >= N) return;
  const size_t row = (static_cast<size_t>(b) * N + idx) * kD;
  const double delta = a.delta[b];
  const double inv = 1.0 / delta;
  const double vt2 = max_trans_vel * max_trans_vel;
  const double vr2 = max_rot_vel * max_rot_vel;
  const double hx[2] = {vt2, vr2};
  JointLimit<kD, 2> J;
  J.aS = s_aS;
  J.vS2 = s_vS2;
  J.hX = hx;
  double p0[kD];
  {
    const double* __restrict__ q = a.q + row;
    const bool has1 = idx < N - 1, has2 = idx + 1 < N - 1;
#pragma unroll
    for (int d = 0; d < kD; ++d) {
      p0[d] = q[d];
      double f1 = 0.0, f2 = 0.0;
      if (has1) {
        const double p1 = q[kD + d];
        f1 = inv * (p1 - p0[d]);
        if (idx >= 1) {
          // first derivative of the next sample (0 at the last one)
          double g1 = 0.0;
          if (has2) g1 = inv * (q[2 * kD + d] - p1);
          f2 = inv * (g1 - f1);
        }
      }
      J.A[d] = f1;
      J.B[d] = f2;
      a.dq[row + d] = f1;
      a.ddq[row + d] = f2;
    }
  }
  double v[6];
  if (kFake7) {
    static_assert(!kFake7 || kD == 7, "the fake robot has 7 joints");
    double dl[12];
    log_so3_derivative(exp_so3(p0 + 3), dl);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      // full row sums like the dense product jacobian * q' (zeros included)
      double s = 0.0;
#pragma unroll
      for (int d = 0; d < kD; ++d) s += ((d == r) ? 1.0 : 0.0) * J.A[d];
      v[r] = s;
      double t = 0.0;
#pragma unroll
      for (int d = 0; d < kD; ++d) t += ((d >= 3) ? dl[r * 4 + (d >= 3 ? d - 3 : 0)] : 0.0) * J.A[d];
      v[3 + r] = t;
    }
  } else {
    const double* __restrict__ Jm = a.jac + (static_cast<size_t>(b) * N + idx) * 6 * kD;
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      double s = 0.0;
#pragma unroll
      for (int d = 0; d < kD; ++d) s += Jm[r * kD + d] * J.A[d];
      v[r] = s;
    }
  }
  J.Bv[kD] = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  J.Bv[kD + 1] = v[3] * v[3] + v[4] * v[4] + v[5] * v[5];
  J.prepare();
  // rows for the sweeps: one 32-byte row = two 16-byte stores
  double2* c = reinterpret_cast<double2*>(a.cons + (static_cast<size_t>(b) * N + idx) * C * 4);
#pragma unroll
  for (int d = 0; d < kD; ++d) {
    c[d * 2 + 0] = make_double2(J.A[d], J.B[d]);
    c[d * 2 + 1] = make_double2(-s_aS[d], s_aS[d]);
    c[(kD + d) * 2 + 0] = make_double2(0.0, J.Bv[d]);
    c[(kD + d) * 2 + 1] = make_double2(0.0, s_vS2[d]);
  }
  c[(2 * kD) * 2 + 0] = make_double2(0.0, J.Bv[kD]);
  c[(2 * kD) * 2 + 1] = make_double2(-vt2, vt2);
  c[(2 * kD + 1) * 2 + 0] = make_double2(0.0, J.Bv[kD + 1]);
  c[(2 * kD + 1) * 2 + 1] = make_double2(-vr2, vr2);
  // CalculateBoundary's per-sample part (time_optimal_path_timing.cc:1366-1377)
  double sd2max, sddmax, sd2zero, smax, smin;
  J.lp(sd2max, sddmax, sd2zero);
  J.sdd_min_max(sd2max, smax, smin);
  const size_t o = static_cast<size_t>(b) * N + idx;
  a.b_raw_sd2_max[o] = sd2max;
  a.b_sd2_zero[o] = sd2zero;
  a.b_raw_sdd_max[o] = smax;
  a.b_raw_sdd_min[o] = smin;
  a.b_at0[o] = fabs(sd2max - sd2zero) < kTiny;
  if (idx == 0) {
    // SetupProblem: s in [parameter_start_, parameter_end_] (:546-547)
    const double path_start = a.path_start[b];
    a.s_start[b] = path_start;
    a.s_end[b] = path_start + delta * (N - 1);
  }
}

}  // namespace

int launch_cart_fit(const Dims& d, const DeviceArrays& a, double rotation_rounding,
                    double translation_rounding, bool auto_delta, cudaStream_t st) {
  const int threads = 64;
  k_cart_fit<<<(d.B + threads - 1) / threads, threads, 0, st>>>(d, a, rotation_rounding,
                                                                translation_rounding, auto_delta);
  return 1;
}

int launch_cart_switch(const Dims& d, const DeviceArrays& a, const SwitchArgs& s,
                       double rotation_rounding, double translation_rounding, cudaStream_t st) {
  k_cart_switch<<<(d.B + 63) / 64, 64, 0, st>>>(d, a, s, rotation_rounding, translation_rounding);
  return 1;
}

int launch_restride_rows(int B, int width, int old_rows, int new_rows, const double* in, double* out,
                         cudaStream_t st) {
  k_restride_rows<<<B, 128, 0, st>>>(width, old_rows, new_rows, in, out);
  return 1;
}

int launch_cart_sample_targets(const Dims& d, const DeviceArrays& a, cudaStream_t st) {
  const size_t smem = sizeof(double) * (static_cast<size_t>(d.max_points) + 3 +
                                        static_cast<size_t>(d.max_points) * (d.D + 7));
  const dim3 grid = path_grid((d.N + 1 + kCartThreads - 1) / kCartThreads, d.B);
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(k_cart_sample_targets, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(smem));
  k_cart_sample_targets<<<grid, kCartThreads, smem, st>>>(d, a);
  return 1;
}

int launch_cart_ik_builtin(const Dims& d, const DeviceArrays& a, int model, cudaStream_t st) {
  if (model != TPB_KIN_FAKE7 || d.D != 7) return -1;
  const dim3 grid = path_grid((d.N + kCartThreads - 1) / kCartThreads, d.B);
  k_cart_ik_fake7<<<grid, kCartThreads, 0, st>>>(d, a);
  return 1;
}

// Rows + per-sample limit curve: one fused launch for 6- and 7-dof arms, else
// k_cart_constraints followed by k_generic_limit.  TPB_VARIANT_CART_UNFUSED
// (tpb_set_kernel_variant) forces the two-kernel path (the parity tests compare the two).
int launch_cart_constraints(const Dims& d, const DeviceArrays& a, double safety,
                            double max_trans_vel, double max_rot_vel, int model, bool have_jac,
                            cudaStream_t st) {
  const dim3 grid = path_grid((d.N + kCartThreads - 1) / kCartThreads, d.B);
  if (!have_jac && (model != TPB_KIN_FAKE7 || d.D != 7)) return -1;
  const bool fused = ((d.variants >> TPB_VARIANT_CART_UNFUSED) & 1u) == 0 && d.C == 2 * d.D + 2;
  if (fused && d.D == 7) {
    if (have_jac)
      k_cart_constraints_limit<false, 7><<<grid, kCartThreads, 0, st>>>(d, a, safety, max_trans_vel, max_rot_vel);
    else
      k_cart_constraints_limit<true, 7><<<grid, kCartThreads, 0, st>>>(d, a, safety, max_trans_vel, max_rot_vel);
    return 1;
  }
  if (fused && d.D == 6 && have_jac) {
    k_cart_constraints_limit<false, 6><<<grid, kCartThreads, 0, st>>>(d, a, safety, max_trans_vel, max_rot_vel);
    return 1;
  }
  if (have_jac) {
    k_cart_constraints<false><<<grid, kCartThreads, 0, st>>>(d, a, safety, max_trans_vel, max_rot_vel);
  } else {
    k_cart_constraints<true><<<grid, kCartThreads, 0, st>>>(d, a, safety, max_trans_vel, max_rot_vel);
  }
  return 1 + launch_generic_limit(d, a, st);
}

}  // namespace tpb
