// Online re-planning geometry of joint paths, one thread per path
// (reference: TimeableJointSplinePath::SwitchToWaypointPath,
// timeable_path_joint_spline.cc:209-250):
//   TruncateSplineAt          splines/bspline.h:402-427 (degree + 1 knot insertions at
//                             keep_path_until, then the second curve section is dropped)
//   knot insertion            splines/bspline.h:244-278 (NURBS Alg. 5.1, the reference's
//                             allocation-free ...Ref variant; its TruncateSplineAt goes
//                             through a QR solve of the same equations)
//   EvalCurve                 splines/bspline.h:514-536
//   ProjectPointOnPath        path_tools.h:59-104 (distance to the SEGMENT, line parameter
//                             clamped to [0, 1]; eigenmath::DistanceFromLineSegment is not
//                             vendored with the reference)
//   PolyLineToBspline3Waypoints  splines/spline_utils.cc:47-102 (tpb_fit.cuh)
//   ExtendWithControlPoints   splines/bspline.h:429-511
// The spline of path b lives in ctrl[b][max_points][D] / knots[b][max_points + 3] with
// the reference's per-path knot capacity max(2 * num_knots_at_fit, 100)
// (timeable_path_joint_spline.cc:262-268) in knot_cap[b].  A few hundred bytes per path,
// a handful of dependent steps: the kernel is launch-latency sized, not a hot path.
#include <cfloat>
#include <cmath>
#include <cstdint>

#include "tpb_engine.h"
#include "tpb_fit.cuh"

namespace tpb {
namespace {

enum : int32_t { kOk = 0, kInvalidArgument = 3, kFailedPrecondition = 9, kOutOfRange = 11,
                 kUnimplemented = 12, kInternal = 13 };
constexpr int kDegree = 2;

// KnotSpan (bspline_base.cc:218-246) on knots this thread is modifying
__device__ int span_of(const double* knots, int num_knots, double u) {
  if (u == knots[num_knots - 1]) return num_knots - kDegree - 2;
  int lo = kDegree, hi = num_knots - kDegree;
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

// InsertKnotAndUpdateControlPointsRef (bspline.h:244-278)
__device__ void insert_knot_ref(double* knots, double* pts, int D, int& num_knots, int& num_points,
                                double knot) {
  double scratch[kDegree][kMaxDofs];
  const int span = span_of(knots, num_knots, knot);
  for (int i = num_points - 1; i >= span; --i) {  // copy_backward
    for (int d = 0; d < D; ++d) pts[(i + 1) * D + d] = pts[i * D + d];
  }
  for (int i = 0; i < kDegree; ++i) {
    const int k = span + i - kDegree + 1;
    const double alpha = (knot - knots[k]) / (knots[k + kDegree] - knots[k]);
    for (int d = 0; d < D; ++d) {
      scratch[i][d] = alpha * pts[k * D + d] + (1.0 - alpha) * pts[(k - 1) * D + d];
    }
  }
  for (int i = 0; i < kDegree; ++i) {
    for (int d = 0; d < D; ++d) pts[(span - kDegree + 1 + i) * D + d] = scratch[i][d];
  }
  // InsertKnotIntoKnotVector (bspline_base.cc:197-214)
  for (int i = num_knots - 1; i >= span + 1; --i) knots[i + 1] = knots[i];
  knots[span + 1] = knot;
  num_knots += 1;
  num_points += 1;
}

__global__ void k_joint_switch(Dims dm, DeviceArrays a, SwitchArgs s, double rounding) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B) return;
  const int D = dm.D;
  const int W = s.num_waypoints ? s.num_waypoints[b] : s.max_waypoints;
  if (W <= 0) {  // this path keeps its spline
    s.status[b] = kOk;
    s.modified[b] = 0;
    return;
  }
  s.modified[b] = 1;  // kModifiedPath, also when the switch fails (.cc:212)
  double* knots = a.knots + static_cast<size_t>(b) * (dm.max_points + 3);
  double* pts = a.ctrl + static_cast<size_t>(b) * dm.max_points * D;
  const double* wps = s.waypoints + static_cast<size_t>(b) * s.max_waypoints * D;
  double* list = a.waypoints + static_cast<size_t>(b) * dm.Wmax * D;  // new waypoint list
  const int cap = a.knot_cap[b];
  int num_points = a.num_points[b];
  int num_knots = (num_points > 0) ? num_points + kDegree + 1 : 0;
  const double keep = s.keep_path_until[b];
  int32_t st = kOk;

  // ---- TruncateSplineAt (bspline.h:402-427)
  if (num_knots > 0 && !(keep >= knots[num_knots - 1])) {
    if (keep <= knots[0]) {
      num_knots = 0;
      num_points = 0;
    } else {
      // CanInsertKnot (bspline_base.cc:166-195); the value is inside the knot range here
      if (num_knots + kDegree + 1 > cap) {
        st = kFailedPrecondition;
      } else {
        for (int i = 0; i < kDegree + 1; ++i) insert_knot_ref(knots, pts, D, num_knots, num_points, keep);
        const int span = span_of(knots, num_knots, keep);
        num_knots = span + 1;
        num_points = num_knots - kDegree - 1;
      }
    }
  }
  // ---- EvalCurve(keep_path_until) (bspline.h:514-536)
  double switch_position[kMaxDofs];
  if (st == kOk) {
    if (num_knots == 0 || keep < knots[0] || keep > knots[num_knots - 1]) {
      st = kOutOfRange;
    } else {
      const int span = span_of(knots, num_knots, keep);
      double basis[kDegree + 1];
      basis_funs<kDegree>(knots, span, keep, basis);
      for (int d = 0; d < D; ++d) switch_position[d] = 0.0;
      for (int i = 0; i <= kDegree; ++i) {
        for (int d = 0; d < D; ++d) switch_position[d] += basis[i] * pts[(span - kDegree + i) * D + d];
      }
    }
  }
  if (st == kOk) {
    // ---- ProjectPointOnPath (path_tools.h:59-104)
    int index = 0;
    double best = DBL_MAX, line_parameter = 0.0;
    double projected[kMaxDofs];
    if (W == 1) {
      for (int d = 0; d < D; ++d) projected[d] = wps[d];
    } else {
      for (int i = 0; i < W - 1; ++i) {
        const double* p0 = wps + i * D;
        const double* p1 = p0 + D;
        double ab2 = 0.0, dot = 0.0;
        for (int d = 0; d < D; ++d) {
          ab2 += (p1[d] - p0[d]) * (p1[d] - p0[d]);
          dot += (switch_position[d] - p0[d]) * (p1[d] - p0[d]);
        }
        double t = (ab2 > 0.0) ? dot / ab2 : 0.0;
        t = (t < 0.0) ? 0.0 : ((t > 1.0) ? 1.0 : t);
        double sum = 0.0;
        for (int d = 0; d < D; ++d) {
          const double e = p0[d] + t * (p1[d] - p0[d]) - switch_position[d];
          sum += e * e;
        }
        const double dist = sqrt(sum);
        if (dist < best) {
          best = dist;
          line_parameter = t;
          index = i;
        }
      }
      const double* p0 = wps + index * D;
      for (int d = 0; d < D; ++d) projected[d] = p0[d] + line_parameter * (p0[D + d] - p0[d]);
    }
    // ---- the new waypoint list (.cc:226-243)
    int n = 0;
    double linf = 0.0;
    for (int d = 0; d < D; ++d) linf = fmax(linf, fabs(switch_position[d] - projected[d]));
    if (linf > 1e-3) {
      for (int d = 0; d < D; ++d) list[d] = projected[d];
      n = 1;
    }
    const int first = (line_parameter >= 0) ? index + 1 : index;
    for (int i = first; i < W; ++i) {
      for (int d = 0; d < D; ++d) list[n * D + d] = wps[i * D + d];
      ++n;
    }
    // ---- PolyLineToBspline3Waypoints + ExtendWithControlPoints (bspline.h:429-511)
    const int nc = (n == 1) ? 4 : 3 * n - 2;
    const int new_num_points = num_points + nc;
    const int added_knots = (nc + 1 + kDegree + 1) - 2 * kDegree;
    const int new_num_knots = num_knots + added_knots;
    if (n == 0) {
      st = kInvalidArgument;  // the reference would index an empty vector
    } else if (num_knots < 2 * kDegree + 2 || new_num_points > cap - kDegree - 1 ||
               new_num_knots > cap) {
      st = kFailedPrecondition;
    } else {
      // the new control points go straight behind the kept ones
      polyline_to_bspline3(list, n, D, rounding, pts + num_points * D);
      const double u_join = knots[num_knots - 1];
      const double old_knot_range = knots[num_knots - 1] - knots[0];
      const int old_inner = num_knots - 2 * kDegree - 1;
      const int new_inner = new_num_knots - 2 * kDegree - 1;
      const double new_knot_range = (old_knot_range * new_inner) / old_inner;
      const int start = num_knots - kDegree - 1;
      const int size = (new_num_knots - kDegree) - start;
      {  // Eigen's setLinSpaced(size, low, high) for |high| >= |low|
        const double low = old_knot_range, high = new_knot_range;
        const int size1 = (size == 1) ? 1 : size - 1;
        const double step = (size == 1) ? 0.0 : (high - low) / static_cast<double>(size - 1);
        for (int i = 0; i < size; ++i) {
          knots[start + i] = (i == size1) ? high : low + static_cast<double>(i) * step;
        }
      }
      for (int i = 0; i < kDegree + 1; ++i) {
        knots[new_num_knots - kDegree - 1 + i] = knots[0] + new_knot_range;
      }
      // the old end point is moved so that the curve keeps its value at the join
      const int join_span = span_of(knots, new_num_knots, u_join);
      const int mod = num_points - 1;
      double basis[kDegree + 1];
      basis_funs<kDegree>(knots, join_span, u_join, basis);
      if (!(basis[1] > 0)) {
        st = kInternal;
      } else {
        for (int d = 0; d < D; ++d) {
          pts[mod * D + d] = 1.0 / (basis[1]) * (pts[mod * D + d] - basis[0] * pts[(mod - 1) * D + d]);
        }
      }
      num_knots = new_num_knots;
      num_points = new_num_points;
    }
  }
  a.num_points[b] = num_points;
  if (num_knots > 0) a.knot_end[b] = knots[num_knots - 1];
  s.status[b] = st;
}

// ctrl / knots re-strided into arrays of a larger per-path capacity
__global__ void k_restride_spline(int B, int D, int old_points, int new_points,
                                  const double* __restrict__ ctrl_in,
                                  const double* __restrict__ knots_in, double* __restrict__ ctrl_out,
                                  double* __restrict__ knots_out) {
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < old_points * D; i += blockDim.x) {
    ctrl_out[static_cast<size_t>(b) * new_points * D + i] =
        ctrl_in[static_cast<size_t>(b) * old_points * D + i];
  }
  for (int i = threadIdx.x; i < old_points + 3; i += blockDim.x) {
    knots_out[static_cast<size_t>(b) * (new_points + 3) + i] =
        knots_in[static_cast<size_t>(b) * (old_points + 3) + i];
  }
  (void)B;
}

}  // namespace

int launch_joint_switch(const Dims& d, const DeviceArrays& a, const SwitchArgs& s, double rounding,
                        cudaStream_t st) {
  k_joint_switch<<<(d.B + 63) / 64, 64, 0, st>>>(d, a, s, rounding);
  return 1;
}

int launch_restride_spline(int B, int D, int old_points, int new_points, const double* ctrl_in,
                           const double* knots_in, double* ctrl_out, double* knots_out,
                           cudaStream_t st) {
  k_restride_spline<<<B, 128, 0, st>>>(B, D, old_points, new_points, ctrl_in, knots_in, ctrl_out,
                                       knots_out);
  return 1;
}

}  // namespace tpb
