// Host-side mirror of trajectory_planning/path_tools.h (:26-117) and path_tools.cc
// (:25-74): ProjectPointOnPath and ComputeStoppingPoint.  Both are a few dozen flops on
// one point, so they run on the host; the batched device form of ProjectPointOnPath is
// part of tpb_switch_to_waypoint_path (csrc/tpb_switch.cu), which
// TimeableJointSplinePath::SwitchToWaypointPath calls.
//
// eigenmath::DistanceFromLineSegment and eigenmath::ScaleDownToLimits are not vendored
// with the reference; they are restated here as the distance to the SEGMENT (line
// parameter clamped to [0, 1]) and as the uniform scaling that brings every component
// within its limit.
#ifndef TRAJECTORY_PLANNING_PATH_TOOLS_H_
#define TRAJECTORY_PLANNING_PATH_TOOLS_H_

#include <algorithm>
#include <cmath>
#include <limits>
#include <string>

#include "tp/base.h"

namespace trajectory_planning {

// path_tools.h:26-36
struct ProjectedPointResult {
  int waypoint_index = 0;
  double distance_to_path = std::numeric_limits<double>::max();
  double line_parameter = 0;
  tp::VectorXd projected_point;
};

namespace path_tools_internal {
inline void DistanceFromLineSegment(const tp::VectorXd& a, const tp::VectorXd& b, const tp::VectorXd& p,
                                    double* distance, double* line_parameter) {
  double ab2 = 0.0, dot = 0.0;
  for (size_t d = 0; d < p.size(); ++d) {
    ab2 += (b[d] - a[d]) * (b[d] - a[d]);
    dot += (p[d] - a[d]) * (b[d] - a[d]);
  }
  double t = (ab2 > 0.0) ? dot / ab2 : 0.0;
  t = (t < 0.0) ? 0.0 : ((t > 1.0) ? 1.0 : t);
  double s = 0.0;
  for (size_t d = 0; d < p.size(); ++d) {
    const double e = a[d] + t * (b[d] - a[d]) - p[d];
    s += e * e;
  }
  *distance = std::sqrt(s);
  *line_parameter = t;
}
}  // namespace path_tools_internal

// ProjectPointOnPath (path_tools.h:59-104): `point` projected on the straight line
// segments between `waypoints`, the closest one wins.
inline tp::StatusOr<ProjectedPointResult> ProjectPointOnPath(tp::Span<const tp::VectorXd> waypoints,
                                                             const tp::VectorXd& point) {
  if (waypoints.empty()) return tp::InvalidArgumentError("No waypoints.");
  ProjectedPointResult result;
  if (waypoints.size() == 1) {
    double s = 0.0;
    const tp::VectorXd& w0 = *waypoints.begin();
    for (size_t d = 0; d < std::min(w0.size(), point.size()); ++d) s += (w0[d] - point[d]) * (w0[d] - point[d]);
    result.waypoint_index = 0;
    result.distance_to_path = std::sqrt(s);
    result.line_parameter = 0.0;
    result.projected_point = w0;
    return result;
  }
  for (const auto& wp : waypoints) {
    if (wp.size() != point.size()) {
      return tp::InvalidArgumentError("Invalid number of joints: point.size()= " + std::to_string(point.size()) +
                                      ", waypoint.size: " + std::to_string(wp.size()));
    }
  }
  const tp::VectorXd* wps = waypoints.data();
  for (size_t i = 0; i + 1 < waypoints.size(); ++i) {
    double distance = 0.0, line_parameter = 0.0;
    path_tools_internal::DistanceFromLineSegment(wps[i], wps[i + 1], point, &distance, &line_parameter);
    if (distance < result.distance_to_path) {
      result.distance_to_path = distance;
      result.line_parameter = line_parameter;
      result.waypoint_index = static_cast<int>(i);
    }
  }
  const tp::VectorXd& a = wps[result.waypoint_index];
  const tp::VectorXd& b = wps[result.waypoint_index + 1];
  result.projected_point = tp::VectorXd(point.size());
  for (size_t d = 0; d < point.size(); ++d) {
    result.projected_point[d] = a[d] + result.line_parameter * (b[d] - a[d]);  // InterpolateLinear
  }
  return result;
}

// PolyLineToBspline3WaypointsCornerOffset (splines/spline_utils.cc:25-45)
inline tp::VectorXd PolyLineToBspline3WaypointsCornerOffset(const tp::VectorXd& delta, double radius) {
  constexpr double kMinNorm = 1e-6;
  constexpr double kMinWaypointSpacingFactor = 4.0;  // spline_utils.h:46
  double s = 0.0;
  for (size_t d = 0; d < delta.size(); ++d) s += delta[d] * delta[d];
  const double norm = std::sqrt(s);
  tp::VectorXd offset(delta.size());
  for (size_t d = 0; d < delta.size(); ++d) offset[d] = (norm > kMinNorm) ? delta[d] / norm : 0.0;
  if (norm > kMinWaypointSpacingFactor * radius) {
    for (size_t d = 0; d < delta.size(); ++d) offset[d] = offset[d] * radius;
  } else {
    for (size_t d = 0; d < delta.size(); ++d) offset[d] = (offset[d] * (1.0 / kMinWaypointSpacingFactor)) * norm;
  }
  return offset;
}

// ComputeStoppingPoint (path_tools.cc:25-74): where the robot could come to a stop with
// a straight-line motion from `position` with `velocity` within `acceleration_limit`,
// pushed further along by the corner rounding.
inline tp::StatusOr<tp::VectorXd> ComputeStoppingPoint(tp::Span<const double> position,
                                                       tp::Span<const double> velocity,
                                                       tp::Span<const double> acceleration_limit,
                                                       double corner_rounding) {
  if (position.size() != velocity.size()) return tp::InvalidArgumentError("position.size != velocity.size.");
  if (position.size() != acceleration_limit.size()) {
    return tp::InvalidArgumentError("position.size != acceleration.size.");
  }
  const size_t n = position.size();
  const double* pos = position.data();
  const double* vel = velocity.data();
  const double* acc = acceleration_limit.data();
  double acc_min = std::numeric_limits<double>::infinity(), acc_max = 0.0, vel_max = 0.0, vel_sq = 0.0;
  for (size_t i = 0; i < n; ++i) {
    acc_min = std::min(acc_min, acc[i]);
    acc_max = std::max(acc_max, acc[i]);
    vel_max = std::max(vel_max, std::abs(vel[i]));
    vel_sq += vel[i] * vel[i];
  }
  if (acc_min <= 0.0) return tp::InvalidArgumentError("Acceleration limits must be strictly positive.");
  tp::VectorXd out(pos, n);
  if (vel_max < std::numeric_limits<double>::epsilon() * 100) return out;
  // the largest deceleration collinear to the velocity within the limits
  const double vel_norm = std::sqrt(vel_sq);
  tp::VectorXd decel(n);
  double scale = 1.0;
  for (size_t i = 0; i < n; ++i) {
    decel[i] = -(vel[i] / vel_norm) * acc_max;
    if (std::abs(decel[i]) > acc[i]) scale = std::min(scale, acc[i] / std::abs(decel[i]));
  }
  double vel_dot_decel = 0.0;
  for (size_t i = 0; i < n; ++i) {
    decel[i] *= scale;  // ScaleDownToLimits
    vel_dot_decel += vel[i] * decel[i];
  }
  const double stop_time = -vel_sq / vel_dot_decel;
  tp::VectorXd stopping_delta(n);
  for (size_t i = 0; i < n; ++i) stopping_delta[i] = stop_time * (vel[i] + 0.5 * stop_time * decel[i]);
  const tp::VectorXd corner = PolyLineToBspline3WaypointsCornerOffset(stopping_delta, corner_rounding);
  for (size_t i = 0; i < n; ++i) out[i] = pos[i] + stopping_delta[i] + corner[i];
  return out;
}

}  // namespace trajectory_planning
#endif  // TRAJECTORY_PLANNING_PATH_TOOLS_H_
