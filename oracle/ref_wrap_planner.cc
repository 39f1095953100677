// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// extern "C" access to the reference's own joint-path planner stack, compiled
// VERBATIM by oracle/Makefile from /root/reference/trajectory_planning against the
// stand-in headers in oracle/shim/:
//   splines/bspline_base.cc, splines/bspline.h, splines/spline_utils.cc  (stage A/B)
//   timeable_path_joint_spline.cc                                        (stage B)
//   time_optimal_path_timing.cc                                          (stage C)
//   path_timing_trajectory.cc                                            (stage D, planner)
// Part of oracle/_ref/libtp_ref.so.  Pins oracle/tp_oracle.c (stages A, B, D),
// tp_oracle_windows.c (moving-horizon Plan), tp_oracle_switch.c (SwitchToWaypointPath)
// and tp_oracle_stop.c (GetPathStopParameter) bit for bit, and is the timed CPU
// baseline ("kind": "reference") of bench.py: tpref_plan_joint_batch runs nothing
// but reference object code per path.
//
// Functions return absl::StatusCode numbers (0 = ok) unless stated otherwise.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <limits>
#include <memory>
#include <string>
#include <thread>
#include <tuple>
#include <utility>
#include <vector>

#include "absl/status/status.h"
#include "absl/status/statusor.h"
#include "absl/time/time.h"
#include "absl/types/span.h"
#include "eigenmath/types.h"

// The planner's and the path's state is private in the reference classes; open it
// for inspection by the tests (this translation unit only; layout is unchanged).
#define private public
#define protected public
#include "trajectory_planning/path_timing_trajectory.h"
#include "trajectory_planning/timeable_path_cartesian_spline.h"
#include "trajectory_planning/timeable_path_joint_spline.h"
#undef private
#undef protected

using eigenmath::VectorXd;
using trajectory_planning::JointPathOptions;
using trajectory_planning::PathTimingTrajectory;
using trajectory_planning::PathTimingTrajectoryOptions;
using trajectory_planning::CartesianPathOptions;
using trajectory_planning::TimeableCartesianSplinePath;
using trajectory_planning::TimeableJointSplinePath;
using trajectory_planning::TimeablePath;

// The user callbacks of a Cartesian path (CartesianPathOptions::PathIKFunc /
// JacobianFunc, timeable_path_cartesian_spline.h:33-41) as plain C functions over
// packed arrays: poses are (tx, ty, tz, qx, qy, qz, qw), the Jacobian is 6 x D row-major.
extern "C" {
typedef void (*tpref_ik_fn)(const double* pose_targets, const double* joint_targets, int M,
                            double* ik);
typedef void (*tpref_jacobian_fn)(const double* joints, double* jacobian);
}

namespace {

std::vector<VectorXd> Rows(const double* p, int n, int D) {
  std::vector<VectorXd> v(n, VectorXd(D));
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < D; ++j) v[i][j] = p[static_cast<size_t>(i) * D + j];
  return v;
}

int Code(const absl::Status& s) { return static_cast<int>(s.code()); }

// knots.back() of the spline the reference fits to these waypoints: a path object
// with the minimum number of samples, asked through its own SetWaypoints.
absl::Status KnotEnd(const std::vector<VectorXd>& wp, int D, double rounding, double* kend) {
  JointPathOptions o;
  o.set_num_dofs(D).set_num_path_samples(3).set_rounding(rounding);
  TimeableJointSplinePath tmp(o);
  absl::Status st = tmp.SetWaypoints(wp);
  if (!st.ok()) return st;
  *kend = tmp.knots_.back();
  return absl::OkStatus();
}

PathTimingTrajectoryOptions PlannerOptions(int N, int D, int64_t dt_ns, int skip_method,
                                           int max_planning_iterations,
                                           double max_initial_velocity_error) {
  PathTimingTrajectoryOptions o;
  o.SetNumDofs(D).SetTimeStep(absl::Nanoseconds(dt_ns));
  o.SetNumPathSamples(N);
  if (max_planning_iterations > 0) o.SetMaxPlanningLoops(max_planning_iterations);
  if (max_initial_velocity_error > 0) o.SetMaxInitialVelocityError(max_initial_velocity_error);
  o.SetTimeSamplingMethod(
      skip_method ? PathTimingTrajectoryOptions::TimeSamplingMethod::kSkipSamplesCloserThanTimeStep
                  : PathTimingTrajectoryOptions::TimeSamplingMethod::kUniformlyInTime);
  return o;
}

struct Planner {
  int N = 0, D = 0;
  double rounding = 0.2, safety = 0.8, delta = -1.0;
  std::vector<double> vmax, amax, init_vel;
  std::shared_ptr<TimeableJointSplinePath> path;      // joint path, or
  std::shared_ptr<TimeableCartesianSplinePath> cart;  // Cartesian path
  std::unique_ptr<PathTimingTrajectory> planner;
  // Cartesian path only
  double translation_rounding = 0.05, max_translational_velocity = 0.0,
         max_rotational_velocity = 0.0;
  tpref_ik_fn ik = nullptr;
  tpref_jacobian_fn jacobian = nullptr;
  TimeablePath* any() const {
    return path ? static_cast<TimeablePath*>(path.get()) : static_cast<TimeablePath*>(cart.get());
  }
  const std::vector<double>* knots() const {
    return path ? &path->knots_ : (cart ? &cart->knots_ : nullptr);
  }
};

}  // namespace

extern "C" {

// delta <= 0: the first tpref_planner_set_waypoints resolves it to
// knots.back() / (N - 1) (the whole path in one planning window).
void* tpref_planner_create(int N, int D, double rounding, double safety, double delta,
                           int64_t dt_ns, int skip_method, int max_planning_iterations,
                           double max_initial_velocity_error) {
  auto* h = new Planner;
  h->N = N;
  h->D = D;
  h->rounding = rounding;
  h->safety = safety;
  h->delta = delta;
  h->vmax.assign(D, 0.0);
  h->amax.assign(D, 0.0);
  h->init_vel.assign(D, 0.0);
  h->planner = std::make_unique<PathTimingTrajectory>(PlannerOptions(
      N, D, dt_ns, skip_method, max_planning_iterations, max_initial_velocity_error));
  return h;
}

void tpref_planner_destroy(void* hv) { delete static_cast<Planner*>(hv); }

int tpref_planner_set_limits(void* hv, const double* vmax, const double* amax) {
  auto* h = static_cast<Planner*>(hv);
  h->vmax.assign(vmax, vmax + h->D);
  h->amax.assign(amax, amax + h->D);
  if (h->any() == nullptr) return 0;
  absl::Status st = h->any()->SetMaxJointVelocity(h->vmax);
  if (!st.ok()) return Code(st);
  return Code(h->any()->SetMaxJointAcceleration(h->amax));
}

// TimeableJointSplinePath::SetWaypoints (+ SetPath on the first call).
int tpref_planner_set_waypoints(void* hv, int W, const double* waypoints) {
  auto* h = static_cast<Planner*>(hv);
  const std::vector<VectorXd> wp = Rows(waypoints, W, h->D);
  if (h->path == nullptr) {
    if (!(h->delta > 0)) {
      double kend = 0.0;
      absl::Status st = KnotEnd(wp, h->D, h->rounding, &kend);
      if (!st.ok()) return Code(st);
      h->delta = kend / (h->N - 1);
    }
    JointPathOptions o;
    o.set_num_dofs(h->D).set_num_path_samples(h->N).set_rounding(h->rounding);
    o.set_constraint_safety(h->safety).set_delta_parameter(h->delta);
    h->path = std::make_shared<TimeableJointSplinePath>(o);
    absl::Status st = h->path->SetMaxJointVelocity(h->vmax);
    if (st.ok()) st = h->path->SetMaxJointAcceleration(h->amax);
    if (st.ok()) st = h->planner->SetPath(h->path);
    if (!st.ok()) return Code(st);
  }
  absl::Status st = h->path->SetWaypoints(wp);
  if (!st.ok()) return Code(st);
  return Code(h->path->SetInitialVelocity(h->init_vel));
}

int tpref_planner_set_initial_velocity(void* hv, const double* velocity) {
  auto* h = static_cast<Planner*>(hv);
  h->init_vel.assign(velocity, velocity + h->D);
  if (h->any() == nullptr) return 0;
  return Code(h->any()->SetInitialVelocity(h->init_vel));
}

// PathTimingTrajectory::Plan(start, horizon)
int tpref_planner_plan(void* hv, int64_t start_ns, int64_t horizon_ns) {
  auto* h = static_cast<Planner*>(hv);
  return Code(h->planner->Plan(absl::FromUnixNanos(start_ns), absl::Nanoseconds(horizon_ns)));
}

// TimeableJointSplinePath::SwitchToWaypointPath
int tpref_planner_switch(void* hv, double keep_path_until, int W, const double* waypoints) {
  auto* h = static_cast<Planner*>(hv);
  if (h->path == nullptr) return static_cast<int>(absl::StatusCode::kFailedPrecondition);
  return Code(h->path->SwitchToWaypointPath(keep_path_until, Rows(waypoints, W, h->D)));
}

// PathTimingTrajectory::GetPathStopParameter
int tpref_planner_stop_parameter(void* hv, int64_t time_ns, double* parameter) {
  auto* h = static_cast<Planner*>(hv);
  auto r = h->planner->GetPathStopParameter(absl::FromUnixNanos(time_ns));
  if (!r.ok()) return Code(r.status());
  *parameter = *r;
  return 0;
}

// out[0..10]: start_time, end_time, final_decel_start (Unix ns), planned_to_end,
// target_reached, TimeablePath::State (0 kNoPath, 1 kNewPath, 2 kModifiedPath,
// 3 kPathWasSampled; -1 without a path), initial_plan, trajectory rows, path samples
// held in the *_at_path_samples_ arrays, GetLastExtremalIndex(), IsTrajectoryAtEnd()
void tpref_planner_state(void* hv, int64_t* out) {
  auto* h = static_cast<Planner*>(hv);
  const PathTimingTrajectory& p = *h->planner;
  out[0] = absl::ToUnixNanos(p.start_time_);
  out[1] = absl::ToUnixNanos(p.end_time_);
  out[2] = absl::ToUnixNanos(p.final_decel_start_);
  out[3] = p.planned_to_end_ ? 1 : 0;
  out[4] = p.target_reached_ ? 1 : 0;
  out[5] = h->any() ? static_cast<int64_t>(h->any()->GetState()) : -1;
  out[6] = p.initial_plan_ ? 1 : 0;
  out[7] = static_cast<int64_t>(p.time_.size());
  out[8] = static_cast<int64_t>(p.time_at_path_samples_.size());
  out[9] = p.profile_.GetLastExtremalIndex();
  out[10] = p.IsTrajectoryAtEnd() ? 1 : 0;
}

// out[0..3]: delta_parameter, knots.back(), path_horizon_, path_start_
void tpref_planner_scalars(void* hv, double* out) {
  auto* h = static_cast<Planner*>(hv);
  out[0] = h->delta;
  out[1] = (h->knots() && !h->knots()->empty()) ? h->knots()->back()
                                                : std::numeric_limits<double>::quiet_NaN();
  out[2] = h->planner->path_horizon_;
  out[3] = h->planner->path_start_;
}

// The trajectory (TrajectoryPlanner getters); every pointer may be NULL.
void tpref_planner_rows(void* hv, double* t, double* pos, double* vel, double* acc, double* s,
                        double* sd, double* sdd) {
  auto* h = static_cast<Planner*>(hv);
  const PathTimingTrajectory& p = *h->planner;
  const size_t M = p.GetTime().size(), D = static_cast<size_t>(h->D);
  for (size_t i = 0; i < M; ++i) {
    if (t) t[i] = p.GetTime()[i];
    if (s) s[i] = p.GetPathParameters()[i];
    if (sd) sd[i] = p.GetPathParameterDerivatives()[i];
    if (sdd) sdd[i] = p.GetSecondPathParameterDerivatives()[i];
    for (size_t d = 0; d < D; ++d) {
      if (pos) pos[i * D + d] = p.GetPositions()[i][d];
      if (vel) vel[i * D + d] = p.GetVelocities()[i][d];
      if (acc) acc[i * D + d] = p.GetAccelerations()[i][d];
    }
  }
}

// The spliced *_at_path_samples_ arrays (tpref_planner_state out[8] samples).
void tpref_planner_path_samples(void* hv, double* time, double* s, double* sd, double* sdd,
                                double* q, double* vel, double* acc) {
  auto* h = static_cast<Planner*>(hv);
  const PathTimingTrajectory& p = *h->planner;
  const size_t n = p.time_at_path_samples_.size(), D = static_cast<size_t>(h->D);
  for (size_t i = 0; i < n; ++i) {
    if (time) time[i] = p.time_at_path_samples_[i];
    if (s) s[i] = p.path_parameter_at_path_samples_[i];
    if (sd) sd[i] = p.path_velocity_at_path_samples_[i];
    if (sdd) sdd[i] = p.path_acceleration_at_path_samples_[i];
    for (size_t d = 0; d < D; ++d) {
      if (q) q[i * D + d] = p.position_at_path_samples_[i][d];
      if (vel) vel[i * D + d] = p.velocity_at_path_samples_[i][d];
      if (acc) acc[i * D + d] = p.acceleration_at_path_samples_[i][d];
    }
  }
}

// The last planning window: SamplePath's q, q', q'' [N][D] and the profile's
// time, s, sd, sdd [N]; every pointer may be NULL.
void tpref_planner_window(void* hv, double* q, double* dq, double* ddq, double* time, double* s,
                          double* sd, double* sdd) {
  auto* h = static_cast<Planner*>(hv);
  if (h->any() == nullptr) return;
  const size_t N = static_cast<size_t>(h->N), D = static_cast<size_t>(h->D);
  const auto& prof = h->planner->profile_;
  for (size_t i = 0; i < N; ++i) {
    for (size_t d = 0; d < D; ++d) {
      if (q) q[i * D + d] = h->any()->GetPathPositionAt(i)[d];
      if (dq) dq[i * D + d] = h->any()->GetFirstPathDerivativeAt(i)[d];
      if (ddq) ddq[i * D + d] = h->any()->GetSecondPathDerivativeAt(i)[d];
    }
  }
  if (prof.GetTimeSamples().size() == static_cast<std::ptrdiff_t>(N)) {
    for (size_t i = 0; i < N; ++i) {
      if (time) time[i] = prof.GetTimeSamples()[i];
      if (s) s[i] = prof.GetPathParameter()[i];
      if (sd) sd[i] = prof.GetPathVelocity()[i];
      if (sdd) sdd[i] = prof.GetPathAcceleration()[i];
    }
  }
}

// ConstraintSetup's rows of the last window, [N][C][4] = (A, B, lower, upper); C = 2D for
// a joint path, 2D + 2 for a Cartesian one
void tpref_planner_constraints(void* hv, double* cons) {
  auto* h = static_cast<Planner*>(hv);
  if (h->any() == nullptr) return;
  const auto& c = h->any()->GetConstraints();
  const size_t C = h->any()->NumConstraints();
  for (size_t i = 0; i < c.size(); ++i) {
    for (size_t r = 0; r < C; ++r) {
      double* row = cons + (i * C + r) * 4;
      row[0] = c[i].a_coefficient(r);
      row[1] = c[i].b_coefficient(r);
      row[2] = c[i].lower(r);
      row[3] = c[i].upper(r);
    }
  }
}

// The path's spline: counts[0] = knots, counts[1] = control points; knots / points
// are filled up to the given capacities (points [cap][D]).
void tpref_planner_spline(void* hv, int* counts, double* knots, int knot_cap, double* points,
                          int point_cap) {
  auto* h = static_cast<Planner*>(hv);
  counts[0] = counts[1] = 0;
  if (h->any() == nullptr) return;
  auto& spline = h->path ? h->path->spline_ : h->cart->joint_spline_;
  const auto kv = spline.GetKnotVector();
  const auto cp = spline.GetControlPoints();
  counts[0] = static_cast<int>(kv.size());
  counts[1] = static_cast<int>(cp.size());
  for (int i = 0; i < counts[0] && i < knot_cap; ++i) knots[i] = kv[i];
  for (int i = 0; i < counts[1] && i < point_cap; ++i)
    for (int d = 0; d < h->D; ++d) points[static_cast<size_t>(i) * h->D + d] = cp[i][d];
}

// BASELINE cfg3 on the host: dense evaluation with derivatives by the reference's own
// BSplineXd (splines/bspline.h:539-568 EvalCurveAndDerivatives over BSplineBase::KnotSpan
// / UpdateBasisAndDerivatives, bspline_base.cc:218-348), one parameter per call as the
// reference's API has it; splines are spread over `nthreads` host threads.
//   knots [S][n + degree + 1], ctrl [S][n][D], u [S][E], out [S][E][nder + 1][D],
//   status [S][E] (absl code per evaluation; may be NULL).  Returns an absl code.
int tpref_spline_eval_batch(int S, int degree, int n, int D, const double* knots,
                            const double* ctrl, int E, const double* u, int nder, int nthreads,
                            double* out, int32_t* status) {
  const int nk = n + degree + 1;
  std::atomic<int> next{0};
  std::atomic<int> first_error{0};
  auto worker = [&]() {
    trajectory_planning::BSplineXd spline;
    std::vector<VectorXd> values(nder + 1, VectorXd(D));
    for (;;) {
      const int s = next.fetch_add(1);
      if (s >= S) break;
      const std::vector<VectorXd> pts = Rows(ctrl + static_cast<size_t>(s) * n * D, n, D);
      const std::vector<double> kn(knots + static_cast<size_t>(s) * nk,
                                   knots + static_cast<size_t>(s + 1) * nk);
      absl::Status st = spline.Init(degree, pts, kn);
      if (!st.ok()) {
        int expected = 0;
        first_error.compare_exchange_strong(expected, Code(st));
        continue;
      }
      for (int e = 0; e < E; ++e) {
        const size_t flat = static_cast<size_t>(s) * E + e;
        st = spline.EvalCurveAndDerivatives(u[flat], absl::MakeSpan(values));
        if (status != nullptr) status[flat] = Code(st);
        if (!st.ok()) continue;
        double* o = out + flat * (nder + 1) * D;
        for (int k = 0; k <= nder; ++k)
          for (int d = 0; d < D; ++d) o[k * D + d] = values[k][d];
      }
    }
  };
  if (nthreads <= 1) {
    worker();
  } else {
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; ++t) pool.emplace_back(worker);
    for (auto& t : pool) t.join();
  }
  return first_error.load();
}

// BASELINE cfg2 / cfg4 on the host: B independent fresh joint paths, each planned
// from its waypoints to its resampled trajectory by the reference's own classes
// (SetWaypoints -> Plan: SamplePath, ConstraintSetup, OptimizePathParameter,
// ResampleTrajectory), the whole path in one window (delta_b = knots.back()/(N-1)),
// on `nthreads` host threads -- the reference is single-threaded, so one independent
// path per task; each thread owns its planner, a path object is built per path
// because its sampling distance is a constructor option.  No Python in the loop.
//   waypoints [B][W][D], vmax / amax [B][D]
//   status [B] absl code, duration [B] (NaN on failure), rows [B]
//   optional: r_offset [B+1] + r_pos [rows][D] -- path b's positions are written at
//   r_offset[b] when it produced exactly r_offset[b+1] - r_offset[b] rows; likewise
//   prof_time [B][N] (the profile's time stamps).
void tpref_plan_joint_batch(int B, int N, int D, int W, const double* waypoints,
                            const double* vmax, const double* amax, double rounding,
                            double safety, int64_t dt_ns, int skip_method, int nthreads,
                            int32_t* status, double* duration, int32_t* rows,
                            const int64_t* r_offset, double* r_pos, double* prof_time) {
  std::atomic<int> next{0};
  auto worker = [&]() {
    PathTimingTrajectory planner(PlannerOptions(N, D, dt_ns, skip_method, 0, 0.0));
    for (;;) {
      const int b = next.fetch_add(1);
      if (b >= B) break;
      status[b] = 0;
      duration[b] = std::numeric_limits<double>::quiet_NaN();
      rows[b] = 0;
      const std::vector<VectorXd> wp = Rows(waypoints + static_cast<size_t>(b) * W * D, W, D);
      double kend = 0.0;
      absl::Status st = KnotEnd(wp, D, rounding, &kend);
      std::shared_ptr<TimeableJointSplinePath> path;
      if (st.ok()) {
        JointPathOptions o;
        o.set_num_dofs(D).set_num_path_samples(N).set_rounding(rounding);
        o.set_constraint_safety(safety).set_delta_parameter(kend / (N - 1));
        path = std::make_shared<TimeableJointSplinePath>(o);
        st = path->SetMaxJointVelocity(
            absl::MakeConstSpan(vmax + static_cast<size_t>(b) * D, D));
      }
      if (st.ok())
        st = path->SetMaxJointAcceleration(
            absl::MakeConstSpan(amax + static_cast<size_t>(b) * D, D));
      if (st.ok()) {
        planner.Reset();
        st = planner.SetPath(path);
      }
      if (st.ok()) st = path->SetWaypoints(wp);
      if (st.ok()) st = planner.Plan(absl::FromUnixNanos(0), absl::Seconds(1));
      status[b] = Code(st);
      if (!st.ok()) continue;
      duration[b] = planner.profile_.GetTotalDuration();
      const size_t M = planner.GetTime().size();
      rows[b] = static_cast<int32_t>(M);
      if (prof_time != nullptr) {
        for (int i = 0; i < N; ++i)
          prof_time[static_cast<size_t>(b) * N + i] = planner.profile_.GetTimeSamples()[i];
      }
      if (r_offset != nullptr && r_pos != nullptr &&
          r_offset[b + 1] - r_offset[b] == static_cast<int64_t>(M)) {
        double* dst = r_pos + static_cast<size_t>(r_offset[b]) * D;
        for (size_t i = 0; i < M; ++i)
          for (int d = 0; d < D; ++d) dst[i * D + d] = planner.GetPositions()[i][d];
      }
    }
  };
  if (nthreads <= 1) {
    worker();
    return;
  }
  std::vector<std::thread> pool;
  for (int t = 0; t < nthreads; ++t) pool.emplace_back(worker);
  for (auto& t : pool) t.join();
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Cartesian path (BASELINE config 5): splines/bsplineq.cc and
// timeable_path_cartesian_spline.cc compiled VERBATIM; pins oracle/tp_oracle_cart.c
// and the Cartesian branch of tp_oracle_windows.c.
// ---------------------------------------------------------------------------
#include "eigenmath/pose3.h"
#include "trajectory_planning/splines/bsplineq.h"
#include "trajectory_planning/splines/spline_utils.h"

namespace {

using eigenmath::Pose3d;
using eigenmath::Quaterniond;
using eigenmath::Vector3d;

Quaterniond QuatOf(const double* q) { return Quaterniond(q[3], q[0], q[1], q[2]); }
void PutQuat(const Quaterniond& q, double* out) {
  out[0] = q.x();
  out[1] = q.y();
  out[2] = q.z();
  out[3] = q.w();
}
Pose3d PoseOf(const double* p) { return Pose3d(QuatOf(p + 3), Vector3d(p[0], p[1], p[2])); }
void PutPose(const Pose3d& p, double* out) {
  out[0] = p.translation()[0];
  out[1] = p.translation()[1];
  out[2] = p.translation()[2];
  PutQuat(p.quaternion(), out + 3);
}
std::vector<Pose3d> Poses(const double* p, int n) {
  std::vector<Pose3d> v;
  v.reserve(n);
  for (int i = 0; i < n; ++i) v.push_back(PoseOf(p + 7 * static_cast<size_t>(i)));
  return v;
}

CartesianPathOptions CartOptions(const Planner* h, int N, double delta) {
  CartesianPathOptions o;
  o.set_num_dofs(h->D).set_num_path_samples(N).set_rounding(h->rounding);
  o.set_constraint_safety(h->safety).set_delta_parameter(delta);
  o.set_translation_rounding(h->translation_rounding);
  const int D = h->D;
  const tpref_ik_fn ik = h->ik;
  const tpref_jacobian_fn jac = h->jacobian;
  o.set_path_ik_func([ik, D](const VectorXd&, const std::vector<Pose3d>& pose,
                             const std::vector<VectorXd>& joint,
                             std::vector<VectorXd>* solution) -> absl::Status {
    const size_t M = pose.size();
    std::vector<double> pt(7 * M), jt(D * M), out(D * M);
    for (size_t i = 0; i < M; ++i) {
      PutPose(pose[i], pt.data() + 7 * i);
      for (int d = 0; d < D; ++d) jt[i * D + d] = joint[i][d];
    }
    ik(pt.data(), jt.data(), static_cast<int>(M), out.data());
    solution->resize(M);
    for (size_t i = 0; i < M; ++i) {
      (*solution)[i].resize(D);
      for (int d = 0; d < D; ++d) (*solution)[i][d] = out[i * D + d];
    }
    return absl::OkStatus();
  });
  o.set_jacobian_func([jac, D](const VectorXd& joints, eigenmath::Matrix6Xd* J) -> absl::Status {
    std::vector<double> q(D), m(6 * static_cast<size_t>(D));
    for (int d = 0; d < D; ++d) q[d] = joints[d];
    jac(q.data(), m.data());
    for (int r = 0; r < 6; ++r)
      for (int d = 0; d < D; ++d) (*J)(r, d) = m[static_cast<size_t>(r) * D + d];
    return absl::OkStatus();
  });
  return o;
}

}  // namespace

extern "C" {

// bsplineq_details (splines/bsplineq.cc:112-160); quaternions are (x, y, z, w)
void tpref_quat_log(const double* q, double* out) {
  PutQuat(trajectory_planning::bsplineq_details::QuatLog(QuatOf(q)), out);
}
void tpref_quat_exp(const double* q, double* out) {
  PutQuat(trajectory_planning::bsplineq_details::QuatExp(QuatOf(q)), out);
}
void tpref_quat_power(const double* q, double power, double* out) {
  PutQuat(trajectory_planning::bsplineq_details::QuatPower(QuatOf(q), power), out);
}
void tpref_quat_power_derivative(const double* q, double power, double* out) {
  PutQuat(trajectory_planning::bsplineq_details::QuatPowerDerivative(QuatOf(q), power), out);
}

// BSplineQ::Init + EvalCurve (dq_out == NULL) or EvalCurveAndDerivative at M parameters.
// status [M]: absl::StatusCode of each evaluation.  Returns Init's code.
int tpref_bsplineq_eval(int degree, int num_knots, const double* knots, int num_points,
                        const double* quats, int M, const double* u, double* q_out,
                        double* dq_out, int* status) {
  std::vector<Quaterniond> pts;
  for (int i = 0; i < num_points; ++i) pts.push_back(QuatOf(quats + 4 * static_cast<size_t>(i)));
  trajectory_planning::BSplineQ spline;
  absl::Status st = spline.Init(degree, pts, absl::MakeConstSpan(knots, num_knots));
  if (!st.ok()) return Code(st);
  for (int i = 0; i < M; ++i) {
    Quaterniond q, dq;
    st = dq_out ? spline.EvalCurveAndDerivative(u[i], q, dq) : spline.EvalCurve(u[i], q);
    status[i] = Code(st);
    PutQuat(q, q_out + 4 * static_cast<size_t>(i));
    if (dq_out) PutQuat(dq, dq_out + 4 * static_cast<size_t>(i));
  }
  return 0;
}

// PolyLineToBspline3Waypoints for poses (splines/spline_utils.cc); returns the number
// of control points written to out [cap][7].
int tpref_pose_polyline_to_bspline3(const double* corners, int W, double translation_radius,
                                    double rotational_radius, double* out, int cap) {
  std::vector<Pose3d> ctrl;
  trajectory_planning::PolyLineToBspline3Waypoints(Poses(corners, W), translation_radius,
                                                   rotational_radius, &ctrl);
  for (size_t i = 0; i < ctrl.size() && static_cast<int>(i) < cap; ++i) PutPose(ctrl[i], out + 7 * i);
  return static_cast<int>(ctrl.size());
}

// Make the planner of tpref_planner_create plan over a TimeableCartesianSplinePath;
// call before the first tpref_planner_set_cart_waypoints.  `rounding` of
// tpref_planner_create is the rotation rounding.
void tpref_planner_set_cartesian(void* hv, double translation_rounding,
                                 double max_translational_velocity,
                                 double max_rotational_velocity, tpref_ik_fn ik,
                                 tpref_jacobian_fn jacobian) {
  auto* h = static_cast<Planner*>(hv);
  h->translation_rounding = translation_rounding;
  h->max_translational_velocity = max_translational_velocity;
  h->max_rotational_velocity = max_rotational_velocity;
  h->ik = ik;
  h->jacobian = jacobian;
}

// TimeableCartesianSplinePath::SetWaypoints (+ SetPath on the first call); poses [W][7]
int tpref_planner_set_cart_waypoints(void* hv, int W, const double* pose_waypoints,
                                     const double* joint_waypoints) {
  auto* h = static_cast<Planner*>(hv);
  const std::vector<VectorXd> jw = Rows(joint_waypoints, W, h->D);
  const std::vector<Pose3d> pw = Poses(pose_waypoints, W);
  if (h->cart == nullptr) {
    if (!(h->delta > 0)) {  // one planning window: delta = knots.back() / (N - 1)
      TimeableCartesianSplinePath tmp(CartOptions(h, 3, 1.0));
      absl::Status st = tmp.SetWaypoints(pw, jw);
      if (!st.ok()) return Code(st);
      h->delta = tmp.knots_.back() / (h->N - 1);
    }
    h->cart = std::make_shared<TimeableCartesianSplinePath>(CartOptions(h, h->N, h->delta));
    absl::Status st = h->cart->SetMaxJointVelocity(h->vmax);
    if (st.ok()) st = h->cart->SetMaxJointAcceleration(h->amax);
    if (st.ok())
      st = h->cart->SetMaxCartesianVelocity(h->max_translational_velocity,
                                            h->max_rotational_velocity);
    if (st.ok()) st = h->planner->SetPath(h->cart);
    if (!st.ok()) return Code(st);
  }
  absl::Status st = h->cart->SetWaypoints(pw, jw);
  if (!st.ok()) return Code(st);
  return Code(h->cart->SetInitialVelocity(h->init_vel));
}

// TimeableCartesianSplinePath::SwitchToWaypointPath
int tpref_planner_cart_switch(void* hv, double keep_path_until, int W,
                              const double* pose_waypoints, const double* joint_waypoints) {
  auto* h = static_cast<Planner*>(hv);
  if (h->cart == nullptr) return static_cast<int>(absl::StatusCode::kFailedPrecondition);
  return Code(h->cart->SwitchToWaypointPath(keep_path_until, Poses(pose_waypoints, W),
                                            Rows(joint_waypoints, W, h->D)));
}

// Control points of the translation / rotation splines as poses [cap][7]; returns their
// number (the knots and the joint control points: tpref_planner_spline).
int tpref_planner_cart_pose_points(void* hv, double* out, int cap) {
  auto* h = static_cast<Planner*>(hv);
  if (h->cart == nullptr) return 0;
  const auto tp = h->cart->translation_spline_.GetControlPoints();
  const auto rp = h->cart->rotation_spline_.GetControlPoints();
  for (size_t i = 0; i < tp.size() && static_cast<int>(i) < cap; ++i)
    PutPose(Pose3d(rp[i], tp[i]), out + 7 * i);
  return static_cast<int>(tp.size());
}

// The IK solution cache path_ik_positions_ [cap][D]; returns its length.
int tpref_planner_cart_ik_positions(void* hv, double* out, int cap) {
  auto* h = static_cast<Planner*>(hv);
  if (h->cart == nullptr) return 0;
  const auto& ik = h->cart->path_ik_positions_;
  for (size_t i = 0; i < ik.size() && static_cast<int>(i) < cap; ++i)
    for (int d = 0; d < h->D; ++d) out[i * h->D + d] = ik[i][d];
  return static_cast<int>(ik.size());
}

// The targets handed to the last PathIKFunc call: poses [cap][7], joints [cap][D];
// returns their number.
int tpref_planner_cart_targets(void* hv, double* pose_out, double* joint_out, int cap) {
  auto* h = static_cast<Planner*>(hv);
  if (h->cart == nullptr) return 0;
  const auto& pt = h->cart->sampled_pose_targets_;
  const auto& jt = h->cart->sampled_joint_targets_;
  for (size_t i = 0; i < pt.size() && static_cast<int>(i) < cap; ++i) {
    if (pose_out) PutPose(pt[i], pose_out + 7 * i);
    if (joint_out)
      for (int d = 0; d < h->D; ++d) joint_out[i * h->D + d] = jt[i][d];
  }
  return static_cast<int>(pt.size());
}

// One call of the path alone (no planner): SamplePath(path_start) + ConstraintSetup().
int tpref_planner_cart_sample(void* hv, double path_start) {
  auto* h = static_cast<Planner*>(hv);
  if (h->cart == nullptr) return static_cast<int>(absl::StatusCode::kFailedPrecondition);
  absl::Status st = h->cart->SamplePath(path_start);
  if (!st.ok()) return Code(st);
  return Code(h->cart->ConstraintSetup());
}

}  // extern "C"

extern "C" {

// BASELINE cfg5 on the host: B independent fresh Cartesian paths (D dofs, W pose + joint
// waypoints each), planned by the reference's own classes (SetWaypoints -> Plan:
// SamplePath with the PathIKFunc, ConstraintSetup with the JacobianFunc,
// OptimizePathParameter, ResampleTrajectory), the whole path in one window, on
// `nthreads` host threads; same conventions as tpref_plan_joint_batch.
//   pose_waypoints [B][W][7], joint_waypoints [B][W][D], vmax / amax [B][D]
//   status [B] absl code, duration [B] (NaN on failure), rows [B];
//   optional prof_time [B][N]
void tpref_plan_cart_batch(int B, int N, int D, int W, const double* pose_waypoints,
                           const double* joint_waypoints, const double* vmax,
                           const double* amax, double rotation_rounding,
                           double translation_rounding, double safety,
                           double max_translational_velocity, double max_rotational_velocity,
                           int64_t dt_ns, int skip_method, int nthreads, tpref_ik_fn ik,
                           tpref_jacobian_fn jacobian, int32_t* status, double* duration,
                           int32_t* rows, double* prof_time) {
  Planner proto;
  proto.D = D;
  proto.rounding = rotation_rounding;
  proto.safety = safety;
  proto.translation_rounding = translation_rounding;
  proto.ik = ik;
  proto.jacobian = jacobian;
  std::atomic<int> next{0};
  auto worker = [&]() {
    PathTimingTrajectory planner(PlannerOptions(N, D, dt_ns, skip_method, 0, 0.0));
    for (;;) {
      const int b = next.fetch_add(1);
      if (b >= B) break;
      status[b] = 0;
      duration[b] = std::numeric_limits<double>::quiet_NaN();
      rows[b] = 0;
      const std::vector<VectorXd> jw = Rows(joint_waypoints + static_cast<size_t>(b) * W * D, W, D);
      const std::vector<Pose3d> pw = Poses(pose_waypoints + static_cast<size_t>(b) * W * 7, W);
      std::shared_ptr<TimeableCartesianSplinePath> path;
      double kend = 0.0;
      absl::Status st;
      {
        TimeableCartesianSplinePath tmp(CartOptions(&proto, 3, 1.0));
        st = tmp.SetWaypoints(pw, jw);
        if (st.ok()) kend = tmp.knots_.back();
      }
      if (st.ok()) {
        path = std::make_shared<TimeableCartesianSplinePath>(
            CartOptions(&proto, N, kend / (N - 1)));
        st = path->SetMaxJointVelocity(absl::MakeConstSpan(vmax + static_cast<size_t>(b) * D, D));
      }
      if (st.ok())
        st = path->SetMaxJointAcceleration(
            absl::MakeConstSpan(amax + static_cast<size_t>(b) * D, D));
      if (st.ok())
        st = path->SetMaxCartesianVelocity(max_translational_velocity, max_rotational_velocity);
      if (st.ok()) {
        planner.Reset();
        st = planner.SetPath(path);
      }
      if (st.ok()) st = path->SetWaypoints(pw, jw);
      if (st.ok()) st = planner.Plan(absl::FromUnixNanos(0), absl::Seconds(1));
      status[b] = Code(st);
      if (!st.ok()) continue;
      duration[b] = planner.profile_.GetTotalDuration();
      rows[b] = static_cast<int32_t>(planner.GetTime().size());
      if (prof_time != nullptr) {
        for (int i = 0; i < N; ++i)
          prof_time[static_cast<size_t>(b) * N + i] = planner.profile_.GetTimeSamples()[i];
      }
    }
  };
  if (nthreads <= 1) {
    worker();
    return;
  }
  std::vector<std::thread> pool;
  for (int t = 0; t < nthreads; ++t) pool.emplace_back(worker);
  for (auto& t : pool) t.join();
}

}  // extern "C"
