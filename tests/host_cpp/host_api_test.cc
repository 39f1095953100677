// Exercises the host C++ mirror of the reference API (host/include/...) the way
// the reference's own tests do (path_timing_trajectory_test.cc:88-173,
// time_optimal_path_timing_test.cc:496-559, splines/bspline_test.cc:730-840) and
// prints the results as one JSON object; tests/test_gpu_host_cpp.py compares
// them with the CPU oracle.  Needs a GPU.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <vector>

#include "trajectory_planning/path_timing_batch.h"
#include "trajectory_planning/path_timing_trajectory.h"
#include "trajectory_planning/path_tools.h"
#include "trajectory_planning/rescale_to_stop.h"
#include "trajectory_planning/trajectory_buffer.h"
#include "trajectory_planning/splines/bspline.h"
#include "trajectory_planning/time_optimal_path_timing.h"
#include "trajectory_planning/timeable_path_cartesian_spline.h"
#include "trajectory_planning/timeable_path_joint_spline.h"

using namespace trajectory_planning;  // NOLINT
using tp::VectorXd;

static int g_fail = 0;
#define EXPECT(cond)                                                     \
  do {                                                                   \
    if (!(cond)) {                                                       \
      fprintf(stderr, "EXPECT failed %s:%d: %s\n", __FILE__, __LINE__, #cond); \
      ++g_fail;                                                          \
    }                                                                    \
  } while (0)

static void PrintArray(const char* name, const std::vector<double>& v, bool comma = true) {
  printf("\"%s\": [", name);
  for (size_t i = 0; i < v.size(); ++i) printf("%s%.17g", i ? "," : "", v[i]);
  printf("]%s\n", comma ? "," : "");
}
static std::vector<double> Flatten(const std::vector<VectorXd>& v) {
  std::vector<double> out;
  for (const auto& x : v) out.insert(out.end(), x.begin(), x.end());
  return out;
}

// ---- the reference test's fake kinematics (path_timing_trajectory_test.cc:548-587)
// as host callbacks: translation = q[0:3], rotation = ExpSO3(q[3:6]).
constexpr int kFakeDofs = 7;
static tp::Quaterniond ExpSO3(double x, double y, double z) {
  const double th2 = x * x + y * y + z * z, th = std::sqrt(th2);
  const double k = (th < 1e-8) ? 0.5 - th2 / 48.0 : std::sin(0.5 * th) / th;
  return tp::Quaterniond(std::cos(0.5 * th), k * x, k * y, k * z);
}
static void LogSO3(const tp::Quaterniond& q, double* v) {
  const double n = std::sqrt(q.x() * q.x() + q.y() * q.y() + q.z() * q.z());
  const double k = (n < 1e-10) ? 2.0 / q.w() : 2.0 * std::atan2(n, q.w()) / n;
  v[0] = k * q.x();
  v[1] = k * q.y();
  v[2] = k * q.z();
}
static tp::Pose3d FakeFK(const VectorXd& j) {
  return tp::Pose3d(ExpSO3(j[3], j[4], j[5]), tp::Vector3d(j[0], j[1], j[2]));
}
static tp::Status FakePathIK(const VectorXd&, const std::vector<tp::Pose3d>& pose,
                             const std::vector<VectorXd>& joint, std::vector<VectorXd>* solution) {
  solution->resize(pose.size());
  for (size_t i = 0; i < pose.size(); ++i) {
    VectorXd s(kFakeDofs);
    for (int k = 0; k < 3; ++k) s[k] = pose[i].translation()[k];
    LogSO3(pose[i].quaternion(), &s[3]);
    s[6] = joint[i][6];
    (*solution)[i] = s;
  }
  return tp::OkStatus();
}
static tp::Status FakeJacobian(const VectorXd& j, tp::Matrix6Xd* jac) {
  const tp::Quaterniond q = ExpSO3(j[3], j[4], j[5]);
  const double qv[3] = {q.x(), q.y(), q.z()};
  const double n2 = qv[0] * qv[0] + qv[1] * qv[1] + qv[2] * qv[2], n = std::sqrt(n2), w = q.w();
  const double den = n2 + w * w;
  double k, c;
  if (n < 1e-6) {
    k = 2.0 / w;
    c = (-4.0 / 3.0) / (w * w * w);
  } else {
    k = 2.0 * std::atan2(n, w) / n;
    c = (2.0 * w / den - k) / n2;
  }
  jac->setZero();
  for (int r = 0; r < 3; ++r) {
    (*jac)(r, r) = 1.0;
    for (int col = 0; col < 3; ++col) (*jac)(3 + r, 3 + col) = ((r == col) ? k : 0.0) + c * qv[r] * qv[col];
    (*jac)(3 + r, 6) = -2.0 * qv[r] / den;
  }
  return tp::OkStatus();
}

int main() {
  constexpr size_t kNDof = 3, kNumSamples = 1000;
  printf("{\n");
  // ---- PathErrors (path_timing_trajectory_test.cc:88-109)
  {
    auto more_dofs = std::make_shared<TimeableJointSplinePath>(
        JointPathOptions().set_num_dofs(kNDof + 1).set_num_path_samples(kNumSamples));
    auto more_samples = std::make_shared<TimeableJointSplinePath>(
        JointPathOptions().set_num_dofs(kNDof).set_num_path_samples(kNumSamples + 1));
    PathTimingTrajectory planner(PathTimingTrajectoryOptions()
                                     .SetTimeStep(tp::Milliseconds(4))
                                     .SetNumDofs(kNDof)
                                     .SetNumPathSamples(kNumSamples));
    EXPECT(planner.SetPath(nullptr).code() == tp::StatusCode::kInvalidArgument);
    EXPECT(planner.SetPath(more_dofs).code() == tp::StatusCode::kInvalidArgument);
    EXPECT(planner.SetPath(more_samples).code() == tp::StatusCode::kInvalidArgument);
    EXPECT(planner.Plan(TimeFromSec(0), tp::Seconds(1)).code() == tp::StatusCode::kFailedPrecondition);
  }
  // ---- RestToRestPlanningWorks (:112-173), with delta chosen so that the whole
  // path fits one planning window
  {
    auto path = std::make_shared<TimeableJointSplinePath>(JointPathOptions()
                                                              .set_num_dofs(kNDof)
                                                              .set_num_path_samples(kNumSamples)
                                                              .set_delta_parameter(0.016));
    PathTimingTrajectory planner(PathTimingTrajectoryOptions()
                                     .SetTimeStep(tp::Milliseconds(4))
                                     .SetNumDofs(kNDof)
                                     .SetNumPathSamples(kNumSamples));
    EXPECT(planner.SetPath(path).ok());
    EXPECT(path->GetState() == TimeablePath::State::kNoPath);
    std::vector<VectorXd> wps = {VectorXd{1, 2, 3}, VectorXd{-1, -2, -3}, VectorXd{1, 2, 3}};
    EXPECT(path->SetWaypoints(wps).ok());
    EXPECT(path->GetState() == TimeablePath::State::kNewPath);
    EXPECT(path->SetWaypoints(std::vector<VectorXd>{VectorXd{1, 2}}).code() ==
           tp::StatusCode::kInvalidArgument);
    EXPECT(path->SetWaypoints(wps).ok());
    VectorXd vmax = VectorXd::Constant(kNDof, 1.0), amax = VectorXd::Constant(kNDof, 2.0);
    EXPECT(path->SetMaxJointVelocity(ToSpan(vmax)).ok());
    EXPECT(path->SetMaxJointAcceleration(ToSpan(amax)).ok());
    EXPECT(path->SetMaxJointVelocity(std::vector<double>{1.0}).code() == tp::StatusCode::kInvalidArgument);
    const tp::Status st = planner.Plan(TimeFromSec(0.0), tp::Seconds(100));
    EXPECT(st.ok());
    if (!st.ok()) fprintf(stderr, "Plan: %s\n", st.message().c_str());
    EXPECT(path->GetState() == TimeablePath::State::kPathWasSampled);
    EXPECT(planner.IsTrajectoryAtEnd());
    const auto& pos = planner.GetPositions();
    EXPECT(!pos.empty());
    if (!pos.empty()) {
      EXPECT(pos.back() == wps.back());  // ends exactly at the last waypoint (:167-172)
      for (size_t d = 0; d < kNDof; ++d) EXPECT(planner.GetVelocities().back()[d] == 0.0);
    }
    // a second Plan() inside the planned span only trims the front (:590-601)
    const size_t m0 = planner.GetNumTimeSamples();
    EXPECT(planner.Plan(TimeFromSec(0.2), tp::Seconds(0.1)).ok());
    EXPECT(planner.GetNumTimeSamples() == m0 - 50);
    EXPECT(std::abs(planner.GetTime().front() - 0.2) < 1e-12);
    printf("\"rest_to_rest\": {\n");
    PrintArray("knots", path->GetKnots());
    PrintArray("control_points", Flatten(path->GetControlPoints()));
    PrintArray("time", planner.GetTime());
    PrintArray("positions", Flatten(planner.GetPositions()));
    PrintArray("velocities", Flatten(planner.GetVelocities()));
    PrintArray("accelerations", Flatten(planner.GetAccelerations()));
    PrintArray("path_parameter", planner.GetPathParameters());
    PrintArray("q0", std::vector<double>(path->GetPathPositionAt(0).begin(), path->GetPathPositionAt(0).end()));
    printf("\"end_time_ns\": %lld, \"final_decel_start_ns\": %lld, \"trimmed\": 50},\n",
           static_cast<long long>(planner.GetEndTime().ns),
           static_cast<long long>(planner.GetFinalDecelStart().ns));
    // constraint rows of sample 10 (ConstraintSetup, timeable_path_joint_spline.cc:320-343)
    const auto& c = path->GetConstraints()[10];
    std::vector<double> rows;
    for (int r = 0; r < c.size(); ++r) {
      rows.push_back(c.a_coefficient(r));
      rows.push_back(c.b_coefficient(r));
      rows.push_back(c.lower(r));
      rows.push_back(c.upper(r));
    }
    PrintArray("constraints_sample10", rows);
  }
  // ---- RestToRestPlanningWorks (:112-173) as the reference runs it: default path
  // sampling distance (several planning windows), Plan() called again every 200 ms
  // with a 750 ms horizon until the trajectory is at its end; both sampling methods
  for (int method = 0; method < 2; ++method) {
    std::vector<VectorXd> wps = {VectorXd{1, 2, 3}, VectorXd{-1, -2, -3}, VectorXd{1, 2, 3}};
    VectorXd vmax = VectorXd::Constant(kNDof, 1.0), amax = VectorXd::Constant(kNDof, 2.0);
    auto path2 = std::make_shared<TimeableJointSplinePath>(
        JointPathOptions().set_num_dofs(kNDof).set_num_path_samples(kNumSamples));
    PathTimingTrajectory planner2(
        PathTimingTrajectoryOptions()
            .SetTimeStep(tp::Milliseconds(4))
            .SetNumDofs(kNDof)
            .SetNumPathSamples(kNumSamples)
            .SetTimeSamplingMethod(method == 0
                                       ? PathTimingTrajectoryOptions::TimeSamplingMethod::kUniformlyInTime
                                       : PathTimingTrajectoryOptions::TimeSamplingMethod::
                                             kSkipSamplesCloserThanTimeStep));
    EXPECT(planner2.SetPath(path2).ok());
    EXPECT(path2->SetWaypoints(wps).ok());
    path2->SetMaxJointVelocity(ToSpan(vmax));
    path2->SetMaxJointAcceleration(ToSpan(amax));
    tp::Time start_time = TimeFromSec(0.0);
    int loops = 0, windows = 0;
    std::vector<double> starts, counts, first_times, end_times;
    for (; !planner2.IsTrajectoryAtEnd() && loops < 500; ++loops) {
      const tp::Status st = planner2.Plan(start_time, tp::Milliseconds(750));
      EXPECT(st.ok());
      if (!st.ok()) {
        fprintf(stderr, "Plan %d: %s\n", loops, st.message().c_str());
        break;
      }
      windows += planner2.LastPlanWindows();
      EXPECT(planner2.GetPositions().size() == planner2.GetVelocities().size());
      EXPECT(planner2.GetPositions().size() == planner2.GetAccelerations().size());
      starts.push_back(static_cast<double>(start_time.ns));
      counts.push_back(static_cast<double>(planner2.GetNumTimeSamples()));
      first_times.push_back(planner2.GetTime().front());
      end_times.push_back(static_cast<double>(planner2.GetEndTime().ns));
      const tp::Time next = start_time + tp::Milliseconds(200);
      start_time = planner2.GetEndTime() < next ? planner2.GetEndTime() : next;
    }
    EXPECT(planner2.IsTrajectoryAtEnd());
    EXPECT(windows >= 3);
    EXPECT(!planner2.GetPositions().empty());
    if (!planner2.GetPositions().empty()) {
      EXPECT(planner2.GetPositions().back() == wps.back());
      for (size_t d = 0; d < kNDof; ++d) EXPECT(planner2.GetVelocities().back()[d] == 0.0);
    }
    // a start past the end of the plan is rejected (HandleTimeArguments, :503-511)
    EXPECT(planner2.Plan(planner2.GetEndTime() + tp::Seconds(1.0), tp::Milliseconds(750)).code() ==
           tp::StatusCode::kOutOfRange);
    printf("\"%s\": {\n", method == 0 ? "replanning" : "replanning_skip");
    PrintArray("starts_ns", starts);
    PrintArray("counts", counts);
    PrintArray("first_times", first_times);
    PrintArray("end_times_ns", end_times);
    PrintArray("time", planner2.GetTime());
    PrintArray("positions", Flatten(planner2.GetPositions()));
    PrintArray("velocities", Flatten(planner2.GetVelocities()));
    PrintArray("accelerations", Flatten(planner2.GetAccelerations()));
    PrintArray("path_parameter", planner2.GetPathParameters());
    printf("\"loops\": %d, \"windows\": %d, \"final_decel_start_ns\": %lld},\n", loops, windows,
           static_cast<long long>(planner2.GetFinalDecelStart().ns));
  }
  // ---- ResetAndReplan (path_timing_trajectory_test.cc:440-563): plan to the end while
  // collecting every window's rows in a TrajectoryBuffer, Reset(), set the same waypoints
  // again, plan to the end again -- both buffers hold the same rows; both sampling methods
  for (int method = 0; method < 2; ++method) {
    std::vector<VectorXd> wps = {VectorXd{1, 2, 3}, VectorXd{-1, -2, -3}, VectorXd{0.5, 1.0, 1.5}};
    VectorXd vmax = VectorXd::Constant(kNDof, 1.0), amax = VectorXd::Constant(kNDof, 2.0);
    auto path3 = std::make_shared<TimeableJointSplinePath>(
        JointPathOptions().set_num_dofs(kNDof).set_num_path_samples(kNumSamples));
    PathTimingTrajectory planner3(
        PathTimingTrajectoryOptions()
            .SetTimeStep(tp::Milliseconds(4))
            .SetNumDofs(kNDof)
            .SetNumPathSamples(kNumSamples)
            .SetTimeSamplingMethod(method == 0
                                       ? PathTimingTrajectoryOptions::TimeSamplingMethod::kUniformlyInTime
                                       : PathTimingTrajectoryOptions::TimeSamplingMethod::
                                             kSkipSamplesCloserThanTimeStep));
    EXPECT(planner3.SetPath(path3).ok());
    EXPECT(path3->SetWaypoints(wps).ok());
    path3->SetMaxJointVelocity(ToSpan(vmax));
    path3->SetMaxJointAcceleration(ToSpan(amax));
    auto plan_to_end = [&](TrajectoryBuffer& buffer) {
      tp::Time start_time = TimeFromSec(0.0);
      int loops = 0;
      for (; !planner3.IsTrajectoryAtEnd() && loops < 500; ++loops) {
        const tp::Status st = planner3.Plan(start_time, tp::Milliseconds(750));
        EXPECT(st.ok());
        if (!st.ok()) break;
        EXPECT(planner3.GetPositions().size() == planner3.GetVelocities().size());
        EXPECT(planner3.GetPositions().size() == planner3.GetAccelerations().size());
        EXPECT(buffer
                   .InsertSegment(planner3.GetTime(), planner3.GetPositions(),
                                  planner3.GetVelocities(), planner3.GetAccelerations())
                   .ok());
        start_time = planner3.GetNextPlanStartTime(start_time + tp::Milliseconds(200));
      }
      return loops;
    };
    auto first = TrajectoryBuffer::Create();
    auto second = TrajectoryBuffer::Create();
    EXPECT(first.ok() && second.ok());
    const int loops1 = plan_to_end(**first);
    EXPECT(planner3.IsTrajectoryAtEnd());
    planner3.Reset();
    EXPECT(!planner3.IsTrajectoryAtEnd());
    EXPECT(planner3.GetNumTimeSamples() == 0);
    EXPECT(path3->SetWaypoints(wps).ok());
    const int loops2 = plan_to_end(**second);
    EXPECT(planner3.IsTrajectoryAtEnd());
    EXPECT(loops1 == loops2 && loops1 >= 3);
    const size_t rows = (*first)->GetNumSamples();
    EXPECT(rows > 100 && rows == (*second)->GetNumSamples());
    bool same = rows == (*second)->GetNumSamples();
    for (size_t i = 0; same && i < rows; ++i) {
      same = (*first)->GetTimes()[i] == (*second)->GetTimes()[i] &&
             (*first)->GetPositions()[i] == (*second)->GetPositions()[i] &&
             (*first)->GetVelocities()[i] == (*second)->GetVelocities()[i] &&
             (*first)->GetAccelerations()[i] == (*second)->GetAccelerations()[i];
    }
    EXPECT(same);
    printf("\"%s\": {\n", method == 0 ? "reset_and_replan" : "reset_and_replan_skip");
    PrintArray("time", std::vector<double>((*second)->GetTimes().begin(), (*second)->GetTimes().end()));
    PrintArray("positions", Flatten(std::vector<VectorXd>((*second)->GetPositions().begin(),
                                                          (*second)->GetPositions().end())));
    printf("\"loops\": %d, \"rows\": %zu, \"same\": %d},\n", loops2, rows, same ? 1 : 0);
  }
  // ---- SwitchToNewJointWaypointPathWorks (path_timing_trajectory_test.cc:299-436):
  // plan until the motion is fast, switch to new waypoints at the stop parameter with the
  // current velocity as the initial velocity, plan to the end of the new path
  {
    std::vector<VectorXd> wps = {VectorXd{1, 2, 3}, VectorXd{-1, -2, -3}, VectorXd{0.5, 1.0, 1.5}};
    VectorXd vmax = VectorXd::Constant(kNDof, 1.0), amax = VectorXd::Constant(kNDof, 2.0);
    auto path = std::make_shared<TimeableJointSplinePath>(JointPathOptions()
                                                              .set_num_dofs(kNDof)
                                                              .set_num_path_samples(kNumSamples)
                                                              .set_delta_parameter(0.001));
    PathTimingTrajectory planner(PathTimingTrajectoryOptions()
                                     .SetTimeStep(tp::Milliseconds(4))
                                     .SetNumDofs(kNDof)
                                     .SetNumPathSamples(kNumSamples));
    EXPECT(planner.SetPath(path).ok());
    EXPECT(path->SetWaypoints(wps).ok());
    path->SetMaxJointVelocity(ToSpan(vmax));
    path->SetMaxJointAcceleration(ToSpan(amax));
    tp::Time start_time = TimeFromSec(0.0);
    int first_calls = 0;
    for (; first_calls < 100; ++first_calls) {
      EXPECT(planner.Plan(start_time, tp::Milliseconds(750)).ok());
      start_time = planner.GetNextPlanStartTime(start_time + tp::Milliseconds(200));
      double closest = 1e9;
      for (size_t d = 0; d < kNDof; ++d) {
        closest = std::min(closest, std::abs(std::abs(planner.GetVelocities().front()[d]) - vmax[d]));
      }
      if (closest < 0.3) break;
    }
    EXPECT(!planner.IsTrajectoryAtEnd());
    const std::vector<VectorXd> new_wps = {VectorXd{1.0, 2.0, 3.0}, VectorXd{0.5, 1.0, 5.5}};
    auto stop = planner.GetPathStopParameter(start_time);
    EXPECT(stop.ok());
    const double stop_parameter = stop.ok() ? *stop : 0.0;
    // the velocity at start_time, interpolated in the planned trajectory
    VectorXd v0(kNDof);
    {
      const auto& t = planner.GetTime();
      const double ts = TimeToSec(start_time);
      size_t hi = 1;
      while (hi + 1 < t.size() && t[hi] < ts) ++hi;
      const double w = (ts - t[hi - 1]) / (t[hi] - t[hi - 1]);
      for (size_t d = 0; d < kNDof; ++d) {
        v0[d] = planner.GetVelocities()[hi - 1][d] +
                w * (planner.GetVelocities()[hi][d] - planner.GetVelocities()[hi - 1][d]);
      }
    }
    const size_t knots_before = path->GetKnots().size();
    EXPECT(path->SwitchToWaypointPath(stop_parameter, new_wps).ok());
    EXPECT(path->GetState() == TimeablePath::State::kModifiedPath);
    EXPECT(path->SetInitialVelocity(ToSpan(v0)).ok());
    EXPECT(!planner.IsTrajectoryAtEnd());
    const std::vector<double> knots_after = path->GetKnots();
    int second_calls = 0;
    for (; !planner.IsTrajectoryAtEnd() && second_calls < 300; ++second_calls) {
      const tp::Status st = planner.Plan(start_time, tp::Milliseconds(750));
      EXPECT(st.ok());
      if (!st.ok()) {
        fprintf(stderr, "Plan after switch %d: %s\n", second_calls, st.message().c_str());
        break;
      }
      start_time = planner.GetNextPlanStartTime(start_time + tp::Milliseconds(200));
    }
    EXPECT(planner.IsTrajectoryAtEnd());
    EXPECT(path->GetState() == TimeablePath::State::kPathWasSampled);
    if (!planner.GetPositions().empty()) {
      for (size_t d = 0; d < kNDof; ++d) {
        EXPECT(std::abs(planner.GetPositions().back()[d] - new_wps.back()[d]) < 1e-10);
        EXPECT(std::abs(planner.GetVelocities().back()[d]) < 1e-10);
      }
    }
    // ProjectPointOnPath / ComputeStoppingPoint (path_tools_test.cc:98-165)
    {
      const std::vector<VectorXd> w3 = {VectorXd{1.0, 1.0}, VectorXd{2.0, 2.0}, VectorXd{-3.0, -3.0}};
      const VectorXd pt{2.0 + 0.4 * -5.0 + 0.1, 2.0 + 0.4 * -5.0 - 0.1};
      auto pr = ProjectPointOnPath(w3, pt);
      EXPECT(pr.ok() && pr->waypoint_index == 1 && std::abs(pr->line_parameter - 0.4) < 1e-14);
      EXPECT(ProjectPointOnPath(std::vector<VectorXd>{}, pt).status().code() ==
             tp::StatusCode::kInvalidArgument);
      const std::vector<double> pos(3, 1.0), vel0(3, 0.0), vel{0.3, -0.2, 0.1}, acc(3, 1.0);
      auto still = ComputeStoppingPoint(pos, vel0, acc, 0.1);
      EXPECT(still.ok() && (*still)[0] == 1.0 && (*still)[2] == 1.0);
      auto stopped = ComputeStoppingPoint(pos, vel, acc, 0.1);
      EXPECT(stopped.ok());
      if (stopped.ok()) {  // on the ray from the position along the velocity
        const double k = ((*stopped)[0] - 1.0) / vel[0];
        EXPECT(k > 0 && std::abs((*stopped)[1] - 1.0 - k * vel[1]) < 1e-12 &&
               std::abs((*stopped)[2] - 1.0 - k * vel[2]) < 1e-12);
      }
      EXPECT(ComputeStoppingPoint(pos, std::vector<double>(2), acc, 0.0).status().code() ==
             tp::StatusCode::kInvalidArgument);
      EXPECT(ComputeStoppingPoint(pos, vel, std::vector<double>(3, -1.0), 0.0).status().code() ==
             tp::StatusCode::kInvalidArgument);
    }
    // BSplineT knot insertion / TruncateSplineAt / ExtendWithControlPoints on the host
    // copy (splines/bspline.h:224-511): same knots and control points as the device's
    // SwitchToWaypointPath building blocks, curve unchanged by insertion
    std::vector<double> bs_knots, bs_ctrl, bs_before, bs_after;
    {
      auto fresh = std::make_shared<TimeableJointSplinePath>(
          JointPathOptions().set_num_dofs(kNDof).set_num_path_samples(kNumSamples));
      EXPECT(fresh->SetWaypoints(wps).ok());
      BSplineXd sp;
      EXPECT(sp.Init(2, 100, kNDof).ok());
      EXPECT(sp.SetKnotVector(fresh->GetKnots()).ok());
      EXPECT(sp.SetControlPoints(fresh->GetControlPoints()).ok());
      const double kend = fresh->GetKnots().back();
      std::vector<double> us;
      for (int i = 0; i <= 40; ++i) us.push_back(0.59 * kend * i / 40.0);
      bs_before.resize(us.size() * kNDof);
      bs_after.resize(us.size() * kNDof);
      EXPECT(sp.EvalCurveBatch(us, bs_before.data(), nullptr).ok());
      EXPECT(sp.InsertKnotAndUpdateControlPointsRef(0.0).code() == tp::StatusCode::kInvalidArgument);
      EXPECT(sp.InsertKnotAndUpdateControlPointsRef(0.3 * kend, 2).ok());
      EXPECT(sp.TruncateSplineAt(0.6 * kend).ok());
      EXPECT(sp.EvalCurveBatch(us, bs_after.data(), nullptr).ok());
      const std::vector<VectorXd> extra = {VectorXd{0.1, 0.2, 0.3}, VectorXd{0.4, 0.1, 0.0},
                                           VectorXd{0.7, 0.0, -0.2}, VectorXd{1.0, 0.0, -0.3}};
      EXPECT(sp.ExtendWithControlPoints(std::vector<VectorXd>{extra[0]}).code() ==
             tp::StatusCode::kUnimplemented);
      EXPECT(sp.ExtendWithControlPoints(extra).ok());
      const auto kv = sp.GetKnotVector();
      bs_knots.assign(kv.begin(), kv.end());
      bs_ctrl = Flatten(sp.GetControlPoints());
    }
    printf("\"bspline_edit\": {\n");
    PrintArray("knots", bs_knots);
    PrintArray("control_points", bs_ctrl);
    PrintArray("before", bs_before);
    PrintArray("after", bs_after, false);
    printf("},\n");
    printf("\"switching\": {\n");
    PrintArray("initial_velocity", std::vector<double>(v0.begin(), v0.end()));
    PrintArray("knots_after_switch", knots_after);
    PrintArray("time", planner.GetTime());
    PrintArray("positions", Flatten(planner.GetPositions()));
    PrintArray("velocities", Flatten(planner.GetVelocities()));
    PrintArray("accelerations", Flatten(planner.GetAccelerations()));
    printf("\"stop_parameter\": %.17g, \"first_calls\": %d, \"second_calls\": %d, "
           "\"knots_before\": %zu, \"switch_start_ns\": %lld},\n",
           stop_parameter, first_calls, second_calls, knots_before,
           0LL);
  }
  // ---- Circle (time_optimal_path_timing_test.cc:496-559)
  {
    const double R = 2.0, vmax = 1.2, amax = 1.0;
    printf("\"circle\": [\n");
    const struct { int n; double sd; } cases[] = {{50, 0.0}, {51, 0.0}, {50, 0.1}, {51, 0.1}};
    for (int ci = 0; ci < 4; ++ci) {
      const int N = cases[ci].n;
      std::vector<TimeOptimalPathProfile::Constraint> cons(N);
      const double ds = M_PI / (N - 1);
      for (int i = 0; i < N; ++i) {
        const double s = i * ds;
        auto& c = cons[i];
        c.resize(4);
        c.a_coefficient(0) = -R * sin(s);
        c.a_coefficient(1) = R * cos(s);
        c.b_coefficient(0) = -R * cos(s);
        c.b_coefficient(1) = -R * sin(s);
        c.b_coefficient(2) = pow(R * sin(s), 2);
        c.b_coefficient(3) = pow(R * cos(s), 2);
        c.upper(0) = c.upper(1) = amax;
        c.lower(0) = c.lower(1) = -amax;
        c.upper(2) = c.upper(3) = pow(vmax, 2);
      }
      TimeOptimalPathTiming opt;  // BASELINE.json's name for TimeOptimalPathProfile
      EXPECT(!opt.OptimizePathParameter());  // not set up yet
      EXPECT(opt.InitSolver(N, 4));
      EXPECT(opt.SetupProblem(cons, 0.0, M_PI, cases[ci].sd, 0.0, 0.0));
      EXPECT(opt.OptimizePathParameter());
      double s0, sd0, sdd0;
      EXPECT(opt.GetPathParameterAndDerivatives(0.0, &s0, &sd0, &sdd0));
      EXPECT(std::abs(sd0 - cases[ci].sd) < 1e-15);
      EXPECT(opt.SolutionSatisfiesConstraints().ok());
      double smx, sddmx, szero;
      opt.FindMaxSd2Simplex(cons[N / 3], &smx, &sddmx, &szero);
      printf("{\"N\": %d, \"sd_start\": %.17g, \"duration\": %.17g, \"last_extremal\": %d, "
             "\"lp_sd2max\": %.17g, \"lp_sd2zero\": %.17g, \"dt_max\": %.17g}%s\n",
             N, cases[ci].sd, opt.GetTotalDuration(), opt.GetLastExtremalIndex(), smx, szero,
             opt.GetMaxTimeIncrement(), ci < 3 ? "," : "");
      // setup errors return false like the reference
      cons[3].lower(1) = cons[3].upper(1);
      EXPECT(!opt.SetupProblem(cons, 0.0, M_PI, 0.0, 0.0, 0.0));
    }
    printf("],\n");
  }
  // ---- BSpline golden configuration (splines/bspline_test.cc:730-840)
  {
    BSplineXd spline;
    std::vector<double> knots = {0, 0, 0, 0, 0.5, 1, 1, 1, 1};
    EXPECT(spline.Init(3, knots.size(), 2).ok());
    EXPECT(spline.SetKnotVector(knots).ok());
    std::vector<VectorXd> poly = {VectorXd{1, 1}, VectorXd{2, 3}, VectorXd{3, -1}, VectorXd{4, 1}, VectorXd{5, 0}};
    EXPECT(spline.SetControlPoints(poly).ok());
    EXPECT(spline.SetControlPoints(std::vector<VectorXd>(4, VectorXd{0, 0})).code() ==
           tp::StatusCode::kInvalidArgument);
    std::vector<double> us(101);
    for (int i = 0; i < 101; ++i) us[i] = 0.0 + i * (1.0 / 100);
    std::vector<double> out(101 * 4 * 2);
    std::vector<int32_t> st(101);
    EXPECT(spline.EvalCurveAndDerivativesBatch(us, 3, out.data(), st.data()).ok());
    VectorXd p(2);
    EXPECT(spline.EvalCurve(0.5, p).ok());
    EXPECT(spline.EvalCurve(1.5, p).code() == tp::StatusCode::kOutOfRange);
    EXPECT(spline.EvalCurve(-0.1, p).code() == tp::StatusCode::kOutOfRange);
    std::vector<VectorXd> pk(4, VectorXd(2));
    EXPECT(spline.EvalCurveAndDerivatives(0.01, tp::Span<VectorXd>(pk.data(), pk.size())).ok());
    printf("\"bspline\": {\n");
    PrintArray("values", out);
    PrintArray("at_0p01", Flatten(pk), false);
    printf("},\n");
    BSplineXd auto_knots;
    EXPECT(auto_knots.Init(1, 7, 1).ok());
    EXPECT(auto_knots.SetUniformKnotVector(7).ok());
    auto kv = auto_knots.GetKnotVector();
    EXPECT(kv.size() == 7 && kv[2] == 0.25 && kv[3] == 0.5 && kv[4] == 0.75);
  }
  // ---- Cartesian path (path_timing_trajectory_test.cc:589-700): corner rounding
  // setters, then a whole path planned with host IK / Jacobian callbacks and with
  // the built-in device kinematics
  {
    constexpr size_t kN = 1000;
    auto make_path = [&](int model) {
      CartesianPathOptions opt;
      opt.set_path_ik_func(FakePathIK).set_jacobian_func(FakeJacobian);
      opt.set_num_dofs(kFakeDofs).set_num_path_samples(kN).set_delta_parameter(0.02);
      opt.set_device_kinematics(model);
      return std::make_shared<TimeableCartesianSplinePath>(opt);
    };
    auto path = make_path(TPB_KIN_HOST);
    EXPECT(path->SetRotationRounding(0.1).ok());
    EXPECT(path->SetTranslationRounding(0.05).ok());
    EXPECT(path->GetRotationRounding() == 0.1 && path->GetTranslationRounding() == 0.05);
    EXPECT(path->SetRotationRounding(0.0).code() == tp::StatusCode::kInvalidArgument);
    EXPECT(path->SetTranslationRounding(-1.0).code() == tp::StatusCode::kInvalidArgument);
    EXPECT(path->SetMaxCartesianVelocity(0.0, 0.25).code() == tp::StatusCode::kInvalidArgument);
    std::vector<VectorXd> jw = {VectorXd{0.1, 0.0, 0.2, 0.0, 0.1, 0.0, 0.3},
                                VectorXd{0.3, 0.1, 0.1, 0.2, -0.1, 0.3, -0.2},
                                VectorXd{0.2, 0.3, 0.0, -0.1, 0.2, 0.1, 0.1}};
    std::vector<tp::Pose3d> pw;
    for (const auto& j : jw) pw.push_back(FakeFK(j));
    printf("\"cartesian\": {\n");
    for (int mode = 0; mode < 2; ++mode) {
      if (mode == 1) {
        path = make_path(TPB_KIN_FAKE7);
        path->SetRotationRounding(0.1);
        path->SetTranslationRounding(0.05);
      }
      PathTimingTrajectory planner(PathTimingTrajectoryOptions()
                                       .SetTimeStep(tp::Milliseconds(4))
                                       .SetNumDofs(kFakeDofs)
                                       .SetNumPathSamples(kN));
      EXPECT(planner.SetPath(path).ok());
      EXPECT(path->SetWaypoints(pw, std::vector<VectorXd>{jw[0]}).code() == tp::StatusCode::kInvalidArgument);
      EXPECT(path->SetWaypoints(pw, jw).ok());
      EXPECT(path->GetState() == TimeablePath::State::kNewPath);
      VectorXd vmax = VectorXd::Constant(kFakeDofs, 1.0), amax = VectorXd::Constant(kFakeDofs, 2.0);
      EXPECT(path->SetMaxJointVelocity(ToSpan(vmax)).ok());
      EXPECT(path->SetMaxJointAcceleration(ToSpan(amax)).ok());
      EXPECT(path->SetMaxCartesianVelocity(0.5, 0.25).ok());
      const tp::Status st = planner.Plan(TimeFromSec(0.0), tp::Seconds(1000));
      EXPECT(st.ok());
      if (!st.ok()) fprintf(stderr, "Cartesian Plan (mode %d): %s\n", mode, st.message().c_str());
      EXPECT(path->GetState() == TimeablePath::State::kPathWasSampled);
      EXPECT(path->NumConstraints() == 16 && path->GetConstraints().size() == kN);
      EXPECT(path->GetSplineIKPosition().size() == kN);
      EXPECT(path->PathIkIndex(0.1) == 5 && path->PathIkParameter(5) == 0.1);
      const auto& pos = planner.GetPositions();
      EXPECT(!pos.empty());
      printf("\"%s\": {\n", mode == 0 ? "host_callbacks" : "device_kinematics");
      PrintArray("knots", path->GetKnots());
      PrintArray("time", planner.GetTime());
      PrintArray("positions", Flatten(planner.GetPositions()));
      PrintArray("velocities", Flatten(planner.GetVelocities()));
      PrintArray("q_last", std::vector<double>(path->GetPathEnd().begin(), path->GetPathEnd().end()));
      const auto& c = path->GetConstraints()[10];
      std::vector<double> rows;
      for (int r = 0; r < c.size(); ++r) {
        rows.push_back(c.a_coefficient(r));
        rows.push_back(c.b_coefficient(r));
        rows.push_back(c.lower(r));
        rows.push_back(c.upper(r));
      }
      PrintArray("constraints_sample10", rows, false);
      printf("}%s\n", mode == 0 ? "," : "");
    }
    // the reference's 5-waypoint Cartesian scenario (path_timing_trajectory_test.cc:695-757):
    // default delta 0.005, Plan() every 200 ms with a 750 ms horizon until the end; with
    // device kinematics every window is sampled and timed on the device; with the user's
    // PathIKFunc / JacobianFunc every window's new IK indices go through the callbacks,
    // warm-started from the previous window (timeable_path_cartesian_spline.cc:464-549)
    for (int mode = 0; mode < 2; ++mode) {
      constexpr size_t kSamples = 1000;
      int ik_calls = 0, ik_samples = 0;
      bool warm_start_ok = true;
      VectorXd last_solution;
      CartesianPathOptions opt;
      opt.set_num_dofs(kFakeDofs).set_num_path_samples(kSamples);
      if (mode == 0) {
        opt.set_device_kinematics(TPB_KIN_FAKE7);
      } else {
        opt.set_jacobian_func(FakeJacobian);
        opt.set_path_ik_func([&](const VectorXd& initial_value, const std::vector<tp::Pose3d>& pose,
                                 const std::vector<VectorXd>& joint, std::vector<VectorXd>* solution) {
          // the warm start of a later window is the last solution of the previous one
          if (ik_calls > 0 && last_solution.size() == initial_value.size()) {
            for (size_t d = 0; d < initial_value.size(); ++d)
              warm_start_ok = warm_start_ok && initial_value[d] == last_solution[d];
          }
          const tp::Status st = FakePathIK(initial_value, pose, joint, solution);
          ++ik_calls;
          ik_samples += static_cast<int>(pose.size());
          if (!solution->empty()) last_solution = solution->back();
          return st;
        });
      }
      auto cpath = std::make_shared<TimeableCartesianSplinePath>(opt);
      PathTimingTrajectory planner(PathTimingTrajectoryOptions()
                                       .SetTimeStep(tp::Milliseconds(4))
                                       .SetNumDofs(kFakeDofs)
                                       .SetNumPathSamples(kSamples));
      EXPECT(planner.SetPath(cpath).ok());
      const std::vector<VectorXd> joints = {VectorXd{1.0, 0, 0, 0, 0, 0, 0}, VectorXd{0, 1.0, 0, 0, 0, 0, 0},
                                            VectorXd{0, 0, 0, 1.0, 0, 0, 0}, VectorXd{0, 0, 0, 0, 1.0, 0, 0},
                                            VectorXd{0, 0, 0, 0, 0, 0, 1.0}};
      std::vector<tp::Pose3d> poses;
      for (const auto& j : joints) poses.push_back(FakeFK(j));
      EXPECT(cpath->SetWaypoints(poses, joints).ok());
      VectorXd vmax = VectorXd::Constant(kFakeDofs, 1.0), amax = VectorXd::Constant(kFakeDofs, 2.0);
      EXPECT(cpath->SetMaxJointVelocity(ToSpan(vmax)).ok());
      EXPECT(cpath->SetMaxJointAcceleration(ToSpan(amax)).ok());
      EXPECT(cpath->SetMaxCartesianVelocity(0.5, 0.25).ok());
      tp::Time start_time = TimeFromSec(0.0);
      int loops = 0, windows = 0;
      std::vector<double> end_times;
      for (; !planner.IsTrajectoryAtEnd() && loops < 2000; ++loops) {
        const tp::Status st = planner.Plan(start_time, tp::Milliseconds(750));
        EXPECT(st.ok());
        if (!st.ok()) {
          fprintf(stderr, "Cartesian re-planning loop %d: %s\n", loops, st.message().c_str());
          break;
        }
        EXPECT(planner.GetVelocities().size() == planner.GetPositions().size());
        windows += planner.LastPlanWindows();
        end_times.push_back(TimeToSec(planner.GetEndTime()));
        const tp::Time next = start_time + tp::Milliseconds(200);
        start_time = planner.GetEndTime() < next ? planner.GetEndTime() : next;
      }
      EXPECT(planner.IsTrajectoryAtEnd() && windows >= 3);
      EXPECT(cpath->GetState() == TimeablePath::State::kPathWasSampled);
      EXPECT(cpath->GetConstraints().size() == kSamples);
      if (mode == 1) {
        // one IK call per planning window, each on the indices the cache does not hold yet
        // plus the re-evaluated first one; never the whole window again
        EXPECT(ik_calls == windows && warm_start_ok);
        EXPECT(ik_samples < windows * static_cast<int>(kSamples + 1) && ik_samples > static_cast<int>(kSamples));
      }
      printf(",\"%s\": {\n\"loops\": %d, \"windows\": %d, \"ik_calls\": %d, \"ik_samples\": %d,\n",
             mode == 0 ? "replanning" : "replanning_host_callbacks", loops, windows, ik_calls, ik_samples);
      PrintArray("end_times", end_times);
      PrintArray("time", planner.GetTime());
      PrintArray("positions", Flatten(planner.GetPositions()));
      PrintArray("velocities", Flatten(planner.GetVelocities()), false);
      printf("}\n");
    }
    // SwitchToNewCartesianWaypointPathWorks (path_timing_trajectory_test.cc:764-920): plan
    // until the motion is fast, switch the three splines to new waypoints at the stop
    // parameter with the current velocity, plan to the end -- with device kinematics and
    // with the user's callbacks (whose IK cache is cut at the switching index)
    for (int mode = 0; mode < 2; ++mode) {
      constexpr size_t kSamples = 1000;
      int ik_calls = 0;
      CartesianPathOptions opt;
      opt.set_num_dofs(kFakeDofs).set_num_path_samples(kSamples).set_delta_parameter(0.005);
      if (mode == 0) {
        opt.set_device_kinematics(TPB_KIN_FAKE7);
      } else {
        opt.set_jacobian_func(FakeJacobian);
        opt.set_path_ik_func([&](const VectorXd& initial_value, const std::vector<tp::Pose3d>& pose,
                                 const std::vector<VectorXd>& joint, std::vector<VectorXd>* solution) {
          ++ik_calls;
          return FakePathIK(initial_value, pose, joint, solution);
        });
      }
      auto cpath = std::make_shared<TimeableCartesianSplinePath>(opt);
      PathTimingTrajectory planner(PathTimingTrajectoryOptions()
                                       .SetTimeStep(tp::Milliseconds(4))
                                       .SetNumDofs(kFakeDofs)
                                       .SetNumPathSamples(kSamples)
                                       .SetMaxInitialVelocityError(1e-2));
      EXPECT(planner.SetPath(cpath).ok());
      const std::vector<VectorXd> joints = {VectorXd{1.0, 0, 0, 0, 0, 0, 0}, VectorXd{0, 1.0, 0, 0, 0, 0, 0},
                                            VectorXd{0, 0, 0, 1.0, 0, 0, 0}, VectorXd{0, 0, 0, 0, 1.0, 0, 0},
                                            VectorXd{0, 0, 0, 0, 0, 0, 1.0}};
      std::vector<tp::Pose3d> poses;
      for (const auto& j : joints) poses.push_back(FakeFK(j));
      EXPECT(cpath->SetWaypoints(poses, joints).ok());
      VectorXd vmax = VectorXd::Constant(kFakeDofs, 1.0), amax = VectorXd::Constant(kFakeDofs, 2.0);
      EXPECT(cpath->SetMaxJointVelocity(ToSpan(vmax)).ok());
      EXPECT(cpath->SetMaxJointAcceleration(ToSpan(amax)).ok());
      EXPECT(cpath->SetMaxCartesianVelocity(0.5, 0.25).ok());
      tp::Time start_time = TimeFromSec(0.0);
      int first_calls = 0;
      for (; first_calls < 200; ++first_calls) {
        const tp::Status st = planner.Plan(start_time, tp::Milliseconds(750));
        EXPECT(st.ok());
        if (!st.ok()) break;
        start_time = planner.GetNextPlanStartTime(start_time + tp::Milliseconds(200));
        double fastest = 0.0;
        for (size_t d = 0; d < kFakeDofs; ++d) {
          fastest = std::max(fastest, std::abs(planner.GetVelocities().front()[d]));
        }
        if (fastest > 0.3 * 0.25) break;
      }
      EXPECT(!planner.IsTrajectoryAtEnd());
      const std::vector<VectorXd> new_joints = {VectorXd::Constant(kFakeDofs, -1.0),
                                                VectorXd::Constant(kFakeDofs, 1.0)};
      std::vector<tp::Pose3d> new_poses;
      for (const auto& j : new_joints) new_poses.push_back(FakeFK(j));
      auto stop = planner.GetPathStopParameter(start_time);
      EXPECT(stop.ok());
      const double stop_parameter = stop.ok() ? *stop : 0.0;
      VectorXd v0(kFakeDofs);
      {
        const auto& t = planner.GetTime();
        const double ts = TimeToSec(start_time);
        size_t hi = 1;
        while (hi + 1 < t.size() && t[hi] < ts) ++hi;
        const double w = (ts - t[hi - 1]) / (t[hi] - t[hi - 1]);
        for (size_t d = 0; d < kFakeDofs; ++d) {
          v0[d] = planner.GetVelocities()[hi - 1][d] +
                  w * (planner.GetVelocities()[hi][d] - planner.GetVelocities()[hi - 1][d]);
        }
      }
      EXPECT(cpath->SwitchToWaypointPath(stop_parameter, new_poses, std::vector<VectorXd>{new_joints[0]}).code() ==
             tp::StatusCode::kInvalidArgument);
      const tp::Status sw = cpath->SwitchToWaypointPath(stop_parameter, new_poses, new_joints);
      EXPECT(sw.ok());
      if (!sw.ok()) fprintf(stderr, "Cartesian switch (mode %d): %s\n", mode, sw.message().c_str());
      EXPECT(cpath->GetState() == TimeablePath::State::kModifiedPath);
      EXPECT(cpath->SetInitialVelocity(ToSpan(v0)).ok());
      const std::vector<double> knots_after = cpath->GetKnots();
      const int ik_calls_before = ik_calls;
      int second_calls = 0;
      for (; !planner.IsTrajectoryAtEnd() && second_calls < 400; ++second_calls) {
        const tp::Status st = planner.Plan(start_time, tp::Milliseconds(750));
        EXPECT(st.ok());
        if (!st.ok()) {
          fprintf(stderr, "Cartesian Plan after switch %d (mode %d): %s\n", second_calls, mode,
                  st.message().c_str());
          break;
        }
        start_time = planner.GetNextPlanStartTime(start_time + tp::Milliseconds(200));
      }
      EXPECT(planner.IsTrajectoryAtEnd());
      EXPECT(mode == 0 || ik_calls > ik_calls_before);
      if (!planner.GetPositions().empty()) {
        for (size_t d = 0; d < kFakeDofs; ++d) {
          EXPECT(std::abs(planner.GetPositions().back()[d] - new_joints.back()[d]) < 1e-8);
          EXPECT(std::abs(planner.GetVelocities().back()[d]) < 1e-10);
        }
      }
      printf(",\"%s\": {\n\"first_calls\": %d, \"second_calls\": %d, \"stop_parameter\": %.17g,\n",
             mode == 0 ? "switch" : "switch_host_callbacks", first_calls, second_calls, stop_parameter);
      PrintArray("initial_velocity", std::vector<double>(v0.begin(), v0.end()));
      PrintArray("knots_after_switch", knots_after);
      PrintArray("time", planner.GetTime());
      PrintArray("positions", Flatten(planner.GetPositions()), false);
      printf("}\n");
    }
    std::vector<double> jflat, pflat;
    for (size_t w = 0; w < jw.size(); ++w) {
      jflat.insert(jflat.end(), jw[w].begin(), jw[w].end());
      for (int k = 0; k < 3; ++k) pflat.push_back(pw[w].translation()[k]);
      for (int k = 0; k < 4; ++k) pflat.push_back(pw[w].quaternion().coeffs()[k]);
    }
    printf(",");
    PrintArray("joint_waypoints", jflat);
    PrintArray("pose_waypoints", pflat, false);
    printf("},\n");
  }
  // ---- more waypoints than the path's engine was created for: the engine is re-created
  // (possibly at the same address) and the planner starts from an empty plan on it
  {
    constexpr size_t kN = 200;
    auto path = std::make_shared<TimeableJointSplinePath>(
        JointPathOptions().set_num_dofs(3).set_num_path_samples(kN));
    PathTimingTrajectory planner(PathTimingTrajectoryOptions()
                                     .SetTimeStep(tp::Milliseconds(4))
                                     .SetNumDofs(3)
                                     .SetNumPathSamples(kN));
    EXPECT(planner.SetPath(path).ok());
    VectorXd vmax = VectorXd::Constant(3, 1.0), amax = VectorXd::Constant(3, 2.0);
    EXPECT(path->SetMaxJointVelocity(ToSpan(vmax)).ok());
    EXPECT(path->SetMaxJointAcceleration(ToSpan(amax)).ok());
    std::vector<VectorXd> few = {VectorXd{0.0, 0.0, 0.0}, VectorXd{0.5, 0.2, -0.1}, VectorXd{0.1, 0.4, 0.3}};
    EXPECT(path->SetWaypoints(few).ok());
    EXPECT(planner.Plan(TimeFromSec(0.0), tp::Seconds(1000)).ok());
    const uint64_t generation = path->DeviceEngineGeneration();
    std::vector<VectorXd> many;
    for (int i = 0; i < 40; ++i) {
      many.push_back(VectorXd{0.1 * std::sin(0.7 * i), 0.05 * i, 0.1 * std::cos(0.3 * i)});
    }
    EXPECT(path->SetWaypoints(many).ok());
    EXPECT(path->DeviceEngineGeneration() == generation + 1);
    const tp::Status st = planner.Plan(TimeFromSec(0.0), tp::Seconds(1000));
    EXPECT(st.ok());
    if (!st.ok()) fprintf(stderr, "Plan on a re-created engine: %s\n", st.message().c_str());
    EXPECT(planner.IsTrajectoryAtEnd() && !planner.GetPositions().empty());
    const auto& last = planner.GetPositions().back();
    for (int d = 0; d < 3; ++d) EXPECT(std::abs(last[d] - many.back()[d]) <= 1e-9);
  }
  // ---- batched front-end
  {
    const int B = 5, W = 4, D = 3, N = 256;
    std::vector<double> wp(B * W * D), vm(B * D, 1.0), am(B * D, 2.0);
    for (size_t i = 0; i < wp.size(); ++i) wp[i] = std::sin(0.37 * static_cast<double>(i) + 0.1) * 2.0;
    std::vector<int32_t> nw = {4, 3, 2, 4, 1};
    PathTimingTrajectoryBatch batch(JointPathOptions().set_num_dofs(D),
                                    PathTimingTrajectoryOptions()
                                        .SetTimeStep(tp::Milliseconds(4))
                                        .SetNumDofs(D)
                                        .SetNumPathSamples(N),
                                    B, W, 1 << 20);
    PathTimingTrajectoryBatch::Result res;
    EXPECT(batch.Plan(B, nw.data(), wp.data(), vm.data(), am.data(), TimeFromSec(0.0), &res).ok());
    printf("\"batch\": {\n");
    PrintArray("waypoints", wp);
    PrintArray("duration", res.duration.to_vector());
    std::vector<double> off(res.offset.begin(), res.offset.end()), st(res.status.begin(), res.status.end());
    PrintArray("offset", off);
    PrintArray("status", st);
    PrintArray("positions", res.positions.to_vector(), false);
    printf("},\n");
    // the same batch over two logical shards of the device, positions + profile only:
    // bit-identical to the single-engine result, nothing else is brought back
    PathTimingTrajectoryBatch sharded(JointPathOptions().set_num_dofs(D),
                                      PathTimingTrajectoryOptions()
                                          .SetTimeStep(tp::Milliseconds(4))
                                          .SetNumDofs(D)
                                          .SetNumPathSamples(N),
                                      B, W, 1 << 20, std::vector<int32_t>{0, 0});
    EXPECT(sharded.num_shards() == 2);
    sharded.set_wanted(PathTimingTrajectoryBatch::kPositions | PathTimingTrajectoryBatch::kProfile);
    PathTimingTrajectoryBatch::Result res2;
    for (int rep = 0; rep < 2; ++rep)
      EXPECT(sharded.Plan(B, nw.data(), wp.data(), vm.data(), am.data(), TimeFromSec(0.0), &res2).ok());
    auto same = [](const auto& x, const auto& y) {  // bytes (a failed path's duration is NaN)
      return x.size() == y.size() &&
             std::memcmp(x.data(), y.data(), x.size() * sizeof(*x.data())) == 0;
    };
    EXPECT(res2.rows == res.rows);
    EXPECT(same(res2.duration, res.duration));
    EXPECT(same(res2.status, res.status));
    EXPECT(same(res2.offset, res.offset));
    EXPECT(same(res2.positions, res.positions));
    EXPECT(res2.velocities.empty() && res2.accelerations.empty() && res2.path_parameter.empty());
    EXPECT(res2.profile_time.size() == static_cast<size_t>(B) * N);
    for (int b = 0; b < B; ++b) {
      if (res.status[b] != 0) continue;
      const double* t = res2.profile_time.data() + static_cast<size_t>(b) * N;
      EXPECT(t[N - 1] - t[0] == res.duration[b]);
    }
    // the profile alone: no resampling, no rows
    sharded.set_wanted(PathTimingTrajectoryBatch::kProfile);
    EXPECT(sharded.Plan(B, nw.data(), wp.data(), vm.data(), am.data(), TimeFromSec(0.0), &res2).ok());
    EXPECT(res2.rows == 0 && res2.positions.empty());
    EXPECT(same(res2.duration, res.duration));
  }
  // ---- RestToRestPlanningWorks with TimeSamplingMethod::kSkipSamplesCloserThanTimeStep
  // (the reference runs its PathTimingTrajectoryTest suite for both methods, :994-1002)
  {
    auto path = std::make_shared<TimeableJointSplinePath>(JointPathOptions()
                                                              .set_num_dofs(kNDof)
                                                              .set_num_path_samples(kNumSamples)
                                                              .set_delta_parameter(0.016));
    PathTimingTrajectory planner(
        PathTimingTrajectoryOptions()
            .SetTimeStep(tp::Milliseconds(4))
            .SetNumDofs(kNDof)
            .SetNumPathSamples(kNumSamples)
            .SetTimeSamplingMethod(
                PathTimingTrajectoryOptions::TimeSamplingMethod::kSkipSamplesCloserThanTimeStep));
    EXPECT(planner.SetPath(path).ok());
    std::vector<VectorXd> wps = {VectorXd{1, 2, 3}, VectorXd{-1, -2, -3}, VectorXd{1, 2, 3}};
    EXPECT(path->SetWaypoints(wps).ok());
    VectorXd vmax = VectorXd::Constant(kNDof, 1.0), amax = VectorXd::Constant(kNDof, 2.0);
    EXPECT(path->SetMaxJointVelocity(ToSpan(vmax)).ok());
    EXPECT(path->SetMaxJointAcceleration(ToSpan(amax)).ok());
    EXPECT(planner.Plan(TimeFromSec(0.0), tp::Seconds(100)).ok());
    EXPECT(planner.IsTrajectoryAtEnd());
    const auto& t = planner.GetTime();
    EXPECT(!t.empty() && t.size() <= kNumSamples + 1);
    if (!t.empty()) {
      EXPECT(planner.GetPositions().back() == wps.back());
      for (size_t d = 0; d < kNDof; ++d) EXPECT(planner.GetVelocities().back()[d] == 0.0);
      for (size_t i = 1; i < t.size(); ++i) EXPECT(t[i] - t[i - 1] >= 0.95 * 0.004 - 1e-15);
    }
    printf("\"rest_to_rest_skip\": {\n");
    PrintArray("time", t);
    PrintArray("positions", Flatten(planner.GetPositions()));
    PrintArray("velocities", Flatten(planner.GetVelocities()), false);
    printf("},\n");
    // a second Plan() inside the planned span trims the front and re-interpolates the
    // first sample (EraseTrajectoryBefore, .cc:545-569)
    const size_t m0 = planner.GetNumTimeSamples();
    EXPECT(planner.Plan(TimeFromSec(0.2), tp::Seconds(0.1)).ok());
    EXPECT(planner.GetNumTimeSamples() < m0 && planner.GetTime().front() == 0.2);
    EXPECT(planner.GetTime()[1] - planner.GetTime()[0] > 0.0);
  }
  // ---- GetPathStopParameterWorksInSimpleCase (path_timing_trajectory_test.cc:942-992),
  // with delta chosen so that the path fits one planning window
  {
    constexpr int kStopSamples = 500;
    auto path = std::make_shared<TimeableJointSplinePath>(JointPathOptions()
                                                              .set_num_dofs(kNDof)
                                                              .set_constraint_safety(1.0)
                                                              .set_num_path_samples(kStopSamples)
                                                              .set_delta_parameter(0.008));
    PathTimingTrajectory planner(PathTimingTrajectoryOptions()
                                     .SetTimeStep(tp::Milliseconds(1))
                                     .SetNumDofs(kNDof)
                                     .SetNumPathSamples(kStopSamples));
    // no plan yet: beginning of the path (.cc:239-242)
    EXPECT(planner.SetPath(path).ok());
    auto none = planner.GetPathStopParameter(TimeFromSec(0.0));
    EXPECT(none.ok() && *none == 0.0);
    EXPECT(path->SetWaypoints(std::vector<VectorXd>{VectorXd{0, 0, 0}, VectorXd{1, 1, 1}, VectorXd{2, 2, 2}}).ok());
    VectorXd vmax = VectorXd::Constant(kNDof, 1.0), amax = VectorXd::Constant(kNDof, 2.0);
    EXPECT(path->SetMaxJointVelocity(ToSpan(vmax)).ok());
    EXPECT(path->SetMaxJointAcceleration(ToSpan(amax)).ok());
    EXPECT(planner.Plan(TimeFromSec(0.0), tp::Seconds(100)).ok());
    // starting from the first sample we can stop (almost) immediately
    auto first = planner.GetPathStopParameter(TimeFromSec(0.0));
    EXPECT(first.ok());
    const auto& ps = planner.GetPathParameters();
    const long stop_index = std::upper_bound(ps.begin(), ps.end(), *first) - ps.begin();
    EXPECT(stop_index <= 3);
    // starting from the last sample the parameter is the final value
    auto last = planner.GetPathStopParameter(TimeFromSec(planner.GetTime().back()));
    EXPECT(last.ok() && *last == ps.back());
    EXPECT(std::upper_bound(ps.begin(), ps.end(), *last) == ps.end());
    // past the end: kInvalidArgument (.cc:244-248)
    EXPECT(planner.GetPathStopParameter(TimeFromSec(planner.GetTime().back() + 1.0)).status().code() ==
           tp::StatusCode::kInvalidArgument);
    const double mid_time = planner.GetTime()[planner.GetTime().size() / 3];
    auto mid = planner.GetPathStopParameter(TimeFromSec(mid_time));
    EXPECT(mid.ok());
    printf("\"stop_query\": {\"first\": %.17g, \"last\": %.17g, \"mid_time\": %.17g, \"mid\": %.17g},\n",
           *first, *last, mid_time, *mid);
  }
  // ---- RescaleTrajectoryBackwardToStop (rescale_to_stop_test.cc:29-84)
  {
    constexpr int kSampleCount = 200, kJointCount = 4;
    constexpr double kTimeStepSec = 8e-3, kMaxAcceleration = 2.0;
    printf("\"stop\": {\n");
    int idx = 0;
    for (double velocity : {-1.0, 1.0}) {
      std::vector<double> times(kSampleCount);
      std::vector<VectorXd> pos(kSampleCount), vel(kSampleCount), acc(kSampleCount);
      for (int i = 0; i < kSampleCount; ++i) {
        times[i] = i * kTimeStepSec;
        acc[i] = VectorXd::Zero(kJointCount);
        vel[i] = VectorXd::Constant(kJointCount, velocity);
        pos[i] = VectorXd::Constant(kJointCount, velocity * times[i]);
      }
      auto result = RescaleTrajectoryBackwardToStop(VectorXd::Constant(kJointCount, kMaxAcceleration),
                                                    times, pos, vel, acc);
      EXPECT(result.ok());
      const SampledTrajectory& r = *result;
      const double ramp_up_time_sec = std::fabs(velocity) / kMaxAcceleration;
      const double stopping_distance = velocity * velocity / (2.0 * kMaxAcceleration);
      EXPECT(!r.positions.empty() && r.times.size() == r.positions.size() &&
             r.times.size() == r.velocities.size() && r.times.size() == r.accelerations.size());
      if (!r.times.empty()) {
        EXPECT(std::fabs((r.times.back() - r.times.front()) - ramp_up_time_sec) <= kTimeStepSec);
        for (int j = 0; j < kJointCount; ++j) {
          EXPECT(r.velocities.back()[j] == 0.0 && r.accelerations.back()[j] == 0.0);
          const double moved = r.positions.back()[j] - r.positions.front()[j];
          EXPECT(std::fabs(moved - stopping_distance * (velocity < 0 ? -1 : 1)) <=
                 std::fabs(velocity * kTimeStepSec));
        }
      }
      printf("\"case%d\": {\n", idx++);
      PrintArray("times", r.times);
      PrintArray("velocities", Flatten(r.velocities));
      PrintArray("accelerations", Flatten(r.accelerations), false);
      printf("},\n");
      // error behaviour: AreInputsValidForSampledTrajectory's statuses come back
      std::vector<double> bad = times;
      bad[5] = bad[4];
      EXPECT(RescaleTrajectoryBackwardToStop(VectorXd::Constant(kJointCount, 2.0), bad, pos, vel, acc)
                 .status().code() == tp::StatusCode::kInvalidArgument);
      vel.back() = VectorXd::Zero(kJointCount);  // already stopped: empty result
      auto empty = RescaleTrajectoryBackwardToStop(VectorXd::Constant(kJointCount, 2.0), times, pos, vel, acc);
      EXPECT(empty.ok() && (*empty).times.empty());
    }
    printf("\"cases\": %d},\n", idx);
  }
  // ---- TrajectoryBuffer (trajectory_buffer_test.cc:99-716)
  {
    constexpr int kNumJoints = 9;
    constexpr double kSampleTimeSec = 8e-3;
    auto test_trajectory = [&](double offset, int count) {  // :80-97
      SampledTrajectory tr;
      for (int i = 0; i < count; ++i) {
        tr.times.push_back(i * kSampleTimeSec + offset);
        tr.positions.push_back(VectorXd::Constant(kNumJoints, i));
        tr.velocities.push_back(VectorXd::Constant(kNumJoints, 10 * i));
        tr.accelerations.push_back(VectorXd::Constant(kNumJoints, 100 * i));
      }
      return tr;
    };
    auto quadratic_trajectory = [&](int count, double dt) {  // :60-77
      SampledTrajectory tr;
      const double final_time = (count - 1) * dt;
      for (int i = 0; i < count; ++i) {
        const double t = dt * i, trel = t - final_time;
        tr.times.push_back(t);
        tr.positions.push_back(VectorXd::Constant(kNumJoints, 0.5 * trel * trel));
        tr.velocities.push_back(VectorXd::Constant(kNumJoints, trel));
        tr.accelerations.push_back(VectorXd::Constant(kNumJoints, 1.0));
      }
      return tr;
    };
    auto insert = [](TrajectoryBuffer& b, const SampledTrajectory& tr) {
      return b.InsertSegment(tr.times, tr.positions, tr.velocities, tr.accelerations);
    };
    printf("\"tbuf\": {\n");
    // CreateChecksOptions (:99-103)
    EXPECT(TrajectoryBuffer::Create({-1}).status().code() == tp::StatusCode::kFailedPrecondition);
    // InsertSegmentWorks (:105-195)
    {
      auto made = TrajectoryBuffer::Create();
      EXPECT(made.ok());
      TrajectoryBuffer& buffer = **made;
      EXPECT(buffer.GetNumSamples() == 0 && buffer.GetStartTime() == TimeFromSec(0) &&
             buffer.GetSequenceNumber() == 0);
      EXPECT(buffer.InsertSegment({}, {}, {}, {}).ok());
      EXPECT(buffer.GetStartTime() == TimeFromSec(0) && buffer.GetEndTime() == TimeFromSec(0));
      EXPECT(buffer.GetSequenceNumber() == 1 && buffer.GetNumSamples() == 0);
      buffer.Clear();
      SampledTrajectory tr = test_trajectory(1.0, 10);
      EXPECT(insert(buffer, tr).ok());
      EXPECT(buffer.GetStartTime() == TimeFromSec(tr.times.front()));
      EXPECT(buffer.GetEndTime() == TimeFromSec(tr.times.back()));
      EXPECT(buffer.GetSequenceNumber() == 0 && buffer.GetNumSamples() == tr.positions.size());
      for (size_t i = 0; i < tr.times.size(); ++i) {
        EXPECT(buffer.GetTimes()[i] == tr.times[i] && buffer.GetPositions()[i] == tr.positions[i] &&
               buffer.GetVelocities()[i] == tr.velocities[i] &&
               buffer.GetAccelerations()[i] == tr.accelerations[i]);
      }
      tr = test_trajectory(1.0 + 3 * kSampleTimeSec, 5);
      EXPECT(insert(buffer, tr).ok());
      EXPECT(buffer.GetStartTime() == TimeFromSec(1) &&
             buffer.GetEndTime() == TimeFromSec(tr.times.back()));
      EXPECT(buffer.GetSequenceNumber() == 1 && buffer.GetNumSamples() == 3 + tr.positions.size());
      const double expect_pos[8] = {0, 1, 2, 0, 1, 2, 3, 4};
      for (int i = 0; i < 8 && buffer.GetNumSamples() == 8; ++i) {
        EXPECT(buffer.GetTimes()[i] == 1.0 + i * kSampleTimeSec);
        EXPECT(buffer.GetPositions()[i] == VectorXd::Constant(kNumJoints, expect_pos[i]));
        EXPECT(buffer.GetVelocities()[i] == VectorXd::Constant(kNumJoints, 10 * expect_pos[i]));
        EXPECT(buffer.GetAccelerations()[i] == VectorXd::Constant(kNumJoints, 100 * expect_pos[i]));
      }
      // InsertSegmentFailsForInvalidArguments (:290-323)
      std::vector<double> short_times(5);
      std::vector<VectorXd> short_rows(5);
      tr = test_trajectory(1.0, 10);
      EXPECT(buffer.InsertSegment(short_times, tr.positions, tr.velocities, tr.accelerations).code() ==
             tp::StatusCode::kInvalidArgument);
      EXPECT(buffer.InsertSegment(tr.times, short_rows, tr.velocities, tr.accelerations).code() ==
             tp::StatusCode::kInvalidArgument);
      EXPECT(buffer.InsertSegment(tr.times, tr.positions, short_rows, tr.accelerations).code() ==
             tp::StatusCode::kInvalidArgument);
      EXPECT(buffer.InsertSegment(tr.times, tr.positions, tr.velocities, short_rows).code() ==
             tp::StatusCode::kInvalidArgument);
      tr.times.front() = TimeToSec(buffer.GetEndTime());
      for (size_t i = 1; i < tr.times.size(); ++i) tr.times[i] = tr.times[i - 1] + kSampleTimeSec;
      EXPECT(insert(buffer, tr).ok());
    }
    // AppendSampleWorks (:197-228)
    {
      auto made = TrajectoryBuffer::Create();
      TrajectoryBuffer& buffer = **made;
      auto c = [&](double x) { return VectorXd::Constant(kNumJoints, x); };
      EXPECT(buffer.AppendSample(1.0, c(1), c(2), c(3)).ok());
      EXPECT(buffer.AppendSample(1.0, c(1), c(2), c(3)).code() == tp::StatusCode::kInvalidArgument);
      EXPECT(buffer.AppendSample(-1.0, c(1), c(2), c(3)).code() == tp::StatusCode::kInvalidArgument);
      EXPECT(buffer.AppendSample(1.1, c(1.1), c(2.1), c(3.1)).ok());
      EXPECT(buffer.GetNumSamples() == 2);
      EXPECT(buffer.GetPositions()[1] == c(1.1) && buffer.GetVelocities()[1] == c(2.1) &&
             buffer.GetAccelerations()[0] == c(3));
    }
    // DiscardWorks (:341-471), GetPositionsUpToTime (:473-501), GetPositionAtTime (:503-528)
    {
      auto made = TrajectoryBuffer::Create();
      TrajectoryBuffer& buffer = **made;
      buffer.DiscardSegmentBefore(TimeFromSec(10));
      EXPECT(buffer.GetSequenceNumber() == 0 && buffer.GetNumSamples() == 0);
      const SampledTrajectory tr = test_trajectory(1.0, 10);
      EXPECT(insert(buffer, tr).ok());
      buffer.DiscardSegmentBefore(buffer.GetStartTime());
      EXPECT(buffer.GetNumSamples() == 10);
      buffer.DiscardSegmentBefore(TimeFromSec(tr.times[4]));
      EXPECT(buffer.GetStartTime() == TimeFromSec(tr.times[4]) && buffer.GetTimes().front() == tr.times[4]);
      EXPECT(buffer.GetEndTime() == TimeFromSec(tr.times.back()) && buffer.GetNumSamples() == 6);
      buffer.DiscardSegmentBefore(TimeFromSec(tr.times.back() + 1.0 / 1e9));
      EXPECT(buffer.GetStartTime() == TimeFromSec(0) && buffer.GetEndTime() == TimeFromSec(0) &&
             buffer.GetNumSamples() == 0);
      buffer.Clear();
      EXPECT(insert(buffer, tr).ok());
      const double between = 0.5 * (tr.times[5] + tr.times[6]);
      buffer.DiscardSegmentBefore(between);
      EXPECT(std::fabs(buffer.GetTimes().front() - between) <= 1e-10);
      PrintArray("discard_between_times", std::vector<double>(buffer.GetTimes().begin(), buffer.GetTimes().end()));
      PrintArray("discard_between_first_position",
                 std::vector<double>(buffer.GetPositions().front().begin(), buffer.GetPositions().front().end()));
      buffer.Clear();
      const SampledTrajectory five = test_trajectory(1.0, 5);
      EXPECT(insert(buffer, five).ok());
      EXPECT(buffer.GetPositionsUpToTime(TimeFromSec(10)).size() == 0);
      EXPECT(buffer.GetPositionsUpToTime(TimeFromSec(-1)).size() == 0);
      EXPECT(buffer.GetPositionsUpToTime(TimeFromSec(five.times[2])).size() == 2);
      EXPECT(buffer.GetPositionsUpToTime(TimeFromSec(std::nextafter(five.times[2], five.times[1]))).size() == 1);
      EXPECT(buffer.GetPositionsUpToTime(TimeFromSec(std::nextafter(five.times[2], five.times[3]))).size() == 2);
      for (size_t i = 0; i < five.times.size(); ++i) {
        auto p = buffer.GetPositionAtTime(TimeFromSec(five.times[i]));
        auto v = buffer.GetVelocityAtTime(TimeFromSec(five.times[i]));
        EXPECT(p.ok() && v.ok());
        for (int j = 0; j < kNumJoints && p.ok() && v.ok(); ++j) {
          EXPECT(std::fabs((*p)[j] - five.positions[i][j]) <= 1e-10);
          EXPECT(std::fabs((*v)[j] - five.velocities[i][j]) <= 1e-8);
        }
      }
      EXPECT(buffer.GetPositionAtTime(100.0).status().code() == tp::StatusCode::kOutOfRange);
      buffer.AddOffsetToTimestamps(0.5);
      EXPECT(buffer.GetTimes().front() == five.times.front() + 0.5);
    }
    // StopAtIndex / StopBeforeTime (:530-716)
    {
      constexpr double kDt = 1e-3;
      constexpr int kNumSamples = 1001, kStopIndex = kNumSamples / 2;
      const SampledTrajectory tr = quadratic_trajectory(kNumSamples, kDt);
      const VectorXd amax = VectorXd::Constant(kNumJoints, 5.0);
      auto made = TrajectoryBuffer::Create();
      TrajectoryBuffer& buffer = **made;
      EXPECT(buffer.StopAtIndex(-1, amax, kDt).code() == tp::StatusCode::kOutOfRange);
      EXPECT(buffer.StopAtIndex(0, amax, kDt).code() == tp::StatusCode::kOutOfRange);
      EXPECT(buffer.StopBeforeTime(1.0, amax, kDt).ok());
      EXPECT(insert(buffer, tr).ok());
      EXPECT(buffer.StopAtIndex(kNumSamples, amax, kDt).code() == tp::StatusCode::kOutOfRange);
      EXPECT(buffer.StopAtIndex(kStopIndex, amax, -0.1).code() == tp::StatusCode::kInvalidArgument);
      EXPECT(buffer.StopAtIndex(kStopIndex, amax, kDt).ok());
      EXPECT(static_cast<int>(buffer.GetNumSamples()) >= kStopIndex + 1);
      for (int j = 0; j < kNumJoints; ++j) EXPECT(std::fabs(buffer.GetVelocities().back()[j]) <= 1e-12);
      for (const auto& acc : buffer.GetAccelerations())
        for (int j = 0; j < kNumJoints; ++j) EXPECT(std::fabs(acc[j]) <= amax[j] + 1e-8);
      printf("\"stop_num_samples\": %d,\n", static_cast<int>(buffer.GetNumSamples()));
      PrintArray("stop_times", std::vector<double>(buffer.GetTimes().begin(), buffer.GetTimes().end()));
      std::vector<VectorXd> vel(buffer.GetVelocities().begin(), buffer.GetVelocities().end());
      PrintArray("stop_velocities", Flatten(vel));
      // infeasible (:593-632)
      buffer.Clear();
      EXPECT(insert(buffer, tr).ok());
      const VectorXd weak = VectorXd::Constant(kNumJoints, 1.0);
      EXPECT(buffer.StopAtIndex(2, weak, kDt).code() == tp::StatusCode::kNotFound);
      EXPECT(buffer.StopBeforeTime(kDt * 2, weak, kDt).code() == tp::StatusCode::kNotFound);
      EXPECT(buffer.StopBeforeTime(kDt * 3.1415, weak, kDt).code() == tp::StatusCode::kNotFound);
      EXPECT(buffer.GetNumSamples() == static_cast<size_t>(kNumSamples));
      // halfway between two samples (:673-685)
      const double halfway = 0.5 * (buffer.GetTimes()[kStopIndex] + buffer.GetTimes()[kStopIndex + 1]);
      EXPECT(buffer.StopBeforeTime(halfway, amax, kDt).ok());
      EXPECT(static_cast<int>(buffer.GetNumSamples()) >= kStopIndex + 1);
      printf("\"stop_halfway_num_samples\": %d,\n", static_cast<int>(buffer.GetNumSamples()));
      // growth beyond the initial capacity keeps rows and sequence number
      auto big = TrajectoryBuffer::Create();
      TrajectoryBuffer& grow = **big;
      EXPECT(insert(grow, quadratic_trajectory(3000, kDt)).ok());
      SampledTrajectory more = quadratic_trajectory(9000, kDt);
      for (auto& t : more.times) t += 2.0;
      EXPECT(insert(grow, more).ok());
      EXPECT(grow.GetNumSamples() == 2000 + 9000 && grow.GetSequenceNumber() == 1);
      EXPECT(grow.GetTimes()[1999] == 1999 * kDt && grow.GetTimes()[2000] == more.times[0]);
    }
    printf("\"done\": 1},\n");
  }
  printf("\"failures\": %d\n}\n", g_fail);
  return g_fail ? 1 : 0;
}
