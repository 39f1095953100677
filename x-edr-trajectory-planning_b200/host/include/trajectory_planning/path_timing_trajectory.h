// Host-side mirror of PathTimingTrajectory
// (trajectory_planning/path_timing_trajectory.h:33-186, .cc:206-902).
//
// Plan() runs on the device.  For a TimeableJointSplinePath and a
// TimeableCartesianSplinePath the whole PathTimingTrajectory::Plan state machine lives
// there (tpb_planner_plan:
// HandleTimeArguments, the moving-horizon loop over ComputeTimingProfile .cc:307-475
// with splicing of the path samples, the "already planned enough" erase, both time
// sampling methods); this class keeps the options, maps the per-path status back to
// the reference's Status codes and downloads the trajectory into the
// TrajectoryPlanner vectors.  Re-planning with a shifted start time (Plan() called
// again while the previous trajectory is being executed) works as in the reference.
//
// A Cartesian path's user PathIKFunc / JacobianFunc are called back from inside
// tpb_planner_plan, once per planning window (tpb_planner_set_host_kinematics).
// GetPathStopParameter runs on the device.
#ifndef TRAJECTORY_PLANNING_PATH_TIMING_TRAJECTORY_H_
#define TRAJECTORY_PLANNING_PATH_TIMING_TRAJECTORY_H_

#include <algorithm>
#include <cmath>
#include <limits>
#include <memory>
#include <string>
#include <vector>

#include "trajectory_planning/timeable_path_joint_spline.h"
#include "trajectory_planning/trajectory_planner.h"

namespace trajectory_planning {

class PathTimingTrajectoryOptions : public TrajectoryPlannerOptions<PathTimingTrajectoryOptions> {
 public:
  enum class TimeSamplingMethod { kUniformlyInTime, kSkipSamplesCloserThanTimeStep };
  size_t GetNumPathSamples() const { return num_path_samples_; }
  PathTimingTrajectoryOptions& SetNumPathSamples(size_t n) {
    num_path_samples_ = n;
    return *this;
  }
  double GetMaxInitialVelocityError() const { return max_initial_velocity_error_; }
  PathTimingTrajectoryOptions& SetMaxInitialVelocityError(double e) {
    max_initial_velocity_error_ = e;
    return *this;
  }
  PathTimingTrajectoryOptions& SetMaxPlanningLoops(int n) {
    max_planning_iterations_ = n;
    return *this;
  }
  int GetMaxPlanningIterations() const { return max_planning_iterations_; }
  PathTimingTrajectoryOptions& SetTimeSamplingMethod(TimeSamplingMethod m) {
    time_sampling_method_ = m;
    return *this;
  }
  TimeSamplingMethod GetTimeSamplingMethod() const { return time_sampling_method_; }

 private:
  size_t num_path_samples_ = 1000;            // path_timing_trajectory.h:71
  double max_initial_velocity_error_ = 1e-2;  // :72
  int max_planning_iterations_ = 200;         // :73
  TimeSamplingMethod time_sampling_method_ = TimeSamplingMethod::kUniformlyInTime;
};

class PathTimingTrajectory : public TrajectoryPlanner {
 public:
  explicit PathTimingTrajectory(const PathTimingTrajectoryOptions& options)
      : options_(options),
        time_step_sec_(TimeToSec(tp::FromUnixNanos(options.GetTimeStep().ns))) {
    Reset();
  }

  // SetPath (.cc:844-864)
  tp::Status SetPath(std::shared_ptr<TimeablePath> path) override {
    if (path == nullptr) return tp::InvalidArgumentError("Path is nullptr.");
    if (path->NumDofs() != options_.GetNumDofs()) {
      return tp::InvalidArgumentError("Path has " + std::to_string(path->NumDofs()) +
                                      " dofs but planner has " + std::to_string(options_.GetNumDofs()));
    }
    if (path->NumPathSamples() != options_.GetNumPathSamples()) {
      return tp::InvalidArgumentError("Path has " + std::to_string(path->NumPathSamples()) +
                                      " samples but planner has " +
                                      std::to_string(options_.GetNumPathSamples()));
    }
    path_ = path;
    return tp::OkStatus();
  }

  // Plan (.cc:579-684)
  tp::Status Plan(tp::Time start, tp::Duration time_horizon) override {
    if (path_ == nullptr) return {tp::StatusCode::kFailedPrecondition, "No path set."};
    if (!path_->SupportsDevicePlanner()) {
      return tp::UnimplementedError(
          "only device-backed paths (TimeableJointSplinePath, TimeableCartesianSplinePath) are timed");
    }
    return PlanOnDevice(start, time_horizon);
  }

  // Extension: capacity (in path samples) of the spliced *_at_path_samples_ arrays on
  // the device; 0 = 64 planning windows.
  void SetPathSampleCapacity(size_t samples) { path_sample_capacity_ = samples; }

 private:
  // Plan (.cc:579-684) through tpb_planner_plan.
  tp::Status PlanOnDevice(tp::Time start, tp::Duration time_horizon) {
    const auto state = path_->GetState();
    if (state == TimeablePath::State::kNoPath) return tp::FailedPreconditionError("No path set");
    if (auto st = path_->BeforeDevicePlan(); !st.ok()) return st;
    tpb_engine* e = path_->DeviceEngine();
    if (e == nullptr) return tp::InternalError("path has no device engine");
    const bool skipping = options_.GetTimeSamplingMethod() ==
                          PathTimingTrajectoryOptions::TimeSamplingMethod::kSkipSamplesCloserThanTimeStep;
    int32_t rc = TPB_OK;
    if (state == TimeablePath::State::kNewPath) {
      const tp::VectorXd& v0 = path_->GetInitialVelocity();
      const bool same_engine =
          device_planner_ == e && device_planner_generation_ == path_->DeviceEngineGeneration();
      if (initial_plan_ && same_engine) {
        rc = tpb_planner_new_path(e, v0.data(), 0);  // the plan so far stays (no Reset())
      } else {
        // a new engine (more waypoints than the old one held) starts from an empty plan
        if (initial_plan_) initial_plan_ = false;
        const size_t cap = path_sample_capacity_ ? path_sample_capacity_ : 64 * options_.GetNumPathSamples();
        rc = tpb_planner_reset(e, static_cast<int32_t>(cap), time_step_sec_, skipping ? 1 : 0,
                               options_.GetMaxPlanningIterations(),
                               options_.GetMaxInitialVelocityError(), v0.data(), 0);
      }
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e));
      device_planner_ = e;
      device_planner_generation_ = path_->DeviceEngineGeneration();
    } else if (state == TimeablePath::State::kModifiedPath) {
      // SwitchToWaypointPath marked the path on the device; the initial velocity set
      // since (SetInitialVelocity) goes to the planner now (.cc:352-385)
      if (device_planner_ != e || device_planner_generation_ != path_->DeviceEngineGeneration()) {
        return tp::FailedPreconditionError("the path was modified, but never planned by this planner");
      }
      rc = tpb_planner_set_initial_velocity(e, path_->GetInitialVelocity().data(), 0);
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e));
    } else if (device_planner_ != e) {
      return tp::FailedPreconditionError("the path was sampled, but not by this planner");
    }
    int64_t total = 0;
    rc = tpb_planner_plan(e, start.ns, time_horizon.ns, &total);
    if (rc != TPB_OK) {
      // a failing PathIKFunc / JacobianFunc ends Plan with its own Status (.cc:518-525)
      if (auto st = path_->TakeCallbackStatus(); !st.ok()) return st;
      return tp::FromCode(rc, tpb_last_error(e));
    }
    // the planner's state after the call
    int32_t status = TPB_PATH_NOT_RUN, loops = 0, to_end = 0, reached = 0, rows_mode = 0;
    int64_t start_ns = 0, end_ns = 0, decel_ns = 0;
    tpb_get(e, TPB_ARR_PLAN_STATUS, &status, sizeof(status), 0);
    tpb_get(e, TPB_ARR_PLAN_LOOPS, &loops, sizeof(loops), 0);
    tpb_get(e, TPB_ARR_PLAN_PLANNED_TO_END, &to_end, sizeof(to_end), 0);
    tpb_get(e, TPB_ARR_PLAN_TARGET_REACHED, &reached, sizeof(reached), 0);
    tpb_get(e, TPB_ARR_PLAN_ERASED, &rows_mode, sizeof(rows_mode), 0);
    tpb_get(e, TPB_ARR_PLAN_START_TIME, &start_ns, sizeof(start_ns), 0);
    tpb_get(e, TPB_ARR_PLAN_END_TIME, &end_ns, sizeof(end_ns), 0);
    tpb_get(e, TPB_ARR_PLAN_FINAL_DECEL_START, &decel_ns, sizeof(decel_ns), 0);
    start_time_ = tp::FromUnixNanos(start_ns);
    end_time_ = tp::FromUnixNanos(end_ns);
    final_decel_start_ = tp::FromUnixNanos(decel_ns);
    planned_to_end_ = to_end != 0;
    target_reached_ = reached != 0;
    last_plan_windows_ = loops;
    if (auto st = path_->AfterDevicePlan(loops > 0); !st.ok()) return st;
    if (auto st = DownloadTrajectory(e, total); !st.ok()) return st;
    if (status != TPB_PATH_OK) return PlanStatus(status);
    if (rows_mode == 0) initial_plan_ = true;
    return tp::OkStatus();
  }

  // the reference's Status for a per-path outcome of tpb_planner_plan
  static tp::Status PlanStatus(int32_t status) {
    switch (status) {
      case TPB_PATH_OK: return tp::OkStatus();
      case TPB_PATH_FAIL_SETUP: return tp::InternalError("Error setting up optimization problem");
      case TPB_PATH_FAIL_DEADLINE:
        return {tp::StatusCode::kDeadlineExceeded, "Reached maximum number of planning loops"};
      case TPB_PATH_FAIL_START_RANGE:
        return {tp::StatusCode::kOutOfRange, "start outside of the previous plan"};
      case TPB_PATH_FAIL_START_ARG:
        return tp::InvalidArgumentError("start time / duration not valid for the previous plan");
      case TPB_PATH_FAIL_START_END: return tp::InternalError("time >= end_time_");
      case TPB_PATH_FAIL_NO_SAMPLES: return tp::FailedPreconditionError("No samples yet.");
      case TPB_PATH_FAIL_INITIAL_VELOCITY:
        return tp::InvalidArgumentError(
            "Could not satisfy initial velocity (probably not parallel to initial tangent)");
      case TPB_PATH_FAIL_CAPACITY:
        return {tp::StatusCode::kResourceExhausted, "path sample capacity exceeded (SetPathSampleCapacity)"};
      default: return tp::InternalError("Error optimizing path parameter");
    }
  }

  // the packed trajectory rows of the (single) path into the TrajectoryPlanner vectors
  tp::Status DownloadTrajectory(tpb_engine* e, int64_t total) {
    const size_t M = static_cast<size_t>(total), D = options_.GetNumDofs();
    time_.resize(M);
    path_parameter_.resize(M);
    path_parameter_derivative_.resize(M);
    second_path_parameter_derivative_.resize(M);
    positions_.resize(M);
    velocities_.resize(M);
    accelerations_.resize(M);
    if (M == 0) return tp::OkStatus();
    std::vector<double> buf(M * D);
    tpb_get(e, TPB_ARR_R_TIME, time_.data(), M * sizeof(double), 0);
    tpb_get(e, TPB_ARR_R_S, path_parameter_.data(), M * sizeof(double), 0);
    tpb_get(e, TPB_ARR_R_SD, path_parameter_derivative_.data(), M * sizeof(double), 0);
    tpb_get(e, TPB_ARR_R_SDD, second_path_parameter_derivative_.data(), M * sizeof(double), 0);
    std::vector<tp::VectorXd>* dst[3] = {&positions_, &velocities_, &accelerations_};
    const tpb_array what[3] = {TPB_ARR_R_POS, TPB_ARR_R_VEL, TPB_ARR_R_ACC};
    for (int k = 0; k < 3; ++k) {
      const int32_t rc = tpb_get(e, what[k], buf.data(), buf.size() * sizeof(double), 0);
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e));
      for (size_t i = 0; i < M; ++i) (*dst[k])[i] = tp::VectorXd(&buf[i * D], D);
    }
    return tp::OkStatus();
  }

 public:
  // Extension: planning windows computed by the last Plan() (device planner).
  int LastPlanWindows() const { return last_plan_windows_; }
  tp::Time GetFinalDecelStart() const { return final_decel_start_; }
  tp::Time GetNextPlanStartTime(tp::Time target_time) {
    return std::min(end_time_, std::max(target_time, start_time_),
                    [](tp::Time a, tp::Time b) { return a < b; });
  }
  void SetProfileDebugVerbosity(int) {}
  // GetPathStopParameter (.cc:235-287): the path parameter at which the path can be
  // stopped when braking from `time` on; ComputeFastestStop (.cc:75-172) runs on
  // the device over the resampled trajectory (tpb_path_stop_parameter).
  tp::StatusOr<double> GetPathStopParameter(tp::Time time) const {
    const double double_time = TimeToSec(time);
    if (!initial_plan_) return 0.0;  // no plan yet: beginning of the path
    const auto time_it = std::lower_bound(time_.begin(), time_.end(), double_time);
    if (time_it == time_.end()) {
      return tp::InvalidArgumentError("Time " + std::to_string(double_time) +
                                      " not in timed path range");
    }
    if (time_it == std::prev(time_.end())) return path_parameter_.back();
    tpb_engine* e = path_ ? path_->DeviceEngine() : nullptr;
    if (e == nullptr) return tp::FailedPreconditionError("path has no device engine");
    double parameter = 0.0;
    int32_t status = TPB_PATH_NOT_RUN;
    const int32_t rc = tpb_path_stop_parameter(e, &double_time, &parameter, nullptr, nullptr, &status, 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e));
    if (status != TPB_PATH_OK) {
      return tp::InvalidArgumentError("Time " + std::to_string(double_time) +
                                      " not in timed path range");
    }
    return parameter;
  }
  const PathTimingTrajectoryOptions& GetOptions() const { return options_; }

 protected:
  void ResetDerived() override {
    initial_plan_ = false;
    path_time_start_ = path_start_ = path_start_velocity_ = path_start_acceleration_ = 0.0;
    path_horizon_ = 0.0;
    planned_to_end_ = true;
    final_decel_start_ = TimeFromSec(0.0);
    time_at_path_samples_.clear();
    device_planner_ = nullptr;
  }

 private:
  const PathTimingTrajectoryOptions options_;
  const double time_step_sec_;
  tp::Time final_decel_start_;
  double path_horizon_ = 0, path_time_start_ = 0, path_start_ = 0, path_start_velocity_ = 0,
         path_start_acceleration_ = 0;
  bool initial_plan_ = false, planned_to_end_ = false;
  std::vector<double> time_at_path_samples_;
  size_t path_sample_capacity_ = 0;
  tpb_engine* device_planner_ = nullptr;  // the engine that holds this planner's state
  uint64_t device_planner_generation_ = 0;
  int last_plan_windows_ = 0;
};

}  // namespace trajectory_planning
#endif  // TRAJECTORY_PLANNING_PATH_TIMING_TRAJECTORY_H_
