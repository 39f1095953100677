// Host-side mirror of TimeableJointSplinePath
// (trajectory_planning/timeable_path_joint_spline.h:28-109, .cc:38-349): same
// public methods and state machine; spline fitting, path sampling, constraint
// rows and the limit curve run on the GPU (tpb_set_joint_waypoints /
// tpb_sample_paths).  One object = a batch of one path on its own engine.
// SwitchToWaypointPath (online replanning) runs on the device too
// (tpb_switch_to_waypoint_path).
#ifndef TRAJECTORY_PLANNING_TIMEABLE_PATH_JOINT_SPLINE_H_
#define TRAJECTORY_PLANNING_TIMEABLE_PATH_JOINT_SPLINE_H_

#include <algorithm>
#include <cmath>
#include <limits>
#include <string>
#include <vector>

#include "trajectory_planning/timeable_path.h"

namespace trajectory_planning {

class JointPathOptions : public PathOptions<JointPathOptions> {
 public:
  JointPathOptions() = default;
};

class TimeableJointSplinePath : public TimeablePath {
 public:
  explicit TimeableJointSplinePath(const JointPathOptions& options, int device = 0)
      : options_(options), num_constraints_(options.num_dofs() * 2), device_(device) {
    // CHECK(options_.num_path_samples() >= 3), timeable_path_joint_spline.cc:45
    if (options_.num_path_samples() < 3) std::abort();
    const size_t N = options_.num_path_samples(), D = options_.num_dofs();
    path_position_.assign(N, tp::VectorXd(D));
    first_path_derivative_.assign(N, tp::VectorXd(D));
    second_path_derivative_.assign(N, tp::VectorXd(D));
    max_joint_velocity_ = tp::VectorXd(D);
    max_joint_acceleration_ = tp::VectorXd(D);
    initial_velocity_ = tp::VectorXd(D);
    Reset();
  }
  ~TimeableJointSplinePath() override { if (e_) tpb_destroy(e_); }
  TimeableJointSplinePath(const TimeableJointSplinePath&) = delete;
  TimeableJointSplinePath& operator=(const TimeableJointSplinePath&) = delete;

  // SetWaypoints (.cc:200-207) + FitSplineToWaypoints (.cc:252-292)
  tp::Status SetWaypoints(tp::Span<const tp::VectorXd> waypoints) {
    for (const auto& wp : waypoints) {
      if (wp.size() != options_.num_dofs()) {
        return tp::InvalidArgumentError("Dimension error, need " + std::to_string(options_.num_dofs()) +
                                        " joint values but waypoint has " + std::to_string(wp.size()));
      }
    }
    if (waypoints.empty()) return tp::InvalidArgumentError("no waypoints");
    waypoints_.assign(waypoints.begin(), waypoints.end());
    path_state_ = State::kNewPath;
    const int W = static_cast<int>(waypoints_.size()), D = static_cast<int>(options_.num_dofs());
    if (!e_ || W > wcap_) {
      if (e_) tpb_destroy(e_);
      e_ = nullptr;
      tpb_config cfg{};
      cfg.device = device_;
      cfg.max_paths = 1;
      cfg.num_dofs = D;
      cfg.max_waypoints = std::max(W, 16);
      cfg.num_samples = static_cast<int32_t>(options_.num_path_samples());
      cfg.max_time_samples_total = int64_t{1} << 20;
      const int32_t rc = tpb_create(&cfg, &e_);
      if (rc != TPB_OK) {
        e_ = nullptr;
        return tp::FromCode(rc, "tpb_create failed (no CUDA device; there is no CPU fallback)");
      }
      wcap_ = cfg.max_waypoints;
      ++engine_generation_;
    }
    std::vector<double> flat(static_cast<size_t>(wcap_) * D, 0.0);
    for (int w = 0; w < W; ++w) std::copy(waypoints_[w].begin(), waypoints_[w].end(), flat.begin() + w * D);
    const int32_t nw = W;
    tpb_set_path_options(e_, options_.constraint_safety(), options_.rounding());
    int32_t rc = tpb_set_joint_waypoints(e_, 1, &nw, flat.data(), 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    const double delta = options_.delta_parameter();
    rc = tpb_set_sampling(e_, &delta, 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    constraints_valid_ = false;
    return DownloadSpline();
  }

  // knots_ and the spline's control points from the device (extension: after a switch
  // GetControlPoints() is the whole spline's, where the reference keeps the appended
  // control polygon only)
  tp::Status DownloadSpline() {
    const int D = static_cast<int>(options_.num_dofs());
    int32_t n = 0;
    size_t knot_bytes = 0, ctrl_bytes = 0;
    void* unused = nullptr;
    int32_t rc = tpb_get(e_, TPB_ARR_NUM_POINTS, &n, sizeof(n), 0);
    if (rc == TPB_OK) rc = tpb_devptr(e_, TPB_ARR_KNOTS, &unused, &knot_bytes);
    if (rc == TPB_OK) rc = tpb_devptr(e_, TPB_ARR_CTRL, &unused, &ctrl_bytes);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    std::vector<double> kn(knot_bytes / sizeof(double)), ct(ctrl_bytes / sizeof(double));
    tpb_get(e_, TPB_ARR_KNOTS, kn.data(), knot_bytes, 0);
    tpb_get(e_, TPB_ARR_CTRL, ct.data(), ctrl_bytes, 0);
    if (n <= 0) {
      knots_.clear();
      control_points_.clear();
      return tp::OkStatus();
    }
    knots_.assign(kn.begin(), kn.begin() + n + 3);
    control_points_.assign(n, tp::VectorXd(D));
    for (int i = 0; i < n; ++i) control_points_[i] = tp::VectorXd(&ct[static_cast<size_t>(i) * D], D);
    return tp::OkStatus();
  }

  // SwitchToWaypointPath (.cc:209-250) on the device (tpb_switch_to_waypoint_path):
  // truncate at keep_path_until, project the switching position onto the new waypoints,
  // append their control polygon.  The path is kModifiedPath afterwards (also when the
  // switch fails, as in the reference).
  tp::Status SwitchToWaypointPath(const double keep_path_until, tp::Span<const tp::VectorXd> waypoints) {
    path_state_ = State::kModifiedPath;
    if (!e_) return tp::FailedPreconditionError("Spline is empty or invalid.");
    for (const auto& wp : waypoints) {
      if (wp.size() != options_.num_dofs()) {
        return tp::InvalidArgumentError("Invalid number of joints: waypoint.size: " + std::to_string(wp.size()));
      }
    }
    if (waypoints.empty()) return tp::InvalidArgumentError("No waypoints.");
    const int W = static_cast<int>(waypoints.size()), D = static_cast<int>(options_.num_dofs());
    if (W > wcap_) {
      return {tp::StatusCode::kResourceExhausted,
              "more waypoints than the engine was created for (" + std::to_string(wcap_) + ")"};
    }
    std::vector<double> flat(static_cast<size_t>(wcap_) * D, 0.0);
    size_t w = 0;
    for (const auto& wp : waypoints) {
      std::copy(wp.begin(), wp.end(), flat.begin() + w * D);
      ++w;
    }
    const int32_t nw = W;
    // SetInitialVelocity may follow (the reference's order); PathTimingTrajectory hands
    // the path's initial velocity to the planner before the next Plan()
    int32_t rc = tpb_switch_to_waypoint_path(e_, &keep_path_until, &nw, flat.data(), nullptr, 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    int32_t status = 0;
    rc = tpb_get(e_, TPB_ARR_SWITCH_STATUS, &status, sizeof(status), 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    if (auto st = DownloadSpline(); !st.ok()) return st;
    constraints_valid_ = false;
    if (status != 0) return tp::FromCode(status, "SwitchToWaypointPath failed (truncate / project / extend)");
    return tp::OkStatus();
  }

  tp::Status SetMaxJointVelocity(tp::Span<const double> max_velocity) override {
    if (max_velocity.size() != options_.num_dofs()) {
      return tp::InvalidArgumentError("Dimension error, need " + std::to_string(options_.num_dofs()) +
                                      " joint values but max_velocity has " + std::to_string(max_velocity.size()));
    }
    max_joint_velocity_ = FromSpan(max_velocity);
    return tp::OkStatus();
  }
  tp::Status SetMaxJointAcceleration(tp::Span<const double> max_acceleration) override {
    if (max_acceleration.size() != options_.num_dofs()) {
      return tp::InvalidArgumentError("Dimension error, need " + std::to_string(options_.num_dofs()) +
                                      " joint values but max_velocity has " + std::to_string(max_acceleration.size()));
    }
    max_joint_acceleration_ = FromSpan(max_acceleration);
    return tp::OkStatus();
  }
  const tp::VectorXd& GetMaxJointVelocity() const override { return max_joint_velocity_; }
  const tp::VectorXd& GetMaxJointAcceleration() const override { return max_joint_acceleration_; }
  tp::Status SetInitialVelocity(tp::Span<const double> velocity) override {
    if (velocity.size() != NumDofs()) {
      return tp::InvalidArgumentError("Velocity dimension doesn't match number of dofs.");
    }
    initial_velocity_ = FromSpan(velocity);
    return tp::OkStatus();
  }
  const tp::VectorXd& GetInitialVelocity() const override { return initial_velocity_; }
  bool CloseToEnd(double parameter) const override {
    constexpr double kSmall = 1e-4;  // .cc:33
    return knots_.empty() || parameter >= knots_.back() - kSmall;
  }
  State GetState() const override { return path_state_; }

  // SamplePath (.cc:294-318); also evaluates the per-sample limit curve on the device.
  tp::Status SamplePath(const double path_start) override {
    if (!e_ || waypoints_.empty()) return tp::FailedPreconditionError("No path set");
    parameter_start_ = path_start;
    parameter_end_ = options_.num_path_samples() * options_.delta_parameter();
    int32_t rc = tpb_set_joint_limits(e_, max_joint_velocity_.data(), max_joint_acceleration_.data(), 0, 0);
    if (rc == TPB_OK) rc = tpb_sample_paths(e_, &path_start, 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    const size_t N = options_.num_path_samples(), D = options_.num_dofs();
    std::vector<double> buf(N * D);
    tp::VectorXd* dst[3] = {path_position_.data(), first_path_derivative_.data(), second_path_derivative_.data()};
    const tpb_array what[3] = {TPB_ARR_Q, TPB_ARR_DQ, TPB_ARR_DDQ};
    for (int k = 0; k < 3; ++k) {
      rc = tpb_get(e_, what[k], buf.data(), buf.size() * sizeof(double), 0);
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
      for (size_t i = 0; i < N; ++i) dst[k][i] = tp::VectorXd(&buf[i * D], D);
    }
    path_state_ = State::kPathWasSampled;
    constraints_valid_ = false;
    return tp::OkStatus();
  }

  // Device planner hooks (timeable_path.h): the limits go to the device before a
  // Plan(); afterwards the last planning window's samples are mirrored on the host
  // and the path counts as sampled.
  bool SupportsDevicePlanner() const override { return true; }
  tp::Status BeforeDevicePlan() override {
    if (!e_ || waypoints_.empty()) return tp::FailedPreconditionError("No path set");
    const int32_t rc =
        tpb_set_joint_limits(e_, max_joint_velocity_.data(), max_joint_acceleration_.data(), 0, 0);
    return rc == TPB_OK ? tp::OkStatus() : tp::FromCode(rc, tpb_last_error(e_));
  }
  tp::Status AfterDevicePlan(bool window_was_sampled) override {
    if (!window_was_sampled) return tp::OkStatus();
    const size_t N = options_.num_path_samples(), D = options_.num_dofs();
    std::vector<double> buf(N * D);
    tp::VectorXd* dst[3] = {path_position_.data(), first_path_derivative_.data(), second_path_derivative_.data()};
    const tpb_array what[3] = {TPB_ARR_Q, TPB_ARR_DQ, TPB_ARR_DDQ};
    for (int k = 0; k < 3; ++k) {
      const int32_t rc = tpb_get(e_, what[k], buf.data(), buf.size() * sizeof(double), 0);
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
      for (size_t i = 0; i < N; ++i) dst[k][i] = tp::VectorXd(&buf[i * D], D);
    }
    std::vector<double> s(N);
    tpb_get(e_, TPB_ARR_S, s.data(), N * sizeof(double), 0);
    parameter_start_ = s[0];
    parameter_end_ = options_.num_path_samples() * options_.delta_parameter();
    path_state_ = State::kPathWasSampled;
    return ConstraintSetup();  // host copy of the rows, as after ComputeTimingProfile
  }

  // ConstraintSetup (.cc:320-343).  The device never materialises these rows;
  // the host copy exists only for GetConstraints().
  tp::Status ConstraintSetup() override {
    const size_t N = options_.num_path_samples(), D = options_.num_dofs();
    constraints_.resize(N);
    const double safety = options_.constraint_safety();
    for (size_t idx = 0; idx < N; ++idx) {
      auto& c = constraints_[idx];
      c.resize(static_cast<int>(num_constraints_));
      for (size_t dof = 0; dof < D; ++dof) {
        const double d1 = first_path_derivative_[idx][dof];
        c.a_coefficient(dof) = d1;
        c.b_coefficient(dof) = second_path_derivative_[idx][dof];
        c.upper(dof) = max_joint_acceleration_[dof] * safety;
        c.lower(dof) = -max_joint_acceleration_[dof] * safety;
        c.a_coefficient(D + dof) = 0.0;
        c.b_coefficient(D + dof) = d1 * d1;
        const double vs = max_joint_velocity_[dof] * safety;
        c.upper(D + dof) = vs * vs;
        c.lower(D + dof) = 0.0;
      }
    }
    constraints_valid_ = true;
    return tp::OkStatus();
  }
  const std::vector<TimeOptimalPathProfile::Constraint>& GetConstraints() const override { return constraints_; }
  size_t NumConstraints() const override { return num_constraints_; }
  size_t NumDofs() const override { return options_.num_dofs(); }
  size_t NumPathSamples() const override { return options_.num_path_samples(); }
  void Reset() override {
    waypoints_.clear();
    control_points_.clear();
    knots_.clear();
    path_state_ = State::kNoPath;
    parameter_start_ = std::numeric_limits<double>::quiet_NaN();
    parameter_end_ = std::numeric_limits<double>::quiet_NaN();
    initial_velocity_.setZero();
  }
  const tp::VectorXd& GetPathStart() const override { return path_position_.front(); }
  const tp::VectorXd& GetPathEnd() const override { return path_position_.back(); }
  const std::vector<tp::VectorXd>& GetWaypoints() const { return waypoints_; }
  double GetParameterStart() const override { return parameter_start_; }
  double GetParameterEnd() const override { return parameter_end_; }
  const tp::VectorXd& GetPathPositionAt(size_t n) const override { return path_position_.at(n); }
  const tp::VectorXd& GetFirstPathDerivativeAt(size_t n) const override { return first_path_derivative_.at(n); }
  const tp::VectorXd& GetSecondPathDerivativeAt(size_t n) const override { return second_path_derivative_.at(n); }
  int GetNumPathSamples() const override { return static_cast<int>(options_.num_path_samples()); }
  double GetPathSamplingDistance() const override { return options_.delta_parameter(); }

  // Extensions: the fitted spline and the engine that owns the device state.
  const std::vector<tp::VectorXd>& GetControlPoints() const { return control_points_; }
  const std::vector<double>& GetKnots() const { return knots_; }
  tpb_engine* engine() { return e_; }
  tpb_engine* DeviceEngine() override { return e_; }
  uint64_t DeviceEngineGeneration() const override { return engine_generation_; }

 private:
  static constexpr int kSplineOrder = 2;  // .h:89
  const JointPathOptions options_;
  const size_t num_constraints_;
  int device_ = 0;
  tpb_engine* e_ = nullptr;
  int wcap_ = 0;
  uint64_t engine_generation_ = 0;
  State path_state_ = State::kNoPath;
  std::vector<tp::VectorXd> waypoints_, control_points_;
  std::vector<tp::VectorXd> path_position_, first_path_derivative_, second_path_derivative_;
  std::vector<double> knots_;
  std::vector<TimeOptimalPathProfile::Constraint> constraints_;
  bool constraints_valid_ = false;
  tp::VectorXd max_joint_velocity_, max_joint_acceleration_, initial_velocity_;
  double parameter_start_ = 0, parameter_end_ = 0;
};

}  // namespace trajectory_planning
#endif  // TRAJECTORY_PLANNING_TIMEABLE_PATH_JOINT_SPLINE_H_
