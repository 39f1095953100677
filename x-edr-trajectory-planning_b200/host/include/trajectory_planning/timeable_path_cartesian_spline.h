// Host-side mirror of TimeableCartesianSplinePath
// (trajectory_planning/timeable_path_cartesian_spline.h:31-190, .cc:39-680): same
// public methods, options and state machine.  Spline fitting (pose + joint
// control polygons, knots), the quaternion / translation / joint spline
// evaluation of SamplePath, the forward-difference path derivatives, the 2D+2
// constraint rows and the limit curve run on the GPU (tpb_set_cart_waypoints,
// tpb_cart_sample_targets, tpb_cart_setup_constraints).
//
// The user's PathIKFunc / JacobianFunc are host callbacks with the reference's
// signatures: SamplePath downloads the pose / joint targets, calls the path IK
// once for the window (sequential warm-start semantics stay the user's), and
// uploads the solution; ConstraintSetup calls the Jacobian per sample and
// uploads the 6 x D matrices.  Extension: set_device_kinematics(TPB_KIN_FAKE7)
// selects a built-in device model instead, then nothing leaves the GPU.
//
// Scope: the reference's whole class -- fresh paths, several planning windows per path with
// the growing path_ik_positions_ cache, SwitchToWaypointPath (online replanning).
#ifndef TRAJECTORY_PLANNING_TIMEABLE_PATH_CARTESIAN_SPLINE_H_
#define TRAJECTORY_PLANNING_TIMEABLE_PATH_CARTESIAN_SPLINE_H_

#include <algorithm>
#include <cmath>
#include <functional>
#include <string>
#include <vector>

#include "trajectory_planning/timeable_path.h"

namespace trajectory_planning {

class CartesianPathOptions : public PathOptions<CartesianPathOptions> {
 public:
  // Path IK function, taking (initial_value, pose_targets, joint_targets, ik_result).
  typedef std::function<tp::Status(const tp::VectorXd&, const std::vector<tp::Pose3d>&,
                                   const std::vector<tp::VectorXd>&, std::vector<tp::VectorXd>*)>
      PathIKFunc;
  typedef std::function<tp::Status(const tp::VectorXd&, tp::Matrix6Xd*)> JacobianFunc;
  double translation_rounding() const { return translational_rounding_; }
  CartesianPathOptions& set_translation_rounding(double rounding) {
    translational_rounding_ = rounding;
    return *this;
  }
  PathIKFunc GetPathIKFunc() const { return path_ik_func_; }
  CartesianPathOptions& set_path_ik_func(PathIKFunc path_ik) {
    path_ik_func_ = path_ik;
    return *this;
  }
  JacobianFunc GetJacobianFunc() const { return jacobian_func_; }
  CartesianPathOptions& set_jacobian_func(JacobianFunc jacobian) {
    jacobian_func_ = jacobian;
    return *this;
  }
  // Extension: built-in device kinematics (TPB_KIN_*) instead of host callbacks.
  int device_kinematics() const { return device_kinematics_; }
  CartesianPathOptions& set_device_kinematics(int model) {
    device_kinematics_ = model;
    return *this;
  }

 private:
  double translational_rounding_ = 0.05;  // .h:50
  PathIKFunc path_ik_func_;
  JacobianFunc jacobian_func_;
  int device_kinematics_ = TPB_KIN_HOST;
};

class TimeableCartesianSplinePath : public TimeablePath {
 public:
  explicit TimeableCartesianSplinePath(const CartesianPathOptions& options, int device = 0)
      : options_(options), num_constraints_(options.num_dofs() * 2 + 2), device_(device) {
    // CHECKs of the reference constructor (.cc:117-122)
    if (options_.num_path_samples() < 3) std::abort();
    if (options_.device_kinematics() == TPB_KIN_HOST &&
        (!options_.GetJacobianFunc() || !options_.GetPathIKFunc())) {
      fprintf(stderr, "Need a Jacobian and a path IK function set in Options.\n");
      std::abort();
    }
    path_ik_func_ = options_.GetPathIKFunc();
    jacobian_func_ = options_.GetJacobianFunc();
    const size_t N = options_.num_path_samples(), D = options_.num_dofs();
    path_position_.assign(N, tp::VectorXd(D));
    first_path_derivative_.assign(N, tp::VectorXd(D));
    second_path_derivative_.assign(N, tp::VectorXd(D));
    max_joint_velocity_ = tp::VectorXd(D);
    max_joint_acceleration_ = tp::VectorXd(D);
    initial_velocity_ = tp::VectorXd(D);
    Reset();
  }
  ~TimeableCartesianSplinePath() override { if (e_) tpb_destroy(e_); }
  TimeableCartesianSplinePath(const TimeableCartesianSplinePath&) = delete;
  TimeableCartesianSplinePath& operator=(const TimeableCartesianSplinePath&) = delete;

  // SetWaypoints (.cc:166-188) + FitSplineToWaypoints (.cc:396-462)
  tp::Status SetWaypoints(tp::Span<const tp::Pose3d> pose_waypoints,
                          tp::Span<const tp::VectorXd> joint_waypoints) {
    path_state_ = State::kNewPath;
    if (joint_waypoints.size() != pose_waypoints.size()) {
      return tp::InvalidArgumentError("'joint_waypoints' and 'pose_waypoints' have different sizes.");
    }
    for (const auto& wp : joint_waypoints) {
      if (wp.size() != options_.num_dofs()) {
        return tp::InvalidArgumentError("Dimension error, need " + std::to_string(options_.num_dofs()) +
                                        " joint values but waypoint has " + std::to_string(wp.size()));
      }
    }
    if (joint_waypoints.empty()) return tp::InvalidArgumentError("no waypoints");
    joint_waypoints_.assign(joint_waypoints.begin(), joint_waypoints.end());
    pose_waypoints_.assign(pose_waypoints.begin(), pose_waypoints.end());
    path_ik_positions_.clear();
    return FitSplineToWaypoints();
  }

  // SwitchToWaypointPath (.cc:183-296) on the device (tpb_switch_to_cart_waypoint_path): the
  // joint, translation and rotation splines are truncated at keep_path_until, the switching
  // position is projected onto the new waypoints, their control polygons are appended.  The
  // path is kModifiedPath afterwards (also when the switch fails, as in the reference).
  tp::Status SwitchToWaypointPath(const double keep_path_until, tp::Span<const tp::Pose3d> pose_waypoints,
                                  tp::Span<const tp::VectorXd> joint_waypoints) {
    // erase the obsolete part of the IK solution (.cc:188-192)
    if (const int ik_index = PathIkIndex(keep_path_until);
        ik_index >= 0 && static_cast<size_t>(ik_index) < path_ik_positions_.size()) {
      path_ik_positions_.resize(ik_index + 1);
    }
    if (joint_waypoints.size() != pose_waypoints.size()) {
      return tp::InvalidArgumentError("'joint_waypoints' and 'pose_waypoints' have different sizes.");
    }
    path_state_ = State::kModifiedPath;
    if (!e_ || joint_waypoints_.empty()) return tp::FailedPreconditionError("Spline is empty or invalid.");
    if (joint_waypoints.empty()) return tp::InvalidArgumentError("No waypoints.");
    const int W = static_cast<int>(joint_waypoints.size()), D = static_cast<int>(options_.num_dofs());
    for (const auto& wp : joint_waypoints) {
      if (static_cast<int>(wp.size()) != D) {
        return tp::InvalidArgumentError("Invalid number of joints: waypoint.size: " + std::to_string(wp.size()));
      }
    }
    if (W > wcap_) {
      return {tp::StatusCode::kResourceExhausted,
              "more waypoints than the engine was created for (" + std::to_string(wcap_) + ")"};
    }
    std::vector<double> jflat(static_cast<size_t>(wcap_) * D, 0.0), pflat(static_cast<size_t>(wcap_) * 7, 0.0);
    for (int w = 0; w < W; ++w) {
      std::copy(joint_waypoints[w].begin(), joint_waypoints[w].end(), jflat.begin() + static_cast<size_t>(w) * D);
      double* o = &pflat[static_cast<size_t>(w) * 7];
      for (int k = 0; k < 3; ++k) o[k] = pose_waypoints[w].translation()[k];
      for (int k = 0; k < 4; ++k) o[3 + k] = pose_waypoints[w].quaternion().coeffs()[k];
    }
    const int32_t nw = W;
    int32_t rc = tpb_set_path_options(e_, options_.constraint_safety(), options_.rounding());
    if (rc == TPB_OK) {
      rc = tpb_set_cart_options(e_, options_.translation_rounding(),
                                max_translational_velocity_ > 0 ? max_translational_velocity_ : 1.0,
                                max_rotational_velocity_ > 0 ? max_rotational_velocity_ : 1.0);
    }
    // SetInitialVelocity may follow (the reference's order); PathTimingTrajectory hands the
    // path's initial velocity to the planner before the next Plan()
    if (rc == TPB_OK) {
      rc = tpb_switch_to_cart_waypoint_path(e_, &keep_path_until, &nw, pflat.data(), jflat.data(), nullptr, 0);
    }
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    int32_t status = 0;
    rc = tpb_get(e_, TPB_ARR_SWITCH_STATUS, &status, sizeof(status), 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    if (auto st = DownloadSplines(); !st.ok()) return st;
    if (status != 0) return tp::FromCode(status, "SwitchToWaypointPath failed (truncate / project / extend)");
    return tp::OkStatus();
  }

  // SetMaxCartesianVelocity (.cc:352-366)
  tp::Status SetMaxCartesianVelocity(double max_translational_velocity, double max_rotational_velocity) {
    if (max_translational_velocity <= 0 || max_rotational_velocity <= 0) {
      return tp::InvalidArgumentError(
          "Velocity limits must be positive, but got max_translational_velocity= " +
          std::to_string(max_translational_velocity) +
          ", max_rotational_velocity= " + std::to_string(max_rotational_velocity));
    }
    max_rotational_velocity_ = max_rotational_velocity;
    max_translational_velocity_ = max_translational_velocity;
    return tp::OkStatus();
  }
  double GetTranslationRounding() const { return options_.translation_rounding(); }
  double GetRotationRounding() const { return options_.rounding(); }
  // SetTranslationRounding / SetRotationRounding (.cc:368-390)
  tp::Status SetTranslationRounding(double translation_rounding) {
    if (translation_rounding <= 0.0) {
      return tp::InvalidArgumentError("translation_rounding needs to be greater than zero, got " +
                                      std::to_string(translation_rounding));
    }
    options_.set_translation_rounding(translation_rounding);
    return tp::OkStatus();
  }
  tp::Status SetRotationRounding(double rotation_rounding) {
    if (rotation_rounding <= 0.0) {
      return tp::InvalidArgumentError("rotation_rounding needs to be greater than zero, got " +
                                      std::to_string(rotation_rounding));
    }
    options_.set_rounding(rotation_rounding);
    return tp::OkStatus();
  }

  // TimeablePath interface ---------------------------------------------------
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
    constexpr double kSmall = 1e-4;  // .cc:37
    return knots_.empty() || parameter >= knots_.back() - kSmall;
  }
  State GetState() const override { return path_state_; }

  // SamplePath (.cc:464-549) for a fresh path: IK indices -1 .. N-1.
  tp::Status SamplePath(const double path_start) override {
    if (!e_ || joint_waypoints_.empty()) return tp::FailedPreconditionError("No path set");
    if (path_start != 0.0 || !path_ik_positions_.empty()) {
      return tp::UnimplementedError("Cartesian paths are sampled from path_start = 0 in one window");
    }
    const size_t N = options_.num_path_samples(), D = options_.num_dofs();
    const double path_horizon = path_start + options_.delta_parameter() * (N - 1);
    int32_t rc = tpb_cart_sample_targets(e_);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    std::vector<double> buf(N * D);
    if (options_.device_kinematics() != TPB_KIN_HOST) {
      rc = tpb_cart_solve_ik(e_, options_.device_kinematics());
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
      rc = tpb_get(e_, TPB_ARR_Q, buf.data(), buf.size() * sizeof(double), 0);
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
      path_ik_positions_.assign(N, tp::VectorXd(D));
      for (size_t i = 0; i < N; ++i) path_ik_positions_[i] = tp::VectorXd(&buf[i * D], D);
    } else {
      // download what the reference hands to the path IK (.cc:470-500)
      std::vector<double> pt((N + 1) * 7), jt((N + 1) * D);
      rc = tpb_get(e_, TPB_ARR_POSE_TARGETS, pt.data(), pt.size() * sizeof(double), 0);
      if (rc == TPB_OK) rc = tpb_get(e_, TPB_ARR_JOINT_TARGETS, jt.data(), jt.size() * sizeof(double), 0);
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
      sampled_pose_targets_.resize(N + 1);
      sampled_joint_targets_.resize(N + 1);
      for (size_t m = 0; m <= N; ++m) {
        const double* p = &pt[m * 7];
        sampled_pose_targets_[m] =
            tp::Pose3d(tp::Quaterniond(p[6], p[3], p[4], p[5]), tp::Vector3d(p[0], p[1], p[2]));
        sampled_joint_targets_[m] = tp::VectorXd(&jt[m * D], D);
      }
      // initial value: the first joint target (current_ik_upper_index == -1, .cc:503-505)
      const tp::VectorXd initial_value = sampled_joint_targets_.front();
      new_ik_path_.clear();
      const tp::Status ik_status =
          path_ik_func_(initial_value, sampled_pose_targets_, sampled_joint_targets_, &new_ik_path_);
      // appended from the second sample on, also on failure (.cc:508-519)
      if (!new_ik_path_.empty()) {
        path_ik_positions_.insert(path_ik_positions_.end(), new_ik_path_.begin() + 1, new_ik_path_.end());
      }
      if (!ik_status.ok()) return ik_status;
      if (path_ik_positions_.size() != N) {
        return tp::InternalError("path IK returned " + std::to_string(new_ik_path_.size()) +
                                 " samples for " + std::to_string(N + 1) + " targets");
      }
      for (size_t i = 0; i < N; ++i) {
        if (path_ik_positions_[i].size() != D) return tp::InternalError("path IK sample has wrong size");
        std::copy(path_ik_positions_[i].begin(), path_ik_positions_[i].end(), buf.begin() + i * D);
      }
      rc = tpb_cart_set_ik_positions(e_, buf.data(), 0);
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    }
    std::copy(path_ik_positions_.begin(), path_ik_positions_.end(), path_position_.begin());
    ComputePathDerivatives();
    path_state_ = State::kPathWasSampled;
    parameter_start_ = path_start;
    parameter_end_ = path_horizon;
    return tp::OkStatus();
  }

  // ConstraintSetup (.cc:551-595): rows and limit curve on the device; the host
  // copy of the rows is downloaded for GetConstraints().
  tp::Status ConstraintSetup() override {
    if (!e_ || path_state_ != State::kPathWasSampled) return tp::FailedPreconditionError("path not sampled");
    const size_t N = options_.num_path_samples(), D = options_.num_dofs();
    int32_t rc = tpb_set_joint_limits(e_, max_joint_velocity_.data(), max_joint_acceleration_.data(), 0, 0);
    if (rc == TPB_OK) {
      rc = tpb_set_cart_options(e_, options_.translation_rounding(), max_translational_velocity_,
                                max_rotational_velocity_);
    }
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    if (options_.device_kinematics() != TPB_KIN_HOST) {
      rc = tpb_cart_setup_constraints(e_, options_.device_kinematics(), nullptr, 0);
    } else {
      std::vector<double> jac(N * 6 * D);
      tp::Matrix6Xd jacobian(6, D);
      for (size_t idx = 0; idx < N; ++idx) {
        jacobian.setZero();
        if (auto st = jacobian_func_(path_position_[idx], &jacobian); !st.ok()) return st;
        if (jacobian.cols() != D) return tp::InternalError("Jacobian has wrong number of columns");
        std::copy(jacobian.data(), jacobian.data() + 6 * D, jac.begin() + idx * 6 * D);
      }
      rc = tpb_cart_setup_constraints(e_, TPB_KIN_HOST, jac.data(), 0);
    }
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    return DownloadConstraints();
  }
  // host copy of the rows of the window the engine holds
  tp::Status DownloadConstraints() {
    const size_t N = options_.num_path_samples();
    std::vector<double> cons(N * num_constraints_ * 4);
    const int32_t rc = tpb_get(e_, TPB_ARR_CONS, cons.data(), cons.size() * sizeof(double), 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    constraints_.resize(N);
    for (size_t idx = 0; idx < N; ++idx) {
      auto& c = constraints_[idx];
      c.resize(static_cast<int>(num_constraints_));
      const double* src = &cons[idx * num_constraints_ * 4];
      for (size_t r = 0; r < num_constraints_; ++r) {
        c.a_coefficient(r) = src[r * 4 + 0];
        c.b_coefficient(r) = src[r * 4 + 1];
        c.lower(r) = src[r * 4 + 2];
        c.upper(r) = src[r * 4 + 3];
      }
    }
    return tp::OkStatus();
  }
  const std::vector<TimeOptimalPathProfile::Constraint>& GetConstraints() const override { return constraints_; }

  // Device planner hooks (timeable_path.h): PathTimingTrajectory plans all windows
  // through tpb_planner_plan.  With built-in device kinematics nothing leaves the GPU;
  // with the user's PathIKFunc / JacobianFunc every window's new targets come back to
  // the host, the callbacks run (warm-started from the previous window's last solution,
  // .cc:503-517) and positions + Jacobians return to the device
  // (tpb_planner_set_host_kinematics).
  bool SupportsDevicePlanner() const override { return true; }
  tp::Status TakeCallbackStatus() override {
    tp::Status st = callback_status_;
    callback_status_ = tp::OkStatus();
    return st;
  }
  tp::Status BeforeDevicePlan() override {
    if (!e_ || joint_waypoints_.empty()) return tp::FailedPreconditionError("No path set");
    int32_t rc = tpb_set_joint_limits(e_, max_joint_velocity_.data(), max_joint_acceleration_.data(), 0, 0);
    if (rc == TPB_OK) {
      if (max_translational_velocity_ <= 0 || max_rotational_velocity_ <= 0) {
        return tp::FailedPreconditionError("Cartesian velocity limits not set");
      }
      rc = tpb_set_cart_options(e_, options_.translation_rounding(), max_translational_velocity_,
                                max_rotational_velocity_);
    }
    if (rc == TPB_OK) {
      rc = options_.device_kinematics() != TPB_KIN_HOST
               ? tpb_planner_set_kinematics(e_, options_.device_kinematics())
               : tpb_planner_set_host_kinematics(e_, &PathIkThunk, &JacobianThunk, this);
    }
    callback_status_ = tp::OkStatus();
    return rc == TPB_OK ? tp::OkStatus() : tp::FromCode(rc, tpb_last_error(e_));
  }
  // tpb_path_ik_fn: the user's PathIKFunc on one window's new targets
  static int32_t PathIkThunk(void* user, int32_t /*path*/, const double* initial_value,
                             int32_t num_samples, const double* pose_targets,
                             const double* joint_targets, double* solution) {
    auto* self = static_cast<TimeableCartesianSplinePath*>(user);
    const size_t D = self->options_.num_dofs(), M = static_cast<size_t>(num_samples);
    self->sampled_pose_targets_.resize(M);
    self->sampled_joint_targets_.resize(M);
    for (size_t m = 0; m < M; ++m) {
      const double* p = pose_targets + m * 7;
      self->sampled_pose_targets_[m] =
          tp::Pose3d(tp::Quaterniond(p[6], p[3], p[4], p[5]), tp::Vector3d(p[0], p[1], p[2]));
      self->sampled_joint_targets_[m] = tp::VectorXd(joint_targets + m * D, D);
    }
    self->new_ik_path_.clear();
    const tp::Status st = self->path_ik_func_(tp::VectorXd(initial_value, D), self->sampled_pose_targets_,
                                              self->sampled_joint_targets_, &self->new_ik_path_);
    for (size_t m = 0; m < M && m < self->new_ik_path_.size(); ++m) {
      if (self->new_ik_path_[m].size() != D) {
        self->callback_status_ = tp::InternalError("path IK sample has wrong size");
        return TPB_INTERNAL;
      }
      std::copy(self->new_ik_path_[m].begin(), self->new_ik_path_[m].end(), solution + m * D);
    }
    if (!st.ok()) {
      self->callback_status_ = st;
      return static_cast<int32_t>(st.code());
    }
    if (self->new_ik_path_.size() != M) {
      self->callback_status_ =
          tp::InternalError("path IK returned " + std::to_string(self->new_ik_path_.size()) +
                            " samples for " + std::to_string(M) + " targets");
      return TPB_INTERNAL;
    }
    return TPB_OK;
  }
  // tpb_jacobian_fn: the user's JacobianFunc at one path sample
  static int32_t JacobianThunk(void* user, int32_t /*path*/, const double* joint_positions,
                               double* jacobian_out) {
    auto* self = static_cast<TimeableCartesianSplinePath*>(user);
    const size_t D = self->options_.num_dofs();
    tp::Matrix6Xd jacobian(6, D);
    jacobian.setZero();
    const tp::Status st = self->jacobian_func_(tp::VectorXd(joint_positions, D), &jacobian);
    if (!st.ok()) {
      self->callback_status_ = st;
      return static_cast<int32_t>(st.code());
    }
    if (jacobian.cols() != D) {
      self->callback_status_ = tp::InternalError("Jacobian has wrong number of columns");
      return TPB_INTERNAL;
    }
    std::copy(jacobian.data(), jacobian.data() + 6 * D, jacobian_out);
    return TPB_OK;
  }
  tp::Status AfterDevicePlan(bool window_was_sampled) override {
    if (!window_was_sampled) return tp::OkStatus();
    // the last window's samples, as SamplePath would have left them (.cc:527-547)
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
    const int32_t rc = tpb_get(e_, TPB_ARR_S, s.data(), N * sizeof(double), 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    parameter_start_ = s[0];
    parameter_end_ = s[0] + options_.delta_parameter() * (N - 1);
    // path_ik_positions_ up to the window's horizon index; entries of windows that were
    // planned and spliced inside the same Plan() call stay on the device (TPB_ARR_W_Q)
    const size_t first = static_cast<size_t>(std::max(0, PathIkIndex(parameter_start_)));
    if (path_ik_positions_.size() < first) path_ik_positions_.resize(first, path_position_.front());
    path_ik_positions_.resize(first + N);
    std::copy(path_position_.begin(), path_position_.end(), path_ik_positions_.begin() + first);
    path_state_ = State::kPathWasSampled;
    return DownloadConstraints();
  }
  size_t NumConstraints() const override { return num_constraints_; }
  size_t NumDofs() const override { return options_.num_dofs(); }
  size_t NumPathSamples() const override { return options_.num_path_samples(); }
  void Reset() override {  // .cc:612-621
    joint_waypoints_.clear();
    pose_waypoints_.clear();
    joint_control_points_.clear();
    pose_control_points_.clear();
    path_ik_positions_.clear();
    knots_.clear();
    path_state_ = State::kNoPath;
    parameter_start_ = -1.0;
    parameter_end_ = -1.0;
  }
  const tp::VectorXd& GetPathStart() const override { return joint_waypoints_.front(); }
  const tp::VectorXd& GetPathEnd() const override { return path_position_.back(); }
  const std::vector<tp::VectorXd>& GetJointWaypoints() const { return joint_waypoints_; }
  const std::vector<tp::Pose3d>& GetPoseWaypoints() const { return pose_waypoints_; }
  double GetParameterStart() const override { return parameter_start_; }
  double GetParameterEnd() const override { return parameter_end_; }
  const tp::VectorXd& GetPathPositionAt(size_t n) const override { return path_position_.at(n); }
  const tp::VectorXd& GetFirstPathDerivativeAt(size_t n) const override { return first_path_derivative_.at(n); }
  const tp::VectorXd& GetSecondPathDerivativeAt(size_t n) const override { return second_path_derivative_.at(n); }
  const std::vector<tp::VectorXd>& GetSplineIKPosition() const { return path_ik_positions_; }
  int GetNumPathSamples() const override { return static_cast<int>(options_.num_path_samples()); }
  double GetPathSamplingDistance() const override { return options_.delta_parameter(); }
  // PathIkIndex / PathIkParameter (.cc:671-678)
  int PathIkIndex(const double path_parameter) const {
    return static_cast<int>(std::round(path_parameter / options_.delta_parameter()));
  }
  double PathIkParameter(const int index) const { return index * options_.delta_parameter(); }

  // Extensions: the fitted splines and the engine that owns the device state.
  const std::vector<tp::Pose3d>& GetPoseControlPoints() const { return pose_control_points_; }
  const std::vector<tp::VectorXd>& GetJointControlPoints() const { return joint_control_points_; }
  const std::vector<double>& GetKnots() const { return knots_; }
  tpb_engine* DeviceEngine() override { return e_; }
  uint64_t DeviceEngineGeneration() const override { return engine_generation_; }

 private:
  // FitSplineToWaypoints (.cc:396-462) on the device.
  tp::Status FitSplineToWaypoints() {
    const int W = static_cast<int>(joint_waypoints_.size()), D = static_cast<int>(options_.num_dofs());
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
      cfg.cartesian = 1;
      const int32_t rc = tpb_create(&cfg, &e_);
      if (rc != TPB_OK) {
        e_ = nullptr;
        return tp::FromCode(rc, "tpb_create failed (no CUDA device; there is no CPU fallback)");
      }
      wcap_ = cfg.max_waypoints;
      ++engine_generation_;
    }
    std::vector<double> jflat(static_cast<size_t>(wcap_) * D, 0.0), pflat(static_cast<size_t>(wcap_) * 7, 0.0);
    for (int w = 0; w < W; ++w) {
      std::copy(joint_waypoints_[w].begin(), joint_waypoints_[w].end(), jflat.begin() + w * D);
      const auto& p = pose_waypoints_[w];
      double* o = &pflat[static_cast<size_t>(w) * 7];
      for (int k = 0; k < 3; ++k) o[k] = p.translation()[k];
      for (int k = 0; k < 4; ++k) o[3 + k] = p.quaternion().coeffs()[k];
    }
    const int32_t nw = W;
    tpb_set_path_options(e_, options_.constraint_safety(), options_.rounding());
    int32_t rc = tpb_set_cart_options(e_, options_.translation_rounding(),
                                      max_translational_velocity_ > 0 ? max_translational_velocity_ : 1.0,
                                      max_rotational_velocity_ > 0 ? max_rotational_velocity_ : 1.0);
    if (rc == TPB_OK) rc = tpb_set_cart_waypoints(e_, 1, &nw, pflat.data(), jflat.data(), 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    const double delta = options_.delta_parameter();
    rc = tpb_set_sampling(e_, &delta, 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    return DownloadSplines();
  }

  // host copies of the knots and of the three splines' control points (their device arrays
  // grow to the reference's knot capacity with the first switch)
  tp::Status DownloadSplines() {
    const size_t D = options_.num_dofs();
    int32_t n = 0;
    int32_t rc = tpb_get(e_, TPB_ARR_NUM_POINTS, &n, sizeof(n), 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    std::vector<double> buf[3];
    const tpb_array what[3] = {TPB_ARR_KNOTS, TPB_ARR_CTRL, TPB_ARR_POSE_CTRL};
    for (int k = 0; k < 3; ++k) {
      void* ptr = nullptr;
      size_t bytes = 0;
      rc = tpb_devptr(e_, what[k], &ptr, &bytes);
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
      buf[k].resize(bytes / sizeof(double));
      rc = tpb_get(e_, what[k], buf[k].data(), bytes, 0);
      if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    }
    knots_.assign(buf[0].begin(), buf[0].begin() + (n > 0 ? n + 3 : 0));
    joint_control_points_.assign(n, tp::VectorXd(D));
    pose_control_points_.resize(n);
    for (int i = 0; i < n; ++i) {
      joint_control_points_[i] = tp::VectorXd(&buf[1][static_cast<size_t>(i) * D], D);
      const double* p = &buf[2][static_cast<size_t>(i) * 7];
      pose_control_points_[i] =
          tp::Pose3d(tp::Quaterniond(p[6], p[3], p[4], p[5]), tp::Vector3d(p[0], p[1], p[2]));
    }
    return tp::OkStatus();
  }

  // ComputePathDerivatives (.cc:39-68): forward differences on the host copy (the
  // device computes the same values for its constraint rows).
  void ComputePathDerivatives() {
    const size_t N = path_position_.size(), D = options_.num_dofs();
    const double inv_delta_parameter = 1.0 / options_.delta_parameter();
    for (size_t idx = 0; idx + 1 < N; ++idx) {
      for (size_t d = 0; d < D; ++d) {
        first_path_derivative_[idx][d] = inv_delta_parameter * (path_position_[idx + 1][d] - path_position_[idx][d]);
      }
    }
    first_path_derivative_.back().setZero();
    for (size_t idx = 1; idx + 1 < N; ++idx) {
      for (size_t d = 0; d < D; ++d) {
        second_path_derivative_[idx][d] =
            inv_delta_parameter * (first_path_derivative_[idx + 1][d] - first_path_derivative_[idx][d]);
      }
    }
    second_path_derivative_.front().setZero();
    second_path_derivative_.back().setZero();
  }

  static constexpr int kSplineOrder = 2;  // .h:137
  CartesianPathOptions options_;
  const size_t num_constraints_;
  int device_ = 0;
  tpb_engine* e_ = nullptr;
  int wcap_ = 0;
  uint64_t engine_generation_ = 0;
  State path_state_ = State::kNoPath;
  std::vector<tp::VectorXd> joint_waypoints_, joint_control_points_;
  std::vector<tp::Pose3d> pose_waypoints_, pose_control_points_;
  std::vector<tp::Pose3d> sampled_pose_targets_;
  std::vector<tp::VectorXd> sampled_joint_targets_;
  std::vector<tp::VectorXd> path_position_, first_path_derivative_, second_path_derivative_;
  std::vector<double> knots_;
  std::vector<tp::VectorXd> path_ik_positions_, new_ik_path_;
  std::vector<TimeOptimalPathProfile::Constraint> constraints_;
  tp::VectorXd max_joint_velocity_, max_joint_acceleration_, initial_velocity_;
  double max_translational_velocity_ = 0.0, max_rotational_velocity_ = 0.0;
  CartesianPathOptions::PathIKFunc path_ik_func_;
  CartesianPathOptions::JacobianFunc jacobian_func_;
  tp::Status callback_status_;  // what a user callback returned inside tpb_planner_plan
  double parameter_start_ = -1.0, parameter_end_ = -1.0;
};

}  // namespace trajectory_planning
#endif  // TRAJECTORY_PLANNING_TIMEABLE_PATH_CARTESIAN_SPLINE_H_
