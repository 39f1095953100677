// Host-side mirror of trajectory_planning/timeable_path.h:44-166: PathOptions
// (same defaults) and the abstract TimeablePath interface PathTimingTrajectory
// consumes.  This is the drop-in seam of the reference.
#ifndef TRAJECTORY_PLANNING_TIMEABLE_PATH_H_
#define TRAJECTORY_PLANNING_TIMEABLE_PATH_H_

#include <cstdint>
#include <type_traits>
#include <vector>

#include "tp/base.h"
#include "trajectory_planning/time_optimal_path_timing.h"

namespace trajectory_planning {

template <typename EigenVector>
tp::Span<const double> ToSpan(const EigenVector& vec) {
  return tp::Span<const double>(vec.data(), vec.size());
}
inline tp::VectorXd FromSpan(tp::Span<const double> span) {
  return tp::VectorXd(span.data(), span.size());
}

// timeable_path.h:44-90 (defaults: safety 0.8, rounding 0.2, 500 samples, 0.005)
template <typename DerivedOptions>
class PathOptions {
 public:
  double constraint_safety() const { return constraint_safety_; }
  DerivedOptions& set_constraint_safety(double safety) {
    constraint_safety_ = safety;
    return static_cast<DerivedOptions&>(*this);
  }
  double rounding() const { return rounding_; }
  DerivedOptions& set_rounding(double rounding) {
    rounding_ = rounding;
    return static_cast<DerivedOptions&>(*this);
  }
  size_t num_dofs() const { return num_dofs_; }
  DerivedOptions& set_num_dofs(size_t dofs) {
    num_dofs_ = dofs;
    return static_cast<DerivedOptions&>(*this);
  }
  size_t num_path_samples() const { return num_path_samples_; }
  DerivedOptions& set_num_path_samples(size_t num_samples) {
    num_path_samples_ = num_samples;
    return static_cast<DerivedOptions&>(*this);
  }
  double delta_parameter() const { return delta_parameter_; }
  DerivedOptions& set_delta_parameter(double delta_parameter) {
    delta_parameter_ = delta_parameter;
    return static_cast<DerivedOptions&>(*this);
  }

 private:
  double constraint_safety_ = 0.8;
  double rounding_ = 0.2;
  size_t num_dofs_ = 0;
  size_t num_path_samples_ = 500;
  double delta_parameter_ = 0.005;
};

class TimeablePath {
 public:
  enum class State { kNoPath, kNewPath, kModifiedPath, kPathWasSampled };
  virtual ~TimeablePath() = default;
  virtual tp::Status SetMaxJointVelocity(tp::Span<const double> max_velocity) = 0;
  virtual tp::Status SetMaxJointAcceleration(tp::Span<const double> max_acceleration) = 0;
  virtual const tp::VectorXd& GetMaxJointVelocity() const = 0;
  virtual const tp::VectorXd& GetMaxJointAcceleration() const = 0;
  virtual tp::Status SetInitialVelocity(tp::Span<const double> velocity) = 0;
  virtual const tp::VectorXd& GetInitialVelocity() const = 0;
  virtual bool CloseToEnd(double parameter) const = 0;
  virtual State GetState() const = 0;
  virtual tp::Status SamplePath(double path_start) = 0;
  virtual int GetNumPathSamples() const = 0;
  virtual double GetPathSamplingDistance() const = 0;
  virtual tp::Status ConstraintSetup() = 0;
  virtual const std::vector<TimeOptimalPathProfile::Constraint>& GetConstraints() const = 0;
  virtual size_t NumConstraints() const = 0;
  virtual size_t NumDofs() const = 0;
  virtual size_t NumPathSamples() const = 0;
  virtual void Reset() = 0;
  virtual const tp::VectorXd& GetPathStart() const = 0;
  virtual const tp::VectorXd& GetPathEnd() const = 0;
  virtual double GetParameterStart() const = 0;
  virtual double GetParameterEnd() const = 0;
  virtual const tp::VectorXd& GetPathPositionAt(size_t n) const = 0;
  virtual const tp::VectorXd& GetFirstPathDerivativeAt(size_t n) const = 0;
  virtual const tp::VectorXd& GetSecondPathDerivativeAt(size_t n) const = 0;
  // Extension (not in the reference): the engine that holds this path's device
  // state once SamplePath + ConstraintSetup have run; PathTimingTrajectory times
  // the path there.  nullptr = not a device-backed path.
  virtual tpb_engine* DeviceEngine() { return nullptr; }
  // Extensions for PathTimingTrajectory's device planner (tpb_planner_plan), which
  // samples and times the planning windows of the path without host round trips:
  // whether the path can be planned that way, upload of what the device needs before
  // a Plan(), and the state SamplePath would have left behind after one.
  virtual bool SupportsDevicePlanner() const { return false; }
  virtual tp::Status BeforeDevicePlan() { return tp::OkStatus(); }
  virtual tp::Status AfterDevicePlan(bool /*window_was_sampled*/) { return tp::OkStatus(); }
  // Counts the engines DeviceEngine() has returned over the object's life: a planner that
  // keeps state on an engine tells a re-created engine from the one it planned on even
  // when the allocator hands the same address back.
  virtual uint64_t DeviceEngineGeneration() const { return 0; }
  // the Status a user callback returned during the device plan (Ok if none failed)
  virtual tp::Status TakeCallbackStatus() { return tp::OkStatus(); }
};

inline constexpr const char* ToString(const TimeablePath::State state) {
  switch (state) {
    case TimeablePath::State::kNoPath: return "kNoPath";
    case TimeablePath::State::kNewPath: return "kNewPath";
    case TimeablePath::State::kModifiedPath: return "kModifiedPath";
    case TimeablePath::State::kPathWasSampled: return "kPathWasSampled";
  }
  return "Invalid TimeablePath::State enum value";
}

}  // namespace trajectory_planning
#endif  // TRAJECTORY_PLANNING_TIMEABLE_PATH_H_
