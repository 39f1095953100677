// Host-side mirror of trajectory_planning/sampled_trajectory.h:27-58 and
// sampled_trajectory.cc:24-44: the sampled-trajectory structs and the input check.
#ifndef TRAJECTORY_PLANNING_SAMPLED_TRAJECTORY_H_
#define TRAJECTORY_PLANNING_SAMPLED_TRAJECTORY_H_

#include <vector>

#include "tp/base.h"

namespace trajectory_planning {

// sampled_trajectory.h:27-36
struct SampledTrajectory {
  std::vector<double> times;
  std::vector<tp::VectorXd> positions;
  std::vector<tp::VectorXd> velocities;
  std::vector<tp::VectorXd> accelerations;
};

// sampled_trajectory.h:39-50
struct UniformlySampledTrajectory {
  double time_step_sec = 0.0;
  double start_time_sec = 0.0;
  std::vector<tp::VectorXd> positions;
  std::vector<tp::VectorXd> velocities;
  std::vector<tp::VectorXd> accelerations;
};

// sampled_trajectory.cc:24-44 (same messages)
inline tp::Status AreInputsValidForSampledTrajectory(tp::Span<const double> times,
                                                     tp::Span<const tp::VectorXd> positions,
                                                     tp::Span<const tp::VectorXd> velocities,
                                                     tp::Span<const tp::VectorXd> accelerations) {
  const int sample_count = static_cast<int>(times.size());
  if (sample_count < 2) return tp::InvalidArgumentError("Need at least two samples.");
  if (static_cast<int>(positions.size()) != sample_count ||
      static_cast<int>(velocities.size()) != sample_count ||
      static_cast<int>(accelerations.size()) != sample_count) {
    return tp::InvalidArgumentError("inconsistent sizes for samples.");
  }
  for (int i = 0; i < sample_count - 1; ++i) {
    if (times[i + 1] <= times[i]) {
      return tp::InvalidArgumentError("Time samples not strictly increasing.");
    }
  }
  return tp::OkStatus();
}

}  // namespace trajectory_planning

#endif  // TRAJECTORY_PLANNING_SAMPLED_TRAJECTORY_H_
