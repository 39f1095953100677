// Host-side mirror of trajectory_planning/rescale_to_stop.h:31-35: same function,
// argument order and error behaviour; the backward integration runs on the device
// (tpb_rescale_to_stop_host, csrc/tpb_stop.cu).  RescaleTrajectoriesBackwardToStop
// is the batched form (many trajectories in one launch).
#ifndef TRAJECTORY_PLANNING_RESCALE_TO_STOP_H_
#define TRAJECTORY_PLANNING_RESCALE_TO_STOP_H_

#include <cstdint>
#include <vector>

#include "tp/base.h"
#include "tpb.h"
#include "trajectory_planning/sampled_trajectory.h"

namespace trajectory_planning {

// One trajectory of a batch (views; all four spans have the same length).
struct SampledTrajectoryView {
  tp::Span<const double> times;
  tp::Span<const tp::VectorXd> positions;
  tp::Span<const tp::VectorXd> velocities;
  tp::Span<const tp::VectorXd> accelerations;
};

// Stopping trajectories for many trajectories at once.  max_accelerations holds
// one vector per trajectory (all of the same size).  results[b] is what
// RescaleTrajectoryBackwardToStop returns for trajectory b; statuses[b] its
// status.  The returned status is only non-OK when the call as a whole failed.
inline tp::Status RescaleTrajectoriesBackwardToStop(
    tp::Span<const tp::VectorXd> max_accelerations, tp::Span<const SampledTrajectoryView> batch,
    std::vector<SampledTrajectory>* results, std::vector<tp::Status>* statuses,
    int32_t device = 0) {
  const size_t T = batch.size();
  results->assign(T, SampledTrajectory{});
  statuses->assign(T, tp::OkStatus());
  if (T == 0) return tp::OkStatus();
  if (max_accelerations.size() != T) {
    return tp::InvalidArgumentError("one max_acceleration vector per trajectory expected");
  }
  const size_t D = max_accelerations[0].size();
  // trajectories that fail the reference's input check never reach the device
  std::vector<int64_t> begin;
  std::vector<int32_t> length;
  std::vector<size_t> which;
  int64_t rows = 0;
  for (size_t b = 0; b < T; ++b) {
    const auto& tr = batch[b];
    (*statuses)[b] = AreInputsValidForSampledTrajectory(tr.times, tr.positions, tr.velocities,
                                                         tr.accelerations);
    if (!(*statuses)[b].ok()) continue;
    if (max_accelerations[b].size() != D) {
      (*statuses)[b] = tp::InvalidArgumentError("max_acceleration sizes differ within the batch");
      continue;
    }
    which.push_back(b);
    begin.push_back(rows);
    length.push_back(static_cast<int32_t>(tr.times.size()));
    rows += static_cast<int64_t>(tr.times.size());
  }
  const size_t K = which.size();
  if (K == 0) return tp::OkStatus();
  std::vector<double> t(rows), v(rows * D), a(rows * D), amax(K * D);
  for (size_t k = 0; k < K; ++k) {
    const auto& tr = batch[which[k]];
    for (size_t j = 0; j < D; ++j) amax[k * D + j] = max_accelerations[which[k]][j];
    for (size_t i = 0; i < tr.times.size(); ++i) {
      const size_t r = static_cast<size_t>(begin[k]) + i;
      t[r] = tr.times[i];
      for (size_t j = 0; j < D; ++j) {
        v[r * D + j] = tr.velocities[i][j];
        a[r * D + j] = tr.accelerations[i][j];
      }
    }
  }
  std::vector<int32_t> count(K), sw(K), st(K);
  std::vector<double> ot(rows), ov(rows * D), oa(rows * D);
  tpb_stop_desc d{};
  d.num_trajectories = static_cast<int32_t>(K);
  d.num_dofs = static_cast<int32_t>(D);
  d.total_rows = rows;
  d.begin = begin.data();
  d.length = length.data();
  d.times = t.data();
  d.velocities = v.data();
  d.accelerations = a.data();
  d.max_acceleration = amax.data();
  d.count = count.data();
  d.switch_index = sw.data();
  d.status = st.data();
  d.out_times = ot.data();
  d.out_velocities = ov.data();
  d.out_accelerations = oa.data();
  const int32_t rc = tpb_rescale_to_stop_host(&d, device);
  if (rc != TPB_OK) return tp::FromCode(rc, "tpb_rescale_to_stop_host failed");
  for (size_t k = 0; k < K; ++k) {
    const size_t b = which[k];
    if (st[k] != TPB_PATH_OK) {
      (*statuses)[b] = tp::InvalidArgumentError("Time samples not strictly increasing.");
      continue;
    }
    SampledTrajectory& r = (*results)[b];
    const size_t m = static_cast<size_t>(count[k]);
    const size_t first = static_cast<size_t>(begin[k]) + static_cast<size_t>(sw[k]);
    r.times.assign(ot.begin() + first, ot.begin() + first + m);
    for (size_t i = 0; i < m; ++i) {
      r.positions.push_back(batch[b].positions[static_cast<size_t>(sw[k]) + i]);
      r.velocities.emplace_back(ov.data() + (first + i) * D, D);
      r.accelerations.emplace_back(oa.data() + (first + i) * D, D);
    }
  }
  return tp::OkStatus();
}

// rescale_to_stop.h:31-35.  Returns a SampledTrajectory that ends in a stop and
// traces the tail of `positions` (empty if the input already ends in a stop), or
// the status of AreInputsValidForSampledTrajectory.
inline tp::StatusOr<SampledTrajectory> RescaleTrajectoryBackwardToStop(
    const tp::VectorXd& max_acceleration, tp::Span<const double> times,
    tp::Span<const tp::VectorXd> positions, tp::Span<const tp::VectorXd> velocities,
    tp::Span<const tp::VectorXd> accelerations) {
  const SampledTrajectoryView view{times, positions, velocities, accelerations};
  std::vector<SampledTrajectory> results;
  std::vector<tp::Status> statuses;
  const tp::Status call = RescaleTrajectoriesBackwardToStop(
      tp::Span<const tp::VectorXd>(&max_acceleration, 1),
      tp::Span<const SampledTrajectoryView>(&view, 1), &results, &statuses);
  if (!call.ok()) return call;
  if (!statuses[0].ok()) return statuses[0];
  return results[0];
}

}  // namespace trajectory_planning

#endif  // TRAJECTORY_PLANNING_RESCALE_TO_STOP_H_
