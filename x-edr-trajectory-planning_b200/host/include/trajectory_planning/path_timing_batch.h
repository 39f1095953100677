// Batched front-end: B independent fresh joint paths per call, i.e. what a user
// of TimeableJointSplinePath::SetWaypoints + PathTimingTrajectory::Plan does in
// a loop over paths, as one GPU pass (tpb_plan_joint_host).  Host buffers in and
// out; the resampled trajectories come back packed (path b owns rows
// [offset[b], offset[b+1]) of positions/velocities/accelerations).
#ifndef TRAJECTORY_PLANNING_PATH_TIMING_BATCH_H_
#define TRAJECTORY_PLANNING_PATH_TIMING_BATCH_H_

#include <vector>

#include "trajectory_planning/path_timing_trajectory.h"

namespace trajectory_planning {

class PathTimingTrajectoryBatch {
 public:
  struct Result {
    std::vector<double> duration;   // [B] (NaN where status != 0)
    std::vector<int32_t> status;    // [B] TPB_PATH_*
    std::vector<int64_t> offset;    // [B+1]
    std::vector<double> positions, velocities, accelerations;  // [rows][D]
    std::vector<double> path_parameter, path_parameter_derivative,
        second_path_parameter_derivative;  // [rows]
  };

  PathTimingTrajectoryBatch(const JointPathOptions& path_options,
                            const PathTimingTrajectoryOptions& planner_options, int max_paths,
                            int max_waypoints, int64_t max_time_samples_total, int device = 0)
      : path_options_(path_options), planner_options_(planner_options) {
    tpb_config cfg{};
    cfg.device = device;
    cfg.max_paths = max_paths;
    cfg.num_dofs = static_cast<int32_t>(path_options.num_dofs());
    cfg.max_waypoints = max_waypoints;
    cfg.num_samples = static_cast<int32_t>(planner_options.GetNumPathSamples());
    cfg.max_time_samples_total = max_time_samples_total;
    max_waypoints_ = max_waypoints;
    create_rc_ = tpb_create(&cfg, &e_);
    if (create_rc_ == TPB_OK) tpb_set_path_options(e_, path_options.constraint_safety(), path_options.rounding());
  }
  ~PathTimingTrajectoryBatch() { if (e_) tpb_destroy(e_); }
  PathTimingTrajectoryBatch(const PathTimingTrajectoryBatch&) = delete;
  PathTimingTrajectoryBatch& operator=(const PathTimingTrajectoryBatch&) = delete;

  // waypoints: [B][max_waypoints][D] (rows past num_waypoints[b] ignored);
  // max_velocity / max_acceleration: [B][D].  use_path_delta = false samples each
  // path over its whole parameter range (delta = knots.back()/(N-1)).
  tp::Status Plan(int num_paths, const int32_t* num_waypoints, const double* waypoints,
                  const double* max_velocity, const double* max_acceleration, tp::Time start,
                  Result* out) {
    if (create_rc_ != TPB_OK) return tp::FromCode(create_rc_, "tpb_create failed (no CUDA device?)");
    const size_t B = num_paths, D = path_options_.num_dofs();
    out->duration.assign(B, 0.0);
    out->status.assign(B, 0);
    out->offset.assign(B + 1, 0);
    tpb_plan_io io{};
    io.num_paths = num_paths;
    io.num_waypoints = num_waypoints;
    io.waypoints = waypoints;
    io.max_velocity = max_velocity;
    io.max_acceleration = max_acceleration;
    io.start_sec = TimeToSec(start);
    io.time_step_sec = TimeToSec(tp::FromUnixNanos(planner_options_.GetTimeStep().ns));
    io.duration = out->duration.data();
    io.status = out->status.data();
    io.r_offset = out->offset.data();
    // first pass sizes the packed output, second call is avoided by reading the
    // rows straight from the engine after the plan
    int32_t rc = tpb_plan_joint_host(e_, &io);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    const size_t rows = static_cast<size_t>(io.r_total);
    out->positions.resize(rows * D);
    out->velocities.resize(rows * D);
    out->accelerations.resize(rows * D);
    out->path_parameter.resize(rows);
    out->path_parameter_derivative.resize(rows);
    out->second_path_parameter_derivative.resize(rows);
    if (rows > 0) {
      tpb_get(e_, TPB_ARR_R_POS, out->positions.data(), rows * D * sizeof(double), 0);
      tpb_get(e_, TPB_ARR_R_VEL, out->velocities.data(), rows * D * sizeof(double), 0);
      tpb_get(e_, TPB_ARR_R_ACC, out->accelerations.data(), rows * D * sizeof(double), 0);
      tpb_get(e_, TPB_ARR_R_S, out->path_parameter.data(), rows * sizeof(double), 0);
      tpb_get(e_, TPB_ARR_R_SD, out->path_parameter_derivative.data(), rows * sizeof(double), 0);
      tpb_get(e_, TPB_ARR_R_SDD, out->second_path_parameter_derivative.data(), rows * sizeof(double), 0);
    }
    return tp::OkStatus();
  }
  tpb_engine* engine() { return e_; }

 private:
  JointPathOptions path_options_;
  PathTimingTrajectoryOptions planner_options_;
  tpb_engine* e_ = nullptr;
  int32_t create_rc_ = TPB_OK;
  int max_waypoints_ = 0;
};

}  // namespace trajectory_planning
#endif  // TRAJECTORY_PLANNING_PATH_TIMING_BATCH_H_
