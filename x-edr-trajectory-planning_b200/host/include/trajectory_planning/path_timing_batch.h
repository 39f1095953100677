// Batched front-end: B independent fresh joint paths per call, i.e. what a user
// of TimeableJointSplinePath::SetWaypoints + PathTimingTrajectory::Plan
// (path_timing_trajectory.cc:579-684) does in a loop over paths, as one pass over one
// or several GPUs (tpb_multi_plan_joint_host: contiguous blocks of the batch per GPU,
// one host thread per shard, no collective).  Host buffers in; the results land in
// page-locked arrays the object owns and reuses (no allocation, no zero fill and no
// pageable staging copy in the steady state), and only the arrays asked for with
// set_wanted() cross PCIe: a caller that reads GetPositions() only moves 8 * D bytes per
// trajectory row instead of 8 * (3 D + 4), one that reads the TimeOptimalPathProfile
// getters 32 * N bytes per path and no rows at all.
// The resampled trajectories come back packed: path b owns rows
// [offset[b], offset[b+1]) of positions / velocities / accelerations / time.
#ifndef TRAJECTORY_PLANNING_PATH_TIMING_BATCH_H_
#define TRAJECTORY_PLANNING_PATH_TIMING_BATCH_H_

#include <cstdint>
#include <vector>

#include "trajectory_planning/path_timing_trajectory.h"

namespace trajectory_planning {

// A growth-only array in page-locked host memory (tpb_host_alloc).
template <typename T>
class PinnedArray {
 public:
  PinnedArray() = default;
  ~PinnedArray() { tpb_host_free(p_); }
  PinnedArray(const PinnedArray&) = delete;
  PinnedArray& operator=(const PinnedArray&) = delete;
  // Makes room for n elements (contents unspecified) and sets size() = n.
  bool resize(size_t n) {
    if (n > cap_) {
      tpb_host_free(p_);
      p_ = nullptr;
      cap_ = 0;
      void* q = nullptr;
      if (tpb_host_alloc(n * sizeof(T), &q) != TPB_OK) {
        n_ = 0;
        return false;
      }
      p_ = static_cast<T*>(q);
      cap_ = n;
    }
    n_ = n;
    return true;
  }
  void shrink(size_t n) { n_ = n < n_ ? n : n_; }
  size_t size() const { return n_; }
  size_t capacity() const { return cap_; }
  bool empty() const { return n_ == 0; }
  T* data() { return p_; }
  const T* data() const { return p_; }
  const T& operator[](size_t i) const { return p_[i]; }
  T& operator[](size_t i) { return p_[i]; }
  const T* begin() const { return p_; }
  const T* end() const { return p_ + n_; }
  std::vector<T> to_vector() const { return std::vector<T>(begin(), end()); }

 private:
  T* p_ = nullptr;
  size_t n_ = 0, cap_ = 0;
};

class PathTimingTrajectoryBatch {
 public:
  // What Plan() brings back to the host (TrajectoryPlanner getters,
  // trajectory_planner.h:86-108, and the TimeOptimalPathProfile getters).
  enum Want : uint32_t {
    kDuration = 1u,        // duration + status per path (always cheap)
    kPositions = 2u,       // GetPositions()
    kVelocities = 4u,      // GetVelocities()
    kAccelerations = 8u,   // GetAccelerations()
    kPathParameters = 16u, // GetPathParameters() + both derivatives
    kTime = 32u,           // GetTime()
    kProfile = 64u,        // the profile at the path samples: time, s, sd, sdd [B][N]
    kTrajectory = kDuration | kPositions | kVelocities | kAccelerations | kPathParameters,
  };
  struct Result {
    PinnedArray<double> duration;   // [B] (NaN where status != 0)
    PinnedArray<int32_t> status;    // [B] TPB_PATH_*
    PinnedArray<int64_t> offset;    // [B+1]
    PinnedArray<double> positions, velocities, accelerations;  // [rows][D]
    PinnedArray<double> path_parameter, path_parameter_derivative,
        second_path_parameter_derivative, time;  // [rows]
    PinnedArray<double> profile_time, profile_s, profile_sd, profile_sdd;  // [B][N]
    int64_t rows = 0;
  };

  // One GPU.
  PathTimingTrajectoryBatch(const JointPathOptions& path_options,
                            const PathTimingTrajectoryOptions& planner_options, int max_paths,
                            int max_waypoints, int64_t max_time_samples_total, int device = 0)
      : PathTimingTrajectoryBatch(path_options, planner_options, max_paths, max_waypoints,
                                  max_time_samples_total, std::vector<int32_t>{device}) {}
  // Several GPUs (or logical shards: an ordinal may repeat); max_paths and
  // max_time_samples_total are totals over all of them.
  PathTimingTrajectoryBatch(const JointPathOptions& path_options,
                            const PathTimingTrajectoryOptions& planner_options, int max_paths,
                            int max_waypoints, int64_t max_time_samples_total,
                            const std::vector<int32_t>& devices)
      : path_options_(path_options), planner_options_(planner_options),
        row_capacity_(max_time_samples_total) {
    tpb_config cfg{};
    cfg.max_paths = max_paths;
    cfg.num_dofs = static_cast<int32_t>(path_options.num_dofs());
    cfg.max_waypoints = max_waypoints;
    cfg.num_samples = static_cast<int32_t>(planner_options.GetNumPathSamples());
    cfg.max_time_samples_total = max_time_samples_total;
    create_rc_ = tpb_multi_create(&cfg, devices.data(), static_cast<int32_t>(devices.size()), &m_);
    if (create_rc_ == TPB_OK) {
      for (int g = 0; g < tpb_multi_num_shards(m_); ++g)
        tpb_set_path_options(shard_engine(g), path_options.constraint_safety(),
                             path_options.rounding());
    }
  }
  ~PathTimingTrajectoryBatch() {
    if (m_) tpb_multi_destroy(m_);
  }
  PathTimingTrajectoryBatch(const PathTimingTrajectoryBatch&) = delete;
  PathTimingTrajectoryBatch& operator=(const PathTimingTrajectoryBatch&) = delete;

  // Which outputs the next Plan() copies to the host (default: the whole trajectory).
  void set_wanted(uint32_t want_mask) { want_ = want_mask | kDuration; }
  uint32_t wanted() const { return want_; }

  // waypoints: [B][max_waypoints][D] (rows past num_waypoints[b] ignored);
  // max_velocity / max_acceleration: [B][D].  Every path is sampled over its whole
  // parameter range (delta = knots.back() / (N - 1)): one planning window per path.
  tp::Status Plan(int num_paths, const int32_t* num_waypoints, const double* waypoints,
                  const double* max_velocity, const double* max_acceleration, tp::Time start,
                  Result* out) {
    if (create_rc_ != TPB_OK)
      return tp::FromCode(create_rc_, "tpb_multi_create failed (no CUDA device?)");
    const size_t B = num_paths, D = path_options_.num_dofs();
    const size_t N = planner_options_.GetNumPathSamples();
    const size_t cap = static_cast<size_t>(row_capacity_);
    const bool rows_wanted =
        (want_ & (kPositions | kVelocities | kAccelerations | kPathParameters | kTime)) != 0;
    bool ok = out->duration.resize(B) && out->status.resize(B) && out->offset.resize(B + 1);
    auto size_rows = [&](PinnedArray<double>& a, bool wanted, size_t width) {
      ok = ok && a.resize(wanted ? cap * width : 0);
      return wanted ? a.data() : nullptr;
    };
    auto size_prof = [&](PinnedArray<double>& a) {
      const bool wanted = (want_ & kProfile) != 0;
      ok = ok && a.resize(wanted ? B * N : 0);
      return wanted ? a.data() : nullptr;
    };
    tpb_plan_io io{};
    io.num_paths = num_paths;
    io.num_waypoints = num_waypoints;
    io.waypoints = waypoints;
    io.max_velocity = max_velocity;
    io.max_acceleration = max_acceleration;
    io.start_sec = TimeToSec(start);
    io.time_step_sec =
        rows_wanted ? TimeToSec(tp::FromUnixNanos(planner_options_.GetTimeStep().ns)) : 0.0;
    io.r_pos = size_rows(out->positions, want_ & kPositions, D);
    io.r_vel = size_rows(out->velocities, want_ & kVelocities, D);
    io.r_acc = size_rows(out->accelerations, want_ & kAccelerations, D);
    io.r_s = size_rows(out->path_parameter, want_ & kPathParameters, 1);
    io.r_sd = size_rows(out->path_parameter_derivative, want_ & kPathParameters, 1);
    io.r_sdd = size_rows(out->second_path_parameter_derivative, want_ & kPathParameters, 1);
    io.r_time = size_rows(out->time, want_ & kTime, 1);
    io.profile_time = size_prof(out->profile_time);
    io.profile_s = size_prof(out->profile_s);
    io.profile_sd = size_prof(out->profile_sd);
    io.profile_sdd = size_prof(out->profile_sdd);
    if (!ok) return tp::FromCode(TPB_INTERNAL, "page-locked host allocation failed");
    io.duration = out->duration.data();
    io.status = out->status.data();
    io.r_offset = out->offset.data();
    io.r_capacity = row_capacity_;
    const int32_t rc = tpb_multi_plan_joint_host(m_, &io);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_multi_last_error(m_));
    out->rows = io.r_total;
    const size_t rows = static_cast<size_t>(io.r_total);
    if (!rows_wanted) {
      for (size_t b = 0; b <= B; ++b) out->offset[b] = 0;
    }
    out->positions.shrink(rows * D);
    out->velocities.shrink(rows * D);
    out->accelerations.shrink(rows * D);
    out->path_parameter.shrink(rows);
    out->path_parameter_derivative.shrink(rows);
    out->second_path_parameter_derivative.shrink(rows);
    out->time.shrink(rows);
    return tp::OkStatus();
  }

  int num_shards() const { return m_ ? tpb_multi_num_shards(m_) : 0; }
  // The engine of shard g (device-resident consumers: tpb_devptr, trajectory buffers).
  tpb_engine* shard_engine(int g) {
    tpb_engine* e = nullptr;
    if (m_) tpb_multi_shard(m_, g, nullptr, nullptr, nullptr, nullptr, nullptr, &e);
    return e;
  }
  tpb_engine* engine() { return shard_engine(0); }
  tpb_multi* multi() { return m_; }

 private:
  JointPathOptions path_options_;
  PathTimingTrajectoryOptions planner_options_;
  tpb_multi* m_ = nullptr;
  int32_t create_rc_ = TPB_OK;
  int64_t row_capacity_ = 0;
  uint32_t want_ = kTrajectory;
};

}  // namespace trajectory_planning
#endif  // TRAJECTORY_PLANNING_PATH_TIMING_BATCH_H_
