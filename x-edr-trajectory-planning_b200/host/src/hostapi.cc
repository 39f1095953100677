// C entry points over the host C++ class PathTimingTrajectoryBatch, so that a harness
// without a C++ compiler at hand (bench.py, the tests: ctypes) drives -- and times --
// exactly the call a C++ user of the drop-in makes: PathTimingTrajectoryBatch::Plan.
// Nothing here adds logic; see trajectory_planning/path_timing_batch.h.
#include <cstdint>
#include <string>
#include <vector>

#include "trajectory_planning/path_timing_batch.h"

using trajectory_planning::JointPathOptions;
using trajectory_planning::PathTimingTrajectoryBatch;
using trajectory_planning::PathTimingTrajectoryOptions;

namespace {
struct Handle {
  PathTimingTrajectoryBatch batch;
  PathTimingTrajectoryBatch::Result result;
  std::string error;
  Handle(const JointPathOptions& po, const PathTimingTrajectoryOptions& to, int max_paths,
         int max_waypoints, int64_t max_rows, const std::vector<int32_t>& devices)
      : batch(po, to, max_paths, max_waypoints, max_rows, devices) {}
};
}  // namespace

extern "C" {

void* tpbh_batch_create(int32_t num_dofs, int32_t num_path_samples, int64_t time_step_ns,
                        double constraint_safety, double rounding, int32_t max_paths,
                        int32_t max_waypoints, int64_t max_time_samples_total,
                        const int32_t* devices, int32_t num_devices) {
  JointPathOptions po;
  po.set_num_dofs(num_dofs).set_constraint_safety(constraint_safety).set_rounding(rounding);
  PathTimingTrajectoryOptions to;
  to.SetTimeStep(tp::Nanoseconds(time_step_ns)).SetNumDofs(num_dofs);
  to.SetNumPathSamples(num_path_samples);
  return new Handle(po, to, max_paths, max_waypoints, max_time_samples_total,
                    std::vector<int32_t>(devices, devices + num_devices));
}

void tpbh_batch_destroy(void* h) { delete static_cast<Handle*>(h); }

// PathTimingTrajectoryBatch::set_wanted + Plan; returns the status code.
int32_t tpbh_batch_plan(void* hv, uint32_t want_mask, int32_t num_paths,
                        const int32_t* num_waypoints, const double* waypoints,
                        const double* max_velocity, const double* max_acceleration,
                        int64_t start_ns) {
  auto* h = static_cast<Handle*>(hv);
  h->batch.set_wanted(want_mask);
  const tp::Status st = h->batch.Plan(num_paths, num_waypoints, waypoints, max_velocity,
                                      max_acceleration, tp::FromUnixNanos(start_ns), &h->result);
  h->error = st.message();
  return static_cast<int32_t>(st.code());
}

const char* tpbh_batch_error(void* hv) { return static_cast<Handle*>(hv)->error.c_str(); }
int64_t tpbh_batch_rows(void* hv) { return static_cast<Handle*>(hv)->result.rows; }

// Address and element count of a result array: 0 duration, 1 status (int32), 2 offset
// (int64), 3 positions, 4 velocities, 5 accelerations, 6 path_parameter, 7 its first and
// 8 second derivative, 9 time, 10..13 profile time / s / sd / sdd.
const void* tpbh_batch_array(void* hv, int32_t which, int64_t* count) {
  auto& r = static_cast<Handle*>(hv)->result;
  const trajectory_planning::PinnedArray<double>* a[] = {
      &r.duration, nullptr, nullptr, &r.positions, &r.velocities, &r.accelerations,
      &r.path_parameter, &r.path_parameter_derivative, &r.second_path_parameter_derivative,
      &r.time, &r.profile_time, &r.profile_s, &r.profile_sd, &r.profile_sdd};
  *count = 0;
  if (which == 1) {
    *count = static_cast<int64_t>(r.status.size());
    return r.status.data();
  }
  if (which == 2) {
    *count = static_cast<int64_t>(r.offset.size());
    return r.offset.data();
  }
  if (which < 0 || which > 13) return nullptr;
  *count = static_cast<int64_t>(a[which]->size());
  return a[which]->data();
}

}  // extern "C"
