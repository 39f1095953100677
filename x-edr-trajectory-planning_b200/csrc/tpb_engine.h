// Device-resident state of one engine (one GPU, one stream) and the kernel
// launchers the C ABI calls.  Layout in HBM (B paths, N samples, D dofs,
// n <= max_points control points, C rows):
//
//   waypoints  [B][Wmax][D]      num_wp [B]          inputs
//   vmax, amax [B][D]                                inputs (broadcast if shared)
//   ctrl       [B][max_points][D] knots [B][max_points+3]  num_points [B]
//   q, dq, ddq [B][N][D]         SamplePath outputs (AoS like the reference's
//                                vector<VectorXd>; one 56-byte row per sample)
//   cons       [B][N][C][4]      explicit constraint rows (generic problems only;
//                                joint paths never materialise them)
//   b_sd2_max, b_sdd_max, b_sdd_min, b_sd2_zero [B][N], b_at0/b_type u8 [B][N]
//   sd2, sdd, sd, s, time [B][N] solver state / outputs
//   status, last_extremal i32 [B]; duration, delta, knot_end, s_start, s_end [B]
//   r_count i32 [B], r_offset i64 [B+1], r_* packed rows
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>

#include <cuda_runtime.h>

#include "../../include/tpb.h"

namespace tpb {

struct DeviceArrays {
  // inputs
  double* waypoints = nullptr;
  int32_t* num_wp = nullptr;
  double* vmax = nullptr;
  double* amax = nullptr;
  double* delta_in = nullptr;
  double* path_start = nullptr;
  double* sd_start = nullptr;
  double* sdd_start = nullptr;
  double* time_start = nullptr;
  // spline
  double* ctrl = nullptr;
  double* knots = nullptr;
  int32_t* num_points = nullptr;
  int32_t* knot_cap = nullptr;  // [B] the reference's knot capacity of the path's spline
  double* delta = nullptr;
  double* knot_end = nullptr;
  // samples
  double* q = nullptr;
  double* dq = nullptr;
  double* ddq = nullptr;
  // Cartesian paths: pose waypoints / control points (tx,ty,tz,qx,qy,qz,qw), the
  // targets SamplePath hands to the path IK, uploaded Jacobians
  double* pose_wp = nullptr;        // [B][Wmax][7]
  double* pose_ctrl = nullptr;      // [B][max_points][7]
  double* pose_targets = nullptr;   // [B][N+1][7]
  double* joint_targets = nullptr;  // [B][N+1][D]
  double* jac = nullptr;            // [B][N][6][D], allocated on first upload
  // generic constraints
  double* cons = nullptr;
  double* s_start = nullptr;
  double* s_end = nullptr;
  // limit curve: per-sample values as computed (raw) + the working copy that
  // CalculateBoundary's passes 2-4 mutate
  double* b_raw_sd2_max = nullptr;
  double* b_raw_sdd_max = nullptr;
  double* b_raw_sdd_min = nullptr;
  double* b_sd2_max = nullptr;
  double* b_sdd_max = nullptr;
  double* b_sdd_min = nullptr;
  double* b_sd2_zero = nullptr;
  uint8_t* b_at0 = nullptr;
  uint8_t* b_type = nullptr;
  // solver
  double* sd2 = nullptr;
  double* sdd = nullptr;
  double* sd = nullptr;
  double* s = nullptr;
  double* time = nullptr;
  double* fix_val = nullptr;  // [B][N] scratch for deferred boundary fixes
  int32_t* status = nullptr;
  int32_t* last_extremal = nullptr;
  double* duration = nullptr;
  // resampled
  int32_t* r_count = nullptr;
  int64_t* r_offset = nullptr;
  double* r_time = nullptr;
  double* r_pos = nullptr;
  double* r_vel = nullptr;
  double* r_acc = nullptr;
  double* r_s = nullptr;
  double* r_sd = nullptr;
  double* r_sdd = nullptr;
  // moving-horizon planner: only paths with active[b] != 0 are sampled / timed
  // (nullptr: all of them)
  const int32_t* active = nullptr;
};

// Per-path overrides of the resampling kernels, used by the moving-horizon planner
// (all nullptr: the plain fresh-path resampling of N samples from start_sec).
struct ResampleView {
  const int32_t* len = nullptr;         // number of path samples held
  const double* origin = nullptr;       // time of row 0 of the uniform grid / of row 0
  const int32_t* grid_skip = nullptr;   // uniform: rows of that grid erased in front
  const int32_t* first_from = nullptr;  // skip method: the lower-index scan of row 0 starts here
  const int32_t* keep_base = nullptr;   // skip method: kept-sample entries erased in front
};

// State of PathTimingTrajectory between Plan() calls for every path of the batch
// (tpb_planner.cu).  P_cap path samples per path.
struct PlannerArrays {
  int P_cap = 0;
  // *_at_path_samples_ (path_timing_trajectory.h): [B][P_cap], [B][P_cap][D]
  double *w_time = nullptr, *w_s = nullptr, *w_sd = nullptr, *w_sdd = nullptr;
  double *w_q = nullptr, *w_dq = nullptr, *w_ddq = nullptr;
  int32_t* keep = nullptr;  // [B][P_cap] sample index of trajectory row k + 1 (skip method)
  // [B]
  // sampled: TimeablePath::State of the path, 0 kNewPath, 1 kPathWasSampled, 2 kModifiedPath
  int32_t *len = nullptr, *active = nullptr, *status = nullptr, *mode = nullptr,
          *initial_plan = nullptr, *sampled = nullptr, *planned_to_end = nullptr,
          *target_reached = nullptr, *loops = nullptr, *offset = nullptr,
          *erase_rows = nullptr, *first_from = nullptr, *keep_base = nullptr;
  int64_t *start_time = nullptr, *end_time = nullptr, *final_decel_start = nullptr,
          *loop_start = nullptr;
  double *path_horizon = nullptr, *decel_time = nullptr, *origin = nullptr,
         *front_time = nullptr, *back_time = nullptr;
  double* init_vel = nullptr;     // [B][D]
  int32_t* any_active = nullptr;  // [1]
};

struct PlannerOptions {
  double dt = 0.0;  // time step
  bool skip = false;
  int max_planning_iterations = 0;
  double max_initial_velocity_error = 0.0;
};

struct Dims {
  int B = 0;        // current batch
  int Bmax = 0;
  int D = 0;
  int Wmax = 0;
  int max_points = 0;  // 3*Wmax-2 (>= 4)
  int N = 0;
  int Cmax = 0;  // capacity of explicit rows
  int C = 0;     // rows of the current problem (2*D for joint paths)
  int64_t r_cap = 0;
  // tpb_set_kernel_variant: bit k set = cross-check variant TPB_VARIANT_<k> is used
  // instead of the default kernel (host-side launch decisions only)
  uint32_t variants = 0;
};

enum class Mode { kNone, kJoint, kGeneric, kCartesian };

// tpb_query_profile (tpb_query.cu): Q query times per path, all pointers on the device,
// outputs [B][Q], any output may be null
struct ProfileQueryArgs {
  int Q = 0;
  bool use_s_range = false;  // the parameter range is s_start / s_end (explicit rows, Cartesian)
  const double* time = nullptr;
  double *s = nullptr, *sd = nullptr, *sdd = nullptr;
  int32_t* prev_index = nullptr;
  double *sk = nullptr, *sdk = nullptr, *sddk = nullptr, *tk = nullptr;
  int32_t* status = nullptr;
};
int launch_profile_query(const struct Dims& d, const struct DeviceArrays& a,
                         const ProfileQueryArgs& p, cudaStream_t st);
int launch_count_violations(const struct Dims& d, const struct DeviceArrays& a, bool joint,
                            double safety, int32_t* count, cudaStream_t st);

// ---- kernel launchers (tpb_kernels.cu); each returns the number of kernels
// it launched.
int launch_fill_f64(double* p, double v, int n, cudaStream_t st);
int launch_fill_i32(int32_t* p, int32_t v, int n, cudaStream_t st);
int launch_joint_fit(const Dims& d, const DeviceArrays& a, double rounding, bool auto_delta,
                     cudaStream_t st);
int launch_joint_sample_limit(const Dims& d, const DeviceArrays& a, double safety,
                              cudaStream_t st);
int launch_generic_limit(const Dims& d, const DeviceArrays& a, cudaStream_t st);
int launch_integrate(const Dims& d, const DeviceArrays& a, Mode mode, double safety,
                     int max_loops, cudaStream_t st,
                     cudaEvent_t* phase_events = nullptr);
// continued: the batch is a chunk following another one; r_offset[0] already
// holds the number of rows before it
// skip: ResampleSkippingSamplesCloserThanTimeStep instead of
// ResampleEquidistantlyInTime
int launch_resample(const Dims& d, const DeviceArrays& a, double start_sec, double dt,
                    int num_sms, cudaStream_t st, bool continued = false, bool skip = false);

// scan of r_count + row assembly only (the counts are already in r_count)
int launch_resample_rows(const Dims& d, const DeviceArrays& a, const ResampleView& v,
                         const int32_t* keep, double start_sec, double dt, int num_sms,
                         cudaStream_t st, bool skip, bool continued = false);

// Moving-horizon planner (tpb_planner.cu).  plan_begin .. plan_end bracket one
// Plan() call; window_begin / window_velocity / window_append bracket the
// sampling + timing of one planning window of the active paths.
int launch_planner_reset(const Dims& d, const PlannerArrays& w, cudaStream_t st);
int launch_planner_new_path(const Dims& d, const PlannerArrays& w, cudaStream_t st);
int launch_plan_begin(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                      const PlannerOptions& o, int64_t start_ns, int64_t horizon_ns,
                      cudaStream_t st);
int launch_window_begin(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                        int64_t start_ns, int64_t horizon_ns, cudaStream_t st);
int launch_window_velocity(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                           const PlannerOptions& o, cudaStream_t st);
int launch_window_append(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                         const PlannerOptions& o, int64_t start_ns, int64_t horizon_ns,
                         cudaStream_t st);
int launch_plan_rows(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                     const PlannerOptions& o, int64_t start_ns, cudaStream_t st);
int launch_plan_end(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                    const PlannerOptions& o, cudaStream_t st);

// Batched TimeableJointSplinePath::SwitchToWaypointPath (tpb_switch.cu); all pointers device.
struct SwitchArgs {
  const double* keep_path_until = nullptr;  // [B]
  const int32_t* num_waypoints = nullptr;   // [B] or nullptr (max_waypoints each); <= 0: path untouched
  const double* waypoints = nullptr;        // [B][max_waypoints][D]
  const double* pose_waypoints = nullptr;   // [B][max_waypoints][7] (Cartesian paths)
  int max_waypoints = 0;                    // <= Dims::Wmax
  int32_t* status = nullptr;                // [B] absl::StatusCode numbers
  int32_t* modified = nullptr;              // [B] 1: the path is TimeablePath::State::kModifiedPath now
};
int launch_joint_switch(const Dims& d, const DeviceArrays& a, const SwitchArgs& s, double rounding,
                        cudaStream_t st);
int launch_cart_switch(const Dims& d, const DeviceArrays& a, const SwitchArgs& s,
                       double rotation_rounding, double translation_rounding, cudaStream_t st);
int launch_restride_rows(int B, int width, int old_rows, int new_rows, const double* in, double* out,
                         cudaStream_t st);
int launch_restride_spline(int B, int D, int old_points, int new_points, const double* ctrl_in,
                           const double* knots_in, double* ctrl_out, double* knots_out,
                           cudaStream_t st);
// planner: paths with modified[b] != 0 become kModifiedPath
int launch_planner_mark_modified(const Dims& d, const PlannerArrays& w, const int32_t* modified,
                                 cudaStream_t st);

// Batched RescaleTrajectoryBackwardToStop (tpb_stop.cu); all pointers device.
struct StopArgs {
  int num_traj, D;
  const int64_t* begin;   // [num_traj] first row of trajectory b in the packed arrays
  const int32_t* length;  // [num_traj] its number of samples
  const double *times, *vel, *acc;  // [rows], [rows][D], [rows][D]
  const double* max_acc;            // [num_traj][D]
  int32_t *count, *switch_index, *status;  // [num_traj]
  double *out_times, *out_vel, *out_acc;   // layout of the inputs
};
int launch_rescale_to_stop(const StopArgs& p, cudaStream_t st);

// Batched GetPathStopParameter over packed resampled trajectories (tpb_stop.cu).
struct StopQueryArgs {
  int B, D;
  const int64_t* offset;                // [B+1] rows of trajectory b: offset[b] .. offset[b+1]-1
  const double *times, *s, *vel, *acc;  // [rows], [rows], [rows][D], [rows][D]
  const double* max_acc;                // [B][D]
  const double* query;                  // [B] query times
  double* parameter;                    // [B] path parameter at which the path can stop
  int32_t* row;                         // [B] its row within the trajectory
  double* duration;                     // [B] duration of the stop
  int32_t* status;                      // [B]
};
int launch_path_stop_parameter(const StopQueryArgs& p, cudaStream_t st);

// The arrays of paths [b0, ...) of a joint-space batch: every per-path pointer
// advanced by b0 paths (the packed r_* rows are addressed through r_offset and
// stay where they are).  Used to run the stage kernels on a chunk of the batch.
inline DeviceArrays shifted_joint(const DeviceArrays& a, const Dims& d, int b0) {
  DeviceArrays s = a;
  const size_t b = static_cast<size_t>(b0), N = d.N, D = d.D;
  auto adv = [&](auto*& p, size_t stride) {
    if (p) p += b * stride;
  };
  adv(s.waypoints, static_cast<size_t>(d.Wmax) * D);
  adv(s.num_wp, 1);
  adv(s.vmax, D);
  adv(s.amax, D);
  adv(s.delta_in, 1);
  adv(s.path_start, 1);
  adv(s.sd_start, 1);
  adv(s.sdd_start, 1);
  adv(s.time_start, 1);
  adv(s.ctrl, static_cast<size_t>(d.max_points) * D);
  adv(s.knots, static_cast<size_t>(d.max_points) + 3);
  adv(s.num_points, 1);
  adv(s.knot_cap, 1);
  adv(s.delta, 1);
  adv(s.knot_end, 1);
  adv(s.q, N * D);
  adv(s.dq, N * D);
  adv(s.ddq, N * D);
  adv(s.b_raw_sd2_max, N);
  adv(s.b_raw_sdd_max, N);
  adv(s.b_raw_sdd_min, N);
  adv(s.b_sd2_max, N);
  adv(s.b_sdd_max, N);
  adv(s.b_sdd_min, N);
  adv(s.b_sd2_zero, N);
  adv(s.b_at0, N);
  adv(s.b_type, N);
  adv(s.sd2, N);
  adv(s.sdd, N);
  adv(s.sd, N);
  adv(s.s, N);
  adv(s.time, N);
  adv(s.fix_val, N);
  adv(s.status, 1);
  adv(s.last_extremal, 1);
  adv(s.duration, 1);
  adv(s.r_count, 1);
  adv(s.r_offset, 1);
  adv(s.active, 1);
  return s;
}

// Cartesian path (tpb_cartesian.cu); -1 = unsupported built-in model
int launch_cart_fit(const Dims& d, const DeviceArrays& a, double rotation_rounding,
                    double translation_rounding, bool auto_delta, cudaStream_t st);
int launch_cart_sample_targets(const Dims& d, const DeviceArrays& a, cudaStream_t st);
int launch_cart_ik_builtin(const Dims& d, const DeviceArrays& a, int model, cudaStream_t st);
// rows AND the per-sample limit curve (fused for 6/7 dofs, else + k_generic_limit)
int launch_cart_constraints(const Dims& d, const DeviceArrays& a, double safety,
                            double max_trans_vel, double max_rot_vel, int model, bool have_jac,
                            cudaStream_t st);

}  // namespace tpb
