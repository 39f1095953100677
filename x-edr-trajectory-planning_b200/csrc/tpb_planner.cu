// Moving-horizon PathTimingTrajectory::Plan for every path of the batch, with the
// planner's state kept on the device between calls
// (reference: trajectory_planning/path_timing_trajectory.cc:307-684).
//
// One Plan() call is
//   plan_begin                 HandleTimeArguments, UpdatePathTrackingStatus, the
//                              "already planned enough" erase, GetTimeOffsetAfter
//   repeat while a path is active (host loop, one 4-byte flag read per window):
//     window_begin             start of the window: sample before the loop start time
//     [joint sample + limits, window_velocity, integrate: the hot-path kernels,
//      masked to the active paths]
//     window_append            splice the window into the *_at_path_samples_ arrays,
//                              loop bookkeeping (:635-658)
//   plan_rows                  row counts (uniform grid / kept samples / erase)
//   [row assembly: k_resample_fill over the spliced arrays]
//   plan_end                   end_time_, final_decel_start_, target_reached_
//
// The bookkeeping kernels are one thread (or one warp) per path: they move a few
// dozen bytes of state per path, the window copy is a plain strided memcpy.
#include <cfloat>
#include <cmath>
#include <cstdint>

#include "tpb_engine.h"

namespace tpb {
namespace {

constexpr unsigned kFullMask = 0xffffffffu;
// w.mode: what plan_rows does with the trajectory rows of the path
constexpr int kModeResample = 0;  // rows are recomputed from the start time
constexpr int kModeKeep = 1;      // rows stay as they are (erased in front or untouched)
constexpr int kModeNone = 2;      // no rows (failed while planning)

// time.h:23-30
__device__ __forceinline__ int64_t time_from_sec(double seconds) {
  return static_cast<int64_t>(seconds * 1000000000.0);
}
__device__ __forceinline__ double time_to_sec(int64_t ns) {
  return static_cast<double>(ns) / 1000000000.0;
}
// ClampToTimeStepMultiple (:229-233)
__device__ __forceinline__ int64_t clamp_to_time_step_multiple(int64_t t, double dt) {
  const int64_t loop_multiple = static_cast<int64_t>(round(time_to_sec(t) / dt));
  return time_from_sec(loop_multiple * dt);
}

// ResetDerived (:214-227)
__global__ void k_planner_reset(Dims dm, PlannerArrays w) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B) return;
  w.len[b] = 0;
  w.active[b] = 0;
  w.status[b] = TPB_PATH_NOT_RUN;
  w.mode[b] = kModeNone;
  w.initial_plan[b] = 0;
  w.sampled[b] = 0;
  w.planned_to_end[b] = 1;
  w.target_reached[b] = 0;
  w.loops[b] = 0;
  w.offset[b] = 0;
  w.erase_rows[b] = 0;
  w.first_from[b] = 0;
  w.keep_base[b] = 0;
  w.start_time[b] = w.end_time[b] = w.loop_start[b] = 0;
  w.final_decel_start[b] = time_from_sec(0.0);
  w.path_horizon[b] = w.decel_time[b] = w.origin[b] = w.front_time[b] = w.back_time[b] = 0.0;
}

// SetWaypoints on a planner that already holds a plan: every path is kNewPath again,
// the planner keeps its trajectory and times.
__global__ void k_planner_new_path(Dims dm, PlannerArrays w) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < dm.B) w.sampled[b] = 0;
}

__global__ void k_planner_mark_modified(Dims dm, PlannerArrays w, const int32_t* modified) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < dm.B && modified[b] != 0) w.sampled[b] = 2;
}

// time of trajectory row i of path b (the rows themselves may lie past r_capacity)
__device__ __forceinline__ double row_time(const PlannerArrays& w, const PlannerOptions& o, int b,
                                           int i) {
  if (!o.skip) return w.origin[b] + o.dt * (i + w.erase_rows[b]);
  if (i == 0) return w.origin[b];
  const size_t base = static_cast<size_t>(b) * w.P_cap;
  return w.w_time[base + w.keep[base + w.keep_base[b] + i - 1]];
}

// Plan (:579-626) up to the planning loop, one thread per path.
__global__ void k_plan_begin(Dims dm, DeviceArrays a, PlannerArrays w, PlannerOptions o,
                             int64_t start, int64_t horizon) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B) return;
  const double start_sec = time_to_sec(start);
  w.loops[b] = 0;
  w.active[b] = 0;
  w.status[b] = TPB_PATH_OK;
  w.mode[b] = kModeKeep;
  const bool initial = w.initial_plan[b] != 0;
  // HandleTimeArguments (:503-538); absl::Seconds(double) counts quarter nanoseconds
  if (initial && 4 * start > 4 * w.end_time[b] + llround(o.dt * 4e9)) {
    w.status[b] = TPB_PATH_FAIL_START_RANGE;
    return;
  }
  if (!initial) {
    w.start_time[b] = start;
    w.end_time[b] = start;
  } else {
    if (start > w.end_time[b] || start < w.start_time[b]) {
      w.status[b] = TPB_PATH_FAIL_START_ARG;
      return;
    }
    w.start_time[b] = start;
  }
  // UpdatePathTrackingStatus (:477-501)
  int target_reached = 0, planned_to_end = 0;
  const bool sampled = w.sampled[b] == 1;  // neither kNewPath nor kModifiedPath
  if (!initial) {
    w.path_horizon[b] = 0.0;
  } else {
    planned_to_end = w.path_horizon[b] >= a.knot_end[b] - 1e-4;  // CloseToEnd
    if (planned_to_end) {
      if (sampled) {
        target_reached = 1;
      } else {
        w.path_horizon[b] = 0.0;
        planned_to_end = 0;
      }
    }
  }
  w.target_reached[b] = target_reached;
  w.planned_to_end[b] = planned_to_end;
  const bool planned_enough = sampled && (w.final_decel_start[b] >= start + horizon);
  const int M = a.r_count[b];
  if (M > 0 && planned_enough) {
    // EraseTrajectoryBefore (:540-577)
    if (start_sec < w.front_time[b]) return;
    if (o.skip) {
      // GetSampleCountUntil (:41-44): rows with a time < start_sec
      int lo = 0, hi = M;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (row_time(w, o, b, mid) < start_sec) {
          lo = mid + 1;
        } else {
          hi = mid;
        }
      }
      const int below = lo;
      // the reference reads time_[below] unchecked; below < M when enough is planned
      const int idx = min(below, M - 1);
      const int until =
          (row_time(w, o, b, idx) < start_sec + 0.95 * o.dt) ? below : below - 1;
      // row 0 becomes InterpolateAtTime(start_sec, below): the scan of the PATH
      // samples starts at the number of TRAJECTORY rows before start_sec (:551-552)
      w.first_from[b] = below;
      w.origin[b] = start_sec;
      w.keep_base[b] += max(until, 0);
      a.r_count[b] = M - max(until, 0);
    } else {
      const int offset = min(static_cast<int>(round((start_sec - w.front_time[b]) / o.dt)), M - 1);
      w.erase_rows[b] += offset;
      a.r_count[b] = M - offset;
    }
    w.front_time[b] = row_time(w, o, b, 0);
    return;
  }
  if (initial) {  // GetTimeOffsetAfter (:289-305)
    if (M == 0) {
      w.status[b] = TPB_PATH_FAIL_NO_SAMPLES;
      return;
    }
    if (start_sec < w.front_time[b]) {
      w.status[b] = TPB_PATH_FAIL_START_RANGE;
      return;
    }
    if (!(w.back_time[b] > start_sec)) {
      w.status[b] = TPB_PATH_FAIL_START_END;
      return;
    }
  }
  w.mode[b] = kModeResample;
  w.loop_start[b] = start;
  w.active[b] = planned_to_end ? 0 : 1;
  if (!planned_to_end) atomicOr(w.any_active, 1);
}

// ComputeTimingProfile (:307-340): where the window starts.
__global__ void k_window_begin(Dims dm, DeviceArrays a, PlannerArrays w, int64_t start,
                               int64_t horizon) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B || w.active[b] == 0) return;
  const int64_t loop_start = w.loop_start[b];
  if (start + horizon - loop_start <= 0) {  // target_duration (:314-318)
    w.status[b] = TPB_PATH_FAIL_START_ARG;
    w.mode[b] = kModeNone;
    w.active[b] = 0;
    return;
  }
  const double start_sec = time_to_sec(loop_start);
  int offset = 0;
  double path_start = 0.0, sd_start = 0.0, time_start = start_sec;
  if (w.sampled[b] != 0 && w.len[b] > 0) {  // kPathWasSampled or kModifiedPath
    // clamp(lower_bound(time_at_path_samples_, start) - begin - 1, 0, n - 1) (:334-338)
    const size_t base = static_cast<size_t>(b) * w.P_cap;
    const int len = w.len[b];
    int lo = 0, hi = len;
    while (lo < hi) {
      const int mid = lo + (hi - lo) / 2;
      if (w.w_time[base + mid] < start_sec) {
        lo = mid + 1;
      } else {
        hi = mid;
      }
    }
    offset = max(0, min(lo - 1, len - 1));
    path_start = w.w_s[base + offset];
    sd_start = w.w_sd[base + offset];
    time_start = w.w_time[base + offset];
  }
  w.offset[b] = offset;
  w.path_horizon[b] = path_start + a.delta[b] * (dm.N - 1);
  a.path_start[b] = path_start;
  a.sd_start[b] = sd_start;
  a.sdd_start[b] = 0.0;
  a.time_start[b] = time_start;
}

// The initial path velocity of a new path (:352-385): projection of the initial joint
// velocity onto the first path tangent.  Runs after the window was sampled.
__global__ void k_window_velocity(Dims dm, DeviceArrays a, PlannerArrays w, PlannerOptions o) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B || w.active[b] == 0) return;
  if (w.sampled[b] == 1) return;  // only new and modified paths
  w.sampled[b] = 1;
  const int D = dm.D;
  const double* d0 = a.dq + static_cast<size_t>(b) * dm.N * D;  // GetFirstPathDerivativeAt(0)
  const double* v0 = w.init_vel + static_cast<size_t>(b) * D;
  double sq = 0.0, dot = 0.0;
  for (int d = 0; d < D; ++d) {
    sq += d0[d] * d0[d];
    dot += v0[d] * d0[d];
  }
  double sd_start = a.sd_start[b];  // 0 for a new path, the previous sample's for a modified one
  if (sq > 100 * DBL_EPSILON) sd_start = fmax(dot / sq, 0.0);
  double err = 0.0;
  for (int d = 0; d < D; ++d) err = fmax(err, fabs(d0[d] * sd_start - v0[d]));
  if (err > o.max_initial_velocity_error) {
    w.status[b] = TPB_PATH_FAIL_INITIAL_VELOCITY;
    w.mode[b] = kModeNone;
    w.active[b] = 0;
    return;
  }
  a.sd_start[b] = sd_start;
}

// Erase the replanned samples and append the window (:417-472), then the loop
// bookkeeping of Plan (:635-658).  One CTA per path.
constexpr int kAppendThreads = 256;
__global__ void __launch_bounds__(kAppendThreads)
k_window_append(Dims dm, DeviceArrays a, PlannerArrays w, PlannerOptions o, int64_t start,
                int64_t horizon) {
  const int b = blockIdx.x;
  if (w.active[b] == 0) return;  // uniform per CTA
  const int N = dm.N, D = dm.D;
  const int solver_status = a.status[b];
  const int offset = w.offset[b];
  int fail = TPB_PATH_OK;
  if (solver_status != TPB_PATH_OK) {
    fail = solver_status;
  } else if (offset + N > w.P_cap) {
    fail = TPB_PATH_FAIL_CAPACITY;
  }
  if (fail != TPB_PATH_OK) {
    if (threadIdx.x == 0) {
      w.status[b] = fail;
      w.mode[b] = kModeNone;
      w.active[b] = 0;
    }
    return;
  }
  const size_t src = static_cast<size_t>(b) * N;
  const size_t dst = static_cast<size_t>(b) * w.P_cap + offset;
  for (int i = threadIdx.x; i < N; i += blockDim.x) {
    w.w_time[dst + i] = a.time[src + i];
    w.w_s[dst + i] = a.s[src + i];
    w.w_sd[dst + i] = a.sd[src + i];
    w.w_sdd[dst + i] = a.sdd[src + i];
  }
  for (int i = threadIdx.x; i < N * D; i += blockDim.x) {
    w.w_q[dst * D + i] = a.q[src * D + i];
    w.w_dq[dst * D + i] = a.dq[src * D + i];
    w.w_ddq[dst * D + i] = a.ddq[src * D + i];
  }
  if (threadIdx.x != 0) return;
  const double* tm = a.time + src;
  w.len[b] = offset + N;
  const int last_extremal = a.last_extremal[b];
  w.decel_time[b] = tm[last_extremal];
  // decel start clamped into the last half of the window (:638-645)
  const int decel_start = max(last_extremal, N / 2);
  const int64_t final_decel_start = time_from_sec(tm[decel_start]);
  w.final_decel_start[b] = final_decel_start;
  const int planned_to_end = w.path_horizon[b] >= a.knot_end[b] - 1e-4;
  w.planned_to_end[b] = planned_to_end;
  const bool horizon_reached =
      (tm[N - 1] - time_to_sec(start)) > (static_cast<double>(horizon) / 1e9);
  const int loop = w.loops[b];
  w.loops[b] = loop + 1;
  if (loop >= o.max_planning_iterations) {  // kDeadlineExceeded (:652-655)
    w.status[b] = TPB_PATH_FAIL_DEADLINE;
    w.mode[b] = kModeNone;
    w.active[b] = 0;
    return;
  }
  w.loop_start[b] = final_decel_start;
  const int again = (!planned_to_end && !horizon_reached) ? 1 : 0;
  w.active[b] = again;
  if (again) atomicOr(w.any_active, 1);
}

// Row counts of ResampleTrajectory (:697-707) over the spliced samples: one warp per
// path.  Uniform: ceil(duration / step) + 1 (:755-757).  Skip method: the greedy scan
// of :818-831 (32 samples per trip; the first lane that passes is kept, the others
// are re-tested against it).
__global__ void k_plan_rows(Dims dm, DeviceArrays a, PlannerArrays w, PlannerOptions o,
                            int64_t start) {
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (b >= dm.B) return;
  const int mode = w.mode[b];
  if (mode == kModeKeep) return;
  if (mode == kModeNone) {
    if (lane == 0) a.r_count[b] = 0;
    return;
  }
  const int len = w.len[b];
  if (len == 0) {  // nothing was ever planned: the reference would read back() of empty vectors
    if (lane == 0) {
      a.r_count[b] = 0;
      w.status[b] = TPB_PATH_FAIL_NO_SAMPLES;
    }
    return;
  }
  const double start_sec = time_to_sec(start);
  const size_t base = static_cast<size_t>(b) * w.P_cap;
  const double* tp = w.w_time + base;
  if (!o.skip) {
    if (lane == 0) {
      const double duration = tp[len - 1] - start_sec;
      const double c = ceil(duration / o.dt) + 1;
      a.r_count[b] = (c > 0 && c < 2.0e9) ? static_cast<int>(c) : 0;
      w.origin[b] = start_sec;
      w.erase_rows[b] = 0;
    }
    return;
  }
  int32_t* kp = w.keep + base;
  // TimeAtPathSamplesLowerIndex(0, start_sec) (:686-695)
  int li = len - 1;
  for (int i0 = 0; i0 < len - 1; i0 += 32) {
    const int i = i0 + lane;
    const unsigned m = __ballot_sync(kFullMask, (i < len - 1) && (tp[i + 1] > start_sec));
    if (m) {
      li = i0 + (__ffs(m) - 1);
      break;
    }
  }
  const double min_delta = 0.95 * o.dt;  // GetMinTimeDeltaToKeep (:893-900)
  double last = start_sec;
  int rows = 1;
  for (int i0 = li + 1; i0 < len; i0 += 32) {
    const int i = i0 + lane;
    const double t = (i < len) ? tp[i] : 0.0;
    unsigned cand = __ballot_sync(kFullMask, (i < len) && !(fabs(t - last) < min_delta));
    while (cand) {
      const int k = __ffs(cand) - 1;
      if (lane == 0) kp[rows - 1] = i0 + k;
      ++rows;
      last = __shfl_sync(kFullMask, t, k);
      cand = __ballot_sync(kFullMask, (i < len) && (lane > k) && !(fabs(t - last) < min_delta));
    }
  }
  if (lane == 0) {
    a.r_count[b] = rows;
    w.origin[b] = start_sec;
    w.first_from[b] = 0;
    w.keep_base[b] = 0;
  }
}

// The end of Plan (:660-676), and the times of the first / last trajectory row for
// the next call's argument checks.
__global__ void k_plan_end(Dims dm, DeviceArrays a, PlannerArrays w, PlannerOptions o) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B) return;
  const int M = a.r_count[b];
  if (M > 0) {
    w.front_time[b] = row_time(w, o, b, 0);
    w.back_time[b] = row_time(w, o, b, M - 1);
  }
  a.status[b] = w.status[b];
  if (w.mode[b] != kModeResample || w.status[b] != TPB_PATH_OK) return;
  w.initial_plan[b] = 1;
  if (M > 0) {
    w.end_time[b] = clamp_to_time_step_multiple(time_from_sec(w.back_time[b]), o.dt);
    w.final_decel_start[b] = clamp_to_time_step_multiple(time_from_sec(w.decel_time[b]), o.dt);
  } else {
    w.end_time[b] = w.start_time[b];
    w.final_decel_start[b] = w.end_time[b];
  }
  w.target_reached[b] = w.planned_to_end[b];
}

inline int blocks_for(int n, int threads) { return (n + threads - 1) / threads; }

}  // namespace

int launch_planner_reset(const Dims& d, const PlannerArrays& w, cudaStream_t st) {
  k_planner_reset<<<blocks_for(d.B, 128), 128, 0, st>>>(d, w);
  return 1;
}
int launch_planner_new_path(const Dims& d, const PlannerArrays& w, cudaStream_t st) {
  k_planner_new_path<<<blocks_for(d.B, 128), 128, 0, st>>>(d, w);
  return 1;
}
int launch_planner_mark_modified(const Dims& d, const PlannerArrays& w, const int32_t* modified,
                                 cudaStream_t st) {
  k_planner_mark_modified<<<blocks_for(d.B, 128), 128, 0, st>>>(d, w, modified);
  return 1;
}
int launch_plan_begin(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                      const PlannerOptions& o, int64_t start_ns, int64_t horizon_ns,
                      cudaStream_t st) {
  k_plan_begin<<<blocks_for(d.B, 128), 128, 0, st>>>(d, a, w, o, start_ns, horizon_ns);
  return 1;
}
int launch_window_begin(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                        int64_t start_ns, int64_t horizon_ns, cudaStream_t st) {
  k_window_begin<<<blocks_for(d.B, 128), 128, 0, st>>>(d, a, w, start_ns, horizon_ns);
  return 1;
}
int launch_window_velocity(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                           const PlannerOptions& o, cudaStream_t st) {
  k_window_velocity<<<blocks_for(d.B, 128), 128, 0, st>>>(d, a, w, o);
  return 1;
}
int launch_window_append(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                         const PlannerOptions& o, int64_t start_ns, int64_t horizon_ns,
                         cudaStream_t st) {
  k_window_append<<<d.B, kAppendThreads, 0, st>>>(d, a, w, o, start_ns, horizon_ns);
  return 1;
}
int launch_plan_rows(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                     const PlannerOptions& o, int64_t start_ns, cudaStream_t st) {
  k_plan_rows<<<blocks_for(d.B * 32, 128), 128, 0, st>>>(d, a, w, o, start_ns);
  return 1;
}
int launch_plan_end(const Dims& d, const DeviceArrays& a, const PlannerArrays& w,
                    const PlannerOptions& o, cudaStream_t st) {
  k_plan_end<<<blocks_for(d.B, 128), 128, 0, st>>>(d, a, w, o);
  return 1;
}

}  // namespace tpb
