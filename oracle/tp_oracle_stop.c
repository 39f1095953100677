/* TEST INFRASTRUCTURE ONLY -- not part of the product.
 *
 * CPU restatement of RescaleTrajectoryBackwardToStop
 * (/root/reference/trajectory_planning/rescale_to_stop.cc:35-161) and of
 * AreInputsValidForSampledTrajectory (sampled_trajectory.cc:24-44), in the
 * reference's order of floating-point operations.  Pinned bit-for-bit against
 * the reference's own object code (oracle/_ref: rescale_to_stop.cc compiled
 * verbatim against oracle/shim) by tests/test_oracle_stop.py, and against the
 * properties the reference's test checks (rescale_to_stop_test.cc:29-84).
 *
 * Layout: trajectories are [n][D] row-major.  The stopping trajectory has
 * `*count` samples and corresponds to input samples switch_index ..
 * switch_index + count - 1 (its positions are that sub-range of the input, as
 * in the reference); it is written to out_* rows 0 .. count-1.
 */
#include <math.h>
#include <stddef.h>

#include "tp_oracle.h"

/* sampled_trajectory.cc:24-44 (the size checks are implied by the layout) */
int orc_sampled_trajectory_inputs_valid(int n, const double* times) {
  if (n < 2) return 0;
  for (int i = 0; i < n - 1; ++i) {
    if (times[i + 1] <= times[i]) return 0;
  }
  return 1;
}

int orc_rescale_to_stop(int n, int D, const double* max_acc, const double* times,
                        const double* vel, const double* acc, int* count, int* switch_index,
                        double* out_times, double* out_vel, double* out_acc) {
  const double kTiny = 1e-8;
  *count = 0;
  *switch_index = 0;
  if (!orc_sampled_trajectory_inputs_valid(n, times)) return ORC_FAIL_SETUP;
  /* already stopped (:60-63): empty result */
  double vinf = 0.0;
  for (int j = 0; j < D; ++j) {
    const double a = fabs(vel[(size_t)(n - 1) * D + j]);
    if (a > vinf) vinf = a;
  }
  if (vinf < kTiny) return ORC_OK;

  /* Samples are produced back to front (:76-78, 84-141) and reversed at the end
   * (:144-146); here they are written at their final place from the back of a
   * scratch range of n rows and shifted down afterwards. */
  int m = 0; /* samples produced so far */
  /* sample 0: zero time, velocity, acceleration */
  out_times[n - 1] = 0.0;
  for (int j = 0; j < D; ++j) {
    out_vel[(size_t)(n - 1) * D + j] = 0.0;
    out_acc[(size_t)(n - 1) * D + j] = 0.0;
  }
  m = 1;
  double rate_squared = 0.0;
  double last_time = 0.0;
  for (int i = n - 1; i > 1; --i) {
    const double* v = vel + (size_t)i * D;
    const double* a = acc + (size_t)i * D;
    double diff_rate_squared = 0.0;
    for (int joint = 0; joint < D; ++joint) {
      if (fabs(v[joint]) < kTiny) continue;
      for (int s = 0; s < 2; ++s) {
        const double sign = s == 0 ? -1.0 : 1.0;
        const double bias_j = a[joint] * rate_squared;
        const double d = -2.0 * (bias_j + sign * max_acc[joint]) / v[joint];
        /* (max - scaled).minCoeff() >= -kTiny && (-max - scaled).maxCoeff() <= kTiny */
        double mn = INFINITY, mx = -INFINITY;
        for (int k = 0; k < D; ++k) {
          const double scaled = a[k] * rate_squared + 0.5 * v[k] * d;
          const double hi = max_acc[k] - scaled;
          const double lo = -max_acc[k] - scaled;
          if (hi < mn) mn = hi;
          if (lo > mx) mx = lo;
        }
        const int valid = (mn >= -kTiny) && (mx <= kTiny);
        if (valid && d < diff_rate_squared) diff_rate_squared = d;
      }
    }
    const double unscaled_dt = times[i] - times[i - 1];
    const double next_rate_squared = rate_squared - diff_rate_squared * unscaled_dt;
    const double clamped = (1.0 < next_rate_squared) ? 1.0 : next_rate_squared; /* std::min */
    const double new_time_delta = 2.0 * unscaled_dt / (sqrt(rate_squared) + sqrt(clamped));
    last_time = last_time - new_time_delta;
    const int row = n - 1 - m;
    out_times[row] = last_time;
    const double sq = sqrt(clamped);
    for (int k = 0; k < D; ++k) {
      out_vel[(size_t)row * D + k] = sq * v[k];
      out_acc[(size_t)row * D + k] = a[k] * rate_squared + 0.5 * v[k] * diff_rate_squared;
    }
    ++m;
    if (next_rate_squared >= 1.0) break;
    rate_squared = next_rate_squared;
  }
  /* time offset (:149-160) */
  const int sw = n - m;
  const double time_offset = times[sw] - out_times[sw];
  for (int r = 0; r < m; ++r) {
    out_times[r] = out_times[sw + r] + time_offset;
    for (int k = 0; k < D; ++k) {
      out_vel[(size_t)r * D + k] = out_vel[(size_t)(sw + r) * D + k];
      out_acc[(size_t)r * D + k] = out_acc[(size_t)(sw + r) * D + k];
    }
  }
  *count = m;
  *switch_index = sw;
  return ORC_OK;
}

/* ---------------------------------------------------------------------------
 * ComputeFastestStop (path_timing_trajectory.cc:75-172) and
 * PathTimingTrajectory::GetPathStopParameter (:235-287).
 *
 * PARITY: restated line by line; the reference file cannot be compiled on its
 * own here (it needs eigenmath / genit), and the reference's test
 * (path_timing_trajectory_test.cc:942-992) states properties, no values: pinned
 * through those properties only.
 * ------------------------------------------------------------------------- */

/* Forward time scaling from rate 1 until the rate reaches 0: returns the index
 * (relative to the first sample) of the last sample used, writes the stopping
 * duration.  Optional series outputs (n+1 capacity each, may be NULL) receive
 * what the reference passes to append_time / append_rate_squared /
 * append_diff_rate_squared; *num_series their count. */
int orc_fastest_stop(int n, int D, const double* stop_acc, const double* times,
                     const double* vel, const double* acc, double* total_duration,
                     double* series_time, double* series_rate2, double* series_diff,
                     int* num_series) {
  double total = 0.0;
  double rate_squared = 1.0;
  double diff_rate_squared = 0.0;
  const double first_time_point = times[0];
  int path_index;
  int ns = 0;
  for (path_index = 0; (path_index < n - 1) && (rate_squared > 0.0); path_index++) {
    const double* v = vel + (size_t)path_index * D;
    const double* a = acc + (size_t)path_index * D;
    double diff_min = 0.0;
    for (int dof = 0; dof < D; dof++) {
      const double kEpsilon = 1e-6;
      if (fabs(v[dof]) < kEpsilon) continue;
      for (int s = 0; s < 2; ++s) {
        const double bias_dof = a[dof] * rate_squared;
        const double d = (s == 0) ? 2.0 * (-bias_dof - stop_acc[dof]) / v[dof]
                                  : 2.0 * (-bias_dof + stop_acc[dof]) / v[dof];
        double mn = INFINITY, mx = -INFINITY;
        for (int k = 0; k < D; ++k) {
          const double ak = a[k] * rate_squared + 0.5 * v[k] * d;
          const double hi = stop_acc[k] - ak;
          const double lo = -stop_acc[k] - ak;
          if (hi < mn) mn = hi;
          if (lo > mx) mx = lo;
        }
        const int valid = (mn >= -1e-10) && (mx <= 1e-10);
        if (valid && d < diff_min) diff_min = d;
      }
    }
    diff_rate_squared = (0.0 < diff_min) ? 0.0 : diff_min; /* std::min(diff_min, 0.0) */
    if (series_time) series_time[ns] = first_time_point + total;
    if (series_rate2) series_rate2[ns] = rate_squared;
    if (series_diff) series_diff[ns] = diff_rate_squared;
    ++ns;
    const double unscaled_dt = times[path_index + 1] - times[path_index];
    const double cand = rate_squared + unscaled_dt * diff_rate_squared;
    const double next_rate_squared = (0.0 < cand) ? cand : 0.0; /* std::max(0.0, cand) */
    const double dt = 2.0 * unscaled_dt / (sqrt(rate_squared) + sqrt(next_rate_squared));
    total += dt;
    rate_squared = next_rate_squared;
  }
  if (series_time) series_time[ns] = first_time_point + total;
  if (series_rate2) series_rate2[ns] = rate_squared;
  if (series_diff) series_diff[ns] = diff_rate_squared;
  ++ns;
  if (num_series) *num_series = ns;
  if (path_index >= n) path_index = n - 1;
  *total_duration = total;
  return path_index;
}

/* GetPathStopParameter on a resampled trajectory of M rows (time, path parameter,
 * velocities, accelerations).  Returns ORC_OK and writes *stop_parameter, the
 * absolute row *stop_row and *stop_duration; ORC_FAIL_SETUP when time is past
 * the trajectory (kInvalidArgument in the reference). */
int orc_path_stop_parameter(int M, int D, const double* time, const double* path_parameter,
                            const double* vel, const double* acc, const double* max_acc,
                            double query_time, double* stop_parameter, int* stop_row,
                            double* stop_duration) {
  /* absl::c_lower_bound(time_, double_time) */
  int lo = 0, hi = M;
  while (lo < hi) {
    const int mid = lo + (hi - lo) / 2;
    if (time[mid] < query_time) lo = mid + 1; else hi = mid;
  }
  *stop_duration = 0.0;
  if (lo == M) return ORC_FAIL_SETUP;
  if (lo == M - 1) {
    *stop_parameter = path_parameter[M - 1];
    *stop_row = M - 1;
    return ORC_OK;
  }
  const int rel = orc_fastest_stop(M - lo, D, max_acc, time + lo, vel + (size_t)lo * D,
                                   acc + (size_t)lo * D, stop_duration, NULL, NULL, NULL, NULL);
  *stop_row = rel + lo;
  *stop_parameter = path_parameter[*stop_row];
  return ORC_OK;
}
