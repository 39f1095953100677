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
