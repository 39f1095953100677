// Batched RescaleTrajectoryBackwardToStop (rescale_to_stop.cc:35-161): for each
// sampled trajectory, the fastest stop that traces its tail -- a backward
// integration of the squared time-scaling rate from 0 at the end until the
// original velocity is reached, with the scaled acceleration
//   0.5 * velocity * d(rate^2) + acceleration * rate^2
// kept inside +-max_acceleration.
//
// One warp per trajectory.  The steps are sequential (each needs the previous
// rate^2); inside a step lane c = 2*joint + (0: sign -1, 1: sign +1) owns one
// of the 2D candidate rates of change (:104-121), tests it against all D joints
// and the minimum over the valid ones is a shuffle reduction -- the same shape
// as FindSddMin in tpb_integrate.cuh.  Results are bit-identical to the
// reference's loops: the minimum does not depend on the order of evaluation,
// everything else follows the reference's order of operations (-fmad=false).
//
// The stopping trajectory of a trajectory of n samples has m <= n-1 samples and
// corresponds to input samples n-m .. n-1 (its positions ARE those input rows).
// Sample r is produced back to front, so it is written straight to row
// begin + n-1-r of the output arrays: no reversal, and the output arrays have
// the layout of the input.  A second, lane-parallel pass adds the time offset
// that lines the segment up with the input (:149-160).
#include <cfloat>

#include "tpb_device.cuh"
#include "tpb_engine.h"

namespace tpb {

namespace {

constexpr int kStopWarpsPerBlock = 4;
constexpr double kStopTiny = 1e-8;  // rescale_to_stop.cc:60, 98

__global__ void __launch_bounds__(kStopWarpsPerBlock * 32)
k_rescale_to_stop(StopArgs p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * kStopWarpsPerBlock + warp;
  if (b >= p.num_traj) return;
  const int D = p.D;
  const int n = p.length[b];
  const size_t row0 = static_cast<size_t>(p.begin[b]);
  const double* times = p.times + row0;
  const double* vel = p.vel + row0 * D;
  const double* acc = p.acc + row0 * D;
  double* out_t = p.out_times + row0;
  double* out_v = p.out_vel + row0 * D;
  double* out_a = p.out_acc + row0 * D;
  const int k = lane < D ? lane : D - 1;  // the joint whose v, a this lane holds
  const double amax_k = p.max_acc[static_cast<size_t>(b) * D + k];

  int status = TPB_PATH_OK, count = 0, sw = 0;
  // AreInputsValidForSampledTrajectory (sampled_trajectory.cc:24-44)
  bool bad = n < 2;
  for (int i = lane; i < n - 1; i += 32) bad |= times[i + 1] <= times[i];
  if (__ballot_sync(kFullMask, bad)) status = TPB_PATH_FAIL_SETUP;

  if (status == TPB_PATH_OK) {
    // already stopped (:60-63)
    const double vlast = fabs(vel[static_cast<size_t>(n - 1) * D + k]);
    const bool moving = !(vlast < kStopTiny);  // lpNorm<Infinity> < kTiny fails if any is not
    if (__ballot_sync(kFullMask, moving && lane < D) != 0u) {
      // sample 0: zero time, velocity, acceleration (:76-78)
      if (lane == 0) out_t[n - 1] = 0.0;
      if (lane < D) {
        out_v[static_cast<size_t>(n - 1) * D + lane] = 0.0;
        out_a[static_cast<size_t>(n - 1) * D + lane] = 0.0;
      }
      int m = 1;
      double rate_squared = 0.0, last_time = 0.0;
      const int joint = lane >> 1;
      const double sign = (lane & 1) ? 1.0 : -1.0;
      const bool is_cand = lane < 2 * D;
      // values of step i; those of step i-1 are requested while step i computes
      double v_k = vel[static_cast<size_t>(n - 1) * D + k];
      double a_k = acc[static_cast<size_t>(n - 1) * D + k];
      double t_i = times[n - 1], t_im = times[n - 2];
      for (int i = n - 1; i > 1; --i) {
        const double v_n = vel[static_cast<size_t>(i - 1) * D + k];
        const double a_n = acc[static_cast<size_t>(i - 1) * D + k];
        const double t_n = times[i - 2];

        const double bias_k = a_k * rate_squared;  // acceleration_bias (:93)
        const double v_j = __shfl_sync(kFullMask, v_k, joint);
        const double bias_j = __shfl_sync(kFullMask, bias_k, joint);
        const double amax_j = __shfl_sync(kFullMask, amax_k, joint);
        const bool skip = !is_cand || (fabs(v_j) < kStopTiny);
        // (:106-108) -2 (bias + sign max) / v; skipped lanes divide 1 by 1
        const double num = -2.0 * (bias_j + sign * amax_j);
        const bool zero = num == 0.0;
        const double quot = guarded_div(skip || zero, num, v_j);
        const double d = zero ? num * v_j : quot;
        // (:109-114) valid for every joint?
        double mn = DBL_MAX, mx = -DBL_MAX;
        for (int q = 0; q < D; ++q) {
          const double bq = __shfl_sync(kFullMask, bias_k, q);
          const double vq = __shfl_sync(kFullMask, v_k, q);
          const double mq = __shfl_sync(kFullMask, amax_k, q);
          const double scaled = bq + 0.5 * vq * d;
          const double hi = mq - scaled, lo = -mq - scaled;
          mn = (hi < mn) ? hi : mn;
          mx = (lo > mx) ? lo : mx;
        }
        const bool valid = !skip && (mn >= -kStopTiny) && (mx <= kStopTiny);
        // (:115-117) the smallest valid candidate below 0
        double diff = (valid && d < 0.0) ? d : 0.0;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
          const double w = __shfl_xor_sync(kFullMask, diff, o);
          diff = (w < diff) ? w : diff;
        }
        // (:122-138)
        const double unscaled_dt = t_i - t_im;
        const double next_rate_squared = rate_squared - diff * unscaled_dt;
        const double clamped = (1.0 < next_rate_squared) ? 1.0 : next_rate_squared;  // std::min
        const double sq = sqrt(clamped);
        const double new_time_delta = 2.0 * unscaled_dt / (sqrt(rate_squared) + sq);
        last_time = last_time - new_time_delta;
        const int row = n - 1 - m;
        if (lane == 0) out_t[row] = last_time;
        if (lane < D) {
          out_v[static_cast<size_t>(row) * D + lane] = sq * v_k;
          out_a[static_cast<size_t>(row) * D + lane] = bias_k + 0.5 * v_k * diff;
        }
        ++m;
        if (next_rate_squared >= 1.0) break;
        rate_squared = next_rate_squared;
        v_k = v_n;
        a_k = a_n;
        t_i = t_im;
        t_im = t_n;
      }
      __syncwarp();
      // line the segment up with the input (:149-160)
      sw = n - m;
      count = m;
      const double time_offset = times[sw] - out_t[sw];
      __syncwarp();
      for (int r = sw + lane; r < n; r += 32) out_t[r] = out_t[r] + time_offset;
    }
  }
  if (lane == 0) {
    p.count[b] = count;
    p.switch_index[b] = sw;
    p.status[b] = status;
  }
}

// ComputeFastestStop (path_timing_trajectory.cc:75-172) under
// PathTimingTrajectory::GetPathStopParameter (:235-287), one warp per trajectory:
// from the first resampled row at or after the query time, scale time forward
// from rate 1 with the largest admissible deceleration until the rate reaches 0;
// the answer is the path parameter of the row where that happens.  Same lane
// layout as above (lane c = 2*dof + sign).
constexpr double kStopQueryEpsilon = 1e-6;     // :111
constexpr double kStopQueryTinyAcc = 1e-10;    // :119

__global__ void __launch_bounds__(kStopWarpsPerBlock * 32)
k_path_stop_parameter(StopQueryArgs p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * kStopWarpsPerBlock + warp;
  if (b >= p.B) return;
  const int D = p.D;
  const size_t row0 = static_cast<size_t>(p.offset[b]);
  const int M = static_cast<int>(p.offset[b + 1] - p.offset[b]);
  const double* times = p.times + row0;
  const double* vel = p.vel + row0 * D;
  const double* acc = p.acc + row0 * D;
  const double query = p.query[b];
  const int k = lane < D ? lane : D - 1;
  const double sa_k = p.max_acc[static_cast<size_t>(b) * D + k];

  // absl::c_lower_bound(time_, double_time) (:243)
  int lo = 0, hi = M;
  while (lo < hi) {
    const int mid = lo + (hi - lo) / 2;
    if (times[mid] < query) {
      lo = mid + 1;
    } else {
      hi = mid;
    }
  }
  int status = TPB_PATH_OK, row = 0;
  double parameter = quiet_nan(), total = 0.0;
  if (M <= 0 || lo == M) {
    status = TPB_PATH_FAIL_SETUP;  // "Time ... not in timed path range" (:244-248)
  } else if (lo == M - 1) {
    row = M - 1;  // (:249-251)
    parameter = p.s[row0 + row];
  } else {
    const int n = M - lo;
    const int dof = lane >> 1;
    const bool is_cand = lane < 2 * D;
    const bool plus = (lane & 1) != 0;  // second candidate: -bias + stop_acc (:125-126)
    double rate_squared = 1.0;
    int path_index = 0;
    double v_k = vel[static_cast<size_t>(lo) * D + k];
    double a_k = acc[static_cast<size_t>(lo) * D + k];
    double t_i = times[lo], t_n = times[lo + 1];
    for (; (path_index < n - 1) && (rate_squared > 0.0); path_index++) {
      // values of the next step (rows stay inside the trajectory)
      const int nx = min(lo + path_index + 1, M - 1), nx2 = min(lo + path_index + 2, M - 1);
      const double v_x = vel[static_cast<size_t>(nx) * D + k];
      const double a_x = acc[static_cast<size_t>(nx) * D + k];
      const double t_x = times[nx2];

      const double bias_k = a_k * rate_squared;  // acc_bias (:103-106)
      const double v_j = __shfl_sync(kFullMask, v_k, dof);
      const double bias_j = __shfl_sync(kFullMask, bias_k, dof);
      const double sa_j = __shfl_sync(kFullMask, sa_k, dof);
      const bool skip = !is_cand || (fabs(v_j) < kStopQueryEpsilon);
      const double num = 2.0 * (plus ? (-bias_j + sa_j) : (-bias_j - sa_j));
      const bool zero = num == 0.0;
      const double quot = guarded_div(skip || zero, num, v_j);
      const double d = zero ? num * v_j : quot;
      double mn = DBL_MAX, mx = -DBL_MAX;
      for (int q = 0; q < D; ++q) {
        const double bq = __shfl_sync(kFullMask, bias_k, q);
        const double vq = __shfl_sync(kFullMask, v_k, q);
        const double sq = __shfl_sync(kFullMask, sa_k, q);
        const double aq = bq + 0.5 * vq * d;
        const double hi_q = sq - aq, lo_q = -sq - aq;
        mn = (hi_q < mn) ? hi_q : mn;
        mx = (lo_q > mx) ? lo_q : mx;
      }
      const bool valid = !skip && (mn >= -kStopQueryTinyAcc) && (mx <= kStopQueryTinyAcc);
      double diff = (valid && d < 0.0) ? d : 0.0;  // min over valid candidates and 0 (:121-131)
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) {
        const double w = __shfl_xor_sync(kFullMask, diff, o);
        diff = (w < diff) ? w : diff;
      }
      // forward Euler in rate^2, trapezoidal time (:137-145)
      const double unscaled_dt = t_n - t_i;
      const double cand = rate_squared + unscaled_dt * diff;
      const double next_rate_squared = (0.0 < cand) ? cand : 0.0;  // std::max(0.0, .)
      const double dt = 2.0 * unscaled_dt / (sqrt(rate_squared) + sqrt(next_rate_squared));
      total += dt;
      rate_squared = next_rate_squared;
      v_k = v_x;
      a_k = a_x;
      t_i = t_n;
      t_n = t_x;
    }
    if (path_index >= n) path_index = n - 1;  // (:150-153)
    row = path_index + lo;
    parameter = p.s[row0 + row];
  }
  if (lane == 0) {
    p.parameter[b] = parameter;
    p.row[b] = row;
    p.duration[b] = total;
    p.status[b] = status;
  }
}

}  // namespace

int launch_path_stop_parameter(const StopQueryArgs& p, cudaStream_t st) {
  const int blocks = (p.B + kStopWarpsPerBlock - 1) / kStopWarpsPerBlock;
  k_path_stop_parameter<<<blocks, kStopWarpsPerBlock * 32, 0, st>>>(p);
  return 1;
}

int launch_rescale_to_stop(const StopArgs& p, cudaStream_t st) {
  const int blocks = (p.num_traj + kStopWarpsPerBlock - 1) / kStopWarpsPerBlock;
  k_rescale_to_stop<<<blocks, kStopWarpsPerBlock * 32, 0, st>>>(p);
  return 1;
}

}  // namespace tpb
