// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// extern "C" access to the reference's own RescaleTrajectoryBackwardToStop
// (/root/reference/trajectory_planning/rescale_to_stop.cc, with
// sampled_trajectory.cc), which oracle/Makefile compiles VERBATIM from where
// they lie against the stand-in headers in oracle/shim/ (eigenmath/types.h,
// absl/...).  Part of oracle/_ref/libtp_ref.so; pins
// oracle/tp_oracle_stop.c bit-for-bit.
#include <cstddef>
#include <vector>

#include "absl/types/span.h"
#include "eigenmath/types.h"
#include "trajectory_planning/rescale_to_stop.h"

using eigenmath::VectorXd;

namespace {
std::vector<VectorXd> Rows(const double* p, int n, int D) {
  std::vector<VectorXd> v(n, VectorXd(D));
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < D; ++j) v[i][j] = p[static_cast<size_t>(i) * D + j];
  return v;
}
}  // namespace

extern "C" {

// Same conventions as orc_rescale_to_stop (oracle/tp_oracle.h).  Returns the
// absl::StatusCode value (0 = ok).  out_pos may be NULL.
int tpref_rescale_to_stop(int n, int D, const double* max_acc, const double* times,
                          const double* pos, const double* vel, const double* acc, int* count,
                          int* switch_index, double* out_times, double* out_pos,
                          double* out_vel, double* out_acc) {
  VectorXd amax(D);
  for (int j = 0; j < D; ++j) amax[j] = max_acc[j];
  const std::vector<double> t(times, times + n);
  const std::vector<VectorXd> p = Rows(pos, n, D), v = Rows(vel, n, D), a = Rows(acc, n, D);
  auto result = trajectory_planning::RescaleTrajectoryBackwardToStop(amax, t, p, v, a);
  *count = 0;
  *switch_index = 0;
  if (!result.ok()) return static_cast<int>(result.status().code());
  const auto& r = *result;
  const int m = static_cast<int>(r.times.size());
  *count = m;
  *switch_index = m > 0 ? n - m : 0;
  for (int i = 0; i < m; ++i) {
    out_times[i] = r.times[i];
    for (int j = 0; j < D; ++j) {
      if (out_pos != nullptr) out_pos[static_cast<size_t>(i) * D + j] = r.positions[i][j];
      out_vel[static_cast<size_t>(i) * D + j] = r.velocities[i][j];
      out_acc[static_cast<size_t>(i) * D + j] = r.accelerations[i][j];
    }
  }
  return 0;
}

}  // extern "C"
