// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// extern "C" access to the reference's own TrajectoryBuffer
// (/root/reference/trajectory_planning/trajectory_buffer.{h,cc}), which
// oracle/Makefile compiles VERBATIM from where it lies against the stand-in
// headers in oracle/shim/ (absl time/status/span/..., eigenmath/types.h and
// eigenmath/interpolation.h -- the latter is UNPINNED, see its header).  Part of
// oracle/_ref/libtp_ref.so; pins oracle/tp_oracle_tbuf.c bit-for-bit.
//
// Same conventions as the orc_tbuf_* functions (oracle/tp_oracle.h): rows are
// [n][D] row-major, return values are absl::StatusCode numbers.
#include <cstddef>
#include <cstdint>
#include <memory>
#include <vector>

#include "absl/time/time.h"
#include "absl/types/span.h"
#include "eigenmath/types.h"
#include "trajectory_planning/trajectory_buffer.h"

using eigenmath::VectorXd;
using trajectory_planning::TrajectoryBuffer;

namespace trajectory_planning {
// time.h defines these without `inline`; trajectory_buffer.cc's object has them.
absl::Time TimeFromSec(double seconds);
double TimeToSec(const absl::Time& time);
}  // namespace trajectory_planning

namespace {
std::vector<VectorXd> Rows(const double* p, int n, int D) {
  std::vector<VectorXd> v(n, VectorXd(D));
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < D; ++j) v[i][j] = p[static_cast<size_t>(i) * D + j];
  return v;
}
VectorXd Row(const double* p, int D) {
  VectorXd v(D);
  for (int j = 0; j < D; ++j) v[j] = p[j];
  return v;
}
TrajectoryBuffer* B(void* h) { return static_cast<TrajectoryBuffer*>(h); }
}  // namespace

extern "C" {

int tpref_tbuf_create(double timestep_tolerance, void** out) {
  auto r = TrajectoryBuffer::Create({.timestep_tolerance = timestep_tolerance});
  *out = nullptr;
  if (!r.ok()) return static_cast<int>(r.status().code());
  *out = (*r).release();
  return 0;
}
void tpref_tbuf_destroy(void* h) { delete B(h); }
void tpref_tbuf_clear(void* h) { B(h)->Clear(); }

int tpref_tbuf_insert_segment(void* h, int n, int D, const double* times, const double* pos,
                              const double* vel, const double* acc) {
  const std::vector<double> t(times, times + n);
  const auto p = Rows(pos, n, D), v = Rows(vel, n, D), a = Rows(acc, n, D);
  return static_cast<int>(B(h)->InsertSegment(t, p, v, a).code());
}
int tpref_tbuf_append_sample(void* h, int D, double time, const double* pos, const double* vel,
                             const double* acc) {
  return static_cast<int>(B(h)->AppendSample(time, Row(pos, D), Row(vel, D), Row(acc, D)).code());
}
void tpref_tbuf_discard_before_sec(void* h, double time_sec) {
  B(h)->DiscardSegmentBefore(time_sec);
}
void tpref_tbuf_discard_before_ns(void* h, int64_t ns) {
  B(h)->DiscardSegmentBefore(absl::FromUnixNanos(ns));
}
int tpref_tbuf_positions_up_to_time_ns(void* h, int64_t ns) {
  return static_cast<int>(B(h)->GetPositionsUpToTime(absl::FromUnixNanos(ns)).size());
}
// which: 0 position, 1 velocity, 2 acceleration
int tpref_tbuf_get_at_time(void* h, int which, double time_sec, double* out) {
  auto r = which == 0   ? B(h)->GetPositionAtTime(time_sec)
           : which == 1 ? B(h)->GetVelocityAtTime(time_sec)
                        : B(h)->GetAccelerationAtTime(time_sec);
  if (!r.ok()) return static_cast<int>(r.status().code());
  for (int j = 0; j < (*r).size(); ++j) out[j] = (*r)[j];
  return 0;
}
int tpref_tbuf_stop_at_index(void* h, int index, int D, const double* max_acc, double time_step) {
  return static_cast<int>(B(h)->StopAtIndex(index, Row(max_acc, D), time_step).code());
}
int tpref_tbuf_stop_before_time(void* h, double time_sec, int D, const double* max_acc,
                                double time_step) {
  return static_cast<int>(B(h)->StopBeforeTime(time_sec, Row(max_acc, D), time_step).code());
}
void tpref_tbuf_add_offset(void* h, double offset) { B(h)->AddOffsetToTimestamps(offset); }

int tpref_tbuf_num_samples(void* h) { return static_cast<int>(B(h)->GetNumSamples()); }
int tpref_tbuf_sequence_number(void* h) { return B(h)->GetSequenceNumber(); }
int64_t tpref_tbuf_start_ns(void* h) { return absl::ToUnixNanos(B(h)->GetStartTime()); }
int64_t tpref_tbuf_end_ns(void* h) { return absl::ToUnixNanos(B(h)->GetEndTime()); }
void tpref_tbuf_rows(void* h, int D, double* times, double* pos, double* vel, double* acc) {
  const TrajectoryBuffer* b = B(h);
  const int n = static_cast<int>(b->GetNumSamples());
  for (int i = 0; i < n; ++i) {
    times[i] = b->GetTimes()[i];
    for (int j = 0; j < D; ++j) {
      pos[static_cast<size_t>(i) * D + j] = b->GetPositions()[i][j];
      vel[static_cast<size_t>(i) * D + j] = b->GetVelocities()[i][j];
      acc[static_cast<size_t>(i) * D + j] = b->GetAccelerations()[i][j];
    }
  }
}
int64_t tpref_time_from_sec_ns(double seconds) {
  return absl::ToUnixNanos(trajectory_planning::TimeFromSec(seconds));
}
double tpref_time_to_sec(int64_t ns) {
  return trajectory_planning::TimeToSec(absl::FromUnixNanos(ns));
}

}  // extern "C"
