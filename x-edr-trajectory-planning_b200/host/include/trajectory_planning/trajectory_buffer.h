// Host-side mirror of trajectory_planning/trajectory_buffer.h:30-184: the same
// class, method names, argument order and error codes; the rows live on the device
// (tpb_tbuf_*, csrc/tpb_tbuf.cu).  TrajectoryBufferBatch is the batched form (many
// buffers, one kernel launch per operation) that device-side consumers use without
// any copy to the host; TrajectoryBuffer is a batch of one.
//
// Differences from the reference (all where it has no defined result): StopAtIndex
// on a standing sample that is not the last one returns kInternal instead of
// CHECK-aborting (trajectory_buffer.cc:338); the number of dofs is fixed by the
// first rows that go in.
#ifndef TRAJECTORY_PLANNING_TRAJECTORY_BUFFER_H_
#define TRAJECTORY_PLANNING_TRAJECTORY_BUFFER_H_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "tp/base.h"
#include "tpb.h"
#include "trajectory_planning/sampled_trajectory.h"

namespace trajectory_planning {

// trajectory_buffer.h:30-33
struct TrajectoryBufferOptions {
  double timestep_tolerance = 1e-6;
};

// B buffers of num_dofs joints on one device.  Per-buffer arguments are vectors of
// B entries; `statuses` (optional) receives one absl-style code per buffer.
class TrajectoryBufferBatch {
 public:
  static constexpr int32_t kSkipIndex = std::numeric_limits<int32_t>::min();
  static double SkipTime() { return std::numeric_limits<double>::quiet_NaN(); }

  static tp::StatusOr<std::unique_ptr<TrajectoryBufferBatch>> Create(
      int num_buffers, int num_dofs, int capacity_rows,
      TrajectoryBufferOptions options = TrajectoryBufferOptions{}, int device = 0) {
    tpb_tbuf_config cfg{};
    cfg.device = device;
    cfg.num_buffers = num_buffers;
    cfg.num_dofs = num_dofs;
    cfg.capacity_rows = capacity_rows;
    cfg.timestep_tolerance = options.timestep_tolerance;
    tpb_tbuf* h = nullptr;
    const int32_t rc = tpb_tbuf_create(&cfg, &h);
    if (rc == TPB_FAILED_PRECONDITION && options.timestep_tolerance <= 0) {
      return tp::FailedPreconditionError("timestep_tolerance (" +
                                         std::to_string(options.timestep_tolerance) +
                                         ") not positive");
    }
    if (rc != TPB_OK) return tp::FromCode(rc, "tpb_tbuf_create failed");
    return std::unique_ptr<TrajectoryBufferBatch>(
        new TrajectoryBufferBatch(h, num_buffers, num_dofs, capacity_rows));
  }
  ~TrajectoryBufferBatch() { tpb_tbuf_destroy(h_); }
  TrajectoryBufferBatch(const TrajectoryBufferBatch&) = delete;
  TrajectoryBufferBatch& operator=(const TrajectoryBufferBatch&) = delete;

  int NumBuffers() const { return B_; }
  int NumDofs() const { return D_; }
  int CapacityRows() const { return cap_; }
  tpb_tbuf* handle() const { return h_; }

  tp::Status Reserve(int capacity_rows) {
    const tp::Status s = Check(tpb_tbuf_reserve(h_, capacity_rows));
    if (s.ok()) cap_ = std::max(cap_, capacity_rows);
    return s;
  }
  tp::Status Clear(const std::vector<int32_t>* select = nullptr) {
    return Check(tpb_tbuf_clear(h_, select ? select->data() : nullptr, 0));
  }
  // Packed segments: segment b = rows begin[b] .. begin[b]+length[b]-1 (length < 0: skip).
  tp::Status InsertSegments(const std::vector<int64_t>& begin, const std::vector<int32_t>& length,
                            const std::vector<double>& times, const std::vector<double>& positions,
                            const std::vector<double>& velocities,
                            const std::vector<double>& accelerations,
                            std::vector<int32_t>* statuses = nullptr) {
    if (static_cast<int>(begin.size()) != B_ || static_cast<int>(length.size()) != B_ ||
        positions.size() != times.size() * D_ || velocities.size() != positions.size() ||
        accelerations.size() != positions.size()) {
      return tp::InvalidArgumentError("InsertSegments: inconsistent sizes");
    }
    return Check(tpb_tbuf_insert_segments(h_, begin.data(), length.data(),
                                          static_cast<int64_t>(times.size()), times.data(),
                                          positions.data(), velocities.data(),
                                          accelerations.data(), Out(statuses), 0));
  }
  // The same with device pointers (e.g. the engine's TPB_ARR_R_* arrays); asynchronous.
  tp::Status InsertSegmentsDevice(const int64_t* begin, const int32_t* length, int64_t total_rows,
                                  const double* times, const double* positions,
                                  const double* velocities, const double* accelerations,
                                  int32_t* statuses) {
    return Check(tpb_tbuf_insert_segments(h_, begin, length, total_rows, times, positions,
                                          velocities, accelerations, statuses, 1));
  }
  tp::Status AppendSamples(const std::vector<double>& time, const std::vector<double>& positions,
                           const std::vector<double>& velocities,
                           const std::vector<double>& accelerations,
                           const std::vector<int32_t>* select = nullptr,
                           std::vector<int32_t>* statuses = nullptr) {
    if (static_cast<int>(time.size()) != B_ || positions.size() != time.size() * D_ ||
        velocities.size() != positions.size() || accelerations.size() != positions.size()) {
      return tp::InvalidArgumentError("AppendSamples: inconsistent sizes");
    }
    return Check(tpb_tbuf_append_samples(h_, select ? select->data() : nullptr, time.data(),
                                         positions.data(), velocities.data(),
                                         accelerations.data(), Out(statuses), 0));
  }
  tp::Status DiscardSegmentsBefore(const std::vector<double>& time_sec) {
    if (static_cast<int>(time_sec.size()) != B_) return tp::InvalidArgumentError("size");
    return Check(tpb_tbuf_discard_before(h_, time_sec.data(), 0));
  }
  tp::Status GetNumPositionsUpToTime(const std::vector<double>& time_sec,
                                     std::vector<int32_t>* num) {
    if (static_cast<int>(time_sec.size()) != B_) return tp::InvalidArgumentError("size");
    num->assign(B_, 0);
    return Check(tpb_tbuf_positions_up_to_time(h_, time_sec.data(), num->data(), 0));
  }
  // time_sec [B][Q]; outputs [B][Q][D] (any may be null), statuses [B][Q].
  tp::Status SampleAtTimes(int queries_per_buffer, const std::vector<double>& time_sec,
                           std::vector<double>* positions, std::vector<double>* velocities,
                           std::vector<double>* accelerations, std::vector<int32_t>* statuses) {
    const size_t n = static_cast<size_t>(B_) * queries_per_buffer;
    if (time_sec.size() != n) return tp::InvalidArgumentError("size");
    for (auto* o : {positions, velocities, accelerations})
      if (o != nullptr) o->assign(n * D_, 0.0);
    if (statuses != nullptr) statuses->assign(n, 0);
    return Check(tpb_tbuf_sample(h_, queries_per_buffer, time_sec.data(),
                                 positions ? positions->data() : nullptr,
                                 velocities ? velocities->data() : nullptr,
                                 accelerations ? accelerations->data() : nullptr,
                                 statuses ? statuses->data() : nullptr, 0));
  }
  tp::Status StopAtIndices(const std::vector<int32_t>& index,
                           const std::vector<double>& max_acceleration, double time_step,
                           std::vector<int32_t>* statuses = nullptr) {
    if (static_cast<int>(index.size()) != B_ ||
        max_acceleration.size() != static_cast<size_t>(B_) * D_) {
      return tp::InvalidArgumentError("StopAtIndices: inconsistent sizes");
    }
    return Check(tpb_tbuf_stop_at_index(h_, index.data(), max_acceleration.data(), time_step,
                                        Out(statuses), 0));
  }
  tp::Status StopBeforeTimes(const std::vector<double>& time_sec,
                             const std::vector<double>& max_acceleration, double time_step,
                             std::vector<int32_t>* statuses = nullptr) {
    if (static_cast<int>(time_sec.size()) != B_ ||
        max_acceleration.size() != static_cast<size_t>(B_) * D_) {
      return tp::InvalidArgumentError("StopBeforeTimes: inconsistent sizes");
    }
    return Check(tpb_tbuf_stop_before_time(h_, time_sec.data(), max_acceleration.data(), time_step,
                                           Out(statuses), 0));
  }
  tp::Status AddOffsetsToTimestamps(const std::vector<double>& offset) {
    if (static_cast<int>(offset.size()) != B_) return tp::InvalidArgumentError("size");
    return Check(tpb_tbuf_add_offset(h_, offset.data(), 0));
  }
  struct State {
    std::vector<int32_t> num_samples, sequence_number;
    std::vector<double> front_time, back_time;
  };
  tp::StatusOr<State> GetState() {
    State s;
    s.num_samples.assign(B_, 0);
    s.sequence_number.assign(B_, 0);
    s.front_time.assign(B_, 0.0);
    s.back_time.assign(B_, 0.0);
    const tp::Status st = Check(tpb_tbuf_get_state(h_, s.num_samples.data(),
                                                   s.sequence_number.data(), s.front_time.data(),
                                                   s.back_time.data(), 0));
    if (!st.ok()) return st;
    return s;
  }
  tp::StatusOr<SampledTrajectory> GetRows(int buffer) {
    auto state = GetState();
    if (!state.ok()) return state.status();
    const int n = (*state).num_samples[buffer];
    std::vector<double> t(n), p(static_cast<size_t>(n) * D_), v(p.size()), a(p.size());
    const tp::Status st =
        Check(tpb_tbuf_get_rows(h_, buffer, 0, n, t.data(), p.data(), v.data(), a.data(), 0));
    if (!st.ok()) return st;
    SampledTrajectory r;
    r.times = std::move(t);
    for (int i = 0; i < n; ++i) {
      r.positions.emplace_back(p.data() + static_cast<size_t>(i) * D_, D_);
      r.velocities.emplace_back(v.data() + static_cast<size_t>(i) * D_, D_);
      r.accelerations.emplace_back(a.data() + static_cast<size_t>(i) * D_, D_);
    }
    return r;
  }

 private:
  TrajectoryBufferBatch(tpb_tbuf* h, int B, int D, int cap) : h_(h), B_(B), D_(D), cap_(cap) {}
  tp::Status Check(int32_t rc) const {
    if (rc == TPB_OK) return tp::OkStatus();
    return tp::FromCode(rc, tpb_tbuf_last_error(h_));
  }
  int32_t* Out(std::vector<int32_t>* statuses) const {
    if (statuses == nullptr) return nullptr;
    statuses->assign(B_, 0);
    return statuses->data();
  }
  tpb_tbuf* h_;
  int B_, D_, cap_;
};

// trajectory_buffer.h:40-184
class TrajectoryBuffer {
 public:
  static tp::StatusOr<std::unique_ptr<TrajectoryBuffer>> Create(
      TrajectoryBufferOptions options = TrajectoryBufferOptions{}) {
    if (options.timestep_tolerance <= 0) {
      return tp::FailedPreconditionError("timestep_tolerance (" +
                                         std::to_string(options.timestep_tolerance) +
                                         ") not positive");
    }
    return std::unique_ptr<TrajectoryBuffer>(new TrajectoryBuffer(options));
  }
  void Clear() {
    if (batch_) Must(batch_->Clear());
    sequence_before_rows_ = 0;
    dirty_ = true;
  }
  void Reserve(size_t size) {
    reserved_ = std::max(reserved_, static_cast<int>(size));
    if (batch_) Must(batch_->Reserve(reserved_));
  }
  tp::Time GetStartTime() const {
    Refresh();
    return times_.empty() ? TimeFromSec(0) : TimeFromSec(times_.front());
  }
  tp::Time GetEndTime() const {
    Refresh();
    return times_.empty() ? tp::Time{0} : TimeFromSec(times_.back());
  }
  size_t GetNumSamples() const {
    Refresh();
    return times_.size();
  }
  int GetSequenceNumber() const {
    Refresh();
    return sequence_number_;
  }
  tp::Span<const tp::VectorXd> GetPositions() const {
    Refresh();
    return positions_;
  }
  tp::Span<const tp::VectorXd> GetVelocities() const {
    Refresh();
    return velocities_;
  }
  tp::Span<const tp::VectorXd> GetAccelerations() const {
    Refresh();
    return accelerations_;
  }
  tp::Span<const double> GetTimes() const {
    Refresh();
    return times_;
  }

  // trajectory_buffer.cc:80-132
  tp::Status InsertSegment(tp::Span<const double> times, tp::Span<const tp::VectorXd> positions,
                           tp::Span<const tp::VectorXd> velocities,
                           tp::Span<const tp::VectorXd> accelerations) {
    if (positions.size() != velocities.size()) {
      return tp::InvalidArgumentError("positions and velocity arguments have different size.");
    }
    if (positions.size() != accelerations.size()) {
      return tp::InvalidArgumentError("positions and accelerations arguments have different size.");
    }
    if (positions.size() != times.size()) {
      return tp::InvalidArgumentError("positions and times arguments have different size.");
    }
    const int n = static_cast<int>(times.size());
    if (n == 0 && !batch_) {  // only counts (:97-101); no rows, no dof count yet
      sequence_before_rows_++;
      dirty_ = true;
      return tp::OkStatus();
    }
    const tp::Status ready = Ensure(static_cast<int>(positions[0].size()), n);
    if (!ready.ok()) return ready;
    const int D = batch_->NumDofs();
    std::vector<double> t(times.begin(), times.end()), p, v, a;
    Flatten(positions, D, &p);
    Flatten(velocities, D, &v);
    Flatten(accelerations, D, &a);
    std::vector<int32_t> st;
    for (int attempt = 0; attempt < 2; ++attempt) {
      const tp::Status call = batch_->InsertSegments({0}, {n}, t, p, v, a, &st);
      if (!call.ok()) return call;
      if (st[0] != TPB_RESOURCE_EXHAUSTED) break;
      const tp::Status grown = batch_->Reserve(2 * (batch_->CapacityRows() + n));
      if (!grown.ok()) return grown;
    }
    dirty_ = true;
    return tp::FromCode(st[0], "InsertSegment");
  }

  // trajectory_buffer.cc:134-148
  tp::Status AppendSample(double time, const tp::VectorXd& positions,
                          const tp::VectorXd& velocities, const tp::VectorXd& accelerations) {
    const tp::Status ready = Ensure(static_cast<int>(positions.size()), 1);
    if (!ready.ok()) return ready;
    const int D = batch_->NumDofs();
    std::vector<double> p(positions.data(), positions.data() + D),
        v(velocities.data(), velocities.data() + D),
        a(accelerations.data(), accelerations.data() + D);
    std::vector<int32_t> st;
    for (int attempt = 0; attempt < 2; ++attempt) {
      const tp::Status call = batch_->AppendSamples({time}, p, v, a, nullptr, &st);
      if (!call.ok()) return call;
      if (st[0] != TPB_RESOURCE_EXHAUSTED) break;
      const tp::Status grown = batch_->Reserve(2 * batch_->CapacityRows());
      if (!grown.ok()) return grown;
    }
    dirty_ = true;
    if (st[0] == TPB_INVALID_ARGUMENT) {
      return tp::InvalidArgumentError("time must be > times_.back().");
    }
    return tp::FromCode(st[0], "AppendSample");
  }

  void DiscardSegmentBefore(tp::Time time) { DiscardSegmentBefore(TimeToSec(time)); }
  void DiscardSegmentBefore(double time_sec) {
    if (!batch_) return;
    Must(batch_->DiscardSegmentsBefore({time_sec}));
    dirty_ = true;
  }

  // trajectory_buffer.cc:209-226
  tp::Span<const tp::VectorXd> GetPositionsUpToTime(tp::Time time) const {
    if (!batch_) return {};
    std::vector<int32_t> num;
    Must(batch_->GetNumPositionsUpToTime({TimeToSec(time)}, &num));
    Refresh();
    return tp::Span<const tp::VectorXd>(positions_.data(), static_cast<size_t>(num[0]));
  }

  // trajectory_buffer.cc:295-358
  tp::Status StopAtIndex(int index, const tp::VectorXd& max_acceleration, double time_step) {
    if (!batch_) {  // no rows: `index > GetNumSamples() - 1` cannot trigger there (:298)
      if (index <= 0) return tp::OutOfRangeError("index out of range");
      return tp::InternalError("StopAtIndex on an empty buffer");
    }
    std::vector<int32_t> st;
    const tp::Status call = batch_->StopAtIndices(
        {index}, std::vector<double>(max_acceleration.data(),
                                     max_acceleration.data() + max_acceleration.size()),
        time_step, &st);
    if (!call.ok()) return call;
    dirty_ = true;
    return StopStatus(st[0]);
  }
  tp::Status StopBeforeTime(tp::Time time, const tp::VectorXd& max_acceleration,
                            double time_step) {
    return StopBeforeTime(TimeToSec(time), max_acceleration, time_step);
  }
  // trajectory_buffer.cc:366-381
  tp::Status StopBeforeTime(double time_sec, const tp::VectorXd& max_acceleration,
                            double time_step) {
    if (!batch_) return tp::OkStatus();
    std::vector<int32_t> st;
    const tp::Status call = batch_->StopBeforeTimes(
        {time_sec}, std::vector<double>(max_acceleration.data(),
                                        max_acceleration.data() + max_acceleration.size()),
        time_step, &st);
    if (!call.ok()) return call;
    dirty_ = true;
    return StopStatus(st[0]);
  }

  tp::StatusOr<tp::VectorXd> GetPositionAtTime(tp::Time time) const {
    return GetPositionAtTime(TimeToSec(time));
  }
  tp::StatusOr<tp::VectorXd> GetPositionAtTime(double time_sec) const {
    return AtTime(0, time_sec);
  }
  tp::StatusOr<tp::VectorXd> GetVelocityAtTime(tp::Time time) const {
    return GetVelocityAtTime(TimeToSec(time));
  }
  tp::StatusOr<tp::VectorXd> GetVelocityAtTime(double time_sec) const {
    return AtTime(1, time_sec);
  }
  tp::StatusOr<tp::VectorXd> GetAccelerationAtTime(tp::Time time) const {
    return GetAccelerationAtTime(TimeToSec(time));
  }
  tp::StatusOr<tp::VectorXd> GetAccelerationAtTime(double time_sec) const {
    return AtTime(2, time_sec);
  }

  // trajectory_buffer.cc:383-390 (Duration / Seconds(1) as a floating-point quotient)
  void AddOffsetToTimestamps(tp::Duration offset) {
    AddOffsetToTimestamps(static_cast<double>(offset.ns) / 1e9);
  }
  void AddOffsetToTimestamps(double offset) {
    if (!batch_) return;
    Must(batch_->AddOffsetsToTimestamps({offset}));
    dirty_ = true;
  }

  // The device batch behind this buffer (null until the first rows went in).
  TrajectoryBufferBatch* batch() const { return batch_.get(); }

 private:
  explicit TrajectoryBuffer(TrajectoryBufferOptions options) : options_(options) {}

  static void Must(const tp::Status& s) {
    if (!s.ok()) {
      std::fprintf(stderr, "TrajectoryBuffer: %s\n", s.message().c_str());
      std::abort();
    }
  }
  static void Flatten(tp::Span<const tp::VectorXd> rows, int D, std::vector<double>* out) {
    out->resize(rows.size() * D);
    for (size_t i = 0; i < rows.size(); ++i)
      for (int j = 0; j < D; ++j) (*out)[i * D + j] = rows[i][j];
  }
  tp::Status Ensure(int num_dofs, int rows) {
    if (batch_) {
      if (num_dofs != batch_->NumDofs()) {
        return tp::InvalidArgumentError("number of dofs differs from the rows in the buffer");
      }
      return tp::OkStatus();
    }
    auto b = TrajectoryBufferBatch::Create(1, num_dofs, std::max({reserved_, 2 * rows, 4096}),
                                           options_);
    if (!b.ok()) return b.status();
    batch_ = std::move(*b);
    // empty InsertSegment calls made before any rows existed only counted (:97-101)
    for (; sequence_before_rows_ > 0; --sequence_before_rows_) {
      const tp::Status s = batch_->InsertSegments({0}, {0}, {}, {}, {}, {});
      if (!s.ok()) return s;
    }
    return tp::OkStatus();
  }
  static tp::Status StopStatus(int32_t code) {
    switch (code) {
      case TPB_OK: return tp::OkStatus();
      case TPB_OUT_OF_RANGE: return tp::OutOfRangeError("index / time out of range");
      case TPB_INVALID_ARGUMENT:
        return tp::InvalidArgumentError(
            "max_acceleration has a non-positive coefficient, `time_step` is not positive, or "
            "the time samples are not strictly increasing");
      case TPB_NOT_FOUND:
        return tp::NotFoundError("No safe stopping trajectory found (likely not enough time).");
      default: return tp::FromCode(code, "StopAtIndex");
    }
  }
  tp::StatusOr<tp::VectorXd> AtTime(int which, double time_sec) const {
    if (!batch_) return tp::FailedPreconditionError("No samples.");
    std::vector<double> out;
    std::vector<int32_t> st;
    const tp::Status call = batch_->SampleAtTimes(1, {time_sec}, which == 0 ? &out : nullptr,
                                                  which == 1 ? &out : nullptr,
                                                  which == 2 ? &out : nullptr, &st);
    if (!call.ok()) return call;
    if (st[0] == TPB_FAILED_PRECONDITION) return tp::FailedPreconditionError("No samples.");
    if (st[0] == TPB_OUT_OF_RANGE) {
      return tp::OutOfRangeError("Time: " + std::to_string(time_sec) + " outside the buffer");
    }
    if (st[0] != TPB_OK) return tp::FromCode(st[0], "GetAtTime");
    return tp::VectorXd(out.data(), out.size());
  }
  // Host copies behind the span getters, refreshed after a mutating call.
  void Refresh() const {
    if (!dirty_) return;
    dirty_ = false;
    times_.clear();
    positions_.clear();
    velocities_.clear();
    accelerations_.clear();
    sequence_number_ = sequence_before_rows_;
    if (!batch_) return;
    auto state = batch_->GetState();
    Must(state.status());
    sequence_number_ = (*state).sequence_number[0];
    auto rows = batch_->GetRows(0);
    Must(rows.status());
    times_ = std::move((*rows).times);
    positions_ = std::move((*rows).positions);
    velocities_ = std::move((*rows).velocities);
    accelerations_ = std::move((*rows).accelerations);
  }

  TrajectoryBufferOptions options_;
  std::unique_ptr<TrajectoryBufferBatch> batch_;
  int reserved_ = 0;
  // empty InsertSegment calls counted before any rows (and so any device buffer) exist
  int sequence_before_rows_ = 0;
  mutable bool dirty_ = true;
  mutable int sequence_number_ = 0;
  mutable std::vector<double> times_;
  mutable std::vector<tp::VectorXd> positions_, velocities_, accelerations_;
};

}  // namespace trajectory_planning

#endif  // TRAJECTORY_PLANNING_TRAJECTORY_BUFFER_H_
