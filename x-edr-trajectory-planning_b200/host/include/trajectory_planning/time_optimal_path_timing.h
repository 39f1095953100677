// Host-side mirror of the reference's TimeOptimalPathProfile
// (trajectory_planning/time_optimal_path_timing.h:36-357): same class, method
// names, argument order and bool/Status conventions; the work runs on the GPU
// through the C ABI (include/tpb.h).  BASELINE.json calls the class
// TimeOptimalPathTiming; that name is exported as an alias.
//
// A TimeOptimalPathProfile is a batch of one.  TimeOptimalPathProfileBatch is
// the same interface over B independent problems (what the GPU is for).
#ifndef TRAJECTORY_PLANNING_TIME_OPTIMAL_PATH_TIMING_H_
#define TRAJECTORY_PLANNING_TIME_OPTIMAL_PATH_TIMING_H_

#include <cmath>
#include <cstdio>
#include <limits>
#include <vector>

#include "tp/base.h"
#include "tpb.h"

namespace trajectory_planning {

class TimeOptimalPathProfileBatch;

class TimeOptimalPathProfile {
 public:
  using Scalar = double;
  enum DebugVerbosity { kNoOutput = 0, kMainAlgorithm, kExtremalLoop, kExtremalControl,
                        kExtremalDetail, kAll };

  // Constraint rows of one sample: lower <= A*sdd + B*sd^2 <= upper
  // (time_optimal_path_timing.h:65-102).  Stored row-major as (A,B,lower,upper),
  // which is the layout tpb_set_constraints consumes.
  class Constraint {
   public:
    void resize(int size) { data_.assign(static_cast<size_t>(size) * 4, 0.0); }
    int size() const { return static_cast<int>(data_.size() / 4); }
    Scalar& lower(int i) { return data_[i * 4 + 2]; }
    const Scalar& lower(int i) const { return data_[i * 4 + 2]; }
    Scalar& upper(int i) { return data_[i * 4 + 3]; }
    const Scalar& upper(int i) const { return data_[i * 4 + 3]; }
    Scalar& a_coefficient(int i) { return data_[i * 4 + 0]; }
    const Scalar& a_coefficient(int i) const { return data_[i * 4 + 0]; }
    Scalar& b_coefficient(int i) { return data_[i * 4 + 1]; }
    const Scalar& b_coefficient(int i) const { return data_[i * 4 + 1]; }
    const Scalar* data() const { return data_.data(); }

   private:
    std::vector<Scalar> data_;
  };

  TimeOptimalPathProfile() = default;
  ~TimeOptimalPathProfile() { if (e_) tpb_destroy(e_); }
  TimeOptimalPathProfile(const TimeOptimalPathProfile&) = delete;
  TimeOptimalPathProfile& operator=(const TimeOptimalPathProfile&) = delete;

  static void SetDebugVerbosity(int) {}  // debug dumps are out of scope
  static void SetDevice(int device) { Device() = device; }

  // InitSolver (time_optimal_path_timing.cc:135-159): allocates (not real-time safe).
  bool InitSolver(int num_samples, int num_constraints) {
    if (e_ && (num_samples != num_samples_ || num_constraints != num_constraints_)) {
      tpb_destroy(e_);
      e_ = nullptr;
    }
    num_samples_ = num_samples;
    num_constraints_ = num_constraints;
    if (!e_) {
      tpb_config cfg{};
      cfg.device = Device();
      cfg.max_paths = 1;
      cfg.num_samples = num_samples;
      cfg.max_constraints = num_constraints;
      if (tpb_create(&cfg, &e_) != TPB_OK) {
        fprintf(stderr, "[E] InitSolver: tpb_create failed (no CUDA device or bad sizes)\n");
        e_ = nullptr;
        return false;
      }
    }
    time_.assign(num_samples, 0.0);
    s_ = sd_ = sdd_ = sd2_ = time_;
    last_extremal_index_ = 0;
    state_ = kAllocated;
    return true;
  }

  // SetupProblem (:161-203).  Validation happens on the device per problem; the
  // result is reported by OptimizePathParameter (status TPB_PATH_FAIL_SETUP).
  bool SetupProblem(const std::vector<Constraint>& constraints, Scalar s_start, Scalar s_end,
                    Scalar sd_start, Scalar sdd_start, Scalar time_start) {
    if (state_ == kInvalidState) return false;
    if (static_cast<int>(constraints.size()) != num_samples_) {
      fprintf(stderr, "[E] Wrong sampling dimension for constraints.\n");
      return false;
    }
    std::vector<double> flat(static_cast<size_t>(num_samples_) * num_constraints_ * 4);
    for (int i = 0; i < num_samples_; ++i) {
      if (constraints[i].size() != num_constraints_) {
        fprintf(stderr, "[E] Constraint size error.\n");
        return false;
      }
      std::copy(constraints[i].data(), constraints[i].data() + num_constraints_ * 4,
                flat.begin() + static_cast<size_t>(i) * num_constraints_ * 4);
    }
    // host-side copies of the cheap argument checks, so that SetupProblem itself
    // returns false like the reference does
    for (size_t r = 0; r < flat.size() / 4; ++r) {
      if (flat[r * 4 + 2] >= flat[r * 4 + 3]) {
        fprintf(stderr, "[E] Constraints must satisfy: lower < upper.\n");
        return false;
      }
    }
    if (s_start >= s_end || sd_start < 0 || num_samples_ < 2) return false;
    if (tpb_set_constraints(e_, 1, num_constraints_, flat.data(), &s_start, &s_end, 0) != TPB_OK) {
      fprintf(stderr, "[E] %s\n", tpb_last_error(e_));
      return false;
    }
    s_start_ = s_start;
    s_end_ = s_end;
    sd_start_ = sd_start;
    sdd_start_ = sdd_start;
    time_start_ = time_start;
    constraints_ = flat;
    state_ = kProblemDefined;
    return true;
  }

  void SetMaxNumSolverLoops(int num_loops) { max_num_loops_ = num_loops; }

  // OptimizePathParameter (:287-490)
  bool OptimizePathParameter() {
    if (state_ != kProblemDefined && state_ != kProblemSolved) {
      fprintf(stderr, "[E] Error, problem not defined/set up.\n");
      return false;
    }
    if (tpb_optimize(e_, &sd_start_, &sdd_start_, &time_start_, max_num_loops_, 0) != TPB_OK) {
      fprintf(stderr, "[E] %s\n", tpb_last_error(e_));
      return false;
    }
    int32_t status = TPB_PATH_NOT_RUN;
    const size_t nb = sizeof(double) * num_samples_;
    bool ok = tpb_get(e_, TPB_ARR_STATUS, &status, sizeof(status), 0) == TPB_OK &&
              tpb_get(e_, TPB_ARR_TIME, time_.data(), nb, 0) == TPB_OK &&
              tpb_get(e_, TPB_ARR_S, s_.data(), nb, 0) == TPB_OK &&
              tpb_get(e_, TPB_ARR_SD, sd_.data(), nb, 0) == TPB_OK &&
              tpb_get(e_, TPB_ARR_SDD, sdd_.data(), nb, 0) == TPB_OK &&
              tpb_get(e_, TPB_ARR_SD2, sd2_.data(), nb, 0) == TPB_OK &&
              tpb_get(e_, TPB_ARR_LAST_EXTREMAL, &last_extremal_index_, sizeof(int32_t), 0) == TPB_OK;
    path_status_ = status;
    if (!ok || status != TPB_PATH_OK) {
      fprintf(stderr, "[E] OptimizePathParameter failed (path status %d)\n", status);
      return false;
    }
    dt_max_ = 0.0;
    for (int i = 1; i < num_samples_; ++i) dt_max_ = std::max(dt_max_, time_[i] - time_[i - 1]);
    state_ = kProblemSolved;
    return true;
  }

  const std::vector<Scalar>& GetTimeSamples() const { return time_; }
  const std::vector<Scalar>& GetPathParameter() const { return s_; }
  const std::vector<Scalar>& GetPathVelocity() const { return sd_; }
  const std::vector<Scalar>& GetPathAcceleration() const { return sdd_; }
  Scalar GetTotalDuration() const { return time_.back() - time_.front(); }
  Scalar GetEndTime() const { return time_.back(); }
  Scalar GetStartTime() const { return time_.front(); }
  int GetLastExtremalIndex() const { return last_extremal_index_; }
  // TPB_PATH_* cause of the last failure (extension; the reference only has bool).
  int GetPathStatus() const { return path_status_; }

  Scalar GetMaxTimeIncrement() const { return state_ == kProblemSolved ? dt_max_ : -1; }

  // SolutionSatisfiesConstraints (:492-518): the audit runs on the device
  // (tpb_count_violations).
  tp::Status SolutionSatisfiesConstraints() {
    if (state_ != kProblemSolved) return tp::FailedPreconditionError("No valid solution.");
    int32_t violations = -1;
    const int32_t rc = tpb_count_violations(e_, &violations, 0);
    if (rc != TPB_OK) return tp::FromCode(rc, tpb_last_error(e_));
    if (violations < 0) return tp::FailedPreconditionError("No valid solution.");
    if (violations > 0) {
      return tp::NotFoundError("Number of constraint violations: " + std::to_string(violations));
    }
    return tp::OkStatus();
  }

  // Result queries (:1497-1668) run on the device (tpb_query_profile, one query of this
  // one profile per call; TimeOptimalPathProfileBatch / the C ABI take whole grids).
  int GetPreviousIndex(Scalar t) const {
    if (state_ != kProblemSolved) return -1;
    int32_t k = -1;
    if (tpb_query_profile(e_, 1, &t, nullptr, nullptr, nullptr, &k, nullptr, nullptr, nullptr,
                          nullptr, nullptr, 0) != TPB_OK)
      return -1;
    return k;
  }
  bool GetPreviousDiscreteValues(Scalar t, Scalar* sk, Scalar* sdk, Scalar* sddk, Scalar* tk) const {
    if (state_ != kProblemSolved) return false;
    int32_t st = 3;
    if (tpb_query_profile(e_, 1, &t, nullptr, nullptr, nullptr, nullptr, sk, sdk, sddk, tk, &st,
                          0) != TPB_OK)
      return false;
    return (st & 2) == 0;
  }
  bool GetPathParameterAndDerivatives(Scalar t, Scalar* s, Scalar* sd, Scalar* sdd) const {
    if (state_ != kProblemSolved) return false;
    int32_t st = 3;
    if (tpb_query_profile(e_, 1, &t, s, sd, sdd, nullptr, nullptr, nullptr, nullptr, nullptr,
                          &st, 0) != TPB_OK)
      return false;
    return (st & 1) == 0;
  }

  // FindMaxSd2Simplex (:1149-1363) for one constraint table, on the device.
  void FindMaxSd2Simplex(const Constraint& constr, Scalar* sd2max, Scalar* sddmax, Scalar* sd2zero) {
    *sd2max = *sddmax = *sd2zero = std::numeric_limits<Scalar>::quiet_NaN();
    tpb_engine* e = nullptr;
    tpb_config cfg{};
    cfg.device = Device();
    cfg.max_paths = 1;
    cfg.num_samples = 2;
    cfg.max_constraints = constr.size();
    if (tpb_create(&cfg, &e) != TPB_OK) return;
    std::vector<double> flat(static_cast<size_t>(2) * constr.size() * 4);
    for (int i = 0; i < 2; ++i)
      std::copy(constr.data(), constr.data() + constr.size() * 4, flat.begin() + i * constr.size() * 4);
    const double s0 = 0.0, s1 = 1.0;
    if (tpb_set_constraints(e, 1, constr.size(), flat.data(), &s0, &s1, 0) == TPB_OK &&
        tpb_optimize(e, nullptr, nullptr, nullptr, 1, 0) == TPB_OK) {
      double v[2];
      // N = 2 has no interior samples, so passes 2-4 leave the limit curve as computed
      if (tpb_get(e, TPB_ARR_B_SD2_MAX, v, sizeof(v), 0) == TPB_OK) *sd2max = v[0];
      if (tpb_get(e, TPB_ARR_B_SD2_ZERO, v, sizeof(v), 0) == TPB_OK) *sd2zero = v[0];
      *sddmax = 0.0;  // not stored by CalculateBoundary either (:1370-1374)
    }
    tpb_destroy(e);
  }

  tpb_engine* engine() { return e_; }

 private:
  enum SolverState { kInvalidState = 0, kAllocated, kProblemDefined, kProblemSolved };
  static int& Device() {
    static int d = 0;
    return d;
  }

  tpb_engine* e_ = nullptr;
  SolverState state_ = kInvalidState;
  int num_constraints_ = 0, num_samples_ = 0, max_num_loops_ = 100;
  int32_t last_extremal_index_ = 0, path_status_ = TPB_PATH_NOT_RUN;
  Scalar s_start_ = 0, s_end_ = 0, sd_start_ = 0, sdd_start_ = 0, time_start_ = 0, dt_max_ = 0;
  std::vector<Scalar> time_, s_, sd_, sdd_, sd2_, constraints_;
};

using TimeOptimalPathTiming = TimeOptimalPathProfile;  // BASELINE.json's name

// B independent problems with the same N and C in one call each.
class TimeOptimalPathProfileBatch {
 public:
  ~TimeOptimalPathProfileBatch() { if (e_) tpb_destroy(e_); }
  tp::Status Init(int max_problems, int num_samples, int num_constraints, int device = 0) {
    if (e_) tpb_destroy(e_);
    e_ = nullptr;
    tpb_config cfg{};
    cfg.device = device;
    cfg.max_paths = max_problems;
    cfg.num_samples = num_samples;
    cfg.max_constraints = num_constraints;
    N_ = num_samples;
    const int32_t rc = tpb_create(&cfg, &e_);
    return tp::FromCode(rc, "tpb_create failed");
  }
  // constraints: [B][N][C][4] rows (A,B,lower,upper); s_start/s_end: [B].
  tp::Status SetupProblems(int B, int C, const double* constraints, const double* s_start,
                           const double* s_end) {
    B_ = B;
    return Ck(tpb_set_constraints(e_, B, C, constraints, s_start, s_end, 0));
  }
  tp::Status OptimizePathParameters(const double* sd_start, const double* sdd_start,
                                    const double* time_start, int max_loops = 100) {
    return Ck(tpb_optimize(e_, sd_start, sdd_start, time_start, max_loops, 0));
  }
  tp::Status Get(tpb_array what, void* dst, size_t bytes) { return Ck(tpb_get(e_, what, dst, bytes, 0)); }
  int num_problems() const { return B_; }
  int num_samples() const { return N_; }
  tpb_engine* engine() { return e_; }

 private:
  tp::Status Ck(int32_t rc) { return tp::FromCode(rc, e_ ? tpb_last_error(e_) : ""); }
  tpb_engine* e_ = nullptr;
  int B_ = 0, N_ = 0;
};

}  // namespace trajectory_planning

#endif  // TRAJECTORY_PLANNING_TIME_OPTIMAL_PATH_TIMING_H_
