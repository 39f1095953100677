// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// extern "C" access to the reference's own TimeOptimalPathProfile
// (/root/reference/trajectory_planning/time_optimal_path_timing.{h,cc}),
// which oracle/Makefile compiles VERBATIM from where it lies against the
// stand-in headers in oracle/shim/.  Output: oracle/_ref/libtp_ref.so.
// Used to (a) pin oracle/tp_oracle.c's stage-C restatement bit-for-bit and
// (b) serve as the timed CPU baseline ("kind": "reference") in bench.py.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <utility>
#include <vector>

#include "absl/status/status.h"
#include "Eigen/Dense"
#include "tp_oracle.h"

// The limit-curve arrays are private in the reference class; open them for
// inspection by the tests (this translation unit only; layout is unchanged).
#define private public
#include "trajectory_planning/time_optimal_path_timing.h"
#undef private

using trajectory_planning::TimeOptimalPathProfile;
using Constraint = TimeOptimalPathProfile::Constraint;

namespace {
// cons: [N][C][4] = (A, B, lower, upper)
std::vector<Constraint> MakeConstraints(int N, int C, const double* cons) {
  std::vector<Constraint> v(N);
  for (int i = 0; i < N; ++i) {
    v[i].resize(C);
    const double* row = cons + static_cast<size_t>(i) * C * 4;
    for (int r = 0; r < C; ++r) {
      v[i].a_coefficient(r) = row[r * 4 + 0];
      v[i].b_coefficient(r) = row[r * 4 + 1];
      v[i].lower(r) = row[r * 4 + 2];
      v[i].upper(r) = row[r * 4 + 3];
    }
  }
  return v;
}

template <typename A>
void CopyOut(double* dst, const A& src, int N) {
  if (dst == nullptr) return;
  for (int i = 0; i < N; ++i) dst[i] = src[i];
}
}  // namespace

extern "C" {

struct tpref_solution {
  double *time, *s, *sd, *sdd, *sd2;
  double *b_sd2_max, *b_sdd_max, *b_sdd_min, *b_sd2_zero;
  unsigned char* b_type;
  int last_extremal_index;
  int satisfies;  // SolutionSatisfiesConstraints().ok()
};

// Returns 1 if InitSolver+SetupProblem+OptimizePathParameter all succeeded,
// 0 if OptimizePathParameter failed, -1 if setup failed.
int tpref_solve(int N, int C, const double* cons, double s_start, double s_end,
                double sd_start, double sdd_start, double time_start,
                int max_loops, tpref_solution* out) {
  TimeOptimalPathProfile p;
  if (!p.InitSolver(N, C)) return -1;
  if (max_loops > 0) p.SetMaxNumSolverLoops(max_loops);
  if (!p.SetupProblem(MakeConstraints(N, C, cons), s_start, s_end, sd_start,
                      sdd_start, time_start)) {
    return -1;
  }
  const bool ok = p.OptimizePathParameter();
  if (out != nullptr) {
    CopyOut(out->time, p.time_, N);
    CopyOut(out->s, p.s_, N);
    CopyOut(out->sd, p.sd_, N);
    CopyOut(out->sdd, p.sdd_, N);
    CopyOut(out->sd2, p.sd2_, N);
    CopyOut(out->b_sd2_max, p.boundary_.sd2_max, N);
    CopyOut(out->b_sdd_max, p.boundary_.sdd_max_for_sd2_max, N);
    CopyOut(out->b_sdd_min, p.boundary_.sdd_min_for_sd2_max, N);
    CopyOut(out->b_sd2_zero, p.boundary_.sd2_max_for_sdd0, N);
    if (out->b_type != nullptr) {
      for (int i = 0; i < N; ++i) out->b_type[i] = p.boundary_.type[i];
    }
    out->last_extremal_index = p.GetLastExtremalIndex();
    out->satisfies = ok ? (p.SolutionSatisfiesConstraints().ok() ? 1 : 0) : 0;
  }
  return ok ? 1 : 0;
}

// One LP (one sample's constraint table [C][4]); out3 = sd2max, sddmax, sd2zero.
void tpref_find_max_sd2(int C, const double* cons_row, int brute_force,
                        double* out3) {
  TimeOptimalPathProfile p;
  p.InitSolver(10, C);
  std::vector<Constraint> v = MakeConstraints(1, C, cons_row);
  if (brute_force) {
    p.FindMaxSd2BruteForce(v[0], &out3[0], &out3[1], &out3[2]);
  } else {
    p.FindMaxSd2Simplex(v[0], &out3[0], &out3[1], &out3[2]);
  }
}

// GetPathParameterAndDerivatives after a solve of the given problem.
int tpref_query(int N, int C, const double* cons, double s_start, double s_end,
                double sd_start, double sdd_start, double time_start, int nq,
                const double* tq, double* s, double* sd, double* sdd) {
  TimeOptimalPathProfile p;
  if (!p.InitSolver(N, C)) return -1;
  if (!p.SetupProblem(MakeConstraints(N, C, cons), s_start, s_end, sd_start,
                      sdd_start, time_start)) {
    return -1;
  }
  if (!p.OptimizePathParameter()) return 0;
  for (int i = 0; i < nq; ++i) {
    if (!p.GetPathParameterAndDerivatives(tq[i], &s[i], &sd[i], &sdd[i])) {
      return 0;
    }
  }
  return 1;
}

// Batch of B independent problems with the same N, C on `nthreads` host
// threads (the reference itself is single-threaded; each thread owns its own
// TimeOptimalPathProfile).  cons: [B][N][C][4].  duration[b] = NaN on failure.
// Used as the timed CPU baseline for stage C.
void tpref_solve_batch(int B, int N, int C, const double* cons,
                       const double* s_end, int max_loops, int nthreads,
                       double* duration, double* time_out /* [B][N] or null */,
                       double* sd_out, double* sdd_out) {
  std::atomic<int> next{0};
  auto worker = [&]() {
    TimeOptimalPathProfile p;
    p.InitSolver(N, C);
    if (max_loops > 0) p.SetMaxNumSolverLoops(max_loops);
    std::vector<Constraint> v(N);
    for (auto& c : v) c.resize(C);
    for (;;) {
      const int b = next.fetch_add(1);
      if (b >= B) break;
      const double* cb = cons + static_cast<size_t>(b) * N * C * 4;
      for (int i = 0; i < N; ++i) {
        const double* row = cb + static_cast<size_t>(i) * C * 4;
        for (int r = 0; r < C; ++r) {
          v[i].a_coefficient(r) = row[r * 4 + 0];
          v[i].b_coefficient(r) = row[r * 4 + 1];
          v[i].lower(r) = row[r * 4 + 2];
          v[i].upper(r) = row[r * 4 + 3];
        }
      }
      bool ok = p.SetupProblem(v, 0.0, s_end[b], 0.0, 0.0, 0.0);
      ok = ok && p.OptimizePathParameter();
      duration[b] = ok ? p.GetTotalDuration()
                       : std::numeric_limits<double>::quiet_NaN();
      if (ok && time_out != nullptr) {
        CopyOut(time_out + static_cast<size_t>(b) * N, p.GetTimeSamples(), N);
      }
      if (ok && sd_out != nullptr) {
        CopyOut(sd_out + static_cast<size_t>(b) * N, p.GetPathVelocity(), N);
      }
      if (ok && sdd_out != nullptr) {
        CopyOut(sdd_out + static_cast<size_t>(b) * N, p.GetPathAcceleration(), N);
      }
    }
  };
  if (nthreads <= 1) {
    worker();
    return;
  }
  std::vector<std::thread> pool;
  for (int t = 0; t < nthreads; ++t) pool.emplace_back(worker);
  for (auto& t : pool) t.join();
}


// Whole pipeline for one fresh joint path in a single window, like
// orc_plan_joint (tp_oracle.c) but with stage C run by the reference's own
// TimeOptimalPathProfile.  Stages A/B/D are the restatement (the reference's
// files for them need Eigen/eigenmath/abseil, which are not available).
int tpref_plan_joint(const orc_plan_cfg* cfg, const double* waypoints,
                     const double* vmax, const double* amax,
                     orc_plan_out* out) {
  const int N = cfg->N, D = cfg->dim, W = cfg->W, C = 2 * D;
  const int nmax = (W == 1) ? 4 : 3 * W - 2;
  std::vector<double> ctrl(static_cast<size_t>(nmax) * D), knots(nmax + 3);
  std::vector<double> cons(static_cast<size_t>(N) * C * 4);
  const int n = orc_joint_fit(waypoints, W, D, cfg->rounding, ctrl.data(),
                              knots.data());
  const int nk = n + 3;
  const double kend = knots[nk - 1];
  const double delta = (cfg->delta > 0) ? cfg->delta : kend / (N - 1);
  out->knot_end = kend;
  out->delta_used = delta;
  orc_joint_sample(knots.data(), nk, ctrl.data(), D, N, 0.0, delta, out->q,
                   out->dq, out->ddq);
  orc_joint_constraints(out->dq, out->ddq, N, D, vmax, amax, cfg->safety,
                        cons.data());
  TimeOptimalPathProfile p;
  out->M = 0;
  out->duration = std::numeric_limits<double>::quiet_NaN();
  out->last_extremal_index = 0;
  if (!p.InitSolver(N, C)) return ORC_FAIL_SETUP;
  p.SetMaxNumSolverLoops(std::max(100, 10 * N));
  if (!p.SetupProblem(MakeConstraints(N, C, cons.data()), 0.0,
                      0.0 + delta * (N - 1), 0.0, 0.0, cfg->start_sec)) {
    return ORC_FAIL_SETUP;
  }
  if (!p.OptimizePathParameter()) return ORC_FAIL_CONNECT;
  CopyOut(out->time, p.GetTimeSamples(), N);
  CopyOut(out->s, p.GetPathParameter(), N);
  CopyOut(out->sd, p.GetPathVelocity(), N);
  CopyOut(out->sdd, p.GetPathAcceleration(), N);
  out->last_extremal_index = p.GetLastExtremalIndex();
  out->duration = p.GetTotalDuration();
  orc_assemble(out->dq, out->ddq, out->sd, out->sdd, amax, N, D, out->vel,
               out->acc);
  if (cfg->dt > 0 && out->r_t != nullptr) {
    int M = orc_resample_count(out->time[N - 1], cfg->start_sec, cfg->dt);
    M = std::min(M, out->M_cap);
    out->M = M;
    orc_resample_uniform(N, D, out->time, out->s, out->sd, out->sdd, out->q,
                         out->vel, out->acc, amax, cfg->start_sec, cfg->dt, M,
                         out->r_t, out->r_pos, out->r_vel, out->r_acc,
                         out->r_s, out->r_sd, out->r_sdd);
  }
  return ORC_OK;
}

}  // extern "C"
