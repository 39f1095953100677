// Path-timing kernels for sm_100a (compiled with -fmad=false, see tpb_device.cuh).
//
//  k_joint_fit            one thread per path: waypoints -> control polygon + knots
//                         (spline_utils.cc:25-102, timeable_path_joint_spline.cc:252-292)
//  k_joint_sample_limit   one thread per (path, sample): SamplePath + ConstraintSetup
//                         (timeable_path_joint_spline.cc:294-343) fused with the
//                         per-sample limit curve (time_optimal_path_timing.cc:1366-1377);
//                         constraint rows exist only in registers/local memory
//  k_generic_limit        same limit-curve step for explicit constraint tables
//  k_integrate            one warp per path: CalculateBoundary passes 2-4 (:1379-1484),
//                         extremals + switching-point loop + time integration (:287-490)
//  k_resample_*           trajectory assembly + ResampleEquidistantlyInTime
//                         (path_timing_trajectory.cc:458-472, :686-783)
#include "tpb_engine.h"
#include "tpb_solver.cuh"

namespace tpb {
namespace {

// ===========================================================================
// K2: joint_fit
// ===========================================================================

// PolyLineToBspline3WaypointsCornerOffset, splines/spline_utils.cc:25-45.
__device__ void corner_offset(const double* delta, int D, double radius, double* off) {
  constexpr double kMinNorm = 1e-6;
  constexpr double kMinWaypointSpacingFactor = 4.0;  // spline_utils.h:46
  double s = 0.0;
  for (int d = 0; d < D; ++d) s += delta[d] * delta[d];
  const double norm = sqrt(s);
  if (norm > kMinNorm) {
    for (int d = 0; d < D; ++d) off[d] = delta[d] / norm;
  } else {
    for (int d = 0; d < D; ++d) off[d] = 0.0;
  }
  if (norm > kMinWaypointSpacingFactor * radius) {
    for (int d = 0; d < D; ++d) off[d] = off[d] * radius;
  } else {
    const double f = 1.0 / kMinWaypointSpacingFactor;
    for (int d = 0; d < D; ++d) off[d] = (off[d] * f) * norm;
  }
}

__global__ void k_joint_fit(Dims dm, DeviceArrays a, double rounding, bool auto_delta) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B) return;
  const int D = dm.D;
  const int W = a.num_wp[b];
  const double* wp = a.waypoints + static_cast<size_t>(b) * dm.Wmax * D;
  double* c = a.ctrl + static_cast<size_t>(b) * dm.max_points * D;
  double* kn = a.knots + static_cast<size_t>(b) * (dm.max_points + 3);
  double delta[kMaxDofs], off[kMaxDofs];

  // PolyLineToBspline3Waypoints (VectorXd), splines/spline_utils.cc:47-102.
  int n;
  if (W == 1) {
    n = 4;
    for (int i = 0; i < 4; ++i)
      for (int d = 0; d < D; ++d) c[i * D + d] = wp[d];
  } else {
    n = 3 * W - 2;
    for (int i = 0; i < W; ++i)
      for (int d = 0; d < D; ++d) c[3 * i * D + d] = wp[i * D + d];
    for (int i = 1; i < W - 1; ++i) {
      const int k = 3 * i, knext = 3 * (i + 1), klast = 3 * (i - 1);
      for (int d = 0; d < D; ++d) delta[d] = c[knext * D + d] - c[k * D + d];
      corner_offset(delta, D, rounding, off);
      for (int d = 0; d < D; ++d) c[(k + 1) * D + d] = c[k * D + d] + off[d];
      for (int d = 0; d < D; ++d) delta[d] = c[klast * D + d] - c[k * D + d];
      corner_offset(delta, D, rounding, off);
      for (int d = 0; d < D; ++d) c[(k - 1) * D + d] = c[k * D + d] + off[d];
    }
    for (int d = 0; d < D; ++d) delta[d] = c[3 * D + d] - c[d];
    corner_offset(delta, D, rounding, off);
    for (int d = 0; d < D; ++d) c[D + d] = c[d] + off[d];
    for (int d = 0; d < D; ++d) delta[d] = c[(n - 4) * D + d] - c[(n - 1) * D + d];
    corner_offset(delta, D, rounding, off);
    for (int d = 0; d < D; ++d) c[(n - 2) * D + d] = c[(n - 1) * D + d] + off[d];
  }

  // MakeUniformKnotVector on [0,1] (bspline_base.cc:350-381), degree 2
  // (kSplineOrder, timeable_path_joint_spline.h:89), then scaled by the control
  // polygon length (timeable_path_joint_spline.cc:271-285).
  constexpr int p = 2;
  const int nk = n + p + 1;
  for (int i = 0; i <= p; ++i) kn[i] = 0.0;
  const double denom = nk - 2.0 * (p + 1.0) + 1.0;
  const double spacing = (1.0 / denom) * (1.0 - 0.0);
  for (int i = p + 1; i < nk - p - 1; ++i) kn[i] = kn[i - 1] + spacing;
  for (int i = nk - p - 1; i < nk; ++i) kn[i] = 1.0;
  double length = 0.0;
  for (int i = 0; i < n - 1; ++i) {
    double s = 0.0;
    for (int d = 0; d < D; ++d) {
      const double df = c[(i + 1) * D + d] - c[i * D + d];
      s += df * df;
    }
    length += sqrt(s);
  }
  constexpr double kMinimumFinalKnotValue = 0.1;
  constexpr double kPathParameterPerPolygonLength = 1.0;
  const double lw = length * kPathParameterPerPolygonLength;
  const double weighted = (lw < kMinimumFinalKnotValue) ? kMinimumFinalKnotValue : lw;
  for (int i = 0; i < nk; ++i) kn[i] *= weighted;

  a.num_points[b] = n;
  const double kend = kn[nk - 1];
  a.knot_end[b] = kend;
  a.delta[b] = auto_delta ? kend / (dm.N - 1) : a.delta_in[b];
}

// ===========================================================================
// K3: joint sample + per-sample limit curve
// ===========================================================================

constexpr int kSampleThreads = 128;

__global__ void __launch_bounds__(kSampleThreads)
k_joint_sample_limit(Dims dm, DeviceArrays a, double safety) {
  extern __shared__ double smem[];
  const int b = blockIdx.y;
  const int D = dm.D;
  const int n = a.num_points[b];
  const int nk = n + 3;
  double* s_knots = smem;                        // [max_points + 3]
  double* s_ctrl = s_knots + dm.max_points + 3;  // [max_points][D]
  double* s_aS = s_ctrl + dm.max_points * D;     // [D]  amax * safety
  double* s_vS2 = s_aS + D;                      // [D]  (vmax * safety)^2
  {
    const double* gk = a.knots + static_cast<size_t>(b) * (dm.max_points + 3);
    const double* gc = a.ctrl + static_cast<size_t>(b) * dm.max_points * D;
    for (int i = threadIdx.x; i < nk; i += blockDim.x) s_knots[i] = gk[i];
    for (int i = threadIdx.x; i < n * D; i += blockDim.x) s_ctrl[i] = gc[i];
    if (threadIdx.x < D) {
      const double am = a.amax[static_cast<size_t>(b) * D + threadIdx.x];
      const double vm = a.vmax[static_cast<size_t>(b) * D + threadIdx.x];
      s_aS[threadIdx.x] = am * safety;
      const double vs = vm * safety;
      s_vS2[threadIdx.x] = vs * vs;
    }
  }
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= dm.N) return;

  // SamplePath, timeable_path_joint_spline.cc:294-318.
  const double delta = a.delta[b];
  const double path_start = a.path_start[b];
  const double parameter = path_start + idx * delta;
  const double k0 = s_knots[0], k1 = s_knots[nk - 1];
  double A[2 * kMaxDofs], Bc[2 * kMaxDofs], LO[2 * kMaxDofs], HI[2 * kMaxDofs];
  const size_t row = (static_cast<size_t>(b) * dm.N + idx) * D;
  if (parameter < k1 + delta) {
    const double u = (parameter < k0) ? k0 : ((k1 < parameter) ? k1 : parameter);
    const int span = knot_span(s_knots, nk, 2, u);
    double ders[3][3];
    basis_ders<2, 2>(s_knots, span, u, ders);
    // EvalCurveAndDerivatives, splines/bspline.h:561-566.
    const double* P0 = s_ctrl + (span - 2) * D;
    for (int d = 0; d < D; ++d) {
      double v0 = 0.0, v1 = 0.0, v2 = 0.0;
#pragma unroll
      for (int j = 0; j <= 2; ++j) {
        const double Pj = P0[j * D + d];
        v0 += ders[0][j] * Pj;
        v1 += ders[1][j] * Pj;
        v2 += ders[2][j] * Pj;
      }
      a.q[row + d] = v0;
      a.dq[row + d] = v1;
      a.ddq[row + d] = v2;
      A[d] = v1;
      Bc[d] = v2;
    }
  } else {
    for (int d = 0; d < D; ++d) {
      a.q[row + d] = s_ctrl[(n - 1) * D + d];
      a.dq[row + d] = 0.0;
      a.ddq[row + d] = 0.0;
      A[d] = 0.0;
      Bc[d] = 0.0;
    }
  }
  // ConstraintSetup, timeable_path_joint_spline.cc:320-343 (rows stay on chip).
  for (int d = 0; d < D; ++d) {
    HI[d] = s_aS[d];
    LO[d] = -s_aS[d];
    A[D + d] = 0.0;
    Bc[D + d] = A[d] * A[d];
    HI[D + d] = s_vS2[d];
    LO[D + d] = 0.0;
  }
  // CalculateBoundary, per-sample part (time_optimal_path_timing.cc:1366-1377).
  double sd2max, sddmax, sd2zero, smax, smin;
  lp_max_sd2(A, Bc, LO, HI, 2 * D, sd2max, sddmax, sd2zero);
  find_sdd_min_max(A, Bc, LO, HI, 2 * D, sd2max, smax, smin);
  const size_t o = static_cast<size_t>(b) * dm.N + idx;
  a.b_raw_sd2_max[o] = sd2max;
  a.b_sd2_zero[o] = sd2zero;
  a.b_raw_sdd_max[o] = smax;
  a.b_raw_sdd_min[o] = smin;
  a.b_at0[o] = fabs(sd2max - sd2zero) < kTiny;
}

__global__ void __launch_bounds__(kSampleThreads)
k_generic_limit(Dims dm, DeviceArrays a) {
  const int b = blockIdx.y;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= dm.N) return;
  const int C = dm.C;
  double A[kMaxRows], Bc[kMaxRows], LO[kMaxRows], HI[kMaxRows];
  const double* c = a.cons + (static_cast<size_t>(b) * dm.N + idx) * C * 4;
  for (int r = 0; r < C; ++r) {
    A[r] = c[r * 4 + 0];
    Bc[r] = c[r * 4 + 1];
    LO[r] = c[r * 4 + 2];
    HI[r] = c[r * 4 + 3];
  }
  double sd2max, sddmax, sd2zero, smax, smin;
  lp_max_sd2(A, Bc, LO, HI, C, sd2max, sddmax, sd2zero);
  find_sdd_min_max(A, Bc, LO, HI, C, sd2max, smax, smin);
  const size_t o = static_cast<size_t>(b) * dm.N + idx;
  a.b_raw_sd2_max[o] = sd2max;
  a.b_sd2_zero[o] = sd2zero;
  a.b_raw_sdd_max[o] = smax;
  a.b_raw_sdd_min[o] = smin;
  a.b_at0[o] = fabs(sd2max - sd2zero) < kTiny;
}

// ===========================================================================
// K4: one warp per path
// ===========================================================================

constexpr int kWarpsPerBlock = 4;

struct WarpSolver {
  int lane, N, C, D;
  bool joint;
  double ds;
  // global, this path
  double *sd2, *sdd, *bmax, *bsmax, *bsmin, *bzero;
  uint8_t *at0, *type;
  const double *dq, *ddq, *cons;
  // shared, this warp: rows of the staged sample
  double *sA, *sB, *sLo, *sHi;
  int staged;

  __device__ __forceinline__ void st(double* p, double v) const {
    if (lane == 0) *p = v;
    __syncwarp();
  }

  // Bring the constraint rows of sample idx on chip.  Joint paths synthesise
  // them from q', q'' (ConstraintSetup); lower/upper were filled once.
  __device__ void stage(int idx) {
    if (staged == idx) return;
    __syncwarp();
    if (joint) {
      if (lane < D) {
        const double d1 = dq[static_cast<size_t>(idx) * D + lane];
        const double d2 = ddq[static_cast<size_t>(idx) * D + lane];
        sA[lane] = d1;
        sB[lane] = d2;
        sA[D + lane] = 0.0;
        sB[D + lane] = d1 * d1;
      }
    } else {
      const double* c = cons + static_cast<size_t>(idx) * C * 4;
      for (int r = lane; r < C; r += 32) {
        sA[r] = c[r * 4 + 0];
        sB[r] = c[r * 4 + 1];
        sLo[r] = c[r * 4 + 2];
        sHi[r] = c[r * 4 + 3];
      }
    }
    __syncwarp();
    staged = idx;
  }

  // AreDerivativesValid (:624-636): lane r tests row r.
  __device__ bool valid(int idx, double sddv, double sd2v) {
    stage(idx);
    bool bad = false;
    if (lane < C) {
      const double v = sA[lane] * sddv + sB[lane] * sd2v;
      bad = (v + kTiny < sLo[lane]) || (v - kTiny > sHi[lane]);
    }
    return __ballot_sync(kFullMask, bad) == 0u;
  }

  // FindSddMax / FindSddMin (:638-695): lane j owns candidate (row j/2, bound j&1)
  // and tests it against all rows; extremum by warp shuffle.
  template <bool kMax>
  __device__ double find_sdd(int idx, double sd2v) {
    stage(idx);
    double best = kMax ? -DBL_MAX : DBL_MAX;
    for (int j = lane; j < 2 * C; j += 32) {
      const int r = j >> 1;
      const double Ar = sA[r];
      if (is_tiny(Ar)) continue;
      const double cand = (((j & 1) ? sHi[r] : sLo[r]) - sB[r] * sd2v) / Ar;
      bool ok = (cand == cand);
      for (int i = 0; i < C && ok; ++i) {
        const double v = sA[i] * cand + sB[i] * sd2v;
        ok = !((v + kTiny < sLo[i]) || (v - kTiny > sHi[i]));
      }
      if (ok) {
        if (kMax) {
          best = (cand > best) ? cand : best;
        } else {
          best = (cand < best) ? cand : best;
        }
      }
    }
    best = kMax ? warp_max(best) : warp_min(best);
    if (best == (kMax ? -DBL_MAX : DBL_MAX)) best = 0.0;
    return best;
  }

  // ComputeSddAtIntersection (:722-751)
  __device__ void sdd_at_intersection(int index) {
    double cand[3];
    int n = 0;
    if (index > 0 && index < N - 1) cand[n++] = 0.25 / ds * (sd2[index + 1] - sd2[index - 1]);
    if (index < N - 1) cand[n++] = 0.5 / ds * (sd2[index + 1] - sd2[index]);
    if (index > 0) cand[n++] = 0.5 / ds * (sd2[index] - sd2[index - 1]);
    const double here = sd2[index];
    double r = 0.0;
    for (int k = 0; k < n; ++k) {
      if (valid(index, cand[k], here)) {
        r = cand[k];
        break;
      }
    }
    st(&sdd[index], r);
  }

  // CalculateBoundary passes 2-4 (:1379-1484).  Pass 2 mutates neighbours in
  // place while scanning, so it is a warp-uniform serial scan; the (rare) fixes
  // call the warp-parallel FindSdd*.  Deferred fixes go to (fix_idx, fix_val).
  __device__ void boundary_passes(double* fix_val, int* fix_idx) {
    if (lane == 0) {
      type[0] = kNone;
      type[N - 1] = kNone;
    }
    int nwork = 0;
    for (int i = 1; i < N - 1; ++i) {
      if (!at0[i - 1] && at0[i] && !at0[i + 1]) {
        const double zm = bzero[i - 1];
        const double a1 = find_sdd<true>(i - 1, zm);
        const double a2 = find_sdd<false>(i - 1, zm);
        const double zp = bzero[i + 1];
        const double a3 = find_sdd<true>(i + 1, zp);
        if (lane == 0) {
          bmax[i - 1] = zm;
          bsmax[i - 1] = a1;
          bsmin[i - 1] = a2;
          bmax[i + 1] = zp;
          bsmax[i + 1] = a3;
          bsmin[i + 1] = a3;  // sic: FindSddMax, time_optimal_path_timing.cc:1394-1395
        }
        __syncwarp();
      }
      const double bi = bmax[i], bp = bmax[i + 1], bm = bmax[i - 1];
      const double sd2p = (bp - bi) / ds;
      const double sd2p_min = 2 * bsmin[i];
      const double smax_i = bsmax[i];
      const double sd2p_max = 2 * smax_i;
      const bool sink_or_source = (sd2p < sd2p_min) || (sd2p > sd2p_max);
      const bool skipped_sdd = (smax_i > 0) && (bsmin[i + 1] < 0);
      const bool skipped_sd2 = (bi > bm - kTiny) && (bi > bp - kTiny);
      if ((skipped_sd2 || skipped_sdd) && sink_or_source) {
        const double fw = bm + 2.0 * ds * find_sdd<true>(i - 1, bm);
        const double bw = bp - 2.0 * ds * find_sdd<false>(i + 1, bp);
        double m = bzero[i];
        if (fw < m) m = fw;
        if (bw < m) m = bw;
        const double v = (0.0 < m) ? m : 0.0;
        if (lane == 0) {
          fix_idx[nwork] = i;
          fix_val[nwork] = v;
        }
        ++nwork;
      }
    }
    __syncwarp();
    for (int w = 0; w < nwork; ++w) {
      const int index = fix_idx[w];
      const double value = fix_val[w];
      const double a1 = find_sdd<true>(index, value);
      const double a2 = find_sdd<false>(index, value);
      if (lane == 0) {
        bmax[index] = value;
        bsmax[index] = a1;
        bsmin[index] = a2;
      }
      if (index > 0) {
        const double z = bzero[index - 1];
        const double c1 = find_sdd<true>(index - 1, z);
        const double c2 = find_sdd<false>(index - 1, z);
        if (lane == 0) {
          bmax[index - 1] = z;
          bsmax[index - 1] = c1;
          bsmin[index - 1] = c2;
        }
      }
      if (index < N - 1) {
        const double z = bzero[index + 1];
        const double c1 = find_sdd<true>(index + 1, z);
        const double c2 = find_sdd<false>(index + 1, z);
        if (lane == 0) {
          bmax[index + 1] = z;
          bsmax[index + 1] = c1;
          bsmin[index + 1] = c2;
        }
      }
      __syncwarp();
    }
    // classification (:1453-1484), lane-parallel
    for (int i = 1 + lane; i < N - 1; i += 32) {
      const double sd2p = (bmax[i + 1] - bmax[i]) / ds;
      const double sd2p_min = 2 * bsmin[i];
      const double sd2p_max = 2 * bsmax[i];
      uint8_t t = kNone;
      if (sd2p < sd2p_min) {
        t = kSink;
      } else if (sd2p > sd2p_max) {
        t = kSource;
      }
      if ((sd2p <= sd2p_max) && (sd2p >= sd2p_min)) t = kTrajectory;
      type[i] = t;
    }
    __syncwarp();
  }

  // AddForwardExtremal (:769-857)
  __device__ int add_forward(int idx_lo) {
    for (int idx = idx_lo; idx < N - 2; ++idx) {
      const double cur = sd2[idx];
      double sd2tmp, sddtmp;
      const bool on_boundary = is_tiny(cur - bmax[idx]);
      if (on_boundary && (type[idx] & kTrajectory) && (idx < N - 1) &&
          (type[idx + 1] & kTrajectory)) {
        sd2tmp = bmax[idx + 1];
        sddtmp = 0.5 * (sd2tmp - cur) / ds;
      } else {
        sddtmp = find_sdd<true>(idx, cur);
        sd2tmp = cur + 2.0 * ds * sddtmp;
      }
      const double nxt = sd2[idx + 1];
      if (!(nxt != nxt) && (nxt < sd2tmp)) {
        sdd_at_intersection(idx);
        return N - 1;
      }
      const double bnext = bmax[idx + 1];
      if (sd2tmp > bnext) {
        const double sdd_bound = 0.5 * (bnext - cur) / ds;
        const bool deriv_invalid = !valid(idx, sdd_bound, bmax[idx]);
        const bool type_invalid = type[idx + 1] & kSink;
        if (type_invalid || deriv_invalid) return idx;
        sd2tmp = bnext;
        sddtmp = sdd_bound;
      }
      if (sd2tmp < 0) {
        sd2tmp = 0.0;
        if (idx <= 1) {
          sddtmp = 0.0;
        } else {
          sddtmp = -sd2[idx - 1] / ds;
        }
      }
      if (lane == 0) {
        sd2[idx + 1] = sd2tmp;
        sdd[idx] = sddtmp;
      }
      __syncwarp();
    }
    return N - 1;
  }

  // AddBackwardExtremal (:859-952)
  __device__ int add_backward(int idx_hi) {
    for (int idx = idx_hi; idx > 1; --idx) {
      const double cur = sd2[idx];
      double sd2tmp, sddtmp;
      const bool on_boundary = is_tiny(cur - bmax[idx]);
      if (on_boundary && (type[idx] & kTrajectory) && (idx > 0) &&
          (type[idx - 1] & kTrajectory)) {
        sd2tmp = bmax[idx - 1];
        sddtmp = 0.5 * (cur - sd2tmp) / ds;
      } else {
        sddtmp = find_sdd<false>(idx, cur);
        sd2tmp = cur - 2.0 * ds * sddtmp;
      }
      const double prv = sd2[idx - 1];
      if (!(prv != prv) && (prv < sd2tmp)) {
        sdd_at_intersection(idx);
        return 0;
      }
      const double bprev = bmax[idx - 1];
      if (sd2tmp > bprev) {
        const double sdd_bound = 0.5 * (cur - bprev) / ds;
        const bool deriv_invalid = !valid(idx, sdd_bound, cur);
        const bool type_invalid = type[idx - 1] & kSource;
        const bool is_connecting = (idx_hi != (N - 1));
        if ((type_invalid || deriv_invalid) && !is_connecting) return idx;
        sd2tmp = bprev;
        sddtmp = sdd_bound;
      }
      if (sd2tmp < 0) {
        sd2tmp = 0.0;
        if (idx < N - 1) {
          sddtmp = sd2[idx + 1] / ds;
        } else {
          sddtmp = 0.0;
        }
      }
      if (lane == 0) {
        sd2[idx - 1] = sd2tmp;
        sdd[idx] = sddtmp;
      }
      __syncwarp();
    }
    return 0;
  }

  // NextCriticalPoint (:697-720), 32 samples per step via ballots.  Sequential
  // meaning: crit = first idx with a Source/Trajectory type; afterwards crit
  // moves to any idx with sd2_max[idx] == sd2_max_for_sdd0[0] (sic, index 0);
  // return crit at the first idx (at or after the first crit) whose sd2 is set.
  __device__ int next_critical(int idx_lo, int idx_hi) const {
    int crit = -1;
    const double bz0 = bzero[0];
    for (int base = idx_lo + 1; base <= idx_hi; base += 32) {
      const int idx = base + lane;
      const bool in = idx <= idx_hi;
      bool f1 = false, f2 = false, f3 = false;
      if (in) {
        f1 = (type[idx] & (kSource | kTrajectory)) != 0;
        f2 = bmax[idx] == bz0;
        const double v = sd2[idx];
        f3 = !(v != v);
      }
      const unsigned m1 = __ballot_sync(kFullMask, f1);
      unsigned m2 = __ballot_sync(kFullMask, f2);
      unsigned m3 = __ballot_sync(kFullMask, f3);
      if (crit < 0) {
        if (m1 == 0u) continue;
        const int p = __ffs(m1) - 1;
        crit = base + p;
        // f2 is only consulted strictly after p; f3 from p on.
        m2 = (p == 31) ? 0u : (m2 & (~0u << (p + 1)));
        m3 &= (~0u << p);
      }
      if (m3) {
        const int t = __ffs(m3) - 1;
        const unsigned upto = (t == 31) ? ~0u : ((1u << (t + 1)) - 1u);
        const unsigned m2r = m2 & upto;
        if (m2r) crit = base + (31 - __clz(m2r));
        return crit;
      }
      if (m2) crit = base + (31 - __clz(m2));
    }
    return -1;
  }
};

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
k_integrate(Dims dm, DeviceArrays a, bool joint, double safety, int max_loops) {
  __shared__ double s_rows[kWarpsPerBlock][4][kMaxRows];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * kWarpsPerBlock + warp;
  if (b >= dm.B) return;
  const int N = dm.N;
  const size_t o = static_cast<size_t>(b) * N;

  WarpSolver S;
  S.lane = lane;
  S.N = N;
  S.C = dm.C;
  S.D = dm.D;
  S.joint = joint;
  S.sd2 = a.sd2 + o;
  S.sdd = a.sdd + o;
  S.bmax = a.b_sd2_max + o;
  S.bsmax = a.b_sdd_max + o;
  S.bsmin = a.b_sdd_min + o;
  S.bzero = a.b_sd2_zero + o;
  S.at0 = a.b_at0 + o;
  S.type = a.b_type + o;
  S.dq = joint ? a.dq + o * dm.D : nullptr;
  S.ddq = joint ? a.ddq + o * dm.D : nullptr;
  S.cons = joint ? nullptr : a.cons + o * dm.C * 4;
  S.sA = s_rows[warp][0];
  S.sB = s_rows[warp][1];
  S.sLo = s_rows[warp][2];
  S.sHi = s_rows[warp][3];
  S.staged = -1;
  double* time = a.time + o;
  double* sv = a.s + o;
  double* sd = a.sd + o;
  double* fix_val = a.fix_val + o;
  int* fix_idx = reinterpret_cast<int*>(a.time + o);  // time[] is free until the end

  const double sd_start = a.sd_start[b];
  const double sdd_start = a.sdd_start[b];
  const double time_start = a.time_start[b];
  double s_start, s_end;
  int status = TPB_PATH_OK;

  // SetupProblem (:161-203) / IsSetupValid (:554-576)
  if (joint) {
    const int D = dm.D;
    bool bad = false;
    if (lane < D) {
      const double aS = a.amax[static_cast<size_t>(b) * D + lane] * safety;
      const double vs = a.vmax[static_cast<size_t>(b) * D + lane] * safety;
      const double vS2 = vs * vs;
      S.sLo[lane] = -aS;
      S.sHi[lane] = aS;
      S.sLo[D + lane] = 0.0;
      S.sHi[D + lane] = vS2;
      bad = (-aS >= aS) || (0.0 >= vS2);
    }
    if (__ballot_sync(kFullMask, bad)) status = TPB_PATH_FAIL_SETUP;
    // ComputeTimingProfile: path_horizon_ = path_start_ + delta * (N - 1)
    // (path_timing_trajectory.cc:341-342)
    s_start = a.path_start[b];
    s_end = s_start + a.delta[b] * (N - 1);
  } else {
    bool bad = false;
    const double* c = S.cons;
    const size_t rows = static_cast<size_t>(N) * dm.C;
    for (size_t r = lane; r < rows; r += 32) bad |= (c[r * 4 + 2] >= c[r * 4 + 3]);
    if (__ballot_sync(kFullMask, bad)) status = TPB_PATH_FAIL_SETUP;
    s_start = a.s_start[b];
    s_end = a.s_end[b];
  }
  if (s_start >= s_end || sd_start < 0 || N < 2) status = TPB_PATH_FAIL_SETUP;
  __syncwarp();

  int last_extremal = 0;
  const double nan = quiet_nan();
  if (status == TPB_PATH_OK) {
    // SetSetupDone (:535-552)
    const double ds = (s_end - s_start) / (N - 1);
    S.ds = ds;
    // The per-sample limit curve is kept as computed (b_raw_*); passes 2-4
    // mutate a working copy, so tpb_optimize can be repeated like the
    // reference's OptimizePathParameter (which recomputes the boundary).
    const double* raw_max = a.b_raw_sd2_max + o;
    const double* raw_smax = a.b_raw_sdd_max + o;
    const double* raw_smin = a.b_raw_sdd_min + o;
    for (int i = lane; i < N; i += 32) {
      sv[i] = (i == N - 1) ? s_end : ds * i + s_start;
      S.sd2[i] = (i == 0) ? sd_start * sd_start : ((i == N - 1) ? 0.0 : nan);
      S.sdd[i] = nan;
      S.bmax[i] = raw_max[i];
      S.bsmax[i] = raw_smax[i];
      S.bsmin[i] = raw_smin[i];
    }
    __syncwarp();

    // ---- OptimizePathParameter (:287-490)
    S.boundary_passes(fix_val, fix_idx);

    int iforw_lo = 0, iback_hi = N - 1, iback_lo, iforw_hi, icrit, icrit_lo, icrit_hi;
    iback_lo = S.add_backward(iback_hi);
    iforw_hi = S.add_forward(iforw_lo);
    icrit_hi = iback_lo;
    if ((iforw_hi < icrit_hi) && ((icrit_hi < N - 2) && (icrit_hi >= 2))) {
      S.st(&S.sd2[icrit_hi], nan);
      icrit_hi++;
      iback_lo++;
    }
    icrit_lo = iforw_hi;
    for (int loop = 0; loop < max_loops; ++loop) {
      if (iforw_hi >= icrit_hi) break;
      icrit = S.next_critical(icrit_lo, icrit_hi);
      if (icrit < 0 || icrit >= N) icrit = static_cast<int>(0.5 * (icrit_lo + icrit_hi));
      if (icrit > 0 && icrit < N - 1) S.st(&S.sd2[icrit], S.bmax[icrit]);
      if (icrit < 1) {
        status = TPB_PATH_FAIL_INDEX;
        break;
      }
      if (S.bmax[icrit - 1] <= S.bmax[icrit]) {
        iback_hi = icrit - 1;
        S.st(&S.sd2[icrit - 1], S.bmax[icrit - 1]);
      } else {
        iback_hi = icrit;
      }
      iback_lo = S.add_backward(iback_hi);
      iforw_lo = icrit;
      iforw_hi = S.add_forward(iforw_lo);
      if (iback_lo > icrit_lo) {
        status = TPB_PATH_FAIL_CONNECT;
        break;
      }
      icrit_lo = iforw_hi;
    }
  }
  if (status == TPB_PATH_OK) {
    // NaN audit + intersection fill (:398-411)
    bool any_nan = false;
    for (int i = lane; i < N; i += 32) {
      const double v = S.sd2[i];
      any_nan |= (v != v);
    }
    if (__ballot_sync(kFullMask, any_nan)) status = TPB_PATH_FAIL_NAN;
  }
  if (status == TPB_PATH_OK) {
    for (int base = 0; base < N; base += 32) {
      const int i = base + lane;
      bool need = false;
      if (i < N) {
        const double v = S.sdd[i];
        need = (v != v);
      }
      unsigned m = __ballot_sync(kFullMask, need);
      while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        S.sdd_at_intersection(base + k);
      }
    }
    if (S.valid(0, sdd_start, S.sd2[0])) S.st(&S.sdd[0], sdd_start);  // (:414-416)
    __syncwarp();
    if (S.sd2[N - 1] != 0) status = TPB_PATH_FAIL_TERMINAL;
  }
  if (status == TPB_PATH_OK) {
    for (int i = lane; i < N; i += 32) sd[i] = sqrt(S.sd2[i]);
    // last_extremal_index_ (:435-445)
    last_extremal = (N - 2 > 1) ? N - 2 : 1;
    while (last_extremal >= 1) {
      if (S.sdd[last_extremal] > 0.0 ||
          fabs(S.sd2[last_extremal] - S.bmax[last_extremal]) < kTiny) {
        break;
      }
      last_extremal--;
    }
    __syncwarp();
    // time integration (:450-467): dt per lane, serial prefix in sample order.
    double t = time_start;
    if (lane == 0) time[0] = t;
    for (int base = 1; base < N; base += 32) {
      const int i = base + lane;
      double dt = 0.0;
      if (i < N) {
        const double s2a = S.sd2[i - 1], s2b = S.sd2[i];
        if ((s2a > 0) || (s2b > 0)) {
          dt = 2.0 * S.ds / (sd[i - 1] + sd[i]);
        } else {
          dt = -1.0;  // marker: both zero
        }
      }
      double mine = 0.0;
      const int cnt = min(32, N - base);
      for (int k = 0; k < cnt; ++k) {
        const double dk = __shfl_sync(kFullMask, dt, k);
        if (dk >= 0.0) t = t + dk;
        if (k == lane) mine = t;
      }
      if (i < N) {
        time[i] = mine;
        if (dt < 0.0) {  // (:456-466) zero velocity on both sides
          S.sdd[i - 1] = 0;
          S.sdd[i] = 0;
        }
      }
    }
    __syncwarp();
  }
  if (lane == 0) {
    a.status[b] = status;
    a.last_extremal[b] = last_extremal;
  }
  __syncwarp();
  if (lane == 0) {
    a.duration[b] = (status == TPB_PATH_OK) ? (time[N - 1] - time[0]) : nan;
  }
}

// ===========================================================================
// K6: assembly + uniform-in-time resampling
// ===========================================================================

// ResampleEquidistantlyInTime sample count (:755-757)
__global__ void k_resample_count(Dims dm, DeviceArrays a, double start_sec, double dt) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B) return;
  int m = 0;
  if (a.status[b] == TPB_PATH_OK) {
    const double duration = a.time[static_cast<size_t>(b) * dm.N + dm.N - 1] - start_sec;
    const double c = ceil(duration / dt) + 1;
    m = (c > 0 && c < 2.0e9) ? static_cast<int>(c) : 0;
  }
  a.r_count[b] = m;
}

// exclusive scan of r_count into r_offset (single block; B is at most ~1e6)
__global__ void k_scan_offsets(int B, const int32_t* __restrict__ count,
                               int64_t* __restrict__ offset) {
  __shared__ long long s_part[1024];
  const int tid = threadIdx.x;
  const int per = (B + blockDim.x - 1) / blockDim.x;
  const int lo = min(B, tid * per), hi = min(B, lo + per);
  long long sum = 0;
  for (int i = lo; i < hi; ++i) sum += count[i];
  s_part[tid] = sum;
  __syncthreads();
  if (tid == 0) {
    long long run = 0;
    for (int i = 0; i < blockDim.x; ++i) {
      const long long v = s_part[i];
      s_part[i] = run;
      run += v;
    }
    offset[B] = run;
  }
  __syncthreads();
  long long run = s_part[tid];
  for (int i = lo; i < hi; ++i) {
    offset[i] = run;
    run += count[i];
  }
}

__device__ __forceinline__ double lerp(double t, double x, double y) {
  // eigenmath::InterpolateLinear is un-vendored; same choice as the oracle.
  return x + t * (y - x);
}
__device__ __forceinline__ double clampa(double v, double lim) {
  v = (v < -lim) ? -lim : v;  // cwiseMax(-amax)
  return (lim < v) ? lim : v;  // cwiseMin(amax)
}

constexpr int kResampleThreads = 128;

// One thread per packed output row (grid-stride; the row count is read from
// r_offset[B] on the device, so no host round trip).  Rows past r_capacity are
// dropped (the host reports kOutOfRange).
__global__ void __launch_bounds__(kResampleThreads)
k_resample_fill(Dims dm, DeviceArrays a, double start_sec, double dt) {
  const int64_t total = min(a.r_offset[dm.B], dm.r_cap);
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; row < total;
       row += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    // path owning this row: last b with r_offset[b] <= row
    int blo = 0, bhi = dm.B - 1;
    while (blo < bhi) {
      const int mid = (blo + bhi + 1) >> 1;
      if (a.r_offset[mid] <= row) {
        blo = mid;
      } else {
        bhi = mid - 1;
      }
    }
    const int b = blo;
    const int M = a.r_count[b];
    const int i = static_cast<int>(row - a.r_offset[b]);
    const int N = dm.N, D = dm.D;
    const size_t o = static_cast<size_t>(b) * N;
    const double* tp = a.time + o;
    const double t = start_sec + dt * i;
    // TimeAtPathSamplesLowerIndex (:686-695): first index with time[index+1] > t,
    // else N-1.  time[] is non-decreasing, so the reference's monotone linear scan
    // equals this binary search.
    int lo = 0, hi = N - 1;  // answer in [0, N-1]
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (tp[mid + 1] > t) {
        hi = mid;
      } else {
        lo = mid + 1;
      }
    }
    const int li = lo;
    const int ui = min(N - 1, li + 1);
    const double tl = tp[li], tu = tp[ui];
    const double at = (fabs(tu - tl) < DBL_EPSILON) ? 0.5 : (t - tl) / (tu - tl);
    const double sdl = a.sd[o + li], sdu = a.sd[o + ui];
    const double sddl = a.sdd[o + li], sddu = a.sdd[o + ui];
    a.r_time[row] = t;
    a.r_s[row] = lerp(at, a.s[o + li], a.s[o + ui]);
    a.r_sd[row] = lerp(at, sdl, sdu);
    a.r_sdd[row] = lerp(at, sddl, sddu);
    const bool last = (i == M - 1);
    const double sdl2 = sdl * sdl, sdu2 = sdu * sdu;  // eigenmath::Square
    for (int d = 0; d < D; ++d) {
      const double am = a.amax[static_cast<size_t>(b) * D + d];
      const size_t rl = (o + li) * D + d, ru = (o + ui) * D + d;
      double pos, vel, acc;
      if (last) {  // (:780-782)
        pos = a.q[(o + N - 1) * D + d];
        vel = 0.0;
        acc = 0.0;
      } else {
        const double d1l = a.dq[rl], d1u = a.dq[ru];
        const double d2l = a.ddq[rl], d2u = a.ddq[ru];
        // per-sample velocity / clamped acceleration (:458-472)
        const double vl = d1l * sdl, vu = d1u * sdu;
        const double al = clampa(d1l * sddl + d2l * sdl2, am);
        const double au = clampa(d1u * sddu + d2u * sdu2, am);
        pos = lerp(at, a.q[rl], a.q[ru]);
        vel = lerp(at, vl, vu);
        acc = clampa(lerp(at, al, au), am);
      }
      a.r_pos[row * D + d] = pos;
      a.r_vel[row * D + d] = vel;
      a.r_acc[row * D + d] = acc;
    }
  }
}

__global__ void k_fill_f64(double* p, double v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
__global__ void k_fill_i32(int32_t* p, int32_t v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

}  // namespace

// ===========================================================================
// launchers
// ===========================================================================

int launch_fill_f64(double* p, double v, int n, cudaStream_t st) {
  k_fill_f64<<<(n + 255) / 256, 256, 0, st>>>(p, v, n);
  return 1;
}
int launch_fill_i32(int32_t* p, int32_t v, int n, cudaStream_t st) {
  k_fill_i32<<<(n + 255) / 256, 256, 0, st>>>(p, v, n);
  return 1;
}

int launch_joint_fit(const Dims& d, const DeviceArrays& a, double rounding, bool auto_delta,
                     cudaStream_t st) {
  const int threads = 64;
  k_joint_fit<<<(d.B + threads - 1) / threads, threads, 0, st>>>(d, a, rounding, auto_delta);
  return 1;
}

int launch_joint_sample_limit(const Dims& d, const DeviceArrays& a, double safety,
                              cudaStream_t st) {
  const size_t smem =
      sizeof(double) * (static_cast<size_t>(d.max_points) + 3 + static_cast<size_t>(d.max_points) * d.D + 2 * d.D);
  dim3 grid((d.N + kSampleThreads - 1) / kSampleThreads, d.B);
  k_joint_sample_limit<<<grid, kSampleThreads, smem, st>>>(d, a, safety);
  return 1;
}

int launch_generic_limit(const Dims& d, const DeviceArrays& a, cudaStream_t st) {
  dim3 grid((d.N + kSampleThreads - 1) / kSampleThreads, d.B);
  k_generic_limit<<<grid, kSampleThreads, 0, st>>>(d, a);
  return 1;
}

int launch_integrate(const Dims& d, const DeviceArrays& a, Mode mode, double safety,
                     int max_loops, cudaStream_t st) {
  const int blocks = (d.B + kWarpsPerBlock - 1) / kWarpsPerBlock;
  k_integrate<<<blocks, kWarpsPerBlock * 32, 0, st>>>(d, a, mode == Mode::kJoint, safety,
                                                       max_loops);
  return 1;
}

int launch_resample(const Dims& d, const DeviceArrays& a, double start_sec, double dt,
                    int num_sms, cudaStream_t st) {
  const int threads = 128;
  k_resample_count<<<(d.B + threads - 1) / threads, threads, 0, st>>>(d, a, start_sec, dt);
  k_scan_offsets<<<1, 1024, 0, st>>>(d.B, a.r_count, a.r_offset);
  k_resample_fill<<<num_sms * 8, kResampleThreads, 0, st>>>(d, a, start_sec, dt);
  return 3;
}

}  // namespace tpb
