// Stage C, sequential part: one warp per path (k_integrate<Policy>).
//
// CalculateBoundary passes 2-4 (time_optimal_path_timing.cc:1379-1484), the
// extremal sweeps AddForwardExtremal / AddBackwardExtremal (:769-952), the
// switching-point loop of OptimizePathParameter (:287-490), NextCriticalPoint
// (:697-720), ComputeSddAtIntersection (:722-751) and the time integration
// (:450-467).
//
// The scalar state machine is replicated in every lane (warp-uniform control
// flow); lanes split the O(C^2) work of FindSddMax/FindSddMin/
// AreDerivativesValid (:624-695) -- how, is the Policy:
//
//   Joint8Policy<D> joint-space paths with D <= 8.  Lane = 8*g + d: group g of 8
//                  lanes tests one candidate per round against acceleration row
//                  d (four candidates per round).  Rows (q', q'' of one sample) live in registers and are
//                  prefetched one step ahead of the sweep; the 2D velocity rows
//                  have A == 0, so they never yield a candidate and their validity
//                  does not depend on the candidate (tested once per call).
//   GenericPolicy  explicit constraint tables (C <= kMaxRows = 64) and joint paths with
//                  D > 8: rows of the current sample staged per warp in shared
//                  memory, lane j owns candidate (row j/2, bound j&1).
//
// Both give results bit-identical to the reference's sequential loops: the
// extremum over the candidates that pass every row's test does not depend on
// evaluation order.
#pragma once

#include "tpb_engine.h"
#include "tpb_device.cuh"

namespace tpb {

constexpr int kWarpsPerBlock = 4;

// ---------------------------------------------------------------------------
// Policies
// ---------------------------------------------------------------------------

struct GenericPolicy {
  struct Rows {
    int idx;
  };
  int lane, C, D;
  bool joint;
  const double *dq, *ddq, *cons;
  double *sA, *sB, *sLo, *sHi;  // shared, this warp
  int staged;

  __device__ __forceinline__ Rows load(int idx) const { return Rows{idx}; }

  // Bring the rows of sample idx on chip.  Joint paths synthesise them from
  // q', q'' (ConstraintSetup); their lower/upper bounds were filled once.
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
  __device__ bool valid(Rows r, double sddv, double sd2v) {
    stage(r.idx);
    bool bad = false;
    for (int r = lane; r < C; r += 32) {  // C <= kMaxRows = 64: up to two rows per lane
      const double v = sA[r] * sddv + sB[r] * sd2v;
      bad = bad || (v + kTiny < sLo[r]) || (v - kTiny > sHi[r]);
    }
    return __ballot_sync(kFullMask, bad) == 0u;
  }

  // FindSddMax / FindSddMin (:638-695)
  template <bool kMax>
  __device__ double find_sdd(Rows rw, double sd2v) {
    stage(rw.idx);
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
};

// kD = number of dofs (1..8), compile-time so that the candidate rounds unroll.
// Padding lanes (d >= kD) hold a copy of row kD-1 and of its limits: a row tested
// twice changes no outcome, so no lane predicate is needed anywhere, and the
// copies keep real operands in the divisions (0/1 would take the slow path of
// the double-precision division).
//
// kCart: the rows of a Cartesian path (timeable_path_cartesian_spline.cc:551-595) are
// the joint rows plus two more of the velocity kind, A = 0, B = |J q'|^2 of the
// translational / rotational part, -v^2 <= . <= v^2.  The padding lanes d == kD of
// groups 0 and 1 (kD <= 7) carry them: their velocity-kind coefficient V is read from
// the stored constraint rows instead of being q'^2, and their bounds are those rows'.
template <int kD, bool kCart = false>
struct Joint8Policy {
  struct Rows {
    double A, B;  // q'_d, q''_d of one sample (d >= kD: a copy of row kD-1)
    double V;     // kCart: B of this lane's velocity-kind row (else unused: it is A*A)
  };
  int lane;
  int g, d;  // lane = 8*g + d
  const double *dq, *ddq;  // this path's rows, already offset by min(d, kD-1)
  double aS, vS2;  // amax_d*safety, (vmax_d*safety)^2 (row bounds: -aS..aS, 0..vS2)
  // kCart only: lanes with xsel read V from bx[idx * xstride]; vlo is their lower bound
  // (0 for the joint velocity rows), vS2 their upper one
  const double* bx;
  int xstride;
  bool xsel;
  double vlo;

  __device__ __forceinline__ Rows load(int idx) const {
    Rows r;
    r.A = dq[idx * kD];
    r.B = ddq[idx * kD];
    if (kCart) {
      r.V = r.A * r.A;
      if (xsel) r.V = bx[static_cast<size_t>(idx) * xstride];
    }
    return r;
  }
  __device__ __forceinline__ double vcoef(const Rows& r) const { return kCart ? r.V : r.A * r.A; }
  __device__ __forceinline__ double vlower() const { return kCart ? vlo : 0.0; }

  // AreDerivativesValid (:624-636)
  __device__ __forceinline__ bool valid(const Rows& r, double sddv, double sd2v) const {
    const double v = r.A * sddv + r.B * sd2v;
    bool bad = outside_sym(v, aS);
    // velocity row (A = 0, B = q'^2, 0 <= . <= vS2): 0*sdd contributes +-0 (or NaN
    // for a non-finite sdd, which then passes every comparison like in the reference)
    const double w = 0.0 * sddv + vcoef(r) * sd2v;
    bad = bad || (w + kTiny < vlower()) || (w - kTiny > vS2);
    return __ballot_sync(kFullMask, bad) == 0u;
  }

  // FindSddMax / FindSddMin (:638-695).  Velocity rows never yield a candidate
  // (A == 0 exactly) and their validity does not depend on a finite candidate, so
  // they are tested once; a non-finite candidate already fails its own row.
  //
  // Row d has two candidates, (-aS - Bs)/A and (aS - Bs)/A; division by A is
  // monotone, so which of the two is the larger one follows from the sign of A.
  // Even groups compute the row's "outer" candidate (the larger one for kMax,
  // the smaller one for kMin), odd groups the other.  The outer candidates of all
  // rows are tested first (two rounds of four); the inner ones can only matter
  // if one of them lies beyond the best valid outer candidate, which a single
  // ballot decides -- then (rarely: infeasible or degenerate samples) they are
  // tested as well.  The extremum over all valid candidates is what the
  // reference's sequential loop returns, whatever the order of evaluation.
  template <bool kMax>
  __device__ __forceinline__ double find_sdd(const Rows& r, double sd2v) const {
    const double Bs = r.B * sd2v;
    const bool pos = r.A > 0.0;
    const bool plus = ((g & 1) == 0) ? (kMax == pos) : (kMax != pos);
    const double num = (plus ? aS : -aS) - Bs;
    // rows without a candidate (|A| < kTiny) divide 1 by 1 instead: x/0 and 0/x
    // would send the whole warp through the division's slow path
    const bool none = is_tiny(r.A);
    // a row exactly on its limit (num == 0, common on the limit curve) yields a
    // zero with the quotient's sign: num * A
    const bool zero = num == 0.0;
    const double quot = guarded_div(none || zero, num, r.A);
    const double cand = none ? quiet_nan() : (zero ? num * r.A : quot);
    const double vv = vcoef(r) * sd2v;
    const bool vbad =
        __ballot_sync(kFullMask, (vv + kTiny < vlower()) || (vv - kTiny > vS2)) != 0u;
    double best = kMax ? -DBL_MAX : DBL_MAX;
    // one round: group g tests the candidate held by lane (src_group, row) against
    // acceleration row d
    auto round = [&](int row, int src_group) {
      const double c = __shfl_sync(kFullMask, cand, (src_group << 3) | (row & 7));
      const double v = r.A * c + Bs;
      const bool bad = outside_sym(v, aS);
      const unsigned m = __ballot_sync(kFullMask, bad);
      // NaN marks "no candidate" (and fails c == c)
      const bool ok = (row < kD) && (c == c) && (((m >> (8 * g)) & 0xffu) == 0u);
      const double pick = kMax ? ((c > best) ? c : best) : ((c < best) ? c : best);
      best = ok ? pick : best;
    };
    auto reduce = [&]() {
#pragma unroll
      for (int o = 8; o <= 16; o <<= 1) {
        const double w = __shfl_xor_sync(kFullMask, best, o);
        best = kMax ? ((w > best) ? w : best) : ((w < best) ? w : best);
      }
    };
#pragma unroll
    for (int j0 = 0; j0 < kD; j0 += 4) round(j0 + g, 0);
    reduce();
    const bool beyond = (g & 1) && (cand == cand) && (kMax ? (cand > best) : (cand < best));
    if (__ballot_sync(kFullMask, beyond) != 0u) {
#pragma unroll
      for (int j0 = 0; j0 < kD; j0 += 4) round(j0 + g, 1);
      reduce();
    }
    if (vbad || best == (kMax ? -DBL_MAX : DBL_MAX)) best = 0.0;
    return best;
  }
};

// ---------------------------------------------------------------------------
// The per-path solver
// ---------------------------------------------------------------------------

template <class P>
struct Solver {
  using Rows = typename P::Rows;
  P p;
  int lane, N;
  double ds;
  double *sd2, *sdd, *bmax, *bsmax, *bsmin, *bzero;
  uint8_t *at0, *type;
#ifdef TPB_PROFILE_PHASES
  int nsteps = 0;  // extremal steps taken (diagnostics)
#endif

  __device__ __forceinline__ void st(double* ptr, double v) const {
    if (lane == 0) *ptr = v;
    __syncwarp();
  }

  // ComputeSddAtIntersection (:722-751)
  __device__ void sdd_at_intersection(int index, const Rows& rows) {
    double cand[3];
    int n = 0;
    if (index > 0 && index < N - 1) cand[n++] = 0.25 / ds * (sd2[index + 1] - sd2[index - 1]);
    if (index < N - 1) cand[n++] = 0.5 / ds * (sd2[index + 1] - sd2[index]);
    if (index > 0) cand[n++] = 0.5 / ds * (sd2[index] - sd2[index - 1]);
    const double here = sd2[index];
    double r = 0.0;
    for (int k = 0; k < n; ++k) {
      if (p.valid(rows, cand[k], here)) {
        r = cand[k];
        break;
      }
    }
    st(&sdd[index], r);
  }

  // One (rare) pass-2 fix of CalculateBoundary at sample i (:1384-1396).
  __device__ void fix_isolated(int i) {
    const double zm = bzero[i - 1];
    const Rows rm = p.load(i - 1);
    const double a1 = p.template find_sdd<true>(rm, zm);
    const double a2 = p.template find_sdd<false>(rm, zm);
    const double zp = bzero[i + 1];
    const Rows rp = p.load(i + 1);
    const double a3 = p.template find_sdd<true>(rp, zp);
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

  // "skipped maximum" value at sample i (:1422-1429)
  __device__ double skipped_value(int i, double bm, double bp) {
    const Rows rm = p.load(i - 1), rp = p.load(i + 1);  // both in flight together
    const double bz = bzero[i];
    const double fw = bm + 2.0 * ds * p.template find_sdd<true>(rm, bm);
    const double bw = bp - 2.0 * ds * p.template find_sdd<false>(rp, bp);
    double m = bz;
    if (fw < m) m = fw;
    if (bw < m) m = bw;
    return (0.0 < m) ? m : 0.0;
  }

  // CalculateBoundary passes 2-4 (:1379-1484).  Pass 2 mutates the neighbours of
  // an isolated critical point (flags at0 = 0,1,0 around i) in place while it
  // scans, and later iterations read the mutated values.  The scan therefore runs
  // 32 samples at a time where that is equivalent to the serial order -- a chunk
  // without an isolated point: everything before it has been applied, nothing
  // inside it mutates -- and sample by sample (warp-uniform) through the rare
  // chunks that contain one.
  __device__ void boundary_passes(double* fix_val, int* fix_idx) {
    if (lane == 0) {
      type[0] = kNone;
      type[N - 1] = kNone;
    }
    int nwork = 0;
    auto test_v = [&](double bm, double bi, double bp, double smax_i, double smin_i,
                      double smin_next) {
      const double sd2p = div_pos(bp - bi, ds);
      const bool sink_or_source = (sd2p < 2 * smin_i) || (sd2p > 2 * smax_i);
      const bool skipped_sdd = (smax_i > 0) && (smin_next < 0);
      const bool skipped_sd2 = (bi > bm - kTiny) && (bi > bp - kTiny);
      return (skipped_sd2 || skipped_sdd) && sink_or_source;
    };
    auto test = [&](int i, double bm, double bi, double bp) {
      return test_v(bm, bi, bp, bsmax[i], bsmin[i], bsmin[i + 1]);
    };
    auto record = [&](int ii, double bm, double bp) {
      const double v = skipped_value(ii, bm, bp);
      if (lane == 0) {
        fix_idx[nwork] = ii;
        fix_val[nwork] = v;
      }
      ++nwork;
    };
    // kGroup chunks at a time: the loads of a whole group are in flight together
    // (one memory round trip per group instead of two per chunk).
    constexpr int kGroup = 2;
    for (int base0 = 1; base0 < N - 1; base0 += 32 * kGroup) {
      bool iso = false;
#pragma unroll
      for (int u = 0; u < kGroup; ++u) {
        const int i = base0 + 32 * u + lane;
        if (i < N - 1) iso |= !at0[i - 1] && at0[i] && !at0[i + 1];
      }
      if (__ballot_sync(kFullMask, iso) == 0u) {
        // nothing in this group mutates the curve
        double bm[kGroup], bi[kGroup], bp[kGroup], sx[kGroup], sn[kGroup], sq[kGroup];
#pragma unroll
        for (int u = 0; u < kGroup; ++u) {
          const int i = base0 + 32 * u + lane;
          if (i < N - 1) {
            bm[u] = bmax[i - 1];
            bi[u] = bmax[i];
            bp[u] = bmax[i + 1];
            sx[u] = bsmax[i];
            sn[u] = bsmin[i];
            sq[u] = bsmin[i + 1];
          }
        }
#pragma unroll
        for (int u = 0; u < kGroup; ++u) {
          const int base = base0 + 32 * u;
          const bool hit = (base + lane < N - 1) && test_v(bm[u], bi[u], bp[u], sx[u], sn[u], sq[u]);
          unsigned m = __ballot_sync(kFullMask, hit);
          while (m) {
            const int k = __ffs(m) - 1;
            m &= m - 1;
            const int ii = base + k;
            record(ii, bmax[ii - 1], bmax[ii + 1]);
          }
        }
        continue;
      }
      for (int base = base0; base < min(base0 + 32 * kGroup, N - 1); base += 32) {
        const int i = base + lane;
        const bool in = i < N - 1;
        const bool iso1 = in && !at0[i - 1] && at0[i] && !at0[i + 1];
        if (__ballot_sync(kFullMask, iso1) == 0u) {
          bool hit = false;
          if (in) hit = test(i, bmax[i - 1], bmax[i], bmax[i + 1]);
          unsigned m = __ballot_sync(kFullMask, hit);
          while (m) {
            const int k = __ffs(m) - 1;
            m &= m - 1;
            const int ii = base + k;
            record(ii, bmax[ii - 1], bmax[ii + 1]);
          }
        } else {
          const int top = min(base + 32, N - 1);
          for (int ii = base; ii < top; ++ii) {
            if (!at0[ii - 1] && at0[ii] && !at0[ii + 1]) fix_isolated(ii);
            const double bm1 = bmax[ii - 1], bp1 = bmax[ii + 1];
            if (test(ii, bm1, bmax[ii], bp1)) record(ii, bm1, bp1);
          }
          __syncwarp();
        }
      }
    }
    __syncwarp();
    // deferred fixes (:1432-1451), in list order
    for (int w = 0; w < nwork; ++w) {
      const int index = fix_idx[w];
      const double value = fix_val[w];
      // the three samples' rows and levels in flight together
      const int im = max(index - 1, 0), ip = min(index + 1, N - 1);
      const Rows r0 = p.load(index), rlo = p.load(im), rhi = p.load(ip);
      const double zlo = bzero[im], zhi = bzero[ip];
      const double a1 = p.template find_sdd<true>(r0, value);
      const double a2 = p.template find_sdd<false>(r0, value);
      if (lane == 0) {
        bmax[index] = value;
        bsmax[index] = a1;
        bsmin[index] = a2;
      }
      if (index > 0) {
        const double z = zlo;
        const Rows r1 = rlo;
        const double c1 = p.template find_sdd<true>(r1, z);
        const double c2 = p.template find_sdd<false>(r1, z);
        if (lane == 0) {
          bmax[index - 1] = z;
          bsmax[index - 1] = c1;
          bsmin[index - 1] = c2;
        }
      }
      if (index < N - 1) {
        const double z = zhi;
        const Rows r1 = rhi;
        const double c1 = p.template find_sdd<true>(r1, z);
        const double c2 = p.template find_sdd<false>(r1, z);
        if (lane == 0) {
          bmax[index + 1] = z;
          bsmax[index + 1] = c1;
          bsmin[index + 1] = c2;
        }
      }
      __syncwarp();
    }
    // classification (:1453-1484), lane-parallel
    for (int i0 = 1 + lane; i0 < N - 1; i0 += 32 * kGroup) {
      double b1[kGroup], b0[kGroup], mn[kGroup], mx[kGroup];
#pragma unroll
      for (int u = 0; u < kGroup; ++u) {
        const int i = i0 + 32 * u;
        if (i < N - 1) {
          b1[u] = bmax[i + 1];
          b0[u] = bmax[i];
          mn[u] = bsmin[i];
          mx[u] = bsmax[i];
        }
      }
#pragma unroll
      for (int u = 0; u < kGroup; ++u) {
        const int i = i0 + 32 * u;
        if (i < N - 1) {
          const double sd2p = div_pos(b1[u] - b0[u], ds);
          const double sd2p_min = 2 * mn[u];
          const double sd2p_max = 2 * mx[u];
          uint8_t t = kNone;
          if (sd2p < sd2p_min) {
            t = kSink;
          } else if (sd2p > sd2p_max) {
            t = kSource;
          }
          if ((sd2p <= sd2p_max) && (sd2p >= sd2p_min)) t = kTrajectory;
          type[i] = t;
        }
      }
    }
    __syncwarp();
  }

  // AddForwardExtremal (:769-857).  The data of step idx+1 (rows, sd2_max,
  // type, previously written sd2) is requested while step idx computes.
  __device__ int add_forward(int idx_lo) {
    if (idx_lo >= N - 2) return N - 1;
    double cur = sd2[idx_lo];
    double bcur = bmax[idx_lo];
    uint8_t tcur = type[idx_lo];
    Rows rows = p.load(idx_lo);
    double bnext = bmax[idx_lo + 1];
    uint8_t tnext = type[idx_lo + 1];
    double nxt = sd2[idx_lo + 1];
    for (int idx = idx_lo; idx < N - 2; ++idx) {
      // prefetch for step idx + 1 (indices stay in range: idx + 2 <= N - 1)
      const Rows rows_n = p.load(idx + 1);
      const double bnext_n = bmax[idx + 2];
      const uint8_t tnext_n = type[idx + 2];
      const double nxt_n = sd2[idx + 2];
#ifdef TPB_PROFILE_PHASES
      ++nsteps;
#endif

      double sd2tmp, sddtmp;
      const bool on_boundary = is_tiny(cur - bcur);
      if (on_boundary && (tcur & kTrajectory) && (tnext & kTrajectory)) {
        sd2tmp = bnext;
        sddtmp = div_pos(0.5 * (sd2tmp - cur), ds);
      } else {
        sddtmp = p.template find_sdd<true>(rows, cur);
        sd2tmp = cur + 2.0 * ds * sddtmp;
      }
      if (!(nxt != nxt) && (nxt < sd2tmp)) {
        sdd_at_intersection(idx, rows);
        return N - 1;
      }
      if (sd2tmp > bnext) {
        const double sdd_bound = div_pos(0.5 * (bnext - cur), ds);
        const bool deriv_invalid = !p.valid(rows, sdd_bound, bcur);
        const bool type_invalid = tnext & kSink;
        if (type_invalid || deriv_invalid) return idx;
        sd2tmp = bnext;
        sddtmp = sdd_bound;
      }
      if (sd2tmp < 0) {
        sd2tmp = 0.0;
        if (idx <= 1) {
          sddtmp = 0.0;
        } else {
          sddtmp = div_pos(-sd2[idx - 1], ds);
        }
      }
      if (lane == 0) {
        sd2[idx + 1] = sd2tmp;
        sdd[idx] = sddtmp;
      }
      __syncwarp();
      cur = sd2tmp;
      bcur = bnext;
      tcur = tnext;
      rows = rows_n;
      bnext = bnext_n;
      tnext = tnext_n;
      nxt = nxt_n;
    }
    return N - 1;
  }

  // AddBackwardExtremal (:859-952)
  __device__ int add_backward(int idx_hi) {
    if (idx_hi <= 1) return 0;
    double cur = sd2[idx_hi];
    double bcur = bmax[idx_hi];
    uint8_t tcur = type[idx_hi];
    Rows rows = p.load(idx_hi);
    double bprev = bmax[idx_hi - 1];
    uint8_t tprev = type[idx_hi - 1];
    double prv = sd2[idx_hi - 1];
    const bool is_connecting = (idx_hi != (N - 1));
    for (int idx = idx_hi; idx > 1; --idx) {
      // prefetch for step idx - 1 (idx - 2 >= 0)
      const Rows rows_n = p.load(idx - 1);
      const double bprev_n = bmax[idx - 2];
      const uint8_t tprev_n = type[idx - 2];
      const double prv_n = sd2[idx - 2];
#ifdef TPB_PROFILE_PHASES
      ++nsteps;
#endif

      double sd2tmp, sddtmp;
      const bool on_boundary = is_tiny(cur - bcur);
      if (on_boundary && (tcur & kTrajectory) && (tprev & kTrajectory)) {
        sd2tmp = bprev;
        sddtmp = div_pos(0.5 * (cur - sd2tmp), ds);
      } else {
        sddtmp = p.template find_sdd<false>(rows, cur);
        sd2tmp = cur - 2.0 * ds * sddtmp;
      }
      if (!(prv != prv) && (prv < sd2tmp)) {
        sdd_at_intersection(idx, rows);
        return 0;
      }
      if (sd2tmp > bprev) {
        const double sdd_bound = div_pos(0.5 * (cur - bprev), ds);
        const bool deriv_invalid = !p.valid(rows, sdd_bound, cur);
        const bool type_invalid = tprev & kSource;
        if ((type_invalid || deriv_invalid) && !is_connecting) return idx;
        sd2tmp = bprev;
        sddtmp = sdd_bound;
      }
      if (sd2tmp < 0) {
        sd2tmp = 0.0;
        if (idx < N - 1) {
          sddtmp = div_pos(sd2[idx + 1], ds);
        } else {
          sddtmp = 0.0;
        }
      }
      if (lane == 0) {
        sd2[idx - 1] = sd2tmp;
        sdd[idx] = sddtmp;
      }
      __syncwarp();
      cur = sd2tmp;
      bcur = bprev;
      tcur = tprev;
      rows = rows_n;
      bprev = bprev_n;
      tprev = tprev_n;
      prv = prv_n;
    }
    return 0;
  }

  // Lookup tables for NextCriticalPoint, built once after the boundary passes
  // (type[] and sd2_max[] do not change afterwards):
  //   next_src[i] = smallest j >= i whose type is Source or Trajectory, else N
  //   prev_eq0[i] = largest j <= i with sd2_max[j] == sd2_max_for_sdd0[0], else -1
  __device__ void build_critical_tables(int* next_src, int* prev_eq0) const {
    const double bz0 = bzero[0];
    constexpr int kGroup = 4;  // chunks whose loads are in flight together
    int carry = N;
    for (int base0 = ((N - 1) / 32) * 32; base0 >= 0; base0 -= 32 * kGroup) {
      bool f[kGroup];
#pragma unroll
      for (int u = 0; u < kGroup; ++u) {
        const int i = base0 - 32 * u + lane;
        f[u] = (i >= 0) && (i < N) && ((type[i] & (kSource | kTrajectory)) != 0);
      }
#pragma unroll
      for (int u = 0; u < kGroup; ++u) {
        const int base = base0 - 32 * u;
        const int i = base + lane;
        const unsigned m = __ballot_sync(kFullMask, f[u]);
        const unsigned mm = m & (~0u << lane);
        if (i >= 0 && i < N) next_src[i] = mm ? base + (__ffs(mm) - 1) : carry;
        if (m) carry = base + (__ffs(m) - 1);
      }
    }
    carry = -1;
    for (int base0 = 0; base0 < N; base0 += 32 * kGroup) {
      bool f[kGroup];
#pragma unroll
      for (int u = 0; u < kGroup; ++u) {
        const int i = base0 + 32 * u + lane;
        f[u] = (i < N) && (bmax[i] == bz0);
      }
#pragma unroll
      for (int u = 0; u < kGroup; ++u) {
        const int base = base0 + 32 * u;
        const int i = base + lane;
        const unsigned m = __ballot_sync(kFullMask, f[u]);
        const unsigned mm = m & (~0u >> (31 - lane));
        if (i < N) prev_eq0[i] = mm ? base + (31 - __clz(mm)) : carry;
        if (m) carry = base + (31 - __clz(m));
      }
    }
    __syncwarp();
  }

  // NextCriticalPoint (:697-720) in O(1).  The sequential scan (kept below as
  // next_critical_scan, and cross-checked against this one in -DTPB_CHECK_SCAN
  // builds) returns at the first sample, at or after the first Source/Trajectory
  // sample q of (idx_lo, idx_hi], whose sd2 is set.  In the switching loop
  // idx_lo is where the last forward extremal stopped and idx_hi the low end of
  // the first backward extremal: sd2 is set on [0, idx_lo] and [idx_hi, N-1] and
  // NaN strictly between (sweeps only ever extend the lower range upwards; the
  // backward sweep from a critical point fills down to the lower range because a
  // connecting sweep ends only on an intersection with set samples).  So the
  // scan always ends at idx_hi, and returns the last sample of (q, idx_hi] on
  // the sdd0 level if there is one, else q.
  __device__ int next_critical(int idx_lo, int idx_hi, const int* next_src,
                               const int* prev_eq0) const {
    if (idx_lo + 1 > idx_hi) return -1;
    const int q = next_src[idx_lo + 1];
    if (q > idx_hi) return -1;
    const int p2 = prev_eq0[idx_hi];
    return (p2 > q) ? p2 : q;
  }

  // NextCriticalPoint (:697-720), 32 samples per step via ballots.  Sequential
  // meaning: crit = first idx with a Source/Trajectory type; afterwards crit
  // moves to any idx with sd2_max[idx] == sd2_max_for_sdd0[0] (sic, index 0);
  // return crit at the first idx (at or after the first crit) whose sd2 is set.
  __device__ int next_critical_scan(int idx_lo, int idx_hi) const {
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
        const int q = __ffs(m1) - 1;
        crit = base + q;
        // f2 is only consulted strictly after q; f3 from q on.
        m2 = (q == 31) ? 0u : (m2 & (~0u << (q + 1)));
        m3 &= (~0u << q);
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

// ---------------------------------------------------------------------------
// Kernel: SetupProblem checks, OptimizePathParameter, outputs
// ---------------------------------------------------------------------------

__device__ __forceinline__ void init_policy(GenericPolicy& p, const Dims& dm,
                                            const DeviceArrays& a, int b, int lane, double safety,
                                            double* smem_rows, bool joint, int* status) {
  const size_t o = static_cast<size_t>(b) * dm.N;
  p.lane = lane;
  p.C = dm.C;
  p.D = dm.D;
  p.joint = joint;
  p.dq = joint ? a.dq + o * dm.D : nullptr;
  p.ddq = joint ? a.ddq + o * dm.D : nullptr;
  p.cons = joint ? nullptr : a.cons + o * dm.C * 4;
  p.sA = smem_rows;
  p.sB = smem_rows + kMaxRows;
  p.sLo = smem_rows + 2 * kMaxRows;
  p.sHi = smem_rows + 3 * kMaxRows;
  p.staged = -1;
  bool bad = false;
  if (joint) {
    const int D = dm.D;
    if (lane < D) {
      const double aS = a.amax[static_cast<size_t>(b) * D + lane] * safety;
      const double vs = a.vmax[static_cast<size_t>(b) * D + lane] * safety;
      const double vS2 = vs * vs;
      p.sLo[lane] = -aS;
      p.sHi[lane] = aS;
      p.sLo[D + lane] = 0.0;
      p.sHi[D + lane] = vS2;
      bad = (-aS >= aS) || (0.0 >= vS2);
    }
  } else {
    // IsSetupValid (:554-561): every row needs lower < upper
    const double* c = p.cons;
    const size_t rows = static_cast<size_t>(dm.N) * dm.C;
    for (size_t r = lane; r < rows; r += 32) bad |= (c[r * 4 + 2] >= c[r * 4 + 3]);
  }
  if (__ballot_sync(kFullMask, bad)) *status = TPB_PATH_FAIL_SETUP;
  __syncwarp();
}

template <int kD, bool kCart>
__device__ __forceinline__ void init_policy(Joint8Policy<kD, kCart>& p, const Dims& dm,
                                            const DeviceArrays& a, int b, int lane, double safety,
                                            double*, bool, int* status) {
  const size_t o = static_cast<size_t>(b) * dm.N;
  p.lane = lane;
  p.g = lane >> 3;
  p.d = lane & 7;
  const int dd = min(p.d, kD - 1);  // padding lanes copy the last row
  p.dq = a.dq + o * kD + dd;
  p.ddq = a.ddq + o * kD + dd;
  p.aS = a.amax[static_cast<size_t>(b) * kD + dd] * safety;
  const double vs = a.vmax[static_cast<size_t>(b) * kD + dd] * safety;
  p.vS2 = vs * vs;
  p.bx = nullptr;
  p.xstride = 0;
  p.xsel = false;
  p.vlo = 0.0;
  if (kCart) {
    static_assert(!kCart || kD <= 7, "the extra rows ride on the padding lanes");
    // rows 2*kD + g of the stored constraints, [sample][C][4] = (A, B, lower, upper)
    p.xsel = (p.d == kD) && (p.g < 2);
    if (p.xsel) {
      const double* row0 = a.cons + (o * dm.C + 2 * kD + p.g) * 4;
      p.bx = row0 + 1;
      p.xstride = dm.C * 4;
      p.vlo = row0[2];
      p.vS2 = row0[3];
    }
  }
  const bool bad = (-p.aS >= p.aS) || (p.vlo >= p.vS2);
  if (__ballot_sync(kFullMask, bad)) *status = TPB_PATH_FAIL_SETUP;
}

// kPhases selects what one launch does: kPhaseSetup (SetSetupDone + boundary
// passes 2-4), kPhaseOptimize (extremal sweeps + switching loop), kPhaseFinalize
// (audits, intersection fill, time integration, outputs).  All state between
// them lives in the per-path arrays and a.status[], so a phase can run in a
// launch of its own (the two-paths-per-warp optimiser of tpb_integrate_pair.cuh
// replaces kPhaseOptimize only).
constexpr int kPhaseSetup = 1, kPhaseOptimize = 2, kPhaseFinalize = 4, kPhaseAll = 7;

template <class P, int kMinBlocks, int kPhases = kPhaseAll>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, kMinBlocks)
k_integrate(Dims dm, DeviceArrays a, bool joint, double safety, int max_loops) {
  __shared__ double s_rows[std::is_same<P, GenericPolicy>::value ? kWarpsPerBlock * 4 * kMaxRows : 1];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * kWarpsPerBlock + warp;
  if (b >= dm.B) return;
  if (a.active != nullptr && a.active[b] == 0) return;  // moving-horizon planner: not in this window
  const int N = dm.N;
  const size_t o = static_cast<size_t>(b) * N;

  int status = TPB_PATH_OK;
  Solver<P> S;
  init_policy(S.p, dm, a, b, lane, safety,
                 s_rows + (std::is_same<P, GenericPolicy>::value ? warp * 4 * kMaxRows : 0), joint,
                 &status);
  S.lane = lane;
  S.N = N;
  S.sd2 = a.sd2 + o;
  S.sdd = a.sdd + o;
  S.bmax = a.b_sd2_max + o;
  S.bsmax = a.b_sdd_max + o;
  S.bsmin = a.b_sdd_min + o;
  S.bzero = a.b_sd2_zero + o;
  S.at0 = a.b_at0 + o;
  S.type = a.b_type + o;
  double* time = a.time + o;
  double* sv = a.s + o;
  double* sd = a.sd + o;
  double* fix_val = a.fix_val + o;
  // time[] is free until the end: first the fix list of the boundary passes, then
  // the two NextCriticalPoint tables (N ints each)
  int* fix_idx = reinterpret_cast<int*>(a.time + o);
  int* next_src = fix_idx;
  int* prev_eq0 = fix_idx + N;

  const double sd_start = a.sd_start[b];
  const double sdd_start = a.sdd_start[b];
  const double time_start = a.time_start[b];
  double s_start, s_end;
  // SetupProblem (:161-203); joint paths: ComputeTimingProfile sets
  // path_horizon_ = path_start_ + delta * (N - 1) (path_timing_trajectory.cc:341-342)
  if (joint) {
    s_start = a.path_start[b];
    s_end = s_start + a.delta[b] * (N - 1);
  } else {
    s_start = a.s_start[b];
    s_end = a.s_end[b];
  }
  if (s_start >= s_end || sd_start < 0 || N < 2) status = TPB_PATH_FAIL_SETUP;
  // a later phase of a split run continues with the status the earlier ones left
  if ((kPhases & kPhaseSetup) == 0 && status == TPB_PATH_OK) status = a.status[b];

  int last_extremal = 0;
  const double nan = quiet_nan();
#ifdef TPB_PROFILE_PHASES
  long long ph[8];
  long long acc_nc = 0, acc_bw = 0, acc_fw = 0, acc_max = 0;
  int steps_bw = 0, steps_fw = 0;
  int nloops = 0;
  for (int k = 0; k < 8; ++k) ph[k] = 0;
  ph[0] = clock64();
#define TPB_STAMP(k) ph[k] = clock64()
#else
#define TPB_STAMP(k)
#endif
  if (status == TPB_PATH_OK) {
    // SetSetupDone (:535-552)
    const double ds = (s_end - s_start) / (N - 1);
    S.ds = ds;
    // The per-sample limit curve is kept as computed (b_raw_*); passes 2-4 mutate
    // a working copy, so tpb_optimize can be repeated like the reference's
    // OptimizePathParameter (which recomputes the boundary).
    if (kPhases & kPhaseSetup) {
    const double* raw_max = a.b_raw_sd2_max + o;
    const double* raw_smax = a.b_raw_sdd_max + o;
    const double* raw_smin = a.b_raw_sdd_min + o;
    for (int i0 = lane; i0 < N; i0 += 32 * 4) {
      double rm[4], rx[4], rn[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + 32 * u;
        if (i < N) {
          rm[u] = raw_max[i];
          rx[u] = raw_smax[i];
          rn[u] = raw_smin[i];
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + 32 * u;
        if (i < N) {
          sv[i] = (i == N - 1) ? s_end : ds * i + s_start;
          S.sd2[i] = (i == 0) ? sd_start * sd_start : ((i == N - 1) ? 0.0 : nan);
          S.sdd[i] = nan;
          S.bmax[i] = rm[u];
          S.bsmax[i] = rx[u];
          S.bsmin[i] = rn[u];
        }
      }
    }
    __syncwarp();

    // ---- OptimizePathParameter (:287-490)
    TPB_STAMP(1);
    S.boundary_passes(fix_val, fix_idx);
    S.build_critical_tables(next_src, prev_eq0);
    TPB_STAMP(2);
    }
    if (kPhases & kPhaseOptimize) {
    int iforw_lo = 0, iback_hi = N - 1, iback_lo, iforw_hi, icrit, icrit_lo, icrit_hi;
    iback_lo = S.add_backward(iback_hi);
    TPB_STAMP(3);
    iforw_hi = S.add_forward(iforw_lo);
    TPB_STAMP(4);
    icrit_hi = iback_lo;
    if ((iforw_hi < icrit_hi) && ((icrit_hi < N - 2) && (icrit_hi >= 2))) {
      S.st(&S.sd2[icrit_hi], nan);
      icrit_hi++;
      iback_lo++;
    }
    icrit_lo = iforw_hi;
    for (int loop = 0; loop < max_loops; ++loop) {
      if (iforw_hi >= icrit_hi) break;
#ifdef TPB_PROFILE_PHASES
      ++nloops;
#endif
#ifdef TPB_PROFILE_PHASES
      const long long c0 = clock64();
#endif
      icrit = S.next_critical(icrit_lo, icrit_hi, next_src, prev_eq0);
#ifdef TPB_CHECK_SCAN
      if (icrit != S.next_critical_scan(icrit_lo, icrit_hi)) {
        status = TPB_PATH_FAIL_INDEX;
        break;
      }
#endif
#ifdef TPB_PROFILE_PHASES
      acc_nc += clock64() - c0;
#endif
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
#ifdef TPB_PROFILE_PHASES
      const long long c1 = clock64();
      const int n1 = S.nsteps;
#endif
      iback_lo = S.add_backward(iback_hi);
#ifdef TPB_PROFILE_PHASES
      const long long c2 = clock64();
      const int n2 = S.nsteps;
      acc_bw += c2 - c1;
      steps_bw += n2 - n1;
#endif
      iforw_lo = icrit;
      iforw_hi = S.add_forward(iforw_lo);
#ifdef TPB_PROFILE_PHASES
      const long long c3 = clock64();
      acc_fw += c3 - c2;
      steps_fw += S.nsteps - n2;
      acc_max += (c3 - c2 > c2 - c1) ? (c3 - c2) : (c2 - c1);
#endif
      if (iback_lo > icrit_lo) {
        status = TPB_PATH_FAIL_CONNECT;
        break;
      }
      icrit_lo = iforw_hi;
    }
    }
  }
  TPB_STAMP(5);
  if ((kPhases & kPhaseFinalize) == 0) {
    if (lane == 0) a.status[b] = status;
    return;
  }
  if (status == TPB_PATH_OK) {
    // NaN audit (:398-403)
    bool any_nan = false;
    for (int i = lane; i < N; i += 32) {
      const double v = S.sd2[i];
      any_nan |= (v != v);
    }
    if (__ballot_sync(kFullMask, any_nan)) status = TPB_PATH_FAIL_NAN;
  }
  if (status == TPB_PATH_OK) {
    // intersection fill (:404-411)
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
        S.sdd_at_intersection(base + k, S.p.load(base + k));
      }
    }
    if (S.p.valid(S.p.load(0), sdd_start, S.sd2[0])) S.st(&S.sdd[0], sdd_start);  // (:414-416)
    __syncwarp();
    if (S.sd2[N - 1] != 0) status = TPB_PATH_FAIL_TERMINAL;
  }
  if (status == TPB_PATH_OK) {
    for (int i = lane; i < N; i += 32) sd[i] = sqrt(S.sd2[i]);
    // last_extremal_index_ (:435-445): highest i in [1, N-2] with sdd > 0 or on
    // the limit curve, else 0; 32 samples per step from the top.
    last_extremal = 0;
    for (int top = (N - 2 > 1) ? N - 2 : 1; top >= 1; top -= 32) {
      const int i = top - lane;
      bool stop = false;
      if (i >= 1) stop = (S.sdd[i] > 0.0) || (fabs(S.sd2[i] - S.bmax[i]) < kTiny);
      const unsigned m = __ballot_sync(kFullMask, stop);
      if (m) {
        last_extremal = top - (__ffs(m) - 1);
        break;
      }
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
#ifdef TPB_PROFILE_PHASES
  // diagnostics build only: per-path phase time stamps (SM cycles) in fix_val
  TPB_STAMP(6);
  if (lane == 0 && N >= 16) {
    for (int k = 0; k < 7; ++k) fix_val[k] = static_cast<double>(ph[k] - ph[0]);
    fix_val[7] = nloops;
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    fix_val[8] = smid;
    fix_val[9] = static_cast<double>(ph[0]);
    fix_val[10] = static_cast<double>(acc_nc);
    fix_val[11] = static_cast<double>(acc_bw);
    fix_val[12] = static_cast<double>(acc_fw);
    fix_val[13] = steps_bw;
    fix_val[14] = steps_fw;
    fix_val[15] = static_cast<double>(acc_max);
    fix_val[16] = S.nsteps;
  }
#endif
#undef TPB_STAMP
}

}  // namespace tpb
