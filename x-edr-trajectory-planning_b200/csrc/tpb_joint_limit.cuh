// Per-sample limit curve of a joint-space path with kD <= 8 dofs, one sample per
// thread, everything in registers (no local-memory arrays).
//
// Same results, bit for bit, as the generic lp_max_sd2 / find_sdd_min_max of
// tpb_solver.cuh (FindMaxSd2Simplex time_optimal_path_timing.cc:1149-1363,
// IsOptimal :1105-1147, FindSddMax/FindSddMin/AreDerivativesValid :624-695)
// applied to the rows TimeableJointSplinePath::ConstraintSetup builds
// (timeable_path_joint_spline.cc:320-343):
//
//   row d      (acceleration)  A = q'_d   B = q''_d    -aS_d <= . <= aS_d
//   row kD + d (velocity)      A = 0      B = q'_d^2    0    <= . <= vS2_d
//
// What the specialisation uses:
//  * loops over rows are compile-time unrolled, so A[], B[] are registers; the
//    reference's constraint_set_ / active-set vectors are 32-bit slot masks
//    (slot = 2*row + (type == lower)) walked in ascending order = list order;
//  * velocity rows have A == 0: along an edge sdd = a + b*sd2 their B
//    coefficient stays q'^2 whenever b is finite, so 1/q'^2 is computed once and
//    reused by every simplex iteration (the identity is checked at run time, the
//    division is redone when it does not hold); their lower slots (bound 0) can
//    never become active because the walk only accepts sd2 > current sd2 > 0;
//  * the common case of exactly one new active row per iteration keeps that row
//    as scalars; ties (two or more rows within kTiny) take a slow path that
//    fetches rows with select chains.
//
// kX > 0 appends kX more rows of the velocity kind with symmetric bounds,
//   row 2*kD + k               A = 0      B = Bx_k     -hX_k <= . <= hX_k
// which is what TimeableCartesianSplinePath::ConstraintSetup adds for the
// translational and rotational speed (timeable_path_cartesian_spline.cc:551-595):
// the caller fills Bv[kD + k] and points hX at the kX bounds.  Their lower slots
// (bound -hX < 0 over B >= 0) can never become active either.
#pragma once

// Code size: k_joint_sample_limit<7> is 60+ KB of straight-line code against a 32 KB
// instruction cache level, and a third of its stall samples are instruction fetch
// (profiles/r2_summary.md).  Measured variants (cfg2, ms per launch):
//   TPB_SL_ROLL_W 0  everything unrolled                       67.7 KB, 122 regs  1.402
//   TPB_SL_ROLL_W 1  the two bounds of a row share one copy of
//                    the candidate test in FindSddMax / Min    62.0 KB, 124 regs  1.296  <- default
//   TPB_SL_ROLL_W 2  + the same in the LP's vertex scan        60.5 KB, 128 regs  1.390
//   TPB_SL_ROLL_W 3  + entry() not inlined                     59.6 KB, 128 regs  1.397
//   TPB_SL_ROLL_ROWS 1  per-row loops not unrolled, rows in
//                    local memory                              33.2 KB,  90 regs  1.616
#ifndef TPB_SL_ROLL_W
#define TPB_SL_ROLL_W 1
#endif
#ifndef TPB_SL_ROLL_ROWS
#define TPB_SL_ROLL_ROWS 0
#endif
#if TPB_SL_ROLL_ROWS
#define TPB_ROW_LOOP _Pragma("unroll 1")
#else
#define TPB_ROW_LOOP _Pragma("unroll")
#endif

#include "tpb_device.cuh"

namespace tpb {

struct LpEntry {
  double A, B, bound;
  int type;  // kUpper / kLower
  int slot;
};

// IsOptimal (:1105-1147) on two explicit rows.
__device__ __forceinline__ bool lp_is_optimal_e(const LpEntry& e1, const LpEntry& e2) {
  const double denom = e2.A * e1.B - e1.A * e2.B;
  if (fabs(denom) < kTiny) return false;
  const double x = denom * e1.A;
  const double y = denom * (-e2.A);
  if (e1.type == kUpper) {
    if (e2.type == kUpper) return x <= 0 && y <= 0;
    return x >= 0 && y <= 0;
  }
  if (e2.type == kUpper) return x <= 0 && y >= 0;
  return x >= 0 && y >= 0;
}

template <int kD, int kX = 0>
struct JointLimit {
  static constexpr int kV = kD + kX;  // rows with A == 0
  double A[kD], B[kD];  // q', q'' of this sample
  double Bv[kV];        // q'^2 (B of the velocity rows), then the kX extra rows' B
  double invBv[kV];     // 1 / Bv
  const double* aS;     // [kD] shared memory: amax * safety
  const double* vS2;    // [kD] shared memory: (vmax * safety)^2
  const double* hX;     // [kX] bounds of the extra rows (unused when kX == 0)

  // bounds of velocity-kind row j (compile-time j in the unrolled loops)
  __device__ __forceinline__ double vhi(int j) const { return (j < kD) ? vS2[j] : hX[j - kD]; }
  __device__ __forceinline__ double vlo(int j) const { return (j < kD) ? 0.0 : -hX[j - kD]; }

  template <int kN>
  __device__ __forceinline__ static double sel(const double (&x)[kN], int i) {
#if TPB_SL_ROLL_ROWS
    return x[i];
#endif
    double r = x[0];
#pragma unroll
    for (int k = 1; k < kN; ++k) r = (i == k) ? x[k] : r;
    return r;
  }

  // row data by slot (slow paths only)
#if TPB_SL_ROLL_W > 2
  __device__ __noinline__ LpEntry entry(int slot) const {
#else
  __device__ LpEntry entry(int slot) const {
#endif
    const int row = slot >> 1;
    const bool lower = slot & 1;
    LpEntry e;
    e.slot = slot;
    e.type = lower ? kLower : kUpper;
    if (row < kD) {
      e.A = sel(A, row);
      e.B = sel(B, row);
      const double s = aS[row];
      e.bound = lower ? -s : s;
    } else {
      const int j = row - kD;
      e.A = 0.0;
      e.B = sel(Bv, j);
      double hi = vS2[(j < kD) ? j : 0], lo = 0.0;
#pragma unroll
      for (int k = 0; k < kX; ++k) {
        hi = (j == kD + k) ? hX[k] : hi;
        lo = (j == kD + k) ? -hX[k] : lo;
      }
      e.bound = lower ? lo : hi;
    }
    return e;
  }

  // (with kX > 0 the caller sets Bv[kD ..] before this)
  __device__ __forceinline__ void prepare() {
    TPB_ROW_LOOP
    for (int d = 0; d < kD; ++d) Bv[d] = A[d] * A[d];
    TPB_ROW_LOOP
    for (int d = 0; d < kV; ++d) {
      // only used where Bv >= kTiny (rows with |Bc| < kTiny are skipped)
      invBv[d] = 1.0 / ((Bv[d] < kTiny) ? 1.0 : Bv[d]);
    }
  }

  // FindMaxSd2Simplex (:1149-1363)
  __device__ void lp(double& sd2max, double& sddmax, double& sd2zero) const {
    constexpr int C = 2 * kD + kX;
    static_assert(C <= 16, "slot masks are 32 bits");
    unsigned act = 0u;
    double sd2 = DBL_MAX;
    double sdd = 0.0;
    LpEntry last;
    last.A = last.B = last.bound = 0.0;
    last.type = kUpper;
    last.slot = 0;

    // Step 2: walk along sdd = 0.
    TPB_ROW_LOOP
    for (int d = 0; d < kD; ++d) {
      const double Bi = B[d];
      const bool up = Bi > kTiny, dn = Bi < -kTiny;
      const double invB = 1.0 / ((up || dn) ? Bi : 1.0);  // unused rows: keep the fast path
      const double bound = up ? aS[d] : -aS[d];
      const double tmp = bound * invB;
      if ((up || dn) && tmp < (sd2 + kTiny) && tmp > 0) {
        if (tmp < sd2 - kTiny) act = 0u;
        const int slot = 2 * d + (up ? 0 : 1);
        act |= 1u << slot;
        sd2 = tmp;
        last.A = A[d];
        last.B = Bi;
        last.bound = bound;
        last.type = up ? kUpper : kLower;
        last.slot = slot;
      }
    }
    TPB_ROW_LOOP
    for (int d = 0; d < kV; ++d) {
      const double Bi = Bv[d];
      const double tmp = vhi(d) * invBv[d];
      if (Bi > kTiny && tmp < (sd2 + kTiny) && tmp > 0) {
        if (tmp < sd2 - kTiny) act = 0u;
        const int slot = 2 * (kD + d);
        act |= 1u << slot;
        sd2 = tmp;
        last.A = 0.0;
        last.B = Bi;
        last.bound = vhi(d);
        last.type = kUpper;
        last.slot = slot;
      }
    }
    if (sd2 > kMaxSd2 || act == 0u) {
      sd2zero = kMaxSd2;
      sd2max = kMaxSd2;
      sddmax = 0.0;
      return;
    }
    sd2zero = sd2;

    // Step 3 on the initial active set.
    LpEntry s = last;
    if (__popc(act) > 1) {
      unsigned m = act;
      double s_slope = DBL_MAX;
      bool first = true;
      while (m) {
        const int f = __ffs(m) - 1;
        m &= m - 1;
        const LpEntry ef = entry(f);
        const double slope = fabs(ef.A * (1.0 / ef.B));
        if (first || slope < s_slope) {
          s = ef;
          s_slope = slope;
          first = false;
        }
        unsigned m2 = m;
        while (m2) {
          const int g = __ffs(m2) - 1;
          m2 &= m2 - 1;
          if (lp_is_optimal_e(ef, entry(g))) {
            sd2max = sd2;
            sddmax = 0.0;
            return;
          }
        }
      }
    }
    unsigned in_set = ((C >= 16) ? ~0u : ((1u << (2 * C)) - 1u)) & ~act;

    for (int loop = 0; loop < C; ++loop) {
      const double As = s.A;
      if (fabs(As) < kTiny) {
        sd2max = sd2;
        sddmax = sdd;
        return;
      }
      const double invA = 1.0 / As;
      const double b = -s.B * invA;
      const double a = s.bound * invA;

      // Step 4: next vertex along sdd = a + b * sd2.
      act = 0u;
      double next_sd2 = DBL_MAX, next_sdd = 0.0;
      TPB_ROW_LOOP
      for (int d = 0; d < kD; ++d) {
        const double Ac = A[d];
        const double Bc = Ac * b + B[d];
        const bool usable = !(fabs(Bc) < kTiny);
        const double invB = 1.0 / (usable ? Bc : 1.0);
        const double Aca = Ac * a;
#if TPB_SL_ROLL_W > 1 || TPB_SL_ROLL_ROWS
#pragma unroll 1
#else
#pragma unroll
#endif
        for (int w = 0; w < 2; ++w) {  // upper slot first, then lower (slot order)
          const int slot = 2 * d + w;
          const double bound = w ? -aS[d] : aS[d];
          const double tmp = (bound - Aca) * invB;
          if (usable && ((in_set >> slot) & 1u) && tmp < (next_sd2 + kTiny) && tmp > sd2) {
            if (tmp < next_sd2 - kTiny) act = 0u;
            act |= 1u << slot;
            next_sd2 = tmp;
            last.A = Ac;
            last.B = B[d];
            last.bound = bound;
            last.type = w ? kLower : kUpper;
            last.slot = slot;
          }
        }
      }
      const double zb = 0.0 * b, za = 0.0 * a;
      TPB_ROW_LOOP
      for (int d = 0; d < kV; ++d) {
        const double Bc = zb + Bv[d];
        const bool usable = !(fabs(Bc) < kTiny);
        const int slot = 2 * (kD + d);
        double invB = invBv[d];
        if (!(Bc == Bv[d])) invB = 1.0 / Bc;  // non-finite b only
        const double tmp = (vhi(d) - za) * invB;
        if (usable && ((in_set >> slot) & 1u) && tmp < (next_sd2 + kTiny) && tmp > sd2) {
          if (tmp < next_sd2 - kTiny) act = 0u;
          act |= 1u << slot;
          next_sd2 = tmp;
          last.A = 0.0;
          last.B = Bv[d];
          last.bound = vhi(d);
          last.type = kUpper;
          last.slot = slot;
        }
        // lower slot: tmp = (lo - za) * invB with lo <= 0 is <= 0 or NaN, never > sd2 > 0
      }
      if (act == 0u) {  // unbounded along this edge
        sd2max = sd2zero;
        sddmax = 0.0;
        return;
      }
      // (the reference assigns this at every acceptance; the last one stands)
      next_sdd = a + b * next_sd2;
      // active list = new rows (slot order) followed by s: pairwise optimality,
      // next edge = lowest slope among the new rows (first one on ties)
      bool optimal = false;
      LpEntry sn = last;
      if (__popc(act) == 1) {
        optimal = lp_is_optimal_e(last, s);
      } else {
        unsigned m = act;
        double s_slope = DBL_MAX;
        while (m) {
          const int f = __ffs(m) - 1;
          m &= m - 1;
          const LpEntry ef = entry(f);
          const double Bc = ef.A * b + ef.B;
          const double slope = fabs(ef.A * (1.0 / Bc));
          if (slope < s_slope) {
            sn = ef;
            s_slope = slope;
          }
          unsigned m2 = m;
          while (m2) {
            const int g = __ffs(m2) - 1;
            m2 &= m2 - 1;
            optimal = optimal || lp_is_optimal_e(ef, entry(g));
          }
          optimal = optimal || lp_is_optimal_e(ef, s);
        }
      }
      if (optimal) {
        sd2max = next_sd2;
        sddmax = next_sdd;
        if (next_sd2 > kMaxSd2) {
          sd2max = kMaxSd2;
          sddmax = 0.0;
        }
        return;
      }
      in_set &= ~act;
      s = sn;
      sd2 = next_sd2;
      sdd = next_sdd;
    }
    sd2max = sd2zero;
    sddmax = 0.0;
  }

  // FindSddMax and FindSddMin at the same sd2 (:638-695) with the validity test
  // (:624-636).  Only acceleration rows yield candidates (velocity rows have
  // A == 0); a velocity row evaluates 0*cand + q'^2*sd2, which does not depend
  // on a finite candidate and is NaN (passes every comparison) otherwise.
  __device__ void sdd_min_max(double sd2, double& sdd_max, double& sdd_min) const {
    double Bs[kD];
    bool vbad = false;
#pragma unroll
    for (int d = 0; d < kD; ++d) {
      Bs[d] = B[d] * sd2;
    }
#pragma unroll
    for (int d = 0; d < kV; ++d) {
      const double w = Bv[d] * sd2;
      vbad = vbad || (w + kTiny < vlo(d)) || (w - kTiny > vhi(d));
    }
    double mx = -DBL_MAX, mn = DBL_MAX;
    TPB_ROW_LOOP
    for (int i = 0; i < kD; ++i) {
      const double Ai = A[i];
      const bool has = !is_tiny(Ai);
#if TPB_SL_ROLL_W || TPB_SL_ROLL_ROWS
#pragma unroll 1
#else
#pragma unroll
#endif
      for (int w = 0; w < 2; ++w) {
        const double cand = ((w == 0 ? -aS[i] : aS[i]) - Bs[i]) / (has ? Ai : 1.0);
        // velocity rows: 0*cand + B*sd2 is NaN (passes) for a non-finite candidate
        bool ok = has && !((fabs(cand) <= DBL_MAX) && vbad);
#pragma unroll
        for (int r = 0; r < kD; ++r) {
          const double v = A[r] * cand + Bs[r];
          ok = ok && !outside_sym(v, aS[r]);
        }
        mx = (ok && cand > mx) ? cand : mx;
        mn = (ok && cand < mn) ? cand : mn;
      }
    }
    sdd_max = (mx == -DBL_MAX) ? 0.0 : mx;
    sdd_min = (mn == DBL_MAX) ? 0.0 : mn;
  }
};

}  // namespace tpb
