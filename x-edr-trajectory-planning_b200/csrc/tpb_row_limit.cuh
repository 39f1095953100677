// Per-sample limit curve for kC <= 16 arbitrary constraint rows, one sample per
// thread, rows and the whole active-set LP in registers (no local-memory arrays).
//
// Same results, bit for bit, as lp_max_sd2 / find_sdd_min_max of tpb_solver.cuh
// (FindMaxSd2Simplex time_optimal_path_timing.cc:1149-1363, IsOptimal :1105-1147,
// FindSddMax/FindSddMin/AreDerivativesValid :624-695): the structure is that of
// JointLimit (tpb_joint_limit.cuh) without its knowledge about the rows -- loops
// over rows are compile-time unrolled, the reference's constraint_set_ and
// active-set vectors are 32-bit slot masks (slot = 2*row + (type == lower)) walked
// in ascending order = list order, the common case of one new active row per
// iteration keeps that row as scalars and ties fetch rows with select chains.
//
// A problem with C < kC rows is padded with rows A = 0, B = 0, -1 <= . <= 1: such a
// row is skipped by every step (|B| < kTiny, |A*b + B| < kTiny or NaN), never
// yields a candidate and passes every validity test.  The simplex loop is bounded
// by the real C like the reference's.
#pragma once

#include "tpb_device.cuh"
#include "tpb_joint_limit.cuh"  // LpEntry, lp_is_optimal_e

namespace tpb {

template <int kC>
struct RowLimit {
  static_assert(kC <= 16, "slot masks are 32 bits");
  double A[kC], B[kC], LO[kC], HI[kC];

  __device__ __forceinline__ static double sel(const double (&x)[kC], int i) {
    double r = x[0];
#pragma unroll
    for (int k = 1; k < kC; ++k) r = (i == k) ? x[k] : r;
    return r;
  }

  // row data by slot (slow paths only)
  __device__ LpEntry entry(int slot) const {
    const int row = slot >> 1;
    const bool lower = slot & 1;
    LpEntry e;
    e.slot = slot;
    e.type = lower ? kLower : kUpper;
    e.A = sel(A, row);
    e.B = sel(B, row);
    e.bound = lower ? sel(LO, row) : sel(HI, row);
    return e;
  }

  // FindMaxSd2Simplex (:1149-1363); C = number of real rows
  __device__ void lp(int C, double& sd2max, double& sddmax, double& sd2zero) const {
    unsigned act = 0u;
    double sd2 = DBL_MAX;
    double sdd = 0.0;
    LpEntry last;
    last.A = last.B = last.bound = 0.0;
    last.type = kUpper;
    last.slot = 0;
    sd2max = 0;
    sddmax = 0;

    // Step 2: walk along sdd = 0.
#pragma unroll
    for (int i = 0; i < kC; ++i) {
      const double Bi = B[i];
      const bool up = Bi > kTiny, dn = Bi < -kTiny;
      const double invB = 1.0 / ((up || dn) ? Bi : 1.0);  // unused rows: keep the fast path
      const double bound = up ? HI[i] : LO[i];
      const double tmp = bound * invB;
      if ((up || dn) && tmp < (sd2 + kTiny) && tmp > 0) {
        if (tmp < sd2 - kTiny) act = 0u;
        const int slot = 2 * i + (up ? 0 : 1);
        act |= 1u << slot;
        sd2 = tmp;
        last.A = A[i];
        last.B = Bi;
        last.bound = bound;
        last.type = up ? kUpper : kLower;
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
    unsigned in_set = ((kC >= 16) ? ~0u : ((1u << (2 * kC)) - 1u)) & ~act;

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
#pragma unroll
      for (int i = 0; i < kC; ++i) {
        const double Ac = A[i];
        const double Bc = Ac * b + B[i];
        const bool usable = !(fabs(Bc) < kTiny);
        const double invB = 1.0 / (usable ? Bc : 1.0);
        const double Aca = Ac * a;
#pragma unroll
        for (int w = 0; w < 2; ++w) {  // upper slot first, then lower (slot order)
          const int slot = 2 * i + w;
          const double bound = w ? LO[i] : HI[i];
          const double tmp = (bound - Aca) * invB;
          if (usable && ((in_set >> slot) & 1u) && tmp < (next_sd2 + kTiny) && tmp > sd2) {
            if (tmp < next_sd2 - kTiny) act = 0u;
            act |= 1u << slot;
            next_sd2 = tmp;
            last.A = Ac;
            last.B = B[i];
            last.bound = bound;
            last.type = w ? kLower : kUpper;
            last.slot = slot;
          }
        }
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
  // (:624-636): extrema over the candidates (bound - B*sd2)/A, |A| >= kTiny, that
  // satisfy every row within kTiny; 0 if none does.
  __device__ void sdd_min_max(double sd2, double& sdd_max, double& sdd_min) const {
    double Bs[kC];
#pragma unroll
    for (int r = 0; r < kC; ++r) Bs[r] = B[r] * sd2;
    double mx = -DBL_MAX, mn = DBL_MAX;
#pragma unroll
    for (int i = 0; i < kC; ++i) {
      const double Ai = A[i];
      const bool has = !is_tiny(Ai);
#pragma unroll
      for (int w = 0; w < 2; ++w) {
        const double cand = ((w == 0 ? LO[i] : HI[i]) - Bs[i]) / (has ? Ai : 1.0);
        bool ok = has;
#pragma unroll
        for (int r = 0; r < kC; ++r) {
          const double v = A[r] * cand + Bs[r];
          ok = ok && !((v + kTiny < LO[r]) || (v - kTiny > HI[r]));
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
