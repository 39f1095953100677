// Per-sample limit-curve pieces of TimeOptimalPathProfile, one sample per
// thread: FindMaxSd2Simplex (time_optimal_path_timing.cc:1149-1363), IsOptimal
// (:1105-1147), FindSddMax / FindSddMin / AreDerivativesValid (:624-695).
//
// Rows live in thread-local arrays A[], B[], LO[], HI[] (C <= kMaxRows).
#pragma once

#include "tpb_device.cuh"

namespace tpb {

// IsOptimal (:1105-1147): KKT sign test for the vertex of rows i1/i2.
__device__ __forceinline__ bool lp_is_optimal(const double* A, const double* B, int i1,
                                              int i2, int t1, int t2) {
  const double denom = A[i2] * B[i1] - A[i1] * B[i2];
  if (fabs(denom) < kTiny) return false;
  const double x = denom * A[i1];
  const double y = denom * (-A[i2]);
  if (t1 == kUpper) {
    if (t2 == kUpper) return x <= 0 && y <= 0;
    return x >= 0 && y <= 0;
  }
  if (t2 == kUpper) return x <= 0 && y >= 0;
  return x >= 0 && y >= 0;
}

// FindMaxSd2Simplex (:1149-1363).  The reference's constraint_set_ vector of
// (row, type) pairs -- ordered (0,U),(0,L),(1,U),... with erase() keeping the
// order -- is a 64-bit membership mask walked in ascending slot order
// (slot = 2*row + (type == lower)).
__device__ inline void lp_max_sd2(const double* A, const double* B, const double* LO,
                                  const double* HI, int C, double& sd2max,
                                  double& sddmax, double& sd2zero) {
  static_assert(kMaxRows <= 64, "in_set holds 128 slots");
  int act_row[2 * kMaxRows + 2];
  int act_type[2 * kMaxRows + 2];
  double act_slope[2 * kMaxRows + 2];
  int nact = 0;
  sd2max = 0;
  sddmax = 0;
  // slots 0..63 in in_set[0], 64..127 in in_set[1] (C <= kMaxRows = 64)
  unsigned long long in_set[2];
  in_set[0] = (C >= 32) ? ~0ull : ((1ull << (2 * C)) - 1ull);
  in_set[1] = (C <= 32) ? 0ull : ((C >= 64) ? ~0ull : ((1ull << (2 * C - 64)) - 1ull));
  auto drop = [&](int row, int type) {
    const int slot = 2 * row + (type == kLower ? 1 : 0);
    in_set[slot >> 6] &= ~(1ull << (slot & 63));
  };

  // Step 2: walk along sdd = 0.
  double sd2 = DBL_MAX;
  double sdd = 0.0;
  for (int i = 0; i < C; ++i) {
    const double Bi = B[i];
    if (fabs(Bi) < kTiny) continue;
    if (Bi > kTiny) {
      const double invB = 1.0 / Bi;
      const double tmp = HI[i] * invB;
      if (tmp < (sd2 + kTiny) && tmp > 0) {
        if (tmp < sd2 - kTiny) nact = 0;
        act_row[nact] = i;
        act_type[nact] = kUpper;
        act_slope[nact] = fabs(A[i] * invB);
        ++nact;
        sd2 = tmp;
      }
    } else if (Bi < -kTiny) {
      const double invB = 1.0 / Bi;
      const double tmp = LO[i] * invB;
      if (tmp < (sd2 + kTiny) && tmp > 0) {
        if (tmp < sd2 - kTiny) nact = 0;
        act_row[nact] = i;
        act_type[nact] = kLower;
        act_slope[nact] = fabs(A[i] * invB);
        ++nact;
        sd2 = tmp;
      }
    }
  }
  if (sd2 > kMaxSd2 || nact == 0) {
    sd2zero = kMaxSd2;
    sd2max = kMaxSd2;
    sddmax = 0.0;
    return;
  }
  sd2zero = sd2;

  // Step 3 on the initial active set.
  int s_row = act_row[0], s_type = act_type[0];
  double s_slope = act_slope[0];
  if (nact >= 2) {
    for (int f = 0; f < nact; ++f) {
      if (act_slope[f] < s_slope) {
        s_row = act_row[f];
        s_type = act_type[f];
        s_slope = act_slope[f];
      }
      for (int s = f + 1; s < nact; ++s) {
        if (lp_is_optimal(A, B, act_row[f], act_row[s], act_type[f], act_type[s])) {
          sd2max = sd2;
          sddmax = 0.0;
          return;
        }
      }
    }
  }
  for (int l = 0; l < nact; ++l) {
    drop(act_row[l], act_type[l]);
  }

  for (int loop = 0; loop < C; ++loop) {
    const double As = A[s_row];
    if (fabs(As) < kTiny) {
      sd2max = sd2;
      sddmax = sdd;
      return;
    }
    const double invA = 1.0 / As;
    const double b = -B[s_row] * invA;
    const double a = (s_type == kUpper) ? HI[s_row] * invA : LO[s_row] * invA;

    // Step 4: next vertex along sdd = a + b * sd2.
    nact = 0;
    double next_sd2 = DBL_MAX, next_sdd = 0.0;
    unsigned long long m = in_set[0];
    int word = 0;
    for (;;) {
      if (m == 0ull) {
        if (word == 1 || C <= 32) break;
        word = 1;
        m = in_set[1];
        continue;
      }
      const int e = 64 * word + __ffsll(static_cast<long long>(m)) - 1;
      m &= m - 1;
      const int ci = e >> 1;
      const bool is_lower = e & 1;
      const double Ac = A[ci];
      const double Bc = Ac * b + B[ci];
      if (fabs(Bc) < kTiny) continue;
      const double invB = 1.0 / Bc;
      const double bound = is_lower ? LO[ci] : HI[ci];
      const double tmp = (bound - Ac * a) * invB;
      if (tmp < (next_sd2 + kTiny) && tmp > sd2) {
        if (tmp < next_sd2 - kTiny) nact = 0;
        act_row[nact] = ci;
        act_type[nact] = is_lower ? kLower : kUpper;
        act_slope[nact] = fabs(Ac * invB);
        ++nact;
        next_sd2 = tmp;
        next_sdd = a + b * next_sd2;
      }
    }
    if (nact == 0) {  // unbounded along this edge
      sd2max = sd2zero;
      sddmax = 0.0;
      return;
    }
    act_row[nact] = s_row;
    act_type[nact] = s_type;
    act_slope[nact] = s_slope;
    ++nact;
    s_row = -1;
    s_type = 0;
    s_slope = DBL_MAX;
    for (int f = 0; f < nact; ++f) {
      if (f != nact - 1 && act_slope[f] < s_slope) {
        s_row = act_row[f];
        s_type = act_type[f];
        s_slope = act_slope[f];
      }
      for (int s = f + 1; s < nact; ++s) {
        if (lp_is_optimal(A, B, act_row[f], act_row[s], act_type[f], act_type[s])) {
          sd2max = next_sd2;
          sddmax = next_sdd;
          if (next_sd2 > kMaxSd2) {
            sd2max = kMaxSd2;
            sddmax = 0.0;
          }
          return;
        }
      }
    }
    --nact;
    for (int l = 0; l < nact; ++l) {
      drop(act_row[l], act_type[l]);
    }
    sd2 = next_sd2;
    sdd = next_sdd;
  }
  sd2max = sd2zero;
  sddmax = 0.0;
}

// FindSddMax and FindSddMin (:638-695) at the same sd2 share candidates and the
// validity test (:624-636): the results are the max / min over the candidates
// (bound - B*sd2)/A, |A| >= kTiny, that satisfy every row within kTiny; 0 if
// none does.  (The reference's running "sddi > sdd &&" test only skips
// candidates that cannot change the extremum.)
__device__ inline void find_sdd_min_max(const double* A, const double* B, const double* LO,
                                        const double* HI, int C, double sd2,
                                        double& sdd_max, double& sdd_min) {
  double mx = -DBL_MAX, mn = DBL_MAX;
  for (int i = 0; i < C; ++i) {
    const double Ai = A[i];
    if (is_tiny(Ai)) continue;
    const double Bs = B[i] * sd2;
#pragma unroll
    for (int w = 0; w < 2; ++w) {
      const double cand = ((w == 0 ? LO[i] : HI[i]) - Bs) / Ai;
      if (!(cand > mx) && !(cand < mn)) continue;
      bool ok = true;
      for (int r = 0; r < C; ++r) {
        const double v = A[r] * cand + B[r] * sd2;
        if (v + kTiny < LO[r] || v - kTiny > HI[r]) {
          ok = false;
          break;
        }
      }
      if (ok) {
        if (cand > mx) mx = cand;
        if (cand < mn) mn = cand;
      }
    }
  }
  sdd_max = (mx == -DBL_MAX) ? 0.0 : mx;
  sdd_min = (mn == DBL_MAX) ? 0.0 : mn;
}

}  // namespace tpb
