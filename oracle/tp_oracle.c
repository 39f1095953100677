/* TEST INFRASTRUCTURE ONLY -- the CPU oracle (see tp_oracle.h).
 *
 * Plain-C restatement of the reference's hot path.  Citations are into
 * /root/reference/trajectory_planning/ (reference commit a5d5aced).
 *
 * PARITY PIN
 *  - stage A (B-spline eval): pinned against the Mathematica table of
 *    splines/bspline_test.cc:97-727 (tests/golden/bspline_mathematica.json,
 *    tolerance 5e-14 as in bspline_test.cc:70-75) and scipy.
 *  - stage C (time-optimal timing): pinned bit-for-bit against the
 *    reference's own time_optimal_path_timing.cc compiled verbatim into
 *    oracle/_ref/libtp_ref.so (tests/test_oracle_vs_ref.py), on the
 *    reference test scenarios (time_optimal_path_timing_test.cc:440-701),
 *    the 5 regression LPs (:744-1088) and random problems.
 *  - stages B and D (control polygon, knots, constraint rows, assembly,
 *    resampling): restated from spline_utils.cc / timeable_path_joint_spline.cc
 *    / path_timing_trajectory.cc; pinned only through the properties the
 *    reference's tests state (spline_utils_test.cc:139-172,
 *    path_timing_trajectory_test.cc:112-173).  Two ulp-level choices are
 *    UNPINNED because they live in un-vendored dependencies: the summation
 *    order of Eigen's norm()/squaredNorm() (sequential here) and
 *    eigenmath::InterpolateLinear (a + t*(b-a) here).
 *
 * Build without FMA contraction (gcc -O2 -ffp-contract=off) so that results
 * are bit-stable and comparable with the GPU build (-fmad=false).
 */
#include "tp_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define ORC_MAX_DEGREE 8
/* time_optimal_path_timing.h:275,279 */
static const double kTiny = DBL_EPSILON * 1e5;
static const double kMaxSd2 = 1e6;

/* ======================================================================
 * Stage A
 * ====================================================================== */

/* splines/bspline_base.cc:218-246 (NURBS A2.1; upper_bound on [p, nk-p)) */
int orc_knot_span(const double* knots, int num_knots, int degree, double u) {
  if (num_knots == 0) return 0;
  const int num_points = num_knots - degree - 1;
  if (u == knots[num_knots - 1]) return num_points - 1;
  int lo = degree, hi = num_knots - degree;
  while (lo < hi) {
    const int mid = lo + (hi - lo) / 2;
    if (knots[mid] <= u) {
      lo = mid + 1;
    } else {
      hi = mid;
    }
  }
  return lo - 1;
}

/* splines/bspline_base.cc:249-265 (NURBS A2.2) */
void orc_basis(const double* knots, int span, int p, double u, double* basis) {
  double left[ORC_MAX_DEGREE + 1], right[ORC_MAX_DEGREE + 1];
  basis[0] = 1.0;
  for (int j = 1; j <= p; ++j) {
    left[j] = u - knots[span + 1 - j];
    right[j] = knots[span + j] - u;
  }
  for (int j = 1; j <= p; ++j) {
    double saved = 0.0;
    for (int r = 0; r < j; ++r) {
      const double tmp = basis[r] / (right[r + 1] + left[j - r]);
      basis[r] = saved + right[r + 1] * tmp;
      saved = left[j - r] * tmp;
    }
    basis[j] = saved;
  }
}

/* splines/bspline_base.cc:268-348 (NURBS A2.3). ders is [(der+1)][(p+1)]. */
void orc_basis_ders(const double* knots, int span, int p, int der, double u,
                    double* ders) {
  double ndu[ORC_MAX_DEGREE + 1][ORC_MAX_DEGREE + 1];
  double a[2][ORC_MAX_DEGREE + 1];
  double left[ORC_MAX_DEGREE + 1], right[ORC_MAX_DEGREE + 1];
  const int w = p + 1;
  memset(a, 0, sizeof(a));
  ndu[0][0] = 1.0;
  for (int j = 1; j <= p; ++j) {
    left[j] = u - knots[span + 1 - j];
    right[j] = knots[span + j] - u;
    double saved = 0.0;
    for (int r = 0; r < j; ++r) {
      ndu[j][r] = right[r + 1] + left[j - r];
      const double tmp = ndu[r][j - 1] / ndu[j][r];
      ndu[r][j] = saved + right[r + 1] * tmp;
      saved = left[j - r] * tmp;
    }
    ndu[j][j] = saved;
  }
  for (int j = 0; j <= p; ++j) ders[0 * w + j] = ndu[j][p];

  for (int r = 0; r <= p; ++r) {
    int s1 = 0, s2 = 1;
    a[0][0] = 1.0;
    for (int k = 1; k <= der; ++k) {
      double d = 0.0;
      const int rk = r - k;
      const int pk = p - k;
      int j1, j2;
      if (r >= k) {
        a[s2][0] = a[s1][0] / ndu[pk + 1][rk];
        d = a[s2][0] * ndu[rk][pk];
      }
      j1 = (rk >= -1) ? 1 : -rk;
      j2 = (r - 1 <= pk) ? k - 1 : p - r;
      for (int j = j1; j <= j2; ++j) {
        a[s2][j] = (a[s1][j] - a[s1][j - 1]) / ndu[pk + 1][rk + j];
        d += a[s2][j] * ndu[rk + j][pk];
      }
      if (r <= pk) {
        a[s2][k] = -a[s1][k - 1] / ndu[pk + 1][r];
        d += a[s2][k] * ndu[r][pk];
      }
      ders[k * w + r] = d;
      const int t = s1;
      s1 = s2;
      s2 = t;
    }
  }
  /* factors p!/(p-k)!  (:341-345) */
  double f = (double)p;
  for (int k = 1; k <= der; ++k) {
    for (int j = 0; j <= p; ++j) ders[k * w + j] *= f;
    f *= (double)(p - k);
  }
}

/* splines/bspline.h:514-536 (NURBS A3.1) */
int orc_eval_curve(const double* knots, int num_knots, const double* ctrl,
                   int dim, int degree, double u, double* value) {
  double basis[ORC_MAX_DEGREE + 1];
  if (u < knots[0] || u > knots[num_knots - 1]) return 11;
  const int span = orc_knot_span(knots, num_knots, degree, u);
  orc_basis(knots, span, degree, u, basis);
  for (int d = 0; d < dim; ++d) value[d] = 0.0;
  for (int i = 0; i <= degree; ++i) {
    const double* P = ctrl + (size_t)(span - degree + i) * dim;
    for (int d = 0; d < dim; ++d) value[d] += basis[i] * P[d];
  }
  return 0;
}

/* splines/bspline.h:539-568 (NURBS A3.2); requires nder <= degree like the
 * reference (CHECK(der <= p), bspline_base.cc:270). */
int orc_eval_curve_ders(const double* knots, int num_knots, const double* ctrl,
                        int dim, int degree, int nder, double u,
                        double* values) {
  double ders[(ORC_MAX_DEGREE + 1) * (ORC_MAX_DEGREE + 1)];
  if (u < knots[0] || u > knots[num_knots - 1]) return 11;
  if (nder > degree) return 3;
  const int span = orc_knot_span(knots, num_knots, degree, u);
  orc_basis_ders(knots, span, degree, nder, u, ders);
  const int w = degree + 1;
  for (int k = 0; k <= nder; ++k) {
    double* v = values + (size_t)k * dim;
    for (int d = 0; d < dim; ++d) v[d] = 0.0;
    for (int j = 0; j <= degree; ++j) {
      const double* P = ctrl + (size_t)(span - degree + j) * dim;
      const double b = ders[k * w + j];
      for (int d = 0; d < dim; ++d) v[d] += b * P[d];
    }
  }
  return 0;
}

/* splines/bspline_base.cc:350-381 (clamped uniform knots by running sum) */
int orc_make_uniform_knots(int num_points, int degree, double lo, double hi,
                           double* knots) {
  if (hi <= lo) return 3;
  if (num_points < degree + 1) return 11;
  const int nk = num_points + degree + 1;
  for (int i = 0; i <= degree; ++i) knots[i] = lo;
  const double denom = nk - 2.0 * (degree + 1.0) + 1.0;
  const double spacing = (1.0 / denom) * (hi - lo);
  for (int i = degree + 1; i < nk - degree - 1; ++i) {
    knots[i] = knots[i - 1] + spacing;
  }
  for (int i = nk - degree - 1; i < nk; ++i) knots[i] = hi;
  return 0;
}

/* ======================================================================
 * Stage B
 * ====================================================================== */

static double vec_norm(const double* v, int n) {
  /* Eigen's norm() = sqrt(squaredNorm()); summation order of Eigen's
   * vectorised reduction is UNPINNED (sequential here). */
  double s = 0.0;
  for (int i = 0; i < n; ++i) s += v[i] * v[i];
  return sqrt(s);
}

/* splines/spline_utils.cc:25-45 */
static void corner_offset(const double* delta, int dim, double radius,
                          double* offset) {
  const double kMinNorm = 1e-6;
  const double kMinWaypointSpacingFactor = 4.0; /* spline_utils.h:46 */
  const double norm = vec_norm(delta, dim);
  if (norm > kMinNorm) {
    for (int d = 0; d < dim; ++d) offset[d] = delta[d] / norm;
  } else {
    for (int d = 0; d < dim; ++d) offset[d] = 0.0;
  }
  if (norm > kMinWaypointSpacingFactor * radius) {
    for (int d = 0; d < dim; ++d) offset[d] = offset[d] * radius;
  } else {
    /* offset * (1.0 / 4.0) * norm, left to right */
    const double f = 1.0 / kMinWaypointSpacingFactor;
    for (int d = 0; d < dim; ++d) offset[d] = (offset[d] * f) * norm;
  }
}

/* splines/spline_utils.cc:47-102 */
int orc_polyline_to_bspline3(const double* corners, int W, int dim,
                             double radius, double* ctrl) {
  double delta[64], off[64];
  if (W == 1) {
    for (int i = 0; i < 4; ++i) memcpy(ctrl + i * dim, corners, sizeof(double) * dim);
    return 4;
  }
  const int n = 3 * W - 2;
  for (int i = 0; i < W; ++i) {
    memcpy(ctrl + (size_t)3 * i * dim, corners + (size_t)i * dim,
           sizeof(double) * dim);
  }
  for (int i = 1; i < W - 1; ++i) {
    const int k = 3 * i, knext = 3 * (i + 1), klast = 3 * (i - 1);
    for (int d = 0; d < dim; ++d) delta[d] = ctrl[knext * dim + d] - ctrl[k * dim + d];
    corner_offset(delta, dim, radius, off);
    for (int d = 0; d < dim; ++d) ctrl[(k + 1) * dim + d] = ctrl[k * dim + d] + off[d];
    for (int d = 0; d < dim; ++d) delta[d] = ctrl[klast * dim + d] - ctrl[k * dim + d];
    corner_offset(delta, dim, radius, off);
    for (int d = 0; d < dim; ++d) ctrl[(k - 1) * dim + d] = ctrl[k * dim + d] + off[d];
  }
  for (int d = 0; d < dim; ++d) delta[d] = ctrl[3 * dim + d] - ctrl[d];
  corner_offset(delta, dim, radius, off);
  for (int d = 0; d < dim; ++d) ctrl[1 * dim + d] = ctrl[d] + off[d];
  for (int d = 0; d < dim; ++d) delta[d] = ctrl[(n - 4) * dim + d] - ctrl[(n - 1) * dim + d];
  corner_offset(delta, dim, radius, off);
  for (int d = 0; d < dim; ++d) ctrl[(n - 2) * dim + d] = ctrl[(n - 1) * dim + d] + off[d];
  return n;
}

/* timeable_path_joint_spline.cc:252-292 (degree kSplineOrder = 2, .h:89) */
int orc_joint_fit(const double* waypoints, int W, int dim, double rounding,
                  double* ctrl, double* knots) {
  double diff[64];
  const int degree = 2;
  const int n = orc_polyline_to_bspline3(waypoints, W, dim, rounding, ctrl);
  const int nk = n + degree + 1;
  orc_make_uniform_knots(n, degree, 0.0, 1.0, knots);
  double length = 0.0;
  for (int i = 0; i < n - 1; ++i) {
    for (int d = 0; d < dim; ++d) diff[d] = ctrl[(i + 1) * dim + d] - ctrl[i * dim + d];
    length += vec_norm(diff, dim);
  }
  const double kMinimumFinalKnotValue = 0.1;
  const double kPathParameterPerPolygonLength = 1.0;
  const double lw = length * kPathParameterPerPolygonLength;
  const double weighted = (lw < kMinimumFinalKnotValue) ? kMinimumFinalKnotValue : lw;
  for (int i = 0; i < nk; ++i) knots[i] *= weighted;
  return n;
}

/* timeable_path_joint_spline.cc:294-318 */
void orc_joint_sample(const double* knots, int num_knots, const double* ctrl,
                      int dim, int N, double path_start, double delta,
                      double* q, double* dq, double* ddq) {
  const int degree = 2;
  const int n = num_knots - degree - 1;
  double vals[3 * 64];
  const double k0 = knots[0], k1 = knots[num_knots - 1];
  for (int idx = 0; idx < N; ++idx) {
    const double parameter = path_start + idx * delta;
    if (parameter < k1 + delta) {
      /* std::clamp(v, lo, hi) */
      const double u = (parameter < k0) ? k0 : ((k1 < parameter) ? k1 : parameter);
      orc_eval_curve_ders(knots, num_knots, ctrl, dim, degree, 2, u, vals);
      memcpy(q + (size_t)idx * dim, vals, sizeof(double) * dim);
      memcpy(dq + (size_t)idx * dim, vals + dim, sizeof(double) * dim);
      memcpy(ddq + (size_t)idx * dim, vals + 2 * dim, sizeof(double) * dim);
    } else {
      memcpy(q + (size_t)idx * dim, ctrl + (size_t)(n - 1) * dim, sizeof(double) * dim);
      for (int d = 0; d < dim; ++d) {
        dq[(size_t)idx * dim + d] = 0.0;
        ddq[(size_t)idx * dim + d] = 0.0;
      }
    }
  }
}

/* timeable_path_joint_spline.cc:320-343; std::pow(x, 2) == x*x exactly. */
void orc_joint_constraints(const double* dq, const double* ddq, int N, int dim,
                           const double* vmax, const double* amax,
                           double safety, double* cons) {
  const int C = 2 * dim;
  for (int idx = 0; idx < N; ++idx) {
    double* row = cons + (size_t)idx * C * 4;
    for (int d = 0; d < dim; ++d) {
      const double d1 = dq[(size_t)idx * dim + d];
      const double d2 = ddq[(size_t)idx * dim + d];
      row[d * 4 + 0] = d1;
      row[d * 4 + 1] = d2;
      row[d * 4 + 3] = amax[d] * safety;
      row[d * 4 + 2] = -amax[d] * safety;
      row[(dim + d) * 4 + 0] = 0.0;
      row[(dim + d) * 4 + 1] = d1 * d1;
      const double vs = vmax[d] * safety;
      row[(dim + d) * 4 + 3] = vs * vs;
      row[(dim + d) * 4 + 2] = 0.0;
    }
  }
}

/* ======================================================================
 * Stage C  (time_optimal_path_timing.cc)
 * ====================================================================== */

#define COL_A 0
#define COL_B 1
#define COL_LO 2
#define COL_HI 3

enum { T_NONE = 0, T_SOURCE = 1, T_SINK = 2, T_TRAJ = 4 }; /* .h:226-231 */
enum { CT_UPPER = 1, CT_LOWER = 2 };                       /* .h:268 */

typedef struct {
  int N, C;
  const double* cons;
  double ds;
  double *sd2, *sdd;
  double *b_sd2_max, *b_sdd_max, *b_sdd_min, *b_sd2_zero;
  unsigned char *b_at0, *b_type;
} solver_t;

static int is_tiny(double v) { return fabs(v) < kTiny; }

/* :624-636 */
static int derivs_valid(const solver_t* S, int idx, double sdd, double sd2) {
  const double* c = S->cons + (size_t)idx * S->C * 4;
  for (int i = 0; i < S->C; ++i) {
    const double v = c[i * 4 + COL_A] * sdd + c[i * 4 + COL_B] * sd2;
    if (v + kTiny < c[i * 4 + COL_LO] || v - kTiny > c[i * 4 + COL_HI]) return 0;
  }
  return 1;
}

/* :638-666 */
static double find_sdd_max(const solver_t* S, int idx, double sd2) {
  double sdd = -DBL_MAX;
  const double* c = S->cons + (size_t)idx * S->C * 4;
  for (int i = 0; i < S->C; ++i) {
    const double A = c[i * 4 + COL_A], B = c[i * 4 + COL_B];
    if (!is_tiny(A)) {
      double cand = (c[i * 4 + COL_LO] - B * sd2) / A;
      if (cand > sdd && derivs_valid(S, idx, cand, sd2)) sdd = cand;
      cand = (c[i * 4 + COL_HI] - B * sd2) / A;
      if (cand > sdd && derivs_valid(S, idx, cand, sd2)) sdd = cand;
    }
  }
  if (sdd == -DBL_MAX) sdd = 0;
  return sdd;
}

/* :668-695 */
static double find_sdd_min(const solver_t* S, int idx, double sd2) {
  double sdd = DBL_MAX;
  const double* c = S->cons + (size_t)idx * S->C * 4;
  for (int i = 0; i < S->C; ++i) {
    const double A = c[i * 4 + COL_A], B = c[i * 4 + COL_B];
    if (!is_tiny(A)) {
      double cand = (c[i * 4 + COL_LO] - B * sd2) / A;
      if (cand < sdd && derivs_valid(S, idx, cand, sd2)) sdd = cand;
      cand = (c[i * 4 + COL_HI] - B * sd2) / A;
      if (cand < sdd && derivs_valid(S, idx, cand, sd2)) sdd = cand;
    }
  }
  if (sdd == DBL_MAX) sdd = 0;
  return sdd;
}

/* :1105-1147 */
static int lp_is_optimal(const double* c, int i1, int i2, int t1, int t2) {
  const double A1 = c[i1 * 4 + COL_A], B1 = c[i1 * 4 + COL_B];
  const double A2 = c[i2 * 4 + COL_A], B2 = c[i2 * 4 + COL_B];
  const double denom = A2 * B1 - A1 * B2;
  if (fabs(denom) < kTiny) return 0;
  const double x = denom * A1, y = denom * (-A2);
  if (t1 == CT_UPPER) {
    if (t2 == CT_UPPER) return x <= 0 && y <= 0;
    return x >= 0 && y <= 0;
  }
  if (t2 == CT_UPPER) return x <= 0 && y >= 0;
  return x >= 0 && y >= 0;
}

typedef struct {
  int index, type;
  double slope;
} active_t;

/* :1149-1363.  The reference's constraint_set_ vector (erase keeps order) is
 * a membership flag per (row, type) slot visited in ascending slot order. */
void orc_find_max_sd2_simplex(const double* c, int C, double* sd2max,
                              double* sddmax, double* sd2zero) {
  unsigned char* in_set = (unsigned char*)malloc((size_t)2 * C);
  active_t* act = (active_t*)malloc(sizeof(active_t) * ((size_t)2 * C + 2));
  int nact = 0;
  *sd2max = 0;
  *sddmax = 0;
  memset(in_set, 1, (size_t)2 * C);

  double sd2 = DBL_MAX, sdd = 0.0;
  for (int i = 0; i < C; ++i) {
    const double B = c[i * 4 + COL_B];
    if (fabs(B) < kTiny) continue;
    if (B > kTiny) {
      const double invB = 1.0 / B;
      const double tmp = c[i * 4 + COL_HI] * invB;
      if (tmp < (sd2 + kTiny) && tmp > 0) {
        if (tmp < sd2 - kTiny) nact = 0;
        act[nact].index = i;
        act[nact].type = CT_UPPER;
        act[nact].slope = fabs(c[i * 4 + COL_A] * invB);
        ++nact;
        sd2 = tmp;
      }
    } else if (B < -kTiny) {
      const double invB = 1.0 / B;
      const double tmp = c[i * 4 + COL_LO] * invB;
      if (tmp < (sd2 + kTiny) && tmp > 0) {
        if (tmp < sd2 - kTiny) nact = 0;
        act[nact].index = i;
        act[nact].type = CT_LOWER;
        act[nact].slope = fabs(c[i * 4 + COL_A] * invB);
        ++nact;
        sd2 = tmp;
      }
    }
  }
  if (sd2 > kMaxSd2 || nact == 0) {
    *sd2zero = kMaxSd2;
    *sd2max = kMaxSd2;
    *sddmax = 0.0;
    goto done;
  }
  *sd2zero = sd2;

  active_t search = act[0];
  if (nact >= 2) {
    for (int f = 0; f < nact; ++f) {
      if (act[f].slope < search.slope) search = act[f];
      for (int s = f + 1; s < nact; ++s) {
        if (lp_is_optimal(c, act[f].index, act[s].index, act[f].type, act[s].type)) {
          *sd2max = sd2;
          *sddmax = 0.0;
          goto done;
        }
      }
    }
  }
  for (int l = 0; l < nact; ++l) {
    in_set[2 * act[l].index + (act[l].type == CT_LOWER ? 1 : 0)] = 0;
  }

  for (int loop = 0; loop < C; ++loop) {
    const double As = c[search.index * 4 + COL_A];
    if (fabs(As) < kTiny) {
      *sd2max = sd2;
      *sddmax = sdd;
      goto done;
    }
    const double invA = 1.0 / As;
    const double b = -c[search.index * 4 + COL_B] * invA;
    const double a = (search.type == CT_UPPER) ? c[search.index * 4 + COL_HI] * invA
                                               : c[search.index * 4 + COL_LO] * invA;
    nact = 0;
    double next_sd2 = DBL_MAX, next_sdd = 0.0;
    for (int e = 0; e < 2 * C; ++e) {
      if (!in_set[e]) continue;
      const int ci = e >> 1;
      const int is_lower = e & 1;
      const double Ac = c[ci * 4 + COL_A];
      const double B = Ac * b + c[ci * 4 + COL_B];
      if (fabs(B) < kTiny) continue;
      const double invB = 1.0 / B;
      const double bound = is_lower ? c[ci * 4 + COL_LO] : c[ci * 4 + COL_HI];
      const double tmp = (bound - Ac * a) * invB;
      if (tmp < (next_sd2 + kTiny) && tmp > sd2) {
        if (tmp < next_sd2 - kTiny) nact = 0;
        act[nact].index = ci;
        act[nact].type = is_lower ? CT_LOWER : CT_UPPER;
        act[nact].slope = fabs(Ac * invB);
        ++nact;
        next_sd2 = tmp;
        next_sdd = a + b * next_sd2;
      }
    }
    if (nact == 0) {
      *sd2max = *sd2zero;
      *sddmax = 0.0;
      goto done;
    }
    act[nact++] = search;
    search.index = -1;
    search.type = 0;
    search.slope = DBL_MAX;
    for (int f = 0; f < nact; ++f) {
      if (f != nact - 1 && act[f].slope < search.slope) search = act[f];
      for (int s = f + 1; s < nact; ++s) {
        if (lp_is_optimal(c, act[f].index, act[s].index, act[f].type, act[s].type)) {
          *sd2max = next_sd2;
          *sddmax = next_sdd;
          if (next_sd2 > kMaxSd2) {
            *sd2max = kMaxSd2;
            *sddmax = 0.0;
          }
          goto done;
        }
      }
    }
    --nact;
    for (int l = 0; l < nact; ++l) {
      in_set[2 * act[l].index + (act[l].type == CT_LOWER ? 1 : 0)] = 0;
    }
    sd2 = next_sd2;
    sdd = next_sdd;
  }
  *sd2max = *sd2zero;
  *sddmax = 0.0;
done:
  free(in_set);
  free(act);
}

/* :1365-1487 */
static void calc_boundary(solver_t* S) {
  const int N = S->N;
  for (int i = 0; i < N; ++i) {
    double m, a, z;
    orc_find_max_sd2_simplex(S->cons + (size_t)i * S->C * 4, S->C, &m, &a, &z);
    S->b_sd2_max[i] = m;
    S->b_sd2_zero[i] = z;
    S->b_sdd_max[i] = find_sdd_max(S, i, m);
    S->b_sdd_min[i] = find_sdd_min(S, i, m);
    S->b_at0[i] = fabs(S->b_sd2_max[i] - S->b_sd2_zero[i]) < kTiny;
  }
  S->b_type[0] = T_NONE;
  S->b_type[N - 1] = T_NONE;

  int nwork = 0;
  int* work_i = (int*)malloc(sizeof(int) * (size_t)(N > 0 ? N : 1));
  double* work_v = (double*)malloc(sizeof(double) * (size_t)(N > 0 ? N : 1));
  for (int i = 1; i < N - 1; ++i) {
    if (!S->b_at0[i - 1] && S->b_at0[i] && !S->b_at0[i + 1]) {
      S->b_sd2_max[i - 1] = S->b_sd2_zero[i - 1];
      S->b_sdd_max[i - 1] = find_sdd_max(S, i - 1, S->b_sd2_max[i - 1]);
      S->b_sdd_min[i - 1] = find_sdd_min(S, i - 1, S->b_sd2_max[i - 1]);
      S->b_sd2_max[i + 1] = S->b_sd2_zero[i + 1];
      S->b_sdd_max[i + 1] = find_sdd_max(S, i + 1, S->b_sd2_max[i + 1]);
      /* sic: the reference assigns FindSddMax to sdd_min here (:1394-1395) */
      S->b_sdd_min[i + 1] = find_sdd_max(S, i + 1, S->b_sd2_max[i + 1]);
    }
    const double sd2p = (S->b_sd2_max[i + 1] - S->b_sd2_max[i]) / S->ds;
    const double sd2p_min = 2 * S->b_sdd_min[i];
    const double sd2p_max = 2 * S->b_sdd_max[i];
    const int sink_or_source = (sd2p < sd2p_min) || (sd2p > sd2p_max);
    const int skipped_sdd = (S->b_sdd_max[i] > 0) && (S->b_sdd_min[i + 1] < 0);
    const int skipped_sd2 = (S->b_sd2_max[i] > S->b_sd2_max[i - 1] - kTiny) &&
                            (S->b_sd2_max[i] > S->b_sd2_max[i + 1] - kTiny);
    if ((skipped_sd2 || skipped_sdd) && sink_or_source) {
      const double sm1 = S->b_sd2_max[i - 1];
      const double fw = sm1 + 2.0 * S->ds * find_sdd_max(S, i - 1, sm1);
      const double sp1 = S->b_sd2_max[i + 1];
      const double bw = sp1 - 2.0 * S->ds * find_sdd_min(S, i + 1, sp1);
      double m = S->b_sd2_zero[i];
      if (fw < m) m = fw;
      if (bw < m) m = bw;
      const double v = (0.0 < m) ? m : 0.0;
      work_i[nwork] = i;
      work_v[nwork] = v;
      ++nwork;
    }
  }
  for (int w = 0; w < nwork; ++w) {
    const int index = work_i[w];
    const double value = work_v[w];
    S->b_sd2_max[index] = value;
    S->b_sdd_max[index] = find_sdd_max(S, index, value);
    S->b_sdd_min[index] = find_sdd_min(S, index, value);
    if (index > 0) {
      S->b_sd2_max[index - 1] = S->b_sd2_zero[index - 1];
      S->b_sdd_max[index - 1] = find_sdd_max(S, index - 1, S->b_sd2_max[index - 1]);
      S->b_sdd_min[index - 1] = find_sdd_min(S, index - 1, S->b_sd2_max[index - 1]);
    }
    if (index < N - 1) {
      S->b_sd2_max[index + 1] = S->b_sd2_zero[index + 1];
      S->b_sdd_max[index + 1] = find_sdd_max(S, index + 1, S->b_sd2_max[index + 1]);
      S->b_sdd_min[index + 1] = find_sdd_min(S, index + 1, S->b_sd2_max[index + 1]);
    }
  }
  free(work_i);
  free(work_v);

  for (int i = 1; i < N - 1; ++i) {
    const double sd2p = (S->b_sd2_max[i + 1] - S->b_sd2_max[i]) / S->ds;
    const double sd2p_min = 2 * S->b_sdd_min[i];
    const double sd2p_max = 2 * S->b_sdd_max[i];
    unsigned char t = T_NONE;
    if (sd2p < sd2p_min) {
      t = T_SINK;
    } else if (sd2p > sd2p_max) {
      t = T_SOURCE;
    }
    if ((sd2p <= sd2p_max) && (sd2p >= sd2p_min)) t = T_TRAJ;
    S->b_type[i] = t;
  }
}

/* :722-751 */
static void sdd_at_intersection(solver_t* S, int index) {
  double cand[3];
  int n = 0;
  const int N = S->N;
  if (index > 0 && index < N - 1) {
    cand[n++] = 0.25 / S->ds * (S->sd2[index + 1] - S->sd2[index - 1]);
  }
  if (index < N - 1) cand[n++] = 0.5 / S->ds * (S->sd2[index + 1] - S->sd2[index]);
  if (index > 0) cand[n++] = 0.5 / S->ds * (S->sd2[index] - S->sd2[index - 1]);
  S->sdd[index] = 0.0;
  for (int k = 0; k < n; ++k) {
    if (derivs_valid(S, index, cand[k], S->sd2[index])) {
      S->sdd[index] = cand[k];
      return;
    }
  }
}

/* :769-857 */
static int add_forward(solver_t* S, int idx_lo) {
  const int N = S->N;
  for (int idx = idx_lo; idx < N - 2; ++idx) {
    double sd2tmp, sddtmp;
    const int on_boundary = is_tiny(S->sd2[idx] - S->b_sd2_max[idx]);
    if (on_boundary && (S->b_type[idx] & T_TRAJ) && (idx < N - 1) &&
        (S->b_type[idx + 1] & T_TRAJ)) {
      sd2tmp = S->b_sd2_max[idx + 1];
      sddtmp = 0.5 * (sd2tmp - S->sd2[idx]) / S->ds;
    } else {
      sddtmp = find_sdd_max(S, idx, S->sd2[idx]);
      sd2tmp = S->sd2[idx] + 2.0 * S->ds * sddtmp;
    }
    if (!isnan(S->sd2[idx + 1]) && (S->sd2[idx + 1] < sd2tmp)) {
      sdd_at_intersection(S, idx);
      return N - 1;
    }
    if (sd2tmp > S->b_sd2_max[idx + 1]) {
      const double sdd_bound = 0.5 * (S->b_sd2_max[idx + 1] - S->sd2[idx]) / S->ds;
      const int deriv_invalid = !derivs_valid(S, idx, sdd_bound, S->b_sd2_max[idx]);
      const int type_invalid = S->b_type[idx + 1] & T_SINK;
      if (type_invalid || deriv_invalid) return idx;
      sd2tmp = S->b_sd2_max[idx + 1];
      sddtmp = sdd_bound;
    }
    if (sd2tmp < 0) {
      sd2tmp = 0.0;
      if (idx <= 1) {
        sddtmp = 0.0;
      } else {
        sddtmp = -S->sd2[idx - 1] / S->ds;
      }
    }
    S->sd2[idx + 1] = sd2tmp;
    S->sdd[idx] = sddtmp;
  }
  return N - 1;
}

/* :859-952 */
static int add_backward(solver_t* S, int idx_hi) {
  const int N = S->N;
  for (int idx = idx_hi; idx > 1; --idx) {
    double sd2tmp, sddtmp;
    const int on_boundary = is_tiny(S->sd2[idx] - S->b_sd2_max[idx]);
    if (on_boundary && (S->b_type[idx] & T_TRAJ) && (idx > 0) &&
        (S->b_type[idx - 1] & T_TRAJ)) {
      sd2tmp = S->b_sd2_max[idx - 1];
      sddtmp = 0.5 * (S->sd2[idx] - sd2tmp) / S->ds;
    } else {
      sddtmp = find_sdd_min(S, idx, S->sd2[idx]);
      sd2tmp = S->sd2[idx] - 2.0 * S->ds * sddtmp;
    }
    if (!isnan(S->sd2[idx - 1]) && (S->sd2[idx - 1] < sd2tmp)) {
      sdd_at_intersection(S, idx);
      return 0;
    }
    if (sd2tmp > S->b_sd2_max[idx - 1]) {
      const double sdd_bound = 0.5 * (S->sd2[idx] - S->b_sd2_max[idx - 1]) / S->ds;
      const int deriv_invalid = !derivs_valid(S, idx, sdd_bound, S->sd2[idx]);
      const int type_invalid = S->b_type[idx - 1] & T_SOURCE;
      const int is_connecting = (idx_hi != (N - 1));
      if ((type_invalid || deriv_invalid) && !is_connecting) return idx;
      sd2tmp = S->b_sd2_max[idx - 1];
      sddtmp = sdd_bound;
    }
    if (sd2tmp < 0) {
      sd2tmp = 0.0;
      if (idx < N - 1) {
        sddtmp = S->sd2[idx + 1] / S->ds;
      } else {
        sddtmp = 0.0;
      }
    }
    S->sd2[idx - 1] = sd2tmp;
    S->sdd[idx] = sddtmp;
  }
  return 0;
}

/* :697-720 (note index 0 in the isolated-point test, sic) */
static int next_critical(const solver_t* S, int idx_lo, int idx_hi) {
  int crit = -1;
  for (int idx = idx_lo + 1; idx <= idx_hi; ++idx) {
    if (crit < 0) {
      if ((S->b_type[idx] & T_SOURCE) || (S->b_type[idx] & T_TRAJ)) crit = idx;
    } else {
      if (S->b_sd2_max[idx] == S->b_sd2_zero[0]) crit = idx;
    }
    if ((crit > 0) && !isnan(S->sd2[idx])) return crit;
  }
  return -1;
}

/* SetupProblem :161-203, SetSetupDone :535-552, IsSetupValid :554-576,
 * OptimizePathParameter :287-490 */
int orc_solve(int N, int C, const double* cons, double s_start, double s_end,
              double sd_start, double sdd_start, double time_start,
              int max_loops, orc_solution* out) {
  if (N < 2) return ORC_FAIL_SETUP;
  for (int i = 0; i < N; ++i) {
    const double* c = cons + (size_t)i * C * 4;
    double mx = c[COL_HI] - c[COL_LO];
    for (int r = 1; r < C; ++r) {
      const double v = c[r * 4 + COL_HI] - c[r * 4 + COL_LO];
      mx = (v > mx) ? v : mx;
    }
    if (mx <= 0) return ORC_FAIL_SETUP;
  }
  if (s_start >= s_end) return ORC_FAIL_SETUP;
  if (sd_start < 0) return ORC_FAIL_SETUP;
  for (int i = 0; i < N; ++i) {
    const double* c = cons + (size_t)i * C * 4;
    for (int r = 0; r < C; ++r) {
      if (c[r * 4 + COL_LO] >= c[r * 4 + COL_HI]) return ORC_FAIL_SETUP;
    }
  }

  solver_t S;
  int status = ORC_OK;
  S.N = N;
  S.C = C;
  S.cons = cons;
  S.ds = (s_end - s_start) / (N - 1);
  const double nan = NAN;
  double* buf = (double*)malloc(sizeof(double) * (size_t)N * 9);
  unsigned char* bbuf = (unsigned char*)malloc((size_t)N * 2);
  S.sd2 = buf;
  S.sdd = buf + N;
  S.b_sd2_max = buf + 2 * (size_t)N;
  S.b_sdd_max = buf + 3 * (size_t)N;
  S.b_sdd_min = buf + 4 * (size_t)N;
  S.b_sd2_zero = buf + 5 * (size_t)N;
  double* sd = buf + 6 * (size_t)N;
  double* tm = buf + 7 * (size_t)N;
  double* sv = buf + 8 * (size_t)N;
  S.b_at0 = bbuf;
  S.b_type = bbuf + N;
  memset(bbuf, 0, (size_t)N * 2);
  for (int i = 0; i < N; ++i) {
    sv[i] = S.ds * i + s_start;
    S.sd2[i] = nan;
    S.sdd[i] = nan;
    S.b_sd2_max[i] = S.b_sdd_max[i] = S.b_sdd_min[i] = S.b_sd2_zero[i] = 0.0;
    sd[i] = nan;
    tm[i] = nan;
  }
  sv[N - 1] = s_end;
  S.sd2[0] = sd_start * sd_start;
  S.sd2[N - 1] = 0;
  int last_extremal = 0, loops_done = 0;

  calc_boundary(&S);

  int iforw_lo = 0, iback_hi = N - 1, iback_lo, iforw_hi, icrit, icrit_lo, icrit_hi;
  iback_lo = add_backward(&S, iback_hi);
  iforw_hi = add_forward(&S, iforw_lo);
  icrit_hi = iback_lo;
  if ((iforw_hi < icrit_hi) && ((icrit_hi < N - 2) && (icrit_hi >= 2))) {
    S.sd2[icrit_hi] = nan;
    icrit_hi++;
    iback_lo++;
  }
  icrit_lo = iforw_hi;
  for (int loop = 0; loop < max_loops; ++loop) {
    if (iforw_hi >= icrit_hi) break;
    ++loops_done;
    icrit = next_critical(&S, icrit_lo, icrit_hi);
    if (icrit < 0 || icrit >= N) icrit = (int)(0.5 * (icrit_lo + icrit_hi));
    if (icrit > 0 && icrit < N - 1) S.sd2[icrit] = S.b_sd2_max[icrit];
    if (icrit < 1) {
      status = ORC_FAIL_INDEX;
      goto finish;
    }
    if (S.b_sd2_max[icrit - 1] <= S.b_sd2_max[icrit]) {
      iback_hi = icrit - 1;
      S.sd2[icrit - 1] = S.b_sd2_max[icrit - 1];
    } else {
      iback_hi = icrit;
    }
    iback_lo = add_backward(&S, iback_hi);
    iforw_lo = icrit;
    iforw_hi = add_forward(&S, iforw_lo);
    if (iback_lo > icrit_lo) {
      status = ORC_FAIL_CONNECT;
      goto finish;
    }
    icrit_lo = iforw_hi;
  }
  for (int idx = 0; idx < N; ++idx) {
    if (isnan(S.sd2[idx])) {
      status = ORC_FAIL_NAN;
      goto finish;
    }
    if (isnan(S.sdd[idx])) sdd_at_intersection(&S, idx);
  }
  if (derivs_valid(&S, 0, sdd_start, S.sd2[0])) S.sdd[0] = sdd_start;

  for (int i = 0; i < N; ++i) sd[i] = sqrt(S.sd2[i]);
  if (S.sd2[N - 1] != 0) {
    status = ORC_FAIL_TERMINAL;
    goto finish;
  }
  last_extremal = (N - 2 > 1) ? N - 2 : 1;
  while (last_extremal >= 1) {
    if (S.sdd[last_extremal] > 0.0 ||
        fabs(S.sd2[last_extremal] - S.b_sd2_max[last_extremal]) < kTiny) {
      break;
    }
    last_extremal--;
  }
  tm[0] = time_start;
  for (int idx = 1; idx < N; ++idx) {
    if ((S.sd2[idx - 1] > 0) || (S.sd2[idx] > 0)) {
      const double dt = 2.0 * S.ds / (sd[idx - 1] + sd[idx]);
      tm[idx] = tm[idx - 1] + dt;
    } else {
      tm[idx] = tm[idx - 1];
      S.sdd[idx - 1] = 0;
      S.sdd[idx] = 0;
    }
  }

finish:
  if (out) {
    if (out->time) memcpy(out->time, tm, sizeof(double) * N);
    if (out->s) memcpy(out->s, sv, sizeof(double) * N);
    if (out->sd) memcpy(out->sd, sd, sizeof(double) * N);
    if (out->sdd) memcpy(out->sdd, S.sdd, sizeof(double) * N);
    if (out->sd2) memcpy(out->sd2, S.sd2, sizeof(double) * N);
    if (out->b_sd2_max) memcpy(out->b_sd2_max, S.b_sd2_max, sizeof(double) * N);
    if (out->b_sdd_max) memcpy(out->b_sdd_max, S.b_sdd_max, sizeof(double) * N);
    if (out->b_sdd_min) memcpy(out->b_sdd_min, S.b_sdd_min, sizeof(double) * N);
    if (out->b_sd2_zero) memcpy(out->b_sd2_zero, S.b_sd2_zero, sizeof(double) * N);
    if (out->b_type) memcpy(out->b_type, S.b_type, (size_t)N);
    out->last_extremal_index = last_extremal;
    out->num_loops = loops_done;
  }
  free(buf);
  free(bbuf);
  return status;
}

/* :492-518 */
int orc_count_violations(int N, int C, const double* cons, const double* sd2,
                         const double* sdd) {
  int count = 0;
  for (int i = 0; i < N; ++i) {
    const double* c = cons + (size_t)i * C * 4;
    for (int r = 0; r < C; ++r) {
      const double v = c[r * 4 + COL_A] * sdd[i] + c[r * 4 + COL_B] * sd2[i];
      if ((v + kTiny < c[r * 4 + COL_LO]) || (v - kTiny > c[r * 4 + COL_HI])) ++count;
    }
  }
  return count;
}

/* SampleIndexFromTime :1497-1524 + GetPathParameterAndDerivatives :1549-1627 */
int orc_query(int N, const double* time, const double* s, const double* sd,
              const double* sd2, double s_start, double s_end, double t,
              double* s_out, double* sd_out, double* sdd_out) {
  const double ds_ = (s_end - s_start) / (N - 1);
  const double inv_ds = 1.0 / ds_;
  if (t <= time[0]) {
    *s_out = s_start;
    *sd_out = sd[0];
    *sdd_out = 0.5 * inv_ds * (sd2[1] - sd2[0]);
    return 1;
  }
  if (t >= time[N - 1]) {
    *s_out = s_end;
    *sd_out = sd[N - 1];
    *sdd_out = 0.0;
    return 1;
  }
  int low = 0;
  while ((time[low] == time[low + 1]) && (low < N - 2)) low++;
  int high = N - 1;
  while ((time[high] == time[high - 1]) && (high >= low)) high--;
  int mid = (low + high) / 2;
  while (t < time[mid] || t >= time[mid + 1]) {
    if (t < time[mid]) {
      high = mid;
    } else {
      low = mid;
    }
    mid = (low + high) / 2;
  }
  while ((time[mid] == time[mid + 1]) && (mid < N - 1)) mid++;
  const int k = mid;
  if (time[k] == time[k + 1]) {
    *s_out = s[k + 1];
    *sd_out = sd[k + 1];
    *sdd_out = 0.5 * inv_ds * (sd2[k + 1] - sd2[k]);
    return 1;
  }
  const double dt = t - time[k];
  if (k > N - 2) return 0;
  const double sda = sd[k], sdb = sd[k + 1], sd2a = sd2[k], sd2b = sd2[k + 1];
  if (sda > 0 || sdb > 0) {
    double ds = sda * dt + dt * dt * 0.25 * inv_ds * (sd2b - sd2a);
    if (ds > ds_) ds = ds_;
    if (dt < 0 || ds < 0) return 0;
    const double sn = s[k] + ds;
    *s_out = (s[k + 1] < sn) ? s[k + 1] : sn;
    *sd_out = sqrt(sd2a + ds * inv_ds * (sd2b - sd2a));
    *sdd_out = 0.5 * inv_ds * (sd2b - sd2a);
  } else {
    *s_out = s[k] + (s[k + 1] - s[k]) * dt / (time[k + 1] - time[k]);
    *sd_out = 0.0;
    *sdd_out = 0.0;
  }
  return 1;
}

/* ======================================================================
 * Stage D  (path_timing_trajectory.cc, time.h)
 * ====================================================================== */

/* time.h:23-30: absl::FromUnixNanos(seconds * 1e9) truncates toward zero. */
int64_t orc_time_from_sec(double seconds) {
  return (int64_t)(seconds * (double)1000000000LL);
}
double orc_time_to_sec(int64_t ns) { return (double)ns / (double)1000000000LL; }

/* path_timing_trajectory.cc:458-472 */
void orc_assemble(const double* dq, const double* ddq, const double* sd,
                  const double* sdd, const double* amax, int N, int dim,
                  double* vel, double* acc) {
  for (int i = 0; i < N; ++i) {
    const double pv = sd[i], pa = sdd[i];
    const double pv2 = pv * pv; /* eigenmath::Square */
    for (int d = 0; d < dim; ++d) {
      const double d1 = dq[(size_t)i * dim + d], d2 = ddq[(size_t)i * dim + d];
      vel[(size_t)i * dim + d] = d1 * pv;
      double a = d1 * pa + d2 * pv2;
      /* .cwiseMax(-amax).cwiseMin(amax) */
      a = (a < -amax[d]) ? -amax[d] : a;
      a = (amax[d] < a) ? amax[d] : a;
      acc[(size_t)i * dim + d] = a;
    }
  }
}

/* :755-757 */
int orc_resample_count(double t_end, double start_sec, double dt) {
  const double duration = t_end - start_sec;
  return (int)(ceil(duration / dt) + 1);
}

static double lerp(double t, double a, double b) {
  /* eigenmath::InterpolateLinear is un-vendored: UNPINNED choice a+t*(b-a) */
  return a + t * (b - a);
}

/* :686-695, :709-753, :755-783 */
void orc_resample_uniform(int N, int dim, const double* time_ps,
                          const double* s_ps, const double* sd_ps,
                          const double* sdd_ps, const double* pos_ps,
                          const double* vel_ps, const double* acc_ps,
                          const double* amax, double start_sec, double dt,
                          int M, double* t, double* pos, double* vel,
                          double* acc, double* s, double* sd, double* sdd) {
  int lower = 0;
  for (int i = 0; i < M; ++i) {
    const double time = start_sec + dt * i;
    int li = N - 1;
    for (int index = lower; index < N - 1; ++index) {
      if (time_ps[index + 1] > time) {
        li = index;
        break;
      }
    }
    lower = li;
    const int ui = (N - 1 < li + 1) ? N - 1 : li + 1;
    const double at = (fabs(time_ps[ui] - time_ps[li]) < DBL_EPSILON)
                          ? 0.5
                          : (time - time_ps[li]) / (time_ps[ui] - time_ps[li]);
    t[i] = time;
    for (int d = 0; d < dim; ++d) {
      pos[(size_t)i * dim + d] = lerp(at, pos_ps[(size_t)li * dim + d], pos_ps[(size_t)ui * dim + d]);
      vel[(size_t)i * dim + d] = lerp(at, vel_ps[(size_t)li * dim + d], vel_ps[(size_t)ui * dim + d]);
      double a = lerp(at, acc_ps[(size_t)li * dim + d], acc_ps[(size_t)ui * dim + d]);
      a = (a < -amax[d]) ? -amax[d] : a;
      a = (amax[d] < a) ? amax[d] : a;
      acc[(size_t)i * dim + d] = a;
    }
    s[i] = lerp(at, s_ps[li], s_ps[ui]);
    sd[i] = lerp(at, sd_ps[li], sd_ps[ui]);
    sdd[i] = lerp(at, sdd_ps[li], sdd_ps[ui]);
  }
  if (M > 0) {
    for (int d = 0; d < dim; ++d) {
      pos[(size_t)(M - 1) * dim + d] = pos_ps[(size_t)(N - 1) * dim + d];
      vel[(size_t)(M - 1) * dim + d] = 0.0;
      acc[(size_t)(M - 1) * dim + d] = 0.0;
    }
  }
}

/* ResampleSkippingSamplesCloserThanTimeStep (:785-836) with InterpolateAtTime
 * (:709-753) for the first row and GetMinTimeDeltaToKeep (:893-900).  Output
 * capacity N + 1 rows; returns the number of rows. */
int orc_resample_skip(int N, int dim, const double* time_ps, const double* s_ps,
                      const double* sd_ps, const double* sdd_ps, const double* pos_ps,
                      const double* vel_ps, const double* acc_ps, const double* amax,
                      double start_sec, double dt, double* t, double* pos, double* vel,
                      double* acc, double* s, double* sd, double* sdd) {
  /* TimeAtPathSamplesLowerIndex(0, start_sec) (:686-695) */
  int li = N - 1;
  for (int index = 0; index < N - 1; ++index) {
    if (time_ps[index + 1] > start_sec) {
      li = index;
      break;
    }
  }
  const double min_delta = 0.95 * dt;
  /* first row: values interpolated at start_sec (:808-816) */
  {
    const int ui = (N - 1 < li + 1) ? N - 1 : li + 1;
    const double at = (fabs(time_ps[ui] - time_ps[li]) < DBL_EPSILON)
                          ? 0.5
                          : (start_sec - time_ps[li]) / (time_ps[ui] - time_ps[li]);
    t[0] = start_sec;
    for (int d = 0; d < dim; ++d) {
      pos[d] = lerp(at, pos_ps[(size_t)li * dim + d], pos_ps[(size_t)ui * dim + d]);
      vel[d] = lerp(at, vel_ps[(size_t)li * dim + d], vel_ps[(size_t)ui * dim + d]);
      double a = lerp(at, acc_ps[(size_t)li * dim + d], acc_ps[(size_t)ui * dim + d]);
      a = (a < -amax[d]) ? -amax[d] : a;
      a = (amax[d] < a) ? amax[d] : a;
      acc[d] = a;
    }
    s[0] = lerp(at, s_ps[li], s_ps[ui]);
    sd[0] = lerp(at, sd_ps[li], sd_ps[ui]);
    sdd[0] = lerp(at, sdd_ps[li], sdd_ps[ui]);
  }
  int M = 1;
  for (int i = li + 1; i < N; ++i) {
    if (fabs(time_ps[i] - t[M - 1]) < min_delta) continue;
    t[M] = time_ps[i];
    for (int d = 0; d < dim; ++d) {
      pos[(size_t)M * dim + d] = pos_ps[(size_t)i * dim + d];
      vel[(size_t)M * dim + d] = vel_ps[(size_t)i * dim + d];
      acc[(size_t)M * dim + d] = acc_ps[(size_t)i * dim + d];
    }
    s[M] = s_ps[i];
    sd[M] = sd_ps[i];
    sdd[M] = sdd_ps[i];
    ++M;
  }
  /* the last KEPT row gets the final position and zero derivatives (:832-835) */
  for (int d = 0; d < dim; ++d) {
    pos[(size_t)(M - 1) * dim + d] = pos_ps[(size_t)(N - 1) * dim + d];
    vel[(size_t)(M - 1) * dim + d] = 0.0;
    acc[(size_t)(M - 1) * dim + d] = 0.0;
  }
  return M;
}

/* SetWaypoints (timeable_path_joint_spline.cc:200-207) + Plan for a fresh
 * path in one window (path_timing_trajectory.cc:579-684 with :307-475):
 * path_start = 0, sd_start = sdd_start = 0, zero initial velocity,
 * max_solver_loops = max(100, 10 N) (:398-400). */
int orc_plan_joint(const orc_plan_cfg* cfg, const double* waypoints,
                   const double* vmax, const double* amax, orc_plan_out* out) {
  const int N = cfg->N, D = cfg->dim, W = cfg->W;
  const int nmax = (W == 1) ? 4 : 3 * W - 2;
  double* ctrl = (double*)malloc(sizeof(double) * (size_t)nmax * D);
  double* knots = (double*)malloc(sizeof(double) * (size_t)(nmax + 3));
  double* cons = (double*)malloc(sizeof(double) * (size_t)N * 2 * D * 4);
  const int n = orc_joint_fit(waypoints, W, D, cfg->rounding, ctrl, knots);
  const int nk = n + 3;
  const double kend = knots[nk - 1];
  const double delta = (cfg->delta > 0) ? cfg->delta : kend / (N - 1);
  out->knot_end = kend;
  out->delta_used = delta;
  orc_joint_sample(knots, nk, ctrl, D, N, 0.0, delta, out->q, out->dq, out->ddq);
  orc_joint_constraints(out->dq, out->ddq, N, D, vmax, amax, cfg->safety, cons);
  const double horizon = 0.0 + delta * (N - 1);
  const int max_loops = (100 > 10 * N) ? 100 : 10 * N;
  orc_solution sol;
  memset(&sol, 0, sizeof(sol));
  sol.time = out->time;
  sol.s = out->s;
  sol.sd = out->sd;
  sol.sdd = out->sdd;
  const int st = orc_solve(N, 2 * D, cons, 0.0, horizon, 0.0, 0.0, cfg->start_sec,
                           max_loops, &sol);
  out->last_extremal_index = sol.last_extremal_index;
  out->M = 0;
  out->duration = NAN;
  if (st == ORC_OK) {
    out->duration = out->time[N - 1] - out->time[0];
    orc_assemble(out->dq, out->ddq, out->sd, out->sdd, amax, N, D, out->vel, out->acc);
    if (cfg->dt > 0 && out->r_t) {
      if (cfg->skip_method && out->M_cap >= N + 1) {
        out->M = orc_resample_skip(N, D, out->time, out->s, out->sd, out->sdd, out->q,
                                   out->vel, out->acc, amax, cfg->start_sec, cfg->dt,
                                   out->r_t, out->r_pos, out->r_vel, out->r_acc, out->r_s,
                                   out->r_sd, out->r_sdd);
        free(ctrl);
        free(knots);
        free(cons);
        return st;
      }
      int M = orc_resample_count(out->time[N - 1], cfg->start_sec, cfg->dt);
      if (M > out->M_cap) M = out->M_cap;
      out->M = M;
      orc_resample_uniform(N, D, out->time, out->s, out->sd, out->sdd, out->q,
                           out->vel, out->acc, amax, cfg->start_sec, cfg->dt, M,
                           out->r_t, out->r_pos, out->r_vel, out->r_acc, out->r_s,
                           out->r_sd, out->r_sdd);
    }
  }
  free(ctrl);
  free(knots);
  free(cons);
  return st;
}
