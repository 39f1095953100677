// Device-side helpers shared by the path-timing kernels (sm_100a).
//
// Everything here is FP64.  The translation units that include this header are
// compiled with -fmad=false: the solver's decisions are thresholded at
// kTiny = 2.2e-11 (time_optimal_path_timing.h:275) and must agree with the
// reference's x86-64 build, which does not contract a*b+c into an FMA.
// Division and sqrt are IEEE-correct (nvcc defaults -prec-div/-prec-sqrt=true).
#pragma once

#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

namespace tpb {

constexpr double kTiny = DBL_EPSILON * 1e5;  // time_optimal_path_timing.h:275
constexpr double kMaxSd2 = 1e6;              // time_optimal_path_timing.h:279
constexpr unsigned kFullMask = 0xffffffffu;

// Boundary::type bits (time_optimal_path_timing.h:226-231)
enum : uint8_t { kNone = 0, kSource = 1, kSink = 2, kTrajectory = 4 };
// ConstraintType (time_optimal_path_timing.h:268)
enum : int { kUpper = 1, kLower = 2 };

// constraint rows per sample handled by the kernels (the reference has no limit; its own
// LP test runs 2..50 rows, time_optimal_path_timing_test.cc:703-736)
constexpr int kMaxRows = 64;
constexpr int kMaxDofs = 16;    // joint dofs (2*D rows <= kMaxRows)
constexpr int kMaxDegree = 5;   // dense spline evaluator

// Kernels with one CTA row per path fold the path index over gridDim.y and
// gridDim.z (gridDim.y alone ends at 65,535 paths; BASELINE cfg4 is 65,536).
constexpr int kMaxGridY = 32768;
inline dim3 path_grid(unsigned tiles, int num_paths) {
  const unsigned gy = static_cast<unsigned>(num_paths < kMaxGridY ? (num_paths > 0 ? num_paths : 1) : kMaxGridY);
  return dim3(tiles, gy, (static_cast<unsigned>(num_paths) + gy - 1) / gy);
}
__device__ __forceinline__ int path_of_block() { return blockIdx.z * gridDim.y + blockIdx.y; }

__device__ __forceinline__ double quiet_nan() {
  return __longlong_as_double(0x7ff8000000000000LL);
}
__device__ __forceinline__ bool is_tiny(double v) { return fabs(v) < kTiny; }
// AreDerivativesValid (:624-636) for a row with symmetric bounds -lim <= v <= lim:
//   (v + kTiny < -lim) || (v - kTiny > lim)
// in one subtraction and one comparison.  Exact: negation commutes with rounding,
// -(v + kTiny) == (-v) - kTiny, so the test is max(v - kTiny, -v - kTiny) > lim, and
// rounding is monotone, so that maximum is |v| - kTiny.  (NaN: false either way.)
__device__ __forceinline__ bool outside_sym(double v, double lim) {
  return fabs(v) - kTiny > lim;
}

// num / den where the quotient is only used when !skip.  Lanes with skip set
// divide 1 by 1 instead: x/0, 0/x and other special operands send the whole warp
// through the slow path of the double-precision division.  The empty asm keeps
// the compiler from folding the guard into the select that discards the result.
__device__ __forceinline__ double guarded_div(bool skip, double num, double den) {
  double n = skip ? 1.0 : num;
  double d = skip ? 1.0 : den;
  asm("" : "+d"(n), "+d"(d));
  return n / d;
}

// x / d for a positive, normal d.  A zero numerator (a flat stretch of the limit
// curve, zero velocity at a path end) would take the division's slow path; the
// quotient is x itself then, sign included.
__device__ __forceinline__ double div_pos(double x, double d) {
  const bool zero = x == 0.0;
  const double q = guarded_div(zero, x, d);
  return zero ? x : q;
}

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double w = __shfl_xor_sync(kFullMask, v, o);
    v = (w > v) ? w : v;
  }
  return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double w = __shfl_xor_sync(kFullMask, v, o);
    v = (w < v) ? w : v;
  }
  return v;
}

// ---------------------------------------------------------------------------
// B-spline basis (NURBS book A2.1 - A2.3) in the reference's operation order.
// ---------------------------------------------------------------------------

// BSplineBase::KnotSpan, splines/bspline_base.cc:218-246: upper_bound over
// [knots+p, knots+num_knots-p) minus one; u == last knot -> num_points-1.
__device__ __forceinline__ int knot_span(const double* __restrict__ knots,
                                         int num_knots, int degree, double u) {
  if (u == knots[num_knots - 1]) return num_knots - degree - 2;
  int lo = degree, hi = num_knots - degree;
  while (lo < hi) {
    const int mid = lo + ((hi - lo) >> 1);
    if (knots[mid] <= u) {
      lo = mid + 1;
    } else {
      hi = mid;
    }
  }
  return lo - 1;
}

// BSplineBase::UpdateBasisAndDerivatives, splines/bspline_base.cc:268-348.
// Fully unrolled so that ndu / a / ders live in registers.
template <int P, int NDER>
__device__ __forceinline__ void basis_ders(const double* __restrict__ knots, int span,
                                           double u, double (&ders)[NDER + 1][P + 1]) {
  static_assert(NDER <= P, "reference CHECKs der <= p (bspline_base.cc:270)");
  double ndu[P + 1][P + 1];
  double left[P + 1], right[P + 1];
  ndu[0][0] = 1.0;
#pragma unroll
  for (int j = 1; j <= P; ++j) {
    left[j] = u - knots[span + 1 - j];
    right[j] = knots[span + j] - u;
    double saved = 0.0;
#pragma unroll
    for (int r = 0; r < j; ++r) {
      ndu[j][r] = right[r + 1] + left[j - r];
      const double tmp = ndu[r][j - 1] / ndu[j][r];
      ndu[r][j] = saved + right[r + 1] * tmp;
      saved = left[j - r] * tmp;
    }
    ndu[j][j] = saved;
  }
#pragma unroll
  for (int j = 0; j <= P; ++j) ders[0][j] = ndu[j][P];

#pragma unroll
  for (int r = 0; r <= P; ++r) {
    double a[2][P + 1];
#pragma unroll
    for (int j = 0; j <= P; ++j) a[0][j] = a[1][j] = 0.0;
    a[0][0] = 1.0;
#pragma unroll
    for (int k = 1; k <= NDER; ++k) {
      const int s1 = (k - 1) & 1, s2 = k & 1;
      double d = 0.0;
      const int rk = r - k, pk = P - k;
      if (r >= k) {
        a[s2][0] = a[s1][0] / ndu[pk + 1][rk];
        d = a[s2][0] * ndu[rk][pk];
      }
      const int j1 = (rk >= -1) ? 1 : -rk;
      const int j2 = (r - 1 <= pk) ? k - 1 : P - r;
#pragma unroll
      for (int j = 1; j <= P; ++j) {
        if (j >= j1 && j <= j2) {
          a[s2][j] = (a[s1][j] - a[s1][j - 1]) / ndu[pk + 1][rk + j];
          d += a[s2][j] * ndu[rk + j][pk];
        }
      }
      if (r <= pk) {
        a[s2][k] = -a[s1][k - 1] / ndu[pk + 1][r];
        d += a[s2][k] * ndu[r][pk];
      }
      ders[k][r] = d;
    }
  }
  double f = static_cast<double>(P);
#pragma unroll
  for (int k = 1; k <= NDER; ++k) {
#pragma unroll
    for (int j = 0; j <= P; ++j) ders[k][j] *= f;
    f *= static_cast<double>(P - k);
  }
}

}  // namespace tpb
