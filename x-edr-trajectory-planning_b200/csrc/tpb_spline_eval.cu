// Dense B-spline evaluation with derivatives (stage A stand-alone, BASELINE
// config 3): BSplineT::EvalCurveAndDerivatives, splines/bspline.h:539-568, over
// many splines x many parameters.
//
// One CTA handles a tile of kTile consecutive parameters of one spline.  The
// spline's knot vector and control points are staged in shared memory once per
// CTA; each thread evaluates one parameter in FP64 registers (NURBS A2.1-A2.3
// in the reference's operation order), the (nder+1) x dim results of the tile
// are transposed through shared memory and leave the SM as one contiguous,
// 16-byte-vectorised stream (the tile's output rows are contiguous in HBM).
#include <string>

#include "tpb_bulk.cuh"
#include "tpb_device.cuh"
#include "../../include/tpb.h"

namespace tpb {
namespace {

constexpr int kTile = 128;  // parameters per CTA == threads per CTA

struct EvalArgs {
  int num_points, dim, num_splines;
  long long evals;
  const double* knots;
  const double* ctrl;
  const double* u;
  const double* u0;
  const double* du;
  double* out;
  int32_t* status;
  bool exact;
  bool bulk;  // polynomial form: tiles leave shared memory as bulk async copies
};

template <int P, int NDER>
// 8 CTAs/SM (64 registers): degree 5 gains 11 % (6.87 -> 6.09 ms per 1e8 evaluations) despite
// small spills, degrees 3 and 4 are unchanged (division-bound either way)
#ifndef TPB_SE_MINBLOCKS
#define TPB_SE_MINBLOCKS 8
#endif
__global__ void __launch_bounds__(kTile, TPB_SE_MINBLOCKS) k_spline_eval(EvalArgs g) {
  extern __shared__ __align__(16) double smem[];
  const int nk = g.num_points + P + 1;
  const int D = g.dim;
  constexpr int K = NDER + 1;
  double* s_knots = smem;                 // [nk]
  double* s_ctrl = s_knots + nk;          // [num_points][D]
  // 16-byte aligned so the tile can be streamed out as double2
  double* s_out = smem + ((nk + g.num_points * D + 1) & ~1);  // [kTile][K*D]
  const int s = blockIdx.y;
  {
    const double* gk = g.knots + static_cast<size_t>(s) * nk;
    const double* gc = g.ctrl + static_cast<size_t>(s) * g.num_points * D;
    for (int i = threadIdx.x; i < nk; i += kTile) s_knots[i] = gk[i];
    for (int i = threadIdx.x; i < g.num_points * D; i += kTile) s_ctrl[i] = gc[i];
  }
  __syncthreads();
  const long long e0 = static_cast<long long>(blockIdx.x) * kTile;
  const long long e = e0 + threadIdx.x;
  const int rowlen = K * D;
  bool ok = false;
  if (e < g.evals) {
    const size_t flat = static_cast<size_t>(s) * g.evals + e;
    const double u = g.u ? g.u[flat] : (g.u0[s] + static_cast<double>(e) * g.du[s]);
    // range check of EvalCurveAndDerivatives (bspline.h:549-552); NaN fails too
    ok = (u >= s_knots[0]) && (u <= s_knots[nk - 1]);
    if (g.status) g.status[flat] = ok ? TPB_OK : TPB_OUT_OF_RANGE;
    if (ok) {
      const int span = knot_span(s_knots, nk, P, u);
      double ders[K][P + 1];
      basis_ders<P, NDER>(s_knots, span, u, ders);
      const double* P0 = s_ctrl + (span - P) * D;
      double* o = s_out + threadIdx.x * rowlen;
      for (int d = 0; d < D; ++d) {
        double acc[K];
#pragma unroll
        for (int k = 0; k < K; ++k) acc[k] = 0.0;
#pragma unroll
        for (int j = 0; j <= P; ++j) {
          const double Pj = P0[j * D + d];
#pragma unroll
          for (int k = 0; k < K; ++k) acc[k] += ders[k][j] * Pj;
        }
#pragma unroll
        for (int k = 0; k < K; ++k) o[k * D + d] = acc[k];
      }
    }
  }
  // Rows that failed the range check must keep the caller's memory: if the
  // whole tile is in range (the common case) stream it out contiguously,
  // otherwise fall back to per-row stores.
  const int all_ok = __syncthreads_and(ok || e >= g.evals);
  const long long n_in_tile = min(static_cast<long long>(kTile), g.evals - e0);
  double* gout = g.out + (static_cast<size_t>(s) * g.evals + e0) * rowlen;
  if (all_ok) {
    const int total = static_cast<int>(n_in_tile) * rowlen;
    if ((reinterpret_cast<uintptr_t>(gout) & 15) == 0 && (total & 1) == 0) {
      const double2* s2 = reinterpret_cast<const double2*>(s_out);
      double2* g2 = reinterpret_cast<double2*>(gout);
      for (int i = threadIdx.x; i < total / 2; i += kTile) g2[i] = s2[i];
    } else {
      for (int i = threadIdx.x; i < total; i += kTile) gout[i] = s_out[i];
    }
  } else if (ok) {
    const double* o = s_out + threadIdx.x * rowlen;
    for (int i = 0; i < rowlen; ++i) gout[threadIdx.x * rowlen + i] = o[i];
  }
}

// ---------------------------------------------------------------------------
// exact == 0: per-span polynomial form.
//
// On one knot span the curve is a polynomial of degree P.  The CTA first builds,
// for every span of its spline, the Taylor coefficients about the span's
// midpoint, c_k = C^(k)(mid) / k!, with the same A2.3 routine (NDER = P), and
// keeps them in shared memory: [span][k][dim].  Every parameter then costs a
// span search plus 3 fused multiply-adds per coefficient and dimension
// (Horner with derivatives) and no division, which makes the kernel a pure
// HBM write stream.  |u - mid| <= h/2 keeps the power basis well conditioned;
// results agree with the exact form to <= 1e-12 (tests/test_gpu_spline_eval.py).
// A CTA evaluates kPolyTiles tiles of kTile parameters so that the table build
// is amortised.
// ---------------------------------------------------------------------------
constexpr int kPolyTiles = 16;

//
// kBulk: the finished tile (kTile x K x dim doubles, contiguous in HBM) leaves shared
// memory as ONE bulk asynchronous copy (cp.async.bulk.global.shared::cta, the 1-D TMA
// path) issued by thread 0 instead of a per-thread 16-byte copy loop; the tile is
// double-buffered so that the copy of tile t overlaps the evaluation of tile t + 1.
// Tiles whose HBM address or size is not a multiple of 16 bytes, and tiles with
// out-of-range parameters, take the per-thread path.
template <int P, int NDER, bool kBulk>
__global__ void __launch_bounds__(kTile) k_spline_eval_poly(EvalArgs g) {
  extern __shared__ __align__(16) double smem[];
  const int n = g.num_points;
  const int nk = n + P + 1;
  const int D = g.dim;
  const int nspan = n - P;
  constexpr int K = NDER + 1;
  double* s_knots = smem;                          // [nk]
  double* s_mid = s_knots + nk;                    // [nspan]
  double* s_coef = s_mid + nspan;                  // [nspan][P+1][D]
  double* s_ctrl = s_coef + nspan * (P + 1) * D;   // [n][D] (table build only)
  // [kTile][K*D], two of them with kBulk; 16-byte aligned
  double* s_out0 = smem + ((nk + nspan + nspan * (P + 1) * D + n * D + 1) & ~1);
  const int tile_doubles = (kTile * K * D + 1) & ~1;
  const int s = blockIdx.y;
  {
    const double* gk = g.knots + static_cast<size_t>(s) * nk;
    const double* gc = g.ctrl + static_cast<size_t>(s) * n * D;
    for (int i = threadIdx.x; i < nk; i += kTile) s_knots[i] = gk[i];
    for (int i = threadIdx.x; i < n * D; i += kTile) s_ctrl[i] = gc[i];
  }
  __syncthreads();
  for (int t = threadIdx.x; t < nspan; t += kTile) {
    const int sp = P + t;
    const double lo = s_knots[sp], hi = s_knots[sp + 1];
    const double mid = 0.5 * (lo + hi);
    s_mid[t] = mid;
    if (hi > lo) {  // empty spans are never selected by KnotSpan
      double ders[P + 1][P + 1];
      basis_ders<P, P>(s_knots, sp, mid, ders);
      double fact = 1.0;
#pragma unroll
      for (int k = 0; k <= P; ++k) {
        if (k > 1) fact *= static_cast<double>(k);
        for (int d = 0; d < D; ++d) {
          double acc = 0.0;
#pragma unroll
          for (int j = 0; j <= P; ++j) acc = fma(ders[k][j], s_ctrl[(sp - P + j) * D + d], acc);
          s_coef[(t * (P + 1) + k) * D + d] = acc / fact;
        }
      }
    }
  }
  __syncthreads();
  const int rowlen = K * D;
  const long long chunk0 = static_cast<long long>(blockIdx.x) * kTile * kPolyTiles;
  for (int tile = 0; tile < kPolyTiles; ++tile) {
    const long long e0 = chunk0 + static_cast<long long>(tile) * kTile;
    if (e0 >= g.evals) break;
    const long long e = e0 + threadIdx.x;
    double* s_out = s_out0 + (kBulk ? (tile & 1) * tile_doubles : 0);
    if (kBulk && tile >= 2) {
      // the bulk copy that read this buffer two tiles ago must be done reading it
      if (threadIdx.x == 0) bulk::bulk_wait_read<1>();
      __syncthreads();
    }
    bool ok = false;
    if (e < g.evals) {
      const size_t flat = static_cast<size_t>(s) * g.evals + e;
      const double u = g.u ? g.u[flat] : (g.u0[s] + static_cast<double>(e) * g.du[s]);
      ok = (u >= s_knots[0]) && (u <= s_knots[nk - 1]);
      if (g.status) g.status[flat] = ok ? TPB_OK : TPB_OUT_OF_RANGE;
      if (ok) {
        const int t = knot_span(s_knots, nk, P, u) - P;
        const double x = u - s_mid[t];
        const double* c = s_coef + t * (P + 1) * D;
        double* o = s_out + threadIdx.x * rowlen;
        for (int d = 0; d < D; ++d) {
          double p0 = c[P * D + d], p1 = 0.0, p2 = 0.0, p3 = 0.0;
#pragma unroll
          for (int k = P - 1; k >= 0; --k) {
            if (NDER >= 3) p3 = fma(p3, x, p2);
            if (NDER >= 2) p2 = fma(p2, x, p1);
            if (NDER >= 1) p1 = fma(p1, x, p0);
            p0 = fma(p0, x, c[k * D + d]);
          }
          o[d] = p0;
          if (NDER >= 1) o[D + d] = p1;
          if (NDER >= 2) o[2 * D + d] = 2.0 * p2;
          if (NDER >= 3) o[3 * D + d] = 6.0 * p3;
        }
      }
    }
    // (with kBulk: the tile's ordinary shared-memory stores become visible to the
    // async proxy before the barrier that precedes the bulk copy)
    if (kBulk) bulk::fence_async_shared();
    const int all_ok = __syncthreads_and(ok || e >= g.evals);
    const long long n_in_tile = min(static_cast<long long>(kTile), g.evals - e0);
    double* gout = g.out + (static_cast<size_t>(s) * g.evals + e0) * rowlen;
    if (all_ok) {
      const int total = static_cast<int>(n_in_tile) * rowlen;
      const bool vec = (reinterpret_cast<uintptr_t>(gout) & 15) == 0 && (total & 1) == 0;
      if (kBulk && vec) {
        if (threadIdx.x == 0) {
          bulk::bulk_store(gout, s_out, static_cast<uint32_t>(total) * sizeof(double));
          bulk::bulk_commit();
        }
      } else if (vec) {
        const double2* s2 = reinterpret_cast<const double2*>(s_out);
        double2* g2 = reinterpret_cast<double2*>(gout);
        for (int i = threadIdx.x; i < total / 2; i += kTile) g2[i] = s2[i];
      } else {
        for (int i = threadIdx.x; i < total; i += kTile) gout[i] = s_out[i];
      }
    } else if (ok) {
      const double* o = s_out + threadIdx.x * rowlen;
      for (int i = 0; i < rowlen; ++i) gout[threadIdx.x * rowlen + i] = o[i];
    }
    if (!kBulk) __syncthreads();  // s_out is reused by the next tile
  }
  // shared memory must stay valid until every bulk copy has read it
  if (kBulk && threadIdx.x == 0) bulk::bulk_wait_read<0>();
}

template <int P, int NDER, bool kBulk>
void launch_one_poly_impl(const EvalArgs& g, cudaStream_t st) {
  const int n = g.num_points, nk = n + P + 1, nspan = n - P;
  const size_t tile = (static_cast<size_t>(kTile) * (NDER + 1) * g.dim + 1) & ~size_t{1};
  const size_t smem = sizeof(double) * (nk + nspan + static_cast<size_t>(nspan) * (P + 1) * g.dim +
                                        static_cast<size_t>(n) * g.dim + 1 +
                                        (kBulk ? 2 : 1) * tile);
  cudaFuncSetAttribute(k_spline_eval_poly<P, NDER, kBulk>,
                       cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  const long long per_cta = static_cast<long long>(kTile) * kPolyTiles;
  dim3 grid(static_cast<unsigned>((g.evals + per_cta - 1) / per_cta), g.num_splines);
  k_spline_eval_poly<P, NDER, kBulk><<<grid, kTile, smem, st>>>(g);
}

template <int P, int NDER>
void launch_one_poly(const EvalArgs& g, cudaStream_t st) {
  if (g.bulk) {
    launch_one_poly_impl<P, NDER, true>(g, st);
  } else {
    launch_one_poly_impl<P, NDER, false>(g, st);
  }
}

template <int P, int NDER>
void launch_one(const EvalArgs& g, cudaStream_t st) {
  const int nk = g.num_points + P + 1;
  const size_t smem = sizeof(double) * (nk + static_cast<size_t>(g.num_points) * g.dim + 1 +
                                        static_cast<size_t>(kTile) * (NDER + 1) * g.dim);
  cudaFuncSetAttribute(k_spline_eval<P, NDER>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       static_cast<int>(smem));
  dim3 grid(static_cast<unsigned>((g.evals + kTile - 1) / kTile), g.num_splines);
  k_spline_eval<P, NDER><<<grid, kTile, smem, st>>>(g);
}

template <int P, int NDER>
void launch_either(const EvalArgs& g, cudaStream_t st) {
  if (g.exact) {
    launch_one<P, NDER>(g, st);
  } else {
    launch_one_poly<P, NDER>(g, st);
  }
}

template <int P>
bool dispatch_nder(int nder, const EvalArgs& g, cudaStream_t st) {
  if (nder == 0) {
    launch_either<P, 0>(g, st);
    return true;
  }
  if constexpr (P >= 1) {
    if (nder == 1) {
      launch_either<P, 1>(g, st);
      return true;
    }
  }
  if constexpr (P >= 2) {
    if (nder == 2) {
      launch_either<P, 2>(g, st);
      return true;
    }
  }
  if constexpr (P >= 3) {
    if (nder == 3) {
      launch_either<P, 3>(g, st);
      return true;
    }
  }
  return false;
}

}  // namespace

int launch_spline_eval(const tpb_spline_eval_desc& d, cudaStream_t st, std::string* err) {
  if (d.degree < 1 || d.degree > kMaxDegree) {
    *err = "degree must be in [1, 5]";
    return TPB_OUT_OF_RANGE;
  }
  if (d.nder < 0 || d.nder > d.degree || d.nder > 3) {
    // reference: CHECK(der <= p), bspline_base.cc:270
    *err = "nder must be <= degree (and <= 3)";
    return TPB_INVALID_ARGUMENT;
  }
  if (d.num_points < d.degree + 1 || d.dim < 1 || d.num_splines < 1 || d.evals_per_spline < 0) {
    *err = "bad spline dimensions";
    return TPB_INVALID_ARGUMENT;
  }
  if (!d.knots || !d.ctrl || !d.out || (!d.u && (!d.u0 || !d.du))) {
    *err = "null pointer";
    return TPB_INVALID_ARGUMENT;
  }
  if (d.num_splines > 65535) {
    *err = "num_splines > 65535";
    return TPB_OUT_OF_RANGE;
  }
  if (d.evals_per_spline == 0) return TPB_OK;
  size_t smem = sizeof(double) * (d.num_points + d.degree + 2 +
                                  static_cast<size_t>(d.num_points) * d.dim +
                                  static_cast<size_t>(kTile) * (d.nder + 1) * d.dim);
  // exact > 0: A2.3; exact == 0: polynomial form, tiles leave as bulk copies;
  // exact < 0: polynomial form with per-thread stores (kept for A/B measurements)
  bool bulk = d.exact == 0;
  if (d.exact <= 0) {  // + span midpoints and the per-span coefficient table
    const size_t nspan = static_cast<size_t>(d.num_points - d.degree);
    smem += sizeof(double) * (nspan + nspan * (d.degree + 1) * d.dim);
    // the second tile buffer of the bulk variant; fall back if it does not fit
    const size_t tile = sizeof(double) * (static_cast<size_t>(kTile) * (d.nder + 1) * d.dim + 2);
    if (bulk && smem + tile <= 200 * 1024) {
      smem += tile;
    } else {
      bulk = false;
    }
  }
  if (smem > 200 * 1024) {
    *err = "spline tables do not fit in shared memory";
    return TPB_OUT_OF_RANGE;
  }
  EvalArgs g{d.num_points, d.dim,  d.num_splines, static_cast<long long>(d.evals_per_spline),
             d.knots,      d.ctrl, d.u,           d.u0,
             d.du,         d.out,  d.status,      d.exact > 0,
             bulk};
  bool ok = false;
  switch (d.degree) {
    case 1: ok = dispatch_nder<1>(d.nder, g, st); break;
    case 2: ok = dispatch_nder<2>(d.nder, g, st); break;
    case 3: ok = dispatch_nder<3>(d.nder, g, st); break;
    case 4: ok = dispatch_nder<4>(d.nder, g, st); break;
    case 5: ok = dispatch_nder<5>(d.nder, g, st); break;
  }
  if (!ok) {
    *err = "unsupported (degree, nder)";
    return TPB_UNIMPLEMENTED;
  }
  const cudaError_t ce = cudaGetLastError();
  if (ce != cudaSuccess) {
    *err = cudaGetErrorString(ce);
    return TPB_INTERNAL;
  }
  return TPB_OK;
}

}  // namespace tpb
