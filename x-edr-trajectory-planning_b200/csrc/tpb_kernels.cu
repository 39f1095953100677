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
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "tpb_bulk.cuh"
#include "tpb_engine.h"
#include "tpb_fit.cuh"
#include "tpb_integrate.cuh"
#include "tpb_integrate_pair.cuh"
#include "tpb_joint_limit.cuh"
#include "tpb_nvtx.h"
#include "tpb_row_limit.cuh"
#include "tpb_solver.cuh"

namespace tpb {
namespace {

// ===========================================================================
// K2: joint_fit
// ===========================================================================

// One warp per path, lane i owns corner i: the two control points around a corner depend on
// the neighbouring corners only (spline_utils.cc:47-102), the polygon length is summed in the
// reference's order by broadcasting the segment lengths one after the other, the knots are
// the running sum of MakeUniformKnotVector (bspline_base.cc:350-381) evaluated per lane.
// (One thread per path took 0.1 ms for 4,096 paths: 64 CTAs walking global memory.)
constexpr int kFitThreads = 128;

__global__ void __launch_bounds__(kFitThreads)
k_joint_fit(Dims dm, DeviceArrays a, double rounding, bool auto_delta) {
  const int b = (blockIdx.x * kFitThreads + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (b >= dm.B) return;  // uniform per warp
  const int D = dm.D;
  // counts that arrive in device memory cannot be range-checked by the host: clamped, so
  // that a bad one stays inside this path's slabs
  const int W = min(max(a.num_wp[b], 1), dm.Wmax);
  const double* wp = a.waypoints + static_cast<size_t>(b) * dm.Wmax * D;
  double* c = a.ctrl + static_cast<size_t>(b) * dm.max_points * D;
  double* kn = a.knots + static_cast<size_t>(b) * (dm.max_points + 3);

  int n = 4;
  if (W == 1) {
    for (int i = lane; i < 4 * D; i += 32) c[i] = wp[i % D];
  } else {
    n = 3 * W - 2;
    for (int i = lane; i < W; i += 32) {
      const double* p = wp + i * D;
      double* ck = c + 3 * i * D;
      double delta[kMaxDofs], off[kMaxDofs];
      for (int d = 0; d < D; ++d) ck[d] = p[d];
      if (i < W - 1) {  // the point behind corner i (the first corner included, :90-94)
        for (int d = 0; d < D; ++d) delta[d] = wp[(i + 1) * D + d] - p[d];
        corner_offset(delta, D, rounding, off);
        for (int d = 0; d < D; ++d) ck[D + d] = p[d] + off[d];
      }
      if (i > 0) {  // the point before corner i (the last corner included, :96-100)
        for (int d = 0; d < D; ++d) delta[d] = wp[(i - 1) * D + d] - p[d];
        corner_offset(delta, D, rounding, off);
        for (int d = 0; d < D; ++d) ck[d - D] = p[d] + off[d];
      }
    }
  }
  __syncwarp();
  // knots scaled by the control polygon length (timeable_path_joint_spline.cc:271-285)
  double length = 0.0;
  for (int base = 0; base < n - 1; base += 32) {
    const int j = base + lane;
    double len = 0.0;
    if (j < n - 1) {
      double s = 0.0;
      for (int d = 0; d < D; ++d) {
        const double df = c[(j + 1) * D + d] - c[j * D + d];
        s += df * df;
      }
      len = sqrt(s);
    }
    const int cnt = min(32, n - 1 - base);
    for (int k = 0; k < cnt; ++k) length += __shfl_sync(kFullMask, len, k);
  }
  constexpr double kMinimumFinalKnotValue = 0.1;
  constexpr double kPathParameterPerPolygonLength = 1.0;
  const double lw = length * kPathParameterPerPolygonLength;
  const double weighted = (lw < kMinimumFinalKnotValue) ? kMinimumFinalKnotValue : lw;
  // MakeUniformKnotVector on [0, 1] for degree 2: three zeros, a running sum of the spacing,
  // three ones (uniform_knots_degree2)
  const int nk = n + 3;
  const double spacing = (1.0 / (nk - 2.0 * 3.0 + 1.0)) * (1.0 - 0.0);
  for (int i = lane; i < nk; i += 32) {
    double v = 0.0;
    if (i >= nk - 3) {
      v = 1.0;
    } else {
      for (int k = 3; k <= i; ++k) v = v + spacing;
    }
    kn[i] = v * weighted;
  }
  if (lane == 0) {
    a.num_points[b] = n;
    // knot capacity the reference gives the spline (timeable_path_joint_spline.cc:257-268)
    a.knot_cap[b] = max(2 * nk, 100);
    const double kend = 1.0 * weighted;
    a.knot_end[b] = kend;
    a.delta[b] = auto_delta ? kend / (dm.N - 1) : a.delta_in[b];
  }
}

// ===========================================================================
// K3: joint sample + per-sample limit curve
// ===========================================================================

constexpr int kSampleThreads = 128;
// CTA size of k_joint_sample_limit<D>: the staging of the path's tables and the limits is
// per CTA, so larger CTAs do it less often.  Measured (cfg2, ms per launch, same 16 warps /
// SM): 64 threads x 8 CTAs/SM 1.321, 128 x 4 1.297, 192 x 2 (138 registers) 1.470,
// **256 x 2 1.228**, 512 x 1 3.864.
#ifndef TPB_JOINT_SAMPLE_THREADS
#define TPB_JOINT_SAMPLE_THREADS 256
#endif
constexpr int kJointSampleThreads = TPB_JOINT_SAMPLE_THREADS;

// Any D <= kMaxDofs: rows in thread-local arrays, generic LP (used for D > 8).
__global__ void __launch_bounds__(kSampleThreads)
k_joint_sample_limit_any(Dims dm, DeviceArrays a, double safety) {
  extern __shared__ double smem[];
  const int b = path_of_block();
  if (b >= dm.B) return;  // uniform per CTA (grid folded over y and z)
  if (a.active != nullptr && a.active[b] == 0) return;  // uniform per CTA
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

// D = kD <= 8: one sample per thread, rows and the whole active-set LP in
// registers (tpb_joint_limit.cuh); q, q', q'' leave through shared memory so that
// each CTA writes its 128 consecutive [sample][D] rows as contiguous runs.
// 4 CTAs/SM (<= 128 registers: 124 used, no spills) measured best: 1.42 ms against 1.54 ms
// unconstrained (140 registers, 3 CTAs/SM) and 1.46 ms at 5 CTAs/SM (96 registers, spills)
// tiles of 256 samples a CTA works through with one staging of the path's tables.  Measured
// (cfg2, ms per launch): 1 tile 1.202, 2 tiles 1.293, 4 tiles 1.406 (128 registers with
// spills, and half / a quarter of the CTAs to balance the SMs with).
#ifndef TPB_SL_TILES
#define TPB_SL_TILES 1
#endif
#ifndef TPB_SL_MINBLOCKS
#define TPB_SL_MINBLOCKS (512 / TPB_JOINT_SAMPLE_THREADS)
#endif
//
// Staging: the path's knot vector and control points are contiguous runs in HBM and come
// in as two bulk asynchronous copies (cp.async.bulk.shared::cluster.global, completion
// counted on an mbarrier), issued by thread 0 while the other threads compute the scaled
// limits; a run that does not start on a 16-byte boundary is fetched from the boundary
// before it (the extra 8 bytes belong to the neighbouring path's row) and addressed with an
// offset.  The three [256][kD] output tiles leave as bulk copies when their HBM address is
// 16-byte aligned (always for even N).  TPB_SL_BULK_IN / TPB_SL_BULK_OUT = 0 build the
// per-thread loops (A/B; profiles/r2_summary.md).
#ifndef TPB_SL_BULK_IN
#define TPB_SL_BULK_IN 1
#endif
#ifndef TPB_SL_BULK_OUT
#define TPB_SL_BULK_OUT 1
#endif
// shared-memory doubles of k_joint_sample_limit: [knots + 2][ctrl + 2][aS][vS2][pad][3 tiles]
__host__ __device__ inline int sample_limit_table_doubles(int max_points, int D) {
  return ((max_points + 3 + 2 + 1) & ~1) + ((max_points * D + 2 + 1) & ~1) + ((2 * D + 1) & ~1);
}

template <int kD>
__global__ void __launch_bounds__(kJointSampleThreads, TPB_SL_MINBLOCKS)
k_joint_sample_limit(Dims dm, DeviceArrays a, double safety) {
  extern __shared__ __align__(16) double smem[];
  __shared__ __align__(8) uint64_t s_bar;
  const int b = path_of_block();
  if (b >= dm.B) return;  // uniform per CTA (grid folded over y and z)
  if (a.active != nullptr && a.active[b] == 0) return;  // uniform per CTA
  const int n = a.num_points[b];
  const int nk = n + 3;
  double* s_kbuf = smem;                                        // [max_points + 3 + 2]
  double* s_cbuf = s_kbuf + ((dm.max_points + 3 + 2 + 1) & ~1);  // [max_points * kD + 2]
  double* s_aS = s_cbuf + ((dm.max_points * kD + 2 + 1) & ~1);   // [kD]  amax * safety
  double* s_vS2 = s_aS + kD;                                    // [kD]  (vmax * safety)^2
  double* s_out = smem + sample_limit_table_doubles(dm.max_points, kD);  // [3][128 * kD]
  const double* gk = a.knots + static_cast<size_t>(b) * (dm.max_points + 3);
  const double* gc = a.ctrl + static_cast<size_t>(b) * dm.max_points * kD;
#if TPB_SL_BULK_IN
  const int kmis = static_cast<int>((reinterpret_cast<uintptr_t>(gk) >> 3) & 1);
  const int cmis = static_cast<int>((reinterpret_cast<uintptr_t>(gc) >> 3) & 1);
  const double* s_knots = s_kbuf + kmis;
  const double* s_ctrl = s_cbuf + cmis;
  if (threadIdx.x == 0) {
    bulk::mbar_init(&s_bar, 1);
    bulk::fence_mbar_init();
    const uint32_t kbytes = static_cast<uint32_t>((nk + kmis + 1) & ~1) * 8u;
    const uint32_t cbytes = static_cast<uint32_t>((n * kD + cmis + 1) & ~1) * 8u;
    bulk::mbar_expect_tx(&s_bar, kbytes + cbytes);
    bulk::bulk_load(s_kbuf, gk - kmis, kbytes, &s_bar);
    bulk::bulk_load(s_cbuf, gc - cmis, cbytes, &s_bar);
  }
#else
  double* s_knots = s_kbuf;
  double* s_ctrl = s_cbuf;
  for (int i = threadIdx.x; i < nk; i += blockDim.x) s_knots[i] = gk[i];
  for (int i = threadIdx.x; i < n * kD; i += blockDim.x) s_ctrl[i] = gc[i];
#endif
  if (threadIdx.x < kD) {
    const double am = a.amax[static_cast<size_t>(b) * kD + threadIdx.x];
    const double vm = a.vmax[static_cast<size_t>(b) * kD + threadIdx.x];
    s_aS[threadIdx.x] = am * safety;
    const double vs = vm * safety;
    s_vS2[threadIdx.x] = vs * vs;
  }
  __syncthreads();  // the barrier is initialised, the limits are written
#if TPB_SL_BULK_IN
  bulk::mbar_wait(&s_bar, 0);  // the tables have landed
#endif
  // From here on the warps are independent: warp w of tile t owns samples
  // [tile * 256 + 32 w, + 32), its own three [32][kD] output tiles in shared memory, and
  // sends them off itself (three bulk copies issued by lane 0) -- no CTA barrier, a warp
  // whose samples need more simplex iterations holds nobody up.
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double* wq = s_out + warp * (3 * 32 * kD);  // this warp's q / q' / q'' tiles
  double* wdq = wq + 32 * kD;
  double* wddq = wdq + 32 * kD;
  JointLimit<kD> J;
  J.aS = s_aS;
  J.vS2 = s_vS2;
  const double delta = a.delta[b];
  const double path_start = a.path_start[b];
  const double k0 = s_knots[0], k1 = s_knots[nk - 1];
#if TPB_SL_BULK_OUT
  bool pending = false;  // lane 0: bulk copies of the previous tile may still read the tiles
#endif
  for (int t = 0; t < TPB_SL_TILES; ++t) {
    const int idx0 = (blockIdx.x * TPB_SL_TILES + t) * kJointSampleThreads + warp * 32;
    if (idx0 >= dm.N) break;  // uniform per warp
    const int idx = idx0 + lane;
    const bool live = idx < dm.N;
#if TPB_SL_BULK_OUT
    if (TPB_SL_TILES > 1) {
      if (pending) bulk::bulk_wait_read<0>();
      __syncwarp();
    }
#endif
    if (live) {
      // SamplePath, timeable_path_joint_spline.cc:294-318.
      const double parameter = path_start + idx * delta;
      double* o0 = wq + lane * kD;
      double* o1 = wdq + lane * kD;
      double* o2 = wddq + lane * kD;
      if (parameter < k1 + delta) {
        const double u = (parameter < k0) ? k0 : ((k1 < parameter) ? k1 : parameter);
        const int span = knot_span(s_knots, nk, 2, u);
        double ders[3][3];
        basis_ders<2, 2>(s_knots, span, u, ders);
        // EvalCurveAndDerivatives, splines/bspline.h:561-566.
        const double* P0 = s_ctrl + (span - 2) * kD;
#pragma unroll
        for (int d = 0; d < kD; ++d) {
          double v0 = 0.0, v1 = 0.0, v2 = 0.0;
#pragma unroll
          for (int j = 0; j <= 2; ++j) {
            const double Pj = P0[j * kD + d];
            v0 += ders[0][j] * Pj;
            v1 += ders[1][j] * Pj;
            v2 += ders[2][j] * Pj;
          }
          o0[d] = v0;
          o1[d] = v1;
          o2[d] = v2;
          J.A[d] = v1;
          J.B[d] = v2;
        }
      } else {
#pragma unroll
        for (int d = 0; d < kD; ++d) {
          o0[d] = s_ctrl[(n - 1) * kD + d];
          o1[d] = 0.0;
          o2[d] = 0.0;
          J.A[d] = 0.0;
          J.B[d] = 0.0;
        }
      }
      // ConstraintSetup (:320-343) + CalculateBoundary's per-sample part
      // (time_optimal_path_timing.cc:1366-1377); rows never leave the registers.
      J.prepare();
      double sd2max, sddmax, sd2zero, smax, smin;
      J.lp(sd2max, sddmax, sd2zero);
      J.sdd_min_max(sd2max, smax, smin);
      const size_t o = static_cast<size_t>(b) * dm.N + idx;
      a.b_raw_sd2_max[o] = sd2max;
      a.b_sd2_zero[o] = sd2zero;
      a.b_raw_sdd_max[o] = smax;
      a.b_raw_sdd_min[o] = smin;
      a.b_at0[o] = fabs(sd2max - sd2zero) < kTiny;
    }
    // copy-out of this warp's rows [idx0, idx0 + cnt): three contiguous runs
    const int cnt = min(32, dm.N - idx0);
    const size_t g0 = (static_cast<size_t>(b) * dm.N + idx0) * kD;
#if TPB_SL_BULK_OUT
    const bool vec = ((cnt * kD) & 1) == 0 && bulk::aligned16(a.q + g0) &&
                     bulk::aligned16(a.dq + g0) && bulk::aligned16(a.ddq + g0);
    if (vec) {
      bulk::fence_async_shared();
      __syncwarp();
      if (lane == 0) {
        const uint32_t bytes = static_cast<uint32_t>(cnt * kD) * 8u;
        bulk::bulk_store(a.q + g0, wq, bytes);
        bulk::bulk_store(a.dq + g0, wdq, bytes);
        bulk::bulk_store(a.ddq + g0, wddq, bytes);
        bulk::bulk_commit();
        pending = true;
      }
      continue;
    }
#endif
    __syncwarp();
    for (int k = lane; k < cnt * kD; k += 32) {
      a.q[g0 + k] = wq[k];
      a.dq[g0 + k] = wdq[k];
      a.ddq[g0 + k] = wddq[k];
    }
    __syncwarp();
  }
#if TPB_SL_BULK_OUT
  // shared memory stays valid until the copies have read it
  if (pending) bulk::bulk_wait_read<0>();
#endif
}

__global__ void __launch_bounds__(kSampleThreads)
k_generic_limit(Dims dm, DeviceArrays a) {
  const int b = path_of_block();
  if (b >= dm.B) return;  // uniform per CTA (grid folded over y and z)
  if (a.active != nullptr && a.active[b] == 0) return;
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

// The same for C <= kC <= 12 rows with everything in registers (tpb_row_limit.cuh);
// rows are read as 16-byte halves (a row is 32 bytes, 32-byte aligned).
template <int kC>
__global__ void __launch_bounds__(kSampleThreads)
k_generic_limit_reg(Dims dm, DeviceArrays a) {
  const int b = path_of_block();
  if (b >= dm.B) return;  // uniform per CTA (grid folded over y and z)
  if (a.active != nullptr && a.active[b] == 0) return;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= dm.N) return;
  const int C = dm.C;
  RowLimit<kC> R;
  const double2* __restrict__ c =
      reinterpret_cast<const double2*>(a.cons + (static_cast<size_t>(b) * dm.N + idx) * C * 4);
#pragma unroll
  for (int r = 0; r < kC; ++r) {
    double2 ab = make_double2(0.0, 0.0), lh = make_double2(-1.0, 1.0);  // padding row
    if (r < C) {
      ab = c[2 * r];
      lh = c[2 * r + 1];
    }
    R.A[r] = ab.x;
    R.B[r] = ab.y;
    R.LO[r] = lh.x;
    R.HI[r] = lh.y;
  }
  double sd2max, sddmax, sd2zero, smax, smin;
  R.lp(C, sd2max, sddmax, sd2zero);
  R.sdd_min_max(sd2max, smax, smin);
  const size_t o = static_cast<size_t>(b) * dm.N + idx;
  a.b_raw_sd2_max[o] = sd2max;
  a.b_sd2_zero[o] = sd2zero;
  a.b_raw_sdd_max[o] = smax;
  a.b_raw_sdd_min[o] = smin;
  a.b_at0[o] = fabs(sd2max - sd2zero) < kTiny;
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

// ResampleSkippingSamplesCloserThanTimeStep (:785-836), the decisions: one warp per
// path.  Row 0 is interpolated at start_sec; after it a path sample is kept iff its
// time is at least GetMinTimeDeltaToKeep() = 0.95 * time step (:893-900) after the
// last kept one.  Lanes test 32 samples against the last kept time; the first lane
// that passes is the next kept sample, whatever the others say (they are re-tested
// against the new time) -- the reference's sequential rule.  keep[b][k] receives
// the sample index of row k + 1, r_count[b] the number of rows.
__global__ void k_resample_skip_count(Dims dm, DeviceArrays a, int32_t* keep, double start_sec,
                                      double dt) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int b = warp;
  if (b >= dm.B) return;
  const int N = dm.N;
  if (a.status[b] != TPB_PATH_OK) {
    if (lane == 0) a.r_count[b] = 0;
    return;
  }
  const double* tp = a.time + static_cast<size_t>(b) * N;
  int32_t* kp = keep + static_cast<size_t>(b) * N;
  // TimeAtPathSamplesLowerIndex(0, start_sec) (:686-695)
  int li = N - 1;
  for (int base = 0; base < N - 1; base += 32) {
    const int i = base + lane;
    const unsigned m = __ballot_sync(kFullMask, (i < N - 1) && (tp[i + 1] > start_sec));
    if (m) {
      li = base + (__ffs(m) - 1);
      break;
    }
  }
  const double min_delta = 0.95 * dt;
  double last = start_sec;
  int rows = 1;
  for (int base = li + 1; base < N; base += 32) {
    const int i = base + lane;
    const double t = (i < N) ? tp[i] : 0.0;
    unsigned cand = __ballot_sync(kFullMask, (i < N) && !(fabs(t - last) < min_delta));
    while (cand) {
      const int k = __ffs(cand) - 1;
      if (lane == 0) kp[rows - 1] = base + k;
      ++rows;
      last = __shfl_sync(kFullMask, t, k);
      cand = __ballot_sync(kFullMask, (i < N) && (lane > k) && !(fabs(t - last) < min_delta));
    }
  }
  if (lane == 0) a.r_count[b] = rows;
}

// exclusive scan of r_count into r_offset (single block; B is at most ~1e6).
// continued != 0: this batch is a chunk that follows another one, offset[0]
// already holds the rows before it (the previous chunk's offset[B]).
__global__ void k_scan_offsets(int B, const int32_t* __restrict__ count,
                               int64_t* __restrict__ offset, int continued) {
  __shared__ long long s_part[1024];
  const int tid = threadIdx.x;
  const int per = (B + blockDim.x - 1) / blockDim.x;
  const int lo = min(B, tid * per), hi = min(B, lo + per);
  long long sum = 0;
  for (int i = lo; i < hi; ++i) sum += count[i];
  s_part[tid] = sum;
  __syncthreads();
  if (tid == 0) {
    long long run = continued ? offset[0] : 0;
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

// 7 warps: 53.6 KB of shared memory per CTA at N = 2048, D = 7.  Measured: 1.11 ms against 1.15 ms
// with 8 warps (59 KB); forcing a fourth CTA per SM (70 registers) takes 1.41 ms -- the gathers
// live on the L1 that the shared-memory carve-out leaves.
constexpr int kResampleThreads = 224;

// One CTA per path (grid-stride over paths): time[] of the path is staged in
// shared memory (the per-row lower-index search runs there), each thread owns
// one output row per pass, and every warp transposes its 32 rows x D values
// through shared memory so that the packed [row][D] arrays are written as
// contiguous 256-byte runs.  Rows past r_capacity are dropped (the host reports
// kOutOfRange).  kSmemTime = false: N too large for shared memory, search in
// global memory.
//
// kSkip: rows follow ResampleSkippingSamplesCloserThanTimeStep instead of the uniform
// grid: row 0 is interpolated at start_sec, row i > 0 IS path sample keep[b][i-1]
// (li = ui = that sample, values taken as they are), so that both methods share the
// assembly and the transposed writes.
//
// ResampleView v (moving-horizon planner): path b holds v.len[b] of the dm.N sample
// slots, its rows start at time v.origin[b], v.grid_skip[b] rows of the uniform grid
// are erased in front (row i is grid row i + grid_skip, so that the times keep the
// bits of the grid they were first computed on), and with kSkip row 0's lower-index
// scan starts at v.first_from[b] and row i > 0 is sample keep[b][keep_base + i - 1].
//
// The three [32][D] tiles of a warp leave shared memory as bulk asynchronous copies
// (cp.async.bulk.global.shared::cta, issued by lane 0) instead of a per-lane copy loop.
// A bulk copy needs 16-byte aligned addresses on both sides; rows are D * 8 bytes, so for
// odd D every other row starts 8 bytes off.  The tile is therefore written into shared
// memory with the same 8-byte offset the destination has (`shift`): the head element (if
// any) and a tail element are stored by lane 0 directly, everything between them is one
// aligned bulk copy per array.
// MEASURED AND NOT ADOPTED (profiles/r2_summary.md): with the bulk copies the kernel takes
// 1.197 ms against 1.121 ms for the per-lane copy loop on the cfg2 batch -- its limit is the
// L1 data pipe of the row gathers, which the copy-out does not touch, and lane 0's three
// issue sequences per pass serialise a warp that otherwise stores with 32 lanes.  The code
// stays behind TPB_RESAMPLE_BULK=1 for that A/B; the default build uses the copy loop.
#ifndef TPB_RESAMPLE_BULK
#define TPB_RESAMPLE_BULK 0
#endif
// doubles of one warp's tile region: 3 x (32 x D + 2 for the shift), kept even
__host__ __device__ inline int resample_tile_doubles(int D) { return 3 * ((32 * D + 3) & ~1); }

template <bool kSmemTime, bool kSkip>
__global__ void __launch_bounds__(kResampleThreads)
k_resample_fill(Dims dm, DeviceArrays a, ResampleView v, const int32_t* keep, double start_sec,
                double dt) {
  extern __shared__ __align__(16) double smem[];
  const int stride = dm.N, D = dm.D;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int plane = (32 * D + 3) & ~1;                      // one array's tile, even
  double* s_time = smem;                                    // [N] (kSmemTime)
  // [3][plane] per warp, 16-byte aligned
  double* s_tile = smem + (kSmemTime ? ((stride + 1) & ~1) : 0) + warp * 3 * plane;
  for (int b = blockIdx.x; b < dm.B; b += gridDim.x) {
    const int M = a.r_count[b];
    if (M == 0) continue;  // uniform per CTA
    const int N = v.len ? v.len[b] : stride;
    const double t0 = v.origin ? v.origin[b] : start_sec;
    const int grid_skip = v.grid_skip ? v.grid_skip[b] : 0;
    const int from = (kSkip && v.first_from) ? v.first_from[b] : 0;
    const int kbase = (kSkip && v.keep_base) ? v.keep_base[b] : 0;
    const int64_t row0 = a.r_offset[b];
    const size_t o = static_cast<size_t>(b) * stride;
    const double* tg = a.time + o;
    if (kSmemTime) {
      __syncthreads();  // previous path's searches are done
      for (int i = threadIdx.x; i < N; i += blockDim.x) s_time[i] = tg[i];
      __syncthreads();
    }
    const double* tp = kSmemTime ? s_time : tg;
    for (int base = warp * 32; base < M; base += kResampleThreads) {
      const int i = base + lane;
      const bool live = i < M;
      // rows [base, base + cnt) of this path are contiguous in the packed arrays
      const int64_t wrow = row0 + base;
      const size_t g0 = static_cast<size_t>(wrow) * D;
      // 8-byte offset of each destination from a 16-byte boundary (caller-provided arrays
      // of tpb_resample_uniform_noalloc may start anywhere on an 8-byte boundary)
      auto off8 = [&](const double* base) {
        return TPB_RESAMPLE_BULK
                   ? static_cast<int>((reinterpret_cast<uintptr_t>(base + g0) >> 3) & 1)
                   : 0;
      };
      const int shp = off8(a.r_pos), shv = off8(a.r_vel), sha = off8(a.r_acc);
      // part 1 (per row): the bracketing samples and the scalar outputs
      bool sample_row = false;
      int li = 0, ui = 0;
      double at = 0.0, sdl = 0.0, sdu = 0.0, sddl = 0.0, sddu = 0.0;
      if (live) {
        const int64_t row = row0 + i;
        sample_row = kSkip && i > 0;
        const int src = sample_row ? keep[o + kbase + i - 1] : 0;
        const double t = sample_row ? tp[src] : (kSkip ? t0 : t0 + dt * (i + grid_skip));
        // TimeAtPathSamplesLowerIndex (:686-695): first index with time[index+1] > t,
        // else N-1.  time[] is non-decreasing, so the reference's monotone linear scan
        // equals this binary search.
        int lo = min(from, N - 1), hi = N - 1;  // answer in [from, N-1]
        while (!sample_row && lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (tp[mid + 1] > t) {
            hi = mid;
          } else {
            lo = mid + 1;
          }
        }
        li = sample_row ? src : lo;
        ui = sample_row ? src : min(N - 1, li + 1);
        const double tl = tp[li], tu = tp[ui];
        at = (fabs(tu - tl) < DBL_EPSILON) ? 0.5 : (t - tl) / (tu - tl);
        sdl = a.sd[o + li];
        sdu = a.sd[o + ui];
        sddl = a.sdd[o + li];
        sddu = a.sdd[o + ui];
        if (row < dm.r_cap) {
          a.r_time[row] = t;
          a.r_s[row] = sample_row ? a.s[o + li] : lerp(at, a.s[o + li], a.s[o + ui]);
          a.r_sd[row] = sample_row ? sdl : lerp(at, sdl, sdu);
          a.r_sdd[row] = sample_row ? sddl : lerp(at, sddl, sddu);
        }
      }
#if TPB_RESAMPLE_BULK
      // the bulk copies of this warp's previous pass must have read the tile before it is
      // rewritten; waiting here (not right after issuing them) lets the searches and loads
      // above overlap them
      if (lane == 0) bulk::bulk_wait_read<0>();
      __syncwarp();
#endif
      // part 2 (per row and dof): the tile
      if (live) {
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
            pos = sample_row ? a.q[rl] : lerp(at, a.q[rl], a.q[ru]);
            vel = sample_row ? vl : lerp(at, vl, vu);
            acc = sample_row ? al : clampa(lerp(at, al, au), am);
          }
          s_tile[shp + lane * D + d] = pos;
          s_tile[plane + shv + lane * D + d] = vel;
          s_tile[2 * plane + sha + lane * D + d] = acc;
        }
      }
      int cnt = min(32, M - base);
      if (wrow + cnt > dm.r_cap) cnt = (wrow < dm.r_cap) ? static_cast<int>(dm.r_cap - wrow) : 0;
      const int total = cnt * D;
#if TPB_RESAMPLE_BULK
      // the tile's ordinary stores become visible to the async proxy, then lane 0 copies
      bulk::fence_async_shared();
      __syncwarp();
      if (lane == 0 && total > 0) {
        double* dst[3] = {a.r_pos + g0, a.r_vel + g0, a.r_acc + g0};
        const int shifts[3] = {shp, shv, sha};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const int head = shifts[k];                // elements before the aligned part
          const int body = (total - head) & ~1;      // even number of doubles
          const double* src = s_tile + k * plane + head;
          if (head) dst[k][0] = src[0];
          if (body > 0)
            bulk::bulk_store(dst[k] + head, src + head, static_cast<uint32_t>(body) * 8u);
          if (head + body < total) dst[k][total - 1] = src[total - 1];
        }
        bulk::bulk_commit();
      }
#else
      __syncwarp();
      for (int k = lane; k < total; k += 32) {
        a.r_pos[g0 + k] = s_tile[k];
        a.r_vel[g0 + k] = s_tile[plane + k];
        a.r_acc[g0 + k] = s_tile[2 * plane + k];
      }
      __syncwarp();
#endif
    }
  }
#if TPB_RESAMPLE_BULK
  if (lane == 0) bulk::bulk_wait_read<0>();  // shared memory stays valid until then
#endif
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
  TPB_NVTX("tpb:joint_fit");
  const int paths_per_cta = kFitThreads / 32;
  k_joint_fit<<<(d.B + paths_per_cta - 1) / paths_per_cta, kFitThreads, 0, st>>>(d, a, rounding,
                                                                                 auto_delta);
  return 1;
}

int launch_joint_sample_limit(const Dims& d, const DeviceArrays& a, double safety,
                              cudaStream_t st) {
  TPB_NVTX("tpb:sample_path+limit_curve");
  const size_t tables =
      sizeof(double) * (static_cast<size_t>(d.max_points) + 3 + static_cast<size_t>(d.max_points) * d.D + 2 * d.D);
  const size_t smem = sizeof(double) * (sample_limit_table_doubles(d.max_points, d.D) +
                                        static_cast<size_t>(3) * kJointSampleThreads * d.D);
  if (d.D <= 8 && smem <= 96 * 1024) {
    const int per_cta = kJointSampleThreads * TPB_SL_TILES;
    const dim3 grid = path_grid((d.N + per_cta - 1) / per_cta, d.B);
    // more than 48 KB of dynamic shared memory needs the opt-in (8 dofs, or the spline at
    // the reference's knot capacity after a switch)
#define TPB_SL(K)                                                                          \
  case K:                                                                                  \
    if (smem > 48 * 1024)                                                                  \
      cudaFuncSetAttribute(k_joint_sample_limit<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                           96 * 1024);                                                     \
    k_joint_sample_limit<K><<<grid, kJointSampleThreads, smem, st>>>(d, a, safety);        \
    break
    switch (d.D) {
      TPB_SL(1); TPB_SL(2); TPB_SL(3); TPB_SL(4); TPB_SL(5); TPB_SL(6); TPB_SL(7); TPB_SL(8);
    }
#undef TPB_SL
  } else {
    const dim3 grid = path_grid((d.N + kSampleThreads - 1) / kSampleThreads, d.B);
    k_joint_sample_limit_any<<<grid, kSampleThreads, tables, st>>>(d, a, safety);
  }
  return 1;
}

// C <= 12: rows in registers, kC = C rounded up to even.  Above that the fully
// predicated candidate tests (2*kC candidates x kC rows) cost more than the local-memory
// kernel's early exits (measured, profiles/r1_summary.md v9).  TPB_VARIANT_GENERIC_LIMIT_LOCAL
// (tpb_set_kernel_variant) keeps the local-memory kernel for every C (the parity tests
// compare the two).
int launch_generic_limit(const Dims& d, const DeviceArrays& a, cudaStream_t st) {
  TPB_NVTX("tpb:limit_curve");
  const dim3 grid = path_grid((d.N + kSampleThreads - 1) / kSampleThreads, d.B);
  const bool local = (d.variants >> TPB_VARIANT_GENERIC_LIMIT_LOCAL) & 1u;
  const int kc = local ? 99 : (d.C + 1) / 2 * 2;
  switch (kc) {
    case 2: k_generic_limit_reg<2><<<grid, kSampleThreads, 0, st>>>(d, a); break;
    case 4: k_generic_limit_reg<4><<<grid, kSampleThreads, 0, st>>>(d, a); break;
    case 6: k_generic_limit_reg<6><<<grid, kSampleThreads, 0, st>>>(d, a); break;
    case 8: k_generic_limit_reg<8><<<grid, kSampleThreads, 0, st>>>(d, a); break;
    case 10: k_generic_limit_reg<10><<<grid, kSampleThreads, 0, st>>>(d, a); break;
    case 12: k_generic_limit_reg<12><<<grid, kSampleThreads, 0, st>>>(d, a); break;
    default: k_generic_limit<<<grid, kSampleThreads, 0, st>>>(d, a); break;
  }
  return 1;
}

int launch_integrate(const Dims& d, const DeviceArrays& a, Mode mode, double safety,
                     int max_loops, cudaStream_t st, cudaEvent_t* phase_events) {
  TPB_NVTX("tpb:optimize_path_parameter");
  // phase_events (4, optional): recorded before setup, before the sweeps, after
  // the sweeps and after finalize; a single-kernel path records the inner two
  // around the whole kernel as well
  auto mark = [&](int k) {
    if (phase_events != nullptr) cudaEventRecord(phase_events[k], st);
  };
  const int blocks = (d.B + kWarpsPerBlock - 1) / kWarpsPerBlock;
  const bool joint = mode == Mode::kJoint;
  // TPB_VARIANT_INTEGRATE_WARP selects the one-path-per-warp kernel
  const bool use_pair = ((d.variants >> TPB_VARIANT_INTEGRATE_WARP) & 1u) == 0;
  if (joint && d.D <= 8 && use_pair) {
    // setup and finalize one warp per path, the sweeps two paths per warp
    const int threads = kWarpsPerBlock * 32;
    const int per_block = kPairWarpsPerBlock * 2;
    const int pblocks = (d.B + per_block - 1) / per_block;
    const int pthreads = kPairWarpsPerBlock * 32;
#define TPB_P8(K)                                                                          \
  case K:                                                                                  \
    mark(0);                                                                               \
    k_integrate<Joint8Policy<K>, 7, kPhaseSetup><<<blocks, threads, 0, st>>>(d, a, true, safety, \
                                                                             max_loops);   \
    mark(1);                                                                               \
    k_optimize_pair<K><<<pblocks, pthreads, 0, st>>>(d, a, safety, max_loops);             \
    mark(2);                                                                               \
    k_integrate<Joint8Policy<K>, 7, kPhaseFinalize><<<blocks, threads, 0, st>>>(           \
        d, a, true, safety, max_loops);                                                    \
    mark(3);                                                                               \
    break
    switch (d.D) {
      TPB_P8(1); TPB_P8(2); TPB_P8(3); TPB_P8(4); TPB_P8(5); TPB_P8(6); TPB_P8(7); TPB_P8(8);
    }
#undef TPB_P8
    return 3;
  }
  // Cartesian paths with 6 or 7 dofs: joint rows + the two speed rows on the padding
  // lanes (Joint8Policy<., true>); TPB_VARIANT_CART_GENERIC keeps the generic rows-in-shared-
  // memory solver (the parity tests compare the two)
  if (mode == Mode::kCartesian && (d.D == 6 || d.D == 7) && d.C == 2 * d.D + 2 &&
      ((d.variants >> TPB_VARIANT_CART_GENERIC) & 1u) == 0) {
    const int threads = kWarpsPerBlock * 32;
    const int per_block = kPairWarpsPerBlock * 2;
    const int pblocks = (d.B + per_block - 1) / per_block;
    const int pthreads = kPairWarpsPerBlock * 32;
#define TPB_C8(K)                                                                              \
  case K:                                                                                      \
    mark(0);                                                                                   \
    k_integrate<Joint8Policy<K, true>, 7, kPhaseSetup><<<blocks, threads, 0, st>>>(            \
        d, a, false, safety, max_loops);                                                       \
    mark(1);                                                                                   \
    k_optimize_pair<K, true><<<pblocks, pthreads, 0, st>>>(d, a, safety, max_loops);           \
    mark(2);                                                                                   \
    k_integrate<Joint8Policy<K, true>, 7, kPhaseFinalize><<<blocks, threads, 0, st>>>(         \
        d, a, false, safety, max_loops);                                                       \
    mark(3);                                                                                   \
    break
    switch (d.D) {
      TPB_C8(6); TPB_C8(7);
    }
#undef TPB_C8
    return 3;
  }
  mark(0);
  mark(1);
  if (joint && d.D <= 8) {
    const int threads = kWarpsPerBlock * 32;
#define TPB_J8(K) \
  case K: k_integrate<Joint8Policy<K>, 7><<<blocks, threads, 0, st>>>(d, a, true, safety, max_loops); break
    switch (d.D) {
      TPB_J8(1); TPB_J8(2); TPB_J8(3); TPB_J8(4); TPB_J8(5); TPB_J8(6); TPB_J8(7); TPB_J8(8);
    }
#undef TPB_J8
  } else {
    k_integrate<GenericPolicy, 4><<<blocks, kWarpsPerBlock * 32, 0, st>>>(d, a, joint, safety, max_loops);
  }
  mark(2);
  mark(3);
  return 1;
}

int launch_resample_rows(const Dims& d, const DeviceArrays& a, const ResampleView& v,
                         const int32_t* keep, double start_sec, double dt, int num_sms,
                         cudaStream_t st, bool skip, bool continued) {
  TPB_NVTX("tpb:resample_rows");
  k_scan_offsets<<<1, 1024, 0, st>>>(d.B, a.r_count, a.r_offset, continued ? 1 : 0);
  const size_t tile = sizeof(double) * (kResampleThreads / 32) * resample_tile_doubles(d.D);
  const size_t with_time = tile + sizeof(double) * ((d.N + 1) & ~1);
  const int grid = std::min(d.B, num_sms * 16);
  if (with_time <= 160 * 1024) {
    // > 48 KB of dynamic shared memory is opt-in (per device, so set it every time)
    if (skip) {
      cudaFuncSetAttribute(k_resample_fill<true, true>,
                           cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
      k_resample_fill<true, true><<<grid, kResampleThreads, with_time, st>>>(d, a, v, keep, start_sec, dt);
    } else {
      cudaFuncSetAttribute(k_resample_fill<true, false>,
                           cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
      k_resample_fill<true, false><<<grid, kResampleThreads, with_time, st>>>(d, a, v, keep, start_sec, dt);
    }
  } else if (skip) {
    // the tile alone exceeds 48 KB from D = 10 on
    cudaFuncSetAttribute(k_resample_fill<false, true>,
                         cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    k_resample_fill<false, true><<<grid, kResampleThreads, tile, st>>>(d, a, v, keep, start_sec, dt);
  } else {
    cudaFuncSetAttribute(k_resample_fill<false, false>,
                         cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    k_resample_fill<false, false><<<grid, kResampleThreads, tile, st>>>(d, a, v, keep, start_sec, dt);
  }
  return 2;
}

int launch_resample(const Dims& d, const DeviceArrays& a, double start_sec, double dt,
                    int num_sms, cudaStream_t st, bool continued, bool skip) {
  TPB_NVTX("tpb:resample");
  const int threads = 128;
  // the kept-sample lists of the skipping method live in the boundary passes' scratch
  int32_t* keep = reinterpret_cast<int32_t*>(a.fix_val);
  if (skip) {
    const int warps_per_block = threads / 32;
    k_resample_skip_count<<<(d.B + warps_per_block - 1) / warps_per_block, threads, 0, st>>>(
        d, a, keep, start_sec, dt);
  } else {
    k_resample_count<<<(d.B + threads - 1) / threads, threads, 0, st>>>(d, a, start_sec, dt);
  }
  return 1 + launch_resample_rows(d, a, ResampleView{}, keep, start_sec, dt, num_sms, st, skip,
                                  continued);
}

}  // namespace tpb

// ===========================================================================
// Allocation-free uniform-in-time resampling (tpb_resample_uniform_noalloc).
// No counterpart in the reference; added for
// callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops).
// The per-path row counts -- the scratch the engine keeps in TPB's r_count array --
// live in a file-scope device array instead, so the call allocates nothing.
// ===========================================================================
// TPB_NOALLOC_MAX_PATHS (include/tpb.h) = 65,536 paths.
// 256 KiB per GPU that loads the library, held until the process exits.  Shared by
// every call: calls that overlap on one GPU race on it.
static __device__ __align__(256) int32_t g_noalloc_row_count[TPB_NOALLOC_MAX_PATHS];

namespace tpb {
namespace {
constexpr int kNoAllocMaxDevices = 64;
// device address of g_noalloc_row_count per device ordinal.  In host code the array's
// name is only a stand-in for the symbol; the kernels need this address.
int32_t* g_noalloc_ptr[kNoAllocMaxDevices] = {};
int g_noalloc_sms[kNoAllocMaxDevices] = {};

// k_resample_count for caller-provided arrays: status may be null (every path timed)
__global__ void k_resample_count_any(Dims dm, DeviceArrays a, double start_sec, double dt) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= dm.B) return;
  int m = 0;
  if (a.status == nullptr || a.status[b] == TPB_PATH_OK) {
    const double duration = a.time[static_cast<size_t>(b) * dm.N + dm.N - 1] - start_sec;
    const double c = ceil(duration / dt) + 1;
    m = (c > 0 && c < 2.0e9) ? static_cast<int>(c) : 0;
  }
  a.r_count[b] = m;
}

void set_error(char* buf, size_t cap, const char* msg) {
  if (buf == nullptr || cap == 0) return;
  std::strncpy(buf, msg, cap - 1);
  buf[cap - 1] = '\0';
}
}  // namespace
}  // namespace tpb

extern "C" int32_t tpb_resample_uniform_noalloc(const tpb_resample_desc* d, void* cuda_stream,
                                                char* error_message, size_t error_capacity) {
  using namespace tpb;
  set_error(error_message, error_capacity, "");
  if (d == nullptr) {
    set_error(error_message, error_capacity, "descriptor is null");
    return TPB_INVALID_ARGUMENT;
  }
  if (d->num_paths < 1 || d->num_paths > TPB_NOALLOC_MAX_PATHS) {
    set_error(error_message, error_capacity,
              "num_paths outside [1, 65536] (TPB_NOALLOC_MAX_PATHS: the static scratch array "
              "holds one row count per path)");
    return TPB_OUT_OF_RANGE;
  }
  if (d->num_samples < 2 || d->num_dofs < 1 || d->num_dofs > kMaxDofs) {
    set_error(error_message, error_capacity, "num_samples < 2 or num_dofs outside [1, 16]");
    return TPB_OUT_OF_RANGE;
  }
  if (!(d->time_step_sec > 0) || d->row_capacity < 0) {
    set_error(error_message, error_capacity, "time step must be positive, row_capacity >= 0");
    return TPB_INVALID_ARGUMENT;
  }
  if (d->device < 0 || d->device >= kNoAllocMaxDevices) {
    set_error(error_message, error_capacity, "device ordinal outside [0, 64)");
    return TPB_OUT_OF_RANGE;
  }
  const void* need[] = {d->time, d->s, d->sd, d->sdd, d->q, d->dq, d->ddq, d->max_acceleration,
                        d->r_offset, d->r_time, d->r_s, d->r_sd, d->r_sdd, d->r_pos, d->r_vel,
                        d->r_acc};
  for (const void* p : need) {
    if (p == nullptr) {
      set_error(error_message, error_capacity, "a required device pointer is null");
      return TPB_INVALID_ARGUMENT;
    }
  }
  if (cudaSetDevice(d->device) != cudaSuccess) {
    set_error(error_message, error_capacity, "cudaSetDevice failed");
    return TPB_FAILED_PRECONDITION;
  }
  if (g_noalloc_ptr[d->device] == nullptr) {
    // once per device, with that device current
    void* ptr = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(&ptr, g_noalloc_row_count);
    if (err != cudaSuccess || ptr == nullptr) {
      set_error(error_message, error_capacity, cudaGetErrorString(err));
      return TPB_INTERNAL;
    }
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, d->device) != cudaSuccess ||
        sms < 1) {
      set_error(error_message, error_capacity, "cudaDeviceGetAttribute failed");
      return TPB_INTERNAL;
    }
    g_noalloc_sms[d->device] = sms;
    g_noalloc_ptr[d->device] = static_cast<int32_t*>(ptr);
  }
  Dims dm;
  dm.B = dm.Bmax = d->num_paths;
  dm.D = d->num_dofs;
  dm.N = d->num_samples;
  dm.r_cap = d->row_capacity;
  DeviceArrays a;
  a.time = const_cast<double*>(d->time);
  a.s = const_cast<double*>(d->s);
  a.sd = const_cast<double*>(d->sd);
  a.sdd = const_cast<double*>(d->sdd);
  a.q = const_cast<double*>(d->q);
  a.dq = const_cast<double*>(d->dq);
  a.ddq = const_cast<double*>(d->ddq);
  a.amax = const_cast<double*>(d->max_acceleration);
  a.status = const_cast<int32_t*>(d->status);
  a.r_count = g_noalloc_ptr[d->device];  // every entry is overwritten: nothing to clear
  a.r_offset = d->r_offset;
  a.r_time = d->r_time;
  a.r_s = d->r_s;
  a.r_sd = d->r_sd;
  a.r_sdd = d->r_sdd;
  a.r_pos = d->r_pos;
  a.r_vel = d->r_vel;
  a.r_acc = d->r_acc;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const int threads = 128;
  k_resample_count_any<<<(dm.B + threads - 1) / threads, threads, 0, st>>>(dm, a, d->start_sec,
                                                                          d->time_step_sec);
  launch_resample_rows(dm, a, ResampleView{}, nullptr, d->start_sec, d->time_step_sec,
                       g_noalloc_sms[d->device], st, /*skip=*/false, /*continued=*/false);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) {
    set_error(error_message, error_capacity, cudaGetErrorString(err));
    return TPB_INTERNAL;
  }
  return TPB_OK;
}
