// Stage C, the extremal sweeps and the switching-point loop, for joint-space
// paths with D <= 8: TWO paths per warp (k_optimize_pair<D>), 16 lanes each.
//
// Same algorithm and the same results, bit for bit, as kPhaseOptimize of
// k_integrate<Joint8Policy> (tpb_integrate.cuh; AddForwardExtremal /
// AddBackwardExtremal time_optimal_path_timing.cc:769-952, the switching-point
// loop :287-490, NextCriticalPoint :697-720 through the lookup tables the setup
// phase leaves behind).  What changes is the mapping onto the warp:
//
//  * a half-warp (lanes 16h .. 16h+15) owns one path;
//  * the sweeps are a per-half state machine that advances ONE extremal step
//    per trip of a common loop.  The forward and the backward step are the same
//    code with a direction sign, and FindSddMax / FindSddMin are one routine
//    that maximises +-candidate (negation is exact).  Whatever the two paths are
//    doing -- sweeping in opposite directions, at different samples -- the
//    halves execute the same instructions, so one instruction stream times two
//    paths;
//  * every warp primitive in the step is issued by all 32 lanes with the full,
//    compile-time mask (a per-half mask is not warp-uniform, and the compiler
//    then serialises the two halves around every shuffle and vote): ballots are
//    taken over the warp and each half reads its own 16 bits, shuffles use
//    width 16, and the rare sub-steps (validity test at the limit curve,
//    ComputeSddAtIntersection) are entered by the whole warp when either half
//    needs them, with the side effects predicated.  Only the driver transitions
//    between sweeps (a few dozen per path) diverge, and they use no primitives.
//
// Lane = 8*g + d within the half: d = acceleration row, g = 0 holds the row's
// outer candidate (the larger one for FindSddMax, the smaller one for
// FindSddMin), g = 1 the inner one.  See Joint8Policy for why testing the outer
// candidates first is exact.
#pragma once

#include "tpb_device.cuh"
#include "tpb_engine.h"

namespace tpb {

constexpr int kPairWarpsPerBlock = 1;  // 2 paths per CTA: 2,048 CTAs spread evenly over the SMs

// kCart: Cartesian rows (see Joint8Policy): the padding lane d == kD of each group of
// the half carries one of the two extra velocity-kind rows; Vk is its B coefficient.
template <int kD, bool kCart = false>
struct PairSolver {
  // lane geometry
  int lane;     // 0..15 within the half
  int g, d;     // lane = 8*g + d
  int shift;    // 0 or 16: position of the half in a warp ballot
  // this half's path
  // Padding lanes (d >= kD) hold a copy of row kD-1 and of its limits: a row
  // tested twice changes no outcome, and the copies keep real operands in the
  // divisions (0/1 would take the slow path of the double-precision division).
  const double *dq, *ddq;  // rows, already offset by min(d, kD-1)
  double aS, vS2;          // amax_d*safety, (vmax_d*safety)^2
  // kCart only: lanes with xsel read Vk from bx[idx * xstride]; vlo = their lower bound
  const double* bx;
  int xstride;
  bool xsel;
  double vlo;
  int N;
  double ds;
  double *sd2, *sdd;
  const double* bmax;
  const uint8_t* type;
  const int *next_src, *prev_eq0;  // Solver::build_critical_tables

  // This half's 16 bits of a warp-wide ballot.  Convergent: all 32 lanes.
  __device__ __forceinline__ unsigned ballot(bool pred) const {
    return (__ballot_sync(kFullMask, pred) >> shift) & 0xffffu;
  }
  __device__ __forceinline__ static bool any(bool pred) {
    return __any_sync(kFullMask, pred);
  }
  __device__ __forceinline__ int clampi(int i) const { return min(max(i, 0), N - 1); }

  // The sweeps work in "key space": with sgn = +1 for a forward and -1 for a
  // backward sweep, rows carry Ak = sgn * q'_d and accelerations are handled as
  // k = sgn * sdd.  Sign flips are exact, so  A*sdd == Ak*k,  (+-aS - Bs)/A ==
  // sgn * ((+-aS - Bs)/Ak), and FindSddMin(sdd) becomes FindSddMax(k): one search
  // routine and one step routine serve both directions without selects.

  // AreDerivativesValid (:624-636) for sdd = sgn * k; see Joint8Policy::valid.
  // Convergent.
  __device__ __forceinline__ bool valid_k(double Ak, double B, double k, double sd2v) const {
    const double v = Ak * k + B * sd2v;
    bool bad = outside_sym(v, aS);
    const double w = 0.0 * k + vcoef(Ak) * sd2v;
    bad = bad || (w + kTiny < vlower()) || (w - kTiny > vS2);
    return ballot(bad) == 0u;
  }
  // B coefficient of this lane's velocity-kind row at the sweep's sample
  __device__ __forceinline__ double vcoef(double Ak) const { return kCart ? Vk : Ak * Ak; }
  __device__ __forceinline__ double vlower() const { return kCart ? vlo : 0.0; }
  __device__ __forceinline__ double load_v(int at, double A) const {
    double v = A * A;
    if (xsel) v = bx[static_cast<size_t>(at) * xstride];
    return v;
  }

  // FindSddMax / FindSddMin (:638-695) as the maximum valid k.  *none_valid
  // tells that no candidate passed (the reference then returns 0.0).
  // Convergent; the data are per half.
  __device__ __forceinline__ double find_k(double Ak, double B, double sd2v,
                                           bool* none_valid) const {
    const double Bs = B * sd2v;
    const bool plus = (g == 0) == (Ak > 0.0);  // g = 0: the row's larger candidate
    const double num = (plus ? aS : -aS) - Bs;
    // rows without a candidate divide 1 by 1 (x/0 and 0/x leave the division's
    // fast path); a row exactly on its limit (num == 0, common on the limit
    // curve) yields a zero with the quotient's sign: num * Ak
    const bool none = is_tiny(Ak);
    const bool zero = num == 0.0;
    const double quot = guarded_div(none || zero, num, Ak);
    const double cand = none ? quiet_nan() : (zero ? num * Ak : quot);
    const double vv = vcoef(Ak) * sd2v;
    const bool vbad = ballot((vv + kTiny < vlower()) || (vv - kTiny > vS2)) != 0u;
    double best = -DBL_MAX;
    const unsigned gmask = 0xffu << (shift + 8 * g);  // this group's rows in a warp ballot
    auto round = [&](int row, int src_group) {
      const double c = __shfl_sync(kFullMask, cand, (src_group << 3) | (row & 7), 16);
      const double v = Ak * c + Bs;
      const bool bad = outside_sym(v, aS);
      const unsigned m = __ballot_sync(kFullMask, bad);
      // NaN marks "no candidate" and fails c > best
      const bool ok = (row < kD) && ((m & gmask) == 0u);
      best = (ok && c > best) ? c : best;
    };
    auto reduce = [&]() {
      const double w = __shfl_xor_sync(kFullMask, best, 8, 16);
      best = (w > best) ? w : best;
    };
#pragma unroll
    for (int j0 = 0; j0 < kD; j0 += 2) round(j0 + g, 0);
    reduce();
    // Inner candidates only matter to a half where one lies beyond its best outer
    // one; testing them in the other half as well changes nothing there (none of
    // its inner candidates exceeds its best).
    const bool beyond = (g == 1) && (cand > best);
    if (any(beyond)) {
#pragma unroll
      for (int j0 = 0; j0 < kD; j0 += 2) round(j0 + g, 1);
      reduce();
    }
    *none_valid = vbad || best == -DBL_MAX;
    return best;
  }

  // ComputeSddAtIntersection (:722-751).  Convergent; stores only where `on`.
  // The reference tries up to three candidates in order and keeps the first
  // valid one; all three are tested here and the first valid one selected.
  __device__ __forceinline__ void sdd_at_intersection(int index, bool on) {
    index = clampi(index);  // a half that is not `on` may have drifted out of range
    const bool has_hi = index < N - 1, has_lo = index > 0;
    const double sp = sd2[clampi(index + 1)];
    const double sm = sd2[clampi(index - 1)];
    const double here = sd2[index];
    const double c1 = 0.25 / ds * (sp - sm);
    const double c2 = 0.5 / ds * (sp - here);
    const double c3 = 0.5 / ds * (here - sm);
    const bool v1 = valid_k(Ak, Bk, sgn * c1, here);
    const bool v2 = valid_k(Ak, Bk, sgn * c2, here);
    const bool v3 = valid_k(Ak, Bk, sgn * c3, here);
    double r = 0.0;
    if (has_lo && v3) r = c3;
    if (has_hi && v2) r = c2;
    if (has_lo && has_hi && v1) r = c1;
    if (on && lane == 0) sdd[index] = r;
    __syncwarp();
  }

  // -------------------------------------------------------------------------
  // Per-half state machine: OptimizePathParameter (:287-490) unrolled into
  // sweeps that advance one extremal step per trip.
  // -------------------------------------------------------------------------
  enum : int { kFirstBack = 0, kFirstFwd = 1, kLoopBack = 2, kLoopFwd = 3 };
  bool sweeping;  // a sweep is in progress (false: this half is done or idle)
  int stage, status;
  // sweep state (AddForwardExtremal :769-857 / AddBackwardExtremal :859-952)
  int dir, idx;
  int left;        // steps the sweep may still take before it runs into the path's end
  int end_ret;     // what the sweep returns there: N - 1 forward, 0 backward
  double sgn;      // dir as a double
  int stop_types;  // types that end a sweep at the limit curve (kSink forward, kSource
                   // backward); 0 for a connecting sweep, which never ends there
  int strict;      // 1: an inadmissible limit-curve step ends the sweep (not connecting)
  double cur, bcur, bnb, nb;  // sd2[idx], sd2_max[idx], sd2_max[idx+dir], sd2[idx+dir]
  int tcur, tnb;
  double Ak, Bk;  // rows of sample idx: sgn * q'_d, q''_d
  double Vk;      // kCart: B of this lane's velocity-kind row at idx
  // driver state
  int iforw_hi, iback_lo, iback_hi, icrit, icrit_lo, icrit_hi, loop, max_loops;

  // A store the same half reads back later in the same (divergent) transition:
  // every lane writes the same value and each then reads its own write.
  __device__ __forceinline__ void st_all(double* ptr, double v) const { *ptr = v; }

  // Starts a sweep at sample `start`.  Returns true (and the sweep's return value
  // in *ret) when the reference's early-out applies, so no step is needed.
  __device__ __forceinline__ bool begin_sweep(int direction, int start, int* ret) {
    dir = direction;
    sgn = static_cast<double>(direction);
    const bool connecting = (direction < 0) && (start != (N - 1));
    stop_types = connecting ? 0 : ((direction > 0) ? kSink : kSource);
    strict = connecting ? 0 : 1;
    left = (direction > 0) ? (N - 2 - start) : (start - 1);
    end_ret = (direction > 0) ? N - 1 : 0;
    if (direction > 0) {
      if (start >= N - 2) {
        *ret = N - 1;
        return true;
      }
    } else {
      if (start <= 1) {
        *ret = 0;
        return true;
      }
    }
    load_state(start);
    return false;
  }

  // The sweep's registers for sample `at` (1 <= at + dir <= N - 1).
  __device__ __forceinline__ void load_state(int at) {
    idx = at;
    cur = sd2[at];
    bcur = bmax[at];
    tcur = type[at];
    Ak = sgn * dq[at * kD];
    Bk = ddq[at * kD];
    if (kCart) Vk = load_v(at, Ak);
    bnb = bmax[at + dir];
    tnb = type[at + dir];
    nb = sd2[at + dir];
  }

  // A run of steps along the limit curve (:790-794 / :882-886): while the sweep
  // sits on the curve and this and the next sample are Trajectory points, a step
  // copies sd2_max[next] and takes sdd from the difference -- no search, and
  // nothing that depends on the step before except where it stands.  Lane k does
  // step k of the run for the halves with `on` set; the run ends before the
  // first step that the one-step routine has to look at (an intersection with an
  // earlier sweep, a non-Trajectory sample, the path's end).  Convergent.
  __device__ __forceinline__ void follow_run(bool on) {
    const int j = clampi(idx + dir * lane), jn = clampi(j + dir);
    const int tj = type[j], tn = type[jn];
    const double bj = bmax[j], bn = bmax[jn], old = sd2[jn];
    const double c = (lane == 0) ? cur : bj;  // sd2 where step k starts
    const bool ok = on && (lane < left) && (lane < 15) && ((tj & tn & kTrajectory) != 0) &&
                    !((old == old) && (old < bn)) && !(bn < 0);
    const int len = __ffs(~ballot(ok)) - 1;  // leading steps that are all plain
    const double sddv = div_pos(0.5 * fma(sgn, bn - c, 0.0), ds);
    if (on && lane < len) {
      sd2[jn] = bn;
      sdd[j] = sddv;
    }
    __syncwarp();
    if (on && len > 0) {
      left -= len;
      load_state(idx + dir * len);
      // Consume the reloaded state here.  A scoreboard wait is attached to the
      // first instruction that reads a loaded register; left to the step below,
      // every trip would pay for it (and wait for that trip's prefetch as well).
      // None of these comparisons can hold.
      const bool never = (cur < -1e300) | (bcur < -1e300) | (bnb < -1e300) | (nb < -1e300) |
                         (Ak != Ak) | (Bk != Bk) | ((tcur + tnb) == 200);
      if (never) left = 0;
    }
  }

  // One iteration of the extremal loop in direction dir, for the halves with
  // `live` set.  Convergent.  Returns true (to a live half) when its sweep has
  // ended, with the sweep's return value in *ret.
  __device__ __forceinline__ bool sweep_step(bool live, int* ret) {
    {
      const bool on_curve = live && (left > 0) && is_tiny(cur - bcur) &&
                            ((tcur & tnb & kTrajectory) != 0);
      if (any(on_curve)) follow_run(on_curve);
    }
    const bool fwd = dir > 0;
    const bool at_end = left <= 0;
    bool run = live && !at_end;
    bool ended = live && at_end;
    int r = end_ret;
    // data of the step after this one (clamped: a half that is not running may
    // sit anywhere)
    const int i1 = clampi(idx + dir), i2 = clampi(idx + 2 * dir);
    const double A_n = dq[i1 * kD];
    const double B_n = ddq[i1 * kD];
    double V_n = 0.0;
    if (kCart) V_n = load_v(i1, A_n);
    const double bnb_n = bmax[i2];
    const int tnb_n = type[i2];
    const double nb_n = sd2[i2];

    // The step that lands on the limit curve, and whether it is admissible
    // (:806-813 / :898-905); independent of the search below.  fwd: 0.5*(bnb -
    // cur)/ds, bwd: 0.5*(cur - bnb)/ds; the fma makes the difference a +0 in
    // both directions when the two are equal.
    const double sb = div_pos(0.5 * fma(sgn, bnb - cur, 0.0), ds);
    const bool sb_ok = valid_k(Ak, Bk, sgn * sb, fwd ? bcur : cur);

    const bool follow = is_tiny(cur - bcur) && ((tcur & tnb & kTrajectory) != 0);
    double k = 0.0;
    bool none_valid = true;
    if (any(run && !follow)) k = find_k(Ak, Bk, cur, &none_valid);
    // fwd: cur + 2ds*sdd, bwd: cur - 2ds*sdd; both are cur + 2ds*k
    double sd2tmp = follow ? bnb : cur + 2.0 * ds * (none_valid ? 0.0 : k);
    double sddtmp = follow ? sb : (none_valid ? 0.0 : sgn * k);
    const bool isect = (nb == nb) && (nb < sd2tmp);
    const bool over = sd2tmp > bnb;
    const bool stop = over && (((tnb & stop_types) | (sb_ok ? 0 : strict)) != 0);
    if (over) {
      sd2tmp = bnb;
      sddtmp = sb;
    }
    const bool neg = sd2tmp < 0;
    // the rare ends of a step: intersection with an earlier sweep, leaving the
    // limit curve's admissible part, a negative sd2
    if (any(ended || (run && (isect || stop || neg)))) {
      if (any(run && isect)) sdd_at_intersection(idx, run && isect);
      if (run && isect) {
        ended = true;
        run = false;
      } else if (run && stop) {
        ended = true;
        run = false;
        r = idx;
      } else if (run && neg) {
        sd2tmp = 0.0;
        if (fwd) {
          sddtmp = (idx <= 1) ? 0.0 : div_pos(-sd2[idx - 1], ds);
        } else {
          sddtmp = (idx < N - 1) ? div_pos(sd2[idx + 1], ds) : 0.0;
        }
      }
    }
    if (run && lane == 0) {
      sd2[idx + dir] = sd2tmp;
      sdd[idx] = sddtmp;
    }
    __syncwarp();
    // (a half that is not running rotates garbage; its next sweep reloads)
    const double A_nx = A_n, B_nx = B_n, bnb_nx = bnb_n, nb_nx = nb_n;
    const int tnb_nx = tnb_n;
    cur = sd2tmp;
    bcur = bnb;
    tcur = tnb;
    Ak = sgn * A_nx;
    Bk = B_nx;
    if (kCart) Vk = V_n;
    bnb = bnb_nx;
    tnb = tnb_nx;
    nb = nb_nx;
    idx += dir;
    --left;
    *ret = r;
    return ended;
  }

  // NextCriticalPoint (:697-720); see Solver::next_critical
  __device__ __forceinline__ int next_critical(int idx_lo, int idx_hi) const {
    if (idx_lo + 1 > idx_hi) return -1;
    const int q = next_src[idx_lo + 1];
    if (q > idx_hi) return -1;
    const int p2 = prev_eq0[idx_hi];
    return (p2 > q) ? p2 : q;
  }

  // Driver transitions (:300-396).  Called by a half whose sweep has ended with
  // value `ret`; runs until the next sweep that needs stepping, or to the end.
  // Divergent between the halves: plain loads and stores only.
  __device__ void advance(int ret) {
    for (;;) {
      if (stage == kFirstBack) {
        iback_lo = ret;
        stage = kFirstFwd;
        if (!begin_sweep(+1, 0, &ret)) return;
        continue;
      }
      if (stage == kLoopBack) {
        iback_lo = ret;
        stage = kLoopFwd;
        if (!begin_sweep(+1, icrit, &ret)) return;
        continue;
      }
      if (stage == kFirstFwd) {
        iforw_hi = ret;
        icrit_hi = iback_lo;
        if ((iforw_hi < icrit_hi) && ((icrit_hi < N - 2) && (icrit_hi >= 2))) {
          st_all(&sd2[icrit_hi], quiet_nan());
          icrit_hi++;
          iback_lo++;
        }
        icrit_lo = iforw_hi;
        loop = 0;
      } else {  // kLoopFwd
        iforw_hi = ret;
        if (iback_lo > icrit_lo) {
          status = TPB_PATH_FAIL_CONNECT;
          sweeping = false;
          return;
        }
        icrit_lo = iforw_hi;
        ++loop;
      }
      // top of the switching loop (:345-352)
      if (loop >= max_loops || iforw_hi >= icrit_hi) {
        sweeping = false;
        return;
      }
      icrit = next_critical(icrit_lo, icrit_hi);
      if (icrit < 0 || icrit >= N) icrit = static_cast<int>(0.5 * (icrit_lo + icrit_hi));
      if (icrit > 0 && icrit < N - 1) st_all(&sd2[icrit], bmax[icrit]);
      if (icrit < 1) {
        status = TPB_PATH_FAIL_INDEX;
        sweeping = false;
        return;
      }
      if (bmax[icrit - 1] <= bmax[icrit]) {
        iback_hi = icrit - 1;
        st_all(&sd2[icrit - 1], bmax[icrit - 1]);
      } else {
        iback_hi = icrit;
      }
      stage = kLoopBack;
      if (!begin_sweep(-1, iback_hi, &ret)) return;
    }
  }

  // Runs both sweeps and the switching-point loop of this half's path to the
  // end, in lock step with the other half (go = false: nothing to do, idle).
  __device__ void run(bool go) {
    // an idle half still executes every step: keep its state harmless
    sweeping = false;
    dir = 1;
    sgn = 1.0;
    stop_types = 0;
    idx = 0;
    left = 0;
    end_ret = 0;
    strict = 0;
    cur = bcur = bnb = nb = 0.0;
    tcur = tnb = kNone;
    Ak = Bk = 0.0;
    Vk = 0.0;
    if (go) {
      stage = kFirstBack;
      sweeping = true;
      int ret = 0;
      if (begin_sweep(-1, N - 1, &ret)) advance(ret);
    }
    __syncwarp();
    bool more = any(sweeping);
    while (more) {
      int ret = 0;
      const bool ended = sweep_step(sweeping, &ret);
      if (any(ended)) {
        if (ended) advance(ret);
        __syncwarp();
        more = any(sweeping);
      }
    }
  }
};

// Phase kPhaseOptimize for two paths per warp.  Runs between
// k_integrate<Joint8Policy<kD>, ., kPhaseSetup> and <., ., kPhaseFinalize>.
template <int kD, bool kCart = false>
__global__ void __launch_bounds__(kPairWarpsPerBlock * 32, 16 / kPairWarpsPerBlock)
k_optimize_pair(Dims dm, DeviceArrays a, double safety, int max_loops) {
  const int warp = threadIdx.x >> 5;
  const int half = (threadIdx.x >> 4) & 1;
  const int lane = threadIdx.x & 15;
  const int b = (blockIdx.x * kPairWarpsPerBlock + warp) * 2 + half;
  const bool has_path = b < dm.B && (a.active == nullptr || a.active[b] != 0);
  const int bb = has_path ? b : 0;  // an absent half computes on path 0 and stores nothing
  const int N = dm.N;
  const size_t o = static_cast<size_t>(bb) * N;

  PairSolver<kD, kCart> S;
  S.lane = lane;
  S.g = lane >> 3;
  S.d = lane & 7;
  S.shift = half * 16;
  const int dd = min(S.d, kD - 1);  // padding lanes copy the last row
  S.dq = a.dq + o * kD + dd;
  S.ddq = a.ddq + o * kD + dd;
  S.N = N;
  S.sd2 = a.sd2 + o;
  S.sdd = a.sdd + o;
  S.bmax = a.b_sd2_max + o;
  S.type = a.b_type + o;
  S.next_src = reinterpret_cast<const int*>(a.time + o);
  S.prev_eq0 = S.next_src + N;
  S.max_loops = max_loops;
  S.aS = a.amax[static_cast<size_t>(bb) * kD + dd] * safety;
  const double vs = a.vmax[static_cast<size_t>(bb) * kD + dd] * safety;
  S.vS2 = vs * vs;
  S.bx = nullptr;
  S.xstride = 0;
  S.xsel = false;
  S.vlo = 0.0;
  if (kCart) {
    // rows 2*kD + g of the stored constraints, [sample][C][4] = (A, B, lower, upper)
    S.xsel = (S.d == kD);
    if (S.xsel) {
      const double* row0 = a.cons + (o * dm.C + 2 * kD + S.g) * 4;
      S.bx = row0 + 1;
      S.xstride = dm.C * 4;
      S.vlo = row0[2];
      S.vS2 = row0[3];
    }
  }
  S.status = has_path ? a.status[bb] : TPB_PATH_FAIL_SETUP;
  const bool go = S.status == TPB_PATH_OK;
  // SetSetupDone (:535-552), as in the setup phase
  S.ds = 1.0;
  if (go) {
    // Cartesian paths: SetupProblem's range as k_cart_constraints* stored it
    const double s_start = kCart ? a.s_start[bb] : a.path_start[bb];
    const double s_end = kCart ? a.s_end[bb] : s_start + a.delta[bb] * (N - 1);
    S.ds = (s_end - s_start) / (N - 1);
  }
  S.run(go);
  if (go && lane == 0) a.status[b] = S.status;
}

}  // namespace tpb
