/* TEST INFRASTRUCTURE ONLY -- not part of the product.
 *
 * CPU restatement of PathTimingTrajectory::Plan with the planner's state kept
 * between calls (path_timing_trajectory.cc:477-684): HandleTimeArguments,
 * UpdatePathTrackingStatus, the "already planned enough" shortcut with
 * EraseTrajectoryBefore, the moving-horizon loop over ComputeTimingProfile
 * (:307-475) that splices the per-path-sample arrays window by window, and the
 * resampling of the spliced arrays, for a joint path (TimeablePath::State
 * kNewPath for the first window, kPathWasSampled afterwards, kModifiedPath after
 * orc_planner_switch = SwitchToWaypointPath, tp_oracle_switch.c).
 *
 * PARITY: path_timing_trajectory.cc cannot be compiled on its own here
 * (eigenmath, genit); restated line by line from stages that are pinned
 * individually (fit, sampling, solver = reference object code in _ref).  The
 * reference's integration tests state properties (ends at the last waypoint at
 * rest, limits hold, no duplicate initial samples:
 * path_timing_trajectory_test.cc:112-260), checked in
 * tests/test_oracle_windows.py.
 */
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "tp_oracle.h"

static double lerp1(double t, double a, double b) { return a + t * (b - a); }

/* TimeAtPathSamplesLowerIndex (:686-695) */
static int lower_index(const orc_planner* p, int starting_index, double time) {
  for (int index = starting_index; index < p->len - 1; ++index) {
    if (p->w_time[index + 1] > time) return index;
  }
  return p->len - 1;
}

/* InterpolateAtTime (:709-753) into row `row` of the trajectory; returns the
 * lower index. */
static int interpolate_at_time(orc_planner* p, double time_sec, int lower_start, int row) {
  const int D = p->dim;
  const int li = lower_index(p, lower_start, time_sec);
  const int ui = (p->len - 1 < li + 1) ? p->len - 1 : li + 1;
  const double at = (fabs(p->w_time[ui] - p->w_time[li]) < DBL_EPSILON)
                        ? 0.5
                        : (time_sec - p->w_time[li]) / (p->w_time[ui] - p->w_time[li]);
  for (int d = 0; d < D; ++d) {
    const size_t l = (size_t)li * D + d, u = (size_t)ui * D + d, o = (size_t)row * D + d;
    p->r_pos[o] = lerp1(at, p->w_q[l], p->w_q[u]);
    p->r_vel[o] = lerp1(at, p->w_vel[l], p->w_vel[u]);
    double a = lerp1(at, p->w_acc[l], p->w_acc[u]);
    a = (a < -p->amax[d]) ? -p->amax[d] : a;
    a = (p->amax[d] < a) ? p->amax[d] : a;
    p->r_acc[o] = a;
  }
  p->r_s[row] = lerp1(at, p->w_s[li], p->w_s[ui]);
  p->r_sd[row] = lerp1(at, p->w_sd[li], p->w_sd[ui]);
  p->r_sdd[row] = lerp1(at, p->w_sdd[li], p->w_sdd[ui]);
  return li;
}

/* EraseSamplesUntil (:876-887) */
static void erase_rows_until(orc_planner* p, int offset) {
  if (offset <= 0) return;
  const int D = p->dim, keep = p->M - offset;
  memmove(p->r_t, p->r_t + offset, sizeof(double) * keep);
  memmove(p->r_s, p->r_s + offset, sizeof(double) * keep);
  memmove(p->r_sd, p->r_sd + offset, sizeof(double) * keep);
  memmove(p->r_sdd, p->r_sdd + offset, sizeof(double) * keep);
  memmove(p->r_pos, p->r_pos + (size_t)offset * D, sizeof(double) * (size_t)keep * D);
  memmove(p->r_vel, p->r_vel + (size_t)offset * D, sizeof(double) * (size_t)keep * D);
  memmove(p->r_acc, p->r_acc + (size_t)offset * D, sizeof(double) * (size_t)keep * D);
  p->M = keep;
}

/* EraseTrajectoryBefore (:540-577) */
static void erase_trajectory_before(orc_planner* p, double time_sec) {
  if (p->M == 0 || time_sec < p->r_t[0]) return;
  if (p->skip_method) {
    /* GetSampleCountUntil (:41-44): rows with a timestamp < time_sec */
    int below = 0;
    while (below < p->M && p->r_t[below] < time_sec) ++below;
    /* the values go to a spare row first: row `below` is still needed */
    const int D = p->dim;
    double* tmp = (double*)malloc(sizeof(double) * (size_t)(3 * D + 3));
    {
      /* InterpolateAtTime(time_sec, samples_smaller_than_time): the scan of the
       * PATH samples starts at the number of TRAJECTORY rows before time_sec,
       * as the reference has it (:551-552) */
      orc_planner q = *p;
      double t0 = 0;
      q.r_t = &t0;
      q.r_pos = tmp;
      q.r_vel = tmp + D;
      q.r_acc = tmp + 2 * D;
      q.r_s = tmp + 3 * D;
      q.r_sd = tmp + 3 * D + 1;
      q.r_sdd = tmp + 3 * D + 2;
      interpolate_at_time(&q, time_sec, below, 0);
    }
    /* the reference reads time_[below] without a bounds check; below < M whenever
     * the caller found "planned enough" */
    const int idx = (below < p->M) ? below : p->M - 1;
    if (p->r_t[idx] < time_sec + 0.95 * p->dt) {
      erase_rows_until(p, below);
    } else {
      erase_rows_until(p, below - 1);
    }
    p->r_t[0] = time_sec;
    memcpy(p->r_pos, tmp, sizeof(double) * D);
    memcpy(p->r_vel, tmp + D, sizeof(double) * D);
    memcpy(p->r_acc, tmp + 2 * D, sizeof(double) * D);
    p->r_s[0] = tmp[3 * D];
    p->r_sd[0] = tmp[3 * D + 1];
    p->r_sdd[0] = tmp[3 * D + 2];
    free(tmp);
  } else {
    int offset = (int)round((time_sec - p->r_t[0]) / p->dt);
    if (offset > p->M - 1) offset = p->M - 1;
    erase_rows_until(p, offset);
  }
}

/* ClampToTimeStepMultiple (:229-233) */
static int64_t clamp_to_time_step_multiple(const orc_planner* p, int64_t t) {
  const int64_t loop_multiple = (int64_t)round(orc_time_to_sec(t) / p->dt);
  return orc_time_from_sec(loop_multiple * p->dt);
}

int orc_planner_init(orc_planner* p) {
  p->spline.degree = 2;
  p->spline.dim = p->dim;
  if (p->cart) { /* FitSplineToWaypoints (timeable_path_cartesian_spline.cc:396-462) */
    p->num_points = orc_cart_fit(p->pose_waypoints, p->waypoints, p->W, p->dim,
                                 p->translation_rounding, p->rounding, p->pose_ctrl,
                                 p->spline.points, p->spline.knots);
  } else {
    p->num_points =
        orc_joint_fit(p->waypoints, p->W, p->dim, p->rounding, p->spline.points, p->spline.knots);
  }
  p->spline.num_points = p->num_points;
  p->spline.num_knots = p->num_points + 3;
  p->knot_end = p->spline.knots[p->num_points + 2];
  p->delta_used = (p->delta > 0) ? p->delta : p->knot_end / (p->N - 1);
  /* ResetDerived (:214-227) */
  p->len = 0;
  p->M = 0;
  p->initial_plan = 0;
  p->sampled = 0;
  p->planned_to_end = 1;
  p->target_reached = 0;
  p->start_time = p->end_time = 0;
  p->final_decel_start = orc_time_from_sec(0.0);
  p->path_horizon = p->path_start = p->decel_time = 0.0;
  p->last_extremal_index = 0;
  p->loops = 0;
  p->erased = 0;
  return ORC_OK;
}

/* ComputeTimingProfile (:307-475).  Scratch: q [5][N][dim], tm [4][N], cons. */
static int compute_timing_profile(orc_planner* p, int64_t start, int64_t target_duration,
                                  double* q, double* tm, double* cons) {
  const int N = p->N, D = p->dim, nk = p->spline.num_knots;
  double *dq = q + (size_t)N * D, *ddq = dq + (size_t)N * D, *vel = ddq + (size_t)N * D,
         *acc = vel + (size_t)N * D;
  double *s = tm + N, *sd = s + N, *sdd = sd + N;
  const double start_sec = orc_time_to_sec(start);
  if (target_duration <= 0) return ORC_FAIL_START_ARG; /* (:314-318) */
  int offset = 0;
  double path_start = 0.0, sd_start = 0.0, time_start = start_sec;
  const int old_state = p->sampled; /* 0 new, 1 sampled, 2 modified */
  if (old_state != 0 && p->len > 0) { /* (the reference CHECKs !time_at_path_samples_.empty()) */
    /* clamp(lower_bound(time_at, start) - begin - 1, 0, n - 1) (:334-338) */
    int lo = 0, hi = p->len;
    while (lo < hi) {
      const int mid = lo + (hi - lo) / 2;
      if (p->w_time[mid] < start_sec) lo = mid + 1; else hi = mid;
    }
    offset = lo - 1;
    if (offset < 0) offset = 0;
    if (offset > p->len - 1) offset = p->len - 1;
    path_start = p->w_s[offset];
    sd_start = p->w_sd[offset];
    time_start = p->w_time[offset];
  }
  p->path_start = path_start;
  p->path_horizon = path_start + p->delta_used * (N - 1);
  const int C = p->cart ? 2 * D + 2 : 2 * D;
  if (p->cart) {
    /* TimeableCartesianSplinePath::SamplePath (:464-549) with FakePathIK, whose
     * solution at an IK index does not depend on the warm start, then
     * ConstraintSetup (:551-595) with FakeJacobian */
    const int start_index = (int)round(path_start / p->delta_used); /* PathIkIndex */
    double* pt = (double*)malloc(sizeof(double) * 7 * (size_t)(N + 1));
    double* jt = (double*)malloc(sizeof(double) * (size_t)D * (N + 1));
    double* ik = (double*)malloc(sizeof(double) * (size_t)D * (N + 1));
    double* jac = (double*)malloc(sizeof(double) * 6 * (size_t)D * N);
    orc_cart_sample_targets_from(p->spline.knots, nk, p->pose_ctrl, p->spline.points, D, N,
                                 p->delta_used, start_index, pt, jt);
    orc_fake_ik(pt, jt, N + 1, ik);
    memcpy(q, ik + D, sizeof(double) * (size_t)D * N);
    orc_cart_derivatives(q, N, D, p->delta_used, dq, ddq);
    for (int i = 0; i < N; ++i) orc_fake_jacobian(q + (size_t)D * i, jac + 6 * (size_t)D * i);
    orc_cart_constraints(dq, ddq, jac, N, D, p->vmax, p->amax, p->safety,
                         p->max_translational_velocity, p->max_rotational_velocity, cons);
    free(pt);
    free(jt);
    free(ik);
    free(jac);
  } else {
    orc_joint_sample(p->spline.knots, nk, p->spline.points, D, N, path_start, p->delta_used, q,
                     dq, ddq);
    orc_joint_constraints(dq, ddq, N, D, p->vmax, p->amax, p->safety, cons);
  }
  p->sampled = 1;
  if (old_state != 1) { /* kNewPath or kModifiedPath (:352-385) */
    double sq = 0.0, dot = 0.0;
    for (int d = 0; d < D; ++d) {
      sq += dq[d] * dq[d];
      dot += (p->initial_velocity ? p->initial_velocity[d] : 0.0) * dq[d];
    }
    if (sq > 100 * DBL_EPSILON) {
      const double v = dot / sq;
      sd_start = (v < 0.0) ? 0.0 : v; /* std::max(v, 0.0) */
    }
    double err = 0.0;
    for (int d = 0; d < D; ++d) {
      const double e =
          fabs(dq[d] * sd_start - (p->initial_velocity ? p->initial_velocity[d] : 0.0));
      if (e > err) err = e;
    }
    if (err > p->max_initial_velocity_error) return ORC_FAIL_INITIAL_VELOCITY;
  }
  const int max_loops = (100 > 10 * N) ? 100 : 10 * N;
  orc_solution sol;
  memset(&sol, 0, sizeof(sol));
  sol.time = tm;
  sol.s = s;
  sol.sd = sd;
  sol.sdd = sdd;
  const int st = orc_solve(N, C, cons, path_start, p->path_horizon, sd_start, 0.0, time_start,
                           max_loops, &sol);
  if (st != ORC_OK) return st;
  p->last_extremal_index = sol.last_extremal_index;
  p->decel_time = tm[sol.last_extremal_index];
  orc_assemble(dq, ddq, sd, sdd, p->amax, N, D, vel, acc);
  /* erase from offset, append the new profile (:417-472) */
  if (offset + N > p->P_cap) return ORC_FAIL_CAPACITY;
  memcpy(p->w_time + offset, tm, sizeof(double) * N);
  memcpy(p->w_s + offset, s, sizeof(double) * N);
  memcpy(p->w_sd + offset, sd, sizeof(double) * N);
  memcpy(p->w_sdd + offset, sdd, sizeof(double) * N);
  memcpy(p->w_q + (size_t)offset * D, q, sizeof(double) * (size_t)N * D);
  memcpy(p->w_vel + (size_t)offset * D, vel, sizeof(double) * (size_t)N * D);
  memcpy(p->w_acc + (size_t)offset * D, acc, sizeof(double) * (size_t)N * D);
  p->len = offset + N;
  return ORC_OK;
}

/* ResampleTrajectory (:697-707) */
static int resample_trajectory(orc_planner* p, double start_sec) {
  if (p->len == 0) { /* the reference would read back() of empty vectors */
    p->M = 0;
    return ORC_FAIL_NO_SAMPLES;
  }
  if (p->skip_method) {
    if (p->M_cap < p->len + 1) return ORC_FAIL_CAPACITY;
    p->M = orc_resample_skip(p->len, p->dim, p->w_time, p->w_s, p->w_sd, p->w_sdd, p->w_q,
                             p->w_vel, p->w_acc, p->amax, start_sec, p->dt, p->r_t, p->r_pos,
                             p->r_vel, p->r_acc, p->r_s, p->r_sd, p->r_sdd);
  } else {
    const int M = orc_resample_count(p->w_time[p->len - 1], start_sec, p->dt);
    if (M > p->M_cap) return ORC_FAIL_CAPACITY;
    p->M = (M > 0) ? M : 0;
    orc_resample_uniform(p->len, p->dim, p->w_time, p->w_s, p->w_sd, p->w_sdd, p->w_q, p->w_vel,
                         p->w_acc, p->amax, start_sec, p->dt, p->M, p->r_t, p->r_pos, p->r_vel,
                         p->r_acc, p->r_s, p->r_sd, p->r_sdd);
  }
  return ORC_OK;
}

/* Plan (:579-684) */
int orc_planner_plan(orc_planner* p, int64_t start, int64_t horizon) {
  const int N = p->N, D = p->dim;
  const double start_sec = orc_time_to_sec(start);
  p->loops = 0;
  p->erased = 0;
  /* HandleTimeArguments (:503-538); absl::Seconds(double) has quarter-ns ticks */
  if (p->initial_plan && 4 * start > 4 * p->end_time + llround(p->dt * 4e9)) {
    return ORC_FAIL_START_RANGE;
  }
  if (!p->initial_plan) {
    p->start_time = start;
    p->end_time = start;
    p->path_start = 0.0;
  } else {
    if (start > p->end_time) return ORC_FAIL_START_ARG;
    if (start < p->start_time) return ORC_FAIL_START_ARG;
    p->start_time = start;
  }
  /* UpdatePathTrackingStatus (:477-501) */
  p->target_reached = 0;
  p->planned_to_end = 0;
  if (!p->initial_plan) {
    p->path_horizon = 0;
    p->path_start = 0;
  } else {
    p->planned_to_end = p->path_horizon >= p->knot_end - 1e-4; /* CloseToEnd */
    if (p->planned_to_end) {
      if (p->sampled == 1) {
        p->target_reached = 1;
      } else {
        p->path_horizon = 0.0;
        p->path_start = 0.0;
        p->planned_to_end = 0;
      }
    }
  }
  const int planned_enough = (p->sampled == 1) && (p->final_decel_start >= start + horizon);
  if (p->M > 0 && planned_enough) { /* (:594-599) */
    erase_trajectory_before(p, start_sec);
    p->erased = 1;
    return ORC_OK;
  }
  if (p->initial_plan) { /* GetTimeOffsetAfter (:289-305); the erase itself is moot */
    if (p->M == 0) return ORC_FAIL_NO_SAMPLES;
    if (start_sec < p->r_t[0]) return ORC_FAIL_START_RANGE;
    if (!(p->r_t[p->M - 1] > start_sec)) return ORC_FAIL_START_END;
  }
  double* cons = (double*)malloc(sizeof(double) * (size_t)N * (2 * D + 2) * 4);
  double* q = (double*)malloc(sizeof(double) * (size_t)N * D * 5);
  double* tm = (double*)malloc(sizeof(double) * (size_t)N * 4);
  int st = ORC_OK;
  int64_t loop_start = start;
  int horizon_reached = 0;
  for (int loop = 0; !p->planned_to_end && !horizon_reached; loop++) {
    st = compute_timing_profile(p, loop_start, start + horizon - loop_start, q, tm, cons);
    if (st != ORC_OK) break;
    p->loops = loop + 1;
    /* (:635-658) */
    const int half = N / 2;
    const int decel_start = (p->last_extremal_index > half) ? p->last_extremal_index : half;
    p->final_decel_start = orc_time_from_sec(tm[decel_start]);
    p->planned_to_end = p->path_horizon >= p->knot_end - 1e-4;
    horizon_reached = (tm[N - 1] - start_sec) > ((double)horizon / 1e9);
    if (loop >= p->max_planning_iterations) {
      st = ORC_FAIL_DEADLINE;
      break;
    }
    loop_start = p->final_decel_start;
  }
  free(cons);
  free(q);
  free(tm);
  if (st != ORC_OK) return st;
  st = resample_trajectory(p, start_sec);
  if (st != ORC_OK) return st;
  p->initial_plan = 1;
  if (p->M > 0) { /* (:662-674) */
    p->end_time = clamp_to_time_step_multiple(p, orc_time_from_sec(p->r_t[p->M - 1]));
    p->final_decel_start = clamp_to_time_step_multiple(p, orc_time_from_sec(p->decel_time));
  } else {
    p->end_time = p->start_time;
    p->final_decel_start = p->end_time;
  }
  p->target_reached = p->planned_to_end;
  return ORC_OK;
}

/* TimeableCartesianSplinePath::SwitchToWaypointPath on the planner's path (cart != 0;
 * pose_ctrl must hold spline.knot_capacity - 3 rows).  The IK cache the reference trims
 * (:188-192) does not exist here: the fake IK does not depend on its warm start. */
int orc_planner_cart_switch(orc_planner* p, double keep_path_until, const double* pose_waypoints,
                            const double* joint_waypoints, int W) {
  p->sampled = 2; /* kModifiedPath */
  const int st = orc_cart_switch(&p->spline, p->pose_ctrl, keep_path_until, pose_waypoints,
                                 joint_waypoints, W, p->translation_rounding, p->rounding);
  if (p->spline.num_knots > 0) p->knot_end = p->spline.knots[p->spline.num_knots - 1];
  p->num_points = p->spline.num_points;
  return st;
}

int orc_planner_switch(orc_planner* p, double keep_path_until, const double* waypoints, int W) {
  double new_ctrl[(3 * (ORC_MAX_SWITCH_WAYPOINTS + 1)) * 64];
  int n = 0;
  p->sampled = 2; /* kModifiedPath, also when the call fails (.cc:212) */
  const int st = orc_joint_switch(&p->spline, keep_path_until, waypoints, W, p->rounding, new_ctrl, &n);
  if (p->spline.num_knots > 0) p->knot_end = p->spline.knots[p->spline.num_knots - 1];
  p->num_points = p->spline.num_points;
  return st;
}
