/* TEST INFRASTRUCTURE ONLY -- the CPU oracle.  Never linked into, imported by,
 * or called from the product (libtpb.so / the host C++ layer).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may use it, and only as the checker or the timed CPU baseline.
 *
 * A plain-C restatement of the reference's algorithm for the hot path
 * (theteamatx/x-edr-trajectory-planning), each function citing the reference
 * file:line it follows.  Parity pin: see the header comment of tp_oracle.c.
 *
 * Layout conventions (all double, row-major):
 *   knots   [num_knots]
 *   ctrl    [num_points][dim]
 *   cons    [N][C][4]  with columns (A, B, lower, upper) of
 *           lower <= A*sdd + B*sd^2 <= upper
 *   q/dq/ddq [N][dim]
 */
#ifndef ORACLE_TP_ORACLE_H_
#define ORACLE_TP_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes of orc_solve (0 == success; the reference only has bool). */
enum {
  ORC_OK = 0,
  ORC_FAIL_SETUP = 1,    /* SetupProblem / IsSetupValid returned false */
  ORC_FAIL_CONNECT = 2,  /* iback_lo > icrit_lo */
  ORC_FAIL_NAN = 3,      /* sd2 still NaN after the extremal loop */
  ORC_FAIL_TERMINAL = 4, /* non-zero terminal velocity */
  ORC_FAIL_INDEX = 5,    /* reference would index sd2_max[-1] (UB there) */
  ORC_FAIL_CAPACITY = 6, /* path samples / rows exceed the caller's capacity */
  ORC_FAIL_DEADLINE = 7, /* kDeadlineExceeded (path_timing_trajectory.cc:652-655) */
  ORC_FAIL_START_RANGE = 8,  /* kOutOfRange (:504-511, :294-296) */
  ORC_FAIL_START_ARG = 9,    /* kInvalidArgument (:517-535, :314-318) */
  ORC_FAIL_START_END = 10,   /* kInternal: start >= last trajectory time (:298-303) */
  ORC_FAIL_NO_SAMPLES = 11,  /* kFailedPrecondition (:291-293) */
  ORC_FAIL_INITIAL_VELOCITY = 12 /* kInvalidArgument (:377-385) */
};

/* ---- stage A: B-splines ------------------------------------------------ */
int orc_knot_span(const double* knots, int num_knots, int degree, double u);
void orc_basis(const double* knots, int span, int p, double u, double* basis);
void orc_basis_ders(const double* knots, int span, int p, int der, double u,
                    double* ders /* [(der+1)][(p+1)] */);
/* returns 0 ok, 11 (kOutOfRange) if u outside [knots[0], knots[last]]. */
int orc_eval_curve(const double* knots, int num_knots, const double* ctrl,
                   int dim, int degree, double u, double* value /* [dim] */);
int orc_eval_curve_ders(const double* knots, int num_knots, const double* ctrl,
                        int dim, int degree, int nder, double u,
                        double* values /* [(nder+1)][dim] */);
int orc_make_uniform_knots(int num_points, int degree, double lo, double hi,
                           double* knots /* [num_points+degree+1] */);

/* ---- stage B: joint-space timeable path -------------------------------- */
/* returns number of control points written (3W-2, or 4 if W==1). */
int orc_polyline_to_bspline3(const double* corners, int W, int dim,
                             double radius, double* ctrl);
/* FitSplineToWaypoints: ctrl [3W-2][dim], knots [3W-2+3]; returns num ctrl. */
int orc_joint_fit(const double* waypoints, int W, int dim, double rounding,
                  double* ctrl, double* knots);
void orc_joint_sample(const double* knots, int num_knots, const double* ctrl,
                      int dim, int N, double path_start, double delta,
                      double* q, double* dq, double* ddq);
void orc_joint_constraints(const double* dq, const double* ddq, int N, int dim,
                           const double* vmax, const double* amax,
                           double safety, double* cons /* [N][2dim][4] */);

/* ---- stage C: Pfeiffer-Johanni time-optimal timing ---------------------- */
void orc_find_max_sd2_simplex(const double* cons_row /* [C][4] */, int C,
                              double* sd2max, double* sddmax, double* sd2zero);
typedef struct {
  /* outputs, all [N] and caller-allocated; any may be NULL */
  double* time;
  double* s;
  double* sd;
  double* sdd;
  double* sd2;
  /* limit curve after CalculateBoundary */
  double* b_sd2_max;
  double* b_sdd_max;
  double* b_sdd_min;
  double* b_sd2_zero;
  unsigned char* b_type;
  int last_extremal_index;
  int num_loops; /* switching-point loop iterations executed */
} orc_solution;
int orc_solve(int N, int C, const double* cons, double s_start, double s_end,
              double sd_start, double sdd_start, double time_start,
              int max_loops, orc_solution* out);
/* SolutionSatisfiesConstraints: number of violated rows. */
int orc_count_violations(int N, int C, const double* cons, const double* sd2,
                         const double* sdd);
/* GetPathParameterAndDerivatives on a solved profile; returns 1 on success. */
int orc_query(int N, const double* time, const double* s, const double* sd,
              const double* sd2, double s_start, double s_end, double t,
              double* s_out, double* sd_out, double* sdd_out);

/* ---- stage D: trajectory assembly + time resampling --------------------- */
int64_t orc_time_from_sec(double seconds);
double orc_time_to_sec(int64_t ns);
/* per-sample joint velocity / clamped acceleration. */
void orc_assemble(const double* dq, const double* ddq, const double* sd,
                  const double* sdd, const double* amax, int N, int dim,
                  double* vel, double* acc);
/* number of uniform samples ResampleEquidistantlyInTime would produce. */
int orc_resample_count(double t_end, double start_sec, double dt);
/* outputs: t[M], pos/vel/acc [M][dim], s/sd/sdd [M] */
int orc_resample_skip(int N, int dim, const double* time_ps, const double* s_ps,
                      const double* sd_ps, const double* sdd_ps, const double* pos_ps,
                      const double* vel_ps, const double* acc_ps, const double* amax,
                      double start_sec, double dt, double* t, double* pos, double* vel,
                      double* acc, double* s, double* sd, double* sdd);
void orc_resample_uniform(int N, int dim, const double* time_ps,
                          const double* s_ps, const double* sd_ps,
                          const double* sdd_ps, const double* pos_ps,
                          const double* vel_ps, const double* acc_ps,
                          const double* amax, double start_sec, double dt,
                          int M, double* t, double* pos, double* vel,
                          double* acc, double* s, double* sd, double* sdd);

/* Whole pipeline for one fresh joint path in a single window
 * (TimeableJointSplinePath::SetWaypoints + PathTimingTrajectory::Plan with a
 * path whose parameter range fits one window).  Caller provides scratch of
 * the sizes documented in tp_oracle.c.  Returns an ORC_* status. */
typedef struct {
  int N, dim, W;
  double rounding, safety, delta; /* delta <= 0: use knots.back()/(N-1) */
  double start_sec;
  double dt; /* <= 0: skip resampling */
  int skip_method; /* != 0: kSkipSamplesCloserThanTimeStep instead of kUniformlyInTime */
} orc_plan_cfg;
typedef struct {
  /* [N] */
  double *time, *s, *sd, *sdd;
  /* [N][dim] */
  double *q, *dq, *ddq, *vel, *acc;
  double duration;
  double delta_used;
  double knot_end;
  int last_extremal_index;
  /* resampled, capacity M_cap rows; M written */
  int M_cap, M;
  double *r_t, *r_pos, *r_vel, *r_acc, *r_s, *r_sd, *r_sdd;
} orc_plan_out;
int orc_plan_joint(const orc_plan_cfg* cfg, const double* waypoints,
                   const double* vmax, const double* amax, orc_plan_out* out);

/* ---- Cartesian timeable path (BASELINE config 5), tp_oracle_cart.c ------- */
/* quaternions: 4 doubles (x, y, z, w); poses: 7 doubles (tx,ty,tz,qx,qy,qz,qw) */
void orc_quat_mul(const double* a, const double* b, double* out);
void orc_quat_log(const double* q, double* out);
void orc_quat_exp(const double* q, double* out);
void orc_quat_power(const double* q, double power, double* out);
int orc_quat_eval_curve(const double* knots, int num_knots, const double* quats,
                        int degree, double u, double* out);
void orc_quat_power_derivative(const double* q, double power, double* out);
/* BSplineQ::EvalCurveAndDerivative (splines/bsplineq.cc:246-307) */
int orc_quat_eval_curve_derivative(const double* knots, int num_knots, const double* quats,
                                   int degree, double u, double* out, double* dout);
int orc_pose_polyline_to_bspline3(const double* corners, int W, double translation_radius,
                                  double rotational_radius, double* out);
int orc_cart_fit(const double* pose_wp, const double* joint_wp, int W, int dim,
                 double translation_rounding, double rotation_rounding, double* pose_ctrl,
                 double* joint_ctrl, double* knots);
void orc_cart_sample_targets(const double* knots, int num_knots, const double* pose_ctrl,
                             const double* joint_ctrl, int dim, int N, double delta,
                             double* pose_targets /* [N+1][7] */,
                             double* joint_targets /* [N+1][dim] */);
void orc_cart_sample_targets_from(const double* knots, int num_knots, const double* pose_ctrl,
                                  const double* joint_ctrl, int dim, int N, double delta,
                                  int start_index, double* pose_targets,
                                  double* joint_targets);
void orc_cart_derivatives(const double* pos, int N, int dim, double delta, double* d1,
                          double* d2);
void orc_cart_constraints(const double* d1, const double* d2, const double* jac, int N,
                          int dim, const double* vmax, const double* amax, double safety,
                          double max_trans_vel, double max_rot_vel,
                          double* cons /* [N][2dim+2][4] */);
/* the reference test's fake kinematics (path_timing_trajectory_test.cc:548-587) */
void orc_exp_so3(const double* v, double* q);
void orc_log_so3(const double* q, double* v);
void orc_log_so3_derivative(const double* q, double* out /* [3][4] */);
void orc_fake_ik(const double* pose_targets, const double* joint_targets, int M, double* ik);
void orc_fake_jacobian(const double* joints, double* J /* [6][7] */);
typedef struct {
  int N, W;
  double translation_rounding, rotation_rounding, safety;
  double delta; /* <= 0: knots.back()/(N-1) */
  double start_sec, dt;
  double max_translational_velocity, max_rotational_velocity;
} orc_cart_cfg;
/* optional outputs (may be NULL): pose_ctrl [3W-2][7], knots [3W+1],
 * pose_targets [N+1][7], cons [N][16][4] */
int orc_plan_cart_fake(const orc_cart_cfg* cfg, const double* pose_wp, const double* joint_wp,
                       const double* vmax, const double* amax, orc_plan_out* out,
                       double* pose_ctrl_out, double* knots_out, double* pose_targets_out,
                       double* cons_out);

/* ---- RescaleTrajectoryBackwardToStop (rescale_to_stop.cc:35-161) ---------- */
/* sampled_trajectory.cc:24-44; 1 = valid */
int orc_sampled_trajectory_inputs_valid(int n, const double* times);
/* Trajectory [n][D] row-major (positions are not needed: the result's positions
 * are input rows switch_index .. switch_index+count-1).  out_* hold n rows of
 * capacity; rows 0..count-1 are the stopping trajectory.  count == 0: the input
 * already ends in a stop.  Returns ORC_OK or ORC_FAIL_SETUP (invalid input). */
int orc_rescale_to_stop(int n, int D, const double* max_acc, const double* times,
                        const double* vel, const double* acc, int* count, int* switch_index,
                        double* out_times, double* out_vel, double* out_acc);

/* ---- ComputeFastestStop / GetPathStopParameter (path_timing_trajectory.cc:75-172,
 * 235-287).  Restated; pinned through the reference test's properties only. */
int orc_fastest_stop(int n, int D, const double* stop_acc, const double* times,
                     const double* vel, const double* acc, double* total_duration,
                     double* series_time, double* series_rate2, double* series_diff,
                     int* num_series);
int orc_path_stop_parameter(int M, int D, const double* time, const double* path_parameter,
                            const double* vel, const double* acc, const double* max_acc,
                            double query_time, double* stop_parameter, int* stop_row,
                            double* stop_duration);

/* ---- TrajectoryBuffer (trajectory_buffer.cc:40-393; tp_oracle_tbuf.c) -------
 * One growable buffer of [n][D] row-major rows.  Return values are
 * absl::StatusCode numbers.  which: 0 position, 1 velocity, 2 acceleration. */
typedef struct orc_tbuf orc_tbuf;
int orc_tbuf_create(double timestep_tolerance, int D, orc_tbuf** out);
void orc_tbuf_destroy(orc_tbuf* b);
void orc_tbuf_clear(orc_tbuf* b);
int orc_tbuf_insert_segment(orc_tbuf* b, int n, const double* times, const double* pos,
                            const double* vel, const double* acc);
int orc_tbuf_append_sample(orc_tbuf* b, double time, const double* pos, const double* vel,
                           const double* acc);
void orc_tbuf_discard_before_sec(orc_tbuf* b, double time_sec);
void orc_tbuf_discard_before_ns(orc_tbuf* b, int64_t ns);
int orc_tbuf_positions_up_to_time_ns(const orc_tbuf* b, int64_t ns);
int orc_tbuf_get_at_time(const orc_tbuf* b, int which, double time_sec, double* out);
int orc_tbuf_stop_at_index(orc_tbuf* b, int index, const double* max_acc, double time_step);
int orc_tbuf_stop_before_time(orc_tbuf* b, double time_sec, const double* max_acc,
                              double time_step);
void orc_tbuf_add_offset(orc_tbuf* b, double offset);
int orc_tbuf_num_samples(const orc_tbuf* b);
int orc_tbuf_sequence_number(const orc_tbuf* b);
int64_t orc_tbuf_start_ns(const orc_tbuf* b);
int64_t orc_tbuf_end_ns(const orc_tbuf* b);
void orc_tbuf_rows(const orc_tbuf* b, double* times, double* pos, double* vel, double* acc);

/* ---- online re-planning geometry of a joint path (tp_oracle_switch.c):
 * BSplineT knot insertion / TruncateSplineAt / ExtendWithControlPoints
 * (splines/bspline.h:224-511), ProjectPointOnPath (path_tools.h:59-104),
 * TimeableJointSplinePath::SwitchToWaypointPath (timeable_path_joint_spline.cc:209-250).
 * Functions return absl::StatusCode numbers (0 = ok). */
#define ORC_MAX_SWITCH_WAYPOINTS 64
typedef struct {
  int degree, dim;
  int knot_capacity; /* knots_.size(); points_.size() = knot_capacity - degree - 1 */
  int num_knots, num_points;
  double* knots;  /* [knot_capacity] */
  double* points; /* [knot_capacity - degree - 1][dim] */
} orc_spline;
int orc_spline_insert_knot(orc_spline* s, double knot, int multiplicity);
int orc_spline_truncate(orc_spline* s, double u_end);
int orc_spline_extend(orc_spline* s, const double* pts, int n);
int orc_project_point_on_path(const double* waypoints, int W, int D, const double* point,
                              int* waypoint_index, double* distance_to_path,
                              double* line_parameter, double* projected_point);
int orc_joint_switch(orc_spline* s, double keep_path_until, const double* waypoints, int W,
                     double rounding, double* new_ctrl, int* num_new_ctrl);

/* ---- PathTimingTrajectory::Plan with its state between calls
 * (path_timing_trajectory.cc:307-475, 477-684) for a joint path that is not
 * modified after the first plan; see tp_oracle_windows.c.  All storage is the
 * caller's. */
typedef struct {
  /* configuration */
  int N, dim, W;
  double rounding, safety, delta; /* delta <= 0: knots.back()/(N-1) */
  double dt;                      /* time step */
  int skip_method;                /* != 0: kSkipSamplesCloserThanTimeStep */
  int max_planning_iterations;
  double max_initial_velocity_error;
  const double *waypoints, *vmax, *amax, *initial_velocity; /* [W][dim], 3 x [dim] */
  /* storage */
  orc_spline spline; /* knot_capacity = max(2 * (max(4,3W-2) + 3), 100)
                        (timeable_path_joint_spline.cc:262-268), storage the caller's */
  int P_cap;            /* capacity of the *_at_path_samples_ arrays */
  double *w_time, *w_s, *w_sd, *w_sdd; /* [P_cap] */
  double *w_q, *w_vel, *w_acc;         /* [P_cap][dim] */
  int M_cap;                           /* capacity of the trajectory */
  double *r_t, *r_pos, *r_vel, *r_acc, *r_s, *r_sd, *r_sdd;
  /* state */
  int num_points;
  double knot_end, delta_used;
  int len; /* path samples held */
  int M;   /* trajectory rows held */
  int initial_plan;
  int sampled; /* TimeablePath::State: 0 kNewPath, 1 kPathWasSampled, 2 kModifiedPath */
  int planned_to_end, target_reached;
  int64_t start_time, end_time, final_decel_start; /* absl::Time as Unix ns */
  double path_horizon, path_start, decel_time;
  int last_extremal_index;
  int loops;  /* planning windows computed by the last call */
  int erased; /* the last call found enough planned and only erased (:594-599) */
  /* Cartesian path with the reference test's fake kinematics (cart != 0; dim 7):
   * TimeableCartesianSplinePath behind the same planner.  `waypoints` are the joint
   * waypoints, spline.points the joint control points; rounding is the rotation
   * rounding. */
  int cart;
  double translation_rounding, max_translational_velocity, max_rotational_velocity;
  const double* pose_waypoints; /* [W][7] (tx,ty,tz,qx,qy,qz,qw) */
  double* pose_ctrl;            /* [max(4,3W-2)][7] storage */
} orc_planner;
int orc_planner_init(orc_planner* p);
int orc_planner_plan(orc_planner* p, int64_t start_ns, int64_t horizon_ns);
/* SwitchToWaypointPath on the planner's path (SetInitialVelocity = write through the
 * initial_velocity pointer); returns an absl::StatusCode number. */
int orc_planner_switch(orc_planner* p, double keep_path_until, const double* waypoints, int W);
/* TimeableCartesianSplinePath::SwitchToWaypointPath (timeable_path_cartesian_spline.cc:183-296);
 * pose_ctrl [joint->knot_capacity - 3][7].  tp_oracle_cart.c */
int orc_cart_switch(orc_spline* joint, double* pose_ctrl, double keep_path_until,
                    const double* pose_waypoints, const double* joint_waypoints, int W,
                    double translation_rounding, double rotation_rounding);
int orc_planner_cart_switch(orc_planner* p, double keep_path_until, const double* pose_waypoints,
                            const double* joint_waypoints, int W);

#ifdef __cplusplus
}
#endif
#endif /* ORACLE_TP_ORACLE_H_ */
