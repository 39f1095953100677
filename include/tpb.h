/* tpb.h -- C ABI of the B200-native batched path-timing engine.
 *
 * This is the drop-in boundary for the hot path of
 * theteamatx/x-edr-trajectory-planning (B-spline evaluation with derivatives,
 * joint-space timeable-path sampling, Pfeiffer-Johanni time-optimal path
 * timing, trajectory assembly / time resampling), batched over many
 * independent paths.  The reference has no FFI of its own: its seam is the C++
 * virtual interface TimeablePath (trajectory_planning/timeable_path.h:92-166)
 * plus the concrete classes named below.  The host C++ layer in
 * x-edr-trajectory-planning_b200/host/ re-creates those classes on top of
 * exactly these entry points; nothing else crosses the boundary.
 *
 * Conventions
 *  - All functions return an int32 equal to absl::StatusCode's numeric value
 *    (0 kOk, 3 kInvalidArgument, 9 kFailedPrecondition, 11 kOutOfRange,
 *    13 kInternal, ...), like the absl::Status the reference returns.
 *  - Every array argument is plain double / int32 memory, row-major.  The
 *    `on_device` flag says whether the pointer is a CUDA device pointer
 *    (consumed in place or copied device-to-device) or host memory (copied
 *    with cudaMemcpyAsync on the engine's stream; pinned memory makes that
 *    copy truly asynchronous).
 *  - An engine owns one device and one stream; all work is stream ordered and
 *    asynchronous until tpb_sync / a host-side getter.  Not thread safe; use
 *    one engine per thread / per GPU.
 *  - A failing path never poisons the batch: per-path results carry an int32
 *    status (TPB_PATH_*), the call itself still returns kOk.
 *  - There is no CPU fallback: if no CUDA device is usable every call fails
 *    with kFailedPrecondition / kInternal.
 */
#ifndef TPB_H_
#define TPB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TPB_ABI_VERSION 4

/* absl::StatusCode values used by this library. */
enum {
  TPB_OK = 0,
  TPB_INVALID_ARGUMENT = 3,
  TPB_DEADLINE_EXCEEDED = 4,
  TPB_NOT_FOUND = 5,
  TPB_RESOURCE_EXHAUSTED = 8,
  TPB_FAILED_PRECONDITION = 9,
  TPB_OUT_OF_RANGE = 11,
  TPB_UNIMPLEMENTED = 12,
  TPB_INTERNAL = 13
};

/* Per-path solver status (TimeOptimalPathProfile::OptimizePathParameter
 * returns only bool; time_optimal_path_timing.cc:287-490 shows the causes). */
enum {
  TPB_PATH_OK = 0,
  TPB_PATH_FAIL_SETUP = 1,    /* SetupProblem/IsSetupValid false (:161-203, :554-576) */
  TPB_PATH_FAIL_CONNECT = 2,  /* iback_lo > icrit_lo (:383-391) */
  TPB_PATH_FAIL_NAN = 3,      /* sd2 still NaN after the loop (:400-403) */
  TPB_PATH_FAIL_TERMINAL = 4, /* non-zero terminal velocity (:422-428) */
  TPB_PATH_FAIL_INDEX = 5,    /* reference would read sd2_max[-1] (:372, UB there) */
  /* tpb_planner_plan (PathTimingTrajectory::Plan, path_timing_trajectory.cc) */
  TPB_PATH_FAIL_CAPACITY = 6,     /* spliced path samples exceed path_sample_capacity */
  TPB_PATH_FAIL_DEADLINE = 7,     /* kDeadlineExceeded: max planning loops (:652-655) */
  TPB_PATH_FAIL_START_RANGE = 8,  /* kOutOfRange: start past the previous plan (:504-511) or
                                     before its first sample (:294-296) */
  TPB_PATH_FAIL_START_ARG = 9,    /* kInvalidArgument: start vs. previous start / end
                                     (:517-535), non-positive target duration (:314-318) */
  TPB_PATH_FAIL_START_END = 10,   /* kInternal: start >= the last trajectory time (:298-303) */
  TPB_PATH_FAIL_NO_SAMPLES = 11,  /* kFailedPrecondition: no samples yet (:291-293) */
  TPB_PATH_FAIL_INITIAL_VELOCITY = 12, /* kInvalidArgument: initial velocity not along the
                                          first path tangent (:377-385) */
  TPB_PATH_NOT_RUN = 100
};

/* Arrays that tpb_get / tpb_devptr can address.  B = current batch size,
 * N = num_samples, D = num_dofs, n = control points of a path. */
typedef enum {
  /* spline of each path: TimeableJointSplinePath::control_points_/knots_ */
  TPB_ARR_NUM_POINTS = 1, /* int32 [B] */
  TPB_ARR_CTRL = 2,       /* double [B][max_points][D] */
  TPB_ARR_KNOTS = 3,      /* double [B][max_points + 3] */
  TPB_ARR_DELTA = 4,      /* double [B] path sampling distance actually used */
  /* Cartesian paths: pose control points, rows (tx,ty,tz,qx,qy,qz,qw) */
  TPB_ARR_POSE_CTRL = 5,  /* double [B][max_points][7] */
  /* SamplePath outputs: GetPathPositionAt / GetFirst.. / GetSecond.. */
  TPB_ARR_Q = 10,   /* double [B][N][D] */
  TPB_ARR_DQ = 11,  /* double [B][N][D] */
  TPB_ARR_DDQ = 12, /* double [B][N][D] */
  /* Cartesian paths: what SamplePath hands to the path IK (IK indices -1..N-1,
   * timeable_path_cartesian_spline.cc:464-500) */
  TPB_ARR_POSE_TARGETS = 13,  /* double [B][N+1][7] */
  TPB_ARR_JOINT_TARGETS = 14, /* double [B][N+1][D] */
  /* TimeOptimalPathProfile outputs (time_optimal_path_timing.h:152-159) */
  TPB_ARR_TIME = 20, /* double [B][N] */
  TPB_ARR_S = 21,    /* double [B][N] */
  TPB_ARR_SD = 22,   /* double [B][N] */
  TPB_ARR_SDD = 23,  /* double [B][N] */
  TPB_ARR_SD2 = 24,  /* double [B][N] */
  /* limit curve (TimeOptimalPathProfile::Boundary, .h:225-255) */
  TPB_ARR_B_SD2_MAX = 30,  /* double [B][N] */
  TPB_ARR_B_SDD_MAX = 31,  /* double [B][N] */
  TPB_ARR_B_SDD_MIN = 32,  /* double [B][N] */
  TPB_ARR_B_SD2_ZERO = 33, /* double [B][N] */
  TPB_ARR_B_TYPE = 34,     /* uint8  [B][N]  (1 source, 2 sink, 4 trajectory) */
  /* constraint rows (A, B, lower, upper) of explicit / Cartesian problems */
  TPB_ARR_CONS = 35,       /* double [B][N][C][4] */
  /* per-path scalars */
  TPB_ARR_STATUS = 40,        /* int32  [B] TPB_PATH_* */
  TPB_ARR_DURATION = 41,      /* double [B] GetTotalDuration (NaN if failed) */
  TPB_ARR_LAST_EXTREMAL = 42, /* int32  [B] GetLastExtremalIndex */
  TPB_ARR_KNOT_END = 43,      /* double [B] knots.back() */
  /* resampled trajectory (TrajectoryPlanner getters, trajectory_planner.h:86-108),
   * ragged and packed: path b owns rows [R_OFFSET[b], R_OFFSET[b+1]) */
  TPB_ARR_R_OFFSET = 50, /* int64  [B+1] */
  TPB_ARR_R_TIME = 51,   /* double [total] */
  TPB_ARR_R_POS = 52,    /* double [total][D] */
  TPB_ARR_R_VEL = 53,    /* double [total][D] */
  TPB_ARR_R_ACC = 54,    /* double [total][D] */
  TPB_ARR_R_S = 55,      /* double [total] */
  TPB_ARR_R_SD = 56,     /* double [total] */
  TPB_ARR_R_SDD = 57,    /* double [total] */
  /* moving-horizon planner (tpb_planner_reset / tpb_planner_plan): the
   * *_at_path_samples_ arrays of PathTimingTrajectory (path_timing_trajectory.h),
   * P = path_sample_capacity slots per path of which W_LEN[b] are held; W_DQ / W_DDQ
   * are the path derivatives (velocity = W_DQ * W_SD etc., :458-472) */
  TPB_ARR_W_LEN = 60,  /* int32  [B] */
  TPB_ARR_W_TIME = 61, /* double [B][P] */
  TPB_ARR_W_S = 62,    /* double [B][P] */
  TPB_ARR_W_SD = 63,   /* double [B][P] */
  TPB_ARR_W_SDD = 64,  /* double [B][P] */
  TPB_ARR_W_Q = 65,    /* double [B][P][D] */
  TPB_ARR_W_DQ = 66,   /* double [B][P][D] */
  TPB_ARR_W_DDQ = 67,  /* double [B][P][D] */
  /* ... and the planner's state after the last tpb_planner_plan */
  TPB_ARR_PLAN_STATUS = 70,            /* int32 [B] TPB_PATH_* of the last call */
  TPB_ARR_PLAN_LOOPS = 71,             /* int32 [B] planning windows computed by it */
  TPB_ARR_PLAN_PLANNED_TO_END = 72,    /* int32 [B] planned_to_end_ */
  TPB_ARR_PLAN_TARGET_REACHED = 73,    /* int32 [B] IsTrajectoryAtEnd() */
  TPB_ARR_PLAN_START_TIME = 74,        /* int64 [B] GetStartTime(), Unix ns */
  TPB_ARR_PLAN_END_TIME = 75,          /* int64 [B] GetEndTime(), Unix ns */
  TPB_ARR_PLAN_FINAL_DECEL_START = 76, /* int64 [B] GetFinalDecelStart(), Unix ns */
  TPB_ARR_PLAN_ERASED = 77,            /* int32 [B] 0: the trajectory was resampled from the
                                          start time, 1: rows kept (enough was planned, only
                                          EraseTrajectoryBefore ran; or the call failed its
                                          argument checks), 2: no rows (failed while planning) */
  TPB_ARR_SWITCH_STATUS = 80, /* int32 [B] absl::StatusCode of the last
                                 tpb_switch_to_waypoint_path per path */
  /* diagnostics: solver scratch [B][N]; holds per-path phase cycle stamps when the
   * library is built with -DTPB_PROFILE_PHASES, otherwise unspecified */
  TPB_ARR_DEBUG = 99
} tpb_array;

typedef struct tpb_engine tpb_engine;

typedef struct {
  int32_t device;          /* CUDA device ordinal */
  int32_t max_paths;       /* batch capacity B */
  int32_t num_dofs;        /* D (JointPathOptions::num_dofs); 0 for constraint-only use */
  int32_t max_waypoints;   /* per path; 0 for constraint-only use */
  int32_t num_samples;     /* N (num_path_samples) */
  int32_t max_constraints; /* C capacity for tpb_set_constraints (<= 64); 0 -> 2*D only */
  int64_t max_time_samples_total; /* capacity of the packed resampled output; 0 -> none */
  int32_t cartesian;       /* != 0: allocate the Cartesian-path buffers (needs num_dofs > 0;
                              the constraint capacity is raised to 2*D + 2) */
} tpb_config;

/* ---- life cycle ---------------------------------------------------------- */
int32_t tpb_abi_version(void);
int32_t tpb_create(const tpb_config* cfg, tpb_engine** out);
int32_t tpb_destroy(tpb_engine* e);
/* Human-readable text of the last non-OK return on this engine ("" if none). */
const char* tpb_last_error(const tpb_engine* e);
/* Use an existing cudaStream_t (e.g. torch's current stream); NULL = own stream. */
int32_t tpb_set_stream(tpb_engine* e, void* cuda_stream);
int32_t tpb_sync(tpb_engine* e);
/* Number of kernels this engine has launched since creation (for accounting). */
int64_t tpb_kernel_launches(const tpb_engine* e);

/* ---- joint-space paths: TimeableJointSplinePath -------------------------- */
/* PathOptions::constraint_safety / rounding (timeable_path.h:44-90; defaults
 * 0.8 / 0.2). */
int32_t tpb_set_path_options(tpb_engine* e, double constraint_safety, double rounding);
/* SetWaypoints (timeable_path_joint_spline.cc:200-207) for `num_paths` paths.
 * waypoints: [num_paths][max_waypoints][D]; num_waypoints: [num_paths] or NULL
 * (all paths use max_waypoints).  Fits the degree-2 spline (FitSplineToWaypoints
 * :252-292).  Sets the batch size B = num_paths. */
int32_t tpb_set_joint_waypoints(tpb_engine* e, int32_t num_paths,
                                const int32_t* num_waypoints,
                                const double* waypoints, int32_t on_device);
/* SetMaxJointVelocity / SetMaxJointAcceleration (:84-108).  per_path = 1:
 * [B][D] each; per_path = 0: one [D] row shared by all paths. */
int32_t tpb_set_joint_limits(tpb_engine* e, const double* max_velocity,
                             const double* max_acceleration, int32_t per_path,
                             int32_t on_device);
/* PathOptions::delta_parameter per path.  delta == NULL: every path uses
 * knots.back() / (N - 1), i.e. the whole path is one planning window. */
int32_t tpb_set_sampling(tpb_engine* e, const double* delta, int32_t on_device);
/* SamplePath(path_start) + ConstraintSetup (:294-343) + the per-sample part of
 * CalculateBoundary (time_optimal_path_timing.cc:1366-1377).  path_start: [B]
 * or NULL (0). */
int32_t tpb_sample_paths(tpb_engine* e, const double* path_start, int32_t on_device);

/* ---- Cartesian paths: TimeableCartesianSplinePath -------------------------- */
/* Built-in device kinematics.  TPB_KIN_FAKE7 is the reference test's fake robot
 * (path_timing_trajectory_test.cc:548-587): translation = q[0:3], rotation =
 * ExpSO3(q[3:6]), q[6] copied from the joint target. */
enum { TPB_KIN_HOST = 0, TPB_KIN_FAKE7 = 1 };
/* CartesianPathOptions::translation_rounding (default 0.05) and
 * SetMaxCartesianVelocity (timeable_path_cartesian_spline.h:31-58, .cc:352-366).
 * The rotation rounding is PathOptions::rounding (tpb_set_path_options). */
int32_t tpb_set_cart_options(tpb_engine* e, double translation_rounding,
                             double max_translational_velocity,
                             double max_rotational_velocity);
/* SetWaypoints (.cc:166-188) + FitSplineToWaypoints (.cc:396-462).
 * pose_waypoints: [num_paths][max_waypoints][7] rows (tx,ty,tz,qx,qy,qz,qw);
 * joint_waypoints: [num_paths][max_waypoints][D]. */
int32_t tpb_set_cart_waypoints(tpb_engine* e, int32_t num_paths, const int32_t* num_waypoints,
                               const double* pose_waypoints, const double* joint_waypoints,
                               int32_t on_device);
/* The spline half of SamplePath(0) (.cc:464-500): fills TPB_ARR_POSE_TARGETS /
 * TPB_ARR_JOINT_TARGETS.  Only fresh paths sampled from path_start = 0 (one
 * planning window) are supported. */
int32_t tpb_cart_sample_targets(tpb_engine* e);
/* Path IK with a built-in device kinematics model: path positions
 * TPB_ARR_Q[b][i] = IK(target i + 1) (.cc:502-513). */
int32_t tpb_cart_solve_ik(tpb_engine* e, int32_t model);
/* Path IK done by the caller (a user PathIKFunc on the host): upload
 * path_ik_positions [B][N][D] (= new_ik_path[1:]). */
int32_t tpb_cart_set_ik_positions(tpb_engine* e, const double* positions, int32_t on_device);
/* The rest of SamplePath (ComputePathDerivatives, .cc:39-68) + ConstraintSetup
 * (.cc:551-595) + the per-sample limit curve.  jacobians: [B][N][6][D] row-major
 * (rows 0-2 linear, 3-5 angular) from the caller's JacobianFunc, or NULL to use
 * the built-in `model`. */
int32_t tpb_cart_setup_constraints(tpb_engine* e, int32_t model, const double* jacobians,
                                   int32_t on_device);

/* ---- explicit constraints: TimeOptimalPathProfile ------------------------ */
/* InitSolver + SetupProblem (time_optimal_path_timing.cc:135-203) for
 * `num_paths` problems of N samples and C rows each.
 * constraints: [num_paths][N][C][4] = (A, B, lower, upper) of
 * lower <= A*sdd + B*sd^2 <= upper.  s_start/s_end: [num_paths].
 * Also evaluates the per-sample limit curve. */
int32_t tpb_set_constraints(tpb_engine* e, int32_t num_paths, int32_t num_constraints,
                            const double* constraints, const double* s_start,
                            const double* s_end, int32_t on_device);

/* ---- stage C: OptimizePathParameter (:287-490) --------------------------- */
/* sd_start, sdd_start, time_start: [B] each or NULL (0).  max_loops <= 0:
 * max(100, 10 N) as PathTimingTrajectory sets it (path_timing_trajectory.cc:398-400). */
int32_t tpb_optimize(tpb_engine* e, const double* sd_start, const double* sdd_start,
                     const double* time_start, int32_t max_loops, int32_t on_device);
/* Device-side timing of tpb_optimize's three phases (diagnostics / benchmarks).
 * With timing on, every tpb_optimize records CUDA events on the engine's stream
 * around (0) the setup phase -- SetSetupDone + CalculateBoundary passes 2-4
 * (time_optimal_path_timing.cc:535-552, 1379-1484), (1) the extremal sweeps and
 * the switching loop (:300-396, 769-952) and (2) the final audits and time
 * integration (:398-467).  tpb_last_phase_ms waits for the last timed call and
 * writes the three durations in milliseconds.  Paths that run all three phases
 * in one kernel (explicit constraint tables, more than 8 dofs) report the whole
 * kernel as phase 1 and zero for the others. */
int32_t tpb_time_phases(tpb_engine* e, int32_t on);
int32_t tpb_last_phase_ms(tpb_engine* e, double* ms3);

/* ---- queries on the solved profiles ------------------------------ */
/* TimeOptimalPathProfile::SolutionSatisfiesConstraints (time_optimal_path_timing.cc:492-518)
 * for every path of the batch, on the device: violations[b] = number of (sample, row) pairs
 * whose value A * sdd + B * sd^2 leaves [lower, upper] by more than kTiny (0 = the reference's
 * OkStatus, > 0 = its kNotFound "Number of constraint violations: n"), -1 for a path without
 * a valid solution (its kFailedPrecondition).  Joint paths use the rows ConstraintSetup would
 * build from q', q''; explicit-row and Cartesian problems the stored rows. */
int32_t tpb_count_violations(tpb_engine* e, int32_t* violations, int32_t on_device);

/* TimeOptimalPathProfile::GetPathParameterAndDerivatives, SampleIndexFromTime,
 * GetPreviousIndex and GetPreviousDiscreteValues (time_optimal_path_timing.cc:
 * 1497-1668) for queries_per_path times per path, on the device, after tpb_optimize
 * (any mode).  time_sec and every output are [B][queries_per_path]; any output may be
 * NULL.  s / sd / sdd: the path parameter and its derivatives at time_sec;
 * previous_index: GetPreviousIndex (-1 where the reference returns -1); sk, sdk, sddk,
 * tk: GetPreviousDiscreteValues.  status: bit 0 = GetPathParameterAndDerivatives returned
 * false, bit 1 = GetPreviousIndex < 0 (3 for every query of a path that was not timed);
 * outputs of a failed query are NaN / -1.  Results equal the reference's bit for bit. */
int32_t tpb_query_profile(tpb_engine* e, int32_t queries_per_path, const double* time_sec,
                          double* s, double* sd, double* sdd, int32_t* previous_index,
                          double* sk, double* sdk, double* sddk, double* tk, int32_t* status,
                          int32_t on_device);

/* Cross-check kernels: for a few stages the library keeps an older, simpler kernel next
 * to the default one; the parity tests run both and compare the bits.  `on` != 0 makes
 * this engine use the variant.  (No environment variables are read anywhere.) */
typedef enum {
  TPB_VARIANT_GENERIC_LIMIT_LOCAL = 0, /* explicit rows: limit curve with rows in local memory */
  TPB_VARIANT_INTEGRATE_WARP = 1,      /* joint paths: sweeps one path per warp (k_integrate) */
  TPB_VARIANT_CART_GENERIC = 2,        /* Cartesian paths: sweeps on the stored 2D+2 rows */
  TPB_VARIANT_CART_UNFUSED = 3         /* Cartesian paths: rows and limit curve in two kernels */
} tpb_kernel_variant;
int32_t tpb_set_kernel_variant(tpb_engine* e, tpb_kernel_variant which, int32_t on);

/* ---- stage D: trajectory assembly + ResampleEquidistantlyInTime ---------- */
/* (path_timing_trajectory.cc:458-472, :755-783).  start_sec = TimeToSec(start),
 * time_step_sec = the planner's time step.  Writes the packed TPB_ARR_R_* arrays;
 * *total_samples (may be NULL) receives R_OFFSET[B] (this synchronises). */
int32_t tpb_resample(tpb_engine* e, double start_sec, double time_step_sec,
                     int64_t* total_samples);
/* The same with TimeSamplingMethod::kSkipSamplesCloserThanTimeStep
 * (ResampleSkippingSamplesCloserThanTimeStep, path_timing_trajectory.cc:785-836):
 * row 0 is interpolated at start_sec, then every path sample whose time lies at
 * least GetMinTimeDeltaToKeep() = 0.95 * time_step_sec (:893-900) after the last
 * kept row is copied; the last kept row gets the final position and zero
 * derivatives.  A path yields at most N + 1 rows. */
int32_t tpb_resample_skipping(tpb_engine* e, double start_sec, double time_step_sec,
                              int64_t* total_samples);

/* ---- allocation-free uniform resampling ------------------------------------- */
/* NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The same kernels and results as tpb_resample
 * (ResampleEquidistantlyInTime), without an engine: every input and output is a
 * caller-provided DEVICE array, the call allocates nothing (no cudaMalloc / cudaFree,
 * no allocating library routine) and is asynchronous on cuda_stream.
 *
 * Scratch: the per-path row counts live in a file-scope `static __device__` array of
 * TPB_NOALLOC_MAX_PATHS = 65,536 int32 (256 KiB, held by every GPU that loads the
 * library until the process exits).  num_paths beyond that is rejected with
 * TPB_OUT_OF_RANGE and a message.  All calls share the array and it keeps the last
 * call's contents (every entry a call uses is overwritten first, so nothing has to be
 * cleared): calls that OVERLAP on one GPU -- different streams or host threads -- race
 * on it; serialise them.
 *
 * Rows beyond row_capacity are dropped; r_offset[num_paths] holds the number of rows
 * the batch needs.  error_message (may be NULL) receives a text for a non-zero return. */
#define TPB_NOALLOC_MAX_PATHS 65536
typedef struct {
  int32_t device;      /* CUDA device ordinal of all pointers below */
  int32_t num_paths;   /* B <= TPB_NOALLOC_MAX_PATHS */
  int32_t num_samples; /* N path samples per path */
  int32_t num_dofs;    /* D <= 16 */
  const double *time, *s, *sd, *sdd; /* [B][N]: the profile (TPB_ARR_TIME, _S, _SD, _SDD) */
  const double *q, *dq, *ddq;        /* [B][N][D]: SamplePath (TPB_ARR_Q, _DQ, _DDQ) */
  const double* max_acceleration;    /* [B][D] */
  const int32_t* status;             /* [B] TPB_PATH_*, paths != 0 get no rows; NULL: all ok */
  double start_sec, time_step_sec;
  int64_t row_capacity;              /* rows the r_* arrays hold */
  int64_t* r_offset;                 /* out [B + 1] */
  double *r_time, *r_s, *r_sd, *r_sdd; /* out [row_capacity] */
  double *r_pos, *r_vel, *r_acc;       /* out [row_capacity][D] */
} tpb_resample_desc;
int32_t tpb_resample_uniform_noalloc(const tpb_resample_desc* d, void* cuda_stream,
                                     char* error_message, size_t error_capacity);

/* ---- PathTimingTrajectory::Plan with its state between calls ----------------- */
/* (path_timing_trajectory.cc:477-684) for every joint path of the batch: the
 * moving-horizon loop over planning windows of num_path_samples samples
 * (ComputeTimingProfile :307-475), splicing of the per-path-sample arrays,
 * the "already planned enough" shortcut with EraseTrajectoryBefore (:540-577,
 * :594-599), the time argument checks (:503-538, :289-305) and the resampling
 * of the spliced samples, all on the device.  The paths are not modified between
 * calls (TimeablePath::State kNewPath for the first window, kPathWasSampled
 * afterwards); tpb_set_joint_waypoints / tpb_resample* end the planner's state.
 *
 * tpb_planner_reset = PathTimingTrajectory::Reset + the options of
 * PathTimingTrajectoryOptions (time step, sampling method, max planning loops,
 * max initial velocity error) + TimeablePath::SetInitialVelocity
 * (initial_velocity [B][num_dofs], NULL = zero).  path_sample_capacity >= N bounds
 * the spliced samples per path (device memory: (5 + 3 * num_dofs) * 8 bytes per slot).
 *
 * tpb_planner_plan = Plan(start, time_horizon), absl::Time / Duration as integer
 * nanoseconds.  Per-path outcome in TPB_ARR_PLAN_STATUS (the reference's Status
 * codes as TPB_PATH_*); the trajectory in TPB_ARR_R_* as after tpb_resample.
 * *total_samples (may be NULL) receives R_OFFSET[B] (this synchronises). */
int32_t tpb_planner_reset(tpb_engine* e, int32_t path_sample_capacity, double time_step_sec,
                          int32_t skip_method, int32_t max_planning_iterations,
                          double max_initial_velocity_error, const double* initial_velocity,
                          int32_t on_device);
/* Cartesian paths (tpb_set_cart_waypoints) behind the planner: every window runs
 * TimeableCartesianSplinePath::SamplePath (targets at the IK indices of the window,
 * path IK, finite differences; timeable_path_cartesian_spline.cc:464-549) and
 * ConstraintSetup (:551-595) on the device with the built-in kinematics `model`
 * (TPB_KIN_FAKE7).  With TPB_KIN_HOST (the default) the kinematics are the user's:
 * tpb_planner_set_host_kinematics. */
int32_t tpb_planner_set_kinematics(tpb_engine* e, int32_t model);
/* CartesianPathOptions::PathIKFunc / JacobianFunc (timeable_path_cartesian_spline.h:33-41)
 * as C callbacks, for Cartesian paths behind the planner.  Every planning window the
 * targets of the IK indices not solved yet come back to the host, `ik` is called once per
 * active path with them -- initial_value [D] is the last solution of the previous window
 * (the first joint target for a fresh path, :503-506); pose_targets [num_samples][7] as
 * (tx,ty,tz,qx,qy,qz,qw), joint_targets / solution [num_samples][D]; solution[0] belongs
 * to initial_value's index and is dropped, the rest extends the path's IK cache
 * (path_ik_positions_, :509-513) -- then `jacobian` once per path sample (jacobian
 * [6][D] row-major, rows 0-2 linear, 3-5 angular), and positions and Jacobians go back
 * to the device for the finite differences, ConstraintSetup and the timing.  A callback
 * returns TPB_OK or a tpb_status, which tpb_planner_plan then returns (the reference
 * returns the PathIKFunc's status from Plan).  The cache is dropped by tpb_planner_reset. */
typedef int32_t (*tpb_path_ik_fn)(void* user, int32_t path, const double* initial_value,
                                  int32_t num_samples, const double* pose_targets,
                                  const double* joint_targets, double* solution);
typedef int32_t (*tpb_jacobian_fn)(void* user, int32_t path, const double* joint_positions,
                                   double* jacobian);
int32_t tpb_planner_set_host_kinematics(tpb_engine* e, tpb_path_ik_fn ik,
                                        tpb_jacobian_fn jacobian, void* user);
int32_t tpb_planner_plan(tpb_engine* e, int64_t start_ns, int64_t horizon_ns,
                         int64_t* total_samples);
/* New waypoints were set (tpb_set_joint_waypoints) on a planner that keeps its plan:
 * every path is TimeablePath::State::kNewPath again (timed from its first sample at
 * the next plan, :322-327, all previous path samples replaced), the trajectory,
 * start / end times stay (no PathTimingTrajectory::Reset).  The batch size must not
 * have changed. */
int32_t tpb_planner_new_path(tpb_engine* e, const double* initial_velocity, int32_t on_device);
/* TimeablePath::SetInitialVelocity for every path ([B][num_dofs], NULL = zero): used by
 * the next window of a new or modified path (:352-385). */
int32_t tpb_planner_set_initial_velocity(tpb_engine* e, const double* initial_velocity,
                                         int32_t on_device);

/* ---- online re-planning geometry ------------------------------------------ */
/* TimeableJointSplinePath::SwitchToWaypointPath (timeable_path_joint_spline.cc:
 * 209-250) for every path b: the spline is truncated at keep_path_until[b]
 * (BSplineT::TruncateSplineAt, splines/bspline.h:402-427, through the knot insertion of
 * :244-278), the switching position is projected onto the new waypoints
 * (ProjectPointOnPath, path_tools.h:59-104), their control polygon
 * (PolyLineToBspline3Waypoints) is appended (ExtendWithControlPoints, :429-511) and
 * the path becomes TimeablePath::State::kModifiedPath: the planner's next
 * tpb_planner_plan re-times it from the sample before the start time with the initial
 * path velocity projected from initial_velocity (path_timing_trajectory.cc:329-385).
 * waypoints is [B][max_waypoints][num_dofs] like tpb_set_joint_waypoints;
 * num_waypoints[b] (NULL = max_waypoints each) <= 0 leaves path b as it is.
 * initial_velocity [B][num_dofs] (NULL = unchanged) = SetInitialVelocity; it is only
 * stored when a planner exists.  The spline keeps the reference's capacity
 * max(2 * knots at SetWaypoints, 100) knots (:257-268); per-path status (absl codes:
 * 0, 3, 9 capacity, 11 keep_path_until before the path) in TPB_ARR_SWITCH_STATUS.
 * Differences to the reference are confined to rounding: its TruncateSplineAt solves
 * the knot insertion equations with a QR decomposition, this one uses the reference's
 * own closed form (InsertKnotAndUpdateControlPointsRef). */
int32_t tpb_switch_to_waypoint_path(tpb_engine* e, const double* keep_path_until,
                                    const int32_t* num_waypoints, const double* waypoints,
                                    const double* initial_velocity, int32_t on_device);
/* TimeableCartesianSplinePath::SwitchToWaypointPath (timeable_path_cartesian_spline.cc:
 * 183-296), same conventions: the joint, translation and rotation splines are truncated at
 * keep_path_until[b] (the rotation spline's knot insertion is slerp of neighbouring control
 * quaternions, splines/bsplineq.cc:355-431, 497-519), the switching joint position and
 * translation are projected onto the new waypoints, the rotation of the projected first
 * waypoint is the slerp of the two pose waypoints around it, the three control polygons are
 * appended (BSplineQ::ExtendWithControlPoints :521-587 for the rotations) and the path
 * becomes kModifiedPath.  pose_waypoints [B][max_waypoints][7] (tx,ty,tz,qx,qy,qz,qw),
 * joint_waypoints [B][max_waypoints][num_dofs].  With user kinematics
 * (tpb_planner_set_host_kinematics) the IK caches are cut at PathIkIndex(keep_path_until)
 * like the reference's (:188-192).  Per-path status in TPB_ARR_SWITCH_STATUS. */
int32_t tpb_switch_to_cart_waypoint_path(tpb_engine* e, const double* keep_path_until,
                                         const int32_t* num_waypoints, const double* pose_waypoints,
                                         const double* joint_waypoints,
                                         const double* initial_velocity, int32_t on_device);

/* Whole pipeline for a batch of fresh joint paths with HOST buffers in and out:
 * set waypoints+limits, sample, optimize, resample, copy back durations,
 * status and (if out_* non-NULL) the packed resampled trajectory.  This is the
 * call a user of PathTimingTrajectory::Plan makes for a batch. */
typedef struct {
  int32_t num_paths;
  const int32_t* num_waypoints; /* [B] or NULL */
  const double* waypoints;      /* [B][max_waypoints][D] */
  const double* max_velocity;   /* [B][D] */
  const double* max_acceleration;
  double start_sec;
  double time_step_sec; /* <= 0: no resampling */
  /* outputs (host); any may be NULL */
  double* duration;      /* [B] */
  int32_t* status;       /* [B] */
  int64_t* r_offset;     /* [B+1] */
  double* r_pos;         /* [cap][D] */
  double* r_vel;         /* [cap][D] */
  double* r_acc;         /* [cap][D] */
  double* r_s;           /* [cap] */
  double* r_sd;          /* [cap] */
  double* r_sdd;         /* [cap] */
  int64_t r_capacity;    /* rows available in r_* */
  int64_t r_total;       /* out: rows written */
  /* ABI 4.  Every output pointer is optional: only the arrays a caller asks for cross
   * PCIe (a GetPositions() consumer needs 8 * D bytes per row instead of 8 * (3 D + 3),
   * a TimeOptimalPathProfile consumer 32 * N bytes per path and no rows at all). */
  double* r_time;        /* [cap] GetTime() */
  double* profile_time;  /* [B][N] GetTimeSamples() of the profile */
  double* profile_s;     /* [B][N] GetPathParameter() */
  double* profile_sd;    /* [B][N] GetPathVelocity() */
  double* profile_sdd;   /* [B][N] GetPathAcceleration() */
} tpb_plan_io;
int32_t tpb_plan_joint_host(tpb_engine* e, tpb_plan_io* io);
/* ---- multi-GPU front-end (csrc/tpb_multi.cu) ------------------------------- */
/* The reference plans one path per PathTimingTrajectory::Plan call
 * (path_timing_trajectory.cc:579-684) and paths are independent, so a batch shards
 * over the GPUs of one box with no exchange step.  Shard g of G owns the
 * contiguous block [g * ceil(B / G), (g + 1) * ceil(B / G)) of the batch
 * (tpb_shard_range; processes that own one GPU each use the same rule), has its own
 * engine and stream and is driven by its own host thread.  cfg->max_paths and
 * cfg->max_time_samples_total are totals over all shards (every shard gets
 * ceil(total / G)); cfg->device is ignored.  devices[g] = CUDA ordinal of shard g
 * (NULL: g modulo the devices present; an ordinal may repeat = logical shards of one
 * GPU).  No collective on the data path: every shard writes its own block of the
 * caller's arrays (tpb_multi_plan_joint_host) or keeps its results on its GPU
 * (tpb_multi_plan_joint_device + tpb_multi_shard / tpb_multi_gather). */
typedef struct tpb_multi tpb_multi;
/* Page-locked host memory (cudaMallocHost / cudaFreeHost) for callers that do not
 * include the CUDA headers: outputs written there leave the GPU at the full PCIe rate. */
int32_t tpb_host_alloc(size_t bytes, void** ptr);
int32_t tpb_host_free(void* ptr);
int32_t tpb_shard_range(int32_t num_paths, int32_t num_shards, int32_t shard, int32_t* first,
                        int32_t* count);
int32_t tpb_multi_create(const tpb_config* cfg, const int32_t* devices, int32_t num_shards,
                         tpb_multi** out);
int32_t tpb_multi_destroy(tpb_multi* m);
const char* tpb_multi_last_error(const tpb_multi* m);
int32_t tpb_multi_num_shards(const tpb_multi* m);
int64_t tpb_multi_kernel_launches(const tpb_multi* m);
/* Where shard `shard` of the last batch lives: device, paths [first_path, first_path +
 * num_paths), rows [first_row, first_row + num_rows) of the packed output, its engine
 * (for tpb_devptr / tpb_tbuf_insert_segments consumers).  Any pointer may be NULL. */
int32_t tpb_multi_shard(const tpb_multi* m, int32_t shard, int32_t* device, int32_t* first_path,
                        int32_t* num_paths, int64_t* first_row, int64_t* num_rows,
                        tpb_engine** engine);
/* tpb_set_joint_waypoints + tpb_set_joint_limits (per path) from HOST arrays of the
 * whole batch, scattered to the shards. */
int32_t tpb_multi_set_joint_paths(tpb_multi* m, int32_t num_paths, const int32_t* num_waypoints,
                                  const double* waypoints, const double* max_velocity,
                                  const double* max_acceleration);
/* tpb_plan_joint_device on every shard at once (returns when all are done);
 * *total_samples (may be NULL) = rows of all shards. */
int32_t tpb_multi_plan_joint_device(tpb_multi* m, double start_sec, double time_step_sec,
                                    int64_t* total_samples);
/* The optional final gather of a per-path array (TPB_ARR_DURATION, _STATUS, _TIME, _Q,
 * ...) of all shards into one host array in path order. */
int32_t tpb_multi_gather(tpb_multi* m, tpb_array what, void* dst, size_t bytes_per_path);
/* tpb_plan_joint_host for a batch spread over the shards: same tpb_plan_io, same
 * outputs (r_offset and the rows are packed in path order over all shards). */
int32_t tpb_multi_plan_joint_host(tpb_multi* m, tpb_plan_io* io);

/* The same hot path for the batch already resident on the device (after
 * tpb_set_joint_waypoints / tpb_set_joint_limits): SamplePath + ConstraintSetup +
 * OptimizePathParameter + (time_step_sec > 0) ResampleTrajectory from start_sec for
 * fresh paths, i.e. tpb_sample_paths(0) + tpb_optimize + tpb_resample, with the same
 * results.  num_chunks > 1 cuts the batch into that many chunks whose stages overlap
 * (sampling on the engine's stream, the latency-bound timing on high-priority
 * streams of its own, resampling on a third); the call is asynchronous on the
 * engine's stream (which waits for the others) unless total_samples is given. */
int32_t tpb_plan_joint_device(tpb_engine* e, double start_sec, double time_step_sec,
                              int32_t num_chunks, int64_t* total_samples);

/* ---- results ------------------------------------------------------------- */
/* Copy array `what` (current batch) to dst (host or device); bytes must be at
 * least the array's size.  Host copies synchronise the stream. */
int32_t tpb_get(tpb_engine* e, tpb_array what, void* dst, size_t bytes, int32_t to_device);
/* Device pointer + size of array `what` for zero-copy consumers. */
int32_t tpb_devptr(tpb_engine* e, tpb_array what, void** ptr, size_t* bytes);

/* ---- stage A stand-alone: dense B-spline evaluation ---------------------- */
/* BSplineT::EvalCurveAndDerivatives (splines/bspline.h:539-568) for
 * num_splines splines that share degree / num_points / dim, evals_per_spline
 * parameters each.  u == NULL: implicit grid u0[s] + i * du[s].
 * out: [num_splines][evals_per_spline][nder+1][dim].
 * status (may be NULL): int32 [num_splines][evals_per_spline], 0 or
 * kOutOfRange (11) when u is outside [knots.front(), knots.back()] (then the
 * output row is left untouched, like the reference).
 * exact != 0: NURBS-book A2.3 operation order (bit-identical to the oracle);
 * exact == 0: per-span polynomial form (<= 1e-12 absolute, faster); finished tiles leave
 *             shared memory as bulk asynchronous copies (the 1-D TMA path);
 * exact < 0:  the same values with per-thread stores (kept for A/B measurements). */
typedef struct {
  int32_t degree, num_points, dim, nder;
  int32_t num_splines;
  int64_t evals_per_spline;
  const double* knots; /* [num_splines][num_points + degree + 1] */
  const double* ctrl;  /* [num_splines][num_points][dim] */
  const double* u;     /* [num_splines][evals_per_spline] or NULL */
  const double* u0;    /* [num_splines] (grid mode) */
  const double* du;    /* [num_splines] (grid mode) */
  double* out;
  int32_t* status;
  int32_t exact;
} tpb_spline_eval_desc;
/* All pointers are device pointers; cuda_stream may be NULL (default stream). */
int32_t tpb_spline_eval_device(const tpb_spline_eval_desc* d, void* cuda_stream);
/* Same with host pointers (allocates, copies in, evaluates, copies out). */
int32_t tpb_spline_eval_host(const tpb_spline_eval_desc* d, int32_t device);

/* ---- stage A stand-alone: quaternion B-spline evaluation ------------------- */
/* BSplineQ::EvalCurve (splines/bsplineq.cc:223-244; dout == NULL) or
 * BSplineQ::EvalCurveAndDerivative (:246-307; dout != NULL) for num_splines quaternion
 * splines that share degree (1..5) and num_points, evals_per_spline parameters each;
 * u == NULL: implicit grid u0[s] + i * du[s].  Quaternions are (x, y, z, w).
 * out / dout: [num_splines][evals_per_spline][4] -- the curve value (normalised with a
 * non-negative real part for EvalCurve; the un-normalised product, sign-flipped when its
 * real part is negative, for EvalCurveAndDerivative, as in the reference) and d/du.
 * status (may be NULL): 0 or kOutOfRange (11) when u is outside the knot range (the
 * output rows are then left untouched).  Values agree with the reference to the
 * accuracy of the device's log / atan2 / sin / cos / exp (<= 1e-12 absolute). */
typedef struct {
  int32_t degree, num_points;
  int32_t num_splines;
  int64_t evals_per_spline;
  const double* knots; /* [num_splines][num_points + degree + 1] */
  const double* quats; /* [num_splines][num_points][4] */
  const double* u;     /* [num_splines][evals_per_spline] or NULL */
  const double* u0;    /* [num_splines] (grid mode) */
  const double* du;    /* [num_splines] (grid mode) */
  double* out;
  double* dout;
  int32_t* status;
} tpb_quat_spline_eval_desc;
/* All pointers are device pointers; cuda_stream may be NULL (default stream). */
int32_t tpb_quat_spline_eval_device(const tpb_quat_spline_eval_desc* d, void* cuda_stream);
/* Same with host pointers (allocates, copies in, evaluates, copies out). */
int32_t tpb_quat_spline_eval_host(const tpb_quat_spline_eval_desc* d, int32_t device);

/* Diagnostics: the FP64 FMA rate of `device` in TFLOP/s (2 flops per FMA), measured
 * with eight independent FMA chains per thread -- the denominator bench.py reports the
 * limit-curve and phase-plane kernels' FP64 FLOP/s against. */
int32_t tpb_measure_fp64_fma_peak(int32_t device, double* tflops);

/* ---- stop queries ----------------------------------------- */
/* Batched PathTimingTrajectory::GetPathStopParameter (path_timing_trajectory.cc:
 * 235-287) with ComputeFastestStop (:75-172) on the engine's resampled
 * trajectories (after tpb_resample / tpb_plan_joint_host): for path b, starting
 * at the first resampled row at or after time_sec[b], time is scaled forward
 * from rate 1 with the largest deceleration the acceleration limits admit until
 * the rate reaches 0.  stop_parameter[b] is the path parameter of the row where
 * that happens (the reference's return value), stop_row[b] that row's index in
 * the path's trajectory, stop_duration[b] the duration of the stop.  status[b] is
 * TPB_PATH_FAIL_SETUP when time_sec[b] lies after the trajectory (the reference
 * returns kInvalidArgument), TPB_PATH_OK otherwise.  Output pointers may be NULL.
 * Uses the acceleration limits as set (no safety factor), like the reference.
 * With host pointers the call returns when the results are in place. */
int32_t tpb_path_stop_parameter(tpb_engine* e, const double* time_sec, double* stop_parameter,
                                int32_t* stop_row, double* stop_duration, int32_t* status,
                                int32_t on_device);

/* ---- stopping trajectories ------------------------------ */
/* Batched RescaleTrajectoryBackwardToStop (rescale_to_stop.cc:35-161, declared
 * in rescale_to_stop.h:31-35): for each sampled trajectory the fastest stop that
 * traces its tail without exceeding max_acceleration, by backward integration of
 * the squared time-scaling rate.  Trajectories are packed: trajectory b is rows
 * begin[b] .. begin[b]+length[b]-1 of times [rows], velocities and accelerations
 * [rows][num_dofs] (the layout of TPB_ARR_R_*; positions are not needed).
 *
 * Results: count[b] samples (0: the trajectory already ends in a stop,
 * rescale_to_stop.cc:60-63), which correspond to input samples switch_index[b] ..
 * length[b]-1 -- the stopping trajectory's positions ARE those input rows -- and
 * are written to the same rows of out_times / out_velocities / out_accelerations
 * (rows begin[b]+switch_index[b] ..; other rows are left untouched).  status[b] is
 * TPB_PATH_OK or TPB_PATH_FAIL_SETUP (AreInputsValidForSampledTrajectory,
 * sampled_trajectory.cc:24-44: fewer than two samples, times not strictly
 * increasing).  num_dofs <= 16. */
typedef struct {
  int32_t num_trajectories;
  int32_t num_dofs;
  int64_t total_rows;             /* rows of the packed arrays */
  const int64_t* begin;           /* [num_trajectories] */
  const int32_t* length;          /* [num_trajectories] */
  const double* times;            /* [total_rows] */
  const double* velocities;       /* [total_rows][num_dofs] */
  const double* accelerations;    /* [total_rows][num_dofs] */
  const double* max_acceleration; /* [num_trajectories][num_dofs] */
  int32_t* count;                 /* [num_trajectories] */
  int32_t* switch_index;          /* [num_trajectories] */
  int32_t* status;                /* [num_trajectories] */
  double* out_times;              /* [total_rows] */
  double* out_velocities;         /* [total_rows][num_dofs] */
  double* out_accelerations;      /* [total_rows][num_dofs] */
} tpb_stop_desc;
/* All pointers are device pointers; asynchronous on cuda_stream. */
int32_t tpb_rescale_to_stop_device(const tpb_stop_desc* d, void* cuda_stream);
/* All pointers are host pointers; copies in, runs on `device`, copies out, returns
 * when done. */
int32_t tpb_rescale_to_stop_host(const tpb_stop_desc* d, int32_t device);

/* ---- trajectory buffers ----------------------------------- */
/* A batch of TrajectoryBuffers (trajectory_buffer.h:30-184) resident on the
 * device: the consumer format of the resampled trajectories.  Buffer b holds
 * count[b] rows (time, position[D], velocity[D], acceleration[D]) at rows
 * head[b] .. head[b]+count[b]-1 of its capacity_rows slots; discarding in front
 * moves head[b], inserting moves the kept rows to the front only when the slots
 * behind them do not suffice.  Every operation is one batched call: per-buffer
 * arguments are arrays of num_buffers entries, per-buffer results are
 * absl::StatusCode values in status[] (may be NULL), and the call's own return
 * value only reports argument / CUDA errors.  A buffer that an operation's
 * "skip" value leaves out keeps its state and gets status TPB_OK.
 *
 * Differences from the reference, all on paths where it has no defined result:
 * rows beyond capacity_rows give TPB_RESOURCE_EXHAUSTED and leave the buffer
 * unchanged; StopAtIndex on a sample that stands still but is not the last one
 * (CHECK-abort at trajectory_buffer.cc:338) and on an empty buffer (out-of-bounds
 * read) give TPB_INTERNAL.  eigenmath::InterpolateLinear is un-vendored; it is
 * y0 + ((x-x0)/(x1-x0))*(y1-y0), y0 when x1 == x0 (unpinned). */
typedef struct tpb_tbuf tpb_tbuf;
typedef struct {
  int32_t device;            /* CUDA device ordinal */
  int32_t num_buffers;       /* B */
  int32_t num_dofs;          /* D; the stop operations need D <= 16 */
  int32_t capacity_rows;     /* slots per buffer */
  double timestep_tolerance; /* TrajectoryBufferOptions::timestep_tolerance (1e-6) */
} tpb_tbuf_config;
typedef enum {
  TPB_TBUF_TIMES = 0,         /* double [B][capacity_rows] */
  TPB_TBUF_POSITIONS = 1,     /* double [B][capacity_rows][D] */
  TPB_TBUF_VELOCITIES = 2,    /* double [B][capacity_rows][D] */
  TPB_TBUF_ACCELERATIONS = 3, /* double [B][capacity_rows][D] */
  TPB_TBUF_HEAD = 4,          /* int32 [B] */
  TPB_TBUF_COUNT = 5,         /* int32 [B] GetNumSamples() */
  TPB_TBUF_SEQUENCE = 6       /* int32 [B] GetSequenceNumber() */
} tpb_tbuf_array;
/* TrajectoryBuffer::Create (trajectory_buffer.cc:40-48): kFailedPrecondition for
 * a non-positive tolerance. */
int32_t tpb_tbuf_create(const tpb_tbuf_config* cfg, tpb_tbuf** out);
int32_t tpb_tbuf_destroy(tpb_tbuf* t);
const char* tpb_tbuf_last_error(const tpb_tbuf* t);
int32_t tpb_tbuf_set_stream(tpb_tbuf* t, void* cuda_stream);
int32_t tpb_tbuf_sync(tpb_tbuf* t);
int64_t tpb_tbuf_kernel_launches(const tpb_tbuf* t);
/* Reserve (:72-77): at least capacity_rows slots per buffer; rows, sequence numbers
 * and contents are kept (device-to-device), the device pointers change. */
int32_t tpb_tbuf_reserve(tpb_tbuf* t, int32_t capacity_rows);
/* Clear (:64-70) of the buffers with select[b] != 0 (NULL: all). */
int32_t tpb_tbuf_clear(tpb_tbuf* t, const int32_t* select, int32_t on_device);
/* InsertSegment (:80-132).  The segments are packed like TPB_ARR_R_*: segment b
 * is rows begin[b] .. begin[b]+length[b]-1 of times [total_rows] and positions /
 * velocities / accelerations [total_rows][D]; with on_device the engine's
 * resampled trajectories go in without leaving HBM (begin = TPB_ARR_R_OFFSET).
 * length[b] < 0 skips buffer b; length[b] == 0 only counts (:97-101). */
int32_t tpb_tbuf_insert_segments(tpb_tbuf* t, const int64_t* begin, const int32_t* length,
                                 int64_t total_rows, const double* times,
                                 const double* positions, const double* velocities,
                                 const double* accelerations, int32_t* status,
                                 int32_t on_device);
/* AppendSample (:134-148) to the buffers with select[b] != 0 (NULL: all): time
 * [B], positions / velocities / accelerations [B][D]. */
int32_t tpb_tbuf_append_samples(tpb_tbuf* t, const int32_t* select, const double* time,
                                const double* positions, const double* velocities,
                                const double* accelerations, int32_t* status,
                                int32_t on_device);
/* DiscardSegmentBefore(double) (:154-207); time_sec[b] NaN skips buffer b.  The
 * absl::Time overload is TimeToSec (time.h:27-30) on the host, then this. */
int32_t tpb_tbuf_discard_before(tpb_tbuf* t, const double* time_sec, int32_t on_device);
/* GetPositionsUpToTime (:209-226): num[b] leading rows lie before time_sec[b]. */
int32_t tpb_tbuf_positions_up_to_time(tpb_tbuf* t, const double* time_sec, int32_t* num,
                                      int32_t on_device);
/* Get{Position,Velocity,Acceleration}AtTime (:228-293), queries_per_buffer of them
 * per buffer: time_sec [B][Q]; outputs [B][Q][D], each may be NULL; status [B][Q]
 * (kFailedPrecondition for an empty buffer, kOutOfRange outside it; such rows of
 * the outputs are left untouched). */
int32_t tpb_tbuf_sample(tpb_tbuf* t, int32_t queries_per_buffer, const double* time_sec,
                        double* positions, double* velocities, double* accelerations,
                        int32_t* status, int32_t on_device);
/* StopAtIndex (:295-358): index [B] (INT32_MIN skips buffer b), max_acceleration
 * [B][D].  Runs the batched RescaleTrajectoryBackwardToStop on rows 0..index of
 * every selected buffer in place, then re-inserts the stopping segments. */
int32_t tpb_tbuf_stop_at_index(tpb_tbuf* t, const int32_t* index, const double* max_acceleration,
                               double time_step, int32_t* status, int32_t on_device);
/* StopBeforeTime(double) (:366-381); time_sec[b] NaN skips buffer b. */
int32_t tpb_tbuf_stop_before_time(tpb_tbuf* t, const double* time_sec,
                                  const double* max_acceleration, double time_step,
                                  int32_t* status, int32_t on_device);
/* AddOffsetToTimestamps(double) (:386-389), offset [B]. */
int32_t tpb_tbuf_add_offset(tpb_tbuf* t, const double* offset, int32_t on_device);
/* GetNumSamples / GetSequenceNumber and the first / last time stamp of every
 * buffer (NaN when empty; GetStartTime / GetEndTime are TimeFromSec of them, 0
 * for an empty buffer, :50-62).  Any pointer may be NULL. */
int32_t tpb_tbuf_get_state(tpb_tbuf* t, int32_t* num_samples, int32_t* sequence_number,
                           double* front_time, double* back_time, int32_t on_device);
/* Rows first_row .. first_row+num_rows-1 of one buffer (GetTimes / GetPositions /
 * GetVelocities / GetAccelerations); any pointer may be NULL. */
int32_t tpb_tbuf_get_rows(tpb_tbuf* t, int32_t buffer, int32_t first_row, int32_t num_rows,
                          double* times, double* positions, double* velocities,
                          double* accelerations, int32_t on_device);
/* Zero-copy access for device-side consumers. */
int32_t tpb_tbuf_devptr(tpb_tbuf* t, tpb_tbuf_array what, void** ptr, size_t* bytes);

#ifdef __cplusplus
}
#endif
#endif /* TPB_H_ */
