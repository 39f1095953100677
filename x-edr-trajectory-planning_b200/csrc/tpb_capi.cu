// C ABI of the engine (include/tpb.h): handle management, stream-ordered
// copies, stage sequencing.  No CPU implementation of any stage lives here: if
// CUDA is unavailable, every entry point fails.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <type_traits>
#include <vector>

#include "tpb_device.cuh"
#include "tpb_engine.h"
#include "tpb_nvtx.h"

namespace tpb {
int launch_spline_eval(const tpb_spline_eval_desc& d, cudaStream_t st, std::string* err);
}

struct tpb_engine {
  tpb::Dims dims;
  tpb::DeviceArrays a;
  tpb::Mode mode = tpb::Mode::kNone;
  int device = 0;
  int num_sms = 148;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  bool cartesian = false;
  double translation_rounding = 0.05;  // CartesianPathOptions (timeable_path_cartesian_spline.h:50)
  double max_trans_vel = 0.0, max_rot_vel = 0.0;
  bool have_targets = false, have_ik = false;
  double safety = 0.8;    // PathOptions::constraint_safety_ (timeable_path.h:80)
  double rounding = 0.2;  // PathOptions::rounding_ (timeable_path.h:82)
  bool auto_delta = true;
  bool have_waypoints = false, have_limits = false, sampled = false, optimized = false,
       resampled = false;
  std::string err;
  int64_t launches = 0;
  std::vector<void*> allocs;
  // tpb_plan_joint_host pipeline: device-to-host copies of finished chunks run on
  // their own stream while later chunks are computed
  static constexpr int kMaxChunksDecl = 16;
  cudaStream_t copy_stream = nullptr;
  // tpb_plan_joint_device: resampling stream, one high-priority stream per chunk for
  // the integrate stage, events between the stages
  cudaStream_t aux_stream = nullptr;
  cudaStream_t integ_stream[kMaxChunksDecl] = {};
  cudaEvent_t join_event = nullptr;
  cudaEvent_t chunk_sampled[kMaxChunksDecl] = {}, chunk_timed[kMaxChunksDecl] = {};
  static constexpr int kMaxChunks = 16;
  cudaEvent_t chunk_done[kMaxChunks] = {};
  cudaEvent_t phase_ev[4] = {};  // tpb_time_phases
  bool time_phases = false;
  int64_t* chunk_rows = nullptr;  // pinned host: rows up to the end of chunk c
  // tpb_path_stop_parameter scratch: [Bmax] each
  double *sq_query = nullptr, *sq_parameter = nullptr, *sq_duration = nullptr;
  int32_t *sq_row = nullptr, *sq_status = nullptr;
  // moving-horizon planner (tpb_planner_reset / tpb_planner_plan)
  tpb::PlannerArrays w;
  tpb::PlannerOptions wopt;
  std::vector<void*> w_big;  // the P_cap-sized arrays (reallocated when the capacity grows)
  bool planner_ready = false;
  int planner_kinematics = TPB_KIN_HOST;  // device kinematics of Cartesian planner windows
  // user kinematics of Cartesian planner windows (tpb_planner_set_host_kinematics) and the
  // reference's path_ik_positions_ cache, one per path: [ik index][D]
  tpb_path_ik_fn host_ik = nullptr;
  tpb_jacobian_fn host_jacobian = nullptr;
  void* host_user = nullptr;
  std::vector<std::vector<double>> ik_cache;
  std::vector<double> hk_pose, hk_joint, hk_start, hk_delta, hk_q, hk_jac, hk_solution;
  std::vector<int32_t> hk_active;
  int32_t* any_active_host = nullptr;  // pinned
  // tpb_switch_to_waypoint_path staging: [Bmax] / [Bmax][Wmax][D]
  double *sw_keep = nullptr, *sw_wp = nullptr, *sw_pose = nullptr;
  int32_t *sw_num = nullptr, *sw_status = nullptr, *sw_modified = nullptr;
  std::vector<void*> spline_big;  // ctrl / knots of the reference's knot capacity
  void* query_scratch = nullptr;  // tpb_query_profile with host arrays
  size_t query_scratch_bytes = 0;
};

namespace {

using tpb::Mode;

int32_t fail(tpb_engine* e, int32_t code, const std::string& msg) {
  if (e) e->err = msg;
  return code;
}

int32_t cuda_fail(tpb_engine* e, cudaError_t err, const char* what) {
  return fail(e, TPB_INTERNAL, std::string(what) + ": " + cudaGetErrorString(err));
}

#define TPB_CUDA(e, call)                                        \
  do {                                                           \
    cudaError_t err__ = (call);                                  \
    if (err__ != cudaSuccess) return cuda_fail(e, err__, #call); \
  } while (0)

template <typename T>
cudaError_t dev_alloc(tpb_engine* e, T** p, size_t count) {
  if (count == 0) count = 1;
  void* q = nullptr;
  // 16 bytes of slack: bulk copies fetch runs rounded up to 16-byte units (tpb_bulk.cuh)
  cudaError_t err = cudaMalloc(&q, count * sizeof(T) + 16);
  if (err != cudaSuccess) return err;
  e->allocs.push_back(q);
  *p = static_cast<T*>(q);
  return cudaSuccess;
}

// copy `bytes` into an engine buffer from host or device memory
cudaError_t copy_in(tpb_engine* e, void* dst, const void* src, size_t bytes, int on_device) {
  return cudaMemcpyAsync(dst, src, bytes,
                         on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                         e->stream);
}

cudaError_t fill_or_copy(tpb_engine* e, double* dst, const double* src, size_t count,
                         int on_device) {
  if (src == nullptr) return cudaMemsetAsync(dst, 0, count * sizeof(double), e->stream);
  return copy_in(e, dst, src, count * sizeof(double), on_device);
}

struct ArrayRef {
  void* ptr = nullptr;
  size_t bytes = 0;
};

bool lookup(tpb_engine* e, tpb_array what, ArrayRef* r) {
  const auto& d = e->dims;
  const auto& a = e->a;
  const size_t B = d.B, N = d.N, D = d.D;
  const size_t BN = B * N * sizeof(double);
  int64_t total = 0;
  auto packed_rows = [&]() -> size_t {
    if (!e->resampled) return 0;
    cudaMemcpyAsync(&total, a.r_offset + d.B, sizeof(int64_t), cudaMemcpyDeviceToHost,
                    e->stream);
    cudaStreamSynchronize(e->stream);
    return static_cast<size_t>(std::min<int64_t>(total, d.r_cap));
  };
  switch (what) {
    case TPB_ARR_NUM_POINTS: *r = {a.num_points, B * sizeof(int32_t)}; return true;
    case TPB_ARR_CTRL: *r = {a.ctrl, B * d.max_points * D * sizeof(double)}; return true;
    case TPB_ARR_KNOTS: *r = {a.knots, B * (d.max_points + 3) * sizeof(double)}; return true;
    case TPB_ARR_DELTA: *r = {a.delta, B * sizeof(double)}; return true;
    case TPB_ARR_POSE_CTRL: *r = {a.pose_ctrl, B * d.max_points * 7 * sizeof(double)}; return true;
    case TPB_ARR_POSE_TARGETS: *r = {a.pose_targets, B * (N + 1) * 7 * sizeof(double)}; return true;
    case TPB_ARR_JOINT_TARGETS: *r = {a.joint_targets, B * (N + 1) * D * sizeof(double)}; return true;
    case TPB_ARR_CONS: *r = {a.cons, BN * d.C * 4}; return true;
    case TPB_ARR_Q: *r = {a.q, BN * D}; return true;
    case TPB_ARR_DQ: *r = {a.dq, BN * D}; return true;
    case TPB_ARR_DDQ: *r = {a.ddq, BN * D}; return true;
    case TPB_ARR_TIME: *r = {a.time, BN}; return true;
    case TPB_ARR_S: *r = {a.s, BN}; return true;
    case TPB_ARR_SD: *r = {a.sd, BN}; return true;
    case TPB_ARR_SDD: *r = {a.sdd, BN}; return true;
    case TPB_ARR_SD2: *r = {a.sd2, BN}; return true;
    case TPB_ARR_B_SD2_MAX: *r = {a.b_sd2_max, BN}; return true;
    case TPB_ARR_B_SDD_MAX: *r = {a.b_sdd_max, BN}; return true;
    case TPB_ARR_B_SDD_MIN: *r = {a.b_sdd_min, BN}; return true;
    case TPB_ARR_B_SD2_ZERO: *r = {a.b_sd2_zero, BN}; return true;
    case TPB_ARR_B_TYPE: *r = {a.b_type, B * N}; return true;
    case TPB_ARR_STATUS: *r = {a.status, B * sizeof(int32_t)}; return true;
    case TPB_ARR_DURATION: *r = {a.duration, B * sizeof(double)}; return true;
    case TPB_ARR_LAST_EXTREMAL: *r = {a.last_extremal, B * sizeof(int32_t)}; return true;
    case TPB_ARR_KNOT_END: *r = {a.knot_end, B * sizeof(double)}; return true;
    case TPB_ARR_DEBUG: *r = {a.fix_val, BN}; return true;
    case TPB_ARR_R_OFFSET: *r = {a.r_offset, (B + 1) * sizeof(int64_t)}; return true;
    case TPB_ARR_R_TIME: *r = {a.r_time, packed_rows() * sizeof(double)}; return true;
    case TPB_ARR_R_POS: *r = {a.r_pos, packed_rows() * D * sizeof(double)}; return true;
    case TPB_ARR_R_VEL: *r = {a.r_vel, packed_rows() * D * sizeof(double)}; return true;
    case TPB_ARR_R_ACC: *r = {a.r_acc, packed_rows() * D * sizeof(double)}; return true;
    case TPB_ARR_R_S: *r = {a.r_s, packed_rows() * sizeof(double)}; return true;
    case TPB_ARR_R_SD: *r = {a.r_sd, packed_rows() * sizeof(double)}; return true;
    case TPB_ARR_R_SDD: *r = {a.r_sdd, packed_rows() * sizeof(double)}; return true;
    default: break;
  }
  if (what == TPB_ARR_SWITCH_STATUS) {
    if (e->sw_status == nullptr) return false;
    *r = {e->sw_status, B * sizeof(int32_t)};
    return true;
  }
  if (what >= TPB_ARR_W_LEN && what <= TPB_ARR_PLAN_ERASED) {
    if (!e->planner_ready) return false;
    const auto& w = e->w;
    const size_t BP = B * static_cast<size_t>(w.P_cap) * sizeof(double);
    const size_t Bi = B * sizeof(int32_t), Bl = B * sizeof(int64_t);
    switch (what) {
      case TPB_ARR_W_LEN: *r = {w.len, Bi}; return true;
      case TPB_ARR_W_TIME: *r = {w.w_time, BP}; return true;
      case TPB_ARR_W_S: *r = {w.w_s, BP}; return true;
      case TPB_ARR_W_SD: *r = {w.w_sd, BP}; return true;
      case TPB_ARR_W_SDD: *r = {w.w_sdd, BP}; return true;
      case TPB_ARR_W_Q: *r = {w.w_q, BP * D}; return true;
      case TPB_ARR_W_DQ: *r = {w.w_dq, BP * D}; return true;
      case TPB_ARR_W_DDQ: *r = {w.w_ddq, BP * D}; return true;
      case TPB_ARR_PLAN_STATUS: *r = {w.status, Bi}; return true;
      case TPB_ARR_PLAN_LOOPS: *r = {w.loops, Bi}; return true;
      case TPB_ARR_PLAN_PLANNED_TO_END: *r = {w.planned_to_end, Bi}; return true;
      case TPB_ARR_PLAN_TARGET_REACHED: *r = {w.target_reached, Bi}; return true;
      case TPB_ARR_PLAN_START_TIME: *r = {w.start_time, Bl}; return true;
      case TPB_ARR_PLAN_END_TIME: *r = {w.end_time, Bl}; return true;
      case TPB_ARR_PLAN_FINAL_DECEL_START: *r = {w.final_decel_start, Bl}; return true;
      case TPB_ARR_PLAN_ERASED: *r = {w.mode, Bi}; return true;
      default: break;
    }
  }
  return false;
}

bool is_cart_array(tpb_array w) {
  return w == TPB_ARR_POSE_CTRL || w == TPB_ARR_POSE_TARGETS || w == TPB_ARR_JOINT_TARGETS;
}

bool is_joint_array(tpb_array w) {
  return w == TPB_ARR_NUM_POINTS || w == TPB_ARR_CTRL || w == TPB_ARR_KNOTS ||
         w == TPB_ARR_DELTA || w == TPB_ARR_Q || w == TPB_ARR_DQ || w == TPB_ARR_DDQ ||
         w == TPB_ARR_KNOT_END || (w >= TPB_ARR_R_OFFSET && w <= TPB_ARR_R_SDD);
}

}  // namespace

extern "C" {

int32_t tpb_abi_version(void) { return TPB_ABI_VERSION; }

int32_t tpb_create(const tpb_config* cfg, tpb_engine** out) {
  if (cfg == nullptr || out == nullptr) return TPB_INVALID_ARGUMENT;
  *out = nullptr;
  if (cfg->max_paths < 1 || cfg->num_samples < 2) return TPB_INVALID_ARGUMENT;
  if (cfg->num_dofs < 0 || cfg->num_dofs > tpb::kMaxDofs) return TPB_OUT_OF_RANGE;
  if (cfg->max_constraints < 0 || cfg->max_constraints > tpb::kMaxRows) return TPB_OUT_OF_RANGE;
  if (cfg->num_dofs == 0 && cfg->max_constraints == 0) return TPB_INVALID_ARGUMENT;
  if (cfg->num_dofs > 0 && cfg->max_waypoints < 1) return TPB_INVALID_ARGUMENT;
  // CHECK(options_.num_path_samples() >= 3), timeable_path_joint_spline.cc:45
  if (cfg->num_dofs > 0 && cfg->num_samples < 3) return TPB_INVALID_ARGUMENT;
  if (cfg->cartesian) {
    if (cfg->num_dofs < 1) return TPB_INVALID_ARGUMENT;
    if (2 * cfg->num_dofs + 2 > tpb::kMaxRows) return TPB_OUT_OF_RANGE;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return TPB_FAILED_PRECONDITION;  // no CUDA device: there is no CPU fallback
  }
  if (cfg->device < 0 || cfg->device >= ndev) return TPB_INVALID_ARGUMENT;
  if (cudaSetDevice(cfg->device) != cudaSuccess) return TPB_INTERNAL;

  tpb_engine* e = new tpb_engine();
  e->device = cfg->device;
  e->cartesian = cfg->cartesian != 0;
  cudaDeviceGetAttribute(&e->num_sms, cudaDevAttrMultiProcessorCount, cfg->device);
  auto& d = e->dims;
  d.Bmax = cfg->max_paths;
  d.B = 0;
  d.D = cfg->num_dofs;
  d.Wmax = cfg->max_waypoints;
  d.max_points = (d.D > 0) ? std::max(3 * d.Wmax - 2, 4) : 0;
  d.N = cfg->num_samples;
  d.Cmax = cfg->max_constraints;
  if (cfg->cartesian) d.Cmax = std::max(d.Cmax, 2 * cfg->num_dofs + 2);
  d.r_cap = (d.D > 0) ? cfg->max_time_samples_total : 0;
  const size_t B = d.Bmax, N = d.N, D = d.D;
  auto& a = e->a;
  cudaError_t err = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
  e->own_stream = true;
#define TPB_ALLOC(field, count) \
  if (err == cudaSuccess) err = dev_alloc(e, &a.field, (count))
  if (D > 0) {
    TPB_ALLOC(waypoints, B * d.Wmax * D);
    TPB_ALLOC(num_wp, B);
    TPB_ALLOC(vmax, B * D);
    TPB_ALLOC(amax, B * D);
    TPB_ALLOC(delta_in, B);
    TPB_ALLOC(ctrl, B * d.max_points * D);
    TPB_ALLOC(knots, B * (d.max_points + 3));
    TPB_ALLOC(num_points, B);
    TPB_ALLOC(knot_cap, B);
    TPB_ALLOC(delta, B);
    TPB_ALLOC(knot_end, B);
    TPB_ALLOC(q, B * N * D);
    TPB_ALLOC(dq, B * N * D);
    TPB_ALLOC(ddq, B * N * D);
  }
  if (e->cartesian) {
    TPB_ALLOC(pose_wp, B * d.Wmax * 7);
    TPB_ALLOC(pose_ctrl, B * d.max_points * 7);
    TPB_ALLOC(pose_targets, B * (N + 1) * 7);
    TPB_ALLOC(joint_targets, B * (N + 1) * D);
  }
  if (d.Cmax > 0) {
    TPB_ALLOC(cons, B * N * d.Cmax * 4);
    TPB_ALLOC(s_start, B);
    TPB_ALLOC(s_end, B);
  }
  TPB_ALLOC(path_start, B);
  TPB_ALLOC(sd_start, B);
  TPB_ALLOC(sdd_start, B);
  TPB_ALLOC(time_start, B);
  TPB_ALLOC(b_raw_sd2_max, B * N);
  TPB_ALLOC(b_raw_sdd_max, B * N);
  TPB_ALLOC(b_raw_sdd_min, B * N);
  TPB_ALLOC(b_sd2_max, B * N);
  TPB_ALLOC(b_sdd_max, B * N);
  TPB_ALLOC(b_sdd_min, B * N);
  TPB_ALLOC(b_sd2_zero, B * N);
  TPB_ALLOC(b_at0, B * N);
  TPB_ALLOC(b_type, B * N);
  TPB_ALLOC(sd2, B * N);
  TPB_ALLOC(sdd, B * N);
  TPB_ALLOC(sd, B * N);
  TPB_ALLOC(s, B * N);
  TPB_ALLOC(time, B * N);
  TPB_ALLOC(fix_val, B * N);
  TPB_ALLOC(status, B);
  TPB_ALLOC(last_extremal, B);
  TPB_ALLOC(duration, B);
  if (d.r_cap > 0) {
    const size_t R = static_cast<size_t>(d.r_cap);
    TPB_ALLOC(r_count, B);
    TPB_ALLOC(r_offset, B + 1);
    TPB_ALLOC(r_time, R);
    TPB_ALLOC(r_pos, R * D);
    TPB_ALLOC(r_vel, R * D);
    TPB_ALLOC(r_acc, R * D);
    TPB_ALLOC(r_s, R);
    TPB_ALLOC(r_sd, R);
    TPB_ALLOC(r_sdd, R);
  }
#undef TPB_ALLOC
  if (err == cudaSuccess) err = cudaMemsetAsync(a.path_start, 0, B * sizeof(double), e->stream);
  if (err == cudaSuccess) err = cudaMemsetAsync(a.sd_start, 0, B * sizeof(double), e->stream);
  if (err == cudaSuccess) err = cudaMemsetAsync(a.sdd_start, 0, B * sizeof(double), e->stream);
  if (err == cudaSuccess) err = cudaMemsetAsync(a.time_start, 0, B * sizeof(double), e->stream);
  if (err != cudaSuccess) {
    fprintf(stderr, "tpb_create: %s\n", cudaGetErrorString(err));
    tpb_destroy(e);
    return err == cudaErrorMemoryAllocation ? 8 /* kResourceExhausted */ : TPB_INTERNAL;
  }
  *out = e;
  return TPB_OK;
}

int32_t tpb_destroy(tpb_engine* e) {
  if (e == nullptr) return TPB_OK;
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (void* p : e->allocs) cudaFree(p);
  for (void* p : e->w_big) cudaFree(p);
  for (void* p : e->spline_big) cudaFree(p);
  if (e->query_scratch) cudaFree(e->query_scratch);
  if (e->any_active_host) cudaFreeHost(e->any_active_host);
  if (e->aux_stream) {
    cudaStreamSynchronize(e->aux_stream);
    cudaStreamDestroy(e->aux_stream);
  }
  for (auto& st : e->integ_stream)
    if (st) {
      cudaStreamSynchronize(st);
      cudaStreamDestroy(st);
    }
  if (e->join_event) cudaEventDestroy(e->join_event);
  for (auto& ev : e->chunk_sampled)
    if (ev) cudaEventDestroy(ev);
  for (auto& ev : e->chunk_timed)
    if (ev) cudaEventDestroy(ev);
  if (e->copy_stream) {
    cudaStreamSynchronize(e->copy_stream);
    cudaStreamDestroy(e->copy_stream);
  }
  for (auto& ev : e->chunk_done)
    if (ev) cudaEventDestroy(ev);
  for (auto& ev : e->phase_ev)
    if (ev) cudaEventDestroy(ev);
  if (e->chunk_rows) cudaFreeHost(e->chunk_rows);
  if (e->own_stream && e->stream) cudaStreamDestroy(e->stream);
  delete e;
  return TPB_OK;
}

const char* tpb_last_error(const tpb_engine* e) { return e ? e->err.c_str() : ""; }

int32_t tpb_set_stream(tpb_engine* e, void* cuda_stream) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  if (e->own_stream && e->stream) cudaStreamDestroy(e->stream);
  if (cuda_stream == nullptr) {
    TPB_CUDA(e, cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    e->own_stream = true;
  } else {
    e->stream = static_cast<cudaStream_t>(cuda_stream);
    e->own_stream = false;
  }
  return TPB_OK;
}

int32_t tpb_sync(tpb_engine* e) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  TPB_CUDA(e, cudaStreamSynchronize(e->stream));
  return TPB_OK;
}

int64_t tpb_kernel_launches(const tpb_engine* e) { return e ? e->launches : 0; }

int32_t tpb_set_path_options(tpb_engine* e, double constraint_safety, double rounding) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  e->safety = constraint_safety;
  e->rounding = rounding;
  return TPB_OK;
}

int32_t tpb_set_joint_waypoints(tpb_engine* e, int32_t num_paths, const int32_t* num_waypoints,
                                const double* waypoints, int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (d.D == 0) return fail(e, TPB_FAILED_PRECONDITION, "engine was created without dofs");
  if (num_paths < 1 || num_paths > d.Bmax)
    return fail(e, TPB_OUT_OF_RANGE, "num_paths exceeds max_paths");
  if (waypoints == nullptr) return fail(e, TPB_INVALID_ARGUMENT, "waypoints is null");
  cudaSetDevice(e->device);
  if (num_waypoints != nullptr && !on_device) {
    for (int b = 0; b < num_paths; ++b) {
      if (num_waypoints[b] < 1 || num_waypoints[b] > d.Wmax)
        return fail(e, TPB_INVALID_ARGUMENT, "num_waypoints out of [1, max_waypoints]");
    }
  }
  d.B = num_paths;
  d.C = 2 * d.D;
  e->mode = Mode::kJoint;
  TPB_CUDA(e, copy_in(e, e->a.waypoints, waypoints,
                      sizeof(double) * static_cast<size_t>(d.B) * d.Wmax * d.D, on_device));
  if (num_waypoints != nullptr) {
    TPB_CUDA(e, copy_in(e, e->a.num_wp, num_waypoints, sizeof(int32_t) * d.B, on_device));
  } else {
    e->launches += tpb::launch_fill_i32(e->a.num_wp, d.Wmax, d.B, e->stream);
  }
  e->launches += tpb::launch_joint_fit(d, e->a, e->rounding, e->auto_delta, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  e->have_waypoints = true;
  e->sampled = e->optimized = e->resampled = false;
  e->planner_ready = false;
  return TPB_OK;
}

int32_t tpb_set_joint_limits(tpb_engine* e, const double* max_velocity,
                             const double* max_acceleration, int32_t per_path,
                             int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (d.D == 0) return fail(e, TPB_FAILED_PRECONDITION, "engine was created without dofs");
  if (max_velocity == nullptr || max_acceleration == nullptr)
    return fail(e, TPB_INVALID_ARGUMENT, "limits are null");
  cudaSetDevice(e->device);
  const size_t rows = per_path ? (d.B > 0 ? d.B : d.Bmax) : 1;
  const size_t row_bytes = sizeof(double) * d.D;
  if (per_path) {
    TPB_CUDA(e, copy_in(e, e->a.vmax, max_velocity, rows * row_bytes, on_device));
    TPB_CUDA(e, copy_in(e, e->a.amax, max_acceleration, rows * row_bytes, on_device));
  } else {
    // broadcast one row to every path
    std::vector<double> hv(d.D), ha(d.D);
    if (on_device) {
      TPB_CUDA(e, cudaMemcpyAsync(hv.data(), max_velocity, row_bytes, cudaMemcpyDeviceToHost, e->stream));
      TPB_CUDA(e, cudaMemcpyAsync(ha.data(), max_acceleration, row_bytes, cudaMemcpyDeviceToHost, e->stream));
      TPB_CUDA(e, cudaStreamSynchronize(e->stream));
    } else {
      std::memcpy(hv.data(), max_velocity, row_bytes);
      std::memcpy(ha.data(), max_acceleration, row_bytes);
    }
    std::vector<double> bv(static_cast<size_t>(d.Bmax) * d.D), ba(bv.size());
    for (int b = 0; b < d.Bmax; ++b) {
      std::memcpy(&bv[static_cast<size_t>(b) * d.D], hv.data(), row_bytes);
      std::memcpy(&ba[static_cast<size_t>(b) * d.D], ha.data(), row_bytes);
    }
    TPB_CUDA(e, cudaMemcpyAsync(e->a.vmax, bv.data(), bv.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    TPB_CUDA(e, cudaMemcpyAsync(e->a.amax, ba.data(), ba.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
  }
  e->have_limits = true;
  e->sampled = e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_set_sampling(tpb_engine* e, const double* delta, int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (d.D == 0) return fail(e, TPB_FAILED_PRECONDITION, "engine was created without dofs");
  cudaSetDevice(e->device);
  if (delta == nullptr) {
    e->auto_delta = true;
  } else {
    if (d.B == 0) return fail(e, TPB_FAILED_PRECONDITION, "set waypoints first");
    e->auto_delta = false;
    TPB_CUDA(e, copy_in(e, e->a.delta_in, delta, sizeof(double) * d.B, on_device));
    TPB_CUDA(e, copy_in(e, e->a.delta, delta, sizeof(double) * d.B, on_device));
  }
  if (e->have_waypoints && delta == nullptr) {
    // recompute the automatic delta for the current splines
    if (e->mode == Mode::kCartesian) {
      e->launches += tpb::launch_cart_fit(d, e->a, e->rounding, e->translation_rounding, true, e->stream);
    } else {
      e->launches += tpb::launch_joint_fit(d, e->a, e->rounding, true, e->stream);
    }
    TPB_CUDA(e, cudaGetLastError());
  }
  e->have_targets = e->have_ik = false;
  e->planner_ready = false;
  e->sampled = e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_sample_paths(tpb_engine* e, const double* path_start, int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (e->mode != Mode::kJoint || !e->have_waypoints)
    return fail(e, TPB_FAILED_PRECONDITION, "No path set");
  if (!e->have_limits) return fail(e, TPB_FAILED_PRECONDITION, "joint limits not set");
  cudaSetDevice(e->device);
  TPB_CUDA(e, fill_or_copy(e, e->a.path_start, path_start, d.B, on_device));
  e->launches += tpb::launch_joint_sample_limit(d, e->a, e->safety, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  e->sampled = true;
  e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_set_cart_options(tpb_engine* e, double translation_rounding,
                             double max_translational_velocity,
                             double max_rotational_velocity) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  if (!e->cartesian) return fail(e, TPB_FAILED_PRECONDITION, "engine was created without cartesian");
  // SetTranslationRounding / SetMaxCartesianVelocity argument checks
  // (timeable_path_cartesian_spline.cc:352-390)
  if (translation_rounding <= 0.0)
    return fail(e, TPB_INVALID_ARGUMENT, "translation_rounding needs to be greater than zero");
  if (max_translational_velocity <= 0 || max_rotational_velocity <= 0)
    return fail(e, TPB_INVALID_ARGUMENT, "Velocity limits must be positive");
  e->translation_rounding = translation_rounding;
  e->max_trans_vel = max_translational_velocity;
  e->max_rot_vel = max_rotational_velocity;
  e->sampled = e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_set_cart_waypoints(tpb_engine* e, int32_t num_paths, const int32_t* num_waypoints,
                               const double* pose_waypoints, const double* joint_waypoints,
                               int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (!e->cartesian) return fail(e, TPB_FAILED_PRECONDITION, "engine was created without cartesian");
  if (num_paths < 1 || num_paths > d.Bmax)
    return fail(e, TPB_OUT_OF_RANGE, "num_paths exceeds max_paths");
  if (pose_waypoints == nullptr || joint_waypoints == nullptr)
    return fail(e, TPB_INVALID_ARGUMENT, "waypoints are null");
  cudaSetDevice(e->device);
  if (num_waypoints != nullptr && !on_device) {
    for (int b = 0; b < num_paths; ++b) {
      if (num_waypoints[b] < 1 || num_waypoints[b] > d.Wmax)
        return fail(e, TPB_INVALID_ARGUMENT, "num_waypoints out of [1, max_waypoints]");
    }
  }
  d.B = num_paths;
  d.C = 2 * d.D + 2;
  e->mode = Mode::kCartesian;
  const size_t BW = static_cast<size_t>(d.B) * d.Wmax;
  TPB_CUDA(e, copy_in(e, e->a.waypoints, joint_waypoints, sizeof(double) * BW * d.D, on_device));
  TPB_CUDA(e, copy_in(e, e->a.pose_wp, pose_waypoints, sizeof(double) * BW * 7, on_device));
  if (num_waypoints != nullptr) {
    TPB_CUDA(e, copy_in(e, e->a.num_wp, num_waypoints, sizeof(int32_t) * d.B, on_device));
  } else {
    e->launches += tpb::launch_fill_i32(e->a.num_wp, d.Wmax, d.B, e->stream);
  }
  e->launches += tpb::launch_cart_fit(d, e->a, e->rounding, e->translation_rounding,
                                      e->auto_delta, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  e->have_waypoints = true;
  e->have_targets = e->have_ik = false;
  e->sampled = e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_cart_sample_targets(tpb_engine* e) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  if (e->mode != Mode::kCartesian || !e->have_waypoints)
    return fail(e, TPB_FAILED_PRECONDITION, "No path set");
  cudaSetDevice(e->device);
  TPB_CUDA(e, cudaMemsetAsync(e->a.path_start, 0, sizeof(double) * e->dims.B, e->stream));
  e->launches += tpb::launch_cart_sample_targets(e->dims, e->a, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  e->have_targets = true;
  e->have_ik = false;
  e->sampled = e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_cart_solve_ik(tpb_engine* e, int32_t model) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  if (e->mode != Mode::kCartesian || !e->have_targets)
    return fail(e, TPB_FAILED_PRECONDITION, "targets not sampled");
  cudaSetDevice(e->device);
  const int n = tpb::launch_cart_ik_builtin(e->dims, e->a, model, e->stream);
  if (n < 0) return fail(e, TPB_INVALID_ARGUMENT, "unknown kinematics model for this dof count");
  e->launches += n;
  TPB_CUDA(e, cudaGetLastError());
  e->have_ik = true;
  e->sampled = e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_cart_set_ik_positions(tpb_engine* e, const double* positions, int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  const auto& d = e->dims;
  if (e->mode != Mode::kCartesian || !e->have_waypoints)
    return fail(e, TPB_FAILED_PRECONDITION, "No path set");
  if (positions == nullptr) return fail(e, TPB_INVALID_ARGUMENT, "positions is null");
  cudaSetDevice(e->device);
  TPB_CUDA(e, copy_in(e, e->a.q, positions, sizeof(double) * static_cast<size_t>(d.B) * d.N * d.D,
                      on_device));
  e->have_ik = true;
  e->sampled = e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_cart_setup_constraints(tpb_engine* e, int32_t model, const double* jacobians,
                                   int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  const auto& d = e->dims;
  if (e->mode != Mode::kCartesian || !e->have_ik)
    return fail(e, TPB_FAILED_PRECONDITION, "path IK positions not set");
  if (!e->have_limits) return fail(e, TPB_FAILED_PRECONDITION, "joint limits not set");
  if (!(e->max_trans_vel > 0) || !(e->max_rot_vel > 0))
    return fail(e, TPB_FAILED_PRECONDITION, "Cartesian velocity limits not set");
  cudaSetDevice(e->device);
  if (jacobians != nullptr) {
    if (e->a.jac == nullptr)
      TPB_CUDA(e, dev_alloc(e, &e->a.jac, static_cast<size_t>(d.Bmax) * d.N * 6 * d.D));
    TPB_CUDA(e, copy_in(e, e->a.jac, jacobians,
                        sizeof(double) * static_cast<size_t>(d.B) * d.N * 6 * d.D, on_device));
  }
  const int n = tpb::launch_cart_constraints(d, e->a, e->safety, e->max_trans_vel, e->max_rot_vel,
                                             model, jacobians != nullptr, e->stream);
  if (n < 0) return fail(e, TPB_INVALID_ARGUMENT, "unknown kinematics model for this dof count");
  e->launches += n;  // rows and the per-sample limit curve
  TPB_CUDA(e, cudaGetLastError());
  e->sampled = true;
  e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_set_constraints(tpb_engine* e, int32_t num_paths, int32_t num_constraints,
                            const double* constraints, const double* s_start,
                            const double* s_end, int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (d.Cmax == 0) return fail(e, TPB_FAILED_PRECONDITION, "engine has no constraint capacity");
  if (num_paths < 1 || num_paths > d.Bmax) return fail(e, TPB_OUT_OF_RANGE, "num_paths");
  if (num_constraints < 1 || num_constraints > d.Cmax)
    return fail(e, TPB_OUT_OF_RANGE, "num_constraints exceeds max_constraints");
  if (!constraints || !s_start || !s_end) return fail(e, TPB_INVALID_ARGUMENT, "null input");
  cudaSetDevice(e->device);
  d.B = num_paths;
  d.C = num_constraints;
  e->mode = Mode::kGeneric;
  TPB_CUDA(e, copy_in(e, e->a.cons, constraints,
                      sizeof(double) * static_cast<size_t>(d.B) * d.N * d.C * 4, on_device));
  TPB_CUDA(e, copy_in(e, e->a.s_start, s_start, sizeof(double) * d.B, on_device));
  TPB_CUDA(e, copy_in(e, e->a.s_end, s_end, sizeof(double) * d.B, on_device));
  e->launches += tpb::launch_generic_limit(d, e->a, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  e->sampled = true;
  e->optimized = e->resampled = false;
  return TPB_OK;
}

int32_t tpb_optimize(tpb_engine* e, const double* sd_start, const double* sdd_start,
                     const double* time_start, int32_t max_loops, int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (!e->sampled) return fail(e, TPB_FAILED_PRECONDITION, "problem not defined/set up");
  cudaSetDevice(e->device);
  TPB_CUDA(e, fill_or_copy(e, e->a.sd_start, sd_start, d.B, on_device));
  TPB_CUDA(e, fill_or_copy(e, e->a.sdd_start, sdd_start, d.B, on_device));
  TPB_CUDA(e, fill_or_copy(e, e->a.time_start, time_start, d.B, on_device));
  if (max_loops <= 0) max_loops = std::max(100, 10 * d.N);
  e->launches += tpb::launch_integrate(d, e->a, e->mode, e->safety, max_loops, e->stream,
                                       e->time_phases ? e->phase_ev : nullptr);
  TPB_CUDA(e, cudaGetLastError());
  e->optimized = true;
  e->resampled = false;
  return TPB_OK;
}

int32_t tpb_time_phases(tpb_engine* e, int32_t on) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(e->device);
  if (on) {
    for (auto& ev : e->phase_ev)
      if (ev == nullptr) TPB_CUDA(e, cudaEventCreate(&ev));
  }
  e->time_phases = on != 0;
  return TPB_OK;
}

int32_t tpb_last_phase_ms(tpb_engine* e, double* ms3) {
  if (e == nullptr || ms3 == nullptr) return TPB_INVALID_ARGUMENT;
  if (!e->time_phases || !e->optimized)
    return fail(e, TPB_FAILED_PRECONDITION, "no timed tpb_optimize on this engine");
  cudaSetDevice(e->device);
  TPB_CUDA(e, cudaEventSynchronize(e->phase_ev[3]));
  for (int k = 0; k < 3; ++k) {
    float ms = 0.f;
    TPB_CUDA(e, cudaEventElapsedTime(&ms, e->phase_ev[k], e->phase_ev[k + 1]));
    ms3[k] = ms;
  }
  return TPB_OK;
}

static int32_t resample_impl(tpb_engine* e, double start_sec, double time_step_sec,
                             int64_t* total_samples, bool skip);

int32_t tpb_resample(tpb_engine* e, double start_sec, double time_step_sec,
                     int64_t* total_samples) {
  return resample_impl(e, start_sec, time_step_sec, total_samples, false);
}

int32_t tpb_resample_skipping(tpb_engine* e, double start_sec, double time_step_sec,
                              int64_t* total_samples) {
  return resample_impl(e, start_sec, time_step_sec, total_samples, true);
}

static int32_t resample_impl(tpb_engine* e, double start_sec, double time_step_sec,
                             int64_t* total_samples, bool skip) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if ((e->mode != Mode::kJoint && e->mode != Mode::kCartesian) || !e->optimized)
    return fail(e, TPB_FAILED_PRECONDITION, "no timed joint paths");
  if (d.r_cap <= 0) return fail(e, TPB_FAILED_PRECONDITION, "max_time_samples_total is 0");
  if (!(time_step_sec > 0)) return fail(e, TPB_INVALID_ARGUMENT, "time step must be positive");
  cudaSetDevice(e->device);
  e->launches += tpb::launch_resample(d, e->a, start_sec, time_step_sec, e->num_sms, e->stream,
                                      false, skip);
  TPB_CUDA(e, cudaGetLastError());
  e->resampled = true;
  e->planner_ready = false;  // the trajectory rows are no longer the planner's
  if (total_samples != nullptr) {
    int64_t total = 0;
    TPB_CUDA(e, cudaMemcpyAsync(&total, e->a.r_offset + d.B, sizeof(int64_t),
                                cudaMemcpyDeviceToHost, e->stream));
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
    *total_samples = total;
    if (total > d.r_cap)
      return fail(e, TPB_OUT_OF_RANGE, "resampled rows exceed max_time_samples_total");
  }
  return TPB_OK;
}

// The whole hot path for the resident batch, software-pipelined.  The sweeps of the
// integrate stage are latency-bound: a path's switching loop takes about as long alone
// as with 4,096 others, so they are worth starting early and hiding behind work that
// is bound by something else.  The batch is cut into chunks; chunk c is sampled on the
// engine's stream (one chunk after the other, each fills the GPU), integrated on its
// own HIGH-PRIORITY stream (its CTAs take the slots the next chunk's sampling CTAs
// free) and resampled on a third stream in chunk order (the continued offset scan).
int32_t tpb_plan_joint_device(tpb_engine* e, double start_sec, double time_step_sec,
                              int32_t num_chunks, int64_t* total_samples) {
  TPB_NVTX("tpb_plan_joint_device");
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (e->mode != Mode::kJoint || !e->have_waypoints)
    return fail(e, TPB_FAILED_PRECONDITION, "No path set");
  if (!e->have_limits) return fail(e, TPB_FAILED_PRECONDITION, "joint limits not set");
  const bool want_rows = time_step_sec > 0;
  if (want_rows && d.r_cap <= 0)
    return fail(e, TPB_FAILED_PRECONDITION, "max_time_samples_total is 0");
  cudaSetDevice(e->device);
  int nchunks = std::max(1, std::min<int>(num_chunks, tpb_engine::kMaxChunks));
  nchunks = std::min(nchunks, std::max(1, d.B / 64));
  TPB_CUDA(e, cudaMemsetAsync(e->a.path_start, 0, sizeof(double) * d.B, e->stream));
  TPB_CUDA(e, cudaMemsetAsync(e->a.sd_start, 0, sizeof(double) * d.B, e->stream));
  TPB_CUDA(e, cudaMemsetAsync(e->a.sdd_start, 0, sizeof(double) * d.B, e->stream));
  e->launches += tpb::launch_fill_f64(e->a.time_start, start_sec, d.B, e->stream);
  const int max_loops = std::max(100, 10 * d.N);
  if (nchunks == 1) {
    e->launches += tpb::launch_joint_sample_limit(d, e->a, e->safety, e->stream);
    e->launches += tpb::launch_integrate(d, e->a, e->mode, e->safety, max_loops, e->stream);
    if (want_rows)
      e->launches += tpb::launch_resample(d, e->a, start_sec, time_step_sec, e->num_sms,
                                          e->stream, false);
  } else {
    int lo_prio = 0, hi_prio = 0;
    TPB_CUDA(e, cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
    for (int c = 0; c < nchunks; ++c) {
      if (e->integ_stream[c] == nullptr)
        TPB_CUDA(e, cudaStreamCreateWithPriority(&e->integ_stream[c], cudaStreamNonBlocking,
                                                 hi_prio));
      if (e->chunk_sampled[c] == nullptr)
        TPB_CUDA(e, cudaEventCreateWithFlags(&e->chunk_sampled[c], cudaEventDisableTiming));
      if (e->chunk_timed[c] == nullptr)
        TPB_CUDA(e, cudaEventCreateWithFlags(&e->chunk_timed[c], cudaEventDisableTiming));
    }
    if (e->aux_stream == nullptr)
      TPB_CUDA(e, cudaStreamCreateWithFlags(&e->aux_stream, cudaStreamNonBlocking));
    if (e->join_event == nullptr)
      TPB_CUDA(e, cudaEventCreateWithFlags(&e->join_event, cudaEventDisableTiming));
    for (int c = 0; c < nchunks; ++c) {
      const int b0 = static_cast<int>(static_cast<int64_t>(d.B) * c / nchunks);
      const int b1 = static_cast<int>(static_cast<int64_t>(d.B) * (c + 1) / nchunks);
      tpb::Dims dc = d;
      dc.B = b1 - b0;
      const tpb::DeviceArrays ac = tpb::shifted_joint(e->a, d, b0);
      e->launches += tpb::launch_joint_sample_limit(dc, ac, e->safety, e->stream);
      TPB_CUDA(e, cudaEventRecord(e->chunk_sampled[c], e->stream));
      TPB_CUDA(e, cudaStreamWaitEvent(e->integ_stream[c], e->chunk_sampled[c], 0));
      e->launches += tpb::launch_integrate(dc, ac, e->mode, e->safety, max_loops,
                                           e->integ_stream[c]);
      TPB_CUDA(e, cudaEventRecord(e->chunk_timed[c], e->integ_stream[c]));
      cudaStream_t last = want_rows ? e->aux_stream : e->stream;
      TPB_CUDA(e, cudaStreamWaitEvent(last, e->chunk_timed[c], 0));
      if (want_rows)
        e->launches += tpb::launch_resample(dc, ac, start_sec, time_step_sec, e->num_sms,
                                            e->aux_stream, /*continued=*/c > 0);
    }
    if (want_rows) {
      TPB_CUDA(e, cudaEventRecord(e->join_event, e->aux_stream));
      TPB_CUDA(e, cudaStreamWaitEvent(e->stream, e->join_event, 0));
    }
  }
  TPB_CUDA(e, cudaGetLastError());
  e->sampled = e->optimized = true;
  e->resampled = want_rows;
  e->planner_ready = false;
  if (want_rows && total_samples != nullptr) {
    int64_t total = 0;
    TPB_CUDA(e, cudaMemcpyAsync(&total, e->a.r_offset + d.B, sizeof(int64_t),
                                cudaMemcpyDeviceToHost, e->stream));
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
    *total_samples = total;
    if (total > d.r_cap)
      return fail(e, TPB_OUT_OF_RANGE, "resampled rows exceed max_time_samples_total");
  }
  return TPB_OK;
}

int32_t tpb_plan_joint_host(tpb_engine* e, tpb_plan_io* io) {
  TPB_NVTX("tpb_plan_joint_host");
  if (e == nullptr || io == nullptr) return TPB_INVALID_ARGUMENT;
  int32_t rc;
  if ((rc = tpb_set_joint_waypoints(e, io->num_paths, io->num_waypoints, io->waypoints, 0))) return rc;
  if ((rc = tpb_set_joint_limits(e, io->max_velocity, io->max_acceleration, 1, 0))) return rc;
  const auto& d = e->dims;
  cudaSetDevice(e->device);
  // fresh path: sd_start = sdd_start = 0, time_start = start (path_timing_trajectory.cc:322-326)
  TPB_CUDA(e, cudaMemsetAsync(e->a.path_start, 0, sizeof(double) * d.B, e->stream));
  TPB_CUDA(e, cudaMemsetAsync(e->a.sd_start, 0, sizeof(double) * d.B, e->stream));
  TPB_CUDA(e, cudaMemsetAsync(e->a.sdd_start, 0, sizeof(double) * d.B, e->stream));
  e->launches += tpb::launch_fill_f64(e->a.time_start, io->start_sec, d.B, e->stream);
  io->r_total = 0;
  const bool want_rows = io->time_step_sec > 0 && d.r_cap > 0;
  const bool any_out =
      io->r_pos || io->r_vel || io->r_acc || io->r_s || io->r_sd || io->r_sdd || io->r_time;
  const bool any_profile =
      io->profile_time || io->profile_s || io->profile_sd || io->profile_sdd;
  const int max_loops = std::max(100, 10 * d.N);

  // Chunks of the batch go through sample -> integrate -> resample one after the
  // other on the engine's stream; the packed rows of a finished chunk are copied
  // to the host on a second stream while the next chunks are computed.  The first
  // chunks are small (their compute time is exposed), later ones double in size.
  int bounds[tpb_engine::kMaxChunks + 1];
  int nchunks = 0;
  bounds[0] = 0;
  // How finely to cut depends on how much leaves the GPU: the sweeps of a small chunk take
  // about as long as those of the whole batch (they are latency-bound), so chunks only pay
  // when the copies they hide are long.  Bytes per trajectory row the caller asked for:
  const int row_bytes =
      want_rows ? 8 * (d.D * ((io->r_pos != nullptr) + (io->r_vel != nullptr) + (io->r_acc != nullptr)) +
                       (io->r_s != nullptr) + (io->r_sd != nullptr) + (io->r_sdd != nullptr) +
                       (io->r_time != nullptr))
                : 0;
  if (row_bytes >= 120 && d.B >= 1024) {
    // most of the trajectory: small first chunks (their compute time is exposed), doubling
    int size = std::max(128, d.B / 16);
    while (bounds[nchunks] < d.B && nchunks < tpb_engine::kMaxChunks - 1) {
      const int next = std::min(d.B, bounds[nchunks] + size);
      bounds[++nchunks] = next;
      if (nchunks >= 2) size *= 2;
    }
    if (bounds[nchunks] < d.B) bounds[++nchunks] = d.B;
  } else if (row_bytes >= 40 && d.B >= 2048) {
    // one array of rows (positions only): two halves
    bounds[++nchunks] = d.B / 2;
    bounds[++nchunks] = d.B;
  } else {
    bounds[++nchunks] = d.B;
  }
  (void)any_profile;
  if (nchunks > 1) {
    if (e->copy_stream == nullptr)
      TPB_CUDA(e, cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
    if (e->chunk_rows == nullptr)
      TPB_CUDA(e, cudaMallocHost(&e->chunk_rows, sizeof(int64_t) * tpb_engine::kMaxChunks));
    for (int c = 0; c < nchunks; ++c)
      if (e->chunk_done[c] == nullptr)
        TPB_CUDA(e, cudaEventCreateWithFlags(&e->chunk_done[c], cudaEventDisableTiming));
  }

  auto launch_chunk = [&](int c) {
    tpb::Dims dc = d;
    dc.B = bounds[c + 1] - bounds[c];
    const tpb::DeviceArrays ac = (nchunks > 1) ? tpb::shifted_joint(e->a, d, bounds[c]) : e->a;
    e->launches += tpb::launch_joint_sample_limit(dc, ac, e->safety, e->stream);
    e->launches += tpb::launch_integrate(dc, ac, e->mode, e->safety, max_loops, e->stream);
    if (want_rows)
      e->launches += tpb::launch_resample(dc, ac, io->start_sec, io->time_step_sec, e->num_sms,
                                          e->stream, /*continued=*/c > 0);
  };
  const size_t D = d.D;
  int64_t rows_before = 0;
  rc = TPB_OK;
  if (nchunks > 1) {
    for (int c = 0; c <= nchunks; ++c) {
      if (c < nchunks) {
        launch_chunk(c);
        TPB_CUDA(e, cudaGetLastError());
        if (want_rows)
          TPB_CUDA(e, cudaMemcpyAsync(&e->chunk_rows[c], e->a.r_offset + bounds[c + 1],
                                      sizeof(int64_t), cudaMemcpyDeviceToHost, e->stream));
        TPB_CUDA(e, cudaEventRecord(e->chunk_done[c], e->stream));
      }
      if (c == 0) continue;
      // chunk c-1 is done (or will be while chunk c runs): copy its rows out
      const int k = c - 1;
      TPB_CUDA(e, cudaEventSynchronize(e->chunk_done[k]));
      {
        // the profiles of chunk k: [bounds[k], bounds[k+1]) x N
        const size_t p0 = static_cast<size_t>(bounds[k]) * d.N;
        const size_t pn = static_cast<size_t>(bounds[k + 1] - bounds[k]) * d.N;
        auto prof = [&](double* dst, const double* src) -> cudaError_t {
          if (dst == nullptr || pn == 0) return cudaSuccess;
          return cudaMemcpyAsync(dst + p0, src + p0, pn * sizeof(double), cudaMemcpyDeviceToHost,
                                 e->copy_stream);
        };
        TPB_CUDA(e, prof(io->profile_time, e->a.time));
        TPB_CUDA(e, prof(io->profile_s, e->a.s));
        TPB_CUDA(e, prof(io->profile_sd, e->a.sd));
        TPB_CUDA(e, prof(io->profile_sdd, e->a.sdd));
      }
      if (!want_rows) continue;
      const int64_t rows_end = e->chunk_rows[k];
      if (rows_end > d.r_cap || (any_out && rows_end > io->r_capacity)) {
        rc = fail(e, TPB_OUT_OF_RANGE, rows_end > d.r_cap
                                           ? "resampled rows exceed max_time_samples_total"
                                           : "resampled rows exceed r_capacity");
        break;
      }
      const size_t r0 = static_cast<size_t>(rows_before), n = static_cast<size_t>(rows_end - rows_before);
      auto out = [&](double* dst, const double* src, size_t width) -> cudaError_t {
        if (dst == nullptr || n == 0) return cudaSuccess;
        return cudaMemcpyAsync(dst + r0 * width, src + r0 * width, n * width * sizeof(double),
                               cudaMemcpyDeviceToHost, e->copy_stream);
      };
      TPB_CUDA(e, out(io->r_pos, e->a.r_pos, D));
      TPB_CUDA(e, out(io->r_vel, e->a.r_vel, D));
      TPB_CUDA(e, out(io->r_acc, e->a.r_acc, D));
      TPB_CUDA(e, out(io->r_s, e->a.r_s, 1));
      TPB_CUDA(e, out(io->r_sd, e->a.r_sd, 1));
      TPB_CUDA(e, out(io->r_sdd, e->a.r_sdd, 1));
      TPB_CUDA(e, out(io->r_time, e->a.r_time, 1));
      rows_before = rows_end;
    }
    io->r_total = rows_before;
  } else {
    launch_chunk(0);
    TPB_CUDA(e, cudaGetLastError());
  }
  e->sampled = e->optimized = true;
  e->resampled = want_rows;
  if (io->duration)
    TPB_CUDA(e, cudaMemcpyAsync(io->duration, e->a.duration, sizeof(double) * d.B, cudaMemcpyDeviceToHost, e->stream));
  if (io->status)
    TPB_CUDA(e, cudaMemcpyAsync(io->status, e->a.status, sizeof(int32_t) * d.B, cudaMemcpyDeviceToHost, e->stream));
  if (want_rows && io->r_offset)
    TPB_CUDA(e, cudaMemcpyAsync(io->r_offset, e->a.r_offset, sizeof(int64_t) * (d.B + 1), cudaMemcpyDeviceToHost, e->stream));
  if (want_rows && nchunks == 1) {
    int64_t total = 0;
    TPB_CUDA(e, cudaMemcpyAsync(&total, e->a.r_offset + d.B, sizeof(int64_t), cudaMemcpyDeviceToHost, e->stream));
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
    io->r_total = total;
    if (total > d.r_cap)
      return fail(e, TPB_OUT_OF_RANGE, "resampled rows exceed max_time_samples_total");
    const size_t rows = static_cast<size_t>(total);
    if (any_out && total > io->r_capacity)
      return fail(e, TPB_OUT_OF_RANGE, "resampled rows exceed r_capacity");
    if (io->r_pos) TPB_CUDA(e, cudaMemcpyAsync(io->r_pos, e->a.r_pos, rows * D * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    if (io->r_vel) TPB_CUDA(e, cudaMemcpyAsync(io->r_vel, e->a.r_vel, rows * D * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    if (io->r_acc) TPB_CUDA(e, cudaMemcpyAsync(io->r_acc, e->a.r_acc, rows * D * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    if (io->r_s) TPB_CUDA(e, cudaMemcpyAsync(io->r_s, e->a.r_s, rows * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    if (io->r_sd) TPB_CUDA(e, cudaMemcpyAsync(io->r_sd, e->a.r_sd, rows * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    if (io->r_sdd) TPB_CUDA(e, cudaMemcpyAsync(io->r_sdd, e->a.r_sdd, rows * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    if (io->r_time) TPB_CUDA(e, cudaMemcpyAsync(io->r_time, e->a.r_time, rows * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  }
  if (nchunks == 1 && any_profile) {
    const size_t pn = static_cast<size_t>(d.B) * d.N * sizeof(double);
    if (io->profile_time) TPB_CUDA(e, cudaMemcpyAsync(io->profile_time, e->a.time, pn, cudaMemcpyDeviceToHost, e->stream));
    if (io->profile_s) TPB_CUDA(e, cudaMemcpyAsync(io->profile_s, e->a.s, pn, cudaMemcpyDeviceToHost, e->stream));
    if (io->profile_sd) TPB_CUDA(e, cudaMemcpyAsync(io->profile_sd, e->a.sd, pn, cudaMemcpyDeviceToHost, e->stream));
    if (io->profile_sdd) TPB_CUDA(e, cudaMemcpyAsync(io->profile_sdd, e->a.sdd, pn, cudaMemcpyDeviceToHost, e->stream));
  }
  TPB_CUDA(e, cudaStreamSynchronize(e->stream));
  if (e->copy_stream) TPB_CUDA(e, cudaStreamSynchronize(e->copy_stream));
  return rc;
}

// ---------------------------------------------------------------- online re-planning
int32_t tpb_switch_to_waypoint_path(tpb_engine* e, const double* keep_path_until,
                                    const int32_t* num_waypoints, const double* waypoints,
                                    const double* initial_velocity, int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (e->mode != Mode::kJoint || !e->have_waypoints)
    return fail(e, TPB_FAILED_PRECONDITION, "No path set");
  if (e->cartesian)
    return fail(e, TPB_UNIMPLEMENTED, "engines created with cartesian = 1 do not switch joint paths");
  if (keep_path_until == nullptr || waypoints == nullptr)
    return fail(e, TPB_INVALID_ARGUMENT, "keep_path_until / waypoints are null");
  cudaSetDevice(e->device);
  const size_t Bm = d.Bmax, D = d.D;
  // the spline arrays grow once to the reference's knot capacity
  const int cap_points = std::max(2 * (std::max(3 * d.Wmax - 2, 4) + 3), 100) - 3;
  if (d.max_points < cap_points) {
    const size_t tables = sizeof(double) * (static_cast<size_t>(cap_points) + 3 +
                                            static_cast<size_t>(cap_points) * D + 2 * D + 8);
    if (tables > 48 * 1024)
      return fail(e, TPB_UNIMPLEMENTED, "spline capacity for switching exceeds the sampling kernel's shared memory");
    double *ctrl = nullptr, *knots = nullptr;
    TPB_CUDA(e, cudaMalloc(reinterpret_cast<void**>(&ctrl), sizeof(double) * Bm * cap_points * D + 16));
    e->spline_big.push_back(ctrl);
    TPB_CUDA(e, cudaMalloc(reinterpret_cast<void**>(&knots), sizeof(double) * Bm * (cap_points + 3) + 16));
    e->spline_big.push_back(knots);
    e->launches += tpb::launch_restride_spline(d.B, d.D, d.max_points, cap_points, e->a.ctrl,
                                               e->a.knots, ctrl, knots, e->stream);
    e->a.ctrl = ctrl;  // the smaller arrays stay allocated until tpb_destroy
    e->a.knots = knots;
    d.max_points = cap_points;
  }
  if (e->sw_keep == nullptr) {
    TPB_CUDA(e, dev_alloc(e, &e->sw_keep, Bm));
    TPB_CUDA(e, dev_alloc(e, &e->sw_num, Bm));
    TPB_CUDA(e, dev_alloc(e, &e->sw_status, Bm));
    TPB_CUDA(e, dev_alloc(e, &e->sw_modified, Bm));
    TPB_CUDA(e, dev_alloc(e, &e->sw_wp, Bm * d.Wmax * D));
  }
  TPB_CUDA(e, copy_in(e, e->sw_keep, keep_path_until, sizeof(double) * d.B, on_device));
  TPB_CUDA(e, copy_in(e, e->sw_wp, waypoints, sizeof(double) * d.B * d.Wmax * D, on_device));
  if (num_waypoints != nullptr)
    TPB_CUDA(e, copy_in(e, e->sw_num, num_waypoints, sizeof(int32_t) * d.B, on_device));
  tpb::SwitchArgs s;
  s.keep_path_until = e->sw_keep;
  s.num_waypoints = num_waypoints ? e->sw_num : nullptr;
  s.waypoints = e->sw_wp;
  s.max_waypoints = d.Wmax;
  s.status = e->sw_status;
  s.modified = e->sw_modified;
  e->launches += tpb::launch_joint_switch(d, e->a, s, e->rounding, e->stream);
  if (e->planner_ready) {
    e->launches += tpb::launch_planner_mark_modified(d, e->w, e->sw_modified, e->stream);
    if (initial_velocity != nullptr)
      TPB_CUDA(e, copy_in(e, e->w.init_vel, initial_velocity, sizeof(double) * d.B * D, on_device));
  }
  TPB_CUDA(e, cudaGetLastError());
  e->sampled = e->optimized = false;  // the sampled window belongs to the old spline
  return TPB_OK;
}

int32_t tpb_switch_to_cart_waypoint_path(tpb_engine* e, const double* keep_path_until,
                                         const int32_t* num_waypoints, const double* pose_waypoints,
                                         const double* joint_waypoints,
                                         const double* initial_velocity, int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (e->mode != Mode::kCartesian || !e->have_waypoints)
    return fail(e, TPB_FAILED_PRECONDITION, "No path set");
  if (keep_path_until == nullptr || pose_waypoints == nullptr || joint_waypoints == nullptr)
    return fail(e, TPB_INVALID_ARGUMENT, "keep_path_until / waypoints are null");
  cudaSetDevice(e->device);
  const size_t Bm = d.Bmax, D = d.D;
  // the spline arrays grow once to the reference's knot capacity
  // (timeable_path_cartesian_spline.cc:404-413)
  const int cap_points = std::max(2 * (std::max(3 * d.Wmax - 2, 4) + 3), 100) - 3;
  if (d.max_points < cap_points) {
    double *ctrl = nullptr, *knots = nullptr, *pose = nullptr;
    TPB_CUDA(e, cudaMalloc(reinterpret_cast<void**>(&ctrl), sizeof(double) * Bm * cap_points * D + 16));
    e->spline_big.push_back(ctrl);
    TPB_CUDA(e, cudaMalloc(reinterpret_cast<void**>(&knots), sizeof(double) * Bm * (cap_points + 3) + 16));
    e->spline_big.push_back(knots);
    TPB_CUDA(e, cudaMalloc(reinterpret_cast<void**>(&pose), sizeof(double) * Bm * cap_points * 7 + 16));
    e->spline_big.push_back(pose);
    e->launches += tpb::launch_restride_spline(d.B, d.D, d.max_points, cap_points, e->a.ctrl,
                                               e->a.knots, ctrl, knots, e->stream);
    e->launches += tpb::launch_restride_rows(d.B, 7, d.max_points, cap_points, e->a.pose_ctrl, pose,
                                             e->stream);
    e->a.ctrl = ctrl;  // the smaller arrays stay allocated until tpb_destroy
    e->a.knots = knots;
    e->a.pose_ctrl = pose;
    d.max_points = cap_points;
  }
  if (e->sw_keep == nullptr) {
    TPB_CUDA(e, dev_alloc(e, &e->sw_keep, Bm));
    TPB_CUDA(e, dev_alloc(e, &e->sw_num, Bm));
    TPB_CUDA(e, dev_alloc(e, &e->sw_status, Bm));
    TPB_CUDA(e, dev_alloc(e, &e->sw_modified, Bm));
    TPB_CUDA(e, dev_alloc(e, &e->sw_wp, Bm * d.Wmax * D));
  }
  if (e->sw_pose == nullptr) TPB_CUDA(e, dev_alloc(e, &e->sw_pose, Bm * d.Wmax * 7));
  TPB_CUDA(e, copy_in(e, e->sw_keep, keep_path_until, sizeof(double) * d.B, on_device));
  TPB_CUDA(e, copy_in(e, e->sw_wp, joint_waypoints, sizeof(double) * d.B * d.Wmax * D, on_device));
  TPB_CUDA(e, copy_in(e, e->sw_pose, pose_waypoints, sizeof(double) * d.B * d.Wmax * 7, on_device));
  if (num_waypoints != nullptr)
    TPB_CUDA(e, copy_in(e, e->sw_num, num_waypoints, sizeof(int32_t) * d.B, on_device));
  tpb::SwitchArgs s;
  s.keep_path_until = e->sw_keep;
  s.num_waypoints = num_waypoints ? e->sw_num : nullptr;
  s.waypoints = e->sw_wp;
  s.pose_waypoints = e->sw_pose;
  s.max_waypoints = d.Wmax;
  s.status = e->sw_status;
  s.modified = e->sw_modified;
  e->launches += tpb::launch_cart_switch(d, e->a, s, e->rounding, e->translation_rounding, e->stream);
  if (e->planner_ready) {
    e->launches += tpb::launch_planner_mark_modified(d, e->w, e->sw_modified, e->stream);
    if (initial_velocity != nullptr)
      TPB_CUDA(e, copy_in(e, e->w.init_vel, initial_velocity, sizeof(double) * d.B * D, on_device));
  }
  TPB_CUDA(e, cudaGetLastError());
  // the IK solutions beyond the kept part belong to the old path (:188-192): the caches of
  // user kinematics (tpb_planner_set_host_kinematics) are cut at PathIkIndex(keep_path_until)
  if (!e->ik_cache.empty()) {
    std::vector<double> keep(d.B), delta(d.B);
    std::vector<int32_t> modified(d.B);
    TPB_CUDA(e, cudaMemcpyAsync(keep.data(), e->sw_keep, sizeof(double) * d.B, cudaMemcpyDeviceToHost,
                                e->stream));
    TPB_CUDA(e, cudaMemcpyAsync(delta.data(), e->a.delta, sizeof(double) * d.B, cudaMemcpyDeviceToHost,
                                e->stream));
    TPB_CUDA(e, cudaMemcpyAsync(modified.data(), e->sw_modified, sizeof(int32_t) * d.B,
                                cudaMemcpyDeviceToHost, e->stream));
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
    for (int b = 0; b < d.B && b < static_cast<int>(e->ik_cache.size()); ++b) {
      if (modified[b] == 0) continue;
      const long ik_index = std::lround(keep[b] / delta[b]);  // PathIkIndex (:671-674)
      auto& cache = e->ik_cache[b];
      if (ik_index >= 0 && static_cast<size_t>(ik_index) < cache.size() / D)
        cache.resize(static_cast<size_t>(ik_index + 1) * D);
    }
  }
  e->sampled = e->optimized = false;  // the sampled window belongs to the old splines
  return TPB_OK;
}

// ---------------------------------------------------------------- moving-horizon planner
int32_t tpb_planner_reset(tpb_engine* e, int32_t path_sample_capacity, double time_step_sec,
                          int32_t skip_method, int32_t max_planning_iterations,
                          double max_initial_velocity_error, const double* initial_velocity,
                          int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if ((e->mode != Mode::kJoint && e->mode != Mode::kCartesian) || !e->have_waypoints)
    return fail(e, TPB_FAILED_PRECONDITION, "No path set");
  if (d.r_cap <= 0) return fail(e, TPB_FAILED_PRECONDITION, "max_time_samples_total is 0");
  if (!(time_step_sec > 0)) return fail(e, TPB_INVALID_ARGUMENT, "time step must be positive");
  if (path_sample_capacity < d.N)
    return fail(e, TPB_INVALID_ARGUMENT, "path_sample_capacity is below num_path_samples");
  cudaSetDevice(e->device);
  auto& w = e->w;
  const size_t Bm = d.Bmax;
  if (w.len == nullptr) {  // the per-path state, once
    int32_t** i32[] = {&w.len, &w.active, &w.status, &w.mode, &w.initial_plan, &w.sampled,
                       &w.planned_to_end, &w.target_reached, &w.loops, &w.offset,
                       &w.erase_rows, &w.first_from, &w.keep_base};
    for (auto p : i32) TPB_CUDA(e, dev_alloc(e, p, Bm));
    int64_t** i64[] = {&w.start_time, &w.end_time, &w.final_decel_start, &w.loop_start};
    for (auto p : i64) TPB_CUDA(e, dev_alloc(e, p, Bm));
    double** f64[] = {&w.path_horizon, &w.decel_time, &w.origin, &w.front_time, &w.back_time};
    for (auto p : f64) TPB_CUDA(e, dev_alloc(e, p, Bm));
    TPB_CUDA(e, dev_alloc(e, &w.init_vel, Bm * d.D));
    TPB_CUDA(e, dev_alloc(e, &w.any_active, 1));
    TPB_CUDA(e, cudaMallocHost(reinterpret_cast<void**>(&e->any_active_host), sizeof(int32_t)));
  }
  if (path_sample_capacity > w.P_cap) {
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
    for (void* p : e->w_big) cudaFree(p);
    e->w_big.clear();
    w.P_cap = 0;
    const size_t BP = Bm * static_cast<size_t>(path_sample_capacity);
    auto big = [&](auto** p, size_t count) -> cudaError_t {
      void* q = nullptr;
      cudaError_t err = cudaMalloc(&q, count * sizeof(**p));
      if (err != cudaSuccess) return err;
      e->w_big.push_back(q);
      *p = static_cast<std::remove_reference_t<decltype(*p)>>(q);
      return cudaSuccess;
    };
    TPB_CUDA(e, big(&w.w_time, BP));
    TPB_CUDA(e, big(&w.w_s, BP));
    TPB_CUDA(e, big(&w.w_sd, BP));
    TPB_CUDA(e, big(&w.w_sdd, BP));
    TPB_CUDA(e, big(&w.w_q, BP * d.D));
    TPB_CUDA(e, big(&w.w_dq, BP * d.D));
    TPB_CUDA(e, big(&w.w_ddq, BP * d.D));
    TPB_CUDA(e, big(&w.keep, BP));
    w.P_cap = path_sample_capacity;
  }
  e->wopt.dt = time_step_sec;
  e->wopt.skip = skip_method != 0;
  e->wopt.max_planning_iterations = max_planning_iterations;
  e->wopt.max_initial_velocity_error = max_initial_velocity_error;
  TPB_CUDA(e, fill_or_copy(e, w.init_vel, initial_velocity, static_cast<size_t>(d.B) * d.D,
                           on_device));
  TPB_CUDA(e, cudaMemsetAsync(e->a.r_count, 0, sizeof(int32_t) * d.B, e->stream));
  TPB_CUDA(e, cudaMemsetAsync(e->a.r_offset, 0, sizeof(int64_t) * (d.B + 1), e->stream));
  e->launches += tpb::launch_planner_reset(d, w, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  for (auto& c : e->ik_cache) c.clear();  // TimeableCartesianSplinePath::Reset / SetWaypoints
  e->planner_ready = true;
  e->resampled = true;  // zero rows so far
  return TPB_OK;
}

int32_t tpb_planner_new_path(tpb_engine* e, const double* initial_velocity, int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (e->mode != Mode::kJoint || !e->have_waypoints)
    return fail(e, TPB_FAILED_PRECONDITION, "No path set");
  if (e->w.len == nullptr || e->w.P_cap < d.N)
    return fail(e, TPB_FAILED_PRECONDITION, "tpb_planner_reset was never called");
  cudaSetDevice(e->device);
  TPB_CUDA(e, fill_or_copy(e, e->w.init_vel, initial_velocity, static_cast<size_t>(d.B) * d.D,
                           on_device));
  e->launches += tpb::launch_planner_new_path(d, e->w, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  e->planner_ready = true;
  e->resampled = true;
  return TPB_OK;
}

int32_t tpb_planner_set_initial_velocity(tpb_engine* e, const double* initial_velocity,
                                         int32_t on_device) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  if (!e->planner_ready)
    return fail(e, TPB_FAILED_PRECONDITION, "tpb_planner_reset was not called for these paths");
  cudaSetDevice(e->device);
  TPB_CUDA(e, fill_or_copy(e, e->w.init_vel, initial_velocity,
                           static_cast<size_t>(e->dims.B) * e->dims.D, on_device));
  return TPB_OK;
}

int32_t tpb_planner_set_kinematics(tpb_engine* e, int32_t model) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  if (model != TPB_KIN_HOST && model != TPB_KIN_FAKE7)
    return fail(e, TPB_INVALID_ARGUMENT, "unknown kinematics model");
  e->planner_kinematics = model;
  return TPB_OK;
}

namespace {

// One planning window of Cartesian paths with user kinematics: SamplePath's IK part
// (timeable_path_cartesian_spline.cc:464-549) and the Jacobians of ConstraintSetup
// (:551-595) on the host.  The device has just written the window's N + 1 pose / joint
// targets (IK indices PathIkIndex(path_start) - 1 ... + N - 1).  Per active path: the
// targets of the IK indices not solved yet go to the PathIKFunc, warm-started from the
// last cached solution (:503-506); its solutions but the first extend the cache
// (:512-513); the window's positions are the cache entries of its indices (:527-540);
// the JacobianFunc is evaluated at each of them.  Positions and Jacobians go back to
// the device (TPB_ARR_Q, the Jacobian table), which does the rest.
int32_t host_kinematics_window(tpb_engine* e) {
  const auto& d = e->dims;
  const size_t B = d.B, N = d.N, D = d.D;
  e->hk_active.resize(B);
  e->hk_start.resize(B);
  e->hk_delta.resize(B);
  e->hk_pose.resize(B * (N + 1) * 7);
  e->hk_joint.resize(B * (N + 1) * D);
  e->hk_q.resize(N * D);
  e->hk_jac.resize(N * 6 * D);
  if (e->ik_cache.size() < B) e->ik_cache.resize(B);
  auto d2h = [&](void* dst, const void* src, size_t bytes) {
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, e->stream);
  };
  TPB_CUDA(e, d2h(e->hk_active.data(), e->w.active, B * sizeof(int32_t)));
  TPB_CUDA(e, d2h(e->hk_start.data(), e->a.path_start, B * sizeof(double)));
  TPB_CUDA(e, d2h(e->hk_delta.data(), e->a.delta, B * sizeof(double)));
  TPB_CUDA(e, d2h(e->hk_pose.data(), e->a.pose_targets, e->hk_pose.size() * sizeof(double)));
  TPB_CUDA(e, d2h(e->hk_joint.data(), e->a.joint_targets, e->hk_joint.size() * sizeof(double)));
  TPB_CUDA(e, cudaStreamSynchronize(e->stream));
  for (size_t b = 0; b < B; ++b) {
    if (e->hk_active[b] == 0) continue;
    std::vector<double>& cache = e->ik_cache[b];
    const double delta = e->hk_delta[b];
    // PathIkIndex (:671-674)
    const int start_index = static_cast<int>(std::round(e->hk_start[b] / delta));
    const double path_horizon = e->hk_start[b] + delta * static_cast<double>(N - 1);
    const int horizon_index = static_cast<int>(std::round(path_horizon / delta));
    const int current_upper = static_cast<int>(cache.size() / D) - 1;
    if (horizon_index - start_index != static_cast<int>(N) - 1 || current_upper < start_index - 1)
      return fail(e, TPB_INTERNAL, "IK index range of the window does not follow the cache");
    if (horizon_index >= current_upper) {
      const int num_new = horizon_index - current_upper + 1;
      const size_t slot = static_cast<size_t>(current_upper - (start_index - 1));
      const double* pt = e->hk_pose.data() + (b * (N + 1) + slot) * 7;
      const double* jt = e->hk_joint.data() + (b * (N + 1) + slot) * D;
      const double* initial =
          current_upper > 0 ? cache.data() + static_cast<size_t>(current_upper) * D : jt;
      e->hk_solution.assign(static_cast<size_t>(num_new) * D, 0.0);
      const int32_t rc = e->host_ik(e->host_user, static_cast<int32_t>(b), initial, num_new, pt,
                                    jt, e->hk_solution.data());
      // partial solutions are kept like the reference keeps them (:509-513)
      cache.insert(cache.end(), e->hk_solution.begin() + D, e->hk_solution.end());
      if (rc != TPB_OK)
        return fail(e, rc, "PathIKFunc failed for path " + std::to_string(b));
    }
    const double* win = cache.data() + static_cast<size_t>(start_index) * D;
    TPB_CUDA(e, cudaMemcpyAsync(e->a.q + b * N * D, win, N * D * sizeof(double),
                                cudaMemcpyHostToDevice, e->stream));
    for (size_t i = 0; i < N; ++i) {
      const int32_t rc = e->host_jacobian(e->host_user, static_cast<int32_t>(b), win + i * D,
                                          e->hk_jac.data() + i * 6 * D);
      if (rc != TPB_OK)
        return fail(e, rc, "JacobianFunc failed for path " + std::to_string(b));
    }
    TPB_CUDA(e, cudaMemcpyAsync(e->a.jac + b * N * 6 * D, e->hk_jac.data(),
                                N * 6 * D * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    // the staging vectors are reused by the next path
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
  }
  return TPB_OK;
}

}  // namespace

int32_t tpb_planner_set_host_kinematics(tpb_engine* e, tpb_path_ik_fn ik, tpb_jacobian_fn jacobian,
                                        void* user) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  if (e->mode != Mode::kCartesian)
    return fail(e, TPB_FAILED_PRECONDITION, "engine was not created for Cartesian paths");
  if (ik == nullptr || jacobian == nullptr)
    return fail(e, TPB_INVALID_ARGUMENT, "PathIKFunc / JacobianFunc is null");
  e->host_ik = ik;
  e->host_jacobian = jacobian;
  e->host_user = user;
  e->planner_kinematics = TPB_KIN_HOST;
  return TPB_OK;
}

int32_t tpb_planner_plan(tpb_engine* e, int64_t start_ns, int64_t horizon_ns,
                         int64_t* total_samples) {
  TPB_NVTX("tpb_planner_plan");
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if ((e->mode != Mode::kJoint && e->mode != Mode::kCartesian) || !e->have_waypoints)
    return fail(e, TPB_FAILED_PRECONDITION, "No path set.");
  if (!e->have_limits) return fail(e, TPB_FAILED_PRECONDITION, "joint limits not set");
  const bool cart = e->mode == Mode::kCartesian;
  const bool host_kinematics = cart && e->planner_kinematics == TPB_KIN_HOST;
  if (cart) {
    if (host_kinematics && (e->host_ik == nullptr || e->host_jacobian == nullptr))
      return fail(e, TPB_FAILED_PRECONDITION,
                  "Cartesian paths behind the planner need kinematics: "
                  "tpb_planner_set_host_kinematics (user PathIKFunc / JacobianFunc) or "
                  "tpb_planner_set_kinematics (built-in device model)");
    if (!host_kinematics && (e->planner_kinematics != TPB_KIN_FAKE7 || d.D != 7))
      return fail(e, TPB_INVALID_ARGUMENT, "unknown kinematics model for this dof count");
    if (!(e->max_trans_vel > 0) || !(e->max_rot_vel > 0))
      return fail(e, TPB_FAILED_PRECONDITION, "Cartesian velocity limits not set");
  }
  if (!e->planner_ready)
    return fail(e, TPB_FAILED_PRECONDITION, "tpb_planner_reset was not called for these paths");
  cudaSetDevice(e->device);
  const auto& w = e->w;
  const auto& o = e->wopt;
  auto read_flag = [&](int32_t* flag) -> cudaError_t {
    cudaError_t err = cudaMemcpyAsync(e->any_active_host, w.any_active, sizeof(int32_t),
                                      cudaMemcpyDeviceToHost, e->stream);
    if (err != cudaSuccess) return err;
    err = cudaStreamSynchronize(e->stream);
    *flag = *e->any_active_host;
    return err;
  };
  TPB_CUDA(e, cudaMemsetAsync(w.any_active, 0, sizeof(int32_t), e->stream));
  e->launches += tpb::launch_plan_begin(d, e->a, w, o, start_ns, horizon_ns, e->stream);
  int32_t more = 0;
  TPB_CUDA(e, read_flag(&more));
  if (host_kinematics && e->a.jac == nullptr)  // the Jacobian table of the host callbacks
    TPB_CUDA(e, dev_alloc(e, &e->a.jac, static_cast<size_t>(d.Bmax) * d.N * 6 * d.D));
  tpb::DeviceArrays am = e->a;  // the stage kernels, masked to the paths that plan a window
  am.active = w.active;
  const int max_solver_loops = std::max(100, 10 * d.N);  // (:399-401)
  for (int window = 0; more != 0; ++window) {
    if (window > o.max_planning_iterations + 1)
      return fail(e, TPB_INTERNAL, "planning loop did not terminate");
    e->launches += tpb::launch_window_begin(d, e->a, w, start_ns, horizon_ns, e->stream);
    if (cart && host_kinematics) {
      e->launches += tpb::launch_cart_sample_targets(d, am, e->stream);
      const int32_t rc = host_kinematics_window(e);  // user PathIKFunc / JacobianFunc
      if (rc != TPB_OK) return rc;
      e->launches += tpb::launch_cart_constraints(d, am, e->safety, e->max_trans_vel,
                                                  e->max_rot_vel, TPB_KIN_HOST, true, e->stream);
    } else if (cart) {  // SamplePath (targets, path IK, finite differences) + ConstraintSetup
      e->launches += tpb::launch_cart_sample_targets(d, am, e->stream);
      e->launches += tpb::launch_cart_ik_builtin(d, am, e->planner_kinematics, e->stream);
      e->launches += tpb::launch_cart_constraints(d, am, e->safety, e->max_trans_vel,
                                                  e->max_rot_vel, e->planner_kinematics, false,
                                                  e->stream);  // + limit curve
    } else {
      e->launches += tpb::launch_joint_sample_limit(d, am, e->safety, e->stream);
    }
    e->launches += tpb::launch_window_velocity(d, e->a, w, o, e->stream);
    e->launches += tpb::launch_integrate(d, am, e->mode, e->safety, max_solver_loops, e->stream);
    TPB_CUDA(e, cudaMemsetAsync(w.any_active, 0, sizeof(int32_t), e->stream));
    e->launches += tpb::launch_window_append(d, e->a, w, o, start_ns, horizon_ns, e->stream);
    TPB_CUDA(e, read_flag(&more));
  }
  e->launches += tpb::launch_plan_rows(d, e->a, w, o, start_ns, e->stream);
  // row assembly over the spliced samples: the resampling kernel on a view
  tpb::Dims dv = d;
  dv.N = w.P_cap;
  tpb::DeviceArrays av = e->a;
  av.time = w.w_time;
  av.s = w.w_s;
  av.sd = w.w_sd;
  av.sdd = w.w_sdd;
  av.q = w.w_q;
  av.dq = w.w_dq;
  av.ddq = w.w_ddq;
  tpb::ResampleView v;
  v.len = w.len;
  v.origin = w.origin;
  v.grid_skip = w.erase_rows;
  v.first_from = w.first_from;
  v.keep_base = w.keep_base;
  e->launches += tpb::launch_resample_rows(dv, av, v, w.keep, 0.0, o.dt, e->num_sms, e->stream,
                                           o.skip);
  e->launches += tpb::launch_plan_end(d, e->a, w, o, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  e->sampled = e->optimized = false;  // the window scratch is not a fresh-path solution
  e->resampled = true;
  if (total_samples != nullptr) {
    int64_t total = 0;
    TPB_CUDA(e, cudaMemcpyAsync(&total, e->a.r_offset + d.B, sizeof(int64_t),
                                cudaMemcpyDeviceToHost, e->stream));
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
    *total_samples = total;
    if (total > d.r_cap)
      return fail(e, TPB_OUT_OF_RANGE, "trajectory rows exceed max_time_samples_total");
  }
  return TPB_OK;
}

int32_t tpb_set_kernel_variant(tpb_engine* e, tpb_kernel_variant which, int32_t on) {
  if (e == nullptr) return TPB_INVALID_ARGUMENT;
  const int bit = static_cast<int>(which);
  if (bit < 0 || bit > 3) return fail(e, TPB_INVALID_ARGUMENT, "unknown kernel variant");
  if (on) {
    e->dims.variants |= (1u << bit);
  } else {
    e->dims.variants &= ~(1u << bit);
  }
  return TPB_OK;
}

int32_t tpb_get(tpb_engine* e, tpb_array what, void* dst, size_t bytes, int32_t to_device) {
  if (e == nullptr || dst == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(e->device);
  if (e->dims.B == 0) return fail(e, TPB_FAILED_PRECONDITION, "no batch set");
  if (is_joint_array(what) && e->mode != Mode::kJoint && e->mode != Mode::kCartesian)
    return fail(e, TPB_FAILED_PRECONDITION, "array only exists for joint paths");
  if (is_cart_array(what) && e->mode != Mode::kCartesian)
    return fail(e, TPB_FAILED_PRECONDITION, "array only exists for Cartesian paths");
  ArrayRef r;
  if (!lookup(e, what, &r) || r.ptr == nullptr) return fail(e, TPB_NOT_FOUND, "unknown array");
  if (bytes < r.bytes) return fail(e, TPB_OUT_OF_RANGE, "destination too small");
  TPB_CUDA(e, cudaMemcpyAsync(dst, r.ptr, r.bytes,
                              to_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                              e->stream));
  if (!to_device) TPB_CUDA(e, cudaStreamSynchronize(e->stream));
  return TPB_OK;
}

int32_t tpb_devptr(tpb_engine* e, tpb_array what, void** ptr, size_t* bytes) {
  if (e == nullptr || ptr == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(e->device);
  ArrayRef r;
  if (!lookup(e, what, &r) || r.ptr == nullptr) return fail(e, TPB_NOT_FOUND, "unknown array");
  *ptr = r.ptr;
  if (bytes) *bytes = r.bytes;
  return TPB_OK;
}

// ---------------------------------------------------------------- profile queries (A16)
int32_t tpb_count_violations(tpb_engine* e, int32_t* violations, int32_t on_device) {
  if (e == nullptr || violations == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (!e->optimized) return fail(e, TPB_FAILED_PRECONDITION, "No valid solution.");
  cudaSetDevice(e->device);
  int32_t* out = violations;
  if (!on_device) {
    const size_t need = sizeof(int32_t) * static_cast<size_t>(d.B);
    if (need > e->query_scratch_bytes) {
      TPB_CUDA(e, cudaStreamSynchronize(e->stream));
      if (e->query_scratch) cudaFree(e->query_scratch);
      e->query_scratch = nullptr;
      e->query_scratch_bytes = 0;
      TPB_CUDA(e, cudaMalloc(&e->query_scratch, need + 16));
      e->query_scratch_bytes = need;
    }
    out = static_cast<int32_t*>(e->query_scratch);
  }
  e->launches += tpb::launch_count_violations(d, e->a, e->mode == Mode::kJoint, e->safety, out,
                                              e->stream);
  TPB_CUDA(e, cudaGetLastError());
  if (!on_device) {
    TPB_CUDA(e, cudaMemcpyAsync(violations, out, sizeof(int32_t) * d.B, cudaMemcpyDeviceToHost,
                                e->stream));
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
  }
  return TPB_OK;
}

int32_t tpb_query_profile(tpb_engine* e, int32_t queries_per_path, const double* time_sec,
                          double* s, double* sd, double* sdd, int32_t* previous_index,
                          double* sk, double* sdk, double* sddk, double* tk, int32_t* status,
                          int32_t on_device) {
  TPB_NVTX("tpb_query_profile");
  if (e == nullptr || time_sec == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (!e->optimized) return fail(e, TPB_FAILED_PRECONDITION, "Error, solution not yet calculated!");
  if (queries_per_path < 1) return fail(e, TPB_INVALID_ARGUMENT, "queries_per_path must be >= 1");
  if (d.B > 65535) return fail(e, TPB_OUT_OF_RANGE, "profile queries need <= 65535 paths");
  cudaSetDevice(e->device);
  const size_t n = static_cast<size_t>(d.B) * queries_per_path;
  tpb::ProfileQueryArgs p;
  p.Q = queries_per_path;
  p.use_s_range = e->mode != Mode::kJoint;
  double* outs[7] = {s, sd, sdd, sk, sdk, sddk, tk};
  if (on_device) {
    p.time = time_sec;
    p.s = s; p.sd = sd; p.sdd = sdd;
    p.sk = sk; p.sdk = sdk; p.sddk = sddk; p.tk = tk;
    p.prev_index = previous_index;
    p.status = status;
  } else {
    // host arrays: staged through a scratch buffer of the engine that only grows
    const size_t need = n * (8 * sizeof(double) + 2 * sizeof(int32_t));
    if (need > e->query_scratch_bytes) {
      TPB_CUDA(e, cudaStreamSynchronize(e->stream));
      if (e->query_scratch) cudaFree(e->query_scratch);
      e->query_scratch = nullptr;
      e->query_scratch_bytes = 0;
      TPB_CUDA(e, cudaMalloc(&e->query_scratch, need + 16));
      e->query_scratch_bytes = need;
    }
    double* base = static_cast<double*>(e->query_scratch);
    TPB_CUDA(e, copy_in(e, base, time_sec, n * sizeof(double), 0));
    p.time = base;
    double** dst[7] = {&p.s, &p.sd, &p.sdd, &p.sk, &p.sdk, &p.sddk, &p.tk};
    for (int k = 0; k < 7; ++k)
      if (outs[k] != nullptr) *dst[k] = base + (k + 1) * n;
    int32_t* ibase = reinterpret_cast<int32_t*>(base + 8 * n);
    if (previous_index) p.prev_index = ibase;
    if (status) p.status = ibase + n;
  }
  e->launches += tpb::launch_profile_query(d, e->a, p, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  if (!on_device) {
    const double* src[7] = {p.s, p.sd, p.sdd, p.sk, p.sdk, p.sddk, p.tk};
    for (int k = 0; k < 7; ++k)
      if (outs[k] != nullptr)
        TPB_CUDA(e, cudaMemcpyAsync(outs[k], src[k], n * sizeof(double), cudaMemcpyDeviceToHost,
                                    e->stream));
    if (previous_index)
      TPB_CUDA(e, cudaMemcpyAsync(previous_index, p.prev_index, n * sizeof(int32_t),
                                  cudaMemcpyDeviceToHost, e->stream));
    if (status)
      TPB_CUDA(e, cudaMemcpyAsync(status, p.status, n * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                  e->stream));
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
  }
  return TPB_OK;
}

int32_t tpb_path_stop_parameter(tpb_engine* e, const double* time_sec, double* stop_parameter,
                                int32_t* stop_row, double* stop_duration, int32_t* status,
                                int32_t on_device) {
  if (e == nullptr || time_sec == nullptr) return TPB_INVALID_ARGUMENT;
  auto& d = e->dims;
  if (!e->resampled) return fail(e, TPB_FAILED_PRECONDITION, "no resampled trajectories");
  if (d.D > 16) return fail(e, TPB_UNIMPLEMENTED, "stop queries need num_dofs <= 16");
  cudaSetDevice(e->device);
  if (e->sq_query == nullptr) {
    TPB_CUDA(e, dev_alloc(e, &e->sq_query, d.Bmax));
    TPB_CUDA(e, dev_alloc(e, &e->sq_parameter, d.Bmax));
    TPB_CUDA(e, dev_alloc(e, &e->sq_duration, d.Bmax));
    TPB_CUDA(e, dev_alloc(e, &e->sq_row, d.Bmax));
    TPB_CUDA(e, dev_alloc(e, &e->sq_status, d.Bmax));
  }
  tpb::StopQueryArgs p;
  p.B = d.B;
  p.D = d.D;
  p.offset = e->a.r_offset;
  p.times = e->a.r_time;
  p.s = e->a.r_s;
  p.vel = e->a.r_vel;
  p.acc = e->a.r_acc;
  p.max_acc = e->a.amax;  // path_->GetMaxJointAcceleration(): the limit itself, no safety factor
  if (on_device) {
    p.query = time_sec;
    p.parameter = stop_parameter ? stop_parameter : e->sq_parameter;
    p.row = stop_row ? stop_row : e->sq_row;
    p.duration = stop_duration ? stop_duration : e->sq_duration;
    p.status = status ? status : e->sq_status;
  } else {
    TPB_CUDA(e, copy_in(e, e->sq_query, time_sec, sizeof(double) * d.B, 0));
    p.query = e->sq_query;
    p.parameter = e->sq_parameter;
    p.row = e->sq_row;
    p.duration = e->sq_duration;
    p.status = e->sq_status;
  }
  e->launches += tpb::launch_path_stop_parameter(p, e->stream);
  TPB_CUDA(e, cudaGetLastError());
  if (!on_device) {
    auto down = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
      if (dst == nullptr) return cudaSuccess;
      return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, e->stream);
    };
    TPB_CUDA(e, down(stop_parameter, e->sq_parameter, sizeof(double) * d.B));
    TPB_CUDA(e, down(stop_row, e->sq_row, sizeof(int32_t) * d.B));
    TPB_CUDA(e, down(stop_duration, e->sq_duration, sizeof(double) * d.B));
    TPB_CUDA(e, down(status, e->sq_status, sizeof(int32_t) * d.B));
    TPB_CUDA(e, cudaStreamSynchronize(e->stream));
  }
  return TPB_OK;
}

int32_t tpb_rescale_to_stop_device(const tpb_stop_desc* d, void* cuda_stream) {
  if (d == nullptr || d->num_trajectories < 0 || d->num_dofs < 1) return TPB_INVALID_ARGUMENT;
  if (d->num_dofs > 16) return TPB_UNIMPLEMENTED;  // one lane per (joint, sign) candidate
  if (d->num_trajectories == 0) return TPB_OK;
  if (!d->begin || !d->length || !d->times || !d->velocities || !d->accelerations ||
      !d->max_acceleration || !d->count || !d->switch_index || !d->status || !d->out_times ||
      !d->out_velocities || !d->out_accelerations)
    return TPB_INVALID_ARGUMENT;
  tpb::StopArgs p;
  p.num_traj = d->num_trajectories;
  p.D = d->num_dofs;
  p.begin = d->begin;
  p.length = d->length;
  p.times = d->times;
  p.vel = d->velocities;
  p.acc = d->accelerations;
  p.max_acc = d->max_acceleration;
  p.count = d->count;
  p.switch_index = d->switch_index;
  p.status = d->status;
  p.out_times = d->out_times;
  p.out_vel = d->out_velocities;
  p.out_acc = d->out_accelerations;
  tpb::launch_rescale_to_stop(p, static_cast<cudaStream_t>(cuda_stream));
  return cudaGetLastError() == cudaSuccess ? TPB_OK : TPB_INTERNAL;
}

int32_t tpb_rescale_to_stop_host(const tpb_stop_desc* hd, int32_t device) {
  if (hd == nullptr || hd->num_trajectories < 0 || hd->num_dofs < 1 || hd->total_rows < 0)
    return TPB_INVALID_ARGUMENT;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return TPB_FAILED_PRECONDITION;
  }
  if (device < 0 || device >= ndev) return TPB_INVALID_ARGUMENT;
  if (cudaSetDevice(device) != cudaSuccess) return TPB_INTERNAL;
  if (hd->num_trajectories == 0) return TPB_OK;
  const size_t T = hd->num_trajectories, D = hd->num_dofs, R = static_cast<size_t>(hd->total_rows);
  // every trajectory must lie inside the packed arrays
  for (size_t b = 0; b < T; ++b) {
    if (hd->begin[b] < 0 || hd->length[b] < 0 ||
        hd->begin[b] + hd->length[b] > hd->total_rows)
      return TPB_INVALID_ARGUMENT;
  }
  std::vector<void*> tmp;
  auto alloc = [&](size_t bytes, void** out) -> cudaError_t {
    cudaError_t err = cudaMalloc(out, bytes ? bytes : 1);
    if (err == cudaSuccess) tmp.push_back(*out);
    return err;
  };
  auto up = [&](const void* src, size_t bytes, const void** dst) -> cudaError_t {
    void* p = nullptr;
    cudaError_t err = alloc(bytes, &p);
    if (err != cudaSuccess) return err;
    *dst = p;
    return cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice);
  };
  tpb_stop_desc d = *hd;
  cudaError_t err = cudaSuccess;
  const void* q = nullptr;
#define TPB_UP(field, bytes)                                        \
  if (err == cudaSuccess) {                                         \
    err = up(hd->field, bytes, &q);                                 \
    d.field = static_cast<decltype(d.field)>(q);                    \
  }
  TPB_UP(begin, T * sizeof(int64_t));
  TPB_UP(length, T * sizeof(int32_t));
  TPB_UP(times, R * sizeof(double));
  TPB_UP(velocities, R * D * sizeof(double));
  TPB_UP(accelerations, R * D * sizeof(double));
  TPB_UP(max_acceleration, T * D * sizeof(double));
#undef TPB_UP
  void* w = nullptr;
#define TPB_OUT(field, bytes)                                       \
  if (err == cudaSuccess) {                                         \
    err = alloc(bytes, &w);                                         \
    d.field = static_cast<decltype(d.field)>(w);                    \
  }
  TPB_OUT(count, T * sizeof(int32_t));
  TPB_OUT(switch_index, T * sizeof(int32_t));
  TPB_OUT(status, T * sizeof(int32_t));
  TPB_OUT(out_times, R * sizeof(double));
  TPB_OUT(out_velocities, R * D * sizeof(double));
  TPB_OUT(out_accelerations, R * D * sizeof(double));
#undef TPB_OUT
  int32_t rc = TPB_OK;
  if (err == cudaSuccess) {
    // rows outside the stopping segments keep what the caller's buffers hold
    err = cudaMemcpy(d.out_times, hd->out_times, R * sizeof(double), cudaMemcpyHostToDevice);
    if (err == cudaSuccess)
      err = cudaMemcpy(d.out_velocities, hd->out_velocities, R * D * sizeof(double),
                       cudaMemcpyHostToDevice);
    if (err == cudaSuccess)
      err = cudaMemcpy(d.out_accelerations, hd->out_accelerations, R * D * sizeof(double),
                       cudaMemcpyHostToDevice);
  }
  if (err == cudaSuccess) {
    rc = tpb_rescale_to_stop_device(&d, nullptr);
    if (rc == TPB_OK) err = cudaDeviceSynchronize();
  }
#define TPB_DOWN(field, bytes) \
  if (rc == TPB_OK && err == cudaSuccess) err = cudaMemcpy(hd->field, d.field, bytes, cudaMemcpyDeviceToHost)
  TPB_DOWN(count, T * sizeof(int32_t));
  TPB_DOWN(switch_index, T * sizeof(int32_t));
  TPB_DOWN(status, T * sizeof(int32_t));
  TPB_DOWN(out_times, R * sizeof(double));
  TPB_DOWN(out_velocities, R * D * sizeof(double));
  TPB_DOWN(out_accelerations, R * D * sizeof(double));
#undef TPB_DOWN
  for (void* p : tmp) cudaFree(p);
  if (err != cudaSuccess) {
    fprintf(stderr, "tpb_rescale_to_stop_host: %s\n", cudaGetErrorString(err));
    return TPB_INTERNAL;
  }
  return rc;
}

int32_t tpb_spline_eval_device(const tpb_spline_eval_desc* d, void* cuda_stream) {
  if (d == nullptr) return TPB_INVALID_ARGUMENT;
  std::string err;
  const int rc = tpb::launch_spline_eval(*d, static_cast<cudaStream_t>(cuda_stream), &err);
  if (rc != TPB_OK) fprintf(stderr, "tpb_spline_eval_device: %s\n", err.c_str());
  return rc;
}

int32_t tpb_spline_eval_host(const tpb_spline_eval_desc* hd, int32_t device) {
  if (hd == nullptr) return TPB_INVALID_ARGUMENT;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return TPB_FAILED_PRECONDITION;
  }
  if (device < 0 || device >= ndev) return TPB_INVALID_ARGUMENT;
  if (cudaSetDevice(device) != cudaSuccess) return TPB_INTERNAL;
  if (hd->num_splines < 1 || hd->evals_per_spline < 0 || hd->num_points < 1 || hd->dim < 1)
    return TPB_INVALID_ARGUMENT;
  tpb_spline_eval_desc d = *hd;
  const size_t S = hd->num_splines, E = static_cast<size_t>(hd->evals_per_spline);
  const size_t nk = static_cast<size_t>(hd->num_points) + hd->degree + 1;
  const size_t nout = S * E * (hd->nder + 1) * hd->dim;
  std::vector<void*> tmp;
  auto up = [&](const void* src, size_t bytes, const void** dst) -> cudaError_t {
    if (src == nullptr) {
      *dst = nullptr;
      return cudaSuccess;
    }
    void* p = nullptr;
    cudaError_t err = cudaMalloc(&p, bytes ? bytes : 1);
    if (err != cudaSuccess) return err;
    tmp.push_back(p);
    *dst = p;
    return cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice);
  };
  cudaError_t err = cudaSuccess;
  const void *dk = nullptr, *dc = nullptr, *du_ = nullptr, *du0 = nullptr, *ddu = nullptr;
  if (err == cudaSuccess) err = up(hd->knots, S * nk * sizeof(double), &dk);
  if (err == cudaSuccess) err = up(hd->ctrl, S * hd->num_points * hd->dim * sizeof(double), &dc);
  if (err == cudaSuccess) err = up(hd->u, S * E * sizeof(double), &du_);
  if (err == cudaSuccess) err = up(hd->u0, S * sizeof(double), &du0);
  if (err == cudaSuccess) err = up(hd->du, S * sizeof(double), &ddu);
  double* dout = nullptr;
  int32_t* dst = nullptr;
  if (err == cudaSuccess) err = cudaMalloc(&dout, std::max<size_t>(nout, 1) * sizeof(double));
  if (err == cudaSuccess) tmp.push_back(dout);
  // rows that fail the range check keep the caller's contents (like the reference)
  if (err == cudaSuccess && hd->out)
    err = cudaMemcpy(dout, hd->out, nout * sizeof(double), cudaMemcpyHostToDevice);
  if (err == cudaSuccess && hd->status) {
    err = cudaMalloc(&dst, std::max<size_t>(S * E, 1) * sizeof(int32_t));
    if (err == cudaSuccess) tmp.push_back(dst);
  }
  int32_t rc = TPB_OK;
  if (err == cudaSuccess) {
    d.knots = static_cast<const double*>(dk);
    d.ctrl = static_cast<const double*>(dc);
    d.u = static_cast<const double*>(du_);
    d.u0 = static_cast<const double*>(du0);
    d.du = static_cast<const double*>(ddu);
    d.out = dout;
    d.status = dst;
    std::string msg;
    rc = tpb::launch_spline_eval(d, nullptr, &msg);
    if (rc == TPB_OK) err = cudaDeviceSynchronize();
    if (rc != TPB_OK) fprintf(stderr, "tpb_spline_eval_host: %s\n", msg.c_str());
  }
  if (rc == TPB_OK && err == cudaSuccess && hd->out)
    err = cudaMemcpy(hd->out, dout, nout * sizeof(double), cudaMemcpyDeviceToHost);
  if (rc == TPB_OK && err == cudaSuccess && hd->status)
    err = cudaMemcpy(hd->status, dst, S * E * sizeof(int32_t), cudaMemcpyDeviceToHost);
  for (void* p : tmp) cudaFree(p);
  if (err != cudaSuccess) {
    fprintf(stderr, "tpb_spline_eval_host: %s\n", cudaGetErrorString(err));
    return TPB_INTERNAL;
  }
  return rc;
}

}  // extern "C"
