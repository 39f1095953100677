"""ctypes binding of include/tpb.h (the C ABI of lib/libtpb.so).

Host arrays are numpy; device arrays are anything with ``data_ptr()`` (torch
tensors) or raw integer device pointers.  No computation happens in Python.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# TPB_LIB: another build of the library (A/B measurements of kernel variants only)
_LIB = os.environ.get("TPB_LIB") or os.path.join(_HERE, "lib", "libtpb.so")

PATH_OK = 0
PATH_FAIL_SETUP = 1


class ARR:
    NUM_POINTS, CTRL, KNOTS, DELTA, POSE_CTRL = 1, 2, 3, 4, 5
    Q, DQ, DDQ, POSE_TARGETS, JOINT_TARGETS = 10, 11, 12, 13, 14
    CONS = 35
    TIME, S, SD, SDD, SD2 = 20, 21, 22, 23, 24
    B_SD2_MAX, B_SDD_MAX, B_SDD_MIN, B_SD2_ZERO, B_TYPE = 30, 31, 32, 33, 34
    STATUS, DURATION, LAST_EXTREMAL, KNOT_END = 40, 41, 42, 43
    R_OFFSET, R_TIME, R_POS, R_VEL, R_ACC, R_S, R_SD, R_SDD = 50, 51, 52, 53, 54, 55, 56, 57
    W_LEN, W_TIME, W_S, W_SD, W_SDD, W_Q, W_DQ, W_DDQ = 60, 61, 62, 63, 64, 65, 66, 67
    (PLAN_STATUS, PLAN_LOOPS, PLAN_PLANNED_TO_END, PLAN_TARGET_REACHED, PLAN_START_TIME,
     PLAN_END_TIME, PLAN_FINAL_DECEL_START, PLAN_ERASED) = 70, 71, 72, 73, 74, 75, 76, 77
    SWITCH_STATUS = 80
    DEBUG = 99


class TpbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"tpb status {code}: {msg}")
        self.code = code


class _Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_paths", C.c_int32), ("num_dofs", C.c_int32),
                ("max_waypoints", C.c_int32), ("num_samples", C.c_int32),
                ("max_constraints", C.c_int32), ("max_time_samples_total", C.c_int64),
                ("cartesian", C.c_int32)]


class _PlanIO(C.Structure):
    _fields_ = [("num_paths", C.c_int32), ("num_waypoints", C.c_void_p),
                ("waypoints", C.c_void_p), ("max_velocity", C.c_void_p),
                ("max_acceleration", C.c_void_p), ("start_sec", C.c_double),
                ("time_step_sec", C.c_double), ("duration", C.c_void_p),
                ("status", C.c_void_p), ("r_offset", C.c_void_p), ("r_pos", C.c_void_p),
                ("r_vel", C.c_void_p), ("r_acc", C.c_void_p), ("r_s", C.c_void_p),
                ("r_sd", C.c_void_p), ("r_sdd", C.c_void_p), ("r_capacity", C.c_int64),
                ("r_total", C.c_int64), ("r_time", C.c_void_p), ("profile_time", C.c_void_p),
                ("profile_s", C.c_void_p), ("profile_sd", C.c_void_p),
                ("profile_sdd", C.c_void_p)]


class _ResampleDesc(C.Structure):
    """tpb_resample_desc (include/tpb.h)"""
    _fields_ = [("device", C.c_int32), ("num_paths", C.c_int32), ("num_samples", C.c_int32),
                ("num_dofs", C.c_int32)] + [(n, C.c_void_p) for n in (
                    "time", "s", "sd", "sdd", "q", "dq", "ddq", "max_acceleration", "status")] + [
        ("start_sec", C.c_double), ("time_step_sec", C.c_double), ("row_capacity", C.c_int64)] + [
        (n, C.c_void_p) for n in ("r_offset", "r_time", "r_s", "r_sd", "r_sdd", "r_pos", "r_vel",
                                  "r_acc")]


class _StopDesc(C.Structure):
    _fields_ = [("num_trajectories", C.c_int32), ("num_dofs", C.c_int32),
                ("total_rows", C.c_int64)] + [(n, C.c_void_p) for n in (
                    "begin", "length", "times", "velocities", "accelerations",
                    "max_acceleration", "count", "switch_index", "status", "out_times",
                    "out_velocities", "out_accelerations")]


class _SplineDesc(C.Structure):
    _fields_ = [("degree", C.c_int32), ("num_points", C.c_int32), ("dim", C.c_int32),
                ("nder", C.c_int32), ("num_splines", C.c_int32),
                ("evals_per_spline", C.c_int64), ("knots", C.c_void_p), ("ctrl", C.c_void_p),
                ("u", C.c_void_p), ("u0", C.c_void_p), ("du", C.c_void_p),
                ("out", C.c_void_p), ("status", C.c_void_p), ("exact", C.c_int32)]


class _TbufConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("num_buffers", C.c_int32), ("num_dofs", C.c_int32),
                ("capacity_rows", C.c_int32), ("timestep_tolerance", C.c_double)]


_lib = None


def lib_path() -> str:
    return _LIB


_dptr = C.POINTER(C.c_double)
# tpb_path_ik_fn / tpb_jacobian_fn (include/tpb.h)
PATH_IK_FN = C.CFUNCTYPE(C.c_int32, C.c_void_p, C.c_int32, _dptr, C.c_int32, _dptr, _dptr, _dptr)
JACOBIAN_FN = C.CFUNCTYPE(C.c_int32, C.c_void_p, C.c_int32, _dptr, _dptr)


def lib():
    """Load libtpb.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            raise FileNotFoundError(
                f"{_LIB} is missing: run __graft_entry__.build() (nvcc, sm_100a). "
                "There is no CPU fallback.")
        L = C.CDLL(_LIB)
        vp, i32, i64, f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
        L.tpb_abi_version.restype = i32
        L.tpb_create.restype = i32
        L.tpb_create.argtypes = [C.POINTER(_Config), C.POINTER(vp)]
        L.tpb_destroy.restype = i32
        L.tpb_destroy.argtypes = [vp]
        L.tpb_last_error.restype = C.c_char_p
        L.tpb_last_error.argtypes = [vp]
        L.tpb_set_stream.restype = i32
        L.tpb_set_stream.argtypes = [vp, vp]
        L.tpb_sync.restype = i32
        L.tpb_sync.argtypes = [vp]
        L.tpb_kernel_launches.restype = i64
        L.tpb_kernel_launches.argtypes = [vp]
        L.tpb_set_path_options.restype = i32
        L.tpb_set_path_options.argtypes = [vp, f64, f64]
        L.tpb_set_joint_waypoints.restype = i32
        L.tpb_set_joint_waypoints.argtypes = [vp, i32, vp, vp, i32]
        L.tpb_set_joint_limits.restype = i32
        L.tpb_set_joint_limits.argtypes = [vp, vp, vp, i32, i32]
        L.tpb_set_sampling.restype = i32
        L.tpb_set_sampling.argtypes = [vp, vp, i32]
        L.tpb_sample_paths.restype = i32
        L.tpb_sample_paths.argtypes = [vp, vp, i32]
        L.tpb_set_cart_options.restype = i32
        L.tpb_set_cart_options.argtypes = [vp, f64, f64, f64]
        L.tpb_set_cart_waypoints.restype = i32
        L.tpb_set_cart_waypoints.argtypes = [vp, i32, vp, vp, vp, i32]
        L.tpb_cart_sample_targets.restype = i32
        L.tpb_cart_sample_targets.argtypes = [vp]
        L.tpb_cart_solve_ik.restype = i32
        L.tpb_cart_solve_ik.argtypes = [vp, i32]
        L.tpb_cart_set_ik_positions.restype = i32
        L.tpb_cart_set_ik_positions.argtypes = [vp, vp, i32]
        L.tpb_cart_setup_constraints.restype = i32
        L.tpb_cart_setup_constraints.argtypes = [vp, i32, vp, i32]
        L.tpb_set_constraints.restype = i32
        L.tpb_set_constraints.argtypes = [vp, i32, i32, vp, vp, vp, i32]
        L.tpb_time_phases.restype = i32
        L.tpb_time_phases.argtypes = [vp, i32]
        L.tpb_last_phase_ms.restype = i32
        L.tpb_last_phase_ms.argtypes = [vp, vp]
        L.tpb_optimize.restype = i32
        L.tpb_optimize.argtypes = [vp, vp, vp, vp, i32, i32]
        L.tpb_resample.restype = i32
        L.tpb_resample.argtypes = [vp, f64, f64, C.POINTER(i64)]
        L.tpb_measure_fp64_fma_peak.restype = i32
        L.tpb_measure_fp64_fma_peak.argtypes = [i32, C.POINTER(f64)]
        L.tpb_plan_joint_device.restype = i32
        L.tpb_plan_joint_device.argtypes = [vp, f64, f64, i32, C.POINTER(i64)]
        L.tpb_plan_joint_host.restype = i32
        L.tpb_plan_joint_host.argtypes = [vp, C.POINTER(_PlanIO)]
        L.tpb_get.restype = i32
        L.tpb_get.argtypes = [vp, i32, vp, C.c_size_t, i32]
        L.tpb_devptr.restype = i32
        L.tpb_devptr.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(C.c_size_t)]
        L.tpb_spline_eval_device.restype = i32
        L.tpb_spline_eval_device.argtypes = [C.POINTER(_SplineDesc), vp]
        L.tpb_spline_eval_host.restype = i32
        L.tpb_spline_eval_host.argtypes = [C.POINTER(_SplineDesc), i32]
        L.tpb_resample_skipping.restype = i32
        L.tpb_resample_skipping.argtypes = [vp, C.c_double, C.c_double, C.POINTER(C.c_int64)]
        L.tpb_switch_to_waypoint_path.restype = i32
        L.tpb_switch_to_waypoint_path.argtypes = [vp, vp, vp, vp, vp, i32]
        L.tpb_switch_to_cart_waypoint_path.restype = i32
        L.tpb_switch_to_cart_waypoint_path.argtypes = [vp, vp, vp, vp, vp, vp, i32]
        L.tpb_planner_reset.restype = i32
        L.tpb_planner_reset.argtypes = [vp, i32, C.c_double, i32, i32, C.c_double, vp, i32]
        L.tpb_planner_new_path.restype = i32
        L.tpb_planner_new_path.argtypes = [vp, vp, i32]
        L.tpb_planner_set_initial_velocity.restype = i32
        L.tpb_planner_set_initial_velocity.argtypes = [vp, vp, i32]
        L.tpb_planner_set_kinematics.restype = i32
        L.tpb_planner_set_kinematics.argtypes = [vp, i32]
        L.tpb_quat_spline_eval_host.restype = i32
        L.tpb_quat_spline_eval_host.argtypes = [vp, i32]
        L.tpb_quat_spline_eval_device.restype = i32
        L.tpb_quat_spline_eval_device.argtypes = [vp, vp]
        L.tpb_count_violations.restype = i32
        L.tpb_count_violations.argtypes = [vp, vp, i32]
        L.tpb_planner_set_host_kinematics.restype = i32
        L.tpb_planner_set_host_kinematics.argtypes = [vp, PATH_IK_FN, JACOBIAN_FN, vp]
        L.tpb_planner_plan.restype = i32
        L.tpb_planner_plan.argtypes = [vp, C.c_int64, C.c_int64, C.POINTER(C.c_int64)]
        L.tpb_path_stop_parameter.restype = i32
        L.tpb_path_stop_parameter.argtypes = [vp, vp, vp, vp, vp, vp, i32]
        L.tpb_rescale_to_stop_device.restype = i32
        L.tpb_rescale_to_stop_device.argtypes = [C.POINTER(_StopDesc), vp]
        # multi-GPU front-end
        L.tpb_shard_range.restype = i32
        L.tpb_shard_range.argtypes = [i32, i32, i32, C.POINTER(i32), C.POINTER(i32)]
        L.tpb_multi_create.restype = i32
        L.tpb_multi_create.argtypes = [C.POINTER(_Config), C.POINTER(i32), i32, C.POINTER(vp)]
        L.tpb_multi_destroy.restype = i32
        L.tpb_multi_destroy.argtypes = [vp]
        L.tpb_multi_last_error.restype = C.c_char_p
        L.tpb_multi_last_error.argtypes = [vp]
        L.tpb_multi_num_shards.restype = i32
        L.tpb_multi_num_shards.argtypes = [vp]
        L.tpb_multi_kernel_launches.restype = C.c_int64
        L.tpb_multi_kernel_launches.argtypes = [vp]
        L.tpb_multi_shard.restype = i32
        L.tpb_multi_shard.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32),
                                      C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(vp)]
        L.tpb_multi_set_joint_paths.restype = i32
        L.tpb_multi_set_joint_paths.argtypes = [vp, i32, vp, vp, vp, vp]
        L.tpb_multi_plan_joint_device.restype = i32
        L.tpb_multi_plan_joint_device.argtypes = [vp, C.c_double, C.c_double,
                                                  C.POINTER(C.c_int64)]
        L.tpb_multi_gather.restype = i32
        L.tpb_multi_gather.argtypes = [vp, i32, vp, C.c_size_t]
        L.tpb_multi_plan_joint_host.restype = i32
        L.tpb_multi_plan_joint_host.argtypes = [vp, C.POINTER(_PlanIO)]
        L.tpb_query_profile.restype = i32
        L.tpb_query_profile.argtypes = [vp, i32] + [vp] * 10 + [i32]
        L.tpb_set_kernel_variant.restype = i32
        L.tpb_set_kernel_variant.argtypes = [vp, i32, i32]
        L.tpb_resample_uniform_noalloc.restype = i32
        L.tpb_resample_uniform_noalloc.argtypes = [C.POINTER(_ResampleDesc), vp, C.c_char_p,
                                                   C.c_size_t]
        L.tpb_rescale_to_stop_host.restype = i32
        L.tpb_rescale_to_stop_host.argtypes = [C.POINTER(_StopDesc), i32]
        # trajectory buffers
        L.tpb_tbuf_create.restype = i32
        L.tpb_tbuf_create.argtypes = [C.POINTER(_TbufConfig), C.POINTER(vp)]
        L.tpb_tbuf_destroy.restype = i32
        L.tpb_tbuf_destroy.argtypes = [vp]
        L.tpb_tbuf_last_error.restype = C.c_char_p
        L.tpb_tbuf_last_error.argtypes = [vp]
        L.tpb_tbuf_set_stream.restype = i32
        L.tpb_tbuf_set_stream.argtypes = [vp, vp]
        L.tpb_tbuf_sync.restype = i32
        L.tpb_tbuf_sync.argtypes = [vp]
        L.tpb_tbuf_kernel_launches.restype = i64
        L.tpb_tbuf_kernel_launches.argtypes = [vp]
        L.tpb_tbuf_reserve.restype = i32
        L.tpb_tbuf_reserve.argtypes = [vp, i32]
        L.tpb_tbuf_clear.restype = i32
        L.tpb_tbuf_clear.argtypes = [vp, vp, i32]
        L.tpb_tbuf_insert_segments.restype = i32
        L.tpb_tbuf_insert_segments.argtypes = [vp, vp, vp, i64, vp, vp, vp, vp, vp, i32]
        L.tpb_tbuf_append_samples.restype = i32
        L.tpb_tbuf_append_samples.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32]
        L.tpb_tbuf_discard_before.restype = i32
        L.tpb_tbuf_discard_before.argtypes = [vp, vp, i32]
        L.tpb_tbuf_positions_up_to_time.restype = i32
        L.tpb_tbuf_positions_up_to_time.argtypes = [vp, vp, vp, i32]
        L.tpb_tbuf_sample.restype = i32
        L.tpb_tbuf_sample.argtypes = [vp, i32, vp, vp, vp, vp, vp, i32]
        L.tpb_tbuf_stop_at_index.restype = i32
        L.tpb_tbuf_stop_at_index.argtypes = [vp, vp, vp, f64, vp, i32]
        L.tpb_tbuf_stop_before_time.restype = i32
        L.tpb_tbuf_stop_before_time.argtypes = [vp, vp, vp, f64, vp, i32]
        L.tpb_tbuf_add_offset.restype = i32
        L.tpb_tbuf_add_offset.argtypes = [vp, vp, i32]
        L.tpb_tbuf_get_state.restype = i32
        L.tpb_tbuf_get_state.argtypes = [vp, vp, vp, vp, vp, i32]
        L.tpb_tbuf_get_rows.restype = i32
        L.tpb_tbuf_get_rows.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp, i32]
        L.tpb_tbuf_devptr.restype = i32
        L.tpb_tbuf_devptr.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(C.c_size_t)]
        _lib = L
    return _lib


def _ptr(x):
    """(pointer, on_device) of a numpy array / torch tensor / None."""
    if x is None:
        return None, 0
    if isinstance(x, np.ndarray):
        assert x.flags["C_CONTIGUOUS"]
        return x.ctypes.data, 0
    if hasattr(x, "data_ptr"):
        assert x.is_contiguous()
        return x.data_ptr(), 1 if x.is_cuda else 0
    raise TypeError(type(x))


def _f64(x):
    if x is None or hasattr(x, "data_ptr"):
        return x
    return np.ascontiguousarray(x, dtype=np.float64)


_DTYPES = {ARR.NUM_POINTS: np.int32, ARR.STATUS: np.int32, ARR.LAST_EXTREMAL: np.int32,
           ARR.B_TYPE: np.uint8, ARR.R_OFFSET: np.int64, ARR.W_LEN: np.int32,
           ARR.SWITCH_STATUS: np.int32,
           ARR.PLAN_STATUS: np.int32, ARR.PLAN_LOOPS: np.int32,
           ARR.PLAN_PLANNED_TO_END: np.int32, ARR.PLAN_TARGET_REACHED: np.int32,
           ARR.PLAN_START_TIME: np.int64, ARR.PLAN_END_TIME: np.int64,
           ARR.PLAN_FINAL_DECEL_START: np.int64, ARR.PLAN_ERASED: np.int32}


class Engine:
    """One engine = one GPU + one stream (tpb_engine)."""

    def __init__(self, max_paths, num_samples, num_dofs=0, max_waypoints=0,
                 max_constraints=0, max_time_samples_total=0, device=0, cartesian=False):
        self._h = C.c_void_p()
        self.max_paths, self.N, self.D = max_paths, num_samples, num_dofs
        self.Wmax = max_waypoints
        self.max_points = max(3 * max_waypoints - 2, 4) if num_dofs else 0
        self.B = 0
        self.C = 0
        cfg = _Config(device, max_paths, num_dofs, max_waypoints, num_samples,
                      max_constraints, max_time_samples_total, 1 if cartesian else 0)
        rc = lib().tpb_create(C.byref(cfg), C.byref(self._h))
        if rc != 0:
            self._h = C.c_void_p()
            raise TpbError(rc, "tpb_create failed (no CUDA device / bad config / out of memory)")

    def close(self):
        if self._h:
            lib().tpb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise TpbError(rc, lib().tpb_last_error(self._h).decode())

    # -- configuration -------------------------------------------------------
    def set_stream(self, cuda_stream_ptr):
        self._ck(lib().tpb_set_stream(self._h, cuda_stream_ptr))

    def sync(self):
        self._ck(lib().tpb_sync(self._h))

    def query_profile(self, times):
        """tpb_query_profile on host times [B][Q]: dict(s, sd, sdd, previous_index, sk, sdk,
        sddk, tk, status), each [B][Q]."""
        t = _f64(times)
        B, Q = t.shape
        assert B == self.B
        out = {k: np.empty((B, Q)) for k in ("s", "sd", "sdd", "sk", "sdk", "sddk", "tk")}
        out["previous_index"] = np.empty((B, Q), np.int32)
        out["status"] = np.empty((B, Q), np.int32)
        self._ck(lib().tpb_query_profile(
            self._h, Q, t.ctypes.data, out["s"].ctypes.data, out["sd"].ctypes.data,
            out["sdd"].ctypes.data, out["previous_index"].ctypes.data, out["sk"].ctypes.data,
            out["sdk"].ctypes.data, out["sddk"].ctypes.data, out["tk"].ctypes.data,
            out["status"].ctypes.data, 0))
        return out

    VARIANTS = dict(generic_limit_local=0, integrate_warp=1, cart_generic=2, cart_unfused=3)

    def set_kernel_variant(self, name, on=True):
        """tpb_set_kernel_variant: use a cross-check kernel instead of the default one."""
        self._ck(lib().tpb_set_kernel_variant(self._h, self.VARIANTS[name], 1 if on else 0))

    @property
    def kernel_launches(self):
        return lib().tpb_kernel_launches(self._h)

    def set_path_options(self, constraint_safety=0.8, rounding=0.2):
        self._ck(lib().tpb_set_path_options(self._h, constraint_safety, rounding))

    # -- joint paths ---------------------------------------------------------
    def set_joint_waypoints(self, waypoints, num_waypoints=None):
        waypoints = _f64(waypoints)
        B = waypoints.shape[0]
        assert tuple(waypoints.shape[1:]) == (self.Wmax, self.D), waypoints.shape
        wp, dev = _ptr(waypoints)
        if num_waypoints is not None and not hasattr(num_waypoints, "data_ptr"):
            num_waypoints = np.ascontiguousarray(num_waypoints, np.int32)
        nw, dev2 = _ptr(num_waypoints)
        assert num_waypoints is None or dev == dev2
        self._ck(lib().tpb_set_joint_waypoints(self._h, B, nw, wp, dev))
        self.B, self.C = B, 2 * self.D

    def set_joint_limits(self, vmax, amax):
        vmax, amax = _f64(vmax), _f64(amax)
        per_path = 1 if vmax.ndim == 2 else 0
        pv, dev = _ptr(vmax)
        pa, dev2 = _ptr(amax)
        assert dev == dev2
        self._ck(lib().tpb_set_joint_limits(self._h, pv, pa, per_path, dev))

    def set_sampling(self, delta=None):
        delta = _f64(delta)
        p, dev = _ptr(delta)
        self._ck(lib().tpb_set_sampling(self._h, p, dev))

    def sample_paths(self, path_start=None):
        path_start = _f64(path_start)
        p, dev = _ptr(path_start)
        self._ck(lib().tpb_sample_paths(self._h, p, dev))

    # -- Cartesian paths -----------------------------------------------------
    def set_cart_options(self, translation_rounding=0.05, max_translational_velocity=0.5,
                         max_rotational_velocity=0.25):
        self._ck(lib().tpb_set_cart_options(self._h, translation_rounding,
                                            max_translational_velocity,
                                            max_rotational_velocity))

    def set_cart_waypoints(self, pose_waypoints, joint_waypoints, num_waypoints=None):
        pose_waypoints, joint_waypoints = _f64(pose_waypoints), _f64(joint_waypoints)
        B = joint_waypoints.shape[0]
        assert tuple(joint_waypoints.shape[1:]) == (self.Wmax, self.D)
        assert tuple(pose_waypoints.shape) == (B, self.Wmax, 7)
        pp, dev = _ptr(pose_waypoints)
        pj, dev2 = _ptr(joint_waypoints)
        assert dev == dev2
        if num_waypoints is not None and not hasattr(num_waypoints, "data_ptr"):
            num_waypoints = np.ascontiguousarray(num_waypoints, np.int32)
        nw, _ = _ptr(num_waypoints)
        self._ck(lib().tpb_set_cart_waypoints(self._h, B, nw, pp, pj, dev))
        self.B, self.C = B, 2 * self.D + 2

    def cart_sample_targets(self):
        self._ck(lib().tpb_cart_sample_targets(self._h))

    def cart_solve_ik(self, model=1):
        self._ck(lib().tpb_cart_solve_ik(self._h, model))

    def cart_set_ik_positions(self, positions):
        positions = _f64(positions)
        p, dev = _ptr(positions)
        self._ck(lib().tpb_cart_set_ik_positions(self._h, p, dev))

    def cart_setup_constraints(self, model=1, jacobians=None):
        jacobians = _f64(jacobians)
        p, dev = _ptr(jacobians)
        self._ck(lib().tpb_cart_setup_constraints(self._h, model, p, dev))

    # -- explicit constraints --------------------------------------------------
    def set_constraints(self, cons, s_start, s_end):
        cons, s_start, s_end = _f64(cons), _f64(s_start), _f64(s_end)
        B, N, Cn, four = cons.shape
        assert N == self.N and four == 4
        pc, dev = _ptr(cons)
        p0, _ = _ptr(s_start)
        p1, _ = _ptr(s_end)
        self._ck(lib().tpb_set_constraints(self._h, B, Cn, pc, p0, p1, dev))
        self.B, self.C = B, Cn

    # -- stages ----------------------------------------------------------------
    def optimize(self, sd_start=None, sdd_start=None, time_start=None, max_loops=0):
        ptrs, dev = [], 0
        for x in (sd_start, sdd_start, time_start):
            p, d = _ptr(_f64(x))
            ptrs.append(p)
            dev = dev or d
        self._ck(lib().tpb_optimize(self._h, *ptrs, max_loops, dev))

    def path_stop_parameter(self, time_sec):
        """Batched GetPathStopParameter on the resampled trajectories (host arrays).
        Returns dict(parameter [B], row [B], duration [B], status [B])."""
        t = _f64(np.broadcast_to(np.asarray(time_sec, np.float64), (self.B,)).copy())
        out = dict(parameter=np.zeros(self.B), row=np.zeros(self.B, np.int32),
                   duration=np.zeros(self.B), status=np.zeros(self.B, np.int32))
        self._ck(lib().tpb_path_stop_parameter(self._h, t.ctypes.data, out["parameter"].ctypes.data,
                                               out["row"].ctypes.data, out["duration"].ctypes.data,
                                               out["status"].ctypes.data, 0))
        return out

    def time_phases(self, on=True):
        """Record CUDA events around the three phases of every optimize()."""
        self._ck(lib().tpb_time_phases(self._h, 1 if on else 0))

    def last_phase_ms(self):
        """(setup, sweeps, finalize) milliseconds of the last timed optimize()."""
        ms = (C.c_double * 3)()
        self._ck(lib().tpb_last_phase_ms(self._h, C.cast(ms, C.c_void_p)))
        return [ms[0], ms[1], ms[2]]

    def resample(self, start_sec, time_step_sec, want_total=True, skip=False):
        """tpb_resample (uniform in time) or, with skip, tpb_resample_skipping."""
        total = C.c_int64(0)
        fn = lib().tpb_resample_skipping if skip else lib().tpb_resample
        self._ck(fn(self._h, start_sec, time_step_sec, C.byref(total) if want_total else None))
        return total.value

    def switch_to_waypoint_path(self, keep_path_until, waypoints, num_waypoints=None,
                                initial_velocity=None):
        """tpb_switch_to_waypoint_path: SwitchToWaypointPath for every path of the batch
        (num_waypoints[b] <= 0 leaves path b alone).  Returns the per-path status codes."""
        keep = _f64(keep_path_until)
        wp = _f64(waypoints)
        assert tuple(wp.shape[1:]) == (self.Wmax, self.D), wp.shape
        if num_waypoints is not None and not hasattr(num_waypoints, "data_ptr"):
            num_waypoints = np.ascontiguousarray(num_waypoints, np.int32)
        v0 = _f64(initial_velocity)
        pk, dev = _ptr(keep)
        self._ck(lib().tpb_switch_to_waypoint_path(self._h, pk, _ptr(num_waypoints)[0],
                                                   _ptr(wp)[0], _ptr(v0)[0], dev))
        return self.get(ARR.SWITCH_STATUS)

    def switch_to_cart_waypoint_path(self, keep_path_until, pose_waypoints, joint_waypoints,
                                     num_waypoints=None, initial_velocity=None):
        """tpb_switch_to_cart_waypoint_path: TimeableCartesianSplinePath::SwitchToWaypointPath
        for every path of the batch.  Returns the per-path status codes."""
        keep = _f64(keep_path_until)
        pw, jw = _f64(pose_waypoints), _f64(joint_waypoints)
        assert tuple(pw.shape[1:]) == (self.Wmax, 7) and tuple(jw.shape[1:]) == (self.Wmax, self.D)
        if num_waypoints is not None and not hasattr(num_waypoints, "data_ptr"):
            num_waypoints = np.ascontiguousarray(num_waypoints, np.int32)
        v0 = _f64(initial_velocity)
        pk, dev = _ptr(keep)
        self._ck(lib().tpb_switch_to_cart_waypoint_path(self._h, pk, _ptr(num_waypoints)[0],
                                                        _ptr(pw)[0], _ptr(jw)[0], _ptr(v0)[0], dev))
        return self.get(ARR.SWITCH_STATUS)

    def planner_reset(self, path_sample_capacity, time_step_sec, skip=False,
                      max_planning_iterations=200, max_initial_velocity_error=1e-4,
                      initial_velocity=None):
        """tpb_planner_reset: PathTimingTrajectory::Reset + its options for every path."""
        v0 = _f64(initial_velocity)  # stays alive across the call
        p, dev = _ptr(v0)
        self._ck(lib().tpb_planner_reset(self._h, int(path_sample_capacity), time_step_sec,
                                         1 if skip else 0, max_planning_iterations,
                                         max_initial_velocity_error, p, dev))
        self.P = int(path_sample_capacity)

    def planner_new_path(self, initial_velocity=None):
        """tpb_planner_new_path: new waypoints on a planner that keeps its plan."""
        v0 = _f64(initial_velocity)
        p, dev = _ptr(v0)
        self._ck(lib().tpb_planner_new_path(self._h, p, dev))

    def planner_set_initial_velocity(self, initial_velocity=None):
        v0 = _f64(initial_velocity)
        p, dev = _ptr(v0)
        self._ck(lib().tpb_planner_set_initial_velocity(self._h, p, dev))

    def planner_set_kinematics(self, model=1):
        self._ck(lib().tpb_planner_set_kinematics(self._h, int(model)))

    def count_violations(self):
        """tpb_count_violations: SolutionSatisfiesConstraints per path (-1: no solution)."""
        out = np.zeros(self.B, np.int32)
        self._ck(lib().tpb_count_violations(self._h, out.ctypes.data, 0))
        return out

    def planner_set_host_kinematics(self, path_ik, jacobian):
        """tpb_planner_set_host_kinematics: the user's PathIKFunc / JacobianFunc for
        Cartesian paths behind the planner, as Python callables
            path_ik(path, initial_value [D], pose_targets [M][7], joint_targets [M][D])
                -> solution [M][D]   (raise to fail the plan)
            jacobian(path, joint_positions [D]) -> [6][D]
        called on the host once per planning window / path sample."""
        D = self.D

        def ik_thunk(_user, path, initial, M, pt, jt, sol):
            try:
                out = path_ik(path, np.ctypeslib.as_array(initial, (D,)),
                              np.ctypeslib.as_array(pt, (M, 7)), np.ctypeslib.as_array(jt, (M, D)))
                np.ctypeslib.as_array(sol, (M, D))[:] = out
                return 0
            except Exception:  # noqa: BLE001 -- a status for the C caller
                return 13

        def jac_thunk(_user, path, q, out):
            try:
                np.ctypeslib.as_array(out, (6, D))[:] = jacobian(path, np.ctypeslib.as_array(q, (D,)))
                return 0
            except Exception:  # noqa: BLE001
                return 13

        self._host_kinematics = (PATH_IK_FN(ik_thunk), JACOBIAN_FN(jac_thunk))  # keep alive
        self._ck(lib().tpb_planner_set_host_kinematics(self._h, self._host_kinematics[0],
                                                       self._host_kinematics[1], None))

    def planner_plan(self, start_ns, horizon_ns, want_total=True):
        """tpb_planner_plan: PathTimingTrajectory::Plan(start, horizon) for every path
        (times in integer nanoseconds).  Returns the number of trajectory rows."""
        total = C.c_int64(0)
        self._ck(lib().tpb_planner_plan(self._h, int(start_ns), int(horizon_ns),
                                        C.byref(total) if want_total else None))
        return total.value

    def plan_joint_device(self, start_sec=0.0, time_step_sec=0.0, num_chunks=1, want_total=True):
        """tpb_plan_joint_device: the whole hot path for the resident batch."""
        total = C.c_int64(0)
        self._ck(lib().tpb_plan_joint_device(self._h, start_sec, time_step_sec, int(num_chunks),
                                             C.byref(total) if want_total else None))
        return total.value

    def plan_joint_host(self, waypoints, vmax, amax, start_sec=0.0, time_step_sec=0.0,
                        num_waypoints=None, out=None):
        """tpb_plan_joint_host.  `out`: dict of preallocated (ideally pinned) host
        arrays keyed duration,status,r_offset,r_pos,r_vel,r_acc,r_s,r_sd,r_sdd,r_time and
        profile_time,profile_s,profile_sd,profile_sdd ([B][N]); only those given are copied."""
        B = waypoints.shape[0]
        out = out or {}
        io = _PlanIO()
        io.num_paths = B
        io.num_waypoints = _ptr(num_waypoints)[0]
        io.waypoints = _ptr(waypoints)[0]
        io.max_velocity = _ptr(vmax)[0]
        io.max_acceleration = _ptr(amax)[0]
        io.start_sec, io.time_step_sec = start_sec, time_step_sec
        cap = None
        for k in ("duration", "status", "r_offset", "r_pos", "r_vel", "r_acc", "r_s",
                  "r_sd", "r_sdd", "r_time", "profile_time", "profile_s", "profile_sd",
                  "profile_sdd"):
            v = out.get(k)
            setattr(io, k, _ptr(v)[0] if v is not None else None)
            if v is not None and k in ("r_pos", "r_vel", "r_acc", "r_s", "r_sd", "r_sdd",
                                       "r_time"):
                rows = v.shape[0]
                cap = rows if cap is None else min(cap, rows)
        io.r_capacity = cap or 0
        self._ck(lib().tpb_plan_joint_host(self._h, C.byref(io)))
        self.B, self.C = B, 2 * self.D
        return io.r_total

    # -- results ---------------------------------------------------------------
    def devptr(self, what):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(lib().tpb_devptr(self._h, what, C.byref(p), C.byref(n)))
        return p.value, n.value

    def get(self, what):
        """Copy array `what` to a new numpy array shaped like the reference's getter."""
        _, nbytes = self.devptr(what)
        dt = np.dtype(_DTYPES.get(what, np.float64))
        out = np.empty(nbytes // dt.itemsize, dt)
        if nbytes:
            self._ck(lib().tpb_get(self._h, what, out.ctypes.data, nbytes, 0))
        B, N, D = self.B, self.N, self.D
        if what in (ARR.Q, ARR.DQ, ARR.DDQ):
            return out.reshape(B, N, D)
        if what == ARR.CTRL:  # the per-path capacity grows with the first switch
            return out.reshape(B, -1, D)
        if what == ARR.KNOTS:
            return out.reshape(B, -1)
        if what == ARR.POSE_CTRL:
            return out.reshape(B, -1, 7)  # grows with the first switch, like CTRL
        if what == ARR.POSE_TARGETS:
            return out.reshape(B, N + 1, 7)
        if what == ARR.JOINT_TARGETS:
            return out.reshape(B, N + 1, D)
        if what == ARR.CONS:
            return out.reshape(B, N, self.C, 4)
        if what in (ARR.R_POS, ARR.R_VEL, ARR.R_ACC):
            return out.reshape(-1, D)
        if what in (ARR.W_Q, ARR.W_DQ, ARR.W_DDQ):
            return out.reshape(B, self.P, D)
        if ARR.W_TIME <= what <= ARR.W_SDD:
            return out.reshape(B, self.P)
        if ARR.TIME <= what <= ARR.B_TYPE:
            return out.reshape(B, N)
        return out


def _desc(degree, num_points, dim, nder, num_splines, evals, knots, ctrl, u, u0, du, out,
          status, exact):
    d = _SplineDesc(degree, num_points, dim, nder, num_splines, evals)
    d.knots, d.ctrl, d.u, d.u0, d.du = (_ptr(x)[0] for x in (knots, ctrl, u, u0, du))
    # exact: True / 1 = A2.3 operation order; False / 0 = polynomial form with bulk-copy
    # tiles; -1 = polynomial form with per-thread stores (A/B)
    mode = exact if (isinstance(exact, int) and not isinstance(exact, bool)) else (1 if exact else 0)
    d.out, d.status, d.exact = _ptr(out)[0], _ptr(status)[0], mode
    return d


def spline_eval_host(knots, ctrl, degree, nder, u=None, u0=None, du=None, evals=None,
                     exact=True, out=None, device=0):
    """tpb_spline_eval_host.  knots [S][n+p+1], ctrl [S][n][dim]; u [S][E] or grid."""
    knots, ctrl = _f64(knots), _f64(ctrl)
    S, n, dim = ctrl.shape
    if u is not None:
        u = _f64(u)
        evals = u.shape[1]
    else:
        u0, du = _f64(u0), _f64(du)
    if out is None:
        out = np.zeros((S, evals, nder + 1, dim))
    status = np.zeros((S, evals), np.int32)
    d = _desc(degree, n, dim, nder, S, evals, knots, ctrl, u, u0, du, out, status, exact)
    rc = lib().tpb_spline_eval_host(C.byref(d), device)
    if rc != 0:
        raise TpbError(rc, "tpb_spline_eval_host")
    return out, status


def spline_eval_device(knots, ctrl, degree, nder, out, u=None, u0=None, du=None, evals=None,
                       status=None, exact=True, stream=None):
    """tpb_spline_eval_device on torch CUDA tensors (asynchronous on `stream`)."""
    S, n, dim = ctrl.shape
    if u is not None:
        evals = u.shape[1]
    d = _desc(degree, n, dim, nder, S, evals, knots, ctrl, u, u0, du, out, status, exact)
    rc = lib().tpb_spline_eval_device(C.byref(d), stream)
    if rc != 0:
        raise TpbError(rc, "tpb_spline_eval_device")


class _QuatSplineDesc(C.Structure):
    _fields_ = [("degree", C.c_int32), ("num_points", C.c_int32), ("num_splines", C.c_int32),
                ("evals_per_spline", C.c_int64), ("knots", C.c_void_p), ("quats", C.c_void_p),
                ("u", C.c_void_p), ("u0", C.c_void_p), ("du", C.c_void_p), ("out", C.c_void_p),
                ("dout", C.c_void_p), ("status", C.c_void_p)]


def quat_spline_eval_host(knots, quats, degree, u=None, u0=None, du=None, evals=None,
                          derivative=False, device=0):
    """tpb_quat_spline_eval_host: BSplineQ::EvalCurve (derivative=False) or
    EvalCurveAndDerivative for knots [S][n+p+1], quats [S][n][4] (x, y, z, w); u [S][E] or
    the grid u0[s] + i * du[s].  Returns (q [S][E][4], dq or None, status [S][E])."""
    knots, quats = _f64(knots), _f64(quats)
    S, n, _ = quats.shape
    if u is not None:
        u = _f64(u)
        evals = u.shape[1]
    else:
        u0, du = _f64(u0), _f64(du)
    out = np.full((S, evals, 4), np.nan)
    dout = np.full((S, evals, 4), np.nan) if derivative else None
    status = np.zeros((S, evals), np.int32)
    d = _QuatSplineDesc(degree, n, S, evals, _addr(knots), _addr(quats), _addr(u), _addr(u0),
                        _addr(du), _addr(out), _addr(dout), _addr(status))
    rc = lib().tpb_quat_spline_eval_host(C.byref(d), device)
    if rc != 0:
        raise TpbError(rc, "tpb_quat_spline_eval_host")
    return out, dout, status


def quat_spline_eval_device(knots, quats, degree, out, dout=None, u=None, u0=None, du=None,
                            evals=None, status=None, stream=None):
    """tpb_quat_spline_eval_device on torch CUDA tensors (asynchronous on `stream`)."""
    S, n, _ = quats.shape
    if u is not None:
        evals = u.shape[1]
    d = _QuatSplineDesc(degree, n, S, evals, _addr(knots), _addr(quats), _addr(u), _addr(u0),
                        _addr(du), _addr(out), _addr(dout), _addr(status))
    rc = lib().tpb_quat_spline_eval_device(C.byref(d), stream)
    if rc != 0:
        raise TpbError(rc, "tpb_quat_spline_eval_device")


def _addr(x):
    """Address of a numpy array or of a torch CUDA tensor."""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    return x.data_ptr()


def rescale_to_stop_host(max_acceleration, begin, length, times, velocities, accelerations,
                         device=0):
    """tpb_rescale_to_stop_host on packed numpy trajectories.

    times [R], velocities/accelerations [R][D]; trajectory b = rows begin[b] ..
    begin[b]+length[b]-1; max_acceleration [T][D].  Returns dict(count, switch_index,
    status, times, velocities, accelerations): the out_* arrays have the layout of the
    inputs, the stopping trajectory of b is rows begin[b]+switch_index[b] ..
    begin[b]+length[b]-1 of them (NaN elsewhere).
    """
    amax = _f64(max_acceleration)
    t, v, a = _f64(times), _f64(velocities), _f64(accelerations)
    begin = np.ascontiguousarray(begin, np.int64)
    length = np.ascontiguousarray(length, np.int32)
    T, D = amax.shape
    R = t.shape[0]
    out = dict(count=np.zeros(T, np.int32), switch_index=np.zeros(T, np.int32),
               status=np.zeros(T, np.int32), times=np.full(R, np.nan),
               velocities=np.full((R, D), np.nan), accelerations=np.full((R, D), np.nan))
    d = _StopDesc(T, D, R, _addr(begin), _addr(length), _addr(t), _addr(v), _addr(a), _addr(amax),
                  _addr(out["count"]), _addr(out["switch_index"]), _addr(out["status"]),
                  _addr(out["times"]), _addr(out["velocities"]), _addr(out["accelerations"]))
    rc = lib().tpb_rescale_to_stop_host(C.byref(d), device)
    if rc != 0:
        raise TpbError(rc, "tpb_rescale_to_stop_host")
    return out


def rescale_to_stop_device(max_acceleration, begin, length, times, velocities, accelerations,
                           count, switch_index, status, out_times, out_velocities,
                           out_accelerations, stream=None):
    """tpb_rescale_to_stop_device on torch CUDA tensors (asynchronous on `stream`)."""
    T, D = max_acceleration.shape
    R = times.shape[0]
    d = _StopDesc(T, D, R, _addr(begin), _addr(length), _addr(times), _addr(velocities),
                  _addr(accelerations), _addr(max_acceleration), _addr(count),
                  _addr(switch_index), _addr(status), _addr(out_times), _addr(out_velocities),
                  _addr(out_accelerations))
    rc = lib().tpb_rescale_to_stop_device(C.byref(d), stream)
    if rc != 0:
        raise TpbError(rc, "tpb_rescale_to_stop_device")


def shard_range(num_paths, num_shards, shard):
    """tpb_shard_range: (first, count) of the contiguous block shard `shard` owns."""
    f, c = C.c_int32(), C.c_int32()
    rc = lib().tpb_shard_range(num_paths, num_shards, shard, C.byref(f), C.byref(c))
    if rc != 0:
        raise TpbError(rc, "tpb_shard_range")
    return f.value, c.value


class MultiEngine:
    """tpb_multi: one batch sharded over several GPUs (or logical shards of one GPU) in
    contiguous blocks, one engine + host thread per shard, no data-path collective."""

    def __init__(self, max_paths, num_samples, num_dofs, max_waypoints, devices=None,
                 num_shards=None, max_time_samples_total=0):
        cfg = _Config(0, max_paths, num_dofs, max_waypoints, num_samples, 0,
                      max_time_samples_total, 0)
        G = num_shards or (len(devices) if devices is not None else 1)
        dev = None
        if devices is not None:
            assert len(devices) == G
            dev = (C.c_int32 * G)(*devices)
        self._h = C.c_void_p()
        rc = lib().tpb_multi_create(C.byref(cfg), dev, G, C.byref(self._h))
        if rc != 0:
            raise TpbError(rc, "tpb_multi_create")
        self.G, self.N, self.D, self.W = G, num_samples, num_dofs, max_waypoints
        self.B = 0

    def _ck(self, rc):
        if rc != 0:
            raise TpbError(rc, (lib().tpb_multi_last_error(self._h) or b"").decode())

    def close(self):
        if self._h:
            lib().tpb_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def kernel_launches(self):
        return lib().tpb_multi_kernel_launches(self._h)

    def shard(self, g):
        dev, f, n = C.c_int32(), C.c_int32(), C.c_int32()
        r0, rn, e = C.c_int64(), C.c_int64(), C.c_void_p()
        self._ck(lib().tpb_multi_shard(self._h, g, C.byref(dev), C.byref(f), C.byref(n),
                                       C.byref(r0), C.byref(rn), C.byref(e)))
        return dict(device=dev.value, first_path=f.value, num_paths=n.value,
                    first_row=r0.value, num_rows=rn.value, engine=e.value)

    def set_joint_paths(self, waypoints, vmax, amax, num_waypoints=None):
        wp, vm, am = _f64(waypoints), _f64(vmax), _f64(amax)
        nw = None if num_waypoints is None else np.ascontiguousarray(num_waypoints, np.int32)
        self._keep = (wp, vm, am, nw)
        self.B = wp.shape[0]
        self._ck(lib().tpb_multi_set_joint_paths(
            self._h, self.B, nw.ctypes.data if nw is not None else None, wp.ctypes.data,
            vm.ctypes.data, am.ctypes.data))

    def plan_device(self, start_sec=0.0, time_step_sec=0.0):
        total = C.c_int64(0)
        self._ck(lib().tpb_multi_plan_joint_device(self._h, start_sec, time_step_sec,
                                                   C.byref(total)))
        return total.value

    def gather(self, what, shape_per_path=(), dtype=np.float64):
        out = np.empty((self.B,) + tuple(shape_per_path), dtype)
        self._ck(lib().tpb_multi_gather(self._h, what, out.ctypes.data,
                                        out.nbytes // max(self.B, 1)))
        return out

    def plan_joint_host(self, waypoints, vmax, amax, start_sec=0.0, time_step_sec=0.0,
                        num_waypoints=None, out=None):
        """tpb_multi_plan_joint_host; arguments as Engine.plan_joint_host."""
        B = waypoints.shape[0]
        out = out or {}
        io = _PlanIO()
        io.num_paths = B
        io.num_waypoints = _ptr(num_waypoints)[0]
        io.waypoints = _ptr(waypoints)[0]
        io.max_velocity = _ptr(vmax)[0]
        io.max_acceleration = _ptr(amax)[0]
        io.start_sec, io.time_step_sec = start_sec, time_step_sec
        cap = None
        for k in ("duration", "status", "r_offset", "r_pos", "r_vel", "r_acc", "r_s",
                  "r_sd", "r_sdd", "r_time", "profile_time", "profile_s", "profile_sd",
                  "profile_sdd"):
            v = out.get(k)
            setattr(io, k, _ptr(v)[0] if v is not None else None)
            if v is not None and k in ("r_pos", "r_vel", "r_acc", "r_s", "r_sd", "r_sdd",
                                       "r_time"):
                rows = v.shape[0]
                cap = rows if cap is None else min(cap, rows)
        io.r_capacity = cap or 0
        self.B = B
        self._ck(lib().tpb_multi_plan_joint_host(self._h, C.byref(io)))
        return io.r_total


_hostapi = None


def hostapi():
    """lib/libtpb_hostapi.so: C entry points over the host C++ class
    PathTimingTrajectoryBatch (host/src/hostapi.cc)."""
    global _hostapi
    if _hostapi is None:
        path = os.path.join(os.path.dirname(lib_path()), "libtpb_hostapi.so")
        if not os.path.exists(path):
            raise TpbError(9, f"{path} is missing: run make -C host")
        H = C.CDLL(path)
        vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
        H.tpbh_batch_create.restype = vp
        H.tpbh_batch_create.argtypes = [i32, i32, i64, C.c_double, C.c_double, i32, i32, i64,
                                        C.POINTER(i32), i32]
        H.tpbh_batch_destroy.argtypes = [vp]
        H.tpbh_batch_plan.restype = i32
        H.tpbh_batch_plan.argtypes = [vp, C.c_uint32, i32, vp, vp, vp, vp, i64]
        H.tpbh_batch_error.restype = C.c_char_p
        H.tpbh_batch_error.argtypes = [vp]
        H.tpbh_batch_rows.restype = i64
        H.tpbh_batch_rows.argtypes = [vp]
        H.tpbh_batch_array.restype = vp
        H.tpbh_batch_array.argtypes = [vp, i32, C.POINTER(i64)]
        _hostapi = H
    return _hostapi


class HostBatch:
    """The host C++ class PathTimingTrajectoryBatch (path_timing_batch.h) through
    libtpb_hostapi.so.  Results are views of the object's page-locked arrays, valid until
    the next plan()."""
    WANT = dict(duration=1, positions=2, velocities=4, accelerations=8, path_parameters=16,
                time=32, profile=64, trajectory=1 | 2 | 4 | 8 | 16)
    _ARRAYS = dict(duration=(0, np.float64), status=(1, np.int32), offset=(2, np.int64),
                   positions=(3, np.float64), velocities=(4, np.float64),
                   accelerations=(5, np.float64), path_parameter=(6, np.float64),
                   path_parameter_derivative=(7, np.float64),
                   second_path_parameter_derivative=(8, np.float64), time=(9, np.float64),
                   profile_time=(10, np.float64), profile_s=(11, np.float64),
                   profile_sd=(12, np.float64), profile_sdd=(13, np.float64))

    def __init__(self, num_dofs, num_path_samples, time_step_ns, max_paths, max_waypoints,
                 max_time_samples_total, devices=(0,), constraint_safety=0.8, rounding=0.2):
        dev = (C.c_int32 * len(devices))(*devices)
        self._h = hostapi().tpbh_batch_create(num_dofs, num_path_samples, time_step_ns,
                                              constraint_safety, rounding, max_paths,
                                              max_waypoints, max_time_samples_total, dev,
                                              len(devices))
        self.D, self.N = num_dofs, num_path_samples

    def close(self):
        if self._h:
            hostapi().tpbh_batch_destroy(self._h)
            self._h = None

    def plan(self, waypoints, vmax, amax, want="trajectory", num_waypoints=None, start_ns=0):
        """PathTimingTrajectoryBatch::set_wanted(want) + Plan(); returns the row count."""
        mask = 0
        for w in ([want] if isinstance(want, str) else want):
            mask |= self.WANT[w]
        rc = hostapi().tpbh_batch_plan(self._h, mask, waypoints.shape[0], _ptr(num_waypoints)[0],
                                       _ptr(waypoints)[0], _ptr(vmax)[0], _ptr(amax)[0], start_ns)
        if rc != 0:
            raise TpbError(rc, (hostapi().tpbh_batch_error(self._h) or b"").decode())
        return hostapi().tpbh_batch_rows(self._h)

    def array(self, name):
        which, dt = self._ARRAYS[name]
        n = C.c_int64(0)
        p = hostapi().tpbh_batch_array(self._h, which, C.byref(n))
        if not p or n.value == 0:
            return np.empty(0, dt)
        buf = (C.c_char * (n.value * np.dtype(dt).itemsize)).from_address(p)
        return np.frombuffer(buf, dtype=dt)


def resample_uniform_noalloc(device, num_paths, num_samples, num_dofs, inputs, outputs, start_sec,
                             time_step_sec, row_capacity, stream=None):
    """tpb_resample_uniform_noalloc: `inputs` / `outputs` map the field names of
    tpb_resample_desc to raw device addresses (ints; status may be None).  Asynchronous on
    `stream`.  Returns (status code, error message); allocates nothing on the device."""
    d = _ResampleDesc(device=device, num_paths=num_paths, num_samples=num_samples,
                      num_dofs=num_dofs, start_sec=start_sec, time_step_sec=time_step_sec,
                      row_capacity=row_capacity, **inputs, **outputs)
    msg = C.create_string_buffer(256)
    rc = lib().tpb_resample_uniform_noalloc(C.byref(d), stream, msg, len(msg))
    return rc, msg.value.decode()


def measure_fp64_fma_peak(device=0):
    """FP64 FMA rate of the device in TFLOP/s (tpb_measure_fp64_fma_peak)."""
    out = C.c_double(0.0)
    rc = lib().tpb_measure_fp64_fma_peak(int(device), C.byref(out))
    if rc != 0:
        raise TpbError(rc, "tpb_measure_fp64_fma_peak")
    return out.value


# ------------------------------------------------------------------ trajectory buffers
SKIP_INDEX = -2 ** 31


def time_from_sec_ns(seconds):
    """time.h:23-25: absl::FromUnixNanos(seconds * 1e9) (truncation toward zero)."""
    return int(np.float64(seconds) * np.float64(1000000000))


def time_to_sec(ns):
    """time.h:27-30"""
    return float(np.float64(int(ns)) / np.float64(1000000000))


class TrajectoryBufferBatch:
    """A batch of device-resident TrajectoryBuffers (tpb_tbuf_*, include/tpb.h).

    Every method is one batched call over all buffers; per-buffer arguments are numpy
    arrays (host) or torch CUDA tensors (device, asynchronous on the stream).  Statuses
    are absl::StatusCode values per buffer.
    """

    def __init__(self, num_buffers, num_dofs, capacity_rows, timestep_tolerance=1e-6, device=0):
        self.B, self.D, self.cap = int(num_buffers), int(num_dofs), int(capacity_rows)
        self._h = C.c_void_p()
        cfg = _TbufConfig(device, self.B, self.D, self.cap, timestep_tolerance)
        rc = lib().tpb_tbuf_create(C.byref(cfg), C.byref(self._h))
        if rc != 0:
            self._h = C.c_void_p()
            raise TpbError(rc, "tpb_tbuf_create")

    def close(self):
        if getattr(self, "_h", None):
            lib().tpb_tbuf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise TpbError(rc, lib().tpb_tbuf_last_error(self._h).decode())

    @staticmethod
    def _dev(*xs):
        """1 when the arguments are device tensors, 0 for numpy; they must agree."""
        kinds = {bool(hasattr(x, "data_ptr") and x.is_cuda) for x in xs if x is not None}
        assert len(kinds) <= 1, "mix of host and device arguments"
        return 1 if kinds == {True} else 0

    def _status(self, dev, shape=None):
        if dev:
            return None
        return np.zeros(self.B if shape is None else shape, np.int32)

    def set_stream(self, cuda_stream_ptr):
        self._ck(lib().tpb_tbuf_set_stream(self._h, cuda_stream_ptr))

    def sync(self):
        self._ck(lib().tpb_tbuf_sync(self._h))

    @property
    def kernel_launches(self):
        return lib().tpb_tbuf_kernel_launches(self._h)

    def reserve(self, capacity_rows):
        self._ck(lib().tpb_tbuf_reserve(self._h, int(capacity_rows)))
        self.cap = max(self.cap, int(capacity_rows))

    def clear(self, select=None):
        if select is not None and not hasattr(select, "data_ptr"):
            select = np.ascontiguousarray(select, np.int32)
        self._ck(lib().tpb_tbuf_clear(self._h, _addr(select), self._dev(select)))

    def insert_segments(self, begin, length, times, positions, velocities, accelerations,
                        status=None):
        """Packed segments (see tpb_tbuf_insert_segments).  Returns status [B] (host calls)."""
        dev = self._dev(begin, length, times, positions, velocities, accelerations)
        if not dev:
            begin = np.ascontiguousarray(begin, np.int64)
            length = np.ascontiguousarray(length, np.int32)
            times, positions, velocities, accelerations = (
                _f64(x) for x in (times, positions, velocities, accelerations))
            status = self._status(dev)
        total = int(times.shape[0])
        self._ck(lib().tpb_tbuf_insert_segments(
            self._h, _addr(begin), _addr(length), total, _addr(times), _addr(positions),
            _addr(velocities), _addr(accelerations), _addr(status), dev))
        return status

    def append_samples(self, time, positions, velocities, accelerations, select=None):
        dev = self._dev(time, positions, velocities, accelerations, select)
        status = self._status(dev)
        if not dev:
            time, positions, velocities, accelerations = (
                _f64(x) for x in (time, positions, velocities, accelerations))
            if select is not None:
                select = np.ascontiguousarray(select, np.int32)
        self._ck(lib().tpb_tbuf_append_samples(
            self._h, _addr(select), _addr(time), _addr(positions), _addr(velocities),
            _addr(accelerations), _addr(status), dev))
        return status

    def discard_before(self, time_sec):
        dev = self._dev(time_sec)
        time_sec = _f64(time_sec)
        self._ck(lib().tpb_tbuf_discard_before(self._h, _addr(time_sec), dev))

    def positions_up_to_time(self, time_sec):
        time_sec = _f64(time_sec)
        num = np.zeros(self.B, np.int32)
        self._ck(lib().tpb_tbuf_positions_up_to_time(self._h, _addr(time_sec), _addr(num), 0))
        return num

    def sample(self, time_sec, positions=True, velocities=True, accelerations=True, out=None):
        """time_sec [B][Q].  Host: returns (status [B][Q], pos, vel, acc [B][Q][D] or None).
        Device: `out` = (pos, vel, acc, status) tensors (any may be None), asynchronous."""
        dev = self._dev(time_sec)
        Q = int(time_sec.shape[1])
        if dev:
            p, v, a, st = out
        else:
            time_sec = _f64(time_sec)
            mk = lambda want: np.zeros((self.B, Q, self.D)) if want else None
            p, v, a = mk(positions), mk(velocities), mk(accelerations)
            st = np.zeros((self.B, Q), np.int32)
        self._ck(lib().tpb_tbuf_sample(self._h, Q, _addr(time_sec), _addr(p), _addr(v), _addr(a),
                                       _addr(st), dev))
        return st, p, v, a

    def stop_at_index(self, index, max_acceleration, time_step):
        dev = self._dev(index, max_acceleration)
        status = self._status(dev)
        if not dev:
            index = np.ascontiguousarray(index, np.int32)
            max_acceleration = _f64(max_acceleration)
        self._ck(lib().tpb_tbuf_stop_at_index(self._h, _addr(index), _addr(max_acceleration),
                                              float(time_step), _addr(status), dev))
        return status

    def stop_before_time(self, time_sec, max_acceleration, time_step):
        dev = self._dev(time_sec, max_acceleration)
        status = self._status(dev)
        if not dev:
            time_sec, max_acceleration = _f64(time_sec), _f64(max_acceleration)
        self._ck(lib().tpb_tbuf_stop_before_time(self._h, _addr(time_sec),
                                                 _addr(max_acceleration), float(time_step),
                                                 _addr(status), dev))
        return status

    def add_offset(self, offset):
        dev = self._dev(offset)
        offset = _f64(offset)
        self._ck(lib().tpb_tbuf_add_offset(self._h, _addr(offset), dev))

    def state(self):
        """dict(num_samples, sequence_number, front_time, back_time), numpy [B]."""
        n, s = np.zeros(self.B, np.int32), np.zeros(self.B, np.int32)
        f, b = np.zeros(self.B), np.zeros(self.B)
        self._ck(lib().tpb_tbuf_get_state(self._h, _addr(n), _addr(s), _addr(f), _addr(b), 0))
        return dict(num_samples=n, sequence_number=s, front_time=f, back_time=b)

    def rows(self, buffer, num_rows=None, first_row=0):
        if num_rows is None:
            num_rows = int(self.state()["num_samples"][buffer]) - first_row
        t = np.zeros(num_rows)
        p, v, a = (np.zeros((num_rows, self.D)) for _ in range(3))
        self._ck(lib().tpb_tbuf_get_rows(self._h, buffer, first_row, num_rows, _addr(t), _addr(p),
                                         _addr(v), _addr(a), 0))
        return t, p, v, a

    def devptr(self, what):
        ptr, nbytes = C.c_void_p(), C.c_size_t()
        self._ck(lib().tpb_tbuf_devptr(self._h, what, C.byref(ptr), C.byref(nbytes)))
        return ptr.value, nbytes.value

    def view(self, buffer):
        return TrajectoryBufferView(self, buffer)


class TrajectoryBufferView:
    """One buffer of a TrajectoryBufferBatch with the single-buffer interface of the
    reference's TrajectoryBuffer (every call is a batched call that skips the others)."""

    def __init__(self, batch, buffer):
        self.batch, self.b = batch, int(buffer)

    def _only(self, value, fill, dtype=np.float64, shape=()):
        x = np.full((self.batch.B,) + shape, fill, dtype)
        x[self.b] = value
        return x

    def _sel(self):
        return self._only(1, 0, np.int32)

    def clear(self):
        self.batch.clear(self._sel())

    def insert_segment(self, times, pos, vel, acc):
        n = len(times)
        D = self.batch.D
        length = self._only(n, -1, np.int32)
        begin = np.zeros(self.batch.B, np.int64)
        reshape = lambda x: np.ascontiguousarray(x, np.float64).reshape(n, D)
        return int(self.batch.insert_segments(begin, length, np.ascontiguousarray(times, np.float64),
                                              reshape(pos), reshape(vel), reshape(acc))[self.b])

    def append_sample(self, time, pos, vel, acc):
        D = self.batch.D
        row = lambda x: self._only(np.asarray(x, np.float64), 0.0, shape=(D,))
        return int(self.batch.append_samples(self._only(time, 0.0), row(pos), row(vel), row(acc),
                                             self._sel())[self.b])

    def discard_before_sec(self, t):
        self.batch.discard_before(self._only(t, np.nan))

    def discard_before_ns(self, ns):
        self.discard_before_sec(time_to_sec(ns))

    def positions_up_to_time_ns(self, ns):
        return int(self.batch.positions_up_to_time(self._only(time_to_sec(ns), np.nan))[self.b])

    def get_at_time(self, which, t):
        want = [which == 0, which == 1, which == 2]
        st, p, v, a = self.batch.sample(self._only(t, np.nan).reshape(-1, 1), *want)
        return int(st[self.b, 0]), (p, v, a)[which][self.b, 0].copy()

    def stop_at_index(self, index, max_acc, time_step):
        D = self.batch.D
        amax = self._only(np.asarray(max_acc, np.float64), 1.0, shape=(D,))
        return int(self.batch.stop_at_index(self._only(index, SKIP_INDEX, np.int32), amax,
                                            time_step)[self.b])

    def stop_before_time(self, t, max_acc, time_step):
        D = self.batch.D
        amax = self._only(np.asarray(max_acc, np.float64), 1.0, shape=(D,))
        return int(self.batch.stop_before_time(self._only(t, np.nan), amax, time_step)[self.b])

    def add_offset(self, offset):
        self.batch.add_offset(self._only(offset, 0.0))

    @property
    def num_samples(self):
        return int(self.batch.state()["num_samples"][self.b])

    @property
    def sequence_number(self):
        return int(self.batch.state()["sequence_number"][self.b])

    @property
    def start_ns(self):
        s = self.batch.state()
        return time_from_sec_ns(0) if s["num_samples"][self.b] == 0 else time_from_sec_ns(
            s["front_time"][self.b])

    @property
    def end_ns(self):
        s = self.batch.state()
        return 0 if s["num_samples"][self.b] == 0 else time_from_sec_ns(s["back_time"][self.b])

    def rows(self):
        return self.batch.rows(self.b)
