"""TEST INFRASTRUCTURE ONLY -- ctypes/numpy access to the CPU oracle.

Loads oracle/_build/liboracle.so (the plain-C restatement, tp_oracle.c) and,
when present, oracle/_ref/libtp_ref.so (the reference's own
time_optimal_path_timing.cc compiled verbatim, see oracle/Makefile).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.  The product never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(HERE, "_build", "liboracle.so")
_REF_SO = os.path.join(HERE, "_ref", "libtp_ref.so")

c_double_p = C.POINTER(C.c_double)
c_ubyte_p = C.POINTER(C.c_ubyte)

ORC_OK = 0
ORC_FAIL_CAPACITY = 6
ORC_FAIL_DEADLINE = 7
ORC_FAIL_START_RANGE = 8
ORC_FAIL_START_ARG = 9
ORC_FAIL_START_END = 10
ORC_FAIL_NO_SAMPLES = 11
ORC_FAIL_INITIAL_VELOCITY = 12


def build(force: bool = False) -> None:
    """Compile the oracle (and oracle/_ref when /root/reference exists)."""
    if force:
        subprocess.run(["make", "-C", HERE, "clean"], check=True, stdout=subprocess.DEVNULL)
    # make is a no-op when everything is up to date; without /root/reference (the GPU
    # box) the Makefile keeps the prebuilt oracle/_ref
    subprocess.run(["make", "-C", HERE, "all"], check=True, stdout=subprocess.DEVNULL)


def _dp(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_double_p)


class _Solution(C.Structure):
    _fields_ = [(n, c_double_p) for n in
                ("time", "s", "sd", "sdd", "sd2", "b_sd2_max", "b_sdd_max",
                 "b_sdd_min", "b_sd2_zero")] + [
        ("b_type", c_ubyte_p), ("last_extremal_index", C.c_int),
        ("extra", C.c_int)]


class _PlanCfg(C.Structure):
    _fields_ = [("N", C.c_int), ("dim", C.c_int), ("W", C.c_int),
                ("rounding", C.c_double), ("safety", C.c_double),
                ("delta", C.c_double), ("start_sec", C.c_double),
                ("dt", C.c_double), ("skip_method", C.c_int)]


class _PlanOut(C.Structure):
    _fields_ = [(n, c_double_p) for n in
                ("time", "s", "sd", "sdd", "q", "dq", "ddq", "vel", "acc")] + [
        ("duration", C.c_double), ("delta_used", C.c_double),
        ("knot_end", C.c_double), ("last_extremal_index", C.c_int),
        ("M_cap", C.c_int), ("M", C.c_int)] + [(n, c_double_p) for n in
        ("r_t", "r_pos", "r_vel", "r_acc", "r_s", "r_sd", "r_sdd")]


class _Spline(C.Structure):
    """orc_spline (tp_oracle.h)"""
    _fields_ = [("degree", C.c_int), ("dim", C.c_int), ("knot_capacity", C.c_int),
                ("num_knots", C.c_int), ("num_points", C.c_int), ("knots", c_double_p),
                ("points", c_double_p)]


class _Planner(C.Structure):
    """orc_planner (tp_oracle.h)"""
    _fields_ = [("N", C.c_int), ("dim", C.c_int), ("W", C.c_int), ("rounding", C.c_double),
                ("safety", C.c_double), ("delta", C.c_double), ("dt", C.c_double),
                ("skip_method", C.c_int), ("max_planning_iterations", C.c_int),
                ("max_initial_velocity_error", C.c_double)] + [
        (n, c_double_p) for n in ("waypoints", "vmax", "amax", "initial_velocity")] + [
        ("spline", _Spline), ("P_cap", C.c_int)] + [
        (n, c_double_p) for n in ("w_time", "w_s", "w_sd", "w_sdd", "w_q", "w_vel", "w_acc")] + [
        ("M_cap", C.c_int)] + [
        (n, c_double_p) for n in ("r_t", "r_pos", "r_vel", "r_acc", "r_s", "r_sd", "r_sdd")] + [
        ("num_points", C.c_int), ("knot_end", C.c_double), ("delta_used", C.c_double),
        ("len", C.c_int), ("M", C.c_int), ("initial_plan", C.c_int), ("sampled", C.c_int),
        ("planned_to_end", C.c_int), ("target_reached", C.c_int), ("start_time", C.c_int64),
        ("end_time", C.c_int64), ("final_decel_start", C.c_int64), ("path_horizon", C.c_double),
        ("path_start", C.c_double), ("decel_time", C.c_double), ("last_extremal_index", C.c_int),
        ("loops", C.c_int), ("erased", C.c_int), ("cart", C.c_int),
        ("translation_rounding", C.c_double), ("max_translational_velocity", C.c_double),
        ("max_rotational_velocity", C.c_double), ("pose_waypoints", c_double_p),
        ("pose_ctrl", c_double_p)]


class _CartCfg(C.Structure):
    _fields_ = [("N", C.c_int), ("W", C.c_int), ("translation_rounding", C.c_double),
                ("rotation_rounding", C.c_double), ("safety", C.c_double),
                ("delta", C.c_double), ("start_sec", C.c_double), ("dt", C.c_double),
                ("max_translational_velocity", C.c_double),
                ("max_rotational_velocity", C.c_double)]


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_ORACLE_SO):
            build()
        L = C.CDLL(_ORACLE_SO)
        L.orc_knot_span.restype = C.c_int
        L.orc_knot_span.argtypes = [c_double_p, C.c_int, C.c_int, C.c_double]
        L.orc_basis.argtypes = [c_double_p, C.c_int, C.c_int, C.c_double, c_double_p]
        L.orc_basis_ders.argtypes = [c_double_p, C.c_int, C.c_int, C.c_int,
                                     C.c_double, c_double_p]
        L.orc_eval_curve.restype = C.c_int
        L.orc_eval_curve.argtypes = [c_double_p, C.c_int, c_double_p, C.c_int,
                                     C.c_int, C.c_double, c_double_p]
        L.orc_eval_curve_ders.restype = C.c_int
        L.orc_eval_curve_ders.argtypes = [c_double_p, C.c_int, c_double_p, C.c_int,
                                          C.c_int, C.c_int, C.c_double, c_double_p]
        L.orc_make_uniform_knots.restype = C.c_int
        L.orc_make_uniform_knots.argtypes = [C.c_int, C.c_int, C.c_double,
                                             C.c_double, c_double_p]
        L.orc_polyline_to_bspline3.restype = C.c_int
        L.orc_polyline_to_bspline3.argtypes = [c_double_p, C.c_int, C.c_int,
                                               C.c_double, c_double_p]
        L.orc_joint_fit.restype = C.c_int
        L.orc_joint_fit.argtypes = [c_double_p, C.c_int, C.c_int, C.c_double,
                                    c_double_p, c_double_p]
        L.orc_joint_sample.argtypes = [c_double_p, C.c_int, c_double_p, C.c_int,
                                       C.c_int, C.c_double, C.c_double,
                                       c_double_p, c_double_p, c_double_p]
        L.orc_joint_constraints.argtypes = [c_double_p, c_double_p, C.c_int,
                                            C.c_int, c_double_p, c_double_p,
                                            C.c_double, c_double_p]
        L.orc_find_max_sd2_simplex.argtypes = [c_double_p, C.c_int, c_double_p,
                                               c_double_p, c_double_p]
        L.orc_solve.restype = C.c_int
        L.orc_solve.argtypes = [C.c_int, C.c_int, c_double_p] + [C.c_double] * 5 + [
            C.c_int, C.POINTER(_Solution)]
        L.orc_count_violations.restype = C.c_int
        L.orc_count_violations.argtypes = [C.c_int, C.c_int, c_double_p,
                                           c_double_p, c_double_p]
        L.orc_query.restype = C.c_int
        L.orc_query.argtypes = [C.c_int] + [c_double_p] * 4 + [C.c_double] * 3 + [
            c_double_p] * 3
        L.orc_time_from_sec.restype = C.c_int64
        L.orc_time_from_sec.argtypes = [C.c_double]
        L.orc_time_to_sec.restype = C.c_double
        L.orc_time_to_sec.argtypes = [C.c_int64]
        L.orc_assemble.argtypes = [c_double_p] * 5 + [C.c_int, C.c_int,
                                                      c_double_p, c_double_p]
        L.orc_resample_count.restype = C.c_int
        L.orc_resample_count.argtypes = [C.c_double] * 3
        L.orc_resample_uniform.argtypes = [C.c_int, C.c_int] + [c_double_p] * 8 + [
            C.c_double, C.c_double, C.c_int] + [c_double_p] * 7
        L.orc_plan_joint.restype = C.c_int
        L.orc_plan_joint.argtypes = [C.POINTER(_PlanCfg), c_double_p, c_double_p,
                                     c_double_p, C.POINTER(_PlanOut)]
        # Cartesian path (tp_oracle_cart.c)
        for fn, n in (("orc_quat_mul", 3), ("orc_quat_log", 2), ("orc_quat_exp", 2),
                      ("orc_exp_so3", 2), ("orc_log_so3", 2), ("orc_log_so3_derivative", 2),
                      ("orc_fake_jacobian", 2)):
            getattr(L, fn).argtypes = [c_double_p] * n
            getattr(L, fn).restype = None
        L.orc_quat_power.argtypes = [c_double_p, C.c_double, c_double_p]
        L.orc_quat_power_derivative.argtypes = [c_double_p, C.c_double, c_double_p]
        L.orc_quat_eval_curve_derivative.argtypes = [c_double_p, C.c_int, c_double_p, C.c_int,
                                                     C.c_double, c_double_p, c_double_p]
        L.orc_quat_eval_curve.restype = C.c_int
        L.orc_quat_eval_curve.argtypes = [c_double_p, C.c_int, c_double_p, C.c_int,
                                          C.c_double, c_double_p]
        L.orc_pose_polyline_to_bspline3.restype = C.c_int
        L.orc_pose_polyline_to_bspline3.argtypes = [c_double_p, C.c_int, C.c_double,
                                                    C.c_double, c_double_p]
        L.orc_cart_fit.restype = C.c_int
        L.orc_cart_fit.argtypes = [c_double_p, c_double_p, C.c_int, C.c_int, C.c_double,
                                   C.c_double, c_double_p, c_double_p, c_double_p]
        L.orc_cart_sample_targets.argtypes = [c_double_p, C.c_int, c_double_p, c_double_p,
                                              C.c_int, C.c_int, C.c_double, c_double_p,
                                              c_double_p]
        L.orc_cart_derivatives.argtypes = [c_double_p, C.c_int, C.c_int, C.c_double,
                                           c_double_p, c_double_p]
        L.orc_cart_constraints.argtypes = [c_double_p, c_double_p, c_double_p, C.c_int,
                                           C.c_int, c_double_p, c_double_p, C.c_double,
                                           C.c_double, C.c_double, c_double_p]
        L.orc_fake_ik.argtypes = [c_double_p, c_double_p, C.c_int, c_double_p]
        L.orc_plan_cart_fake.restype = C.c_int
        L.orc_plan_cart_fake.argtypes = [C.POINTER(_CartCfg), c_double_p, c_double_p,
                                         c_double_p, c_double_p, C.POINTER(_PlanOut),
                                         c_double_p, c_double_p, c_double_p, c_double_p]
        L.orc_resample_skip.restype = C.c_int
        L.orc_resample_skip.argtypes = [C.c_int, C.c_int] + [c_double_p] * 8 + [
            C.c_double, C.c_double] + [c_double_p] * 7
        L.orc_planner_init.restype = C.c_int
        L.orc_planner_init.argtypes = [C.POINTER(_Planner)]
        L.orc_planner_plan.restype = C.c_int
        L.orc_planner_plan.argtypes = [C.POINTER(_Planner), C.c_int64, C.c_int64]
        L.orc_planner_switch.restype = C.c_int
        L.orc_planner_switch.argtypes = [C.POINTER(_Planner), C.c_double, c_double_p, C.c_int]
        # online re-planning geometry (tp_oracle_switch.c)
        L.orc_spline_insert_knot.restype = C.c_int
        L.orc_spline_insert_knot.argtypes = [C.POINTER(_Spline), C.c_double, C.c_int]
        L.orc_spline_truncate.restype = C.c_int
        L.orc_spline_truncate.argtypes = [C.POINTER(_Spline), C.c_double]
        L.orc_spline_extend.restype = C.c_int
        L.orc_spline_extend.argtypes = [C.POINTER(_Spline), c_double_p, C.c_int]
        L.orc_project_point_on_path.restype = C.c_int
        L.orc_project_point_on_path.argtypes = [c_double_p, C.c_int, C.c_int, c_double_p,
                                                C.POINTER(C.c_int), c_double_p, c_double_p,
                                                c_double_p]
        L.orc_joint_switch.restype = C.c_int
        L.orc_joint_switch.argtypes = [C.POINTER(_Spline), C.c_double, c_double_p, C.c_int,
                                       C.c_double, c_double_p, C.POINTER(C.c_int)]
        # stopping trajectories (tp_oracle_stop.c)
        L.orc_sampled_trajectory_inputs_valid.restype = C.c_int
        L.orc_sampled_trajectory_inputs_valid.argtypes = [C.c_int, c_double_p]
        L.orc_rescale_to_stop.restype = C.c_int
        L.orc_rescale_to_stop.argtypes = [C.c_int, C.c_int] + [c_double_p] * 4 + [
            C.POINTER(C.c_int), C.POINTER(C.c_int)] + [c_double_p] * 3
        L.orc_fastest_stop.restype = C.c_int
        L.orc_fastest_stop.argtypes = [C.c_int, C.c_int] + [c_double_p] * 8 + [C.POINTER(C.c_int)]
        L.orc_path_stop_parameter.restype = C.c_int
        L.orc_path_stop_parameter.argtypes = [C.c_int, C.c_int] + [c_double_p] * 5 + [
            C.c_double, c_double_p, C.POINTER(C.c_int), c_double_p]
        # trajectory buffer (tp_oracle_tbuf.c)
        _register_tbuf(L, "orc_tbuf_", port=True)
        _lib = L
    return _lib


def _register_tbuf(L, pre, port):
    """ctypes signatures of the TrajectoryBuffer functions; the port (orc_tbuf_*)
    keeps the dof count in the buffer, the reference wrapper (tpref_tbuf_*) takes
    it per call."""
    vp, dbl, i32, i64 = C.c_void_p, C.c_double, C.c_int, C.c_int64
    D = [] if port else [i32]
    f = lambda name: getattr(L, pre + name)
    f("create").restype = i32
    f("create").argtypes = ([dbl, i32] if port else [dbl]) + [C.POINTER(vp)]
    f("destroy").restype = None
    f("destroy").argtypes = [vp]
    f("clear").restype = None
    f("clear").argtypes = [vp]
    f("insert_segment").restype = i32
    f("insert_segment").argtypes = [vp, i32] + D + [c_double_p] * 4
    f("append_sample").restype = i32
    f("append_sample").argtypes = [vp] + D + [dbl] + [c_double_p] * 3
    f("discard_before_sec").restype = None
    f("discard_before_sec").argtypes = [vp, dbl]
    f("discard_before_ns").restype = None
    f("discard_before_ns").argtypes = [vp, i64]
    f("positions_up_to_time_ns").restype = i32
    f("positions_up_to_time_ns").argtypes = [vp, i64]
    f("get_at_time").restype = i32
    f("get_at_time").argtypes = [vp, i32, dbl, c_double_p]
    f("stop_at_index").restype = i32
    f("stop_at_index").argtypes = [vp, i32] + D + [c_double_p, dbl]
    f("stop_before_time").restype = i32
    f("stop_before_time").argtypes = [vp, dbl] + D + [c_double_p, dbl]
    f("add_offset").restype = None
    f("add_offset").argtypes = [vp, dbl]
    f("num_samples").restype = i32
    f("num_samples").argtypes = [vp]
    f("sequence_number").restype = i32
    f("sequence_number").argtypes = [vp]
    f("start_ns").restype = i64
    f("start_ns").argtypes = [vp]
    f("end_ns").restype = i64
    f("end_ns").argtypes = [vp]
    f("rows").restype = None
    f("rows").argtypes = [vp] + D + [c_double_p] * 4


def _register_ref_planner(R):
    """The reference's own TimeableJointSplinePath + PathTimingTrajectory
    (oracle/ref_wrap_planner.cc)."""
    vp, i64p, ip = C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int)
    i32p = C.POINTER(C.c_int32)
    R.tpref_planner_create.restype = vp
    R.tpref_planner_create.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                       C.c_int64, C.c_int, C.c_int, C.c_double]
    R.tpref_planner_destroy.restype = None
    R.tpref_planner_destroy.argtypes = [vp]
    for name, args in (("set_limits", [c_double_p, c_double_p]),
                       ("set_waypoints", [C.c_int, c_double_p]),
                       ("set_initial_velocity", [c_double_p]),
                       ("plan", [C.c_int64, C.c_int64]),
                       ("switch", [C.c_double, C.c_int, c_double_p]),
                       ("stop_parameter", [C.c_int64, c_double_p])):
        f = getattr(R, "tpref_planner_" + name)
        f.restype = C.c_int
        f.argtypes = [vp] + args
    for name, args in (("state", [i64p]), ("scalars", [c_double_p]),
                       ("rows", [c_double_p] * 7), ("path_samples", [c_double_p] * 7),
                       ("window", [c_double_p] * 7), ("constraints", [c_double_p]),
                       ("spline", [ip, c_double_p, C.c_int, c_double_p, C.c_int])):
        f = getattr(R, "tpref_planner_" + name)
        f.restype = None
        f.argtypes = [vp] + args
    R.tpref_spline_eval_batch.restype = C.c_int
    R.tpref_spline_eval_batch.argtypes = [C.c_int] * 4 + [c_double_p, c_double_p, C.c_int,
                                                          c_double_p, C.c_int, C.c_int,
                                                          c_double_p, i32p]
    # Cartesian path (splines/bsplineq.cc, timeable_path_cartesian_spline.cc verbatim)
    for name in ("log", "exp"):
        f = getattr(R, "tpref_quat_" + name)
        f.restype, f.argtypes = None, [c_double_p, c_double_p]
    for name in ("power", "power_derivative"):
        f = getattr(R, "tpref_quat_" + name)
        f.restype, f.argtypes = None, [c_double_p, C.c_double, c_double_p]
    R.tpref_bsplineq_eval.restype = C.c_int
    R.tpref_bsplineq_eval.argtypes = [C.c_int, C.c_int, c_double_p, C.c_int, c_double_p, C.c_int,
                                      c_double_p, c_double_p, c_double_p, i32p]
    R.tpref_pose_polyline_to_bspline3.restype = C.c_int
    R.tpref_pose_polyline_to_bspline3.argtypes = [c_double_p, C.c_int, C.c_double, C.c_double,
                                                  c_double_p, C.c_int]
    R.tpref_planner_set_cartesian.restype = None
    R.tpref_planner_set_cartesian.argtypes = [vp, C.c_double, C.c_double, C.c_double, vp, vp]
    R.tpref_planner_set_cart_waypoints.restype = C.c_int
    R.tpref_planner_set_cart_waypoints.argtypes = [vp, C.c_int, c_double_p, c_double_p]
    R.tpref_planner_cart_switch.restype = C.c_int
    R.tpref_planner_cart_switch.argtypes = [vp, C.c_double, C.c_int, c_double_p, c_double_p]
    for name in ("pose_points", "ik_positions"):
        f = getattr(R, "tpref_planner_cart_" + name)
        f.restype, f.argtypes = C.c_int, [vp, c_double_p, C.c_int]
    R.tpref_planner_cart_targets.restype = C.c_int
    R.tpref_planner_cart_targets.argtypes = [vp, c_double_p, c_double_p, C.c_int]
    R.tpref_planner_cart_sample.restype = C.c_int
    R.tpref_planner_cart_sample.argtypes = [vp, C.c_double]
    R.tpref_plan_cart_batch.restype = None
    R.tpref_plan_cart_batch.argtypes = [
        C.c_int, C.c_int, C.c_int, C.c_int, c_double_p, c_double_p, c_double_p, c_double_p,
        C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int64, C.c_int, C.c_int,
        vp, vp, i32p, c_double_p, i32p, c_double_p]
    R.tpref_plan_joint_batch.restype = None
    R.tpref_plan_joint_batch.argtypes = [
        C.c_int, C.c_int, C.c_int, C.c_int, c_double_p, c_double_p, c_double_p, C.c_double,
        C.c_double, C.c_int64, C.c_int, C.c_int, i32p, c_double_p, i32p, i64p, c_double_p,
        c_double_p]


def have_ref() -> bool:
    return os.path.exists(_REF_SO)


def ref():
    global _ref
    if _ref is None:
        R = C.CDLL(_REF_SO)
        R.tpref_solve.restype = C.c_int
        R.tpref_solve.argtypes = [C.c_int, C.c_int, c_double_p] + [C.c_double] * 5 + [
            C.c_int, C.POINTER(_Solution)]
        R.tpref_find_max_sd2.argtypes = [C.c_int, c_double_p, C.c_int, c_double_p]
        R.tpref_query.restype = C.c_int
        R.tpref_query.argtypes = [C.c_int, C.c_int, c_double_p] + [C.c_double] * 5 + [
            C.c_int, c_double_p, c_double_p, c_double_p, c_double_p]
        R.tpref_solve_batch.argtypes = [C.c_int, C.c_int, C.c_int, c_double_p,
                                        c_double_p, C.c_int, C.c_int, c_double_p,
                                        c_double_p, c_double_p, c_double_p]
        R.tpref_plan_joint.restype = C.c_int
        R.tpref_plan_joint.argtypes = [C.POINTER(_PlanCfg), c_double_p, c_double_p,
                                       c_double_p, C.POINTER(_PlanOut)]
        R.tpref_rescale_to_stop.restype = C.c_int
        R.tpref_rescale_to_stop.argtypes = [C.c_int, C.c_int] + [c_double_p] * 5 + [
            C.POINTER(C.c_int), C.POINTER(C.c_int)] + [c_double_p] * 4
        _register_tbuf(R, "tpref_tbuf_", port=False)
        R.tpref_time_from_sec_ns.restype = C.c_int64
        R.tpref_time_from_sec_ns.argtypes = [C.c_double]
        R.tpref_time_to_sec.restype = C.c_double
        R.tpref_time_to_sec.argtypes = [C.c_int64]
        _register_ref_planner(R)
        _ref = R
    return _ref


# ----------------------------------------------------------------- stage A
def knot_span(knots, degree, u):
    knots = np.ascontiguousarray(knots, np.float64)
    return lib().orc_knot_span(_dp(knots), len(knots), degree, float(u))


def basis(knots, span, p, u):
    knots = np.ascontiguousarray(knots, np.float64)
    out = np.zeros(p + 1)
    lib().orc_basis(_dp(knots), span, p, float(u), _dp(out))
    return out


def basis_ders(knots, span, p, der, u):
    knots = np.ascontiguousarray(knots, np.float64)
    out = np.zeros((der + 1, p + 1))
    lib().orc_basis_ders(_dp(knots), span, p, der, float(u), _dp(out))
    return out


def eval_curve(knots, ctrl, degree, u):
    knots = np.ascontiguousarray(knots, np.float64)
    ctrl = np.ascontiguousarray(ctrl, np.float64)
    out = np.zeros(ctrl.shape[1])
    st = lib().orc_eval_curve(_dp(knots), len(knots), _dp(ctrl), ctrl.shape[1],
                              degree, float(u), _dp(out))
    return st, out


def eval_curve_ders(knots, ctrl, degree, nder, us):
    """Evaluate at many parameters; returns (status[n], values[n][nder+1][dim])."""
    knots = np.ascontiguousarray(knots, np.float64)
    ctrl = np.ascontiguousarray(ctrl, np.float64)
    us = np.atleast_1d(np.asarray(us, np.float64))
    dim = ctrl.shape[1]
    out = np.zeros((len(us), nder + 1, dim))
    st = np.zeros(len(us), np.int32)
    L = lib()
    for i, u in enumerate(us):
        st[i] = L.orc_eval_curve_ders(_dp(knots), len(knots), _dp(ctrl), dim,
                                      degree, nder, float(u), _dp(out[i]))
    return st, out


def make_uniform_knots(num_points, degree, lo=0.0, hi=1.0):
    k = np.zeros(num_points + degree + 1)
    st = lib().orc_make_uniform_knots(num_points, degree, lo, hi, _dp(k))
    assert st == 0
    return k


# ----------------------------------------------------------------- stage B
def polyline_to_bspline3(corners, radius):
    corners = np.ascontiguousarray(corners, np.float64)
    W, D = corners.shape
    out = np.zeros((max(3 * W - 2, 4), D))
    n = lib().orc_polyline_to_bspline3(_dp(corners), W, D, float(radius), _dp(out))
    return out[:n].copy()


def joint_fit(waypoints, rounding):
    wp = np.ascontiguousarray(waypoints, np.float64)
    W, D = wp.shape
    ctrl = np.zeros((max(3 * W - 2, 4), D))
    knots = np.zeros(max(3 * W - 2, 4) + 3)
    n = lib().orc_joint_fit(_dp(wp), W, D, float(rounding), _dp(ctrl), _dp(knots))
    return ctrl[:n].copy(), knots[:n + 3].copy()


def joint_sample(knots, ctrl, N, path_start, delta):
    knots = np.ascontiguousarray(knots, np.float64)
    ctrl = np.ascontiguousarray(ctrl, np.float64)
    D = ctrl.shape[1]
    q, dq, ddq = np.zeros((N, D)), np.zeros((N, D)), np.zeros((N, D))
    lib().orc_joint_sample(_dp(knots), len(knots), _dp(ctrl), D, N,
                           float(path_start), float(delta), _dp(q), _dp(dq), _dp(ddq))
    return q, dq, ddq


def joint_constraints(dq, ddq, vmax, amax, safety):
    dq = np.ascontiguousarray(dq, np.float64)
    ddq = np.ascontiguousarray(ddq, np.float64)
    vmax = np.ascontiguousarray(vmax, np.float64)
    amax = np.ascontiguousarray(amax, np.float64)
    N, D = dq.shape
    cons = np.zeros((N, 2 * D, 4))
    lib().orc_joint_constraints(_dp(dq), _dp(ddq), N, D, _dp(vmax), _dp(amax),
                                float(safety), _dp(cons))
    return cons


# ----------------------------------------------------------------- stage C
def find_max_sd2(cons_row, impl="port"):
    cons_row = np.ascontiguousarray(cons_row, np.float64)
    Cn = cons_row.shape[0]
    if impl == "port":
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        lib().orc_find_max_sd2_simplex(_dp(cons_row), Cn, C.byref(a), C.byref(b),
                                       C.byref(c))
        return np.array([a.value, b.value, c.value])
    out = np.zeros(3)
    ref().tpref_find_max_sd2(Cn, _dp(cons_row), 1 if impl == "ref_brute" else 0,
                             _dp(out))
    return out


def solve(cons, s_start, s_end, sd_start=0.0, sdd_start=0.0, time_start=0.0,
          max_loops=100, impl="port"):
    """Returns dict(status_ok, status, time, s, sd, sdd, sd2, b_*, last_extremal_index)."""
    cons = np.ascontiguousarray(cons, np.float64)
    N, Cn, four = cons.shape
    assert four == 4
    arrs = {k: np.full(N, np.nan) for k in
            ("time", "s", "sd", "sdd", "sd2", "b_sd2_max", "b_sdd_max",
             "b_sdd_min", "b_sd2_zero")}
    btype = np.zeros(N, np.uint8)
    sol = _Solution(**{k: _dp(v) for k, v in arrs.items()},
                    b_type=btype.ctypes.data_as(c_ubyte_p))
    if impl == "port":
        st = lib().orc_solve(N, Cn, _dp(cons), s_start, s_end, sd_start, sdd_start,
                             time_start, max_loops, C.byref(sol))
        ok = st == ORC_OK
    else:
        r = ref().tpref_solve(N, Cn, _dp(cons), s_start, s_end, sd_start, sdd_start,
                              time_start, max_loops, C.byref(sol))
        st = 0 if r == 1 else (1 if r == -1 else 2)
        ok = r == 1
    out = dict(arrs)
    out.update(ok=ok, status=st, b_type=btype,
               last_extremal_index=sol.last_extremal_index, extra=sol.extra)
    return out


def count_violations(cons, sd2, sdd):
    cons = np.ascontiguousarray(cons, np.float64)
    N, Cn, _ = cons.shape
    return lib().orc_count_violations(N, Cn, _dp(cons), _dp(np.ascontiguousarray(sd2)),
                                      _dp(np.ascontiguousarray(sdd)))


def query(sol, s_start, s_end, t):
    N = len(sol["time"])
    a, b, c = C.c_double(), C.c_double(), C.c_double()
    ok = lib().orc_query(N, _dp(sol["time"]), _dp(sol["s"]), _dp(sol["sd"]),
                         _dp(sol["sd2"]), s_start, s_end, float(t), C.byref(a),
                         C.byref(b), C.byref(c))
    return ok, a.value, b.value, c.value


# ----------------------------------------------------------------- stage D
def plan_joint(waypoints, vmax, amax, N, rounding=0.2, safety=0.8, delta=-1.0,
               start_sec=0.0, dt=0.0, M_cap=0, impl="port", skip=False):
    """One fresh joint path, single window.  impl: 'port' | 'ref'.  skip: resample with
    kSkipSamplesCloserThanTimeStep (port only; M_cap >= N + 1)."""
    wp = np.ascontiguousarray(waypoints, np.float64)
    vmax = np.ascontiguousarray(vmax, np.float64)
    amax = np.ascontiguousarray(amax, np.float64)
    W, D = wp.shape
    cfg = _PlanCfg(N, D, W, rounding, safety, delta, start_sec, dt, 1 if skip else 0)
    assert not (skip and impl != "port")
    o = {k: np.full(N, np.nan) for k in ("time", "s", "sd", "sdd")}
    o.update({k: np.zeros((N, D)) for k in ("q", "dq", "ddq", "vel", "acc")})
    r = {"r_t": np.zeros(M_cap), "r_s": np.zeros(M_cap), "r_sd": np.zeros(M_cap),
         "r_sdd": np.zeros(M_cap), "r_pos": np.zeros((M_cap, D)),
         "r_vel": np.zeros((M_cap, D)), "r_acc": np.zeros((M_cap, D))}
    out = _PlanOut(**{k: _dp(v) for k, v in o.items()}, M_cap=M_cap,
                   **{k: (_dp(v) if M_cap > 0 else None) for k, v in r.items()})
    fn = lib().orc_plan_joint if impl == "port" else ref().tpref_plan_joint
    st = fn(C.byref(cfg), _dp(wp), _dp(vmax), _dp(amax), C.byref(out))
    res = dict(o)
    M = out.M
    res.update({k: v[:M] for k, v in r.items()})
    res.update(status=st, ok=(st == ORC_OK), duration=out.duration,
               delta=out.delta_used, knot_end=out.knot_end, M=M,
               last_extremal_index=out.last_extremal_index)
    return res


class Planner:
    """PathTimingTrajectory with its state between Plan() calls, for one joint path
    (orc_planner, tp_oracle_windows.c).  Times are absl::Time / Duration as integer
    nanoseconds."""

    def __init__(self, waypoints, vmax, amax, N, delta, dt, rounding=0.2, safety=0.8, skip=False,
                 max_planning_iterations=200, initial_velocity=None,
                 max_initial_velocity_error=1e-4, P_cap=None, M_cap=200000, pose_waypoints=None,
                 translation_rounding=0.05, max_translational_velocity=0.5,
                 max_rotational_velocity=0.25):
        """pose_waypoints [W][7] (tx,ty,tz,qx,qy,qz,qw): a Cartesian path with the reference
        test's fake kinematics (`waypoints` are then the joint waypoints, 7 dofs)."""
        wp = np.ascontiguousarray(waypoints, np.float64)
        W, D = wp.shape
        P = P_cap or 64 * N
        n = max(4, 3 * W - 2)
        cap = max(2 * (n + 3), 100)  # timeable_path_joint_spline.cc:262-268
        self._keep = dict(
            waypoints=wp, vmax=np.ascontiguousarray(vmax, np.float64),
            amax=np.ascontiguousarray(amax, np.float64),
            initial_velocity=np.ascontiguousarray(
                np.zeros(D) if initial_velocity is None else initial_velocity, np.float64).copy())
        self._knots, self._points = np.zeros(cap), np.zeros((cap - 3, D))
        spline = _Spline(2, D, cap, 0, 0, _dp(self._knots), _dp(self._points))
        cart = {}
        if pose_waypoints is not None:
            self._pose_wp = np.ascontiguousarray(pose_waypoints, np.float64)
            self._pose_ctrl = np.zeros((cap - 3, 7))
            cart = dict(cart=1, translation_rounding=translation_rounding,
                        max_translational_velocity=max_translational_velocity,
                        max_rotational_velocity=max_rotational_velocity,
                        pose_waypoints=_dp(self._pose_wp), pose_ctrl=_dp(self._pose_ctrl))
        self.w = {k: np.full(P, np.nan) for k in ("w_time", "w_s", "w_sd", "w_sdd")}
        self.w.update({k: np.zeros((P, D)) for k in ("w_q", "w_vel", "w_acc")})
        self.r = {k: np.zeros(M_cap) for k in ("r_t", "r_s", "r_sd", "r_sdd")}
        self.r.update({k: np.zeros((M_cap, D)) for k in ("r_pos", "r_vel", "r_acc")})
        self.c = _Planner(N=N, dim=D, W=W, rounding=rounding, safety=safety, delta=delta, dt=dt,
                          skip_method=1 if skip else 0,
                          max_planning_iterations=max_planning_iterations,
                          max_initial_velocity_error=max_initial_velocity_error, P_cap=P,
                          M_cap=M_cap, spline=spline, **cart,
                          **{k: _dp(v) for k, v in self._keep.items()},
                          **{k: _dp(v) for k, v in self.w.items()},
                          **{k: _dp(v) for k, v in self.r.items()})
        lib().orc_planner_init(C.byref(self.c))

    def plan(self, start_ns, horizon_ns):
        """Plan(start, horizon); returns the ORC_* status."""
        return lib().orc_planner_plan(C.byref(self.c), int(start_ns), int(horizon_ns))

    def switch(self, keep_path_until, waypoints, initial_velocity=None, pose_waypoints=None):
        """SwitchToWaypointPath (+ SetInitialVelocity); returns the absl status code."""
        wp = np.ascontiguousarray(waypoints, np.float64)
        if initial_velocity is not None:
            self._keep["initial_velocity"][:] = initial_velocity
        if pose_waypoints is not None:
            pw = np.ascontiguousarray(pose_waypoints, np.float64)
            L = lib()
            L.orc_planner_cart_switch.restype = C.c_int
            L.orc_planner_cart_switch.argtypes = [C.POINTER(_Planner), C.c_double, c_double_p,
                                                  c_double_p, C.c_int]
            return L.orc_planner_cart_switch(C.byref(self.c), float(keep_path_until), _dp(pw),
                                             _dp(wp), wp.shape[0])
        return lib().orc_planner_switch(C.byref(self.c), float(keep_path_until), _dp(wp),
                                        wp.shape[0])

    def spline(self):
        """(knots, control points) of the path's spline."""
        sp = self.c.spline
        return self._knots[:sp.num_knots].copy(), self._points[:sp.num_points].copy()

    def trajectory(self):
        M = self.c.M
        return {k: v[:M].copy() for k, v in self.r.items()}

    def path_samples(self):
        n = self.c.len
        return {k: v[:n].copy() for k, v in self.w.items()}

    def __getattr__(self, name):  # state fields: len, M, end_time, planned_to_end, ...
        return getattr(self.__dict__["c"], name)


#: absl::StatusCode an ORC_* planner status corresponds to (ref_wrap_planner.cc returns
#: the reference's own codes)
ABSL_OF_ORC = {0: 0, 1: 13, 2: 13, 3: 13, 4: 13, 5: 13, ORC_FAIL_DEADLINE: 4,
               ORC_FAIL_START_RANGE: 11, ORC_FAIL_START_ARG: 3, ORC_FAIL_START_END: 13,
               ORC_FAIL_NO_SAMPLES: 9, ORC_FAIL_INITIAL_VELOCITY: 3}
#: TimeablePath::State of the reference -> the oracle's `sampled` field
#: (0 kNewPath, 1 kPathWasSampled, 2 kModifiedPath)
ORC_STATE_OF_REF = {1: 0, 3: 1, 2: 2}


class RefPlanner:
    """The REFERENCE's own TimeableJointSplinePath + PathTimingTrajectory object code
    (oracle/_ref, ref_wrap_planner.cc) behind the interface of `Planner`.  plan() and
    switch() return absl::StatusCode numbers."""

    _STATE = ("start_time", "end_time", "final_decel_start", "planned_to_end", "target_reached",
              "path_state", "initial_plan", "M", "len", "last_extremal_index", "at_end")

    def __init__(self, waypoints, vmax, amax, N, delta, dt, rounding=0.2, safety=0.8, skip=False,
                 max_planning_iterations=200, initial_velocity=None,
                 max_initial_velocity_error=1e-4, pose_waypoints=None,
                 translation_rounding=0.05, max_trans_vel=0.5, max_rot_vel=0.25,
                 ik_fn=None, jacobian_fn=None):
        """pose_waypoints [W][7] (tx,ty,tz,qx,qy,qz,qw): a TimeableCartesianSplinePath
        (`waypoints` are then the joint waypoints, `rounding` the rotation rounding) whose
        PathIKFunc / JacobianFunc are the C functions ik_fn / jacobian_fn (default: the
        reference test's fake kinematics as restated in oracle/tp_oracle_cart.c -- user
        callbacks, not part of the path)."""
        wp = np.ascontiguousarray(waypoints, np.float64)
        self.W, self.D = wp.shape
        self.N = N
        self.cart = pose_waypoints is not None
        dt_ns = int(round(dt * 1e9))
        assert dt_ns * 1e-9 == dt or abs(dt_ns / 1e9 - dt) == 0.0, "dt must be whole ns"
        self._h = ref().tpref_planner_create(N, self.D, rounding, safety, float(delta), dt_ns,
                                             1 if skip else 0, max_planning_iterations,
                                             max_initial_velocity_error)
        R = ref()
        if self.cart:
            assert self.D == 7 or (ik_fn is not None and jacobian_fn is not None)
            ik = ik_fn if ik_fn is not None else C.cast(lib().orc_fake_ik, C.c_void_p)
            jac = jacobian_fn if jacobian_fn is not None else C.cast(lib().orc_fake_jacobian,
                                                                    C.c_void_p)
            self._callbacks = (ik_fn, jacobian_fn)  # keep ctypes callbacks alive
            R.tpref_planner_set_cartesian(self._h, float(translation_rounding),
                                          float(max_trans_vel), float(max_rot_vel), ik, jac)
        st = R.tpref_planner_set_limits(self._h, _dp(np.ascontiguousarray(vmax, np.float64)),
                                        _dp(np.ascontiguousarray(amax, np.float64)))
        assert st == 0, st
        if initial_velocity is not None:
            iv = np.ascontiguousarray(initial_velocity, np.float64)
            R.tpref_planner_set_initial_velocity(self._h, _dp(iv))
        if self.cart:
            pw = np.ascontiguousarray(pose_waypoints, np.float64)
            self.set_status = R.tpref_planner_set_cart_waypoints(self._h, self.W, _dp(pw), _dp(wp))
        else:
            self.set_status = R.tpref_planner_set_waypoints(self._h, self.W, _dp(wp))

    def close(self):
        if self._h:
            ref().tpref_planner_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # interpreter shutdown
            pass

    def set_waypoints(self, waypoints):
        wp = np.ascontiguousarray(waypoints, np.float64)
        return ref().tpref_planner_set_waypoints(self._h, wp.shape[0], _dp(wp))

    def plan(self, start_ns, horizon_ns):
        return ref().tpref_planner_plan(self._h, int(start_ns), int(horizon_ns))

    def switch(self, keep_path_until, waypoints, initial_velocity=None, pose_waypoints=None):
        wp = np.ascontiguousarray(waypoints, np.float64)
        if self.cart:
            pw = np.ascontiguousarray(pose_waypoints, np.float64)
            st = ref().tpref_planner_cart_switch(self._h, float(keep_path_until), wp.shape[0],
                                                 _dp(pw), _dp(wp))
        else:
            st = ref().tpref_planner_switch(self._h, float(keep_path_until), wp.shape[0],
                                            _dp(wp))
        if initial_velocity is not None:
            iv = np.ascontiguousarray(initial_velocity, np.float64)
            st2 = ref().tpref_planner_set_initial_velocity(self._h, _dp(iv))
            st = st or st2
        return st

    # Cartesian path only
    def pose_points(self):
        out = np.zeros((4096, 7))
        n = ref().tpref_planner_cart_pose_points(self._h, _dp(out), 4096)
        return out[:n].copy()

    def ik_positions(self):
        out = np.zeros((1 << 16, self.D))
        n = ref().tpref_planner_cart_ik_positions(self._h, _dp(out), 1 << 16)
        return out[:n].copy()

    def targets(self):
        pt, jt = np.zeros((1 << 16, 7)), np.zeros((1 << 16, self.D))
        n = ref().tpref_planner_cart_targets(self._h, _dp(pt), _dp(jt), 1 << 16)
        return pt[:n].copy(), jt[:n].copy()

    def sample(self, path_start):
        """SamplePath(path_start) + ConstraintSetup() on the path alone."""
        return ref().tpref_planner_cart_sample(self._h, float(path_start))

    def stop_parameter(self, time_ns):
        out = C.c_double(float("nan"))
        st = ref().tpref_planner_stop_parameter(self._h, int(time_ns), C.byref(out))
        return st, out.value

    def state(self):
        a = np.zeros(11, np.int64)
        ref().tpref_planner_state(self._h, a.ctypes.data_as(C.POINTER(C.c_int64)))
        return dict(zip(self._STATE, (int(x) for x in a)))

    def scalars(self):
        a = np.zeros(4)
        ref().tpref_planner_scalars(self._h, _dp(a))
        return dict(delta=a[0], knot_end=a[1], path_horizon=a[2], path_start=a[3])

    def spline(self):
        cnt = (C.c_int * 2)()
        kn, pts = np.zeros(4096), np.zeros((4096, self.D))
        ref().tpref_planner_spline(self._h, cnt, _dp(kn), 4096, _dp(pts), 4096)
        return kn[:cnt[0]].copy(), pts[:cnt[1]].copy()

    def trajectory(self):
        M, D = self.state()["M"], self.D
        r = {k: np.zeros(M) for k in ("r_t", "r_s", "r_sd", "r_sdd")}
        r.update({k: np.zeros((M, D)) for k in ("r_pos", "r_vel", "r_acc")})
        ref().tpref_planner_rows(self._h, _dp(r["r_t"]), _dp(r["r_pos"]), _dp(r["r_vel"]),
                                 _dp(r["r_acc"]), _dp(r["r_s"]), _dp(r["r_sd"]), _dp(r["r_sdd"]))
        return r

    def path_samples(self):
        n, D = self.state()["len"], self.D
        w = {k: np.zeros(n) for k in ("w_time", "w_s", "w_sd", "w_sdd")}
        w.update({k: np.zeros((n, D)) for k in ("w_q", "w_vel", "w_acc")})
        ref().tpref_planner_path_samples(self._h, _dp(w["w_time"]), _dp(w["w_s"]), _dp(w["w_sd"]),
                                         _dp(w["w_sdd"]), _dp(w["w_q"]), _dp(w["w_vel"]),
                                         _dp(w["w_acc"]))
        return w

    def window(self):
        N, D = self.N, self.D
        o = {k: np.full(N, np.nan) for k in ("time", "s", "sd", "sdd")}
        o.update({k: np.zeros((N, D)) for k in ("q", "dq", "ddq")})
        ref().tpref_planner_window(self._h, _dp(o["q"]), _dp(o["dq"]), _dp(o["ddq"]),
                                   _dp(o["time"]), _dp(o["s"]), _dp(o["sd"]), _dp(o["sdd"]))
        return o

    def constraints(self):
        c = np.zeros((self.N, 2 * self.D + (2 if self.cart else 0), 4))
        ref().tpref_planner_constraints(self._h, _dp(c))
        return c


def plan_joint_ref(waypoints, vmax, amax, N, rounding=0.2, safety=0.8, delta=-1.0, dt=0.004,
                   skip=False, start_ns=0):
    """plan_joint() run by NOTHING BUT the reference's object code: SetWaypoints + one
    Plan() of a fresh path (whole path in one window when delta <= 0).  Same keys as
    plan_joint(); `status` is the absl code."""
    p = RefPlanner(waypoints, vmax, amax, N, delta, dt, rounding, safety, skip)
    # a 1 ns horizon: Plan() stops after the first window also when that window ends
    # inside the path (time_horizon_reached, path_timing_trajectory.cc:645-647)
    st = p.set_status or p.plan(start_ns, 1)
    res = p.window()
    w = p.path_samples()
    sc, state = p.scalars(), p.state()
    res.update(p.trajectory())
    ok = st == 0
    res.update(status=st, ok=ok, delta=sc["delta"], knot_end=sc["knot_end"], M=state["M"],
               last_extremal_index=state["last_extremal_index"],
               duration=(res["time"][-1] - res["time"][0]) if ok else float("nan"),
               vel=w["w_vel"], acc=w["w_acc"], cons=p.constraints(), spline=p.spline())
    p.close()
    return res


def spline_eval_batch_ref(knots, ctrl, degree, nder, u, threads=1):
    """BASELINE cfg3 by the reference's own BSplineXd::EvalCurveAndDerivatives object code:
    knots [S][n+degree+1], ctrl [S][n][D], u [S][E] -> (out [S][E][nder+1][D], status [S][E])."""
    knots = np.ascontiguousarray(knots, np.float64)
    ctrl = np.ascontiguousarray(ctrl, np.float64)
    u = np.ascontiguousarray(u, np.float64)
    S, n, D = ctrl.shape
    E = u.shape[1]
    out = np.full((S, E, nder + 1, D), np.nan)
    status = np.zeros((S, E), np.int32)
    rc = ref().tpref_spline_eval_batch(S, degree, n, D, _dp(knots), _dp(ctrl), E, _dp(u), nder,
                                       int(threads), _dp(out),
                                       status.ctypes.data_as(C.POINTER(C.c_int32)))
    assert rc == 0, rc
    return out, status


def plan_joint_batch_ref(waypoints, vmax, amax, N, dt=0.004, rounding=0.2, safety=0.8,
                         skip=False, threads=1, r_offset=None, want_time=False):
    """B fresh joint paths through the reference's own classes on `threads` host threads
    (tpref_plan_joint_batch: no Python in the loop).  Returns status [B] (absl codes),
    duration [B], rows [B] and, on request, the positions packed at r_offset and the
    profiles' time stamps."""
    wp = np.ascontiguousarray(waypoints, np.float64)
    B, W, D = wp.shape
    vmax = np.ascontiguousarray(vmax, np.float64)
    amax = np.ascontiguousarray(amax, np.float64)
    status, rows = np.zeros(B, np.int32), np.zeros(B, np.int32)
    duration = np.zeros(B)
    off = r_pos = None
    if r_offset is not None:
        off = np.ascontiguousarray(r_offset, np.int64)
        r_pos = np.full((int(off[-1]), D), np.nan)
    ptime = np.full((B, N), np.nan) if want_time else None
    i32 = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))
    ref().tpref_plan_joint_batch(
        B, N, D, W, _dp(wp), _dp(vmax), _dp(amax), rounding, safety, int(round(dt * 1e9)),
        1 if skip else 0, int(threads), i32(status), _dp(duration), i32(rows),
        off.ctypes.data_as(C.POINTER(C.c_int64)) if off is not None else None, _dp(r_pos),
        _dp(ptime))
    out = dict(status=status, duration=duration, rows=rows)
    if r_pos is not None:
        out["r_pos"] = r_pos
    if ptime is not None:
        out["time"] = ptime
    return out


def plan_cart_batch_ref(pose_waypoints, joint_waypoints, vmax, amax, N, dt=0.004,
                        rotation_rounding=0.2, translation_rounding=0.05, safety=0.8,
                        max_trans_vel=0.5, max_rot_vel=0.25, skip=False, threads=1,
                        want_time=False):
    """B fresh Cartesian paths (7 dofs, the reference test's fake kinematics as the user
    callbacks) through the reference's own classes on `threads` host threads
    (tpref_plan_cart_batch).  Returns status [B] (absl codes), duration [B], rows [B]."""
    pw = np.ascontiguousarray(pose_waypoints, np.float64)
    jw = np.ascontiguousarray(joint_waypoints, np.float64)
    B, W, D = jw.shape
    assert D == 7 and pw.shape == (B, W, 7)
    vmax = np.ascontiguousarray(vmax, np.float64)
    amax = np.ascontiguousarray(amax, np.float64)
    status, rows = np.zeros(B, np.int32), np.zeros(B, np.int32)
    duration = np.zeros(B)
    ptime = np.full((B, N), np.nan) if want_time else None
    i32 = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))
    ref().tpref_plan_cart_batch(
        B, N, D, W, _dp(pw), _dp(jw), _dp(vmax), _dp(amax), rotation_rounding,
        translation_rounding, safety, max_trans_vel, max_rot_vel, int(round(dt * 1e9)),
        1 if skip else 0, int(threads), C.cast(lib().orc_fake_ik, C.c_void_p),
        C.cast(lib().orc_fake_jacobian, C.c_void_p), i32(status), _dp(duration), i32(rows),
        _dp(ptime))
    out = dict(status=status, duration=duration, rows=rows)
    if ptime is not None:
        out["time"] = ptime
    return out


class Spline:
    """A degree-2 BSplineT with the reference's capacity rules (orc_spline)."""

    def __init__(self, knots, points, knot_capacity=None):
        knots = np.asarray(knots, np.float64)
        points = np.asarray(points, np.float64)
        D = points.shape[1]
        cap = knot_capacity or max(2 * len(knots), 100)
        self._knots, self._points = np.zeros(cap), np.zeros((cap - 3, D))
        self._knots[:len(knots)] = knots
        self._points[:len(points)] = points
        self.c = _Spline(2, D, cap, len(knots), len(points), _dp(self._knots), _dp(self._points))

    @property
    def knots(self):
        return self._knots[:self.c.num_knots].copy()

    @property
    def points(self):
        return self._points[:self.c.num_points].copy()

    def insert_knot(self, knot, multiplicity=1):
        return lib().orc_spline_insert_knot(C.byref(self.c), float(knot), int(multiplicity))

    def truncate(self, u_end):
        return lib().orc_spline_truncate(C.byref(self.c), float(u_end))

    def extend(self, points):
        pts = np.ascontiguousarray(points, np.float64)
        return lib().orc_spline_extend(C.byref(self.c), _dp(pts), pts.shape[0])

    def switch(self, keep_path_until, waypoints, rounding=0.2):
        wp = np.ascontiguousarray(waypoints, np.float64)
        new_ctrl = np.zeros((3 * (wp.shape[0] + 1), wp.shape[1]))
        n = C.c_int(0)
        st = lib().orc_joint_switch(C.byref(self.c), float(keep_path_until), _dp(wp), wp.shape[0],
                                    float(rounding), _dp(new_ctrl), C.byref(n))
        return st, new_ctrl[:n.value].copy()

    def eval(self, u, nder=0):
        return eval_curve_ders(self.knots, self.points, 2, nder, u)


def project_point_on_path(waypoints, point):
    """ProjectPointOnPath -> dict(status, waypoint_index, distance_to_path, line_parameter,
    projected_point)."""
    wp = np.ascontiguousarray(waypoints, np.float64)
    pt = np.ascontiguousarray(point, np.float64)
    W, D = wp.shape if wp.size else (0, len(pt))
    idx, dist, t, proj = C.c_int(0), C.c_double(0), C.c_double(0), np.zeros(D)
    st = lib().orc_project_point_on_path(_dp(wp), W, D, _dp(pt), C.byref(idx), C.byref(dist),
                                         C.byref(t), _dp(proj))
    return dict(status=st, waypoint_index=idx.value, distance_to_path=dist.value,
                line_parameter=t.value, projected_point=proj)


def resample_skip(plan, amax, start_sec, dt):
    """ResampleSkippingSamplesCloserThanTimeStep on the path samples of plan_joint's result."""
    N, D = plan["q"].shape
    amax = np.ascontiguousarray(amax, np.float64)
    r = {"r_t": np.zeros(N + 1), "r_s": np.zeros(N + 1), "r_sd": np.zeros(N + 1),
         "r_sdd": np.zeros(N + 1), "r_pos": np.zeros((N + 1, D)), "r_vel": np.zeros((N + 1, D)),
         "r_acc": np.zeros((N + 1, D))}
    M = lib().orc_resample_skip(N, D, _dp(plan["time"]), _dp(plan["s"]), _dp(plan["sd"]),
                                _dp(plan["sdd"]), _dp(plan["q"]), _dp(plan["vel"]),
                                _dp(plan["acc"]), _dp(amax), float(start_sec), float(dt),
                                _dp(r["r_t"]), _dp(r["r_pos"]), _dp(r["r_vel"]), _dp(r["r_acc"]),
                                _dp(r["r_s"]), _dp(r["r_sd"]), _dp(r["r_sdd"]))
    out = {k: v[:M].copy() for k, v in r.items()}
    out["M"] = M
    return out


# ----------------------------------------------------------------- stopping trajectory
def rescale_to_stop(max_acc, times, positions, velocities, accelerations, impl="port"):
    """RescaleTrajectoryBackwardToStop.  impl: 'port' | 'ref'.
    Returns dict(status, count, switch_index, times, positions, velocities, accelerations)."""
    t = np.ascontiguousarray(times, np.float64)
    p = np.ascontiguousarray(positions, np.float64)
    v = np.ascontiguousarray(velocities, np.float64)
    a = np.ascontiguousarray(accelerations, np.float64)
    amax = np.ascontiguousarray(max_acc, np.float64)
    n, D = v.shape
    ot, ov, oa = np.zeros(n), np.zeros((n, D)), np.zeros((n, D))
    cnt, sw = C.c_int(0), C.c_int(0)
    if impl == "port":
        st = lib().orc_rescale_to_stop(n, D, _dp(amax), _dp(t), _dp(v), _dp(a), C.byref(cnt),
                                       C.byref(sw), _dp(ot), _dp(ov), _dp(oa))
        m = cnt.value
        op = p[sw.value:sw.value + m].copy()
    else:
        op = np.zeros((n, D))
        st = ref().tpref_rescale_to_stop(n, D, _dp(amax), _dp(t), _dp(p), _dp(v), _dp(a),
                                         C.byref(cnt), C.byref(sw), _dp(ot), _dp(op),
                                         _dp(ov), _dp(oa))
        m = cnt.value
        op = op[:m].copy()
    return dict(status=st, count=m, switch_index=sw.value, times=ot[:m].copy(), positions=op,
                velocities=ov[:m].copy(), accelerations=oa[:m].copy())


def fastest_stop(stop_acc, times, velocities, accelerations):
    """ComputeFastestStop: dict(stop_index, duration, time, rate2, diff)."""
    t = np.ascontiguousarray(times, np.float64)
    v = np.ascontiguousarray(velocities, np.float64)
    a = np.ascontiguousarray(accelerations, np.float64)
    sa = np.ascontiguousarray(stop_acc, np.float64)
    n, D = v.shape
    dur = np.zeros(1)
    st, sr, sd = np.zeros(n + 1), np.zeros(n + 1), np.zeros(n + 1)
    ns = C.c_int(0)
    idx = lib().orc_fastest_stop(n, D, _dp(sa), _dp(t), _dp(v), _dp(a), _dp(dur), _dp(st),
                                 _dp(sr), _dp(sd), C.byref(ns))
    k = ns.value
    return dict(stop_index=idx, duration=float(dur[0]), time=st[:k].copy(), rate2=sr[:k].copy(),
                diff=sd[:k].copy())


def path_stop_parameter(time, path_parameter, velocities, accelerations, max_acc, query_time):
    """GetPathStopParameter on a resampled trajectory: dict(status, parameter, row, duration)."""
    t = np.ascontiguousarray(time, np.float64)
    s = np.ascontiguousarray(path_parameter, np.float64)
    v = np.ascontiguousarray(velocities, np.float64)
    a = np.ascontiguousarray(accelerations, np.float64)
    am = np.ascontiguousarray(max_acc, np.float64)
    M, D = v.shape
    par, dur = np.zeros(1), np.zeros(1)
    row = C.c_int(0)
    st = lib().orc_path_stop_parameter(M, D, _dp(t), _dp(s), _dp(v), _dp(a), _dp(am),
                                       float(query_time), _dp(par), C.byref(row), _dp(dur))
    return dict(status=st, parameter=float(par[0]), row=row.value, duration=float(dur[0]))


# ----------------------------------------------------------------- Cartesian path
def _q(a):
    return np.ascontiguousarray(a, np.float64)


def quat_mul(a, b):
    out = np.zeros(4)
    lib().orc_quat_mul(_dp(_q(a)), _dp(_q(b)), _dp(out))
    return out


def quat_exp(q):
    out = np.zeros(4)
    lib().orc_quat_exp(_dp(_q(q)), _dp(out))
    return out


def quat_log(q):
    out = np.zeros(4)
    lib().orc_quat_log(_dp(_q(q)), _dp(out))
    return out


def quat_power(q, power):
    out = np.zeros(4)
    lib().orc_quat_power(_dp(_q(q)), float(power), _dp(out))
    return out


def quat_power_derivative(q, power):
    out = np.zeros(4)
    lib().orc_quat_power_derivative(_dp(_q(q)), float(power), _dp(out))
    return out


def quat_eval_curve_derivative(knots, quats, degree, u):
    """BSplineQ::EvalCurveAndDerivative; returns (status, q, dq/du)."""
    knots, quats = _q(knots), _q(quats)
    out, dout = np.zeros(4), np.zeros(4)
    st = lib().orc_quat_eval_curve_derivative(_dp(knots), len(knots), _dp(quats), degree, float(u),
                                          _dp(out), _dp(dout))
    return st, out, dout


def quat_eval_curve(knots, quats, degree, u):
    """BSplineQ::EvalCurve; quats [n][4] in (x, y, z, w) order."""
    knots, quats = _q(knots), _q(quats)
    out = np.zeros(4)
    st = lib().orc_quat_eval_curve(_dp(knots), len(knots), _dp(quats), degree, float(u),
                                   _dp(out))
    return st, out


def quat_ref(kind, q, power=None):
    """bsplineq_details::QuatLog / QuatExp / QuatPower / QuatPowerDerivative, reference
    object code (kind: "log", "exp", "power", "power_derivative")."""
    out = np.zeros(4)
    f = getattr(ref(), "tpref_quat_" + kind)
    if power is None:
        f(_dp(_q(q)), _dp(out))
    else:
        f(_dp(_q(q)), float(power), _dp(out))
    return out


def bsplineq_eval_ref(knots, quats, degree, us, derivative=False):
    """BSplineQ::EvalCurve / EvalCurveAndDerivative, reference object code.  Returns
    (status [M], q [M][4], dq [M][4] or None)."""
    knots, quats, us = _q(knots), _q(quats), _q(np.atleast_1d(us))
    M = len(us)
    q, dq, st = np.zeros((M, 4)), np.zeros((M, 4)), np.zeros(M, np.int32)
    rc = ref().tpref_bsplineq_eval(degree, len(knots), _dp(knots), quats.shape[0], _dp(quats), M,
                                   _dp(us), _dp(q), _dp(dq) if derivative else None,
                                   st.ctypes.data_as(C.POINTER(C.c_int32)))
    assert rc == 0, rc
    return st, q, (dq if derivative else None)


def pose_polyline_to_bspline3_ref(corners, translation_radius, rotational_radius):
    corners = _q(corners)
    W = corners.shape[0]
    out = np.zeros((max(3 * W - 2, 4) + 4, 7))
    n = ref().tpref_pose_polyline_to_bspline3(_dp(corners), W, float(translation_radius),
                                              float(rotational_radius), _dp(out), out.shape[0])
    return out[:n]


def pose_polyline_to_bspline3(corners, translation_radius, rotational_radius):
    corners = _q(corners)
    W = corners.shape[0]
    out = np.zeros((max(3 * W - 2, 4), 7))
    n = lib().orc_pose_polyline_to_bspline3(_dp(corners), W, float(translation_radius),
                                            float(rotational_radius), _dp(out))
    return out[:n]


def exp_so3(v):
    out = np.zeros(4)
    lib().orc_exp_so3(_dp(_q(v)), _dp(out))
    return out


def log_so3(q):
    out = np.zeros(3)
    lib().orc_log_so3(_dp(_q(q)), _dp(out))
    return out


def log_so3_derivative(q):
    out = np.zeros((3, 4))
    lib().orc_log_so3_derivative(_dp(_q(q)), _dp(out))
    return out


def fake_jacobian(joints):
    out = np.zeros((6, 7))
    lib().orc_fake_jacobian(_dp(_q(joints)), _dp(out))
    return out


def fake_ik(pose_targets, joint_targets):
    pt, jt = _q(pose_targets), _q(joint_targets)
    out = np.zeros((pt.shape[0], 7))
    lib().orc_fake_ik(_dp(pt), _dp(jt), pt.shape[0], _dp(out))
    return out


def cart_fit(pose_wp, joint_wp, translation_rounding=0.05, rotation_rounding=0.2):
    pose_wp, joint_wp = _q(pose_wp), _q(joint_wp)
    W, D = joint_wp.shape
    nmax = max(3 * W - 2, 4)
    pc, jc, kn = np.zeros((nmax, 7)), np.zeros((nmax, D)), np.zeros(nmax + 3)
    n = lib().orc_cart_fit(_dp(pose_wp), _dp(joint_wp), W, D, float(translation_rounding),
                           float(rotation_rounding), _dp(pc), _dp(jc), _dp(kn))
    return pc[:n], jc[:n], kn[:n + 3]


def cart_sample_targets(knots, pose_ctrl, joint_ctrl, N, delta):
    knots, pose_ctrl, joint_ctrl = _q(knots), _q(pose_ctrl), _q(joint_ctrl)
    D = joint_ctrl.shape[1]
    pt, jt = np.zeros((N + 1, 7)), np.zeros((N + 1, D))
    lib().orc_cart_sample_targets(_dp(knots), len(knots), _dp(pose_ctrl), _dp(joint_ctrl), D,
                                  N, float(delta), _dp(pt), _dp(jt))
    return pt, jt


def cart_derivatives(pos, delta):
    pos = _q(pos)
    N, D = pos.shape
    d1, d2 = np.zeros((N, D)), np.zeros((N, D))
    lib().orc_cart_derivatives(_dp(pos), N, D, float(delta), _dp(d1), _dp(d2))
    return d1, d2


def cart_constraints(d1, d2, jac, vmax, amax, safety, max_trans_vel, max_rot_vel):
    d1, d2, jac, vmax, amax = _q(d1), _q(d2), _q(jac), _q(vmax), _q(amax)
    N, D = d1.shape
    cons = np.zeros((N, 2 * D + 2, 4))
    lib().orc_cart_constraints(_dp(d1), _dp(d2), _dp(jac), N, D, _dp(vmax), _dp(amax),
                               float(safety), float(max_trans_vel), float(max_rot_vel),
                               _dp(cons))
    return cons


def plan_cart_fake(pose_wp, joint_wp, vmax, amax, N, translation_rounding=0.05,
                   rotation_rounding=0.2, safety=0.8, delta=-1.0, start_sec=0.0, dt=0.0,
                   M_cap=0, max_trans_vel=0.5, max_rot_vel=0.25):
    """One fresh 7-dof Cartesian path with the reference test's fake kinematics."""
    pose_wp, joint_wp, vmax, amax = _q(pose_wp), _q(joint_wp), _q(vmax), _q(amax)
    W, D = joint_wp.shape
    assert D == 7
    cfg = _CartCfg(N, W, translation_rounding, rotation_rounding, safety, delta, start_sec,
                   dt, max_trans_vel, max_rot_vel)
    o = {k: np.full(N, np.nan) for k in ("time", "s", "sd", "sdd")}
    o.update({k: np.zeros((N, D)) for k in ("q", "dq", "ddq", "vel", "acc")})
    r = {"r_t": np.zeros(M_cap), "r_s": np.zeros(M_cap), "r_sd": np.zeros(M_cap),
         "r_sdd": np.zeros(M_cap), "r_pos": np.zeros((M_cap, D)),
         "r_vel": np.zeros((M_cap, D)), "r_acc": np.zeros((M_cap, D))}
    out = _PlanOut(**{k: _dp(v) for k, v in o.items()}, M_cap=M_cap,
                   **{k: (_dp(v) if M_cap > 0 else None) for k, v in r.items()})
    nmax = max(3 * W - 2, 4)
    pose_ctrl, knots = np.zeros((nmax, 7)), np.zeros(nmax + 3)
    pose_targets, cons = np.zeros((N + 1, 7)), np.zeros((N, 2 * D + 2, 4))
    st = lib().orc_plan_cart_fake(C.byref(cfg), _dp(pose_wp), _dp(joint_wp), _dp(vmax),
                                  _dp(amax), C.byref(out), _dp(pose_ctrl), _dp(knots),
                                  _dp(pose_targets), _dp(cons))
    res = dict(o)
    M = out.M
    res.update({k: v[:M] for k, v in r.items()})
    n = 4 if W == 1 else 3 * W - 2
    res.update(status=st, ok=(st == ORC_OK), duration=out.duration, delta=out.delta_used,
               knot_end=out.knot_end, M=M, last_extremal_index=out.last_extremal_index,
               pose_ctrl=pose_ctrl[:n], knots=knots[:n + 3], pose_targets=pose_targets,
               cons=cons)
    return res


def time_from_sec(x):
    return lib().orc_time_from_sec(float(x))


def time_to_sec(ns):
    return lib().orc_time_to_sec(int(ns))


# ----------------------------------------------------------------- trajectory buffer
class TrajectoryBuffer:
    """One TrajectoryBuffer (trajectory_buffer.h:30-184).  impl: 'port' (the C
    restatement) | 'ref' (the reference's object code).  Status returns are
    absl::StatusCode numbers."""

    def __init__(self, num_dofs, timestep_tolerance=1e-6, impl="port"):
        self.D, self.port = int(num_dofs), impl == "port"
        self.L = lib() if self.port else ref()
        self.pre = "orc_tbuf_" if self.port else "tpref_tbuf_"
        self.h = C.c_void_p()
        args = (timestep_tolerance, self.D) if self.port else (timestep_tolerance,)
        self.create_status = self._f("create")(*args, C.byref(self.h))

    def _f(self, name):
        return getattr(self.L, self.pre + name)

    def _d(self):
        return () if self.port else (self.D,)

    def close(self):
        if self.h:
            self._f("destroy")(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        self.close()

    def clear(self):
        self._f("clear")(self.h)

    def insert_segment(self, times, pos, vel, acc):
        t = np.ascontiguousarray(times, np.float64)
        p, v, a = (np.ascontiguousarray(x, np.float64).reshape(len(t), self.D)
                   for x in (pos, vel, acc))
        return self._f("insert_segment")(self.h, len(t), *self._d(), _dp(t), _dp(p), _dp(v), _dp(a))

    def append_sample(self, time, pos, vel, acc):
        p, v, a = (np.ascontiguousarray(x, np.float64) for x in (pos, vel, acc))
        return self._f("append_sample")(self.h, *self._d(), float(time), _dp(p), _dp(v), _dp(a))

    def discard_before_sec(self, t):
        self._f("discard_before_sec")(self.h, float(t))

    def discard_before_ns(self, ns):
        self._f("discard_before_ns")(self.h, int(ns))

    def positions_up_to_time_ns(self, ns):
        return self._f("positions_up_to_time_ns")(self.h, int(ns))

    def get_at_time(self, which, t):
        out = np.zeros(self.D)
        st = self._f("get_at_time")(self.h, int(which), float(t), _dp(out))
        return st, out

    def stop_at_index(self, index, max_acc, time_step):
        am = np.ascontiguousarray(max_acc, np.float64)
        return self._f("stop_at_index")(self.h, int(index), *self._d(), _dp(am), float(time_step))

    def stop_before_time(self, t, max_acc, time_step):
        am = np.ascontiguousarray(max_acc, np.float64)
        return self._f("stop_before_time")(self.h, float(t), *self._d(), _dp(am), float(time_step))

    def add_offset(self, offset):
        self._f("add_offset")(self.h, float(offset))

    @property
    def num_samples(self):
        return self._f("num_samples")(self.h)

    @property
    def sequence_number(self):
        return self._f("sequence_number")(self.h)

    @property
    def start_ns(self):
        return self._f("start_ns")(self.h)

    @property
    def end_ns(self):
        return self._f("end_ns")(self.h)

    def rows(self):
        n = self.num_samples
        t = np.zeros(n)
        p, v, a = np.zeros((n, self.D)), np.zeros((n, self.D)), np.zeros((n, self.D))
        self._f("rows")(self.h, *self._d(), _dp(t), _dp(p), _dp(v), _dp(a))
        return t, p, v, a
