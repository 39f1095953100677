"""Moving-horizon PathTimingTrajectory::Plan on the GPU (tpb_planner_reset /
tpb_planner_plan): every call of a re-planning loop compared bit for bit with the CPU
oracle's planner (oracle/tp_oracle_windows.c), per path of a batch, plus the
properties of the reference's integration tests
(path_timing_trajectory_test.cc:112-297)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as orc  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
ARR = pkg.ARR
pytestmark = pytest.mark.gpu

MS = 1_000_000
REF_WAYPOINTS = np.array([[1, 2, 3], [-1, -2, -3], [1, 2, 3.0]])  # _test.cc:127-128


def device_state(eng):
    keys = dict(status=ARR.PLAN_STATUS, loops=ARR.PLAN_LOOPS, planned_to_end=ARR.PLAN_PLANNED_TO_END,
                target_reached=ARR.PLAN_TARGET_REACHED, start_time=ARR.PLAN_START_TIME,
                end_time=ARR.PLAN_END_TIME, final_decel_start=ARR.PLAN_FINAL_DECEL_START,
                erased=ARR.PLAN_ERASED, len=ARR.W_LEN)
    return {k: eng.get(v) for k, v in keys.items()}


def device_rows(eng):
    off = eng.get(ARR.R_OFFSET)
    rows = dict(r_t=eng.get(ARR.R_TIME), r_pos=eng.get(ARR.R_POS), r_vel=eng.get(ARR.R_VEL),
                r_acc=eng.get(ARR.R_ACC), r_s=eng.get(ARR.R_S), r_sd=eng.get(ARR.R_SD),
                r_sdd=eng.get(ARR.R_SDD))
    return off, rows


def compare_call(eng, planners, statuses, what):
    st = device_state(eng)
    off, rows = device_rows(eng)
    for b, (pl, ost) in enumerate(zip(planners, statuses)):
        tag = f"{what} path {b}"
        assert st["status"][b] == ost, tag
        assert st["len"][b] == pl.len, tag
        assert st["start_time"][b] == pl.start_time, tag
        assert st["end_time"][b] == pl.end_time, tag
        assert st["final_decel_start"][b] == pl.final_decel_start, tag
        assert st["planned_to_end"][b] == pl.planned_to_end, tag
        assert st["target_reached"][b] == pl.target_reached, tag
        if ost == orc.ORC_OK:
            assert st["loops"][b] == pl.loops, tag
            assert (st["erased"][b] == 1) == bool(pl.erased), tag
        tr = pl.trajectory()
        assert off[b + 1] - off[b] == pl.M, tag
        for k, v in tr.items():
            assert rows[k][off[b]:off[b + 1]].tobytes() == v.tobytes(), f"{tag} {k}"
    return st


def compare_path_samples(eng, planners):
    wt, ws, wsd, wsdd = (eng.get(a) for a in (ARR.W_TIME, ARR.W_S, ARR.W_SD, ARR.W_SDD))
    wq, wdq = eng.get(ARR.W_Q), eng.get(ARR.W_DQ)
    for b, pl in enumerate(planners):
        ps, n = pl.path_samples(), pl.len
        assert wt[b, :n].tobytes() == ps["w_time"].tobytes()
        assert ws[b, :n].tobytes() == ps["w_s"].tobytes()
        assert wsd[b, :n].tobytes() == ps["w_sd"].tobytes()
        assert wsdd[b, :n].tobytes() == ps["w_sdd"].tobytes()
        assert wq[b, :n].tobytes() == ps["w_q"].tobytes()
        assert (wdq[b, :n] * wsd[b, :n, None]).tobytes() == ps["w_vel"].tobytes()


def run_replanning(wp, vmax, amax, N, delta, dt, skip, horizon, replan, P_cap, max_calls=400,
                   initial_velocity=None):
    """Plan() at start, start + replan, ... on the device batch and per path in the
    oracle until every path is at its end; every call compared."""
    B, W, D = wp.shape
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 40000)
    eng.set_joint_limits(vmax, amax)
    eng.set_joint_waypoints(wp)
    eng.set_sampling(np.full(B, delta))
    eng.planner_reset(P_cap, dt, skip=skip, initial_velocity=initial_velocity)
    planners = [orc.Planner(wp[b], vmax[b], amax[b], N, delta, dt, skip=skip, P_cap=P_cap,
                            initial_velocity=None if initial_velocity is None
                            else initial_velocity[b]) for b in range(B)]
    start, calls, erased, windows = 0, 0, 0, 0
    done = np.zeros(B, bool)
    while not done.all():
        statuses = [pl.plan(start, horizon) for pl in planners]
        total = eng.planner_plan(start, horizon)
        st = compare_call(eng, planners, statuses, f"call {calls} start {start}")
        assert total == sum(pl.M for pl in planners)
        erased += int((st["erased"] == 1).sum())
        windows += int(st["loops"].sum())
        done |= (st["target_reached"] == 1) | (st["status"] != 0)
        start += replan
        calls += 1
        assert calls < max_calls
    compare_path_samples(eng, planners)
    return eng, planners, dict(calls=calls, erased=erased, windows=windows)


@pytest.mark.parametrize("skip", [False, True])
def test_reference_rest_to_rest_replanning(skip):
    """RestToRestPlanningWorks (_test.cc:112-173), N = 1000, 4 ms, horizon 750 ms,
    re-planned every 200 ms."""
    wp = REF_WAYPOINTS[None]
    eng, planners, n = run_replanning(wp, np.ones((1, 3)), np.full((1, 3), 2.0), 1000, 0.005,
                                      0.004, skip, 750 * MS, 200 * MS, 8192)
    assert n["erased"] > 10 and n["windows"] >= 3
    off, rows = device_rows(eng)
    np.testing.assert_allclose(rows["r_pos"][-1], REF_WAYPOINTS[-1], atol=1e-12)
    np.testing.assert_array_equal(rows["r_vel"][-1], 0.0)
    assert np.all(np.abs(rows["r_vel"]) <= 1.0 + 1e-9) and np.all(np.abs(rows["r_acc"]) <= 2.0 + 1e-9)


@pytest.mark.parametrize("skip", [False, True])
def test_batch_of_seven_dof_paths(skip):
    """A batch of different 7-dof paths re-planned in lockstep; the paths need different
    numbers of windows and finish at different times (late calls fail the time argument
    checks for the finished ones exactly as the reference's would)."""
    B = 12
    wp, vmax, amax = pkg.workloads.joint_paths(B, 6, 7, seed=77)
    knot_end = max(orc.joint_fit(wp[b], 0.2)[1][-1] for b in range(B))
    N = 256
    delta = knot_end / (N - 1) / 3.2
    eng, planners, n = run_replanning(wp, vmax, amax, N, delta, 0.008, skip, 600 * MS, 250 * MS,
                                      16 * N)
    assert n["windows"] >= 3 * B and n["erased"] > 0
    assert all(pl.target_reached for pl in planners)


def test_one_call_long_horizon_and_statuses():
    """One Plan() with a horizon longer than the motion loops over the windows on its
    own (:628-658); capacity / deadline / argument failures are reported per path."""
    B = 6
    wp, vmax, amax = pkg.workloads.joint_paths(B, 5, 7, seed=5)
    knot_end = [orc.joint_fit(wp[b], 0.2)[1][-1] for b in range(B)]
    N, dt = 192, 0.004
    delta = max(knot_end) / (N - 1) / 4.0
    long_horizon = 3600 * 1000 * MS

    def engine(P_cap, iters):
        eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=7, max_waypoints=5,
                         max_time_samples_total=B * 60000)
        eng.set_joint_limits(vmax, amax)
        eng.set_joint_waypoints(wp)
        eng.set_sampling(np.full(B, delta))
        eng.planner_reset(P_cap, dt, max_planning_iterations=iters)
        return eng

    eng = engine(16 * N, 200)
    planners = [orc.Planner(wp[b], vmax[b], amax[b], N, delta, dt, P_cap=16 * N) for b in range(B)]
    statuses = [pl.plan(5 * MS, long_horizon) for pl in planners]
    eng.planner_plan(5 * MS, long_horizon)
    st = compare_call(eng, planners, statuses, "long horizon")
    assert np.all(st["status"] == 0) and np.all(st["loops"] >= 3) and np.all(st["target_reached"] == 1)
    compare_path_samples(eng, planners)
    # calls that fail HandleTimeArguments leave everything as it was
    for start in (int(max(pl.end_time for pl in planners)) + 50 * MS, 1 * MS):
        statuses = [pl.plan(start, long_horizon) for pl in planners]
        eng.planner_plan(start, long_horizon)
        st = compare_call(eng, planners, statuses, f"bad start {start}")
        assert np.all(st["status"] != 0)
    # too few planning loops / too little room for the spliced samples
    for P_cap, iters, expect in ((16 * N, 1, orc.ORC_FAIL_DEADLINE), (N + N // 2, 200, orc.ORC_FAIL_CAPACITY)):
        eng = engine(P_cap, iters)
        planners = [orc.Planner(wp[b], vmax[b], amax[b], N, delta, dt, P_cap=P_cap,
                                max_planning_iterations=iters) for b in range(B)]
        statuses = [pl.plan(0, long_horizon) for pl in planners]
        eng.planner_plan(0, long_horizon)
        got = eng.get(ARR.PLAN_STATUS)
        assert list(got) == statuses and expect in statuses


def test_initial_velocity():
    """(:352-385) per path: projection onto the first tangent, rejection otherwise."""
    B = 4
    wp, vmax, amax = pkg.workloads.joint_paths(B, 4, 7, seed=9)
    v0 = np.zeros((B, 7))
    for b in range(B):
        t = wp[b, 1] - wp[b, 0]
        v0[b] = 0.03 * t / np.linalg.norm(t)
    v0[2] = 0.0
    v0[3, :] = [0.05, 0, 0, 0, 0, 0, 0]  # not along the tangent
    N, dt = 256, 0.004
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=7, max_waypoints=4,
                     max_time_samples_total=B * 40000)
    eng.set_joint_limits(vmax, amax)
    eng.set_joint_waypoints(wp)
    delta = eng.get(ARR.DELTA)
    eng.planner_reset(4 * N, dt, initial_velocity=v0)
    planners = [orc.Planner(wp[b], vmax[b], amax[b], N, 0.0, dt, P_cap=4 * N,
                            initial_velocity=v0[b]) for b in range(B)]
    statuses = [pl.plan(0, 500 * MS) for pl in planners]
    assert statuses[3] == orc.ORC_FAIL_INITIAL_VELOCITY and statuses[:3] == [0, 0, 0]
    eng.planner_plan(0, 500 * MS)
    compare_call(eng, planners, statuses, "initial velocity")
    off, rows = device_rows(eng)
    for b in range(2):
        np.testing.assert_allclose(rows["r_vel"][off[b]], v0[b], atol=1e-9)
    assert np.all(delta > 0)


# ---- Cartesian paths behind the planner (device kinematics TPB_KIN_FAKE7) -------------
def run_cart_replanning(pose_wp, joint_wp, N, delta, dt, skip, horizon, replan, P_cap,
                        max_calls=600, host_callbacks=None):
    """Like run_replanning for TimeableCartesianSplinePath with the reference test's fake
    kinematics on the device.  The device evaluates sin/cos/atan2 of the quaternion
    spline and of the fake IK with its own libm, so values are compared within the
    BASELINE tolerances (durations 1e-6 relative, positions 1e-9) and everything
    discrete (statuses, window counts, spliced sample counts, flags) exactly."""
    B, W, D = joint_wp.shape
    vmax, amax = np.full((B, D), 1.0), np.full((B, D), 2.0)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 60000, cartesian=True)
    eng.set_cart_options(0.05, 0.5, 0.25)
    eng.set_cart_waypoints(pose_wp, joint_wp)
    eng.set_joint_limits(vmax, amax)
    eng.set_sampling(np.full(B, delta))
    eng.planner_reset(P_cap, dt, skip=skip, max_initial_velocity_error=1e-2)
    with pytest.raises(pkg.TpbError) as ei:  # neither callbacks nor a device model yet
        eng.planner_plan(0, horizon)
    assert ei.value.code == 9
    if host_callbacks is None:
        eng.planner_set_kinematics(1)
    else:
        eng.planner_set_host_kinematics(*host_callbacks)
    planners = [orc.Planner(joint_wp[b], vmax[b], amax[b], N, delta, dt, skip=skip, P_cap=P_cap,
                            pose_waypoints=pose_wp[b], max_initial_velocity_error=1e-2)
                for b in range(B)]
    start, calls, windows = 0, 0, 0
    done = np.zeros(B, bool)
    while not done.all():
        statuses = [pl.plan(start, horizon) for pl in planners]
        eng.planner_plan(start, horizon)
        st = device_state(eng)
        off, rows = device_rows(eng)
        for b, (pl, ost) in enumerate(zip(planners, statuses)):
            tag = f"call {calls} path {b}"
            assert st["status"][b] == ost, tag
            assert st["len"][b] == pl.len, tag
            assert st["planned_to_end"][b] == pl.planned_to_end, tag
            assert st["target_reached"][b] == pl.target_reached, tag
            assert abs(st["end_time"][b] - pl.end_time) <= max(1e-6 * abs(pl.end_time), 1), tag
            if ost == orc.ORC_OK:
                assert st["loops"][b] == pl.loops, tag
            tr = pl.trajectory()
            M = off[b + 1] - off[b]
            assert abs(M - pl.M) <= 1, tag
            m = min(M, pl.M) - 1
            if m > 0:
                sl = slice(off[b], off[b] + m)
                assert np.max(np.abs(rows["r_t"][sl] - tr["r_t"][:m])) <= 1e-9, tag
                assert np.max(np.abs(rows["r_pos"][sl] - tr["r_pos"][:m])) <= 1e-9, tag
                assert np.max(np.abs(rows["r_vel"][sl] - tr["r_vel"][:m])) <= 1e-7, tag
        windows += int(st["loops"].sum())
        done |= (st["target_reached"] == 1) | (st["status"] != 0)
        start += replan
        calls += 1
        assert calls < max_calls
    return eng, planners, dict(calls=calls, windows=windows)


@pytest.mark.parametrize("skip", [False, True])
def test_reference_cartesian_replanning(skip):
    """The 5-waypoint Cartesian path of path_timing_trajectory_test.cc:695-757: N = 1000,
    delta 0.005, 4 ms, horizon 750 ms, re-planned every 200 ms until the end."""
    import test_oracle_windows as TW
    joint = TW.CART_WAYPOINTS[None]
    pose = np.stack([TW.fake_fk_pose(q) for q in TW.CART_WAYPOINTS])[None]
    eng, planners, n = run_cart_replanning(pose, joint, 1000, 0.005, 0.004, skip, 750 * MS,
                                           200 * MS, 16384)
    assert n["windows"] >= 3 and planners[0].target_reached
    off, rows = device_rows(eng)
    np.testing.assert_allclose(rows["r_pos"][-1], TW.CART_WAYPOINTS[-1], atol=1e-9)
    np.testing.assert_array_equal(rows["r_vel"][-1], 0.0)
    assert np.all(np.abs(rows["r_vel"]) <= 1.0 + 1e-9) and np.all(np.abs(rows["r_acc"]) <= 2.0 + 1e-9)


class CountingFakeKinematics:
    """The reference test's FakePathIK / FakeJacobian (the oracle's C restatement) as the
    user's host callbacks; records what tpb_planner_plan asks for."""

    def __init__(self, B):
        self.calls = np.zeros(B, int)
        self.samples = np.zeros(B, int)
        self.last = [None] * B
        self.warm_ok = True

    def path_ik(self, path, initial, pose_targets, joint_targets):
        if self.calls[path] > 0:  # warm start = last solution of the previous window
            self.warm_ok = self.warm_ok and np.array_equal(initial, self.last[path])
        else:                     # fresh path: the first joint target (:503-505)
            self.warm_ok = self.warm_ok and np.array_equal(initial, joint_targets[0])
        sol = orc.fake_ik(pose_targets, joint_targets)
        self.calls[path] += 1
        self.samples[path] += len(sol)
        self.last[path] = sol[-1].copy()
        return sol

    def jacobian(self, path, q):
        return orc.fake_jacobian(q)


@pytest.mark.parametrize("skip", [False, True])
def test_cartesian_replanning_with_user_kinematics(skip):
    """a user PathIKFunc / JacobianFunc over several planning windows
    (timeable_path_cartesian_spline.cc:464-549 with path_start > 0 and the IK cache
    :504-517), driven from tpb_planner_plan: the reference's 5-waypoint scenario
    (path_timing_trajectory_test.cc:695-757) equals the oracle after every Plan() (within
    the tolerances of run_cart_replanning: the targets handed to the callbacks come from
    the device's quaternion spline, i.e. CUDA's sin / cos / atan2); the IK is asked once
    per window, for the indices it has not solved yet, warm-started from the last one."""
    import test_oracle_windows as TW
    joint = TW.CART_WAYPOINTS[None]
    pose = np.stack([TW.fake_fk_pose(q) for q in TW.CART_WAYPOINTS])[None]
    kin = CountingFakeKinematics(1)
    eng, planners, n = run_cart_replanning(pose, joint, 1000, 0.005, 0.004, skip, 750 * MS,
                                           200 * MS, 16384,
                                           host_callbacks=(kin.path_ik, kin.jacobian))
    assert n["windows"] >= 3 and planners[0].target_reached
    assert kin.warm_ok and kin.calls[0] == n["windows"]
    assert 1001 < kin.samples[0] < 1001 * n["windows"]


def test_batch_of_cartesian_paths_with_user_kinematics():
    """Eight Cartesian paths in lockstep, user kinematics: per-path IK caches."""
    B = 8
    pose_wp, joint_wp, _, _ = pkg.workloads.cart_paths(B, 5, seed=5300)
    N = 256
    knot_end = max(orc.plan_cart_fake(pose_wp[b], joint_wp[b], np.ones(7), np.full(7, 2.0),
                                      64)["knot_end"] for b in range(B))
    delta = knot_end / (N - 1) / 2.5
    kin = CountingFakeKinematics(B)
    eng, planners, n = run_cart_replanning(pose_wp, joint_wp, N, delta, 0.008, False, 1000 * MS,
                                           400 * MS, 16 * N,
                                           host_callbacks=(kin.path_ik, kin.jacobian))
    assert n["windows"] >= 2 * B and all(pl.target_reached for pl in planners)
    assert kin.warm_ok and kin.calls.sum() == n["windows"]
    # a failing PathIKFunc ends the call with its status
    eng.planner_reset(16 * N, 0.008, max_initial_velocity_error=1e-2)

    def bad_ik(path, initial, pose_targets, joint_targets):
        raise RuntimeError("no solution")

    eng.planner_set_host_kinematics(bad_ik, kin.jacobian)
    with pytest.raises(pkg.TpbError) as ei:
        eng.planner_plan(0, 1000 * MS)
    assert ei.value.code == 13 and "PathIKFunc" in str(ei.value)


def test_batch_of_cartesian_paths():
    """cfg5-style Cartesian paths re-planned in lockstep over several windows each."""
    B = 8
    pose_wp, joint_wp, _, _ = pkg.workloads.cart_paths(B, 5, seed=5300)
    N = 256
    knot_end = max(orc.plan_cart_fake(pose_wp[b], joint_wp[b], np.ones(7), np.full(7, 2.0),
                                      64)["knot_end"] for b in range(B))
    delta = knot_end / (N - 1) / 2.5
    eng, planners, n = run_cart_replanning(pose_wp, joint_wp, N, delta, 0.008, False, 1000 * MS,
                                           400 * MS, 16 * N)
    assert n["windows"] >= 2 * B
    assert all(pl.target_reached for pl in planners)


@pytest.mark.parametrize("D,skip", [(12, False), (16, False), (12, True)])
def test_replanning_many_dofs_with_the_default_capacity(D, skip):
    """num_dofs 10..16 with the host classes' default path-sample capacity of 64 * N: the
    planner resamples views of P_cap samples, so the time stamps do not fit shared memory
    and the row tile alone exceeds 48 KB (k_resample_fill<false, .> with its opt-in to
    large dynamic shared memory)."""
    B, W, N = 3, 4, 160
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=70 + D)
    ends = [orc.joint_fit(wp[b], 0.2)[1][-1] for b in range(B)]
    delta = min(ends) / (N - 1) / 2.5
    eng, planners, stats = run_replanning(wp, vmax, amax, N, delta, 0.004, skip, 750 * MS, 200 * MS,
                                          P_cap=64 * N)
    assert stats["windows"] >= 2 * B
    assert all(pl.target_reached for pl in planners)
    eng.close()
