"""Online re-planning on the GPU: tpb_switch_to_waypoint_path (SwitchToWaypointPath for a
batch of joint paths) and the re-timing of the modified paths by the device planner,
bit for bit against the CPU oracle (oracle/tp_oracle_switch.c, tp_oracle_windows.c)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import pyoracle as orc  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
import test_gpu_planner as P  # noqa: E402

pkg = load_package()
ARR = pkg.ARR
pytestmark = pytest.mark.gpu
MS = 1_000_000


def device_splines(eng):
    n = eng.get(ARR.NUM_POINTS)
    ctrl, knots = eng.get(ARR.CTRL), eng.get(ARR.KNOTS)
    return [(knots[b, :n[b] + 3].copy() if n[b] else np.zeros(0), ctrl[b, :n[b]].copy())
            for b in range(len(n))]


def test_switch_matches_oracle_per_path():
    B, W, D = 16, 5, 7
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=21)
    rng = np.random.default_rng(8)
    eng = pkg.Engine(max_paths=B, num_samples=64, num_dofs=D, max_waypoints=W)
    eng.set_joint_limits(vmax, amax)
    eng.set_joint_waypoints(wp)
    splines = []
    for b in range(B):
        ctrl, knots = orc.joint_fit(wp[b], 0.2)
        splines.append(orc.Spline(knots, ctrl))
    seen = set()
    for round_ in range(24):  # repeated switches until the knot capacity (100) runs out
        kend = np.array([s.knots[-1] if len(s.knots) else 0.0 for s in splines])
        keep = rng.uniform(0.2, 0.9, B) * kend
        keep[1] = kend[1] + 1.0   # nothing to truncate, EvalCurve out of range
        keep[2] = -0.5            # clears the curve: kOutOfRange from then on
        new_wp = rng.uniform(-2, 2, size=(B, W, D))
        num = rng.integers(2, W + 1, size=B).astype(np.int32)
        num[3] = 0                # untouched
        num[4] = 1                # single waypoint
        keep[6:8] = 0.97 * kend[6:8]  # these two grow by 3 W - 2 points per switch
        num[6:8] = W
        # path 5: new waypoints start exactly at the switching position (no projected point)
        st5, pos5 = orc.eval_curve(splines[5].knots, splines[5].points, 2, keep[5])
        if st5 == 0:
            new_wp[5, 0] = pos5
        status = eng.switch_to_waypoint_path(keep, new_wp, num_waypoints=num)
        got = device_splines(eng)
        for b in range(B):
            expect = 0 if num[b] <= 0 else splines[b].switch(keep[b], new_wp[b, :num[b]])[0]
            assert status[b] == expect, (round_, b)
            seen.add(int(expect))
            # a failed switch leaves the same partially modified spline on both sides
            assert got[b][0].tobytes() == splines[b].knots.tobytes(), (round_, b)
            assert got[b][1].tobytes() == splines[b].points.tobytes(), (round_, b)
        if round_ == 0:
            assert status[1] == 11 and status[2] == 11 and status[3] == 0
            assert (status[[0, 4, 5, 6, 7]] == 0).all()
    assert seen >= {0, 9, 11}  # ok, capacity exhausted, outside the path


@pytest.mark.parametrize("skip", [False, True])
def test_switch_and_replan_matches_oracle(skip):
    """SwitchToNewJointWaypointPathWorks (path_timing_trajectory_test.cc:299-436) for a
    batch: plan, switch every path to new waypoints at its stop parameter with its current
    velocity, plan to the end; every Plan() call compared with the oracle."""
    B, W, D, N, dt = 6, 4, 3, 1000, 0.004
    rng = np.random.default_rng(11)
    wp = np.tile(np.array([[1, 2, 3], [-1, -2, -3], [0.5, 1, 1.5], [0, 0, 0.0]]), (B, 1, 1))
    wp[1:, :3] += rng.uniform(-0.3, 0.3, size=(B - 1, 3, D))
    num_wp = np.full(B, 3, np.int32)
    vmax, amax = np.ones((B, D)), np.full((B, D), 2.0)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 40000)
    eng.set_joint_limits(vmax, amax)
    eng.set_joint_waypoints(wp, num_waypoints=num_wp)
    eng.set_sampling(np.full(B, 0.001))
    eng.planner_reset(16 * N, dt, skip=skip, max_initial_velocity_error=1e-2)
    planners = [orc.Planner(wp[b, :3], vmax[b], amax[b], N, 0.001, dt, skip=skip, P_cap=16 * N,
                            max_initial_velocity_error=1e-2) for b in range(B)]
    start = 0
    for call in range(4):
        statuses = [pl.plan(start, 750 * MS) for pl in planners]
        eng.planner_plan(start, 750 * MS)
        P.compare_call(eng, planners, statuses, f"before switch {call}")
        start += 200 * MS
    # stop parameters (device, checked against the oracle) and current velocities
    stop = eng.path_stop_parameter(start / 1e9)
    new_wp = np.zeros((B, W, D))
    v0 = np.zeros((B, D))
    for b, pl in enumerate(planners):
        tr = pl.trajectory()
        o = orc.path_stop_parameter(tr["r_t"], tr["r_s"], tr["r_vel"], tr["r_acc"], amax[b], start / 1e9)
        assert stop["parameter"][b] == o["parameter"]
        new_wp[b, 0] = wp[b, 0]
        new_wp[b, 1] = [0.5, 1.0, 5.5] + rng.uniform(-0.2, 0.2, D)
        v0[b] = [np.interp(start / 1e9, tr["r_t"], tr["r_vel"][:, d]) for d in range(D)]
    num_new = np.full(B, 2, np.int32)
    num_new[B - 1] = 0  # the last path keeps its waypoints
    status = eng.switch_to_waypoint_path(stop["parameter"], new_wp, num_waypoints=num_new,
                                         initial_velocity=v0)
    for b, pl in enumerate(planners):
        if num_new[b] > 0:
            assert pl.switch(stop["parameter"][b], new_wp[b, :2], initial_velocity=v0[b]) == status[b] == 0
    got = device_splines(eng)
    for b, pl in enumerate(planners):
        k, p = pl.spline()
        assert got[b][0].tobytes() == k.tobytes() and got[b][1].tobytes() == p.tobytes()
    done = np.zeros(B, bool)
    calls = 0
    while not done.all():
        statuses = [pl.plan(start, 750 * MS) for pl in planners]
        eng.planner_plan(start, 750 * MS)
        st = P.compare_call(eng, planners, statuses, f"after switch {calls}")
        done |= (st["target_reached"] == 1) | (st["status"] != 0)
        start += 200 * MS
        calls += 1
        assert calls < 300
    assert all(pl.target_reached for pl in planners)
    P.compare_path_samples(eng, planners)
    for b, pl in enumerate(planners[:-1]):
        np.testing.assert_allclose(pl.trajectory()["r_pos"][-1], new_wp[b, 1], atol=1e-10)
