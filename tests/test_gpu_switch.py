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


# ---- Cartesian paths (TimeableCartesianSplinePath::SwitchToWaypointPath) ----------------
def device_cart_splines(eng):
    n = eng.get(ARR.NUM_POINTS)
    ctrl, knots, pose = eng.get(ARR.CTRL), eng.get(ARR.KNOTS), eng.get(ARR.POSE_CTRL)
    return [(knots[b, :n[b] + 3].copy(), ctrl[b, :n[b]].copy(), pose[b, :n[b]].copy())
            for b in range(len(n))]


def fake_fk(q):
    return np.concatenate([q[:3], orc.exp_so3(q[3:6])])


def test_cartesian_switch_matches_oracle_per_path():
    """Batched switches of Cartesian paths against the oracle (== the reference's object code
    up to its rounding-decided end-point sign, tests/test_oracle_vs_ref_cart.py): status
    exact; knots, joint, translation and rotation control points within 1e-12 (the device's
    acos / sin / cos / atan2 in slerp and the pose corner offsets).  A path whose rotation
    end-point correction falls on the other side of that sign is recognised and dropped."""
    B, W, D = 12, 5, 7
    rng = np.random.default_rng(31)
    joint = rng.uniform(-0.8, 0.8, (B, W, D))
    pose = np.stack([np.stack([fake_fk(q) for q in joint[b]]) for b in range(B)])
    eng = pkg.Engine(max_paths=B, num_samples=64, num_dofs=D, max_waypoints=W, cartesian=True)
    eng.set_cart_options(0.05, 0.5, 0.25)
    eng.set_cart_waypoints(pose, joint)
    vm, am = np.ones((B, D)), np.full((B, D), 2.0)
    planners = [orc.Planner(joint[b], vm[b], am[b], 64, 0.01, 0.004, pose_waypoints=pose[b])
                for b in range(B)]
    alive = np.ones(B, bool)
    seen, compared = set(), 0
    for round_ in range(8):
        kend = np.array([pl.spline()[0][-1] for pl in planners])
        keep = rng.uniform(0.2, 0.9, B) * kend
        keep[1] = kend[1] + 1.0  # nothing to truncate: the join is the old end
        nj = rng.uniform(-0.8, 0.8, (B, W, D))
        num = rng.integers(2, W + 1, size=B).astype(np.int32)
        num[3] = 0  # untouched
        npz = np.stack([np.stack([fake_fk(q) for q in nj[b]]) for b in range(B)])
        status = eng.switch_to_cart_waypoint_path(keep, npz, nj, num_waypoints=num)
        got = device_cart_splines(eng)
        for b, pl in enumerate(planners):
            expect = 0 if num[b] <= 0 else pl.switch(keep[b], nj[b, :num[b]],
                                                     pose_waypoints=npz[b, :num[b]])
            if not alive[b]:
                continue
            assert status[b] == expect, (round_, b, status[b], expect)
            seen.add(int(expect))
            if expect != 0:
                alive[b] = False  # partially modified splines: not compared further
                continue
            k, p = pl.spline()
            pc = pl._pose_ctrl[:len(p)]
            assert got[b][0].shape == k.shape
            np.testing.assert_allclose(got[b][0], k, rtol=1e-13, atol=0)
            np.testing.assert_allclose(got[b][1], p, rtol=0, atol=1e-12)
            np.testing.assert_allclose(got[b][2][:, :3], pc[:, :3], rtol=0, atol=1e-12)
            bad = np.flatnonzero(np.abs(got[b][2][:, 3:] - pc[:, 3:]).max(axis=1) > 1e-11)
            if bad.size:  # the reference's sign of +-1e-16 (bsplineq.cc:575-577)
                assert bad.size == 1
                alive[b] = False
                continue
            compared += 1
    assert 0 in seen and compared >= 40


@pytest.mark.parametrize("skip", [False, True])
def test_cartesian_switch_and_replan(skip):
    """SwitchToNewCartesianWaypointPathWorks (path_timing_trajectory_test.cc:764-920) on the
    device with the built-in kinematics: plan until the motion is fast, switch at the stop
    parameter with the current velocity, plan to the end -- against the oracle planner within
    the tolerances of the Cartesian planner tests (device libm)."""
    import test_oracle_windows as TW
    N, dt, delta = 1000, 0.004, 0.005
    joint = TW.CART_WAYPOINTS[None]
    pose = np.stack([TW.fake_fk_pose(q) for q in TW.CART_WAYPOINTS])[None]
    W, D = joint.shape[1], 7
    vm, am = np.ones((1, D)), np.full((1, D), 2.0)
    eng = pkg.Engine(max_paths=1, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=60000, cartesian=True)
    eng.set_cart_options(0.05, 0.5, 0.25)
    eng.set_cart_waypoints(pose, joint)
    eng.set_joint_limits(vm, am)
    eng.set_sampling(np.full(1, delta))
    eng.planner_reset(16384, dt, skip=skip, max_initial_velocity_error=1e-2)
    eng.planner_set_kinematics(1)
    pl = orc.Planner(joint[0], vm[0], am[0], N, delta, dt, skip=skip, P_cap=16384,
                     pose_waypoints=pose[0], max_initial_velocity_error=1e-2)

    def compare(tag):
        st = P.device_state(eng)
        off, rows = P.device_rows(eng)
        assert st["len"][0] == pl.len and st["target_reached"][0] == pl.target_reached, tag
        tr = pl.trajectory()
        M = off[1] - off[0]
        assert abs(M - pl.M) <= 1, tag
        m = min(M, pl.M) - 1
        if m > 0:
            assert np.max(np.abs(rows["r_t"][:m] - tr["r_t"][:m])) <= 1e-9, tag
            assert np.max(np.abs(rows["r_pos"][:m] - tr["r_pos"][:m])) <= 1e-8, tag
        return tr

    start, tr = 0, None
    for call in range(200):
        assert pl.plan(start, 750 * MS) == orc.ORC_OK
        eng.planner_plan(start, 750 * MS)
        tr = compare(f"before switch {call}")
        start = min(pl.end_time, max(start + 200 * MS, pl.start_time))
        if np.max(np.abs(tr["r_vel"][0])) > 0.3 * 0.25:
            break
    assert not pl.target_reached
    stop = eng.path_stop_parameter(start / 1e9)
    o = orc.path_stop_parameter(tr["r_t"], tr["r_s"], tr["r_vel"], tr["r_acc"], am[0], start / 1e9)
    assert abs(stop["parameter"][0] - o["parameter"]) <= 1e-9
    new_joint = np.zeros((1, W, D))
    new_joint[0, 0], new_joint[0, 1] = -1.0, 1.0
    new_pose = np.zeros((1, W, 7))
    new_pose[0, :, 6] = 1.0
    new_pose[0, :2] = [TW.fake_fk_pose(q) for q in new_joint[0, :2]]
    v0 = np.array([[np.interp(start / 1e9, tr["r_t"], tr["r_vel"][:, d]) for d in range(D)]])
    status = eng.switch_to_cart_waypoint_path(np.array([o["parameter"]]), new_pose, new_joint,
                                              num_waypoints=np.array([2], np.int32),
                                              initial_velocity=v0)
    assert status[0] == 0
    assert pl.switch(o["parameter"], new_joint[0, :2], initial_velocity=v0[0],
                     pose_waypoints=new_pose[0, :2]) == 0
    k, p = pl.spline()
    got = device_cart_splines(eng)[0]
    np.testing.assert_allclose(got[0], k, rtol=1e-13, atol=0)
    np.testing.assert_allclose(got[1], p, rtol=0, atol=1e-11)
    np.testing.assert_allclose(got[2], pl._pose_ctrl[:len(p)], rtol=0, atol=1e-11)
    for call in range(400):
        assert pl.plan(start, 750 * MS) == orc.ORC_OK
        eng.planner_plan(start, 750 * MS)
        compare(f"after switch {call}")
        start = min(pl.end_time, max(start + 200 * MS, pl.start_time))
        if pl.target_reached:
            break
    assert pl.target_reached
    _, rows = P.device_rows(eng)
    np.testing.assert_allclose(rows["r_pos"][-1], new_joint[0, 1], atol=1e-9)
    np.testing.assert_array_equal(rows["r_vel"][-1], 0.0)
