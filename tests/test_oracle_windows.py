"""PathTimingTrajectory::Plan with its state between calls (moving horizon, splicing,
"planned enough" erase): the CPU restatement oracle/tp_oracle_windows.c against the
properties the reference's integration tests state
(path_timing_trajectory_test.cc:112-260).  CPU only."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as orc  # noqa: E402

MS = 1_000_000
WAYPOINTS = np.array([[1, 2, 3], [-1, -2, -3], [1, 2, 3.0]])  # _test.cc:127-128
VMAX, AMAX = np.full(3, 1.0), np.full(3, 2.0)
N, DELTA, DT = 1000, 0.005, 0.004  # kNumSamples, JointPathOptions default, kTimeStep
HORIZON, REPLAN = 750 * MS, 200 * MS


def make(skip, **kw):
    return orc.Planner(WAYPOINTS, VMAX, AMAX, N, DELTA, DT, skip=skip, **kw)


@pytest.mark.parametrize("skip", [False, True])
def test_rest_to_rest_replanning(skip):
    """RestToRestPlanningWorks (_test.cc:112-173): Plan() repeatedly while shifting the
    start; the executed samples respect the limits and end at the last waypoint at rest."""
    pl = make(skip)
    start, loops, erased, windows = 0, 0, 0, 0
    executed_t = []
    while not pl.target_reached:
        assert pl.plan(start, HORIZON) == orc.ORC_OK
        tr = pl.trajectory()
        assert len(tr["r_t"]) == len(tr["r_pos"]) == len(tr["r_vel"]) == len(tr["r_acc"]) > 0
        assert tr["r_t"][0] == pytest.approx(start / 1e9, abs=1e-9)
        assert np.all(np.diff(tr["r_t"]) > 0)
        assert np.all(np.abs(tr["r_vel"]) <= VMAX + 1e-9)
        assert np.all(np.abs(tr["r_acc"]) <= AMAX + 1e-9)
        erased += pl.erased
        windows += pl.loops
        nxt = min(pl.end_time, start + REPLAN)
        executed_t.append(tr["r_t"][tr["r_t"] < nxt / 1e9])
        start = nxt
        loops += 1
        assert loops < 500
    assert erased > 10 and windows >= 3  # both branches of Plan were exercised
    assert pl.len > N  # windows were spliced
    np.testing.assert_allclose(tr["r_pos"][-1], WAYPOINTS[-1], atol=1e-12)
    np.testing.assert_array_equal(tr["r_vel"][-1], 0.0)
    ps = pl.path_samples()
    assert np.all(np.diff(ps["w_time"]) >= 0) and np.all(np.diff(ps["w_s"]) > 0)
    assert np.all(np.diff(np.concatenate(executed_t)) > 0)


@pytest.mark.parametrize("skip", [False, True])
def test_one_call_with_long_horizon_plans_to_the_end(skip):
    """The loop inside one Plan() call (:628-658): a horizon longer than the motion
    makes it plan window after window until the path ends."""
    pl = make(skip)
    assert pl.plan(0, 3600 * 1000 * MS) == orc.ORC_OK
    assert pl.loops >= 3 and pl.planned_to_end and pl.target_reached
    tr = pl.trajectory()
    np.testing.assert_allclose(tr["r_pos"][-1], WAYPOINTS[-1], atol=1e-12)
    # close to (never faster than 0.1 % of) the one-window solution of the whole path
    whole = orc.plan_joint(WAYPOINTS, VMAX, AMAX, 4000, dt=DT, M_cap=20000)
    assert whole["ok"]
    assert tr["r_t"][-1] == pytest.approx(whole["r_t"][-1], rel=5e-3)
    # one window that covers the whole path is the plain fresh-path plan, bit for bit
    one = orc.Planner(WAYPOINTS, VMAX, AMAX, 4000, 0.0, DT, skip=skip)
    assert one.plan(0, 3600 * 1000 * MS) == orc.ORC_OK and one.loops == 1
    ref = orc.plan_joint(WAYPOINTS, VMAX, AMAX, 4000, dt=DT, M_cap=20000, skip=skip)
    for k in ("r_t", "r_pos", "r_vel", "r_acc", "r_s", "r_sd", "r_sdd"):
        np.testing.assert_array_equal(one.trajectory()[k], ref[k][:ref["M"]])


@pytest.mark.parametrize("skip", [False, True])
def test_no_duplicate_initial_samples(skip):
    """NoDuplicateInitialSamples (_test.cc:175-255): re-planning from a time at / next
    to an existing sample keeps the first two samples apart."""
    pl = make(skip)
    assert pl.plan(0, HORIZON) == orc.ORC_OK
    eps = 0.01 * DT
    for sample in (0, 2, 1, 3):
        t = pl.trajectory()["r_t"]
        start = int(t[sample] * 1e9)
        assert pl.plan(start, HORIZON) == orc.ORC_OK
        t = pl.trajectory()["r_t"]
        assert pl.start_time == start and len(t) >= 5
        assert t[0] == pytest.approx(start / 1e9, abs=1e-9)
        assert t[1] >= t[0] + eps


@pytest.mark.parametrize("skip", [False, True])
def test_invariant_to_starting_time(skip):
    """IsInvariantToStartingTime (_test.cc:257-297)"""
    a, b = make(skip), make(skip)
    assert a.plan(0, HORIZON) == orc.ORC_OK and b.plan(42_000 * MS, HORIZON) == orc.ORC_OK
    pa, pb = a.trajectory()["r_pos"], b.trajectory()["r_pos"]
    assert pa.shape == pb.shape
    np.testing.assert_allclose(pa, pb, atol=1e-10)


def test_time_argument_errors():
    """HandleTimeArguments (:503-538), GetTimeOffsetAfter (:289-305),
    kDeadlineExceeded (:652-655)"""
    pl = make(False)
    assert pl.plan(1000 * MS, HORIZON) == orc.ORC_OK
    end = pl.end_time
    assert pl.plan(end + 5 * MS, HORIZON) == orc.ORC_FAIL_START_RANGE  # > end + time step
    assert pl.plan(end + 2 * MS, HORIZON) == orc.ORC_FAIL_START_ARG    # > end
    assert pl.plan(500 * MS, HORIZON) == orc.ORC_FAIL_START_ARG        # < previous start
    assert pl.plan(1000 * MS, HORIZON) == orc.ORC_OK
    # a horizon that cannot be reached within the allowed number of windows
    few = make(False, max_planning_iterations=1)
    assert few.plan(0, 3600 * 1000 * MS) == orc.ORC_FAIL_DEADLINE
    # not enough room for the spliced samples (not a reference condition)
    small = make(False, P_cap=N + 10)
    assert small.plan(0, 3600 * 1000 * MS) == orc.ORC_FAIL_CAPACITY


def test_initial_velocity_projection():
    """ComputeTimingProfile (:352-385): the initial path velocity is the projection of
    the initial joint velocity onto the first path tangent; a velocity that is not
    parallel to it is rejected."""
    tangent = WAYPOINTS[1] - WAYPOINTS[0]
    v0 = 0.05 * tangent / np.linalg.norm(tangent)
    pl = make(False, initial_velocity=v0)
    assert pl.plan(0, HORIZON) == orc.ORC_OK
    np.testing.assert_allclose(pl.trajectory()["r_vel"][0], v0, atol=1e-9)
    bad = make(False, initial_velocity=np.array([0.05, 0.0, 0.0]))
    assert bad.plan(0, HORIZON) == orc.ORC_FAIL_INITIAL_VELOCITY


# ---- Cartesian paths behind the same planner (path_timing_trajectory_test.cc:613-757)
def fake_fk_pose(q):
    """FakeFK (_test.cc:548-552): translation q[0:3], rotation ExpSO3(q[3:6])."""
    out = np.zeros(7)
    out[:3] = q[:3]
    quat = np.zeros(4)
    v = np.ascontiguousarray(q[3:6], np.float64)
    orc.lib().orc_exp_so3(v.ctypes.data_as(orc.c_double_p), quat.ctypes.data_as(orc.c_double_p))
    out[3:] = quat
    return out


CART_WAYPOINTS = np.array([[1.0, 0, 0, 0, 0, 0, 0], [0, 1.0, 0, 0, 0, 0, 0], [0, 0, 0, 1.0, 0, 0, 0],
                           [0, 0, 0, 0, 1.0, 0, 0], [0, 0, 0, 0, 0, 0, 1.0]])  # _test.cc:711-716


def make_cart(joint_wp, skip=False, N=1000, delta=0.005, **kw):
    joint_wp = np.asarray(joint_wp, np.float64)
    poses = np.stack([fake_fk_pose(q) for q in joint_wp])
    return orc.Planner(joint_wp, np.full(7, 1.0), np.full(7, 2.0), N, delta, DT, skip=skip,
                       pose_waypoints=poses, max_initial_velocity_error=1e-2, **kw)


@pytest.mark.parametrize("skip", [False, True])
def test_cartesian_path_replanning(skip):
    """The 5-waypoint Cartesian path of _test.cc:695-757, Plan() every 200 ms with a
    750 ms horizon until the trajectory is at its end: several windows are spliced, the
    limits hold and the motion ends at the last joint waypoint at rest."""
    pl = make_cart(CART_WAYPOINTS, skip)
    start, loops, windows = 0, 0, 0
    while not pl.target_reached:
        assert pl.plan(start, HORIZON) == orc.ORC_OK
        tr = pl.trajectory()
        assert len(tr["r_t"]) > 0 and np.all(np.diff(tr["r_t"]) > 0)
        assert np.all(np.abs(tr["r_vel"]) <= 1.0 + 1e-9) and np.all(np.abs(tr["r_acc"]) <= 2.0 + 1e-9)
        windows += pl.loops
        start = min(pl.end_time, start + REPLAN)
        loops += 1
        assert loops < 2000
    assert windows >= 3 and pl.len > N
    np.testing.assert_allclose(tr["r_pos"][-1], CART_WAYPOINTS[-1], atol=1e-9)
    np.testing.assert_array_equal(tr["r_vel"][-1], 0.0)
    ps = pl.path_samples()
    assert np.all(np.diff(ps["w_s"]) > 0)
    # Cartesian speed limits (0.5 m/s, 0.25 rad/s through the fake Jacobian's linear part)
    assert np.all(np.linalg.norm(ps["w_vel"][:, :3], axis=1) <= 0.5 + 1e-6)


@pytest.mark.parametrize("skip", [False, True])
def test_zero_length_cartesian_path(skip):
    """SimpleCartesianPath (_test.cc:613-693): two identical waypoints give one sample at
    the waypoint, at rest."""
    wp = np.array([[1.0, 0, 0, 0, 0, 0, 0], [1.0, 0, 0, 0, 0, 0, 0]])
    pl = make_cart(wp, skip)
    start, loops = 0, 0
    while not pl.target_reached:
        assert pl.plan(start, HORIZON) == orc.ORC_OK
        start = min(pl.end_time, start + REPLAN)
        loops += 1
        assert loops < 50
    tr = pl.trajectory()
    assert len(tr["r_pos"]) == 1 and len(tr["r_vel"]) == 1
    np.testing.assert_allclose(tr["r_pos"][0], wp[0], atol=1e-12)
    np.testing.assert_array_equal(tr["r_vel"][0], 0.0)


def test_cartesian_single_window_equals_the_fresh_path_plan():
    """A window that covers the whole path is the one-window plan of tp_oracle_cart.c
    (pinned to the reference's golden vectors in tests/test_oracle_cart.py)."""
    joint = CART_WAYPOINTS
    poses = np.stack([fake_fk_pose(q) for q in joint])
    N = 400
    one = orc.plan_cart_fake(poses, joint, np.full(7, 1.0), np.full(7, 2.0), N, dt=DT, M_cap=20000)
    assert one["ok"]
    pl = make_cart(joint, N=N, delta=-1.0)
    assert pl.plan(0, 60_000 * MS) == orc.ORC_OK
    assert pl.loops == 1 and pl.len == N and pl.target_reached
    ps, tr = pl.path_samples(), pl.trajectory()
    assert ps["w_time"].tobytes() == one["time"].tobytes() and ps["w_q"].tobytes() == one["q"].tobytes()
    assert tr["r_pos"].tobytes() == one["r_pos"].tobytes() and tr["r_t"].tobytes() == one["r_t"].tobytes()
