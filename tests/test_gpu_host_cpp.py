"""GPU test: the host C++ mirror of the reference API (host/include/...) driven by
tests/host_cpp/host_api_test.cc, compared with the CPU oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "x-edr-trajectory-planning_b200", "host", "bin", "host_api_test")


@pytest.fixture(scope="module")
def result():
    if not os.path.exists(BIN):
        subprocess.run(["make", "-C", os.path.dirname(os.path.dirname(BIN)), "all"], check=True)
    p = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    r = json.loads(p.stdout)
    assert r["failures"] == 0, p.stderr[-2000:]
    return r


def test_rest_to_rest_plan_matches_oracle(result, po):
    r = result["rest_to_rest"]
    rows10 = result["constraints_sample10"]
    wp = np.array([[1, 2, 3], [-1, -2, -3], [1, 2, 3]], float)
    vmax, amax = np.full(3, 1.0), np.full(3, 2.0)
    o = po.plan_joint(wp, vmax, amax, 1000, delta=0.016, dt=0.004, M_cap=20000)
    assert o["ok"]
    ctrl, knots = po.joint_fit(wp, 0.2)
    np.testing.assert_array_equal(np.array(r["knots"]), knots)
    np.testing.assert_array_equal(np.array(r["control_points"]).reshape(-1, 3), ctrl)
    k = r["trimmed"]  # the second Plan() erased the first 50 samples
    np.testing.assert_array_equal(np.array(r["time"]), o["r_t"][k:])
    assert np.max(np.abs(np.array(r["positions"]).reshape(-1, 3) - o["r_pos"][k:])) <= 1e-9
    assert np.max(np.abs(np.array(r["velocities"]).reshape(-1, 3) - o["r_vel"][k:])) <= 1e-9
    assert np.max(np.abs(np.array(r["accelerations"]).reshape(-1, 3) - o["r_acc"][k:])) <= 1e-9
    np.testing.assert_allclose(np.array(r["path_parameter"]), o["r_s"][k:], atol=1e-12)
    # end_time_ = TimeFromSec(time_.back()) clamped to a time-step multiple
    # (path_timing_trajectory.cc:662-666, 229-233)
    end_ns = po.time_from_sec(round(po.time_to_sec(po.time_from_sec(o["r_t"][-1])) / 0.004) * 0.004)
    assert r["end_time_ns"] == end_ns
    cons = po.joint_constraints(o["dq"], o["ddq"], vmax, amax, 0.8)
    np.testing.assert_array_equal(np.array(rows10).reshape(6, 4), cons[10])


def test_circle_profile_matches_oracle(result, po, pkg):
    for c in result["circle"]:
        cons = pkg.workloads.circle_constraints(c["N"])
        o = po.solve(cons, 0.0, np.pi, c["sd_start"])
        assert o["ok"]
        assert c["duration"] == o["time"][-1] - o["time"][0]
        assert c["last_extremal"] == o["last_extremal_index"]
        lp = po.find_max_sd2(cons[c["N"] // 3])
        assert c["lp_sd2max"] == lp[0] and c["lp_sd2zero"] == lp[2]
        assert c["dt_max"] == pytest.approx(np.max(np.diff(o["time"])), abs=1e-15)


def test_bspline_matches_golden_and_oracle(result, po):
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "bspline_mathematica.json")))
    v = np.array(result["bspline"]["values"]).reshape(101, 4, 2)
    for k, (nx, ny) in enumerate((("X", "Y"), ("Xp", "Yp"), ("Xpp", "Ypp"), ("Xppp", "Yppp"))):
        assert np.sum((v[:, k, 0] - g[nx]) ** 2) <= 5e-14
        assert np.sum((v[:, k, 1] - g[ny]) ** 2) <= 5e-14
    us = 0.0 + np.arange(101) * (1.0 / 100)
    _, ref = po.eval_curve_ders(np.array(g["knots"], float), np.array(g["control_points"], float),
                                3, 3, us)
    np.testing.assert_array_equal(v, ref)
    np.testing.assert_array_equal(np.array(result["bspline"]["at_0p01"]).reshape(4, 2), ref[1])


def test_batch_front_end_matches_oracle(result, po):
    b = result["batch"]
    wp = np.array(b["waypoints"]).reshape(5, 4, 3)
    nw = [4, 3, 2, 4, 1]
    off = np.array(b["offset"], np.int64)
    pos = np.array(b["positions"]).reshape(-1, 3)
    for i in range(5):
        o = po.plan_joint(wp[i, :nw[i]], np.full(3, 1.0), np.full(3, 2.0), 256, dt=0.004,
                          M_cap=20000)
        assert (b["status"][i] == 0) == o["ok"]
        if not o["ok"]:
            continue
        assert abs(b["duration"][i] - o["duration"]) <= 1e-6 * max(o["duration"], 1e-300)
        assert off[i + 1] - off[i] == o["M"]
        assert np.max(np.abs(pos[off[i]:off[i + 1]] - o["r_pos"])) <= 1e-9


def test_cartesian_path_matches_oracle(result, po):
    """TimeableCartesianSplinePath + PathTimingTrajectory::Plan with the reference
    test's fake kinematics: as host PathIKFunc / JacobianFunc callbacks and as the
    built-in device model, against the oracle's whole-path restatement."""
    c = result["cartesian"]
    jw = np.array(c["joint_waypoints"]).reshape(3, 7)
    pw = np.array(c["pose_waypoints"]).reshape(3, 7)
    vmax, amax = np.full(7, 1.0), np.full(7, 2.0)
    o = po.plan_cart_fake(pw, jw, vmax, amax, 1000, translation_rounding=0.05,
                          rotation_rounding=0.1, delta=0.02, dt=0.004, M_cap=40000)
    assert o["ok"]
    for mode in ("host_callbacks", "device_kinematics"):
        r = c[mode]
        assert np.max(np.abs(np.array(r["knots"]) - o["knots"])) <= 1e-12 * o["knots"][-1]
        t = np.array(r["time"])
        pos = np.array(r["positions"]).reshape(-1, 7)
        vel = np.array(r["velocities"]).reshape(-1, 7)
        assert abs(len(t) - o["M"]) <= 1
        assert abs(t[-1] - o["r_t"][-1]) <= 1e-6 * o["r_t"][-1] + 0.004
        m = min(len(t), o["M"]) - 1
        assert np.max(np.abs(pos[:m] - o["r_pos"][:m])) <= 1e-9
        assert np.max(np.abs(vel[:m] - o["r_vel"][:m])) <= 1e-6
        # ends at the last waypoint (fake IK of the last pose), at rest
        assert np.max(np.abs(pos[-1] - jw[-1])) <= 1e-9
        assert (vel[-1] == 0).all()
        assert np.max(np.abs(np.array(r["q_last"]) - jw[-1])) <= 1e-12
        rows = np.array(r["constraints_sample10"]).reshape(16, 4)
        assert np.max(np.abs(rows - o["cons"][10])) <= 1e-8
        assert (rows[:7, 3] == 1.6).all() and (rows[14, 3] == 0.25) and (rows[15, 3] == 0.0625)


def test_rescale_to_stop_matches_oracle(result, po):
    """host RescaleTrajectoryBackwardToStop on rescale_to_stop_test.cc's scenario."""
    import test_oracle_stop as T
    assert result["stop"]["cases"] == 2
    for k, velocity in enumerate((-1.0, 1.0)):
        t, pos, vel, acc = T.constant_velocity_case(velocity)
        o = po.rescale_to_stop(np.full(4, 2.0), t, pos, vel, acc)
        c = result["stop"][f"case{k}"]
        assert np.array_equal(np.array(c["times"]), o["times"])
        assert np.array_equal(np.array(c["velocities"]).reshape(-1, 4), o["velocities"])
        assert np.array_equal(np.array(c["accelerations"]).reshape(-1, 4), o["accelerations"])


def test_path_stop_parameter_matches_oracle(result, po):
    """host GetPathStopParameter on path_timing_trajectory_test.cc:942-992's scenario."""
    wp = np.array([[0.0, 0, 0], [1, 1, 1], [2, 2, 2]])
    amax = np.full(3, 2.0)
    o = po.plan_joint(wp, np.ones(3), amax, 500, safety=1.0, delta=0.008, dt=0.001, M_cap=20000)
    assert o["ok"]
    args = (o["r_t"], o["r_s"], o["r_vel"], o["r_acc"], amax)
    q = result["stop_query"]
    assert q["first"] == po.path_stop_parameter(*args, 0.0)["parameter"]
    assert q["last"] == o["r_s"][-1]
    # TimeFromSec quantises the query time to nanoseconds
    t = po.time_to_sec(po.time_from_sec(q["mid_time"]))
    assert q["mid"] == po.path_stop_parameter(*args, t)["parameter"]


def test_skip_method_plan_matches_oracle(result, po):
    """host Plan() with kSkipSamplesCloserThanTimeStep vs the oracle's restatement."""
    wp = np.array([[1.0, 2, 3], [-1, -2, -3], [1, 2, 3]])
    o = po.plan_joint(wp, np.ones(3), np.full(3, 2.0), 1000, delta=0.016, dt=0.004, M_cap=1001,
                      skip=True)
    assert o["ok"]
    r = result["rest_to_rest_skip"]
    assert np.array_equal(np.array(r["time"]), o["r_t"])
    assert np.array_equal(np.array(r["positions"]).reshape(-1, 3), o["r_pos"])
    assert np.array_equal(np.array(r["velocities"]).reshape(-1, 3), o["r_vel"])


@pytest.mark.parametrize("key,skip", [("replanning", False), ("replanning_skip", True)])
def test_moving_horizon_replanning_matches_oracle(result, po, key, skip):
    """host Plan() called again every 200 ms with a 750 ms horizon over several planning
    windows (RestToRestPlanningWorks, path_timing_trajectory_test.cc:112-173) against the
    oracle's planner, every call's sample count / first time / end time and the final
    trajectory bit for bit."""
    wp = np.array([[1.0, 2, 3], [-1, -2, -3], [1, 2, 3]])
    r = result[key]
    pl = po.Planner(wp, np.ones(3), np.full(3, 2.0), 1000, 0.005, 0.004, skip=skip,
                    max_initial_velocity_error=1e-2)
    start, counts, firsts, ends, windows = 0, [], [], [], 0
    while not pl.target_reached:
        assert pl.plan(start, 750_000_000) == po.ORC_OK
        counts.append(pl.M)
        firsts.append(pl.trajectory()["r_t"][0])
        ends.append(pl.end_time)
        windows += pl.loops
        start = min(pl.end_time, start + 200_000_000)
    assert r["loops"] == len(counts) and r["windows"] == windows
    assert [int(c) for c in r["counts"]] == counts
    assert np.array_equal(np.array(r["first_times"]), np.array(firsts))
    assert [int(x) for x in r["end_times_ns"]] == ends
    tr = pl.trajectory()
    assert np.array_equal(np.array(r["time"]), tr["r_t"])
    assert np.array_equal(np.array(r["positions"]).reshape(-1, 3), tr["r_pos"])
    assert np.array_equal(np.array(r["velocities"]).reshape(-1, 3), tr["r_vel"])
    assert np.array_equal(np.array(r["accelerations"]).reshape(-1, 3), tr["r_acc"])
    assert np.array_equal(np.array(r["path_parameter"]), tr["r_s"])
    assert r["final_decel_start_ns"] == pl.final_decel_start


@pytest.mark.parametrize("key,skip", [("reset_and_replan", False), ("reset_and_replan_skip", True)])
def test_reset_and_replan_matches_oracle(result, po, key, skip):
    """ResetAndReplan (path_timing_trajectory_test.cc:440-563): the host classes plan to the
    end into a TrajectoryBuffer, Reset(), and plan again -- the C++ side checks both buffers
    are equal; here the second one is compared bit for bit with the oracle's planner and
    the oracle's TrajectoryBuffer driven the same way."""
    r = result[key]
    assert r["same"] == 1
    wp = np.array([[1.0, 2, 3], [-1, -2, -3], [0.5, 1.0, 1.5]])
    pl = po.Planner(wp, np.ones(3), np.full(3, 2.0), 1000, 0.005, 0.004, skip=skip,
                    max_initial_velocity_error=1e-2)
    buf = po.TrajectoryBuffer(3)
    start, loops = 0, 0
    while not pl.target_reached:
        assert pl.plan(start, 750_000_000) == po.ORC_OK
        tr = pl.trajectory()
        assert buf.insert_segment(tr["r_t"], tr["r_pos"], tr["r_vel"], tr["r_acc"]) == 0
        # GetNextPlanStartTime (path_timing_trajectory.cc:870-872)
        start = min(pl.end_time, max(start + 200_000_000, pl.start_time))
        loops += 1
    assert r["loops"] == loops and r["rows"] == buf.num_samples
    t, p, _, _ = buf.rows()
    assert np.array_equal(np.array(r["time"]), t)
    assert np.array_equal(np.array(r["positions"]).reshape(-1, 3), p)


def test_switch_to_waypoint_path_matches_oracle(result, po):
    """host SwitchToWaypointPath + re-planning of the modified path
    (SwitchToNewJointWaypointPathWorks, path_timing_trajectory_test.cc:299-436) against the
    oracle's planner driven the same way."""
    r = result["switching"]
    wp = np.array([[1.0, 2, 3], [-1, -2, -3], [0.5, 1.0, 1.5]])
    vmax, amax = np.ones(3), np.full(3, 2.0)
    pl = po.Planner(wp, vmax, amax, 1000, 0.001, 0.004, max_initial_velocity_error=1e-2)
    start, first_calls = 0, 0
    for first_calls in range(100):
        assert pl.plan(start, 750_000_000) == po.ORC_OK
        tr = pl.trajectory()
        start = min(pl.end_time, max(start + 200_000_000, pl.start_time))
        if np.min(np.abs(np.abs(tr["r_vel"][0]) - vmax)) < 0.3:
            break
    assert r["first_calls"] == first_calls
    sp = po.path_stop_parameter(tr["r_t"], tr["r_s"], tr["r_vel"], tr["r_acc"], amax, start / 1e9)
    assert r["stop_parameter"] == sp["parameter"]
    v0 = np.array(r["initial_velocity"])
    ts = start / 1e9
    hi = 1
    while hi + 1 < len(tr["r_t"]) and tr["r_t"][hi] < ts:
        hi += 1
    w = (ts - tr["r_t"][hi - 1]) / (tr["r_t"][hi] - tr["r_t"][hi - 1])
    assert np.array_equal(v0, tr["r_vel"][hi - 1] + w * (tr["r_vel"][hi] - tr["r_vel"][hi - 1]))
    assert pl.switch(sp["parameter"], np.array([[1.0, 2.0, 3.0], [0.5, 1.0, 5.5]]),
                     initial_velocity=v0) == 0
    assert np.array_equal(np.array(r["knots_after_switch"]), pl.spline()[0])
    second_calls = 0
    while not pl.target_reached:
        assert pl.plan(start, 750_000_000) == po.ORC_OK
        start = min(pl.end_time, max(start + 200_000_000, pl.start_time))
        second_calls += 1
    assert r["second_calls"] == second_calls
    tr = pl.trajectory()
    assert np.array_equal(np.array(r["time"]), tr["r_t"])
    assert np.array_equal(np.array(r["positions"]).reshape(-1, 3), tr["r_pos"])
    assert np.array_equal(np.array(r["velocities"]).reshape(-1, 3), tr["r_vel"])
    assert np.array_equal(np.array(r["accelerations"]).reshape(-1, 3), tr["r_acc"])


def test_bspline_knot_insertion_truncate_extend_match_oracle(result, po):
    """host BSplineXd::InsertKnotAndUpdateControlPointsRef / TruncateSplineAt /
    ExtendWithControlPoints (splines/bspline.h:224-511) against the oracle's restatement,
    and the curve before the truncation point is what it was."""
    r = result["bspline_edit"]
    wp = np.array([[1.0, 2, 3], [-1, -2, -3], [0.5, 1.0, 1.5]])
    ctrl, knots = po.joint_fit(wp, 0.2)
    sp = po.Spline(knots, ctrl, knot_capacity=100)
    kend = knots[-1]
    assert sp.insert_knot(0.3 * kend, 2) == 0
    assert sp.truncate(0.6 * kend) == 0
    extra = np.array([[0.1, 0.2, 0.3], [0.4, 0.1, 0.0], [0.7, 0.0, -0.2], [1.0, 0.0, -0.3]])
    assert sp.extend(extra) == 0
    assert np.array_equal(np.array(r["knots"]), sp.knots)
    assert np.array_equal(np.array(r["control_points"]).reshape(-1, 3), sp.points)
    np.testing.assert_allclose(np.array(r["after"]), np.array(r["before"]), atol=1e-12)


def test_trajectory_buffer_matches_oracle(result, po):
    """host TrajectoryBuffer on trajectory_buffer_test.cc's scenarios (the EXPECTs run in
    the exerciser); the rows it prints against the oracle, bit for bit."""
    import tbuf_cases as TC
    r = result["tbuf"]
    assert r["done"] == 1
    tr = TC.test_trajectory(1.0, 10)
    b = po.TrajectoryBuffer(TC.K_NUM_JOINTS)
    assert b.insert_segment(*tr) == 0
    b.discard_before_sec(0.5 * (tr[0][5] + tr[0][6]))
    t, p, _, _ = b.rows()
    assert np.array_equal(np.array(r["discard_between_times"]), t)
    assert np.array_equal(np.array(r["discard_between_first_position"]), p[0])
    quad = TC.quadratic_test_trajectory(1001, 1e-3)
    s = po.TrajectoryBuffer(TC.K_NUM_JOINTS)
    assert s.insert_segment(*quad) == 0
    assert s.stop_at_index(500, np.full(TC.K_NUM_JOINTS, 5.0), 1e-3) == 0
    t, _, v, _ = s.rows()
    assert r["stop_num_samples"] == s.num_samples
    assert np.array_equal(np.array(r["stop_times"]), t)
    assert np.array_equal(np.array(r["stop_velocities"]).reshape(-1, TC.K_NUM_JOINTS), v)
    h = po.TrajectoryBuffer(TC.K_NUM_JOINTS)
    assert h.insert_segment(*quad) == 0
    assert h.stop_before_time(0.5 * (quad[0][500] + quad[0][501]), np.full(TC.K_NUM_JOINTS, 5.0), 1e-3) == 0
    assert r["stop_halfway_num_samples"] == h.num_samples


@pytest.mark.parametrize("key", ["replanning", "replanning_host_callbacks"])
def test_cartesian_replanning_matches_oracle(result, po, key):
    """The reference's 5-waypoint Cartesian re-planning scenario
    (path_timing_trajectory_test.cc:695-757) through the host classes, against the oracle
    planner's loop (== the reference's object code, tests/test_oracle_vs_ref_cart.py):
    with device kinematics (all windows on the device) and with the user's PathIKFunc /
    JacobianFunc as host callbacks driven window by window from tpb_planner_plan (the IK
    cache grows by the new indices only, timeable_path_cartesian_spline.cc:464-549)."""
    import test_oracle_windows as TW
    r = result["cartesian"][key]
    pl = TW.make_cart(TW.CART_WAYPOINTS)
    start, loops, windows, end_times = 0, 0, 0, []
    while not pl.target_reached:
        assert pl.plan(start, TW.HORIZON) == po.ORC_OK
        windows += pl.loops
        end_times.append(pl.end_time / 1e9)
        start = min(pl.end_time, start + TW.REPLAN)
        loops += 1
    assert r["loops"] == loops and r["windows"] == windows
    assert np.max(np.abs(np.array(r["end_times"]) - np.array(end_times))) <= 0.004 + 1e-9
    tr = pl.trajectory()
    t = np.array(r["time"])
    pos = np.array(r["positions"]).reshape(-1, 7)
    vel = np.array(r["velocities"]).reshape(-1, 7)
    assert abs(len(t) - len(tr["r_t"])) <= 1
    m = min(len(t), len(tr["r_t"])) - 1
    assert np.max(np.abs(pos[:m] - tr["r_pos"][:m])) <= 1e-9
    assert np.max(np.abs(vel[:m] - tr["r_vel"][:m])) <= 1e-6
    assert np.max(np.abs(pos[-1] - TW.CART_WAYPOINTS[-1])) <= 1e-9 and (vel[-1] == 0).all()
    if key == "replanning_host_callbacks":
        assert r["ik_calls"] == windows
        assert 1001 < r["ik_samples"] < 1001 * windows
        if po.have_ref():
            # every IK index is solved once (+ the re-evaluated first sample of each call):
            # the reference's own path_ik_positions_ cache after the same loop
            import test_oracle_vs_ref_cart as TC
            _, ref = TC.cart_pair(TW.CART_WAYPOINTS)
            start = 0
            while not ref.state()["target_reached"]:
                assert ref.plan(start, TW.HORIZON) == 0
                start = min(ref.state()["end_time"], start + TW.REPLAN)
            assert r["ik_samples"] == len(ref.ik_positions()) + windows


@pytest.mark.parametrize("key", ["switch", "switch_host_callbacks"])
def test_cartesian_switch_matches_oracle(result, po, key):
    """SwitchToNewCartesianWaypointPathWorks (path_timing_trajectory_test.cc:764-920) through
    the host classes -- device kinematics and the user's callbacks -- against the oracle
    planner driven through the same sequence (== the reference's object code,
    tests/test_oracle_vs_ref_cart.py): stop parameter, knots after the switch, number of
    Plan() calls on both sides of the switch, the final trajectory."""
    import test_oracle_windows as TW
    r = result["cartesian"][key]
    pl = TW.make_cart(TW.CART_WAYPOINTS)
    start, tr = 0, None
    for first_calls in range(200):
        assert pl.plan(start, TW.HORIZON) == po.ORC_OK
        tr = pl.trajectory()
        start = min(pl.end_time, max(start + TW.REPLAN, pl.start_time))
        if np.max(np.abs(tr["r_vel"][0])) > 0.3 * 0.25:
            break
    assert r["first_calls"] == first_calls
    sp = po.path_stop_parameter(tr["r_t"], tr["r_s"], tr["r_vel"], tr["r_acc"], np.full(7, 2.0), start / 1e9)
    assert sp["status"] == 0 and abs(r["stop_parameter"] - sp["parameter"]) <= 1e-9
    v0 = np.array([np.interp(start / 1e9, tr["r_t"], tr["r_vel"][:, d]) for d in range(7)])
    assert np.max(np.abs(np.array(r["initial_velocity"]) - v0)) <= 1e-7
    new_joint = np.array([[-1.0] * 7, [1.0] * 7])
    new_pose = np.stack([TW.fake_fk_pose(q) for q in new_joint])
    assert pl.switch(sp["parameter"], new_joint, initial_velocity=v0, pose_waypoints=new_pose) == 0
    k, _ = pl.spline()
    assert np.max(np.abs(np.array(r["knots_after_switch"]) - k)) <= 1e-9 * k[-1]
    for second_calls in range(1, 401):
        assert pl.plan(start, TW.HORIZON) == po.ORC_OK
        start = min(pl.end_time, max(start + TW.REPLAN, pl.start_time))
        if pl.target_reached:
            break
    assert r["second_calls"] == second_calls
    tr = pl.trajectory()
    t = np.array(r["time"])
    pos = np.array(r["positions"]).reshape(-1, 7)
    assert abs(len(t) - len(tr["r_t"])) <= 1
    m = min(len(t), len(tr["r_t"])) - 1
    assert np.max(np.abs(pos[:m] - tr["r_pos"][:m])) <= 1e-7
    assert np.max(np.abs(pos[-1] - new_joint[-1])) <= 1e-8
