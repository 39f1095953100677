"""Online re-planning geometry (oracle/tp_oracle_switch.c): knot insertion, TruncateSplineAt,
ExtendWithControlPoints, ProjectPointOnPath, SwitchToWaypointPath and the re-timing of the
modified path, against the properties the reference's tests state (splines/bspline_test.cc
knot insertion / truncation / extension cases, path_tools_test.cc:41-115,
path_timing_trajectory_test.cc:299-436).  CPU only."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as orc  # noqa: E402

MS = 1_000_000


def random_spline(rng, W=5, D=4):
    wp = rng.uniform(-2, 2, size=(W, D))
    ctrl, knots = orc.joint_fit(wp, 0.2)
    return orc.Spline(knots, ctrl), wp


def curve(sp, us):
    st, v = sp.eval(us, 1)
    assert np.all(st == 0)
    return v


def test_knot_insertion_keeps_the_curve():
    rng = np.random.default_rng(3)
    for _ in range(10):
        sp, _ = random_spline(rng)
        us = np.linspace(0, sp.knots[-1], 200)
        before = curve(sp, us)
        n0 = len(sp.points)
        u = rng.uniform(0.05, 0.95) * sp.knots[-1]
        assert sp.insert_knot(u, 2) == 0
        assert len(sp.points) == n0 + 2 and len(sp.knots) == n0 + 2 + 3
        assert np.all(np.diff(sp.knots) >= 0) and np.sum(sp.knots == u) == 2
        np.testing.assert_allclose(curve(sp, us), before, atol=1e-12)
    # CanInsertKnot (bspline_base.cc:166-195)
    sp, _ = random_spline(rng)
    assert sp.insert_knot(0.0) == 3 and sp.insert_knot(sp.knots[-1]) == 3
    assert sp.insert_knot(0.5 * sp.knots[-1], 4) == 3
    small = orc.Spline(sp.knots, sp.points, knot_capacity=len(sp.knots) + 1)
    assert small.insert_knot(0.5 * sp.knots[-1], 2) == 9
    assert small.insert_knot(0.5 * sp.knots[-1], 1) == 0


def test_truncate_spline():
    rng = np.random.default_rng(4)
    sp, _ = random_spline(rng)
    kend = sp.knots[-1]
    ref_knots, ref_points = sp.knots, sp.points
    assert sp.truncate(kend + 1.0) == 0  # nothing to do
    assert np.array_equal(sp.knots, ref_knots) and np.array_equal(sp.points, ref_points)
    u_end = 0.37 * kend
    us = np.linspace(0, u_end, 100)
    before = curve(sp, us)
    assert sp.truncate(u_end) == 0
    assert sp.knots[-1] == u_end and np.sum(sp.knots == u_end) == 3
    assert len(sp.knots) == len(sp.points) + 3
    np.testing.assert_allclose(curve(sp, us), before, atol=1e-12)
    assert sp.eval(np.array([u_end + 1e-9]))[0][0] != 0  # out of range now
    assert sp.truncate(-1.0) == 0 and len(sp.knots) == 0 and len(sp.points) == 0  # cleared


def test_extend_with_control_points():
    rng = np.random.default_rng(5)
    sp, _ = random_spline(rng, W=4, D=3)
    u_join = sp.knots[-1]
    us = np.linspace(0, u_join, 150)
    before = curve(sp, us)
    end_point = sp.points[-1].copy()
    new_pts = end_point + np.cumsum(rng.uniform(0.1, 0.5, size=(4, 3)), axis=0)
    n0 = len(sp.points)
    assert sp.extend(new_pts) == 0
    assert len(sp.points) == n0 + 4 and len(sp.knots) == n0 + 4 + 3
    assert np.all(np.diff(sp.knots) >= 0) and sp.knots[-1] > u_join
    assert np.array_equal(sp.points[-4:], new_pts)
    # positions before the join are unchanged except in the last old span, where the
    # modified end point keeps the VALUE at the join (bspline.h:489-505)
    last_span_start = sp.knots[np.searchsorted(sp.knots, u_join) - 1]
    keep = us <= last_span_start
    np.testing.assert_allclose(curve(sp, us[keep]), before[keep], atol=1e-12)
    np.testing.assert_allclose(sp.eval(np.array([u_join]))[1][0, 0], end_point, atol=1e-12)
    # continuous first derivative across the join
    lo, hi = sp.eval(np.array([u_join - 1e-7]), 1)[1][0, 1], sp.eval(np.array([u_join + 1e-7]), 1)[1][0, 1]
    np.testing.assert_allclose(lo, hi, atol=1e-5)
    assert sp.extend(new_pts[:1]) == 12  # needs >= 2 points
    tiny = orc.Spline(sp.knots, sp.points, knot_capacity=len(sp.knots) + 2)
    assert tiny.extend(new_pts) == 9


def test_project_point_on_path():
    """path_tools_test.cc:41-115"""
    assert orc.project_point_on_path(np.zeros((0, 2)), np.zeros(2))["status"] == 3
    r = orc.project_point_on_path([[1.0, 1.0]], [1.0, 1.0])
    assert (r["waypoint_index"], r["distance_to_path"], r["line_parameter"]) == (0, 0.0, 0.0)
    r = orc.project_point_on_path([[1.0, 1.0], [2.0, 2.0]], [1.0, 1.0])
    assert (r["waypoint_index"], r["distance_to_path"], r["line_parameter"]) == (0, 0.0, 0.0)
    r = orc.project_point_on_path([[1.0, 1.0], [2.0, 2.0]], [2.0, 2.0])
    assert (r["waypoint_index"], r["distance_to_path"], r["line_parameter"]) == (0, 0.0, 1.0)
    np.testing.assert_allclose(r["projected_point"], [2.0, 2.0], atol=1e-10)
    wps = np.array([[1.0, 1.0], [2.0, 2.0], [-3.0, -3.0]])
    proj = wps[1] + 0.4 * (wps[2] - wps[1])
    r = orc.project_point_on_path(wps, proj + np.array([0.1, -0.1]))
    assert r["waypoint_index"] == 1
    assert r["distance_to_path"] == pytest.approx(np.linalg.norm([0.1, -0.1]), rel=1e-14)
    assert r["line_parameter"] == pytest.approx(0.4, rel=1e-14)
    np.testing.assert_allclose(r["projected_point"], proj, atol=1e-10)


@pytest.mark.parametrize("skip", [False, True])
def test_switch_to_new_waypoint_path(skip):
    """SwitchToNewJointWaypointPathWorks (path_timing_trajectory_test.cc:299-436): plan
    until the motion is fast, switch to new waypoints at the stop parameter with the
    current velocity as initial velocity, plan to the end of the new path."""
    wp = np.array([[1, 2, 3], [-1, -2, -3], [0.5, 1, 1.5]])
    vmax, amax = np.ones(3), np.full(3, 2.0)
    pl = orc.Planner(wp, vmax, amax, 1000, 0.001, 0.004, skip=skip, max_initial_velocity_error=1e-2)
    start = 0
    executed = []

    def advance(tr):
        nonlocal start
        nxt = min(pl.end_time, max(start + 200 * MS, pl.start_time))  # GetNextPlanStartTime
        # (half a nanosecond of slack: an erased trajectory keeps the bits of its old grid)
        keep = tr["r_t"] < nxt / 1e9 - 5e-10 if not pl.target_reached else np.ones(len(tr["r_t"]), bool)
        executed.append({k: v[keep] for k, v in tr.items()})
        start = nxt

    for _ in range(100):
        assert pl.plan(start, 750 * MS) == orc.ORC_OK
        tr = pl.trajectory()
        advance(tr)
        if np.min(np.abs(np.abs(tr["r_vel"][0]) - vmax)) < 0.3:
            break
    assert not pl.target_reached
    new_wp = np.array([[1.0, 2.0, 3.0], [0.5, 1.0, 5.5]])
    sp = orc.path_stop_parameter(tr["r_t"], tr["r_s"], tr["r_vel"], tr["r_acc"], amax, start / 1e9)
    k0, p0 = pl.spline()
    v0 = np.array([np.interp(start / 1e9, tr["r_t"], tr["r_vel"][:, d]) for d in range(3)])
    assert pl.switch(sp["parameter"], new_wp, initial_velocity=v0) == 0
    k1, p1 = pl.spline()
    us = np.linspace(0, sp["parameter"], 100)
    np.testing.assert_allclose(orc.eval_curve_ders(k1, p1, 2, 0, us)[1],
                               orc.eval_curve_ders(k0, p0, 2, 0, us)[1], atol=1e-12)
    np.testing.assert_allclose(orc.eval_curve(k1, p1, 2, k1[-1])[1], new_wp[-1], atol=1e-12)
    for _ in range(300):
        assert pl.plan(start, 750 * MS) == orc.ORC_OK
        tr = pl.trajectory()
        advance(tr)
        if pl.target_reached:
            break
    assert pl.target_reached
    np.testing.assert_allclose(tr["r_pos"][-1], new_wp[-1], atol=1e-10)
    np.testing.assert_allclose(tr["r_vel"][-1], 0.0, atol=1e-10)
    # the executed trajectory is smooth across the switch: finite differences of the
    # positions match the velocities (_test.cc:409-435)
    t = np.concatenate([e["r_t"] for e in executed])
    pos = np.concatenate([e["r_pos"] for e in executed])
    vel = np.concatenate([e["r_vel"] for e in executed])
    acc = np.concatenate([e["r_acc"] for e in executed])
    assert np.all(np.diff(t) > 0)
    assert np.all(np.abs(vel) <= vmax + 1e-9) and np.all(np.abs(acc) <= amax + 1e-9)
    fd = (pos[2:] - pos[:-2]) / (t[2:] - t[:-2])[:, None]
    err = np.abs(fd - vel[1:-1])
    assert err[:-20].max() < 1e-2 and err[-20:].max() < 1e-1
