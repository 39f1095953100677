"""CPU tests: pin the oracle against every golden vector / known-answer test the
reference holds for the hot path.  No GPU needed."""
import json
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


# ---------------------------------------------------------------- stage A
def test_bspline_mathematica_table(po):
    """BSplineGoldenTest.CompareToReference, splines/bspline_test.cc:730-840:
    L2 norm of value errors and squared L2 norm of derivative errors <= 5e-14."""
    g = _load("bspline_mathematica.json")
    knots = np.array(g["knots"], float)
    ctrl = np.array(g["control_points"], float)
    du = (knots[-1] - knots[0]) / (g["num_samples"] - 1)
    us = knots[0] + np.arange(g["num_samples"]) * du
    eps = g["tolerance"]
    pos = np.array([po.eval_curve(knots, ctrl, 3, u)[1] for u in us])
    assert np.linalg.norm(pos[:, 0] - g["X"]) <= eps
    assert np.linalg.norm(pos[:, 1] - g["Y"]) <= eps
    st, vals = po.eval_curve_ders(knots, ctrl, 3, 3, us)
    assert (st == 0).all()
    for k, (nx, ny) in enumerate((("X", "Y"), ("Xp", "Yp"), ("Xpp", "Ypp"), ("Xppp", "Yppp"))):
        assert np.sum((vals[:, k, 0] - g[nx]) ** 2) <= eps, nx
        assert np.sum((vals[:, k, 1] - g[ny]) ** 2) <= eps, ny
    # EvalCurve and EvalCurveAndDerivatives agree on the position
    assert np.max(np.abs(vals[:, 0] - pos)) <= 1e-15


def test_bspline_against_scipy(po):
    from scipy.interpolate import BSpline
    rng = np.random.default_rng(0)
    for p in (1, 2, 3, 4, 5):
        n = 9
        knots = np.concatenate([np.zeros(p), np.sort(rng.uniform(0, 3, n - p + 1)), ])
        knots = np.concatenate([[knots[p]] * 0, knots])
        knots = np.concatenate([np.full(p + 1, 0.0), np.sort(rng.uniform(0.2, 2.8, n - p - 1)),
                                np.full(p + 1, 3.0)])
        ctrl = rng.uniform(-2, 2, (n, 3))
        us = np.concatenate([rng.uniform(0, 3, 200), knots, [0.0, 3.0]])
        nder = min(p, 3)
        st, vals = po.eval_curve_ders(knots, ctrl, p, nder, us)
        assert (st == 0).all()
        spl = BSpline(knots, ctrl, p)
        for k in range(nder + 1):
            ref = spl.derivative(k)(us) if k else spl(us)
            # scipy is left-continuous only at the very last knot, like KnotSpan
            scale = max(1.0, np.max(np.abs(ref)))
            assert np.max(np.abs(vals[:, k] - ref)) <= 1e-11 * scale, (p, k)


def test_partition_of_unity_and_knot_span(po):
    """BSplineNdTest.BasisFunctions, bspline_test.cc:1091-1129."""
    knots = np.array([0, 0, 0, 1, 2, 3, 4, 4, 5, 5, 5], float)
    p = 2
    npts = len(knots) - p - 1
    u = knots[0]
    while u <= knots[-1]:
        span = po.knot_span(knots, p, u)
        assert knots[span] <= u and (u < knots[span + 1] or u == knots[-1])
        b = po.basis(knots, span, p, u)
        assert (b >= 0).all()
        assert abs(b.sum() - 1.0) <= 4 * np.finfo(float).eps
        u += 0.01
    # KnotSpan semantics (bspline_base.cc:218-246): right-continuous at interior
    # knots, last knot -> num_points - 1.
    assert po.knot_span(knots, p, 5.0) == npts - 1
    assert po.knot_span(knots, p, 1.0) == 3
    assert po.knot_span(knots, p, 4.0) == 7  # double knot: span after the multiple
    assert po.knot_span(knots, p, 0.0) == 2


def test_uniform_knot_vectors(po):
    """BSplineTest.AutoKnotVector, bspline_test.cc:1270-1299 (exact values)."""
    assert list(po.make_uniform_knots(4, 3)) == [0, 0, 0, 0, 1, 1, 1, 1]
    assert list(po.make_uniform_knots(5, 3)) == [0, 0, 0, 0, 0.5, 1, 1, 1, 1]
    assert list(po.make_uniform_knots(5, 1)) == [0, 0, 0.25, 0.5, 0.75, 1, 1]


def test_out_of_range_parameter(po):
    knots = np.array([0, 0, 0, 0, 0.5, 1, 1, 1, 1], float)
    ctrl = np.arange(10, dtype=float).reshape(5, 2)
    assert po.eval_curve(knots, ctrl, 3, -1e-9)[0] == 11  # kOutOfRange
    assert po.eval_curve(knots, ctrl, 3, 1.0 + 1e-9)[0] == 11
    assert po.eval_curve(knots, ctrl, 3, 1.0)[0] == 0
    st, v = po.eval_curve(knots, ctrl, 3, 0.0)
    assert st == 0 and np.array_equal(v, ctrl[0])  # end-point interpolation
    assert np.array_equal(po.eval_curve(knots, ctrl, 3, 1.0)[1], ctrl[-1])


# ---------------------------------------------------------------- stage B
def test_polyline_control_points_properties(po):
    """PolyLineToBspline3Waypoints (spline_utils.cc:47-102): corners kept, inner
    points at min(radius, dist/4) along each segment; the maximum deviation
    formula of spline_utils_test.cc:139-172 (degree-3 uniform spline at u=0.5)."""
    for wd in np.arange(0.5, 6, 0.5):
        corners = np.array([[0, 0, 0], [wd, 0, 0], [0, 0, 0]], float)
        for radius in np.arange(0, wd, 0.1):
            c = po.polyline_to_bspline3(corners, radius)
            assert c.shape == (7, 3)
            np.testing.assert_array_equal(c[[0, 3, 6]], corners)
            eff = min(radius, wd / 4.0)
            assert abs(np.linalg.norm(c[2] - c[3]) - eff) <= 1e-12
            assert abs(np.linalg.norm(c[4] - c[3]) - eff) <= 1e-12
            knots = po.make_uniform_knots(7, 3)
            st, pt = po.eval_curve(knots, c, 3, 0.5)
            assert st == 0
            assert abs(np.max(np.abs(pt - corners[1])) - eff / 3.0) <= 1e-10
    one = po.polyline_to_bspline3(np.array([[1.0, 2.0]]), 0.2)
    np.testing.assert_array_equal(one, np.tile([1.0, 2.0], (4, 1)))


def test_joint_fit_knots_and_sampling(po):
    """FitSplineToWaypoints / SamplePath (timeable_path_joint_spline.cc:252-318)."""
    rng = np.random.default_rng(3)
    wp = rng.uniform(-2, 2, (5, 4))
    ctrl, knots = po.joint_fit(wp, 0.2)
    assert ctrl.shape == (13, 4) and knots.shape == (16,)
    length = np.sum(np.linalg.norm(np.diff(ctrl, axis=0), axis=1))
    assert abs(knots[-1] - length) <= 1e-12 * length
    assert (knots[:3] == 0).all() and (knots[-3:] == knots[-1]).all()
    assert (np.diff(knots) >= 0).all()
    N = 200
    delta = knots[-1] / (N - 1) * 1.1  # tail beyond the end
    q, dq, ddq = po.joint_sample(knots, ctrl, N, 0.0, delta)
    np.testing.assert_array_equal(q[0], wp[0])
    tail = np.arange(N) * delta >= knots[-1] + delta
    assert tail.any()
    np.testing.assert_array_equal(q[tail], np.tile(wp[-1], (tail.sum(), 1)))
    assert (dq[tail] == 0).all() and (ddq[tail] == 0).all()
    # first derivative consistent with finite differences of the position
    h = 1e-6
    u = 0.37 * knots[-1]
    _, v = po.eval_curve_ders(knots, ctrl, 2, 2, [u - h, u, u + h])
    np.testing.assert_allclose((v[2, 0] - v[0, 0]) / (2 * h), v[1, 1], atol=1e-6)
    # tiny paths get the minimum final knot value 0.1 (:278-282)
    _, k2 = po.joint_fit(wp[:1].repeat(3, 0) + 1e-4 * rng.random((3, 4)), 0.2)
    assert k2[-1] == pytest.approx(0.1)


def test_constraint_rows(po):
    rng = np.random.default_rng(5)
    dq, ddq = rng.normal(size=(10, 3)), rng.normal(size=(10, 3))
    vmax, amax = np.array([1.0, 2.0, 3.0]), np.array([4.0, 5.0, 6.0])
    c = po.joint_constraints(dq, ddq, vmax, amax, 0.8)
    assert c.shape == (10, 6, 4)
    np.testing.assert_array_equal(c[:, :3, 0], dq)
    np.testing.assert_array_equal(c[:, :3, 1], ddq)
    np.testing.assert_array_equal(c[:, :3, 3], np.tile(amax * 0.8, (10, 1)))
    np.testing.assert_array_equal(c[:, :3, 2], -c[:, :3, 3])
    np.testing.assert_array_equal(c[:, 3:, 0], 0 * dq)
    np.testing.assert_array_equal(c[:, 3:, 1], dq * dq)
    np.testing.assert_array_equal(c[:, 3:, 3], np.tile((vmax * 0.8) ** 2, (10, 1)))


# ---------------------------------------------------------------- stage C
def test_lp_regression_cases(po):
    """FindMaxSd2Regression, time_optimal_path_timing_test.cc:744-1088: simplex ==
    brute force within 1e-8; and the port reproduces the reference's own simplex
    output bit for bit."""
    for lp in _load("lp_regression.json")["lps"]:
        rows = np.stack([lp["a"], lp["b"], lp["lower"], lp["upper"]], 1)
        got = po.find_max_sd2(rows, "port")
        np.testing.assert_array_equal(got, lp["ref_simplex"])
        assert abs(got[0] - lp["ref_brute"][0]) <= 1e-8
        assert abs(got[2] - lp["ref_brute"][2]) <= 1e-8


def test_reference_scenarios_golden(po):
    """Durations / profiles produced by the reference's own solver (compiled
    verbatim, tests/golden/make_golden.py) on its test scenarios."""
    from __graft_entry__ import load_package
    W = load_package().workloads
    for sc in _load("ref_solver_scenarios.json")["scenarios"]:
        name, N = sc["name"], sc["N"]
        if name.startswith("circle"):
            cons = W.circle_constraints(N)
        elif name.startswith("sine"):
            cons = W.sine_constraints(N, sc["s_start"], sc["s_end"])
        elif name.startswith("line"):
            cons = W.line_constraints(N)
        elif name.startswith("scalar_straight"):
            cons = W.scalar_straight_constraints(N)
        else:
            cons = W.scalar_curved_constraints(N)
        r = po.solve(cons, sc["s_start"], sc["s_end"], sc["sd_start"])
        assert r["ok"], name
        assert r["time"][-1] - r["time"][0] == sc["duration"], name
        assert r["last_extremal_index"] == sc["last_extremal_index"]
        np.testing.assert_array_equal(r["time"], sc["time"])
        np.testing.assert_array_equal(r["sd"], sc["sd"])
        np.testing.assert_array_equal(r["sdd"], sc["sdd"])
        # the reference's own acceptance: constraints respected within kTiny
        assert po.count_violations(cons, r["sd2"], r["sdd"]) == 0
        assert sc["satisfies"] == 1
    # anchor from SURVEY.md section 8c
    circ = [s for s in _load("ref_solver_scenarios.json")["scenarios"]
            if s["name"] == "circle_N50_sd0.0"][0]
    assert circ["duration"] == 6.0003326777697756


def test_circle_limits_respected(po):
    """VerifyCircleExampleSolution, time_optimal_path_timing_test.cc:164-210."""
    from __graft_entry__ import load_package
    W = load_package().workloads
    R, vmax, amax, ktiny = 2.0, 1.2, 1.0, np.finfo(float).eps * 1e5
    for N in (50, 51):
        r = po.solve(W.circle_constraints(N), 0.0, np.pi)
        s, sd, sdd = r["s"], r["sd"], r["sdd"]
        xd, yd = -R * np.sin(s) * sd, R * np.cos(s) * sd
        xdd = -R * np.sin(s) * sdd - R * np.cos(s) * sd * sd
        ydd = R * np.cos(s) * sdd - R * np.sin(s) * sd * sd
        for arr, lim in ((xd, vmax), (yd, vmax), (xdd, amax), (ydd, amax)):
            assert np.max(np.abs(arr) - lim) < ktiny


def test_query_interpolation(po):
    from __graft_entry__ import load_package
    W = load_package().workloads
    r = po.solve(W.circle_constraints(50), 0.0, np.pi, sd_start=0.1)
    ok, s, sd, sdd = po.query(r, 0.0, np.pi, 0.0)
    assert ok and s == 0.0 and sd == pytest.approx(0.1)
    ok, s, sd, sdd = po.query(r, 0.0, np.pi, 1e9)
    assert ok and s == np.pi and sd == 0.0 and sdd == 0.0
    ts = np.linspace(r["time"][0], r["time"][-1], 300)
    ss = [po.query(r, 0.0, np.pi, t)[1] for t in ts]
    assert (np.diff(ss) >= 0).all()


# ---------------------------------------------------------------- stage D
def test_time_conversion(po):
    """time.h:23-30: seconds * 1e9 truncated toward zero to int64 nanoseconds."""
    assert po.time_from_sec(1.5) == 1500000000
    assert po.time_from_sec(0.0049999999999) == 4999999
    assert po.time_from_sec(-0.0000000015) == -1
    assert po.time_to_sec(1500000000) == 1.5


def test_plan_properties(po):
    """path_timing_trajectory_test.cc:112-173: the resampled trajectory starts at
    the first waypoint, ends exactly at the last one with zero velocity and
    acceleration, respects the limits, and velocity ~ finite differences."""
    from __graft_entry__ import load_package
    wp, vmax, amax = load_package().workloads.joint_paths(6, 5, 3, seed=11)
    for b in range(6):
        r = po.plan_joint(wp[b], vmax[b], amax[b], 1000, dt=0.004, M_cap=20000)
        assert r["ok"]
        M = r["M"]
        assert M == int(np.ceil(r["duration"] / 0.004)) + 1
        np.testing.assert_allclose(r["r_pos"][0], wp[b, 0], atol=1e-14)
        np.testing.assert_array_equal(r["r_pos"][-1], wp[b, -1])
        assert (r["r_vel"][-1] == 0).all() and (r["r_acc"][-1] == 0).all()
        assert (np.abs(r["r_acc"]) <= amax[b]).all()
        assert (np.abs(r["r_vel"]) <= vmax[b] * 0.8 + 1e-6).all()
        fd = np.diff(r["r_pos"][:-1], axis=0) / 0.004
        assert np.max(np.abs(fd - r["r_vel"][:-2])) <= 0.1  # lerp between path samples
        assert (np.diff(r["r_s"][:-1]) >= -1e-12).all()


def test_joint_pipeline_golden(po):
    """Whole pipeline vs outputs recorded with the reference's solver in the loop."""
    from __graft_entry__ import load_package
    wp, vmax, amax = load_package().workloads.joint_paths(16, 10, 7, seed=1000)
    for g in _load("ref_joint_paths.json")["paths"][:6]:
        b = g["index"]
        r = po.plan_joint(wp[b], vmax[b], amax[b], 2048, dt=0.004, M_cap=16384)
        assert r["ok"] and r["duration"] == g["duration"] and r["M"] == g["M"]
        assert r["last_extremal_index"] == g["last_extremal_index"]
        assert float(np.sum(r["time"])) == g["sum_time"]
        assert float(np.sum(r["r_pos"])) == g["sum_r_pos"]
