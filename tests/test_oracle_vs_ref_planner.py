"""Stages A, B, D and the moving-horizon planner of the CPU oracle against the
REFERENCE'S OWN OBJECT CODE: splines/bspline_base.cc, splines/bspline.h,
splines/spline_utils.cc, timeable_path_joint_spline.cc, time_optimal_path_timing.cc and
path_timing_trajectory.cc compiled verbatim into oracle/_ref (oracle/Makefile,
oracle/ref_wrap_planner.cc).  Port (oracle/tp_oracle*.c) and reference run the
scenarios of path_timing_trajectory_test.cc:112-546 and seeded random paths side by
side and are compared BIT FOR BIT after every Plan().  CPU only.

What is still an assumption after this (un-vendored Eigen / eigenmath, see
oracle/shim): the summation order of norm()/dot(), eigenmath::InterpolateLinear,
DistanceFromLineSegment, and Eigen's colPivHouseholderQr inside TruncateSplineAt (the
shim restates Eigen's algorithm; the port uses the closed-form knot insertion, so the
switch test holds control points to 1e-12 instead of bit equality)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as orc  # noqa: E402

pytestmark = pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built")

MS = 1_000_000
WAYPOINTS = np.array([[1, 2, 3], [-1, -2, -3], [1, 2, 3.0]])  # _test.cc:127-128
VMAX, AMAX = np.full(3, 1.0), np.full(3, 2.0)
N, DELTA, DT = 1000, 0.005, 0.004  # kNumSamples, JointPathOptions default, kTimeStep
HORIZON, REPLAN = 750 * MS, 200 * MS
ROW_KEYS = ("r_t", "r_pos", "r_vel", "r_acc", "r_s", "r_sd", "r_sdd")
SAMPLE_KEYS = ("w_time", "w_s", "w_sd", "w_sdd", "w_q", "w_vel", "w_acc")


def pair(wp=WAYPOINTS, vmax=VMAX, amax=AMAX, n=N, delta=DELTA, skip=False, **kw):
    port = orc.Planner(wp, vmax, amax, n, delta, DT, skip=skip, **kw)
    kw.pop("P_cap", None)
    kw.pop("M_cap", None)
    ref = orc.RefPlanner(wp, vmax, amax, n, delta, DT, skip=skip, **kw)
    assert ref.set_status == 0
    return port, ref


def same_bits(a, b, what):
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if a.tobytes() != b.tobytes():
        bad = np.flatnonzero(a.ravel() != b.ravel())
        raise AssertionError(f"{what}: {bad.size} of {a.size} values differ, first at {bad[0]}: "
                             f"{a.ravel()[bad[0]]!r} vs {b.ravel()[bad[0]]!r}")


def check_equal(port, ref, st_port, st_ref, spline_bits=True):
    """Status, planner state, trajectory and spliced path samples after a Plan()."""
    assert orc.ABSL_OF_ORC[st_port] == st_ref, (st_port, st_ref)
    s = ref.state()
    for k in ("start_time", "end_time", "final_decel_start", "planned_to_end", "target_reached",
              "initial_plan", "M", "len"):
        assert getattr(port, k) == s[k], (k, getattr(port, k), s[k])
    assert port.sampled == orc.ORC_STATE_OF_REF[s["path_state"]]
    tp, tr = port.trajectory(), ref.trajectory()
    for k in ROW_KEYS:
        same_bits(tp[k], tr[k], k)
    wp_, wr = port.path_samples(), ref.path_samples()
    for k in SAMPLE_KEYS:
        same_bits(wp_[k], wr[k], k)
    if spline_bits:
        (kp, pp), (kr, pr) = port.spline(), ref.spline()
        same_bits(kp, kr, "knots")
        same_bits(pp, pr, "control points")
    return tp


@pytest.mark.parametrize("skip", [False, True])
def test_rest_to_rest_replanning(skip):
    """RestToRestPlanningWorks (_test.cc:112-173): Plan() every 200 ms with a 750 ms
    horizon until the target is reached -- several windows spliced, most calls only
    erase.  Equal after every call."""
    port, ref = pair(skip=skip)
    start, calls = 0, 0
    while not port.target_reached:
        check_equal(port, ref, port.plan(start, HORIZON), ref.plan(start, HORIZON))
        start = min(port.end_time, start + REPLAN)
        calls += 1
        assert calls < 500
    assert calls > 20 and port.len > N and ref.state()["at_end"] == 1


@pytest.mark.parametrize("skip", [False, True])
def test_one_call_with_long_horizon(skip):
    """The window loop inside one Plan() (:628-658)."""
    port, ref = pair(skip=skip)
    h = 3600 * 1000 * MS
    check_equal(port, ref, port.plan(0, h), ref.plan(0, h))
    assert port.loops >= 3 and port.target_reached


@pytest.mark.parametrize("skip", [False, True])
def test_no_duplicate_initial_samples(skip):
    """NoDuplicateInitialSamples (_test.cc:175-255): re-plan from times at / next to
    existing samples."""
    port, ref = pair(skip=skip)
    check_equal(port, ref, port.plan(0, HORIZON), ref.plan(0, HORIZON))
    for sample in (0, 2, 1, 3):
        t = port.trajectory()["r_t"]
        start = int(t[sample] * 1e9)
        check_equal(port, ref, port.plan(start, HORIZON), ref.plan(start, HORIZON))
    # start times between the samples (:229-254)
    for frac in (0.25, 0.5, 0.99):
        t = port.trajectory()["r_t"]
        start = int((t[1] + frac * (t[2] - t[1])) * 1e9)
        check_equal(port, ref, port.plan(start, HORIZON), ref.plan(start, HORIZON))


@pytest.mark.parametrize("skip", [False, True])
def test_invariant_to_starting_time(skip):
    """IsInvariantToStartingTime (_test.cc:257-297), both start times on both sides."""
    for start in (0, 42_000 * MS):
        port, ref = pair(skip=skip)
        check_equal(port, ref, port.plan(start, HORIZON), ref.plan(start, HORIZON))


def test_time_argument_errors():
    """HandleTimeArguments (:503-538), GetTimeOffsetAfter (:289-305), kDeadlineExceeded
    (:652-655): the same error codes, and the state survives them identically."""
    port, ref = pair()
    check_equal(port, ref, port.plan(1000 * MS, HORIZON), ref.plan(1000 * MS, HORIZON))
    end = port.end_time
    for start in (end + 5 * MS, end + 2 * MS, 500 * MS, 1000 * MS):
        check_equal(port, ref, port.plan(start, HORIZON), ref.plan(start, HORIZON))
    port, ref = pair(max_planning_iterations=1)
    h = 3600 * 1000 * MS
    st_p, st_r = port.plan(0, h), ref.plan(0, h)
    assert st_p == orc.ORC_FAIL_DEADLINE and st_r == 4


def test_initial_velocity_projection():
    """ComputeTimingProfile (:352-385)"""
    tangent = WAYPOINTS[1] - WAYPOINTS[0]
    v0 = 0.05 * tangent / np.linalg.norm(tangent)
    port, ref = pair(initial_velocity=v0)
    check_equal(port, ref, port.plan(0, HORIZON), ref.plan(0, HORIZON))
    port, ref = pair(initial_velocity=np.array([0.05, 0.0, 0.0]))
    assert port.plan(0, HORIZON) == orc.ORC_FAIL_INITIAL_VELOCITY and ref.plan(0, HORIZON) == 3


def random_path(rng, W, D):
    """cfg2-style waypoints and limits."""
    wp = np.zeros((W, D))
    wp[0] = rng.uniform(-np.pi, np.pi, D)
    for w in range(1, W):
        wp[w] = np.clip(wp[w - 1] + rng.uniform(-1, 1, D), -np.pi, np.pi)
    return wp, rng.uniform(0.5, 2.0, D), rng.uniform(1.0, 4.0, D)


FRESH_KEYS = ("q", "dq", "ddq", "time", "s", "sd", "sdd", "vel", "acc") + ROW_KEYS


@pytest.mark.parametrize("skip", [False, True])
def test_fresh_paths_all_stages(skip):
    """100 seeded random fresh paths, whole path in one window: control points, knots,
    SamplePath (q, q', q''), ConstraintSetup rows, the profile, per-sample velocity /
    acceleration and the resampled trajectory of the port equal the reference's, bit
    for bit; D in 1..8, W in 2..12, N in {64..700}."""
    rng = np.random.default_rng(20260 + int(skip))
    n_ok = 0
    for case in range(100):
        D, W = int(rng.integers(1, 9)), int(rng.integers(2, 13))
        n = int(rng.choice([64, 100, 257, 400, 700]))
        wp, vmax, amax = random_path(rng, W, D)
        rounding = float(rng.choice([0.2, 0.05, 0.5]))
        safety = float(rng.choice([0.8, 1.0, 0.6]))
        o = orc.plan_joint(wp, vmax, amax, n, rounding=rounding, safety=safety, dt=DT,
                           M_cap=40000, skip=skip)
        r = orc.plan_joint_ref(wp, vmax, amax, n, rounding=rounding, safety=safety, dt=DT,
                               skip=skip)
        assert orc.ABSL_OF_ORC[o["status"]] == r["status"], (case, o["status"], r["status"])
        ctrl, kn = orc.joint_fit(wp, rounding)
        same_bits(kn, r["spline"][0], "knots")
        same_bits(ctrl, r["spline"][1], "control points")
        assert o["delta"] == r["delta"] and o["knot_end"] == r["knot_end"]
        same_bits(o["q"], r["q"], "q")
        same_bits(o["dq"], r["dq"], "dq")
        same_bits(o["ddq"], r["ddq"], "ddq")
        same_bits(orc.joint_constraints(o["dq"], o["ddq"], vmax, amax, safety), r["cons"], "rows")
        if not o["ok"]:
            continue
        n_ok += 1
        assert o["M"] == r["M"] and o["last_extremal_index"] == r["last_extremal_index"]
        for k in FRESH_KEYS:
            same_bits(o[k][:o["M"]] if k in ROW_KEYS else o[k], r[k], f"case {case} {k}")
    assert n_ok >= 90


def test_fresh_path_edge_cases():
    """One waypoint (:37-41 of spline_utils.cc), duplicate waypoints, sub-kMinNorm
    segments, a non-default sampling distance that ends beyond the path (tail rule of
    SamplePath :301-315)."""
    vmax, amax = np.full(3, 1.0), np.full(3, 2.0)
    cases = [
        (np.array([[0.3, -0.2, 0.1]]), -1.0),
        (np.array([[0, 0, 0], [0, 0, 0], [1, 1, 1.0]]), -1.0),
        (np.array([[0, 0, 0], [1e-7, 0, 0], [1, 1, 1.0], [1, 1, 1.0]]), -1.0),
        (WAYPOINTS, 0.02),   # 200 samples x 0.02 = 4 > knots.back(): tail samples
        (WAYPOINTS, 0.001),  # window ends inside the path
    ]
    for wp, delta in cases:
        o = orc.plan_joint(wp, vmax, amax, 200, delta=delta, dt=DT, M_cap=40000)
        r = orc.plan_joint_ref(wp, vmax, amax, 200, delta=delta, dt=DT)
        assert orc.ABSL_OF_ORC[o["status"]] == r["status"]
        for k in ("q", "dq", "ddq"):
            same_bits(o[k], r[k], k)
        if o["ok"]:
            for k in FRESH_KEYS:
                same_bits(o[k][:o["M"]] if k in ROW_KEYS else o[k], r[k], k)


@pytest.mark.parametrize("skip", [False, True])
def test_random_paths_replanned_in_windows(skip):
    """Seeded random 7-dof paths re-planned every 200 ms in windows that cover about a
    quarter of the path: equal after every Plan()."""
    rng = np.random.default_rng(777 + int(skip))
    for case in range(6):
        wp, vmax, amax = random_path(rng, int(rng.integers(3, 9)), 7)
        _, kn = orc.joint_fit(wp, 0.2)
        n = 256
        delta = kn[-1] / (n - 1) / 4.0
        port, ref = pair(wp, vmax, amax, n, delta, skip, P_cap=64 * n)
        start, calls = 0, 0
        while not port.target_reached:
            st_p, st_r = port.plan(start, HORIZON), ref.plan(start, HORIZON)
            check_equal(port, ref, st_p, st_r)
            assert st_p == orc.ORC_OK
            start = min(port.end_time, start + REPLAN)
            calls += 1
            assert calls < 1000
        assert port.len > n


def test_stop_parameter():
    """GetPathStopParameter / ComputeFastestStop (:75-172, 235-287) on the planned
    trajectory, at times before, inside, at the end of and beyond it."""
    port, ref = pair()
    assert ref.stop_parameter(0) == (0, 0.0)  # no plan yet
    check_equal(port, ref, port.plan(0, HORIZON), ref.plan(0, HORIZON))
    tr = port.trajectory()
    t = tr["r_t"]
    queries = [0, int(t[3] * 1e9), int(t[len(t) // 2] * 1e9), int(t[len(t) // 2] * 1e9) + 1_500_000,
               int(t[-2] * 1e9), int(t[-1] * 1e9), int(t[-1] * 1e9) + 10 * MS]
    for q in queries:
        st_r, par_r = ref.stop_parameter(q)
        o = orc.path_stop_parameter(tr["r_t"], tr["r_s"], tr["r_vel"], tr["r_acc"], AMAX, q / 1e9)
        # the port reports "time past the trajectory" as ORC_FAIL_SETUP (1); the reference
        # returns kInvalidArgument (3)
        assert {0: 0, 1: 3}[o["status"]] == st_r, (q, o["status"], st_r)
        if st_r == 0:
            assert o["parameter"] == par_r, (q, o["parameter"], par_r)


@pytest.mark.parametrize("skip", [False, True])
def test_switch_to_new_waypoint_path(skip):
    """SwitchToNewJointWaypointPathWorks (_test.cc:299-436): plan until the motion is
    fast, switch at the stop parameter with the current velocity, plan to the end.
    TruncateSplineAt goes through the shim's restatement of Eigen's QR in the reference
    and through the closed form in the port: knots equal bit for bit, control points
    and everything planned after the switch within 1e-12 / the BASELINE tolerances."""
    wp = np.array([[1, 2, 3], [-1, -2, -3], [0.5, 1, 1.5]])
    vmax, amax = np.ones(3), np.full(3, 2.0)
    port, ref = pair(wp, vmax, amax, 1000, 0.001, skip, max_initial_velocity_error=1e-2)
    start = 0
    for _ in range(100):
        tr = check_equal(port, ref, port.plan(start, HORIZON), ref.plan(start, HORIZON))
        start = min(port.end_time, max(start + REPLAN, port.start_time))
        if np.min(np.abs(np.abs(tr["r_vel"][0]) - vmax)) < 0.3:
            break
    assert not port.target_reached
    new_wp = np.array([[1.0, 2.0, 3.0], [0.5, 1.0, 5.5]])
    st_r, par_r = ref.stop_parameter(start)
    sp = orc.path_stop_parameter(tr["r_t"], tr["r_s"], tr["r_vel"], tr["r_acc"], amax, start / 1e9)
    assert st_r == 0 and sp["status"] == 0 and sp["parameter"] == par_r
    v0 = np.array([np.interp(start / 1e9, tr["r_t"], tr["r_vel"][:, d]) for d in range(3)])
    assert port.switch(par_r, new_wp, initial_velocity=v0) == 0
    assert ref.switch(par_r, new_wp, initial_velocity=v0) == 0
    (kp, pp), (kr, pr) = port.spline(), ref.spline()
    same_bits(kp, kr, "knots after the switch")
    assert pp.shape == pr.shape
    np.testing.assert_allclose(pp, pr, rtol=0, atol=1e-12)
    assert ref.state()["path_state"] == 2 and port.sampled == 2  # kModifiedPath
    for _ in range(300):
        st_p, st_r = port.plan(start, HORIZON), ref.plan(start, HORIZON)
        assert st_p == orc.ORC_OK and st_r == 0
        s = ref.state()
        assert (port.M, port.len, port.target_reached) == (s["M"], s["len"], s["target_reached"])
        assert abs(port.end_time - s["end_time"]) <= 4 * MS
        tp, tr = port.trajectory(), ref.trajectory()
        np.testing.assert_allclose(tp["r_t"], tr["r_t"], rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(tp["r_pos"], tr["r_pos"], rtol=0, atol=1e-9)
        start = min(port.end_time, max(start + REPLAN, port.start_time))
        if port.target_reached:
            break
    assert port.target_reached
    np.testing.assert_allclose(tr["r_pos"][-1], new_wp[-1], atol=1e-10)



@pytest.mark.parametrize("degree", [1, 2, 3, 4, 5])
def test_spline_evaluation_equals_reference_object_code(degree):
    """Stage A (KnotSpan, UpdateBasisAndDerivatives, EvalCurveAndDerivatives): the port's
    values and derivatives of every order up to min(degree, 3) equal the reference's
    BSplineXd bit for bit -- uniform grid, the knots themselves, both ends, shuffled
    parameters, out-of-range parameters (kOutOfRange, output untouched)."""
    rng = np.random.default_rng(degree)
    S, n, D, E = 3, 12 + degree, 5, 400
    nder = min(degree, 3)
    knots = np.stack([orc.make_uniform_knots(n, degree, 0.0, 1.0) * s for s in (1.0, 8.0, 0.37)])
    ctrl = rng.uniform(-np.pi, np.pi, (S, n, D))
    u = np.stack([np.concatenate([np.linspace(k[0], k[-1], E - 2 * len(k) - 4), k, k[::-1],
                                  [k[0] - 0.1, k[-1] + 1e-9, np.nextafter(k[-1], 0), k[0]]])
                  for k in knots])
    u[1] = rng.permutation(u[1])
    ref_out, ref_st = orc.spline_eval_batch_ref(knots, ctrl, degree, nder, u, threads=2)
    for s in range(S):
        st, out = orc.eval_curve_ders(knots[s], ctrl[s], degree, nder, u[s])
        assert np.array_equal(np.asarray(st), ref_st[s])
        ok = ref_st[s] == 0
        assert (~ok).sum() == 2 and (ref_st[s][~ok] == 11).all()
        same_bits(np.asarray(out)[ok], ref_out[s][ok], f"degree {degree} spline {s}")
