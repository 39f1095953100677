"""The Cartesian path (BASELINE config 5) of the CPU oracle against the REFERENCE'S OWN
OBJECT CODE: splines/bsplineq.cc and timeable_path_cartesian_spline.cc compiled verbatim
into oracle/_ref (oracle/Makefile, oracle/ref_wrap_planner.cc), behind the reference's
path_timing_trajectory.cc.  Port (oracle/tp_oracle_cart.c, the Cartesian branch of
tp_oracle_windows.c) and reference run the scenarios of path_timing_trajectory_test.cc:
613-757, bsplineq_test.cc and seeded random paths side by side and are compared BIT FOR
BIT.  CPU only.

The PathIKFunc / JacobianFunc are user callbacks, not part of the path: both sides call
the same C functions (the reference test's FakePathIK / FakeJacobian as restated in
tp_oracle_cart.c).

Still an assumption after this (un-vendored Eigen / eigenmath, oracle/shim): the scalar
formulas and summation order of Eigen's quaternion product, squaredNorm, stableNorm,
slerp and of the 6 x D matrix-vector product, and eigenmath::PoseErrorBetween's
translation part."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as orc  # noqa: E402

pytestmark = pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built")

MS = 1_000_000
DT = 0.004
HORIZON, REPLAN = 750 * MS, 200 * MS
ROW_KEYS = ("r_t", "r_pos", "r_vel", "r_acc", "r_s", "r_sd", "r_sdd")
SAMPLE_KEYS = ("w_time", "w_s", "w_sd", "w_sdd", "w_q", "w_vel", "w_acc")
CART_WAYPOINTS = np.array([[1.0, 0, 0, 0, 0, 0, 0], [0, 1.0, 0, 0, 0, 0, 0], [0, 0, 0, 1.0, 0, 0, 0],
                           [0, 0, 0, 0, 1.0, 0, 0], [0, 0, 0, 0, 0, 0, 1.0]])  # _test.cc:711-716


def same_bits(a, b, what):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if a.tobytes() != b.tobytes():
        bad = np.flatnonzero(a.ravel() != b.ravel())
        raise AssertionError(f"{what}: {bad.size} of {a.size} values differ, first at {bad[0]}: "
                             f"{a.ravel()[bad[0]]!r} vs {b.ravel()[bad[0]]!r}")


def rand_quat(rng, n=None):
    q = rng.normal(size=(4,) if n is None else (n, 4))
    return q / np.linalg.norm(q, axis=-1, keepdims=True)


def fake_fk_pose(q):
    """FakeFK (_test.cc:548-552): translation q[0:3], rotation ExpSO3(q[3:6])."""
    return np.concatenate([q[:3], orc.exp_so3(q[3:6])])


# ---------------------------------------------------------------- quaternion pieces
def test_quat_log_exp_power_bits():
    """bsplineq_details (bsplineq.cc:112-160) on unit, non-unit, negative-real, nearly
    real and exactly real quaternions."""
    rng = np.random.default_rng(1)
    cases = [rand_quat(rng) for _ in range(200)]
    cases += [rng.normal(size=4) * s for s in (1e-3, 0.1, 3.0, 50.0) for _ in range(20)]
    cases += [np.array([0, 0, 0, 1.0]), np.array([0, 0, 0, -1.0]), np.array([0, 0, 0, 2.5]),
              np.array([1e-14, 0, 0, 1.0]), np.array([1e-13, -2e-13, 1e-14, 1.0]),
              np.array([1.0, 0, 0, 0]), np.array([0, -1.0, 0, 1e-300]),
              np.array([3e-12, 0, 0, -1.0])]
    for q in cases:
        same_bits(orc.quat_log(q), orc.quat_ref("log", q), ("log", q))
        same_bits(orc.quat_exp(q), orc.quat_ref("exp", q), ("exp", q))
        for p in (0.0, 1.0, 0.5, 0.123456789, 1e-9, 0.999999):
            same_bits(orc.quat_power(q, p), orc.quat_ref("power", q, p), ("power", q, p))


@pytest.mark.parametrize("degree", [1, 2, 3])
def test_bsplineq_eval_curve_bits(degree):
    """BSplineQ::EvalCurve (bsplineq.cc:223-244), including both ends of the knot range,
    parameters on knots and out-of-range parameters (kOutOfRange)."""
    rng = np.random.default_rng(20 + degree)
    for n in (degree + 1, degree + 3, 12):
        knots = orc.make_uniform_knots(n, degree, 0.0, 2.5)
        quats = rand_quat(rng, n)
        quats[rng.random(n) < 0.3] *= -1.0  # both signs of the same rotation
        us = np.concatenate([rng.uniform(0, 2.5, 64), knots, [0.0, 2.5, -0.1, 2.6]])
        st, q, _ = orc.bsplineq_eval_ref(knots, quats, degree, us)
        for i, u in enumerate(us):
            st_p, q_p = orc.quat_eval_curve(knots, quats, degree, u)
            assert st_p == st[i], (u, st_p, st[i])
            if st_p == 0:
                same_bits(q_p, q[i], ("EvalCurve", degree, n, u))
        assert (st[-2:] == 11).all() and (st[:-2] == 0).all()


@pytest.mark.parametrize("degree", [1, 2, 3, 4])
def test_bsplineq_eval_curve_and_derivative_bits(degree):
    """BSplineQ::EvalCurveAndDerivative (bsplineq.cc:246-307) and QuatPowerDerivative
    (:148-160); the derivative also against a central difference of EvalCurve."""
    rng = np.random.default_rng(40 + degree)
    for q in [rand_quat(rng) for _ in range(50)] + [np.array([0, 0, 0, 1.0])]:
        for p in (0.0, 0.3, 1.0):
            same_bits(orc.quat_power_derivative(q, p), orc.quat_ref("power_derivative", q, p),
                      ("power derivative", q, p))
    for n in (degree + 1, degree + 4, 11):
        knots = orc.make_uniform_knots(n, degree, 0.0, 1.7)
        quats = rand_quat(rng, n)
        quats[rng.random(n) < 0.3] *= -1.0
        us = np.concatenate([rng.uniform(0, 1.7, 48), knots, [-1e-9, 1.7000001]])
        st, q, dq = orc.bsplineq_eval_ref(knots, quats, degree, us, derivative=True)
        for i, u in enumerate(us):
            st_p, q_p, dq_p = orc.quat_eval_curve_derivative(knots, quats, degree, u)
            assert st_p == st[i]
            if st_p == 0:
                same_bits(q_p, q[i], ("value", degree, n, u))
                same_bits(dq_p, dq[i], ("derivative", degree, n, u))
        assert (st[-2:] == 11).all() and (st[:-2] == 0).all()
    # the derivative is d(curve)/du: central difference of the normalised EvalCurve value
    # where consecutive control points are close (the sign of the value is then stable)
    base = rand_quat(rng)
    n = degree + 5
    quats = np.stack([orc.quat_mul(base, orc.exp_so3(0.2 * rng.normal(size=3))) for _ in range(n)])
    knots = orc.make_uniform_knots(n, degree, 0.0, 1.0)
    h = 1e-6
    for u in rng.uniform(0.05, 0.95, 16):
        _, q0, d0 = orc.quat_eval_curve_derivative(knots, quats, degree, u)
        _, qa = orc.quat_eval_curve(knots, quats, degree, u - h)
        _, qb = orc.quat_eval_curve(knots, quats, degree, u + h)
        np.testing.assert_allclose(d0, (qb - qa) / (2 * h), atol=2e-6)


def test_pose_polyline_bits():
    """PolyLineToBspline3Waypoints for poses (spline_utils.cc), 1..6 corners, with
    repeated corners and half-turn rotations."""
    rng = np.random.default_rng(3)
    for W in (1, 2, 3, 4, 6):
        for rep in range(6):
            corners = np.concatenate([rng.uniform(-1, 1, (W, 3)), rand_quat(rng, W)], axis=1)
            if rep == 1 and W > 1:
                corners[1] = corners[0]
            if rep == 2 and W > 2:
                corners[2, 3:] = -corners[1, 3:]  # same rotation, other sign
            for tr, rr in ((0.05, 0.2), (0.5, 0.01), (1e-3, 1.0)):
                same_bits(orc.pose_polyline_to_bspline3(corners, tr, rr),
                          orc.pose_polyline_to_bspline3_ref(corners, tr, rr), ("pose polyline", W, rep))


# ---------------------------------------------------------------- the path object
def cart_pair(joint_wp, N=1000, delta=0.005, skip=False, vmax=1.0, amax=2.0, **kw):
    joint_wp = np.asarray(joint_wp, np.float64)
    poses = np.stack([fake_fk_pose(q) for q in joint_wp])
    args = (joint_wp, np.full(7, vmax), np.full(7, amax), N, delta, DT)
    port = orc.Planner(*args, skip=skip, pose_waypoints=poses, max_initial_velocity_error=1e-2, **kw)
    kw.pop("P_cap", None)
    kw.pop("M_cap", None)
    ref = orc.RefPlanner(*args, skip=skip, pose_waypoints=poses, max_initial_velocity_error=1e-2,
                         **kw)
    assert ref.set_status == 0
    return port, ref


def check_equal(port, ref, st_port, st_ref):
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
    (kp, pp), (kr, pr) = port.spline(), ref.spline()
    same_bits(kp, kr, "knots")
    same_bits(pp, pr, "joint control points")
    same_bits(port._pose_ctrl[:len(pr)], ref.pose_points(), "pose control points")


def test_fit_sample_and_rows_one_window():
    """FitSplineToWaypoints, SamplePath(0) and ConstraintSetup (:396-595) of a fresh
    path: knots, control points, the N + 1 targets handed to the IK, q / q' / q'' and all
    2 D + 2 constraint rows per sample."""
    rng = np.random.default_rng(5)
    for trial in range(6):
        W = int(rng.integers(2, 6))
        joint = rng.uniform(-1.0, 1.0, (W, 7))
        poses = np.stack([fake_fk_pose(q) for q in joint])
        N = int(rng.integers(50, 400))
        vmax, amax = rng.uniform(0.5, 2.0, 7), rng.uniform(1.0, 4.0, 7)
        one = orc.plan_cart_fake(poses, joint, vmax, amax, N, dt=DT, M_cap=100000,
                                 translation_rounding=0.03, rotation_rounding=0.15,
                                 max_trans_vel=0.4, max_rot_vel=0.3)
        ref = orc.RefPlanner(joint, vmax, amax, N, -1.0, DT, rounding=0.15, pose_waypoints=poses,
                             translation_rounding=0.03, max_trans_vel=0.4, max_rot_vel=0.3)
        assert ref.set_status == 0
        st = ref.plan(0, 1)
        assert (st == 0) == one["ok"], (trial, st, one["status"])
        kn, jc = ref.spline()
        same_bits(one["knots"], kn, "knots")
        same_bits(one["pose_ctrl"], ref.pose_points(), "pose control points")
        pt, _ = ref.targets()
        same_bits(one["pose_targets"], pt, "pose targets")
        same_bits(one["q"], ref.ik_positions()[:N], "ik positions")
        w = ref.window()
        for k in ("q", "dq", "ddq"):
            same_bits(one[k], w[k], k)
        same_bits(one["cons"], ref.constraints(), "constraint rows")
        if one["ok"]:
            for k in ("time", "s", "sd", "sdd"):
                same_bits(one[k], w[k], k)
            tr = ref.trajectory()
            for k in ROW_KEYS:
                same_bits(one[k], tr[k], k)
        ref.close()


@pytest.mark.parametrize("skip", [False, True])
def test_cartesian_path_replanning(skip):
    """CartesianPathReplanning (_test.cc:695-757): the 5-waypoint path, Plan() every
    200 ms with a 750 ms horizon; several windows (path_start > 0, the IK cache grows
    window by window, :464-549) -- equal after every call."""
    port, ref = cart_pair(CART_WAYPOINTS, skip=skip)
    start, calls = 0, 0
    while not port.target_reached:
        check_equal(port, ref, port.plan(start, HORIZON), ref.plan(start, HORIZON))
        start = min(port.end_time, start + REPLAN)
        calls += 1
        assert calls < 2000
    assert port.len > 1000 and ref.state()["at_end"] == 1
    # the reference's IK cache holds every index up to the last window's horizon
    assert len(ref.ik_positions()) >= port.len


@pytest.mark.parametrize("skip", [False, True])
def test_simple_cartesian_path(skip):
    """SimpleCartesianPath (_test.cc:613-693): two identical waypoints."""
    wp = np.array([[1.0, 0, 0, 0, 0, 0, 0], [1.0, 0, 0, 0, 0, 0, 0]])
    port, ref = cart_pair(wp, skip=skip)
    start, calls = 0, 0
    while not port.target_reached:
        check_equal(port, ref, port.plan(start, HORIZON), ref.plan(start, HORIZON))
        start = min(port.end_time, start + REPLAN)
        calls += 1
        assert calls < 50


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_random_cartesian_paths_long_horizon(seed):
    """Seeded random waypoints, one Plan() with an hour's horizon: the window loop
    (path_timing_trajectory.cc:628-658) over a Cartesian path."""
    rng = np.random.default_rng(100 + seed)
    W = int(rng.integers(2, 6))
    joint = rng.uniform(-0.8, 0.8, (W, 7))
    port, ref = cart_pair(joint, N=int(rng.integers(200, 500)), delta=float(rng.uniform(0.004, 0.02)),
                          skip=bool(seed & 1), vmax=float(rng.uniform(0.5, 1.5)),
                          amax=float(rng.uniform(1.0, 3.0)))
    h = 3600 * 1000 * MS
    check_equal(port, ref, port.plan(0, h), ref.plan(0, h))
    assert port.target_reached


# ---------------------------------------------------------------- online re-planning
@pytest.mark.parametrize("skip", [False, True])
def test_switch_to_new_cartesian_waypoint_path(skip):
    """SwitchToNewCartesianWaypointPathWorks (_test.cc:764-920): plan until the motion is
    fast, switch all three splines at the stop parameter with the current velocity, plan
    to the end.  The joint and the translation spline are truncated through the shim's
    restatement of Eigen's QR in the reference and through the closed form in the port
    (knots bit for bit, control points 1e-12); the rotation spline's knot insertion is
    slerp on both sides."""
    joint = CART_WAYPOINTS
    port, ref = cart_pair(joint, skip=skip)
    start, tr = 0, None
    for _ in range(200):
        check_equal(port, ref, port.plan(start, HORIZON), ref.plan(start, HORIZON))
        tr = port.trajectory()
        start = min(port.end_time, max(start + REPLAN, port.start_time))
        v = np.max(np.abs(tr["r_vel"][0]))
        if v > 0.3 * 0.25:  # kFractionToStopAt of the smallest limit
            break
    assert not port.target_reached
    new_joint = np.array([[-1.0] * 7, [1.0] * 7])
    new_pose = np.stack([fake_fk_pose(q) for q in new_joint])
    st_r, par = ref.stop_parameter(start)
    assert st_r == 0
    v0 = np.array([np.interp(start / 1e9, tr["r_t"], tr["r_vel"][:, d]) for d in range(7)])
    assert port.switch(par, new_joint, initial_velocity=v0, pose_waypoints=new_pose) == 0
    assert ref.switch(par, new_joint, initial_velocity=v0, pose_waypoints=new_pose) == 0
    (kp, pp), (kr, pr) = port.spline(), ref.spline()
    same_bits(kp, kr, "knots after the switch")
    assert pp.shape == pr.shape
    np.testing.assert_allclose(pp, pr, rtol=0, atol=1e-12)
    poses_r = ref.pose_points()
    poses_p = port._pose_ctrl[:len(poses_r)]
    np.testing.assert_allclose(poses_p[:, :3], poses_r[:, :3], rtol=0, atol=1e-12)
    np.testing.assert_allclose(poses_p[:, 3:], poses_r[:, 3:], rtol=0, atol=1e-12)
    assert ref.state()["path_state"] == 2 and port.sampled == 2  # kModifiedPath
    for _ in range(400):
        st_p, st_r = port.plan(start, HORIZON), ref.plan(start, HORIZON)
        assert st_p == orc.ORC_OK and st_r == 0
        s = ref.state()
        assert (port.len, port.target_reached) == (s["len"], s["target_reached"])
        assert abs(port.M - s["M"]) <= 1 and abs(port.end_time - s["end_time"]) <= 4 * MS
        tp, trr = port.trajectory(), ref.trajectory()
        m = min(len(tp["r_t"]), len(trr["r_t"])) - 1
        np.testing.assert_allclose(tp["r_t"][:m], trr["r_t"][:m], rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(tp["r_pos"][:m], trr["r_pos"][:m], rtol=0, atol=1e-8)
        start = min(port.end_time, max(start + REPLAN, port.start_time))
        if port.target_reached:
            break
    assert port.target_reached
    np.testing.assert_allclose(tp["r_pos"][-1], new_joint[-1], atol=1e-9)


def test_cartesian_switch_geometry_random():
    """Seeded switches on fresh fits: truncation points inside the path, two to five new
    waypoints, repeated switches until the knot capacity is exhausted -- status and knots bit
    for bit, control points of the joint and the translation spline within 1e-11 (QR in the
    reference, closed form in the port).  The rotation spline agrees to 1e-9 except where the
    reference's own end-point correction is decided by rounding noise:
    BSplineQ::ExtendWithControlPoints negates the whole logarithm of a relative rotation when
    its real part -- 0.5 log |q|^2 / B, i.e. +-1e-16 for a unit quaternion -- is negative
    (bsplineq.cc:573-577), which turns the correction into its inverse.  Port and reference
    then pick the two alternatives whenever their inputs differ in the last bit (they do
    after the first switch, through the translation spline); the test recognises that case
    (relative rotations inverse to each other) and ends the trial there."""
    rng = np.random.default_rng(11)
    checked = flipped = 0
    for trial in range(12):
        W = int(rng.integers(2, 5))
        joint = rng.uniform(-0.8, 0.8, (W, 7))
        port, ref = cart_pair(joint, N=100, delta=0.01)
        for rep in range(8):
            kn, _ = ref.spline()
            keep = float(rng.uniform(0.15, 0.9) * kn[-1])
            Wn = int(rng.integers(2, 6))
            nj = rng.uniform(-0.8, 0.8, (Wn, 7))
            npz = np.stack([fake_fk_pose(q) for q in nj])
            st_p = port.switch(keep, nj, pose_waypoints=npz)
            st_r = ref.switch(keep, nj, pose_waypoints=npz)
            assert st_p == st_r, (trial, rep, st_p, st_r)
            if st_p != 0:
                assert st_p == 9  # kFailedPrecondition: capacity
                break
            (kp, pp), (kr, pr) = port.spline(), ref.spline()
            same_bits(kp, kr, ("knots", trial, rep))
            np.testing.assert_allclose(pp, pr, rtol=0, atol=1e-11)
            poses_r = ref.pose_points()
            poses_p = port._pose_ctrl[:len(poses_r)]
            assert len(poses_r) == len(pr)
            np.testing.assert_allclose(poses_p[:, :3], poses_r[:, :3], rtol=0, atol=1e-11)
            np.testing.assert_allclose(np.linalg.norm(poses_r[:, 3:], axis=1), 1.0, atol=1e-9)
            bad = np.flatnonzero(np.abs(poses_p[:, 3:] - poses_r[:, 3:]).max(axis=1) > 1e-9)
            checked += 1
            if bad.size:
                r = bad[0]  # the corrected end point of the kept part
                assert bad.size == 1 and r >= 1
                inv = lambda q: np.array([-q[0], -q[1], -q[2], q[3]])
                rel_p = orc.quat_mul(inv(poses_p[r - 1, 3:]), poses_p[r, 3:])
                rel_r = orc.quat_mul(inv(poses_r[r - 1, 3:]), poses_r[r, 3:])
                both = orc.quat_mul(rel_p, rel_r)
                assert np.max(np.abs(both[:3])) <= 1e-8 and abs(abs(both[3]) - 1.0) <= 1e-8
                flipped += 1
                break
    assert checked >= 30 and flipped <= checked // 4
