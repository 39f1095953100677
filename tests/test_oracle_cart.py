"""CPU pins of the Cartesian-path oracle (oracle/tp_oracle_cart.c) against the
golden values and identities the reference's own tests hold:
QuatExp Mathematica table (splines/bsplineq_test.cc:106-168), degree-1
quaternion spline == piecewise slerp (:309-344), AABB control points -> constant
rotation axis, monotone angle (:805-865), pose control polygon values
(splines/spline_utils_test.cc:30-137), fake kinematics consistency
(path_timing_trajectory_test.cc:548-587), ComputePathDerivatives
(timeable_path_cartesian_spline.cc:39-68)."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _wxyz(w, x, y, z):
    return np.array([x, y, z, w], float)


def _slerp(a, b, t):
    """Eigen's QuaternionBase::slerp (shortest arc)."""
    d = float(np.dot(a, b))
    ad = abs(d)
    if ad >= 1.0 - np.finfo(float).eps:
        s0, s1 = 1.0 - t, t
    else:
        th = np.arccos(ad)
        s0, s1 = np.sin((1.0 - t) * th) / np.sin(th), np.sin(t * th) / np.sin(th)
    if d < 0:
        s1 = -s1
    return s0 * a + s1 * b


def _same_rotation(a, b, tol):
    return min(np.max(np.abs(a - b)), np.max(np.abs(a + b))) <= tol


def test_quat_exp_mathematica_table(po):
    g = json.load(open(os.path.join(HERE, "golden", "quat_exp_mathematica.json")))
    assert g["order"] == "w,x,y,z" and len(g["inputs"]) == 15
    for qi, qo in zip(g["inputs"], g["outputs"]):
        got = po.quat_exp(_wxyz(*qi))
        want = _wxyz(*qo)
        # reference tolerance: IsApprox (relative 1e-12 in Eigen's sense)
        assert np.linalg.norm(got - want) <= 1e-12 * np.linalg.norm(want)
        # log inverts exp where the vector part's norm is below pi
        if np.linalg.norm(qi[1:]) < np.pi:
            back = po.quat_log(got)
            assert np.max(np.abs(back - _wxyz(*qi))) <= 1e-9


def test_quat_power_identities(po):
    rng = np.random.default_rng(3)
    for _ in range(50):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        assert _same_rotation(po.quat_power(q, 1.0), q, 1e-14)
        assert np.max(np.abs(po.quat_power(q, 0.0) - [0, 0, 0, 1])) <= 1e-15
        h = po.quat_power(q, 0.5)
        assert _same_rotation(po.quat_mul(h, h), q, 1e-14)


def test_linear_quaternion_spline_is_slerp(po):
    """BSplineQTest.LinearCaseWorks (bsplineq_test.cc:309-344)."""
    r = np.sqrt(0.5)
    pts = np.array([_wxyz(1, 0, 0, 0), _wxyz(0, 1, 0, 0), _wxyz(0, 0, 1, 0), _wxyz(0, 0, 0, 1),
                    _wxyz(r, r, 0, 0), _wxyz(r, 0, r, 0), _wxyz(r, 0, 0, r)])
    knots = po.make_uniform_knots(len(pts), 1)
    for i in range(1, len(knots) - 1):
        st, q = po.quat_eval_curve(knots, pts, 1, knots[i])
        assert st == 0 and _same_rotation(q, pts[i - 1], 1e-12)
        if i < len(pts) - 2:
            for t in (0.1, 0.3, 0.5, 0.7, 0.9):
                u = knots[i] + t * (knots[i + 1] - knots[i])
                st, q = po.quat_eval_curve(knots, pts, 1, u)
                assert st == 0 and _same_rotation(q, _slerp(pts[i - 1], pts[i], t), 1e-12)
    assert po.quat_eval_curve(knots, pts, 1, -0.1)[0] == 11   # kOutOfRange
    assert po.quat_eval_curve(knots, pts, 1, knots[-1] + 0.1)[0] == 11


@pytest.mark.parametrize("qa,qb", [
    ((.007, -.707, .707, .007), (.007, .707, -.707, .007)),
    ((.707, .707, 0, 0), (.707, 0, .707, 0)),
    ((.707, 0, .707, 0), (.707, .707, 0, 0)),
    ((.707, 0, .707, 0), (.707, 0, 0, .707))])
def test_aabb_control_points_rotate_about_one_axis(po, qa, qb):
    """BSplineQTest.SlerpInterpolationForAABBCase (bsplineq_test.cc:805-865)."""
    a, b = _wxyz(*qa), _wxyz(*qb)
    a, b = a / np.linalg.norm(a), b / np.linalg.norm(b)
    pts = np.array([a, a, b, b])
    knots = po.make_uniform_knots(4, 3)

    def diff_axis_angle(p, q):
        conj = np.array([-p[0], -p[1], -p[2], p[3]])
        d = po.quat_mul(conj, q)
        n = np.linalg.norm(d[:3])
        ang = 2 * np.arctan2(n, abs(d[3]))
        ax = d[:3] / (n if d[3] >= 0 else -n) if n > 0 else np.array([1.0, 0, 0])
        return ax, ang
    axis_all, ang_all = diff_axis_angle(a, b)
    prev = 0.0
    for s in np.arange(0.1, 1.0 + 1e-12, 0.1):
        st, q = po.quat_eval_curve(knots, pts, 3, min(s, 1.0))
        assert st == 0
        ax, ang = diff_axis_angle(a, q)
        assert ang >= prev - 1e-12
        assert np.max(np.abs(ax - axis_all)) <= 1e-6
        prev = ang
    assert abs(prev - ang_all) <= 1e-6


def test_pose_polyline_golden_values(po):
    ident = [0, 0, 0, 1.0]
    # OneCorner (:30-40)
    out = po.pose_polyline_to_bspline3(np.array([[1, 2, 3.0] + ident]), 0, 0)
    assert out.shape == (4, 7) and (out == [1, 2, 3, 0, 0, 0, 1]).all()
    # Translation (:42-66)
    c = np.array([[1, 0, 0.0] + ident, [2, 0, 0.0] + ident, [2, 1, 0.0] + ident])
    out = po.pose_polyline_to_bspline3(c, 0.1, 0.1)
    want = [[1, 0, 0], [1.1, 0, 0], [1.9, 0, 0], [2, 0, 0], [2, .1, 0], [2, .9, 0], [2, 1, 0]]
    assert np.max(np.abs(out[:, :3] - want)) <= 1e-15
    assert np.max(np.abs(out[:, 3:] - ident)) <= 1e-15
    # Rotation (:68-99)
    ax = np.array([1, 2, 3.0]) / np.sqrt(14.0)

    def aa(t):
        return np.r_[np.zeros(3), np.sin(t / 2) * ax, np.cos(t / 2)]
    out = po.pose_polyline_to_bspline3(np.array([aa(1.0), aa(2.0), aa(3.0)]), 0.1, 0.1)
    for k, t in enumerate([1.0, 1.1, 1.9, 2.0, 2.1, 2.9, 3.0]):
        assert np.max(np.abs(out[k] - aa(t))) <= 1e-15
    # RadiusOutOfBounds (:101-119): clamp at a quarter of the segment
    c = np.array([[1, 0, 0.0] + ident, [2, 0, 0.0] + ident])
    out = po.pose_polyline_to_bspline3(c, 0.6, 0.1)
    assert np.max(np.abs(out[:, 0] - [1, 1.25, 1.75, 2])) <= 1e-15
    # ZeroRadius (:121-137)
    out = po.pose_polyline_to_bspline3(c, 0.0, 0.1)
    assert (out[:, 0] == [1, 1, 2, 2]).all()


def test_fake_kinematics_consistency(po):
    rng = np.random.default_rng(9)
    for _ in range(30):
        v = rng.uniform(-1, 1, 3)
        q = po.exp_so3(v)
        assert abs(np.linalg.norm(q) - 1) <= 1e-15
        assert np.max(np.abs(po.log_so3(q) - v)) <= 1e-14
        # d LogSO3 / d quaternion coefficients against central differences
        J = po.log_so3_derivative(q)
        for c in range(4):
            e = np.zeros(4)
            e[c] = 1e-6
            fd = (po.log_so3(q + e) - po.log_so3(q - e)) / 2e-6
            assert np.max(np.abs(fd - J[:, c])) <= 1e-8
    # small-angle branches are continuous
    for s in (1e-5, 1e-7, 1e-9, 1e-12, 0.0):
        q = po.exp_so3(np.array([s, 0, 0]))
        assert np.max(np.abs(po.log_so3(q) - [s, 0, 0])) <= 1e-15
        J = po.log_so3_derivative(q)
        assert np.max(np.abs(J[:, :3] - 2 * np.eye(3))) <= 1e-9 and np.max(np.abs(J[:, 3])) <= 3 * s
    jac = po.fake_jacobian(np.array([.1, .2, .3, .4, -.5, .6, .7]))
    assert (jac[:3, :3] == np.eye(3)).all() and (jac[:3, 3:] == 0).all() and (jac[3:, :3] == 0).all()


def test_path_derivatives_and_rows(po):
    """ComputePathDerivatives: forward differences, zero at the ends
    (timeable_path_cartesian_spline.cc:39-68); ConstraintSetup rows (:551-595)."""
    rng = np.random.default_rng(4)
    N, D, delta = 50, 7, 0.005
    pos = np.cumsum(rng.normal(size=(N, D)) * 1e-3, axis=0)
    d1, d2 = po.cart_derivatives(pos, delta)
    inv = 1.0 / delta
    assert np.array_equal(d1[:-1], inv * (pos[1:] - pos[:-1])) and (d1[-1] == 0).all()
    assert np.array_equal(d2[1:-1], inv * (d1[2:] - d1[1:-1]))
    assert (d2[0] == 0).all() and (d2[-1] == 0).all()
    jac = np.stack([po.fake_jacobian(p) for p in pos])
    vmax, amax = np.full(D, 1.0), np.full(D, 2.0)
    cons = po.cart_constraints(d1, d2, jac, vmax, amax, 0.8, 0.5, 0.25)
    assert cons.shape == (N, 16, 4)
    assert np.array_equal(cons[:, :7, 0], d1) and np.array_equal(cons[:, :7, 1], d2)
    assert (cons[:, :7, 3] == 1.6).all() and (cons[:, :7, 2] == -1.6).all()
    assert np.array_equal(cons[:, 7:14, 1], d1 * d1) and (cons[:, 7:14, 3] == 0.8 ** 2).all()
    v = np.einsum("nrd,nd->nr", jac, d1)
    assert np.allclose(cons[:, 14, 1], (v[:, :3] ** 2).sum(1), rtol=1e-13, atol=0)
    assert np.allclose(cons[:, 15, 1], (v[:, 3:] ** 2).sum(1), rtol=1e-13, atol=1e-300)
    assert (cons[:, 14, 3] == 0.25).all() and (cons[:, 14, 2] == -0.25).all()
    assert (cons[:, 15, 3] == 0.0625).all() and (cons[:, 15, 2] == -0.0625).all()


def test_whole_cartesian_path_properties(po, pkg):
    """End at the last waypoint with zero velocity, limits respected
    (path_timing_trajectory_test.cc:589-700 state these for SimpleCartesianPath)."""
    pw, jw, vm, am = pkg.workloads.cart_paths(6, 5, seed=5300)
    for b in range(6):
        o = po.plan_cart_fake(pw[b], jw[b], vm[b], am[b], 1000, dt=0.004, M_cap=40000)
        assert o["ok"], o["status"]
        # knot scaling: 10 * max(2 * translation polygon length, 0.1) (sic, .cc:432-446)
        pc = o["pose_ctrl"]
        L = np.linalg.norm(np.diff(pc[:, :3], axis=0), axis=1).sum()
        assert abs(o["knot_end"] - 10.0 * max(2 * L, 0.1)) <= 1e-12 * o["knot_end"]
        assert np.max(np.abs(o["r_pos"][-1] - jw[b, -1])) <= 1e-12
        assert (o["r_vel"][-1] == 0).all() and (o["r_acc"][-1] == 0).all()
        assert np.max(np.abs(o["r_vel"])) <= 1.0 * 0.8 + 1e-6
        assert np.max(np.abs(o["r_acc"])) <= 2.0 + 1e-9
        # Cartesian speed limits: |J q'| sd <= limit
        sd = o["sd"]
        assert np.max(np.sqrt(o["cons"][:, 14, 1]) * sd) <= 0.5 + 1e-6
        assert np.max(np.sqrt(o["cons"][:, 15, 1]) * sd) <= 0.25 + 1e-6
