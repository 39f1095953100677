"""GPU parity: quaternion B-spline evaluation (BSplineQ::EvalCurve, bsplineq.cc:223-244, and
BSplineQ::EvalCurveAndDerivative, :246-307) through the C ABI (tpb_quat_spline_eval_*)
against the CPU oracle, which equals the reference's object code bit for bit
(tests/test_oracle_vs_ref_cart.py).  The device's log / atan2 / sin / cos / exp differ from
glibc's by rounding, so values are held to 1e-12 absolute (BASELINE.json's tolerance for
spline evaluation); statuses exactly."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12


def rand_quats(rng, S, n):
    q = rng.normal(size=(S, n, 4))
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    q[rng.random((S, n)) < 0.3] *= -1.0  # both signs of the same rotation
    return q


def smooth_quats(rng, po, S, n, step=0.3):
    """Control quaternions a moderate rotation apart (a path a robot would follow)."""
    out = np.zeros((S, n, 4))
    for s in range(S):
        q = rand_quats(rng, 1, 1)[0, 0]
        for i in range(n):
            out[s, i] = q
            q = po.quat_mul(q, po.exp_so3(step * rng.normal(size=3)))
    return out


@pytest.mark.parametrize("degree", [1, 2, 3, 4, 5])
def test_eval_curve_matches_oracle(pkg, po, degree):
    rng = np.random.default_rng(60 + degree)
    S, n, E = 3, degree + 6, 257
    knots = np.stack([po.make_uniform_knots(n, degree, 0.0, 1.0 + s) for s in range(S)])
    quats = smooth_quats(rng, po, S, n)
    quats[2] = rand_quats(rng, 1, n)[0]  # arbitrary control points too
    u = np.stack([np.concatenate([rng.uniform(0, 1.0 + s, E - 6), knots[s][[0, -1, degree + 1]],
                                  [-0.25, 1.0 + s + 1e-9, np.nan]]) for s in range(S)])
    q, dq, st = pkg.quat_spline_eval_host(knots, quats, degree, u=u)
    assert dq is None
    for s in range(S):
        for i in range(E):
            st_o, q_o = po.quat_eval_curve(knots[s], quats[s], degree, u[s, i]) \
                if np.isfinite(u[s, i]) else (11, None)
            assert st[s, i] == st_o, (s, i, u[s, i])
            if st_o == 0:
                assert np.max(np.abs(q[s, i] - q_o)) <= TOL, (s, i)
                assert abs(np.linalg.norm(q[s, i]) - 1.0) <= 1e-12 and q[s, i, 3] >= 0
            else:
                assert np.isnan(q[s, i]).all()  # the row is left untouched
    assert (st[:, -3:] == 11).all() and (st[:, :-3] == 0).all()


@pytest.mark.parametrize("degree", [1, 2, 3, 4, 5])
def test_eval_curve_and_derivative_matches_oracle(pkg, po, degree):
    rng = np.random.default_rng(80 + degree)
    S, n, E = 2, degree + 5, 200
    knots = np.stack([po.make_uniform_knots(n, degree, 0.0, 2.0) for _ in range(S)])
    quats = smooth_quats(rng, po, S, n)
    u = np.sort(rng.uniform(0, 2.0, (S, E)), axis=1)
    u[:, 0], u[:, -1] = 0.0, 2.0
    q, dq, st = pkg.quat_spline_eval_host(knots, quats, degree, u=u, derivative=True)
    assert (st == 0).all()
    for s in range(S):
        for i in range(E):
            st_o, q_o, dq_o = po.quat_eval_curve_derivative(knots[s], quats[s], degree, u[s, i])
            assert st_o == 0
            assert np.max(np.abs(q[s, i] - q_o)) <= TOL, (s, i)
            assert np.max(np.abs(dq[s, i] - dq_o)) <= 1e-11 * max(1.0, np.max(np.abs(dq_o))), (s, i)
    # the derivative is the derivative: central differences of the value
    h = 1e-6
    um = np.clip(u[:, 5:-5], h, 2.0 - h)
    qa, _, _ = pkg.quat_spline_eval_host(knots, quats, degree, u=um - h, derivative=True)
    qb, _, _ = pkg.quat_spline_eval_host(knots, quats, degree, u=um + h, derivative=True)
    _, dm, _ = pkg.quat_spline_eval_host(knots, quats, degree, u=um, derivative=True)
    scale = max(1.0, np.max(np.abs(dm)))
    assert np.max(np.abs((qb - qa) / (2 * h) - dm)) <= 1e-5 * scale


def test_grid_mode_device_pointers_and_large_batch(pkg, po):
    """Implicit parameter grid on device tensors, 2,048 splines x 1,024 parameters (the
    rotation splines of cfg5-sized Cartesian paths): spot-checked against the oracle."""
    import torch
    rng = np.random.default_rng(7)
    S, n, E, degree = 2048, 13, 1024, 2
    knots = np.tile(po.make_uniform_knots(n, degree, 0.0, 1.0), (S, 1)) * rng.uniform(0.5, 3.0, (S, 1))
    quats = rand_quats(rng, S, n)
    u0 = np.zeros(S)
    du = knots[:, -1] / (E - 1) * (1.0 - 1e-12)
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    d_k, d_q, d_u0, d_du = t(knots), t(quats), t(u0), t(du)
    out = torch.full((S, E, 4), float("nan"), dtype=torch.float64, device=dev)
    dout = torch.full((S, E, 4), float("nan"), dtype=torch.float64, device=dev)
    status = torch.full((S, E), -1, dtype=torch.int32, device=dev)
    pkg.quat_spline_eval_device(d_k, d_q, degree, out, dout=dout, u0=d_u0, du=d_du, evals=E,
                                status=status)
    torch.cuda.synchronize()
    q, dq, st = out.cpu().numpy(), dout.cpu().numpy(), status.cpu().numpy()
    assert (st == 0).all() and np.isfinite(q).all() and np.isfinite(dq).all()
    for s, i in zip(rng.integers(0, S, 200), rng.integers(0, E, 200)):
        _, q_o, dq_o = po.quat_eval_curve_derivative(knots[s], quats[s], degree, u0[s] + i * du[s])
        assert np.max(np.abs(q[s, i] - q_o)) <= TOL
        assert np.max(np.abs(dq[s, i] - dq_o)) <= 1e-11 * max(1.0, np.max(np.abs(dq_o)))


def test_argument_checks(pkg):
    knots = np.zeros((1, 10))
    quats = np.zeros((1, 3, 4))
    with pytest.raises(pkg.TpbError) as ei:  # degree 6
        pkg.quat_spline_eval_host(knots, quats, 6, u=np.zeros((1, 4)))
    assert ei.value.code == 11
    with pytest.raises(pkg.TpbError) as ei:  # fewer control points than degree + 1
        pkg.quat_spline_eval_host(knots, quats, 3, u=np.zeros((1, 4)))
    assert ei.value.code == 3
