"""GPU parity: dense B-spline evaluation with derivatives (BASELINE config 3)
through tpb_spline_eval_* vs the oracle's BSplineT::EvalCurveAndDerivatives
restatement (splines/bspline.h:539-568, bspline_base.cc:218-348).

exact = 1: NURBS-book A2.3 in the reference's operation order -> bit-identical.
exact = 0: per-span polynomial form -> within BASELINE.json's 1e-12 absolute.
"""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-12


def _oracle(po, knots, ctrl, P, nder, u):
    S = ctrl.shape[0]
    out = np.zeros((S, u.shape[1], nder + 1, ctrl.shape[2]))
    st = np.zeros((S, u.shape[1]), np.int32)
    for s in range(S):
        inside = (u[s] >= knots[s, 0]) & (u[s] <= knots[s, -1])
        st[s] = np.where(inside, 0, 11)
        _, vals = po.eval_curve_ders(knots[s], ctrl[s], P, nder, u[s][inside])
        out[s][inside] = vals
    return out, st


@pytest.mark.parametrize("P", [1, 2, 3, 4, 5])
@pytest.mark.parametrize("D", [1, 6, 7])
def test_exact_mode_is_bit_identical(pkg, po, P, D):
    nder = min(P, 2)
    knots, ctrl = pkg.workloads.dense_splines(5, P, D, num_points=12, knot_scale=3.0)
    rng = np.random.default_rng(P * 10 + D)
    u = np.sort(rng.uniform(0.0, 3.0, (5, 700)), axis=1)
    u[:, 0] = 0.0          # first knot
    u[:, -1] = 3.0         # last knot: span = num_points - 1 (bspline_base.cc:223-225)
    u[:, 5] = knots[0, P + 2]  # exactly on an interior knot
    out, st = pkg.spline_eval_host(knots, ctrl, P, nder, u=u, exact=True)
    ref, rst = _oracle(po, knots, ctrl, P, nder, u)
    assert (st == 0).all() and (rst == 0).all()
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("P,nder", [(3, 3), (4, 3), (5, 3), (3, 0), (2, 1)])
def test_exact_mode_other_derivative_orders(pkg, po, P, nder):
    knots, ctrl = pkg.workloads.dense_splines(3, P, 7, num_points=20, knot_scale=8.0)
    u = np.tile(np.linspace(0.0, 8.0, 513), (3, 1))
    out, st = pkg.spline_eval_host(knots, ctrl, P, nder, u=u, exact=True)
    ref, _ = _oracle(po, knots, ctrl, P, nder, u)
    assert (st == 0).all()
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("P", [1, 2, 3, 4, 5])
@pytest.mark.parametrize("D", [6, 7])
def test_polynomial_mode_within_tolerance(pkg, po, P, D):
    """cfg3 shape (32 control points, knots on [0, 8]); uniform grid, shuffled
    parameters and the implicit grid u0 + i*du."""
    nder = min(P, 2)
    knots, ctrl = pkg.workloads.dense_splines(4, P, D, num_points=32, knot_scale=8.0)
    E = 4099  # not a multiple of the tile
    rng = np.random.default_rng(P + D)
    u = np.tile(np.linspace(0.0, 8.0, E), (4, 1))
    u[1] = rng.permutation(u[1])
    u[2] = np.sort(rng.uniform(0, 8, E))
    out, st = pkg.spline_eval_host(knots, ctrl, P, nder, u=u, exact=False)
    ref, _ = _oracle(po, knots, ctrl, P, nder, u)
    assert (st == 0).all()
    for k in range(nder + 1):
        err = np.max(np.abs(out[:, :, k] - ref[:, :, k]))
        assert err <= TOL, (P, D, k, err)
    # implicit grid (no parameter array read): same parameters as u[0] up to an ulp
    out2, st2 = pkg.spline_eval_host(knots, ctrl, P, nder, u0=np.zeros(4),
                                     du=np.full(4, 8.0 / (E - 1)), evals=E, exact=False)
    ug = np.tile(0.0 + np.arange(E) * (8.0 / (E - 1)), (4, 1))
    ug = np.minimum(ug, 8.0)
    ref2, rst2 = _oracle(po, knots, ctrl, P, nder, ug)
    inside = rst2 == 0
    assert np.array_equal(st2 == 0, inside)
    assert np.max(np.abs(out2[inside] - ref2[inside])) <= TOL


def test_mathematica_golden_both_modes(pkg):
    """splines/bspline_test.cc:97-727 (degree 3, 101 samples, 3 derivatives; the
    reference's tolerance is a squared-error sum <= 5e-14, :70-75)."""
    g = json.load(open(os.path.join(HERE, "golden", "bspline_mathematica.json")))
    knots = np.array([g["knots"]], float)
    ctrl = np.array([g["control_points"]], float)
    us = (0.0 + np.arange(101) * (1.0 / 100))[None, :]
    for exact in (True, False):
        out, st = pkg.spline_eval_host(knots, ctrl, 3, 3, u=us, exact=exact)
        assert (st == 0).all()
        v = out[0]
        for k, (nx, ny) in enumerate((("X", "Y"), ("Xp", "Yp"), ("Xpp", "Ypp"), ("Xppp", "Yppp"))):
            assert np.sum((v[:, k, 0] - g[nx]) ** 2) <= 5e-14
            assert np.sum((v[:, k, 1] - g[ny]) ** 2) <= 5e-14


@pytest.mark.parametrize("exact", [True, False])
def test_out_of_range_rows_are_untouched(pkg, exact):
    """u outside [knots.front(), knots.back()] -> kOutOfRange (11), the output row
    keeps the caller's contents (bspline.h:549-552); NaN fails the check too."""
    knots, ctrl = pkg.workloads.dense_splines(2, 3, 7, num_points=10, knot_scale=2.0)
    u = np.tile(np.linspace(-0.5, 2.5, 300), (2, 1))
    u[1, 17] = np.nan
    sentinel = np.full((2, 300, 3, 7), -77.0)
    out, st = pkg.spline_eval_host(knots, ctrl, 3, 2, u=u, exact=exact, out=sentinel.copy())
    bad = ~((u >= 0.0) & (u <= 2.0))
    assert np.array_equal(st == 11, bad) and np.array_equal(st == 0, ~bad)
    assert (out[bad] == -77.0).all()
    assert (out[~bad] != -77.0).any()


def test_argument_errors(pkg):
    knots, ctrl = pkg.workloads.dense_splines(1, 3, 7, num_points=10, knot_scale=2.0)
    u = np.zeros((1, 4))
    with pytest.raises(pkg.TpbError) as ei:
        pkg.spline_eval_host(knots, ctrl, 6, 2, u=u)  # degree > 5
    assert ei.value.code == 11
    with pytest.raises(pkg.TpbError) as ei:
        pkg.spline_eval_host(knots, ctrl, 2, 3, u=u)  # CHECK(der <= p), bspline_base.cc:270
    assert ei.value.code == 3


@pytest.mark.parametrize("P,D,E", [(3, 7, 4096), (3, 7, 4099), (5, 6, 6000), (2, 3, 2050), (4, 7, 777)])
def test_bulk_copy_tiles_equal_thread_stores(pkg, P, D, E):
    """The polynomial-form kernel writes its tiles with bulk asynchronous copies
    (exact = 0) or with per-thread stores (exact = -1): the same bits, also for ragged
    last tiles, odd tile sizes (16-byte misalignment -> fallback path), rows that fail the
    range check (those tiles take the per-row path and leave the caller's memory alone) and
    an explicit shuffled parameter array."""
    nder = min(P, 2)
    S = 5
    knots, ctrl = pkg.workloads.dense_splines(S, P, D, num_points=32, knot_scale=8.0)
    rng = np.random.default_rng(E)
    u = np.tile(np.linspace(0.0, 8.0, E), (S, 1))
    u[1] = rng.permutation(u[1])
    u[2, 300:310] = 9.5   # out of range
    u[3, -1] = -0.25
    a, sa = pkg.spline_eval_host(knots, ctrl, P, nder, u=u, exact=0)
    b, sb = pkg.spline_eval_host(knots, ctrl, P, nder, u=u, exact=-1)
    assert np.array_equal(sa, sb) and (sa != 0).sum() == 11
    assert a.tobytes() == b.tobytes()
    a2, _ = pkg.spline_eval_host(knots, ctrl, P, nder, u0=np.zeros(S), du=np.full(S, 8.0 / E),
                                 evals=E, exact=0)
    b2, _ = pkg.spline_eval_host(knots, ctrl, P, nder, u0=np.zeros(S), du=np.full(S, 8.0 / E),
                                 evals=E, exact=-1)
    assert a2.tobytes() == b2.tobytes()
