"""CPU tests: the stage-C restatement (oracle/tp_oracle.c) must reproduce the
reference's own time_optimal_path_timing.cc (compiled verbatim into
oracle/_ref/libtp_ref.so) bit for bit.  Skipped where oracle/_ref is absent
(it is built wherever /root/reference exists and travels with the repo)."""
import numpy as np
import pytest

KEYS = ("time", "s", "sd", "sdd", "sd2", "b_sd2_max", "b_sdd_max", "b_sdd_min",
        "b_sd2_zero", "b_type")


@pytest.fixture(scope="module")
def ref_available(po):
    if not po.have_ref():
        pytest.skip("oracle/_ref/libtp_ref.so not built (no /root/reference here)")
    return True


def _same(po, cons, s0, s1, sd0=0.0, sdd0=0.0, t0=0.0, max_loops=100):
    a = po.solve(cons, s0, s1, sd0, sdd0, t0, max_loops, impl="port")
    b = po.solve(cons, s0, s1, sd0, sdd0, t0, max_loops, impl="ref")
    assert a["ok"] == b["ok"]
    if a["ok"]:
        assert a["last_extremal_index"] == b["last_extremal_index"]
        for k in KEYS:
            assert np.array_equal(a[k], b[k], equal_nan=True), k
    return a["ok"]


def test_reference_scenarios_bitwise(po, pkg, ref_available):
    W = pkg.workloads
    for N in (50, 51, 1000, 2048):
        for sd0 in (0.0, 0.1):
            assert _same(po, W.circle_constraints(N), 0.0, np.pi, sd0)
    for off in (0.0, np.pi / 2, np.pi / 4):
        for N in (30, 31, 100, 111):
            assert _same(po, W.sine_constraints(N, off, np.pi + off), off, np.pi + off)
    assert _same(po, W.line_constraints(30), 0.0, np.pi)
    assert _same(po, W.scalar_straight_constraints(30), 0.0, 1.0)
    assert _same(po, W.scalar_curved_constraints(100), -3.0, 1.0)


def test_random_problems_bitwise(po, pkg, ref_available):
    rng = np.random.default_rng(2024)
    n_ok = 0
    for trial in range(150):
        N = int(rng.integers(2, 400))
        C = int(rng.integers(1, 33))
        cons = pkg.workloads.random_constraints(rng, N, C)
        s0 = float(rng.uniform(-1, 1))
        n_ok += _same(po, cons, s0, s0 + float(rng.uniform(0.3, 6)),
                      float(rng.uniform(0, 0.4)) * (trial % 2), float(rng.uniform(-1, 1)),
                      float(rng.uniform(0, 5)), max_loops=int(rng.choice([3, 100, 4000])))
    assert n_ok > 50


def test_lp_random_vs_reference_and_bruteforce(po, ref_available):
    """FindMaxSd2Random, time_optimal_path_timing_test.cc:703-736 (sizes 2-50,
    A,B ~ U(-100,100), lower ~ U(-10,0), upper ~ U(0,10)): port == reference simplex
    exactly, and simplex == brute force within 1e-8."""
    rng = np.random.default_rng(7)
    for _ in range(3000):
        C = int(rng.integers(2, 51))
        rows = np.stack([rng.uniform(-100, 100, C), rng.uniform(-100, 100, C),
                         rng.uniform(-10, 0, C), rng.uniform(0, 10, C)], 1)
        a = po.find_max_sd2(rows, "port")
        b = po.find_max_sd2(rows, "ref_simplex")
        c = po.find_max_sd2(rows, "ref_brute")
        np.testing.assert_array_equal(a, b)
        np.testing.assert_allclose(a, c, atol=1e-8, rtol=0)


def test_joint_pipeline_bitwise(po, pkg, ref_available):
    wp, vmax, amax = pkg.workloads.joint_paths(12, 10, 7, seed=1000)
    for b in range(12):
        x = po.plan_joint(wp[b], vmax[b], amax[b], 1024, dt=0.004, M_cap=16384, impl="port")
        y = po.plan_joint(wp[b], vmax[b], amax[b], 1024, dt=0.004, M_cap=16384, impl="ref")
        assert x["ok"] and y["ok"]
        for k in ("time", "sd", "sdd", "r_pos", "r_vel", "r_acc"):
            np.testing.assert_array_equal(x[k], y[k])


def test_query_matches_reference(po, pkg, ref_available):
    import ctypes as C
    cons = pkg.workloads.circle_constraints(51)
    r = po.solve(cons, 0.0, np.pi, 0.1)
    tq = np.concatenate([np.linspace(-0.5, r["time"][-1] + 0.5, 97), r["time"][:10]])
    s, sd, sdd = np.zeros_like(tq), np.zeros_like(tq), np.zeros_like(tq)
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    ok = po.ref().tpref_query(51, 4, dp(np.ascontiguousarray(cons)), 0.0, np.pi, 0.1, 0.0, 0.0,
                              len(tq), dp(tq), dp(s), dp(sd), dp(sdd))
    assert ok == 1
    for i, t in enumerate(tq):
        o = po.query(r, 0.0, np.pi, t)
        assert o[0] == 1 and (o[1], o[2], o[3]) == (s[i], sd[i], sdd[i]), (i, t)
