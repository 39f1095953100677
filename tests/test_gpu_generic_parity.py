"""GPU parity: explicit-constraint problems (the TimeOptimalPathProfile API)
through the C ABI vs the CPU oracle, on the reference's own test scenarios
(time_optimal_path_timing_test.cc:440-701) and seeded random problems."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

KEYS = (("TIME", "time"), ("S", "s"), ("SD", "sd"), ("SDD", "sdd"), ("SD2", "sd2"),
        ("B_SD2_MAX", "b_sd2_max"), ("B_SDD_MAX", "b_sdd_max"), ("B_SDD_MIN", "b_sdd_min"),
        ("B_SD2_ZERO", "b_sd2_zero"), ("B_TYPE", "b_type"))


def _solve_batch(pkg, cons, s0, s1, sd0=None, sdd0=None, t0=None, max_loops=100, variants=()):
    B, N, C, _ = cons.shape
    eng = pkg.Engine(max_paths=B, num_samples=N, max_constraints=C)
    for v in variants:  # cross-check kernels (tpb_set_kernel_variant)
        eng.set_kernel_variant(v)
    eng.set_constraints(cons, s0, s1)
    eng.optimize(sd0, sdd0, t0, max_loops=max_loops)
    A = pkg.ARR
    res = {k: eng.get(getattr(A, k)) for k, _ in KEYS}
    for k in ("STATUS", "DURATION", "LAST_EXTREMAL"):
        res[k] = eng.get(getattr(A, k))
    eng.close()
    return res


def _compare(po, res, cons, s0, s1, sd0=None, sdd0=None, t0=None, max_loops=100,
             expect_all_ok=False):
    B = cons.shape[0]
    for b in range(B):
        o = po.solve(cons[b], s0[b], s1[b], 0.0 if sd0 is None else sd0[b],
                     0.0 if sdd0 is None else sdd0[b], 0.0 if t0 is None else t0[b],
                     max_loops=max_loops)
        assert (res["STATUS"][b] == 0) == o["ok"], (b, res["STATUS"][b], o["status"])
        assert res["STATUS"][b] == o["status"], (b, res["STATUS"][b], o["status"])
        if expect_all_ok:
            assert o["ok"]
        if not o["ok"]:
            continue
        assert res["LAST_EXTREMAL"][b] == o["last_extremal_index"]
        assert abs(res["DURATION"][b] - (o["time"][-1] - o["time"][0])) <= 1e-6 * abs(o["time"][-1] - o["time"][0]) + 1e-300
        for k, ok in KEYS:
            # integer/bit-exact: same operation order, no FMA contraction
            assert np.array_equal(res[k][b], o[ok], equal_nan=True), (b, k)
        assert po.count_violations(cons[b], res["SD2"][b], res["SDD"][b]) == \
            po.count_violations(cons[b], o["sd2"], o["sdd"])


def test_circle_scenarios(pkg, po):
    W = pkg.workloads
    for N in (50, 51, 1000, 2048):
        cons = np.stack([W.circle_constraints(N), W.circle_constraints(N)])
        s0, s1 = np.zeros(2), np.full(2, np.pi)
        sd0 = np.array([0.0, 0.1])
        res = _solve_batch(pkg, cons, s0, s1, sd0)
        _compare(po, res, cons, s0, s1, sd0, expect_all_ok=True)
    # golden duration of the reference itself (tests/golden, N=50, sd_start=0)
    cons = W.circle_constraints(50)[None]
    res = _solve_batch(pkg, cons, np.zeros(1), np.full(1, np.pi))
    assert abs(res["DURATION"][0] - 6.0003326777697756) <= 1e-6 * 6.0


def test_sine_scenarios(pkg, po):
    W = pkg.workloads
    for N in (30, 31, 100, 111):
        offs = np.array([0.0, np.pi / 2, np.pi / 4])
        cons = np.stack([W.sine_constraints(N, o, np.pi + o) for o in offs])
        res = _solve_batch(pkg, cons, offs, offs + np.pi)
        _compare(po, res, cons, offs, offs + np.pi, expect_all_ok=True)


def test_line_scalar_scenarios(pkg, po):
    W = pkg.workloads
    cons = W.line_constraints(30)[None]
    res = _solve_batch(pkg, cons, np.zeros(1), np.full(1, np.pi))
    _compare(po, res, cons, np.zeros(1), np.full(1, np.pi), expect_all_ok=True)
    cons = W.scalar_straight_constraints(30)[None]
    res = _solve_batch(pkg, cons, np.zeros(1), np.ones(1))
    _compare(po, res, cons, np.zeros(1), np.ones(1), expect_all_ok=True)
    cons = W.scalar_curved_constraints(100)[None]
    res = _solve_batch(pkg, cons, np.full(1, -3.0), np.ones(1))
    _compare(po, res, cons, np.full(1, -3.0), np.ones(1), expect_all_ok=True)


@pytest.mark.parametrize("N,C", [(64, 2), (200, 6), (333, 14), (128, 32), (2, 3), (3, 4),
                                 (300, 9), (150, 12), (100, 16), (90, 17), (257, 1),
                                 (80, 33), (60, 50), (48, 64)])
def test_random_problems(pkg, po, N, C):
    rng = np.random.default_rng(100 * N + C)
    B = 24
    cons = np.stack([pkg.workloads.random_constraints(rng, N, C) for _ in range(B)])
    s0 = rng.uniform(-1, 1, B)
    s1 = s0 + rng.uniform(0.5, 5, B)
    sd0 = rng.uniform(0, 0.3, B) * (rng.random(B) < 0.5)
    sdd0 = rng.uniform(-0.5, 0.5, B)
    t0 = rng.uniform(0, 10, B)
    res = _solve_batch(pkg, cons, s0, s1, sd0, sdd0, t0)
    _compare(po, res, cons, s0, s1, sd0, sdd0, t0)


def test_find_max_sd2_random_sizes_2_to_50(pkg, po):
    """FindMaxSd2Random (time_optimal_path_timing_test.cc:703-736): random LPs with 2..50
    rows, A, B ~ U(-100, 100), lower ~ U(-10, 0), upper ~ U(0, 10).  One LP per path (the
    same rows at both samples of an N = 2 problem, so the boundary passes leave the limit
    curve as computed): sd2_max and sd2_max_for_sdd0 equal the oracle's FindMaxSd2Simplex
    (== the reference's object code, tests/test_oracle_vs_ref.py) bit for bit and its brute
    force within 1e-8."""
    rng = np.random.default_rng(703)
    for C in list(range(2, 51, 3)) + [50]:
        B = 40
        rows = np.stack([rng.uniform(-100, 100, (B, C)), rng.uniform(-100, 100, (B, C)),
                         rng.uniform(-10, 0, (B, C)), rng.uniform(0, 10, (B, C))], -1)
        cons = np.ascontiguousarray(np.repeat(rows[:, None], 2, axis=1))  # [B][2][C][4]
        eng = pkg.Engine(max_paths=B, num_samples=2, max_constraints=C)
        eng.set_constraints(cons, np.zeros(B), np.ones(B))
        eng.optimize()
        sd2max = eng.get(pkg.ARR.B_SD2_MAX)[:, 0]
        sd2zero = eng.get(pkg.ARR.B_SD2_ZERO)[:, 0]
        eng.close()
        for b in range(B):
            a = po.find_max_sd2(rows[b], "port")
            assert (sd2max[b], sd2zero[b]) == (a[0], a[2]), (C, b)
            if po.have_ref():
                c = po.find_max_sd2(rows[b], "ref_brute")
                assert abs(sd2max[b] - c[0]) <= 1e-8


@pytest.mark.parametrize("C", [4, 7, 10, 16])
def test_tied_rows_register_and_local_limit_kernels(pkg, po, monkeypatch, C):
    """Duplicate, mirrored and scaled copies of rows put two or more rows on every vertex
    of the limit-curve LP (the tie branches of FindMaxSd2Simplex): oracle bits, and the
    register kernel (k_generic_limit_reg) against the local-memory one."""
    rng = np.random.default_rng(900 + C)
    B, N = 16, 120
    half = (C + 1) // 2
    base = np.stack([pkg.workloads.random_constraints(rng, N, half) for _ in range(B)])
    cons = np.empty((B, N, C, 4))
    cons[:, :, :half] = base
    for r in range(half, C):
        src = base[:, :, r - half]
        kind = (r - half) % 3
        if kind == 0:      # exact duplicate
            cons[:, :, r] = src
        elif kind == 1:    # mirrored: -A, -B, [-hi, -lo]
            cons[:, :, r, 0] = -src[..., 0]
            cons[:, :, r, 1] = -src[..., 1]
            cons[:, :, r, 2] = -src[..., 3]
            cons[:, :, r, 3] = -src[..., 2]
        else:              # scaled by 2 (exact in binary)
            cons[:, :, r] = 2.0 * src
    s0 = np.zeros(B)
    s1 = rng.uniform(0.5, 5, B)
    reg = _solve_batch(pkg, cons, s0, s1)
    _compare(po, reg, cons, s0, s1)
    loc = _solve_batch(pkg, cons, s0, s1, variants=("generic_limit_local",))
    for k in reg:
        assert reg[k].tobytes() == loc[k].tobytes(), k


def test_setup_failures_are_per_path(pkg, po):
    """SetupProblem failures (lower >= upper, s_start >= s_end, sd_start < 0)
    mark only the offending path (time_optimal_path_timing.cc:161-203, :554-576)."""
    N = 50
    cons = np.stack([pkg.workloads.circle_constraints(N) for _ in range(4)])
    cons[1, 10, 2, 2] = cons[1, 10, 2, 3]            # lower == upper
    s0, s1 = np.zeros(4), np.full(4, np.pi)
    s1[2] = 0.0                                       # s_start >= s_end
    sd0 = np.array([0.0, 0.0, 0.0, -0.1])             # negative start velocity
    res = _solve_batch(pkg, cons, s0, s1, sd0)
    assert list(res["STATUS"]) == [0, 1, 1, 1]
    assert np.isnan(res["DURATION"][1:]).all()
    _compare(po, res, cons, s0, s1, sd0)
