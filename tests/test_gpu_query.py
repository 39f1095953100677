"""tpb_query_profile: TimeOptimalPathProfile::GetPathParameterAndDerivatives,
SampleIndexFromTime, GetPreviousIndex and GetPreviousDiscreteValues
(time_optimal_path_timing.cc:1497-1668) batched on the device, against the CPU oracle
(orc_query == the reference's object code, tests/test_oracle_vs_ref.py:83-95), bit for
bit, over grids of times that include t < t0, t > t_end, the time stamps themselves and
their neighbours one ulp away."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu


def _times(time, rng, extra=24):
    t0, t1 = time[0], time[-1]
    stamps = time[rng.integers(0, len(time), 12)]
    return np.concatenate([
        [t0 - 1.0, t0 - 1e-9, t0, np.nextafter(t0, 9e9), t1, np.nextafter(t1, -9e9),
         np.nextafter(t1, 9e9), t1 + 3.0],
        stamps, np.nextafter(stamps, 9e9), np.nextafter(stamps, -9e9),
        rng.uniform(t0, t1, extra)])


def _previous_index(time, t):
    """GetPreviousIndex (:1628-1645) with SampleIndexFromTime (:1497-1524)."""
    N = len(time)
    if t < time[0]:
        return -1
    if t > time[-1]:
        return N - 1
    if t <= time[0]:
        return 0
    if t >= time[-1]:
        return N - 2
    k = int(np.searchsorted(time, t, side="right")) - 1  # time[k] <= t < time[k+1]
    while k < N - 1 and time[k] == time[k + 1]:
        k += 1
    return k


def _check(po, got, b, tq, sol, s_start, s_end):
    for i, t in enumerate(tq):
        ok, s, sd, sdd = po.query(sol, s_start, s_end, t)
        assert (got["status"][b, i] & 1) == (0 if ok else 1), (b, i, t)
        if ok:
            assert (got["s"][b, i], got["sd"][b, i], got["sdd"][b, i]) == (s, sd, sdd), (b, i, t)
        k = _previous_index(sol["time"], t)
        assert got["previous_index"][b, i] == k, (b, i, t, k)
        assert ((got["status"][b, i] >> 1) & 1) == (1 if k < 0 else 0)
        if k >= 0:
            assert got["sk"][b, i] == sol["s"][k] and got["sdk"][b, i] == sol["sd"][k]
            assert got["sddk"][b, i] == sol["sdd"][k] and got["tk"][b, i] == sol["time"][k]
        else:
            assert np.isnan(got["sk"][b, i])


def test_queries_on_explicit_constraint_profiles(pkg, po):
    """The reference's Circle scenario (time_optimal_path_timing_test.cc:125-158) with and
    without an initial velocity, a shifted start time, and random problems of which some
    fail: queries of a failed path all report status 3."""
    W = pkg.workloads
    rng = np.random.default_rng(5)
    N, C = 51, 4
    cons = np.stack([W.circle_constraints(N), W.circle_constraints(N)] +
                    [W.random_constraints(rng, N, C) for _ in range(6)])
    B = cons.shape[0]
    s0 = np.zeros(B)
    s1 = np.concatenate([[np.pi, np.pi], rng.uniform(0.5, 5, B - 2)])
    sd0 = np.concatenate([[0.0, 0.1], np.zeros(B - 2)])
    t0 = np.concatenate([[0.0, 42.5], np.zeros(B - 2)])
    eng = pkg.Engine(max_paths=B, num_samples=N, max_constraints=C)
    eng.set_constraints(cons, s0, s1)
    eng.optimize(sd0, None, t0)
    status = eng.get(pkg.ARR.STATUS)
    sols = [po.solve(cons[b], s0[b], s1[b], sd0[b], 0.0, t0[b]) for b in range(B)]
    Q = 68
    tq = np.zeros((B, Q))
    for b in range(B):
        base = sols[b]["time"] if sols[b]["ok"] else np.linspace(0.0, 1.0, N)
        tq[b] = _times(base, rng)
    got = eng.query_profile(tq)
    assert (status == 0).sum() >= 3
    for b in range(B):
        assert (status[b] == 0) == sols[b]["ok"]
        if status[b] != 0:
            assert (got["status"][b] == 3).all() and (got["previous_index"][b] == -1).all()
            assert np.isnan(got["s"][b]).all()
            continue
        _check(po, got, b, tq[b], sols[b], s0[b], s1[b])
    eng.close()


def test_queries_on_joint_path_profiles(pkg, po):
    """Joint paths (parameter range from the path's sampling distance), including a
    one-waypoint path and a path sampled beyond its end."""
    B, W, D, N = 9, 6, 7, 300
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=31)
    nw = np.full(B, W, np.int32)
    nw[3] = 1
    nw[5] = 2
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W)
    eng.set_joint_waypoints(wp, nw)
    eng.set_joint_limits(vmax, amax)
    eng.sample_paths()
    eng.optimize()
    A = pkg.ARR
    arr = {k: eng.get(getattr(A, k.upper())) for k in ("time", "s", "sd", "sdd", "sd2")}
    status, delta = eng.get(A.STATUS), eng.get(A.DELTA)
    rng = np.random.default_rng(6)
    Q = 68
    tq = np.stack([_times(arr["time"][b] if status[b] == 0 else np.linspace(0, 1, N), rng)
                   for b in range(B)])
    got = eng.query_profile(tq)
    assert (status == 0).sum() >= B - 2
    for b in range(B):
        if status[b] != 0:
            assert (got["status"][b] == 3).all()
            continue
        sol = {k: arr[k][b] for k in arr}
        _check(po, got, b, tq[b], sol, 0.0, 0.0 + delta[b] * (N - 1))
    # before the profiles exist the reference returns false
    fresh = pkg.Engine(max_paths=2, num_samples=N, num_dofs=D, max_waypoints=W)
    fresh.set_joint_waypoints(wp[:2])
    fresh.set_joint_limits(vmax[:2], amax[:2])
    with pytest.raises(pkg.TpbError) as ei:
        fresh.query_profile(np.zeros((2, 4)))
    assert ei.value.code == 9
    fresh.close()
    eng.close()


def test_solution_satisfies_constraints_on_the_device(pkg, po):
    """tpb_count_violations = TimeOptimalPathProfile::SolutionSatisfiesConstraints
    (time_optimal_path_timing.cc:492-518) for a batch: explicit rows (with a path that has
    no solution) and joint paths (rows synthesised from q', q''), against the oracle's
    count (the reference's loop restated)."""
    rng = np.random.default_rng(5)
    B, N, C = 12, 120, 6
    cons = np.stack([pkg.workloads.random_constraints(rng, N, C) for _ in range(B)])
    s0 = np.zeros(B)
    s1 = rng.uniform(0.5, 4.0, B)
    s1[3] = -1.0  # setup failure: no valid solution for this path
    eng = pkg.Engine(max_paths=B, num_samples=N, max_constraints=C)
    eng.set_constraints(cons, s0, s1)
    eng.optimize()
    got = eng.count_violations()
    st, sd2, sdd = eng.get(pkg.ARR.STATUS), eng.get(pkg.ARR.SD2), eng.get(pkg.ARR.SDD)
    for b in range(B):
        if st[b] != 0:
            assert got[b] == -1
        else:
            assert got[b] == po.count_violations(cons[b], sd2[b], sdd[b]), b
    assert got[3] == -1 and (got[st == 0] >= 0).all()
    eng.close()
    # joint paths: the rows ConstraintSetup would build; tighter limits -> violations
    Bj, W, D, Nj = 6, 6, 7, 256
    wp, vm, am = pkg.workloads.joint_paths(Bj, W, D, seed=77)
    eng = pkg.Engine(max_paths=Bj, num_samples=Nj, num_dofs=D, max_waypoints=W)
    eng.set_joint_limits(vm, am)
    eng.set_joint_waypoints(wp)
    eng.sample_paths()
    eng.optimize()
    got = eng.count_violations()
    dq, ddq = eng.get(pkg.ARR.DQ), eng.get(pkg.ARR.DDQ)
    sd2, sdd = eng.get(pkg.ARR.SD2), eng.get(pkg.ARR.SDD)
    for b in range(Bj):
        rows = po.joint_constraints(dq[b], ddq[b], vm[b], am[b], 0.8)
        assert got[b] == po.count_violations(rows, sd2[b], sdd[b]) == 0
    # new limits invalidate the solution: the audit refuses like the reference does
    eng.set_joint_limits(0.5 * vm, 0.5 * am)
    with pytest.raises(pkg.TpbError) as ei:
        eng.count_violations()
    assert ei.value.code == 9
    eng.close()
