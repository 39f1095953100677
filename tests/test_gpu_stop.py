"""Batched RescaleTrajectoryBackwardToStop on the GPU (tpb_rescale_to_stop_*)
against the CPU oracle (bit for bit), the committed golden vectors written from the
reference's own object code, and the properties of rescale_to_stop_test.cc."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import pyoracle as orc  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
import test_oracle_stop as T  # noqa: E402

pkg = load_package()
pytestmark = pytest.mark.gpu


def pack(cases):
    """[(amax, t, pos, vel, acc)] with a common D -> packed arrays."""
    D = cases[0][3].shape[1]
    begin, length, row = [], [], 0
    for c in cases:
        begin.append(row)
        length.append(len(c[1]))
        row += len(c[1]) + 3  # gaps between trajectories must stay untouched
    t = np.zeros(row)
    v = np.zeros((row, D))
    a = np.zeros((row, D))
    for (amax, tt, pos, vel, acc), b in zip(cases, begin):
        t[b:b + len(tt)] = tt
        v[b:b + len(tt)] = vel
        a[b:b + len(tt)] = acc
    amax = np.stack([c[0] for c in cases])
    return amax, np.array(begin, np.int64), np.array(length, np.int32), t, v, a


def check_against_oracle(cases):
    amax, begin, length, t, v, a = pack(cases)
    out = pkg.rescale_to_stop_host(amax, begin, length, t, v, a)
    touched = np.zeros(len(t), bool)
    for b, (am, tt, pos, vel, acc) in enumerate(cases):
        r = orc.rescale_to_stop(am, tt, pos, vel, acc)
        assert out["status"][b] == (0 if r["status"] == 0 else 1)
        assert out["count"][b] == r["count"]
        assert out["switch_index"][b] == r["switch_index"]
        lo = begin[b] + r["switch_index"]
        hi = lo + r["count"]
        touched[lo:hi] = True
        assert out["times"][lo:hi].tobytes() == r["times"].tobytes()
        assert out["velocities"][lo:hi].tobytes() == r["velocities"].tobytes()
        assert out["accelerations"][lo:hi].tobytes() == r["accelerations"].tobytes()
    # rows outside the stopping segments are not written
    assert np.all(np.isnan(out["times"][~touched]))
    assert np.all(np.isnan(out["velocities"][~touched]))
    return out


def test_reference_test_scenarios_bitwise_and_properties():
    cases = [(np.full(4, 2.0),) + T.constant_velocity_case(v) for v in (-1.0, 1.0, 0.3)]
    out = check_against_oracle(cases)
    # rescale_to_stop_test.cc:54-83 on the GPU results
    for b, velocity in enumerate((-1.0, 1.0)):
        m, sw = out["count"][b], out["switch_index"][b]
        assert m > 0
        _, begin, _, t, _, _ = pack(cases)
        ts = out["times"][begin[b] + sw:begin[b] + sw + m]
        assert abs((ts[-1] - ts[0]) - abs(velocity) / 2.0) <= 8e-3
        assert np.all(out["velocities"][begin[b] + sw + m - 1] == 0)


def test_timed_trajectories_bitwise():
    check_against_oracle(T.timed_trajectories(num=8))


def test_random_trajectories_all_dofs_bitwise():
    rng = np.random.default_rng(11)
    for D in (1, 2, 3, 5, 7, 8, 12, 16):
        cases = []
        for _ in range(12):
            n = int(rng.integers(2, 500))
            t = np.cumsum(rng.uniform(1e-3, 2e-2, n))
            vel = np.cumsum(rng.normal(0, 0.05, (n, D)), axis=0) + rng.normal(0, 1, D)
            acc = rng.normal(0, 1.0, (n, D))
            if rng.random() < 0.2:
                vel[:, int(rng.integers(0, D))] = 0.0  # a joint at rest: no candidate
            cases.append((rng.uniform(0.5, 4.0, D), t, np.cumsum(vel, axis=0) * 0.01, vel, acc))
        check_against_oracle(cases)


def test_edge_cases():
    t, pos, vel, acc = T.constant_velocity_case(1.0)
    stopped = vel.copy()
    stopped[-1] = 0.0
    bad_t = t.copy()
    bad_t[10] = bad_t[9]
    cases = [(np.full(4, 2.0), t, pos, stopped, acc),         # already stopped: empty
             (np.full(4, 2.0), bad_t, pos, vel, acc),         # times not increasing
             (np.full(4, 2.0), t[:1], pos[:1], vel[:1], acc[:1]),  # one sample
             (np.full(4, 2.0), t[:2], pos[:2], vel[:2], acc[:2]),  # two samples: only the zero sample
             (np.full(4, 1e-3), t, pos, vel, acc)]            # cannot stop within the trajectory
    out = check_against_oracle(cases)
    assert list(out["count"][:4]) == [0, 0, 0, 1]
    assert out["status"][1] != 0 and out["status"][2] != 0
    assert out["count"][4] == len(t) - 1


def test_golden_vectors_from_reference_object_code():
    g = np.load(T.GOLDEN)
    cases = [(g[f"amax_{c}"], g[f"t_{c}"], g[f"pos_{c}"], g[f"vel_{c}"], g[f"acc_{c}"])
             for c in range(int(g["num_cases"]))]
    for D in sorted({c[3].shape[1] for c in cases}):
        sub = [(i, c) for i, c in enumerate(cases) if c[3].shape[1] == D]
        amax, begin, length, t, v, a = pack([c for _, c in sub])
        out = pkg.rescale_to_stop_host(amax, begin, length, t, v, a)
        for k, (i, _) in enumerate(sub):
            m, sw = int(g[f"count_{i}"]), int(g[f"switch_{i}"])
            assert out["count"][k] == m and out["switch_index"][k] == sw
            lo = begin[k] + sw
            assert out["times"][lo:lo + m].tobytes() == g[f"out_times_{i}"].tobytes()
            assert out["velocities"][lo:lo + m].tobytes() == g[f"out_velocities_{i}"].tobytes()
            assert out["accelerations"][lo:lo + m].tobytes() == g[f"out_accelerations_{i}"].tobytes()


def test_too_many_dofs_is_unimplemented():
    n, D = 10, 17
    t = np.arange(n) * 0.01
    with pytest.raises(pkg.TpbError) as ei:
        pkg.rescale_to_stop_host(np.ones((1, D)), [0], [n], t, np.ones((n, D)), np.zeros((n, D)))
    assert ei.value.code == 12


# ---- GetPathStopParameter / ComputeFastestStop on the engine's resampled batch ----
def test_path_stop_parameter_matches_oracle_bitwise():
    B, W, D, N, dt = 24, 6, 7, 400, 0.004
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=777)
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 8000)
    eng.set_joint_waypoints(wp)
    eng.set_joint_limits(vmax, amax)
    eng.sample_paths()
    eng.optimize()
    eng.resample(0.0, dt)
    off = eng.get(pkg.ARR.R_OFFSET)
    rt, rs = eng.get(pkg.ARR.R_TIME), eng.get(pkg.ARR.R_S)
    rv, ra = eng.get(pkg.ARR.R_VEL).reshape(-1, D), eng.get(pkg.ARR.R_ACC).reshape(-1, D)
    rng = np.random.default_rng(3)
    for trial in range(4):
        if trial == 0:
            query = np.zeros(B)                        # from the start
        elif trial == 1:
            query = np.array([rt[off[b + 1] - 1] for b in range(B)])   # from the last sample
        elif trial == 2:
            query = np.array([rt[off[b + 1] - 1] + 1.0 for b in range(B)])  # past the end
        else:
            query = np.array([rng.uniform(0, rt[off[b + 1] - 1]) for b in range(B)])
        got = eng.path_stop_parameter(query)
        for b in range(B):
            sl = slice(off[b], off[b + 1])
            o = orc.path_stop_parameter(rt[sl], rs[sl], rv[sl], ra[sl], amax[b], query[b])
            assert (got["status"][b] == 0) == (o["status"] == 0)
            if o["status"] != 0:
                continue
            assert got["row"][b] == o["row"]
            assert got["parameter"][b] == o["parameter"]
            assert got["duration"][b] == o["duration"]
    eng.close()


def test_path_stop_parameter_reference_test_scenario():
    """path_timing_trajectory_test.cc:942-992 through the engine."""
    wp = np.array([[[0.0, 0, 0], [1, 1, 1], [2, 2, 2]]])
    eng = pkg.Engine(max_paths=1, num_samples=500, num_dofs=3, max_waypoints=3,
                     max_time_samples_total=20000)
    eng.set_path_options(constraint_safety=1.0)
    eng.set_joint_waypoints(wp)
    eng.set_joint_limits(np.ones((1, 3)), np.full((1, 3), 2.0))
    eng.sample_paths()
    eng.optimize()
    eng.resample(0.0, 0.001)
    rs, rt = eng.get(pkg.ARR.R_S), eng.get(pkg.ARR.R_TIME)
    q = eng.path_stop_parameter([0.0])
    assert q["status"][0] == 0
    assert np.searchsorted(rs, q["parameter"][0], side="right") <= 3
    q = eng.path_stop_parameter([rt[-1]])
    assert q["parameter"][0] == rs[-1]
    eng.close()
