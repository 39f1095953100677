"""RescaleTrajectoryBackwardToStop: the CPU restatement (oracle/tp_oracle_stop.c)
against the reference's own object code (oracle/_ref, rescale_to_stop.cc compiled
verbatim) and against the properties the reference's test checks
(rescale_to_stop_test.cc:29-84).  CPU only."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as orc  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
GOLDEN = os.path.join(ROOT, "tests", "golden", "rescale_to_stop.npz")


def constant_velocity_case(velocity, n=200, dt=8e-3, D=4):
    """rescale_to_stop_test.cc:30-52"""
    t = np.arange(n) * dt
    vel = np.full((n, D), velocity)
    acc = np.zeros((n, D))
    pos = velocity * t[:, None] * np.ones((1, D))
    return t, pos, vel, acc


def timed_trajectories(num=6, N=300, dt=0.004):
    """Resampled trajectories of timed joint paths (cfg2 generator), cut before the
    end so that they do not end in a stop."""
    wp, vmax, amax = pkg.workloads.joint_paths(num, 6, 7, seed=4242)
    out = []
    for b in range(num):
        r = orc.plan_joint(wp[b], vmax[b], amax[b], N, dt=dt, M_cap=20000)
        assert r["ok"]
        M = r["M"]
        cut = int(M * (0.35 + 0.1 * b))
        out.append((amax[b], r["r_t"][:cut], r["r_pos"][:cut], r["r_vel"][:cut], r["r_acc"][:cut]))
    return out


@pytest.mark.parametrize("velocity", [-1.0, 1.0])
def test_reference_test_properties(velocity):
    """rescale_to_stop_test.cc:54-83"""
    n, dt, D, amax = 200, 8e-3, 4, 2.0
    t, pos, vel, acc = constant_velocity_case(velocity, n, dt, D)
    r = orc.rescale_to_stop(np.full(D, amax), t, pos, vel, acc)
    assert r["status"] == 0 and r["count"] > 0
    ramp = abs(velocity) / amax
    stop_dist = velocity * velocity / (2.0 * amax)
    assert abs((r["times"][-1] - r["times"][0]) - ramp) <= dt
    assert np.all(r["accelerations"][-1] == 0) and np.all(r["velocities"][-1] == 0)
    moved = r["positions"][-1] - r["positions"][0]
    assert np.all(np.abs(moved - stop_dist * np.sign(velocity)) <= abs(velocity * dt))
    # the rescaled accelerations respect the limit
    assert np.all(np.abs(r["accelerations"]) <= amax + 1e-8)


def test_already_stopped_and_invalid_inputs():
    t, pos, vel, acc = constant_velocity_case(1.0)
    vel[-1] = 0.0
    r = orc.rescale_to_stop(np.full(4, 2.0), t, pos, vel, acc)
    assert r["status"] == 0 and r["count"] == 0
    t2 = t.copy()
    t2[10] = t2[9]
    assert orc.rescale_to_stop(np.full(4, 2.0), t2, pos, vel, acc)["status"] != 0
    assert orc.rescale_to_stop(np.full(4, 2.0), t[:1], pos[:1], vel[:1], acc[:1])["status"] != 0


@pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built")
def test_port_is_bit_identical_to_reference_object_code():
    cases = [(np.full(4, 2.0),) + constant_velocity_case(v) for v in (-1.0, 1.0, 0.3)]
    cases += timed_trajectories()
    rng = np.random.default_rng(5)
    for _ in range(20):  # random smooth trajectories, random limits
        n, D = int(rng.integers(3, 400)), int(rng.integers(1, 9))
        t = np.cumsum(rng.uniform(1e-3, 2e-2, n))
        vel = np.cumsum(rng.normal(0, 0.05, (n, D)), axis=0) + rng.normal(0, 1, D)
        acc = rng.normal(0, 1.0, (n, D))
        pos = np.cumsum(vel, axis=0) * 0.01
        cases.append((rng.uniform(0.5, 4.0, D), t, pos, vel, acc))
    nonempty = 0
    for amax, t, pos, vel, acc in cases:
        a = orc.rescale_to_stop(amax, t, pos, vel, acc, impl="port")
        b = orc.rescale_to_stop(amax, t, pos, vel, acc, impl="ref")
        assert a["status"] == b["status"] == 0
        assert a["count"] == b["count"] and a["switch_index"] == b["switch_index"]
        nonempty += a["count"] > 0
        for k in ("times", "positions", "velocities", "accelerations"):
            assert a[k].tobytes() == b[k].tobytes(), k
    assert nonempty >= 20


def test_port_matches_committed_golden_vectors():
    """tests/golden/rescale_to_stop.npz was written by tests/golden/make_rescale_to_stop.py
    from the reference object code in this container."""
    g = np.load(GOLDEN)
    for c in range(int(g["num_cases"])):
        r = orc.rescale_to_stop(g[f"amax_{c}"], g[f"t_{c}"], g[f"pos_{c}"], g[f"vel_{c}"],
                                g[f"acc_{c}"])
        assert r["count"] == int(g[f"count_{c}"]) and r["switch_index"] == int(g[f"switch_{c}"])
        for k in ("times", "velocities", "accelerations"):
            assert r[k].tobytes() == g[f"out_{k}_{c}"].tobytes(), (c, k)


# ---- ComputeFastestStop / GetPathStopParameter --------------------------------
def test_path_stop_parameter_reference_test_properties():
    """path_timing_trajectory_test.cc:942-992 (GetPathStopParameterWorksInSimpleCase):
    collinear 3-dof waypoints, safety 1, 500 samples, 1 ms time step."""
    wp = np.array([[0.0, 0, 0], [1, 1, 1], [2, 2, 2]])
    amax = np.full(3, 2.0)
    r = orc.plan_joint(wp, np.ones(3), amax, 500, safety=1.0, dt=0.001, M_cap=20000)
    assert r["ok"]
    args = (r["r_t"], r["r_s"], r["r_vel"], r["r_acc"], amax)
    # from the first sample the path can stop (almost) immediately
    q = orc.path_stop_parameter(*args, 0.0)
    assert q["status"] == 0
    assert np.searchsorted(r["r_s"], q["parameter"], side="right") <= 3
    # from the last sample the parameter is the final one
    q = orc.path_stop_parameter(*args, r["r_t"][-1])
    assert q["status"] == 0 and q["parameter"] == r["r_s"][-1]
    assert np.searchsorted(r["r_s"], q["parameter"], side="right") == len(r["r_s"])
    # past the trajectory: the reference returns kInvalidArgument
    assert orc.path_stop_parameter(*args, r["r_t"][-1] + 1.0)["status"] != 0
    # in between: the stop lasts about |v| / a_max and ends where the scaled trajectory rests
    k = r["M"] // 3
    q = orc.path_stop_parameter(*args, r["r_t"][k])
    v = np.abs(r["r_vel"][k]).max()
    assert q["row"] > k and abs(q["duration"] - v / 2.0) < 0.05 * v / 2.0 + 2e-3


def test_fastest_stop_series_is_consistent():
    t, pos, vel, acc = constant_velocity_case(1.0)
    f = orc.fastest_stop(np.full(4, 2.0), t, vel, acc)
    assert f["rate2"][0] == 1.0 and f["rate2"][-1] == 0.0
    assert np.all(np.diff(f["rate2"]) <= 0) and np.all(f["diff"] <= 0)
    # v / a_max = 0.5 in the continuum; the trapezoidal time of the last step, where the
    # rate reaches 0, overshoots by O(sqrt(time step))
    assert 0.5 <= f["duration"] <= 0.5 + 2.0 * np.sqrt(8e-3)
    assert abs(f["time"][-1] - (t[0] + f["duration"])) < 1e-15
    assert f["stop_index"] == len(f["time"]) - 1
