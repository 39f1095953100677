"""The CPU oracle (oracle/tp_oracle*.c) against committed golden vectors written from the
REFERENCE'S OWN OBJECT CODE (tests/golden/planner_ref.npz, generator
tests/golden/make_planner_ref.py: TimeableJointSplinePath + PathTimingTrajectory compiled
verbatim into oracle/_ref).  Runs without oracle/_ref and without /root/reference.  CPU only."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as orc  # noqa: E402

MS = 1_000_000
WAYPOINTS = np.array([[1, 2, 3], [-1, -2, -3], [1, 2, 3.0]])  # _test.cc:127-128
VMAX, AMAX = np.full(3, 1.0), np.full(3, 2.0)
N, DELTA, DT = 1000, 0.005, 0.004
HORIZON = 750 * MS
ROW_KEYS = ("r_t", "r_pos", "r_vel", "r_acc", "r_s", "r_sd", "r_sdd")
FRESH_KEYS = ("q", "dq", "ddq", "time", "s", "sd", "sdd", "vel", "acc") + ROW_KEYS


def same_bits(a, b, what):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert a.tobytes() == b.tobytes(), f"{what}: {(a != b).sum()} of {a.size} values differ"


GOLDEN = os.path.join(ROOT, "tests", "golden", "planner_ref.npz")


def test_port_reproduces_golden_fresh_paths():
    g = np.load(GOLDEN)
    for c in range(int(g["num_fresh"])):
        wp, vmax, amax = g[f"f{c}_wp"], g[f"f{c}_vmax"], g[f"f{c}_amax"]
        o = orc.plan_joint(wp, vmax, amax, 192, dt=float(g["fresh_dt"]), M_cap=40000)
        assert o["ok"]
        ctrl, kn = orc.joint_fit(wp, 0.2)
        same_bits(kn, g[f"f{c}_knots"], "knots")
        same_bits(ctrl, g[f"f{c}_ctrl"], "control points")
        for k in FRESH_KEYS:
            same_bits(o[k][:o["M"]] if k in ROW_KEYS else o[k], g[f"f{c}_{k}"], f"path {c} {k}")


@pytest.mark.parametrize("skip", [0, 1])
def test_port_reproduces_golden_replanning(skip):
    g = np.load(GOLDEN)
    states, digests = g[f"r{skip}_states"], g[f"r{skip}_digests"]
    port = orc.Planner(WAYPOINTS, VMAX, AMAX, N, DELTA, DT, skip=bool(skip))
    for call in range(len(states)):
        start = int(states[call][0])
        assert port.plan(start, HORIZON) == orc.ORC_OK
        got = [port.start_time, port.end_time, port.final_decel_start, port.planned_to_end,
               port.target_reached, port.M, port.len]
        assert got == [int(x) for x in states[call][1:]], (call, got)
        tr = port.trajectory()
        M = len(tr["r_t"])
        rows = [0, M // 2, M - 1]
        d = np.concatenate([[M, tr["r_t"].sum(), tr["r_pos"].sum(), tr["r_vel"].sum(),
                             tr["r_acc"].sum()], tr["r_t"][rows], tr["r_pos"][rows].ravel(),
                            tr["r_vel"][rows].ravel()])
        same_bits(d, digests[call], f"digest of call {call}")
    assert port.target_reached
    for k, v in port.trajectory().items():
        same_bits(v, g[f"r{skip}_final_{k}"], k)
