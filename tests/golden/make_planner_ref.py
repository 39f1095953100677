"""Writes tests/golden/planner_ref.npz: outputs of the reference's own
TimeableJointSplinePath + PathTimingTrajectory (oracle/_ref = bspline_base.cc,
spline_utils.cc, timeable_path_joint_spline.cc, time_optimal_path_timing.cc and
path_timing_trajectory.cc compiled verbatim from /root/reference against oracle/shim).
Run in the build container:
    python tests/golden/make_planner_ref.py

Contents: (a) six seeded fresh 7-dof paths (N = 192, whole path in one window): control
points, knots, q / q' / q'', profile, rows resampled at 16 ms; (b) the RestToRest re-planning
scenario of path_timing_trajectory_test.cc:112-173 for both time sampling methods: the
planner state after every Plan() and a digest (row count, sum, first / middle / last
row) of every trajectory, plus the full final trajectory."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import pyoracle as orc  # noqa: E402
import test_oracle_vs_ref_planner as T  # noqa: E402

assert orc.have_ref(), "build oracle/_ref first (make -C oracle)"
out = {}
rng = np.random.default_rng(4242)
NF = 6
FRESH_DT = 0.016  # keeps the fixture small
for c in range(NF):
    wp, vmax, amax = T.random_path(rng, 4 + c, 7)
    r = orc.plan_joint_ref(wp, vmax, amax, 192, dt=FRESH_DT)
    assert r["ok"]
    out.update({f"f{c}_wp": wp, f"f{c}_vmax": vmax, f"f{c}_amax": amax,
                f"f{c}_knots": r["spline"][0], f"f{c}_ctrl": r["spline"][1]})
    for k in T.FRESH_KEYS:
        out[f"f{c}_{k}"] = r[k]
out["num_fresh"] = NF
out["fresh_dt"] = FRESH_DT


def digest(tr):
    M = len(tr["r_t"])
    rows = [0, M // 2, M - 1]
    return np.concatenate([[M, tr["r_t"].sum(), tr["r_pos"].sum(), tr["r_vel"].sum(),
                            tr["r_acc"].sum()], tr["r_t"][rows], tr["r_pos"][rows].ravel(),
                           tr["r_vel"][rows].ravel()])


for skip in (0, 1):
    ref = orc.RefPlanner(T.WAYPOINTS, T.VMAX, T.AMAX, T.N, T.DELTA, T.DT, skip=bool(skip))
    start, states, digests = 0, [], []
    while True:
        assert ref.plan(start, T.HORIZON) == 0
        s = ref.state()
        states.append([start] + [s[k] for k in ("start_time", "end_time", "final_decel_start",
                                                "planned_to_end", "target_reached", "M", "len")])
        digests.append(digest(ref.trajectory()))
        if s["target_reached"]:
            break
        start = min(s["end_time"], start + T.REPLAN)
    out[f"r{skip}_states"] = np.array(states, np.int64)
    out[f"r{skip}_digests"] = np.array(digests)
    for k, v in ref.trajectory().items():
        out[f"r{skip}_final_{k}"] = v
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "planner_ref.npz"), **out)
print("wrote", NF, "fresh paths and", [len(out[f"r{s}_states"]) for s in (0, 1)], "plan calls")
