"""Writes tests/golden/rescale_to_stop.npz: inputs and outputs of the reference's own
RescaleTrajectoryBackwardToStop (oracle/_ref = rescale_to_stop.cc compiled verbatim
from /root/reference against oracle/shim).  Run in the build container:
    python tests/golden/make_rescale_to_stop.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import pyoracle as orc  # noqa: E402
import test_oracle_stop as T  # noqa: E402

assert orc.have_ref(), "build oracle/_ref first (make -C oracle)"
cases = [(np.full(4, 2.0),) + T.constant_velocity_case(v) for v in (-1.0, 1.0)]
cases += T.timed_trajectories(num=3)
out = {"num_cases": len(cases)}
for c, (amax, t, pos, vel, acc) in enumerate(cases):
    r = orc.rescale_to_stop(amax, t, pos, vel, acc, impl="ref")
    assert r["status"] == 0 and r["count"] > 0
    out.update({f"amax_{c}": amax, f"t_{c}": t, f"pos_{c}": pos, f"vel_{c}": vel, f"acc_{c}": acc,
                f"count_{c}": r["count"], f"switch_{c}": r["switch_index"],
                f"out_times_{c}": r["times"], f"out_velocities_{c}": r["velocities"],
                f"out_accelerations_{c}": r["accelerations"]})
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "rescale_to_stop.npz"), **out)
print("wrote", len(cases), "cases")
