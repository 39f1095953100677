"""Writes tests/golden/trajectory_buffer.npz: what the reference's own
TrajectoryBuffer (oracle/_ref = trajectory_buffer.cc compiled verbatim from
/root/reference against oracle/shim) shows after every operation of seeded
operation sequences (tests/tbuf_cases.py): per-operation status / sample count /
sequence number / start time, the final rows, and a SHA-256 over all intermediate
rows.  Run in the build container:
    python tests/golden/make_trajectory_buffer.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import pyoracle as orc  # noqa: E402
import tbuf_cases as TC  # noqa: E402

assert orc.have_ref(), "build oracle/_ref first (make -C oracle)"
out = {}
specs = [TC.ops_spec(100 + c, D=[7, 9, 3][c % 3], num_ops=50, tol=[1e-6, 1e-5][c % 2])
         for c in range(6)]
for c, spec in enumerate(specs):
    # the sequence is materialised on the reference itself
    ops = TC.generate_ops(spec, orc.TrajectoryBuffer(spec["D"], spec["tol"], impl="ref"))
    obs = TC.run_ops(orc.TrajectoryBuffer(spec["D"], spec["tol"], impl="ref"), ops)
    status = [-1 if o["result"] is None else o["result"] if isinstance(o["result"], int)
              else o["result"][0] for o in obs]
    t, p, v, a = (np.frombuffer(obs[-1][k]).copy() for k in "tpva")
    out.update({f"seed_{c}": spec["seed"], f"D_{c}": spec["D"], f"num_ops_{c}": spec["num_ops"],
                f"dt_{c}": spec["dt"], f"tol_{c}": spec["tol"],
                f"n_{c}": np.array([o["n"] for o in obs]), f"seq_{c}": np.array([o["seq"] for o in obs]),
                f"start_{c}": np.array([o["start"] for o in obs], dtype=np.int64),
                f"status_{c}": np.array(status), f"t_{c}": t, f"p_{c}": p, f"v_{c}": v, f"a_{c}": a,
                f"digest_{c}": TC.digest(obs)})
out["num_cases"] = len(specs)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "trajectory_buffer.npz"), **out)
print("wrote", len(specs), "cases")
