"""Times tpb_plan_joint_device on the cfg2 batch for several chunk counts (CUDA events
on the engine's stream) and checks that every variant returns the rows of the three
separate stage calls bit for bit.  Run on the GPU box:
    python profiles/tools/pipeline_sweep.py"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
B, W, D, N, DT = 4096, 10, 7, 2048, 0.004
wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=1000)
stream = torch.cuda.Stream()
eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                 max_time_samples_total=B * 6000)
eng.set_stream(stream.cuda_stream)
eng.set_joint_waypoints(wp)
eng.set_joint_limits(vmax, amax)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def three_calls():
    eng.sample_paths()
    eng.optimize()
    eng.resample(0.0, DT, want_total=False)


out = {"three_calls_ms": timed(three_calls)}
keys = ("R_OFFSET", "R_TIME", "R_POS", "R_VEL", "R_ACC", "R_S", "DURATION", "STATUS")
ref = {k: eng.get(getattr(pkg.ARR, k)).copy() for k in keys}
for chunks in (1, 2, 3, 4, 6, 8, 12, 16):
    ms = timed(lambda: eng.plan_joint_device(0.0, DT, chunks, want_total=False))
    same = all(np.array_equal(eng.get(getattr(pkg.ARR, k)), ref[k], equal_nan=True)
               if ref[k].dtype.kind == "f" else np.array_equal(eng.get(getattr(pkg.ARR, k)), ref[k])
               for k in keys)
    out[f"chunks_{chunks}"] = {"ms": ms, "paths_per_s": B / ms * 1e3, "identical": bool(same)}
print(json.dumps(out, indent=1))
