#!/usr/bin/env python
"""Per-stage device time of the hot path against the number of paths per launch
(diagnostics for BASELINE cfg4: 65,536 paths on ONE GPU; not the benchmark).

    python profiles/tools/batch_scaling.py [B ...]      # default 4096 16384 65536

Inputs resident in HBM, CUDA events on the launching stream, 3 warm-ups + 5 timed
steps per size.  Prints one JSON line per size.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
N, D, W, DT = 2048, 7, 10, 0.004
sizes = [int(x) for x in sys.argv[1:]] or [4096, 16384, 65536]
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
for B in sizes:
    wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=1000)
    d_wp, d_vm, d_am = (torch.from_numpy(x).to(dev) for x in (wp, vmax, amax))
    eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                     max_time_samples_total=B * 4608, device=0)
    eng.set_stream(stream.cuda_stream)
    eng.time_phases(True)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    stage = np.zeros(4)
    phase = np.zeros(3)
    reps = 5
    for it in range(3 + reps):
        evs[0].record(stream)
        eng.set_joint_waypoints(d_wp)
        eng.set_joint_limits(d_vm, d_am)
        evs[1].record(stream)
        eng.sample_paths()
        evs[2].record(stream)
        eng.optimize()
        evs[3].record(stream)
        eng.resample(0.0, DT, want_total=False)
        evs[4].record(stream)
        torch.cuda.synchronize()
        if it >= 3:
            stage += [evs[i].elapsed_time(evs[i + 1]) / reps for i in range(4)]
            phase += np.array(eng.last_phase_ms()) / reps
    status = eng.get(pkg.ARR.STATUS)
    rows = int(eng.get(pkg.ARR.R_OFFSET)[-1])
    total = float(stage.sum())
    print(json.dumps({"paths": B, "ms_step": total, "paths_per_s": B / (total / 1e3),
                      "fit": stage[0], "sample_limit": stage[1], "integrate": stage[2],
                      "resample": stage[3], "setup": phase[0], "sweeps": phase[1],
                      "finalize": phase[2], "ok": int((status == 0).sum()), "rows": rows,
                      "us_per_path_sweeps": 1e3 * phase[1] / B}), flush=True)
    eng.close()
    del d_wp, d_vm, d_am
    torch.cuda.empty_cache()
