#!/usr/bin/env python
"""Time tpb_set_constraints (device-resident rows: a D2D copy + the per-sample limit-curve
kernel) for explicit rows, register kernel vs TPB_GENERIC_LIMIT_LOCAL=1.
  python profiles/tools/generic_limit_time.py [C] [B] [N]"""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
pkg = importlib.import_module("x-edr-trajectory-planning_b200")

C = int(sys.argv[1]) if len(sys.argv) > 1 else 16
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
N = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
rng = np.random.default_rng(1)
base = np.stack([pkg.workloads.random_constraints(rng, N, C) for _ in range(16)])
cons = torch.from_numpy(np.ascontiguousarray(np.tile(base, (B // 16, 1, 1, 1)))).cuda()
s0 = torch.zeros(B, dtype=torch.float64, device="cuda")
s1 = torch.ones(B, dtype=torch.float64, device="cuda")
for local in ("0", "1"):
    os.environ["TPB_GENERIC_LIMIT_LOCAL"] = local
    eng = pkg.Engine(max_paths=B, num_samples=N, max_constraints=C)
    eng.set_stream(torch.cuda.current_stream().cuda_stream) if hasattr(eng, "set_stream") else None
    for _ in range(3):
        eng.set_constraints(cons, s0, s1)
    eng.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        eng.set_constraints(cons, s0, s1)
    eng.sync()
    e1.record()
    torch.cuda.synchronize()
    eng.optimize()
    sd2 = eng.get(pkg.ARR.B_SD2_MAX)
    print(f"C={C} B={B} N={N} local={local}: {e0.elapsed_time(e1) / 10:.3f} ms per call "
          f"(rows {cons.numel() * 8 / 1e9:.2f} GB), checksum {float(np.nansum(sd2)):.17g}")
    eng.close()
