#!/usr/bin/env python
"""Per-path phase timing of k_integrate (diagnostics, not a benchmark).

Needs a library built with -DTPB_PROFILE_PHASES:
    make -C x-edr-trajectory-planning_b200/csrc all LIB=$PWD/x-edr-trajectory-planning_b200/lib_ab/phases EXTRA=-DTPB_PROFILE_PHASES
    TPB_LIB=$PWD/x-edr-trajectory-planning_b200/lib_ab/phases/libtpb.so python profiles/tools/phase_profile.py
Prints the distribution over paths of the SM-cycle count of every phase.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
B, N, D, W = 4096, 2048, 7, 10
wp, vmax, amax = pkg.workloads.joint_paths(B, W, D, seed=1000)
eng = pkg.Engine(max_paths=B, num_samples=N, num_dofs=D, max_waypoints=W,
                 max_time_samples_total=B * 6144)
eng.set_kernel_variant("integrate_warp")  # the counters live in the one-path-per-warp kernel
eng.set_joint_waypoints(wp)
eng.set_joint_limits(vmax, amax)
eng.sample_paths()
for _ in range(3):
    eng.optimize()
dbg = eng.get(pkg.ARR.DEBUG).reshape(B, N)[:, :20]
names = ["init", "boundary", "first_back", "first_fwd", "loop", "finalize"]
ph = np.diff(dbg[:, :7], axis=1)
tot = dbg[:, 6]
print("cycles per path: phase  mean  p50  p90  p99  max")
for k, n in enumerate(names):
    c = ph[:, k]
    print(f"{n:12s} {c.mean():10.0f} {np.percentile(c, 50):10.0f} {np.percentile(c, 90):10.0f} "
          f"{np.percentile(c, 99):10.0f} {c.max():10.0f}")
print(f"{'total':12s} {tot.mean():10.0f} {np.percentile(tot, 50):10.0f} {np.percentile(tot, 90):10.0f} "
      f"{np.percentile(tot, 99):10.0f} {tot.max():10.0f}")
print("switching loops: mean %.1f max %d" % (dbg[:, 7].mean(), dbg[:, 7].max()))
for k, n in ((10, "next_critical"), (11, "add_backward"), (12, "add_forward")):
    c = dbg[:, k]
    print(f"  loop/{n:14s} mean {c.mean():10.0f} p99 {np.percentile(c, 99):10.0f}")
print("  loop steps backward mean %.0f forward mean %.0f; cycles/step bw %.0f fw %.0f" % (
    dbg[:, 13].mean(), dbg[:, 14].mean(), dbg[:, 11].sum() / max(dbg[:, 13].sum(), 1),
    dbg[:, 12].sum() / max(dbg[:, 14].sum(), 1)))
print("  loop: sum of max(backward, forward) cycles mean %.0f (vs sum %.0f); all extremal steps mean %.0f max %d" % (
    dbg[:, 15].mean(), (dbg[:, 11] + dbg[:, 12]).mean(), dbg[:, 16].mean(), dbg[:, 16].max()))
start = dbg[:, 9] - dbg[:, 9].min()
print("kernel span (cycles): %.0f; last start %.0f" % ((start + tot).max(), start.max()))
worst = np.argsort(-tot)[:5]
for b in worst:
    print("path", b, "phases", ph[b].astype(int).tolist(), "loops", int(dbg[b, 7]))
steps = dbg[:, 16]
print("extremal steps per path: mean %.0f p50 %.0f p90 %.0f p99 %.0f max %.0f" % (
    steps.mean(), np.percentile(steps, 50), np.percentile(steps, 90), np.percentile(steps, 99), steps.max()))
pair = np.maximum(steps[0::2], steps[1::2])
print("max over the two paths of a warp (k_optimize_pair): mean %.0f p90 %.0f p99 %.0f max %.0f" % (
    pair.mean(), np.percentile(pair, 90), np.percentile(pair, 99), pair.max()))
srt = np.sort(steps)
pair_sorted = np.maximum(srt[0::2], srt[1::2])
print("the same with paths paired by step count: mean %.0f max %.0f" % (pair_sorted.mean(), pair_sorted.max()))
np.save(os.path.join(ROOT, "gpurun_out", "phase_dbg.npy"), dbg)
