#!/usr/bin/env python
"""fp64_ops.json / traffic.json entries from an `ncu --page raw --csv` export of one
`--set full` capture (first launch of every kernel name counts).
  ncu -i cap.ncu-rep --page raw --csv > raw.csv
  python profiles/tools/ops_from_raw.py raw.csv <source note> [--write]
  -> prints both JSON fragments, or writes profiles/fp64_ops.json and traffic.json (with the
     hash of csrc/ so that bench.py can tell when they are stale)"""
import csv
import json
import sys

NAMES = (("k_joint_sample_limit", "k_joint_sample_limit"), ("kPhaseSetup|, 7, 1>", "k_integrate_setup"),
         ("k_optimize_pair", "k_optimize_pair"), (", 7, 4>", "k_integrate_finalize"),
         ("k_resample_fill", "k_resample_fill"))


def num(x):
    return float(x.replace(",", ""))


rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]
col = {n: i for i, n in enumerate(h)}
ops, traffic = {}, {}
for r in rows[2:]:
    if len(r) < len(h):
        continue
    kn = r[col["Kernel Name"]]
    key = next((k for pats, k in NAMES if any(p in kn for p in pats.split("|"))), None)
    if key is None or key in ops:
        continue
    cyc = num(r[col["smsp__cycles_elapsed.avg"]])
    nsmsp = 148 * 4
    o = {}
    for op in ("dadd", "dmul", "dfma"):
        name = f"smsp__sass_thread_inst_executed_op_{op}_pred_on.sum.per_cycle_elapsed"
        o[op] = int(round(num(r[col[name]]) * cyc))
    o["flops_per_launch"] = o["dadd"] + o["dmul"] + 2 * o["dfma"]
    o["fp64_pipe_pct"] = num(r[col["sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"]])
    t = num(r[col["gpu__time_duration.sum"]])
    unit = rows[1][col["gpu__time_duration.sum"]]
    o["ncu_ms"] = t / 1e3 if unit in ("us", "usecond") else (t / 1e6 if unit in ("ns", "nsecond") else t)
    ops[key] = o

    def gb(name):
        v, u = num(r[col[name]]), rows[1][col[name]]
        return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[u]
    rd, wr = gb("dram__bytes_read.sum"), gb("dram__bytes_write.sum")
    traffic[key] = {"dram_bytes_per_launch": int(rd + wr), "read": rd, "write": wr}
note = sys.argv[2] if len(sys.argv) > 2 else ""
# the hash of the CUDA sources the capture was taken from (bench.py prints "stale" when
# the tree's differs)
import hashlib
import os
root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
HOT_KERNEL_SOURCES = ("tpb_kernels.cu", "tpb_integrate.cuh", "tpb_integrate_pair.cuh",
                      "tpb_joint_limit.cuh", "tpb_row_limit.cuh", "tpb_solver.cuh", "tpb_device.cuh",
                      "tpb_bulk.cuh", "tpb_fit.cuh")  # as in bench.py
hsh = hashlib.sha256()
for f in HOT_KERNEL_SOURCES:
    hsh.update(open(os.path.join(root, "x-edr-trajectory-planning_b200", "csrc", f), "rb").read())
digest = hsh.hexdigest()[:16]
if len(sys.argv) > 3 and sys.argv[3] == "--write":
    with open(os.path.join(root, "profiles", "fp64_ops.json"), "w") as f:
        json.dump({"_source": note, "_csrc_hash": digest, **ops}, f, indent=1)
    with open(os.path.join(root, "profiles", "traffic.json"), "w") as f:
        json.dump({"_source": note, "_csrc_hash": digest, **traffic}, f, indent=1)
    print("wrote profiles/fp64_ops.json and profiles/traffic.json, csrc hash", digest)
else:
    print(json.dumps({"_source": note, "_csrc_hash": digest, **ops}, indent=1))
    print(json.dumps({"_csrc_hash": digest, **traffic}, indent=1))
