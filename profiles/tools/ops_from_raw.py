#!/usr/bin/env python
"""fp64_ops.json / traffic.json entries from an `ncu --page raw --csv` export of one
`--set full` capture (first launch of every kernel name counts).
  ncu -i cap.ncu-rep --page raw --csv > raw.csv
  python profiles/tools/ops_from_raw.py raw.csv <source note>  ->  prints both JSON fragments"""
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
print(json.dumps({"_source": note, **ops}, indent=1))
print(json.dumps(traffic, indent=1))
