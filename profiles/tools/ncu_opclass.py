#!/usr/bin/env python
"""Per-opcode-class instruction and stall-cycle budget of one kernel's hot loop.

  python ncu_opclass.py src.csv kernels.sass <mangled-kernel-substring> <kernel cycles> <warps> [trips]

`kernel cycles` = sm__cycles_elapsed.max, `warps` = warps launched; stall samples
are scaled to warp cycles per trip of the hot loop (trips defaults to the most
frequent execution count / 2, pass it when known).
"""
import csv
import sys
from collections import defaultdict

sys.path.insert(0, __file__.rsplit("/", 1)[0])
from ncu_by_line import sass_lines  # noqa: E402

src_csv, sass, ksub, cycles, warps = sys.argv[1:6]
rows = list(csv.reader(open(src_csv)))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
sl = sass_lines(sass, ksub)
body = [r for r in rows[h + 1:] if len(r) > ci][:len(sl)]
ts = sum(float(r[cs] or 0) for r in body)
ex_all = sorted(float(r[ci] or 0) for r in body)
trips = float(sys.argv[6]) if len(sys.argv) > 6 else ex_all[len(ex_all) * 3 // 4]
f = float(warps) * float(cycles) / ts
cls = defaultdict(lambda: [0.0, 0.0])
for (tag, op), r in zip(sl, body):
    o = op.split()
    name = (o[1] if o[0].startswith("@") else o[0]).split(".")[0]
    cls[name][0] += float(r[ci] or 0) / trips
    cls[name][1] += float(r[cs] or 0) * f / trips
print(f"trips {trips:.0f}; per trip: {sum(v[0] for v in cls.values()):.1f} instructions, "
      f"{sum(v[1] for v in cls.values()):.0f} warp cycles")
for k, v in sorted(cls.items(), key=lambda kv: -kv[1][1])[:24]:
    print(f"{k:10s} exec/trip {v[0]:6.1f}  cycles/trip {v[1]:7.1f}  cyc/inst {v[1] / max(v[0], 1e-9):5.1f}")
