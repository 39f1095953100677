#!/usr/bin/env python
"""Per-source-line instruction counts and stall samples of one kernel from an ncu report
captured with --set full --import-source on (the evidence behind profiles/r2_summary.md
section 6).

    ncu -i cap.ncu-rep --page source --csv --kernel-name regex:<kernel> --print-source sass,cuda > src.csv
    python profiles/tools/ncu_source_by_line.py src.csv [top_n]

Prints: total warp instructions, the stall-reason breakdown over all samples, static code
size by file, and the top lines by executed instructions with their share of the samples."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr_idx = [i for i, r in enumerate(rows) if r and r[0] == "Line No"] + [len(rows)]
kernel = next((r[1] for r in rows if r and r[0] == "Function Name"), "?")
agg, stall, static = [], collections.Counter(), collections.Counter()
for a, b in zip(hdr_idx[:-1], hdr_idx[1:]):
    f = rows[a - 2][1].split("/")[-1]
    hdr = rows[a]
    ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
    cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    for r in rows[a + 1:b]:
        if len(r) < len(hdr):
            continue
        if r[2] == "-":  # a source line
            try:
                agg.append((f, int(r[0]), r[1].strip()[:90], int(r[ci]), int(r[cs])))
            except ValueError:
                continue
            for i, h in cols:
                try:
                    stall[h] += int(r[i])
                except ValueError:
                    pass
        elif r[3] and r[3] != "...":
            static[f] += 1  # a SASS row (listed once per view: halve)
ti, ts = sum(x[3] for x in agg), sum(x[4] for x in agg)
print("kernel:", kernel[:140])
print("warp instructions executed: %d, stall samples: %d" % (ti, ts))
ss = sum(stall.values())
print("stall samples: " + ", ".join("%s %.1f %%" % (k[6:], 100.0 * v / ss) for k, v in stall.most_common(9)))
print("static SASS instructions by file: " +
      ", ".join("%s %d (%.1f KB)" % (k, v // 2, v // 2 * 16 / 1024) for k, v in static.most_common()))
agg.sort(key=lambda x: -x[3])
print("%6s %6s  line" % ("inst%", "smp%"))
for f, l, src, n, s in agg[:top]:
    print("%6.1f %6.1f  %s:%d  %s" % (100.0 * n / ti, 100.0 * s / max(ts, 1), f, l, src))
