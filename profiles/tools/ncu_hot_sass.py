#!/usr/bin/env python
"""Per-SASS-instruction stall cycles of one kernel from `ncu --page source --csv`.

  python ncu_hot_sass.py src.csv kernels.sass <mangled-kernel-substring> [top]

Stall samples are scaled so that the column reads "warp cycles spent waiting to
issue this instruction, per execution of the hottest loop body" (the execution
count of the most frequent instruction is taken as the trip count).
"""
import csv
import sys
from collections import defaultdict

sys.path.insert(0, __file__.rsplit("/", 1)[0])
from ncu_by_line import sass_lines  # noqa: E402


def main():
    src_csv, sass, ksub = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(src_csv)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    ci, ct, cs = (hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"),
                  hdr.index("# Samples"))
    stall_cols = [i for i, hh in enumerate(hdr) if hh.startswith("stall_") and "Not Issued" not in hh]
    sl = sass_lines(sass, ksub)
    body = [r for r in rows[h + 1:] if len(r) > ci][:len(sl)]
    ts = sum(float(r[cs] or 0) for r in body)
    trips = max(float(r[ci] or 0) for r in body)
    print(f"instructions {len(body)}, samples {ts:.0f}, trips (max exec count) {trips:.0f}")
    per_line = defaultdict(float)
    items = []
    for k, ((tag, op), r) in enumerate(zip(sl, body)):
        s = float(r[cs] or 0)
        per_line[tag] += s
        best = max(((float(r[i] or 0), hdr[i]) for i in stall_cols), default=(0, ""))
        items.append((s, k, tag, op, float(r[ci] or 0), float(r[ct] or 0), best[1]))
    items.sort(reverse=True)
    print("share%  idx  line  exec/trip thr  top-stall  sass")
    for s, k, tag, op, ex, th, why in items[:top]:
        print(f"{100 * s / ts:6.2f} {k:5d} {tag.split('/')[-1]:26s} {ex / trips:5.2f} {th / max(ex, 1):5.1f} "
              f"{why:22s} {op[:70]}")
    print("\nby line (share% of stall samples):")
    for tag, s in sorted(per_line.items(), key=lambda kv: -kv[1])[:25]:
        print(f"{100 * s / ts:6.2f} {tag}")


if __name__ == "__main__":
    main()
