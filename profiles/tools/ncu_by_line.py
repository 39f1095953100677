#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` SASS listing by CUDA source line.

ncu's CSV export of the source page carries per-SASS-instruction counters but
no line numbers; `nvdisasm -g -c` of the same cubin carries the line of every
instruction.  The two listings hold the same instructions in the same order, so
they are joined by position.

  cuobjdump -xelf all lib/libtpb.so                      # -> *.cubin
  nvdisasm -g -c tpb_kernels.sm_100a.cubin > kernels.sass
  ncu -i prof.ncu-rep --page source --csv --kernel-name regex:k_integrate > src.csv
  python ncu_by_line.py src.csv kernels.sass <mangled-kernel-substring> [min_pct]
"""
import csv
import re
import sys
from collections import defaultdict


def sass_lines(path, kernel_sub):
    """[(line_tag, opcode text)] for the kernel whose .text section matches."""
    out, inside, cur = [], False, "?"
    ins = re.compile(r"^\s+/\*([0-9a-f]{4,})\*/\s+(.*?);")
    for ln in open(path, errors="replace"):
        if ln.startswith("//---") and ".text." in ln:
            inside = kernel_sub in ln
            cur = "?"
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            cur = f"{m.group(1).rsplit('/', 1)[-1]}:{m.group(2)}"
            continue
        m = ins.match(ln)
        if m:
            out.append((cur, m.group(2)))
    return out


def main():
    src_csv, sass, ksub = sys.argv[1:4]
    min_pct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
    rows = list(csv.reader(open(src_csv)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
    body = [r for r in rows[h + 1:] if len(r) > ci]
    sl = sass_lines(sass, ksub)
    if len(sl) != len(body):
        print(f"warning: {len(body)} profiled instructions vs {len(sl)} disassembled", file=sys.stderr)
    inst, samp = defaultdict(float), defaultdict(float)
    for (tag, _), r in zip(sl, body):
        inst[tag] += float(r[ci] or 0)
        samp[tag] += float(r[cs] or 0)
    ti, ts = sum(inst.values()) or 1.0, sum(samp.values()) or 1.0
    print(f"total warp instructions {ti:.4g}, stall samples {ts:.4g}")
    print(f"{'line':28s} {'inst %':>8s} {'samples %':>10s}")

    def key(t):
        f, _, n = t.partition(":")
        return (f, int(n) if n.isdigit() else 0)
    for tag in sorted(inst, key=key):
        pi, ps = 100 * inst[tag] / ti, 100 * samp[tag] / ts
        if pi >= min_pct or ps >= min_pct:
            print(f"{tag:28s} {pi:8.2f} {ps:10.2f}")


if __name__ == "__main__":
    main()
