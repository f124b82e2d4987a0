#!/usr/bin/env python
"""Per-source-line instruction and stall-sample shares from `ncu --page source --csv --print-source cuda,sass` output."""
import collections
import csv
import sys


def main(path, top=40):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
    hdr = rows[hi]
    ix = {h: i for i, h in enumerate(hdr)}
    agg = collections.defaultdict(lambda: [0, 0, ""])
    tot = tots = 0
    for r in rows[hi + 1:]:
        if len(r) < len(hdr):
            continue
        try:
            ln = int(r[0]); ie = int(r[ix["Instructions Executed"]]); s = int(r[ix["# Samples"]])
        except ValueError:
            continue
        a = agg[ln]; a[0] += ie; a[1] += s; a[2] = r[1]
        tot += ie; tots += s
    print("total warp instructions", tot, "samples", tots)
    for ln, (ie, s, src) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{ln:5d} {ie / tot * 100:5.1f}% inst {s / tots * 100:5.1f}% smp  {src.strip()[:120]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
