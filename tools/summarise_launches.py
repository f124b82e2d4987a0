#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launch count, total and mean device time,
share of the listed launches.   python tools/summarise_launches.py gpurun_out/launches.csv > profiles/<name>.md"""
import collections
import csv
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = None
    agg = collections.OrderedDict()
    for r in rows:
        if len(r) > 5 and r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            if d.get("Metric Name") != "gpu__time_duration.sum":
                continue
            k = d["Kernel Name"].split("(")[0]
            v = float(d["Metric Value"].replace(",", ""))
            if d["Metric Unit"] == "us":
                v *= 1e3
            a = agg.setdefault(k, [0, 0.0, d["Grid Size"], d["Block Size"]])
            a[0] += 1
            a[1] += v
    tot = sum(a[1] for a in agg.values()) or 1.0
    print(f"source: {path} (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised launches)\n")
    print("| kernel | launches | total us | mean us | share | grid | block |")
    print("|---|---:|---:|---:|---:|---|---|")
    for k, (c, t, g, b) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"| {k} | {c} | {t / 1e3:.1f} | {t / c / 1e3:.2f} | {100 * t / tot:.1f}% | {g} | {b} |")


if __name__ == "__main__":
    main(sys.argv[1])
