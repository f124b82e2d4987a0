#!/usr/bin/env python
"""Summarises an `ncu --set full` report: per-kernel duration, DRAM traffic, issue utilisation, occupancy, hit rates.

  ncu -i gpurun_out/X.ncu-rep --page raw --csv > /tmp/raw.csv
  python tools/summarise_full.py /tmp/raw.csv > profiles/<name>.md          (table, mean over the captured launches)
  python tools/summarise_full.py /tmp/raw.csv --traffic profiles/traffic.json   (also writes the per-launch DRAM bytes)
"""
import collections
import csv
import json
import sys

COLS = [
    ("gpu__time_duration.sum", "time us", 1.0),
    ("dram__bytes_read.sum", "dram rd MB", 1.0),
    ("dram__bytes_write.sum", "dram wr MB", 1.0),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %", 1.0),
    ("smsp__inst_executed.sum", "warp inst", 1.0),
    ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue %", 1.0),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps act %", 1.0),
    ("launch__registers_per_thread", "regs", 1.0),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %", 1.0),
    ("lts__t_sector_hit_rate.pct", "L2 hit %", 1.0),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64 pipe %", 1.0),
]
UNIT_SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}


def main(path, traffic_out=None):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    agg = collections.OrderedDict()
    for d in data:
        if len(d) != len(hdr):
            continue
        k = d[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("pcp::", "")
        vals = []
        for m, _, _ in COLS:
            if m not in ix:
                vals.append(float("nan"))
                continue
            v = float(d[ix[m]].replace(",", "") or "nan")
            vals.append(v * UNIT_SCALE.get(units[ix[m]], 1.0))
        a = agg.setdefault(k, [0, [0.0] * len(COLS)])
        a[0] += 1
        a[1] = [x + y for x, y in zip(a[1], vals)]
    print(f"source: {path} (ncu --set full --clock-control none; mean over the captured launches of each kernel)\n")
    print("| kernel | launches | " + " | ".join(c[1] for c in COLS) + " |")
    print("|---|---:|" + "---:|" * len(COLS))
    out = {}
    for k, (n, s) in agg.items():
        m = [x / n for x in s]
        print(f"| {k} | {n} | " + " | ".join(f"{x:.2f}" if x < 1e5 else f"{x:.3e}" for x in m) + " |")
        out[k] = {"dram_bytes_per_launch": (m[1] + m[2]) * 1e6, "time_us": m[0]}
    # stall reasons: warps stalled per issue-active cycle, by reason (smsp__average_warps_issue_stalled_*_per_issue_active.ratio)
    stall_cols = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and "_not_issued" not in h]
    if stall_cols:
        sagg = collections.OrderedDict()
        for d in data:
            if len(d) != len(hdr):
                continue
            k = d[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("pcp::", "")
            a = sagg.setdefault(k, [0, [0.0] * len(stall_cols)])
            a[0] += 1
            a[1] = [x + float(d[ix[c]].replace(",", "") or 0.0) for x, c in zip(a[1], stall_cols)]
        print("\nStall reasons (warps stalled per issue-active cycle; the five largest per kernel):\n")
        print("| kernel | " + " | ".join(f"reason {i + 1}" for i in range(5)) + " |")
        print("|---|" + "---|" * 5)
        for k, (n, s) in sagg.items():
            m = sorted(((x / n, c[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]) for x, c in zip(s, stall_cols)), reverse=True)[:5]
            print(f"| {k} | " + " | ".join(f"{nm} {v:.2f}" for v, nm in m) + " |")
    if traffic_out:
        pass1 = [v for k, v in out.items() if k.startswith("scan") or k.startswith("classify")]  # pass 1 = classify + scan
        doc = {"source": path, "kernels": out}
        if pass1:
            doc["scan_kernel_dram_bytes_per_launch"] = sum(v["dram_bytes_per_launch"] for v in pass1)
            doc["source_note"] = "profiles/traffic.json: dram__bytes_read.sum + dram__bytes_write.sum per frame of classify_kernel + scan kernel, from " + path
        json.dump(doc, open(traffic_out, "w"), indent=1)


if __name__ == "__main__":
    t = None
    if "--traffic" in sys.argv:
        t = sys.argv[sys.argv.index("--traffic") + 1]
    main(sys.argv[1], t)
