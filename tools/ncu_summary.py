#!/usr/bin/env python
"""Summaries for profiles/: per-kernel shares from an ncu launch list (csv) and key metrics from an .ncu-rep.

    python tools/ncu_summary.py launches gpurun_out/launches.csv  > profiles/rNN_launches.md
    python tools/ncu_summary.py metrics  gpurun_out/prof.ncu-rep  > profiles/rNN_metrics.md
"""
import collections
import csv
import subprocess
import sys

mode, path = sys.argv[1], sys.argv[2]
if mode == "launches":
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            agg.setdefault(row["Kernel Name"].split("(")[0], []).append(float(row["Metric Value"].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    print("| kernel | launches | mean ns | share of profiled time |\n|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print("| `%s` | %d | %.0f | %.3f |" % (k, len(v), sum(v) / len(v), sum(v) / tot))
else:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    hdr = r[0]
    want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "smsp__inst_executed.sum", "lts__t_sectors.sum"]
    idx = [hdr.index(w) for w in want if w in hdr]
    print("| " + " | ".join(hdr[i] for i in idx) + " |")
    print("| " + " | ".join(r[1][i] for i in idx) + " |")
    print("|" + "---|" * len(idx))
    for row in r[2:]:
        print("| " + " | ".join(row[i].split("(")[0][:40] for i in idx) + " |")
