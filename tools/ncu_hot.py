#!/usr/bin/env python
"""Per-source-line hot spots of one kernel in an .ncu-rep (needs -lineinfo + --import-source on).

    python tools/ncu_hot.py gpurun_out/prof.ncu-rep proj_write [top]
Sums 'Instructions Executed' and stall samples of the SASS lines under every CUDA source line.
"""
import collections
import csv
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
by = 1 if (len(sys.argv) > 4 and sys.argv[4] == "stall") else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                      "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr_i = next(i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r)
hdr = rows[hdr_i]
ci = {n: hdr.index(n) for n in ("Source", "Instructions Executed", "# Samples", "Address")}
agg = collections.OrderedDict()
cur = None
first_kernel_done = False
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr):
        if r and r[0] == "Kernel Name" and agg:
            break                           # only the first launch of the kernel
        continue
    src, addr = r[ci["Source"]], r[ci["Address"]]
    try:
        ins, smp = float(r[ci["Instructions Executed"]] or 0), float(r[ci["# Samples"]] or 0)
    except ValueError:
        continue
    if not addr.startswith("0x"):           # a CUDA source line: "file:line" in Address
        cur = (addr, src.strip())
        agg.setdefault(cur, [0.0, 0.0])
    elif cur is not None:
        agg[cur][0] += ins
        agg[cur][1] += smp
tot_i = sum(v[0] for v in agg.values()) or 1
tot_s = sum(v[1] for v in agg.values()) or 1
print("total warp instructions %.3g, samples %.0f" % (tot_i, tot_s))
for (addr, src), (ins, smp) in sorted(agg.items(), key=lambda kv: -kv[1][by])[:top]:
    print("%5.1f%% inst %5.1f%% stall  %-28s %s" % (100 * ins / tot_i, 100 * smp / tot_s, addr[-28:], src[:110]))
