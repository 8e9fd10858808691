#!/usr/bin/env python
"""Everything profiles/ holds for one build, regenerated from the captures of ONE gpurun call.

    python tools/ncu_profiles.py TAG gpurun_out/TAG_path.ncu-rep [gpurun_out/TAG_launches.csv] [--frames 300]

writes profiles/TAG_metrics.md   per (kernel, grid) means of the `ncu --set full` capture (duration, DRAM bytes,
                                  DRAM %, issue-active %, occupancy, registers, warp instructions)
       profiles/TAG_stalls.md    per kernel, the source lines with the most stall samples / instructions
       profiles/TAG_launches.md  the launch list (gpu__time_duration pass) separated by grid size
       profiles/TAG_sass.md      opcode histogram of libsvb200.so (cuobjdump -sass): UBLKCP / SYNCS / UCGABAR / REDUX ...
       profiles/traffic.json     DRAM read / write bytes per launch of the path kernels (what bench.py's roofline.traffic reads)
"""
import collections
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH_KERNELS = {"fill_seq": "fill_seq", "proj_write": "proj_write", "proj_scan": "proj_scan", "ransac_score": "ransac_score",
                "ransac_finalize": "ransac_finalize", "ransac_replay": "ransac_replay", "road_sites_near": "road_sites",
                "road_sites_scatter": "road_sites_scatter", "road_sites_fused": "road_sites"}


MAX_ID = None          # --max-id N: only the first N + 1 captured launches (one step) enter the metrics table / traffic.json


def short(name):
    n = name.split("(")[0]
    n = re.sub(r"^void\s+", "", n)
    return n.replace("<unnamed>::", "")


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return float("nan")


def metrics(rep, tag, frames):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    col = {n: i for i, n in enumerate(hdr)}
    want = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "MB read"), ("dram__bytes_write.sum", "MB written"),
            ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
            ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
            ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
            ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp instr"),
            ("lts__t_sectors.sum", "L2 sectors"),
            ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts")]
    want = [(m, l) for m, l in want if m in col]
    agg = collections.OrderedDict()
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        if MAX_ID is not None and int(r[col["ID"]]) > MAX_ID:
            continue
        key = (short(r[col["Kernel Name"]]), r[col["Grid Size"]].replace(" ", ""), r[col["launch__block_size"]])
        vals = []
        for m, _ in want:
            v = num(r[col[m]])
            u = units[col[m]].lower()
            if m.startswith("dram__bytes"):
                v *= {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}.get(u, 1.0)
            if m == "gpu__time_duration.sum":
                v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
            vals.append(v)
        agg.setdefault(key, []).append(vals)
    lines = ["# %s: ncu --set full --clock-control none (per launch, mean over the captured launches; %d frames per launch for the path kernels)" % (tag, frames), "",
             "| kernel | grid | block | launches | " + " | ".join(l for _, l in want) + " |", "|---|---|---|---|" + "---|" * len(want)]
    traffic = {}
    for (k, grid, blk), v in agg.items():
        n = len(v)
        mean = [sum(x[i] for x in v) / n for i in range(len(want))]
        lines.append("| `%s` | %s | %s | %d | " % (k, grid, blk, n) + " | ".join("%.4g" % x for x in mean) + " |")
        for pat, name in PATH_KERNELS.items():
            if k.startswith(pat) and name not in traffic:
                d = dict(zip([m for m, _ in want], mean))
                traffic[name] = {"kernel": k, "grid": grid, "dram_read_bytes": d["dram__bytes_read.sum"] * 1e6,
                                 "dram_write_bytes": d["dram__bytes_write.sum"] * 1e6, "ncu_duration_us": d["gpu__time_duration.sum"],
                                 "warp_instructions": d.get("smsp__inst_executed.sum"),
                                 "issue_active_pct": d.get("smsp__issue_active.avg.pct_of_peak_sustained_active")}
    open(os.path.join(ROOT, "profiles", tag + "_metrics.md"), "w").write("\n".join(lines) + "\n")
    json.dump({"source": "profiles/%s_metrics.md (ncu --set full --clock-control none --import-source on, python bench.py --steps 1 "
                         "--warmup 3 --path-only --no-e2e --no-cpu-baseline, %d frames per launch)" % (tag, frames),
               "frames": frames, "kernels": traffic}, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    return [k for (k, _, _) in agg]


def stalls(rep, tag, kernels):
    lines = ["# %s: source lines by stall samples (ncu --import-source on, -lineinfo), first captured launch of each kernel" % tag, ""]
    seen = set()
    for k in kernels:
        base = re.split(r"[<(]", k)[0]
        if base in seen or base.startswith("at::") or "elementwise" in base or "reduce_kernel" in base:
            continue
        seen.add(base)
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_hot.py"), rep, base, "8", "stall"], capture_output=True, text=True)
        if r.returncode == 0 and r.stdout.strip():
            lines += ["## `%s`" % base, "```", r.stdout.rstrip(), "```", ""]
    open(os.path.join(ROOT, "profiles", tag + "_stalls.md"), "w").write("\n".join(lines) + "\n")


def launches(path, tag):
    rows = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(rows):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            key = (short(row["Kernel Name"]), row.get("Grid Size", "").replace(" ", ""))
            v = num(row["Metric Value"]) * {"ns": 1.0, "us": 1e3, "ms": 1e6}.get(row.get("Metric Unit", "ns"), 1.0)
            agg.setdefault(key, []).append(v)
    tot = sum(sum(v) for v in agg.values())
    lines = ["# %s: launch list (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised: shares, not absolutes)" % tag, "",
             "| kernel | grid | launches | mean ns | share of profiled time |", "|---|---|---|---|---|"]
    for (k, g), v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        lines.append("| `%s` | %s | %d | %.0f | %.3f |" % (k, g, len(v), sum(v) / len(v), sum(v) / tot))
    open(os.path.join(ROOT, "profiles", tag + "_launches.md"), "w").write("\n".join(lines) + "\n")


def sass(tag):
    so = os.path.join(ROOT, "stereo.vision_b200", "libsvb200.so")
    out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    per, cur = collections.OrderedDict(), None
    for l in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", l)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            per[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
        if m and cur:
            per[cur][m.group(1).split(".")[0]] += 1
    watch = ["UBLKCP", "SYNCS", "UCGABAR", "REDUX", "MATCH", "VIMNMX3", "VIMNMX", "ATOMS", "RED", "ATOMG", "LDGSTS", "UTMALDG", "UTCMMA", "LDTM",
             "DFMA", "DMUL", "DADD", "IMAD", "LDG", "STG", "LDS", "STS", "SHFL", "PRMT", "BAR", "FFMA"]
    lines = ["# %s: SASS opcode counts per kernel of libsvb200.so (cuobjdump -sass, sm_100a; static instruction counts)" % tag, "",
             "`UBLKCP` = bulk async copy (cp.async.bulk), `SYNCS` = mbarrier, `UCGABAR` = cluster barrier, `REDUX` = warp reduce, `MATCH` = match.any.",
             "No `UTMALDG` / `UTCMMA` / `LDTM`: the path has no contraction and no 2-D tile movement (DESIGN 3).", "",
             "| kernel | total | " + " | ".join(watch) + " |", "|---|---|" + "---|" * len(watch)]
    for k, c in per.items():
        lines.append("| `%s` | %d | " % (k[:60], sum(c.values())) + " | ".join(str(c.get(w, 0)) for w in watch) + " |")
    open(os.path.join(ROOT, "profiles", tag + "_sass.md"), "w").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    tag, rep = sys.argv[1], sys.argv[2]
    frames = int(sys.argv[sys.argv.index("--frames") + 1]) if "--frames" in sys.argv else 300
    if "--max-id" in sys.argv:
        MAX_ID = int(sys.argv[sys.argv.index("--max-id") + 1])
    ks = metrics(rep, tag, frames)
    stalls(rep, tag, ks)
    csvs = [a for a in sys.argv[3:] if a.endswith(".csv")]
    if csvs:
        launches(csvs[0], tag)
    sass(tag)
    print("profiles/%s_* written" % tag)
