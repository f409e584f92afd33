#!/usr/bin/env python
"""Turn the raw ncu outputs a GPU-box pass left in gpurun_out/ into the small tracked summaries under profiles/.

    python scripts/summarize_profiles.py <tag> [workload]      e.g.  r01_v4 base16

  gpurun_out/launches_<wl>.csv   (ncu --metrics gpu__time_duration.sum --csv)  -> profiles/<tag>_launches_<wl>.csv (copy)
                                                                                   profiles/<tag>_launch_summary_<wl>.csv
  gpurun_out/prof_<wl>.ncu-rep   (ncu --set full)                               -> profiles/<tag>_ncu_full_<wl>.csv
  gpurun_out/bench_<wl>.json                                                     -> profiles/<tag>_bench_<wl>.json
"""
import csv
import io
import os
import re
import shutil
import subprocess
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GO, PR = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
FULL_METRICS = [
    "gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sectors.sum.pct_of_peak_sustained_elapsed", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "sm__cycles_elapsed.avg.per_second",
]


def short(name):
    name = re.sub(r"^void\s+", "", name)
    name = name.replace("unnamed>::", "").replace("ufm::<", "").replace("(anonymous namespace)::", "")
    return re.sub(r"\(.*\)$", "", name).strip()


def launches(tag, wl):
    src = os.path.join(GO, f"launches_{wl}.csv")
    if not os.path.exists(src):
        return
    shutil.copy(src, os.path.join(PR, f"{tag}_launches_{wl}.csv"))
    rows = [l for l in open(src) if l.startswith('"')]
    agg = OrderedDict()
    for r in csv.DictReader(io.StringIO("".join(rows))):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        a = agg.setdefault(short(r["Kernel Name"]), [0, 0.0])
        a[0] += 1
        a[1] += float(r["Metric Value"]) / 1e3
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(PR, f"{tag}_launch_summary_{wl}.csv"), "w") as f:
        f.write(f"# {tag}: ncu launch list of one forward pass, workload {wl} (cold-cache, serialised per-launch times: compare SHARES)\n")
        f.write("# command: ncu --metrics gpu__time_duration.sum --clock-control none -s 2L -c L python scripts/one_step.py --workload " + wl + "\n")
        f.write("kernel,launches,total_us,share\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k},{a[0]},{a[1]:.1f},{a[1] / tot:.3f}\n")
        f.write(f"total,,{tot:.1f},1.000\n")
    print("launch summary:", len(agg), "kernels", f"{tot / 1e3:.2f} ms")


def full(tag, wl, rep=None, out=None):
    rep = rep or os.path.join(GO, f"prof_{wl}.ncu-rep")
    if not os.path.exists(rep):
        return
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    cols = [m for m in FULL_METRICS if m in ix]
    out = out or os.path.join(PR, f"{tag}_ncu_full_{wl}.csv")
    with open(out, "w") as f:
        f.write(f"# {tag}: ncu --set full --clock-control none --import-source on ({os.path.basename(rep)}); one row per captured launch\n")
        w = csv.writer(f)
        w.writerow(["kernel"] + [f"{c} [{units[ix[c]]}]" for c in cols])
        for r in data:
            w.writerow([short(r[ix["Kernel Name"]])] + [r[ix[c]] for c in cols])
    print("ncu full summary:", len(data), "launches ->", os.path.relpath(out, ROOT))


def main():
    tag = sys.argv[1]
    wl = sys.argv[2] if len(sys.argv) > 2 else "base16"
    os.makedirs(PR, exist_ok=True)
    launches(tag, wl)
    full(tag, wl)
    b = os.path.join(GO, f"bench_{wl}.json")
    if os.path.exists(b):
        line = [l for l in open(b).read().splitlines() if l.startswith("{")][-1]
        open(os.path.join(PR, f"{tag}_bench_{wl}.json"), "w").write(line + "\n")
        print("bench line copied")
    for extra in sys.argv[3:]:  # extra .ncu-rep files: name=path
        name, path = extra.split("=")
        full(tag, wl, path, os.path.join(PR, f"{tag}_ncu_full_{name}.csv"))


if __name__ == "__main__":
    main()
