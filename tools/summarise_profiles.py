#!/usr/bin/env python
"""Turn the ncu artefacts a GPU visit left in gpurun_out/ into tracked summaries under
profiles/.  Usage: tools/summarise_profiles.py <tag> [<report-name>:<kernel-regex> ...]
  launches_<tag>.csv            -> profiles/<tag>_launches.md   (per-kernel time shares)
  prof_<name>_<tag>.ncu-rep     -> profiles/<tag>_<name>.md     (key raw metrics + stall summary)
"""
import collections
import csv
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
GO = os.path.join(ROOT, "gpurun_out")

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "sm__cycles_elapsed.avg.per_second"]


def launches(tag):
    path = os.path.join(GO, f"launches_{tag}.csv")
    if not os.path.exists(path):
        return
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    ix = {h: i for i, h in enumerate(rows[0])}
    agg = collections.OrderedDict()
    for r in rows[1:]:
        a = agg.setdefault(r[ix["Kernel Name"]], [0, 0.0, r[ix["Grid Size"]], r[ix["Block Size"]]])
        a[0] += 1
        a[1] += float(r[ix["Metric Value"]].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(OUT, f"{tag}_launches.md"), "w") as f:
        f.write(f"# ncu launch list, {tag}\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` over "
                "`python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-render-e2e --no-e2e` "
                "(cold-cache, serialised: compare shares, not absolutes).\n\n"
                "| launches | total ms | share | avg us | grid | block | kernel |\n|---|---|---|---|---|---|---|\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| {a[0]} | {a[1] / 1e6:.3f} | {100 * a[1] / tot:.1f}% | {a[1] / a[0] / 1e3:.1f} | "
                    f"{a[2]} | {a[3]} | `{k[:110]}` |\n")


def report(tag, name):
    rep = os.path.join(GO, f"prof_{name}_{tag}.ncu-rep")
    if not os.path.exists(rep):
        return
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    with open(os.path.join(OUT, f"{tag}_{name}.md"), "w") as f:
        f.write(f"# ncu --set full, {name}, {tag}\n\n`ncu --set full --clock-control none --import-source on` "
                f"(report gpurun_out/prof_{name}_{tag}.ncu-rep, not tracked).\n\n")
        for r in rows[2:]:
            f.write(f"## `{r[hdr.index('Kernel Name')][:120]}`\n\n| metric | value | unit |\n|---|---|---|\n")
            for k in KEYS:
                if k in hdr:
                    f.write(f"| {k} | {r[hdr.index(k)]} | {units[hdr.index(k)]} |\n")
            f.write("\n")
        st = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_stalls.py"), rep, "30"],
                            capture_output=True, text=True).stdout
        f.write("## warp-stall sampling, first profiled launch (tools/ncu_stalls.py)\n\n```\n" + st + "```\n")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    tag = sys.argv[1]
    launches(tag)
    for name in sys.argv[2:] or ["extent", "render"]:
        report(tag, name)
    print(sorted(os.listdir(OUT)))
