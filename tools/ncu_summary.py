#!/usr/bin/env python3
"""Summarise ncu output for profiles/: per-kernel shares from a launch list (gpu__time_duration.sum CSV)
and the key metrics of a `--set full` report.

    python tools/ncu_summary.py --launches gpurun_out/launches.csv --rep gpurun_out/prof.ncu-rep > profiles/rNN_x.md
"""
import argparse
import collections
import csv
import re
import subprocess

KEYS = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|launch__registers_per_thread|launch__occupancy_limit_\w+|"
    r"launch__grid_size|launch__block_size|launch__shared_mem_per_block_dynamic|"
    r"sm__warps_active\.avg\.pct_of_peak_sustained_active|smsp__issue_active\.avg\.pct_of_peak_sustained_active|"
    r"smsp__inst_executed\.sum|sm__inst_executed_pipe_\w+\.sum|sm__pipe_(alu|fma|fmaheavy|fmalite|fp64)_cycles_active\.avg\.pct_of_peak_sustained_active|"
    r"sm__inst_executed_pipe_\w+\.avg\.pct_of_peak_sustained_active|"
    r"smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio|sass__inst_executed_local_(loads|stores)|"
    r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"smsp__cycles_active\.avg|sm__cycles_elapsed\.max|gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed)$")


def launches(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    r = list(csv.reader(lines))
    hdr = r[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for x in r[1:]:
        if len(x) <= vi:
            continue
        name = re.sub(r"\(.*", "", x[ki])
        v = float(x[vi].replace(",", ""))
        v *= {"us": 1e-3, "ns": 1e-6, "ms": 1.0, "s": 1e3}.get(x[ui], 1.0)
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    print("## Launch list (ncu --metrics gpu__time_duration.sum --clock-control none): per-kernel share\n")
    print("| kernel | launches | total ms | share | avg ms |\n|---|---:|---:|---:|---:|")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| `%s` | %d | %.3f | %.1f%% | %.4f |" % (k[:70], v[0], v[1], 100 * v[1] / tot, v[1] / v[0]))
    print()


def full(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    print("## `ncu --set full --clock-control none` key metrics (%s)\n" % path)
    names = [re.sub(r"\(.*", "", d[ki]).replace("void ", "") for d in data]
    print("| metric | unit | " + " | ".join("`%s`" % n[:40] for n in names) + " |")
    print("|---|---|" + "---:|" * len(names))
    for i, h in enumerate(hdr):
        if KEYS.match(h):
            vals = [d[i] for d in data]
            if all(v in ("0", "0.000000", "") for v in vals) and "stalled" in h:
                continue
            print("| %s | %s | %s |" % (h, units[i], " | ".join(vals)))
    print()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--launches")
    ap.add_argument("--rep")
    a = ap.parse_args()
    if a.launches:
        launches(a.launches)
    if a.rep:
        full(a.rep)
