#!/usr/bin/env python
"""Summarise ncu output brought back from the GPU box (gpurun_out/) into small tracked text files.

    python profiles/summarize.py launches gpurun_out/launches.csv  > profiles/rNN_launches.txt
    python profiles/summarize.py full gpurun_out/prof.ncu-rep      > profiles/rNN_<kernel>_full.txt

`launches`: the `--metrics gpu__time_duration.sum` pass -> per-kernel count, total/mean time, share.
`full`    : the `--set full` capture -> the raw-page metrics the roofline uses (needs ncu on PATH).
"""
import csv
import re
import subprocess
import sys
from collections import OrderedDict

KEEP = re.compile(
    r"dram__bytes_(read|write)\.sum$|dram__cycles_active|gpu__dram_throughput|gpu__time_duration\.sum|"
    r"sm__pipe_tensor.*cycles_active|sm__warps_active\.avg\.pct|launch__registers_per_thread|launch__grid_size|"
    r"launch__block_size|launch__shared_mem_per_block|sm__throughput\.avg\.pct|lts__t_bytes\.sum$|"
    r"l1tex__data_bank_conflicts|smsp__inst_executed\.sum$|sm__inst_executed_pipe_(fma|alu|fp64|lsu|uniform|tensor).*\.sum$|"
    r"smsp__cycles_active\.avg$|launch__occupancy_limit|sm__cycles_elapsed\.max$|lts__t_sector_hit_rate|"
    r"smsp__average_warp.*issue_stalled.*_per_warp_active|smsp__issue_active\.avg\.pct")


def short(name):
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("void ", "").replace("unnamed>::", "").replace("yupvq::", "").replace("<unnamed>::", "")
    return name.strip()


def launches(path):
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r["Metric Name"] == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", ""))
            if r["Metric Unit"] in ("us", "usecond"):
                v *= 1e3
            elif r["Metric Unit"] in ("ms", "msecond"):
                v *= 1e6
            rows.append((short(r["Kernel Name"]), r["Grid Size"], r["Block Size"], v))
    agg = OrderedDict()
    for name, grid, block, ns in rows:
        a = agg.setdefault(name, {"n": 0, "ns": 0.0, "grid": grid, "block": block})
        a["n"] += 1
        a["ns"] += ns
    total = sum(a["ns"] for a in agg.values()) or 1.0
    print(f"# {path}: {len(rows)} launches, {total / 1e6:.3f} ms summed device time (ncu: cold-cache, serialised)")
    print(f"{'kernel':58s} {'launches':>8s} {'total_ms':>10s} {'mean_us':>10s} {'share':>7s}  grid block")
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["ns"]):
        print(f"{name[:58]:58s} {a['n']:8d} {a['ns'] / 1e6:10.3f} {a['ns'] / a['n'] / 1e3:10.2f} "
              f"{100 * a['ns'] / total:6.1f}%  {a['grid']} {a['block']}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    lines = [ln for ln in out.splitlines() if ln.startswith('"')]
    rd = list(csv.reader(lines))
    head, units, data = rd[0], rd[1], rd[2:]
    for row in data:
        kn = row[head.index("Kernel Name")]
        print(f"## {short(kn)}  grid={row[head.index('Grid Size')]} block={row[head.index('Block Size')]}")
        for h, u, v in zip(head, units, row):
            if KEEP.search(h):
                print(f"{h:90s} {v:>18s} {u}")
        print()


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
