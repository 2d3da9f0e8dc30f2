#!/usr/bin/env python3
"""Turn ncu outputs into the small text summaries committed under profiles/.
  python tools/ncu_summary.py launches <launches.csv>          per-kernel launch durations
  python tools/ncu_summary.py full <report.ncu-rep>            key metrics + top stalls per kernel"""
import csv
import subprocess
import sys
from collections import defaultdict


def launches(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = defaultdict(list)
    for row in csv.DictReader(lines):
        agg[row["Kernel Name"][:70]].append(float(row["Metric Value"].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    print(f"{'kernel':72s} {'n':>4s} {'mean us':>10s} {'min us':>10s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k:72s} {len(v):4d} {sum(v)/len(v)/1e3:10.2f} {min(v)/1e3:10.2f} {sum(v)/tot:7.1%}")


WANT = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__waves_per_multiprocessor",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.sum",
        "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_fmaheavy.sum",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed"]


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")]
    for r in rows[2:]:
        print("----", r[idx["Kernel Name"]][:110])
        for w in WANT:
            if w in idx:
                print(f"  {w:66s} {r[idx[w]]} {units[idx[w]]}")
        for v, h in sorted(((float(r[idx[h]]), h) for h in stall), reverse=True)[:6]:
            print(f"  stall/issue {h[34:-23]:28s} {v:.2f}")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
