#!/usr/bin/env python
"""Turns ncu output into the text summaries kept under profiles/ (the .ncu-rep files stay in gpurun_out/).

    python profiles/summarize.py launches gpurun_out/launches.csv          # per-kernel share of a step
    python profiles/summarize.py full gpurun_out/prof.ncu-rep [...]        # key metrics of `ncu --set full`
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]


def launches(path):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr = rows[hi]
    d = collections.OrderedDict()
    for r in rows[hi + 1:]:
        if len(r) < len(hdr):
            continue
        rec = dict(zip(hdr, r))
        try:
            v = float(rec["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        if rec.get("Metric Unit", "") in ("us", "usecond"):
            v *= 1e3
        elif rec.get("Metric Unit", "") in ("ms", "msecond"):
            v *= 1e6
        d.setdefault(rec["Kernel Name"].split("(")[0][:70], []).append(v)
    tot = sum(sum(v) for v in d.values())
    print(f"# {path}: {sum(len(v) for v in d.values())} launches, {tot / 1e6:.2f} ms of device time (ncu, serialised, cold cache)")
    print(f"{'kernel':70s} {'n':>5s} {'avg us':>10s} {'total ms':>10s} {'share':>7s}")
    for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k:70s} {len(v):5d} {sum(v) / len(v) / 1e3:10.1f} {sum(v) / 1e6:10.2f} {100 * sum(v) / tot:6.1f}%")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {path}")
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print(f"## {d['Kernel Name'][:100]}  (launch id {d.get('ID')})")
        for k in KEYS:
            if k in d and d[k] != "":
                print(f"  {k:90s} {d[k]:>18s} {units[hdr.index(k)]}")


if __name__ == "__main__":
    mode = sys.argv[1]
    for p in sys.argv[2:]:
        launches(p) if mode == "launches" else full(p)
