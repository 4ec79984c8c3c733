#!/usr/bin/env python
"""Summarise ncu output for profiles/: (1) a launch list CSV (--metrics gpu__time_duration.sum)
-> per-kernel totals and shares; (2) a --set full report (.ncu-rep) -> the roofline counters of
each captured launch.  Usage:
    python tools/ncu_summary.py launches gpurun_out/x_launches.csv
    python tools/ncu_summary.py full gpurun_out/x_prof.ncu-rep
"""
import collections
import csv
import io
import re
import subprocess
import sys

FULL = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio",
    "smsp__average_warp_latency_issue_stalled_barrier.ratio",
    "smsp__average_warp_latency_issue_stalled_wait.ratio",
    "smsp__average_warp_latency_issue_stalled_not_selected.ratio",
    "smsp__average_warp_latency_issue_stalled_lg_throttle.ratio",
    "smsp__average_warp_latency_issue_stalled_mio_throttle.ratio",
]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        name = re.sub(r"^void ", "", row["Kernel Name"])
        name = name.replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        m = re.match(r"(b200gat::k_\w+)(<[^>]*>)?", name)        # template arguments kept for our kernels (tile shapes)
        name = (m.group(1) + (m.group(2) or "")) if m else re.sub(r"[<(].*", "", name)
        name = name[:72]
        tot[name] += v
        cnt[name] += 1
    T = sum(tot.values())
    print(f"# {path}: {sum(cnt.values())} launches, {T:.1f} us of kernel time (cold-cache, serialised)")
    print(f"{'total_us':>12} {'share':>7} {'n':>6} {'avg_us':>10}  kernel")
    for k, v in sorted(tot.items(), key=lambda x: -x[1]):
        print(f"{v:12.1f} {100 * v / T:6.1f}% {cnt[k]:6d} {v / cnt[k]:10.1f}  {k}")


def full(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    print(f"# {path}: {len(data)} captured launch(es), ncu --set full --clock-control none")
    names = [re.sub(r"\(.*", "", d[col["Kernel Name"]])[:40] for d in data]
    print("metric | unit | " + " | ".join(names))
    for m in FULL:
        if m in col:
            i = col[m]
            print(f"{m} | {units[i]} | " + " | ".join(d[i] for d in data))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
