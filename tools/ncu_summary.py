#!/usr/bin/env python
"""Condenses an ncu report (.ncu-rep) into the metrics profiles/README.md quotes. Runs without a GPU.
  python tools/ncu_summary.py gpurun_out/ncu/*.ncu-rep > profiles/ncu_round2.md
"""
import csv
import io
import os
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__shared_mem_per_block_dynamic", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def summarise(path):
    if path.endswith(".csv"):  # already condensed on the GPU box (tools/prof_all.sh)
        text = open(path).read()
    else:
        r = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True)
        if r.returncode != 0:
            return "%s: ncu -i failed: %s" % (path, r.stderr[-200:])
        text = r.stdout
    rows = list(csv.reader(io.StringIO(text)))
    if len(rows) < 3:
        return "%s: empty report" % path
    head, units, vals = rows[0], rows[1], rows[2]
    col = {h: i for i, h in enumerate(head)}
    out = ["### %s" % os.path.basename(path).replace(".ncu-rep", "").replace(".raw.csv", ""), "", "| metric | value |", "|---|---|"]
    out.append("| kernel | `%s` |" % vals[col.get("Kernel Name", 4)][:150])
    for m in METRICS:
        if m in col:
            out.append("| %s | %s %s |" % (m, vals[col[m]], units[col[m]]))
    return "\n".join(out) + "\n"


if __name__ == "__main__":
    print("# ncu --set full captures (one launch per kernel family, emulated ranks on one B200)\n")
    print("Produced by tools/prof_all.sh on a gpurun box, condensed here by tools/ncu_summary.py. Never a timing")
    print("source: the durations below are under the profiler (cold caches, serialised); timings are in README.md.\n")
    for p in sys.argv[1:]:
        print(summarise(p))
