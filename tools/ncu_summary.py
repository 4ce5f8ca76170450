#!/usr/bin/env python3
"""Development aid: per-kernel averages of an `ncu --page raw --csv` export.
    ncu -i X.ncu-rep --page raw --csv > raw.csv ; tools/ncu_summary.py raw.csv out_summary.json
bench.py reads profiles/r*_full_summary.json for the `roofline.traffic` figure."""
import collections
import csv
import json
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__grid_size",
        "sm__inst_executed.avg.per_cycle_elapsed", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        # integer-pipe utilisation against the SM's peak (north_star), the other pipes, issue rate, cache hit rates
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_active.avg",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio"]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    want = [w for w in WANT if w in hdr]
    agg = collections.defaultdict(lambda: collections.defaultdict(list))
    for r in rows[2:]:
        k = r[hdr.index("Kernel Name")].split("(")[0]
        for w in want:
            try:
                agg[k][w].append(float(r[hdr.index(w)].replace(",", "")))
            except ValueError:
                pass
    out = {}
    for k, d in agg.items():
        out[k] = {"%s [%s]" % (w, units[hdr.index(w)]): round(sum(v) / len(v), 3) for w, v in d.items() if v}
        out[k]["launches_captured"] = len(d[want[0]])
    json.dump(out, open(sys.argv[2], "w"), indent=1)
    for k, v in out.items():
        print(k, v)


if __name__ == "__main__":
    main()
