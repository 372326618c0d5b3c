#!/usr/bin/env python
"""Extracts the counters SURVEY.md section 8(d) asks for from an `ncu --set full` report, one column per profiled launch.

    python tools/ncu_counters.py gpurun_out/x.ncu-rep profiles/r02_k3_sort_counters.csv
"""
import csv
import subprocess
import sys

METRICS = """gpu__time_duration.sum launch__grid_size launch__block_size launch__registers_per_thread smsp__inst_executed.sum
smsp__issue_active.avg.pct_of_peak_sustained_active sm__inst_executed.avg.per_cycle_active
sm__warps_active.avg.pct_of_peak_sustained_active smsp__warps_eligible.avg.per_cycle_active
l1tex__data_pipe_lsu_wavefronts_mem_shared.sum smsp__inst_executed_op_shared_atom.sum
l1tex__throughput.avg.pct_of_peak_sustained_elapsed lts__throughput.avg.pct_of_peak_sustained_elapsed dram__bytes_read.sum
dram__bytes_write.sum lts__t_sector_hit_rate.pct l1tex__t_sector_hit_rate.pct smsp__average_warp_latency_per_inst_issued.ratio
smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio
smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio""".split()


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    head, units, launches = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(head)}
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [r[col["Kernel Name"]].split("::")[-1] for r in launches])
        w.writerow(["Kernel Name", ""] + [r[col["Kernel Name"]] for r in launches])
        for m in METRICS:
            if m in col:
                w.writerow([m, units[col[m]]] + [r[col[m]] for r in launches])
    print(open(out).read())


if __name__ == "__main__":
    main()
