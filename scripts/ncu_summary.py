#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into the few lines profiles/*.ncu_summary.txt hold: duration,
occupancy limiters, pipe utilisation, memory traffic, the top warp-stall reasons, tcgen05/TMA instruction counts."""
import csv
import sys

KEEP = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.avg", "smsp__issue_active.avg.per_cycle_active",
        "sm__inst_executed.avg.per_cycle_elapsed"]


def main():
    rows = list(csv.reader(open(sys.argv[1], newline="")))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"]
    if not hdr:
        sys.exit("no header row")
    h = hdr[0]
    names, units = rows[h], rows[h + 1]
    for r in rows[h + 2:]:
        if len(r) != len(names):
            continue
        d = dict(zip(names, r))
        u = dict(zip(names, units))
        print("----")
        print("  Kernel Name ", d.get("Kernel Name"))
        for k in KEEP:
            if k in d and d[k] != "":
                print(f"  {k} {u.get(k, '')} {d[k]}")
        stalls = []
        for k, v in d.items():
            if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio"):
                try:
                    stalls.append((float(v.replace(",", "")), k[len("smsp__average_warps_issue_stalled_"):]))
                except ValueError:
                    pass
        for v, k in sorted(stalls, reverse=True)[:8]:
            print(f"   stall {k:<48s} {v:.2f}")


if __name__ == "__main__":
    main()
