#!/usr/bin/env python
"""Condense `ncu -i rep --page raw --csv` of ONE launch into the metric/unit/value table kept under profiles/.

  python tools/ncu_summary.py gpurun_out/x.ncu-rep "header comment" > profiles/x.raw_summary.csv
"""
import csv, io, subprocess, sys

KEEP = ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor",
        "sm__pipe_tensor", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg")

def main():
    rep, note = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    print(f"# {note}")
    print("metric,unit,value")
    for name in ("Kernel Name", "Block Size", "Grid Size"):
        if name in hdr:
            print(f'{name},,"{vals[hdr.index(name)]}"')
    for h, u, v in zip(hdr, units, vals):
        if any(h.startswith(k) for k in KEEP):
            print(f'{h},{u},"{v}"')

if __name__ == "__main__":
    main()
