"""Compact per-kernel summary (CSV on stdout) of an .ncu-rep captured with `ncu --set full`:
    python tools/ncu_summarise.py gpurun_out/x.ncu-rep > profiles/x_summary.csv
Duration, DRAM bytes, pipe utilisation, occupancy and the warp-stall breakdown (warps stalled per issue-active cycle)."""
import csv, io, subprocess, sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
] + [f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio" for s in (
    "long_scoreboard", "short_scoreboard", "barrier", "mio_throttle", "lg_throttle", "math_pipe_throttle", "wait",
    "not_selected", "dispatch_stall", "no_instruction", "branch_resolving", "drain", "membar", "sleeping")]

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
out = csv.writer(sys.stdout)
keys = [k for k in KEYS if k in hdr]
out.writerow(["kernel"] + [f"{k} [{units[hdr.index(k)]}]" for k in keys])
for r in rows[2:]:
    out.writerow([r[hdr.index("Kernel Name")]] + [r[hdr.index(k)] for k in keys])
