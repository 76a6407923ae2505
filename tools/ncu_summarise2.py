"""Summarise an .ncu-rep (raw page) per kernel: time, occupancy, pipe utilisation, DRAM bytes, shared-memory bank
conflicts and the stall mix.  usage: tools/ncu_summarise2.py report.ncu-rep [out.json]"""
import csv, json, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
keep = ['Kernel Name', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread', 'gpu__time_duration.sum',
        'smsp__inst_executed.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']
stalls = [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio')]
out = []
for r in rows[2:]:
    d = {k: r[idx[k]] for k in keep if k in idx}
    d['stalls_per_issue'] = {h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]: round(float(r[idx[h]]), 3)
                             for h in stalls if float(r[idx[h]] or 0) > 0.05}
    out.append(d)
if len(sys.argv) > 2:
    json.dump({"source": "ncu --set full --clock-control none: " + sys.argv[1], "kernels": out}, open(sys.argv[2], "w"), indent=1)
for d in out:
    print(json.dumps(d, indent=1))
