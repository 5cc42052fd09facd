#!/usr/bin/env python3
"""Print the metrics we care about from an .ncu-rep (run where ncu is installed; no GPU needed)."""
import csv
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.pct',
        'l1tex__average_t_sectors_per_request_pipe_lsu_mem_global_op_ld.ratio', 'l1tex__average_t_sectors_per_request_pipe_lsu_mem_global_op_st.ratio']
for r in rows[2:]:
    print("==", r[hdr.index('Kernel Name')][:100])
    for m in want:
        if m in hdr:
            print("   %-78s %s %s" % (m, r[hdr.index(m)], units[hdr.index(m)]))
    st = [(float(r[i]), hdr[i]) for i in range(len(hdr)) if 'smsp__pcsamp_warps_issue_stalled' in hdr[i] and 'not_issued' not in hdr[i] and r[i] not in ('', 'n/a')]
    st.sort(reverse=True)
    tot = sum(v for v, _ in st) or 1
    print("   stalls: " + ", ".join("%s %.0f%%" % (n.replace('smsp__pcsamp_warps_issue_stalled_', ''), 100 * v / tot) for v, n in st[:7]))
