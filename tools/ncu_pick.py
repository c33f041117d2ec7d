#!/usr/bin/env python3
"""Print a focused subset of an `ncu --page raw --csv` dump, one block per captured launch."""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
pat = sys.argv[2] if len(sys.argv) > 2 else (
    r'^(Kernel Name|gpu__time_duration.sum|l1tex__throughput.avg.pct_of_peak_sustained_elapsed|l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed'
    r'|l1tex__t_sectors_pipe_lsu_mem_(global|local)_op_(ld|st)\.sum|l1tex__t_requests_pipe_lsu_mem_(global|local)_op_(ld|st)\.sum'
    r'|smsp__inst_executed_op_local_(ld|st).sum|smsp__inst_executed_op_global_(ld|st).sum|smsp__inst_executed.sum'
    r'|smsp__thread_inst_executed_per_inst_executed.ratio|sm__warps_active.avg.pct_of_peak_sustained_active|launch__registers_per_thread'
    r'|smsp__average_warps_issue_stalled_.*_per_issue_active.ratio|smsp__issue_active.avg.pct_of_peak_sustained_active'
    r'|l1tex__t_sector_hit_rate.pct|lts__t_sector_hit_rate.pct|l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed'
    r'|.*breakdown.*|sm__inst_executed_pipe_.*pct_of_peak_sustained_active|dram__bytes_(read|write).sum'
    r'|sm__throughput.avg.pct_of_peak_sustained_elapsed|launch__occupancy_limit.*|sm__cycles_elapsed.avg)$')
keys = [i for i, h in enumerate(hdr) if re.search(pat, h)]
for r in data:
    print('-----')
    for i in keys:
        v = r[i]
        if v not in ('', '0', 'n/a'):
            print(hdr[i], '=', v[:160], units[i])
