#!/usr/bin/env python
"""Side-by-side of the headline ncu raw metrics of two `ncu --page raw --csv` dumps (first kernel of each)."""
import csv, re, sys
def load(f):
    rows = list(csv.reader(open(f)))
    return dict(zip(rows[0], rows[2]))
a, b = load(sys.argv[1]), load(sys.argv[2])
pat = (r'issue_stalled.*per_issue_active|smsp__inst_executed.sum$|gpu__time_duration.sum|sm__inst_executed_pipe_(alu|fma|lsu|xu).sum.pct_of_peak_sustained_active'
       r'|l1tex__data_pipe_lsu_wavefronts.sum.pct|smsp__issue_active.avg.pct|l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate|dram__bytes_read.sum$|dram__bytes_write.sum$'
       r'|lts__t_sector_hit_rate.pct|registers_per_thread$|warps_active.avg.pct')
for k in a:
    if re.search(pat, k):
        print('%-90s %14s %14s' % (k, a.get(k), b.get(k)))
