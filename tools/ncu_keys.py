#!/usr/bin/env python
"""Headline metrics of the first kernel of an `ncu --page raw --csv` dump.   python tools/ncu_keys.py raw.csv [raw2.csv]"""
import csv, re, sys
def load(f):
    rows = list(csv.reader(open(f)))
    return dict(zip(rows[0], rows[2]))
ds = [load(f) for f in sys.argv[1:]]
pat = (r'issue_stalled.*per_issue_active|smsp__inst_executed.sum$|gpu__time_duration.sum|sm__inst_executed_pipe_(alu|fma|lsu|xu).sum.pct_of_peak_sustained_active'
       r'|l1tex__data_pipe_lsu_wavefronts.sum.pct|smsp__issue_active.avg.pct|l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate|dram__bytes_read.sum$|dram__bytes_write.sum$'
       r'|lts__t_sector_hit_rate.pct|registers_per_thread$|warps_active.avg.pct|sass__inst_executed_local|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum$|shared_mem_per_block')
for k in ds[0]:
    if re.search(pat, k):
        print('%-88s %s' % (k, '  '.join('%14s' % d.get(k) for d in ds)))
