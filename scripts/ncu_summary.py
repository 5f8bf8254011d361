#!/usr/bin/env python
"""Per-kernel summary of an `ncu --page raw --csv` export (profiles/*_raw.csv): time, issue-active, lanes per instruction,
DRAM bytes, occupancy, hit rates, top stall reasons."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]
want = ['gpu__time_duration.sum', 'sm__issue_active.avg.pct_of_peak_sustained_elapsed', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'launch__registers_per_thread']
names = ['time', 'issue%', 'lanes', 'rd', 'wr', 'occ%', 'L1hit', 'L2hit', 'regs']
ix = [h.index(w) for w in want]
stall = [(i, x.split('stalled_')[1].split('_per_issue')[0]) for i, x in enumerate(h) if x.startswith('smsp__average_warps_issue_stalled_') and x.endswith('_per_issue_active.ratio')]
print('units:', [rows[1][i] for i in ix])
print('kernel'.ljust(22), *[n.rjust(9) for n in names], ' top stalls (warps per issue)')
for r in rows[2:]:
    st = sorted(((float(r[i].replace(',', '') or 0), n) for i, n in stall), reverse=True)[:3]
    print(r[h.index('Kernel Name')].split('(')[0][-22:].ljust(22), *[r[i][:9].rjust(9) for i in ix], ' ', ' '.join('%s=%.1f' % (n, v) for v, n in st))
