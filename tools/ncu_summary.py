#!/usr/bin/env python
"""Summarise an ncu export: tools/ncu_summary.py raw.csv source.csv  (csv pages 'raw' and 'source')."""
import csv, sys
raw, src = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'sm__cycles_elapsed.max',
        'smsp__inst_executed.sum', 'l1tex__data_pipe_lsu_wavefronts.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_lg.sum']
for k in keys:
    if k in hdr:
        i = hdr.index(k)
        print(k, units[i], [r[i][:44] for r in rows[2:]])
rows = list(csv.reader(open(src)))
sec = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
seen = set()
for si, s in enumerate(sec):
    name = rows[s][1]
    if name in seen: continue
    seen.add(name)
    h = rows[s + 1]; e = sec[si + 1] if si + 1 < len(sec) else len(rows); body = rows[s + 2:e]
    cols = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
    tot = {c: sum(int(r[h.index(c)] or 0) for r in body) for c in cols}
    T = sum(tot.values())
    print(name[:60], 'samples', T)
    print('   ', ', '.join('%s %.1f%%' % (c[6:], 100 * v / T) for c, v in sorted(tot.items(), key=lambda x: -x[1])[:9]))
    isrc = h.index('Source'); ins = h.index('# Samples')
    top = sorted(enumerate(body), key=lambda kr: -int(kr[1][ins] or 0))[:int(sys.argv[3]) if len(sys.argv) > 3 else 12]
    for k, r in top:
        st = max(cols, key=lambda c: int(r[h.index(c)] or 0))
        print('    %5d %6s %-10s %s' % (k, r[ins], st[6:], r[isrc][:90]))
