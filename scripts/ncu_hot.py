#!/usr/bin/env python
"""Top SASS instructions by stall samples for one kernel of an .ncu-rep.
usage: ncu_hot.py rep kernel_regex [topN] [instance]"""
import csv, subprocess, sys, collections, re
rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
inst = int(sys.argv[4]) if len(sys.argv) > 4 else 0
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + rx],
                     capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(out.splitlines()):
    if row and row[0] == 'Kernel Name':
        cur = {'name': row[1], 'rows': [], 'hdr': None}
        blocks.append(cur)
    elif cur is not None and cur['hdr'] is None:
        cur['hdr'] = row
    elif cur is not None:
        cur['rows'].append(row)
b = blocks[inst]
h = b['hdr']
iS, iSamp, iEx = h.index('Source'), h.index('# Samples'), h.index('Instructions Executed')
rows = b['rows']
tot = sum(int(r[iSamp] or 0) for r in rows)
print(b['name'][:100], ' total samples', tot, ' instrs', len(rows))
byop = collections.Counter()
for r in rows:
    op = r[iS].split()[0] if r[iS].split() else '?'
    if op.startswith('@'):
        op = r[iS].split()[1]
    byop[op.split('.')[0]] += int(r[iSamp] or 0)
print('by opcode:', ', '.join(f'{k}:{100*v/tot:.1f}%' for k, v in byop.most_common(14)))
idx = sorted(range(len(rows)), key=lambda i: -int(rows[i][iSamp] or 0))[:top]
for i in sorted(idx):
    r = rows[i]
    print(f'{i:5d} {100*int(r[iSamp] or 0)/tot:5.1f}%  ex={r[iEx]:>9s}  {r[iS].strip()[:110]}')
