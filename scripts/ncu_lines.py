#!/usr/bin/env python
"""Executed warp instructions and stall samples per CUDA source line for one kernel of an .ncu-rep
(needs -lineinfo and --import-source on).  usage: ncu_lines.py rep kernel_regex [topN] [instance] [file suffix, default: the kernel's own file]"""
import csv, subprocess, sys, collections
rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
inst = int(sys.argv[4]) if len(sys.argv) > 4 else 0
want_file = sys.argv[5] if len(sys.argv) > 5 else None
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass',
                      '--kernel-name', 'regex:' + rx], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# kernels are separated by "Kernel Name" rows; files by "File Name" rows; the metric header starts with "Line No"
# the page is a sequence of (File Path, Function Name, table) blocks; a launch contributes one block per source
# file that has inlined code in it, the kernel's own file first.  Only that first block is used: its lines carry
# the inlined instructions too.
kern, k, fpath, first, launch = [], None, '?', None, -1
for r in rows:
    if r and r[0] == 'File Path':
        fpath = r[1]
        first = first or fpath
        if fpath == first:
            launch += 1
        k = None
    elif r and r[0] == 'Function Name':
        if (fpath.endswith(want_file) if want_file else fpath == first):
            k = [['File Name', fpath]]
            kern.append((r[1], k))
    elif k is not None:
        k.append(r)
name, body = kern[inst]
hdr, fname, cur = None, '?', None
per = collections.OrderedDict()
for r in body:
    if not r:
        continue
    if r[0] == 'File Name':
        fname = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
        iEx = hdr.index('Instructions Executed')
        iS = hdr.index('# Samples') if '# Samples' in hdr else hdr.index('Warp Stall Sampling (All Samples)')
    elif r[0] != '':
        cur = (fname, int(r[0]))
        per.setdefault(cur, [r[1].strip(), 0, 0])
    elif hdr and cur and len(r) > iEx and r[iEx] not in ('-', ''):
        per[cur][1] += int(r[iEx])
        per[cur][2] += int(r[iS] or 0)
totE = sum(v[1] for v in per.values()) or 1
totS = sum(v[2] for v in per.values()) or 1
print(name[:110], ' warp-instr', totE, ' samples', totS)
keys = sorted(per, key=lambda q: -per[q][1])[:top]
for q in sorted(keys):
    s, e, sm = per[q]
    print(f'{q[0]:>12s}:{q[1]:<5d} ex {100*e/totE:5.1f}%  st {100*sm/totS:5.1f}%  {s[:100]}')
