#!/usr/bin/env python
"""Executed-instruction profile of one `ncu --page source --csv` export (SASS rows): splits the
kernel at every change of the per-row execution count into regions and prints the regions that
carry the most executed warp instructions, with their opcode mix and stall samples.
usage: ncu_hot_regions.py gpurun_out/p2_mg_source.csv [top]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
iS, iN, iSm = h.index('Source'), h.index('Instructions Executed'), h.index('# Samples')
body = [(r[iS].strip(), int(r[iN] or 0), int(r[iSm] or 0)) for r in rows[2:] if len(r) > iN]
total = sum(n for _, n, _ in body)
regions, cur = [], None
for k, (s, n, sm) in enumerate(body):
    if cur is None or n != cur['n']:
        cur = dict(start=k, n=n, count=0, samples=0, ops=collections.Counter(), first=s)
        regions.append(cur)
    cur['count'] += 1
    cur['samples'] += sm
    p = s.split()
    op = p[1] if p and p[0].startswith('@') and len(p) > 1 else (p[0] if p else '?')
    cur['ops'][op.split('.')[0]] += 1
print('instructions (static) %d, executed warp instructions %d, samples %d' %
      (len(body), total, sum(sm for _, _, sm in body)))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 14
for r in sorted(regions, key=lambda r: -r['n'] * r['count'])[:top]:
    print('rows %5d..%5d  x%-9d = %5.1f%% of executed, %5.1f%% of samples | %s | %s' % (
        r['start'], r['start'] + r['count'] - 1, r['n'], 100.0 * r['n'] * r['count'] / total,
        100.0 * r['samples'] / max(1, sum(sm for _, _, sm in body)),
        ', '.join('%s %d' % kv for kv in r['ops'].most_common(6)), r['first'][:40]))
