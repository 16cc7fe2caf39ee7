#!/usr/bin/env python
"""Top stalled SASS instructions of an `ncu --page source --csv` export."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
h = rows[hi[0]]
end = hi[1] - 1 if len(hi) > 1 else len(rows)
body = [r for r in rows[hi[0] + 1:end] if len(r) == len(h)]
iS, iE, iN = h.index('Source'), h.index('Instructions Executed'), h.index('# Samples')
cols = [i for i, c in enumerate(h) if c.startswith('stall_') and 'Not' not in c]
tot = sum(int(r[iN] or 0) for r in body)
agg = collections.Counter()
for r in body:
    for i in cols:
        agg[h[i]] += int(r[i] or 0)
print('samples', tot, ' '.join('%s=%.1f%%' % (k[6:], 100.0 * v / max(tot, 1)) for k, v in agg.most_common(10)))
print('dynamic instructions', sum(int(r[iE] or 0) for r in body))
for r in sorted(body, key=lambda r: -int(r[iN] or 0))[:n]:
    st = sorted(((int(r[i] or 0), h[i][6:]) for i in cols), reverse=True)[:2]
    print(r[0][-5:], '%5s %5.1f%%' % (r[iN], 100.0 * int(r[iN] or 0) / max(tot, 1)), r[iE].rjust(9),
          r[iS][:60].ljust(60), st)
