#!/usr/bin/env python
"""Static SASS loop census: for every backward branch of a kernel print the
loop's instruction count and opcode mix (cuobjdump -sass output on stdin or a
file). Used to track instructions per step of the fused kernels offline."""
import collections
import re
import sys

txt = open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()
ins = []
for m in re.finditer(r'/\*([0-9a-f]{4,})\*/\s+(.*?);', txt):
    ins.append((int(m.group(1), 16), m.group(2).strip()))
addr_index = {a: i for i, (a, _) in enumerate(ins)}
print('instructions:', len(ins))
for i, (a, s) in enumerate(ins):
    m = re.search(r'\bBRA(?:\.\w+)*\s+(?:\w+,\s*)?`?\(?\.?L?_?x?_?([0-9a-fx]+)\)?', s)
    if 'BRA' in s:
        t = re.search(r'0x([0-9a-f]+)', s)
        if t:
            ta = int(t.group(1), 16)
            if ta <= a and ta in addr_index:
                j = addr_index[ta]
                body = ins[j:i + 1]
                ops = collections.Counter()
                for _, b in body:
                    p = b.split()
                    op = p[1] if p[0].startswith('@') else p[0]
                    ops[op.split('.')[0]] += 1
                if len(body) > 20:
                    print('loop 0x%x..0x%x: %d instr' % (ta, a, len(body)))
                    print('   ', ', '.join('%s %d' % kv for kv in ops.most_common(24)))
