#!/usr/bin/env python3
"""sass_funcs.py stage.sass KERNEL_SUBSTR : device functions inside one kernel's section of an `nvdisasm -g -c` listing,
with instruction counts and opcode mix (which non-inlined function the hot 4 KB chunks of an ncu source page belong to)."""
import re, sys
from collections import Counter
path, want = sys.argv[1], sys.argv[2]
lines = open(path).read().split('\n')
ins = re.compile(r'/\*([0-9a-f]{4,6})\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)')
sec = None; cur = None; funcs = []
for l in lines:
    if l.startswith('.text.'):
        sec = l[6:-1]; cur = None
        if want in sec: cur = [sec, None, None, Counter()]; funcs.append(cur)
        continue
    if sec is None or want not in sec: continue
    m = re.match(r'^(\$?[A-Za-z_$][\w$]*):$', l)
    if m and not m.group(1).startswith('.L'):
        name = m.group(1).split('$')[-1]
        cur = [name, None, None, Counter()]; funcs.append(cur); continue
    m = ins.search(l)
    if m and cur is not None:
        a = int(m.group(1), 16)
        if cur[1] is None: cur[1] = a
        cur[2] = a
        cur[3][m.group(2).split('.')[0]] += 1
for name, a0, a1, ops in funcs:
    if a0 is None: continue
    n = sum(ops.values())
    print("%-60s 0x%05x-0x%05x %5d instr  %s" % (name[:60], a0, a1, n, ' '.join('%s:%d' % kv for kv in ops.most_common(8))))
