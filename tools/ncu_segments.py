#!/usr/bin/env python3
"""ncu_segments.py source.csv : split the SASS of one kernel (ncu --page source --csv) at RET instructions (= the kernel body
and the non-inlined device functions behind it) and print, per segment, executed warp instructions, share, stall samples,
average active threads and the opcode mix — where a kernel's instructions go when the source view has no line mapping."""
import csv, sys
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ia, isrc, iex, ith, ismp = (hdr.index(n) for n in ("Address", "Source", "Instructions Executed", "Thread Instructions Executed", "# Samples"))
segs, cur = [], None
base = int(rows[2][ia], 16)
for r in rows[2:]:
    if r and r[0] == "Kernel Name": break  # (first instance only)
    if len(r) <= ith or r[ia] == "Address": continue
    a = int(r[ia], 16) - base
    op = r[isrc].split()
    op = op[1] if op and op[0].startswith('@') else (op[0] if op else '?')
    ex, th, smp = int(r[iex]), int(r[ith]), int(r[ismp])
    if cur is None: cur = dict(a0=a, ex=0, th=0, smp=0, n=0, ops=Counter(), exops=Counter())
    cur['a1'] = a; cur['ex'] += ex; cur['th'] += th; cur['smp'] += smp; cur['n'] += 1
    cur['ops'][op.split('.')[0]] += 1; cur['exops'][op.split('.')[0]] += ex
    if op.startswith('RET') or op.startswith('EXIT') and False:
        segs.append(cur); cur = None
if cur: segs.append(cur)
tot = sum(s['ex'] for s in segs); tsmp = sum(s['smp'] for s in segs)
print("total warp instructions %d, samples %d" % (tot, tsmp))
for s in segs:
    if s['ex'] == 0: continue
    print("0x%05x-0x%05x %5d sass  %5.1f%% instr %5.1f%% samples  thr/inst %4.1f  static VABS %d | executed: %s" % (
        s['a0'], s['a1'], s['n'], 100 * s['ex'] / tot, 100 * s['smp'] / max(tsmp, 1), s['th'] / max(s['ex'], 1), s['ops']['VABSDIFF4'],
        ' '.join('%s:%.1f%%' % (k, 100 * v / tot) for k, v in s['exops'].most_common(7))))
