#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export: executed warp instructions per SASS opcode and the hottest
instructions.  usage: sass_hist.py src.csv [top_n]"""
import csv, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
data = [r for r in rows[1:] if r[:2] != hdr[:2]]
iA, iS, iT, iSm = (hdr.index(k) for k in ('Instructions Executed', 'Source', 'Thread Instructions Executed', '# Samples'))
tot = sum(int(r[iA]) for r in data)
tth = sum(int(r[iT]) for r in data)
print('warp instructions', tot, 'thread instructions', tth, 'avg active threads %.2f' % (tth / max(tot, 1)), 'sass lines', len(data))
h, hs = collections.Counter(), collections.Counter()
for r in data:
    s = r[iS].strip()
    if s.startswith('@'):
        s = s.split(None, 1)[1]
    op = s.split()[0].split('.')[0]
    h[op] += int(r[iA]); hs[op] += int(r[iSm])
ts = sum(hs.values())
for op, c in h.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 25):
    print(f'{op:10s} {c:12d} {100*c/tot:5.1f}%  stall samples {100*hs[op]/max(ts,1):5.1f}%')
