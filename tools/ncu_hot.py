#!/usr/bin/env python
"""Top stalled SASS instructions of an ncu report: ncu -i X.ncu-rep --page source --csv | python tools/ncu_hot.py [N]"""
import csv, sys
n = int(sys.argv[1]) if len(sys.argv) > 1 else 25
rows = list(csv.reader(sys.stdin))
h = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[h]
body = rows[h + 1:]
si = hdr.index('# Samples'); src = hdr.index('Source'); ie = hdr.index('Instructions Executed')
stall_cols = [i for i, c in enumerate(hdr) if c.startswith('stall_') and 'Not Issued' not in c]
tot = sum(int(r[si] or 0) for r in body)
print('total samples', tot)
order = sorted(range(len(body)), key=lambda i: -int(body[i][si] or 0))[:n]
for i in sorted(order):
    r = body[i]
    st = sorted(((int(r[c] or 0), hdr[c][6:]) for c in stall_cols), reverse=True)[:3]
    print(f"{i:5d} {int(r[si]):7d} {100.0*int(r[si])/tot:5.1f}% ex={r[ie]:>9} {r[src].strip()[:90]:90s} {[f'{b}:{a}' for a,b in st if a]}")
