#!/usr/bin/env python3
"""Summarise an `ncu --page source --csv --print-source cuda,sass` dump: stall mix and the hottest source lines."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
cur = None; H = None; data = []
for r in rows:
    if r and r[0] == 'File Path': cur = r[1]; continue
    if r and r[0] == 'Line No': H = r; continue
    if r and r[0].isdigit() and H: data.append((cur, int(r[0]), r))
iS = H.index('# Samples'); iI = H.index('Instructions Executed')
num = lambda x: int(x) if x.isdigit() else 0
totI = sum(num(r[iI]) for _, _, r in data); totS = sum(num(r[iS]) for _, _, r in data)
stall = [i for i, h in enumerate(H) if h.startswith('stall_') and 'Not Issued' not in h]
agg = {}
for _, _, r in data:
    for i in stall: agg[H[i]] = agg.get(H[i], 0) + num(r[i])
print("warp-inst", totI, "samples", totS, [(k[6:], round(100 * v / totS, 1)) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]])
for f, l, r in sorted(data, key=lambda x: -num(x[2][iS]))[:top]:
    st = sorted(((num(r[i]), H[i][6:]) for i in stall if num(r[i]) > 0), reverse=True)[:2]
    print(f"{100*num(r[iS])/totS:5.1f}% smp  inst {100*num(r[iI])/totI:5.1f}%  {f.split('/')[-1][:22]}:{l:<4} {r[1][:88]}  {st}")
