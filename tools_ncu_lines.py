#!/usr/bin/env python
"""Ranks CUDA source lines of an ncu report by stall samples / executed instructions.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv; python tools_ncu_lines.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
data = []
fname = ''
for r in rows:
    if r and r[0] == 'File Path': fname = r[1].split('/')[-1]
    if r and r[0] == 'Line No' and len(r) > 5:
        hdr = r; si = hdr.index('# Samples'); ii = hdr.index('Instructions Executed'); ti = hdr.index('Thread Instructions Executed'); continue
    if hdr and len(r) > ii and r[0].isdigit() and r[2] == '-':
        try: data.append((int(r[si] or 0), int(r[ii] or 0), int(r[ti] or 0), fname, r[0], r[1].strip()[:100]))
        except ValueError: pass
ts = sum(d[0] for d in data) or 1; tin = sum(d[1] for d in data) or 1
print('total samples', ts, 'warp-inst', tin, 'thread-inst', sum(d[2] for d in data))
for d in sorted(data, reverse=True)[:N]:
    print(f"{d[0]:7d} {100*d[0]/ts:5.1f}% | inst {100*d[1]/tin:5.1f}% thr/inst {d[2]/max(d[1],1):4.1f} | {d[3]}:{d[4]}: {d[5]}")
