#!/usr/bin/env python
"""Warp-instructions and stall samples of an ncu source page by STAGE of the hot path.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv
       python tools_stage_budget.py src.csv [read_bases]
Line ranges are those of tksm_b200/csrc/kernels.cuh at commit 400533e (the state profiles/r2v_* were taken on)."""
import csv, sys
STAGES = [   # (name, file, first line, last line)
    ("philox", "kernels.cuh", 20, 60),
    ("identity / crossed / slot weights", "kernels.cuh", 61, 239),
    ("stop time: histogram scan", "kernels.cuh", 240, 303),
    ("qscore row lookup (hash, trims)", "kernels.cuh", 304, 336),
    ("compaction queues (WarpQueue)", "kernels.cuh", 337, 384),
    ("inject: set-up", "kernels.cuh", 385, 426),
    ("inject: materialise", "kernels.cuh", 427, 502),
    ("inject: target, epoch class", "kernels.cuh", 503, 569),
    ("inject: trials", "kernels.cuh", 572, 675),
    ("inject: arrivals", "kernels.cuh", 570, 705),
    ("inject: stop-time selection", "kernels.cuh", 706, 780),
    ("inject: hand-off to mid", "kernels.cuh", 781, 795),
    ("team alignment (FIN_FULL)", "kernels.cuh", 796, 947),
    ("finish: set-up, trims", "kernels.cuh", 948, 994),
    ("finish: expand", "kernels.cuh", 995, 1142),
    ("finish: qscore alias draw", "kernels.cuh", 1143, 1203),
    ("finish: qscore slow path", "kernels.cuh", 1204, 1219),
    ("finish: qscore fast path", "kernels.cuh", 1220, 1296),
    ("kernel frame, tickets", "kernels.cuh", 1297, 1400),
    ("align: byte streams", "kernels.cuh", 1401, 1450),
    ("align: forward pass", "kernels.cuh", 1451, 1597),
    ("align: walk", "kernels.cuh", 1598, 1697),
    ("align: kernel frame", "kernels.cuh", 1698, 1770),
    ("align.cuh helpers", "align.cuh", 1, 100000),
]
rows = list(csv.reader(open(sys.argv[1])))
bases = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
fname, hdr, data = "", None, []
for r in rows:
    if r and r[0] == "File Path": fname = r[1].split("/")[-1]
    if r and r[0] == "Line No" and len(r) > 5:
        hdr = r; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed"); ti = hdr.index("Thread Instructions Executed"); continue
    if hdr and len(r) > ii and r[0].isdigit() and r[2] == "-":
        try: data.append((fname, int(r[0]), int(r[si] or 0), int(r[ii] or 0), int(r[ti] or 0)))
        except ValueError: pass
tot_s = sum(d[2] for d in data) or 1; tot_i = sum(d[3] for d in data) or 1
acc = {}
for f, ln, s, i, t in data:
    name = "other (" + f + ")"
    for st, sf, a, b in STAGES:   # trials before arrivals: the first match wins
        if f == sf and a <= ln <= b:
            name = st; break
    x = acc.setdefault(name, [0, 0, 0]); x[0] += s; x[1] += i; x[2] += t
print(f"total warp-instructions {tot_i/1e6:.1f} M, stall samples {tot_s}" + (f", {tot_i/bases:.1f} warp-instructions per read base" if bases else ""))
print(f"{'stage':40s} {'warp-inst':>10s} {'share':>6s} {'thr/inst':>8s} {'samples':>7s}" + ("  per base" if bases else ""))
for name, (s, i, t) in sorted(acc.items(), key=lambda kv: -kv[1][1]):
    if i == 0 and s == 0: continue
    print(f"{name:40s} {i/1e6:9.1f}M {100*i/tot_i:5.1f}% {t/max(i,1):8.1f} {100*s/tot_s:6.1f}%" + (f"  {i/bases:7.2f}" if bases else ""))
