#!/usr/bin/env python
"""ncu_lines.py — per CUDA source line: warp-stall samples and instructions executed of the first kernel in an
.ncu-rep captured with --import-source on (kernel compiled with -lineinfo).  Usage: python tools/ncu_lines.py rep [top]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fname = ""
hdr = None
agg = []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = {h: i for i, h in enumerate(r)}; si = r.index("# Samples"); ii = r.index("Instructions Executed"); continue
    if hdr is None or len(r) < 8 or not r[0]: continue
    try: agg.append((int(r[si]), int(r[ii]), fname, r[0], r[1].strip()[:110]))
    except ValueError: pass
tot = sum(a[0] for a in agg)
print("total samples", tot)
for s, n, f, ln, src in sorted(agg, reverse=True)[:top]:
    print("%7d %5.1f%% %11d  %s:%s  %s" % (s, 100.0 * s / max(tot, 1), n, f, ln, src))
