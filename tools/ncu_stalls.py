#!/usr/bin/env python
"""ncu_stalls.py — summarise an .ncu-rep (run where ncu is installed; no GPU needed): stall mix and the
hottest SASS instructions of the first kernel in the report.  Usage: python tools/ncu_stalls.py rep [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(out)))
print(r[0][1][:100] if len(r[0]) > 1 else r[0])
hdr, rows = r[1], r[2:]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {s: 0 for s in stalls}
N = 0
def num(x):
    try: return int(x)
    except Exception: return 0
for row in rows:
    n = num(row[ix["# Samples"]]); N += n
    for s in stalls: tot[s] += num(row[ix[s]])
print("total samples", N, " instructions executed", sum(num(row[ix["Instructions Executed"]]) for row in rows))
for s, v in sorted(tot.items(), key=lambda x: -x[1])[:10]:
    print("  %-26s %8d %5.1f%%" % (s, v, 100.0 * v / max(N, 1)))
for i, row in sorted(enumerate(rows), key=lambda x: -num(x[1][ix["# Samples"]]))[:top]:
    st = sorted([(num(row[ix[s]]), s[6:]) for s in stalls], reverse=True)[:2]
    print("%6d %9s  #%-5d %-62s %s" % (num(row[ix["# Samples"]]), row[ix["Instructions Executed"]], i, row[ix["Source"]][:62], st))
