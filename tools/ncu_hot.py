#!/usr/bin/env python
"""Top SASS instructions by warp-stall samples of an ncu report (source page), with neighbours.
Usage: tools/ncu_hot.py report.ncu-rep [top_n] [context]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25; ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
c = {n: i for i, n in enumerate(hdr)}
S = c["# Samples"]; X = c["Instructions Executed"]
tot = sum(int(r[S]) for r in data); totx = sum(int(r[X]) for r in data)
print("total samples %d, warp instructions %d, sass lines %d" % (tot, totx, len(data)))
stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
agg = {n: sum(int(r[c[n]]) for r in data) for n in stall_cols}
print("stall totals:", ", ".join("%s %d" % (n[6:], v) for n, v in sorted(agg.items(), key=lambda t: -t[1]) if v))
order = sorted(range(len(data)), key=lambda i: -int(data[i][S]))[:top]
for i in sorted(order):
    for j in range(max(0, i - ctx), min(len(data), i + ctx + 1)):
        r = data[j]
        st = sorted(((int(r[c[n]]), n[6:]) for n in stall_cols), reverse=True)[:2]
        print("%s%5d %6s smp %8s ex  %-70s %s" % ("*" if j == i else " ", j, r[S], r[X], r[c["Source"]].strip()[:70], " ".join("%s=%d" % (n, v) for v, n in st if v)))
    if ctx: print()
