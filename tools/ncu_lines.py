#!/usr/bin/env python
"""Aggregate an ncu report's per-SASS samples / executed counts by CUDA source line, using the line
table of the in-tree cubin (nvdisasm -g) aligned by instruction index.
Usage: tools/ncu_lines.py report.ncu-rep <cubin> <mangled-kernel-substring> [top_n]"""
import csv, io, re, subprocess, sys, collections
rep, cubin, ksub = sys.argv[1:4]; top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
dis = subprocess.run(["nvdisasm", "-gi", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
lines, cur, infun = [], None, False
for l in dis:
    if l.startswith(".text."):
        infun = ksub in l
        continue
    if not infun:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        inl = re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3))
        cur = (m.group(1).split("/")[-1], int(m.group(2)), tuple((a.split("/")[-1], int(b)) for a, b in inl))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        lines.append(cur)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
c = {n: i for i, n in enumerate(hdr)}
print("sass rows in report %d, in cubin %d" % (len(data), len(lines)))
n = min(len(data), len(lines))
agg = collections.defaultdict(lambda: [0, 0])
for i in range(n):
    loc = lines[i]
    # attribute to the outermost frame inside the kernel file (last inlined-at), keep the leaf for detail
    key = loc[2][-1] if loc and loc[2] else (loc[:2] if loc else ("?", 0))
    agg[key][0] += int(data[i][c["# Samples"]]); agg[key][1] += int(data[i][c["Instructions Executed"]])
tots = sum(v[0] for v in agg.values()); totx = sum(v[1] for v in agg.values())
print("by kernel-level source line (outermost inline frame): samples %d, warp instr %d" % (tots, totx))
src = {}
for (f, ln), v in sorted(agg.items(), key=lambda t: -t[1][1])[:top]:
    if f not in src:
        try: src[f] = open("bts_b200/csrc/" + f).read().splitlines()
        except OSError: src[f] = []
    text = src[f][ln - 1].strip()[:90] if 0 < ln <= len(src[f]) else ""
    print("%-22s %5d  %6.2f%% ex  %6.2f%% smp   %s" % (f, ln, 100.0 * v[1] / totx, 100.0 * v[0] / tots, text))
