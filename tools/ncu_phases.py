#!/usr/bin/env python3
"""Instruction / stall-sample shares of a kernel between source markers (development aid).
   python tools/ncu_phases.py rep kernel file 'marker1' 'marker2' ..."""
import csv, subprocess, sys
rep, kern, path = sys.argv[1:4]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kern,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines())); hdr = None; L = {}; S = {}
def num(x):
    try: return int(x)
    except ValueError: return 0
for r in rows:
    if "Instructions Executed" in r: hdr = r; continue
    if hdr and len(r) > 8 and r[0].isdigit():
        L[int(r[0])] = L.get(int(r[0]), 0) + num(r[hdr.index("Instructions Executed")])
        S[int(r[0])] = S.get(int(r[0]), 0) + num(r[hdr.index("# Samples")])
tot = sum(L.values()) or 1; ts = sum(S.values()) or 1
src = open(path).read().split("\n")
marks = [("(start)", 0)]
for m in sys.argv[4:]:
    marks.append((m[:28], next(i + 1 for i, l in enumerate(src) if m in l)))
marks.append(("(end)", len(src) + 1))
print("total warp instr", tot)
for (n, a), (_, b) in zip(marks, marks[1:]):
    print("%-30s L%4d-%4d inst %.3f samples %.3f" % (n, a, b, sum(c for l, c in L.items() if a <= l < b) / tot, sum(c for l, c in S.items() if a <= l < b) / ts))
