#!/usr/bin/env python3
"""Per-CUDA-source-line executed-instruction shares from an .ncu-rep (needs -lineinfo).
   python tools/ncu_lines.py rep kernel_name [min_share]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
minshare = float(sys.argv[3]) if len(sys.argv) > 3 else 0.01
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kern,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
lines, hdr = [], None
for r in rows:
    if "Instructions Executed" in r:
        hdr = r
        continue
    if hdr and len(r) > 8 and r[0].isdigit():
        ie, ist = hdr.index("Instructions Executed"), hdr.index("# Samples")
        try:
            lines.append((int(r[ie]), int(r[ist] or 0), int(r[0]), r[1].strip()[:120]))
        except ValueError:
            pass
tot = sum(l[0] for l in lines) or 1
ts = sum(l[1] for l in lines) or 1
print("kernel %s: %d warp instructions, %d stall samples" % (kern, tot, ts))
for v, st, ln, src in sorted(lines, key=lambda t: -t[0]):
    if v / tot >= minshare:
        print("%6.3f inst %6.3f samp  L%-4d %s" % (v / tot, st / ts, ln, src))
