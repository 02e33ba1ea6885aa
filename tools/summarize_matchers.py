#!/usr/bin/env python3
"""profiles/r2_matchers_full.txt from the raw-page CSVs of the matcher kernels' ncu --set full captures (exported on the GPU box with
`ncu -i x.ncu-rep --page raw --csv`; a later CSV replaces the kernels it holds):  python tools/summarize_matchers.py out.txt a.csv [b.csv ...]"""
import collections
import csv
import sys

COLS = [("us", "gpu__time_duration.sum"), ("issue%", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        ("warps%", "sm__warps_active.avg.pct_of_peak_sustained_active"), ("regs", "launch__registers_per_thread"),
        ("smem_dyn_KB", "launch__shared_mem_per_block_dynamic"), ("smem_static_KB", "launch__shared_mem_per_block_static"),
        ("cta/SM smem", "launch__occupancy_limit_shared_mem"), ("cta/SM regs", "launch__occupancy_limit_registers"),
        ("winst", "smsp__inst_executed.sum"), ("dram_rd_MB", "dram__bytes_read.sum"), ("dram_wr_MB", "dram__bytes_write.sum")]
last = collections.OrderedDict()
for path in sys.argv[2:]:
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
        vals = {}
        for label, key in COLS:
            v, u = float(r[idx[key]].replace(",", "")), units[idx[key]]
            if label == "us":
                v *= {"ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}[u]
            if label.startswith("dram"):
                v *= {"Mbyte": 1.0, "Kbyte": 1e-3, "Gbyte": 1e3, "byte": 1e-6}[u]
            vals[label] = v
        last["%s grid %s x %s" % (name.split("<")[0] if "search_by_bow" in name else name, r[idx["launch__grid_size"]], r[idx["launch__block_size"]])] = vals
with open(sys.argv[1], "w") as f:
    f.write("# ncu --set full --clock-control none on bench.py --steps 2 --warmup 3 --no-cpu-baseline, kernels of the matcher legs (the last launch of each form)\n")
    f.write("%-52s" % "kernel, grid x block" + "".join("%15s" % c for c, _ in COLS) + "\n")
    for k, v in last.items():
        f.write("%-52s" % k + "".join("%15.3f" % v[c] if v[c] < 1e6 else "%15.4g" % v[c] for c, _ in COLS) + "\n")
print(open(sys.argv[1]).read())
