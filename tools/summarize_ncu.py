#!/usr/bin/env python3
"""Summarise ncu outputs brought back in gpurun_out/ into small text files under profiles/.

  python tools/summarize_ncu.py launches gpurun_out/launches_r1.csv profiles/r1_launches.txt
  python tools/summarize_ncu.py full gpurun_out/prof_fast_r1.ncu-rep profiles/r1_fast_full.txt
"""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "lts__t_bytes.sum", "sm__cycles_elapsed.max"]


def launches(src, dst):
    rows = [r for r in csv.reader(open(src)) if len(r) > 5]
    hdr = rows[0]
    ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        a = agg.setdefault(r[ki].split("(")[0], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(dst, "w") as f:
        f.write("# per-kernel device time from `ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache,\n"
                "# serialised launches: compare SHARES, not absolutes).  source: %s\n" % src)
        f.write("%-28s %6s %12s %7s\n" % ("kernel", "n", "total_us", "share"))
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("%-28s %6d %12.1f %7.3f\n" % (k, n, t / 1e3, t / tot))
    print(open(dst).read())


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write("# ncu --set full --clock-control none, raw page, selected metrics.  source: %s\n" % src)
        ki = hdr.index("Kernel Name")
        for r in rows[2:]:
            f.write("kernel %s  grid %s block %s\n" % (r[ki], r[hdr.index("Grid Size")], r[hdr.index("Block Size")]))
            for k in KEYS:
                if k in hdr:
                    f.write("  %-68s %s %s\n" % (k, r[hdr.index(k)], units[hdr.index(k)]))
            stalls = [(float(r[i]), h) for i, h in enumerate(hdr) if h.startswith("smsp__average_warp") and h.endswith("_per_issue_active.ratio") or
                      (h.startswith("smsp__average_warps_issue_stalled") and "ratio" in h) for _ in [0] if r[i] not in ("", "n/a")]
            for v, h in sorted(stalls, reverse=True)[:6]:
                f.write("  stall %-62s %.3f\n" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("smsp__average_warp_latency_issue_stalled_", ""), v))
    print(open(dst).read())




def table(src, dst, frames):
    """All kernels of one captured step: duration, DRAM traffic, issue/pipe utilisation, occupancy, top stalls;
    plus profiles/<round>_traffic.json = DRAM bytes per frame per stage (what bench.py reports as roofline.traffic)."""
    import json
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[0]
    c = hdr.index
    cols = [("us", "gpu__time_duration.sum"), ("dram_rd_MB", "dram__bytes_read.sum"), ("dram_wr_MB", "dram__bytes_write.sum"),
            ("issue%", "smsp__issue_active.avg.pct_of_peak_sustained_active"), ("alu%", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
            ("xu%", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"), ("lsu%", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
            ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), ("warps%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
            ("fma%", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
            ("regs", "launch__registers_per_thread"), ("winst", "smsp__inst_executed.sum")]
    stage_of = {"k_pyramid_copy0": "pyramid", "k_pyramid_level": "pyramid", "k_pyramid_borders": "pyramid", "k_pyramid_fringe": "pyramid", "k_fast_stream": "fast",
                "k_fast_nms": "fast", "k_fast_cells_list": "fast", "k_blur7": "blur", "k_blur7_tma": "blur", "k_quadtree": "quadtree",
                "void k_quadtree<0>": "quadtree", "void k_quadtree<1>": "quadtree", "k_assemble": "assemble",
                "k_orient_describe": "orient_describe", "k_stereo_bucket": "stereo", "k_stereo_match": "stereo", "k_stereo_median": "stereo"}
    seen, traffic, times, winst = {}, collections.OrderedDict(), collections.OrderedDict(), collections.OrderedDict()
    per_kernel = collections.OrderedDict()
    with open(dst, "w") as f:
        f.write("# ncu --set full --clock-control none, one whole step of bench.py (%d frames per launch); source: %s\n" % (frames, src))
        f.write("%-20s %-15s " % ("kernel", "grid") + " ".join("%10s" % n for n, _ in cols) + "  top stalls (per issue)\n")
        for r in rows[2:]:
            name = r[c("Kernel Name")].split("(")[0]
            key = (name, r[c("Grid Size")])
            if key in seen:
                continue
            seen[key] = 1
            st = sorted(((float(r[i]), h) for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled") and r[i] not in ("", "n/a")), reverse=True)[:3]
            stalls = ", ".join("%s %.1f" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v) for v, h in st)
            f.write("%-20s %-15s " % (name, r[c("Grid Size")].replace(" ", "")) + " ".join("%10.10s" % r[c(m)] for _, m in cols) + "  " + stalls + "\n")
            stg = stage_of.get(name)
            if stg:
                mb = float(r[c("dram__bytes_read.sum")]) + float(r[c("dram__bytes_write.sum")])
                traffic[stg] = traffic.get(stg, 0.0) + mb * 1e6 / frames
                times[stg] = times.get(stg, 0.0) + float(r[c("gpu__time_duration.sum")])
                wi = float(r[c("smsp__inst_executed.sum")])
                winst[stg] = winst.get(stg, 0.0) + wi / frames
                k = per_kernel.setdefault(name, {"warp_instructions_per_frame": 0.0, "us_under_ncu": 0.0, "launches": 0})
                k["warp_instructions_per_frame"] += wi / frames
                k["us_under_ncu"] += float(r[c("gpu__time_duration.sum")])
                k["launches"] += 1
                for label, metric in (("issue_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active"), ("alu_pct", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                                      ("fma_pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"), ("lsu_pct", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
                                      ("xu_pct", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active")):
                    k[label] = max(k.get(label, 0.0), float(r[c(metric)]))   # of the largest launch of that kernel
    json.dump({"frames_per_launch": frames, "dram_bytes_per_frame": traffic, "stage_us_under_ncu": times,
               "warp_instructions_per_frame": winst, "per_kernel": per_kernel, "source": src},
              open(dst.replace("_kernels_full.txt", "_traffic.json"), "w"), indent=1)
    print(open(dst).read())


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "table":
    table(sys.argv[2], sys.argv[3], int(sys.argv[4]))
if __name__ == "__main__" and sys.argv[1] in ("launches", "full"):
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
