#!/usr/bin/env python3
"""Print the interesting fields of bench.py's JSON line (development aid): python bench.py ... | python tools/show_bench.py"""
import json
import signal
import sys
signal.signal(signal.SIGPIPE, signal.SIG_DFL)
for l in sys.stdin:
    if not l.startswith("{"):
        print(l.rstrip()[:300])
        continue
    d = json.loads(l)
    print("value %.0f %s  ms/step %.3f  e2e %.0f  n_gpus %d" % (d["value"], d["unit"], d["ms_per_step"], d["e2e"]["value"], d["n_gpus"]))
    r = d.get("roofline") or {}
    print("stages ms:", {k: round(v, 3) for k, v in (r.get("stage_ms_per_step") or {}).items()})
    print("roofline top kernel %s frac %.3f achieved %.0f GB/s | whole-step frac %.3f" % (r.get("kernel"), r.get("frac", 0), r.get("achieved", 0), (r.get("whole_step") or {}).get("frac", 0)))
    h = d.get("hamming") or {}
    print("hamming %.3e pairs/s  ms %.3f  frac %.3f" % (h.get("value", 0), h.get("ms_per_step", 0), (h.get("roofline") or {}).get("frac", 0)))
    print("clocks", d.get("clocks"), "cpu", d.get("cpu_baseline"))
    print("projection", d.get("projection"), "| rectify %.0f frames/s" % (d.get("rectify") or {}).get("value", 0))
    print("bow", d.get("bow"))
    print("initialization", d.get("initialization"), "| serial step ms", (d.get("config") or {}).get("serial_ms_per_step"))
    print("consumers", d.get("consumers"))
    print("cpu consumers", {k: (d.get("cpu_baseline") or {}).get(k) for k in ("search_by_bow", "search_for_triangulation", "distinctive_descriptors", "initialization")})
    for k, v in ((d.get("cpu_baseline") or {}).get("one_call_latency") or {}).items():
        print("one call %-34s gpu host call %.3f ms | cpu one thread %s ms" % (k, v["gpu_host_call_ms"], "%.3f" % v["cpu_one_thread_ms"] if v["cpu_one_thread_ms"] else "-"))
