"""Experiment: device-resident stereo step on one stream vs left / right extractors on their own streams vs 2 x half batches."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import bench
from orb_slam3_detailed_comments_kor_b200 import ORBextractor
from orb_slam3_detailed_comments_kor_b200 import matcher as M

bench.set_workload("euroc")
W, H, PARAMS, LAP, MBF, FX = bench.W, bench.H, bench.PARAMS, bench.LAP, bench.MBF, bench.FX
P = 256
hL, hR = bench.gen_pairs(P, 0, pin=True, distinct=64)
dL, dR = hL.cuda(), hR.cuda()
fb = W * H

def timeit(fn, streams, steps=20, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(streams[0])
    for s in streams[1:]: s.wait_event(e0)
    for _ in range(steps): fn()
    for s in streams[1:]:
        ev = torch.cuda.Event(); ev.record(s); streams[0].wait_event(ev)
    e1.record(streams[0])
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps

def variant(nsplit, two_streams):
    SUB = P // nsplit
    exs = [(ORBextractor(*PARAMS, width=W, height=H, max_batch=SUB), ORBextractor(*PARAMS, width=W, height=H, max_batch=SUB)) for _ in range(nsplit)]
    sA = [torch.cuda.Stream() for _ in range(nsplit)]
    sB = [torch.cuda.Stream() for _ in range(nsplit)] if two_streams else sA
    def fn():
        for i in range(nsplit):
            exL, exR = exs[i]
            exL.extract_batch_device(dL.data_ptr() + i * SUB * fb, SUB, LAP, sA[i].cuda_stream)
            exR.extract_batch_device(dR.data_ptr() + i * SUB * fb, SUB, LAP, sB[i].cuda_stream)
            if two_streams:
                ev = torch.cuda.Event(); ev.record(sB[i]); sA[i].wait_event(ev)
            M.compute_stereo_matches(exL, exR, SUB, MBF, 0.0, FX, sA[i].cuda_stream)
            if two_streams:
                ev2 = torch.cuda.Event(); ev2.record(sA[i]); sB[i].wait_event(ev2)   # the next right extraction must not overwrite what the matcher reads
    streams = list(dict.fromkeys(sA + sB))
    ms = timeit(fn, streams)
    for a, b in exs: a.close(); b.close()
    return ms

def variant_api(nsplit):
    SUB = P // nsplit
    exs = [(ORBextractor(*PARAMS, width=W, height=H, max_batch=SUB), ORBextractor(*PARAMS, width=W, height=H, max_batch=SUB)) for _ in range(nsplit)]
    main = torch.cuda.Stream()
    def fn():
        for i in range(nsplit):
            M.stereo_frames_batch_device(exs[i][0], exs[i][1], dL.data_ptr() + i * SUB * fb, dR.data_ptr() + i * SUB * fb, SUB, MBF, 0.0, FX, LAP, None)
    def timeit2(steps=50, warm=3):
        import time
        for _ in range(warm): fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps): fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3 / steps
    ms = timeit2()
    for a, b in exs: a.close(); b.close()
    return ms

for nsplit in (1, 2):
    ms = variant_api(nsplit)
    print("orbfe_stereo_frames_batch_device, %d sub-batches own streams, wall clock: %.3f ms/step  %.1f k frames/s" % (nsplit, ms, 2 * P / ms), flush=True)

exL = ORBextractor(*PARAMS, width=W, height=H, max_batch=P); exR = ORBextractor(*PARAMS, width=W, height=H, max_batch=P)
main = torch.cuda.Stream()
for rep in range(3):
    for name, sarg in (("user stream", main.cuda_stream), ("no user stream", None)):
        import time
        fn = lambda: M.stereo_frames_batch_device(exL, exR, dL.data_ptr(), dR.data_ptr(), P, MBF, 0.0, FX, LAP, sarg)
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record(main)
        for _ in range(20): fn()
        e1.record(main); torch.cuda.synchronize(); wall = (time.perf_counter() - t0) * 1e3 / 20
        print("%s: events %.3f ms  wall %.3f ms" % (name, e0.elapsed_time(e1) / 20, wall), flush=True)

print("with the bench's clock sampler running")
smp = bench.ClockSampler(0); smp.start()
for rep in range(2):
    fn = lambda: M.stereo_frames_batch_device(exL, exR, dL.data_ptr(), dR.data_ptr(), P, MBF, 0.0, FX, LAP, main.cuda_stream)
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    for _ in range(20): fn()
    e1.record(main); torch.cuda.synchronize()
    print("user stream + sampler: events %.3f ms" % (e0.elapsed_time(e1) / 20), flush=True)
print(smp.stop())
hL2, hR2 = bench.gen_pairs(P, 0, pin=True)
dL2, dR2 = hL2.cuda(), hR2.cuda()
for rep in range(2):
    fn = lambda: M.stereo_frames_batch_device(exL, exR, dL2.data_ptr(), dR2.data_ptr(), P, MBF, 0.0, FX, LAP, main.cuda_stream)
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    for _ in range(20): fn()
    e1.record(main); torch.cuda.synchronize()
    print("256 distinct pairs: events %.3f ms" % (e0.elapsed_time(e1) / 20), flush=True)
