"""Why does bench.py see 3.48 ms where tools/exp/streams.py sees 3.34 ms?  Replays bench.py's sequence step by step."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import bench
from orb_slam3_detailed_comments_kor_b200 import ORBextractor
from orb_slam3_detailed_comments_kor_b200 import matcher as M
bench.set_workload("euroc")
W, H, PARAMS, LAP, MBF, FX = bench.W, bench.H, bench.PARAMS, bench.LAP, bench.MBF, bench.FX
P = 256
hL, hR = bench.gen_pairs(P, 0, pin=True)
dL, dR = hL.cuda(non_blocking=True), hR.cuda(non_blocking=True)
exL = ORBextractor(*PARAMS, width=W, height=H, device=0, max_batch=P)
exR = ORBextractor(*PARAMS, width=W, height=H, device=0, max_batch=P)
mode = sys.argv[1] if len(sys.argv) > 1 else "a"
tstream = torch.cuda.Stream()
if "s" in mode:
    torch.cuda.set_stream(tstream)
st = tstream.cuda_stream
def step_serial():
    exL.extract_batch_device(dL.data_ptr(), P, (0, 0), st)
    exR.extract_batch_device(dR.data_ptr(), P, (0, 0), st)
    M.compute_stereo_matches(exL, exR, P, MBF, 0.0, FX, st)
def step():
    M.stereo_frames_batch_device(exL, exR, dL.data_ptr(), dR.data_ptr(), P, MBF, 0.0, FX, (0, 0), st)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(tstream)
    for _ in range(n): fn()
    e1.record(tstream); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
if "1" in mode:
    M.stereo_frames_batch_device(exL, exR, dL.data_ptr(), dR.data_ptr(), 1, MBF, 0.0, FX, (0, 0), st); torch.cuda.synchronize()
if "x" in mode:
    exL.extract_batch_device(dL.data_ptr(), 1, (0, 0), st); torch.cuda.synchronize()
if "y" in mode:
    exL.extract_batch_device(dL.data_ptr(), 1, (0, 0), st); exR.extract_batch_device(dR.data_ptr(), 1, (0, 0), st); torch.cuda.synchronize()
if "z" in mode:
    exL.extract_batch_device(dL.data_ptr(), P, (0, 0), st); torch.cuda.synchronize()
if "r" in mode:
    print("serial %.3f" % t(step_serial))
print("concurrent %.3f" % t(step))
print("concurrent %.3f" % t(step))
print("serial %.3f" % t(step_serial))
print("concurrent %.3f" % t(step))
