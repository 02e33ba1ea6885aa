"""Per-kernel shared-memory carve-out sweep: serial stage times and the concurrent step."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import bench
from orb_slam3_detailed_comments_kor_b200 import ORBextractor, _lib
from orb_slam3_detailed_comments_kor_b200 import matcher as M
if os.environ.get('ORBFE_EXP_LIB'): _lib.LIB_PATH = os.environ['ORBFE_EXP_LIB']
bench.set_workload("euroc")
W, H, PARAMS, LAP, MBF, FX = bench.W, bench.H, bench.PARAMS, bench.LAP, bench.MBF, bench.FX
P = 256
hL, hR = bench.gen_pairs(P, 0, pin=True, distinct=64)
dL, dR = hL.cuda(), hR.cuda()
exL = ORBextractor(*PARAMS, width=W, height=H, device=0, max_batch=P)
exR = ORBextractor(*PARAMS, width=W, height=H, device=0, max_batch=P)
L = _lib.lib()
ts = torch.cuda.Stream(); st = ts.cuda_stream
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
    e0.record(ts)
    for _ in range(n): fn()
    e1.record(ts); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
conc = t(step)
for ex in (exL, exR): _lib.check(L.orbfe_set_profiling(ex.handle, 1))
for _ in range(10): step_serial()
torch.cuda.synchronize()
ms = np.zeros(8)
for ex in (exL, exR):
    buf = (C.c_double * 8)(); n = C.c_int64(0)
    _lib.check(L.orbfe_stage_times(ex.handle, buf, C.byref(n))); ms += np.array(list(buf))
print("%-28s conc %.3f | " % (os.environ.get("ORBFE_EXP_CARVEOUT", "default"), conc) + " ".join("%.3f" % (v / 10) for v in ms[:7]), flush=True)
