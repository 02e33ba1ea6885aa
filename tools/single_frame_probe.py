"""Per-stage device time of one 752x480 frame (development aid)."""
import os, sys, time, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth, _lib
import bench
L = _lib.lib()
img = synth.frame(0, bench.W, bench.H)
ex = ORBextractor(*bench.PARAMS, width=bench.W, height=bench.H, max_batch=1)
d = torch.from_numpy(img).cuda(); s = torch.cuda.Stream()
for _ in range(5): ex(img)
t = time.perf_counter()
for _ in range(50): ex(img)
print("orbfe_extract host call: %.1f us" % ((time.perf_counter() - t) / 50 * 1e6))
_lib.check(L.orbfe_set_profiling(ex.handle, 1))
N = 50
for _ in range(N):
    ex.extract_batch_device(d.data_ptr(), 1, (0, 0), s.cuda_stream)
torch.cuda.synchronize()
buf = (C.c_double * 8)(); n = C.c_int64(0)
_lib.check(L.orbfe_stage_times(ex.handle, buf, C.byref(n)))
print("device us per frame:", {k: round(v / N * 1e3, 1) for k, v in zip(bench.STAGES, buf)}, "sum %.1f" % (sum(buf) / N * 1e3))
