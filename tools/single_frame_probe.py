"""Single-frame latency of orbfe_extract (development aid): python tools/single_frame_probe.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth
W, H = 752, 480
ex = ORBextractor(1200, 1.2, 8, 20, 7, width=W, height=H, max_batch=1)
imgs = [synth.frame(i, W, H) for i in range(8)]
for i in range(10):
    ex(imgs[i % 8])
t = time.perf_counter()
for i in range(200):
    ex(imgs[i % 8])
print("graph", os.environ.get("ORBFE_SINGLE_FRAME_GRAPH", "1"), "single-frame latency %.1f us" % ((time.perf_counter() - t) / 200 * 1e6))
# per-stage device time of one frame (library-side CUDA events)
import ctypes as C
from orb_slam3_detailed_comments_kor_b200 import _lib
L = _lib.lib()
_lib.check(L.orbfe_set_profiling(ex.handle, 1))
buf = (C.c_double * 8)()
nl = C.c_int64(0)
for i in range(50):
    ex(imgs[i % 8])
_lib.check(L.orbfe_stage_times(ex.handle, buf, C.byref(nl)))
print("stage us per frame:", dict(zip(["pyramid", "fast", "blur", "quadtree", "assemble", "orient", "st", "stm"], [round(v * 1000 / 50, 1) for v in buf])))
