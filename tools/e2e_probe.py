"""Where the end-to-end step spends its time (development aid)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth, _lib
import bench
P = 256
hL, hR = bench.gen_pairs(P, 0, pin=True)
ex = ORBextractor(*bench.PARAMS, width=bench.W, height=bench.H, max_batch=P)
cap = ex.capacity
out = (np.zeros(P, np.int32), np.zeros(P, np.int32), torch.empty((P, cap, 28), dtype=torch.uint8, pin_memory=True).numpy().view(_lib.KP_DTYPE).reshape(P, cap),
       torch.empty((P, cap, 32), dtype=torch.uint8, pin_memory=True).numpy())
x = hL.numpy()
for _ in range(3): ex.extract_batch_host(x, (0, 0), out)
t = time.perf_counter()
for _ in range(10): ex.extract_batch_host(x, (0, 0), out)
print("chunk", os.environ.get("ORBFE_HOST_CHUNK", "default"), "extract_batch_host ms per 256 frames: %.3f" % ((time.perf_counter() - t) * 100))
