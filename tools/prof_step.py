#!/usr/bin/env python3
"""Smallest program that runs the hot path a few times (for ncu captures under gpurun):
   python tools/prof_step.py [pairs=64] [steps=2] [workload=euroc|kitti|tumvi]
Left and right extraction + stereo matching on device-resident synthetic frames, as bench.py's timed region does."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth
from orb_slam3_detailed_comments_kor_b200 import matcher as M

SHAPES = {"euroc": (752, 480, (1200, 1.2, 8, 20, 7), 47.90639, 435.2047), "kitti": (1241, 376, (2000, 1.2, 8, 20, 7), 386.1448, 718.856),
          "tumvi": (512, 512, (1500, 1.2, 8, 20, 7), 0, 0)}
pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 64
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
W, H, params, mbf, fx = SHAPES[sys.argv[3] if len(sys.argv) > 3 else "euroc"]
uniq = min(pairs, 16)
L, R = zip(*[synth.stereo_pair(1000 + i, W, H) for i in range(uniq)])
dL = torch.from_numpy(np.stack([L[i % uniq] for i in range(pairs)])).cuda()
dR = torch.from_numpy(np.stack([R[i % uniq] for i in range(pairs)])).cuda()
exL = ORBextractor(*params, width=W, height=H, max_batch=pairs)
exR = ORBextractor(*params, width=W, height=H, max_batch=pairs)
st = torch.cuda.Stream()
torch.cuda.synchronize()
for _ in range(steps):
    exL.extract_batch_device(dL.data_ptr(), pairs, (0, 0), st.cuda_stream)
    exR.extract_batch_device(dR.data_ptr(), pairs, (0, 0), st.cuda_stream)
    if mbf:
        M.compute_stereo_matches(exL, exR, pairs, mbf, 0.0, fx, stream=st.cuda_stream)
    st.synchronize()
cnt, _ = exL.batch_counts(pairs)
print("ok: %d pairs x %d steps, %.1f keypoints per left frame" % (pairs, steps, float(cnt.mean())))
