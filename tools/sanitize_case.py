"""Small end-to-end case for compute-sanitizer (memcheck / racecheck / initcheck), run on the GPU box:
   compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from orb_slam3_detailed_comments_kor_b200 import ORBextractor, ORBmatcher, synth
from orb_slam3_detailed_comments_kor_b200 import matcher as M

W, H, params, nb = 400, 300, (600, 1.2, 6, 20, 7), 3
pairs = [synth.stereo_pair(500 + i, W, H) for i in range(nb)]
Ls = np.stack([p[0] for p in pairs]); Rs = np.stack([p[1] for p in pairs])
Ls[1] = (Ls[1] // 6 + 100).astype(np.uint8)          # low contrast: exercises the minTh fallback kernel
exL = ORBextractor(*params, width=W, height=H, max_batch=nb)
exR = ORBextractor(*params, width=W, height=H, max_batch=nb)
cl, _, kl, dl = exL.extract_batch_host(Ls, (0, 200))
cr, _, kr, dr = exR.extract_batch_host(Rs, (0, 200))
M.compute_stereo_matches(exL, exR, nb, 40.0, 0.0, 300.0)
u, d = M.stereo_result(exL, 0)
res = ORBmatcher(0.9).search_top2(dl[0, :cl[0]], dr[0, :cr[0]])
m, d0, d1 = M.knn2_ratio(dl[0, :cl[0]], dr[0, :cr[0]], 0.7)
print("ok", cl, cr, int((u >= 0).sum()), int(res["accepted"].sum()), int((m >= 0).sum()))
