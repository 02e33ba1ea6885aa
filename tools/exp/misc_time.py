"""Experiment (not a product path): device times of the entries bench.py has no leg for — the rig branch of SearchForTriangulation through
its host call, CLAHE and grey conversion device to device — to see whether any of them hides a serial chain like the ones the
SearchByBoW / SearchByProjection legs uncovered."""
import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from test_bow import _rig_inputs  # noqa: E402
from orb_slam3_detailed_comments_kor_b200 import bow as B
from orb_slam3_detailed_comments_kor_b200.conditioning import CLAHE, to_gray_device


def wall(fn, n=20):
    fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / n * 1e3


def dev(fn, n=20):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


F, ep = np.zeros(9, np.float32), np.zeros(2, np.float32)
for n, up in ((1000, 2), (3000, 2), (2000, 1), (3000, 1), (2000, 0)):
    d1, k1, mp1, d2, k2, mp2, rig, geo, fv1, fv2, scale, sigma2 = _rig_inputs(10, 3, up, 52, n)
    for coarse in (False, True):
        ms = wall(lambda: B.search_for_triangulation(d1, k1, None, mp1, fv1, d2, k2, None, mp2, fv2, F, ep, scale, sigma2, 50, False, coarse, True, rig=rig))
        print("rig triangulation host call n=%d levelsup=%d (%d nodes) coarse=%d: %.3f ms (rows %d / %d)" % (n, up, len(fv1), coarse, ms, len(d1), len(d2)))

# host entries: the per-call overhead (workspace, copies) beside the kernel
from test_bow import make_vocabulary, make_pair  # noqa: E402
from test_projection import _run  # noqa: E402
from test_initialization import random_case  # noqa: E402
from oracle import orb_oracle as O  # noqa: E402
from orb_slam3_detailed_comments_kor_b200 import projection as P  # noqa: E402
from orb_slam3_detailed_comments_kor_b200.synth import projection_cases as PC  # noqa: E402
c = PC.make_case(seed=21, n=2000, nq=1500, th=1.0)
print("search_by_projection host call n=2000 nq=1500: %.3f ms" % wall(lambda: _run(P.search_by_projection, c)))
c = PC.make_case(seed=21, n=200, nq=150, th=1.0)
print("search_by_projection host call n=200 nq=150:   %.3f ms" % wall(lambda: _run(P.search_by_projection, c)))
rows = make_vocabulary(7, 10, 3)
voc = O.OracleVocabulary(10, 3, 0, 0, *rows)
d_kf, k_kf, has_mp, d_f, k_f = make_pair(7, rows, 1500, 1500)
fv_kf, fv_f = voc.transform(d_kf, 1)["feat_vec"], voc.transform(d_f, 1)["feat_vec"]
print("search_by_bow host call 1500 x 1500: %.3f ms" % wall(lambda: B.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 50, 0.7, True)))
k1, d1, k2, d2, g, prev = random_case(5, 2000, 2000)
print("search_for_initialization host call 2000 x 2000: %.3f ms" % wall(lambda: P.search_for_initialization(k1, d1, k2, d2, g, prev, 100, 0.9, True)))

W, H, NF = 752, 480, 512
src = torch.randint(0, 256, (NF, H, W), dtype=torch.uint8, device="cuda")
dst = torch.empty_like(src)
cl = CLAHE(W, H, 3.0, (8, 8), max_batch=NF)
st = torch.cuda.Stream()
torch.cuda.set_stream(st)
s = st.cuda_stream
ms = dev(lambda: cl.batch_device(src.data_ptr(), dst.data_ptr(), NF, s))
print("CLAHE %d frames %dx%d: %.3f ms  (%.0f GB/s read+read+write)" % (NF, W, H, ms, NF * W * H * 3 / ms / 1e6))
rgb = torch.randint(0, 256, (NF, H, W, 3), dtype=torch.uint8, device="cuda")
ms = dev(lambda: to_gray_device(rgb.data_ptr(), 3, True, W, H, NF, dst.data_ptr(), 0, s))
print("gray  %d frames %dx%d x3: %.3f ms  (%.0f GB/s read+write)" % (NF, W, H, ms, NF * W * H * 4 / ms / 1e6))
