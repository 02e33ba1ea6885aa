"""Experiment: which of the CLAHE parity cases differ between the GPU and the oracle, and where (development aid for the banded apply kernel)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from test_conditioning import _clahe_cases, GOLD
from oracle import orb_oracle as O
from orb_slam3_detailed_comments_kor_b200.conditioning import CLAHE
g = np.load(GOLD)
for name, src, clip, tiles, want in _clahe_cases(g):
    c = CLAHE(src.shape[1], src.shape[0], clip, tiles)
    out = c.apply(src)
    ref = O.clahe(src, clip, tiles)
    bad = np.argwhere(out != ref)
    print(name, src.shape, tiles, "bad", len(bad), "first", bad[:5].tolist(), "rows", sorted(set(bad[:, 0].tolist()))[:12], "cols", sorted(set(bad[:, 1].tolist()))[:12])
    c.close()
