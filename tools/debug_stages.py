"""Stage-by-stage GPU vs oracle comparison with verbose diagnostics (development aid; run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import orb_oracle as O
from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth

W, H, seed = 752, 480, 0
params = (1200, 1.2, 8, 20, 7)
if len(sys.argv) > 3:
    W, H, seed = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
if len(sys.argv) > 4:
    params = (int(sys.argv[4]), 1.2, int(sys.argv[5]) if len(sys.argv) > 5 else 8, 20, 7)
img = synth.frame(seed, W, H)
ex = ORBextractor(*params, width=W, height=H)
orc = O.OracleExtractor(*params)
mono, kps, desc = ex(img, (0, 0))
omono, okps, odesc = orc(img, (0, 0))
print("counts", len(kps), len(okps), "mono", mono, omono)
for l in range(params[2]):
    gp, op = ex.pyramid_level(l, 0, True), orc.pyramid_level(l, True)
    nbad = int((gp != op).sum())
    gc, oc = ex.fast_candidates(l), orc.fast_candidates(l)
    same_c = len(gc) == len(oc) and all(np.array_equal(gc[f], oc[f]) for f in ("x", "y", "response"))
    ob = orc.blurred_level(l)
    nb = int((ex.blurred_level(l) != ob).sum()) if ob is not None else -1
    gk = kps[kps["octave"] == l]; ok = okps[okps["octave"] == l]
    gs = set(zip(gk["x"].tolist(), gk["y"].tolist())); os_ = set(zip(ok["x"].tolist(), ok["y"].tolist()))
    print("level %d: pyr bad px %d | fast %d vs %d same=%s | blur bad %d | kps %d vs %d, set diff %d/%d, order same=%s" % (
        l, nbad, len(gc), len(oc), same_c, nb, len(gk), len(ok), len(gs - os_), len(os_ - gs),
        len(gk) == len(ok) and np.array_equal(gk["x"], ok["x"]) and np.array_equal(gk["y"], ok["y"])))
    if not same_c:
        gset = set(zip(gc["x"].tolist(), gc["y"].tolist(), gc["response"].tolist()))
        oset = set(zip(oc["x"].tolist(), oc["y"].tolist(), oc["response"].tolist()))
        print("   fast only-gpu", sorted(gset - oset)[:8], "only-oracle", sorted(oset - gset)[:8])
if len(kps) == len(okps):
    print("angle max diff", np.abs(kps["angle"] - okps["angle"]).max(), "desc rows differing", int((desc != odesc).any(1).sum()))

# first cell of level 0, in detail
import math
l = 0
pyr = orc.pyramid_level(l)
w, h = orc.level_size(l)
width, height = w - 32, h - 32
nC, nR = int(width / 35), int(height / 35)
wC, hC = math.ceil(width / nC), math.ceil(height / nR)
for (ci, cj) in [(0, 0), (0, 1), (5, 0)]:
    x0, y0 = 16 + cj * wC, 16 + ci * hC
    roi = pyr[y0:min(y0 + hC + 6, h - 16), x0:min(x0 + wC + 6, w - 16)]
    o = O.fast9_16(roi, 20)
    gc = ex.fast_candidates(l)
    sel = gc[(gc["x"] >= cj * wC + 3) & (gc["x"] < (cj + 1) * wC + 3) & (gc["y"] >= ci * hC + 3) & (gc["y"] < (ci + 1) * hC + 3)]
    print("cell", ci, cj, "roi", roi.shape, "x0,y0", x0, y0)
    print("  oracle:", [(int(a), int(b), int(c)) for a, b, c in zip(o["x"] + cj * wC, o["y"] + ci * hC, o["response"])])
    print("  gpu   :", [(int(a), int(b), int(c)) for a, b, c in zip(sel["x"], sel["y"], sel["response"])])
