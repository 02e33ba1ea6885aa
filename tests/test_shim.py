"""The C++ drop-in shim (include/orbfe_shim): compiles as the reference's callers use it (CPU), and on a GPU
produces the oracle's results when driven like Frame.cc drives ORBextractor / ComputeStereoMatches."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "orb_slam3_detailed_comments_kor_b200")


def test_shim_compiles_without_opencv():
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-fsyntax-only", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "shim_smoke.cpp")])


REFERENCE_HEADERS = ["Atlas.h", "Config.h", "Converter.h", "Frame.h", "FrameDrawer.h", "G2oTypes.h", "ImuTypes.h", "Initializer.h", "KeyFrame.h",
                     "KeyFrameDatabase.h", "LocalMapping.h", "LoopClosing.h", "MLPnPsolver.h", "Map.h", "MapDrawer.h", "MapPoint.h", "ORBVocabulary.h",
                     "ORBextractor.h", "ORBmatcher.h", "OptimizableTypes.h", "Optimizer.h", "PnPsolver.h", "Sim3Solver.h", "System.h", "Tracking.h",
                     "TwoViewReconstruction.h", "Viewer.h"]   # /root/reference/include (absent on the GPU box, hence the list)


def test_shim_directory_shadows_only_the_header_it_fully_replaces():
    """INTEGRATION.md puts include/orbfe_shim in front of the reference's include/.  Every file there that carries the name of a
    reference header replaces that header for ALL of its includers, so only a complete stand-in may use such a name: that is
    ORBextractor.h (whole public surface of include/ORBextractor.h:43-109) and nothing else — in particular no ORBmatcher.h."""
    ref_inc = "/root/reference/include"
    names = set(REFERENCE_HEADERS)
    if os.path.isdir(ref_inc):
        found = {f for f in os.listdir(ref_inc) if f.endswith(".h")}
        assert found == names, "reference headers the list above does not know: %s" % sorted(found - names)
    shim = set(os.listdir(os.path.join(ROOT, "include", "orbfe_shim")))
    assert shim & names == {"ORBextractor.h"}
    src = open(os.path.join(ROOT, "include", "orbfe_shim", "ORBextractor.h")).read()
    for member in ("ORBextractor(int nfeatures, float scaleFactor, int nlevels", "operator()", "GetLevels", "GetScaleFactor", "GetScaleFactors",
                   "GetInverseScaleFactors", "GetScaleSigmaSquares", "GetInverseScaleSigmaSquares", "mvImagePyramid"):
        assert member in src, member


@pytest.mark.gpu
def test_shim_matches_oracle(tmp_path):
    from oracle import orb_oracle as O
    from orb_slam3_detailed_comments_kor_b200 import synth
    exe = str(tmp_path / "shim_smoke")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "shim_smoke.cpp"),
                           "-o", exe, "-L" + PKG, "-lorbfe", "-L" + os.path.join(PKG, "synth"), "-lsynth",
                           "-Wl,-rpath," + PKG, "-Wl,-rpath," + os.path.join(PKG, "synth")])
    out = subprocess.check_output([exe], text=True)
    f = dict(re.findall(r"(\w+)=([-\d.x(),]+)", out))
    params = (1200, 1.2, 8, 20, 7)
    L, R = synth.stereo_pair(0, 752, 480)
    oL, oR = O.OracleExtractor(*params), O.OracleExtractor(*params)
    mono, kL, dL = oL(L)
    monoR, kR, dR = oR(R)
    kept, u, d = O.stereo_match(oL, oR, kL, dL, kR, dR, 47.90639, 0.0, 435.2047)
    top2 = O.hamming_top2(dL, dR, 50, 0.9, False)
    assert int(f["n"]) == len(kL) and int(f["mono"]) == mono and int(f["monoR"]) == monoR
    assert f["pyr0"] == "752x480" and int(f["levels"]) == 8 and int(f["empty"]) == -1
    assert int(f["stereo"]) == kept
    assert int(f["d01"]) == O.descriptor_distance(dL[0], dL[1])
    assert int(f["top2acc"]) == int(top2["accepted"].sum())
    pq = np.zeros(len(kL), O.PROJ_QUERY_DTYPE)
    scale = np.asarray(oL.scale_factors(), np.float32)[kL["octave"]]
    pq["x"], pq["y"], pq["radius"], pq["xr"] = kL["x"] + np.float32(0.5), kL["y"] - np.float32(0.5), np.float32(4.0) * scale, u
    pq["min_level"], pq["max_level"], pq["flags"] = kL["octave"] - 1, kL["octave"], 2
    grid = np.array([0, 0, np.float32(64) / np.float32(752), np.float32(48) / np.float32(480)], np.float32)
    pm, _, nm = O.search_by_projection(kL, dL, pq, dL, grid, u_right=u, nnratio=0.8)
    assert int(f["proj"]) == nm and int(f["projself"]) == int(((pm["best_idx"] == np.arange(len(kL))) & (pm["accepted"] == 1)).sum())
    assert nm > 1000
    im, _, inm = O.search_for_initialization(kL, dL, kR, dR, grid, np.stack([kL["x"], kL["y"]], 1), 100, 0.9, True)
    assert int(f["init"]) == inm and int(f["initcount"]) == int((im >= 0).sum()) and inm > 10
    # SearchByBoW / SearchForTriangulation through the std::map wrappers: features grouped by the top six bits of their first descriptor byte
    fv = lambda d: {int(g): np.nonzero((d[:, 0] >> 2) == g)[0].astype(np.int32) for g in sorted(set((d[:, 0] >> 2).tolist()))}  # noqa: E731
    fvL, fvR = fv(dL), fv(dR)
    bm, _, bnm = O.search_by_bow(dL, kL, fvL, np.ones(len(kL), np.uint8), dR, kR, fvR, 50, 0.7, True)
    assert int(f["bow"]) == bnm, (f["bow"], bnm)
    assert int(f["bowsum"]) == int(((np.arange(len(bm)) + 1).astype(np.int64) * (bm.astype(np.int64) + 1)).sum())
    assert bnm > 5, bnm
    sf = np.asarray(oL.scale_factors(), np.float32)
    tm, tnm = O.search_for_triangulation(dL, kL, None, None, fvL, dR, kR, None, None, fvR, np.array([0, 0, 0, 0, 0, -1, 0, 1, 0], np.float32),
                                         np.array([-3000.0, 240.0], np.float32), sf, (sf * sf).astype(np.float32), 50, False, False, True)
    assert int(f["tri"]) == tnm, (f["tri"], tnm)
    assert int(f["trisum"]) == int(((np.arange(len(tm)) + 1).astype(np.int64) * (tm.astype(np.int64) + 1)).sum())
    assert tnm > 5, tnm
    # the fisheye rig of the probe: extraction with the lapping interval, knn-2 + 0.7, TriangulateMatches, depth test (src/Frame.cc:1119-1159)
    fL, fR = synth.stereo_pair(800, 512, 512)
    fp = (1500, 1.2, 8, 20, 7)
    xl, xr = O.OracleExtractor(*fp), O.OracleExtractor(*fp)
    ml, fkl, fdl = xl(fL, (0, 511)); mr, fkr, fdr = xr(fR, (0, 511))
    knn = O.knn2_ratio(fdl[ml:], fdr[mr:], 0.7)
    knn = knn[0] if isinstance(knn, tuple) else knn
    sig = np.asarray(xl.scale_factors(), np.float32) ** 2
    cam = np.array([190, 190, 256, 256, 0, 0, 0, 0], np.float32)
    qi = np.nonzero(knn >= 0)[0]
    li, ri = qi + ml, knn[qi] + mr
    dep, _ = O.kb8_triangulate_matches(cam, cam, np.eye(3, dtype=np.float32), np.array([0.1, 0, 0], np.float32), np.stack([fkl["x"][li], fkl["y"][li]], 1),
                                       np.stack([fkr["x"][ri], fkr["y"][ri]], 1), sig[fkl["octave"][li]], sig[fkr["octave"][ri]])
    keep = dep > 0.0001
    assert int(f["fish"]) == int(keep.sum()) and int(f["fishsum"]) == int(((li[keep] + 1).astype(np.int64) * (ri[keep] + 1)).sum())
    assert keep.sum() >= 3
    first = "(%.1f,%.1f,%.3f,%d)" % (kL[0]["x"], kL[0]["y"], kL[0]["angle"], kL[0]["octave"])
    assert f["first"] == first
