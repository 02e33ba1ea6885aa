"""CPU tests: the oracle (oracle/liborb_oracle.so) against the cv2-4.13-generated goldens in tests/golden/.

The goldens were produced by tools/make_goldens.py, where cv2 itself executes every OpenCV primitive the
reference calls (src/ORBextractor.cc:808,827,1115,1165-1173; src/Frame.cc:911,1137) and the glue is an
independent Python restatement.  Bit-exact everywhere except angles (tolerance 1e-3 degrees, and in practice 0).
"""
import hashlib
import os

import numpy as np
import pytest

from oracle import orb_oracle as O
from orb_slam3_detailed_comments_kor_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["euroc", "tumvi", "kitti", "small", "noise", "lowc"]
ANGLE_TOL_DEG = 1e-3


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load_case(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    W, H = [int(v) for v in g["wh"]]
    if "imgL" in g:
        L, R = g["imgL"], g["imgR"]
    else:
        L, R = synth.stereo_pair(int(g["seed"][0]), W, H)
        assert sha(L) == str(g["imgL_sha"][0]) and sha(R) == str(g["imgR_sha"][0]), "synthetic generator drifted"
    p = g["params"]
    params = (int(p[0]), float(p[1]), int(p[2]), int(p[3]), int(p[4]))
    return g, L, R, params


def kps_matrix(k):
    return np.stack([k["x"], k["y"], k["size"], k["angle"], k["response"], k["octave"].astype(np.float32)], 1)


def assert_kps_equal(got, want):
    assert got.shape == want.shape
    cols = [0, 1, 2, 4, 5]
    assert np.array_equal(got[:, cols], want[:, cols]), "keypoint x/y/size/response/octave differ"
    assert np.max(np.abs(got[:, 3] - want[:, 3]), initial=0) <= ANGLE_TOL_DEG


@pytest.mark.parametrize("name", CASES)
def test_extractor_matches_cv2_pipeline(name):
    g, L, R, params = load_case(name)
    ex = O.OracleExtractor(*params)
    assert ex.quotas() == list(g["quota"])
    assert ex.umax() == list(g["umax"])
    assert np.array_equal(np.array(ex.scale_factors(), np.float32), g["scale"])
    li = 0
    while "lap%d" % li in g:
        lap = tuple(int(v) for v in g["lap%d" % li])
        mono, kps, desc = ex(L, lap)
        assert mono == int(g["mono%d" % li][0])
        assert_kps_equal(kps_matrix(kps), g["kps%d" % li])
        assert np.array_equal(desc, g["desc%d" % li])
        li += 1
    for l in range(params[2]):
        assert list(ex.level_size(l)) == list(g["level_wh"][l])
        assert sha(ex.pyramid_level(l, padded=True)) == str(g["pyr_sha"][l]), "pyramid level %d" % l
        b = ex.blurred_level(l)
        assert (sha(b) if b is not None else "") == str(g["blur_sha"][l]), "blurred level %d" % l
        c = ex.fast_candidates(l)
        want = g["cand%d" % l]
        got = np.stack([c["x"], c["y"], c["response"]], 1).astype(np.int32).reshape(-1, 3)
        assert np.array_equal(got, want), "FAST candidates level %d" % l
        lk = kps_matrix(ex.level_keypoints(l)).reshape(-1, 6)
        assert_kps_equal(lk, g["lkp%d" % l])


@pytest.mark.parametrize("name", ["euroc", "kitti", "lowc"])
def test_stereo_matches_cv2_pipeline(name):
    g, L, R, params = load_case(name)
    li = max(i for i in range(4) if "lap%d" % i in g)
    lap = tuple(int(v) for v in g["lap%d" % li])
    exL, exR = O.OracleExtractor(*params), O.OracleExtractor(*params)
    _, kL, dL = exL(L, lap)
    monoR, kR, dR = exR(R, lap)
    assert monoR == int(g["monoR"][0])
    assert_kps_equal(kps_matrix(kR), g["kpsR"])
    assert np.array_equal(dR, g["descR"])
    mbf, maxD = [float(v) for v in g["stereo_params"]]
    kept, u, d = O.stereo_match(exL, exR, kL, dL, kR, dR, mbf, 0.0, maxD)
    assert kept == int((g["uRight"] >= 0).sum()) and kept > 50
    assert np.array_equal(u, g["uRight"])
    assert np.array_equal(d, g["depth"])


@pytest.mark.parametrize("name", ["tumvi", "noise"])
def test_knn2_ratio_matches_bfmatcher(name):
    g, L, R, params = load_case(name)
    li = max(i for i in range(4) if "lap%d" % i in g)
    lap = tuple(int(v) for v in g["lap%d" % li])
    exL, exR = O.OracleExtractor(*params), O.OracleExtractor(*params)
    monoL, kL, dL = exL(L, lap)
    monoR, kR, dR = exR(R, lap)
    m, d0, d1 = O.knn2_ratio(dL[monoL:], dR[monoR:], 0.7)
    assert len(m) == len(g["knn_match"]) and len(m) > 100
    assert np.array_equal(m, g["knn_match"])
    assert np.array_equal(np.stack([d0, d1], 1), g["knn_d01"])
    assert (m >= 0).sum() > 0


def test_primitives_match_cv2():
    g = np.load(os.path.join(GOLD, "primitives.npz"))
    src = g["resize_src"]
    for k in range(4):
        dw, dh = [int(v) for v in g["resize_wh%d" % k]]
        assert np.array_equal(O.resize_linear(src, dw, dh), g["resize_out%d" % k])
    assert np.array_equal(O.gaussian7(src), g["blur_out"])
    for t in (20, 7):
        c = O.fast9_16(g["fast_roi"], t)
        got = np.stack([c["x"], c["y"], c["response"]], 1).astype(np.int32).reshape(-1, 3)
        assert np.array_equal(got, g["fast_out%d" % t]) and len(got) > 0
    got = np.array([O.fast_atan2(y, x) for y, x in g["atan_yx"]], np.float32)
    assert np.array_equal(got, g["atan_out"])


def test_known_answers():
    z = np.zeros(32, np.uint8)
    o = np.full(32, 255, np.uint8)
    one = z.copy(); one[17] = 0x10
    assert O.descriptor_distance(z, z) == 0
    assert O.descriptor_distance(z, o) == 256
    assert O.descriptor_distance(z, one) == 1
    ex = O.OracleExtractor(1200, 1.2, 8, 20, 7)
    assert ex.quotas() == [261, 217, 181, 151, 126, 105, 87, 72]
    assert ex.umax() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert O.OracleExtractor(1000, 1.2, 8, 20, 7).quotas() == [217, 181, 151, 126, 105, 87, 73, 60]
    mono, kps, desc = ex(np.zeros((0, 0), np.uint8))
    assert mono == -1                                                       # ORBextractor.cc:1072-1073


def test_top2_semantics():
    rng = np.random.RandomState(3)
    db = rng.randint(0, 256, (500, 32)).astype(np.uint8)
    q = db[[5, 100, 7]].copy()
    q[1, 0] ^= 0x0F                      # 4 bits off row 100
    db[300] = db[5]                      # exact duplicate -> best1 == best2 == 0, idx = lowest (5)
    out = O.hamming_top2(q, db, 50, 0.9, False)
    assert out["idx"][0] == 5 and out["best1"][0] == 0 and out["best2"][0] == 0 and out["accepted"][0] == 0
    assert out["idx"][1] == 100 and out["best1"][1] == 4 and out["accepted"][1] == 1
    # brute-force cross-check
    x = np.unpackbits(q[:, None, :] ^ db[None, :, :], axis=2).sum(2)
    srt = np.sort(x, axis=1)
    assert np.array_equal(out["best1"], srt[:, 0]) and np.array_equal(out["best2"], srt[:, 1])
    assert np.array_equal(out["idx"], np.argmin(x, axis=1))
    # TH boundary: `<=` (KF-Frame, ORBmatcher.cc:373) vs `<` (KF-KF, ORBmatcher.cc:906)
    q2 = db[[10]].copy()
    bits = np.unpackbits(q2[0]); bits[:50] ^= 1; q2[0] = np.packbits(bits)
    assert O.hamming_top2(q2, db, 50, 0.9, False)["accepted"][0] == 1
    assert O.hamming_top2(q2, db, 50, 0.9, True)["accepted"][0] == 0
