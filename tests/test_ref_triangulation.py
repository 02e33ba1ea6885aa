"""The reference's own KannalaBrandt8::TriangulateMatches (src/CameraModels/KannalaBrandt8.cpp:338-407, with Triangulate, unproject, project)
and ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:965-1206, both the pinhole and the two-camera branch, with
Pinhole::epipolarConstrain), cut out of their files and compiled into oracle/_ref against the float-matrix stand-in of
oracle/refstub/cvstub_float.hpp, against the oracle.  CPU only.

What the stand-in supplies is the arithmetic of the OpenCV routines those bodies call (gemm, addWeighted, dot, norm, SVD), each in the
form that was checked against cv2 4.13 (tests/test_triangulate.py); what is compared here is everything the reference's own text adds:
the order of operations, the conversions between float and double, the tests and their thresholds, the loop and its tie rules."""
import importlib.util
import os

import numpy as np
import pytest

from oracle import orb_oracle as O
from oracle import ref_oracle as R

HERE = os.path.dirname(os.path.abspath(__file__))
pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built (needs /root/reference)")


def _bow_helpers():
    spec = importlib.util.spec_from_file_location("_test_bow_helpers", os.path.join(HERE, "test_bow.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_triangulate_matches_equals_reference():
    """Accept / reject identical, depth and point identical to the bit: same libm (this host's), same OpenCV arithmetic models."""
    g = np.load(os.path.join(HERE, "golden", "triangulate.npz"))
    d, p = O.kb8_triangulate_matches(g["cam1"], g["cam2"], g["R12"], g["t12"], g["kp1"], g["kp2"], g["sigma1"], g["sigma2"])
    rd, rp = R.kb8_triangulate_matches(g["cam1"], g["cam2"], g["R12"], g["t12"], g["kp1"], g["kp2"], g["sigma1"], g["sigma2"])
    assert np.array_equal(d > 0, rd > 0)
    assert 100 < int((d > 0).sum()) < len(d) - 100
    assert np.array_equal(d, rd) and np.array_equal(p, rp)
    # a second rig and random pixel pairs, many of them unusable
    rng = np.random.RandomState(5)
    cam1 = np.array([190.978, 190.973, 254.93, 256.90, 0.00348, 0.000715, -0.00205, 0.000203], np.float32)
    cam2 = np.array([190.442, 190.434, 252.60, 254.92, 0.00340, 0.001766, -0.00266, 0.00033], np.float32)
    ang = 0.05
    R12 = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]], np.float32)
    t12 = np.array([0.101, 0.002, 0.001], np.float32)
    X = np.stack([rng.uniform(-2, 2, 3000), rng.uniform(-2, 2, 3000), rng.uniform(0.3, 6, 3000)], 1).astype(np.float32)
    uv1 = np.array([O.kb8_project(cam1, x) for x in X])
    X2 = ((X - t12) @ R12).astype(np.float32)                               # x_1 = R12 x_2 + t12
    uv2 = np.array([O.kb8_project(cam2, x) for x in X2]) + rng.normal(0, 0.7, (3000, 2)).astype(np.float32)
    s = (np.float32(1.2) ** rng.randint(0, 8, 3000)).astype(np.float32) ** 2
    d, p = O.kb8_triangulate_matches(cam1, cam2, R12, t12, uv1, uv2, s, s)
    rd, rp = R.kb8_triangulate_matches(cam1, cam2, R12, t12, uv1, uv2, s, s)
    assert np.array_equal(d, rd) and np.array_equal(p, rp)
    assert 300 < int((d > 0).sum()) < 2700


def test_kb8_project_and_unproject_equal_reference():
    rng = np.random.RandomState(8)
    cam = np.array([190.978, 190.973, 254.93, 256.90, 0.00348, 0.000715, -0.00205, 0.000203], np.float32)
    X = np.stack([rng.uniform(-3, 3, 20000), rng.uniform(-3, 3, 20000), rng.uniform(-0.5, 6, 20000)], 1).astype(np.float32)     # behind the camera too
    want = R.kb8_project(cam, X)
    got = np.array([O.kb8_project(cam, x) for x in X])
    assert np.array_equal(got, want)
    uv = rng.uniform(-50, 562, (20000, 2)).astype(np.float32)
    assert np.array_equal(np.array([O.kb8_unproject(cam, u[0], u[1]) for u in uv]), R.kb8_unproject(cam, uv))


PINHOLE_POSES = {   # keyframe 2 from the world (= keyframe 1): x_2 = R x_w + t
    "sideways": (0.0, [-0.11, 0.004, 0.02]),                                # the epipole far outside the image
    "forward": (0.02, [-0.03, 0.01, -0.35]),                                # the epipole inside the image: the :1089-1097 test removes candidates
}


@pytest.mark.parametrize("pose", sorted(PINHOLE_POSES))
@pytest.mark.parametrize("cfg", [(4, 3, 1, 31, 600), (10, 3, 2, 32, 1500), (3, 4, 4, 33, 900)])
def test_search_for_triangulation_pinhole_equals_reference(cfg, pose):
    tb = _bow_helpers()
    (d1, k1, u1, mp1, d2, k2, u2, mp2, _, _, scale, sigma2), fv1, fv2 = tb._triangulation_inputs(*cfg)
    ang, t2 = PINHOLE_POSES[pose]
    R2 = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]], np.float32)
    poses = np.concatenate([np.eye(3, dtype=np.float32).reshape(9), np.zeros(3, np.float32), R2.reshape(9), np.asarray(t2, np.float32)])
    K = np.array([458.654, 457.296, 367.215, 248.375, 0, 0, 0, 0], np.float32)
    cams = np.stack([K, K, K, K])
    seen = []
    for only_stereo, coarse, ori in ((False, False, True), (True, False, True), (False, True, False), (False, False, False)):
        rm, rnm, (F12, ep) = R.search_for_triangulation(d1, k1, u1, mp1, -1, O._fv_arrays(fv1), d2, k2, u2, mp2, -1, O._fv_arrays(fv2), cams, False, poses,
                                                        scale, sigma2, only_stereo, coarse, ori)
        m, nm = O.search_for_triangulation(d1, k1, u1, mp1, fv1, d2, k2, u2, mp2, fv2, F12, ep, scale, sigma2, 50, only_stereo, coarse, ori)
        assert nm == rnm and np.array_equal(m, rm), (cfg, pose, only_stereo, coarse, ori)
        seen.append(nm)
    assert seen[2] > seen[3] > 0                                            # the epipolar test cuts, and lets pairs through
    assert np.all(np.isfinite(ep))
    if pose == "forward":
        assert 0 < ep[0] < 752 and 0 < ep[1] < 480


@pytest.mark.parametrize("cfg", [(4, 3, 1, 51, 500), (10, 3, 2, 52, 700)])
def test_search_for_triangulation_rig_equals_reference(cfg):
    tb = _bow_helpers()
    d1, k1, mp1, d2, k2, mp2, _, geo, fv1, fv2, scale, sigma2 = tb._rig_inputs(*cfg)
    camL, camR, _, nl1, nl2 = geo
    f32 = np.float32
    # the generator's geometry as keyframe poses: world = keyframe 1's left camera; x_2 = x_1 + T21; x_left = x_right + tlr
    tlr = np.array([0.101, 0.002, 0.001], f32)
    T21 = np.array([-0.25, 0.0, 0.02], f32)
    I = np.eye(3, dtype=f32).reshape(9)
    left1, left2 = np.concatenate([I, np.zeros(3, f32)]), np.concatenate([I, T21])
    right1, right2 = np.concatenate([I, -tlr]), np.concatenate([I, (T21 - tlr).astype(f32)])      # x_r = x_l - tlr
    poses = np.concatenate([left1, left2, right1, right2])
    cams = np.stack([camL, camR, camL, camR])
    F, ep = np.zeros(9, f32), np.zeros(2, f32)
    counts = []
    for coarse, ori in ((False, True), (False, False), (True, False)):
        rm, rnm, (Rs, ts) = R.search_for_triangulation(d1, k1, None, mp1, nl1, O._fv_arrays(fv1), d2, k2, None, mp2, nl2, O._fv_arrays(fv2), cams, True, poses,
                                                       scale, sigma2, False, coarse, ori)
        rig = np.concatenate([np.concatenate([cams.reshape(32), Rs.reshape(36), ts.reshape(12)]).astype(f32).view(np.uint8), np.array([nl1, nl2], np.int32).view(np.uint8)])
        m, nm = O.search_for_triangulation(d1, k1, None, mp1, fv1, d2, k2, None, mp2, fv2, F, ep, scale, sigma2, 50, False, coarse, ori, rig=rig)
        assert nm == rnm and np.array_equal(m, rm), (cfg, coarse, ori)
        counts.append(nm)
    assert counts[2] > counts[1] > 20


@pytest.mark.parametrize("seed,lap", [(800, (0, 511)), (801, (100, 400)), (802, (0, 511))])
def test_fisheye_stereo_frame_equals_reference(seed, lap):
    """Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1119-1159) — the reference's own body, compiled (its BFMatcher::knnMatch is OpenCV's code:
    the stand-in keeps, per query, the two smallest Hamming distances in ascending order, the lower train row first among equals, which the knn
    goldens pin to cv2.BFMatcher) — against what the oracle composes for the same frame: knn-2 with `d0 < d1 * 0.7` in double over the rows
    inside the lapping interval, TriangulateMatches per survivor, `depth > 0.0001f`.  Every output vector identical, depths and points to the bit."""
    from orb_slam3_detailed_comments_kor_b200 import synth
    fL, fR = synth.stereo_pair(seed, 512, 512)
    params = (1500, 1.2, 8, 20, 7)
    xl, xr = O.OracleExtractor(*params), O.OracleExtractor(*params)
    ml, kl, dl = xl(fL, lap)
    mr, kr, dr = xr(fR, lap)
    cam1 = np.array([190.978, 190.973, 254.93, 256.90, 0.00348, 0.000715, -0.00205, 0.000203], np.float32)
    cam2 = np.array([190.442, 190.434, 252.60, 254.92, 0.00340, 0.001766, -0.00266, 0.00033], np.float32)
    if seed == 802:                                                       # the shim probe's rig: no distortion, a pure baseline
        cam1 = cam2 = np.array([190, 190, 256, 256, 0, 0, 0, 0], np.float32)
    Rlr, tlr = np.eye(3, dtype=np.float32), np.array([0.1, 0.0, 0.0], np.float32)
    sig = (np.asarray(xl.scale_factors(), np.float32) ** 2).astype(np.float32)
    l2r, r2l, depth, p3d, kept = R.fisheye_stereo(kl, dl, ml, kr, dr, mr, cam1, cam2, Rlr, tlr, sig)
    knn = O.knn2_ratio(dl[ml:], dr[mr:], 0.7)
    knn = knn[0] if isinstance(knn, tuple) else knn
    want_l2r, want_r2l = np.full(len(kl), -1, np.int32), np.full(len(kr), -1, np.int32)
    want_depth, want_p3d = np.full(len(kl), -1, np.float32), np.zeros((len(kl), 3), np.float32)
    qi = np.nonzero(knn >= 0)[0]
    if len(qi):
        li, ri = qi + ml, knn[qi] + mr
        d, p = O.kb8_triangulate_matches(cam1, cam2, Rlr, tlr, np.stack([kl["x"][li], kl["y"][li]], 1), np.stack([kr["x"][ri], kr["y"][ri]], 1),
                                         sig[kl["octave"][li]], sig[kr["octave"][ri]])
        for a, b, dd, pp in zip(li, ri, d, p):                            # in query order: a later query may overwrite mvRightToLeftMatch of the same right row
            if dd > np.float32(0.0001):
                want_l2r[a], want_r2l[b], want_depth[a], want_p3d[a] = b, a, dd, pp
    assert np.array_equal(l2r, want_l2r) and np.array_equal(r2l, want_r2l)
    assert np.array_equal(depth, want_depth) and np.array_equal(p3d, want_p3d)
    assert kept == int((want_l2r >= 0).sum()) and (kept >= 3 or seed == 801)
