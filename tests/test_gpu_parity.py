"""GPU parity tests (-m gpu): liborbfe.so, called through its C ABI, against the CPU oracle and the cv2 goldens.

Bar: bit-exact pyramid pixels, FAST candidates (x, y, score), blurred pixels, keypoints (x, y, size, response,
octave, order), descriptors, monoIndex, match indices/distances, stereo uRight/depth; angles within 1e-3 degrees
(north_star), in practice identical bits."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import orb_oracle as O                                        # noqa: E402
from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth        # noqa: E402
from orb_slam3_detailed_comments_kor_b200 import matcher as M               # noqa: E402
from orb_slam3_detailed_comments_kor_b200._lib import KP_DTYPE, OrbfeError  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ANGLE_TOL_DEG = 1e-3

SHAPES = {
    # name: (W, H, seed, (nfeatures, scale, levels, iniTh, minTh), lappings)
    "euroc": (752, 480, 0, (1200, 1.2, 8, 20, 7), [(0, 0), (200, 400)]),
    "tumvi": (512, 512, 1, (1500, 1.2, 8, 20, 7), [(0, 511)]),
    "kitti": (1241, 376, 2, (2000, 1.2, 8, 20, 7), [(0, 0), (0, 1000)]),
    "small": (160, 120, 3, (300, 1.2, 4, 20, 7), [(0, 0)]),
    "init5x": (752, 480, 11, (6000, 1.2, 8, 20, 7), [(0, 1000)]),     # mpIniORBextractor, src/Tracking.cc:1157
    # shapes and parameters off the beaten path: widths that are not multiples of 4 / 16, other scale factors, level
    # counts and FAST thresholds (exercise the byte-wise copy path, the partial last column group of the resize,
    # wider source windows, other cell grids)
    "odd641": (641, 481, 12, (800, 1.2, 6, 20, 7), [(0, 0), (100, 300)]),
    "scale11": (503, 377, 13, (900, 1.1, 8, 25, 9), [(0, 0)]),
    "scale15": (701, 499, 14, (700, 1.5, 4, 30, 10), [(0, 350)]),
    "one_level": (322, 242, 15, (400, 1.2, 1, 20, 7), [(0, 0)]),
    "tall": (300, 500, 16, (500, 1.25, 5, 15, 5), [(0, 0)]),
    # a large feature budget (the per-level quota near the 2045 the on-chip quadtree holds) on 2016 cells ...
    "big1980": (1980, 1270, 17, (5000, 1.2, 8, 20, 7), [(0, 1000)]),
    # mpIniORBextractor of the monocular KITTI settings (5 x 2000 features, src/Tracking.cc:1157): 2171 features on level 0, more than
    # the 2048 quadtree nodes shared memory holds -> the node arrays move to global scratch
    "kitti_init": (1241, 376, 20, (10000, 1.2, 8, 20, 7), [(0, 0)]),
    # ... and the largest image the library takes: 9690 cells on level 0, gathered in slices by the quadtree kernel
    "max4000": (4000, 3000, 19, (3000, 1.2, 8, 20, 7), [(0, 2000)]),
    # the smallest image the reference's cell grid accepts at 2 levels (a 35 px grid needs > 66 px on the coarsest level)
    "tiny": (96, 84, 18, (100, 1.2, 2, 20, 7), [(0, 0)]),
}


def kmat(k):
    return np.stack([k["x"], k["y"], k["size"], k["angle"], k["response"], k["octave"].astype(np.float32)], 1)


def assert_same_keypoints(got, want, what=""):
    assert len(got) == len(want), "%s: %d vs %d keypoints" % (what, len(got), len(want))
    g, w = kmat(got), kmat(want)
    cols = [0, 1, 2, 4, 5]
    bad = np.nonzero((g[:, cols] != w[:, cols]).any(1))[0]
    assert len(bad) == 0, "%s: first differing keypoint %d: %s vs %s" % (what, bad[0], g[bad[0]], w[bad[0]])
    assert np.max(np.abs(g[:, 3] - w[:, 3]), initial=0) <= ANGLE_TOL_DEG, what
    assert np.all(got["class_id"] == -1)


def compare_stages(ex, orc, nlevels, frame=0):
    for l in range(nlevels):
        assert ex.level_size(l) == orc.level_size(l)
        assert np.array_equal(ex.pyramid_level(l, frame, padded=True), orc.pyramid_level(l, padded=True)), "pyramid level %d" % l
        gc, oc = ex.fast_candidates(l, frame), orc.fast_candidates(l)
        assert len(gc) == len(oc), "FAST count level %d: %d vs %d" % (l, len(gc), len(oc))
        for fld in ("x", "y", "response"):
            assert np.array_equal(gc[fld], oc[fld]), "FAST %s level %d" % (fld, l)
        ob = orc.blurred_level(l)
        if ob is not None:
            assert np.array_equal(ex.blurred_level(l, frame), ob), "blur level %d" % l


@pytest.mark.parametrize("name", list(SHAPES))
def test_extractor_matches_oracle(name):
    W, H, seed, params, laps = SHAPES[name]
    img = synth.frame(seed, W, H)
    ex = ORBextractor(*params, width=W, height=H)
    orc = O.OracleExtractor(*params)
    assert list(ex.features_per_level()) == orc.quotas()
    assert list(ex.umax()) == orc.umax()
    assert np.array_equal(ex.GetScaleFactors(), np.array(orc.scale_factors(), np.float32))
    for lap in laps:
        mono, kps, desc = ex(img, lap)
        omono, okps, odesc = orc(img, lap)
        assert mono == omono
        assert_same_keypoints(kps, okps, "%s lap %s" % (name, lap))
        assert np.array_equal(desc, odesc)
    compare_stages(ex, orc, params[2])
    ex.close()


@pytest.mark.parametrize("name", ["euroc", "tumvi", "kitti", "small", "noise", "lowc"])
def test_extractor_matches_cv2_goldens(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    W, H = [int(v) for v in g["wh"]]
    img = g["imgL"] if "imgL" in g else synth.stereo_pair(int(g["seed"][0]), W, H)[0]
    p = g["params"]
    ex = ORBextractor(int(p[0]), float(p[1]), int(p[2]), int(p[3]), int(p[4]), width=W, height=H)
    li = 0
    while "lap%d" % li in g:
        lap = tuple(int(v) for v in g["lap%d" % li])
        mono, kps, desc = ex(img, lap)
        assert mono == int(g["mono%d" % li][0])
        want = g["kps%d" % li]
        got = kmat(kps)
        assert got.shape == want.shape
        assert np.array_equal(got[:, [0, 1, 2, 4, 5]], want[:, [0, 1, 2, 4, 5]])
        assert np.max(np.abs(got[:, 3] - want[:, 3]), initial=0) <= ANGLE_TOL_DEG
        assert np.array_equal(desc, g["desc%d" % li])
        li += 1
    for l in range(int(p[2])):
        c = ex.fast_candidates(l)
        got = np.stack([c["x"], c["y"], c["response"]], 1).astype(np.int32).reshape(-1, 3)
        assert np.array_equal(got, g["cand%d" % l]), "FAST candidates level %d" % l
    ex.close()


@pytest.mark.parametrize("pattern", ["dots", "checker", "stripes_noise"])
def test_regular_patterns_stress_ties(pattern):
    """Regular images produce many corners with equal scores (NMS ties: equal neighbours suppress each other) and many
    quadtree nodes of equal size (the creation-order tie-break and the partial rounds decide who survives)."""
    W, H, params = 640, 480, (1000, 1.2, 8, 20, 7)
    yy, xx = np.mgrid[0:H, 0:W]
    if pattern == "dots":
        img = np.where(((xx % 12) < 3) & ((yy % 12) < 3), 220, 40).astype(np.uint8)
    elif pattern == "checker":
        img = np.where(((xx // 16) + (yy // 16)) % 2 == 0, 200, 60).astype(np.uint8)
    else:
        rng = np.random.RandomState(5)
        img = (np.where((xx // 9) % 2 == 0, 180, 70) + rng.randint(-12, 13, (H, W))).clip(0, 255).astype(np.uint8)
    ex = ORBextractor(*params, width=W, height=H)
    orc = O.OracleExtractor(*params)
    for lap in [(0, 0), (0, 1000)]:
        mono, kps, desc = ex(img, lap)
        omono, okps, odesc = orc(img, lap)
        assert mono == omono
        assert_same_keypoints(kps, okps, "%s lap %s" % (pattern, lap))
        assert np.array_equal(desc, odesc)
    compare_stages(ex, orc, params[2])
    assert len(kps) > 300
    ex.close()


@pytest.mark.parametrize("name,stereo", [("euroc", (47.90639, 435.2047)), ("kitti", (386.1448, 718.856)), ("tumvi", None), ("init5x", None)])
def test_gpu_equals_compiled_reference(name, stereo):
    """CUDA output against the reference's own ORBextractor.cc / Frame::ComputeStereoMatches body compiled by oracle/Makefile
    (oracle/_ref/liborb_ref.so, prebuilt where /root/reference exists; see tests/test_ref_pinned.py): keypoints, descriptors,
    monoIndex, padded pyramid, uRight and depth bit-exact; angles within 1e-3 degrees."""
    from oracle import ref_oracle as Rf
    if not Rf.available():
        pytest.skip("oracle/_ref not shipped to this machine")
    W, H, seed, params, laps = SHAPES[name]
    L, R = synth.stereo_pair(seed, W, H)
    exL, exR = ORBextractor(*params, width=W, height=H), ORBextractor(*params, width=W, height=H)
    ref = Rf.ReferenceExtractor(*params)
    for lap in laps:
        mono, kps, desc = exL(L, lap)
        rmono, rkps, rdesc, pyr = ref(L, lap, want_pyramid=True)
        assert mono == rmono
        assert_same_keypoints(kps, rkps, "%s lap %s" % (name, lap))
        assert np.array_equal(desc, rdesc)
    for l in range(params[2]):
        assert np.array_equal(exL.pyramid_level(l, padded=True), pyr[l]), "padded pyramid level %d" % l
    if stereo:
        mbf, fx = stereo
        mb = np.float32(mbf) / np.float32(fx)              # Frame constructor, src/Frame.cc:173; maxD = mbf/mb, :827-829
        kL, dL, kR, dR, u, d = ref.stereo(L, R, mbf, float(mb))
        _, gkL, gdL = exL(L, (0, 0))
        _, gkR, gdR = exR(R, (0, 0))
        M.compute_stereo_matches(exL, exR, 1, mbf, 0.0, float(np.float32(mbf) / mb))
        gu, gd = M.stereo_result(exL, 0)
        assert_same_keypoints(gkR, kR, "right")
        assert np.array_equal(gdL, dL) and np.array_equal(gdR, dR)
        assert np.array_equal(gu, u) and np.array_equal(gd, d) and int((u >= 0).sum()) > 50
    exL.close(); exR.close()


def test_batch_paths_agree_with_oracle():
    import torch
    W, H, params, nb = 752, 480, (1200, 1.2, 8, 20, 7), 6
    imgs = np.stack([synth.frame(100 + i, W, H) for i in range(nb)])
    imgs[3] = 128                                        # a featureless frame inside the batch -> 0 keypoints
    ex = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cnt, mono, kps, desc = ex.extract_batch_host(imgs, (0, 0))
    d_imgs = torch.from_numpy(imgs).cuda()
    torch.cuda.synchronize()
    ex.extract_batch_device(d_imgs.data_ptr(), nb, (0, 0), torch.cuda.Stream().cuda_stream)
    torch.cuda.synchronize()
    cnt2, mono2 = ex.batch_counts(nb)
    assert np.array_equal(cnt, cnt2) and np.array_equal(mono, mono2)
    orc = O.OracleExtractor(*params)
    for f in range(nb):
        omono, okps, odesc = orc(imgs[f], (0, 0))
        assert cnt[f] == len(okps) and mono[f] == omono
        assert_same_keypoints(kps[f, :cnt[f]], okps, "frame %d" % f)
        assert np.array_equal(desc[f, :cnt[f]], odesc)
        m2, k2, d2 = ex.batch_result(f)
        assert m2 == omono and np.array_equal(d2, odesc)
        assert_same_keypoints(k2, okps, "device frame %d" % f)
        if f in (0, nb - 1):
            compare_stages(ex, orc, params[2], frame=f)
    assert cnt[3] == 0
    ex.close()


def test_host_batch_pipeline_over_several_chunks():
    """orbfe_extract_batch_host pipelines chunks of 64 frames (upload / kernels / download overlap): 150 frames = 3 chunks
    must give exactly what the one-launch device path gives, and what the oracle gives on sampled frames."""
    import torch
    W, H, params, nb = 320, 240, (500, 1.2, 8, 20, 7), 150
    imgs = np.stack([synth.frame(300 + i, W, H) for i in range(nb)])
    ex = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cnt, mono, kps, desc = ex.extract_batch_host(imgs, (0, 160))
    kps, desc = kps.copy(), desc.copy()
    d_imgs = torch.from_numpy(imgs).cuda()
    torch.cuda.synchronize()
    ex.extract_batch_device(d_imgs.data_ptr(), nb, (0, 160), torch.cuda.Stream().cuda_stream)
    torch.cuda.synchronize()
    cnt2, mono2 = ex.batch_counts(nb)
    assert np.array_equal(cnt, cnt2) and np.array_equal(mono, mono2) and cnt.min() > 50
    orc = O.OracleExtractor(*params)
    for f in range(nb):
        m2, k2, d2 = ex.batch_result(f)
        assert np.array_equal(kps[f, :cnt[f]], k2) and np.array_equal(desc[f, :cnt[f]], d2), "frame %d" % f
        if f in (0, 63, 64, 127, 128, 149):
            omono, okps, odesc = orc(imgs[f], (0, 160))
            assert omono == mono[f]
            assert_same_keypoints(k2, okps, "frame %d" % f)
            assert np.array_equal(d2, odesc)
    ex.close()


def test_quadtree_above_8192_nodes_per_level():
    """A per-level quota above 8192 (50000 features: 10860 on level 0): the partial rounds sort more than 8192 nodes, which
    the sort key has to index without aliasing.  Dense noise gives level 0 far more candidates than its quota."""
    W, H, params = 1600, 1200, (50000, 1.2, 8, 20, 7)
    rng = np.random.RandomState(9)
    img = rng.randint(0, 256, (H // 2, W // 2)).astype(np.uint8).repeat(2, 0).repeat(2, 1)
    img = (img.astype(np.int32) + rng.randint(-20, 21, (H, W))).clip(0, 255).astype(np.uint8)
    ex = ORBextractor(*params, width=W, height=H)
    orc = O.OracleExtractor(*params)
    mono, kps, desc = ex(img, (0, 0))
    omono, okps, odesc = orc(img, (0, 0))
    assert int((okps["octave"] == 0).sum()) > 8192, "the case must exercise more than 8192 nodes on a level"
    assert mono == omono
    assert_same_keypoints(kps, okps, "50000 features")
    assert np.array_equal(desc, odesc)
    ex.close()


def test_extractors_of_different_sizes_side_by_side():
    """Tracking keeps mpIniORBextractor (5 x nFeatures) next to the per-frame extractors (src/Tracking.cc:1151-1157), and
    the kernels' dynamic shared-memory limits belong to the device, not to an object: creating a small extractor after a
    large one must not take the large one's launches away.  Large -> small -> large again, single-frame (graph) and batch paths."""
    W, H = 752, 480
    big_p, small_p = (5000, 1.2, 8, 20, 7), (300, 1.2, 8, 20, 7)
    img = synth.frame(31, W, H)
    big = ORBextractor(*big_p, width=W, height=H, max_batch=2)
    _, k0, d0 = big(img, (0, 0))
    small = ORBextractor(*small_p, width=W, height=H)
    small2 = ORBextractor(*small_p, width=W // 2, height=H // 2)          # other cell sizes -> other FAST tile size
    _, ks, ds = small(img, (0, 0))
    _, k1, d1 = big(img, (0, 0))
    assert np.array_equal(k0, k1) and np.array_equal(d0, d1)
    cnt, mono, kb, db = big.extract_batch_host(np.stack([img, img]), (0, 0))
    assert cnt[0] == len(k0) and np.array_equal(db[0][:cnt[0]], d0) and np.array_equal(db[1][:cnt[1]], d0)
    okb, oks = O.OracleExtractor(*big_p)(img, (0, 0)), O.OracleExtractor(*small_p)(img, (0, 0))
    assert_same_keypoints(k1, okb[1], "big")
    assert_same_keypoints(ks, oks[1], "small")
    assert np.array_equal(d1, okb[2]) and np.array_equal(ds, oks[2])
    big.close(); small.close(); small2.close()


def test_capacity_limits_are_refused_loudly():
    """What the library cannot hold (more than 16384 quadtree nodes on a level, a side above 4000 px) is an error at creation, not
    a wrong result."""
    for W, H, nfeat in ((752, 480, 80000), (4001, 480, 1000), (480, 4001, 1000)):
        with pytest.raises(OrbfeError):
            ORBextractor(nfeat, 1.2, 8, 20, 7, width=W, height=H)


def test_error_behaviour():
    ex = ORBextractor(1200, 1.2, 8, 20, 7, width=752, height=480)
    mono, kps, desc = ex(np.zeros((0, 0), np.uint8))
    assert mono == -1 and len(kps) == 0                 # src/ORBextractor.cc:1072-1073
    with pytest.raises(OrbfeError):
        ex(np.zeros((100, 100), np.uint8))               # shape fixed at creation
    with pytest.raises(OrbfeError):
        ORBextractor(1200, 1.2, 8, 20, 7, width=100, height=80)   # level 7 too small for the 35 px grid
    ex.close()


@pytest.mark.parametrize("name,stereo", [("euroc", (47.90639, 435.2047)), ("kitti", (386.1448, 718.856))])
def test_stereo_matches_oracle_and_golden(name, stereo):
    W, H, seed, params, laps = SHAPES[name]
    L, R = synth.stereo_pair(seed, W, H)
    exL = ORBextractor(*params, width=W, height=H)
    exR = ORBextractor(*params, width=W, height=H)
    oL, oR = O.OracleExtractor(*params), O.OracleExtractor(*params)
    g = np.load(os.path.join(GOLD, name + ".npz"))
    for lap in laps:                                     # the goldens were made with the last lapping interval
        monoL, kL, dL = exL(L, lap)
        monoR, kR, dR = exR(R, lap)
        M.compute_stereo_matches(exL, exR, 1, stereo[0], 0.0, stereo[1])
        u, d = M.stereo_result(exL, 0)
        _, okL, odL = oL(L, lap)
        _, okR, odR = oR(R, lap)
        kept, ou, od = O.stereo_match(oL, oR, okL, odL, okR, odR, stereo[0], 0.0, stereo[1])
        assert kept > 50
        assert np.array_equal(u, ou), "uRight: %d differ" % int((u != ou).sum())
        assert np.array_equal(d, od)
    assert np.array_equal(u, g["uRight"]) and np.array_equal(d, g["depth"])
    exL.close(); exR.close()


def test_stereo_batch_and_no_match_frame():
    W, H, params, nb = 752, 480, (1200, 1.2, 8, 20, 7), 3
    pairs = [synth.stereo_pair(200 + i, W, H) for i in range(nb)]
    Ls = np.stack([p[0] for p in pairs]); Rs = np.stack([p[1] for p in pairs])
    Rs[1] = 128                                          # right image without features: no matches, no median cut
    exL = ORBextractor(*params, width=W, height=H, max_batch=nb)
    exR = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cntL, _, kL, dL = exL.extract_batch_host(Ls)
    cntR, _, kR, dR = exR.extract_batch_host(Rs)
    M.compute_stereo_matches(exL, exR, nb, 47.90639, 0.0, 435.2047)
    for f in range(nb):
        u, d = M.stereo_result(exL, f)
        oL, oR = O.OracleExtractor(*params), O.OracleExtractor(*params)
        _, okL, odL = oL(Ls[f]); _, okR, odR = oR(Rs[f])
        kept, ou, od = O.stereo_match(oL, oR, okL, odL, okR, odR, 47.90639, 0.0, 435.2047)
        assert np.array_equal(u, ou) and np.array_equal(d, od), "frame %d" % f
        if f == 1:
            assert kept == 0 and np.all(u == -1)
    exL.close(); exR.close()


def test_stereo_frames_pipeline_matches_separate_calls():
    """orbfe_stereo_frames_batch_host (upload, both extractions, stereo matching, download pipelined over chunks in one call) must
    return exactly what extract_batch_host x 2 + compute_stereo_matches return; 150 pairs = three chunks, one of them short."""
    W, H, params, nb = 320, 240, (500, 1.2, 8, 20, 7), 150
    pairs = [synth.stereo_pair(400 + i, W, H) for i in range(nb)]
    Ls = np.stack([p[0] for p in pairs]); Rs = np.stack([p[1] for p in pairs])
    exL = ORBextractor(*params, width=W, height=H, max_batch=nb)
    exR = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cl, ml, kl, dl, cr, mr, kr, dr, u, d = [a.copy() for a in M.stereo_frames_batch_host(exL, exR, Ls, Rs, 40.0, 0.0, 300.0, (0, 0))]
    wl = [a.copy() for a in exL.extract_batch_host(Ls, (0, 0))]
    wr = [a.copy() for a in exR.extract_batch_host(Rs, (0, 0))]
    M.compute_stereo_matches(exL, exR, nb, 40.0, 0.0, 300.0)
    assert np.array_equal(cl, wl[0]) and np.array_equal(ml, wl[1]) and np.array_equal(cr, wr[0]) and np.array_equal(mr, wr[1])
    matched = 0
    for f in range(nb):
        assert np.array_equal(kl[f, :cl[f]], wl[2][f, :cl[f]]) and np.array_equal(dl[f, :cl[f]], wl[3][f, :cl[f]]), "left frame %d" % f
        assert np.array_equal(kr[f, :cr[f]], wr[2][f, :cr[f]]) and np.array_equal(dr[f, :cr[f]], wr[3][f, :cr[f]]), "right frame %d" % f
        wu, wd = M.stereo_result(exL, f)
        assert np.array_equal(u[f, :cl[f]], wu) and np.array_equal(d[f, :cl[f]], wd), "stereo of pair %d" % f
        matched += int((wu >= 0).sum())
    assert matched > 20 * nb
    # and against the oracle on one pair of the last (short) chunk
    oL, oR = O.OracleExtractor(*params), O.OracleExtractor(*params)
    _, okL, odL = oL(Ls[149]); _, okR, odR = oR(Rs[149])
    kept, ou, od = O.stereo_match(oL, oR, okL, odL, okR, odR, 40.0, 0.0, 300.0)
    assert np.array_equal(u[149, :cl[149]], ou) and np.array_equal(d[149, :cl[149]], od)
    exL.close(); exR.close()


def test_stereo_frames_submit_wait_keeps_two_batches_in_flight():
    """orbfe_stereo_frames_submit / _wait: four different batches streamed with two in flight (two alternating output sets) give,
    batch by batch, what the synchronous call gives; a third submit without a wait is refused; wait without a batch is an error."""
    import torch
    W, H, params, nb = 320, 240, (500, 1.2, 8, 20, 7), 70                  # 70 pairs = two chunks, one of them short
    batches = []
    for bi in range(4):
        pairs = [synth.stereo_pair(900 + 100 * bi + i, W, H) for i in range(nb)]
        batches.append((torch.from_numpy(np.stack([p[0] for p in pairs])).pin_memory().numpy(), torch.from_numpy(np.stack([p[1] for p in pairs])).pin_memory().numpy()))
    exL = ORBextractor(*params, width=W, height=H, max_batch=nb)
    exR = ORBextractor(*params, width=W, height=H, max_batch=nb)
    want = [[a.copy() for a in M.stereo_frames_batch_host(exL, exR, Lb, Rb, 40.0, 0.0, 300.0, (0, 0))] for Lb, Rb in batches]
    cap = exL.capacity

    def out_set():
        side = lambda: (np.zeros(nb, np.int32), np.zeros(nb, np.int32), np.zeros((nb, cap), KP_DTYPE), np.zeros((nb, cap, 32), np.uint8))  # noqa: E731
        return side() + side() + (np.zeros((nb, cap), np.float32), np.zeros((nb, cap), np.float32))
    outs = [out_set(), out_set()]
    got = []
    M.stereo_frames_submit(exL, exR, batches[0][0], batches[0][1], 40.0, 0.0, 300.0, (0, 0), outs[0])
    for k in range(1, 4):
        M.stereo_frames_submit(exL, exR, batches[k][0], batches[k][1], 40.0, 0.0, 300.0, (0, 0), outs[k & 1])
        if k == 1:
            with pytest.raises(OrbfeError):
                M.stereo_frames_submit(exL, exR, batches[2][0], batches[2][1], 40.0, 0.0, 300.0, (0, 0), outs[0])
        got.append([a.copy() for a in M.stereo_frames_wait(exL, exR, outs[(k - 1) & 1])])
    got.append([a.copy() for a in M.stereo_frames_wait(exL, exR, outs[1])])
    with pytest.raises(OrbfeError):
        M.stereo_frames_wait(exL, exR, outs[0])
    for bi in range(4):
        cl, ml, kl, dl, cr, mr, kr, dr, u, d = got[bi]
        wl = want[bi]
        assert np.array_equal(cl, wl[0]) and np.array_equal(ml, wl[1]) and np.array_equal(cr, wl[4]) and np.array_equal(mr, wl[5]), "batch %d" % bi
        for f in range(nb):
            assert np.array_equal(kl[f, :cl[f]], wl[2][f, :cl[f]]) and np.array_equal(dl[f, :cl[f]], wl[3][f, :cl[f]]), "batch %d left %d" % (bi, f)
            assert np.array_equal(kr[f, :cr[f]], wl[6][f, :cr[f]]) and np.array_equal(dr[f, :cr[f]], wl[7][f, :cr[f]]), "batch %d right %d" % (bi, f)
            assert np.array_equal(u[f, :cl[f]], wl[8][f, :cl[f]]) and np.array_equal(d[f, :cl[f]], wl[9][f, :cl[f]]), "batch %d stereo %d" % (bi, f)
    assert not np.array_equal(got[0][3], got[1][3])                      # the batches really differ
    exL.close(); exR.close()


def test_stereo_frames_batch_device_left_right_side_by_side():
    """orbfe_stereo_frames_batch_device / orbfe_extract_pair_batch_device (left and right on their own streams, the blur beside the
    quadtree) against the oracle for the first pairs and against the one-stream calls for all; repeated, on the caller's stream and
    without one, with a consumer enqueued right behind on the caller's stream."""
    import torch
    W, H, params, nb = 320, 240, (500, 1.2, 8, 20, 7), 24
    pairs = [synth.stereo_pair(4100 + i, W, H) for i in range(nb)]
    dLs = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    dRs = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    exL = ORBextractor(*params, width=W, height=H, max_batch=nb)
    exR = ORBextractor(*params, width=W, height=H, max_batch=nb)
    st = torch.cuda.Stream()
    exL.extract_batch_device(dLs.data_ptr(), nb, (0, 0), st.cuda_stream)
    exR.extract_batch_device(dRs.data_ptr(), nb, (0, 0), st.cuda_stream)
    M.compute_stereo_matches(exL, exR, nb, 40.0, 0.0, 300.0, st.cuda_stream)
    torch.cuda.synchronize()

    def snapshot():
        out = []
        for f in range(nb):
            ml, kl, dl = exL.batch_result(f)
            mr, kr, dr = exR.batch_result(f)
            u, d = M.stereo_result(exL, f)
            out.append((ml, kl.copy(), dl.copy(), mr, kr.copy(), dr.copy(), u.copy(), d.copy()))
        return out
    want = snapshot()
    oL, oR = O.OracleExtractor(*params), O.OracleExtractor(*params)
    for f in range(3):
        _, okl, odl = oL(pairs[f][0], (0, 0))
        _, okr, odr = oR(pairs[f][1], (0, 0))
        assert_same_keypoints(want[f][1], okl)
        assert np.array_equal(want[f][2], odl)
        _, ou, od = O.stereo_match(oL, oR, okl, odl, okr, odr, 40.0, 0.0, 300.0)
        assert np.array_equal(want[f][6], ou) and np.array_equal(want[f][7], od)
    for rep, stream in enumerate([st, None, st]):
        with torch.cuda.stream(st):
            dLs2, dRs2 = dLs.clone(), dRs.clone()                       # produced on the caller's stream right before the call
        M.stereo_frames_batch_device(exL, exR, dLs2.data_ptr(), dRs2.data_ptr(), nb, 40.0, 0.0, 300.0, (0, 0), stream.cuda_stream if stream else None)
        if stream is None:
            torch.cuda.synchronize()                                        # no caller stream: the getters wait for the objects themselves
        got = snapshot()
        for f in range(nb):
            for a, b in zip(got[f], want[f]):
                assert np.array_equal(a, b), "rep %d frame %d" % (rep, f)
    # the rig variant: both extractions, then a consumer (the stereo matcher as a separate call) on the caller's stream
    M.extract_pair_batch_device(exL, exR, dLs.data_ptr(), dRs.data_ptr(), nb, (0, 0), st.cuda_stream)
    M.compute_stereo_matches(exL, exR, nb, 40.0, 0.0, 300.0, st.cuda_stream)
    got = snapshot()
    for f in range(nb):
        for a, b in zip(got[f], want[f]):
            assert np.array_equal(a, b), "pair call, frame %d" % f
    with pytest.raises(OrbfeError):
        M.stereo_frames_batch_device(exL, exL, dLs.data_ptr(), dRs.data_ptr(), nb, 40.0, 0.0, 300.0)
    with pytest.raises(OrbfeError):
        M.stereo_frames_batch_device(exL, exR, dLs.data_ptr(), dRs.data_ptr(), nb + 1, 40.0, 0.0, 300.0)
    exL.close(); exR.close()


def test_entry_points_interleaved_on_the_same_objects():
    """The same two extractor objects driven through every way in, one after the other and repeatedly: device pair call on a caller's
    stream, pipelined host batches, synchronous host batch, single frame (CUDA graph), device pair call without a caller stream.  The
    calls share streams, events and staging buffers inside the objects; every one must give what the first gave."""
    import torch
    W, H, params, nb = 320, 240, (500, 1.2, 8, 20, 7), 40
    pairs = [synth.stereo_pair(5200 + i, W, H) for i in range(nb)]
    hL = torch.from_numpy(np.stack([p[0] for p in pairs])).pin_memory(); hR = torch.from_numpy(np.stack([p[1] for p in pairs])).pin_memory()
    dLs, dRs = hL.cuda(), hR.cuda()
    exL = ORBextractor(*params, width=W, height=H, max_batch=nb)
    exR = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cap = exL.capacity
    want = [a.copy() for a in M.stereo_frames_batch_host(exL, exR, hL.numpy(), hR.numpy(), 40.0, 0.0, 300.0, (0, 0))]
    st = torch.cuda.Stream()

    def device_snapshot():
        cl, _ = exL.batch_counts(nb); cr, _ = exR.batch_counts(nb)
        assert np.array_equal(cl, want[0]) and np.array_equal(cr, want[4])
        for f in (0, nb // 2, nb - 1):
            ml, kl, dl = exL.batch_result(f); mr, kr, dr = exR.batch_result(f)
            u, d = M.stereo_result(exL, f)
            assert np.array_equal(kl, want[2][f, :cl[f]]) and np.array_equal(dl, want[3][f, :cl[f]]) and np.array_equal(kr, want[6][f, :cr[f]])
            assert np.array_equal(u, want[8][f, :cl[f]]) and np.array_equal(d, want[9][f, :cl[f]])

    def out_set():
        side = lambda: (np.zeros(nb, np.int32), np.zeros(nb, np.int32), np.zeros((nb, cap), KP_DTYPE), np.zeros((nb, cap, 32), np.uint8))  # noqa: E731
        return side() + side() + (np.zeros((nb, cap), np.float32), np.zeros((nb, cap), np.float32))
    outs = [out_set(), out_set()]
    for rep in range(3):
        M.stereo_frames_batch_device(exL, exR, dLs.data_ptr(), dRs.data_ptr(), nb, 40.0, 0.0, 300.0, (0, 0), st.cuda_stream)
        device_snapshot()
        M.stereo_frames_submit(exL, exR, hL.numpy(), hR.numpy(), 40.0, 0.0, 300.0, (0, 0), outs[0])
        M.stereo_frames_submit(exL, exR, hL.numpy(), hR.numpy(), 40.0, 0.0, 300.0, (0, 0), outs[1])
        with pytest.raises(OrbfeError):                                     # host batches in flight: the device pair call is refused
            M.stereo_frames_batch_device(exL, exR, dLs.data_ptr(), dRs.data_ptr(), nb, 40.0, 0.0, 300.0)
        for o in outs:
            got = M.stereo_frames_wait(exL, exR, o)
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[4], want[4])
            for f in range(nb):
                n = want[0][f]
                assert np.array_equal(got[2][f, :n], want[2][f, :n]) and np.array_equal(got[3][f, :n], want[3][f, :n])
                assert np.array_equal(got[8][f, :n], want[8][f, :n]) and np.array_equal(got[9][f, :n], want[9][f, :n])
        mono, k, dsc = exL(pairs[3][0], (0, 0))                             # single frame: frame 0 of the object is now pair 3's left image
        assert np.array_equal(k, want[2][3, :want[0][3]]) and np.array_equal(dsc, want[3][3, :want[0][3]])
        M.stereo_frames_batch_device(exL, exR, dLs.data_ptr(), dRs.data_ptr(), nb, 40.0, 0.0, 300.0)   # own streams only
        device_snapshot()
        got = M.stereo_frames_batch_host(exL, exR, hL.numpy(), hR.numpy(), 40.0, 0.0, 300.0, (0, 0))
        for a, b in zip(got, want):
            assert np.array_equal(a[:nb], b[:nb]) if a.ndim == 1 else True
    exL.close(); exR.close()


def test_kitti_full_size_batch_properties():
    """BASELINE config 4 at full size: 4096 KITTI-shaped images (2048 stereo pairs, 1241x376, 2000 features) in one device
    batch per camera.  The batch is 16 distinct pairs repeated in a scrambled order: every frame must give exactly what its
    first occurrence gives (position independence across the whole batch), and the 16 distinct pairs must give the oracle's
    keypoints, descriptors and stereo matches."""
    import torch
    W, H, params, mbf, fx = 1241, 376, (2000, 1.2, 8, 20, 7), 386.1448, 718.856
    ND, NP = 16, 2048
    pairs = [synth.stereo_pair(700 + i, W, H) for i in range(ND)]
    order = (np.arange(NP) * 7 + (np.arange(NP) // ND) * 3) % ND            # every source appears 128 times, scattered
    order[:ND] = np.arange(ND)                                             # ... and first in its natural position
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        dL = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()[torch.from_numpy(order).cuda()].contiguous()
        dR = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()[torch.from_numpy(order).cuda()].contiguous()
    st.synchronize()
    exL = ORBextractor(*params, width=W, height=H, max_batch=NP)
    exR = ORBextractor(*params, width=W, height=H, max_batch=NP)
    exL.extract_batch_device(dL.data_ptr(), NP, (0, 0), st.cuda_stream)
    exR.extract_batch_device(dR.data_ptr(), NP, (0, 0), st.cuda_stream)
    M.compute_stereo_matches(exL, exR, NP, mbf, 0.0, fx, stream=st.cuda_stream)
    st.synchronize()
    cntL, monoL = exL.batch_counts(NP)
    cntR, monoR = exR.batch_counts(NP)
    assert np.array_equal(cntL, cntL[order]) and np.array_equal(cntR, cntR[order]) and np.array_equal(monoL, cntL)
    assert cntL.min() > 1500
    first = {}
    for f in range(NP):
        _, k, d = exL.batch_result(f)
        _, kr, dr = exR.batch_result(f)
        u, dep = M.stereo_result(exL, f)
        s = int(order[f])
        if f < ND:
            first[s] = (k, d, kr, dr, u, dep)
        else:
            w = first[s]
            assert np.array_equal(k, w[0]) and np.array_equal(d, w[1]) and np.array_equal(kr, w[2]) and np.array_equal(dr, w[3]), "frame %d" % f
            assert np.array_equal(u, w[4]) and np.array_equal(dep, w[5]), "stereo of frame %d" % f
    oL, oR = O.OracleExtractor(*params), O.OracleExtractor(*params)
    for s_ in range(ND):
        _, okL, odL = oL(pairs[s_][0]); _, okR, odR = oR(pairs[s_][1])
        kept, ou, od = O.stereo_match(oL, oR, okL, odL, okR, odR, mbf, 0.0, fx)
        k, d, kr, dr, u, dep = first[s_]
        assert_same_keypoints(k, okL, "pair %d left" % s_); assert_same_keypoints(kr, okR, "pair %d right" % s_)
        assert np.array_equal(d, odL) and np.array_equal(dr, odR)
        assert kept > 100 and np.array_equal(u, ou) and np.array_equal(dep, od), "pair %d stereo" % s_
    exL.close(); exR.close()


def test_descriptor_distance_known_answers():
    z = np.zeros(32, np.uint8); o = np.full(32, 255, np.uint8)
    one = z.copy(); one[9] = 0x40
    assert M.ORBmatcher.DescriptorDistance(z, z) == 0
    assert M.ORBmatcher.DescriptorDistance(z, o) == 256
    assert M.ORBmatcher.DescriptorDistance(z, one) == 1
    rng = np.random.RandomState(0)
    a = rng.randint(0, 256, (1000, 32)).astype(np.uint8); b = rng.randint(0, 256, (1000, 32)).astype(np.uint8)
    want = np.unpackbits(a ^ b, axis=1).sum(1)
    assert np.array_equal(M.ORBmatcher.DescriptorDistance(a, b), want)


@pytest.mark.parametrize("nq,ndb", [(1, 1), (3, 2), (37, 300), (1024, 5000), (1500, 70001), (4096, 20000)])
@pytest.mark.parametrize("strict", [False, True])
def test_top2_matches_oracle(nq, ndb, strict):
    db, q = synth.descriptors(nq * 7 + ndb, max(ndb, 2), nq)
    db = db[:ndb].copy()
    if ndb > 100:
        db[ndb // 2] = db[3]                             # planted duplicate rows: ties -> lowest index
        q[0] = db[3]
    m = M.ORBmatcher(0.9, True)
    got = m.search_top2(q, db, kf_kf=strict)
    want = O.hamming_top2(q, db, 50, 0.9, strict)
    for fld in ("best1", "best2", "idx", "accepted"):
        assert np.array_equal(got[fld], want[fld]), fld
    if ndb > 100:
        assert got["idx"][0] == 3 and got["best1"][0] == 0 and got["best2"][0] == 0


def test_top2_full_size_properties_and_sharding():
    """Config 5 size (4096 x 2^20): a sample of queries against the oracle, planted matches recovered, and the
    8-shard + merge path identical to the single-database path (emulates the NCCL merge on one GPU)."""
    import torch
    nq, ndb, shards = 4096, 1 << 20, 8
    db, q = synth.descriptors(5, ndb, nq)
    whole = M.DescriptorDatabase(ndb, nq)
    whole.set(db)
    full = whole.top2(q, 50, 0.9, False)
    sample = np.arange(0, nq, 128)
    want = O.hamming_top2(q[sample], db, 50, 0.9, False)
    for fld in ("best1", "best2", "idx", "accepted"):
        assert np.array_equal(full[fld][sample], want[fld]), fld
    planted = full[0::4]
    assert np.all(planted["best1"] <= 60) and planted["accepted"].sum() > 500
    assert np.all(full["best2"] >= full["best1"]) and np.all(full["idx"] >= 0)
    # sharded: per-shard top-2 on contiguous row ranges, then the merge kernel
    rows = ndb // shards
    d_q = torch.from_numpy(q).cuda()
    d_parts = torch.zeros((shards, nq, 4), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ts = torch.cuda.Stream()                             # one explicit stream orders the shard searches and the merge
    st = ts.cuda_stream
    dbs = []
    for s in range(shards):
        sd = M.DescriptorDatabase(rows, nq)
        sd.set(db[s * rows:(s + 1) * rows], index_base=s * rows)
        sd.top2_device(d_q.data_ptr(), nq, 50, 0.9, False, d_parts[s].data_ptr(), st)
        dbs.append(sd)
    d_out = torch.zeros((nq, 4), dtype=torch.int32, device="cuda")
    M.merge_top2_device(0, d_parts.data_ptr(), shards, nq, 50, 0.9, False, d_out.data_ptr(), st)
    torch.cuda.synchronize()
    merged = d_out.cpu().numpy()
    assert np.array_equal(merged[:, 0], full["best1"]) and np.array_equal(merged[:, 1], full["best2"])
    assert np.array_equal(merged[:, 2], full["idx"]) and np.array_equal(merged[:, 3], full["accepted"])
    host = O.merge_top2_host(d_parts.cpu().numpy().view(O.MATCH_DTYPE).reshape(shards, nq), 50, 0.9, False)
    assert np.array_equal(host["idx"], full["idx"]) and np.array_equal(host["accepted"], full["accepted"])


def test_sharded_search_entry_point_single_rank():
    """orbfe_hamming_top2_sharded on one GPU: without a communicator (plain search) and through a real one-rank NCCL communicator
    (dlopen of NCCL, ncclAllGather of the shard's records, merge) — both equal the oracle, index_base included.  The multi-rank
    case is checked by bench.py under torchrun against the torch.distributed path (profiles/r2_bench_2gpu.json)."""
    import ctypes as C
    import torch
    from orb_slam3_detailed_comments_kor_b200 import _lib
    nq, ndb, base = 700, 50000, 123456
    db, q = synth.descriptors(77, ndb, nq)
    want = O.hamming_top2(q, db, 50, 0.9, True)
    want["idx"] += np.where(want["idx"] >= 0, base, 0).astype(np.int32)
    shard = M.DescriptorDatabase(ndb, nq)
    shard.set(db, index_base=base)
    d_q, d_out = torch.from_numpy(q).cuda(), torch.zeros((nq, 4), dtype=torch.int32, device="cuda")
    L = _lib.lib()
    ident, comm = (C.c_uint8 * 128)(), C.c_void_p()
    _lib.check(L.orbfe_nccl_unique_id(ident))
    _lib.check(L.orbfe_nccl_comm_create(0, 1, 0, ident, C.byref(comm)))
    for handle in (None, comm.value):
        d_out.zero_()
        torch.cuda.synchronize()
        shard.top2_sharded(handle, 1, d_q.data_ptr(), nq, 50, 0.9, True, d_out.data_ptr())
        torch.cuda.synchronize()
        got = d_out.cpu().numpy().view(O.MATCH_DTYPE).reshape(nq)
        for fld in ("best1", "best2", "idx", "accepted"):
            assert np.array_equal(got[fld], want[fld]), (fld, handle)
    _lib.check(L.orbfe_nccl_comm_destroy(comm))
    with pytest.raises(OrbfeError):
        shard.top2_sharded(None, 2, d_q.data_ptr(), nq, 50, 0.9, True, d_out.data_ptr())   # two ranks need a communicator
    shard.close()


@pytest.mark.parametrize("name", ["tumvi", "noise"])
def test_knn2_matches_oracle_and_golden(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    W, H = [int(v) for v in g["wh"]]
    L, R = (g["imgL"], g["imgR"]) if "imgL" in g else synth.stereo_pair(int(g["seed"][0]), W, H)
    p = g["params"]
    params = (int(p[0]), float(p[1]), int(p[2]), int(p[3]), int(p[4]))
    li = max(i for i in range(4) if "lap%d" % i in g)
    lap = tuple(int(v) for v in g["lap%d" % li])
    exL = ORBextractor(*params, width=W, height=H); exR = ORBextractor(*params, width=W, height=H)
    monoL, kL, dL = exL(L, lap); monoR, kR, dR = exR(R, lap)
    m, d0, d1 = M.knn2_ratio(dL[monoL:], dR[monoR:], 0.7)
    om, od0, od1 = O.knn2_ratio(dL[monoL:], dR[monoR:], 0.7)
    assert np.array_equal(m, om) and np.array_equal(d0, od0) and np.array_equal(d1, od1)
    assert np.array_equal(m, g["knn_match"]) and np.array_equal(np.stack([d0, d1], 1), g["knn_d01"])
    exL.close(); exR.close()


def test_fisheye_match_batch_on_device():
    """orbfe_fisheye_match_batch_device: ComputeStereoFishEyeMatches' knn-2 + 0.7 ratio step (src/Frame.cc:1119-1144) for a batch of
    TUM-VI-shaped pairs straight from the extractors' device buffers, against the oracle's knn-2 on the downloaded descriptors.
    One lapping interval puts every keypoint in the stereo part (mono = 0), the other splits the frame; one pair has a blank
    right image (no train rows), one a blank left image (no queries)."""
    import torch
    W, H, params, nb = 512, 512, (1500, 1.2, 8, 20, 7), 5
    pairs = [synth.stereo_pair(700 + i, W, H) for i in range(nb)]
    Ls = np.stack([p[0] for p in pairs]); Rs = np.stack([p[1] for p in pairs])
    Rs[3] = 90; Ls[4] = 17
    exL = ORBextractor(*params, width=W, height=H, max_batch=nb)
    exR = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cap = exL.capacity
    grand = []
    for lap in [(0, 511), (150, 400)]:
        cl, ml, kl, dl = [a.copy() for a in exL.extract_batch_host(Ls, lap)]
        cr, mr, kr, dr = [a.copy() for a in exR.extract_batch_host(Rs, lap)]
        d_match = torch.full((nb, cap), -7, dtype=torch.int32, device="cuda")
        d_dist = torch.zeros((nb, cap, 2), dtype=torch.int32, device="cuda")
        M.fisheye_match_batch_device(exL, exR, nb, 0.7, d_match.data_ptr(), d_dist.data_ptr())
        torch.cuda.synchronize()
        got, gd = d_match.cpu().numpy(), d_dist.cpu().numpy()
        total = 0
        for f in range(nb):
            q, t = dl[f, ml[f]:cl[f]], dr[f, mr[f]:cr[f]]
            want = np.full(cl[f], -1, np.int32)
            if len(q):
                m, d0, d1 = O.knn2_ratio(q, t, 0.7) if len(t) else (np.full(len(q), -1, np.int32), None, None)
                want[ml[f]:] = np.where(m >= 0, m + mr[f], -1)
                if len(t) >= 2:
                    assert np.array_equal(gd[f, ml[f]:cl[f], 0], d0) and np.array_equal(gd[f, ml[f]:cl[f], 1], d1), "distances frame %d" % f
            assert np.array_equal(got[f, :cl[f]], want), "frame %d lap %s" % (f, lap)
            total += int((want >= 0).sum())
        grand.append(total)
        assert cr[3] == 0 and cl[4] == 0 and (got[3, :cl[3]] == -1).all()
    assert grand[0] > 20 and grand[1] > 0, grand       # knn-2 over whole frames with the 0.7 ratio keeps few, but not no, pairs
    exL.close(); exR.close()


def test_knn2_degenerate_train_sizes():
    q = np.random.RandomState(1).randint(0, 256, (5, 32)).astype(np.uint8)
    m, d0, d1 = M.knn2_ratio(q, q[:1], 0.7)            # one train row: knnMatch returns <2 -> rejected (Frame.cc:1144)
    om, od0, od1 = O.knn2_ratio(q, q[:1], 0.7)
    assert np.array_equal(m, om) and np.all(m == -1) and np.array_equal(d0, od0) and np.array_equal(d1, od1)
