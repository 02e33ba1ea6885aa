"""ORBmatcher::SearchForInitialization (/root/reference/src/ORBmatcher.cc:706-821; caller Tracking::MonocularInitialization,
src/Tracking.cc:2462): the oracle against the reference's own body compiled into oracle/_ref (CPU), CUDA against the oracle (-m gpu).
Integers only: everything is compared exactly (vnMatches12, nmatches, vbPrevMatched)."""
import numpy as np
import pytest

from oracle import orb_oracle as O
from oracle import ref_oracle as R

KP = O.KP_DTYPE


def grid_for(W, H):
    """Frame::mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv of an undistorted W x H image (src/Frame.cc:1006-1018)."""
    return np.array([0.0, 0.0, 64.0 / W, 48.0 / H], np.float32)


def extracted_case(seed, W=640, H=480, nfeatures=2000, shift=(9.0, -4.0)):
    """Two frames the way the monocular initialiser sees them: the initial extractor's keypoints of an image and of a displaced view."""
    from orb_slam3_detailed_comments_kor_b200 import synth
    a, b = synth.stereo_pair(seed, W, H)
    ex = O.OracleExtractor(nfeatures, 1.2, 8, 20, 7)
    _, k1, d1 = ex(a, (0, 0))
    _, k2, d2 = O.OracleExtractor(nfeatures, 1.2, 8, 20, 7)(b, (0, 0))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    return k1, d1, k2, d2, grid_for(W, H), prev


def random_case(seed, n1, n2, W=752, H=480, clusters=40, noise=12, level0=0.7, far=0.0):
    """Random keypoints; descriptors are noisy copies of a few cluster centres so that many F1 keypoints compete for the same F2
    keypoint (displacements, equal distances, ratio failures); a share of the previous matches lies outside the image."""
    rng = np.random.RandomState(seed)
    centres = rng.randint(0, 256, (clusters, 32)).astype(np.uint8)

    def frame(n):
        k = np.zeros(n, KP)
        k["x"] = rng.uniform(0, W, n).astype(np.float32)
        k["y"] = rng.uniform(0, H, n).astype(np.float32)
        k["octave"] = np.where(rng.rand(n) < level0, 0, rng.randint(1, 8, n))
        k["angle"] = rng.uniform(0, 360, n).astype(np.float32)
        k["size"] = 31.0
        c = rng.randint(0, clusters, n)
        d = centres[c].copy()
        flips = rng.randint(0, noise + 1, n)
        for i in range(n):
            for bit in rng.randint(0, 256, flips[i]):
                d[i, bit >> 3] ^= 1 << (bit & 7)
        return k, d
    k1, d1 = frame(n1)
    k2, d2 = frame(n2)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32) + rng.uniform(-30, 30, (n1, 2)).astype(np.float32)
    out = rng.rand(n1) < far
    prev[out] += np.float32(2000.0)
    return k1, d1, k2, d2, grid_for(W, H), prev


RANDOM_CASES = {
    "dense_conflicts": dict(seed=1, n1=1500, n2=1500, clusters=25, noise=10),
    "few_clusters": dict(seed=2, n1=800, n2=600, clusters=4, noise=6),
    "identical_descriptors": dict(seed=3, n1=300, n2=300, clusters=3, noise=0),
    "sparse": dict(seed=4, n1=200, n2=150, clusters=200, noise=30),
    "outside_windows": dict(seed=5, n1=1000, n2=1000, clusters=30, noise=10, far=0.3),
    "no_level0": dict(seed=6, n1=100, n2=100, level0=0.0),
    "big": dict(seed=7, n1=4096, n2=4096, clusters=60, noise=14, level0=0.9),
}


@pytest.mark.skipif(not R.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("name", sorted(RANDOM_CASES))
@pytest.mark.parametrize("window,ratio,ori", [(100, 0.9, True), (30, 0.9, False), (100, 0.6, True)])
def test_oracle_equals_reference_random(name, window, ratio, ori):
    k1, d1, k2, d2, g, prev = random_case(**RANDOM_CASES[name])
    m, p, nm = O.search_for_initialization(k1, d1, k2, d2, g, prev, window, ratio, ori)
    rm, rp, rnm = R.search_for_initialization(k1, d1, k2, d2, g, prev, window, ratio, ori)
    assert nm == rnm and np.array_equal(m, rm) and np.array_equal(p, rp)
    assert nm == int((m >= 0).sum())
    if name in ("dense_conflicts", "few_clusters", "big"):
        assert nm > 10


@pytest.mark.skipif(not R.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_equals_reference_on_extracted_frames_and_second_call():
    k1, d1, k2, d2, g, prev = extracted_case(31)
    m, p, nm = O.search_for_initialization(k1, d1, k2, d2, g, prev, 100, 0.9, True)
    rm, rp, rnm = R.search_for_initialization(k1, d1, k2, d2, g, prev, 100, 0.9, True)
    assert nm == rnm and np.array_equal(m, rm) and np.array_equal(p, rp)
    assert nm > 10
    # the initialiser calls it again on the next frame with the updated vbPrevMatched (src/Tracking.cc:2462)
    m2, p2, nm2 = O.search_for_initialization(k1, d1, k2, d2, g, p, 100, 0.9, True)
    rm2, rp2, rnm2 = R.search_for_initialization(k1, d1, k2, d2, g, rp, 100, 0.9, True)
    assert nm2 == rnm2 and np.array_equal(m2, rm2) and np.array_equal(p2, rp2)


def test_oracle_hand_cases():
    """Two F1 keypoints want the same F2 keypoint: the later one wins only with a strictly smaller distance, and the earlier one loses its match."""
    g = grid_for(640, 480)
    k1 = np.zeros(2, KP); k2 = np.zeros(2, KP)
    k1["x"] = [100, 102]; k1["y"] = [100, 100]; k2["x"] = [101, 400]; k2["y"] = [100, 300]
    d2 = np.zeros((2, 32), np.uint8); d2[1] = 0xFF
    d1 = np.zeros((2, 32), np.uint8)
    d1[0, 0] = 0b111          # distance 3 to F2[0]
    d1[1, 0] = 0b1            # distance 1: displaces the first
    prev = np.stack([k1["x"], k1["y"]], 1)
    m, p, nm = O.search_for_initialization(k1, d1, k2, d2, g, prev, 100, 0.9, False)
    assert list(m) == [-1, 0] and nm == 1
    assert np.array_equal(p[0], prev[0]) and list(p[1]) == [101, 100]
    d1[1, 0] = 0b111          # equal distance: vMatchedDistance <= dist hides the keypoint, the first keeps it
    m, p, nm = O.search_for_initialization(k1, d1, k2, d2, g, prev, 100, 0.9, False)
    assert list(m) == [0, -1] and nm == 1
    k1["octave"] = [0, 1]     # only level 0 is searched
    d1[1, 0] = 0
    m, p, nm = O.search_for_initialization(k1, d1, k2, d2, g, prev, 100, 0.9, False)
    assert list(m) == [0, -1] and nm == 1


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(RANDOM_CASES))
@pytest.mark.parametrize("window,ratio,ori", [(100, 0.9, True), (30, 0.9, False), (100, 0.6, True)])
def test_gpu_equals_oracle_random(name, window, ratio, ori):
    from orb_slam3_detailed_comments_kor_b200 import projection as P
    k1, d1, k2, d2, g, prev = random_case(**RANDOM_CASES[name])
    m, p, nm = O.search_for_initialization(k1, d1, k2, d2, g, prev, window, ratio, ori)
    gm, gp, gnm = P.search_for_initialization(k1, d1, k2, d2, g, prev, window, ratio, ori)
    assert gnm == nm and np.array_equal(gm, m) and np.array_equal(gp, p)


@pytest.mark.gpu
def test_gpu_batch_of_pairs_on_the_device_and_limits():
    """Several pairs in one launch (extracted frames, random ones, an empty frame, one beyond the row limit), results per pair as the
    oracle's; the second call with the updated vbPrevMatched; rounds reported."""
    import torch
    from orb_slam3_detailed_comments_kor_b200 import projection as P
    from orb_slam3_detailed_comments_kor_b200._lib import OrbfeError
    cases = [extracted_case(31), random_case(**RANDOM_CASES["dense_conflicts"]), random_case(seed=9, n1=0, n2=50), random_case(seed=10, n1=60, n2=0),
             random_case(seed=11, n1=8192, n2=8192, clusters=80, noise=14, level0=0.5), random_case(seed=12, n1=8193, n2=10), extracted_case(32, nfeatures=5000)]
    off1 = np.cumsum([0] + [len(c[0]) for c in cases]).astype(np.int32)
    off2 = np.cumsum([0] + [len(c[2]) for c in cases]).astype(np.int32)
    cat = lambda i, dt: np.ascontiguousarray(np.concatenate([np.asarray(c[i]).reshape(len(c[i]), -1) if dt is None else c[i] for c in cases]))  # noqa: E731
    k1 = np.concatenate([c[0] for c in cases]); d1 = np.concatenate([c[1] for c in cases])
    k2 = np.concatenate([c[2] for c in cases]); d2 = np.concatenate([c[3] for c in cases])
    grids = np.stack([c[4] for c in cases]); prev = np.concatenate([c[5] for c in cases]).astype(np.float32)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()  # noqa: E731
    t_k1, t_d1, t_k2, t_d2, t_o1, t_o2, t_g, t_prev = dev(k1), dev(d1), dev(k2), dev(d2), dev(off1), dev(off2), dev(grids), dev(prev)
    t_m = torch.full((len(k1),), -7, dtype=torch.int32, device="cuda")
    t_nm = torch.zeros(len(cases), dtype=torch.int32, device="cuda")
    t_rounds = torch.zeros(len(cases), dtype=torch.int32, device="cuda")
    for call in range(2):
        P.search_for_initialization_device(len(cases), t_k1.data_ptr(), t_d1.data_ptr(), t_o1.data_ptr(), t_k2.data_ptr(), t_d2.data_ptr(), t_o2.data_ptr(),
                                           t_g.data_ptr(), t_prev.data_ptr(), t_m.data_ptr(), t_nm.data_ptr(), 100, 0.9, True, 50, t_rounds.data_ptr())
        torch.cuda.synchronize()
        gm, gnm, gprev = t_m.cpu().numpy(), t_nm.cpu().numpy(), t_prev.cpu().numpy().view(np.float32).reshape(-1, 2)
        rounds = t_rounds.cpu().numpy()
        for i, c in enumerate(cases):
            a, b = off1[i], off1[i + 1]
            if len(c[0]) > 8192:
                assert gnm[i] == -1
                continue
            m, p, nm = O.search_for_initialization(c[0], c[1], c[2], c[3], c[4], prev[a:b], 100, 0.9, True)
            assert gnm[i] == nm, "call %d pair %d" % (call, i)
            assert np.array_equal(gm[a:b], m) and np.array_equal(gprev[a:b], p), "call %d pair %d" % (call, i)
            prev[a:b] = p
        assert rounds[0] >= 1 and rounds[4] >= 2
        print("rounds per pair:", rounds.tolist(), "matches:", gnm.tolist())
    with pytest.raises(OrbfeError):
        c = cases[5]
        P.search_for_initialization(c[0], c[1], c[2], c[3], c[4], c[5])
