"""KannalaBrandt8::TriangulateMatches (/root/reference/src/CameraModels/KannalaBrandt8.cpp:338-407, 499-513; called per knn-2
survivor at src/Frame.cc:1149): oracle against the cv2-4.13-backed goldens (CPU), CUDA against the oracle (-m gpu).

Tolerances, and why they are not zero everywhere:
  * the 4x4 SVD (cv::SVD::compute = OpenCV's Jacobi) is BIT-EXACT between the oracle and cv2.SVDecomp, and so are the products
    (cv2.gemm) and the rows of the system (cv2.addWeighted);
  * tanf / atan2f / cosf / sinf come from three different math libraries (glibc in the oracle, numpy's double routines rounded to
    float in the golden script, CUDA's on the device), each within 1-2 ulp; their differences pass through the SVD, so depth and
    point are compared at REL_TOL = 1e-4 (measured: <= 3e-6), and the accept / reject decision must be identical except for pairs
    whose reprojection error or parallax lies within 1e-3 relative of its threshold (none in the goldens).
"""
import os

import numpy as np
import pytest

from oracle import orb_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "triangulate.npz")
REL_TOL = 1e-4


def load():
    return np.load(GOLD)


def test_svd_4x4_is_bit_exact_with_cv2():
    g = load()
    for A, w, vt in zip(g["svd_A"], g["svd_w"], g["svd_vt"]):
        ow, ovt = O.svd4x4_vt(A)
        assert np.array_equal(ow, w) and np.array_equal(ovt, vt)
    assert len(g["svd_A"]) > 600


def test_unproject_and_project_match_restatement():
    g = load()
    ray = np.array([O.kb8_unproject(g["cam1"], *u) for u in g["unproject_uv"]])
    assert np.max(np.abs(ray - g["unproject_ray"])) <= 5e-7                      # one ulp of a value below 4
    uv = np.array([O.kb8_project(g["cam2"], p) for p in g["project_p"]])
    assert np.max(np.abs(uv - g["project_uv"])) <= 2e-4                          # a few ulp of a pixel coordinate around 500
    c, done = g["cam1"], 0                                                       # project(unproject(p)) returns p inside the field of view
    for u in g["unproject_uv"]:
        if np.hypot(u[0] - c[2], u[1] - c[3]) < 250:                                # beyond theta_d = pi/2 unproject clamps (KannalaBrandt8.cpp:100)
            back = O.kb8_project(c, O.kb8_unproject(c, *u))
            assert np.max(np.abs(back - u)) < 2e-3
            done += 1
    assert done > 50


def test_triangulate_matches_oracle_vs_cv2_goldens():
    g = load()
    depth, p3d = O.kb8_triangulate_matches(g["cam1"], g["cam2"], g["R12"], g["t12"], g["kp1"], g["kp2"], g["sigma1"], g["sigma2"])
    acc, want = depth > 0, g["depth"] > 0
    assert np.array_equal(acc, want)
    assert 100 < int(acc.sum()) < len(acc) - 100                                   # both outcomes are exercised
    assert np.all(depth[~acc] == -1) and np.all(p3d[~acc] == 0)
    assert np.max(np.abs(depth[acc] - g["depth"][acc]) / g["depth"][acc]) <= REL_TOL
    assert np.max(np.abs(p3d[acc] - g["p3d"][acc]) / np.abs(g["p3d"][acc]).max(1, keepdims=True)) <= REL_TOL
    assert np.mean(depth[acc] == g["depth"][acc]) > 0.8                            # in practice most are identical


def test_triangulation_recovers_planted_points():
    """Independent of any restatement: noiseless projections of known points must triangulate back to them."""
    g = load()
    rng = np.random.RandomState(4)
    R12, t12 = g["R12"].astype(np.float64), g["t12"].astype(np.float64)
    R21, t21 = R12.T, -R12.T @ t12
    # within 3 m of a 10 cm baseline, so that the parallax test (cos < 0.9998, i.e. more than 1.15 degrees) lets the points through
    X1 = np.stack([rng.uniform(-1, 1, 300), rng.uniform(-1, 1, 300), rng.uniform(0.5, 3.0, 300)], 1)
    X2 = X1 @ R21.T + t21
    uv1 = np.array([O.kb8_project(g["cam1"], p) for p in X1.astype(np.float32)])
    uv2 = np.array([O.kb8_project(g["cam2"], p) for p in X2.astype(np.float32)])
    depth, p3d = O.kb8_triangulate_matches(g["cam1"], g["cam2"], g["R12"], g["t12"], uv1, uv2, np.ones(300, np.float32), np.ones(300, np.float32))
    assert (depth > 0).mean() > 0.95
    ok = depth > 0
    assert np.median(np.abs(p3d[ok] - X1[ok]).max(1) / X1[ok, 2]) < 2e-3          # float pixels: a few mm per metre of depth


@pytest.mark.gpu
def test_gpu_triangulate_pairs_matches_oracle():
    import torch
    from orb_slam3_detailed_comments_kor_b200 import matcher as M
    g = load()
    rig = M.make_fisheye_rig(g["cam1"], g["cam2"], g["R12"], g["t12"])
    n = len(g["kp1"])
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda()  # noqa: E731
    uv1, uv2, s1, s2 = d(g["kp1"]), d(g["kp2"]), d(g["sigma1"]), d(g["sigma2"])
    depth, p3d = torch.zeros(n, device="cuda"), torch.zeros((n, 3), device="cuda")
    M.kb8_triangulate_pairs_device(rig, uv1.data_ptr(), uv2.data_ptr(), s1.data_ptr(), s2.data_ptr(), n, depth.data_ptr(), p3d.data_ptr())
    torch.cuda.synchronize()
    got_d, got_p = depth.cpu().numpy(), p3d.cpu().numpy()
    want_d, want_p = O.kb8_triangulate_matches(g["cam1"], g["cam2"], g["R12"], g["t12"], g["kp1"], g["kp2"], g["sigma1"], g["sigma2"])
    acc = want_d > 0
    assert np.array_equal(got_d > 0, acc)
    assert np.all(got_d[~acc] == -1) and np.all(got_p[~acc] == 0)
    assert np.max(np.abs(got_d[acc] - want_d[acc]) / want_d[acc]) <= REL_TOL
    assert np.max(np.abs(got_p[acc] - want_p[acc]) / np.abs(want_p[acc]).max(1, keepdims=True)) <= REL_TOL


@pytest.mark.gpu
def test_gpu_fisheye_stereo_frame_on_device():
    """Frame::ComputeStereoFishEyeMatches end to end on the device for a batch of 512x512 pairs: extraction with the lapping
    interval, knn-2 + ratio, TriangulateMatches, depth test; mvLeftToRightMatch / mvRightToLeftMatch / mvDepth / mvStereo3Dpoints
    against the oracle driven the way the reference's loop runs (src/Frame.cc:1119-1157)."""
    import torch
    from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth
    from orb_slam3_detailed_comments_kor_b200 import matcher as M
    W, H, params, nb = 512, 512, (1500, 1.2, 8, 20, 7), 4
    pairs = [synth.stereo_pair(800 + i, W, H) for i in range(nb)]
    Ls, Rs = np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])
    # a rig that agrees with the synthetic generator (the right image is the left one shifted along x): identity rotation, baseline along x
    cam = np.array([190.0, 190.0, 256.0, 256.0, 0.0, 0.0, 0.0, 0.0], np.float32)
    R12, t12 = np.eye(3, dtype=np.float32), np.array([0.1, 0.0, 0.0], np.float32)
    exL = ORBextractor(*params, width=W, height=H, max_batch=nb)
    exR = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cap = exL.capacity
    cl, ml, kl, dl = [a.copy() for a in exL.extract_batch_host(Ls, (0, 511))]
    cr, mr, kr, dr = [a.copy() for a in exR.extract_batch_host(Rs, (0, 511))]
    d_lr = torch.empty((nb, cap), dtype=torch.int32, device="cuda")
    d_rl = torch.empty((nb, cap), dtype=torch.int32, device="cuda")
    d_depth, d_p3d = torch.empty((nb, cap), device="cuda"), torch.empty((nb, cap, 3), device="cuda")
    M.fisheye_match_batch_device(exL, exR, nb, 0.7, d_lr.data_ptr())
    torch.cuda.synchronize()                                                       # the kernel ran on the extractor's own stream
    knn = d_lr.cpu().numpy().copy()
    M.fisheye_triangulate_batch_device(exL, exR, nb, cam, cam, R12, t12, d_lr.data_ptr(), d_rl.data_ptr(), d_depth.data_ptr(), d_p3d.data_ptr())
    torch.cuda.synchronize()
    lr, rl, depth, p3d = d_lr.cpu().numpy(), d_rl.cpu().numpy(), d_depth.cpu().numpy(), d_p3d.cpu().numpy()
    sig = np.array(O.OracleExtractor(*params).scale_factors(), np.float32) ** 2
    kept = 0
    for f in range(nb):
        q = np.nonzero(knn[f, :cl[f]] >= 0)[0]
        t = knn[f, q]
        uv1 = np.stack([kl[f]["x"][q], kl[f]["y"][q]], 1)
        uv2 = np.stack([kr[f]["x"][t], kr[f]["y"][t]], 1)
        wd, wp = O.kb8_triangulate_matches(cam, cam, R12, t12, uv1, uv2, sig[kl[f]["octave"][q]], sig[kr[f]["octave"][t]])
        want_lr, want_rl = np.full(cl[f], -1, np.int32), np.full(cr[f], -1, np.int32)
        want_depth = np.full(cl[f], -1, np.float32)
        for j in range(len(q)):                                                    # the reference's loop order: the last writer keeps a right keypoint
            if wd[j] > 0.0001:
                want_lr[q[j]] = t[j]; want_rl[t[j]] = q[j]; want_depth[q[j]] = wd[j]
        assert np.array_equal(lr[f, :cl[f]], want_lr) and np.array_equal(rl[f, :cr[f]], want_rl), "frame %d" % f
        ok = want_depth > 0
        assert np.array_equal(depth[f, :cl[f]] > 0, ok)
        assert np.all(depth[f, :cl[f]][~ok] == -1)
        if ok.any():
            assert np.max(np.abs(depth[f, :cl[f]][ok] - want_depth[ok]) / want_depth[ok]) <= REL_TOL
            wp_full = np.zeros((cl[f], 3), np.float32); wp_full[q] = wp
            assert np.max(np.abs(p3d[f, :cl[f]][ok] - wp_full[ok]) / np.abs(wp_full[ok]).max(1, keepdims=True)) <= REL_TOL
        kept += int(ok.sum())
    assert kept > 10, kept
    # the host-output call for the same frames (what the C++ shim's ComputeStereoFishEyeMatches uses)
    h_lr, h_rl, h_d, h_p = M.fisheye_stereo_frames(exL, exR, nb, M.make_fisheye_rig(cam, cam, R12, t12), 0.7)
    for f in range(nb):
        assert np.array_equal(h_lr[f, :cl[f]], lr[f, :cl[f]]) and np.array_equal(h_rl[f, :cr[f]], rl[f, :cr[f]])
        ok = lr[f, :cl[f]] >= 0
        assert np.array_equal(h_d[f, :cl[f]][ok], depth[f, :cl[f]][ok]) and np.array_equal(h_p[f, :cl[f]][ok], p3d[f, :cl[f]][ok])
    exL.close(); exR.close()


@pytest.mark.gpu
def test_gpu_fisheye_frames_submit_wait_equals_the_device_calls():
    """orbfe_fisheye_frames_submit / orbfe_stereo_frames_wait (host images in, everything of the rig constructor out, two batches in
    flight, chunked) against the per-step device calls on the same pairs."""
    import torch
    from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth
    from orb_slam3_detailed_comments_kor_b200 import matcher as M
    from orb_slam3_detailed_comments_kor_b200._lib import KP_DTYPE
    W, H, params, nb = 512, 512, (1500, 1.2, 8, 20, 7), 70                       # two chunks, one of them short
    cam = np.array([190.0, 190.0, 256.0, 256.0, 0.0, 0.0, 0.0, 0.0], np.float32)
    rig = M.make_fisheye_rig(cam, cam, np.eye(3, dtype=np.float32), np.array([0.1, 0.0, 0.0], np.float32))
    exL = ORBextractor(*params, width=W, height=H, max_batch=nb)
    exR = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cap = exL.capacity
    batches, want = [], []
    for bi in range(3):
        pairs = [synth.stereo_pair(7000 + 100 * bi + i, W, H) for i in range(nb)]
        Ls = torch.from_numpy(np.stack([p[0] for p in pairs])).pin_memory().numpy(); Rs = torch.from_numpy(np.stack([p[1] for p in pairs])).pin_memory().numpy()
        batches.append((Ls, Rs))
        cl, ml, kl, dl = [a.copy() for a in exL.extract_batch_host(Ls, (0, 511))]
        cr, mr, kr, dr = [a.copy() for a in exR.extract_batch_host(Rs, (0, 511))]
        d_lr = torch.empty((nb, cap), dtype=torch.int32, device="cuda"); d_rl = torch.empty((nb, cap), dtype=torch.int32, device="cuda")
        d_depth, d_p3d = torch.empty((nb, cap), device="cuda"), torch.empty((nb, cap, 3), device="cuda")
        M.fisheye_match_batch_device(exL, exR, nb, 0.7, d_lr.data_ptr())
        M.fisheye_triangulate_batch_device(exL, exR, nb, cam, cam, np.eye(3, dtype=np.float32), np.array([0.1, 0.0, 0.0], np.float32),
                                           d_lr.data_ptr(), d_rl.data_ptr(), d_depth.data_ptr(), d_p3d.data_ptr())
        torch.cuda.synchronize()
        want.append((cl, ml, kl, dl, cr, mr, kr, dr, d_lr.cpu().numpy(), d_rl.cpu().numpy(), d_depth.cpu().numpy(), d_p3d.cpu().numpy()))

    def out_set():
        side = lambda: (np.zeros(nb, np.int32), np.zeros(nb, np.int32), np.zeros((nb, cap), KP_DTYPE), np.zeros((nb, cap, 32), np.uint8))  # noqa: E731
        return side() + side() + (np.zeros((nb, cap), np.int32), np.zeros((nb, cap), np.int32), np.zeros((nb, cap), np.float32), np.zeros((nb, cap, 3), np.float32))
    outs = [out_set(), out_set()]
    got = []
    M.fisheye_frames_submit(exL, exR, batches[0][0], batches[0][1], rig, 0.7, (0, 511), outs[0])
    for k in range(1, 3):
        M.fisheye_frames_submit(exL, exR, batches[k][0], batches[k][1], rig, 0.7, (0, 511), outs[k & 1])
        got.append([a.copy() for a in M.fisheye_frames_wait(exL, exR, outs[(k - 1) & 1])])
    got.append([a.copy() for a in M.fisheye_frames_wait(exL, exR, outs[0])])
    kept = 0
    for bi in range(3):
        g, w = got[bi], want[bi]
        for j in (0, 1, 4, 5):
            assert np.array_equal(g[j], w[j]), "batch %d counts" % bi
        for f in range(nb):
            nl, nr = w[0][f], w[4][f]
            assert np.array_equal(g[2][f, :nl], w[2][f, :nl]) and np.array_equal(g[3][f, :nl], w[3][f, :nl])
            assert np.array_equal(g[6][f, :nr], w[6][f, :nr]) and np.array_equal(g[7][f, :nr], w[7][f, :nr])
            assert np.array_equal(g[8][f, :nl], w[8][f, :nl]) and np.array_equal(g[9][f, :nr], w[9][f, :nr]), "batch %d frame %d matches" % (bi, f)
            ok = w[8][f, :nl] >= 0
            assert np.array_equal(g[10][f, :nl][ok], w[10][f, :nl][ok]) and np.array_equal(g[11][f, :nl][ok], w[11][f, :nl][ok])
            kept += int(ok.sum())
    assert kept > 50
    exL.close(); exR.close()
