"""Input conditioning: cv::remap rectification and colour-to-gray.  CPU: the oracle against cv2's own
outputs (tests/golden/conditioning.npz); GPU: the kernels against the oracle and the goldens, bit-exact."""
import os

import numpy as np
import pytest

from oracle import orb_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "conditioning.npz")


def test_oracle_remap_and_gray_match_cv2():
    g = np.load(GOLD)
    assert np.array_equal(O.remap_linear(g["src"], g["mapx"], g["mapy"]), g["remap_out"])
    assert (g["remap_out"] == 0).sum() > 100          # the map leaves the image in places: constant-0 border exercised
    col = g["color"]
    assert np.array_equal(O.to_gray(np.ascontiguousarray(col[..., :3]), True), g["gray_rgb"])
    assert np.array_equal(O.to_gray(np.ascontiguousarray(col[..., :3]), False), g["gray_bgr"])
    assert np.array_equal(O.to_gray(col, True), g["gray_rgba"])
    assert np.array_equal(O.to_gray(col, False), g["gray_bgra"])


@pytest.mark.gpu
def test_gpu_remap_matches_oracle_and_cv2():
    import torch
    from orb_slam3_detailed_comments_kor_b200 import synth
    from orb_slam3_detailed_comments_kor_b200.conditioning import Rectifier
    g = np.load(GOLD)
    r = Rectifier(g["mapx"], g["mapy"])
    assert np.array_equal(r(g["src"]), g["remap_out"])
    r.close()
    # EuRoC-sized frames, a batch, device to device, odd width
    for (W, H, nb) in [(752, 480, 5), (641, 481, 2)]:
        yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
        xn, yn = (xx - W / 2) / 458.0, (yy - H / 2) / 457.0
        r2 = xn * xn + yn * yn
        fac = (1 - 0.283 * r2 + 0.074 * r2 * r2).astype(np.float32)      # EuRoC-like radial distortion (EuRoC.yaml k1, k2)
        mx, my = (xn * fac * 458.0 + W / 2 + 3.1).astype(np.float32), (yn * fac * 457.0 + H / 2 - 2.2).astype(np.float32)
        imgs = np.stack([synth.frame(700 + i, W, H) for i in range(nb)])
        rect = Rectifier(mx, my)
        d_src = torch.from_numpy(imgs).cuda()
        d_dst = torch.zeros_like(d_src)
        st = torch.cuda.Stream()
        torch.cuda.synchronize()
        rect.batch_device(d_src.data_ptr(), d_dst.data_ptr(), nb, st.cuda_stream)
        st.synchronize()
        out = d_dst.cpu().numpy()
        for f in range(nb):
            assert np.array_equal(out[f], O.remap_linear(imgs[f], mx, my)), "frame %d" % f
        assert np.array_equal(rect(imgs[0]), out[0])
        rect.close()


@pytest.mark.gpu
def test_gpu_gray_matches_oracle_and_cv2():
    import torch
    from orb_slam3_detailed_comments_kor_b200.conditioning import to_gray_device
    g = np.load(GOLD)
    col = g["color"]
    h, w = col.shape[:2]
    for ch, rgb, key in [(3, True, "gray_rgb"), (3, False, "gray_bgr"), (4, True, "gray_rgba"), (4, False, "gray_bgra")]:
        src = np.ascontiguousarray(col[..., :ch])
        batch = np.stack([src, src[::-1].copy()])
        d_src = torch.from_numpy(batch).cuda()
        d_dst = torch.zeros((2, h, w), dtype=torch.uint8, device="cuda")
        st = torch.cuda.Stream()
        torch.cuda.synchronize()
        to_gray_device(d_src.data_ptr(), ch, rgb, w, h, 2, d_dst.data_ptr(), stream=st.cuda_stream)
        st.synchronize()
        out = d_dst.cpu().numpy()
        assert np.array_equal(out[0], g[key]), key
        assert np.array_equal(out[1], O.to_gray(batch[1], rgb)), key
