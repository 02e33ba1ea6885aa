"""Input conditioning: cv::remap rectification, colour-to-gray and CLAHE.  CPU: the oracle against cv2's own
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


def _sha(a):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _clahe_cases(g):
    """(name, source image, clip, tiles(across, down), cv2 output or its sha)"""
    from orb_slam3_detailed_comments_kor_b200 import synth
    L, _ = synth.stereo_pair(1, 512, 512)
    dark = (L // 3 + 20).astype(np.uint8)
    assert _sha(dark) == str(g["clahe_tumvi_src_sha"][0]), "synthetic generator drifted"
    E, _ = synth.stereo_pair(0, 752, 480)
    cases = [("tumvi", dark, 3.0, (8, 8), str(g["clahe_tumvi_sha"][0])), ("euroc", E, 3.0, (8, 8), str(g["clahe_euroc_sha"][0]))]
    for tag in "abc":
        cases.append((tag, g["clahe_src_" + tag], 3.0, (8, 8), g["clahe_out_" + tag]))
    cases.append(("flat", np.full((64, 64), 77, np.uint8), 3.0, (8, 8), g["clahe_flat_out"]))
    cases.append(("d", g["clahe_src_a"], 40.0, (4, 6), g["clahe_out_d"]))
    return cases


def _same(out, want):
    return _sha(out) == want if isinstance(want, str) else np.array_equal(out, want)


def test_oracle_clahe_matches_cv2():
    g = np.load(GOLD)
    for name, src, clip, tiles, want in _clahe_cases(g):
        assert _same(O.clahe(src, clip, tiles), want), name
    with pytest.raises(ValueError):
        O.clahe(np.zeros((3, 100), np.uint8))           # 3 rows cannot be reflected out to 8 tile rows


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
        # a pixel count that is not a multiple of four (the word kernel plus the byte kernel for the rest), and pointers off the word grid
        odd = np.ascontiguousarray(src[:7, :9])
        flat = torch.zeros(odd.size + 8, dtype=torch.uint8, device="cuda")
        o_dst = torch.zeros(7 * 9 + 8, dtype=torch.uint8, device="cuda")
        for shift_src, shift_dst in ((0, 0), (1, 0), (0, 2), (3, 1)):
            flat[shift_src:shift_src + odd.size] = torch.from_numpy(odd.reshape(-1)).cuda()
            o_dst.zero_()
            torch.cuda.synchronize()
            to_gray_device(flat.data_ptr() + shift_src, ch, rgb, 9, 7, 1, o_dst.data_ptr() + shift_dst, stream=st.cuda_stream)
            st.synchronize()
            got = o_dst.cpu().numpy()
            assert np.array_equal(got[shift_dst:shift_dst + 63].reshape(7, 9), O.to_gray(odd, rgb)), (key, shift_src, shift_dst)
            assert not got[:shift_dst].any() and not got[shift_dst + 63:].any()


@pytest.mark.gpu
def test_gpu_clahe_matches_oracle_and_cv2():
    import torch
    from orb_slam3_detailed_comments_kor_b200 import synth
    from orb_slam3_detailed_comments_kor_b200._lib import OrbfeError
    from orb_slam3_detailed_comments_kor_b200.conditioning import CLAHE
    g = np.load(GOLD)
    for name, src, clip, tiles, want in _clahe_cases(g):
        c = CLAHE(src.shape[1], src.shape[0], clip, tiles)
        out = c.apply(src)
        ref, lut = O.clahe(src, clip, tiles, return_lut=True)
        assert np.array_equal(c.lut(0), lut), name + ": lookup tables"
        assert np.array_equal(out, ref) and _same(out, want), name
        c.close()
    # shapes that reach the banded kernel with padding in both directions, the most tiles across it takes, and one more (the general kernel)
    rng = np.random.RandomState(3)
    for (h, w), clip, tiles in (((470, 644), 3.0, (8, 8)), ((480, 752), 2.0, (16, 4)), ((331, 680), 4.0, (17, 10)), ((97, 256), 3.0, (2, 3))):
        src = np.clip(rng.normal(110, 50, (h, w)) + 40 * np.sin(np.arange(w) / 37.0)[None, :] + 30 * np.cos(np.arange(h) / 23.0)[:, None], 0, 255).astype(np.uint8)
        c = CLAHE(w, h, clip, tiles)
        ref, lut = O.clahe(src, clip, tiles, return_lut=True)
        assert np.array_equal(c.apply(src), ref), ((h, w), tiles)
        assert np.array_equal(c.lut(0), lut), ((h, w), tiles, "lookup tables")
        c.close()
    # a TUM-VI sized batch, device to device, in place, frames at different brightness
    W = H = 512
    frames = []
    for f in range(5):
        L, R = synth.stereo_pair(40 + f, W, H)
        frames += [(L // (f + 1)).astype(np.uint8), R]
    batch = np.stack(frames)
    c = CLAHE(W, H, 3.0, (8, 8), max_batch=len(frames))
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        d = torch.from_numpy(batch).cuda()
        c.batch_device(d.data_ptr(), d.data_ptr(), len(frames), stream=st.cuda_stream)
        out = d.cpu().numpy()
    st.synchronize()
    for f in range(len(frames)):
        assert np.array_equal(out[f], O.clahe(batch[f])), "frame %d" % f
    with pytest.raises(OrbfeError):
        c.batch_device(d.data_ptr(), d.data_ptr(), len(frames) + 1, stream=st.cuda_stream)
    c.close()
    with pytest.raises(OrbfeError):
        CLAHE(100, 3)
