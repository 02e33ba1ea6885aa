"""CPU tests of the boundary: the C-ABI library loads and exports every symbol include/orbfe.h declares, and
fails loudly (no CPU fallback) when no CUDA device is usable."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from orb_slam3_detailed_comments_kor_b200 import _lib
from oracle.orb_oracle import merge_top2_host

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "orbfe.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(orbfe_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    syms = declared_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(L, s), "liborbfe.so does not export " + s
        assert s in _lib.SIGNATURES, "python binding lacks " + s


def test_no_cpu_fallback_without_gpu():
    L = _lib.lib()
    if L.orbfe_device_count() > 0:
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    p = _lib.OrbfeParams(1200, 1.2, 8, 20, 7)
    rc = L.orbfe_extractor_create(C.byref(p), 0, 752, 480, 1, C.byref(h))
    assert rc == -3 and not h.value and b"no CPU path" in L.orbfe_last_error()
    rc = L.orbfe_matcher_create(0, 16, 16, C.byref(h))
    assert rc == -3


def test_keypoint_layout_is_cv_keypoint():
    assert _lib.KP_DTYPE.itemsize == 28 and _lib.MATCH_DTYPE.itemsize == 16


def test_product_package_never_touches_the_oracle():
    import glob
    pkg = os.path.join(ROOT, "orb_slam3_detailed_comments_kor_b200")
    for path in glob.glob(os.path.join(pkg, "**", "*"), recursive=True):
        if path.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
            text = open(path, errors="ignore").read()
            assert "orb_oracle" not in text and "liborb_oracle" not in text and "from oracle" not in text, path


def test_host_merge_rule():
    parts = np.zeros((3, 2), _lib.MATCH_DTYPE)
    parts[0] = [(10, 30, 5, 0), (256, 256, -1, 0)]
    parts[1] = [(10, 12, 105, 0), (40, 45, 130, 0)]
    parts[2] = [(50, 60, 205, 0), (41, 256, 230, 0)]
    out = merge_top2_host(parts, 50, 0.9, False)
    assert tuple(out[0]) == (10, 10, 5, 0)      # tie on best1 -> lowest shard keeps idx; second = the other 10
    assert tuple(out[1]) == (40, 41, 130, 0)    # 40 < 0.9*41 = 36.9 is false


def test_host_libm_reproduces_the_trig_exception_table():
    """The 96 rotation angles whose cosf / sinf the device takes from a table: on this host (the image's glibc) the table must be the
    libm's own values; elsewhere the count says how many angles can differ from that host's CPU build."""
    import ctypes as C
    L = _lib.lib()
    n = C.c_int(0)
    assert L.orbfe_check_host_libm(C.byref(n)) == 0
    assert n.value == 96


def test_frames_entry_points_refuse_malformed_feature_vectors_before_touching_a_device():
    """orbfe_search_by_bow_frames / orbfe_search_for_triangulation_frames index device arrays with what the caller's FeatureVector says: rows
    outside the frame, node ids out of order, offsets that do not start at 0 and pairs outside the frame list are refused on the host."""
    import pytest
    from orb_slam3_detailed_comments_kor_b200 import bow as B
    from orb_slam3_detailed_comments_kor_b200._lib import OrbfeError
    rng = np.random.RandomState(1)
    desc = rng.randint(0, 256, (6, 32)).astype(np.uint8)
    kps = np.zeros(6, _lib.KP_DTYPE)
    good = {3: np.array([0, 1], np.int32), 7: np.array([2, 5], np.int32)}
    for bad in ({3: np.array([0, 6], np.int32)},                                   # a row beyond the frame
                {3: np.array([0, -1], np.int32)},
                {7: np.array([0], np.int32), 3: np.array([1], np.int32)}):         # node ids not ascending
        with pytest.raises(OrbfeError, match="malformed"):
            B.search_by_bow_frames([dict(desc=desc, kps=kps, feat_vec=good), dict(desc=desc, kps=kps, feat_vec=bad)], [0], [1])
    with pytest.raises(OrbfeError, match="outside the list"):
        B.search_by_bow_frames([dict(desc=desc, kps=kps, feat_vec=good)], [0], [1])
    with pytest.raises(OrbfeError, match="malformed"):
        B.search_for_triangulation_frames([dict(desc=desc, kps=kps, feat_vec=good), dict(desc=desc, kps=kps, feat_vec={3: np.array([9], np.int32)})], [0], [1],
                                          np.zeros((1, 9), np.float32), np.zeros((1, 2), np.float32), np.ones(8, np.float32), np.ones(8, np.float32))
