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
