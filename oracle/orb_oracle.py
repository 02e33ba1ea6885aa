"""ctypes loader for the CPU oracle (oracle/liborb_oracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this
module; the product package never does.  See oracle/orb_oracle.h for what each function restates.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liborb_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
MATCH_DTYPE = np.dtype([("best1", "<i4"), ("best2", "<i4"), ("idx", "<i4"), ("accepted", "<i4")])


def build(force=False):
    """Compile the oracle with its Makefile (gcc only; no reference sources are read)."""
    srcs = [os.path.join(_HERE, f) for f in ("orb_oracle.cpp", "match_oracle.cpp", "condition_oracle.cpp", "project_oracle.cpp", "bow_oracle.cpp",
                                             "triangulate_oracle.cpp", "orb_oracle.h", "Makefile")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "liborb_oracle.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(_SO)
    u8p, i32p, f32p, vp = C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_float), C.c_void_p
    L.orc_extractor_create.restype = vp
    L.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    L.orc_extractor_destroy.argtypes = [vp]
    L.orc_extract.restype = C.c_int
    L.orc_extract.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, i32p]
    L.orc_num_levels.argtypes = [vp]
    L.orc_scale_factor.restype = C.c_float
    L.orc_scale_factor.argtypes = [vp, C.c_int]
    L.orc_features_per_level.argtypes = [vp, C.c_int]
    L.orc_umax.argtypes = [vp, C.c_int]
    L.orc_level_size.argtypes = [vp, C.c_int, i32p, i32p]
    L.orc_pyramid_level.argtypes = [vp, C.c_int, C.c_int, vp]
    L.orc_blurred_level.argtypes = [vp, C.c_int, vp]
    L.orc_fast_candidates.argtypes = [vp, C.c_int, vp, C.c_int]
    L.orc_level_keypoints.argtypes = [vp, C.c_int, vp, C.c_int]
    L.orc_resize_linear_u8.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int]
    L.orc_gaussian7_u8.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int]
    L.orc_fast9_16.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int]
    L.orc_fast_atan2.restype = C.c_float
    L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
    L.orc_descriptor_distance.argtypes = [vp, vp]
    L.orc_hamming_top2.argtypes = [vp, C.c_int, vp, C.c_int64, C.c_int, C.c_float, C.c_int, vp]
    L.orc_stereo_match.restype = C.c_int
    L.orc_stereo_match.argtypes = [vp, vp, vp, vp, C.c_int, vp, vp, C.c_int, C.c_float, C.c_float, C.c_float, vp, vp]
    L.orc_knn2_ratio.argtypes = [vp, C.c_int, vp, C.c_int, C.c_double, vp, vp, vp]
    L.orc_remap_linear_u8.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int]
    L.orc_to_gray.argtypes = [vp, C.c_longlong, C.c_int, C.c_int, vp]
    L.orc_vocabulary_create.restype = vp
    L.orc_vocabulary_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp]
    L.orc_vocabulary_destroy.argtypes = [vp]
    L.orc_vocabulary_words.argtypes = [vp]
    L.orc_bow_transform.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.orc_search_by_bow.argtypes = [vp, vp, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, vp, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, vp, vp]
    L.orc_search_by_bow.restype = C.c_int
    L.orc_search_for_triangulation.argtypes = [vp, vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, vp,
                                               C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
    L.orc_search_for_triangulation.restype = C.c_int
    L.orc_distinctive_descriptors.argtypes = [vp, vp, C.c_int, vp, vp]
    L.orc_search_by_projection.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_float, C.c_int, vp, vp]
    L.orc_search_by_projection.restype = C.c_int
    L.orc_search_for_initialization.argtypes = [vp, vp, C.c_int, vp, vp, C.c_int, vp, vp, C.c_int, C.c_float, C.c_int, C.c_int, vp]
    L.orc_search_for_initialization.restype = C.c_int
    L.orc_clahe_u8.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, vp, C.c_int, vp]
    L.orc_clahe_u8.restype = C.c_int
    L.orc_svd4x4_vt.argtypes = [vp, vp, vp]
    L.orc_kb8_unproject.argtypes = [vp, C.c_float, C.c_float, vp]
    L.orc_kb8_project.argtypes = [vp, vp, vp]
    L.orc_kb8_triangulate_matches.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, C.c_int, vp, vp]
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleExtractor:
    """Mirror of ORB_SLAM3::ORBextractor (include/ORBextractor.h:43-109) over the CPU oracle."""

    def __init__(self, nfeatures=1200, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.L = lib()
        self.h = self.L.orc_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        if not self.h:
            raise ValueError("bad extractor parameters")
        self.nfeatures, self.nlevels = nfeatures, nlevels

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_extractor_destroy(self.h)
            self.h = None

    def __call__(self, image, lapping=(0, 0)):
        """Returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32] u8)."""
        img = np.ascontiguousarray(image, dtype=np.uint8)
        cap = self.nfeatures * 2 + 64 * self.nlevels
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int32(0)
        rows, cols = (img.shape if img.ndim == 2 else (0, 0))
        mono = self.L.orc_extract(self.h, _ptr(img) if img.size else None, rows, cols, cols, lapping[0], lapping[1],
                                  _ptr(kps), _ptr(desc), cap, C.byref(n))
        if mono < 0:
            return mono, kps[:0], desc[:0]
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def quotas(self):
        return [self.L.orc_features_per_level(self.h, l) for l in range(self.nlevels)]

    def scale_factors(self):
        return [self.L.orc_scale_factor(self.h, l) for l in range(self.nlevels)]

    def umax(self):
        return [self.L.orc_umax(self.h, v) for v in range(16)]

    def level_size(self, level):
        w, h = C.c_int32(0), C.c_int32(0)
        self.L.orc_level_size(self.h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def pyramid_level(self, level, padded=False):
        w, h = self.level_size(level)
        b = 19 if padded else 0
        out = np.zeros((h + 2 * b, w + 2 * b), np.uint8)
        self.L.orc_pyramid_level(self.h, level, int(padded), _ptr(out))
        return out

    def blurred_level(self, level):
        w, h = self.level_size(level)
        out = np.zeros((h, w), np.uint8)
        rc = self.L.orc_blurred_level(self.h, level, _ptr(out))
        return out if rc == 0 else None

    def fast_candidates(self, level):
        n = self.L.orc_fast_candidates(self.h, level, None, 0)
        out = np.zeros(max(n, 1), KP_DTYPE)
        self.L.orc_fast_candidates(self.h, level, _ptr(out), n)
        return out[:n]

    def level_keypoints(self, level):
        n = self.L.orc_level_keypoints(self.h, level, None, 0)
        out = np.zeros(max(n, 1), KP_DTYPE)
        self.L.orc_level_keypoints(self.h, level, _ptr(out), n)
        return out[:n]


def resize_linear(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((dh, dw), np.uint8)
    lib().orc_resize_linear_u8(_ptr(src), src.shape[1], src.shape[0], src.shape[1], _ptr(dst), dw, dh, dw)
    return dst


def gaussian7(src):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros_like(src)
    lib().orc_gaussian7_u8(_ptr(src), src.shape[1], src.shape[0], src.shape[1], _ptr(dst), src.shape[1])
    return dst


def fast9_16(roi, t):
    roi = np.ascontiguousarray(roi, np.uint8)
    cap = roi.size
    out = np.zeros(max(cap, 1), KP_DTYPE)
    n = lib().orc_fast9_16(_ptr(roi), roi.shape[1], roi.shape[0], roi.shape[1], t, _ptr(out), cap)
    return out[:n]


def fast_atan2(y, x):
    return lib().orc_fast_atan2(float(y), float(x))


def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().orc_descriptor_distance(_ptr(a), _ptr(b))


def hamming_top2(q, db, th=50, ratio=0.9, strict_lt=False):
    q = np.ascontiguousarray(q, np.uint8)
    db = np.ascontiguousarray(db, np.uint8)
    out = np.zeros(q.shape[0], MATCH_DTYPE)
    lib().orc_hamming_top2(_ptr(q), q.shape[0], _ptr(db), db.shape[0], th, ratio, int(strict_lt), _ptr(out))
    return out


def stereo_match(exL, exR, kpsL, descL, kpsR, descR, mbf, minD, maxD):
    kpsL = np.ascontiguousarray(kpsL); kpsR = np.ascontiguousarray(kpsR)
    descL = np.ascontiguousarray(descL, np.uint8); descR = np.ascontiguousarray(descR, np.uint8)
    u = np.zeros(len(kpsL), np.float32)
    d = np.zeros(len(kpsL), np.float32)
    kept = lib().orc_stereo_match(exL.h, exR.h, _ptr(kpsL), _ptr(descL), len(kpsL), _ptr(kpsR), _ptr(descR),
                                  len(kpsR), mbf, minD, maxD, _ptr(u), _ptr(d))
    return kept, u, d


def knn2_ratio(q, t, ratio=0.7):
    q = np.ascontiguousarray(q, np.uint8)
    t = np.ascontiguousarray(t, np.uint8)
    m = np.zeros(q.shape[0], np.int32)
    d0 = np.zeros(q.shape[0], np.int32)
    d1 = np.zeros(q.shape[0], np.int32)
    lib().orc_knn2_ratio(_ptr(q), q.shape[0], _ptr(t), t.shape[0], ratio, _ptr(m), _ptr(d0), _ptr(d1))
    return m, d0, d1


def remap_linear(src, mapx, mapy):
    """cv::remap(src, dst, mapx, mapy, INTER_LINEAR), 8UC1, float maps, constant-0 border."""
    src = np.ascontiguousarray(src, np.uint8)
    mx = np.ascontiguousarray(mapx, np.float32); my = np.ascontiguousarray(mapy, np.float32)
    dst = np.zeros_like(src)
    lib().orc_remap_linear_u8(_ptr(src), src.shape[1], src.shape[0], src.shape[1], _ptr(mx), _ptr(my), _ptr(dst), src.shape[1])
    return dst


def to_gray(img, rgb):
    """cv::cvtColor(img, COLOR_{RGB,BGR,RGBA,BGRA}2GRAY); img is [h, w, 3 or 4] u8, rgb = first channel is red."""
    img = np.ascontiguousarray(img, np.uint8)
    dst = np.zeros(img.shape[:2], np.uint8)
    lib().orc_to_gray(_ptr(img), img.shape[0] * img.shape[1], img.shape[2], int(rgb), _ptr(dst))
    return dst


class OracleVocabulary:
    """DBoW2 TemplatedVocabulary<FORB> built from text-file rows; transform() = Frame::ComputeBoW's call."""

    def __init__(self, k, L, scoring, weighting, parent, is_leaf, desc, weight):
        parent = np.ascontiguousarray(parent, np.int32); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8); weight = np.ascontiguousarray(weight, np.float64)
        self.L = lib()
        self.h = self.L.orc_vocabulary_create(k, L, scoring, weighting, parent.shape[0], _ptr(parent), _ptr(is_leaf), _ptr(desc), _ptr(weight))
        if not self.h:
            raise ValueError("bad vocabulary rows")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_vocabulary_destroy(self.h)
            self.h = None

    def words(self):
        return self.L.orc_vocabulary_words(self.h)

    def transform(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8)
        n = desc.shape[0]
        wid, nid = np.zeros(n, np.int32), np.zeros(n, np.int32)
        bw, bv = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        fn, fs, ff = np.zeros(max(n, 1), np.int32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nb, nf = C.c_int(0), C.c_int(0)
        self.L.orc_bow_transform(self.h, _ptr(desc), n, levelsup, _ptr(wid), _ptr(nid), _ptr(bw), _ptr(bv), C.byref(nb), _ptr(fn), _ptr(fs), _ptr(ff),
                                 C.byref(nf))
        nb, nf = nb.value, nf.value
        return {"word_ids": wid, "node_ids": nid, "bow_words": bw[:nb].copy(), "bow_values": bv[:nb].copy(),
                "feat_vec": {int(fn[e]): ff[fs[e]:fs[e + 1]].copy() for e in range(nf)}}


def _fv_arrays(feat_vec):
    """{node: feature rows} (key order) -> (nodes, starts, feats) int32 arrays"""
    nodes = np.array(list(feat_vec.keys()), np.int32)
    lens = [len(v) for v in feat_vec.values()]
    starts = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    feats = np.concatenate([np.asarray(v, np.int32) for v in feat_vec.values()]).astype(np.int32) if lens else np.zeros(0, np.int32)
    return nodes, starts, feats


def search_by_bow(desc_kf, kps_kf, fv_kf, has_map_point, desc_f, kps_f, fv_f, th_low=50, nnratio=0.7, check_orientation=True, kf_kf=False,
                  has_map_point_f=None, n_left_f=-1):
    """ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches); fv_* = {node: rows} dicts.  Returns (match_f, kf_out, nmatches)."""
    desc_kf = np.ascontiguousarray(desc_kf, np.uint8); desc_f = np.ascontiguousarray(desc_f, np.uint8)
    kps_kf = np.ascontiguousarray(kps_kf, KP_DTYPE); kps_f = np.ascontiguousarray(kps_f, KP_DTYPE)
    na, sa, fa = _fv_arrays(fv_kf)
    nb, sb, fb = _fv_arrays(fv_f)
    mp = None if has_map_point is None else np.ascontiguousarray(has_map_point, np.uint8)
    mpf = None if has_map_point_f is None else np.ascontiguousarray(has_map_point_f, np.uint8)
    match_f = np.zeros(max(desc_f.shape[0], 1), np.int32)
    kf_out = np.zeros(max(desc_kf.shape[0], 1), MATCH_DTYPE)
    nm = lib().orc_search_by_bow(_ptr(desc_kf), _ptr(kps_kf), desc_kf.shape[0], _ptr(na), _ptr(sa), _ptr(fa), len(na), _ptr(mp) if mp is not None else None,
                                 _ptr(desc_f), _ptr(kps_f), desc_f.shape[0], _ptr(nb), _ptr(sb), _ptr(fb), len(nb), _ptr(mpf) if mpf is not None else None, int(n_left_f), th_low, nnratio,
                                 int(check_orientation), int(kf_kf), _ptr(match_f), _ptr(kf_out))
    return match_f[:desc_f.shape[0]], kf_out[:desc_kf.shape[0]], nm


def search_for_triangulation(desc_1, kps_1, u_right_1, has_mp_1, fv_1, desc_2, kps_2, u_right_2, has_mp_2, fv_2, f12, epipole, scale_factors,
                             level_sigma2, th_low=50, only_stereo=False, coarse=False, check_orientation=True, rig=None):
    """ORBmatcher::SearchForTriangulation; rig = the bytes of an orbfe_triangulation_rig for two-camera keyframes, None for pinhole ones.
    Returns (vMatches12, nmatches)."""
    c = lambda a, t: None if a is None else np.ascontiguousarray(a, t)  # noqa: E731
    d1, d2, k1, k2 = c(desc_1, np.uint8), c(desc_2, np.uint8), c(kps_1, KP_DTYPE), c(kps_2, KP_DTYPE)
    u1, u2, m1, m2 = c(u_right_1, np.float32), c(u_right_2, np.float32), c(has_mp_1, np.uint8), c(has_mp_2, np.uint8)
    n1, s1, f1 = _fv_arrays(fv_1)
    n2, s2, f2 = _fv_arrays(fv_2)
    F, ep = c(np.asarray(f12).reshape(9), np.float32), c(epipole, np.float32)
    sf, ls = c(np.resize(np.asarray(scale_factors, np.float32), 16), np.float32), c(np.resize(np.asarray(level_sigma2, np.float32), 16), np.float32)
    out = np.zeros(max(d1.shape[0], 1), np.int32)
    p = lambda a: _ptr(a) if a is not None else None  # noqa: E731
    nm = lib().orc_search_for_triangulation(p(d1), p(k1), p(u1), p(m1), d1.shape[0], p(n1), p(s1), p(f1), len(n1), p(d2), p(k2), p(u2), p(m2), d2.shape[0],
                                            p(n2), p(s2), p(f2), len(n2), p(F), p(ep), p(sf), p(ls), th_low, int(only_stereo), int(coarse),
                                            int(check_orientation), p(out), p(rig))
    return out[:d1.shape[0]], nm


def distinctive_descriptors(desc, offsets):
    """MapPoint::ComputeDistinctiveDescriptors for many map points; returns (BestIdx, BestMedian) per point."""
    desc = np.ascontiguousarray(desc, np.uint8)
    off = np.ascontiguousarray(offsets, np.int32)
    idx = np.zeros(off.shape[0] - 1, np.int32)
    med = np.zeros(off.shape[0] - 1, np.int32)
    lib().orc_distinctive_descriptors(_ptr(desc), _ptr(off), off.shape[0] - 1, _ptr(idx), _ptr(med))
    return idx, med


PROJ_QUERY_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("radius", "<f4"), ("xr", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"),
                             ("flags", "<i4"), ("angle", "<f4")])
PROJ_MATCH_DTYPE = np.dtype([("best_idx", "<i4"), ("best_dist", "<i4"), ("best_level", "<i4"), ("second_dist", "<i4"),
                             ("second_level", "<i4"), ("accepted", "<i4"), ("status", "<i4")])


def search_by_projection(kps, desc, queries, qdesc, grid, n_left=-1, u_right=None, partner=None, taken=None, th_high=100, nnratio=0.8,
                         check_orientation=False, inv_level_sigma2=None):
    """ORBmatcher::SearchByProjection(Frame&, vpMapPoints, th) / (CurrentFrame, LastFrame, th, bMono) / Fuse on arrays, serial.
    Returns (matches, taken_after, nmatches)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc = np.ascontiguousarray(desc, np.uint8)
    queries = np.ascontiguousarray(queries, PROJ_QUERY_DTYPE)
    qdesc = np.ascontiguousarray(qdesc, np.uint8)
    grid = np.ascontiguousarray(grid, np.float32)
    ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
    pt = None if partner is None else np.ascontiguousarray(partner, np.int32)
    tk = np.zeros(kps.shape[0], np.uint8) if taken is None else np.array(taken, np.uint8)
    out = np.zeros(queries.shape[0], PROJ_MATCH_DTYPE)
    sig = None if inv_level_sigma2 is None else np.ascontiguousarray(np.resize(np.asarray(inv_level_sigma2, np.float32), 16))
    nm = lib().orc_search_by_projection(_ptr(kps), _ptr(desc), kps.shape[0], n_left, _ptr(ur) if ur is not None else None,
                                        _ptr(pt) if pt is not None else None, _ptr(tk), _ptr(grid), _ptr(queries), _ptr(qdesc),
                                        queries.shape[0], th_high, nnratio, int(check_orientation), _ptr(sig) if sig is not None else None, _ptr(out))
    return out, tk, nm


def search_for_initialization(kps1, desc1, kps2, desc2, grid, prev_matched, window=100, nnratio=0.9, check_orientation=True, th_low=50):
    """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:706-821) on arrays.  Returns (vnMatches12, vbPrevMatched afterwards, nmatches)."""
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    desc1 = np.ascontiguousarray(desc1, np.uint8); desc2 = np.ascontiguousarray(desc2, np.uint8)
    grid = np.ascontiguousarray(grid, np.float32)
    prev = np.array(prev_matched, np.float32).reshape(kps1.shape[0], 2).copy()
    m12 = np.zeros(kps1.shape[0], np.int32)
    nm = lib().orc_search_for_initialization(_ptr(kps1), _ptr(desc1), kps1.shape[0], _ptr(kps2), _ptr(desc2), kps2.shape[0], _ptr(grid), _ptr(prev),
                                             int(window), float(nnratio), int(check_orientation), int(th_low), _ptr(m12))
    return m12, prev, nm


def clahe(img, clip=3.0, tiles=(8, 8), return_lut=False):
    """cv::createCLAHE(clip, cv::Size(tiles[0], tiles[1]))->apply(img); tiles = (across, down)."""
    img = np.ascontiguousarray(img, np.uint8)
    dst = np.zeros_like(img)
    lut = np.zeros((tiles[1], tiles[0], 256), np.uint8)
    rc = lib().orc_clahe_u8(_ptr(img), img.shape[1], img.shape[0], img.shape[1], float(clip), tiles[0], tiles[1], _ptr(dst), img.shape[1], _ptr(lut))
    if rc:
        raise ValueError("image too small for the tile grid")
    return (dst, lut) if return_lut else dst


def merge_top2_host(parts, th, ratio, strict_lt):
    """Oracle of the shard merge (k_top2_merge / orbfe_top2_merge_device): parts is [nshards, nq] MATCH_DTYPE in ascending
    shard (= global row) order; scanning the shards in that order with the reference's strict `<` update
    (src/ORBmatcher.cc:330-340) gives what one scan over the whole database gives."""
    parts = np.asarray(parts)
    ns, nq = parts.shape
    out = np.zeros(nq, MATCH_DTYPE)
    b1 = np.full(nq, 256, np.int32); b2 = b1.copy(); bi = np.full(nq, -1, np.int32)
    for s in range(ns):
        d1, d2, ix = parts[s]["best1"], parts[s]["best2"], parts[s]["idx"]
        lt = d1 < b1
        b2 = np.where(lt, b1, np.where(d1 < b2, d1, b2))
        bi = np.where(lt, ix, bi)
        b1 = np.where(lt, d1, b1)
        b2 = np.minimum(b2, d2)
    out["best1"], out["best2"], out["idx"] = b1, b2, bi
    th_ok = (b1 < th) if strict_lt else (b1 <= th)
    out["accepted"] = th_ok & (b1.astype(np.float32) < np.float32(ratio) * b2.astype(np.float32))
    return out


def svd4x4_vt(A):
    """cv::SVD::compute(A, w, u, vt, MODIFY_A | FULL_UV) for a 4x4 float32 matrix: (w[4], vt[4, 4])."""
    A = np.ascontiguousarray(A, np.float32)
    w, vt = np.zeros(4, np.float32), np.zeros((4, 4), np.float32)
    lib().orc_svd4x4_vt(_ptr(A), _ptr(w), _ptr(vt))
    return w, vt


def kb8_unproject(params, u, v):
    p = np.ascontiguousarray(params, np.float32)
    out = np.zeros(3, np.float32)
    lib().orc_kb8_unproject(_ptr(p), float(u), float(v), _ptr(out))
    return out


def kb8_project(params, p3):
    p, x = np.ascontiguousarray(params, np.float32), np.ascontiguousarray(p3, np.float32)
    out = np.zeros(2, np.float32)
    lib().orc_kb8_project(_ptr(p), _ptr(x), _ptr(out))
    return out


def kb8_triangulate_matches(cam1, cam2, R12, t12, uv1, uv2, sigma1, sigma2):
    """KannalaBrandt8::TriangulateMatches per keypoint pair: (depth[n] (-1 = rejected), p3d[n, 3])."""
    c = lambda a: np.ascontiguousarray(a, np.float32)  # noqa: E731
    cam1, cam2, R12, t12, uv1, uv2, sigma1, sigma2 = c(cam1), c(cam2), c(R12), c(t12), c(uv1), c(uv2), c(sigma1), c(sigma2)
    n = uv1.shape[0]
    depth, p3d = np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
    lib().orc_kb8_triangulate_matches(_ptr(cam1), _ptr(cam2), _ptr(R12), _ptr(t12), _ptr(uv1), _ptr(uv2), _ptr(sigma1), _ptr(sigma2), n, _ptr(depth), _ptr(p3d))
    return depth, p3d
