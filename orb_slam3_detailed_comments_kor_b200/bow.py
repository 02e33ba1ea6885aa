"""Host-side mirror of the bag-of-words assignment over the C ABI: ORBVocabulary (DBoW2
TemplatedVocabulary<FORB::TDescriptor, FORB>) and its transform(features, BowVector, FeatureVector, levelsup) as called by
Frame::ComputeBoW (/root/reference/src/Frame.cc:724-731; Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1125-1260)."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, MATCH_DTYPE, check, ptr

L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = range(6)   # DBoW2::ScoringType (BowVector.h:48-56)
TF_IDF, TF, IDF, BINARY = range(4)                                       # DBoW2::WeightingType (BowVector.h:39-45)


class ORBVocabulary:
    """Built from the rows of the vocabulary text file (parent, is_leaf, 32 descriptor bytes, weight per node, file order)."""

    def __init__(self, k, L, scoring, weighting, parent, is_leaf, desc, weight, max_features=4096, max_batch=1, device=0):
        parent = np.ascontiguousarray(parent, np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8)
        weight = np.ascontiguousarray(weight, np.float64)
        self.max_features, self.max_batch = max_features, max_batch
        self._L = _lib.lib()
        self._h = C.c_void_p()
        check(self._L.orbfe_vocabulary_create(device, k, L, scoring, weighting, parent.shape[0], ptr(parent), ptr(is_leaf), ptr(desc), ptr(weight),
                                              max_features, max_batch, C.byref(self._h)))

    @property
    def handle(self):
        return self._h

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbfe_vocabulary_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def words(self):
        return self._L.orbfe_vocabulary_words(self._h)

    def _fetch(self, frame, n):
        F = self.max_features
        wid, nid = np.zeros(n, np.int32), np.zeros(n, np.int32)
        bw, bv = np.zeros(F, np.int32), np.zeros(F, np.float64)
        fn, fs, ff = np.zeros(F, np.int32), np.zeros(F + 1, np.int32), np.zeros(F, np.int32)
        nb, nf = C.c_int(0), C.c_int(0)
        check(self._L.orbfe_bow_result(self._h, frame, n, ptr(wid), ptr(nid), ptr(bw), ptr(bv), C.byref(nb), ptr(fn), ptr(fs), ptr(ff), C.byref(nf)))
        nb, nf = nb.value, nf.value
        feat_vec = {int(fn[e]): ff[fs[e]:fs[e + 1]].copy() for e in range(nf)}
        return {"word_ids": wid, "node_ids": nid, "bow_words": bw[:nb].copy(), "bow_values": bv[:nb].copy(), "feat_vec": feat_vec}

    def transform(self, desc, levelsup=4):
        """transform(features, BowVec, FeatVec, levelsup) for one frame's descriptors [n, 32] on the host."""
        desc = np.ascontiguousarray(desc, np.uint8)
        n = desc.shape[0]
        check(self._L.orbfe_bow_transform(self._h, ptr(desc) if n else None, n, levelsup, None, None, None, None, None, None, None, None, None))
        return self._fetch(0, n)

    def transform_batch_device(self, d_desc, row_stride, d_counts, nframes, levelsup=4, stream=None):
        check(self._L.orbfe_bow_transform_batch_device(self._h, C.c_void_p(d_desc), row_stride, C.c_void_p(d_counts), nframes, levelsup,
                                                       C.c_void_p(stream) if stream else None))

    def result(self, frame, nfeatures):
        return self._fetch(frame, nfeatures)


def _fv_arrays(feat_vec):
    nodes = np.array(list(feat_vec.keys()), np.int32)
    lens = [len(v) for v in feat_vec.values()]
    starts = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    feats = np.concatenate([np.asarray(v, np.int32) for v in feat_vec.values()]).astype(np.int32) if lens else np.zeros(0, np.int32)
    return nodes, starts, feats


def search_by_bow(desc_kf, kps_kf, fv_kf, has_map_point, desc_f, kps_f, fv_f, th_low=50, nnratio=0.7, check_orientation=True, device=0,
                  kf_kf=False, has_map_point_f=None, n_left_f=-1):
    """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:269-470) — or, with kf_kf, the
    keyframe-keyframe overload (:823-963) — for one pair on the host; fv_* are the feat_vec dicts of ORBVocabulary.transform.
    Returns (match_f[nF] keyframe row or -1, kf_out, nmatches)."""
    desc_kf = np.ascontiguousarray(desc_kf, np.uint8); desc_f = np.ascontiguousarray(desc_f, np.uint8)
    kps_kf = np.ascontiguousarray(kps_kf, KP_DTYPE); kps_f = np.ascontiguousarray(kps_f, KP_DTYPE)
    na, sa, fa = _fv_arrays(fv_kf)
    nb, sb, fb = _fv_arrays(fv_f)
    mp = None if has_map_point is None else np.ascontiguousarray(has_map_point, np.uint8)
    mpf = None if has_map_point_f is None else np.ascontiguousarray(has_map_point_f, np.uint8)
    match_f = np.zeros(max(desc_f.shape[0], 1), np.int32)
    kf_out = np.zeros(max(desc_kf.shape[0], 1), MATCH_DTYPE)
    nm = C.c_int(0)
    check(_lib.lib().orbfe_search_by_bow(device, ptr(desc_kf), ptr(kps_kf), desc_kf.shape[0], ptr(na), ptr(sa), ptr(fa), len(na),
                                         ptr(mp) if mp is not None else None, ptr(desc_f), ptr(kps_f), desc_f.shape[0], ptr(nb), ptr(sb), ptr(fb), len(nb),
                                         ptr(mpf) if mpf is not None else None, int(n_left_f), th_low, C.c_float(nnratio), int(check_orientation),
                                         int(kf_kf),
                                         ptr(match_f), ptr(kf_out), C.byref(nm)))
    return match_f[:desc_f.shape[0]], kf_out[:desc_kf.shape[0]], nm.value


def make_triangulation_rig(cams, Rs, ts, n_left_1, n_left_2):
    """orbfe_triangulation_rig: cams = 4 x 8 (camera 1 left / right, camera 2 left / right), Rs = 4 x 3 x 3 and ts = 4 x 3 (ll, lr, rl, rr)."""
    buf = np.concatenate([np.asarray(cams, np.float32).reshape(32), np.asarray(Rs, np.float32).reshape(36), np.asarray(ts, np.float32).reshape(12)]).astype(np.float32)
    return np.concatenate([buf.view(np.uint8), np.array([n_left_1, n_left_2], np.int32).view(np.uint8)])


def search_for_triangulation(desc_1, kps_1, u_right_1, has_mp_1, fv_1, desc_2, kps_2, u_right_2, has_mp_2, fv_2, f12, epipole, scale_factors,
                             level_sigma2, th_low=50, only_stereo=False, coarse=False, check_orientation=True, device=0, rig=None):
    """ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, ..., bOnlyStereo, bCoarse) (src/ORBmatcher.cc:965-1206) for one pinhole
    keyframe pair on the host.  Returns (vMatches12[n1], nmatches)."""
    c = lambda a, t: None if a is None else np.ascontiguousarray(a, t)  # noqa: E731
    d1, d2, k1, k2 = c(desc_1, np.uint8), c(desc_2, np.uint8), c(kps_1, KP_DTYPE), c(kps_2, KP_DTYPE)
    u1, u2, m1, m2 = c(u_right_1, np.float32), c(u_right_2, np.float32), c(has_mp_1, np.uint8), c(has_mp_2, np.uint8)
    n1, s1, f1 = _fv_arrays(fv_1)
    n2, s2, f2 = _fv_arrays(fv_2)
    F, ep = c(np.asarray(f12).reshape(9), np.float32), c(epipole, np.float32)
    sf, ls = c(scale_factors, np.float32), c(level_sigma2, np.float32)
    out = np.zeros(max(d1.shape[0], 1), np.int32)
    nm = C.c_int(0)
    p = lambda a: ptr(a) if a is not None else None  # noqa: E731
    check(_lib.lib().orbfe_search_for_triangulation(device, p(d1), p(k1), p(u1), p(m1), d1.shape[0], p(n1), p(s1), p(f1), len(n1), p(d2), p(k2), p(u2), p(m2),
                                                    d2.shape[0], p(n2), p(s2), p(f2), len(n2), p(F), p(ep), p(sf), p(ls), len(sf), th_low, int(only_stereo),
                                                    int(coarse), int(check_orientation), p(out), C.byref(nm), p(rig)))
    return out[:d1.shape[0]], nm.value


class _HostFrame(C.Structure):
    """orbfe_host_frame"""
    _fields_ = [("desc", C.c_void_p), ("kps", C.c_void_p), ("u_right", C.c_void_p), ("has_map_point", C.c_void_p), ("n", C.c_int32),
                ("fv_nodes", C.c_void_p), ("fv_starts", C.c_void_p), ("fv_feats", C.c_void_p), ("nfv", C.c_int32), ("n_left", C.c_int32)]


def _host_frames(frames):
    """frames: dicts with desc [n, 32], kps, feat_vec and optionally u_right, has_map_point, n_left.  Returns (ctypes array, the numpy arrays it
    points into, which must outlive the call)."""
    arr = (_HostFrame * len(frames))()
    keep = []
    for i, f in enumerate(frames):
        c = lambda a, t: None if a is None else np.ascontiguousarray(a, t)  # noqa: E731
        d, k = c(f["desc"], np.uint8), c(f.get("kps"), KP_DTYPE)
        u, m = c(f.get("u_right"), np.float32), c(f.get("has_map_point"), np.uint8)
        n, s, ff = _fv_arrays(f["feat_vec"])
        keep += [d, k, u, m, n, s, ff]
        a = lambda x: x.ctypes.data if x is not None and x.size else None  # noqa: E731
        arr[i] = _HostFrame(a(d), a(k), a(u), a(m), d.shape[0], a(n), s.ctypes.data if len(n) else None, a(ff), len(n), int(f.get("n_left", -1)))
    return arr, keep


def search_for_triangulation_frames(frames, pair_1, pair_2, f12, epipole, scale_factors, level_sigma2, th_low=50, only_stereo=False, coarse=False,
                                    check_orientation=True, rigs=None, device=0):
    """orbfe_search_for_triangulation_frames: SearchForTriangulation for many keyframe pairs in one call (LocalMapping::CreateNewMapPoints' loop over
    the neighbours).  f12: [npairs, 9], epipole: [npairs, 2], rigs: list of make_triangulation_rig() byte arrays or None.  Returns (list of
    vMatches12 arrays, nmatches[npairs])."""
    arr, keep = _host_frames(frames)
    p1, p2 = np.ascontiguousarray(pair_1, np.int32), np.ascontiguousarray(pair_2, np.int32)
    F, ep = np.ascontiguousarray(f12, np.float32).reshape(len(p1), 9), np.ascontiguousarray(epipole, np.float32).reshape(len(p1), 2)
    sf, ls = np.ascontiguousarray(scale_factors, np.float32), np.ascontiguousarray(level_sigma2, np.float32)
    stride = max(max(f["desc"].shape[0] for f in frames), 1)
    out, nm = np.zeros((len(p1), stride), np.int32), np.zeros(len(p1), np.int32)
    rg = None if rigs is None else np.ascontiguousarray(np.concatenate([np.asarray(r, np.uint8) for r in rigs]))
    check(_lib.lib().orbfe_search_for_triangulation_frames(device, C.addressof(arr), len(frames), ptr(p1), ptr(p2), len(p1), ptr(F), ptr(ep), ptr(sf), ptr(ls), len(sf),
                                                           th_low, int(only_stereo), int(coarse), int(check_orientation), ptr(out), stride, ptr(nm),
                                                           ptr(rg) if rg is not None else None))
    return [out[p, :frames[int(p1[p])]["desc"].shape[0]] for p in range(len(p1))], nm


def search_by_bow_frames(frames, pair_kf, pair_f, th_low=50, nnratio=0.7, check_orientation=True, kf_kf=False, want_kf_out=False, device=0):
    """orbfe_search_by_bow_frames: SearchByBoW for many (keyframe, frame) pairs in one call (the candidate loops of Tracking::Relocalization and
    LoopClosing).  Returns (list of per-frame-row match arrays, list of kf_out arrays or None, nmatches[npairs])."""
    arr, keep = _host_frames(frames)
    pk, pf = np.ascontiguousarray(pair_kf, np.int32), np.ascontiguousarray(pair_f, np.int32)
    stride = max(max(f["desc"].shape[0] for f in frames), 1)
    out, nm = np.zeros((len(pk), stride), np.int32), np.zeros(len(pk), np.int32)
    ko = np.zeros((len(pk), stride), MATCH_DTYPE) if want_kf_out else None
    check(_lib.lib().orbfe_search_by_bow_frames(device, C.addressof(arr), len(frames), ptr(pk), ptr(pf), len(pk), th_low, C.c_float(nnratio), int(check_orientation),
                                                int(kf_kf), ptr(out), stride, ptr(ko) if ko is not None else None, ptr(nm)))
    n = lambda i: frames[int(i)]["desc"].shape[0]  # noqa: E731
    return [out[p, :n(pf[p])] for p in range(len(pk))], ([ko[p, :n(pk[p])] for p in range(len(pk))] if want_kf_out else None), nm
