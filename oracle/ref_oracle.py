"""ctypes loader for oracle/_ref/liborb_ref.so: the REFERENCE'S OWN ORBextractor.cc (compiled unmodified) and the bodies of
Frame::ComputeStereoMatches / ORBmatcher::DescriptorDistance, built by oracle/Makefile.  TEST INFRASTRUCTURE ONLY.

The library can only be built where /root/reference exists (this container); the GPU box receives the prebuilt file.
Only tests/ and bench.py's cpu_baseline / --impl reference legs import this module.  See oracle/ref_shim.cpp.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from .orb_oracle import KP_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("ORBFE_REFERENCE", "/root/reference")
_libs = {}


def so_path(fma=False):
    return os.path.join(_HERE, "_ref", "liborb_ref_fma.so" if fma else "liborb_ref.so")


def build():
    """Run the committed recipe (oracle/Makefile) when the reference sources are present; otherwise keep what is there."""
    if os.path.exists(os.path.join(REF_SRC, "src", "ORBextractor.cc")):
        subprocess.check_call(["make", "-C", _HERE, "REF=" + REF_SRC, "_ref/liborb_ref.so", "_ref/liborb_ref_fma.so", "_ref/libdbow_ref.so"],
                              stdout=subprocess.DEVNULL)
    return so_path() if os.path.exists(so_path()) else None


def available(fma=False):
    return os.path.exists(so_path(fma)) or build() is not None


def lib(fma=False):
    if fma in _libs:
        return _libs[fma]
    if not available(fma):
        raise RuntimeError("oracle/_ref is not built and %s is absent" % REF_SRC)
    L = C.CDLL(so_path(fma))
    vp, i32p = C.c_void_p, C.POINTER(C.c_int32)
    par = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    L.ref_set_arena.argtypes = [C.c_int]
    L.ref_arena_peak.restype = C.c_size_t
    L.ref_tables.argtypes = par + [vp, vp, vp]
    L.ref_extract.argtypes = par + [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, i32p, vp]
    L.ref_level_keypoints.argtypes = par + [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, vp]
    L.ref_stereo.argtypes = par + [vp, vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, vp, vp, i32p, vp, vp, i32p, C.c_int, vp, vp]
    L.ref_descriptor_distance.argtypes = [vp, vp]
    L.ref_search_by_bow.argtypes = [vp, vp, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, vp, C.c_int, C.c_float, C.c_int, C.c_int, vp]
    L.ref_distinctive_descriptor.argtypes = [vp, C.c_int, vp]
    L.ref_search_for_initialization.argtypes = [vp, vp, C.c_int, vp, vp, C.c_int, vp, vp, C.c_int, C.c_float, C.c_int, vp]
    L.ref_search_for_initialization.restype = C.c_int
    L.ref_search_by_projection_last_frame.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_float, C.c_float, vp, vp, vp, vp, C.c_int, C.c_int,
                                                      vp, vp, vp, vp, vp, vp, C.c_float, C.c_int, C.c_int, vp]
    L.ref_search_by_projection_last_frame.restype = C.c_int
    L.ref_fisheye_stereo.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_int, vp, vp, vp, vp]
    L.ref_fisheye_stereo.restype = C.c_int
    L.ref_search_by_projection_reloc.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, C.c_float, C.c_int,
                                                 C.c_int, vp]
    L.ref_search_by_projection_reloc.restype = C.c_int
    L.ref_fuse.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_float, vp, vp, C.c_int] + [vp] * 10 + [C.c_float, C.c_int, vp, vp]
    L.ref_fuse.restype = C.c_int
    L.ref_search_by_projection_sim3.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp, C.c_int, vp, vp, C.c_int, vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_float, vp, vp, vp]
    L.ref_search_by_projection_sim3.restype = C.c_int
    L.ref_fuse_sim3.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp, C.c_int, vp, vp, C.c_int, vp, vp, vp, vp, vp, vp, vp, C.c_float, vp, vp]
    L.ref_fuse_sim3.restype = C.c_int
    L.ref_search_by_sim3.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp, vp, vp] * 2 + [vp, vp, vp, vp, C.c_int, vp, C.c_float, vp, vp, C.c_float, vp]
    L.ref_search_by_sim3.restype = C.c_int
    L.ref_kb8_project.argtypes = [vp, vp, C.c_int, vp]
    L.ref_kb8_project.restype = None
    L.ref_kb8_unproject.argtypes = [vp, vp, C.c_int, vp]
    L.ref_kb8_unproject.restype = None
    L.ref_kb8_triangulate_matches.argtypes = [vp] * 8 + [C.c_int, vp, vp]
    L.ref_kb8_triangulate_matches.restype = None
    L.ref_search_for_triangulation.argtypes = ([vp, vp, vp, vp, C.c_int, C.c_int, vp, vp, vp, C.c_int] * 2
                                               + [vp, C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp])
    L.ref_search_for_triangulation.restype = C.c_int
    L.ref_search_by_projection.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int] + [vp] * 12 + [C.c_float, C.c_float, vp]
    _libs[fma] = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class ReferenceExtractor:
    """ORB_SLAM3::ORBextractor itself (include/ORBextractor.h:43-109); a fresh C++ object per call, as the tables are cheap."""

    def __init__(self, nfeatures=1200, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, fma=False):
        self.L = lib(fma)
        self.params = (int(nfeatures), float(scale_factor), int(nlevels), int(ini_th), int(min_th))
        self.nfeatures, self.nlevels = int(nfeatures), int(nlevels)

    def tables(self):
        q, u, s = np.zeros(self.nlevels, np.int32), np.zeros(16, np.int32), np.zeros(self.nlevels, np.float32)
        self.L.ref_tables(*self.params, _ptr(q), _ptr(u), _ptr(s))
        return list(q), list(u), s

    def level_sizes(self, width, height):
        """cvRound((float)cols*scale) per level, src/ORBextractor.cc:1156-1157 (needed to slice the pyramid dump)."""
        _, _, s = self.tables()
        inv = (np.float32(1.0) / s).astype(np.float32)
        r = lambda v: int(np.rint(np.float32(v)))  # noqa: E731
        return [(r(np.float32(width) * inv[l]), r(np.float32(height) * inv[l])) for l in range(self.nlevels)]

    def __call__(self, image, lapping=(0, 0), want_pyramid=False):
        img = np.ascontiguousarray(image, dtype=np.uint8)
        cap = self.nfeatures * 2 + 64 * self.nlevels
        kps, desc, n = np.zeros(cap, KP_DTYPE), np.zeros((cap, 32), np.uint8), C.c_int32(0)
        rows, cols = (img.shape if img.ndim == 2 else (0, 0))
        pyr = None
        if want_pyramid:
            sizes = self.level_sizes(cols, rows)
            pyr = np.zeros(sum((w + 38) * (h + 38) for w, h in sizes), np.uint8)
        mono = self.L.ref_extract(*self.params, _ptr(img) if img.size else None, rows, cols, cols, lapping[0], lapping[1], _ptr(kps), _ptr(desc),
                                  cap, C.byref(n), _ptr(pyr) if pyr is not None else None)
        if mono < 0:
            return (mono, kps[:0], desc[:0]) + ((None,) if want_pyramid else ())
        out = (mono, kps[:n.value].copy(), desc[:n.value].copy())
        if want_pyramid:
            levels, o = [], 0
            for w, h in sizes:
                levels.append(pyr[o:o + (w + 38) * (h + 38)].reshape(h + 38, w + 38))
                o += (w + 38) * (h + 38)
            out += (levels,)
        return out

    def level_keypoints(self, image):
        """Per level: keypoints after DistributeOctTree and computeOrientation, in level coordinates."""
        img = np.ascontiguousarray(image, dtype=np.uint8)
        cap = self.nfeatures * 2 + 64 * self.nlevels
        out, counts = np.zeros(cap, KP_DTYPE), np.zeros(self.nlevels, np.int32)
        total = self.L.ref_level_keypoints(*self.params, _ptr(img), img.shape[0], img.shape[1], img.shape[1], _ptr(out), cap, _ptr(counts))
        assert total >= 0
        off = np.concatenate([[0], np.cumsum(counts)])
        return [out[off[l]:off[l + 1]].copy() for l in range(self.nlevels)]

    def stereo(self, left, right, mbf, mb):
        """Both extractions + Frame::ComputeStereoMatches; returns (kpsL, descL, kpsR, descR, mvuRight, mvDepth)."""
        a, b = np.ascontiguousarray(left, np.uint8), np.ascontiguousarray(right, np.uint8)
        cap = self.nfeatures * 2 + 64 * self.nlevels
        kL, dL, kR, dR = np.zeros(cap, KP_DTYPE), np.zeros((cap, 32), np.uint8), np.zeros(cap, KP_DTYPE), np.zeros((cap, 32), np.uint8)
        nL, nR, u, d = C.c_int32(0), C.c_int32(0), np.zeros(cap, np.float32), np.zeros(cap, np.float32)
        rc = self.L.ref_stereo(*self.params, _ptr(a), _ptr(b), a.shape[0], a.shape[1], a.shape[1], mbf, mb, _ptr(kL), _ptr(dL), C.byref(nL),
                               _ptr(kR), _ptr(dR), C.byref(nR), cap, _ptr(u), _ptr(d))
        assert rc >= 0
        return kL[:nL.value].copy(), dL[:nL.value].copy(), kR[:nR.value].copy(), dR[:nR.value].copy(), u[:nL.value].copy(), d[:nL.value].copy()


def descriptor_distance(a, b, fma=False):
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    return lib(fma).ref_descriptor_distance(_ptr(a), _ptr(b))


def search_by_bow(desc_kf, kps_kf, fv_kf, has_map_point, desc_f, kps_f, fv_f, nnratio=0.7, check_orientation=True, kf_kf=False, has_map_point_f=None,
                  n_left_f=-1):
    """The reference's own ORBmatcher::SearchByBoW (KeyFrame*, Frame&) — or, with kf_kf, (KeyFrame*, KeyFrame*) — on arrays; same arguments as
    orb_oracle.search_by_bow (has_map_point entries: 0 none, 1 good, 2 bad).  Returns (match, nmatches)."""
    from .orb_oracle import _fv_arrays
    desc_kf, desc_f = np.ascontiguousarray(desc_kf, np.uint8), np.ascontiguousarray(desc_f, np.uint8)
    kps_kf, kps_f = np.ascontiguousarray(kps_kf, KP_DTYPE), np.ascontiguousarray(kps_f, KP_DTYPE)
    na, sa, fa = _fv_arrays(fv_kf)
    nb, sb, fb = _fv_arrays(fv_f)
    mp = np.ascontiguousarray(has_map_point, np.uint8)
    mpf = None if has_map_point_f is None else np.ascontiguousarray(has_map_point_f, np.uint8)
    out = np.zeros(max(desc_kf.shape[0] if kf_kf else desc_f.shape[0], 1), np.int32)
    nm = lib().ref_search_by_bow(_ptr(desc_kf), _ptr(kps_kf), desc_kf.shape[0], _ptr(na), _ptr(sa), _ptr(fa), len(na), _ptr(mp), _ptr(desc_f), _ptr(kps_f),
                                 desc_f.shape[0], _ptr(nb), _ptr(sb), _ptr(fb), len(nb), _ptr(mpf) if mpf is not None else None, int(n_left_f), nnratio,
                                 int(check_orientation), int(kf_kf), _ptr(out))
    return out[:(desc_kf.shape[0] if kf_kf else desc_f.shape[0])], nm


def search_by_projection(kps, desc, n_left, u_right, l2r, r2l, taken, grid4, scale_factors, mp, th, nnratio):
    """The reference's own ORBmatcher::SearchByProjection(Frame&, vpMapPoints, th) on arrays; `mp` is a dict of per-map-point arrays named
    like the MapPoint fields (in_view, in_view_r, level, level_r, view_cos, view_cos_r, proj_x, proj_y, proj_xr, proj_yr, obs, desc).
    Returns (owner[n] = map point index per keypoint row, -2 the initial one, -1 none; nmatches)."""
    kps, desc = np.ascontiguousarray(kps, KP_DTYPE), np.ascontiguousarray(desc, np.uint8)
    n = kps.shape[0]
    f = lambda a, t: np.ascontiguousarray(a, t)  # noqa: E731
    ur = None if u_right is None else f(u_right, np.float32)
    a_l2r = None if l2r is None else f(l2r, np.int32)
    a_r2l = None if r2l is None else f(r2l, np.int32)
    tk, g, sf = f(taken, np.uint8), f(grid4, np.float32), f(scale_factors, np.float32)
    arrs = [f(mp["in_view"], np.uint8), f(mp["in_view_r"], np.uint8), f(mp["level"], np.int32), f(mp["level_r"], np.int32), f(mp["view_cos"], np.float32),
            f(mp["view_cos_r"], np.float32), f(mp["proj_x"], np.float32), f(mp["proj_y"], np.float32), f(mp["proj_xr"], np.float32),
            f(mp["proj_yr"], np.float32), f(mp["obs"], np.int32), f(mp["desc"], np.uint8)]
    owner = np.zeros(max(n, 1), np.int32)
    p = lambda a: _ptr(a) if a is not None else None  # noqa: E731
    nm = lib().ref_search_by_projection(_ptr(kps), _ptr(desc), n, int(n_left), p(ur), p(a_l2r), p(a_r2l), _ptr(tk), _ptr(g), _ptr(sf), len(sf), len(arrs[0]),
                                        *[_ptr(a) for a in arrs], float(th), float(nnratio), _ptr(owner))
    return owner[:n], nm


def search_for_initialization(kps1, desc1, kps2, desc2, grid4, prev_matched, window=100, nnratio=0.9, check_orientation=True):
    """The reference's own ORBmatcher::SearchForInitialization on arrays.  Returns (vnMatches12, vbPrevMatched afterwards, nmatches)."""
    kps1, kps2 = np.ascontiguousarray(kps1, KP_DTYPE), np.ascontiguousarray(kps2, KP_DTYPE)
    desc1, desc2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    g = np.ascontiguousarray(grid4, np.float32)
    prev = np.array(prev_matched, np.float32).reshape(kps1.shape[0], 2).copy()
    m12 = np.zeros(max(kps1.shape[0], 1), np.int32)
    nm = lib().ref_search_for_initialization(_ptr(kps1), _ptr(desc1), kps1.shape[0], _ptr(kps2), _ptr(desc2), kps2.shape[0], _ptr(g), _ptr(prev), int(window),
                                             float(nnratio), int(check_orientation), _ptr(m12))
    return m12[:kps1.shape[0]], prev, nm


def search_by_projection_last_frame(kps, desc, n_left, u_right, obs_cur, grid4, bounds4, scale_factors, mbf, mb, Tcw, Trl, cam4, kps_last, n_left_last, has_mp,
                                    outlier, world, obs_mp, desc_mp, Tlw, th, mono, check_orientation):
    """The reference's own ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) on arrays (see ref_shim.cpp).
    Returns (owner[n], nmatches)."""
    c = lambda a, t: None if a is None else np.ascontiguousarray(a, t)  # noqa: E731
    p = lambda a: _ptr(a) if a is not None else None  # noqa: E731
    f = np.float32
    kps, desc, ur, oc = c(kps, KP_DTYPE), c(desc, np.uint8), c(u_right, f), c(obs_cur, np.int32)
    g, b, sf, Tcw, Trl, cam4 = c(grid4, f), c(bounds4, f), c(scale_factors, f), c(np.asarray(Tcw, f).reshape(16), f), c(np.asarray(Trl, f).reshape(12), f), c(cam4, f)
    kl, hm, ol, w, om, dm, Tlw = c(kps_last, KP_DTYPE), c(has_mp, np.uint8), c(outlier, np.uint8), c(world, f), c(obs_mp, np.int32), c(desc_mp, np.uint8), c(np.asarray(Tlw, f).reshape(16), f)
    owner = np.zeros(max(len(kps), 1), np.int32)
    nm = lib().ref_search_by_projection_last_frame(p(kps), p(desc), len(kps), int(n_left), p(ur), p(oc), p(g), p(b), p(sf), len(sf), float(mbf), float(mb), p(Tcw), p(Trl),
                                                   p(cam4), p(kl), len(kl), int(n_left_last), p(hm), p(ol), p(w), p(om), p(dm), p(Tlw), float(th), int(mono),
                                                   int(check_orientation), p(owner))
    return owner[:len(kps)], nm


def fisheye_stereo(kps_l, desc_l, mono_l, kps_r, desc_r, mono_r, cam1, cam2, Rlr, tlr, level_sigma2):
    """The reference's own Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1119-1159) on the two images' extractor outputs.  Returns
    (mvLeftToRightMatch[nl], mvRightToLeftMatch[nr], mvDepth[nl], mvStereo3Dpoints[nl, 3] with 0 where empty, matches kept)."""
    c = lambda a, t: np.ascontiguousarray(a, t)  # noqa: E731
    f = np.float32
    kl, dl, kr, dr = c(kps_l, KP_DTYPE), c(desc_l, np.uint8), c(kps_r, KP_DTYPE), c(desc_r, np.uint8)
    c1, c2, R, t, ls = c(cam1, f), c(cam2, f), c(np.asarray(Rlr, f).reshape(9), f), c(tlr, f), c(level_sigma2, f)
    l2r, r2l = np.zeros(max(len(kl), 1), np.int32), np.zeros(max(len(kr), 1), np.int32)
    depth, p3d = np.zeros(max(len(kl), 1), f), np.zeros((max(len(kl), 1), 3), f)
    kept = lib().ref_fisheye_stereo(_ptr(kl), _ptr(dl), len(kl), int(mono_l), _ptr(kr), _ptr(dr), len(kr), int(mono_r), _ptr(c1), _ptr(c2), _ptr(R), _ptr(t), _ptr(ls),
                                    len(ls), _ptr(l2r), _ptr(r2l), _ptr(depth), _ptr(p3d))
    return l2r[:len(kl)], r2l[:len(kr)], depth[:len(kl)], p3d[:len(kl)], kept


def search_by_projection_reloc(kps, desc, obs_cur, grid4, bounds4, scale_factors, Tcw, cam4, kps_kf, has_mp, bad, found, world, min_dist, max_dist, level, desc_mp,
                               th, orb_dist, check_orientation):
    """The reference's own ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (Tracking::Relocalization) on arrays (see
    ref_shim.cpp).  Returns (owner[n], nmatches)."""
    c = lambda a, t: np.ascontiguousarray(a, t)  # noqa: E731
    f = np.float32
    kps, desc, oc = c(kps, KP_DTYPE), c(desc, np.uint8), c(obs_cur, np.int32)
    g, b, sf, Tcw, cam4 = c(grid4, f), c(bounds4, f), c(scale_factors, f), c(np.asarray(Tcw, f).reshape(16), f), c(cam4, f)
    kk, hm, bd, fd, w = c(kps_kf, KP_DTYPE), c(has_mp, np.uint8), c(bad, np.uint8), c(found, np.uint8), c(world, f)
    mn, mx, lv, dm = c(min_dist, f), c(max_dist, f), c(level, np.int32), c(desc_mp, np.uint8)
    owner = np.zeros(max(len(kps), 1), np.int32)
    nm = lib().ref_search_by_projection_reloc(_ptr(kps), _ptr(desc), len(kps), _ptr(oc), _ptr(g), _ptr(b), _ptr(sf), len(sf), _ptr(Tcw), _ptr(cam4), _ptr(kk), len(kk),
                                              _ptr(hm), _ptr(bd), _ptr(fd), _ptr(w), _ptr(mn), _ptr(mx), _ptr(lv), _ptr(dm), float(th), int(orb_dist),
                                              int(check_orientation), _ptr(owner))
    return owner[:len(kps)], nm


def fuse(kps, desc, n_left, u_right, row_obs, row_bad, grid4, bounds4, scale_factors, inv_level_sigma2, bf, cam4, pose, mp, th, right):
    """The reference's own ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight) on arrays; `mp` = dict of per-map-point arrays (present, bad, in_kf,
    world, normal, min_dist, max_dist, level, obs, desc).  Returns (fused_row[nmp], fused_how[nmp], nFused)."""
    c = lambda a, t: None if a is None else np.ascontiguousarray(a, t)  # noqa: E731
    p = lambda a: _ptr(a) if a is not None else None  # noqa: E731
    f = np.float32
    kps, desc, ur, ro, rb = c(kps, KP_DTYPE), c(desc, np.uint8), c(u_right, f), c(row_obs, np.int32), c(row_bad, np.uint8)
    g, b, sf, isig, cam4, pose = c(grid4, f), c(bounds4, f), c(scale_factors, f), c(inv_level_sigma2, f), c(cam4, f), c(np.asarray(pose, f).reshape(12), f)
    arrs = [c(mp["present"], np.uint8), c(mp["bad"], np.uint8), c(mp["in_kf"], np.uint8), c(mp["world"], f), c(mp["normal"], f), c(mp["min_dist"], f),
            c(mp["max_dist"], f), c(mp["level"], np.int32), c(mp["obs"], np.int32), c(mp["desc"], np.uint8)]
    nmp = len(arrs[0])
    row, how = np.zeros(max(nmp, 1), np.int32), np.zeros(max(nmp, 1), np.int32)
    nf = lib().ref_fuse(p(kps), p(desc), len(kps), int(n_left), p(ur), p(ro), p(rb), p(g), p(b), p(sf), p(isig), len(sf), float(bf), p(cam4), p(pose), nmp,
                        *[p(a) for a in arrs], float(th), int(right), p(row), p(how))
    return row[:nmp], how[:nmp], nf


def search_by_projection_sim3(kps, desc, matched_in, grid4, bounds4, scale_factors, cam4, Scw, mp, th, ratio_hamming, point_kf=None):
    """The reference's own ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming) on arrays.  Returns (owner[n], nmatches);
    with point_kf (the keyframe number of every candidate) the second overload (vpPointsKFs / vpMatchedKF) and (owner, owner_kf[n], nmatches)."""
    c = lambda a, t: np.ascontiguousarray(a, t)  # noqa: E731
    f = np.float32
    kps, desc, mi, g, b, sf, cam4, Scw = c(kps, KP_DTYPE), c(desc, np.uint8), c(matched_in, np.int32), c(grid4, f), c(bounds4, f), c(scale_factors, f), c(cam4, f), c(np.asarray(Scw, f).reshape(16), f)
    arrs = [c(mp["bad"], np.uint8), c(mp["world"], f), c(mp["normal"], f), c(mp["min_dist"], f), c(mp["max_dist"], f), c(mp["level"], np.int32), c(mp["desc"], np.uint8)]
    owner = np.zeros(max(len(kps), 1), np.int32)
    pk = None if point_kf is None else c(point_kf, np.int32)
    okf = np.full(max(len(kps), 1), -1, np.int32)
    nm = lib().ref_search_by_projection_sim3(_ptr(kps), _ptr(desc), len(kps), _ptr(mi), _ptr(g), _ptr(b), _ptr(sf), len(sf), _ptr(cam4), _ptr(Scw), len(arrs[0]),
                                             *[_ptr(a) for a in arrs], int(th), float(ratio_hamming), _ptr(owner), _ptr(pk) if pk is not None else None,
                                             _ptr(okf) if pk is not None else None)
    if pk is not None:
        return owner[:len(kps)], okf[:len(kps)], nm
    return owner[:len(kps)], nm


def fuse_sim3(kps, desc, row_state, grid4, bounds4, scale_factors, cam4, Scw, mp, th):
    """The reference's own ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) on arrays.  Returns (fused_row[nmp], replaced[nmp], nFused)."""
    c = lambda a, t: np.ascontiguousarray(a, t)  # noqa: E731
    f = np.float32
    kps, desc, rs, g, b, sf, cam4, Scw = c(kps, KP_DTYPE), c(desc, np.uint8), c(row_state, np.int32), c(grid4, f), c(bounds4, f), c(scale_factors, f), c(cam4, f), c(np.asarray(Scw, f).reshape(16), f)
    arrs = [c(mp["bad"], np.uint8), c(mp["world"], f), c(mp["normal"], f), c(mp["min_dist"], f), c(mp["max_dist"], f), c(mp["level"], np.int32), c(mp["desc"], np.uint8)]
    nmp = len(arrs[0])
    row, rep = np.zeros(max(nmp, 1), np.int32), np.zeros(max(nmp, 1), np.int32)
    nf = lib().ref_fuse_sim3(_ptr(kps), _ptr(desc), len(kps), _ptr(rs), _ptr(g), _ptr(b), _ptr(sf), len(sf), _ptr(cam4), _ptr(Scw), nmp, *[_ptr(a) for a in arrs],
                             float(th), _ptr(row), _ptr(rep))
    return row[:nmp], rep[:nmp], nf


def search_by_sim3(kf1, kf2, pre12, grid4, bounds4, scale_factors, cam4, s12, R12, t12, th):
    """The reference's own ORBmatcher::SearchBySim3 on arrays; kf = dict(kps, desc, has_mp, bad, world, min_dist, max_dist, level).
    Returns (match12[n1], nFound)."""
    c = lambda a, t: np.ascontiguousarray(a, t)  # noqa: E731
    f = np.float32
    def side(k):
        return [c(k["kps"], KP_DTYPE), c(k["desc"], np.uint8), None, c(k["has_mp"], np.uint8), c(k["bad"], np.uint8), c(k["world"], f), c(k["min_dist"], f),
                c(k["max_dist"], f), c(k["level"], np.int32)]
    a, b = side(kf1), side(kf2)
    n1, n2 = len(a[0]), len(b[0])
    pre, g, bd, sf, cam4, R12, t12 = c(pre12, np.int32), c(grid4, f), c(bounds4, f), c(scale_factors, f), c(cam4, f), c(np.asarray(R12, f).reshape(9), f), c(t12, f)
    out = np.zeros(max(n1, 1), np.int32)
    pa = [_ptr(a[0]), _ptr(a[1]), n1] + [_ptr(v) for v in a[3:]]
    pb = [_ptr(b[0]), _ptr(b[1]), n2] + [_ptr(v) for v in b[3:]]
    nf = lib().ref_search_by_sim3(*pa, *pb, _ptr(pre), _ptr(g), _ptr(bd), _ptr(sf), len(sf), _ptr(cam4), float(s12), _ptr(R12), _ptr(t12), float(th), _ptr(out))
    return out[:n1], nf


def kb8_project(cam, p3):
    cam, p3 = np.ascontiguousarray(cam, np.float32), np.ascontiguousarray(p3, np.float32).reshape(-1, 3)
    uv = np.zeros((p3.shape[0], 2), np.float32)
    lib().ref_kb8_project(_ptr(cam), _ptr(p3), p3.shape[0], _ptr(uv))
    return uv


def kb8_unproject(cam, uv):
    cam, uv = np.ascontiguousarray(cam, np.float32), np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
    ray = np.zeros((uv.shape[0], 3), np.float32)
    lib().ref_kb8_unproject(_ptr(cam), _ptr(uv), uv.shape[0], _ptr(ray))
    return ray


def kb8_triangulate_matches(cam1, cam2, R12, t12, uv1, uv2, sigma1, sigma2):
    """The reference's own KannalaBrandt8::TriangulateMatches for n keypoint pairs: (depth[n] = return value, p3d[n, 3])."""
    f = lambda a: np.ascontiguousarray(a, np.float32)  # noqa: E731
    cam1, cam2, R12, t12, uv1, uv2, sigma1, sigma2 = f(cam1), f(cam2), f(R12), f(t12), f(uv1), f(uv2), f(sigma1), f(sigma2)
    n = uv1.shape[0]
    depth, p3d = np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
    lib().ref_kb8_triangulate_matches(_ptr(cam1), _ptr(cam2), _ptr(R12), _ptr(t12), _ptr(uv1), _ptr(uv2), _ptr(sigma1), _ptr(sigma2), n, _ptr(depth), _ptr(p3d))
    return depth, p3d


def search_for_triangulation(desc_1, kps_1, u_right_1, has_mp_1, n_left_1, fv_1, desc_2, kps_2, u_right_2, has_mp_2, n_left_2, fv_2, cams, rig, poses,
                             scale_factors, level_sigma2, only_stereo=False, coarse=False, check_orientation=True):
    """The reference's own ORBmatcher::SearchForTriangulation on arrays (oracle/ref_shim.cpp::ref_search_for_triangulation explains them).
    fv = (nodes, starts, feats) arrays.  Returns (vMatches12, nmatches, geometry): geometry = (F12[3, 3], ep[2]) for pinhole keyframes,
    (R[4, 3, 3], t[4, 3]) for a rig: the inputs of the library's entry point, evaluated by the reference's expressions."""
    c = lambda a, t: None if a is None else np.ascontiguousarray(a, t)  # noqa: E731
    p = lambda a: _ptr(a) if a is not None else None  # noqa: E731
    d1, d2, k1, k2 = c(desc_1, np.uint8), c(desc_2, np.uint8), c(kps_1, KP_DTYPE), c(kps_2, KP_DTYPE)
    u1, u2, m1, m2 = c(u_right_1, np.float32), c(u_right_2, np.float32), c(has_mp_1, np.uint8), c(has_mp_2, np.uint8)
    (n1, s1, f1), (n2, s2, f2) = [tuple(c(a, np.int32) for a in fv) for fv in (fv_1, fv_2)]
    cams, poses = c(np.asarray(cams, np.float32).reshape(32), np.float32), c(np.asarray(poses, np.float32).reshape(-1), np.float32)
    assert poses.shape[0] == (48 if rig else 24)
    if not rig:
        poses = np.concatenate([poses, np.zeros(24, np.float32)])
    sf, ls = c(scale_factors, np.float32), c(level_sigma2, np.float32)
    out = np.zeros(max(d1.shape[0], 1), np.int32)
    geom = np.zeros(48, np.float32)
    nm = lib().ref_search_for_triangulation(p(d1), p(k1), p(u1), p(m1), d1.shape[0], int(n_left_1), p(n1), p(s1), p(f1), len(n1),
                                            p(d2), p(k2), p(u2), p(m2), d2.shape[0], int(n_left_2), p(n2), p(s2), p(f2), len(n2),
                                            p(cams), int(bool(rig)), p(poses), p(sf), p(ls), len(sf), int(only_stereo), int(coarse), int(check_orientation),
                                            p(out), p(geom))
    geometry = (geom[:36].reshape(4, 3, 3).copy(), geom[36:48].reshape(4, 3).copy()) if rig else (geom[:9].reshape(3, 3).copy(), geom[9:11].copy())
    return out[:d1.shape[0]], nm, geometry


def distinctive_descriptor(desc, bad_kf=None):
    """The reference's own MapPoint::ComputeDistinctiveDescriptors for one map point: row of `desc` it selects (-1: none)."""
    desc = np.ascontiguousarray(desc, np.uint8)
    bad = None if bad_kf is None else np.ascontiguousarray(bad_kf, np.uint8)
    return lib().ref_distinctive_descriptor(_ptr(desc), desc.shape[0], _ptr(bad) if bad is not None else None)


def set_arena(on, fma=False):
    """on=False: plain malloc/free instead of the monotonic arena (to see what heap-address ties do)."""
    lib(fma).ref_set_arena(int(bool(on)))


def arena_peak(fma=False):
    return lib(fma).ref_arena_peak()


# ---- the reference's vendored DBoW2 (Thirdparty/DBoW2), oracle/ref_dbow_shim.cpp -------------------------------------------------
_dbow = None


def dbow_path():
    return os.path.join(_HERE, "_ref", "libdbow_ref.so")


def dbow_available():
    return os.path.exists(dbow_path()) or (build() is not None and os.path.exists(dbow_path()))


def dbow_lib():
    global _dbow
    if _dbow is None:
        if not dbow_available():
            raise RuntimeError("oracle/_ref/libdbow_ref.so is not built and %s is absent" % REF_SRC)
        L = C.CDLL(dbow_path())
        vp = C.c_void_p
        L.ref_voc_load_text.restype = vp
        L.ref_voc_load_text.argtypes = [C.c_char_p]
        L.ref_voc_destroy.argtypes = [vp]
        L.ref_voc_words.argtypes = [vp]
        L.ref_voc_transform.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, vp]
        L.ref_voc_score.restype = C.c_double
        L.ref_voc_score.argtypes = [vp, vp, vp, C.c_int, vp, vp, C.c_int]
        _dbow = L
    return _dbow


def write_vocabulary_text(path, k, L, scoring, weighting, parent, is_leaf, desc, weight):
    """The file ORBVocabulary::loadFromTextFile reads (TemplatedVocabulary.h:1338-1426): `k L scoring weighting`, then one line per
    node `parent isLeaf d0 .. d31 weight`; no newline after the last node (the reader would turn an empty last line into a node)."""
    lines = ["%d %d %d %d" % (k, L, scoring, weighting)]
    for p, leaf, d, w in zip(parent, is_leaf, desc, weight):
        lines.append("%d %d %s %s" % (int(p), int(leaf), " ".join(str(int(b)) for b in d), repr(float(w))))
    with open(path, "w") as f:
        f.write("\n".join(lines))


class ReferenceVocabulary:
    """ORB_SLAM3::ORBVocabulary (DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>) itself, loaded from a text file."""

    def __init__(self, path):
        self.L = dbow_lib()
        self.h = self.L.ref_voc_load_text(path.encode())
        if not self.h:
            raise ValueError("loadFromTextFile failed")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_voc_destroy(self.h)
            self.h = None

    def words(self):
        return self.L.ref_voc_words(self.h)

    def transform(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8)
        n = desc.shape[0]
        bw, bv = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        fn, fs, ff = np.zeros(max(n, 1), np.int32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nb, nf = C.c_int32(0), C.c_int32(0)
        self.L.ref_voc_transform(self.h, _ptr(desc), n, levelsup, _ptr(bw), _ptr(bv), C.byref(nb), _ptr(fn), _ptr(fs), _ptr(ff), C.byref(nf))
        nb, nf = nb.value, nf.value
        return {"bow_words": bw[:nb].copy(), "bow_values": bv[:nb].copy(), "feat_vec": {int(fn[e]): ff[fs[e]:fs[e + 1]].copy() for e in range(nf)}}
