"""Host-side mirror of the projection-gated search over the C ABI:
ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints)
(/root/reference/src/ORBmatcher.cc:44-214) and SearchByProjection(CurrentFrame, LastFrame, th, bMono) (:2193-2419), with Frame::AssignFeaturesToGrid / GetFeaturesInArea
(/root/reference/src/Frame.cc:380-411, 643-709) inside."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, check, ptr

PROJ_QUERY_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("radius", "<f4"), ("xr", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"),
                             ("flags", "<i4"), ("angle", "<f4")])
PROJ_MATCH_DTYPE = np.dtype([("best_idx", "<i4"), ("best_dist", "<i4"), ("best_level", "<i4"), ("second_dist", "<i4"),
                             ("second_level", "<i4"), ("accepted", "<i4"), ("status", "<i4")])
PQ_RIGHT, PQ_BLOCKS, PQ_SKIP, PQ_AFTER_LEFT, PQ_BEST_ONLY, PQ_NEEDS_LEFT_WINDOW, PQ_ERROR_GATE = 1, 2, 4, 8, 16, 32, 64
(PS_NO_CANDIDATES, PS_ACCEPTED, PS_OVER_THRESHOLD, PS_RATIO_REJECTED, PS_SKIPPED, PS_CAPACITY, PS_DROPPED, PS_ROTATION,
 PS_EMPTY_WINDOW) = range(9)
TH_HIGH = 100   # ORBmatcher.cc:36
FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48   # include/Frame.h:37-38


def radius_by_viewing_cos(view_cos):
    """ORBmatcher::RadiusByViewingCos (src/ORBmatcher.cc:216-222)"""
    return 2.5 if view_cos > 0.998 else 4.0


def frame_grid(min_x, max_x, min_y, max_y):
    """(mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv) as the Frame constructors set them (src/Frame.cc:158-159)"""
    f32 = np.float32
    return np.array([min_x, min_y, f32(FRAME_GRID_COLS) / (f32(max_x) - f32(min_x)), f32(FRAME_GRID_ROWS) / (f32(max_y) - f32(min_y))], np.float32)


def search_by_projection(kps, desc, queries, qdesc, grid, n_left=-1, u_right=None, partner=None, taken=None, th_high=TH_HIGH, nnratio=0.8,
                         check_orientation=False, inv_level_sigma2=None, device=0):
    """One frame, host arrays.  Returns (matches[PROJ_MATCH_DTYPE], taken_after u8[n], nmatches)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc = np.ascontiguousarray(desc, np.uint8)
    queries = np.ascontiguousarray(queries, PROJ_QUERY_DTYPE)
    qdesc = np.ascontiguousarray(qdesc, np.uint8)
    grid = np.ascontiguousarray(grid, np.float32)
    n, nq = kps.shape[0], queries.shape[0]
    assert desc.shape == (n, 32) and qdesc.shape == (nq, 32) and grid.shape == (4,)
    ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
    pt = None if partner is None else np.ascontiguousarray(partner, np.int32)
    tk = np.zeros(n, np.uint8) if taken is None else np.array(taken, np.uint8)
    out = np.zeros(nq, PROJ_MATCH_DTYPE)
    sig = None if inv_level_sigma2 is None else np.ascontiguousarray(np.resize(np.asarray(inv_level_sigma2, np.float32), 16))
    nm = C.c_int(0)
    check(_lib.lib().orbfe_search_by_projection(device, ptr(kps), ptr(desc), n, n_left, ptr(ur) if ur is not None else None,
                                                ptr(pt) if pt is not None else None, ptr(tk), ptr(grid), ptr(queries), ptr(qdesc), nq,
                                                th_high, C.c_float(nnratio), int(check_orientation), ptr(sig) if sig is not None else None, ptr(out),
                                                C.byref(nm)))
    return out, tk, nm.value


def search_by_projection_device(nframes, max_keypoints, d_kps, d_desc, d_kp_offsets, d_grids, d_queries, d_qdesc, d_q_offsets, d_out,
                                d_n_left=None, d_u_right=None, d_partner=None, d_taken_in=None, d_taken_out=None, th_high=TH_HIGH,
                                nnratio=0.8, check_orientation=False, d_inv_level_sigma2=None, device=0, stream=None):
    """A batch of frames, device pointers (ints), asynchronous on `stream`."""
    vp = lambda p: C.c_void_p(p) if p else None  # noqa: E731
    check(_lib.lib().orbfe_search_by_projection_device(device, nframes, max_keypoints, vp(d_kps), vp(d_desc), vp(d_kp_offsets), vp(d_n_left),
                                                       vp(d_u_right), vp(d_partner), vp(d_taken_in), vp(d_grids), vp(d_queries), vp(d_qdesc),
                                                       vp(d_q_offsets), th_high, C.c_float(nnratio), int(check_orientation), vp(d_inv_level_sigma2), vp(d_out),
                                                       vp(d_taken_out),
                                                       vp(stream)))


def search_for_initialization(kps1, desc1, kps2, desc2, grid, prev_matched, window=100, nnratio=0.9, check_orientation=True, th_low=50, device=0):
    """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:706-821), one pair, host arrays.
    Returns (vnMatches12, vbPrevMatched afterwards, nmatches)."""
    kps1, kps2 = np.ascontiguousarray(kps1, KP_DTYPE), np.ascontiguousarray(kps2, KP_DTYPE)
    desc1, desc2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    grid = np.ascontiguousarray(grid, np.float32)
    n1, n2 = kps1.shape[0], kps2.shape[0]
    assert desc1.shape == (n1, 32) and desc2.shape == (n2, 32) and grid.shape == (4,)
    prev = np.array(prev_matched, np.float32).reshape(n1, 2).copy()
    m12 = np.full(n1, -1, np.int32)
    nm = C.c_int(0)
    check(_lib.lib().orbfe_search_for_initialization(device, ptr(kps1), ptr(desc1), n1, ptr(kps2), ptr(desc2), n2, ptr(grid), ptr(prev), int(window),
                                                     C.c_float(nnratio), int(check_orientation), int(th_low), ptr(m12), C.byref(nm)))
    return m12, prev, nm.value


def search_for_initialization_device(npairs, d_kps1, d_desc1, d_off1, d_kps2, d_desc2, d_off2, d_grids, d_prev_matched, d_matches12, d_nmatches,
                                     window=100, nnratio=0.9, check_orientation=True, th_low=50, d_rounds=None, device=0, stream=None):
    """A batch of frame pairs, device pointers (ints), asynchronous on `stream`."""
    vp = lambda p: C.c_void_p(p) if p else None  # noqa: E731
    check(_lib.lib().orbfe_search_for_initialization_device(device, npairs, vp(d_kps1), vp(d_desc1), vp(d_off1), vp(d_kps2), vp(d_desc2), vp(d_off2),
                                                            vp(d_grids), vp(d_prev_matched), int(window), C.c_float(nnratio), int(check_orientation),
                                                            int(th_low), vp(d_matches12), vp(d_nmatches), vp(d_rounds), vp(stream)))
