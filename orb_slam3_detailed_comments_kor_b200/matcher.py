"""Host-side mirror of the descriptor-matching core of ORB_SLAM3::ORBmatcher and Frame over the C ABI.

Reference surface: ORBmatcher(nnratio, checkOri), TH_LOW/TH_HIGH, static DescriptorDistance
(/root/reference/include/ORBmatcher.h:35-111, src/ORBmatcher.cc:36-42, 2591-2607); the top-2 core of SearchByBoW
(src/ORBmatcher.cc:311-375, 870-908); Frame::ComputeStereoMatches / ComputeStereoFishEyeMatches
(src/Frame.cc:797-967, 1119-1159)."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import MATCH_DTYPE, check, ptr


class DescriptorDatabase:
    """A shard of map-point descriptors resident on one GPU (rows index_base .. index_base+rows)."""

    def __init__(self, max_rows, max_queries=4096, device=0):
        self._L = _lib.lib()
        self._h = C.c_void_p()
        check(self._L.orbfe_matcher_create(device, max_queries, max_rows, C.byref(self._h)))
        self.device, self.max_queries = device, max_queries

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbfe_matcher_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set(self, db, index_base=0):
        db = np.ascontiguousarray(db, np.uint8)
        check(self._L.orbfe_matcher_set_db(self._h, ptr(db), db.shape[0], index_base))

    def set_device(self, d_ptr, rows, index_base=0):
        check(self._L.orbfe_matcher_set_db_device(self._h, C.c_void_p(d_ptr), rows, index_base))

    def top2(self, queries, th, ratio, strict_lt):
        q = np.ascontiguousarray(queries, np.uint8)
        out = np.zeros(q.shape[0], MATCH_DTYPE)
        check(self._L.orbfe_matcher_top2(self._h, ptr(q), q.shape[0], th, ratio, int(strict_lt), ptr(out)))
        return out

    def top2_device(self, d_q, nq, th, ratio, strict_lt, d_out, stream=None):
        check(self._L.orbfe_matcher_top2_device(self._h, C.c_void_p(d_q), nq, th, ratio, int(strict_lt), C.c_void_p(d_out),
                                                C.c_void_p(stream) if stream else None))


    def top2_sharded(self, comm, nranks, d_q, nq, th, ratio, strict_lt, d_out, stream=None):
        """orbfe_hamming_top2_sharded: this shard's search, ncclAllGather of every rank's records and the merge, all on `stream`."""
        check(self._L.orbfe_hamming_top2_sharded(self._h, C.c_void_p(comm) if comm else None, nranks, C.c_void_p(d_q), nq, th, ratio, int(strict_lt),
                                                 C.c_void_p(d_out), C.c_void_p(stream) if stream else None))


class ORBmatcher:
    TH_HIGH = 100       # src/ORBmatcher.cc:36
    TH_LOW = 50         # src/ORBmatcher.cc:37
    HISTO_LENGTH = 30   # src/ORBmatcher.cc:38

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self.mfNNratio = float(nnratio)
        self.mbCheckOrientation = bool(checkOri)
        self.device = device

    @staticmethod
    def DescriptorDistance(a, b, device=0):
        """Hamming distance of 32-byte rows; a, b may be single rows or [n,32] batches (GPU, src/ORBmatcher.cc:2591)."""
        a2 = np.ascontiguousarray(np.atleast_2d(a), np.uint8)
        b2 = np.ascontiguousarray(np.atleast_2d(b), np.uint8)
        out = np.zeros(a2.shape[0], np.int32)
        check(_lib.lib().orbfe_descriptor_distances(device, ptr(a2), ptr(b2), a2.shape[0], ptr(out)))
        return int(out[0]) if np.ndim(a) == 1 else out

    def search_top2(self, queries, database, kf_kf=False):
        """Brute-force top-2 of SearchByBoW: kf_kf=False -> `best1 <= TH_LOW` (KF-Frame), True -> `<` (KF-KF)."""
        q = np.ascontiguousarray(queries, np.uint8)
        db = np.ascontiguousarray(database, np.uint8)
        out = np.zeros(q.shape[0], MATCH_DTYPE)
        check(_lib.lib().orbfe_hamming_top2(self.device, ptr(q), q.shape[0], ptr(db), db.shape[0], self.TH_LOW,
                                            self.mfNNratio, int(kf_kf), ptr(out)))
        return out


def merge_top2_device(device, d_parts, nshards, nq, th, ratio, strict_lt, d_out, stream=None):
    check(_lib.lib().orbfe_top2_merge_device(device, C.c_void_p(d_parts), nshards, nq, th, ratio, int(strict_lt),
                                             C.c_void_p(d_out), C.c_void_p(stream) if stream else None))


def compute_stereo_matches(exL, exR, nframes, mbf, minD, maxD, stream=None):
    """Frame::ComputeStereoMatches on the frames exL/exR extracted last (device resident)."""
    check(_lib.lib().orbfe_stereo_match_batch_device(exL.handle, exR.handle, nframes, mbf, minD, maxD,
                                                     C.c_void_p(stream) if stream else None))


def stereo_frames_batch_device(exL, exR, d_left_ptr, d_right_ptr, nframes, mbf, minD, maxD, vLappingArea=(0, 0), stream=None):
    """Frame's stereo constructor for a batch of device-resident pairs (src/Frame.cc:104-170): left and right extraction side by
    side on the two objects' streams, ComputeStereoMatches behind them; ordered after and before the caller's `stream`."""
    check(_lib.lib().orbfe_stereo_frames_batch_device(exL.handle, exR.handle, C.c_void_p(d_left_ptr), C.c_void_p(d_right_ptr), nframes,
                                                      int(vLappingArea[0]), int(vLappingArea[1]), mbf, minD, maxD, C.c_void_p(stream) if stream else None))


def extract_pair_batch_device(exL, exR, d_left_ptr, d_right_ptr, nframes, vLappingArea=(0, 0), stream=None):
    """The two ExtractORB threads of Frame's rig constructor (src/Frame.cc:1051-1054) for a batch of device-resident pairs."""
    check(_lib.lib().orbfe_extract_pair_batch_device(exL.handle, exR.handle, C.c_void_p(d_left_ptr), C.c_void_p(d_right_ptr), nframes,
                                                     int(vLappingArea[0]), int(vLappingArea[1]), C.c_void_p(stream) if stream else None))


def stereo_frames_batch_host(exL, exR, left, right, mbf, minD, maxD, vLappingArea=(0, 0), out=None):
    """Frame's stereo constructor for a batch of pairs (src/Frame.cc:104-178): both extractions and ComputeStereoMatches,
    pipelined over chunks inside one call.  left / right: [n, H, W] u8 host arrays.  Returns
    (cntL, monoL, kpsL, descL, cntR, monoR, kpsR, descR, uRight[n, cap], depth[n, cap])."""
    from ._lib import KP_DTYPE
    L = np.ascontiguousarray(left, np.uint8); R = np.ascontiguousarray(right, np.uint8)
    nf, cap = L.shape[0], exL.capacity
    if out is None:
        side = lambda: (np.zeros(nf, np.int32), np.zeros(nf, np.int32), np.zeros((nf, cap), KP_DTYPE), np.zeros((nf, cap, 32), np.uint8))  # noqa: E731
        out = side() + side() + (np.zeros((nf, cap), np.float32), np.zeros((nf, cap), np.float32))
    cl, ml, kl, dl, cr, mr, kr, dr, u, d = out
    check(_lib.lib().orbfe_stereo_frames_batch_host(exL.handle, exR.handle, ptr(L), ptr(R), nf, int(vLappingArea[0]), int(vLappingArea[1]),
                                                    C.c_float(mbf), C.c_float(minD), C.c_float(maxD), ptr(kl), ptr(dl), ptr(cl), ptr(ml),
                                                    ptr(kr), ptr(dr), ptr(cr), ptr(mr), cap, ptr(u), ptr(d)))
    return out


def stereo_frames_submit(exL, exR, left, right, mbf, minD, maxD, vLappingArea, out):
    """orbfe_stereo_frames_submit: enqueue one batch and return (at most two in flight).  left / right / every array of `out` (the
    tuple stereo_frames_batch_host returns) must stay alive and untouched until the matching stereo_frames_wait; pinned memory
    makes the copies asynchronous."""
    nf, cap = left.shape[0], exL.capacity
    cl, ml, kl, dl, cr, mr, kr, dr, u, d = out
    check(_lib.lib().orbfe_stereo_frames_submit(exL.handle, exR.handle, ptr(left), ptr(right), nf, int(vLappingArea[0]), int(vLappingArea[1]),
                                                C.c_float(mbf), C.c_float(minD), C.c_float(maxD), ptr(kl), ptr(dl), ptr(kr), ptr(dr), cap, ptr(u), ptr(d)))


def fisheye_stereo_frames(exL, exR, nframes, rig, ratio=0.7):
    """Frame::ComputeStereoFishEyeMatches for the frames both extractors processed last; returns host arrays
    (match_lr[n, cap], match_rl[n, cap], depth[n, cap], p3d[n, cap, 3])."""
    cap = exL.capacity
    lr, rl = np.full((nframes, cap), -1, np.int32), np.full((nframes, cap), -1, np.int32)
    d, p3 = np.full((nframes, cap), -1, np.float32), np.zeros((nframes, cap, 3), np.float32)
    rig = np.ascontiguousarray(rig, np.float32)
    check(_lib.lib().orbfe_fisheye_stereo_frames(exL.handle, exR.handle, nframes, ptr(rig), C.c_double(ratio), ptr(lr), ptr(rl), ptr(d), ptr(p3), cap))
    return lr, rl, d, p3


def fisheye_frames_submit(exL, exR, left, right, rig, ratio, vLappingArea, out):
    """orbfe_fisheye_frames_submit: like stereo_frames_submit for a fisheye rig.  `out` = (cntL, monoL, kpsL, descL, cntR, monoR, kpsR,
    descR, match_lr[n, cap] i32, match_rl[n, cap] i32, depth[n, cap] f32, p3d[n, cap, 3] f32); collect with fisheye_frames_wait."""
    nf, cap = left.shape[0], exL.capacity
    cl, ml, kl, dl, cr, mr, kr, dr, lr, rl, d, p3 = out
    rig = np.ascontiguousarray(rig, np.float32)
    check(_lib.lib().orbfe_fisheye_frames_submit(exL.handle, exR.handle, ptr(left), ptr(right), nf, int(vLappingArea[0]), int(vLappingArea[1]), ptr(rig),
                                                 C.c_double(ratio), ptr(kl), ptr(dl), ptr(kr), ptr(dr), cap, ptr(lr), ptr(rl), ptr(d), ptr(p3)))


def fisheye_frames_wait(exL, exR, out):
    check(_lib.lib().orbfe_stereo_frames_wait(exL.handle, exR.handle, ptr(out[0]), ptr(out[1]), ptr(out[4]), ptr(out[5])))
    return out


def stereo_frames_wait(exL, exR, out):
    """orbfe_stereo_frames_wait: block until the oldest submitted batch is on the host; fills the count arrays of `out`."""
    cl, ml, kl, dl, cr, mr, kr, dr, u, d = out
    check(_lib.lib().orbfe_stereo_frames_wait(exL.handle, exR.handle, ptr(cl), ptr(ml), ptr(cr), ptr(mr)))
    return out


def fisheye_match_batch_device(exL, exR, nframes, ratio, d_match_lr, d_dist01=None, stream=None):
    """Matching step of Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1119-1144) for the frames exL / exR extracted last."""
    check(_lib.lib().orbfe_fisheye_match_batch_device(exL.handle, exR.handle, nframes, float(ratio), C.c_void_p(d_match_lr),
                                                      C.c_void_p(d_dist01) if d_dist01 else None, C.c_void_p(stream) if stream else None))


def fisheye_triangulate_batch_device(exL, exR, nframes, cam1, cam2, R12, t12, d_match_lr, d_match_rl, d_depth, d_p3d, stream=None):
    """KannalaBrandt8::TriangulateMatches + the depth test of Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1145-1157) on the
    matches fisheye_match_batch_device produced; cam1 / cam2 = mvParameters (8 floats), R12 / t12 = mRlr, mtlr."""
    rig = np.concatenate([np.asarray(cam1, np.float32).reshape(8), np.asarray(cam2, np.float32).reshape(8), np.asarray(R12, np.float32).reshape(9),
                          np.asarray(t12, np.float32).reshape(3)]).astype(np.float32)
    check(_lib.lib().orbfe_fisheye_triangulate_batch_device(exL.handle, exR.handle, nframes, ptr(rig), C.c_void_p(d_match_lr), C.c_void_p(d_match_rl),
                                                            C.c_void_p(d_depth), C.c_void_p(d_p3d), C.c_void_p(stream) if stream else None))


def make_fisheye_rig(cam1, cam2, R12, t12):
    return np.concatenate([np.asarray(cam1, np.float32).reshape(8), np.asarray(cam2, np.float32).reshape(8), np.asarray(R12, np.float32).reshape(9),
                           np.asarray(t12, np.float32).reshape(3)]).astype(np.float32)


def kb8_triangulate_pairs_device(rig, d_uv1, d_uv2, d_sigma1, d_sigma2, n, d_depth, d_p3d, device=0, stream=None):
    check(_lib.lib().orbfe_kb8_triangulate_pairs_device(device, ptr(rig), C.c_void_p(d_uv1), C.c_void_p(d_uv2), C.c_void_p(d_sigma1), C.c_void_p(d_sigma2), n,
                                                        C.c_void_p(d_depth), C.c_void_p(d_p3d), C.c_void_p(stream) if stream else None))


def stereo_result(exL, frame):
    u = np.zeros(exL.capacity, np.float32)
    d = np.zeros(exL.capacity, np.float32)
    n = C.c_int(0)
    check(_lib.lib().orbfe_stereo_result(exL.handle, frame, ptr(u), ptr(d), exL.capacity, C.byref(n)))
    return u[:n.value], d[:n.value]


def knn2_ratio(queries, train, ratio=0.7, device=0):
    """BFMatcher(NORM_HAMMING).knnMatch(k=2) + Lowe ratio of ComputeStereoFishEyeMatches (src/Frame.cc:1137-1144)."""
    q = np.ascontiguousarray(queries, np.uint8)
    t = np.ascontiguousarray(train, np.uint8)
    m = np.zeros(q.shape[0], np.int32); d0 = m.copy(); d1 = m.copy()
    check(_lib.lib().orbfe_knn2_ratio(device, ptr(q), q.shape[0], ptr(t), t.shape[0], float(ratio), ptr(m), ptr(d0), ptr(d1)))
    return m, d0, d1


def distinctive_descriptors(desc, offsets, device=0):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:349-423) for many map points: point p owns rows
    offsets[p]:offsets[p+1] of desc.  Returns (best_idx, best_median) per point (best_idx relative to the point's first row)."""
    desc = np.ascontiguousarray(desc, np.uint8)
    off = np.ascontiguousarray(offsets, np.int32)
    npts = off.shape[0] - 1
    idx = np.zeros(npts, np.int32)
    med = np.zeros(npts, np.int32)
    check(_lib.lib().orbfe_distinctive_descriptors(device, ptr(desc), ptr(off), npts, ptr(idx), ptr(med)))
    return idx, med
