"""Deterministic synthetic inputs.  Thin ctypes wrapper over synth.cpp (host-only C++)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libsynth.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "synth.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(src) > os.path.getmtime(_SO):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", _SO, src])
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orbfe_synth_stereo.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        _lib.orbfe_synth_descriptors.argtypes = [C.c_uint64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int]
    return _lib


def stereo_pair(frame_index, width, height, out_left=None, out_right=None):
    """(left, right) u8 images of one synthetic stereo frame."""
    L = out_left if out_left is not None else np.empty((height, width), np.uint8)
    R = out_right if out_right is not None else np.empty((height, width), np.uint8)
    rc = _load().orbfe_synth_stereo(frame_index, width, height, L.ctypes.data, R.ctypes.data)
    if rc:
        raise ValueError("synth_stereo failed")
    return L, R


def frame(frame_index, width, height):
    L = np.empty((height, width), np.uint8)
    rc = _load().orbfe_synth_stereo(frame_index, width, height, L.ctypes.data, None)
    if rc:
        raise ValueError("synth failed")
    return L


def descriptors(seed, ndb, nq):
    """(db[ndb,32], q[nq,32]) u8 for the top-2 matching configs."""
    db = np.empty((ndb, 32), np.uint8)
    q = np.empty((nq, 32), np.uint8)
    rc = _load().orbfe_synth_descriptors(seed, db.ctypes.data, ndb, q.ctypes.data, nq)
    if rc:
        raise ValueError("synth_descriptors failed")
    return db, q
