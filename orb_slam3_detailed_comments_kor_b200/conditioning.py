"""Input conditioning in front of the extractor, over the C ABI: stereo rectification (cv::remap with the CV_32F maps of
cv::initUndistortRectifyMap, /root/reference/Examples/Stereo-Inertial/stereo_inertial_euroc.cc:150-151,201-202) and the
colour-to-gray conversion of Tracking::GrabImage* (/root/reference/src/Tracking.cc:1303-1328), and the CLAHE of the TUM-VI
drivers (/root/reference/Examples/Stereo-Inertial/stereo_inertial_tum_vi.cc:128-138)."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check, ptr


class Rectifier:
    """cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) for fixed maps M1 (x) and M2 (y), both [H, W] float32."""

    def __init__(self, map_x, map_y, device=0):
        mx = np.ascontiguousarray(map_x, np.float32)
        my = np.ascontiguousarray(map_y, np.float32)
        assert mx.shape == my.shape and mx.ndim == 2
        self.height, self.width = mx.shape
        self._L = _lib.lib()
        self._h = C.c_void_p()
        check(self._L.orbfe_rectifier_create(device, self.width, self.height, ptr(mx), ptr(my), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbfe_rectifier_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __call__(self, image):
        img = np.ascontiguousarray(image, np.uint8)
        assert img.shape == (self.height, self.width)
        out = np.empty_like(img)
        check(self._L.orbfe_rectify(self._h, ptr(img), img.strides[0], ptr(out), out.strides[0]))
        return out

    def batch_device(self, d_src, d_dst, nframes, stream=None):
        check(self._L.orbfe_rectify_batch_device(self._h, C.c_void_p(d_src), C.c_void_p(d_dst), nframes,
                                                 C.c_void_p(stream) if stream else None))


class CLAHE:
    """cv::createCLAHE(clip_limit, cv::Size(tiles[0], tiles[1])) for width x height 8-bit images; tiles = (across, down)."""

    def __init__(self, width, height, clip_limit=3.0, tiles=(8, 8), max_batch=1, device=0):
        self.width, self.height, self.tiles = width, height, tuple(tiles)
        self._L = _lib.lib()
        self._h = C.c_void_p()
        check(self._L.orbfe_clahe_create(device, width, height, float(clip_limit), tiles[0], tiles[1], max_batch, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbfe_clahe_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def apply(self, image):
        img = np.ascontiguousarray(image, np.uint8)
        assert img.shape == (self.height, self.width)
        out = np.empty_like(img)
        check(self._L.orbfe_clahe_apply(self._h, ptr(img), img.strides[0], ptr(out), out.strides[0]))
        return out

    def batch_device(self, d_src, d_dst, nframes, stream=None):
        check(self._L.orbfe_clahe_apply_batch_device(self._h, C.c_void_p(d_src), C.c_void_p(d_dst), nframes,
                                                     C.c_void_p(stream) if stream else None))

    def lut(self, frame=0):
        out = np.empty((self.tiles[1], self.tiles[0], 256), np.uint8)
        check(self._L.orbfe_clahe_lut(self._h, frame, ptr(out)))
        return out


def to_gray_device(d_src, channels, rgb_order, width, height, nframes, d_dst, device=0, stream=None):
    check(_lib.lib().orbfe_to_gray_device(device, C.c_void_p(d_src), channels, int(rgb_order), width, height, nframes,
                                          C.c_void_p(d_dst), C.c_void_p(stream) if stream else None))
