"""Host-side mirror of ORB_SLAM3::ORBextractor (/root/reference/include/ORBextractor.h:43-109) over the C ABI.

Same constructor arguments, same call semantics (`mono_index = ex(image, keypoints, descriptors, lapping)`),
same getters; `mvImagePyramid` is served lazily from the device-resident pyramid.  Python is the test/bench
harness language here; the C++ drop-in shim with the reference's exact signatures is include/orbfe_shim/."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, OrbfeParams, check, ptr


class ORBextractor:
    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, *, width, height, device=0, max_batch=1):
        self._L = _lib.lib()
        self._h = C.c_void_p()
        p = OrbfeParams(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        check(self._L.orbfe_extractor_create(C.byref(p), device, width, height, max_batch, C.byref(self._h)))
        self.width, self.height, self.nlevels, self.max_batch, self.device = width, height, nlevels, max_batch, device
        self.capacity = self._L.orbfe_keypoint_capacity(self._h)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbfe_extractor_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    # ---- getters, include/ORBextractor.h:61-81
    def GetLevels(self):
        return self._L.orbfe_get_levels(self._h)

    def _factors(self, which):
        arrs = [np.zeros(self.nlevels, np.float32) for _ in range(4)]
        check(self._L.orbfe_get_scale_factors(self._h, *[ptr(a) for a in arrs]))
        return arrs[which]

    def GetScaleFactor(self):
        return float(self._factors(0)[1]) if self.nlevels > 1 else 1.0

    def GetScaleFactors(self):
        return self._factors(0)

    def GetInverseScaleFactors(self):
        return self._factors(1)

    def GetScaleSigmaSquares(self):
        return self._factors(2)

    def GetInverseScaleSigmaSquares(self):
        return self._factors(3)

    def features_per_level(self):
        q = np.zeros(self.nlevels, np.int32)
        check(self._L.orbfe_get_features_per_level(self._h, ptr(q)))
        return q

    def umax(self):
        u = np.zeros(16, np.int32)
        check(self._L.orbfe_get_umax(self._h, ptr(u)))
        return u

    def level_size(self, level):
        w, h = C.c_int(0), C.c_int(0)
        check(self._L.orbfe_level_size(self._h, level, C.byref(w), C.byref(h)))
        return w.value, h.value

    # ---- operator(), src/ORBextractor.cc:1068-1150
    def __call__(self, image, vLappingArea=(0, 0)):
        """Returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32]).  monoIndex == -1 for an empty image."""
        img = np.asarray(image)
        if img.size == 0:
            rc = self._L.orbfe_extract(self._h, None, 0, 0, 0, 0, 0, None, None, 0, None, None)
            assert rc == -1
            return -1, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert img.dtype == np.uint8 and img.ndim == 2, "CV_8UC1 expected (src/ORBextractor.cc:1076)"
        if img.strides[1] != 1:
            img = np.ascontiguousarray(img)
        kps = np.zeros(self.capacity, KP_DTYPE)
        desc = np.zeros((self.capacity, 32), np.uint8)
        n, mono = C.c_int(0), C.c_int(0)
        check(self._L.orbfe_extract(self._h, ptr(img), img.shape[0], img.shape[1], img.strides[0], int(vLappingArea[0]),
                                    int(vLappingArea[1]), ptr(kps), ptr(desc), self.capacity, C.byref(n), C.byref(mono)))
        return mono.value, kps[:n.value], desc[:n.value]

    # ---- batch entry points
    def extract_batch_host(self, images, vLappingArea=(0, 0), out=None):
        """images: [n, H, W] u8 host array.  Returns (counts, mono, kps[n,cap], desc[n,cap,32])."""
        imgs = np.ascontiguousarray(images, np.uint8)
        nf = imgs.shape[0]
        if out is None:
            out = (np.zeros(nf, np.int32), np.zeros(nf, np.int32), np.zeros((nf, self.capacity), KP_DTYPE),
                   np.zeros((nf, self.capacity, 32), np.uint8))
        cnt, mono, kps, desc = out
        check(self._L.orbfe_extract_batch_host(self._h, ptr(imgs), nf, int(vLappingArea[0]), int(vLappingArea[1]),
                                               ptr(kps), ptr(desc), self.capacity, ptr(cnt), ptr(mono)))
        return cnt, mono, kps, desc

    def extract_batch_device(self, d_images_ptr, nframes, vLappingArea=(0, 0), stream=None):
        check(self._L.orbfe_extract_batch_device(self._h, C.c_void_p(d_images_ptr), nframes, int(vLappingArea[0]),
                                                 int(vLappingArea[1]), C.c_void_p(stream) if stream else None))

    def batch_counts(self, nframes):
        n, m = np.zeros(nframes, np.int32), np.zeros(nframes, np.int32)
        check(self._L.orbfe_batch_counts(self._h, nframes, ptr(n), ptr(m)))
        return n, m

    def batch_result(self, frame):
        kps = np.zeros(self.capacity, KP_DTYPE)
        desc = np.zeros((self.capacity, 32), np.uint8)
        n, mono = C.c_int(0), C.c_int(0)
        check(self._L.orbfe_batch_result(self._h, frame, ptr(kps), ptr(desc), self.capacity, C.byref(n), C.byref(mono)))
        return mono.value, kps[:n.value], desc[:n.value]

    # ---- mvImagePyramid (include/ORBextractor.h:83) and stage views
    def pyramid_level(self, level, frame=0, padded=False):
        w, h = self.level_size(level)
        b = 19 if padded else 0
        out = np.zeros((h + 2 * b, w + 2 * b), np.uint8)
        check(self._L.orbfe_pyramid_level(self._h, frame, level, int(padded), ptr(out)))
        return out

    @property
    def mvImagePyramid(self):
        return [self.pyramid_level(l) for l in range(self.nlevels)]

    def blurred_level(self, level, frame=0):
        w, h = self.level_size(level)
        out = np.zeros((h, w), np.uint8)
        check(self._L.orbfe_blurred_level(self._h, frame, level, ptr(out)))
        return out

    def fast_candidates(self, level, frame=0):
        n = C.c_int(0)
        check(self._L.orbfe_fast_candidates(self._h, frame, level, None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 1), KP_DTYPE)
        check(self._L.orbfe_fast_candidates(self._h, frame, level, ptr(out), n.value, C.byref(n)))
        return out[:n.value]
