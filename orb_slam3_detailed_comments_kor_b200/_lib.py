"""ctypes binding of liborbfe.so (the C ABI declared in include/orbfe.h).

There is no CPU fallback: if the library is missing or no CUDA device is usable, calls raise."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liborbfe.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
MATCH_DTYPE = np.dtype([("best1", "<i4"), ("best2", "<i4"), ("idx", "<i4"), ("accepted", "<i4")])


class OrbfeParams(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32)]


class OrbfeError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("orbfe error %d: %s" % (code, text))
        self.code = code


_lib = None

# name -> (restype, argtypes); every symbol include/orbfe.h declares
vp, i32, i64, f32, f64, sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double, C.c_size_t
SIGNATURES = {
    "orbfe_last_error": (C.c_char_p, []),
    "orbfe_thread_release": (None, []),
    "orbfe_device_count": (i32, []),
    "orbfe_version": (C.c_char_p, []),
    "orbfe_extractor_create": (i32, [C.POINTER(OrbfeParams), i32, i32, i32, i32, C.POINTER(vp)]),
    "orbfe_extractor_destroy": (None, [vp]),
    "orbfe_get_levels": (i32, [vp]),
    "orbfe_get_scale_factors": (i32, [vp, vp, vp, vp, vp]),
    "orbfe_get_features_per_level": (i32, [vp, vp]),
    "orbfe_get_umax": (i32, [vp, vp]),
    "orbfe_keypoint_capacity": (i32, [vp]),
    "orbfe_level_size": (i32, [vp, i32, C.POINTER(i32), C.POINTER(i32)]),
    "orbfe_extract": (i32, [vp, vp, i32, i32, sz, i32, i32, vp, vp, i32, C.POINTER(i32), C.POINTER(i32)]),
    "orbfe_extract_batch_device": (i32, [vp, vp, i32, i32, i32, vp]),
    "orbfe_extract_batch_host": (i32, [vp, vp, i32, i32, i32, vp, vp, i32, vp, vp]),
    "orbfe_stereo_frames_batch_host": (i32, [vp, vp, vp, vp, i32, i32, i32, C.c_float, C.c_float, C.c_float, vp, vp, vp, vp, vp, vp, vp, vp, i32, vp, vp]),
    "orbfe_stereo_frames_submit": (i32, [vp, vp, vp, vp, i32, i32, i32, C.c_float, C.c_float, C.c_float, vp, vp, vp, vp, i32, vp, vp]),
    "orbfe_stereo_frames_wait": (i32, [vp, vp, vp, vp, vp, vp]),
    "orbfe_hamming_top2_sharded": (i32, [vp, vp, i32, vp, i32, i32, f32, i32, vp, vp]),
    "orbfe_nccl_unique_id": (i32, [vp]),
    "orbfe_nccl_comm_create": (i32, [i32, i32, i32, vp, C.POINTER(vp)]),
    "orbfe_nccl_comm_destroy": (i32, [vp]),
    "orbfe_batch_counts": (i32, [vp, i32, vp, vp]),
    "orbfe_batch_result": (i32, [vp, i32, vp, vp, i32, C.POINTER(i32), C.POINTER(i32)]),
    "orbfe_batch_device_views": (i32, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
    "orbfe_pyramid_level": (i32, [vp, i32, i32, i32, vp]),
    "orbfe_pyramid_levels": (i32, [vp, i32, i32, vp]),
    "orbfe_probe_popc": (i32, [i32, C.POINTER(C.c_double)]),
    "orbfe_fisheye_match_batch_device": (i32, [vp, vp, i32, C.c_double, vp, vp, vp]),
    "orbfe_fisheye_triangulate_batch_device": (i32, [vp, vp, i32, vp, vp, vp, vp, vp, vp]),
    "orbfe_kb8_triangulate_pairs_device": (i32, [i32, vp, vp, vp, vp, vp, i32, vp, vp, vp]),
    "orbfe_pin_host": (i32, [vp, sz]),
    "orbfe_unpin_host": (i32, [vp]),
    "orbfe_blurred_level": (i32, [vp, i32, i32, vp]),
    "orbfe_fast_candidates": (i32, [vp, i32, i32, vp, i32, C.POINTER(i32)]),
    "orbfe_stereo_match_batch_device": (i32, [vp, vp, i32, f32, f32, f32, vp]),
    "orbfe_stereo_frames_batch_device": (i32, [vp, vp, vp, vp, i32, i32, i32, f32, f32, f32, vp]),
    "orbfe_extract_pair_batch_device": (i32, [vp, vp, vp, vp, i32, i32, i32, vp]),
    "orbfe_stereo_result": (i32, [vp, i32, vp, vp, i32, C.POINTER(i32)]),
    "orbfe_stereo_result_batch": (i32, [vp, i32, vp, vp, i32]),
    "orbfe_set_profiling": (i32, [vp, i32]),
    "orbfe_stage_times": (i32, [vp, vp, C.POINTER(i64)]),
    "orbfe_stereo_device_views": (i32, [vp, C.POINTER(vp), C.POINTER(vp)]),
    "orbfe_descriptor_distances": (i32, [i32, vp, vp, i64, vp]),
    "orbfe_matcher_create": (i32, [i32, i32, i64, C.POINTER(vp)]),
    "orbfe_matcher_destroy": (None, [vp]),
    "orbfe_matcher_set_db": (i32, [vp, vp, i64, i64]),
    "orbfe_matcher_set_db_device": (i32, [vp, vp, i64, i64]),
    "orbfe_matcher_top2": (i32, [vp, vp, i32, i32, f32, i32, vp]),
    "orbfe_matcher_top2_device": (i32, [vp, vp, i32, i32, f32, i32, vp, vp]),
    "orbfe_top2_merge_device": (i32, [i32, vp, i32, i32, i32, f32, i32, vp, vp]),
    "orbfe_hamming_top2_static": (i32, [i32, vp, i32, vp, i64, i64, i32, f32, i32, vp, vp]),
    "orbfe_hamming_top2": (i32, [i32, vp, i32, vp, i64, i32, f32, i32, vp]),
    "orbfe_rectifier_create": (i32, [i32, i32, i32, vp, vp, C.POINTER(vp)]),
    "orbfe_rectifier_destroy": (None, [vp]),
    "orbfe_rectify_batch_device": (i32, [vp, vp, vp, i32, vp]),
    "orbfe_rectify": (i32, [vp, vp, sz, vp, sz]),
    "orbfe_to_gray_device": (i32, [i32, vp, i32, i32, i32, i32, i32, vp, vp]),
    "orbfe_vocabulary_create": (i32, [i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, i32, i32, C.POINTER(vp)]),
    "orbfe_vocabulary_destroy": (None, [vp]),
    "orbfe_vocabulary_words": (i32, [vp]),
    "orbfe_bow_transform_batch_device": (i32, [vp, vp, i32, vp, i32, i32, vp]),
    "orbfe_bow_result": (i32, [vp, i32, i32, vp, vp, vp, vp, C.POINTER(i32), vp, vp, vp, C.POINTER(i32)]),
    "orbfe_bow_transform": (i32, [vp, vp, i32, i32, vp, vp, vp, vp, C.POINTER(i32), vp, vp, vp, C.POINTER(i32)]),
    "orbfe_bow_device_views": (i32, [vp] + [C.POINTER(vp)] * 9),
    "orbfe_search_by_bow_device": (i32, [i32, i32, i32, vp, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, i32, C.c_float, i32, i32, vp, vp, vp, vp]),
    "orbfe_search_by_bow": (i32, [i32, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, i32, vp, i32, i32, C.c_float, i32, i32, vp, vp, C.POINTER(i32)]),
    "orbfe_search_for_triangulation_device": (i32, [i32, i32, i32, vp, vp, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp, vp, vp, vp]),
    "orbfe_search_for_triangulation_frames": (i32, [i32, vp, i32, vp, vp, i32, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, i32, vp, vp]),
    "orbfe_search_by_bow_frames": (i32, [i32, vp, i32, vp, vp, i32, i32, C.c_float, i32, i32, vp, i32, vp, vp]),
    "orbfe_search_for_triangulation": (i32, [i32, vp, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp, vp, vp, i32, i32, i32, i32,
                                             i32, vp, C.POINTER(i32), vp]),
    "orbfe_distinctive_descriptors_device": (i32, [i32, vp, vp, i32, vp, vp, vp]),
    "orbfe_distinctive_descriptors": (i32, [i32, vp, vp, i32, vp, vp]),
    "orbfe_search_by_projection_device": (i32, [i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, C.c_float, i32, vp, vp, vp, vp]),
    "orbfe_search_for_initialization_device": (i32, [i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, i32, C.c_float, i32, i32, vp, vp, vp, vp]),
    "orbfe_search_for_initialization": (i32, [i32, vp, vp, i32, vp, vp, i32, vp, vp, i32, C.c_float, i32, i32, vp, C.POINTER(i32)]),
    "orbfe_check_host_libm": (i32, [C.POINTER(i32)]),
    "orbfe_fisheye_stereo_frames": (i32, [vp, vp, i32, vp, C.c_double, vp, vp, vp, vp, i32]),
    "orbfe_fisheye_frames_submit": (i32, [vp, vp, vp, vp, i32, i32, i32, vp, C.c_double, vp, vp, vp, vp, i32, vp, vp, vp, vp]),
    "orbfe_debug_projection_rounds": (i32, [C.POINTER(i32), C.POINTER(i32)]),
    "orbfe_search_by_projection": (i32, [i32, vp, vp, i32, i32, vp, vp, vp, vp, vp, vp, i32, i32, C.c_float, i32, vp, vp, C.POINTER(i32)]),
    "orbfe_clahe_create": (i32, [i32, i32, i32, C.c_double, i32, i32, i32, C.POINTER(vp)]),
    "orbfe_clahe_destroy": (None, [vp]),
    "orbfe_clahe_apply_batch_device": (i32, [vp, vp, vp, i32, vp]),
    "orbfe_clahe_apply": (i32, [vp, vp, sz, vp, sz]),
    "orbfe_clahe_lut": (i32, [vp, i32, vp]),
    "orbfe_knn2_ratio": (i32, [i32, vp, i32, vp, i32, f64, vp, vp, vp]),
}


def lib():
    """Load liborbfe.so; raises if it has not been built (python -m orb_slam3_detailed_comments_kor_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OrbfeError(-3, "liborbfe.so is not built (%s); run __graft_entry__.build()" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc):
    if rc < 0:
        raise OrbfeError(rc, lib().orbfe_last_error().decode())
    return rc


def ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None
