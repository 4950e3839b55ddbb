"""ctypes binding of libkfslam_frontend.so (include/kfslam_frontend.h).  No CPU fallback: if the shared
library is missing this raises, and every compute entry point fails loudly without a CUDA device."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# KFS_LIB selects another build of the same library (kernel-variant experiments); the default is the in-tree product
LIB_PATH = os.environ.get("KFS_LIB") or os.path.join(HERE, "libkfslam_frontend.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
DM_DTYPE = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])
assert KP_DTYPE.itemsize == 28 and DM_DTYPE.itemsize == 16

KFS_OK, KFS_ERR_INVALID_ARG, KFS_ERR_UNSUPPORTED, KFS_ERR_CAPACITY, KFS_ERR_CUDA, KFS_ERR_NO_DEVICE = range(6)
KFS_MAX_WINDOW = 64  # include/kfslam_frontend.h
STATUS_NAMES = ["KFS_OK", "KFS_ERR_INVALID_ARG", "KFS_ERR_UNSUPPORTED", "KFS_ERR_CAPACITY", "KFS_ERR_CUDA", "KFS_ERR_NO_DEVICE"]
STAGE_NAMES = ["ingest", "pyramid", "fast", "select_harris", "select_sort", "blur", "describe"]
NORM_HAMMING = 6
ORB_HARRIS_SCORE, ORB_FAST_SCORE = 0, 1


class OrbParams(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("edge_threshold", C.c_int32), ("first_level", C.c_int32), ("wta_k", C.c_int32),
                ("score_type", C.c_int32), ("patch_size", C.c_int32), ("fast_threshold", C.c_int32)]


class BfPair(C.Structure):
    _fields_ = [("query", C.c_void_p), ("nq", C.c_int32), ("d_nq", C.c_void_p),
                ("train", C.c_void_p), ("nt", C.c_int32), ("d_nt", C.c_void_p),
                ("out", C.c_void_p), ("counts", C.c_void_p)]


class ReplayParams(C.Structure):
    _fields_ = [("orb", OrbParams), ("max_batch", C.c_int32), ("keyframe_slots", C.c_int32), ("ratio", C.c_float),
                ("chunk_frames", C.c_int32), ("arenas", C.c_int32)]


class ReplayFrame(C.Structure):
    _fields_ = [("keyframe", C.c_int32), ("store_slot", C.c_int32), ("flags", C.c_int32)]


class ReplayIO(C.Structure):
    _fields_ = [("gray", C.c_void_p), ("step", C.c_size_t), ("frame_stride", C.c_size_t), ("w", C.c_int32), ("h", C.c_int32),
                ("kp", C.c_void_p), ("desc", C.c_void_p), ("capacity", C.c_int32), ("n_kp", C.c_void_p),
                ("knn", C.c_void_p), ("n_query", C.c_void_p), ("good", C.c_void_p), ("n_good", C.c_void_p),
                ("mask", C.c_void_p), ("mask_step", C.c_size_t), ("mask_frame_stride", C.c_size_t), ("channels", C.c_int32)]


REPLAY_NO_MATCH = -2**31
REPLAY_SKIP_ROWS = 1

# name -> (restype, argtypes); this table is also what tests/test_abi.py checks against the header.
_vp, _i, _sz, _f = C.c_void_p, C.c_int, C.c_size_t, C.c_float
_ip = C.POINTER(C.c_int)
SIGNATURES = {
    "kfs_version": (_i, []),
    "kfs_last_error": (C.c_char_p, []),
    "kfs_device_count": (_i, []),
    "kfs_orb_default_params": (None, [C.POINTER(OrbParams)]),
    "kfs_orb_create": (_i, [C.POINTER(OrbParams), _i, C.POINTER(_vp)]),
    "kfs_orb_destroy": (None, [_vp]),
    "kfs_orb_capacity": (_i, [_vp]),
    "kfs_orb_set_capacity": (_i, [_vp, _i]),
    "kfs_orb_detect_and_compute": (_i, [_vp, _vp, _sz, _i, _i, _vp, _sz, _vp, _vp, _i, _ip]),
    "kfs_orb_detect_and_compute_batch": (_i, [_vp, _vp, _sz, _sz, _i, _i, _i, _vp, _sz, _sz, _vp, _vp, _i, _ip]),
    "kfs_orb_extract_device": (_i, [_vp, _vp, _sz, _sz, _i, _i, _i, _vp, _sz, _sz]),
    "kfs_orb_device_results": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), _ip, _ip]),
    "kfs_orb_fetch_results": (_i, [_vp, _vp, _vp, _i, _ip]),
    "kfs_orb_sync": (_i, [_vp]),
    "kfs_orb_set_stream": (_i, [_vp, _vp]),
    "kfs_orb_get_stream": (_vp, [_vp]),
    "kfs_orb_set_pipeline_chunk": (_i, [_vp, _i]),
    "kfs_orb_level_info": (_i, [_vp, _i, _ip, _ip, _ip, C.POINTER(_f)]),
    "kfs_orb_download_level": (_i, [_vp, _i, _i, _vp, _sz]),
    "kfs_orb_set_blur_fma": (_i, [_vp, _i]),
    "kfs_orb_download_blurred_level": (_i, [_vp, _i, _i, _vp, _sz]),
    "kfs_orb_download_fast_corners": (_i, [_vp, _i, _i, _vp, _i, _ip]),
    "kfs_orb_set_profiling": (_i, [_vp, _i]),
    "kfs_orb_stage_times": (_i, [_vp, C.POINTER(_f)]),
    "kfs_orb_last_launches": (_i, [_vp]),
    "kfs_bf_create": (_i, [_i, _i, _i, C.POINTER(_vp)]),
    "kfs_bf_destroy": (None, [_vp]),
    "kfs_bf_knn_match": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp]),
    "kfs_bf_knn_match_pairs": (_i, [_vp, C.POINTER(BfPair), _i, _i]),
    "kfs_bf_knn_match_pairs_device": (_i, [_vp, C.POINTER(BfPair), _i, _i]),
    "kfs_bf_knn_match_ratio": (_i, [_vp, _vp, _i, _vp, _i, _f, _vp, _ip]),
    "kfs_bf_set_stream": (_i, [_vp, _vp]),
    "kfs_bf_get_stream": (_vp, [_vp]),
    "kfs_bf_sync": (_i, [_vp]),
    "kfs_bf_last_launches": (_i, [_vp]),
    "kfs_bf_attach_orb": (_i, [_vp, _vp]),
    "kfs_bf_last_aliased": (_i, [_vp]),
    "kfs_bf_knn_match_device_static": (_i, [_i, _vp, _vp, _i, _vp, _i, _i, _vp, _vp]),
    "kfs_replay_default_params": (None, [C.POINTER(ReplayParams)]),
    "kfs_replay_create": (_i, [C.POINTER(ReplayParams), _i, C.POINTER(_vp)]),
    "kfs_replay_destroy": (None, [_vp]),
    "kfs_replay_capacity": (_i, [_vp]),
    "kfs_replay_submit": (_i, [_vp, C.POINTER(ReplayIO), C.POINTER(ReplayFrame), _i, _ip]),
    "kfs_replay_wait": (_i, [_vp, _i]),
    "kfs_replay_last_launches": (_i, [_vp]),
    "kfs_replay_arenas": (_i, [_vp]),
    "kfs_frontend_create": (_i, [C.POINTER(OrbParams), _i, C.POINTER(_vp)]),
    "kfs_frontend_destroy": (None, [_vp]),
    "kfs_frontend_capacity": (_i, [_vp]),
    "kfs_frontend_set_capacity": (_i, [_vp, _i]),
    "kfs_frontend_set_blur_fma": (_i, [_vp, _i]),
    "kfs_frontend_extract_features": (_i, [_vp, _vp, _sz, _i, _i, _i, _vp, _sz, _vp, _vp, _i, _ip]),
    "kfs_frontend_promote_keyframe": (_i, [_vp, _vp, _sz]),
    "kfs_frontend_set_keyframe": (_i, [_vp, _vp, _vp, _i, _vp, _sz, _i, _i]),
    "kfs_frontend_match_features": (_i, [_vp, _vp, _vp, _i, C.POINTER(C.c_double), _f, _vp, _vp, _vp, _i, _ip]),
    "kfs_frontend_track": (_i, [_vp, _vp, _sz, _i, _i, _i, _vp, _sz, C.POINTER(C.c_double), _f, _vp, _vp, _i, _ip, _vp, _vp, _vp, _i, _ip]),
    "kfs_frontend_set_window": (_i, [_vp, _i]),
    "kfs_frontend_window": (_i, [_vp]),
    "kfs_frontend_window_count": (_i, [_vp]),
    "kfs_frontend_match_window": (_i, [_vp, _vp, _i, _f, _ip, _vp, _i, _vp, _ip, _i, _ip]),
    "kfs_frontend_last_launches": (_i, [_vp]),
    "kfs_pipe_peak": (_i, [_i, _i, C.POINTER(C.c_double)]),
}

_LIB = None


def build(verbose=False):
    """Compile csrc/*.cu for sm_100a into libkfslam_frontend.so (nvcc cross-compiles without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", os.path.join(HERE, "csrc"), "-s", "-j4"], stdout=out, stderr=out)
    return LIB_PATH


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback for this path)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


class KfsError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("%s: %s" % (STATUS_NAMES[status] if 0 <= status < len(STATUS_NAMES) else status, msg))
        self.status = status


def check(status, allow=()):
    if status != KFS_OK and status not in allow:
        raise KfsError(status, lib().kfs_last_error().decode("utf-8", "replace"))
    return status
