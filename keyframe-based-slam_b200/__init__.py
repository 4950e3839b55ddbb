"""keyframe-based-slam_b200 — B200-native ORB + Hamming-kNN front-end for angelo-nutu/Keyframe-Based-SLAM.

Only what the hot path needs lives here: csrc/ (sm_100a kernels + the C ABI of include/kfslam_frontend.h),
the ctypes binding (_native), the cv2-shaped host mirror (frontend), the replay/sharding helpers (replay)
and the synthetic frame source (synth).  Import as
    kfs = importlib.import_module("keyframe-based-slam_b200")
(the directory name is not a Python identifier; `kfslam_b200.py` at the repo root re-exports it).
"""
from . import _native, replay, synth  # noqa: F401
from ._native import KP_DTYPE, DM_DTYPE, NORM_HAMMING, ORB_HARRIS_SCORE, KfsError, build, lib  # noqa: F401
from .frontend import ORB, ORB_create, BFMatcher, BFMatcher_create, Replay, TrackerFrontEnd, pipe_peak, knn_match_device_static, STATIC_MAX_QUERIES  # noqa: F401
