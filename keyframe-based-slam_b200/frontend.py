"""Host-side mirror of the reference's operator interface for the hot path, over the C ABI.

The reference tracker (C++) calls, per frame (/root/reference/src/lib/VisualOdometry.cpp):
    ptrExtractor = cv::ORB::create(3000, 1.2f, 8, 31, 0, 2, cv::ORB::HARRIS_SCORE, 31, 20)      :10-16
    ptrExtractor->detectAndCompute(gray, mask, kp, desc)                                         :30
    ptrMatcher   = cv::BFMatcher::create(cv::NORM_HAMMING)                                       :18
    ptrMatcher->knnMatch(keyframe->matDescriptors, dpCurrImg, knnMatches, 2)                     :39
The C++ drop-in for those is include/kfslam_frontend_cv.hpp.  This module is the same surface for
Python callers and for the parity tests (names and argument meaning follow cv2's: ORB_create,
detectAndCompute, BFMatcher_create, knnMatch), so a test written against cv2 reads the same here.
Keypoints come back as a structured array with cv::KeyPoint's fields; matches as cv::DMatch's.
"""
import ctypes as C

import numpy as np

from . import _native as N
from ._native import KP_DTYPE, DM_DTYPE, NORM_HAMMING, ORB_HARRIS_SCORE, KfsError  # noqa: F401


def _ptr(a):
    return a.ctypes.data if a is not None else None


class ORB:
    """cv::ORB stand-in.  One instance per host thread / GPU (not thread-safe, like the reference's use)."""

    def __init__(self, nfeatures=500, scaleFactor=1.2, nlevels=8, edgeThreshold=31, firstLevel=0, WTA_K=2,
                 scoreType=ORB_HARRIS_SCORE, patchSize=31, fastThreshold=20, device=0):
        self._h = C.c_void_p()
        self.params = N.OrbParams(nfeatures, scaleFactor, nlevels, edgeThreshold, firstLevel, WTA_K, scoreType,
                                  patchSize, fastThreshold)
        N.check(N.lib().kfs_orb_create(C.byref(self.params), device, C.byref(self._h)))
        self.device = device

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().kfs_orb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):   # at interpreter shutdown module globals may already be gone: nothing to free then
        try:
            self.close()
        except Exception:
            pass

    # -- cv2-shaped calls ---------------------------------------------------------------------------
    def detectAndCompute(self, image, mask=None, capacity=None):
        """-> (keypoints[KP_DTYPE], descriptors uint8[n,32]); descriptors is None when n == 0 (as cv2)."""
        kps, descs = self.detectAndComputeBatch(np.asarray(image)[None], None if mask is None else np.asarray(mask)[None], capacity)
        return kps[0], (descs[0] if len(kps[0]) else None)

    def detectAndComputeBatch(self, images, masks=None, capacity=None):
        """images uint8 [B,H,W] (host) -> lists of per-frame (keypoints, descriptors)."""
        images = np.ascontiguousarray(images, dtype=np.uint8)
        if images.ndim != 3:
            raise ValueError("images must be [B,H,W] uint8 (gray)")
        B, H, W = images.shape
        if masks is not None:
            masks = np.ascontiguousarray(masks, dtype=np.uint8)
            if masks.shape != images.shape:
                raise ValueError("mask shape must equal image shape")
        cap = capacity or self.capacity
        while True:
            kp = np.zeros((B, cap), KP_DTYPE)
            desc = np.zeros((B, cap, 32), np.uint8)
            n = (C.c_int * B)()
            st = N.lib().kfs_orb_detect_and_compute_batch(
                self._h, _ptr(images), W, H * W, W, H, B,
                _ptr(masks), W if masks is not None else 0, H * W if masks is not None else 0,
                _ptr(kp), _ptr(desc), cap, n)
            N.check(st, allow=(N.KFS_ERR_CAPACITY,))
            if st == N.KFS_ERR_CAPACITY and capacity is None:
                need = max(n)
                if need > self.capacity:
                    N.check(N.lib().kfs_orb_set_capacity(self._h, need + 64))
                cap = self.capacity
                continue
            break
        counts = [min(int(c), cap) for c in n]
        return [kp[f, :counts[f]].copy() for f in range(B)], [desc[f, :counts[f]].copy() for f in range(B)]

    # -- device-resident path -----------------------------------------------------------------------
    def extract_device(self, d_ptr, step, frame_stride, w, h, batch, d_mask=None, mask_step=0, mask_frame_stride=0):
        N.check(N.lib().kfs_orb_extract_device(self._h, d_ptr, step, frame_stride, w, h, batch, d_mask, mask_step, mask_frame_stride))

    def device_results(self):
        kp, desc, cnt = C.c_void_p(), C.c_void_p(), C.c_void_p()
        cap, batch = C.c_int(), C.c_int()
        N.check(N.lib().kfs_orb_device_results(self._h, C.byref(kp), C.byref(desc), C.byref(cnt), C.byref(cap), C.byref(batch)))
        return kp.value, desc.value, cnt.value, cap.value, batch.value

    def fetch_results(self, batch, capacity=None):
        cap = capacity or self.capacity
        kp = np.zeros((batch, cap), KP_DTYPE)
        desc = np.zeros((batch, cap, 32), np.uint8)
        n = (C.c_int * batch)()
        N.check(N.lib().kfs_orb_fetch_results(self._h, _ptr(kp), _ptr(desc), cap, n), allow=(N.KFS_ERR_CAPACITY,))
        counts = [min(int(c), cap) for c in n]
        return [kp[f, :counts[f]].copy() for f in range(batch)], [desc[f, :counts[f]].copy() for f in range(batch)]

    def sync(self):
        N.check(N.lib().kfs_orb_sync(self._h))

    def set_stream(self, cuda_stream):
        N.check(N.lib().kfs_orb_set_stream(self._h, cuda_stream))

    def set_pipeline_chunk(self, frames):
        """Frames per chunk of the host-buffer pipeline (0 = automatic)."""
        N.check(N.lib().kfs_orb_set_pipeline_chunk(self._h, int(frames)))

    @property
    def capacity(self):
        return N.lib().kfs_orb_capacity(self._h)

    def set_capacity(self, cap):
        N.check(N.lib().kfs_orb_set_capacity(self._h, cap))

    # -- stage taps ---------------------------------------------------------------------------------
    def level_info(self, level):
        w, h, q, s = C.c_int(), C.c_int(), C.c_int(), C.c_float()
        N.check(N.lib().kfs_orb_level_info(self._h, level, C.byref(w), C.byref(h), C.byref(q), C.byref(s)))
        return w.value, h.value, q.value, s.value

    def download_level(self, frame, level):
        w, h, _, _ = self.level_info(level)
        out = np.empty((h, w), np.uint8)
        N.check(N.lib().kfs_orb_download_level(self._h, frame, level, _ptr(out), w))
        return out

    def set_blur_fma(self, fma=True):
        """a1.9: True = OpenCV's default FMA dispatch of the 7x7 blur, False = its plain mul + add build (cv2.setUseOptimized(False))."""
        N.check(N.lib().kfs_orb_set_blur_fma(self._h, int(bool(fma))))

    def download_blurred_level(self, frame, level):
        w, h, _, _ = self.level_info(level)
        out = np.zeros((h, w), np.uint8)
        N.check(N.lib().kfs_orb_download_blurred_level(self._h, frame, level, out.ctypes.data, out.strides[0]))
        return out

    def download_fast_corners(self, frame, level):
        """-> int32 [n,3] rows (x, y, score), sorted by (y, x)."""
        w, h, _, _ = self.level_info(level)
        cap = max(16, (w // 2 + 2) * (h // 2 + 2))
        out = np.zeros(cap, np.uint32)
        n = C.c_int()
        N.check(N.lib().kfs_orb_download_fast_corners(self._h, frame, level, _ptr(out), cap, C.byref(n)))
        v = out[:n.value]
        arr = np.stack([(v & 0xfff).astype(np.int32), ((v >> 12) & 0xfff).astype(np.int32), (v >> 24).astype(np.int32)], 1)
        return arr[np.lexsort((arr[:, 0], arr[:, 1]))]

    def set_profiling(self, on=True):
        N.check(N.lib().kfs_orb_set_profiling(self._h, int(on)))

    def stage_times(self):
        ms = (C.c_float * len(N.STAGE_NAMES))()
        N.check(N.lib().kfs_orb_stage_times(self._h, ms))
        return dict(zip(N.STAGE_NAMES, [float(x) for x in ms]))

    @property
    def last_launches(self):
        return N.lib().kfs_orb_last_launches(self._h)


def ORB_create(nfeatures=500, scaleFactor=1.2, nlevels=8, edgeThreshold=31, firstLevel=0, WTA_K=2,
               scoreType=ORB_HARRIS_SCORE, patchSize=31, fastThreshold=20, device=0):
    """cv2.ORB_create / cv::ORB::create with the same positional arguments."""
    return ORB(nfeatures, scaleFactor, nlevels, edgeThreshold, firstLevel, WTA_K, scoreType, patchSize, fastThreshold, device)


class BFMatcher:
    """cv::BFMatcher stand-in (NORM_HAMMING, crossCheck=false, k <= 2)."""

    def __init__(self, normType=NORM_HAMMING, crossCheck=False, device=0):
        self._h = C.c_void_p()
        N.check(N.lib().kfs_bf_create(normType, int(crossCheck), device, C.byref(self._h)))
        self.device = device

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().kfs_bf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):   # at interpreter shutdown module globals may already be gone: nothing to free then
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _desc(a):
        a = np.ascontiguousarray(a, dtype=np.uint8)
        if a.size == 0:
            return np.zeros((0, 32), np.uint8)
        if a.ndim != 2 or a.shape[1] != 32:
            raise ValueError("descriptors must be [n,32] uint8")
        return a

    def knnMatch(self, queryDescriptors, trainDescriptors, k=2):
        """-> (matches[DM_DTYPE] shape [nq,k], counts int32[nq]); row q holds counts[q] = min(k, nt) entries."""
        q, t = self._desc(queryDescriptors), self._desc(trainDescriptors)
        out = np.zeros((len(q), k), DM_DTYPE)
        counts = np.zeros(len(q), np.int32)
        N.check(N.lib().kfs_bf_knn_match(self._h, _ptr(q), len(q), _ptr(t), len(t), k, _ptr(out), _ptr(counts)))
        return out, counts

    def knnMatchPairs(self, pairs, k=2):
        """pairs: list of (query, train) host arrays -> list of (matches, counts); one launch for all pairs."""
        qs = [self._desc(p[0]) for p in pairs]
        ts = [self._desc(p[1]) for p in pairs]
        # identical input arrays share one upload (window matching: one query set vs many train sets)
        outs = [np.zeros((len(q), k), DM_DTYPE) for q in qs]
        cnts = [np.zeros(len(q), np.int32) for q in qs]
        arr = (N.BfPair * len(pairs))()
        ids = {}
        for i, (p, q, t) in enumerate(zip(pairs, qs, ts)):
            kq, kt = id(p[0]), id(p[1])
            q = ids.setdefault(("q", kq), q)
            t = ids.setdefault(("t", kt), t)
            arr[i] = N.BfPair(_ptr(q), len(q), None, _ptr(t), len(t), None, _ptr(outs[i]), _ptr(cnts[i]))
        N.check(N.lib().kfs_bf_knn_match_pairs(self._h, arr, len(pairs), k))
        return list(zip(outs, cnts))

    def knnMatchPairsDevice(self, pairs, k=2):
        """pairs: list of dict(query=ptr, nq=int, train=ptr, nt=int, out=ptr, counts=ptr|None, d_nq=ptr|None, d_nt=ptr|None)."""
        arr = (N.BfPair * len(pairs))()
        for i, p in enumerate(pairs):
            arr[i] = N.BfPair(p["query"], p["nq"], p.get("d_nq"), p["train"], p["nt"], p.get("d_nt"), p["out"], p.get("counts"))
        N.check(N.lib().kfs_bf_knn_match_pairs_device(self._h, arr, len(pairs), k))

    def knnMatchRatio(self, queryDescriptors, trainDescriptors, ratio=0.6):
        """knnMatch(k=2) + the reference's ratio test (VisualOdometry.cpp:46) fused on the device."""
        q, t = self._desc(queryDescriptors), self._desc(trainDescriptors)
        good = np.zeros(max(len(q), 1), DM_DTYPE)
        n = C.c_int()
        N.check(N.lib().kfs_bf_knn_match_ratio(self._h, _ptr(q), len(q), _ptr(t), len(t), ratio, _ptr(good), C.byref(n)))
        return good[:n.value].copy()

    def set_stream(self, cuda_stream):
        N.check(N.lib().kfs_bf_set_stream(self._h, cuda_stream))

    def sync(self):
        N.check(N.lib().kfs_bf_sync(self._h))

    @property
    def last_launches(self):
        return N.lib().kfs_bf_last_launches(self._h)


def BFMatcher_create(normType=NORM_HAMMING, crossCheck=False, device=0):
    """cv2.BFMatcher_create / cv::BFMatcher::create."""
    return BFMatcher(normType, crossCheck, device)


def knn_match_device_static(query_ptr, nq, train_ptr, nt, out_ptr, counts_ptr=None, k=2, device=0, stream=None):
    """kfs_bf_knn_match_device_static: one (query, train) problem in DEVICE memory, matched without creating a handle and
    without any allocation (the split launch's scratch is a ~2 MiB array compiled into the library).  For hosts that may not
    call cudaMalloc while they run; see include/kfslam_frontend.h for the rules (nq <= STATIC_MAX_QUERIES, one call at a time
    per GPU, asynchronous on `stream`).  No counterpart in the reference."""
    N.check(N.lib().kfs_bf_knn_match_device_static(device, stream, query_ptr, nq, train_ptr, nt, k, out_ptr, counts_ptr))


STATIC_MAX_QUERIES = 16384   # KFS_BF_STATIC_MAX_QUERIES


def pipe_peak(device=0, what="popc"):
    """Measured integer-pipe issue rate in 1e9 ops/s (what = 'popc' | 'lop3'); matcher roofline denominator."""
    g = C.c_double()
    N.check(N.lib().kfs_pipe_peak(device, 0 if what == "popc" else 1, C.byref(g)))
    return g.value


class Replay:
    """Replay executor (kfs_replay_*): batches of host frames -> keypoints, descriptors, kNN matches against each
    frame's keyframe and ratio-test survivors, double-buffered on the device (VisualOdometry::Track's front half,
    /root/reference/src/lib/VisualOdometry.cpp:129-172).  submit() only enqueues; wait() blocks.  All buffers
    passed to submit() must stay alive (and unmodified) until the matching wait()."""
    NO_MATCH = N.REPLAY_NO_MATCH

    def __init__(self, nfeatures=500, scaleFactor=1.2, nlevels=8, fastThreshold=20, max_batch=64, keyframe_slots=8,
                 ratio=0.6, chunk_frames=0, device=0, arenas=0):
        self._h = C.c_void_p()
        p = N.ReplayParams()
        N.lib().kfs_replay_default_params(C.byref(p))
        p.orb.nfeatures, p.orb.scale_factor, p.orb.nlevels, p.orb.fast_threshold = nfeatures, scaleFactor, nlevels, fastThreshold
        p.max_batch, p.keyframe_slots, p.ratio, p.chunk_frames, p.arenas = max_batch, keyframe_slots, ratio, chunk_frames, arenas
        N.check(N.lib().kfs_replay_create(C.byref(p), device, C.byref(self._h)))
        self.max_batch = max_batch
        self._keep = {}

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().kfs_replay_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):   # at interpreter shutdown module globals may already be gone: nothing to free then
        try:
            self.close()
        except Exception:
            pass

    @property
    def capacity(self):
        return N.lib().kfs_replay_capacity(self._h)

    @property
    def last_launches(self):
        return N.lib().kfs_replay_last_launches(self._h)

    def alloc_outputs(self, batch, pinned=None):
        """Output buffers for one submit: dict of numpy arrays (pass pinned=callable(shape, dtype) to place them)."""
        cap = self.capacity
        mk = pinned or (lambda shape, dtype: np.zeros(shape, dtype))
        return dict(kp=mk((batch, cap), KP_DTYPE), desc=mk((batch, cap, 32), np.uint8), n_kp=mk((batch,), np.int32),
                    knn=mk((batch, cap, 2), DM_DTYPE), n_query=mk((batch,), np.int32),
                    good=mk((batch, cap), DM_DTYPE), n_good=mk((batch,), np.int32))

    @property
    def arenas(self):
        return N.lib().kfs_replay_arenas(self._h)

    def submit(self, frames, plan, out, masks=None):
        """frames uint8 [B,H,W] gray or [B,H,W,3] R,G,B (host, C-contiguous); plan: list of (keyframe, store_slot) or
        (keyframe, store_slot, flags); out: alloc_outputs(); masks: None or uint8 [B,H,W], non-zero = keep."""
        B, H, W = frames.shape[:3]
        ch = 3 if frames.ndim == 4 else 1
        assert frames.dtype == np.uint8 and frames.flags.c_contiguous and (frames.ndim == 3 or frames.shape[3] == 3)
        if masks is not None:
            assert masks.dtype == np.uint8 and masks.flags.c_contiguous and masks.shape == (B, H, W)
        io = N.ReplayIO(frames.ctypes.data, W * ch, W * H * ch, W, H, out["kp"].ctypes.data, out["desc"].ctypes.data, out["kp"].shape[1],
                        out["n_kp"].ctypes.data, _ptr(out.get("knn")), _ptr(out.get("n_query")), _ptr(out.get("good")), _ptr(out.get("n_good")),
                        _ptr(masks), W, W * H, ch)
        pl = (N.ReplayFrame * B)(*[N.ReplayFrame(int(e[0]), int(e[1]), int(e[2]) if len(e) > 2 else 0) for e in plan])
        t = C.c_int(-1)
        N.check(N.lib().kfs_replay_submit(self._h, C.byref(io), pl, B, C.byref(t)))
        self._keep[t.value] = (frames, out, io, pl, masks)
        return t.value

    def wait(self, ticket):
        st = N.check(N.lib().kfs_replay_wait(self._h, ticket), allow=(N.KFS_ERR_CAPACITY,))
        self._keep.pop(ticket, None)
        return st


class TrackerFrontEnd:
    """VisualOdometry::ExtractFeatures / MatchFeatures as the reference's tracker calls them
    (/root/reference/src/lib/VisualOdometry.cpp:21-33, :35-62) over kfs_frontend_*: colour -> gray on the device, the last
    keyframe (descriptors, keypoints, depth; KeyFrame.hpp:35-39) cached on the device, and kNN + ratio test + depth
    look-up + pin-hole back-projection returning the (points3D, points2D, matches) the PnP stage consumes."""

    def __init__(self, nfeatures=3000, scaleFactor=1.2, nlevels=8, fastThreshold=20, device=0):
        self._h = C.c_void_p()
        p = N.OrbParams()
        N.lib().kfs_orb_default_params(C.byref(p))
        p.nfeatures, p.scale_factor, p.nlevels, p.fast_threshold = nfeatures, scaleFactor, nlevels, fastThreshold
        N.check(N.lib().kfs_frontend_create(C.byref(p), device, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().kfs_frontend_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):   # at interpreter shutdown module globals may already be gone: nothing to free then
        try:
            self.close()
        except Exception:
            pass

    @property
    def capacity(self):
        return N.lib().kfs_frontend_capacity(self._h)

    @property
    def last_launches(self):
        return N.lib().kfs_frontend_last_launches(self._h)

    def extractFeatures(self, rgb, mask=None):
        """rgb: uint8 [H,W,3] (R,G,B byte order, as the reference receives it) or [H,W] gray -> (keypoints, descriptors)."""
        rgb = np.asarray(rgb)
        assert rgb.dtype == np.uint8 and rgb.ndim in (2, 3) and rgb.strides[-1] == 1
        ch = 1 if rgb.ndim == 2 else rgb.shape[2]
        if rgb.ndim == 3 and rgb.strides[1] != ch:
            rgb = np.ascontiguousarray(rgb)
        h, w = rgb.shape[:2]
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.uint8)
        cap = self.capacity
        kp = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int()
        N.check(N.lib().kfs_frontend_extract_features(self._h, rgb.ctypes.data, rgb.strides[0], w, h, ch, _ptr(mask),
                                                      mask.strides[0] if mask is not None else 0, kp.ctypes.data, desc.ctypes.data, cap, C.byref(n)),
                allow=(N.KFS_ERR_CAPACITY,))
        m = min(n.value, cap)
        return kp[:m].copy(), desc[:m].copy()

    def promoteKeyframe(self, depth):
        depth = np.ascontiguousarray(depth, dtype=np.uint16)
        N.check(N.lib().kfs_frontend_promote_keyframe(self._h, depth.ctypes.data, depth.strides[0]))

    def setKeyframe(self, desc, kp, depth):
        desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        kp = np.ascontiguousarray(kp, dtype=KP_DTYPE)
        depth = np.ascontiguousarray(depth, dtype=np.uint16)
        N.check(N.lib().kfs_frontend_set_keyframe(self._h, desc.ctypes.data, kp.ctypes.data, len(kp), depth.ctypes.data, depth.strides[0],
                                                  depth.shape[1], depth.shape[0]))

    def matchFeatures(self, K, dpCurrImg=None, kpCurrImg=None, ratio=0.6):
        """K: 3x3 camera matrix.  -> (points3D float64 [n,3], points2D float64 [n,2], matches DM_DTYPE [n]) in the order of the
        reference's loop over the keyframe's descriptors.  With dpCurrImg/kpCurrImg None the device-resident result of the
        last extractFeatures is matched."""
        K = np.asarray(K, dtype=np.float64)
        intr = (C.c_double * 4)(K[0, 0], K[1, 1], K[0, 2], K[1, 2])
        cap = self.capacity
        m = np.zeros(cap, DM_DTYPE)
        p3 = np.zeros((cap, 3), np.float64)
        p2 = np.zeros((cap, 2), np.float64)
        n = C.c_int()
        if dpCurrImg is not None:
            dpCurrImg = np.ascontiguousarray(dpCurrImg, dtype=np.uint8).reshape(-1, 32)
            kpCurrImg = np.ascontiguousarray(kpCurrImg, dtype=KP_DTYPE)
            assert len(dpCurrImg) == len(kpCurrImg)
        N.check(N.lib().kfs_frontend_match_features(self._h, _ptr(dpCurrImg), _ptr(kpCurrImg), 0 if dpCurrImg is None else len(dpCurrImg),
                                                    intr, ratio, m.ctypes.data, p3.ctypes.data, p2.ctypes.data, cap, C.byref(n)))
        return p3[:n.value].copy(), p2[:n.value].copy(), m[:n.value].copy()

    def track(self, rgb, K, mask=None, ratio=0.6):
        """ExtractFeatures + MatchFeatures of one frame in one call / one synchronisation (VisualOdometry::Track's front half).
        -> (keypoints, descriptors, points3D, points2D, matches); the last three are None while no keyframe is cached."""
        rgb = np.asarray(rgb)
        assert rgb.dtype == np.uint8 and rgb.ndim in (2, 3)
        rgb = np.ascontiguousarray(rgb)
        ch = 1 if rgb.ndim == 2 else rgb.shape[2]
        h, w = rgb.shape[:2]
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.uint8)
        K = np.asarray(K, dtype=np.float64)
        intr = (C.c_double * 4)(K[0, 0], K[1, 1], K[0, 2], K[1, 2])
        cap = self.capacity
        kp = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        m = np.zeros(cap, DM_DTYPE)
        p3 = np.zeros((cap, 3), np.float64)
        p2 = np.zeros((cap, 2), np.float64)
        n, nm = C.c_int(), C.c_int()
        N.check(N.lib().kfs_frontend_track(self._h, rgb.ctypes.data, rgb.strides[0], w, h, ch, _ptr(mask), mask.strides[0] if mask is not None else 0,
                                           intr, ratio, kp.ctypes.data, desc.ctypes.data, cap, C.byref(n),
                                           m.ctypes.data, p3.ctypes.data, p2.ctypes.data, cap, C.byref(nm)))
        k = min(n.value, cap)
        if nm.value < 0:
            return kp[:k].copy(), desc[:k].copy(), None, None, None
        return kp[:k].copy(), desc[:k].copy(), p3[:nm.value].copy(), p2[:nm.value].copy(), m[:nm.value].copy()

    def set_blur_fma(self, fma):
        """Blur variant of the extractor: True = OpenCV's FMA dispatch (default), False = multiply then add."""
        N.check(N.lib().kfs_frontend_set_blur_fma(self._h, int(bool(fma))))

    # ---- keyframe window (Map::GetNKeyFrames, /root/reference/src/lib/Map.cpp:72-78) ----
    def setWindow(self, window):
        """Keep the last `window` keyframes on the device (drops the cached ones)."""
        N.check(N.lib().kfs_frontend_set_window(self._h, window))

    @property
    def window(self):
        return N.lib().kfs_frontend_window(self._h)

    @property
    def windowCount(self):
        return N.lib().kfs_frontend_window_count(self._h)

    def matchWindow(self, dpCurrImg=None, ratio=0.6, want_knn=False):
        """Current frame against every cached keyframe, oldest first (GetNKeyFrames' order), one kNN launch.
        -> list of per-keyframe survivor arrays (DM_DTYPE, query = keyframe row, train = current row); with want_knn also
        the raw kNN rows [rows_i, 2] per keyframe."""
        cap = self.capacity
        W = N.KFS_MAX_WINDOW
        rows = (C.c_int * W)()
        ngood = (C.c_int * W)()
        nkf = C.c_int()
        cnt = self.windowCount
        good = np.zeros((max(cnt, 1), cap), DM_DTYPE)
        knn = np.zeros((max(cnt, 1), cap, 2), DM_DTYPE) if want_knn else None
        if dpCurrImg is not None:
            dpCurrImg = np.ascontiguousarray(dpCurrImg, dtype=np.uint8).reshape(-1, 32)
        N.check(N.lib().kfs_frontend_match_window(self._h, _ptr(dpCurrImg), 0 if dpCurrImg is None else len(dpCurrImg), ratio,
                                                  rows, _ptr(knn), cap if want_knn else 0, good.ctypes.data, ngood, cap, C.byref(nkf)))
        out = [good[i, :ngood[i]].copy() for i in range(nkf.value)]
        if want_knn:
            return out, [knn[i, :rows[i]].copy() for i in range(nkf.value)]
        return out
