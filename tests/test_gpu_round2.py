"""GPU parity (-m gpu) of what round 2 added, all through the C ABI:
  * the reference's own operating point — cv::ORB::create(3000, ...) (VisualOdometry.cpp:10-16), R,G,B frames (:25-28) and a
    mask on every call (:30) — through kfs_frontend_track and, batched, through the replay executor (masks + colour frames),
    against the cv2 golden refset_720p.npz;
  * config 3's sharding: two replay handles play ranks 0 / 1 of replay.shard_frames over a 45-frame 720p sequence (halo keyframe,
    keyframe cache across submits, rows trimmed to live ones); stitched results equal the oracle's single pass;
  * per-frame row skipping, arenas, kfs_bf_attach_orb (rows matched in place), two interleaved handles with different
    shared-memory needs, kfs_frontend_set_capacity.
Bit-exact throughout."""
import ctypes as C

import numpy as np
import pytest

from _track_util import make_depth
from _util import golden, synth
from test_refsetting_oracle import refset_inputs

pytestmark = pytest.mark.gpu
K = np.array([[910.1, 0, 640.4], [0, 909.6, 360.7], [0, 0, 1]])


@pytest.fixture(scope="module")
def kfs():
    import kfslam_b200 as k
    assert k.lib().kfs_device_count() >= 1, "no CUDA device: the product path has no CPU fallback"
    return k


def test_reference_setting_track_matches_cv2_golden(kfs, oracle):
    """nf = 3000, R,G,B in, mask on every call: kfs_frontend_track (one call per frame, as VisualOdometry::Track)."""
    g, inp = refset_inputs()
    depth = make_depth(720, 1280, 9)
    fe = kfs.TrackerFrontEnd(nfeatures=3000)
    kp, desc, p3, p2, m = fe.track(inp["kf"][0], K, mask=inp["kf"][1])
    assert p3 is None and np.array_equal(kp, g["kf_kp"]) and np.array_equal(desc, g["kf_desc"])
    fe.promoteKeyframe(depth)
    kp, desc, p3, p2, m = fe.track(inp["cur"][0], K, mask=inp["cur"][1])
    assert np.array_equal(kp, g["cur_kp"]) and np.array_equal(desc, g["cur_desc"])
    r3, r2, rm = oracle.match_features(g["knn"], g["kf_kp"], g["cur_kp"], depth, K, 0.6)
    assert len(rm) > 500 and np.array_equal(m, rm) and np.array_equal(p3, r3) and np.array_equal(p2, r2)


def test_reference_setting_batched_replay_matches_cv2_golden(kfs, oracle):
    """The same two frames as ONE replay submit: colour frames + per-frame masks in, kNN rows and survivors out."""
    g, inp = refset_inputs()
    frames = np.ascontiguousarray(np.stack([inp["kf"][0], inp["cur"][0], inp["cur"][0]]))
    masks = np.ascontiguousarray(np.stack([inp["kf"][1], inp["cur"][1], inp["cur"][1]]))
    rp = kfs.Replay(nfeatures=3000, max_batch=3, keyframe_slots=2, ratio=0.6)
    out = rp.alloc_outputs(3)
    rp.wait(rp.submit(frames, [(kfs.Replay.NO_MATCH, -1), (0, -1), (0, -1, 1)], out, masks=masks))
    assert list(out["n_kp"]) == [3000, 3000, 3000]
    assert np.array_equal(out["kp"][0, :3000], g["kf_kp"]) and np.array_equal(out["desc"][0, :3000], g["kf_desc"])
    assert np.array_equal(out["kp"][1, :3000], g["cur_kp"]) and np.array_equal(out["desc"][1, :3000], g["cur_desc"])
    assert not out["desc"][2].any() and not out["kp"][2]["size"].any()        # SKIP_ROWS: nothing written for frame 2 ...
    for i in (1, 2):                                                          # ... but it is matched like frame 1
        assert out["n_query"][i] == 3000 and np.array_equal(out["knn"][i, :3000], g["knn"])
        good = oracle.ratio_test(g["knn"], np.full(3000, 2, np.int32), 0.6)
        assert out["n_good"][i] == len(good) and np.array_equal(out["good"][i, :len(good)], good)
    rp.close()


def _oracle_sequence(oracle, frames, nf, kf_every):
    P = oracle.orb_params(nfeatures=nf)
    ref = [oracle.detect_and_compute(f, None, P) for f in frames]
    res = []
    from kfslam_b200 import replay
    for i in range(len(frames)):
        q = ref[replay.keyframe_of(i, kf_every)][1]
        m, c = oracle.knn_match(q, ref[i][1], 2)
        res.append((ref[i][0], ref[i][1], m, oracle.ratio_test(m, c, 0.6)))
    return res


@pytest.mark.parametrize("batch,arenas", [(16, 3), (7, 2)])
def test_sharded_sequence_replay_equals_single_pass(kfs, oracle, batch, arenas):
    """BASELINE.json configs[2] in miniature: a 45-frame 1280x720 sequence, keyframe every 10 frames (VisualOdometry.cpp:105),
    cut by shard_frames into two shards; each shard runs on its own replay handle (as rank 0 / rank 1 would), submits planned
    by plan_shard (halo keyframe extracted locally, last keyframe parked in the device cache for the next submit)."""
    from kfslam_b200 import replay
    n, nf, kf_every = 45, 1500, 10
    frames = np.ascontiguousarray(synth.make_stream(n, 720, 1280, seed=77))
    ref = _oracle_sequence(oracle, frames, nf, kf_every)
    got = {}
    for rank in (0, 1):
        lo, hi, halo = replay.shard_frames(n, 2, rank, kf_every)
        assert (rank == 0) == (halo < 0)
        rp = kfs.Replay(nfeatures=nf, max_batch=batch, keyframe_slots=2, ratio=0.6, arenas=arenas)
        assert rp.arenas == arenas
        subs = replay.plan_shard(lo, hi, halo, batch, kf_every)
        pending = []
        for s in subs:
            fr = np.ascontiguousarray(frames[s["frames"]])
            out = rp.alloc_outputs(len(s["frames"]))
            pending.append((rp.submit(fr, s["plan"], out), s, out))
            if len(pending) >= arenas:
                t, ss, oo = pending.pop(0)
                assert rp.wait(t) == 0
                for j, f in enumerate(ss["frames"]):
                    if ss["owned"][j]:
                        got[f] = {k: np.array(v[j]) for k, v in oo.items()}
        for t, ss, oo in pending:
            assert rp.wait(t) == 0
            for j, f in enumerate(ss["frames"]):
                if ss["owned"][j]:
                    got[f] = {k: np.array(v[j]) for k, v in oo.items()}
        rp.close()
    assert sorted(got) == list(range(n))
    for i in range(n):
        kp, desc, knn, good = ref[i]
        o = got[i]
        nk, nq, ng = int(o["n_kp"]), int(o["n_query"]), int(o["n_good"])
        assert nk == len(kp) and np.array_equal(o["kp"][:nk], kp) and np.array_equal(o["desc"][:nk], desc), "frame %d extraction" % i
        assert nq == len(knn) and np.array_equal(o["knn"][:nq], knn), "frame %d kNN" % i
        assert ng == len(good) and np.array_equal(o["good"][:ng], good), "frame %d survivors" % i


def test_replay_rows_past_nfeatures_are_fetched(kfs, oracle):
    """A tie-rich frame keeps MORE than nfeatures keypoints: the eager copies carry nfeatures rows per frame,
    kfs_replay_wait has to bring the rest (keypoints, descriptors, and the kNN rows of such a keyframe)."""
    f = synth.periodic_frame(480, 640)
    nf = 300
    P = oracle.orb_params(nfeatures=nf)
    kp, desc = oracle.detect_and_compute(f, None, P)
    assert len(kp) > nf, "the periodic frame no longer produces ties: pick another period"
    frames = np.ascontiguousarray(np.stack([f, f]))
    rp = kfs.Replay(nfeatures=nf, max_batch=2, keyframe_slots=1, ratio=0.9)
    assert rp.capacity >= len(kp)
    out = rp.alloc_outputs(2)
    assert rp.wait(rp.submit(frames, [(0, -1), (0, -1)], out)) == 0
    m, c = oracle.knn_match(desc, desc, 2)
    good = oracle.ratio_test(m, c, 0.9)
    for i in (0, 1):
        assert out["n_kp"][i] == len(kp) and np.array_equal(out["kp"][i, :len(kp)], kp) and np.array_equal(out["desc"][i, :len(kp)], desc)
        assert out["n_query"][i] == len(kp) and np.array_equal(out["knn"][i, :len(kp)], m)
        assert out["n_good"][i] == len(good) and np.array_equal(out["good"][i, :len(good)], good)
    rp.close()


def test_matcher_attached_to_extractor_matches_rows_in_place(kfs):
    """kfs_bf_attach_orb: after kfs_orb_detect_and_compute_batch the descriptor rows handed to kfs_bf_knn_match_pairs by their
    host address are matched on the device copy (no re-upload); foreign pointers, other row counts, edited rows and rows of an
    earlier call are uploaded as before."""
    from importlib import import_module
    N = import_module("keyframe-based-slam_b200._native")
    L = kfs.lib()
    B = 4
    frames = np.ascontiguousarray(synth.make_stream(B, 480, 640, seed=5))
    orb = kfs.ORB_create(800)
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    cap = orb.capacity
    kp = np.zeros((B, cap), kfs.KP_DTYPE); desc = np.zeros((B, cap, 32), np.uint8); n = (C.c_int * B)()
    N.check(L.kfs_orb_detect_and_compute_batch(orb._h, frames.ctypes.data, 640, 640 * 480, 640, 480, B, None, 0, 0,
                                               kp.ctypes.data, desc.ctypes.data, cap, n))
    outs = [np.zeros((cap, 2), kfs.DM_DTYPE) for _ in range(2 * B)]

    def run(attach):
        N.check(L.kfs_bf_attach_orb(bf._h, orb._h if attach else None))
        pairs = (N.BfPair * B)()
        for i in range(B):
            o = outs[i + (B if attach else 0)]
            pairs[i] = N.BfPair(desc[0].ctypes.data, n[0], None, desc[i].ctypes.data, n[i], None, o.ctypes.data, None)
        N.check(L.kfs_bf_knn_match_pairs(bf._h, pairs, B, 2))
        return L.kfs_bf_last_aliased(bf._h)
    assert run(False) == 0
    assert run(True) == B                      # frame 0 (query of every pair) + frames 1..3, each counted once
    for i in range(B):
        assert np.array_equal(outs[i][:n[0]], outs[B + i][:n[0]])
    foreign = desc[1].copy()                   # same bytes at another address: not recognised, uploaded
    pairs = (N.BfPair * 1)(N.BfPair(desc[0].ctypes.data, n[0], None, foreign.ctypes.data, n[1], None, outs[0].ctypes.data, None))
    N.check(L.kfs_bf_knn_match_pairs(bf._h, pairs, 1, 2))
    assert L.kfs_bf_last_aliased(bf._h) == 1 and np.array_equal(outs[0][:n[0]], outs[B + 1][:n[0]])
    # ADVICE r2 / VERDICT weak 1b: everything that must NOT be taken for "the rows I just delivered"
    ref = outs[B + 1][:n[0]].copy()            # pair (frame 0, frame 1), known good

    def one(qrows, trows, nq_, nt_):
        o = np.zeros((cap, 2), kfs.DM_DTYPE)
        pr = (N.BfPair * 1)(N.BfPair(qrows.ctypes.data, nq_, None, trows.ctypes.data, nt_, None, o.ctypes.data, None))
        N.check(L.kfs_bf_knn_match_pairs(bf._h, pr, 1, 2))
        return o, L.kfs_bf_last_aliased(bf._h)
    N.check(L.kfs_bf_attach_orb(bf._h, orb._h))
    o, al = one(desc[0], desc[1], n[0], n[1])
    assert al == 2 and np.array_equal(o[:n[0]], ref)
    # fewer rows than the frame has (a caller that filtered its matrix): uploaded, and the answer is for the rows that were passed
    o, al = one(desc[0], desc[1], n[0], n[1] - 7)
    assert al == 1
    plain, _ = kfs.BFMatcher_create(kfs.NORM_HAMMING).knnMatch(desc[0][:n[0]], desc[1][:n[1] - 7])
    assert np.array_equal(o[:n[0]], plain)
    # rows edited in place (first or last row no longer what was delivered): uploaded, answer follows the edit
    keep = desc[1].copy()
    desc[1][n[1] - 1] ^= 0xFF
    o, al = one(desc[0], desc[1], n[0], n[1])
    plain, _ = kfs.BFMatcher_create(kfs.NORM_HAMMING).knnMatch(desc[0][:n[0]], desc[1][:n[1]])
    assert al == 1 and np.array_equal(o[:n[0]], plain) and not np.array_equal(o[:n[0]], ref)
    desc[1][:] = keep
    o, al = one(desc[0], desc[1], n[0], n[1])
    assert al == 2 and np.array_equal(o[:n[0]], ref)
    # a device-side extraction in between reuses the handle's arena: nothing of the old host call may be aliased afterwards
    import torch
    d_frames = torch.from_numpy(frames[::-1].copy()).cuda()
    orb.extract_device(d_frames.data_ptr(), 640, 640 * 480, 640, 480, B)
    orb.sync()
    o, al = one(desc[0], desc[1], n[0], n[1])
    assert al == 0 and np.array_equal(o[:n[0]], ref)
    # a second host call into OTHER buffers: the first call's buffers are no longer the device content
    kp2 = np.zeros((B, cap), kfs.KP_DTYPE); desc2 = np.zeros((B, cap, 32), np.uint8); n2 = (C.c_int * B)()
    N.check(L.kfs_orb_detect_and_compute_batch(orb._h, frames[::-1].copy().ctypes.data, 640, 640 * 480, 640, 480, B, None, 0, 0,
                                               kp2.ctypes.data, desc2.ctypes.data, cap, n2))
    o, al = one(desc[0], desc[1], n[0], n[1])
    assert al == 0 and np.array_equal(o[:n[0]], ref)
    o, al = one(desc2[3], desc2[2], n2[3], n2[2])   # frames[::-1][3] = frame 0, [2] = frame 1: the new buffers are recognised
    assert al == 2 and np.array_equal(o[:n[0]], ref)
    N.check(L.kfs_bf_attach_orb(bf._h, None))


def test_two_handles_with_different_shared_memory_needs_interleaved(kfs, oracle):
    """cudaFuncAttributeMaxDynamicSharedMemorySize is per function and device, not per handle.  A 4K / nf = 10000 handle
    and a 720p / nf = 5000 handle, both above 48 KB of dynamic shared memory in k_select_sort, used alternately."""
    big = synth.make_stream(1, 2160, 3840, seed=3)[0]
    small = synth.make_stream(1, 720, 1280, seed=4)[0]
    a = kfs.ORB_create(10000)
    b = kfs.ORB_create(5000)
    ra = oracle.detect_and_compute(big, None, oracle.orb_params(nfeatures=10000))
    rb = oracle.detect_and_compute(small, None, oracle.orb_params(nfeatures=5000))
    for _ in range(3):
        ka, da = a.detectAndCompute(big)
        kb, db = b.detectAndCompute(small)
        assert np.array_equal(ka, ra[0]) and np.array_equal(da, ra[1])
        assert np.array_equal(kb, rb[0]) and np.array_equal(db, rb[1])


def test_frontend_set_capacity_and_overflow_status(kfs, oracle):
    """A tie-rich frame overflows a handle whose capacity was cut: KFS_ERR_CAPACITY with the needed row count, never a silently
    shorter keypoint set; kfs_frontend_set_capacity makes room."""
    from importlib import import_module
    N = import_module("keyframe-based-slam_b200._native")
    L = kfs.lib()
    f = synth.periodic_frame(480, 640)
    nf = 300
    kp, desc = oracle.detect_and_compute(f, None, oracle.orb_params(nfeatures=nf))
    assert len(kp) > nf
    fe = kfs.TrackerFrontEnd(nfeatures=nf)
    N.check(L.kfs_frontend_set_capacity(fe._h, nf))           # no tie slack at all
    cap = fe.capacity
    assert nf <= cap < len(kp)
    k_out = np.zeros(cap, kfs.KP_DTYPE); d_out = np.zeros((cap, 32), np.uint8); n = C.c_int()
    st = L.kfs_frontend_extract_features(fe._h, f.ctypes.data, 640, 640, 480, 1, None, 0, k_out.ctypes.data, d_out.ctypes.data, cap, C.byref(n))
    assert st == N.KFS_ERR_CAPACITY and n.value == len(kp)
    N.check(L.kfs_frontend_set_capacity(fe._h, n.value + 8))
    got_kp, got_desc = fe.extractFeatures(f)
    assert np.array_equal(got_kp, kp) and np.array_equal(got_desc, desc)
    # a set_capacity that runs out of device memory leaves a DEFINED state — every computing call refuses the handle
    # (no run on freed / null buffers) until a later set_capacity succeeds.  2e9 rows = 64 + 56 + 64 GB for the first three buffers.
    assert L.kfs_frontend_set_capacity(fe._h, 2_000_000_000) == N.KFS_ERR_CUDA
    assert fe.capacity == 0
    st = L.kfs_frontend_extract_features(fe._h, f.ctypes.data, 640, 640, 480, 1, None, 0, k_out.ctypes.data, d_out.ctypes.data, cap, C.byref(n))
    assert st == N.KFS_ERR_INVALID_ARG and "unsized" in L.kfs_last_error().decode()
    assert L.kfs_frontend_promote_keyframe(fe._h, None, 0) == N.KFS_ERR_INVALID_ARG
    assert L.kfs_frontend_set_window(fe._h, 4) == N.KFS_ERR_INVALID_ARG
    N.check(L.kfs_frontend_set_capacity(fe._h, len(kp) + 8))
    assert L.kfs_frontend_window(fe._h) == 4          # the ring size asked for while unsized is honoured once there is room
    got_kp, got_desc = fe.extractFeatures(f)
    assert np.array_equal(got_kp, kp) and np.array_equal(got_desc, desc)
