"""GPU parity tests (-m gpu) of the replay executor (kfs_replay_*, include/kfslam_frontend.h) and of the chunked
host-buffer pipeline of kfs_orb_detect_and_compute_batch: both must give, frame by frame, exactly what the
oracle's detectAndCompute + knnMatch + ratio test give (VisualOdometry.cpp:30, :39, :46), whatever the chunking,
double buffering or keyframe-cache path taken."""
import numpy as np
import pytest

from _util import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def kfs():
    import kfslam_b200 as k
    assert k.lib().kfs_device_count() >= 1, "no CUDA device: the product path has no CPU fallback"
    return k


def _oracle_frames(oracle, frames, nf):
    P = oracle.orb_params(nfeatures=nf)
    return [oracle.detect_and_compute(f, None, P) for f in frames]


def _check_batch(oracle, out, B, ref, kf_desc_of, ratio=0.6):
    """ref[i] = oracle (kp, desc) of frame i; kf_desc_of(i) = oracle descriptors of frame i's keyframe or None."""
    for i in range(B):
        n = int(out["n_kp"][i])
        rk, rd = ref[i]
        assert n == len(rk)
        assert np.array_equal(out["kp"][i, :n], rk) and np.array_equal(out["desc"][i, :n], rd), "frame %d extraction differs" % i
        q = kf_desc_of(i)
        if q is None:
            assert out["n_query"][i] == 0 and out["n_good"][i] == 0
            continue
        rm, rc = oracle.knn_match(q, rd, 2)
        assert out["n_query"][i] == len(q)
        assert np.array_equal(out["knn"][i, :len(q)], rm), "frame %d kNN differs" % i
        good = oracle.ratio_test(rm, rc, ratio)
        assert out["n_good"][i] == len(good)
        assert np.array_equal(out["good"][i, :len(good)], good), "frame %d ratio-test survivors differ" % i


def test_replay_in_batch_keyframes_and_double_buffering(kfs, oracle):
    B, nf = 6, 500
    NB = 5
    frames = [np.ascontiguousarray(synth.make_stream(B, 240, 320, seed=30 + s)) for s in range(NB)]
    rp = kfs.Replay(nfeatures=nf, max_batch=B, keyframe_slots=2, ratio=0.6)
    assert kfs.lib().kfs_replay_arenas(rp._h) == 3
    plan = [(0, -1)] * B  # every frame (the keyframe itself included) against frame 0 of its batch
    outs = [rp.alloc_outputs(B) for _ in range(NB)]
    tickets = [rp.submit(frames[s], plan, outs[s]) for s in range(NB)]   # three batches in flight; the 4th and 5th submit
    for t in tickets:                                                     # complete the oldest ones internally
        assert rp.wait(t) == 0
    for s in range(NB):
        ref = _oracle_frames(oracle, frames[s], nf)
        _check_batch(oracle, outs[s], B, ref, lambda i: ref[0][1])
    assert rp.last_launches > 0
    rp.close()


def test_replay_keyframe_cache_across_batches(kfs, oracle):
    B, nf = 4, 400
    stream = np.ascontiguousarray(synth.make_stream(2 * B, 200, 333, seed=41))
    rp = kfs.Replay(nfeatures=nf, max_batch=B, keyframe_slots=3, ratio=0.6)
    ref = _oracle_frames(oracle, stream, nf)
    # batch 0: frame 0 is the keyframe (stored in slot 1); frame 3 becomes the next keyframe (slot 2) and is NOT matched
    plan0 = [(kfs.Replay.NO_MATCH, 1), (0, -1), (0, -1), (0, 2)]
    out0 = rp.alloc_outputs(B)
    rp.wait(rp.submit(stream[:B], plan0, out0))
    _check_batch(oracle, out0, B, ref[:B], lambda i: None if i == 0 else ref[0][1])
    # batch 1: frames 4,5 against cached slot 1 (= frame 0), frames 6,7 against cached slot 2 (= frame 3);
    # frame 5 overwrites slot 1 in this batch, which must not affect this batch's own matches
    plan1 = [(-(1 + 1), -1), (-(1 + 1), 1), (-(1 + 2), -1), (-(1 + 2), -1)]
    out1 = rp.alloc_outputs(B)
    rp.wait(rp.submit(stream[B:], plan1, out1))
    _check_batch(oracle, out1, B, ref[B:], lambda i: ref[0][1] if i < 2 else ref[3][1])
    # batch 2: slot 1 now holds frame 5
    out2 = rp.alloc_outputs(B)
    rp.wait(rp.submit(stream[:B], [(-(1 + 1), -1)] * B, out2))
    _check_batch(oracle, out2, B, ref[:B], lambda i: ref[5][1])
    rp.close()


def test_replay_rejects_bad_plans(kfs):
    rp = kfs.Replay(nfeatures=100, max_batch=2, keyframe_slots=1)
    frames = np.zeros((2, 120, 160), np.uint8)
    out = rp.alloc_outputs(2)
    for plan in ([(2, -1), (0, -1)], [(0, 1), (0, -1)], [(-3, -1), (0, -1)]):
        with pytest.raises(kfs.KfsError):
            rp.submit(frames, plan, out)
    with pytest.raises(kfs.KfsError):
        rp.submit(np.zeros((3, 120, 160), np.uint8), [(0, -1)] * 3, rp.alloc_outputs(3))
    rp.close()


@pytest.mark.parametrize("chunk", [1, 2, 3, 7, 0])
def test_chunked_host_pipeline_equals_unchunked(kfs, oracle, chunk):
    """kfs_orb_detect_and_compute_batch cuts the batch into chunks over three streams; any chunking = same result."""
    B, nf = 9, 300
    frames = np.ascontiguousarray(synth.make_stream(B, 200, 333, seed=52))
    masks = np.zeros_like(frames)
    masks[:, 30:170, 30:300] = 255
    masks[4] = 0
    orb = kfs.ORB_create(nf)
    orb.set_pipeline_chunk(chunk)
    P = oracle.orb_params(nfeatures=nf)
    for use_mask in (False, True):
        kps, descs = orb.detectAndComputeBatch(frames, masks if use_mask else None)
        for f in range(B):
            rk, rd = oracle.detect_and_compute(frames[f], masks[f] if use_mask else None, P)
            assert np.array_equal(kps[f], rk) and np.array_equal(descs[f], rd), (chunk, use_mask, f)


def test_tapered_chunk_schedule_of_large_batches(kfs, oracle):
    """Batches of 32 frames and more take the tapered automatic schedule (three quarter-size chunks, then 1/8, 1/16, 1/16 of the
    batch): same rows as the oracle for every frame, with and without masks, for a batch that the taper divides unevenly."""
    nf = 200
    P = oracle.orb_params(nfeatures=nf)
    orb = kfs.ORB_create(nf)
    for B in (32, 45):
        frames = np.ascontiguousarray(synth.make_stream(B, 120, 173, seed=60 + B, pad=8))
        masks = np.zeros_like(frames)
        masks[:, 10:100, 20:160] = 255
        for use_mask in (False, True):
            kps, descs = orb.detectAndComputeBatch(frames, masks if use_mask else None)
            for f in range(B):
                rk, rd = oracle.detect_and_compute(frames[f], masks[f] if use_mask else None, P)
                assert np.array_equal(kps[f], rk) and np.array_equal(descs[f], rd), (B, use_mask, f)
