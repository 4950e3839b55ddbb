"""Randomised parity sweep: frame sizes (odd ones included), ORB parameters, masks, one frame per call and small batches — the GPU
through the C ABI against the oracle, bit for bit, followed by the k=2 matcher on the descriptors just extracted (ragged row counts).
The oracle is pinned to cv2 by tests/test_oracle_golden.py and tests/test_params_variants.py; with cv2 importable on the host
(the GPU boxes have it) every fifth case is also checked against cv2 directly.
KFS_FUZZ_CASES sets the number of cases (default 24; tools/fuzz_parity.py runs a few hundred and records the outcome)."""
import os

import numpy as np
import pytest

from _util import synth

pytestmark = pytest.mark.gpu


def _sorted(kp, desc):
    o = np.lexsort((kp["x"], kp["y"], kp["octave"]))
    return kp[o], desc[o]


def run_cases(kfs, oracle, ncases, seed=2024, cv2=None, log=None, scale=1.0):
    rng = np.random.default_rng(seed)
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    stats = dict(cases=0, frames=0, keypoints=0, knn_rows=0, cv2_checked=0, cv2_rejected=0, empty=0)
    for c in range(ncases):
        w, h = int(rng.integers(70, 900) * scale), int(rng.integers(70, 700) * scale)
        if c % 7 == 0:                                   # tiny frames: levels vanish, lists stay empty
            w, h = int(rng.integers(16, 90)), int(rng.integers(16, 90))
        nf = int(rng.choice([50, 300, 500, 1000, 2000, 3000]))
        sf = float(rng.choice([1.1, 1.2, 1.2, 1.2, 1.3, 1.5, 2.0]))
        nl = int(rng.integers(1, 11))
        thr = int(rng.choice([5, 10, 20, 20, 20, 35]))
        nb = int(rng.choice([1, 1, 2, 3, 5]))
        frames = np.ascontiguousarray(synth.make_stream(nb, h, w, seed=int(rng.integers(1 << 30)), pad=8))
        mask = None
        if rng.random() < 0.35:
            mask = np.zeros((h, w), np.uint8)
            y0, x0 = int(rng.integers(0, h // 2)), int(rng.integers(0, w // 2))
            mask[y0:y0 + int(rng.integers(h // 4, h)), x0:x0 + int(rng.integers(w // 4, w))] = int(rng.choice([255, 1, 77]))
        P = oracle.orb_params(nfeatures=nf, scale_factor=sf, nlevels=nl, fast_threshold=thr)
        orb = kfs.ORB_create(nf, sf, nl, 31, 0, 2, kfs.ORB_HARRIS_SCORE, 31, thr)
        tag = "case %d: %dx%d nf=%d sf=%.1f nl=%d thr=%d nb=%d mask=%s" % (c, w, h, nf, sf, nl, thr, nb, mask is not None)
        descs = []
        if nb == 1 or mask is not None:
            got = [orb.detectAndCompute(f, mask) for f in frames]
        else:
            kps, ds = orb.detectAndComputeBatch(frames)
            got = list(zip(kps, ds))
        for f, (kp, desc) in zip(frames, got):
            rk, rd = oracle.detect_and_compute(f, mask, P)
            assert len(kp) == len(rk), tag
            if len(rk):
                assert np.array_equal(kp, rk) and np.array_equal(desc, rd), tag
            else:
                stats["empty"] += 1
            descs.append(rd if len(rk) else np.zeros((0, 32), np.uint8))
            stats["frames"] += 1
            stats["keypoints"] += len(rk)
        if cv2 is not None and c % 5 == 0 and len(descs[0]):
            try:
                ck, cd = cv2.ORB_create(nf, sf, nl, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, thr).detectAndCompute(frames[0], mask)
            except cv2.error:   # a frame so small that one of the nl levels has zero size: OpenCV asserts in resize(); the
                stats["cv2_rejected"] += 1   # library (and the oracle) skip the levels that vanish
                ck = None
        if cv2 is not None and c % 5 == 0 and len(descs[0]) and ck is not None:
            a = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave, k.class_id) for k in ck], dtype=kfs.KP_DTYPE)
            sk, sd = _sorted(a, cd)
            gk, gd = _sorted(*got[0])
            assert np.array_equal(sk, gk) and np.array_equal(sd, gd), tag + " (cv2)"
            stats["cv2_checked"] += 1
        orb.close()
        q, t = descs[0], descs[-1][::-1].copy() if len(descs[-1]) else descs[-1]
        if len(q) and len(t) >= 1:
            m, cnt = bf.knnMatch(q, t, 2)
            rm, rc = oracle.knn_match(q, t, 2)
            kk = min(2, len(t))
            assert np.array_equal(cnt, rc), tag
            assert np.array_equal(m["trainIdx"][:, :kk], rm["trainIdx"][:, :kk]) and np.array_equal(m["distance"][:, :kk], rm["distance"][:, :kk]), tag
            stats["knn_rows"] += len(q)
        stats["cases"] += 1
        if log:
            log(tag + " ok")
    bf.close()
    return stats


def test_random_parameter_and_size_sweep(oracle):
    import kfslam_b200 as kfs
    assert kfs.lib().kfs_device_count() >= 1, "no CUDA device: the product path has no CPU fallback"
    try:
        import cv2
    except Exception:
        cv2 = None
    n = int(os.environ.get("KFS_FUZZ_CASES", "24"))
    stats = run_cases(kfs, oracle, n, cv2=cv2)
    print("\n[fuzz] %s" % stats)
    assert stats["cases"] == n and stats["keypoints"] > 0


def run_track_cases(kfs, oracle, ncases, seed=99, log=None):
    """Random tracker sequences through kfs_frontend_track (one call per frame: R,G,B or gray in, keypoints + descriptors + ratio-test
    survivors + back-projected points out), keyframe promotions at random frames, a random keyframe window matched at the end —
    against detect_and_compute / knn_match / match_features / match_window of the oracle."""
    from _track_util import make_depth, make_rgb
    rng = np.random.default_rng(seed)
    stats = dict(cases=0, frames=0, matches=0, window_pairs=0)
    for c in range(ncases):
        w, h = int(rng.integers(160, 520)), int(rng.integers(120, 420))
        nf = int(rng.choice([300, 800, 1500]))
        win = int(rng.integers(1, 5))
        ratio = float(rng.choice([0.6, 0.75, 0.9]))
        K = np.array([[500.0 + rng.random() * 50, 0, w / 2 + rng.random()], [0, 500.0 + rng.random() * 50, h / 2 + rng.random()], [0, 0, 1]])
        gray_in = rng.random() < 0.3
        mask = None
        if rng.random() < 0.4:
            mask = np.zeros((h, w), np.uint8)
            mask[h // 8:h - h // 8, w // 10:w - w // 12] = 255
        fe = kfs.TrackerFrontEnd(nfeatures=nf)
        fe.setWindow(win)
        P = oracle.orb_params(nfeatures=nf)
        tag = "track case %d: %dx%d nf=%d win=%d ratio=%.2f gray=%s mask=%s" % (c, w, h, nf, win, ratio, gray_in, mask is not None)
        s0 = int(rng.integers(1, 1 << 20))
        kf, kfs_seen, shift = None, [], 0
        nframes = int(rng.integers(3, 7))
        for i in range(nframes):
            shift += int(rng.integers(1, 5))
            rgb = make_rgb(h, w, s0, shift=shift)
            img = oracle.rgb2gray(rgb) if gray_in else rgb
            kp, desc, p3, p2, m = fe.track(img, K, mask=mask, ratio=ratio)
            rk, rd = oracle.detect_and_compute(oracle.rgb2gray(rgb), mask, P)
            assert np.array_equal(kp, rk) and np.array_equal(desc, rd), tag
            if kf is None:
                assert m is None, tag
            else:
                knn, _ = oracle.knn_match(kf[1], rd, 2)
                r3, r2, rm = oracle.match_features(knn, kf[0], rk, kf[2], K, ratio)
                assert np.array_equal(m, rm) and np.array_equal(p3, r3) and np.array_equal(p2, r2), tag
                stats["matches"] += len(rm)
            if kf is None or rng.random() < 0.4:
                depth = make_depth(h, w, int(rng.integers(1 << 20)))
                fe.promoteKeyframe(depth)
                kf = (rk, rd, depth)
                kfs_seen.append(rd)
            stats["frames"] += 1
        good = fe.matchWindow(None, ratio)                 # the last frame against the window, oldest keyframe first
        _, ref = oracle.match_window(kfs_seen, rd, win, ratio)
        assert len(good) == len(ref), tag
        for a, b in zip(good, ref):
            assert np.array_equal(a, b), tag
        stats["window_pairs"] += len(ref)
        fe.close()
        stats["cases"] += 1
        if log:
            log(tag + " ok")
    return stats


def test_random_tracker_sequences(oracle):
    import kfslam_b200 as kfs
    n = int(os.environ.get("KFS_FUZZ_TRACK_CASES", "6"))
    stats = run_track_cases(kfs, oracle, n)
    print("\n[fuzz track] %s" % stats)
    assert stats["cases"] == n and stats["matches"] > 0


def run_replay_cases(kfs, oracle, ncases, seed=5, log=None):
    """Random plans through the replay executor: several submits per case (batches of random length in flight over the arenas), every
    frame matched against a random earlier frame of its submit, a random cache slot an earlier submit filled, or nothing
    (KFS_REPLAY_NO_MATCH); random store_slot writes (a slot written in a submit must not affect that submit's own matches), random
    KFS_REPLAY_SKIP_ROWS flags, gray or R,G,B frames, optional masks.  A host model of the cache gives the oracle its keyframes."""
    from _track_util import make_rgb
    rng = np.random.default_rng(seed)
    stats = dict(cases=0, submits=0, frames=0, knn_rows=0, survivors=0)
    for c in range(ncases):
        w, h = int(rng.integers(120, 420)), int(rng.integers(100, 320))
        nf = int(rng.choice([200, 500, 900]))
        mb = int(rng.integers(2, 9))
        slots = int(rng.integers(1, 4))
        ratio = float(rng.choice([0.6, 0.75]))
        arenas = int(rng.choice([0, 2, 3, 4]))
        rgb = rng.random() < 0.3
        use_mask = rng.random() < 0.3
        rp = kfs.Replay(nfeatures=nf, max_batch=mb, keyframe_slots=slots, ratio=ratio, arenas=arenas)
        P = oracle.orb_params(nfeatures=nf)
        tag = "replay case %d: %dx%d nf=%d max_batch=%d slots=%d ratio=%.2f arenas=%d rgb=%s mask=%s" % (c, w, h, nf, mb, slots, ratio, arenas, rgb, use_mask)
        cache = [None] * slots                       # host model: descriptors a slot holds
        pending = []
        nsub = int(rng.integers(2, 6))
        for s_ in range(nsub):
            B = int(rng.integers(1, mb + 1))
            s0 = int(rng.integers(1 << 20))
            if rgb:
                frames = np.ascontiguousarray(np.stack([make_rgb(h, w, s0, shift=2 * i) for i in range(B)]))
                grays = [oracle.rgb2gray(f) for f in frames]
            else:
                frames = np.ascontiguousarray(synth.make_stream(B, h, w, seed=s0, pad=8))
                grays = list(frames)
            masks = None
            if use_mask:
                masks = np.zeros((B, h, w), np.uint8)
                masks[:, h // 10:h - h // 7, w // 9:w - w // 8] = 255
            ref = [oracle.detect_and_compute(grays[i], masks[i] if use_mask else None, P) for i in range(B)]
            plan, kf_desc = [], []
            for i in range(B):
                choices = ["none"] + (["batch"] if i > 0 else []) + (["cache"] if any(x is not None for x in cache) else [])
                pick = str(rng.choice(choices))
                if pick == "batch":
                    k = int(rng.integers(0, i)); plan_k = k; q = ref[k][1]
                elif pick == "cache":
                    sl = int(rng.choice([j for j in range(slots) if cache[j] is not None])); plan_k = -(1 + sl); q = cache[sl]
                else:
                    plan_k = kfs.Replay.NO_MATCH; q = None
                plan.append([plan_k, -1, 1 if rng.random() < 0.25 else 0])
                kf_desc.append(q)
            used = set()
            for i in range(B):                          # stores take effect AFTER this submit's matching
                if rng.random() < 0.3:
                    sl = int(rng.integers(0, slots))
                    if sl not in used:
                        plan[i][1] = sl; used.add(sl)
            out = rp.alloc_outputs(B)
            pending.append((rp.submit(frames, [tuple(e) for e in plan], out, masks), out, ref, kf_desc, plan, B))
            for i in range(B):
                if plan[i][1] >= 0:
                    cache[plan[i][1]] = ref[i][1]
            stats["submits"] += 1
        for t, out, ref, kf_desc, plan, B in pending:
            assert rp.wait(t) == 0, tag
            for i in range(B):
                rk, rd = ref[i]
                assert int(out["n_kp"][i]) == len(rk), tag
                if not plan[i][2]:                      # rows of a SKIP_ROWS frame stay on the device
                    assert np.array_equal(out["kp"][i, :len(rk)], rk) and np.array_equal(out["desc"][i, :len(rk)], rd), tag
                q = kf_desc[i]
                if q is None:
                    assert out["n_query"][i] == 0 and out["n_good"][i] == 0, tag
                    continue
                rm, rc = oracle.knn_match(q, rd, 2)
                assert out["n_query"][i] == len(q), tag
                kk = min(2, len(rd))
                if len(q) and kk:
                    assert np.array_equal(out["knn"][i, :len(q)]["trainIdx"][:, :kk], rm["trainIdx"][:, :kk]), tag
                    assert np.array_equal(out["knn"][i, :len(q)]["distance"][:, :kk], rm["distance"][:, :kk]), tag
                good = oracle.ratio_test(rm, rc, ratio)
                assert out["n_good"][i] == len(good) and np.array_equal(out["good"][i, :len(good)], good), tag
                stats["knn_rows"] += len(q); stats["survivors"] += len(good)
                stats["frames"] += 1
        rp.close()
        stats["cases"] += 1
        if log:
            log(tag + " ok")
    return stats


def test_random_replay_plans(oracle):
    import kfslam_b200 as kfs
    n = int(os.environ.get("KFS_FUZZ_REPLAY_CASES", "8"))
    stats = run_replay_cases(kfs, oracle, n)
    print("\n[fuzz replay] %s" % stats)
    assert stats["cases"] == n and stats["knn_rows"] > 0
