"""Every A/B switch of the extraction chain (CUDA graphs, TMA tile loads, one-launch pyramid, half-height FAST
tiles, programmatic dependent launch) must leave the results untouched: the mechanisms differ, the bytes do not."""
import os

import numpy as np
import pytest

from _util import synth

pytestmark = pytest.mark.gpu

SWITCHES = ["KFS_ORB_GRAPHS", "KFS_ORB_TMA", "KFS_ORB_FUSED_PYRAMID", "KFS_ORB_FAST_HALF", "KFS_ORB_PDL", "KFS_ORB_DESCRIBE2", "KFS_ORB_RESIZE_TMA"]


@pytest.fixture(scope="module")
def kfs():
    import kfslam_b200 as k
    assert k.lib().kfs_device_count() >= 1, "no CUDA device: the product path has no CPU fallback"
    return k


def extract(kfs, frames, env):
    saved = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        orb = kfs.ORB_create(1200)
        out = []
        for rep in range(3):                       # eager, capture, replay
            out.append([orb.detectAndCompute(f) for f in frames[:2]])
        out.append(orb.detectAndComputeBatch(frames))
        launches = orb.last_launches
        orb.close()
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return out, launches


def same(a, b):
    for x, y in zip(a[:3], b[:3]):
        for (kx, dx), (ky, dy) in zip(x, y):
            assert np.array_equal(kx, ky) and np.array_equal(dx, dy)
    for f in range(len(a[3][0])):
        assert np.array_equal(a[3][0][f], b[3][0][f]) and np.array_equal(a[3][1][f], b[3][1][f])


def test_switches_do_not_change_results(kfs):
    frames = np.ascontiguousarray(synth.make_stream(5, 480, 640, seed=31))
    ref, _ = extract(kfs, frames, {})
    assert len(ref[0][0][0]) > 500
    for name in SWITCHES:
        got, _ = extract(kfs, frames, {name: "0"})
        same(ref, got)
    got, _ = extract(kfs, frames, {name: "0" for name in SWITCHES})   # everything off at once: the plain chain
    same(ref, got)
    got, _ = extract(kfs, frames, {"KFS_ORB_FUSED_PYRAMID": "2"})     # one-launch pyramid for the batch as well
    same(ref, got)


def _knn_with_env(kfs, q, t, env):
    saved = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)      # KFS_BF_SPLIT / KFS_BF_CLUSTER are read when the handle is created
        out = [bf.knnMatch(q, t, 2) for _ in range(2)]   # twice: the arrival counters of the global merge must return to zero
        out.append(bf.knnMatchPairs([(q, t), (q[:130], t), (q, t[:129])], 2))
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return out


@pytest.mark.parametrize("nq,nt", [(300, 2000), (2000, 2000), (700, 100), (129, 1025)])
def test_knn_split_and_cluster_merge_do_not_change_results(kfs, nq, nt):
    """Train-row splits over CTAs (KFS_BF_SPLIT = 1..16) and both ways of merging them — a thread-block cluster's distributed
    shared memory (default) or global scratch + last arriver (KFS_BF_CLUSTER=0) — return the oracle's rows."""
    from oracle import oracle as O
    q = synth.random_descriptors(nq, seed=3 * nq + 1)
    t = synth.random_descriptors(nt, seed=5 * nt + 2)
    rm, rc = O.knn_match(q, t, 2)
    for cluster in ("2", "1", "0"):
        for split in (None, "1", "2", "3", "5", "8", "13", "16"):
            env = {"KFS_BF_CLUSTER": cluster}
            if split:
                env["KFS_BF_SPLIT"] = split
            a, b, pairs = _knn_with_env(kfs, q, t, env)
            for m, c in (a, b, pairs[0]):
                assert np.array_equal(c, rc), (cluster, split)
                assert np.array_equal(m["trainIdx"], rm["trainIdx"]) and np.array_equal(m["distance"], rm["distance"]), (cluster, split)
            m1, c1 = pairs[1]
            assert np.array_equal(m1["trainIdx"], rm["trainIdx"][:130]) and np.array_equal(m1["distance"], rm["distance"][:130])
            r2, _ = O.knn_match(q, t[:129], 2)
            assert np.array_equal(pairs[2][0]["trainIdx"], r2["trainIdx"]) and np.array_equal(pairs[2][0]["distance"], r2["distance"])


@pytest.mark.parametrize("live_nt", [0, 1, 3, 130, 2047])
def test_knn_cluster_splits_with_tiny_device_side_train_count(kfs, live_nt):
    """the launch is sized from the host-side capacity (2048 train rows -> 16 splits), the kernel clamps to the device-side
    count, so with few live rows most CTAs of a cluster have NO train rows and reach the exchange through rank 0's shared memory
    right after they start.  Rows must still be the oracle's, for the cluster merge and the global one."""
    import torch
    from oracle import oracle as O
    N = kfs._native
    nq, cap = 300, 2048
    q = synth.random_descriptors(nq, seed=77)
    t = synth.random_descriptors(cap, seed=78)
    dev = torch.device("cuda:0")
    dq, dt = torch.from_numpy(q).to(dev), torch.from_numpy(t).to(dev)
    dn = torch.tensor([live_nt], dtype=torch.int32, device=dev)
    out = torch.zeros(nq * 2 * 16, dtype=torch.uint8, device=dev)
    cnt = torch.zeros(nq, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    rm, rc = O.knn_match(q, t[:live_nt], 2)
    for cluster in ("2", "1", "0"):
        saved = {k: os.environ.get(k) for k in ("KFS_BF_CLUSTER", "KFS_BF_SPLIT")}
        os.environ.update({"KFS_BF_CLUSTER": cluster, "KFS_BF_SPLIT": "16"})
        try:
            bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
        finally:
            for k, v in saved.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v
        for _ in range(3):
            out.zero_(); cnt.fill_(-1)
            torch.cuda.synchronize()
            bf.knnMatchPairsDevice([dict(query=dq.data_ptr(), nq=nq, train=dt.data_ptr(), nt=cap, d_nt=dn.data_ptr(), out=out.data_ptr(),
                                         counts=cnt.data_ptr())], 2)
            bf.sync()
            m = out.cpu().numpy().view(N.DM_DTYPE).reshape(nq, 2)
            assert np.array_equal(cnt.cpu().numpy(), rc), cluster
            live = np.arange(2)[None, :] < rc[:, None]      # row q holds rc[q] = min(2, live_nt) entries; the tail is {q, -1, -1, FLT_MAX}
            assert np.array_equal(m["trainIdx"][live], rm["trainIdx"][live]) and np.array_equal(m["distance"][live], rm["distance"][live]), cluster
            assert (m["trainIdx"][~live] == -1).all() and (m["queryIdx"] == np.arange(nq)[:, None]).all(), cluster
        bf.close()
