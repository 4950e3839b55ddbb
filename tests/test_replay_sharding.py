"""CPU suite: host-side sharding logic of the multi-GPU replay path, including a world_size-2 gloo run
(the N>1 path of bench.py minus the kernels: shard -> process -> SUM frames / MAX time, no data-path collective)."""
import os
import socket
import sys

import pytest

from _util import ROOT

import kfslam_b200 as kfs
from kfslam_b200 import replay


def test_keyframe_of():
    assert [replay.keyframe_of(i) for i in (0, 1, 9, 10, 11, 19, 20, 25)] == [0, 0, 0, 0, 10, 10, 10, 20]
    assert replay.keyframe_of(7, 5) == 5 and replay.keyframe_of(5, 5) == 0


@pytest.mark.parametrize("n,world", [(10000, 1), (10000, 2), (10000, 4), (10000, 8), (95, 8), (7, 4), (0, 2), (64, 3)])
def test_shards_cover_and_colocate(n, world):
    seen = []
    for r in range(world):
        lo, hi, halo = replay.shard_frames(n, world, r)
        assert 0 <= lo <= hi <= n
        seen += list(range(lo, hi))
        local = set(range(lo, hi)) | ({halo} if halo >= 0 else set())
        for i, k in replay.local_pairs(lo, hi):
            assert k in local, (i, k, lo, hi, halo)   # a frame's keyframe is extracted on the same GPU
    assert seen == list(range(n))                      # disjoint, ordered cover
    sizes = [replay.shard_frames(n, world, r)[1] - replay.shard_frames(n, world, r)[0] for r in range(world)]
    assert max(sizes) - min(sizes) < 20     # at most one group, plus the ragged last group


def test_shard_pairs_round_robin():
    got = sorted(sum((replay.shard_pairs(32, 8, r) for r in range(8)), []))
    assert got == list(range(32))
    assert replay.shard_pairs(10, 4, 1) == [1, 5, 9]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import kfslam_b200 as k
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi, halo = k.replay.shard_frames(1000, world, rank)
    pairs = k.replay.local_pairs(lo, hi)
    dist.barrier()
    fake_seconds = 1.0 + rank          # rank-dependent "device time"
    t = k.replay.max_over_ranks(fake_seconds, dist)
    total = k.replay.sum_over_ranks(float(len(pairs)), dist)
    dist.barrier()
    q.put((rank, lo, hi, halo, t, total))
    dist.destroy_process_group()


def test_gloo_world2_replay_reduction():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, lo0, hi0, h0, t0, n0), (r1, lo1, hi1, h1, t1, n1) = res
    assert (lo0, hi0, lo1, hi1) == (0, 500, 500, 1000) and h0 == -1 and h1 == 490
    assert t0 == t1 == 2.0             # MAX over ranks
    assert n0 == n1 == 1000.0          # every frame processed exactly once, whole-job count


def test_abi_exports_every_declared_symbol():
    """The C-ABI library loads and exports every symbol include/kfslam_frontend.h declares (no compute calls)."""
    import ctypes, re
    hdr = open(os.path.join(ROOT, "include", "kfslam_frontend.h")).read()
    declared = set(re.findall(r"KFS_API\s+[\w\s\*]+?\b(kfs_\w+)\s*\(", hdr))
    assert len(declared) >= 30
    lib = ctypes.CDLL(os.path.join(ROOT, "keyframe-based-slam_b200", "libkfslam_frontend.so"))
    for name in declared:
        assert hasattr(lib, name), "missing export: " + name
    N = kfs._native
    assert declared == set(N.SIGNATURES), declared ^ set(N.SIGNATURES)
    assert kfs.lib().kfs_version() >= 100
    assert ctypes.sizeof(N.OrbParams) == 36 and N.KP_DTYPE.itemsize == 28 and N.DM_DTYPE.itemsize == 16


def test_no_device_fails_loudly():
    """No CPU fallback: on a box without a GPU every constructor refuses (KFS_ERR_NO_DEVICE)."""
    if kfs.lib().kfs_device_count() > 0:
        pytest.skip("GPU present")
    with pytest.raises(kfs.KfsError) as e:
        kfs.ORB_create(500)
    assert e.value.status == 5
    with pytest.raises(kfs.KfsError) as e:
        kfs.BFMatcher_create(kfs.NORM_HAMMING)
    assert e.value.status == 5


def test_argument_validation_without_device():
    with pytest.raises(kfs.KfsError) as e:
        kfs.ORB_create(500, 1.2, 8, 31, 0, 3)      # WTA_K=3: unsupported is reported before the device probe
    assert e.value.status == 2
    with pytest.raises(kfs.KfsError) as e:
        kfs.ORB_create(0)
    assert e.value.status == 1
    with pytest.raises(kfs.KfsError) as e:
        kfs.BFMatcher_create(4)
    assert e.value.status == 2


@pytest.mark.parametrize("n,world,batch", [(10000, 1, 128), (10000, 8, 128), (95, 2, 7), (45, 2, 16), (45, 2, 5), (33, 1, 4), (200, 3, 10)])
def test_plan_shard_resolves_every_keyframe(n, world, batch):
    """replay.plan_shard: every owned frame is matched against keyframe_of(frame) — through an in-submit index or a cache slot an
    EARLIER submit stored (simulated here) — the halo leads the first submit unmatched, owned frames cover the shard in order."""
    for r in range(world):
        lo, hi, halo = replay.shard_frames(n, world, r)
        subs = replay.plan_shard(lo, hi, halo, batch)
        cache, seen = {}, []
        for s in subs:
            assert 0 < len(s["frames"]) <= batch
            for f, (kf, st, fl), own in zip(s["frames"], s["plan"], s["owned"]):
                if own:
                    seen.append(f)
                    got = s["frames"][kf] if kf >= 0 else cache[-(kf + 1)]
                    assert got == replay.keyframe_of(f) and fl == 0
                else:
                    assert f == halo and kf == replay.NO_MATCH and fl == replay.SKIP_ROWS
            for f, (kf, st, fl) in zip(s["frames"], s["plan"]):
                assert st in (-1, 0, 1)
                if st >= 0:
                    cache[st] = f
        assert seen == list(range(lo, hi))
    rows = replay.plan_shard(0, 30, -1, 16, keyframes_only_rows=True)
    assert [fl for s in rows for (_, _, fl) in s["plan"]] == [0 if f % 10 == 0 else replay.SKIP_ROWS for f in range(30)]
