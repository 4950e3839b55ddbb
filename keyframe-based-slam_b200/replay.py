"""Offline-replay sharding for the front-end (BASELINE.json configs[2..4]).

Frames are independent for extraction and every (frame, keyframe) pair is independent for matching, so the
multi-GPU story is a static partition with NO collective on the data path: one process per GPU, each with its
own kfs_orb / kfs_bf handles, each owning a contiguous run of keyframe groups so that a frame and the keyframe
it is matched against (VisualOdometry.cpp:36-39 `GetLastKeyFrame`) live on the same GPU.  torch.distributed is
used only for the barrier around the timed region and the MAX-over-ranks of the device time.
"""
from typing import List, Tuple


def keyframe_of(i: int, kf_every: int = 10) -> int:
    """Index of the keyframe frame i is matched against: the most recent keyframe strictly before it
    (keyframes are frames 0, kf_every, 2*kf_every, ... — the '>= 10 frames since last KF' rule of
    VisualOdometry.cpp:105 with the motion terms always true); frame 0 is matched against itself."""
    k = (i // kf_every) * kf_every
    if k == i:
        k = max(0, i - kf_every)
    return k


def shard_frames(n_frames: int, world: int, rank: int, kf_every: int = 10) -> Tuple[int, int, int]:
    """Contiguous shard [lo, hi) of frames for `rank`, cut on keyframe-group boundaries, plus `halo_kf`:
    the index of the one keyframe of the previous shard that must ALSO be extracted locally (frames that are
    themselves keyframes are matched against the previous keyframe), or -1.  Shards are disjoint, cover
    [0, n_frames) and differ by at most one group in size."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    groups = (n_frames + kf_every - 1) // kf_every
    base, rem = divmod(groups, world)
    g_lo = rank * base + min(rank, rem)
    g_hi = g_lo + base + (1 if rank < rem else 0)
    lo, hi = min(g_lo * kf_every, n_frames), min(g_hi * kf_every, n_frames)
    halo = lo - kf_every if (lo > 0 and hi > lo) else -1
    return lo, hi, halo


def shard_pairs(n_pairs: int, world: int, rank: int) -> List[int]:
    """Round-robin shard of independent (query, train) problems (window matching, configs[3..4])."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    return list(range(rank, n_pairs, world))


def local_pairs(lo: int, hi: int, kf_every: int = 10) -> List[Tuple[int, int]]:
    """(frame, keyframe) pairs a shard owns; every keyframe index is either inside [lo, hi) or the halo."""
    return [(i, keyframe_of(i, kf_every)) for i in range(lo, hi)]


def max_over_ranks(seconds: float, dist=None, device=None) -> float:
    """Device-time reduction used by bench.py: MAX over ranks (never wall clock on one rank)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return seconds
    import torch
    t = torch.tensor([seconds], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, dist=None, device=None) -> float:
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return value
    import torch
    t = torch.tensor([value], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
