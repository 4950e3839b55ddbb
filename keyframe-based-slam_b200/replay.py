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


NO_MATCH = -2 ** 31   # kfs_replay_frame.keyframe: extract only (KFS_REPLAY_NO_MATCH)
SKIP_ROWS = 1         # kfs_replay_frame.flags: keypoints / descriptors of this frame stay on the device (KFS_REPLAY_SKIP_ROWS)


def plan_shard(lo: int, hi: int, halo: int, batch: int, kf_every: int = 10, keyframes_only_rows: bool = False):
    """Cuts a shard [lo, hi) (+ its halo keyframe, shard_frames) into replay-executor submits of at most `batch` frames.

    Returns a list of dicts, one per submit:
      frames  global frame indices the submit carries, in order (the halo keyframe, when there is one, leads the first)
      plan    [(keyframe, store_slot, flags)] per carried frame, in kfs_replay_frame terms: keyframe >= 0 is an index INTO
              THIS SUBMIT, -(1 + slot) a keyframe an earlier submit left in the device cache, NO_MATCH for the halo (its
              matches belong to the previous shard); store_slot parks the submit's last keyframe for the next submit
      owned   per carried frame: True when the shard owns the frame's results (False only for the halo)
    Two cache slots are enough: the only earlier keyframe a submit can need is the last keyframe before its first frame
    (VisualOdometry.cpp:36-39: a frame is matched against the most recent keyframe; a keyframe against the one before it).
    With keyframes_only_rows the rows of non-keyframes are flagged SKIP_ROWS (the tracker keeps keypoints / descriptors of
    keyframes only, VisualOdometry.cpp:155-165)."""
    if batch < 2:
        raise ValueError("batch must be >= 2")
    seq = ([halo] if halo >= 0 else []) + list(range(lo, hi))
    out, slot_of, next_slot = [], {}, 0
    for b0 in range(0, len(seq), batch):
        frames = seq[b0:b0 + batch]
        pos = {f: j for j, f in enumerate(frames)}
        kfs_here = [f for f in frames if f % kf_every == 0]
        # the keyframe the NEXT submit may need: the last one seen so far
        park = kfs_here[-1] if (kfs_here and b0 + batch < len(seq)) else None
        plan, owned = [], []
        for f in frames:
            own = not (f == halo and halo >= 0)
            k = keyframe_of(f, kf_every)
            if not own:
                kf = NO_MATCH
            elif k in pos:
                kf = pos[k]
            elif k in slot_of:
                kf = -(1 + slot_of[k])
            else:
                raise AssertionError("frame %d needs keyframe %d, which is neither in its submit nor parked" % (f, k))
            store = -1
            if f == park:
                store = next_slot
            flags = SKIP_ROWS if (not own or (keyframes_only_rows and f % kf_every != 0)) else 0
            plan.append((kf, store, flags))
            owned.append(own)
        if park is not None:
            slot_of = {k: s for k, s in slot_of.items() if s != next_slot}
            slot_of[park] = next_slot
            next_slot ^= 1
        out.append(dict(frames=frames, plan=plan, owned=owned))
    return out


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
