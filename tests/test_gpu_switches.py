"""Every A/B switch of the extraction chain (CUDA graphs, TMA tile loads, one-launch pyramid, half-height FAST
tiles, programmatic dependent launch) must leave the results untouched: the mechanisms differ, the bytes do not."""
import os

import numpy as np
import pytest

from _util import synth

pytestmark = pytest.mark.gpu

SWITCHES = ["KFS_ORB_GRAPHS", "KFS_ORB_TMA", "KFS_ORB_FUSED_PYRAMID", "KFS_ORB_FAST_HALF", "KFS_ORB_PDL"]


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
