"""k_pyramid (the whole pyramid in one launch) against the per-level resize kernels: every level's bytes,
and the keypoints / descriptors that come out of the chain, have to be identical for any size, scale factor, level count,
batch and mask.  KFS_ORB_FUSED_PYRAMID is read when a handle plans its arena: 0 = per-level kernels, 2 = one launch for any
batch (the default, 1, uses it for calls of at most 4 frames)."""
import os

import numpy as np
import pytest

from _util import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def kfs():
    import kfslam_b200 as k
    assert k.lib().kfs_device_count() >= 1, "no CUDA device: the product path has no CPU fallback"
    return k


def run(kfs, mode, frames, masks, nfeatures, scale, nlevels):
    old = os.environ.get("KFS_ORB_FUSED_PYRAMID")
    os.environ["KFS_ORB_FUSED_PYRAMID"] = str(mode)
    try:
        orb = kfs.ORB_create(nfeatures, scale, nlevels)
        kps, descs = orb.detectAndComputeBatch(frames, masks)
        nl = orb.params.nlevels
        levels = [[orb.download_level(f, l).copy() for l in range(nl) if orb.level_info(l)[0] > 0] for f in range(len(frames))]
        launches = orb.last_launches
        orb.close()
    finally:
        if old is None:
            os.environ.pop("KFS_ORB_FUSED_PYRAMID", None)
        else:
            os.environ["KFS_ORB_FUSED_PYRAMID"] = old
    return kps, descs, levels, launches


@pytest.mark.parametrize("h,w,scale,nlevels,batch", [
    (720, 1280, 1.2, 8, 1), (720, 1280, 1.2, 8, 6), (480, 640, 1.2, 8, 2), (241, 323, 1.2, 8, 3), (300, 401, 1.5, 5, 2),
    (480, 640, 1.1, 12, 1), (600, 800, 2.0, 4, 2), (1080, 1920, 1.2, 8, 1), (97, 131, 1.2, 4, 2), (720, 1280, 1.3, 3, 5)])
def test_fused_pyramid_equals_per_level(kfs, h, w, scale, nlevels, batch):
    frames = np.ascontiguousarray(synth.make_stream(batch, h, w, seed=h + w))
    a = run(kfs, 0, frames, None, 800, scale, nlevels)
    b = run(kfs, 2, frames, None, 800, scale, nlevels)
    for f in range(batch):
        assert len(a[2][f]) == len(b[2][f])
        for l, (x, y) in enumerate(zip(a[2][f], b[2][f])):
            assert np.array_equal(x, y), "frame %d level %d differs" % (f, l)
        assert np.array_equal(a[0][f], b[0][f]) and np.array_equal(a[1][f], b[1][f])
    assert sum(len(k) for k in a[0]) > 0


def test_fused_pyramid_with_masks(kfs):
    frames = np.ascontiguousarray(synth.make_stream(5, 480, 640, seed=9))
    masks = np.zeros((5, 480, 640), np.uint8)
    rng = np.random.default_rng(4)
    for f in range(5):
        masks[f, 20 + 7 * f:400 + 11 * f, 33 + 5 * f:600 - 9 * f] = 255
        masks[f][rng.random((480, 640)) < 0.02] = 0          # isolated holes: the resized mask keeps only exact 255s
        masks[f, 100:140, 200:260] = 7                       # non-zero, non-255 input counts as set (threshold > 0)
    a = run(kfs, 0, frames, masks, 600, 1.2, 8)
    b = run(kfs, 2, frames, masks, 600, 1.2, 8)
    for f in range(5):
        assert len(a[0][f]) > 50
        assert np.array_equal(a[0][f], b[0][f]) and np.array_equal(a[1][f], b[1][f])
        for x, y in zip(a[2][f], b[2][f]):
            assert np.array_equal(x, y)


def test_fused_pyramid_is_the_single_frame_default(kfs):
    """Default policy: one frame per call takes the one-launch pyramid (6 launches fewer than the per-level chain)."""
    if os.environ.get("KFS_ORB_TMA") == "0":
        pytest.skip("TMA switched off: the pyramid stays on the per-level kernels")
    frames = np.ascontiguousarray(synth.make_stream(1, 480, 640, seed=2))
    per_level = run(kfs, 0, frames, None, 500, 1.2, 8)[3]
    default = run(kfs, 1, frames, None, 500, 1.2, 8)[3]
    assert per_level - default == 6
