"""GPU parity (-m gpu) on images where most pixels pass FAST's quick test: the kernel's over-half-full queue path,
quotas overflowing on every level, massive score ties.  The oracle was checked against cv2 4.13 on exactly these
three images (identical keypoint fields and descriptors) when the test was written."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def dense_image(kind, h=300, w=420):
    rng = np.random.default_rng(5)
    if kind == "noise":
        return rng.integers(0, 256, (h, w), dtype=np.uint8)
    if kind == "checker2":
        yy, xx = np.mgrid[0:h, 0:w]
        return ((((yy >> 1) + (xx >> 1)) & 1) * 200 + 20).astype(np.uint8)
    img = np.zeros((h, w), np.uint8)
    img[rng.random((h, w)) < 0.08] = 255
    return img


@pytest.mark.parametrize("kind", ["noise", "checker2", "salt"])
def test_orb_dense_corner_images_vs_oracle(oracle, kind):
    import kfslam_b200 as kfs
    img = dense_image(kind)
    orb = kfs.ORB_create(1500)
    orb.set_capacity(8000)
    kp, desc = orb.detectAndCompute(img)
    rk, rd = oracle.detect_and_compute(img, None, oracle.orb_params(nfeatures=1500), capacity=8000)
    if desc is None:
        desc = np.zeros((0, 32), np.uint8)
    assert len(rk) == len(kp) and len(kp) > 400
    assert np.array_equal(kp, rk) and np.array_equal(desc, rd)
    for l in range(8):
        wl, hl, _, _ = orb.level_info(l)
        if wl <= 62 or hl <= 62:
            continue
        ref = oracle.fast9_detect(orb.download_level(0, l), 20)
        keep = (ref[:, 0] >= 31) & (ref[:, 0] < wl - 31) & (ref[:, 1] >= 31) & (ref[:, 1] < hl - 31)
        ref = ref[keep]
        ref = ref[np.lexsort((ref[:, 0], ref[:, 1]))]
        assert np.array_equal(orb.download_fast_corners(0, l), ref), (kind, l)
