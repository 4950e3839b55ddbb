"""CPU suite: pins the oracle (oracle/*.c) against golden vectors made by the real reference code
(cv2 4.13.0 = OpenCV behind VisualOdometry.cpp:30,39) — tools/make_golden.py."""
import numpy as np
import pytest

from _util import ORB_GOLDENS, golden, golden_input, sha, compare_keypoints


@pytest.mark.parametrize("name", ORB_GOLDENS)
def test_orb_oracle_matches_cv2_golden(oracle, name):
    g = golden(name)
    gray, mask = golden_input(name)
    meta = g["meta"]
    kp, desc, pyr = oracle.detect_and_compute(gray, mask, oracle.orb_params(nfeatures=meta["nfeatures"]), want_pyramid=True)
    # pyramid pixels bit-exact (a1.1)
    assert [sha(p) for p in pyr] == meta["pyr_sha"]
    # keypoints: identical set, identical fields, identical descriptors (a1.3-a1.10)
    assert len(kp) == meta["n"]
    r = compare_keypoints(g["kp"], g["desc"], kp, desc)
    assert r["n_common"] == meta["n"] and r["recall"] == 1.0
    if meta["n"]:
        assert r["exact_fields"] == 6
        assert r["angle_exact"] == meta["n"]
        assert r["desc_equal"] == meta["n"]
        # oracle output order is (octave, y, x) == the golden's order
        assert np.array_equal(kp, g["kp"]) and np.array_equal(desc, g["desc"])


@pytest.mark.parametrize("name", ["cfg1_480p_a", "ties_periodic", "small_64x64", "odd_333x257"])
def test_fast_oracle_matches_cv2_golden(oracle, name):
    meta = golden(name)["meta"]
    gray, _ = golden_input(name)
    for l, img in enumerate(oracle.pyramid(gray)):
        if img.shape[0] < 7 or img.shape[1] < 7:
            continue
        c = oracle.fast9_detect(img, 20)
        c = np.array(sorted((int(y), int(x), int(s)) for x, y, s in c.tolist()), np.int32).reshape(-1, 3)
        assert len(c) == meta["fast_counts"][l]
        assert sha(c) == meta["fast_sha"][l]


def test_pyramid_pixels_golden(oracle):
    g = golden("cfg1_480p_a")
    gray, _ = golden_input("cfg1_480p_a")
    pyr = oracle.pyramid(gray)
    assert np.array_equal(pyr[1], g["pyr1"]) and np.array_equal(pyr[7], g["pyr7"])


def test_geometry_tables(oracle):
    # SURVEY §8 header [probe]
    assert oracle.level_sizes(1280, 720) == [(1280, 720), (1067, 600), (889, 500), (741, 417), (617, 347), (514, 289), (429, 241), (357, 201)]
    assert oracle.level_sizes(3840, 2160)[-1] == (1072, 603)
    assert oracle.level_quotas(1000) == [217, 181, 151, 126, 105, 87, 73, 60]
    assert oracle.level_quotas(2000) == [434, 362, 302, 251, 209, 175, 145, 122]
    assert oracle.level_quotas(3000) == [652, 543, 452, 377, 314, 262, 218, 182]
    assert oracle.level_quotas(10000) == [2172, 1810, 1508, 1257, 1047, 873, 727, 606]


def test_knn_oracle_matches_cv2_golden(oracle):
    z = golden("knn_cases")
    for case in ("rand", "ties", "nt1", "nt2", "nq1"):
        q, t, idx, dist = z[case + "_q"], z[case + "_t"], z[case + "_idx"], z[case + "_dist"]
        m, counts = oracle.knn_match(q, t, 2)
        k = min(2, len(t))
        assert (counts == k).all()
        assert np.array_equal(m["trainIdx"][:, :k], idx[:, :k])
        assert np.array_equal(m["distance"][:, :k], dist[:, :k])
        assert (m["queryIdx"][:, 0] == np.arange(len(q))).all() and (m["imgIdx"] == 0).all()


def test_knn_oracle_cfg1_pair(oracle):
    a, b, z = golden("cfg1_480p_a"), golden("cfg1_480p_b"), golden("cfg1_knn")
    m, counts = oracle.knn_match(a["desc"], b["desc"], 2)
    assert np.array_equal(m["trainIdx"], z["idx"]) and np.array_equal(m["distance"], z["dist"])
    good = oracle.ratio_test(m, counts, 0.6)
    ref = [(i, int(z["idx"][i, 0])) for i in range(len(z["idx"])) if z["dist"][i, 0] < np.float32(0.6) * z["dist"][i, 1]]
    assert [(int(g["queryIdx"]), int(g["trainIdx"])) for g in good] == ref
    assert len(ref) > 100  # the synthetic pair must give the ratio test survivors


def test_knn_oracle_empty(oracle):
    q = np.zeros((5, 32), np.uint8)
    m, c = oracle.knn_match(q, np.zeros((0, 32), np.uint8), 2)
    assert (c == 0).all()
    m, c = oracle.knn_match(np.zeros((0, 32), np.uint8), q, 2)
    assert m.shape == (0, 2)
