"""kfs_bf_knn_match_device_static — the matcher entry point for hosts that are not allowed to allocate.

Who that is: a tracker thread with a frame deadline that shares the GPU with a mapping thread (cudaMalloc / cudaFree stall every
stream of the device), a process whose HBM already belongs to some other allocator (cudaMalloc would simply fail), or code that
only ever receives a stream and device pointers.  The entry point takes no handle; its scratch is a fixed array inside the
library (include/kfslam_frontend.h has the rules).  The reference has nothing like it.

What is checked: through the C ABI, on two calls in a row (different shapes, so that the second one meets whatever the first
left in the shared array), the rows are byte-for-byte those of the existing entry points — kfs_bf_knn_match (host buffers) and
kfs_bf_knn_match_pairs_device (device buffers).  Integer work, so the tolerance is zero, as in every other kNN test here.
The descriptors are tie-rich on purpose: "lower trainIdx wins" is decided in the merge of the partial lists, which is exactly
the part that runs through the scratch.

This file sorts last so that a failure here cannot hide the other GPU tests from `pytest -x`; it must also pass on its own in
a fresh process (the first call then does the per-device set-up).
"""
import ctypes as C

import numpy as np
import pytest

import kfslam_b200 as kfs

N = kfs._native


def _tie_rich(rows, seed):
    a = np.random.default_rng(seed).integers(0, 256, (rows, 32), dtype=np.uint8)
    a[:, 8:] &= 0x11   # few live bits: equal distances are common
    return a


def test_static_entry_rejects_oversize_without_a_gpu():
    """Sizes beyond the compiled-in scratch are refused with a message that names the limit; argument checks come before any
    device work, so this holds on a box without a GPU too."""
    L = kfs.lib()
    assert kfs.STATIC_MAX_QUERIES == 16384
    st = L.kfs_bf_knn_match_device_static(0, None, C.c_void_p(256), kfs.STATIC_MAX_QUERIES + 1, C.c_void_p(256), 100, 2, C.c_void_p(256), None)
    assert st == N.KFS_ERR_INVALID_ARG
    msg = L.kfs_last_error().decode()
    assert "KFS_BF_STATIC_MAX_QUERIES" in msg and "16384" in msg
    assert L.kfs_bf_knn_match_device_static(0, None, C.c_void_p(256), 10, C.c_void_p(256), 1 << 18, 2, C.c_void_p(256), None) == N.KFS_ERR_INVALID_ARG
    assert L.kfs_bf_knn_match_device_static(0, None, C.c_void_p(256), 10, C.c_void_p(256), 10, 3, C.c_void_p(256), None) == N.KFS_ERR_UNSUPPORTED
    assert L.kfs_bf_knn_match_device_static(0, None, C.c_void_p(260), 10, C.c_void_p(256), 10, 2, C.c_void_p(256), None) == N.KFS_ERR_INVALID_ARG


@pytest.mark.gpu
def test_static_entry_matches_existing_entry_points_twice_in_a_row():
    import torch
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    # (nq, nt, k): a full-size tracker pair (16 tiles x 16 splits), then a ragged one with another split count and k, then the
    # largest query count the array allows against a short train set, then one too short to split (scratch untouched)
    shapes = [(2000, 2000, 2), (777, 3001, 2), (kfs.STATIC_MAX_QUERIES, 1300, 2), (1500, 2500, 1), (300, 100, 2)]
    problems = []
    for i, (nq, nt, k) in enumerate(shapes):   # every device buffer exists before the first call: nothing below allocates
        q, t = _tie_rich(nq, 100 + i), _tie_rich(nt, 200 + i)
        problems.append(dict(q=q, t=t, k=k, dq=torch.from_numpy(q).to(dev), dt=torch.from_numpy(t).to(dev),
                             out=torch.zeros(nq * k * 16, dtype=torch.uint8, device=dev), cnt=torch.zeros(nq, dtype=torch.int32, device=dev),
                             out2=torch.zeros(nq * k * 16, dtype=torch.uint8, device=dev), cnt2=torch.zeros(nq, dtype=torch.int32, device=dev)))
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    for rep in range(2):          # the whole sequence twice: every call follows another one's leftovers
        for i, p in enumerate(problems):
            nq, nt, k = len(p["q"]), len(p["t"]), p["k"]
            p["out"].fill_(0xAB); p["cnt"].fill_(-7)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            kfs.knn_match_device_static(p["dq"].data_ptr(), nq, p["dt"].data_ptr(), nt, p["out"].data_ptr(), p["cnt"].data_ptr(), k=k,
                                        device=0, stream=stream)
            torch.cuda.synchronize()
            if rep == 1:   # (the very first call of a process may load the library's kernels; after that the driver's books do not move)
                assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the device's free memory"
            got = p["out"].cpu().numpy().view(N.DM_DTYPE).reshape(nq, k)
            got_cnt = p["cnt"].cpu().numpy()
            ref, ref_cnt = bf.knnMatch(p["q"], p["t"], k=k)                      # existing entry point, host buffers
            assert np.array_equal(got.view(np.uint8), ref.view(np.uint8)), "call %d/%d differs from kfs_bf_knn_match" % (rep, i)
            assert np.array_equal(got_cnt, ref_cnt)
            bf.knnMatchPairsDevice([dict(query=p["dq"].data_ptr(), nq=nq, train=p["dt"].data_ptr(), nt=nt, out=p["out2"].data_ptr(),
                                         counts=p["cnt2"].data_ptr())], k=k)     # existing entry point, device buffers
            bf.sync()
            assert torch.equal(p["out"], p["out2"]) and torch.equal(p["cnt"], p["cnt2"])
    bf.close()


@pytest.mark.gpu
def test_static_entry_back_to_back_on_one_stream_without_sync():
    """Calls on ONE stream are ordered, so they may be queued without waiting in between even though they share the array."""
    import torch
    dev = torch.device("cuda:0")
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING)
    q1, t1, q2, t2 = _tie_rich(2000, 1), _tie_rich(1999, 2), _tie_rich(901, 3), _tie_rich(4000, 4)
    d = [torch.from_numpy(a).to(dev) for a in (q1, t1, q2, t2)]
    o1 = torch.zeros(2000 * 2 * 16, dtype=torch.uint8, device=dev)
    o2 = torch.zeros(901 * 2 * 16, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        kfs.knn_match_device_static(d[0].data_ptr(), 2000, d[1].data_ptr(), 1999, o1.data_ptr(), None, stream=stream)
        kfs.knn_match_device_static(d[2].data_ptr(), 901, d[3].data_ptr(), 4000, o2.data_ptr(), None, stream=stream)
    torch.cuda.synchronize()
    r1, _ = bf.knnMatch(q1, t1)
    r2, _ = bf.knnMatch(q2, t2)
    assert np.array_equal(o1.cpu().numpy(), r1.view(np.uint8).reshape(-1))
    assert np.array_equal(o2.cpu().numpy(), r2.view(np.uint8).reshape(-1))
    bf.close()
