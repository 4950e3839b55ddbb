#!/usr/bin/env python3
"""Device time of ONE 2000 x 2000 k=2 match (the tracker's per-frame call) for the train-row split the library picks and for
forced splits (KFS_BF_SPLIT is read when a matcher handle is created).  python tools/latency_knn1.py"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kfslam_b200 as kfs

rng = np.random.default_rng(5)
q = torch.from_numpy(rng.integers(0, 256, (2000, 32), dtype=np.uint8)).cuda()
t = torch.from_numpy(rng.integers(0, 256, (2000, 32), dtype=np.uint8)).cuda()
out = torch.zeros((2000, 2, 4), dtype=torch.int32, device="cuda")  # kfs_dmatch rows (16 bytes)
pair = [dict(query=q.data_ptr(), nq=2000, train=t.data_ptr(), nt=2000, out=out.data_ptr())]
for split in ("auto", "1", "2", "4", "8", "16"):
    if split == "auto":
        os.environ.pop("KFS_BF_SPLIT", None)
    else:
        os.environ["KFS_BF_SPLIT"] = split
    bf = kfs.BFMatcher()
    stream = torch.cuda.Stream()
    bf.set_stream(stream.cuda_stream)
    with torch.cuda.stream(stream):
        for _ in range(5):
            bf.knnMatchPairsDevice(pair)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tot = 0.0
        for _ in range(50):
            a.record(stream)
            bf.knnMatchPairsDevice(pair)
            b.record(stream)
            torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        dev = 0.0
        for _ in range(50):                      # the host runs ahead behind a 100 us spin kernel: device time alone
            torch.cuda._sleep(200000)
            a.record(stream)
            bf.knnMatchPairsDevice(pair)
            b.record(stream)
            torch.cuda.synchronize()
            dev += a.elapsed_time(b)
    print("split %-4s: %.1f us per synchronised call, %.1f us on the device alone" % (split, tot / 50 * 1e3, dev / 50 * 1e3), flush=True)
    bf.close()
