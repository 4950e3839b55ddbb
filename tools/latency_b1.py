#!/usr/bin/env python3
"""Stage times of ONE 1280x720 frame (the tracker's streaming call pattern): per-stage CUDA-event times in profiling mode,
and the end-to-end device time of the chain with and without CUDA graphs.  python tools/latency_b1.py"""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kfslam_b200 as kfs

W, H = 1280, 720
frames = kfs.synth.make_stream(4, H, W, seed=100)
dev = torch.device("cuda", 0)
fd = torch.from_numpy(frames).to(dev)
stream = torch.cuda.Stream()
orb = kfs.ORB_create(2000)
orb.set_stream(stream.cuda_stream)
with torch.cuda.stream(stream):
    for i in range(4):
        orb.extract_device(fd[0].data_ptr(), W, W * H, W, H, 1)
    torch.cuda.synchronize()
    orb.set_profiling(True)
    acc = {}
    for i in range(10):
        orb.extract_device(fd[0].data_ptr(), W, W * H, W, H, 1)
        orb.sync()
        for k, v in orb.stage_times().items():
            acc[k] = acc.get(k, 0) + v / 10
    orb.set_profiling(False)
    print("stage us (serialised, profiling):", {k: round(v * 1e3, 1) for k, v in acc.items()}, "sum", round(sum(acc.values()) * 1e3, 1))
    for i in range(4):
        orb.extract_device(fd[0].data_ptr(), W, W * H, W, H, 1)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0
    for i in range(50):
        a.record(stream)
        orb.extract_device(fd[0].data_ptr(), W, W * H, W, H, 1)
        b.record(stream)
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    print("chain per frame: %.1f us (graphs %s, %d launches)" % (tot / 50 * 1e3, os.environ.get("KFS_ORB_GRAPHS", "1"), orb.last_launches))
    dev = 0
    for i in range(50):                          # the host runs ahead behind a 100 us spin kernel: device time alone
        torch.cuda._sleep(200000)
        a.record(stream)
        orb.extract_device(fd[0].data_ptr(), W, W * H, W, H, 1)
        b.record(stream)
        torch.cuda.synchronize()
        dev += a.elapsed_time(b)
    print("chain on the device alone (host enqueue hidden): %.1f us" % (dev / 50 * 1e3))
orb.close()
