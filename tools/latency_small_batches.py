#!/usr/bin/env python3
"""Device time of one extraction call on nb = 1 .. 16 device-resident 1280x720 frames, for the small-call mechanisms on and off
(KFS_ORB_FUSED_PYRAMID, KFS_ORB_FAST_HALF): where the thresholds PYR_FUSED_MAX_FRAMES / FAST_HALF_MAX_FRAMES should sit.
    python tools/latency_small_batches.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kfslam_b200 as kfs

W, H = 1280, 720
frames = kfs.synth.make_stream(16, H, W, seed=100)
fd = torch.from_numpy(frames).cuda()
settings = [("default", {}), ("pyramid per level", {"KFS_ORB_FUSED_PYRAMID": "0"}), ("pyramid fused always", {"KFS_ORB_FUSED_PYRAMID": "2"}),
            ("fast 62-row tiles", {"KFS_ORB_FAST_HALF": "0"})]
for name, env in settings:
    for k in ("KFS_ORB_FUSED_PYRAMID", "KFS_ORB_FAST_HALF"):
        os.environ.pop(k, None)
    os.environ.update(env)
    row = []
    for nb in (1, 2, 3, 4, 6, 8, 16):
        orb = kfs.ORB_create(2000)
        stream = torch.cuda.Stream()
        orb.set_stream(stream.cuda_stream)
        with torch.cuda.stream(stream):
            for _ in range(4):
                orb.extract_device(fd.data_ptr(), W, W * H, W, H, nb)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            tot = 0.0
            for _ in range(30):
                a.record(stream)
                orb.extract_device(fd.data_ptr(), W, W * H, W, H, nb)
                b.record(stream)
                torch.cuda.synchronize()
                tot += a.elapsed_time(b)
        row.append("%d: %.1f" % (nb, tot / 30 * 1e3))
        orb.close()
    print("%-22s us per call  %s" % (name, "  ".join(row)), flush=True)
