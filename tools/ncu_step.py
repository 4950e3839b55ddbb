#!/usr/bin/env python3
"""The resident step of bench.py (128 frames of 1280x720: extraction chain + one kNN launch for the 128 pairs) run a few times
and nothing else — the short command the ncu passes wrap (tools/gpu_ncu.sh):
  python tools/ncu_step.py [steps] [batch] [frames_per_call]
frames_per_call < batch runs the chain on that many frames per call (1 = the tracker's one frame per call)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import kfslam_b200 as kfs  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
B = int(sys.argv[2]) if len(sys.argv) > 2 else 128
per_call = int(sys.argv[3]) if len(sys.argv) > 3 else B
W, H, NF = 1280, 720, 2000
dev = torch.device("cuda", 0)
frames = torch.from_numpy(kfs.synth.make_stream(B, H, W, seed=100)).to(dev)
stream = torch.cuda.Stream(device=dev)
orb = kfs.ORB_create(NF, 1.2, 8, 31, 0, 2, kfs.ORB_HARRIS_SCORE, 31, 20, device=0)
bf = kfs.BFMatcher_create(kfs.NORM_HAMMING, device=0)
orb.set_stream(stream.cuda_stream)
bf.set_stream(stream.cuda_stream)
cap = orb.capacity
matches = torch.empty((B, cap, 2, 16), dtype=torch.uint8, device=dev)
orb.extract_device(frames.data_ptr(), W, W * H, W, H, per_call)
_, d_desc, d_cnt, _, _ = orb.device_results()
pairs = []
for i in range(per_call):
    k = (i // 10) * 10
    pairs.append(dict(query=d_desc + k * cap * 32, nq=cap, d_nq=d_cnt + 4 * k, train=d_desc + i * cap * 32, nt=cap, d_nt=d_cnt + 4 * i,
                      out=matches[i].data_ptr(), counts=None))
torch.cuda.synchronize()
t0 = time.perf_counter()
with torch.cuda.stream(stream):
    for s in range(steps):
        for f0 in range(0, B, per_call):
            orb.extract_device(frames[f0].data_ptr(), W, W * H, W, H, per_call)
            bf.knnMatchPairsDevice(pairs, 2)
torch.cuda.synchronize()
kps, _ = orb.fetch_results(per_call)
print("ncu_step: %d steps of %d frames (%d per call) in %.1f ms, %d keypoints in frame 0"
      % (steps, B, per_call, (time.perf_counter() - t0) * 1e3, len(kps[0])))
