#!/usr/bin/env python3
"""A few hundred random (size, parameters, mask, batch) cases through tests/test_gpu_fuzz.py's sweep, on the GPU box:
  python tools/fuzz_parity.py [cases] [seed] [size scale] > gpurun_out/<tag>_fuzz.txt"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import kfslam_b200 as kfs  # noqa: E402
from oracle import oracle as O  # noqa: E402  (the checker; never on the product path)
import test_gpu_fuzz as F  # noqa: E402

try:
    import cv2
except Exception:
    cv2 = None
n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 7
scale = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
t0 = time.time()
stats = F.run_cases(kfs, O, n, seed=seed, cv2=cv2, log=lambda s: print(s, flush=True), scale=scale)
print("fuzz (seed %d, size scale %.1f): %d cases, %d frames, %d keypoints, %d kNN rows bit-exact against the oracle; %d cases also against cv2 %s directly (%d more rejected by cv2 itself: a level of zero size); %d empty frames; %.0f s"
      % (seed, scale, stats["cases"], stats["frames"], stats["keypoints"], stats["knn_rows"], stats["cv2_checked"], getattr(cv2, "__version__", "n/a"), stats["cv2_rejected"], stats["empty"], time.time() - t0))
if len(sys.argv) > 4:   # fourth argument: number of random tracker sequences (kfs_frontend_track + promotions + window matching)
    t0 = time.time()
    ts = F.run_track_cases(kfs, O, int(sys.argv[4]), seed=seed + 1, log=lambda s: print(s, flush=True))
    print("fuzz track (seed %d): %d sequences, %d frames, %d ratio-test survivors with back-projected points, %d window pairs bit-exact against the oracle; %.0f s"
          % (seed + 1, ts["cases"], ts["frames"], ts["matches"], ts["window_pairs"], time.time() - t0))
if len(sys.argv) > 5:   # fifth argument: number of random replay-executor cases (plans, cache slots, skip flags, arenas)
    t0 = time.time()
    rs = F.run_replay_cases(kfs, O, int(sys.argv[5]), seed=seed + 2, log=lambda s: print(s, flush=True))
    print("fuzz replay (seed %d): %d cases, %d submits, %d matched frames, %d kNN rows, %d ratio-test survivors bit-exact against the oracle; %.0f s"
          % (seed + 2, rs["cases"], rs["submits"], rs["frames"], rs["knn_rows"], rs["survivors"], time.time() - t0))
