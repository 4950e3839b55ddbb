#!/usr/bin/env python3
"""Host<->device copy ceiling of the box for bench.py's e2e number: pinned 1280x720 gray frames up, result-sized buffers
down, on N GPUs at once (one process per GPU, all copying during the same 2-second window).
    python tools/pcie_ceiling.py [--gpus N]
Prints the aggregate H2D GB/s and the frames/s ceiling it implies (0.92 MB per frame), alone and with D2H in flight."""
import argparse
import multiprocessing as mp
import time


def worker(gpu, both, t_start, seconds, q):
    import torch
    torch.cuda.set_device(gpu)
    frames = 64
    x = torch.empty(frames * 720 * 1280, dtype=torch.uint8).pin_memory()
    d = torch.empty_like(x, device="cuda")
    y = torch.empty(17750784, dtype=torch.uint8).pin_memory()
    dy = torch.empty(17750784, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(3):
        d.copy_(x, non_blocking=True)
    torch.cuda.synchronize()
    while time.time() < t_start:
        time.sleep(0.001)
    n = 0
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        for _ in range(4):
            with torch.cuda.stream(s1):
                d.copy_(x, non_blocking=True)
            if both:
                with torch.cuda.stream(s2):
                    y.copy_(dy, non_blocking=True)
            n += 1
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    q.put((gpu, n * x.numel() / dt / 1e9, n * frames / dt))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--seconds", type=float, default=2.0)
    a = ap.parse_args()
    ctx = mp.get_context("spawn")
    for both in (0, 1):
        q = ctx.Queue()
        t_start = time.time() + 20.0 if a.gpus > 1 else time.time() + 8.0   # every process is up (torch import, pinning) by then
        ps = [ctx.Process(target=worker, args=(g, both, t_start, a.seconds, q)) for g in range(a.gpus)]
        for p in ps:
            p.start()
        res = sorted(q.get() for _ in ps)
        for p in ps:
            p.join()
        gbs, fps = sum(r[1] for r in res), sum(r[2] for r in res)
        print("%d GPU(s), %s: %.1f GB/s H2D aggregate (%s per GPU) -> ceiling %.0f frames/s" %
              (a.gpus, "H2D + D2H in flight" if both else "H2D only", gbs, "/".join("%.0f" % r[1] for r in res), fps), flush=True)


if __name__ == "__main__":
    main()
