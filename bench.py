#!/usr/bin/env python3
"""bench.py — headline benchmark of the ORB + Hamming-kNN front-end (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W             # our arm (sm_100a kernels through the C ABI)
  python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU implementation (OpenCV via cv2,
                                                            # else the oracle port), all host cores
A "step" = one batch of `--batch` synthetic 1280x720 frames per GPU through the whole hot path:
ORB detect+describe (nfeatures=2000, 8 levels, 1.2) and brute-force k=2 Hamming matching of every frame against
its keyframe (keyframe every 10th frame, VisualOdometry.cpp:105), i.e. BASELINE.json configs[1] batched as in
configs[2].  `value` = frames/s with the frames already resident in HBM; `e2e` = the same work through the
host-buffer C-ABI calls (H2D of frames, D2H of keypoints/descriptors/matches inside the timed region).
One process per GPU (torchrun for N>1); frames are sharded, there is no collective on the data path.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "ORB+kNN frames/s @1280x720, 2000 kp; Hamming Gcmp/s vs POPC peak"
W, H, NF = 1280, 720, 2000
KF_EVERY = 10


def level_sizes(w, h, n=8, sf=1.2):
    import numpy as np
    out = []
    for l in range(n):
        s = np.float32(np.float64(np.float32(sf)) ** l)
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float64(np.float32(w) * inv))), int(np.rint(np.float64(np.float32(h) * inv)))))
    return out


def keyframe_of(i):
    """Frame i is matched against the most recent keyframe before it (keyframes: every 10th frame)."""
    k = (i // KF_EVERY) * KF_EVERY
    if k == i:
        k = max(0, i - KF_EVERY)
    return k


# ------------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 8:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2])); pw.append(float(p[3]))
            except ValueError:
                continue
            for n, v in zip(names, p[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": max(pw)}


# ------------------------------------------------------------------------------------------------------
def _cpu_worker_init(kind, seed):
    global _W
    import importlib
    import numpy as np
    synth = importlib.import_module("keyframe-based-slam_b200.synth")
    frames = synth.make_stream(KF_EVERY + 1, H, W, seed=seed)
    _W = {"kind": kind, "frames": frames, "np": np}
    if kind == "reference":
        import cv2
        cv2.setNumThreads(1)
        _W["orb"] = cv2.ORB_create(NF, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20)
        _W["bf"] = cv2.BFMatcher.create(cv2.NORM_HAMMING)
        _, _W["kf_desc"] = _W["orb"].detectAndCompute(frames[0], None)
    else:
        from oracle import oracle as O
        _W["O"] = O
        _W["P"] = O.orb_params(nfeatures=NF)
        _, _W["kf_desc"] = O.detect_and_compute(frames[0], None, _W["P"])


def _cpu_worker_run(nframes):
    """ORB + kNN(k=2, query = keyframe) on `nframes` frames; returns (frames, keypoints, ratio survivors)."""
    w = _W
    nk = ng = 0
    for i in range(nframes):
        f = w["frames"][1 + i % KF_EVERY]
        if w["kind"] == "reference":
            kp, desc = w["orb"].detectAndCompute(f, None)
            knn = w["bf"].knnMatch(w["kf_desc"], desc, 2)
            ng += sum(1 for r in knn if len(r) == 2 and r[0].distance < 0.6 * r[1].distance)
            nk += len(kp)
        else:
            kp, desc = w["O"].detect_and_compute(f, None, w["P"])
            m, c = w["O"].knn_match(w["kf_desc"], desc, 2)
            ng += len(w["O"].ratio_test(m, c, 0.6))
            nk += len(kp)
    return nframes, nk, ng


class CpuArm:
    """The reference's CPU implementation of the path on all host cores: one worker process per core, one
    thread each (OpenCV's ORB does not scale with its own threads: SURVEY §6), distinct frames per worker."""

    def __init__(self):
        import multiprocessing as mp
        try:
            import cv2  # noqa: F401
            self.kind = "reference"
        except Exception:
            self.kind = "port"
        self.cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        ctx = mp.get_context("spawn")
        self.pool = ctx.Pool(self.cores, initializer=_cpu_worker_init, initargs=(self.kind, 1))

    def run(self, frames_per_worker):
        t0 = time.perf_counter()
        res = self.pool.map(_cpu_worker_run, [frames_per_worker] * self.cores)
        dt = time.perf_counter() - t0
        n = sum(r[0] for r in res)
        return n / dt, dt, n, sum(r[1] for r in res) / max(n, 1), sum(r[2] for r in res) / max(n, 1)

    def close(self):
        self.pool.close()
        self.pool.join()


def cpu_latency_lines(cores, nframes=6):
    """SURVEY 8(d) 'CPU baseline timing' items (i), (ii), (iv): per-frame latency of the path (ORB + 2000x2000 kNN vs the keyframe)
    with cv2 on one thread, cv2 with setNumThreads(all cores), and the scalar C oracle port; the CPU model."""
    out = {}
    try:
        import importlib
        synth = importlib.import_module("keyframe-based-slam_b200.synth")
        frames = synth.make_stream(nframes + 1, H, W, seed=7)
        try:
            for line in open("/proc/cpuinfo"):
                if line.startswith("model name"):
                    out["cpu_model"] = line.split(":", 1)[1].strip()
                    break
        except OSError:
            pass

        def timed(fn):
            kf = fn(frames[0], None)
            t0 = time.perf_counter()
            for f in frames[1:]:
                fn(f, kf)
            return (time.perf_counter() - t0) / nframes * 1e3
        try:
            import cv2
            orb = cv2.ORB_create(NF, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20)
            bf = cv2.BFMatcher.create(cv2.NORM_HAMMING)

            def cv_frame(f, kf):
                _, d = orb.detectAndCompute(f, None)
                if kf is not None:
                    bf.knnMatch(kf, d, 2)
                return d
            cv2.setNumThreads(1)
            out["cv2_1_thread_ms_per_frame"] = round(timed(cv_frame), 2)
            cv2.setNumThreads(cores)
            out["cv2_%d_threads_ms_per_frame" % cores] = round(timed(cv_frame), 2)
            cv2.setNumThreads(1)
        except ImportError:
            pass
        from oracle import oracle as O
        P = O.orb_params(nfeatures=NF)

        def port_frame(f, kf):
            _, d = O.detect_and_compute(f, None, P)
            if kf is not None:
                O.knn_match(kf, d, 2)
            return d
        out["c_port_1_thread_ms_per_frame"] = round(timed(port_frame), 2)
    except Exception as e:  # informative only
        out["latency_error"] = str(e)[:160]
    return out


def cpu_baseline_block(arm, seconds_target=12.0):
    arm.run(1)  # warm-up (also builds the keyframe)
    fps, dt, n, _, _ = arm.run(2)
    per = max(2, int(seconds_target / max(dt / 2, 1e-3)))
    per = min(per, 200)
    fps, dt, n, nk, ng = arm.run(per)
    extra = cpu_latency_lines(arm.cores)
    return {**extra, "value": round(fps, 2), "unit": "frames/s", "cores": arm.cores, "kind": arm.kind,
            "sample": "%d frames of 1280x720 (nf=2000 ORB + 2000x2000 kNN k=2 vs keyframe) over %d worker processes x 1 thread, %.1f s wall, "
                      "%s" % (n, arm.cores, dt, "cv2 %s (OpenCV, the code behind VisualOdometry.cpp:30,39)" % __import__("cv2").__version__
                              if arm.kind == "reference" else "oracle/ C port, scalar"),
            "keypoints_per_frame": round(nk, 1), "ratio_survivors_per_frame": round(ng, 1)}


# ------------------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    arm = CpuArm()
    arm.run(1)
    per = max(1, args.ref_frames_per_worker)
    for _ in range(args.warmup):
        arm.run(per)
    t0 = time.perf_counter()
    tot = 0
    for _ in range(args.steps):
        _, _, n, nk, ng = arm.run(per)
        tot += n
    dt = time.perf_counter() - t0
    fps = tot / dt
    arm.close()
    line = {"impl": "reference", "metric": METRIC, "value": round(fps, 2), "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "cfg2_720p_nf2000_stream_batched", "frame": "%dx%d" % (W, H), "nfeatures": NF, "knn": "2000x2000 k=2 vs keyframe",
                       "frames_per_step": per * arm.cores},
            "cpu_baseline": {"value": round(fps, 2), "unit": "frames/s", "cores": arm.cores, "kind": arm.kind,
                             "sample": "%d frames/step (%d per worker x %d workers x 1 thread)" % (per * arm.cores, per, arm.cores)},
            "e2e": {"value": round(fps, 2), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit_line(line)
    return 0


# ------------------------------------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    import kfslam_b200 as kfs

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or kfs.lib().kfs_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device — the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    affinity = bind_to_gpu_numa_node(local_rank, int(os.environ.get("LOCAL_WORLD_SIZE", world))) if world > 1 else None
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    B, K, Wm = args.batch, args.steps, args.warmup
    nrot = 3  # rotating input batches: 3 x B x 0.92 MB (177 MB at B=64) > 126 MB L2

    # ---- synthetic input (seeded; different stream per rank)
    t0 = time.perf_counter()
    frames_h = kfs.synth.make_stream(nrot * B, H, W, seed=100 + rank)
    gen_s = time.perf_counter() - t0
    pinned = torch.empty((nrot, B, H, W), dtype=torch.uint8).pin_memory()
    pinned.numpy()[...] = frames_h.reshape(nrot, B, H, W)
    frames_d = pinned.to(dev, non_blocking=False)

    stream = torch.cuda.Stream(device=dev)
    orb = kfs.ORB_create(NF, 1.2, 8, 31, 0, 2, kfs.ORB_HARRIS_SCORE, 31, 20, device=local_rank)
    bf = kfs.BFMatcher_create(kfs.NORM_HAMMING, device=local_rank)
    orb.set_stream(stream.cuda_stream)
    bf.set_stream(stream.cuda_stream)
    cap = orb.capacity
    matches_d = torch.empty((B, cap, 2, 16), dtype=torch.uint8, device=dev)

    def extract(r):
        orb.extract_device(frames_d[r].data_ptr(), W, W * H, W, H, B)

    extract(0)
    d_kp, d_desc, d_cnt, cap2, _ = orb.device_results()
    assert cap2 == cap
    pairs = []
    for i in range(B):
        k = keyframe_of(i)
        pairs.append(dict(query=d_desc + k * cap * 32, nq=cap, d_nq=d_cnt + 4 * k,
                          train=d_desc + i * cap * 32, nt=cap, d_nt=d_cnt + 4 * i,
                          out=matches_d[i].data_ptr(), counts=None))

    def step(s):
        extract(s % nrot)
        bf.knnMatchPairsDevice(pairs, 2)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- value: device-resident inputs, CUDA events on the launching stream
    with torch.cuda.stream(stream):
        for s in range(Wm):
            step(s)
        barrier()
        sampler = ClockSampler(local_rank) if rank == 0 else None
        if sampler:
            sampler.start()
            time.sleep(0.3)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        launches = 0
        for s in range(K):
            step(Wm + s)
            launches += orb.last_launches + bf.last_launches
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None
    t_ms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms_max = float(t_ms.item())
    value = world * B * K / (ms_max * 1e-3)

    # ---- result sanity on the last batch (not timed)
    kps, descs = orb.fetch_results(B)
    counts = np.array([len(k) for k in kps])
    mh = matches_d.cpu().numpy().view(kfs.DM_DTYPE).reshape(B, cap, 2)
    good = []
    for i in range(B):
        nq = counts[keyframe_of(i)]
        m = mh[i, :nq]
        good.append(int(((m["trainIdx"][:, 1] >= 0) & (m["distance"][:, 0] < np.float32(0.6) * m["distance"][:, 1])).sum()))
    fast_counts = [int(len(orb.download_fast_corners(0, l))) for l in range(8)]

    # ---- e2e: host buffers through the public C-ABI calls (H2D + D2H inside the timed region)
    kp_h = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory().numpy()
    desc_h = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory().numpy()
    match_h = torch.empty((B, cap, 2, 16), dtype=torch.uint8).pin_memory().numpy()
    n_h = (C.c_int * B)()
    L = kfs.lib()
    from importlib import import_module
    N = import_module("keyframe-based-slam_b200._native")
    pair_arr = (N.BfPair * B)()

    e2e_t = {"orb": 0.0, "knn": 0.0}
    if args.e2e_chunk:
        orb.set_pipeline_chunk(args.e2e_chunk)

    def e2e_step(s):
        src = pinned[s % nrot].numpy()
        ta = time.perf_counter()
        N.check(L.kfs_orb_detect_and_compute_batch(orb._h, src.ctypes.data, W, W * H, W, H, B, None, 0, 0,
                                                   kp_h.ctypes.data, desc_h.ctypes.data, cap, n_h))
        tb = time.perf_counter()
        e2e_t["orb"] += tb - ta
        up = 0
        for i in range(B):
            k = keyframe_of(i)
            pair_arr[i] = N.BfPair(desc_h[k].ctypes.data, min(n_h[k], cap), None, desc_h[i].ctypes.data, min(n_h[i], cap), None,
                                   match_h[i].ctypes.data, None)
        tc = time.perf_counter()
        N.check(L.kfs_bf_knn_match_pairs(bf._h, pair_arr, B, 2))
        e2e_t["knn"] += time.perf_counter() - tc
        nq = [min(n_h[keyframe_of(i)], cap) for i in range(B)]
        h2d = B * W * H + 32 * (sum(min(n_h[i], cap) for i in range(B)) + sum(min(n_h[k], cap) for k in set(keyframe_of(i) for i in range(B))))
        d2h = B * cap * 60 + 4 * B + sum(nq) * 32
        return h2d, d2h

    Ke = max(1, min(K, args.e2e_steps))
    for s in range(min(Wm, 2)):
        e2e_step(s)
    barrier()
    e2e_t["orb"] = e2e_t["knn"] = 0.0
    t0 = time.perf_counter()
    for s in range(Ke):
        h2d, d2h = e2e_step(s)
    torch.cuda.synchronize(dev)
    e2e_dt = time.perf_counter() - t0
    t_e = torch.tensor([e2e_dt], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
    dropin_value = world * B * Ke / float(t_e.item())
    dropin = {"value": round(dropin_value, 1), "unit": "frames/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
              "api": "kfs_orb_detect_and_compute_batch + kfs_bf_knn_match_pairs (the two drop-in calls, each blocking) on pinned host buffers",
              "ms_per_step_orb_call": round(e2e_t["orb"] / Ke * 1e3, 3), "ms_per_step_knn_call": round(e2e_t["knn"] / Ke * 1e3, 3),
              "pipeline_chunk_frames": args.e2e_chunk or "auto (batch/4)"}

    # ---- e2e through the replay executor (kfs_replay_*): same host buffers in, keypoints + descriptors + kNN matches +
    # ratio-test survivors out, descriptors matched where they were produced; batch s+1 is submitted before batch s is awaited
    rp = C.c_void_p()
    rpar = N.ReplayParams()
    L.kfs_replay_default_params(C.byref(rpar))
    rpar.orb.nfeatures, rpar.max_batch, rpar.keyframe_slots, rpar.ratio, rpar.chunk_frames = NF, B, 2, 0.6, args.replay_chunk
    N.check(L.kfs_replay_create(C.byref(rpar), local_rank, C.byref(rp)))
    rcap = L.kfs_replay_capacity(rp)
    plan = (N.ReplayFrame * B)(*[N.ReplayFrame(keyframe_of(i), -1) for i in range(B)])
    r_out, r_io = [], []
    n_arena = L.kfs_replay_arenas(rp)
    for b in range(n_arena):
        o = dict(kp=torch.empty((B, rcap, 28), dtype=torch.uint8).pin_memory().numpy(),
                 desc=torch.empty((B, rcap, 32), dtype=torch.uint8).pin_memory().numpy(),
                 knn=torch.empty((B, rcap, 2, 16), dtype=torch.uint8).pin_memory().numpy(),
                 good=torch.empty((B, rcap, 16), dtype=torch.uint8).pin_memory().numpy(),
                 n_kp=np.zeros(B, np.int32), n_query=np.zeros(B, np.int32), n_good=np.zeros(B, np.int32))
        r_out.append(o)
    def r_submit(s):
        o = r_out[s % n_arena]
        io = N.ReplayIO(pinned[s % nrot].data_ptr(), W, W * H, W, H, o["kp"].ctypes.data, o["desc"].ctypes.data, rcap, o["n_kp"].ctypes.data,
                        o["knn"].ctypes.data, o["n_query"].ctypes.data, o["good"].ctypes.data, o["n_good"].ctypes.data)
        t = C.c_int(-1)
        N.check(L.kfs_replay_submit(rp, C.byref(io), plan, B, C.byref(t)))
        return t.value

    def r_run(n):
        tickets = []
        for s in range(n):
            tickets.append(r_submit(s))
            if len(tickets) >= n_arena:              # every arena busy: take the oldest batch's results
                N.check(L.kfs_replay_wait(rp, tickets.pop(0)))
        for t in tickets:
            N.check(L.kfs_replay_wait(rp, t))

    Kr = max(2, Ke)
    r_run(2 * n_arena + 1)   # warm-up: every arena has run its chain eagerly and captured its CUDA graph
    barrier()
    t0 = time.perf_counter()
    r_run(Kr)
    rep_dt = time.perf_counter() - t0
    t_r = torch.tensor([rep_dt], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_r, op=dist.ReduceOp.MAX)
    e2e_value = world * B * Kr / float(t_r.item())
    rep_launches = L.kfs_replay_last_launches(rp)
    rep_h2d = B * W * H
    rep_d2h = B * rcap * (28 + 32 + 32 + 16) + 3 * 4 * B
    # sanity: the executor's last batch equals the drop-in calls' results on the same frames
    last = r_out[(Kr - 1) % n_arena]
    src_last = pinned[(Kr - 1) % nrot].numpy()
    N.check(L.kfs_orb_detect_and_compute_batch(orb._h, src_last.ctypes.data, W, W * H, W, H, B, None, 0, 0,
                                               kp_h.ctypes.data, desc_h.ctypes.data, cap, n_h))
    for i in (0, 1, B - 1):
        n_i = int(last["n_kp"][i])
        assert n_i == n_h[i] and np.array_equal(last["desc"][i, :n_i], desc_h[i, :n_i]) and np.array_equal(last["kp"][i, :n_i], kp_h[i, :n_i]), \
            "replay executor and drop-in call disagree on frame %d" % i
    rep_good = float(np.mean(last["n_good"]))
    L.kfs_replay_destroy(rp)

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- per-kernel times (rank 0, outside the timed region): cudaEvent pairs between stages + kNN alone
    orb.set_profiling(True)
    acc = {}
    reps = 5
    knn_ms = 0.0
    with torch.cuda.stream(stream):
        for r in range(reps):
            extract(r % nrot)
            orb.sync()
            for k_, v in orb.stage_times().items():
                acc[k_] = acc.get(k_, 0.0) + v / reps
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            bf.knnMatchPairsDevice(pairs, 2)   # keeps the GPU busy while the host prepares the timed launches
            a.record(stream)
            for _ in range(3):
                bf.knnMatchPairsDevice(pairs, 2)
            b_.record(stream)
            torch.cuda.synchronize(dev)
            knn_ms += a.elapsed_time(b_) / 3 / reps
    orb.set_profiling(False)

    # ---- streaming latency (configs[1] as the tracker runs it: ONE frame per call): extract + 2000x2000 kNN vs the
    #      keyframe, device-resident, CUDA events around 20 back-to-back frames
    lat = None
    try:
        orb1 = kfs.ORB_create(NF, 1.2, 8, 31, 0, 2, kfs.ORB_HARRIS_SCORE, 31, 20, device=local_rank)
        orb1.set_stream(stream.cuda_stream)
        one = frames_d[0]
        orb1.extract_device(one[0].data_ptr(), W, W * H, W, H, 1)
        _, d1, c1, cap1, _ = orb1.device_results()
        kf_desc = d_desc  # keyframe descriptors of the last batch (frame 0), resident
        m1 = torch.empty((cap1, 2, 16), dtype=torch.uint8, device=dev)
        pair1 = [dict(query=kf_desc, nq=cap, d_nq=d_cnt, train=d1, nt=cap1, d_nt=c1, out=m1.data_ptr(), counts=None)]
        with torch.cuda.stream(stream):
            RING = 4  # the camera side of a tracker hands frames over in a small ring of device buffers
            for i in range(3 * RING):
                orb1.extract_device(one[i % RING].data_ptr(), W, W * H, W, H, 1)
                bf.knnMatchPairsDevice(pair1, 2)
            torch.cuda.synchronize(dev)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            nlat = 20
            ext_ms = knn1_ms = 0.0
            for i in range(nlat):
                ev[0].record(stream)
                orb1.extract_device(one[i % RING].data_ptr(), W, W * H, W, H, 1)
                ev[1].record(stream)
                bf.knnMatchPairsDevice(pair1, 2)
                ev[2].record(stream)
                torch.cuda.synchronize(dev)
                ext_ms += ev[0].elapsed_time(ev[1]) / nlat
                knn1_ms += ev[1].elapsed_time(ev[2]) / nlat
        lat = {"frames_per_call": 1, "extract_us": round(ext_ms * 1e3, 1), "knn_us": round(knn1_ms * 1e3, 1),
               "frame_us": round((ext_ms + knn1_ms) * 1e3, 1), "note": "device-resident single frame out of a ring of 4 input buffers, each frame synchronised (latency, not throughput); "
                       "the kernel chain replays as a CUDA graph"}
        orb1.close()
    except Exception as ex:  # latency is informative only
        lat = {"error": str(ex)[:200]}

    # ---- the tracker's own call pattern through the C ABI with HOST buffers (VisualOdometry::Track, :129-172): per frame
    #      kfs_frontend_extract_features (pinned gray frame in, keypoints + descriptors out) + kfs_frontend_match_features
    #      against the device-resident keyframe (ratio test + back-projection, 3-D / 2-D lists out); blocking calls, one frame
    try:
        fe = C.c_void_p()
        fpar = N.OrbParams()
        L.kfs_orb_default_params(C.byref(fpar))
        fpar.nfeatures = NF
        N.check(L.kfs_frontend_create(C.byref(fpar), local_rank, C.byref(fe)))
        fcap = L.kfs_frontend_capacity(fe)
        f_kp = torch.empty((fcap, 28), dtype=torch.uint8).pin_memory().numpy()
        f_desc = torch.empty((fcap, 32), dtype=torch.uint8).pin_memory().numpy()
        f_m = torch.empty((fcap, 16), dtype=torch.uint8).pin_memory().numpy()
        f_p3 = torch.empty((fcap, 3), dtype=torch.float64).pin_memory().numpy()
        f_p2 = torch.empty((fcap, 2), dtype=torch.float64).pin_memory().numpy()
        depth_h = torch.full((H, W), 1500, dtype=torch.int16).pin_memory().numpy().view(np.uint16)
        intr = (C.c_double * 4)(910.1, 909.6, 640.4, 360.7)
        fn, fg = C.c_int(), C.c_int()
        src = pinned[0].numpy()

        def track(i):
            N.check(L.kfs_frontend_extract_features(fe, src[i % B].ctypes.data, W, W, H, 1, None, 0, f_kp.ctypes.data, f_desc.ctypes.data, fcap, C.byref(fn)))
            if i % KF_EVERY == 0:
                N.check(L.kfs_frontend_promote_keyframe(fe, depth_h.ctypes.data, 2 * W))
            N.check(L.kfs_frontend_match_features(fe, None, None, 0, intr, 0.6, f_m.ctypes.data, f_p3.ctypes.data, f_p2.ctypes.data, fcap, C.byref(fg)))
        for i in range(5):
            track(i)
        nh = 40
        t0 = time.perf_counter()
        for i in range(nh):
            track(i)
        host_us = (time.perf_counter() - t0) / nh * 1e6
        # the same in ONE call / one synchronisation per frame (kfs_frontend_track)
        fm = C.c_int()

        def track1(i):
            N.check(L.kfs_frontend_track(fe, src[i % B].ctypes.data, W, W, H, 1, None, 0, intr, 0.6, f_kp.ctypes.data, f_desc.ctypes.data, fcap, C.byref(fn),
                                         f_m.ctypes.data, f_p3.ctypes.data, f_p2.ctypes.data, fcap, C.byref(fm)))
            if i % KF_EVERY == 0:
                N.check(L.kfs_frontend_promote_keyframe(fe, depth_h.ctypes.data, 2 * W))
        for i in range(5):
            track1(i)
        t0 = time.perf_counter()
        for i in range(nh):
            track1(i)
        track_us = (time.perf_counter() - t0) / nh * 1e6
        if lat is not None and "error" not in lat:
            lat["host_api_track_call_us"] = round(track_us, 1)
            lat["host_api_track_frames_per_s"] = round(1e6 / track_us, 1)
            lat["host_api_frame_us"] = round(host_us, 1)
            lat["host_api_frames_per_s"] = round(1e6 / host_us, 1)
            lat["host_api"] = ("kfs_frontend_extract_features + kfs_frontend_match_features per frame, pinned host frame in (%d B), keypoints + "
                               "descriptors + matches + 3-D/2-D points out, keyframe promoted every %d frames, blocking" % (W * H, KF_EVERY))
        L.kfs_frontend_destroy(fe)
    except Exception as ex:
        if isinstance(lat, dict):
            lat["host_api_error"] = str(ex)[:200]

    # ---- window matching (configs[3]: current frame vs the 10 keyframes of a local-BA window = 10 pairs of 2000 x 2000 per
    #      frame; kfs_frontend_match_window issues exactly this launch): (a) one frame per call, (b) every frame of the batch
    win = None
    try:
        WIN = 10
        wout = torch.empty((B * WIN, cap, 2, 16), dtype=torch.uint8, device=dev)

        def wpairs(frames):
            pp = []
            for i in frames:
                for j in range(WIN):
                    k = (i + 1 + j) % B     # ten other frames of the batch stand in for the window's keyframes
                    pp.append(dict(query=d_desc + k * cap * 32, nq=cap, d_nq=d_cnt + 4 * k, train=d_desc + i * cap * 32, nt=cap,
                                   d_nt=d_cnt + 4 * i, out=wout[len(pp)].data_ptr(), counts=None))
            return pp
        one_f, all_f = wpairs([0]), wpairs(range(B))
        res = {}
        with torch.cuda.stream(stream):
            for name, pp, reps_w in (("one_frame", one_f, 20), ("batch", all_f, 3)):
                bf.knnMatchPairsDevice(pp, 2)
                torch.cuda.synchronize(dev)
                a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                for _ in range(reps_w):
                    bf.knnMatchPairsDevice(pp, 2)
                b_.record(stream)
                torch.cuda.synchronize(dev)
                nfr = len(pp) // WIN
                cmp_ = float(sum(counts[(i + 1 + j) % B] * counts[i] for i in range(nfr) for j in range(WIN)))
                ms_w = a.elapsed_time(b_) / reps_w
                res[name] = {"frames": nfr, "pairs": len(pp), "us_per_frame": round(ms_w * 1e3 / nfr, 2), "gcmp_s": round(cmp_ / (ms_w * 1e-3) / 1e9, 1)}
        win = {"workload": "cfg4: 2000 x (10 x 2000) per frame, k=2", "window": WIN, **res}
        del wout
    except Exception as ex:  # informative only
        win = {"error": str(ex)[:200]}

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    sizes = level_sizes(W, H)
    px = [w_ * h_ for w_, h_ in sizes]
    nkp = float(counts.mean())
    quotas = [434, 362, 302, 251, 209, 175, 145, 122]
    # algorithmic bytes per frame
    alg = {
        "pyramid": sum(px[:7]) + sum(px[1:]),
        "fast": sum(px),
        "select_harris": 2 * sum(quotas) * 81 + 4 * sum(fast_counts),
        "select_sort": 2 * sum(quotas) * 16,
        "blur": 2 * sum(px),
        "describe": nkp * (709 + 512 + 32 + 28),
    }
    kernels = []
    for k_ in ("pyramid", "fast", "select_harris", "select_sort", "blur", "describe"):
        t = acc.get(k_, 0.0)
        gbs = alg[k_] * B / (t * 1e-3) / 1e9 if t > 0 else 0.0
        kernels.append({"stage": k_, "ms_per_batch": round(t, 4), "us_per_frame": round(t * 1e3 / B, 3), "alg_bytes_per_frame": int(alg[k_]),
                        "achieved_gbs": round(gbs, 1), "frac_hbm": round(gbs / hbm_peak, 4)})
    dom = max(kernels, key=lambda d: d["ms_per_batch"])
    dom_kernel = {"pyramid": "k_resize", "fast": "k_fast", "select_harris": "k_select_fast+k_harris", "select_sort": "k_select_sort",
                  "blur": "k_blur7", "describe": "k_describe"}[dom["stage"]]
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(dom_kernel)
    except Exception:
        pass
    roofline = {"kernel": dom_kernel, "bound": "hbm", "achieved": dom["achieved_gbs"], "peak": hbm_peak, "unit": "GB/s",
                "frac": dom["frac_hbm"], "traffic": traffic, "peak_source": peak_src,
                "note": "dominant extraction kernel by device time; per-launch algorithmic bytes = %d B/frame x %d frames" % (dom["alg_bytes_per_frame"], B)}
    try:
        popc = kfs.pipe_peak(local_rank, "popc")
        lop3 = kfs.pipe_peak(local_rank, "lop3")
    except Exception:
        popc = lop3 = None
    ncmp = float(sum(counts[keyframe_of(i)] * counts[i] for i in range(B)))
    gcmp = ncmp / (knn_ms * 1e-3) / 1e9 if knn_ms > 0 else 0.0
    roofline_popc = {"kernel": "k_hamming_knn2", "bound": "popc", "achieved": round(gcmp, 1), "unit": "Gcmp/s",
                     "peak": round(popc / 8.0, 1) if popc else None, "frac": round(gcmp / (popc / 8.0), 4) if popc else None,
                     "popc_gops_measured": round(popc, 1) if popc else None, "lop3_gops_measured": round(lop3, 1) if lop3 else None,
                     "ms_per_batch": round(knn_ms, 4), "us_per_frame": round(knn_ms * 1e3 / B, 3),
                     "note": "peak = measured POPC.32 issue rate / 8 (8 POPC per 256-bit compare, SURVEY §8d)"}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            arm = CpuArm()
            cpu = cpu_baseline_block(arm, args.cpu_seconds)
            arm.close()
        except Exception as e:  # never lose the GPU line to a CPU-arm hiccup
            cpu = {"value": None, "unit": "frames/s", "cores": 0, "kind": "port", "sample": "failed: %r" % (e,)}

    line = {"metric": METRIC, "value": round(value, 1), "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": round(ms_max / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic",
            "config": {"workload": "cfg2_720p_nf2000_stream_batched", "frame": "%dx%d gray u8" % (W, H), "nfeatures": NF, "nlevels": 8,
                       "batch_per_gpu": B, "knn": "per frame: keyframe desc (query) x frame desc (train), k=2, keyframe every %d frames" % KF_EVERY,
                       "l2": "inputs rotate over %d batches (%d MB > 126 MB L2); pyramid+lists per batch %d MB" % (nrot, nrot * B * W * H // 2**20, B * (sum(px) + 13 * 10**6) // 2**20),
                       "sharding": "frames sharded across ranks, no collective on the data path",
                       "keypoints_per_frame": round(nkp, 1), "ratio_survivors_per_frame": round(float(np.mean(good)), 1),
                       "fast_corners_per_level_frame0": fast_counts, "synth_gen_s": round(gen_s, 2), "rank0_cpu_affinity": affinity},
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 1), "unit": "frames/s", "h2d_bytes_per_step": int(rep_h2d), "d2h_bytes_per_step": int(rep_d2h),
                    "steps": Kr, "api": "kfs_replay_submit/kfs_replay_wait (C ABI replay executor): pinned host frames in; keypoints, descriptors, "
                                        "k=2 matches vs keyframe and ratio-test survivors out to pinned host buffers; three batches in flight (upload | kernels | download)",
                    "ratio_survivors_per_frame": round(rep_good, 1), "launches_per_step": int(rep_launches),
                    "replay_chunk_frames": args.replay_chunk or B,
                    "drop_in_calls": dropin},
            "gpu_launches": int(launches),
            "roofline": roofline, "roofline_popc": roofline_popc, "kernels": kernels, "stream_latency": lat,
            "window_matching": win, "cpu_baseline": cpu}
    emit_line(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


_REAL_STDOUT = None


def guard_stdout():
    """Only the JSON line may reach stdout: libraries that print there (NCCL's version banner, torchrun children) are sent to
    stderr by pointing fd 1 at fd 2; emit_line() writes to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def bind_to_gpu_numa_node(gpu_index, local_world):
    """Pin this rank (and the pinned host buffers it is about to allocate: first touch) to the CPUs next to its GPU, so that
    8 ranks do not pull their frames across the socket interconnect.  Falls back to an even split of the visible cores."""
    info = {}
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        near = [c for c in range(ncpu) if (int(words[c // 64]) >> (c % 64)) & 1]
        allowed = sorted(os.sched_getaffinity(0))
        near = [c for c in near if c in allowed] or allowed
        # ranks that share a NUMA node split its cores
        sharers = [g for g in range(local_world) if list(pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(g), (ncpu + 63) // 64)) == list(words)]
        k, n = sharers.index(gpu_index) if gpu_index in sharers else 0, max(1, len(sharers))
        per = max(1, len(near) // n)
        mine = near[k * per:(k + 1) * per] or near
        os.sched_setaffinity(0, mine)
        info = {"cpus": "%d-%d" % (mine[0], mine[-1]), "numa_cpus": len(near), "ranks_on_node": n}
    except Exception as e:  # affinity is an optimisation only
        info = {"error": str(e)[:120]}
    return info


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--batch", type=int, default=128, help="frames per GPU per step")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--replay-chunk", type=int, default=0, help="frames per chunk inside one replay batch (0 = whole batch)")
    ap.add_argument("--e2e-chunk", type=int, default=0, help="frames per chunk of the host-buffer pipeline (0 = library default)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--ref-frames-per-worker", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3  # timing rule: >= 3 warm-up steps
    guard_stdout()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
