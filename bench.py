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


def shared_config(batch):
    """The `config` object BOTH arms print, key for key: the workload BASELINE.json's metric is quoted on (configs[1], batched as
    configs[2]).  Anything an arm measured about it (keypoint counts, its own step size ...) goes to `workload_stats` instead."""
    return {"workload": "cfg2_720p_nf2000_stream_batched", "frame": "%dx%d gray u8" % (W, H), "nfeatures": NF, "nlevels": 8, "scale_factor": 1.2,
            "fast_threshold": 20, "knn": "per frame: keyframe desc (query) x frame desc (train), k=2, ratio 0.6", "keyframe_every": KF_EVERY,
            "batch_per_gpu": batch,
            "l2": "inputs rotate over 3 batches of batch_per_gpu frames (%d MB > 126 MB L2)" % (3 * batch * W * H // 2**20),
            "sharding": "frames sharded across ranks, no collective on the data path"}


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
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,timestamp"

    def __init__(self, gpu_index, period_ms=100):
        self.idx = gpu_index
        self.period_ms = period_ms
        self.proc = None
        self.lines = []
        self.t_begin = self.t_end = None

    def mark_begin(self):
        """Wall-clock start of the timed region: samples are attributed to the region by nvidia-smi's own timestamp."""
        import datetime
        self.t_begin = datetime.datetime.now()

    def mark_end(self):
        import datetime
        self.t_end = datetime.datetime.now()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
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
        import datetime
        sm, mx, reasons, pw, sm_in = [], [], set(), [], []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        slack = datetime.timedelta(milliseconds=self.period_ms)
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 8:
                continue
            try:
                v_sm, v_mx, v_pw = float(p[1]), float(p[2]), float(p[3])
            except ValueError:
                continue
            sm.append(v_sm); mx.append(v_mx); pw.append(v_pw)
            for n, v in zip(names, p[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
            if self.t_begin and self.t_end and len(p) >= 9:
                try:
                    ts = datetime.datetime.strptime(p[8], "%Y/%m/%d %H:%M:%S.%f")
                    if self.t_begin - slack <= ts <= self.t_end + slack:
                        sm_in.append(v_sm)
                except ValueError:
                    pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort(); sm_in.sort()
        out = {"sm_mhz": (sm_in[len(sm_in) // 2] if sm_in else sm[len(sm) // 2]), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
               "samples_in_region": len(sm_in), "period_ms": self.period_ms, "power_w_max": max(pw)}
        if not sm_in:
            out["note"] = "no sample carried a timestamp inside the timed region: median over the whole sampling window (incl. the idle lead-in)"
        return out


# ------------------------------------------------------------------------------------------------------
def _cpu_worker_init(kind, seed, counter):
    global _W
    import importlib
    import numpy as np
    synth = importlib.import_module("keyframe-based-slam_b200.synth")
    with counter.get_lock():          # worker index -> its own stream of frames (seed + index)
        idx = counter.value
        counter.value += 1
    frames = synth.make_stream(KF_EVERY + 1, H, W, seed=seed + idx)
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
            if i == 0:   # the ratio test is a Python loop over 2000 DMatch tuples here (C++ in the reference, where it costs nothing): only
                         # the first frame of a call pays it, as the workload check; the rest of the timed loop is pure cv2
                ng += nframes * sum(1 for r in knn if len(r) == 2 and r[0].distance < 0.6 * r[1].distance)
            nk += len(kp)
        else:
            kp, desc = w["O"].detect_and_compute(f, None, w["P"])
            m, c = w["O"].knn_match(w["kf_desc"], desc, 2)
            ng += len(w["O"].ratio_test(m, c, 0.6))
            nk += len(kp)
    return nframes, nk, ng


class CpuArm:
    """The reference's CPU implementation of the path on all host cores: one worker process per core, one
    thread each (OpenCV's ORB does not scale with its own threads: SURVEY §6); worker i has its own 11-frame stream (seed 1 + i:
    one keyframe + the 10 frames matched against it, VisualOdometry.cpp:105), which it cycles through."""

    def __init__(self):
        import multiprocessing as mp
        try:
            import cv2  # noqa: F401
            self.kind = "reference"
        except Exception:
            self.kind = "port"
        self.cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        ctx = mp.get_context("spawn")
        self.pool = ctx.Pool(self.cores, initializer=_cpu_worker_init, initargs=(self.kind, 1, ctx.Value("i", 0)))

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


def cpu_side_configs():
    """cv2 (the reference's implementation) timed on a FEW frames of the two side workloads, on all host cores it can use:
    config 5 (4K, nf = 10000, 32 x knnMatch 10000 x 10000 + ratio 0.75) and the reference's own setting (nf = 3000, RGB, mask)."""
    out = {}
    try:
        import cv2
        import importlib
        import numpy as np
        synth = importlib.import_module("keyframe-based-slam_b200.synth")
        cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        cv2.setNumThreads(cores)
        f = synth.make_stream(2, 2160, 3840, seed=500)
        orb = cv2.ORB_create(10000, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20)
        bf = cv2.BFMatcher.create(cv2.NORM_HAMMING)
        _, dk = orb.detectAndCompute(f[1], None)
        t0 = time.perf_counter()
        _, dc = orb.detectAndCompute(f[0], None)
        t_ext = time.perf_counter() - t0
        t0 = time.perf_counter()
        npairs = 4
        surv = 0
        for _ in range(npairs):
            rows = bf.knnMatch(dk, dc, 2)
            surv = sum(1 for r in rows if len(r) == 2 and r[0].distance < 0.75 * r[1].distance)
        t_pair = (time.perf_counter() - t0) / npairs
        out["cv2_cfg5"] = {"extract_ms": round(t_ext * 1e3, 1), "knn_pair_ms": round(t_pair * 1e3, 1), "ms_per_frame_32_pairs": round((t_ext + 32 * t_pair) * 1e3, 1),
                           "frames_per_s": round(1.0 / (t_ext + 32 * t_pair), 3), "threads": cores, "survivors_per_pair": surv,
                           "sample": "1 extraction of a 3840x2160 frame (nf=10000) + %d knnMatch(10000 x 10000, k=2) + ratio 0.75, cv2 %s with setNumThreads(%d); "
                                     "32 pairs extrapolated from the per-pair time" % (npairs, cv2.__version__, cores)}
        g = synth.make_stream(7, 720, 1280, seed=9)
        rgb = np.stack([g, np.roll(g, 3, axis=2), 255 - np.roll(g, 5, axis=1)], axis=3)
        mask = np.full((720, 1280), 255, np.uint8)
        mask[:, :40] = 0; mask[:25, :] = 0; mask[300:420, 700:900] = 0
        orb3 = cv2.ORB_create(3000, 1.2, 8, 31, 0, 2, cv2.ORB_HARRIS_SCORE, 31, 20)

        def one(i, kf):
            gray = cv2.cvtColor(cv2.cvtColor(rgb[i], cv2.COLOR_RGB2BGR), cv2.COLOR_BGR2GRAY)   # VisualOdometry.cpp:25-28
            _, d = orb3.detectAndCompute(gray, mask)                                           # :30
            if kf is not None:
                bf.knnMatch(kf, d, 2)                                                          # :39
            return d
        kf = one(0, None)
        t0 = time.perf_counter()
        for i in range(1, 7):
            one(i, kf)
        out["cv2_reference_setting"] = {"ms_per_frame": round((time.perf_counter() - t0) / 6 * 1e3, 2), "threads": cores,
                                        "sample": "6 frames, cvtColor x2 + detectAndCompute(gray, mask) nf=3000 + knnMatch vs keyframe, one process, cv2 threads = %d" % cores}
        cv2.setNumThreads(1)
    except Exception as e:
        out["error"] = str(e)[:200]
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
                      "each worker cycles its own 10 frames (cache-hot, in the CPU's favour); ratio test counted on the first frame of a call only "
                      "(a Python loop here, free in the reference's C++); %s" % (n, arm.cores, dt, "cv2 %s (OpenCV, the code behind VisualOdometry.cpp:30,39)" % __import__("cv2").__version__
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
            "config": shared_config(args.batch),
            "workload_stats": {"frames_per_step": per * arm.cores, "keypoints_per_frame": round(nk, 1), "ratio_survivors_per_frame": round(ng, 1)},
            "cpu_baseline": {"value": round(fps, 2), "unit": "frames/s", "cores": arm.cores, "kind": arm.kind,
                             "sample": "%d frames/step (%d per worker x %d workers x 1 thread), each step a bounded sample of the batched stream; each worker cycles its "
                                       "own 10 cache-hot frames; the Python ratio-test loop runs on the first frame of a step only (extrapolated count); "
                                       "at --gpus N > 1 this arm is still rank 0 alone on ALL %d host cores of the box (the whole host, not 1/N of it)"
                                       % (per * arm.cores, per, arm.cores, arm.cores)},
            "e2e": {"value": round(fps, 2), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit_line(line)
    return 0


# ------------------------------------------------------------------------------------------------------
class ReplayRunner:
    """kfs_replay_* handle + one set of pinned output buffers per arena + the submit / wait loop every end-to-end figure uses
    (pinned host frames in -> keypoints, descriptors, k=2 matches vs the keyframe, ratio-test survivors out)."""

    def __init__(self, N, L, torch, np, device, nfeatures, max_batch, ratio=0.6, arenas=0, chunk=0, slots=2):
        self.N, self.L, self.np = N, L, np
        self.h = C.c_void_p()
        par = N.ReplayParams()
        L.kfs_replay_default_params(C.byref(par))
        par.orb.nfeatures, par.max_batch, par.keyframe_slots, par.ratio, par.chunk_frames, par.arenas = nfeatures, max_batch, slots, ratio, chunk, arenas
        N.check(L.kfs_replay_create(C.byref(par), device, C.byref(self.h)))
        self.cap = L.kfs_replay_capacity(self.h)
        self.arenas = L.kfs_replay_arenas(self.h)
        self.nf, self.B = nfeatures, max_batch
        self.out = []
        for _ in range(self.arenas):
            self.out.append(dict(kp=torch.empty((max_batch, self.cap, 28), dtype=torch.uint8).pin_memory().numpy(),
                                 desc=torch.empty((max_batch, self.cap, 32), dtype=torch.uint8).pin_memory().numpy(),
                                 knn=torch.empty((max_batch, self.cap, 2, 16), dtype=torch.uint8).pin_memory().numpy(),
                                 good=torch.empty((max_batch, self.cap, 16), dtype=torch.uint8).pin_memory().numpy(),
                                 n_kp=np.zeros(max_batch, np.int32), n_query=np.zeros(max_batch, np.int32), n_good=np.zeros(max_batch, np.int32)))
        self.seq = 0
        self.tickets = []

    def submit(self, frames_ptr, w, h, plan, nb, mask_ptr=None, channels=1):
        o = self.out[self.seq % self.arenas]
        self.seq += 1
        io = self.N.ReplayIO(frames_ptr, w * channels, w * h * channels, w, h, o["kp"].ctypes.data, o["desc"].ctypes.data, self.cap, o["n_kp"].ctypes.data,
                             o["knn"].ctypes.data, o["n_query"].ctypes.data, o["good"].ctypes.data, o["n_good"].ctypes.data,
                             mask_ptr, w, w * h, channels)
        t = C.c_int(-1)
        self.N.check(self.L.kfs_replay_submit(self.h, C.byref(io), plan, nb, C.byref(t)))
        self.tickets.append(t.value)
        if len(self.tickets) >= self.arenas:          # every arena busy: take the oldest batch's results
            self.N.check(self.L.kfs_replay_wait(self.h, self.tickets.pop(0)))
        return o

    def drain(self):
        for t in self.tickets:
            self.N.check(self.L.kfs_replay_wait(self.h, t))
        self.tickets = []

    def d2h_bytes(self, nb):
        """Bytes the executor copies device -> host for a batch of nb frames: min(capacity, nfeatures) rows of keypoints (28 B),
        descriptors (32 B), kNN rows (2 x 16 B) and survivors (16 B) per frame + three counts (kfs_replay.cu / kfs_orb.cu)."""
        rows = min(self.cap, self.nf)
        return nb * (rows * (28 + 32 + 32 + 16) + 12)

    def launches(self):
        return self.L.kfs_replay_last_launches(self.h)

    def close(self):
        self.L.kfs_replay_destroy(self.h)


def copy_ceiling(torch, dev, pinned_in, d2h_bytes_per_batch, nbatches, barrier):
    """What the host <-> device link gives THIS rank for the e2e traffic pattern with no kernel at all: per batch one H2D of the
    frames and one D2H of the result bytes, on two streams, everything in flight at once (all ranks run this at the same
    time, so at N > 1 it includes the sharing of the host's PCIe uplinks).  Returns (seconds for nbatches with both directions in
    flight, seconds for the uploads alone).  The first is what an unpaced pipeline gets; the executor can beat it (its downloads are
    shorter than its uploads and staggered against them), so the fraction is quoted against the uploads alone as well."""
    nrot, B = pinned_in.shape[0], pinned_in.shape[1]
    d_in = [torch.empty_like(pinned_in[0], device=dev) for _ in range(2)]
    d_out = torch.empty(d2h_bytes_per_batch, dtype=torch.uint8, device=dev)
    h_out = [torch.empty(d2h_bytes_per_batch, dtype=torch.uint8).pin_memory() for _ in range(2)]
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def run(n, down=True):
        for i in range(n):
            with torch.cuda.stream(s_in):
                d_in[i & 1].copy_(pinned_in[i % nrot], non_blocking=True)
            if down:
                with torch.cuda.stream(s_out):
                    h_out[i & 1].copy_(d_out, non_blocking=True)
        s_in.synchronize(); s_out.synchronize()
    run(3)
    barrier()
    t0 = time.perf_counter()
    run(nbatches)
    both = time.perf_counter() - t0
    barrier()
    t0 = time.perf_counter()
    run(nbatches, down=False)      # the uploads alone: no schedule of the same frames can be faster than this
    up = time.perf_counter() - t0
    return both, up


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
        sampler = ClockSampler(local_rank, period_ms=10) if rank == 0 else None
        if sampler:
            sampler.start()
            time.sleep(0.3)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        if sampler:
            sampler.mark_begin()
        e0.record(stream)
        launches = 0
        for s in range(K):
            step(Wm + s)
            launches += orb.last_launches + bf.last_launches
        e1.record(stream)
        barrier()
        if sampler:
            sampler.mark_end()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None
    t_ms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms_max = float(t_ms.item())
    value = world * B * K / (ms_max * 1e-3)

    # ---- --sustain S: the same resident step loop for S seconds (CUDA events around the whole run, SM clock sampled every 20 ms
    #      INSIDE the region): what a long replay sees once the burst is over
    sustain = None
    if args.sustain > 0:
        n_s = int(max(K, args.sustain / max(ms_max / K * 1e-3, 1e-5)))
        smp = ClockSampler(local_rank, period_ms=20) if rank == 0 else None
        with torch.cuda.stream(stream):
            barrier()
            if smp:
                smp.start()
                time.sleep(0.1)
                smp.mark_begin()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record(stream)
            for s_ in range(n_s):
                step(s_)
                if s_ % 64 == 63:
                    torch.cuda.synchronize(dev)   # keep the launch queue bounded
            s1.record(stream)
            barrier()
            if smp:
                smp.mark_end()
            t_s = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t_s, op=dist.ReduceOp.MAX)
        sustain = {"seconds": round(float(t_s.item()) * 1e-3, 3), "steps": n_s, "frames_per_s": round(world * B * n_s / (float(t_s.item()) * 1e-3), 1),
                   "clocks": smp.stop() if smp else None}

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

    # the matcher is attached to the extractor (kfs_bf_attach_orb): descriptor rows the ORB call just delivered are matched where
    # they still are on the device instead of being uploaded again; a plain (detached) pass is timed beside it
    io = {"kp": kp_h, "desc": desc_h, "match": match_h, "src": None}   # swapped for pageable arrays by the `pageable` pass below

    def e2e_step(s):
        src = io["src"] if io["src"] is not None else pinned[s % nrot].numpy()
        kp_h, desc_h, match_h = io["kp"], io["desc"], io["match"]
        ta = time.perf_counter()
        N.check(L.kfs_orb_detect_and_compute_batch(orb._h, src.ctypes.data, W, W * H, W, H, B, None, 0, 0,
                                                   kp_h.ctypes.data, desc_h.ctypes.data, cap, n_h))
        tb = time.perf_counter()
        e2e_t["orb"] += tb - ta
        for i in range(B):
            k = keyframe_of(i)
            pair_arr[i] = N.BfPair(desc_h[k].ctypes.data, min(n_h[k], cap), None, desc_h[i].ctypes.data, min(n_h[i], cap), None,
                                   match_h[i].ctypes.data, None)
        tc = time.perf_counter()
        N.check(L.kfs_bf_knn_match_pairs(bf._h, pair_arr, B, 2))
        e2e_t["knn"] += time.perf_counter() - tc
        nq = [min(n_h[keyframe_of(i)], cap) for i in range(B)]
        uploaded = 0 if L.kfs_bf_last_aliased(bf._h) == B else sum(min(n_h[i], cap) for i in range(B))  # every frame is train of its own pair
        h2d = B * W * H + 32 * uploaded
        d2h = B * (min(cap, NF) * 60 + 4) + sum(nq) * 32
        return h2d, d2h

    def dropin_pass(attach, seconds):
        N.check(L.kfs_bf_attach_orb(bf._h, orb._h if attach else None))
        for s_ in range(2):
            e2e_step(s_)
        barrier()
        t0 = time.perf_counter()
        e2e_step(0); e2e_step(1)
        n = int(max(2, min(400, seconds / max((time.perf_counter() - t0) / 2, 1e-4))))
        t_n = torch.tensor([n], dtype=torch.int64, device=dev)
        if dist is not None:
            dist.all_reduce(t_n, op=dist.ReduceOp.MAX)
        n = int(t_n.item())
        barrier()
        e2e_t["orb"] = e2e_t["knn"] = 0.0
        t0 = time.perf_counter()
        for s_ in range(n):
            hb, db = e2e_step(s_)
        torch.cuda.synchronize(dev)
        t_e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
        return world * B * n / float(t_e.item()), n, hb, db, e2e_t["orb"] / n * 1e3, e2e_t["knn"] / n * 1e3

    Ke = max(1, min(K, args.e2e_steps))
    plain_value, _, plain_h2d, _, _, plain_knn_ms = dropin_pass(False, 0.15)
    dropin_value, Kd, h2d, d2h, orb_ms, knn_ms_call = dropin_pass(True, max(0.15, args.e2e_seconds / 2))
    # the same two calls on PAGEABLE buffers (plain allocations: what cv::Mat clones and std::vector outputs are in the unmodified tracker,
    # main.cpp:34, VisualOdometry.cpp:21-39): the copies are staged by the driver and no longer overlap the kernels
    pageable = None
    try:
        io.update(kp=np.empty((B, cap, 28), np.uint8), desc=np.empty((B, cap, 32), np.uint8), match=np.empty((B, cap, 2, 16), np.uint8),
                  src=np.array(pinned[0].numpy()))
        pg_value, pg_n, _, _, pg_orb_ms, pg_knn_ms = dropin_pass(True, 0.3)
        pageable = {"value": round(pg_value, 1), "steps": pg_n, "ms_per_step_orb_call": round(pg_orb_ms, 3), "ms_per_step_knn_call": round(pg_knn_ms, 3),
                    "vs_pinned": round(pg_value / dropin_value, 3),
                    "note": "every host buffer pageable (numpy allocations), same frames every step; INTEGRATION.md section 1 says how a tracker pins its cv::Mat buffers"}
    except Exception as ex:
        pageable = {"error": str(ex)[:200]}
    io.update(kp=kp_h, desc=desc_h, match=match_h, src=None)
    N.check(L.kfs_bf_attach_orb(bf._h, None))
    dropin = {"value": round(dropin_value, 1), "pageable": pageable, "unit": "frames/s", "steps": Kd, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
              "api": "kfs_orb_detect_and_compute_batch + kfs_bf_knn_match_pairs (the two drop-in calls, each blocking) on pinned host buffers, the matcher "
                     "attached to the extractor (kfs_bf_attach_orb: rows just extracted are matched on the device, not uploaded again)",
              "ms_per_step_orb_call": round(orb_ms, 3), "ms_per_step_knn_call": round(knn_ms_call, 3),
              "detached": {"value": round(plain_value, 1), "h2d_bytes_per_step": int(plain_h2d), "ms_per_step_knn_call": round(plain_knn_ms, 3)},
              "pipeline_chunk_frames": args.e2e_chunk or "auto (3 x batch/4, then batch/8, batch/16, batch/16)"}

    # ---- e2e through the replay executor (kfs_replay_*): same host buffers in, keypoints + descriptors + kNN matches +
    # ratio-test survivors out, descriptors matched where they were produced; up to `arenas` batches in flight.
    # The timed region is sized to >= args.e2e_seconds of wall clock (a 25 ms burst says nothing about a long replay).
    from kfslam_b200 import replay as RP
    rr = ReplayRunner(N, L, torch, np, local_rank, NF, B, ratio=0.6, arenas=args.arenas, chunk=args.replay_chunk)
    rcap, n_arena = rr.cap, rr.arenas
    plan = (N.ReplayFrame * B)(*[N.ReplayFrame(keyframe_of(i), -1, 0) for i in range(B)])

    def r_run(n, first=0):
        last = None
        for s_ in range(first, first + n):
            last = rr.submit(pinned[s_ % nrot].data_ptr(), W, H, plan, B)
        rr.drain()
        return last

    r_run(2 * n_arena + 1)   # warm-up: every arena has run its chain eagerly and captured its CUDA graph
    barrier()
    t0 = time.perf_counter()
    r_run(12)
    probe = (time.perf_counter() - t0) / 12
    # the probe pays pipeline fill and drain once per 12 batches, the timed run once per Kr: size Kr with 25 % headroom so that the
    # region really lasts --e2e-seconds (round 2's 4-batch probe gave a 0.49 s region for a 0.6 s floor)
    Kr = int(max(Ke, 2, min(4000, 1.25 * args.e2e_seconds / max(probe, 1e-4))))
    t_k = torch.tensor([Kr], dtype=torch.int64, device=dev)
    if dist is not None:
        dist.all_reduce(t_k, op=dist.ReduceOp.MAX)   # every rank times the same number of batches
    Kr = int(t_k.item())
    barrier()
    e2e_sampler = ClockSampler(local_rank, period_ms=20) if rank == 0 else None
    if e2e_sampler:
        e2e_sampler.start()
        time.sleep(0.1)
        e2e_sampler.mark_begin()
    t0 = time.perf_counter()
    last = r_run(Kr)
    rep_dt = time.perf_counter() - t0
    if e2e_sampler:
        e2e_sampler.mark_end()
    e2e_clocks = e2e_sampler.stop() if e2e_sampler else None
    t_r = torch.tensor([rep_dt], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_r, op=dist.ReduceOp.MAX)
    e2e_dt = float(t_r.item())
    e2e_value = world * B * Kr / e2e_dt
    rep_launches = rr.launches()
    rep_h2d = B * W * H
    rep_d2h = rr.d2h_bytes(B)
    # sanity: the executor's last batch equals the drop-in calls' results on the same frames
    src_last = pinned[(Kr - 1) % nrot].numpy()
    N.check(L.kfs_orb_detect_and_compute_batch(orb._h, src_last.ctypes.data, W, W * H, W, H, B, None, 0, 0,
                                               kp_h.ctypes.data, desc_h.ctypes.data, cap, n_h))
    for i in (0, 1, B - 1):
        n_i = int(last["n_kp"][i])
        assert n_i == n_h[i] and np.array_equal(last["desc"][i, :n_i], desc_h[i, :n_i]) and np.array_equal(last["kp"][i, :n_i], kp_h[i, :n_i]), \
            "replay executor and drop-in call disagree on frame %d" % i
    rep_good = float(np.mean(last["n_good"][:B]))
    live_d2h = int(sum(int(last["n_kp"][i]) * 60 + int(last["n_query"][i]) * 32 + int(last["n_good"][i]) * 16 + 12 for i in range(B)))

    # ---- the tracker's outputs only (VisualOdometry.cpp:155-165 keeps keypoints / descriptors of keyframes; every frame keeps its
    #      ratio-test survivors): a SECOND end-to-end figure, never the headline
    plan_t = (N.ReplayFrame * B)(*[N.ReplayFrame(keyframe_of(i), -1, 0 if i % KF_EVERY == 0 else N.REPLAY_SKIP_ROWS) for i in range(B)])
    plan_full, plan = plan, plan_t
    r_run(n_arena)
    barrier()
    Kt = max(2, Kr // 2)
    t0 = time.perf_counter()
    r_run(Kt)
    t_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_t, op=dist.ReduceOp.MAX)
    tracker_value = world * B * Kt / float(t_t.item())
    n_kf_rows = sum(1 for i in range(B) if i % KF_EVERY == 0)
    tracker_d2h = n_kf_rows * min(rcap, NF) * 60 + B * (min(rcap, NF) * 48 + 12)
    plan = plan_full

    # ---- copy ceiling of THIS run: the same bytes per batch over the same pinned frames, no kernels (all ranks at once)
    ceil_dt, up_dt = copy_ceiling(torch, dev, pinned, rep_d2h, Kr, barrier)
    t_c = torch.tensor([ceil_dt, up_dt], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_c, op=dist.ReduceOp.MAX)
    ceiling_value = world * B * Kr / float(t_c[0].item())
    h2d_only_value = world * B * Kr / float(t_c[1].item())

    # ---- cfg3 (BASELINE.json configs[2]): a sequence of args.cfg3_frames 1280x720 frames, keyframe every 10 frames
    #      (VisualOdometry.cpp:105), cut by replay.shard_frames over the ranks (STRONG scaling: the total is fixed); every shard
    #      extracts its halo keyframe locally and carries its last keyframe from submit to submit in the device cache
    #      (replay.plan_shard).  Host frames in, results out, wall clock, max over ranks.
    lo, hi, halo = RP.shard_frames(args.cfg3_frames, world, rank, KF_EVERY)
    subs = RP.plan_shard(lo, hi, halo, B, KF_EVERY)
    c_plans = [((N.ReplayFrame * len(sb["frames"]))(*[N.ReplayFrame(*e) for e in sb["plan"]]), len(sb["frames"])) for sb in subs]

    cfg3_stamps = []

    def cfg3_run(nsub):
        del cfg3_stamps[:]
        for j in range(nsub):
            pl, nb_ = c_plans[j]
            rr.submit(pinned[j % nrot].data_ptr(), W, H, pl, nb_)   # frame i of the sequence = frame of the rotating synthetic pool
            cfg3_stamps.append(time.perf_counter())
        rr.drain()
        cfg3_stamps.append(time.perf_counter())
    cfg3_run(min(len(c_plans), n_arena + 1))
    barrier()
    t0 = time.perf_counter()
    cfg3_run(len(c_plans))
    t_3 = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_3, op=dist.ReduceOp.MAX)
    cfg3_s = float(t_3.item())
    cfg3 = {"workload": "cfg3: %d-frame 1280x720 sequence, nf=%d, keyframe every %d frames, sharded by replay.shard_frames over %d rank(s)"
                        % (args.cfg3_frames, NF, KF_EVERY, world),
            "frames": args.cfg3_frames, "frames_per_s": round(args.cfg3_frames / cfg3_s, 1), "seconds": round(cfg3_s, 4), "scaling": "strong",
            "rank0_shard": [lo, hi, halo], "rank0_submits": len(c_plans), "batch": B,
            "rank0_submit_ms": {"median": round(float(np.median(np.diff(cfg3_stamps))) * 1e3, 3) if len(cfg3_stamps) > 2 else None,
                                "max": round(float(np.max(np.diff(cfg3_stamps))) * 1e3, 3) if len(cfg3_stamps) > 2 else None,
                                "first8": [round(float(x) * 1e3, 2) for x in np.diff(cfg3_stamps)[:8]],
                                "last4": [round(float(x) * 1e3, 2) for x in np.diff(cfg3_stamps)[-4:]]},
            "api": "kfs_replay_submit/wait, plans from replay.plan_shard (halo keyframe extracted locally and unmatched, last keyframe of a submit parked "
                   "in the device keyframe cache for the next); pinned host frames in, keypoints + descriptors + kNN rows + survivors out; wall clock, max over ranks",
            "data": "frame i of the sequence is frame (i mod %d) of this rank's synthetic pool" % (nrot * B)}
    rr.close()

    # ---- cfg5 (BASELINE.json configs[4], the stress case): one 3840x2160 frame, nf = 10000, k = 2 + ratio 0.75 against a window of 32
    #      keyframes (Map::GetNKeyFrames, Map.cpp:72-78).  The 32 (keyframe, frame) pairs are dealt to the ranks by replay.shard_pairs:
    #      keyframe k lives in the ring of rank k mod N; every rank extracts the current frame itself (no collective on the data path).
    #      Per frame through the host API: kfs_frontend_extract_features (pinned 4K frame in, keypoints + descriptors out) +
    #      kfs_frontend_match_window (this rank's share of the window, survivors out).
    cfg5 = None
    try:
        W5, H5, NF5, WIN5, R5 = 3840, 2160, 10000, 32, 0.75
        mine = RP.shard_pairs(WIN5, world, rank)
        f5 = kfs.synth.make_stream(5, H5, W5, seed=500)       # frame 0 = current frame, 1..4 = keyframe sources (cycled)
        p5 = torch.empty((5, H5, W5), dtype=torch.uint8).pin_memory()
        p5.numpy()[...] = f5
        fe5 = C.c_void_p()
        par5 = N.OrbParams()
        L.kfs_orb_default_params(C.byref(par5))
        par5.nfeatures = NF5
        N.check(L.kfs_frontend_create(C.byref(par5), local_rank, C.byref(fe5)))
        N.check(L.kfs_frontend_set_window(fe5, max(1, len(mine))))
        cap5 = L.kfs_frontend_capacity(fe5)
        kp5 = torch.empty((cap5, 28), dtype=torch.uint8).pin_memory().numpy()
        de5 = torch.empty((cap5, 32), dtype=torch.uint8).pin_memory().numpy()
        good5 = torch.empty((max(1, len(mine)), cap5, 16), dtype=torch.uint8).pin_memory().numpy()
        depth5 = torch.full((H5, W5), 1500, dtype=torch.int16).pin_memory().numpy().view(np.uint16)
        n5, nkf5 = C.c_int(), C.c_int()
        ng5 = (C.c_int * N.KFS_MAX_WINDOW)()
        rows5 = (C.c_int * N.KFS_MAX_WINDOW)()
        src5 = p5.numpy()
        for k_ in mine:                                        # this rank's keyframes of the window
            N.check(L.kfs_frontend_extract_features(fe5, src5[1 + k_ % 4].ctypes.data, W5, W5, H5, 1, None, 0, kp5.ctypes.data, de5.ctypes.data, cap5, C.byref(n5)))
            N.check(L.kfs_frontend_promote_keyframe(fe5, depth5.ctypes.data, 2 * W5))

        def frame5():
            N.check(L.kfs_frontend_extract_features(fe5, src5[0].ctypes.data, W5, W5, H5, 1, None, 0, kp5.ctypes.data, de5.ctypes.data, cap5, C.byref(n5)))
            if mine:
                N.check(L.kfs_frontend_match_window(fe5, None, 0, R5, rows5, None, 0, good5.ctypes.data, ng5, cap5, C.byref(nkf5)))
        for _ in range(3):
            frame5()
        barrier()
        n5f = args.cfg5_frames
        t0 = time.perf_counter()
        for _ in range(n5f):
            frame5()
        t_5 = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t_5, op=dist.ReduceOp.MAX)
        sec5 = float(t_5.item()) / n5f
        cmp5 = float(sum(rows5[i] for i in range(nkf5.value)) * n5.value) if mine else 0.0
        t_cmp = torch.tensor([cmp5], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t_cmp, op=dist.ReduceOp.SUM)
        cfg5 = {"workload": "cfg5: %dx%d frame, nf=%d, k=2 + ratio %.2f vs a %d-keyframe window, pairs dealt by replay.shard_pairs over %d rank(s)"
                            % (W5, H5, NF5, R5, WIN5, world),
                "frames_per_s": round(1.0 / sec5, 2), "ms_per_frame": round(sec5 * 1e3, 3), "pairs_per_rank": len(mine), "frames_timed": n5f,
                "keypoints": int(n5.value), "survivors_per_pair_rank0": [int(ng5[i]) for i in range(min(nkf5.value, 4))],
                "gcmp_s_e2e": round(float(t_cmp.item()) / sec5 / 1e9, 1),
                "api": "kfs_frontend_extract_features (pinned host frame in, keypoints + descriptors out) + kfs_frontend_match_window (ONE kNN launch + ONE "
                       "ratio launch for this rank's pairs, survivors out) per frame, blocking; wall clock, max over ranks; every rank extracts the frame itself"}
        L.kfs_frontend_destroy(fe5)
        if rank == 0:
            # device-only figures on rank 0: per-stage extraction times of the 4K frame (roofline vs 133.9 MB/frame algorithmic)
            # and the matcher alone on this rank's share of the 32 pairs
            orb5 = kfs.ORB_create(NF5, 1.2, 8, 31, 0, 2, kfs.ORB_HARRIS_SCORE, 31, 20, device=local_rank)
            orb5.set_stream(stream.cuda_stream)
            d5 = p5[:1].to(dev)
            with torch.cuda.stream(stream):
                for _ in range(3):
                    orb5.extract_device(d5.data_ptr(), W5, W5 * H5, W5, H5, 1)
                orb5.set_profiling(True)
                acc5 = {}
                for _ in range(5):
                    orb5.extract_device(d5.data_ptr(), W5, W5 * H5, W5, H5, 1)
                    orb5.sync()
                    for k_, v in orb5.stage_times().items():
                        acc5[k_] = acc5.get(k_, 0.0) + v / 5
                orb5.set_profiling(False)
                _, dd5, dc5, c5cap, _ = orb5.device_results()
                np5 = max(1, len(mine))
                m5 = torch.empty((np5, c5cap, 2, 16), dtype=torch.uint8, device=dev)
                pr5 = [dict(query=dd5, nq=c5cap, d_nq=dc5, train=dd5, nt=c5cap, d_nt=dc5, out=m5[i].data_ptr(), counts=None) for i in range(np5)]
                bf.knnMatchPairsDevice(pr5, 2)
                torch.cuda.synchronize(dev)
                a5, b5 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a5.record(stream)
                for _ in range(5):
                    bf.knnMatchPairsDevice(pr5, 2)
                b5.record(stream)
                torch.cuda.synchronize(dev)
                knn5_ms = a5.elapsed_time(b5) / 5
            k5, _ = orb5.fetch_results(1)
            nk5 = len(k5[0])
            px5 = [w_ * h_ for w_, h_ in level_sizes(W5, H5)]
            alg5 = sum(px5[:7]) + sum(px5[1:]) + sum(px5) + 2 * sum(px5) + nk5 * (709 + 512 + 32 + 28)   # pyramid + FAST + blur + describe
            ext_ms = sum(acc5.get(k_, 0.0) for k_ in ("pyramid", "fast", "select_harris", "select_sort", "blur", "describe"))
            cfg5["device"] = {"extract_stage_ms": {k_: round(v, 4) for k_, v in acc5.items()}, "extract_ms": round(ext_ms, 4),
                              "extract_alg_mb_per_frame": round(alg5 / 1e6, 1),
                              "extract_gbs": round(alg5 / (ext_ms * 1e-3) / 1e9, 1) if ext_ms > 0 else None,
                              "knn_pairs": np5, "knn_ms": round(knn5_ms, 4), "knn_gcmp_s": round(np5 * float(nk5) ** 2 / (knn5_ms * 1e-3) / 1e9, 1),
                              "note": "one frame per call (latency regime: 8.3 Mpx fill the GPU for ~0.1 ms per stage); stages timed serially by cudaEvent pairs"}
            orb5.close()
            del m5, d5
        del p5
    except Exception as ex:  # informative block: never lose the headline line to it
        cfg5 = {"error": str(ex)[:300]}

    # ---- cfg4 (BASELINE.json configs[3]): the current 1280x720 frame against a sliding window of 10 local-BA keyframes
    #      (Map::GetNKeyFrames, /root/reference/src/lib/Map.cpp:72-78): 2000 x (10 x 2000) Hamming kNN per frame + ratio 0.6, the 10 pairs
    #      dealt over the ranks by replay.shard_pairs.  Same host API as cfg5, at every N (round 2 ran this shape on rank 0 only).
    cfg4 = None
    try:
        WIN4, R4 = 10, 0.6
        mine4 = RP.shard_pairs(WIN4, world, rank)
        f4 = kfs.synth.make_stream(5, H, W, seed=400)
        p4 = torch.empty((5, H, W), dtype=torch.uint8).pin_memory()
        p4.numpy()[...] = f4
        fe4 = C.c_void_p()
        par4 = N.OrbParams()
        L.kfs_orb_default_params(C.byref(par4))
        par4.nfeatures = NF
        N.check(L.kfs_frontend_create(C.byref(par4), local_rank, C.byref(fe4)))
        N.check(L.kfs_frontend_set_window(fe4, max(1, len(mine4))))
        cap4 = L.kfs_frontend_capacity(fe4)
        kp4 = torch.empty((cap4, 28), dtype=torch.uint8).pin_memory().numpy()
        de4 = torch.empty((cap4, 32), dtype=torch.uint8).pin_memory().numpy()
        good4 = torch.empty((max(1, len(mine4)), cap4, 16), dtype=torch.uint8).pin_memory().numpy()
        depth4 = torch.full((H, W), 1500, dtype=torch.int16).pin_memory().numpy().view(np.uint16)
        n4, nkf4 = C.c_int(), C.c_int()
        ng4 = (C.c_int * N.KFS_MAX_WINDOW)()
        rows4 = (C.c_int * N.KFS_MAX_WINDOW)()
        src4 = p4.numpy()
        for k_ in mine4:
            N.check(L.kfs_frontend_extract_features(fe4, src4[1 + k_ % 4].ctypes.data, W, W, H, 1, None, 0, kp4.ctypes.data, de4.ctypes.data, cap4, C.byref(n4)))
            N.check(L.kfs_frontend_promote_keyframe(fe4, depth4.ctypes.data, 2 * W))

        def frame4():
            N.check(L.kfs_frontend_extract_features(fe4, src4[0].ctypes.data, W, W, H, 1, None, 0, kp4.ctypes.data, de4.ctypes.data, cap4, C.byref(n4)))
            if mine4:
                N.check(L.kfs_frontend_match_window(fe4, None, 0, R4, rows4, None, 0, good4.ctypes.data, ng4, cap4, C.byref(nkf4)))
        for _ in range(5):
            frame4()
        barrier()
        n4f = args.cfg4_frames
        t0 = time.perf_counter()
        for _ in range(n4f):
            frame4()
        t_4 = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t_4, op=dist.ReduceOp.MAX)
        sec4 = float(t_4.item()) / n4f
        cmp4 = float(sum(rows4[i] for i in range(nkf4.value)) * n4.value) if mine4 else 0.0
        t_cmp4 = torch.tensor([cmp4], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t_cmp4, op=dist.ReduceOp.SUM)
        cfg4 = {"workload": "cfg4: %dx%d frame, nf=%d, k=2 + ratio %.2f vs a %d-keyframe window (2000 x 20000 per frame), pairs dealt by replay.shard_pairs "
                            "over %d rank(s)" % (W, H, NF, R4, WIN4, world),
                "frames_per_s": round(1.0 / sec4, 1), "us_per_frame": round(sec4 * 1e6, 1), "pairs_per_rank": len(mine4), "frames_timed": n4f,
                "keypoints": int(n4.value), "survivors_per_pair_rank0": [int(ng4[i]) for i in range(min(nkf4.value, 4))],
                "gcmp_s_e2e": round(float(t_cmp4.item()) / sec4 / 1e9, 1), "scaling": "strong",
                "api": "kfs_frontend_extract_features + kfs_frontend_match_window per frame, blocking, pinned host frame in, keypoints + descriptors + "
                       "survivors out; wall clock, max over ranks; every rank extracts the frame itself (the device-only rate of this shape is "
                       "window_matching)"}
        L.kfs_frontend_destroy(fe4)
        del p4
    except Exception as ex:  # informative block: never lose the headline line to it
        cfg4 = {"error": str(ex)[:300]}

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
        # the same two variants on PAGEABLE memory: what the unmodified tracker hands over (cv::Mat clones, main.cpp:34; std::vector /
        # cv::Mat outputs, VisualOdometry.cpp:22-23,38) — plain numpy allocations here, inputs and every output
        pg_src = np.array(src[:KF_EVERY + 1])
        pg_kp, pg_desc, pg_m = np.empty((fcap, 28), np.uint8), np.empty((fcap, 32), np.uint8), np.empty((fcap, 16), np.uint8)
        pg_p3, pg_p2 = np.empty((fcap, 3), np.float64), np.empty((fcap, 2), np.float64)
        pg_depth = np.array(depth_h)

        def track_pg(i):
            N.check(L.kfs_frontend_extract_features(fe, pg_src[i % (KF_EVERY + 1)].ctypes.data, W, W, H, 1, None, 0, pg_kp.ctypes.data, pg_desc.ctypes.data, fcap, C.byref(fn)))
            if i % KF_EVERY == 0:
                N.check(L.kfs_frontend_promote_keyframe(fe, pg_depth.ctypes.data, 2 * W))
            N.check(L.kfs_frontend_match_features(fe, None, None, 0, intr, 0.6, pg_m.ctypes.data, pg_p3.ctypes.data, pg_p2.ctypes.data, fcap, C.byref(fg)))

        def track1_pg(i):
            N.check(L.kfs_frontend_track(fe, pg_src[i % (KF_EVERY + 1)].ctypes.data, W, W, H, 1, None, 0, intr, 0.6, pg_kp.ctypes.data, pg_desc.ctypes.data, fcap,
                                         C.byref(fn), pg_m.ctypes.data, pg_p3.ctypes.data, pg_p2.ctypes.data, fcap, C.byref(fm)))
            if i % KF_EVERY == 0:
                N.check(L.kfs_frontend_promote_keyframe(fe, pg_depth.ctypes.data, 2 * W))
        pg = {}
        for name_, fn_ in (("host_api_pageable_frame_us", track_pg), ("host_api_pageable_track_call_us", track1_pg)):
            for i in range(5):
                fn_(i)
            t0 = time.perf_counter()
            for i in range(nh):
                fn_(i)
            pg[name_] = round((time.perf_counter() - t0) / nh * 1e6, 1)
        if lat is not None and "error" not in lat:
            lat.update(pg)
            lat["host_api_pageable"] = ("the same calls with every host buffer pageable (plain allocations, as the unmodified tracker's cv::Mat clones and "
                                        "std::vector outputs are); ratio to pinned: %.2fx per kfs_frontend_track call" % (pg["host_api_pageable_track_call_us"] / track_us))
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

    # ---- the reference's OWN operating point (not the headline workload): cv::ORB::create(3000, ...) (VisualOdometry.cpp:10-16),
    #      R,G,B frames converted at :25-28, a mask on EVERY call (:30, main.cpp:31-34).  (a) one frame per kfs_frontend_track call as
    #      VisualOdometry::Track does it, (b) batched through the replay executor with colour frames + per-frame masks.
    refset = None
    try:
        NFR, BR = 3000, 32
        g_ = pinned[0].numpy()[:2 * BR]
        rgb_p = torch.empty((2, BR, H, W, 3), dtype=torch.uint8).pin_memory()
        rv = rgb_p.numpy().reshape(2 * BR, H, W, 3)
        rv[..., 0] = g_; rv[..., 1] = np.roll(g_, 3, axis=2); rv[..., 2] = 255 - np.roll(g_, 5, axis=1)
        mask_p = torch.empty((2, BR, H, W), dtype=torch.uint8).pin_memory()
        mv = mask_p.numpy()
        mv[...] = 255
        mv[..., :, :40] = 0; mv[..., :25, :] = 0; mv[..., 300:420, 700:900] = 0
        feR = C.c_void_p()
        parR = N.OrbParams()
        L.kfs_orb_default_params(C.byref(parR))
        parR.nfeatures = NFR
        N.check(L.kfs_frontend_create(C.byref(parR), local_rank, C.byref(feR)))
        capR = L.kfs_frontend_capacity(feR)
        r_kp = torch.empty((capR, 28), dtype=torch.uint8).pin_memory().numpy()
        r_de = torch.empty((capR, 32), dtype=torch.uint8).pin_memory().numpy()
        r_m = torch.empty((capR, 16), dtype=torch.uint8).pin_memory().numpy()
        r_p3 = torch.empty((capR, 3), dtype=torch.float64).pin_memory().numpy()
        r_p2 = torch.empty((capR, 2), dtype=torch.float64).pin_memory().numpy()
        depthR = torch.full((H, W), 1500, dtype=torch.int16).pin_memory().numpy().view(np.uint16)
        intrR = (C.c_double * 4)(910.1, 909.6, 640.4, 360.7)
        rn, rm = C.c_int(), C.c_int()
        mflat = mv.reshape(2 * BR, H, W)

        def trackR(i):
            j = i % (2 * BR)
            N.check(L.kfs_frontend_track(feR, rv[j].ctypes.data, 3 * W, W, H, 3, mflat[j].ctypes.data, W, intrR, 0.6, r_kp.ctypes.data, r_de.ctypes.data, capR,
                                         C.byref(rn), r_m.ctypes.data, r_p3.ctypes.data, r_p2.ctypes.data, capR, C.byref(rm)))
            if i % KF_EVERY == 0:
                N.check(L.kfs_frontend_promote_keyframe(feR, depthR.ctypes.data, 2 * W))
        for i in range(6):
            trackR(i)
        nR = 60
        t0 = time.perf_counter()
        for i in range(nR):
            trackR(i)
        track_us_R = (time.perf_counter() - t0) / nR * 1e6
        launches_R = L.kfs_frontend_last_launches(feR)
        kpR, survR = int(rn.value), int(rm.value)
        L.kfs_frontend_destroy(feR)
        rrR = ReplayRunner(N, L, torch, np, local_rank, NFR, BR, ratio=0.6, arenas=args.arenas)
        planR = (N.ReplayFrame * BR)(*[N.ReplayFrame(keyframe_of(i), -1, 0) for i in range(BR)])

        def runR(n):
            o = None
            for s_ in range(n):
                o = rrR.submit(rgb_p[s_ % 2].data_ptr(), W, H, planR, BR, mask_ptr=mask_p[s_ % 2].data_ptr(), channels=3)
            rrR.drain()
            return o
        runR(2 * rrR.arenas + 1)
        nbR = 24
        t0 = time.perf_counter()
        oR = runR(nbR)
        dtR = time.perf_counter() - t0
        refset = {"workload": "reference setting: 1280x720 R,G,B u8 frames, ORB nfeatures=3000, a mask on every call, k=2 kNN vs keyframe + ratio 0.6",
                  "track_call_us": round(track_us_R, 1), "track_frames_per_s": round(1e6 / track_us_R, 1), "track_launches": int(launches_R),
                  "keypoints_per_frame": kpR, "ratio_survivors_last_frame": survR,
                  "batched_frames_per_s": round(nbR * BR / dtR, 1), "batched_batch": BR, "batched_launches_per_batch": int(rrR.launches()),
                  "batched_h2d_bytes_per_batch": BR * W * H * 4, "batched_d2h_bytes_per_batch": rrR.d2h_bytes(BR),
                  "batched_keypoints_per_frame": round(float(np.mean(oR["n_kp"][:BR])), 1), "batched_survivors_per_frame": round(float(np.mean(oR["n_good"][:BR])), 1),
                  "api": "(a) kfs_frontend_track: one blocking call per frame, pinned colour frame + mask in, keypoints + descriptors + matches + 3-D/2-D points out, "
                         "keyframe promoted every %d frames; (b) kfs_replay_submit/wait with io.channels = 3 and io.mask" % KF_EVERY}
        rrR.close()
        del rgb_p, mask_p
    except Exception as ex:
        refset = {"error": str(ex)[:300]}

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
    dom_kernel = {"pyramid": "k_resize_tma", "fast": "k_fast", "select_harris": "k_select_fast+k_harris", "select_sort": "k_select_sort",
                  "blur": "k_blur7", "describe": "k_describe2"}[dom["stage"]]
    # DRAM bytes per launch of that kernel from the committed `ncu --set full` capture (profiles/ncu_traffic.json: demangled kernel
    # name -> dram__bytes_read.sum + dram__bytes_write.sum of one launch over `frames_per_launch` frames), scaled to this batch
    traffic = traffic_ratio = traffic_src = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        per_launch_frames = float(tj.get("frames_per_launch", 128))
        traffic_src = tj.get("source", "profiles/ncu_traffic.json (committed ncu --set full capture of an earlier run, not this run)")
        for name, v in tj.items():
            base = name.replace("void ", "").split("<")[0].split("(")[0].strip()
            if base == dom_kernel.split("+")[0] and isinstance(v, (int, float)):
                if base.startswith("k_resize") and not tj.get(base + "_launches_summed"):
                    break                                  # one level's launch only: not the pyramid stage's traffic, report none
                traffic = int(v * B / per_launch_frames)   # pyramid: the file holds the sum of the 7 per-level launches
                traffic_ratio = round(traffic / float(dom["alg_bytes_per_frame"] * B), 3)
                break
    except Exception:
        pass
    # ncu (profiles/r02*_kernels.md) shows every extraction kernel issue-bound, not HBM-bound: the binding utilisation of the
    # dominant kernel is printed beside the HBM fraction (north_star's 60 % HBM target is NOT met by any extraction kernel)
    issue = None
    try:
        issue = json.load(open(os.path.join(ROOT, "profiles", "ncu_issue.json"))).get(dom_kernel.split("+")[0])
    except Exception:
        pass
    roofline = {"kernel": dom_kernel, "bound": "hbm", "achieved": dom["achieved_gbs"], "peak": hbm_peak, "unit": "GB/s",
                "frac": dom["frac_hbm"], "traffic": traffic, "traffic_over_algorithmic": traffic_ratio, "traffic_source": traffic_src, "peak_source": peak_src,
                "binding_utilisation_ncu": issue,
                "note": "dominant extraction kernel by device time; per-launch algorithmic bytes = %d B/frame x %d frames; the kernel is integer-issue-bound "
                        "(binding_utilisation_ncu), the 60 %% HBM target of north_star is not met" % (dom["alg_bytes_per_frame"], B)}
    try:
        popc = kfs.pipe_peak(local_rank, "popc")
        lop3 = kfs.pipe_peak(local_rank, "lop3")
    except Exception:
        popc = lop3 = None
    ncmp = float(sum(counts[keyframe_of(i)] * counts[i] for i in range(B)))
    gcmp = ncmp / (knn_ms * 1e-3) / 1e9 if knn_ms > 0 else 0.0
    # the matcher's ceiling: the tighter of the POPC floor (the carry-save compare issues 4 POPC.32, kfs_bf.cu) and the ALU-dispatch
    # floor (16.5 ALU-pipe slots per compare: 14 LOP3 + 2.5 VIMNMX of the two-row top-2 update, at the measured LOP3
    # rate; the weighted sum runs as IMADs on the FMA pipe)
    POPC_PER_CMP, ALU_PER_CMP = 4.0, 16.5
    peak_popc4 = popc / POPC_PER_CMP if popc else None
    peak_alu = lop3 / ALU_PER_CMP if lop3 else None
    peak_cmp = min(peak_popc4, peak_alu) if popc and lop3 else (peak_popc4 or peak_alu)
    roofline_popc = {"kernel": "k_hamming_knn2", "bound": "popc", "achieved": round(gcmp, 1), "unit": "Gcmp/s",
                     "peak": round(peak_cmp, 1) if peak_cmp else None, "frac": round(gcmp / peak_cmp, 4) if peak_cmp else None,
                     "peak_popc_floor_4_per_compare": round(peak_popc4, 1) if peak_popc4 else None,
                     "peak_alu_floor_16.5_slots_per_compare": round(peak_alu, 1) if peak_alu else None,
                     "frac_of_survey_8_popc_definition": round(gcmp / (popc / 8.0), 4) if popc else None,
                     "popc_gops_measured": round(popc, 1) if popc else None, "lop3_gops_measured": round(lop3, 1) if lop3 else None,
                     "ms_per_batch": round(knn_ms, 4), "us_per_frame": round(knn_ms * 1e3 / B, 3),
                     "note": "peak = min(measured POPC.32 rate / 4, measured LOP3 rate / 16.5): what this kernel's instruction mix allows (round 1's compare "
                             "needed 5 POPC + 21.3 ALU slots: its ceiling was 863 Gcmp/s); SURVEY 8d's definition (8 POPC per compare) is kept beside it "
                             "and exceeds 1 because the compare needs only 4"}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            arm = CpuArm()
            cpu = cpu_baseline_block(arm, args.cpu_seconds)
            arm.close()
        except Exception as e:  # never lose the GPU line to a CPU-arm hiccup
            cpu = {"value": None, "unit": "frames/s", "cores": 0, "kind": "port", "sample": "failed: %r" % (e,)}

    if world == 1 and not args.no_cpu_baseline:
        side = cpu_side_configs()
        if isinstance(cfg5, dict) and "cv2_cfg5" in side:
            cfg5["cpu_reference"] = side["cv2_cfg5"]
        if isinstance(refset, dict) and "cv2_reference_setting" in side:
            refset["cpu_reference"] = side["cv2_reference_setting"]

    line = {"metric": METRIC, "value": round(value, 1), "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": round(ms_max / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic",
            "config": shared_config(B),
            "workload_stats": {"keypoints_per_frame": round(nkp, 1), "ratio_survivors_per_frame": round(float(np.mean(good)), 1),
                               "fast_corners_per_level_frame0": fast_counts, "synth_gen_s": round(gen_s, 2), "rank0_cpu_affinity": affinity,
                               "device_arena_mb_per_batch": B * (sum(px) + 13 * 10**6) // 2**20},
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 1), "unit": "frames/s", "h2d_bytes_per_step": int(rep_h2d), "d2h_bytes_per_step": int(rep_d2h),
                    "steps": Kr, "seconds": round(e2e_dt, 3),
                    "api": "kfs_replay_submit/kfs_replay_wait (C ABI replay executor): pinned host frames in; keypoints, descriptors, "
                           "k=2 matches vs keyframe and ratio-test survivors out to pinned host buffers; %d batches in flight (upload | kernels | download)" % n_arena,
                    "d2h_rows": "min(capacity, nfeatures) = %d rows per frame and array (capacity %d incl. tie slack is never shipped; live bytes of the last "
                                "batch: %d)" % (min(rcap, NF), rcap, live_d2h),
                    "copy_ceiling_frames_s": round(ceiling_value, 1), "frac_of_ceiling": round(e2e_value / ceiling_value, 4),
                    "h2d_only_frames_s": round(h2d_only_value, 1), "frac_of_h2d_only": round(e2e_value / h2d_only_value, 4),
                    "copy_ceiling_note": "copy_ceiling = the same pinned frames H2D + the same D2H bytes per batch, two streams, no kernels, all ranks at once, "
                                         "same number of batches (an unpaced pipeline: the executor's staggered, shorter downloads can beat it at N > 1); "
                                         "h2d_only = the uploads alone, a strict upper bound for any schedule of these frames",
                    "clocks_in_region": e2e_clocks,
                    "ratio_survivors_per_frame": round(rep_good, 1), "launches_per_step": int(rep_launches),
                    "replay_chunk_frames": args.replay_chunk or B, "arenas": n_arena,
                    "e2e_tracker_outputs": {"value": round(tracker_value, 1), "unit": "frames/s", "d2h_bytes_per_step": int(tracker_d2h),
                                            "note": "NOT the headline: keypoints / descriptors only of the frames that become keyframes (VisualOdometry.cpp:155-165), "
                                                    "kNN rows and survivors of every frame"},
                    "drop_in_calls": dropin},
            "gpu_launches": int(launches),
            "roofline": roofline, "roofline_popc": roofline_popc, "kernels": kernels, "stream_latency": lat,
            "window_matching": win, "cfg3": cfg3, "cfg4": cfg4, "cfg5": cfg5, "reference_setting": refset, "sustain": sustain,
            "cpu_baseline": cpu}
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
    ap.add_argument("--e2e-seconds", type=float, default=0.6, help="minimum wall-clock length of the end-to-end timed region")
    ap.add_argument("--arenas", type=int, default=0, help="batches in flight in the replay executor (0 = library default)")
    ap.add_argument("--cfg4-frames", type=int, default=300, help="frames of the sharded window-matching leg (BASELINE.json configs[3])")
    ap.add_argument("--cfg3-frames", type=int, default=10000, help="length of the sharded sequence replay (BASELINE.json configs[2])")
    ap.add_argument("--cfg5-frames", type=int, default=12, help="4K frames timed against the 32-keyframe window (configs[4])")
    ap.add_argument("--sustain", type=float, default=1.0, help="also run the resident step loop for this many seconds (clocks sampled every 20 ms)")
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
