#!/usr/bin/env python3
"""bench.py — stereo frames/s of ORB extraction + stereo matching (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A step is one pass of the hot path over one batch of B synthetic KITTI-shaped stereo pairs
(1241x376, nFeatures 2000, 8 levels, 1.2; BASELINE.json configs[1]) on every rank:
pyramid -> FAST cells -> quad-tree -> orientation+blur+rBRIEF (both eyes) -> ComputeStereoMatches.
Independent frames are sharded over ranks (weak scaling, no data-path collective); NCCL only
gathers counts / checksums / timings after the timed regions.

  value : whole-job stereo frames/s, inputs already resident in HBM
  e2e   : the same through the C ABI with pinned HOST buffers: H2D of every image and D2H of all
          results (keypoints, descriptors, uRight, depth, counts) inside the timed region; beside it
          the bare-copy ceiling of this box measured in the same job (orb_slam2_b200/hostfeed.py)
  verified    : after each timed region (outside it) every rank re-computes a seeded choice of the pairs it
                just processed with the CPU oracle and compares bytes; any mismatch fails the run
  roofline    : the dominant kernel, algorithmic bytes / CUDA-event time vs measured HBM peak
  cpu_baseline: the reference's own CPU code for the path on this box's host cores (N=1, rank 0)
  secondary   : BASELINE.json configs 1, 3, 4, 5 (and the SURVEY 8f rows) at every N, each with a CPU twin
  latency     : the reference's call pattern - one image / one stereo frame per call (N=1)
  --impl reference: the CPU path alone, all host threads, same metric and config (rank 0 only)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from orb_slam2_b200 import synth  # noqa: E402

CFG = synth.KITTI
PARAMS = (CFG["nfeatures"], 1.2, 8, 20, 7)
NSETS = 3            # input sets rotated between steps so inputs are never L2-resident from the last step
L2_MB = 126
KITTI_CAM = (718.856, 718.856, 607.1928, 185.2157)
EUROC_CAM = (435.2047, 435.2047, 367.4517, 252.2008)
TUM_CAM = (517.306408, 516.469215, 318.643040, 255.313989)
METRIC = "stereo_frames_per_sec_orb_extract_plus_stereo_match"


def bench_config(batch, world):
    """The `config` object of the JSON line: identical in both arms (--impl ours / reference) by construction."""
    w, h = CFG["w"], CFG["h"]
    return {"workload": "KITTI-shaped 1241x376 stereo, nFeatures=2000, 8 levels, scale 1.2, ORB extraction + ComputeStereoMatches (BASELINE.json configs[1])",
            "stereo_pairs_per_step_per_gpu": batch, "sharding": f"independent frames, {world} rank(s), no data-path collective",
            "l2": f"GPU arm: inputs rotate over {NSETS} sets of {2 * batch * w * h / 1e6:.0f} MB (> {L2_MB} MB L2 in total); per-step intermediates {2 * batch * 1.8:.0f} MB",
            "bf": CFG["bf"], "fx": CFG["fx"]}


def level_sizes(w, h):
    s = np.float32(1.0); out = []
    for l in range(8):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        s = np.float32(np.float64(s) * np.float64(np.float32(1.2)))
    return out


def algorithmic_bytes(w, h, nf, cands_per_eye, kps_per_eye):
    """Compulsory bytes per STEREO frame, per stage and for the whole path."""
    lv = level_sizes(w, h)
    raw = sum(a * b for a, b in lv)
    bordered = sum((a + 38) * (b + 38) for a, b in lv)
    st = {
        "pyramid": 2 * (w * h + bordered),
        "fast": 2 * (raw + 4 * cands_per_eye),
        "quadtree": 2 * (4 * cands_per_eye + 4 * kps_per_eye),
        "blur": 2 * (bordered + raw),                                  # K4a reads the bordered planes, writes the blurred ROIs
        "describe": 2 * kps_per_eye * (48 * 31 + 64 * 37 + 60),        # K4b: raw moments box + blurred tap box per keypoint, 60 B out
        "stereo_match": 2 * kps_per_eye * 60 + kps_per_eye * (11 * 32 + 8),
        "stereo_median": kps_per_eye * 12,
    }
    path = 2 * w * h + 2 * bordered + 2 * nf * 60 + nf * 8      # 4 666 350 B for KITTI
    return st, path


def make_inputs(rank, batch):
    """NSETS x (2*batch, h, w): first `batch` left eyes then `batch` right eyes.  Distinct seeded pairs;
    odd sets reuse the pairs flipped vertically (keeps the horizontal disparity structure)."""
    w, h = CFG["w"], CFG["h"]
    need = batch * ((NSETS + 1) // 2)
    pairs = [synth.stereo_pair(100000 * rank + i, w, h)[:2] for i in range(need)]
    sets = []
    for s in range(NSETS):
        chunk = pairs[(s // 2) * batch:(s // 2 + 1) * batch]
        flip = s % 2 == 1
        L = [np.flipud(p[0]) if flip else p[0] for p in chunk]
        R = [np.flipud(p[1]) if flip else p[1] for p in chunk]
        sets.append(np.ascontiguousarray(np.stack(L + R)))
    return sets


class ClockSampler:
    """SM clock and slowdown reasons DURING the timed regions: NVML polled every 10 ms from a thread (a 0.1 s region
    still yields ~10 samples); `nvidia-smi -lms` as the fall-back."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu, uuid=None):
        self.gpu, self.uuid, self.proc, self.lines, self.nv = gpu, uuid, None, [], None
        self.sm, self.mx, self.reasons, self.stop_flag = [], [], set(), False

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            if self.uuid:
                for u in (self.uuid, "GPU-" + self.uuid):
                    try:
                        h = pynvml.nvmlDeviceGetHandleByUUID(u); break
                    except Exception:
                        pass
            self.h = h or pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            self.nv = pynvml
            self.mx = [float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))]
            self.t = threading.Thread(target=self._poll, daemon=True); self.t.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.01)

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.nv is not None:
            self.stop_flag = True
            self.t.join(timeout=1.0)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml, 100 Hz"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi -lms 100"}


def measured_hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(stage, batch):
    """DRAM bytes per launch of a stage from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if t.get("stereo_pairs_per_step") == batch:
            return t["dram_bytes_per_step"].get(stage)
    except Exception:
        pass
    return None


def measured_issue(stage):
    """Issue-slot / ALU-pipe / L1 utilisation of a stage from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return dict(t["issue"][stage], source=t["issue"]["source"])
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------------------------
# CPU legs (the reference's own code compiled in place, oracle/_ref/orbref_slam; the oracle port when it is absent)
# ------------------------------------------------------------------------------------------------------------------
def cpu_sample(nframes, distinct=32):
    """nframes KITTI-shaped stereo pairs for the CPU legs (`distinct` painted ones and their variants)."""
    return synth.stereo_set(900000, min(nframes, distinct), nframes, CFG["w"], CFG["h"])


def reference_built():
    """oracle/_ref/orbref_slam: the reference's own Frame.cc / ORBextractor.cc / ORBmatcher.cc compiled in place (it travels
    with the snapshot; built by `make -C oracle ref` where /root/reference exists)."""
    from oracle import pyoracle as O
    return O.ref_slam_available()


def cpu_run(L, R, threads, reps=1):
    """`reps` passes over the sample on the host cores.  With oracle/_ref present this is the REFERENCE's stereo Frame constructor
    (src/Frame.cc:61-117: both extractors in two threads, then ComputeStereoMatches), threads/2 frames in flight so that all
    `threads` cores are busy (kind "reference"; its OpenCV calls are served by oracle/cv_prims.hpp).  Otherwise the oracle
    port, one stereo frame per std::thread (kind "port").  Returns (seconds, kind)."""
    from oracle import pyoracle as O
    if reference_built():
        secs, _, _ = O.ref_slam_bench(L, R, CFG["nfeatures"], KITTI_CAM, CFG["bf"], max(1, threads // 2), reps=reps)
        return secs, "reference"
    secs, _, _ = O.bench_stereo(L, R, PARAMS, CFG["bf"], CFG["fx"], threads, reps=reps)
    return secs, "port"


def cpu_how(kind, threads):
    return ("the reference's stereo Frame constructor compiled in place (oracle/_ref/orbref_slam: src/Frame.cc, src/ORBextractor.cc, "
            f"src/ORBmatcher.cc; OpenCV calls served by oracle/cv_prims.hpp), {max(1, threads // 2)} frames in flight x 2 extractor threads"
            if kind == "reference" else "oracle port, one pair per std::thread")


def cpu_baseline(nframes, threads):
    """nframes stereo frames of CPU work: 4 * threads distinct pairs (the sample travels to the reference binary through a
    file), passed over nframes / (4 * threads) times."""
    base = min(nframes, 4 * threads)
    reps = max(1, nframes // base)
    L, R = cpu_sample(base)
    cpu_run(L[:threads], R[:threads], threads)                                  # warm caches / threads
    secs, kind = cpu_run(L, R, threads, reps)
    return base * reps / secs, secs, kind


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the host cores (see cpu_run), on the same config:
    every step is one batch of `--batch` KITTI-shaped stereo pairs."""
    if rank != 0:
        return
    from oracle import pyoracle as O
    O.build()
    threads = O.hardware_threads()
    per_step = args.batch
    base = min(per_step, 64)                                                    # distinct pairs (the sample goes to the binary through a file)
    reps = max(1, per_step // base)
    per_step = base * reps
    L, R = cpu_sample(base)
    kind = "port"
    for _ in range(args.warmup):
        cpu_run(L[:2 * threads], R[:2 * threads], threads)                      # warm-up steps are short: a CPU has no clocks to ramp
    secs = 0.0
    for _ in range(args.steps):
        s1, kind = cpu_run(L, R, threads, reps)
        secs += s1
    val = per_step * args.steps / secs
    sample = f"{per_step} KITTI-shaped stereo pairs per step, {cpu_how(kind, threads)}, {threads} host threads"
    line = {"impl": "reference", "metric": METRIC, "value": val,
            "unit": "stereo frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000 * secs / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic", "config": bench_config(args.batch, world),
            "cpu_baseline": {"value": val, "unit": "stereo frames/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": "stereo frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_twins(threads):
    """CPU figures beside the `secondary` GPU numbers, each a few seconds of the REFERENCE's own code on all host threads
    (oracle/_ref/orbref_slam bench modes; nothing when the binary did not travel)."""
    from oracle import pyoracle as O
    if not reference_built():
        return {"unavailable": "oracle/_ref/orbref_slam not present"}
    out = {"cores": threads, "kind": "reference", "how": "oracle/_ref/orbref_slam bench: the reference's own Frame / ORBmatcher members, one unit per std::thread"}
    half = max(1, threads // 2)
    # config 3: EuRoC stereo Frame constructor
    c = synth.EUROC
    L, R = synth.stereo_set(910000, 8, 4 * half, c["w"], c["h"])
    s, _, _ = O.ref_slam_bench(L, R, c["nfeatures"], EUROC_CAM, c["bf"], half)
    out["euroc_4096_sequence_stereo_frames_per_s"] = len(L) / s
    # config 1: TUM monocular Frame constructor (ORBextractor(1000) + UndistortKeyPoints + AssignFeaturesToGrid)
    F = np.stack([synth.variant(synth.mono_frame(920000 + i % 8, 640, 480), i // 8) for i in range(6 * threads)])
    s, _, _ = O.ref_slam_bench(F, None, 1000, TUM_CAM, 0.0, threads, mode="mono")
    out["tum_mono_frames_per_s"] = len(F) / s
    # config 4: SearchForInitialization on frame pairs built beforehand
    mp = [synth.mono_pair(700000 + i, 640, 480)[:2] for i in range(8)]
    A = np.stack([mp[i % 8][0] for i in range(threads)]); B = np.stack([mp[i % 8][1] for i in range(threads)])
    reps = 40
    s, _, m, lat = O.ref_slam_bench(A, B, 1000, TUM_CAM, 0.0, threads, reps=reps, mode="init", ratio=0.9, window=100, detail=True)
    out["search_for_initialization_pairs_per_s"] = threads * reps / s
    out["search_for_initialization_ms_per_call"] = lat[0]
    # config 5: SearchByBoW(KF,KF), one bucket of 2000 x 2000
    mp = [synth.mono_pair(800000 + i, 1241, 376)[:2] for i in range(8)]
    A = np.stack([mp[i % 8][0] for i in range(threads)]); B = np.stack([mp[i % 8][1] for i in range(threads)])
    s, _, m, lat = O.ref_slam_bench(A, B, 2000, KITTI_CAM, 0.0, threads, reps=2, mode="bow", ratio=0.75, detail=True)
    out["knn_ratio_2000x2000_pairs_per_s"] = threads * 2 / s
    out["knn_ratio_ms_per_call"] = lat[0]
    return out


def cv2_crosscheck():
    """What real OpenCV (the cv2 4.13 wheel: SSE/AVX dispatch, IPP) buys over oracle/cv_prims.hpp on the OpenCV share of the
    path: per KITTI image, one thread, the chained resize + copyMakeBorder pyramid, FAST-9 (threshold 20, NMS) and the 7x7
    sigma-2 blur on all 8 levels.  Labelled separately: it bounds how much faster a production build of the reference
    could be than the shim-served reference arm; it is not the reference's path (no cells, quad-tree, descriptors, stereo)."""
    try:
        import cv2
        from oracle import pyoracle as O
        cv2.setNumThreads(1)
        imgs = [synth.mono_frame(930000 + i, CFG["w"], CFG["h"]) for i in range(4)]
        lv = level_sizes(CFG["w"], CFG["h"])
        fast = cv2.FastFeatureDetector_create(20, True)

        def with_cv2(im):
            cur = im
            for l, (w, h) in enumerate(lv):
                if l:
                    cur = cv2.resize(cur, (w, h), interpolation=cv2.INTER_LINEAR)
                cv2.copyMakeBorder(cur, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
                fast.detect(cur)
                cv2.GaussianBlur(cur, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)

        def with_shim(im):
            cur = im
            for l, (w, h) in enumerate(lv):
                if l:
                    cur = O.resize(cur, w, h)
                O.border(cur, 19)
                O.fast(cur, 20)
                O.blur(cur)

        def t(fn):
            fn(imgs[0])
            t0 = time.perf_counter()
            for _ in range(3):
                for im in imgs:
                    fn(im)
            return (time.perf_counter() - t0) / (3 * len(imgs)) * 1e3
        a, b = t(with_cv2), t(with_shim)
        return {"cv2_version": cv2.__version__, "cv2_ms_per_image": a, "shim_ms_per_image": b, "shim_over_cv2": b / a,
                "what": "pyramid (resize+border) + whole-level FAST-9(20, NMS) + GaussianBlur 7x7 on 8 levels of one 1241x376 image, single thread"}
    except Exception as exc:
        return {"unavailable": repr(exc)}


# ------------------------------------------------------------------------------------------------------------------
# checks against the CPU oracle, outside the timed regions
# ------------------------------------------------------------------------------------------------------------------
def verify_pairs(images, sel, kps, desc, cnt, ur, dp, nm, batch, cap, params, bf, fx):
    """images (2*batch, h, w) host array of the step; sel = indices of the pairs to check; kps/desc/cnt [2*batch][cap] and
    ur/dp/nm [batch] = what the CUDA path returned for the step (numpy).  Returns the number of mismatching pairs."""
    from oracle import pyoracle as O
    L = np.stack([images[i] for i in sel]); R = np.stack([images[batch + i] for i in sel])
    okps, odesc, ocnt, our, odp, onm = O.batch_stereo(L, R, params, bf, fx, cap)
    n = len(sel)
    bad = 0
    for j, i in enumerate(sel):
        ok = True
        for gi, oi in ((i, j), (batch + i, n + j)):
            c = int(ocnt[oi])
            ok &= int(cnt[gi]) == c and kps[gi, :c].tobytes() == okps[oi, :c].tobytes() and desc[gi, :c].tobytes() == odesc[oi, :c].tobytes()
        c = int(ocnt[j])
        ok &= int(nm[i]) == int(onm[j]) and ur[i, :c].tobytes() == our[j, :c].tobytes() and dp[i, :c].tobytes() == odp[j, :c].tobytes()
        bad += not ok
    return bad


# ------------------------------------------------------------------------------------------------------------------
# secondary workloads (BASELINE.json configs 1, 3, 4, 5 + SURVEY 8f rows), every rank, any world size
# ------------------------------------------------------------------------------------------------------------------
def secondary_workloads(gx, torch, local, rank, world):
    """Returns ({name: (units, ms)} for the timed workloads of THIS rank, extras dict).  Device-resident inputs, CUDA events."""
    import ctypes as C
    from orb_slam2_b200 import sharding
    from orb_slam2_b200.api import KP_DTYPE
    out, extra = {}, {}
    ev = lambda: torch.cuda.Event(enable_timing=True)
    st = torch.cuda.current_stream().cuda_stream

    def timed(fn, reps, warm=2):
        for _ in range(warm):
            fn()
        a, b = ev(), ev(); a.record()
        for _ in range(reps):
            fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b)

    # --- config 3: a 4096-frame EuRoC-shaped stereo sequence in contiguous blocks over the ranks (strong scaling) ---
    cfg = synth.EUROC
    w, h, total, B = cfg["w"], cfg["h"], 4096, 256
    lo, hi = sharding.shard_range(total, world, rank)
    mine = hi - lo
    nb = -(-mine // B)
    # the resident part of the shard: 512 distinct images per eye (16 painted pairs per rank and their variants), cycled by block
    Ls, Rs = synth.stereo_set(500000 + 1000 * rank, 16, min(mine, 512), w, h)
    dev = [torch.from_numpy(np.concatenate([Ls[i:i + B], Rs[i:i + B]])).cuda() for i in range(0, len(Ls), B)]
    G = gx.ORBextractor(cfg["nfeatures"], 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2 * B, device=local)
    cap = G.kp_cap
    ur = torch.empty((B, cap), dtype=torch.float32, device="cuda"); dp = torch.empty_like(ur)
    nm = torch.empty(B, dtype=torch.int32, device="cuda")
    def euroc_block(i):
        d = dev[i % len(dev)]
        nb_ = d.shape[0] // 2
        G.extract_batch(d, 0, st)
        gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=nb_, slotL0=0, slotR0=nb_, out=(ur, dp, nm), stream=st)
    for i in range(min(2, nb)):
        euroc_block(i)
    torch.cuda.synchronize()
    a, b = ev(), ev(); a.record()
    for i in range(nb):
        euroc_block(i)
    b.record(); torch.cuda.synchronize()
    out["euroc_4096_sequence_stereo_frames_per_s"] = (mine, a.elapsed_time(b))
    # self-check of the last block against the oracle (4 pairs per rank)
    last = (nb - 1) % len(dev)
    nlast = dev[last].shape[0] // 2
    res_k = torch.empty((2 * nlast, cap, 28), dtype=torch.uint8, device="cuda"); res_d = torch.empty((2 * nlast, cap, 32), dtype=torch.uint8, device="cuda")
    res_c = torch.empty(2 * nlast, dtype=torch.int32, device="cuda")
    G.fetch_into(0, 2 * nlast, res_k, res_d, res_c, st)
    torch.cuda.synchronize()
    sel = sorted(np.random.default_rng(77 + rank).choice(nlast, 4, replace=False).tolist())
    bad = verify_pairs(dev[last].cpu().numpy(), sel, res_k.cpu().numpy().view(KP_DTYPE).reshape(2 * nlast, cap), res_d.cpu().numpy(), res_c.cpu().numpy(),
                       ur.cpu().numpy(), dp.cpu().numpy(), nm.cpu().numpy(), nlast, cap, (cfg["nfeatures"], 1.2, 8, 20, 7), cfg["bf"], cfg["fx"])
    extra["euroc_verified"] = (len(sel), bad)
    G.close(); del dev, ur, dp, nm, res_k, res_d, res_c
    # --- config 1: TUM-shaped monocular frames, ORBextractor(1000) ---
    w, h, P = 640, 480, 512
    G = gx.ORBextractor(1000, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=P, device=local)
    frames = np.stack([synth.variant(synth.mono_frame(600000 + 100 * rank + i % 8, w, h), i // 8) for i in range(P)])
    dframes = torch.from_numpy(frames).cuda()
    out["tum_mono_frames_per_s"] = (5 * P, timed(lambda: G.extract_batch(dframes, 0, st), 5))
    G.close(); del dframes
    # --- config 4: SearchForInitialization, 640x480 pairs, 1000 features, ratio 0.9, window 100 ---
    P = 256
    G = gx.ORBextractor(1000, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2 * P, device=local)
    cap = G.kp_cap
    mp = [synth.mono_pair(700000 + i, w, h)[:2] for i in range(8)]
    batch = np.stack([mp[i % 8][0] for i in range(P)] + [mp[i % 8][1] for i in range(P)])
    G.extract_batch(torch.from_numpy(batch).cuda(), 0, st)
    kps = torch.empty((2 * P, cap, 28), dtype=torch.uint8, device="cuda"); desc = torch.empty((2 * P, cap, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.empty(2 * P, dtype=torch.int32, device="cuda")
    G.fetch_into(0, 2 * P, kps, desc, cnt, st)
    torch.cuda.synchronize()
    prev0 = kps.view(torch.float32).reshape(2 * P, cap, 7)[:P, :, :2].contiguous()
    m12 = torch.empty((P, cap), dtype=torch.int32, device="cuda"); nmm = torch.empty(P, dtype=torch.int32, device="cuda")
    L = gx.lib()
    gb = torch.tensor([0, w, 0, h, 64.0 / w, 48.0 / h], dtype=torch.float32, device="cuda")     # undistorted camera: Frame's bounds = the image
    def init_step():
        prev = prev0.clone()
        rc = L.orbx_search_init_batch(kps.data_ptr(), desc.data_ptr(), cnt.data_ptr(), kps[P:].data_ptr(), desc[P:].data_ptr(), cnt[P:].data_ptr(),
                                      P, cap, gb.data_ptr(), prev.data_ptr(), 0.9, 100, 1, m12.data_ptr(), nmm.data_ptr(), 1, st)
        assert rc == 0, L.orbx_last_error_string()
    out["search_for_initialization_pairs_per_s"] = (5 * P, timed(init_step, 5))
    extra["search_for_initialization_matches_per_pair"] = float(nmm.float().mean())
    G.close()
    # --- config 5: brute-force 256-bit kNN + ratio, 2000 x 2000 (KITTI-shaped frames), ratio 0.75 ---
    w, h, P = 1241, 376, 128
    G = gx.ORBextractor(2000, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2 * P, device=local)
    cap = G.kp_cap
    mp = [synth.mono_pair(800000 + i, w, h)[:2] for i in range(8)]
    batch = np.stack([mp[i % 8][0] for i in range(P)] + [mp[i % 8][1] for i in range(P)])
    G.extract_batch(torch.from_numpy(batch).cuda(), 0, st)
    kps = torch.empty((2 * P, cap, 28), dtype=torch.uint8, device="cuda"); desc = torch.empty((2 * P, cap, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.empty(2 * P, dtype=torch.int32, device="cuda")
    G.fetch_into(0, 2 * P, kps, desc, cnt, st)
    torch.cuda.synchronize()
    ang = kps.view(torch.float32).reshape(2 * P, cap, 7)[:, :, 3].contiguous()
    m12 = torch.empty((P, cap), dtype=torch.int32, device="cuda"); nmm = torch.empty(P, dtype=torch.int32, device="cuda")
    def knn_step():
        rc = L.orbx_knn_ratio_batch(desc.data_ptr(), ang.data_ptr(), cnt.data_ptr(), desc[P:].data_ptr(), ang[P:].data_ptr(), cnt[P:].data_ptr(),
                                    P, cap, 50, 0.75, 1, 1, m12.data_ptr(), nmm.data_ptr(), 1, st)
        assert rc == 0, L.orbx_last_error_string()
    out["knn_ratio_2000x2000_pairs_per_s"] = (3 * P, timed(knn_step, 3))
    extra["knn_ratio_matches_per_pair"] = float(nmm.float().mean())
    extra["_knn_distances_per_pair"] = float((cnt[:P].double() * cnt[P:].double()).mean())
    # --- SearchByBoW(KF,KF) over vocabulary buckets; stand-in vocabulary: 100 nodes from descriptor bits ---
    node = ((desc[:, :, 0].int() >> 4) * 7 + (desc[:, :, 1].int() >> 5)).contiguous()          # 0..111, equal for most true matches
    valid = torch.ones((2 * P, cap), dtype=torch.uint8, device="cuda")
    def bow_step():
        rc = L.orbx_search_by_bow_batch(desc.data_ptr(), ang.data_ptr(), valid.data_ptr(), node.data_ptr(), cnt.data_ptr(),
                                        desc[P:].data_ptr(), ang[P:].data_ptr(), valid[P:].data_ptr(), node[P:].data_ptr(), cnt[P:].data_ptr(),
                                        P, cap, 50, 0.75, 1, 1, m12.data_ptr(), nmm.data_ptr(), 1, st)
        assert rc == 0, L.orbx_last_error_string()
    out["search_by_bow_buckets_pairs_per_s"] = (3 * P, timed(bow_step, 3))
    extra["search_by_bow_matches_per_pair"] = float(nmm.float().mean())
    # --- DBoW2 transform (Frame::ComputeBoW) of the same frames against a synthetic vocabulary of
    #     ORBvoc.txt's shape (k = 10, L = 6: 1.1 M nodes, 60 Hamming distances per feature), levelsup = 4 ---
    V = synth.vocabulary_fast(4242, 10, 6)
    voc = gx.ORBVocabulary(V, weighting=0, norm=1, device=local)
    F2 = 2 * P
    wo = torch.empty((F2, cap), dtype=torch.int32, device="cuda"); no = torch.empty_like(wo); bw = torch.empty_like(wo)
    bv = torch.empty((F2, cap), dtype=torch.float64, device="cuda"); bc = torch.empty(F2, dtype=torch.int32, device="cuda")
    def voc_step():
        rc = L.orbx_bow_transform_batch(voc.h, desc.data_ptr(), cnt.data_ptr(), F2, cap, 4, wo.data_ptr(), no.data_ptr(), bw.data_ptr(),
                                        bv.data_ptr(), bc.data_ptr(), 1, st)
        assert rc == 0, L.orbx_last_error_string()
    out["bow_transform_frames_per_s"] = (3 * F2, timed(voc_step, 3))
    extra["bow_transform_words_per_frame"] = float(bc.float().mean())
    voc.close()
    # --- UndistortKeyPoints + image bounds + grid for the frames above (TUM1 distortion), and
    #     GetFeaturesInArea queries against them (window 100 px around every keypoint of the first frames) ---
    cam = (C.c_float * 9)(517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
    F = 2 * P
    kun = torch.empty_like(kps); bounds = torch.empty(6, dtype=torch.float32, device="cuda")
    cstart = torch.empty((F, 64 * 48 + 1), dtype=torch.int32, device="cuda"); citems = torch.empty((F, cap), dtype=torch.int32, device="cuda")
    def grid_step():
        rc = L.orbx_undistort_grid_batch(kps.data_ptr(), cnt.data_ptr(), F, cap, cam, w, h, kun.data_ptr(), bounds.data_ptr(),
                                         cstart.data_ptr(), citems.data_ptr(), 1, st)
        assert rc == 0, L.orbx_last_error_string()
    out["undistort_grid_frames_per_s"] = (5 * F, timed(grid_step, 5))
    nq = 64 * 2000
    kf = kun.view(torch.float32).reshape(F, cap, 7)
    qf = torch.arange(64, dtype=torch.int32, device="cuda").repeat_interleave(2000)
    qi = torch.arange(2000, device="cuda").repeat(64)
    qx = kf[qf.long(), qi, 0].contiguous(); qy = kf[qf.long(), qi, 1].contiguous()
    qr = torch.full((nq,), 100.0, dtype=torch.float32, device="cuda")
    qlo = torch.zeros(nq, dtype=torch.int32, device="cuda"); qhi = torch.zeros(nq, dtype=torch.int32, device="cuda")
    oidx = torch.empty((nq, 256), dtype=torch.int32, device="cuda"); ocnt = torch.empty(nq, dtype=torch.int32, device="cuda")
    def area_step():
        rc = L.orbx_features_in_area_batch(kun.data_ptr(), cnt.data_ptr(), F, cap, bounds.data_ptr(), cstart.data_ptr(), citems.data_ptr(),
                                           qf.data_ptr(), qx.data_ptr(), qy.data_ptr(), qr.data_ptr(), qlo.data_ptr(), qhi.data_ptr(),
                                           nq, 256, oidx.data_ptr(), ocnt.data_ptr(), 1, st)
        assert rc == 0, L.orbx_last_error_string()
    out["features_in_area_queries_per_s"] = (5 * nq, timed(area_step, 5))
    extra["features_in_area_hits_per_query"] = float(ocnt.float().mean())
    G.close()
    return out, extra


# ------------------------------------------------------------------------------------------------------------------
# latency: the reference's call pattern (one image per operator(), one stereo frame per constructor)
# ------------------------------------------------------------------------------------------------------------------
def pcts(ms):
    a = np.sort(np.asarray(ms, np.float64))
    return {"p50_ms": float(a[len(a) // 2]), "p99_ms": float(a[min(len(a) - 1, int(0.99 * len(a)))]), "calls": int(len(a))}


def latency_block(gx, torch, local):
    """(1) orbx_extract on one host KITTI image, (2) a whole stereo frame the way src/Frame.cc:78-90 runs it - two handles on
    two host threads and two streams, then orbx_stereo_match_batch(n=1) -, (3) the reference's stereo Frame constructor itself
    over the drop-in (oracle/_ref/orbref_slam_gpustereo: dropin/ORBextractor.h + dropin/FrameStereo.cc) and (4) the same
    constructor all-reference on the CPU (2 extractor threads), same images."""
    from concurrent.futures import ThreadPoolExecutor
    import ctypes as C
    out = {}
    w, h, nf = CFG["w"], CFG["h"], CFG["nfeatures"]
    Ls, Rs = synth.stereo_set(940000, 8, 16, w, h)
    L = gx.lib()
    GL = gx.ORBextractor(*PARAMS, max_w=w, max_h=h, max_batch=1, device=local)
    GR = gx.ORBextractor(*PARAMS, max_w=w, max_h=h, max_batch=1, device=local)
    sL, sR = C.c_void_p(), C.c_void_p()
    assert L.orbx_stream_create(local, C.byref(sL)) == 0 and L.orbx_stream_create(local, C.byref(sR)) == 0
    for i in range(8):
        GL(Ls[i], stream=sL)
    t = []
    for i in range(200):
        t0 = time.perf_counter(); GL(Ls[i % 16], stream=sL); t.append((time.perf_counter() - t0) * 1e3)
    out["extract_one_image"] = dict(pcts(t), what="ORBextractor::operator() on one host 1241x376 image: H2D, 13 launches, D2H of keypoints + descriptors, synchronous")
    pool = ThreadPoolExecutor(2)
    cap = GL.kp_cap
    ur = np.zeros((1, cap), np.float32); dp = np.zeros((1, cap), np.float32); nm = np.zeros(1, np.int32)
    def frame(i):
        fl = pool.submit(lambda: GL(Ls[i % 16], stream=sL)); fr = pool.submit(lambda: GR(Rs[i % 16], stream=sR))
        fl.result(); fr.result()
        rc = L.orbx_stereo_match_batch(GL._h, 0, GR._h, 0, 1, CFG["bf"], CFG["fx"], ur.ctypes.data, dp.ctypes.data, nm.ctypes.data, cap, 0, sL)
        assert rc == 0
        L.orbx_stream_synchronize(sL)
    for i in range(8):
        frame(i)
    t = []
    for i in range(200):
        t0 = time.perf_counter(); frame(i); t.append((time.perf_counter() - t0) * 1e3)
    out["stereo_frame_two_handles"] = dict(pcts(t), what="left + right operator() on two host threads / two streams, then ComputeStereoMatches (n=1), results on the host")
    pool.shutdown()
    GL.close(); GR.close()
    L.orbx_stream_destroy(sL); L.orbx_stream_destroy(sR)
    try:
        from oracle import pyoracle as O
        gpu_bin = os.path.join(os.path.dirname(O.REF_SLAM_BIN), "orbref_slam_gpustereo")
        if O.ref_slam_available():
            s, _, m, lat = O.ref_slam_bench(Ls, Rs, nf, KITTI_CAM, CFG["bf"], 1, reps=4, detail=True)
            out["frame_constructor_reference_cpu"] = {"p50_ms": lat[0], "p99_ms": lat[1], "calls": 64,
                                                      "what": "the reference's stereo Frame constructor (src/Frame.cc:61-117), 2 extractor threads, one frame in flight"}
        if os.path.exists(gpu_bin):
            s, _, m2, lat = O.ref_slam_bench(Ls, Rs, nf, KITTI_CAM, CFG["bf"], 1, reps=13, detail=True, binary=gpu_bin)
            out["frame_constructor_dropin_gpu"] = {"p50_ms": lat[0], "p99_ms": lat[1], "calls": 208,
                                                   "what": "the same constructor, unmodified src/Frame.cc, with dropin/ORBextractor.h + dropin/FrameStereo.cc over liborbx.so"}
        # per-call latency of the ORBmatcher members on real KeyFrames / MapPoints (the kf case of the parity tests): the same driver
        # source compiled with the reference's src/ORBmatcher.cc (CPU) and with dropin/ORBmatcher.cc over liborbx.so
        drop_bin = os.path.join(os.path.dirname(O.REF_SLAM_BIN), "orbref_slam_gpu")
        if O.ref_slam_available() and os.path.exists(drop_bin):
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import gen_golden_slam as gs
            ims, V, T1, T2, F12 = gs.kf_case_inputs(5, 5)
            kw = dict(reps=40, check_ori=True, **gs.KF_PARAMS)
            ref, sizes = O.ref_slam_kfbench(*ims, nf, KITTI_CAM, CFG["bf"], V, T1, T2, F12, **kw)
            gpu, _ = O.ref_slam_kfbench(*ims, nf, KITTI_CAM, CFG["bf"], V, T1, T2, F12, binary=drop_bin, **kw)
            out["matcher_members"] = {"what": "p50 / p99 ms of one call on the kf case (2004-feature KITTI-shaped keyframes, %d / %d map points), 40 calls each; "
                                              "reference = src/ORBmatcher.cc on the CPU, dropin = dropin/ORBmatcher.cc (flatten + one C-ABI call + write-back)" % sizes[2:],
                                      "members": {k: {"reference_cpu_p50_ms": ref[k][0], "reference_cpu_p99_ms": ref[k][1], "dropin_gpu_p50_ms": gpu[k][0],
                                                      "dropin_gpu_p99_ms": gpu[k][1], "matches": [ref[k][2], gpu[k][2]]} for k in ref if k in gpu}}
    except Exception as exc:
        out["frame_constructor_error"] = repr(exc)
    return out


def spread_device(torch, local, world):
    """The GPU of local rank `local`.  When the job sees more GPUs than it has ranks, the ranks are spread evenly over them
    (rank r -> device r * ndev / world) instead of taking the first `world`: on this pool's 8-GPU boxes GPUs 0-3 and 4-7
    hang off two host bridges with separate upload ceilings (profiles/: host-feed probe - four uploads through one bridge get
    29 GB/s each, two get 55), and the end-to-end leg is bound by exactly that.  With as many GPUs as ranks: the identity."""
    try:
        ndev = torch.cuda.device_count()
        if os.environ.get("ORBX_BENCH_SPREAD", "1") != "0" and ndev > world and ndev % world == 0:
            return local * (ndev // world)
    except Exception:
        pass
    return local


def bind_near_gpu(torch, local):
    """Pins this rank's threads to the CPU cores NVML reports as local to its GPU, so that the pinned host
    buffers of the end-to-end leg are allocated (first touched) on the GPU's own NUMA node.  Best effort."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(local).uuid)
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(local)
        before = len(os.sched_getaffinity(0))
        pynvml.nvmlDeviceSetCpuAffinity(h)
        after = len(os.sched_getaffinity(0))
        return f"nvml gpu-local cpus: {after} of {before}"
    except Exception as exc:
        return f"unchanged ({type(exc).__name__})"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="stereo pairs per step per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the configs 1/3/4/5 side measurements, the latency block and the cv2 cross-check")
    ap.add_argument("--no-verify", action="store_true", help="skip the oracle self-check after the timed regions")
    ap.add_argument("--verify-pairs", type=int, default=8, help="pairs per rank and leg re-computed by the CPU oracle")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import orb_slam2_b200 as gx
    from orb_slam2_b200 import hostfeed, sharding
    from orb_slam2_b200.api import KP_DTYPE
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the CUDA path has no CPU fallback")
    local = spread_device(torch, local, world)
    torch.cuda.set_device(local)
    all_cpus = os.sched_getaffinity(0)
    affinity = bind_near_gpu(torch, local)      # before any pinned allocation: first touch decides the NUMA node
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dd = dist if world > 1 else None
    B, w, h, nf = args.batch, CFG["w"], CFG["h"], CFG["nfeatures"]
    bf, fx = CFG["bf"], CFG["fx"]

    sets = make_inputs(rank, B)
    host = [torch.from_numpy(s).pin_memory() for s in sets]
    dev = [t.cuda(non_blocking=True) for t in host]
    NPIPE = max(2, int(os.environ.get("ORBX_BENCH_PIPE", "3")))   # handles / streams / pinned result sets in flight in the end-to-end leg
    G = [gx.ORBextractor(*PARAMS, max_w=w, max_h=h, max_batch=2 * B, device=local) for _ in range(NPIPE)]
    cap = G[0].kp_cap
    d_ur = torch.empty((B, cap), dtype=torch.float32, device="cuda"); d_dp = torch.empty_like(d_ur)
    d_nm = torch.empty(B, dtype=torch.int32, device="cuda")
    streams = [torch.cuda.Stream() for _ in range(NPIPE)]
    # pinned result buffers for the end-to-end leg, one set per pipeline stage
    h_out = [dict(pk=torch.empty((2 * B, cap), dtype=torch.int64).pin_memory(),          # 8-byte keypoints as they cross the bus
                  lc=torch.empty((2 * B, 8), dtype=torch.int32).pin_memory(),
                  kps=np.zeros((2 * B, cap), KP_DTYPE), cnt=np.zeros(2 * B, np.int32), # the caller's cv::KeyPoint arrays, filled on the host ("packed")
                  kps_pinned=torch.empty((2 * B, cap, 28), dtype=torch.uint8).pin_memory(),   # ... or by the copy itself ("full")
                  cnt_pinned=torch.empty(2 * B, dtype=torch.int32).pin_memory(),
                  done=torch.cuda.Event(blocking=True),           # the worker thread sleeps on it instead of polling the driver
                  desc=torch.empty((2 * B, cap, 32), dtype=torch.uint8).pin_memory(),
                  ur=torch.empty((B, cap), dtype=torch.float32).pin_memory(),
                  dp=torch.empty((B, cap), dtype=torch.float32).pin_memory(),
                  nm=torch.empty(B, dtype=torch.int32).pin_memory()) for _ in range(NPIPE)]
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()

    def step_resident(i, g, st):
        g.extract_batch(dev[i % NSETS], 0, st.cuda_stream)
        gx.ComputeStereoMatches(g, g, bf, fx, n=B, slotL0=0, slotR0=B, out=(d_ur, d_dp, d_nm), stream=st.cuda_stream)

    # How the keypoints come back: "full" = the 28-byte cv::KeyPoint records by D2H copy (default), "packed" = 8 bytes per
    # keypoint over the bus + orbx_unpack_keypoints on the host (inside the timed region).  Measured on this pool's boxes
    # (profiles/README.md, round 2): packed loses 3-6 % at 1 GPU and 3 % at 8 - the download is not what holds the leg, and the
    # unpack costs 5 ms of one core per step and rank - so it is an option (ORBX_BENCH_D2H=packed) for hosts with a slow
    # device-to-host path, not the default.
    e2e_mode = os.environ.get("ORBX_BENCH_D2H", "full")

    def step_e2e(i, g, st, o):
        s = st.cuda_stream
        g.extract_batch(host[i % NSETS], 0, s)                                    # H2D of 2B images inside
        gx.ComputeStereoMatches(g, g, bf, fx, n=B, slotL0=0, slotR0=B, out=(o["ur"], o["dp"], o["nm"]), stream=s)   # + D2H uRight/depth
        if e2e_mode == "packed":
            g.fetch_packed_into(0, 2 * B, o["pk"], o["desc"], o["lc"], s)         # D2H keypoints (8 bytes each), descriptors, per-level counts
        else:
            g.fetch_into(0, 2 * B, o["kps_pinned"], o["desc"], o["cnt_pinned"], s)   # D2H cv::KeyPoint records, descriptors, counts
        o["done"].record(st)

    def finish_e2e(o):
        """Host side of a step: wait for its copies; packed mode rebuilds the cv::KeyPoint records in the caller's arrays."""
        o["done"].synchronize()
        if e2e_mode == "packed":
            G[0].unpack_keypoints(o["pk"].numpy(), o["lc"].numpy(), o["kps"], o["cnt"])

    # ---- device-resident throughput ("value") ----
    st0 = streams[0]
    for i in range(args.warmup):
        step_resident(i, G[0], st0)
    torch.cuda.synchronize()
    G[0].set_profiling(True)
    try:
        uuid = str(torch.cuda.get_device_properties(local).uuid)
    except Exception:
        uuid = None
    sampler = ClockSampler(local, uuid); sampler.start()
    launches0 = gx.kernel_launch_count()
    barrier(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(st0)
    for i in range(args.steps):
        step_resident(i, G[0], st0)
    e1.record(st0)
    torch.cuda.synchronize(); barrier()
    ms = e0.elapsed_time(e1)
    launches = gx.kernel_launch_count() - launches0
    prof = G[0].get_profile()
    G[0].set_profiling(False)
    # results of the last resident step, as arrays [2B][cap]
    r_kps = np.zeros((2 * B, cap), KP_DTYPE); r_desc = np.zeros((2 * B, cap, 32), np.uint8); r_cnt = np.zeros(2 * B, np.int32)
    G[0].fetch_into(0, 2 * B, r_kps, r_desc, r_cnt)
    torch.cuda.synchronize()
    nkp = r_cnt.astype(np.int64); matches = d_nm.cpu().numpy()
    cands = 0
    for l in range(8):
        cands += len(G[0].debug_candidates(0, l))
    checksum = int(np.bitwise_xor.reduce(np.frombuffer(r_desc[0, :r_cnt[0]].tobytes(), np.uint64))) if r_cnt[0] else 0
    ver_pairs = ver_bad = 0
    if not args.no_verify:
        sel = sorted(np.random.default_rng(1234 + rank).choice(B, min(args.verify_pairs, B), replace=False).tolist())
        ver_bad += verify_pairs(sets[(args.steps - 1) % NSETS], sel, r_kps, r_desc, r_cnt, d_ur.cpu().numpy(), d_dp.cpu().numpy(), matches, B, cap, PARAMS, bf, fx)
        ver_pairs += len(sel)

    # ---- end to end through the C ABI with host buffers, NPIPE handles/streams in flight ----
    # Per step: H2D of the images, the kernels, D2H of the results, and on the host the unpacking of the keypoints into
    # cv::KeyPoint records - done for step i - NPIPE right before its pipeline slot is reused, and for the last NPIPE steps
    # before the closing event is recorded, so all of it lies inside the timed region.
    from concurrent.futures import ThreadPoolExecutor
    pool = ThreadPoolExecutor(NPIPE)                      # the host side of a step runs beside the enqueueing thread (ctypes drops the GIL)
    fut = [None] * NPIPE
    for i in range(max(args.warmup, NPIPE)):
        if fut[i % NPIPE] is not None:
            fut[i % NPIPE].result()
        step_e2e(i, G[i % NPIPE], streams[i % NPIPE], h_out[i % NPIPE])
        fut[i % NPIPE] = pool.submit(finish_e2e, h_out[i % NPIPE])
    for f in fut:
        f.result()
    torch.cuda.synchronize(); barrier()
    ee0 = torch.cuda.Event(enable_timing=True); ee1 = torch.cuda.Event(enable_timing=True)
    ee0.record(streams[0])
    for k in range(1, NPIPE):
        streams[k].wait_stream(streams[0])               # nothing of the timed steps starts before ee0
    for i in range(args.steps):
        if i >= NPIPE:
            fut[i % NPIPE].result()                      # the slot's previous step has been unpacked into the caller's arrays
        step_e2e(i, G[i % NPIPE], streams[i % NPIPE], h_out[i % NPIPE])
        fut[i % NPIPE] = pool.submit(finish_e2e, h_out[i % NPIPE])
    for f in fut:
        f.result()                                       # the last steps' host work, still before the closing event
    for k in range(1, NPIPE):
        streams[0].wait_stream(streams[k])
    ee1.record(streams[0])
    torch.cuda.synchronize(); barrier()
    ms_e2e = ee0.elapsed_time(ee1)
    pool.shutdown()
    clocks = sampler.stop()
    last = h_out[(args.steps - 1) % NPIPE]
    if e2e_mode != "packed":
        last["kps"] = last["kps_pinned"].numpy().view(KP_DTYPE).reshape(2 * B, cap); last["cnt"] = last["cnt_pinned"].numpy()
    assert int(last["cnt"].sum()) > 0 and int(last["nm"].sum()) > 0
    if not args.no_verify:
        sel = sorted(np.random.default_rng(4321 + rank).choice(B, min(args.verify_pairs, B), replace=False).tolist())
        ver_bad += verify_pairs(sets[(args.steps - 1) % NSETS], sel, last["kps"], last["desc"].numpy(),
                                last["cnt"], last["ur"].numpy(), last["dp"].numpy(), last["nm"].numpy(), B, cap, PARAMS, bf, fx)
        ver_pairs += len(sel)

    # ---- the host-feed ceiling of this box, same job, no kernels: bare pinned copies of the step's sizes on all ranks at once ----
    h2d = 2 * B * w * h
    if e2e_mode == "packed":
        d2h = 2 * B * cap * 40 + 2 * B * 32 + 2 * B * cap * 4 + B * 4  # 8-byte keypoints + descriptors + per-level counts, uRight + depth, match counts
    else:
        d2h = 2 * B * cap * 60 + 2 * B * 4 + 2 * B * cap * 4 + B * 4   # cv::KeyPoint records + descriptors + counts, uRight + depth, match counts
    for g in G:
        g.close()
    del dev
    feed = hostfeed.probe(torch, h2d, d2h, streams=(1, 3), reps=4, dist=dd)
    feed_all = hostfeed.gather(torch, dd, [feed["h2d_gbs"][1], feed["h2d_gbs"][3], feed["d2h_gbs"][3], *feed["duplex_gbs"][3]])

    # ---- max over ranks, totals over ranks ----
    ms, ms_e2e = sharding.max_over_ranks([ms, ms_e2e], device="cuda")
    rec = sharding.make_record(B * args.steps, int(nkp.sum()), int(matches.sum()), checksum)
    rec_all = sharding.gather_records(rec, device="cuda")     # counts + checksums
    ver_all = sharding.gather_records(sharding.make_record(ver_pairs, ver_bad, 0, 0), device="cuda")
    frames_total = int(rec_all[:, 0].sum())

    # ---- side workloads on every rank (BASELINE.json configs 1, 3, 4, 5) ----
    sec_local, sec_extra, sec_err = {}, {}, None
    if not args.no_secondary:
        try:
            sec_local, sec_extra = secondary_workloads(gx, torch, local, rank, world)
        except Exception as exc:                       # side measurements must never cost the headline
            sec_err = repr(exc)
            try:
                torch.cuda.synchronize()
            except Exception:
                pass
    names = sorted(sec_local)
    sec_all = None
    if not args.no_secondary:
        flags = hostfeed.gather(torch, dd, [0.0 if sec_err else 1.0, float(len(names))])
        if all(f[0] == 1.0 and int(f[1]) == len(names) for f in flags) and names:
            sec_all = hostfeed.gather(torch, dd, [v for n in names for v in sec_local[n]] + list(sec_extra.get("euroc_verified", (0, 0))))

    if rank == 0:
        value = frames_total / (ms / 1000.0)
        e2e_val = frames_total / (ms_e2e / 1000.0)
        kps_eye = float(nkp.mean()); cands_eye = float(cands)
        stage_bytes, path_bytes = algorithmic_bytes(w, h, nf, cands_eye, kps_eye)
        stage_ms = {s: (v[0] / v[1] if v[1] else 0.0) for s, v in prof.items()}
        top = max(stage_ms, key=stage_ms.get)
        peak, peak_src = measured_hbm_peak()
        ach = stage_bytes[top] * B / (stage_ms[top] / 1000.0) / 1e9 if stage_ms[top] > 0 else 0.0
        e2e_step_s = ms_e2e / args.steps / 1000.0
        ceil_h2d = min(max(v[0], v[1]) for v in feed_all)           # slowest rank's bare H2D rate (best of 1 / 3 streams), all ranks copying at once
        ceil_dup = min(v[3] for v in feed_all)                      # the same with the D2H of a step's results running beside it
        line = {
            "metric": METRIC, "value": value, "unit": "stereo frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": bench_config(B, world),
            "e2e": {"value": e2e_val, "unit": "stereo frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "pipeline": f"{NPIPE} handles x {NPIPE} streams, pinned host buffers; a pipeline slot is reused only after the host has its step's results",
                    "d2h_mode": e2e_mode + (": keypoints cross the bus as 8-byte records and are rebuilt as cv::KeyPoint on the host inside the timed region" if e2e_mode == "packed"
                                            else ": cv::KeyPoint records copied as they are"), "cpu_affinity": affinity,
                    "h2d_gbs": h2d / e2e_step_s / 1e9, "d2h_gbs": d2h / e2e_step_s / 1e9,
                    "ceiling_gbs": ceil_h2d, "ceiling_duplex_h2d_gbs": ceil_dup, "frac_of_ceiling": h2d / e2e_step_s / 1e9 / ceil_h2d,
                    "frac_of_duplex_ceiling": h2d / e2e_step_s / 1e9 / ceil_dup,   # against uploads WITH a step's downloads running beside them: what the leg does
                    "ceiling": {"what": "bare cudaMemcpyAsync of one step's bytes from/to pinned host memory, no kernels, all ranks at once (orb_slam2_b200/hostfeed.py); per-GPU GB/s, [rank]",
                                "h2d_1stream": [round(v[0], 2) for v in feed_all], "h2d_3streams": [round(v[1], 2) for v in feed_all],
                                "d2h_3streams": [round(v[2], 2) for v in feed_all], "duplex3_h2d": [round(v[3], 2) for v in feed_all],
                                "duplex3_d2h": [round(v[4], 2) for v in feed_all]}},
            "gpu_launches": int(launches),
            "devices": {"rank0_device": local, "visible": torch.cuda.device_count(), "placement": "ranks spread evenly over the visible GPUs (spread_device)"},
            "clocks": clocks,
            "verified": {"pairs": int(ver_all[:, 0].sum()), "mismatches": int(ver_all[:, 1].sum()), "per_rank": ver_all[:, :2].tolist(),
                         "what": "after each timed region every rank re-computes a seeded choice of the pairs of its last step with the CPU oracle: keypoint bytes, "
                                 "descriptor bytes, uRight / depth bits and match counts of both legs (device-resident and end to end)"},
            "roofline": {"bound": "hbm", "kernel": top, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": measured_traffic(top, B), "peak_source": peak_src, "issue": measured_issue(top),
                         "algorithmic_bytes_per_launch": stage_bytes[top] * B,
                         "note": "front end is instruction-issue / latency bound, not HBM bound (DESIGN.md); see profiles/ for ncu issue utilisation"},
            "stage_ms_per_step": stage_ms, "stage_calls": {s: v[1] for s, v in prof.items()},
            "path": {"algorithmic_bytes_per_stereo_frame": path_bytes, "achieved_gbs": path_bytes * B / (ms / args.steps / 1000.0) / 1e9,
                     "frac_of_hbm_peak": path_bytes * B / (ms / args.steps / 1000.0) / 1e9 / peak},
            "results": {"keypoints_per_eye": kps_eye, "fast_candidates_per_eye": cands_eye, "stereo_matches_per_frame": float(matches.mean()),
                        "per_rank": rec_all.tolist()},
        }
        if not args.no_secondary:
            if sec_all is not None:
                sec = {}
                for j, n in enumerate(names):
                    units = sum(v[2 * j] for v in sec_all); tmax = max(v[2 * j + 1] for v in sec_all)
                    sec[n] = units / (tmax / 1e3)
                sec.update({k: v for k, v in sec_extra.items() if not k.startswith("_") and k != "euroc_verified"})
                sec["euroc_4096_sequence"] = {"frames": 4096, "shards": [list(sharding.shard_range(4096, world, r)) for r in range(world)], "scaling": "strong",
                                              "verified": {"pairs": int(sum(v[-2] for v in sec_all)), "mismatches": int(sum(v[-1] for v in sec_all))}}
                # K7 is POPC-issue bound: 4 POPC per 256-bit distance (carry-save tree) against the measured
                # 0.5 warp-POPC / clk / SM (tools/pipe_probe.cu, profiles/r1d_pipe_rates.jsonl)
                mhz = (clocks or {}).get("sm_mhz") or 1920.0
                sec["knn_popc_issue_frac"] = sec["knn_ratio_2000x2000_pairs_per_s"] / world * sec_extra["_knn_distances_per_pair"] * 4 / (148 * 16 * mhz * 1e6)
                sec["note"] = "whole-job rates: units of all ranks / slowest rank's CUDA-event time; device-resident inputs"
                line["secondary"] = sec
            else:
                line["secondary"] = {"error": sec_err or "a rank failed"}
        os.sched_setaffinity(0, all_cpus)               # the CPU legs get every core again
        if not args.no_secondary:
            try:
                from oracle import pyoracle as O
                O.build()
                line["secondary_cpu"] = cpu_twins(O.hardware_threads())
            except Exception as exc:
                line["secondary_cpu"] = {"error": repr(exc)}
        if world == 1 and not args.no_cpu_baseline:
            from oracle import pyoracle as O
            O.build()
            threads = O.hardware_threads()
            nfr = 64 * threads                              # ~4 s of wall time on all cores (>= 60 core-seconds)
            v, secs, kind = cpu_baseline(nfr, threads)
            line["cpu_baseline"] = {"value": v, "unit": "stereo frames/s", "cores": threads, "kind": kind,
                                    "sample": f"{nfr} KITTI-shaped stereo pairs, {cpu_how(kind, threads)}, {threads} host threads, {secs:.1f} s wall = {secs * threads:.0f} core-seconds"}
            if kind == "reference":                          # the faster restatement beside it, for scale
                pv = nfr / O.bench_stereo(*cpu_sample(4 * threads), PARAMS, CFG["bf"], CFG["fx"], threads, reps=16)[0]
                line["cpu_baseline"]["oracle_port_value"] = pv
            if not args.no_secondary:
                line["cpu_cv2_crosscheck"] = cv2_crosscheck()
        else:
            line["cpu_baseline"] = None
        if world == 1 and not args.no_secondary:
            try:
                line["latency"] = latency_block(gx, torch, local)
            except Exception as exc:
                line["latency"] = {"error": repr(exc)}
        print(json.dumps(line), flush=True)
    barrier()
    if world > 1:
        dist.destroy_process_group()
    if int(ver_all[:, 1].sum()) != 0:
        raise SystemExit(f"bench.py: {int(ver_all[:, 1].sum())} of {int(ver_all[:, 0].sum())} re-computed stereo pairs differ from the CPU oracle")


if __name__ == "__main__":
    main()
