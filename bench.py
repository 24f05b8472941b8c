#!/usr/bin/env python3
"""bench.py — stereo frames/s of ORB extraction + stereo matching (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A step is one pass of the hot path over one batch of B synthetic KITTI-shaped stereo pairs
(1241x376, nFeatures 2000, 8 levels, 1.2; BASELINE.json configs[1]) on every rank:
pyramid -> FAST cells -> quad-tree -> orientation+blur+rBRIEF (both eyes) -> ComputeStereoMatches.
Independent frames are sharded over ranks (weak scaling, no data-path collective); NCCL only
gathers {frames, keypoints, matches, checksum} after the timed region.

  value : whole-job stereo frames/s, inputs already resident in HBM
  e2e   : the same through the C ABI with pinned HOST buffers: H2D of every image and D2H of all
          results (keypoints, descriptors, uRight, depth, counts) inside the timed region
  roofline    : the dominant kernel, algorithmic bytes / CUDA-event time vs measured HBM peak
  cpu_baseline: the CPU oracle (port of the reference's path) on this box's host cores
  --impl reference: the CPU path alone, all host threads, same metric (rank 0 only)
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
        "describe": 2 * kps_per_eye * (43 * 43 + 60),
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
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu):
        self.gpu, self.proc, self.lines = gpu, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
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
                "reasons": sorted(reasons), "samples": len(sm)}


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


def cpu_sample(nframes, distinct=32):
    """nframes KITTI-shaped stereo pairs for the CPU legs (at most `distinct` different ones, tiled)."""
    w, h = CFG["w"], CFG["h"]
    pairs = [synth.stereo_pair(900000 + i, w, h)[:2] for i in range(min(nframes, distinct))]
    idx = [i % len(pairs) for i in range(nframes)]
    return np.stack([pairs[i][0] for i in idx]), np.stack([pairs[i][1] for i in idx])


KITTI_CAM = (718.856, 718.856, 607.1928, 185.2157)


def reference_built():
    """oracle/_ref/orbref_slam: the reference's own Frame.cc / ORBextractor.cc / ORBmatcher.cc compiled in place (it travels
    with the snapshot; built by `make -C oracle ref` where /root/reference exists)."""
    from oracle import pyoracle as O
    return O.ref_slam_available()


def cpu_run(L, R, threads):
    """One pass over the sample on the host cores.  With oracle/_ref present this is the REFERENCE's stereo Frame constructor
    (src/Frame.cc:61-117: both extractors in two threads, then ComputeStereoMatches), threads/2 frames in flight so that all
    `threads` cores are busy (kind "reference"; its OpenCV calls are served by oracle/cv_prims.hpp).  Otherwise the oracle
    port, one stereo frame per std::thread (kind "port").  Returns (seconds, kind)."""
    from oracle import pyoracle as O
    if reference_built():
        secs, _, _ = O.ref_slam_bench(L, R, CFG["nfeatures"], KITTI_CAM, CFG["bf"], max(1, threads // 2))
        return secs, "reference"
    secs, _, _ = O.bench_stereo(L, R, PARAMS, CFG["bf"], CFG["fx"], threads)
    return secs, "port"


def cpu_baseline(nframes, threads):
    L, R = cpu_sample(nframes)
    cpu_run(L[:threads], R[:threads], threads)                                  # warm caches / threads
    secs, kind = cpu_run(L, R, threads)
    return nframes / secs, secs, kind


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the host cores (see cpu_run)."""
    if rank != 0:
        return
    from oracle import pyoracle as O
    O.build()
    threads = O.hardware_threads()
    per_step = 2 * threads
    L, R = cpu_sample(per_step)
    kind = "port"
    for _ in range(args.warmup):
        cpu_run(L, R, threads)
    secs = 0.0
    for _ in range(args.steps):
        s1, kind = cpu_run(L, R, threads)
        secs += s1
    val = per_step * args.steps / secs
    how = ("the reference's stereo Frame constructor (oracle/_ref/orbref_slam), " + str(max(1, threads // 2)) + " frames in flight x 2 extractor threads"
           if kind == "reference" else "oracle port, one pair per std::thread")
    sample = f"{per_step} KITTI-shaped stereo pairs per step, {how}, {threads} host threads"
    line = {"impl": "reference", "metric": "stereo_frames_per_sec_orb_extract_plus_stereo_match", "value": val,
            "unit": "stereo frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000 * secs / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": "KITTI-shaped 1241x376 stereo, nFeatures=2000, 8 levels, scale 1.2, ORB extraction + ComputeStereoMatches",
                       "frames_per_step": per_step},
            "cpu_baseline": {"value": val, "unit": "stereo frames/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": "stereo frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def secondary_workloads(gx, torch, local):
    """BASELINE.json configs 3-5 as side measurements (single GPU, device-resident, CUDA events):
    EuRoC-shaped stereo frames/s, SearchForInitialization pairs/s (TUM shape, 1000 features),
    brute-force kNN+ratio pairs/s (2000 x 2000 descriptors).  Not the headline; reported for coverage."""
    import ctypes as C
    out = {}
    ev = lambda: torch.cuda.Event(enable_timing=True)
    st = torch.cuda.current_stream().cuda_stream
    # --- config 3: EuRoC 752x480 stereo, nFeatures 1200 ---
    cfg = synth.EUROC
    w, h, B = cfg["w"], cfg["h"], 128
    pairs = [synth.stereo_pair(500000 + i, w, h)[:2] for i in range(16)]
    batch = np.stack([pairs[i % 16][0] for i in range(B)] + [pairs[i % 16][1] for i in range(B)])
    dev = torch.from_numpy(batch).cuda()
    G = gx.ORBextractor(cfg["nfeatures"], 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2 * B, device=local)
    cap = G.kp_cap
    ur = torch.empty((B, cap), dtype=torch.float32, device="cuda"); dp = torch.empty_like(ur)
    nm = torch.empty(B, dtype=torch.int32, device="cuda")
    def step():
        G.extract_batch(dev, 0, st)
        gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=B, slotL0=0, slotR0=B, out=(ur, dp, nm), stream=st)
    for _ in range(3):
        step()
    a, b = ev(), ev(); a.record(); [step() for _ in range(5)]; b.record(); torch.cuda.synchronize()
    out["euroc_stereo_frames_per_s"] = 5 * B / (a.elapsed_time(b) / 1e3)
    G.close()
    # --- config 4: SearchForInitialization, 640x480 pairs, 1000 features, ratio 0.9, window 100 ---
    w, h, P = 640, 480, 256
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
    for _ in range(2):
        init_step()
    a, b = ev(), ev(); a.record(); [init_step() for _ in range(5)]; b.record(); torch.cuda.synchronize()
    out["search_for_initialization_pairs_per_s"] = 5 * P / (a.elapsed_time(b) / 1e3)
    out["search_for_initialization_matches_per_pair"] = float(nmm.float().mean())
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
    for _ in range(2):
        knn_step()
    a, b = ev(), ev(); a.record(); [knn_step() for _ in range(3)]; b.record(); torch.cuda.synchronize()
    out["knn_ratio_2000x2000_pairs_per_s"] = 3 * P / (a.elapsed_time(b) / 1e3)
    out["knn_ratio_matches_per_pair"] = float(nmm.float().mean())
    out["_knn_distances_per_pair"] = float((cnt[:P].double() * cnt[P:].double()).mean())
    # --- SearchByBoW(KF,KF) over vocabulary buckets; stand-in vocabulary: 100 nodes from descriptor bits ---
    node = ((desc[:, :, 0].int() >> 4) * 7 + (desc[:, :, 1].int() >> 5)).contiguous()          # 0..111, equal for most true matches
    valid = torch.ones((2 * P, cap), dtype=torch.uint8, device="cuda")
    def bow_step():
        rc = L.orbx_search_by_bow_batch(desc.data_ptr(), ang.data_ptr(), valid.data_ptr(), node.data_ptr(), cnt.data_ptr(),
                                        desc[P:].data_ptr(), ang[P:].data_ptr(), valid[P:].data_ptr(), node[P:].data_ptr(), cnt[P:].data_ptr(),
                                        P, cap, 50, 0.75, 1, 1, m12.data_ptr(), nmm.data_ptr(), 1, st)
        assert rc == 0, L.orbx_last_error_string()
    for _ in range(2):
        bow_step()
    a, b = ev(), ev(); a.record(); [bow_step() for _ in range(3)]; b.record(); torch.cuda.synchronize()
    out["search_by_bow_buckets_pairs_per_s"] = 3 * P / (a.elapsed_time(b) / 1e3)
    out["search_by_bow_matches_per_pair"] = float(nmm.float().mean())
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
    for _ in range(2):
        voc_step()
    a, b = ev(), ev(); a.record(); [voc_step() for _ in range(3)]; b.record(); torch.cuda.synchronize()
    out["bow_transform_frames_per_s"] = 3 * F2 / (a.elapsed_time(b) / 1e3)
    out["bow_transform_words_per_frame"] = float(bc.float().mean())
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
    for _ in range(2):
        grid_step()
    a, b = ev(), ev(); a.record(); [grid_step() for _ in range(5)]; b.record(); torch.cuda.synchronize()
    out["undistort_grid_frames_per_s"] = 5 * F / (a.elapsed_time(b) / 1e3)
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
    for _ in range(2):
        area_step()
    a, b = ev(), ev(); a.record(); [area_step() for _ in range(5)]; b.record(); torch.cuda.synchronize()
    out["features_in_area_queries_per_s"] = 5 * nq / (a.elapsed_time(b) / 1e3)
    out["features_in_area_hits_per_query"] = float(ocnt.float().mean())
    G.close()
    return out


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
    ap.add_argument("--no-secondary", action="store_true", help="skip the configs 3-5 side measurements")
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
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the CUDA path has no CPU fallback")
    torch.cuda.set_device(local)
    all_cpus = os.sched_getaffinity(0)
    affinity = bind_near_gpu(torch, local)      # before any pinned allocation: first touch decides the NUMA node
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
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
    h_out = [dict(kps=torch.empty((2 * B, cap, 28), dtype=torch.uint8).pin_memory(),
                  desc=torch.empty((2 * B, cap, 32), dtype=torch.uint8).pin_memory(),
                  cnt=torch.empty(2 * B, dtype=torch.int32).pin_memory(),
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

    def step_e2e(i, g, st, o):
        s = st.cuda_stream
        g.extract_batch(host[i % NSETS], 0, s)                                    # H2D of 2B images inside
        gx.ComputeStereoMatches(g, g, bf, fx, n=B, slotL0=0, slotR0=B, out=(o["ur"], o["dp"], o["nm"]), stream=s)   # + D2H uRight/depth
        g.fetch_into(0, 2 * B, o["kps"], o["desc"], o["cnt"], s)                  # D2H keypoints, descriptors, counts

    # ---- device-resident throughput ("value") ----
    st0 = streams[0]
    for i in range(args.warmup):
        step_resident(i, G[0], st0)
    torch.cuda.synchronize()
    G[0].set_profiling(True)
    sampler = ClockSampler(local); sampler.start()
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
    res = G[0].fetch(0, 2 * B)
    nkp = np.array([len(r[0]) for r in res]); matches = d_nm.cpu().numpy()
    cands = 0
    for l in range(8):
        cands += len(G[0].debug_candidates(0, l))
    checksum = int(np.bitwise_xor.reduce(np.frombuffer(res[0][1].tobytes(), np.uint64))) if len(res[0][1]) else 0

    # ---- end to end through the C ABI with host buffers, two handles/streams ping-pong ----
    for i in range(max(args.warmup, NPIPE)):
        step_e2e(i, G[i % NPIPE], streams[i % NPIPE], h_out[i % NPIPE])
    torch.cuda.synchronize(); barrier()
    ee0 = torch.cuda.Event(enable_timing=True); ee1 = torch.cuda.Event(enable_timing=True)
    ee0.record(streams[0])
    for k in range(1, NPIPE):
        streams[k].wait_stream(streams[0])               # nothing of the timed steps starts before ee0
    for i in range(args.steps):
        step_e2e(i, G[i % NPIPE], streams[i % NPIPE], h_out[i % NPIPE])
    for k in range(1, NPIPE):
        streams[0].wait_stream(streams[k])
    ee1.record(streams[0])
    torch.cuda.synchronize(); barrier()
    ms_e2e = ee0.elapsed_time(ee1)
    clocks = sampler.stop()
    last = h_out[(args.steps - 1) % NPIPE]
    assert int(last["cnt"].sum()) > 0 and int(last["nm"].sum()) > 0

    # ---- max over ranks, totals over ranks ----
    from orb_slam2_b200 import sharding
    ms, ms_e2e = sharding.max_over_ranks([ms, ms_e2e], device="cuda")
    rec = sharding.make_record(B * args.steps, int(nkp.sum()), int(matches.sum()), checksum)
    rec_all = sharding.gather_records(rec, device="cuda")     # the only collective of the job: counts + checksums
    frames_total = int(rec_all[:, 0].sum())

    if rank == 0:
        value = frames_total / (ms / 1000.0)
        e2e_val = frames_total / (ms_e2e / 1000.0)
        h2d = 2 * B * w * h
        d2h = 2 * B * cap * 60 + 2 * B * 4 + 2 * B * cap * 4 + B * 4
        kps_eye = float(nkp.mean()); cands_eye = float(cands)
        stage_bytes, path_bytes = algorithmic_bytes(w, h, nf, cands_eye, kps_eye)
        stage_ms = {s: (v[0] / v[1] if v[1] else 0.0) for s, v in prof.items()}
        top = max(stage_ms, key=stage_ms.get)
        peak, peak_src = measured_hbm_peak()
        ach = stage_bytes[top] * B / (stage_ms[top] / 1000.0) / 1e9 if stage_ms[top] > 0 else 0.0
        line = {
            "metric": "stereo_frames_per_sec_orb_extract_plus_stereo_match", "value": value, "unit": "stereo frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "KITTI-shaped 1241x376 stereo, nFeatures=2000, 8 levels, scale 1.2, ORB extraction + ComputeStereoMatches (BASELINE.json configs[1])",
                       "stereo_pairs_per_step_per_gpu": B, "sharding": f"independent frames, {world} rank(s), no data-path collective",
                       "l2": f"inputs rotate over {NSETS} sets of {2 * B * w * h / 1e6:.0f} MB (> {L2_MB} MB L2 in total); per-step intermediates {2 * B * 1.8:.0f} MB",
                       "bf": bf, "fx": fx},
            "e2e": {"value": e2e_val, "unit": "stereo frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "pipeline": f"{NPIPE} handles x {NPIPE} streams, pinned host buffers", "cpu_affinity": affinity},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": top, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": measured_traffic(top, B), "peak_source": peak_src, "issue": measured_issue(top),
                         "algorithmic_bytes_per_launch": stage_bytes[top] * B,
                         "note": "front end is instruction-issue / latency bound, not HBM bound (DESIGN.md); see profiles/ for ncu issue utilisation"},
            "stage_ms_per_step": stage_ms,
            "path": {"algorithmic_bytes_per_stereo_frame": path_bytes, "achieved_gbs": path_bytes * B / (ms / args.steps / 1000.0) / 1e9,
                     "frac_of_hbm_peak": path_bytes * B / (ms / args.steps / 1000.0) / 1e9 / peak},
            "results": {"keypoints_per_eye": kps_eye, "fast_candidates_per_eye": cands_eye, "stereo_matches_per_frame": float(matches.mean()),
                        "per_rank": rec_all.tolist()},
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import pyoracle as O
            O.build()
            os.sched_setaffinity(0, all_cpus)               # the CPU leg gets every core again
            threads = O.hardware_threads()
            nfr = 16 * threads                              # ~10-30 s of CPU work
            v, secs, kind = cpu_baseline(nfr, threads)
            how = ("the reference's stereo Frame constructor compiled in place (oracle/_ref/orbref_slam: src/Frame.cc, src/ORBextractor.cc, "
                   f"src/ORBmatcher.cc; OpenCV calls served by oracle/cv_prims.hpp), {max(1, threads // 2)} frames in flight x 2 extractor threads"
                   if kind == "reference" else "oracle port, one pair per std::thread")
            line["cpu_baseline"] = {"value": v, "unit": "stereo frames/s", "cores": threads, "kind": kind,
                                    "sample": f"{nfr} KITTI-shaped stereo pairs, {how}, {threads} host threads, {secs:.1f} s"}
            if kind == "reference":                          # the faster restatement beside it, for scale
                pv = nfr / O.bench_stereo(*cpu_sample(nfr), PARAMS, CFG["bf"], CFG["fx"], threads)[0]
                line["cpu_baseline"]["oracle_port_value"] = pv
        else:
            line["cpu_baseline"] = None
        if world == 1 and not args.no_secondary:
            try:
                for g in G:
                    g.close()
                sec = secondary_workloads(gx, torch, local)
                # K7 is POPC-issue bound: 4 POPC per 256-bit distance (carry-save tree) against the measured
                # 0.5 warp-POPC / clk / SM (tools/pipe_probe.cu, profiles/r1d_pipe_rates.jsonl)
                mhz = (line.get("clocks") or {}).get("sm_mhz") or 1920.0
                popc_peak = 148 * 16 * mhz * 1e6
                sec["knn_popc_issue_frac"] = sec["knn_ratio_2000x2000_pairs_per_s"] * sec.pop("_knn_distances_per_pair") * 4 / popc_peak
                line["secondary"] = sec
            except Exception as exc:                       # side measurements must never cost the headline
                line["secondary"] = {"error": repr(exc)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
