"""Where the end-to-end leg loses time against the bare upload rate (one GPU): the same pipelined steps with parts removed.
python tools/exp_e2e.py  ->  ms per 256-pair step for: upload only | upload + kernels | + download | + host wait per slot"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import orb_slam2_b200 as gx
from orb_slam2_b200 import synth
cfg = synth.KITTI; w, h, B = cfg["w"], cfg["h"], 256
L, R = synth.stereo_set(100, 32, B, w, h)
host = [torch.from_numpy(np.concatenate([L, R])).pin_memory() for _ in range(3)]
def run(npipe, kernels, download, hostwait, steps=20):
    G = [gx.ORBextractor(2000, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2 * B) for _ in range(npipe)]
    cap = G[0].kp_cap
    st = [torch.cuda.Stream() for _ in range(npipe)]
    dev = [torch.empty((2 * B, h, w), dtype=torch.uint8, device="cuda") for _ in range(npipe)]
    out = [dict(kps=torch.empty((2 * B, cap, 28), dtype=torch.uint8).pin_memory(), desc=torch.empty((2 * B, cap, 32), dtype=torch.uint8).pin_memory(),
                cnt=torch.empty(2 * B, dtype=torch.int32).pin_memory(), ur=torch.empty((B, cap), dtype=torch.float32).pin_memory(),
                dp=torch.empty((B, cap), dtype=torch.float32).pin_memory(), nm=torch.empty(B, dtype=torch.int32).pin_memory(),
                ev=torch.cuda.Event(blocking=True)) for _ in range(npipe)]
    d_ur = torch.empty((B, cap), dtype=torch.float32, device="cuda"); d_dp = torch.empty_like(d_ur); d_nm = torch.empty(B, dtype=torch.int32, device="cuda")
    def step(i):
        k = i % npipe; s = st[k].cuda_stream; o = out[k]
        if hostwait and i >= npipe:
            o["ev"].synchronize()
        if kernels:
            G[k].extract_batch(host[i % 3], 0, s)
            gx.ComputeStereoMatches(G[k], G[k], cfg["bf"], cfg["fx"], n=B, slotL0=0, slotR0=B, out=(o["ur"], o["dp"], o["nm"]) if download else (d_ur, d_dp, d_nm), stream=s)
            if download:
                G[k].fetch_into(0, 2 * B, o["kps"], o["desc"], o["cnt"], s)
        else:
            with torch.cuda.stream(st[k]):
                dev[k].copy_(host[i % 3], non_blocking=True)
        o["ev"].record(st[k])
    for i in range(6): step(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(steps): step(i)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps * 1e3
    for g in G: g.close()
    return dt
for name, a in [("upload only", (3, False, False, False)), ("upload + kernels", (3, True, False, False)), ("upload + kernels + download", (3, True, True, False)),
                ("... + host waits for the slot's previous step", (3, True, True, True)), ("the same, 4 slots", (4, True, True, True)), ("upload + kernels, 2 slots", (2, True, False, False))]:
    ms = run(*a)
    print(f"{name:55s} {ms:6.3f} ms/step  {B / ms:6.1f} k stereo frames/s  upload {2 * B * w * h / ms / 1e6:5.1f} GB/s")
