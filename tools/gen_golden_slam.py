#!/usr/bin/env python3
"""Generates tests/golden/ref_slam.json from the REFERENCE ITSELF: oracle/_ref/orbref_slam is the reference's
src/Frame.cc, src/ORBmatcher.cc, src/ORBextractor.cc (+ KeyFrame, MapPoint, Map, DBoW2) compiled in place
(make -C oracle ref).  Seeded synthetic frames go through the reference's own Frame constructors and matcher
members; the file stores digests of what they produce.  Run in the build container only."""
import hashlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orb_slam2_b200 import synth
from oracle import pyoracle as O

sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
TUM1 = (517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
EUROC_RAW = (458.654, 457.296, 367.215, 248.375, -0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0)
KITTI_CAM = (718.856, 718.856, 607.1928, 185.2157)
STEREO = [("kitti", synth.KITTI, KITTI_CAM, [5, 6]), ("euroc", synth.EUROC, (435.2047, 435.2047, 367.4517, 252.2008), [5])]
INIT = [(60, 100, 0.9, True), (61, 40, 0.9, False), (62, 100, 0.8, True)]
GRID = [("tum1", 640, 480, 1000, TUM1, 300), ("euroc_raw", 752, 480, 1200, EUROC_RAW, 301), ("kitti", 1241, 376, 2000, KITTI_CAM, 302)]


def grid_queries(seed, w, h, n=200):
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        lo, hi = (-1, -1) if rng.random() < 0.4 else (int(rng.integers(0, 3)), int(rng.integers(-1, 8)))
        out.append((float(rng.uniform(-20, w + 20)), float(rng.uniform(-20, h + 20)), float(rng.uniform(2, 120)), lo, hi))
    return out


if __name__ == "__main__":
    out = {"generator": "tools/gen_golden_slam.py", "source": "oracle/_ref/orbref_slam (reference src/Frame.cc, src/ORBmatcher.cc, ... compiled in place)",
           "stereo": [], "init": [], "grid": []}
    for name, cfg, cam, seeds in STEREO:
        for s in seeds:
            L, R, _ = synth.stereo_pair(s, cfg["w"], cfg["h"])
            F = O.ref_slam_stereo(L, R, cfg["nfeatures"], cam, cfg["bf"])
            out["stereo"].append(dict(name=name, seed=s, n=len(F["keys"]), n_right=len(F["keys_right"]), matches=int((F["u_right"] >= 0).sum()),
                                      keys_sha256=sha(F["keys"]), u_right_sha256=sha(F["u_right"]), depth_sha256=sha(F["depth"])))
    for seed, win, ratio, ori in INIT:
        a, b, _ = synth.mono_pair(seed, 640, 480)
        F1, F2, n, m12, prev = O.ref_slam_init(a, b, 1000, TUM1[:4], ratio, win, ori)
        out["init"].append(dict(seed=seed, window=win, ratio=ratio, check_ori=ori, nmatches=n, n1=len(F1["keys"]), n2=len(F2["keys"]),
                                matches12_sha256=sha(m12), prev_sha256=sha(prev)))
    for name, w, h, nf, cam, frame in GRID:
        qs = grid_queries(frame, w, h)
        F, b, cells, hits = O.ref_slam_grid(synth.mono_frame(frame, w, h), nf, cam, qs)
        out["grid"].append(dict(name=name, w=w, h=h, nfeatures=nf, cam=list(cam), frame=frame, n=len(F["keys"]), keys_un_sha256=sha(F["keys_un"]),
                                bounds_hex=[float(v).hex() for v in b], cells_sha256=sha(np.concatenate([np.array([len(c)], np.int32) for c in cells] + cells)),
                                hits_sha256=sha(np.concatenate([np.array([len(x)], np.int32) for x in hits] + hits)), total_hits=int(sum(len(x) for x in hits))))
    json.dump(out, open(os.path.join(ROOT, "tests/golden/ref_slam.json"), "w"), indent=1)
    print("wrote", {k: len(v) for k, v in out.items() if isinstance(v, list)})
