"""How much of a drop-in matcher call is the C-ABI call itself (staging + kernel + sync): K8 / K9 with host arrays, one frame."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import orb_slam2_b200 as gx
from orb_slam2_b200 import synth
cfg = synth.KITTI; w, h, nf = cfg["w"], cfg["h"], cfg["nfeatures"]
G = gx.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2)
L, R, _ = synth.stereo_pair(90, w, h)
G.extract_batch(np.stack([L, R]))
ur, dp, nm = gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=1, slotL0=0, slotR0=1)
(kf, df), _ = G.fetch(0, 2)
n = len(kf); rng = np.random.default_rng(1); m = 400
sel = rng.choice(n, m, replace=False)
mp = dict(inview=np.ones(m, np.uint8), projx=(kf["x"][sel] + rng.normal(0, 2, m)).astype(np.float32), projy=(kf["y"][sel] + rng.normal(0, 2, m)).astype(np.float32),
          level=kf["octave"][sel].astype(np.int32), viewcos=np.full(m, 0.9, np.float32), desc=df[sel].copy(), obs=np.ones(m, np.int32))
mp["projxr"] = (mp["projx"] - 10).astype(np.float32)
scale = G.GetScaleFactors(); F = gx.Frame(kf, df, w, h); M = gx.ORBmatcher(0.8, True)
for env in ("", "1"):
    if env: os.environ["ORBX_PROJ_SERIAL"] = "1"
    for _ in range(5): M.SearchByProjection(F, ur[0, :n], scale, mp, 3.0)
    t = []
    for _ in range(100):
        t0 = time.perf_counter(); r = M.SearchByProjection(F, ur[0, :n], scale, mp, 3.0); t.append((time.perf_counter() - t0) * 1e3)
    print("SearchByProjection(Frame, 400 map points) python wrapper + C-ABI call, %s kernel: p50 %.3f ms, matches %d" % ("one-warp" if env else "8-warp two-phase", np.median(t), r[0]))
