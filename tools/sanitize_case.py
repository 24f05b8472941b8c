"""Small end-to-end case for compute-sanitizer (memcheck / racecheck / initcheck / synccheck), one tool per run:
   compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import orb_slam2_b200 as gx
from orb_slam2_b200 import synth

w, h, nf = 400, 300, 500
L, R, _ = synth.stereo_pair(1, w, h)
G = gx.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=4)
G.extract_batch(np.stack([L, L, R, R]))
ur, dp, nm = gx.ComputeStereoMatches(G, G, 40.0, 300.0, n=2, slotL0=0, slotR0=2)
res = G.fetch(0, 4)
a, b, _ = synth.mono_pair(2, w, h)
k1, d1 = G(a); k2, d2 = G(b)
prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
n1, m12 = gx.ORBmatcher(0.9, True).SearchForInitialization(gx.Frame(k1, d1, w, h), gx.Frame(k2, d2, w, h), prev, 100)
n2, m = gx.ORBmatcher(0.75, True).SearchByBoW(d1, k1["angle"], d2, k2["angle"])
print("sanitize case ok:", [len(r[0]) for r in res], nm.tolist(), n1, n2)
