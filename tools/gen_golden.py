#!/usr/bin/env python3
"""Generates tests/golden/ref_extract.json from the REFERENCE ITSELF: oracle/_ref/orbref is the
reference's src/ORBextractor.cc compiled in place (make -C oracle ref).  For each seeded synthetic
frame the file stores SHA-256 digests of the keypoint stream and the descriptor matrix, plus the
count and a few literal keypoints so a failure is readable.  Run in the build container only."""
import hashlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orb_slam2_b200 import synth
from oracle import pyoracle as O
import cv2

CASES = [("kitti", 1241, 376, 2000, [0, 1, 2]), ("euroc", 752, 480, 1200, [0, 1]),
         ("tum", 640, 480, 1000, [0, 1]), ("tum_ini2x", 640, 480, 2000, [0])]
out = {"generator": "tools/gen_golden.py", "source": "oracle/_ref/orbref (reference src/ORBextractor.cc, bump allocator)",
       "cv2_version_primitives_pinned_to": cv2.__version__, "base_seed": synth.BASE_SEED, "cases": []}
for name, w, h, nf, frames in CASES:
    imgs = np.stack([synth.mono_frame(f, w, h) for f in frames])
    res, _ = O.ref_extract(imgs, nf)
    for f, (k, d) in zip(frames, res):
        out["cases"].append(dict(name=name, w=w, h=h, nfeatures=nf, frame=f, count=len(k),
                                 img_sha256=hashlib.sha256(imgs[frames.index(f)].tobytes()).hexdigest(),
                                 kps_sha256=hashlib.sha256(k.tobytes()).hexdigest(),
                                 desc_sha256=hashlib.sha256(d.tobytes()).hexdigest(),
                                 per_level=[int((k["octave"] == l).sum()) for l in range(8)],
                                 first_kps=[[float(k[i]["x"]), float(k[i]["y"]), float(k[i]["angle"]), float(k[i]["response"]), int(k[i]["octave"])] for i in range(3)],
                                 first_desc=d[0].tolist()))
os.makedirs(os.path.join(ROOT, "tests/golden"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "tests/golden/ref_extract.json"), "w"), indent=1)
print("wrote", len(out["cases"]), "cases")
