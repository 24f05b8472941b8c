#!/usr/bin/env python3
"""Generates tests/golden/ref_dist.json (+ ref_kfd_*.npz) from the REFERENCE ITSELF (oracle/_ref/orbref_slam, see
tools/gen_golden_slam.py) for DISTORTED cameras: there Frame::mnMinX / mnMinY are not 0, undistorted keypoints may lie left
of / above the image, and the 64x48 feature grid is laid over [mnMinX, mnMaxX) x [mnMinY, mnMaxY) (src/Frame.cc:207-213,
382-392, 327-346).  Covers the three members that read Frame's bounds: SearchForInitialization (src/ORBmatcher.cc:405),
SearchByProjection(Frame, MapPoints) (:45) and SearchByProjection(Current, Last) (:1328, image test :1373-1376).
Run in the build container only."""
import hashlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from orb_slam2_b200 import synth
from oracle import pyoracle as O
import gen_golden_slam as gs

sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
# KITTI intrinsics with a mild barrel distortion: the undistorted corners move outwards, mnMinX ~ -90 px, so keypoints
# near the left / top border get negative undistorted coordinates (with mnMinX = 0 assumed they would fall out of the grid)
KITTI_DIST = gs.KITTI_CAM + (-0.12, 0.02, 0.001, -0.0005, 0.0)
INIT = [(63, 100, 0.9, True), (64, 60, 0.9, False)]        # (seed, window, ratio, check_ori) with the full TUM1 distortion
KFD = [(8, 5, True), (9, 5, False)]                          # (seed, vocabulary levels, check_ori) with KITTI_DIST


if __name__ == "__main__":
    out = {"generator": "tools/gen_golden_dist.py", "source": "oracle/_ref/orbref_slam (reference sources compiled in place)", "init": [], "kf": []}
    for seed, win, ratio, ori in INIT:
        a, b, _ = synth.mono_pair(seed, 640, 480)
        F1, F2, n, m12, prev = O.ref_slam_init(a, b, 1000, gs.TUM1, ratio, win, ori)
        out["init"].append(dict(seed=seed, window=win, ratio=ratio, check_ori=ori, cam=list(gs.TUM1), nmatches=n, n1=len(F1["keys"]), n2=len(F2["keys"]),
                                keys_un2_sha256=sha(F2["keys_un"]), matches12_sha256=sha(m12), prev_sha256=sha(prev)))
    cfg = synth.KITTI
    for seed, L, ori in KFD:
        ims, V, T1, T2, F12 = gs.kf_case_inputs(seed, L)
        o = O.ref_slam_kf(*ims, cfg["nfeatures"], KITTI_DIST, cfg["bf"], V, T1, T2, F12, check_ori=ori, **gs.KF_PARAMS)
        name = f"ref_kfd_{seed}.npz"
        np.savez_compressed(os.path.join(ROOT, "tests/golden", name), mps1=o["mps1"], track=o["track"], has1=o["has1"], kf_bounds=o["kf_bounds"],
                            proj=o["proj"][1], proj_last=o["proj_last"][1])
        out["kf"].append(dict(seed=seed, L=L, check_ori=ori, cam=list(KITTI_DIST), fixture=name, n1=len(o["F1"]["keys"]), n2=len(o["F2"]["keys"]),
                              keys_un1_sha256=sha(o["F1"]["keys_un"]), keys_un2_sha256=sha(o["F2"]["keys_un"]), u_right2_sha256=sha(o["F2"]["u_right"]),
                              negative_x=int((o["F2"]["keys_un"]["x"] < 0).sum()), n_proj=o["proj"][0], n_proj_last=o["proj_last"][0]))
    json.dump(out, open(os.path.join(ROOT, "tests/golden/ref_dist.json"), "w"), indent=1)
    print(json.dumps(out["kf"], indent=1)[:1500])
    print([(c["nmatches"], c["n1"]) for c in out["init"]])
