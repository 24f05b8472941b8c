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


# (seed, vocabulary levels L, only_stereo, mono, check_ori): L = 3 with levelsup 4 puts every feature in the root node,
# i.e. SearchByBoW degenerates to the one-bucket brute-force search (K7)
KF_CASES = [(5, 5, False, False, True), (7, 3, False, True, False), (6, 5, True, False, True)]
KF_PARAMS = dict(pct1=60, pct2=50, ratio_kk=0.75, ratio_kf=0.7, th_proj=3.0, th_last=15.0)


def shifted(img, dx, dy, seed):
    rng = np.random.default_rng(seed)
    out = np.roll(np.roll(img, dy, 0), dx, 1).astype(np.int16) + rng.integers(-2, 3, img.shape)
    return np.clip(out, 0, 255).astype(np.uint8)


def kf_case_inputs(seed, L):
    """Two stereo frames (the second = the first shifted by (3, 1) px plus noise), a synthetic vocabulary, poses and F12."""
    cfg = synth.KITTI
    L1, R1, _ = synth.stereo_pair(seed, cfg["w"], cfg["h"])
    L2, R2 = shifted(L1, 3, 1, seed + 1000), shifted(R1, 3, 1, seed + 2000)
    V = synth.vocabulary(77 + seed, 10, L, stop_fraction=0.0)
    T1 = np.eye(4, dtype=np.float32); T2 = np.eye(4, dtype=np.float32); T2[:3, 3] = (0.05, -0.02, 0.0)
    K = np.array([[KITTI_CAM[0], 0, KITTI_CAM[2]], [0, KITTI_CAM[1], KITTI_CAM[3]], [0, 0, 1]], np.float64)
    t12 = -(T2[:3, 3].astype(np.float64))
    tx = np.array([[0, -t12[2], t12[1]], [t12[2], 0, -t12[0]], [-t12[1], t12[0], 0]])
    F12 = (np.linalg.inv(K).T @ tx @ np.linalg.inv(K)).astype(np.float32)
    return (L1, R1, L2, R2), V, T1, T2, F12


def distinct_lists(seed=21):
    rng = np.random.default_rng(seed)
    lists = []
    for n in [1, 2, 3, 4, 5, 8, 31, 32, 33, 64, 100] + list(rng.integers(2, 60, 30)):
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        lists.append(base[None, :] ^ np.packbits(rng.random((n, 256)) < rng.uniform(0.02, 0.25), axis=1))
    lists.append(np.repeat(rng.integers(0, 256, (1, 32), dtype=np.uint8), 7, 0))
    return lists


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
    # KeyFrame / MapPoint based searches: the reference-made MapPoint fields, isInFrustum tracking fields and results go to
    # small .npz fixtures (they are inputs the restatements cannot derive without the reference's KeyFrame machinery)
    out["kf"] = []
    cfg = synth.KITTI
    for seed, L, only_stereo, mono, ori in KF_CASES:
        ims, V, T1, T2, F12 = kf_case_inputs(seed, L)
        o = O.ref_slam_kf(*ims, cfg["nfeatures"], KITTI_CAM, cfg["bf"], V, T1, T2, F12, check_ori=ori, only_stereo=only_stereo, mono=mono, **KF_PARAMS)
        o2 = O.ref_slam_kf(*ims, cfg["nfeatures"], KITTI_CAM, cfg["bf"], V, T1, T2, F12, check_ori=ori, only_stereo=only_stereo, mono=mono, fuse_scw=True, **KF_PARAMS)
        name = f"ref_kf_{seed}.npz"
        np.savez_compressed(os.path.join(ROOT, "tests/golden", name), mps1=o["mps1"], track=o["track"], Ow1=o["Ow1"],
                            node1=o["node1"], node2=o["node2"], nodec=o["nodec"], has1=o["has1"], has2=o["has2"],
                            bow_kk=o["bow_kk"][1], bow_kf=o["bow_kf"][1], tri=o["tri"][1], proj=o["proj"][1], proj_last=o["proj_last"][1],
                            mps2=o["mps2"], log_scale=np.float32(o["log_scale"]), kf_bounds=o["kf_bounds"], inv_sigma2=o["inv_sigma2"],
                            sim3_xf=np.concatenate([o["sim3_xf"][k].reshape(-1) for k in ("sR12", "t12", "sR21", "t21")]).astype(np.float32),
                            sim3=o["sim3"][1], fuse=o["fuse"][1], proj_scw=o["proj_scw"][1], proj_kf=o["proj_kf"][1], reloc_Ow=o["reloc_Ow"],
                            scw_xf=np.concatenate([o["scw_xf"][k].reshape(-1) for k in ("Rcw", "tcw", "Ow")]).astype(np.float32), fuse_scw=o2["fuse"][1])
        out["kf"].append(dict(seed=seed, L=L, only_stereo=only_stereo, mono=mono, check_ori=ori, fixture=name, n1=len(o["F1"]["keys"]),
                              n2=len(o["F2"]["keys"]), map_points=len(o["mps1"]), keys1_sha256=sha(o["F1"]["keys"]), u_right2_sha256=sha(o["F2"]["u_right"]),
                              n_bow_kk=o["bow_kk"][0], n_bow_kf=o["bow_kf"][0], n_tri=o["tri"][0], n_proj=o["proj"][0], n_proj_last=o["proj_last"][0], n_sim3=o["sim3"][0], n_fuse=o["fuse"][0], n_proj_scw=o["proj_scw"][0], n_proj_kf=o["proj_kf"][0], n_fuse_scw=o2["fuse"][0]))
    lists = distinct_lists()
    chosen = O.ref_slam_distinct(lists, synth.mono_frame(1, 1241, 376))
    out["distinct"] = dict(seed=21, n=len(lists), chosen_sha256=sha(chosen))
    json.dump(out, open(os.path.join(ROOT, "tests/golden/ref_slam.json"), "w"), indent=1)
    print("wrote", {k: len(v) for k, v in out.items() if isinstance(v, list)})
