"""Shared by the CPU and GPU tests of the KeyFrame / MapPoint based searches: rebuilds the inputs of a tests/golden/ref_kf_*.npz
case (frames by the oracle's extraction + stereo matching, feature-vector nodes by the oracle's DBoW2 transform, MapPoint and
isInFrustum fields from the fixture the REFERENCE wrote) and names the reference's results to compare with."""
import importlib.util
import json
import os

import numpy as np

from orb_slam2_b200 import synth

HERE = os.path.dirname(__file__)
_spec = importlib.util.spec_from_file_location("gen_golden_slam", os.path.join(HERE, "..", "tools", "gen_golden_slam.py"))
gs = importlib.util.module_from_spec(_spec); _spec.loader.exec_module(gs)
GOLD = json.load(open(os.path.join(HERE, "golden", "ref_slam.json")))


def build(oracle, case):
    cfg = synth.KITTI
    nf = cfg["nfeatures"]
    ims, V, T1, T2, F12 = gs.kf_case_inputs(case["seed"], case["L"])
    fx = np.load(os.path.join(HERE, "golden", case["fixture"]))
    frames = []
    for L, R in ((ims[0], ims[1]), (ims[2], ims[3])):
        EL, ER = oracle.Extractor(nf), oracle.Extractor(nf)
        kl, dl = EL.extract(L); kr, dr = ER.extract(R)
        ur, dp, _, _ = oracle.stereo(EL, ER, kl, dl, kr, dr, cfg["bf"], cfg["fx"])
        frames.append(dict(keys=kl, desc=dl, u_right=ur, depth=dp))
    tb = oracle.Extractor(nf).tables()
    node = [oracle.voc_transform(V, f["desc"], 4, 0, 1)[1] for f in frames]
    has = [np.array([f["depth"][i] > 0 and oracle.kf_rule(i, s, p) for i in range(len(f["keys"]))], np.uint8)
           for f, s, p in ((frames[0], 0, gs.KF_PARAMS["pct1"]), (frames[1], 1, gs.KF_PARAMS["pct2"]))]
    return dict(cfg=cfg, F1=frames[0], F2=frames[1], node1=node[0], node2=node[1], has1=has[0], has2=has[1], fx=fx, T2=T2, F12=F12,
                scale=tb["scale"], sigma2=tb["sigma2"], cam=gs.KITTI_CAM)


GOLD_DIST = json.load(open(os.path.join(HERE, "golden", "ref_dist.json")))


def build_dist(extract_stereo, undistort, case):
    """Inputs of a tests/golden/ref_kfd_*.npz case (a DISTORTED stereo camera: Frame::mnMinX / mnMinY < 0).
    extract_stereo(L, R) -> dict(keys, desc, u_right, depth); undistort(cam, keys, w, h) -> (keys_un, bounds6).
    Returns what SearchByProjection(Frame, MapPoints) and SearchByProjection(Current, Last) read."""
    cfg = synth.KITTI
    w, h = cfg["w"], cfg["h"]
    ims, _, _, T2, _ = gs.kf_case_inputs(case["seed"], case["L"])
    fx = np.load(os.path.join(HERE, "golden", case["fixture"]))
    F1, F2 = extract_stereo(ims[0], ims[1]), extract_stereo(ims[2], ims[3])
    for F in (F1, F2):
        F["keys_un"], F["bounds"] = undistort(case["cam"], F["keys"], w, h)
    mp, tr = fx["mps1"], fx["track"]
    mpd = dict(inview=tr["inview"].astype(np.uint8), projx=tr["projx"], projy=tr["projy"], projxr=tr["projxr"], level=tr["level"], viewcos=tr["viewcos"],
               desc=mp["desc"], obs=mp["obs"])
    n1 = len(F1["keys"])
    last = dict(valid=fx["has1"], xw=np.zeros((n1, 3), np.float32), octave=F1["keys"]["octave"], angle=F1["keys_un"]["angle"], desc=np.zeros((n1, 32), np.uint8),
                obs=np.zeros(n1, np.int32))
    last["xw"][mp["idx"]] = mp["xw"]; last["desc"][mp["idx"]] = mp["desc"]; last["obs"][mp["idx"]] = mp["obs"]
    return dict(cfg=cfg, F1=F1, F2=F2, fx=fx, T2=T2, mpd=mpd, last=last, cam=tuple(case["cam"][:4]))
