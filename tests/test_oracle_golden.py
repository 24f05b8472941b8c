"""Oracle vs (a) golden vectors produced by the reference's own ORBextractor.cc
(tests/golden/ref_extract.json, made by tools/gen_golden.py from oracle/_ref/orbref), (b) the
oracle/_ref binary itself when it travelled with the snapshot, and (c) known-answer tests for
the constants SURVEY.md §8c lists."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import missing_reference_binary

from orb_slam2_b200 import synth

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "ref_extract.json")))


@pytest.mark.parametrize("case", GOLD["cases"], ids=lambda c: f"{c['name']}-{c['frame']}")
def test_oracle_matches_reference_golden(oracle, case):
    img = synth.mono_frame(case["frame"], case["w"], case["h"])
    assert hashlib.sha256(img.tobytes()).hexdigest() == case["img_sha256"], "synthetic generator drifted"
    k, d = oracle.Extractor(case["nfeatures"]).extract(img)
    assert len(k) == case["count"]
    assert [int((k["octave"] == l).sum()) for l in range(8)] == case["per_level"]
    assert hashlib.sha256(k.tobytes()).hexdigest() == case["kps_sha256"]
    assert hashlib.sha256(d.tobytes()).hexdigest() == case["desc_sha256"]
    assert d[0].tolist() == case["first_desc"]


def test_oracle_matches_reference_binary(oracle):
    if not oracle.ref_available():
        missing_reference_binary("oracle/_ref/orbref")
    for w, h, nf in [(1241, 376, 2000), (640, 480, 1000)]:
        imgs = np.stack([synth.synth_image(900 + i, w, h) for i in range(2)])
        ref, _ = oracle.ref_extract(imgs, nf)
        E = oracle.Extractor(nf)
        for i in range(2):
            k, d = E.extract(imgs[i])
            assert k.tobytes() == ref[i][0].tobytes()
            assert np.array_equal(d, ref[i][1])


def test_constants_kat(oracle):
    t = oracle.Extractor(2000).tables()
    assert t["umax"].tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert t["quota"].tolist() == [434, 362, 302, 251, 209, 175, 145, 122]
    assert oracle.Extractor(1200).tables()["quota"].tolist() == [261, 217, 181, 151, 126, 105, 87, 72]
    assert oracle.Extractor(1000).tables()["quota"].tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert [int(31 * s) for s in t["scale"]] == [31, 37, 44, 53, 64, 77, 92, 111]
    assert t["scale"][1] == np.float32(1.2) and t["inv"][0] == 1.0


def test_keypoint_fields(oracle):
    k, d = oracle.Extractor(1000).extract(synth.mono_frame(3, 640, 480))
    sizes = {0: 31, 1: 37, 2: 44, 3: 53, 4: 64, 5: 77, 6: 92, 7: 111}
    assert all(k["size"][i] == sizes[int(k["octave"][i])] for i in range(len(k)))
    assert (k["class_id"] == -1).all() and (k["angle"] >= 0).all() and (k["angle"] < 360).all()
    assert np.all(np.diff(k["octave"]) >= 0)                      # level-major output
    assert 1000 <= len(k) <= 1000 + 3 * 8 and d.shape == (len(k), 32)


def test_hamming_kat(oracle):
    z = np.zeros(32, np.uint8); o = np.full(32, 255, np.uint8)
    assert oracle.hamming256(z, z) == 0 and oracle.hamming256(z, o) == 256
    one = z.copy(); one[17] = 0x10
    assert oracle.hamming256(z, one) == 1 and oracle.hamming256(o, one) == 255
    rng = np.random.default_rng(0)
    a, b = rng.integers(0, 256, 32, dtype=np.uint8), rng.integers(0, 256, 32, dtype=np.uint8)
    assert oracle.hamming256(a, b) == int(np.unpackbits(a ^ b).sum())




def test_quadtree_tie_rule_h1(oracle):
    """Two equally sized open nodes, room for only one more split: the node created LAST is the
    one split (reference under a bump allocator: sort by (size, address), walk from the back)."""
    KP = oracle.KP_DTYPE
    pts = [(10, 10, 5), (30, 30, 6),        # child n1 of the root: 2 keys
           (90, 10, 7), (110, 30, 8),       # child n2: 2 keys
           (10, 90, 9)]                     # child n3: singleton
    cand = np.zeros(len(pts), KP)
    for i, (x, y, r) in enumerate(pts):
        cand[i]["x"], cand[i]["y"], cand[i]["response"] = x, y, r
    # root 128x128 -> 3 children (2 open, sizes equal).  N=4: 3 + 3*2 > 4 -> careful phase; the
    # first careful split already reaches 4 nodes, so exactly one of the two open nodes is split.
    out = oracle.quadtree(cand, 0, 128, 0, 128, 4)
    got = sorted((float(k["x"]), float(k["y"])) for k in out)
    # n2 (created after n1) is split: both of its keys survive; n1 keeps only its best (30,30).
    assert got == sorted([(30.0, 30.0), (90.0, 10.0), (110.0, 30.0), (10.0, 90.0)])
    # list order: newest children first
    assert [(float(k["x"]), float(k["y"])) for k in out] == [(110.0, 30.0), (90.0, 10.0), (10.0, 90.0), (30.0, 30.0)]


def test_stereo_and_matchers_behave(oracle):
    cfg = synth.KITTI
    L, R, rowd = synth.stereo_pair(5, cfg["w"], cfg["h"])
    EL, ER = oracle.Extractor(cfg["nfeatures"]), oracle.Extractor(cfg["nfeatures"])
    kl, dl = EL.extract(L); kr, dr = ER.extract(R)
    ur, dp, sad, m = oracle.stereo(EL, ER, kl, dl, kr, dr, cfg["bf"], cfg["fx"])
    ok = ur >= 0
    assert m == ok.sum() and m > 300
    assert np.all(dp[ok] > 0) and np.all(dp[~ok] == -1)
    err = np.abs((kl["x"] - ur)[ok] - rowd[kl["y"].astype(int)][ok])
    assert np.median(err) < 1.0                                   # planted disparity recovered
    a, b, (dx, dy) = synth.mono_pair(1, 640, 480)
    E = oracle.Extractor(1000)
    k1, d1 = E.extract(a); k2, d2 = E.extract(b)
    prev = np.stack([k1["x"], k1["y"]], 1)
    m12, n, prev2 = oracle.search_init(k1, d1, k2, d2, 640, 480, prev)
    assert n == (m12 >= 0).sum() and n > 30
    hit = [abs(k2["x"][j] - k1["x"][i] - dx) < 2 and abs(k2["y"][j] - k1["y"][i] - dy) < 2 for i, j in enumerate(m12) if j >= 0]
    assert np.mean(hit) > 0.8
    assert np.all(k1["octave"][m12 >= 0] == 0)
    mm, nn = oracle.knn_ratio(d1, k1["angle"], d2, k2["angle"])
    assert nn == (mm >= 0).sum() and nn > 100
    used = mm[mm >= 0]
    assert len(set(used.tolist())) == len(used)                   # greedy exclusion: B side unique


# ---- DBoW2 transform (K15): the oracle against the reference's own Thirdparty/DBoW2 ----------------------------
import importlib.util as _ilu

_spec = _ilu.spec_from_file_location("gen_golden_bow", os.path.join(os.path.dirname(__file__), "..", "tools", "gen_golden_bow.py"))
_gb = _ilu.module_from_spec(_spec); _spec.loader.exec_module(_gb)
GOLD_BOW = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "ref_bow.json")))


@pytest.mark.parametrize("case", GOLD_BOW["cases"], ids=lambda c: f"k{c['k']}L{c['L']}w{c['weighting']}s{c['scoring']}")
def test_oracle_bow_transform_matches_reference_golden(oracle, case):
    """tests/golden/ref_bow.json was written by the reference's DBoW2 compiled in place (tools/gen_golden_bow.py)."""
    V = synth.vocabulary(case["seed"], case["k"], case["L"], stop_fraction=0.05)
    assert hashlib.sha256(V["desc"].tobytes() + V["weight"].tobytes()).hexdigest() == case["voc_sha256"], "synthetic vocabulary drifted"
    D = _gb.descriptors(case["seed"] + 100, V)
    assert hashlib.sha256(D.tobytes()).hexdigest() == case["desc_sha256"]
    wo, no, bw, bv = oracle.voc_transform(V, D, case["levelsup"], case["weighting"], case["norm"])
    assert len(bw) == case["bow_size"] and int((no < 0).sum()) == case["stopped"]
    assert bw[:4].tolist() == case["first_words"] and [float(v).hex() for v in bv[:4]] == case["first_values_hex"]
    assert hashlib.sha256(bw.tobytes()).hexdigest() == case["bow_words_sha256"]
    assert hashlib.sha256(bv.tobytes()).hexdigest() == case["bow_values_sha256"]
    assert hashlib.sha256(no.tobytes()).hexdigest() == case["nodes_sha256"]


def test_oracle_bow_transform_matches_reference_binary(oracle):
    if not oracle.ref_dbow_available():
        missing_reference_binary("oracle/_ref/orbref_dbow")
    for seed, k, L, levelsup, scoring, weighting, norm in [(31, 10, 3, 2, 0, 0, 1), (32, 7, 4, 4, 1, 1, 2), (33, 9, 3, 1, 5, 2, 0)]:
        V = synth.vocabulary(seed, k, L, stop_fraction=0.1)
        D = _gb.descriptors(seed, V, 900)
        node, bw, bv = oracle.ref_dbow_transform(V, D, levelsup, scoring, weighting)
        wo, no, obw, obv = oracle.voc_transform(V, D, levelsup, weighting, norm)
        assert np.array_equal(no, node) and np.array_equal(obw, bw)
        assert np.array_equal(obv.view(np.uint64), bv.view(np.uint64))


# ---- Frame.cc / ORBmatcher.cc: the oracle against the reference's own sources compiled in place (oracle/_ref/orbref_slam) ----
_spec2 = _ilu.spec_from_file_location("gen_golden_slam", os.path.join(os.path.dirname(__file__), "..", "tools", "gen_golden_slam.py"))
_gs = _ilu.module_from_spec(_spec2); _spec2.loader.exec_module(_gs)
GOLD_SLAM = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "ref_slam.json")))
_sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("case", GOLD_SLAM["stereo"], ids=lambda c: f"{c['name']}-{c['seed']}")
def test_oracle_stereo_matches_reference_golden(oracle, case):
    """uRight / depth of the reference's stereo Frame constructor (Frame::ComputeStereoMatches, src/Frame.cc:466-640)."""
    cfg = dict(kitti=synth.KITTI, euroc=synth.EUROC)[case["name"]]
    L, R, _ = synth.stereo_pair(case["seed"], cfg["w"], cfg["h"])
    EL, ER = oracle.Extractor(cfg["nfeatures"]), oracle.Extractor(cfg["nfeatures"])
    kl, dl = EL.extract(L); kr, dr = ER.extract(R)
    assert len(kl) == case["n"] and len(kr) == case["n_right"] and _sha(kl) == case["keys_sha256"]
    ur, dp, _, _ = oracle.stereo(EL, ER, kl, dl, kr, dr, cfg["bf"], cfg["fx"])
    assert int((ur >= 0).sum()) == case["matches"]
    assert _sha(ur) == case["u_right_sha256"] and _sha(dp) == case["depth_sha256"]


@pytest.mark.parametrize("case", GOLD_SLAM["init"], ids=lambda c: f"{c['seed']}-w{c['window']}")
def test_oracle_search_for_initialization_matches_reference_golden(oracle, case):
    a, b, _ = synth.mono_pair(case["seed"], 640, 480)
    E = oracle.Extractor(1000)
    k1, d1 = E.extract(a); k2, d2 = E.extract(b)
    assert len(k1) == case["n1"] and len(k2) == case["n2"]
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    m12, n, prev = oracle.search_init(k1, d1, k2, d2, 640, 480, prev, case["ratio"], case["window"], case["check_ori"])
    assert n == case["nmatches"] and _sha(m12) == case["matches12_sha256"] and _sha(prev) == case["prev_sha256"]


@pytest.mark.parametrize("case", GOLD_SLAM["grid"], ids=lambda c: c["name"])
def test_oracle_frame_grid_matches_reference_golden(oracle, case):
    """mvKeysUn, image bounds, mGrid and GetFeaturesInArea of the reference's monocular Frame constructor."""
    w, h = case["w"], case["h"]
    k, _ = oracle.Extractor(case["nfeatures"]).extract(synth.mono_frame(case["frame"], w, h))
    kun, b, cs, items = oracle.undistort_grid(case["cam"], k, w, h)
    assert len(k) == case["n"] and _sha(kun) == case["keys_un_sha256"]
    assert [float(v).hex() for v in b] == case["bounds_hex"]
    cells = [items[cs[c]:cs[c + 1]] for c in range(64 * 48)]
    assert _sha(np.concatenate([np.array([len(c)], np.int32) for c in cells] + cells)) == case["cells_sha256"]
    hits = [oracle.features_in_area(kun, b, *q) for q in _gs.grid_queries(case["frame"], w, h)]
    assert sum(len(x) for x in hits) == case["total_hits"]
    assert _sha(np.concatenate([np.array([len(x)], np.int32) for x in hits] + hits)) == case["hits_sha256"]


def test_oracle_matches_reference_slam_binary(oracle):
    if not oracle.ref_slam_available():
        missing_reference_binary("oracle/_ref/orbref_slam")
    cfg = synth.KITTI
    L, R, _ = synth.stereo_pair(41, cfg["w"], cfg["h"])
    F = oracle.ref_slam_stereo(L, R, cfg["nfeatures"], _gs.KITTI_CAM, cfg["bf"])
    EL, ER = oracle.Extractor(cfg["nfeatures"]), oracle.Extractor(cfg["nfeatures"])
    kl, dl = EL.extract(L); kr, dr = ER.extract(R)
    ur, dp, _, _ = oracle.stereo(EL, ER, kl, dl, kr, dr, cfg["bf"], cfg["fx"])
    assert kl.tobytes() == F["keys"].tobytes() and ur.tobytes() == F["u_right"].tobytes() and dp.tobytes() == F["depth"].tobytes()
    a, b, _ = synth.mono_pair(42, 640, 480)
    F1, F2, n, m12, prev = oracle.ref_slam_init(a, b, 1000, _gs.TUM1[:4], 0.9, 100, True)
    p0 = np.stack([F1["keys_un"]["x"], F1["keys_un"]["y"]], 1).astype(np.float32)
    om, on, oprev = oracle.search_init(F1["keys_un"], F1["desc"], F2["keys_un"], F2["desc"], 640, 480, p0, 0.9, 100, True)
    assert on == n and np.array_equal(om, m12) and oprev.tobytes() == prev.tobytes()


# ---- KeyFrame / MapPoint based searches: the oracle against results the reference's own ORBmatcher.cc produced on real
#      KeyFrames and MapPoints (tests/golden/ref_kf_*.npz written by oracle/_ref/orbref_slam) --------------------------------
_spec3 = _ilu.spec_from_file_location("kf_cases", os.path.join(os.path.dirname(__file__), "kf_cases.py"))
_kc = _ilu.module_from_spec(_spec3); _spec3.loader.exec_module(_kc)


@pytest.mark.parametrize("case", GOLD_SLAM["kf"], ids=lambda c: f"kf{c['seed']}-L{c['L']}")
def test_oracle_keyframe_searches_match_reference(oracle, case):
    c = _kc.build(oracle, case)
    F1, F2, fx = c["F1"], c["F2"], c["fx"]
    w, h = c["cfg"]["w"], c["cfg"]["h"]
    ori, only_stereo = case["check_ori"], case["only_stereo"]
    assert _sha(F1["keys"]) == case["keys1_sha256"] and _sha(F2["u_right"]) == case["u_right2_sha256"]
    # what the restatements derive themselves agrees with what the reference's KeyFrames held
    assert np.array_equal(c["node1"], fx["node1"]) and np.array_equal(c["node2"], fx["node2"]) and np.array_equal(c["node2"], fx["nodec"])
    assert np.array_equal(c["has1"], fx["has1"]) and np.array_equal(c["has2"], fx["has2"])
    P = _kc.gs.KF_PARAMS
    # SearchByBoW(KF, KF)  src/ORBmatcher.cc:522
    m, n = oracle.search_by_bow(F1["desc"], F1["keys"]["angle"], c["has1"], c["node1"], F2["desc"], F2["keys"]["angle"], c["has2"], c["node2"],
                                50, P["ratio_kk"], ori, True)
    assert n == case["n_bow_kk"] and np.array_equal(m, fx["bow_kk"])
    if case["L"] <= 4:                                   # one bucket: the brute-force form (K7) gives the same as the bucketed one
        va, vb = c["has1"].astype(bool), c["has2"].astype(bool)
        mk, nk = oracle.knn_ratio(F1["desc"][va], F1["keys"]["angle"][va], F2["desc"][vb], F2["keys"]["angle"][vb], 50, P["ratio_kk"], ori, True)
        full = np.full(len(va), -1, np.int32); full[np.nonzero(va)[0][mk >= 0]] = np.nonzero(vb)[0][mk[mk >= 0]]
        assert nk == n and np.array_equal(full, m)
    # SearchByBoW(KF, Frame)  :159  (the reference indexes its result by frame feature)
    m, n = oracle.search_by_bow(F1["desc"], F1["keys"]["angle"], c["has1"], c["node1"], F2["desc"], F2["keys"]["angle"], np.ones(len(F2["desc"]), np.uint8),
                                c["node2"], 50, P["ratio_kf"], ori, False)
    inv = np.full(len(F2["desc"]), -1, np.int32); ok = m >= 0; inv[m[ok]] = np.nonzero(ok)[0]
    assert n == case["n_bow_kf"] and np.array_equal(inv, fx["bow_kf"])
    # SearchForTriangulation  :657
    T2 = c["T2"]
    m, n = oracle.search_for_triangulation(F1["keys"], F1["desc"], c["has1"], F1["u_right"], c["node1"], F2["keys"], F2["desc"], c["has2"], F2["u_right"],
                                           c["node2"], c["F12"], np.concatenate([T2[:3, :3].reshape(9), T2[:3, 3]]), fx["Ow1"], np.array(c["cam"], np.float32),
                                           c["scale"], c["sigma2"], only_stereo, ori)
    assert n == case["n_tri"] and np.array_equal(np.stack([np.nonzero(m >= 0)[0], m[m >= 0]], 1).astype(np.int32), fx["tri"])
    # SearchByProjection(Frame, MapPoints)  :45, fed with the tracking fields the reference's Frame::isInFrustum produced
    mp, tr = fx["mps1"], fx["track"]
    mpd = dict(inview=tr["inview"].astype(np.uint8), projx=tr["projx"], projy=tr["projy"], projxr=tr["projxr"], level=tr["level"], viewcos=tr["viewcos"],
               desc=mp["desc"], obs=mp["obs"])
    a, n, _ = oracle.search_projection(F2["keys"], F2["desc"], F2["u_right"], w, h, c["scale"], mpd, P["th_proj"], 0.8)
    assert n == case["n_proj"] and np.array_equal(a, fx["proj"])
    # SearchByProjection(CurrentFrame, LastFrame)  :1328
    n1 = len(F1["keys"])
    last = dict(valid=c["has1"], xw=np.zeros((n1, 3), np.float32), octave=F1["keys"]["octave"], angle=F1["keys"]["angle"], desc=np.zeros((n1, 32), np.uint8),
                obs=np.zeros(n1, np.int32))
    last["xw"][mp["idx"]] = mp["xw"]; last["desc"][mp["idx"]] = mp["desc"]; last["obs"][mp["idx"]] = mp["obs"]
    a, n, _ = oracle.search_projection_frame(F2["keys"], F2["desc"], F2["u_right"], w, h, c["scale"], T2[:3, :3], T2[:3, 3], c["cam"] + (c["cfg"]["bf"],), last,
                                             P["th_last"], 0, ori)
    assert n == case["n_proj_last"] and np.array_equal(a, fx["proj_last"])
    assert min(case["n_bow_kk"], case["n_bow_kf"], case["n_tri"], case["n_proj"], case["n_proj_last"]) > 20


def test_oracle_distinctive_descriptors_match_reference(oracle):
    """MapPoint::ComputeDistinctiveDescriptors of the reference, run on real MapPoints observed by up to 100 real KeyFrames."""
    lists = _kc.gs.distinct_lists(GOLD_SLAM["distinct"]["seed"])
    chosen = np.stack([d[oracle.distinctive_descriptor(d)] for d in lists])
    assert _sha(chosen) == GOLD_SLAM["distinct"]["chosen_sha256"]
    if oracle.ref_slam_available():
        assert np.array_equal(oracle.ref_slam_distinct(lists, synth.mono_frame(1, 1241, 376)), chosen)


@pytest.mark.parametrize("case", GOLD_SLAM["kf"], ids=lambda c: f"kf{c['seed']}-L{c['L']}")
def test_oracle_fuse_and_sim3_match_reference(oracle, case):
    """ORBmatcher::Fuse (search) and SearchBySim3 of the reference on the same real KeyFrames / MapPoints."""
    c = _kc.build(oracle, case)
    F1, F2, fx = c["F1"], c["F2"], c["fx"]
    cam, bf, T2 = c["cam"], c["cfg"]["bf"], c["T2"]
    kfv = (cam[0], cam[1], cam[2], cam[3], bf, float(fx["log_scale"]), *fx["kf_bounds"])
    mp1, mp2 = fx["mps1"], fx["mps2"]
    R2, t2 = T2[:3, :3], T2[:3, 3]
    mpd = dict(valid=np.ones(len(mp1), np.uint8), xw=mp1["xw"], normal=mp1["normal"], dmin=mp1["dmin"], dmax=mp1["dmax"], desc=mp1["desc"])
    best = oracle.project_search(F2["keys"], F2["desc"], F2["u_right"], kfv, c["scale"], fx["inv_sigma2"], dict(R1=R2, t1=t2, Ow=(-R2.T @ t2).astype(np.float32)),
                                 mpd, 3.0, oracle.FUSE_MODE, 50)
    assert int((best >= 0).sum()) == case["n_fuse"] and np.array_equal(best, fx["fuse"])
    x = fx["sim3_xf"]
    sR12, t12, sR21, t21 = x[:9].reshape(3, 3), x[9:12], x[12:21].reshape(3, 3), x[21:24]
    one = lambda mp: dict(valid=np.ones(len(mp), np.uint8), xw=mp["xw"], normal=mp["normal"], dmin=mp["dmin"], dmax=mp["dmax"], desc=mp["desc"])
    b1 = oracle.project_search(F2["keys"], F2["desc"], F2["u_right"], kfv, c["scale"], fx["inv_sigma2"], dict(R1=np.eye(3), t1=np.zeros(3), R2=sR21, t2=t21), one(mp1), 7.5,
                               oracle.SIM3_MODE, 100)
    b2 = oracle.project_search(F1["keys"], F1["desc"], F1["u_right"], kfv, c["scale"], fx["inv_sigma2"], dict(R1=R2, t1=t2, R2=sR12, t2=t12), one(mp2), 7.5,
                               oracle.SIM3_MODE, 100)
    n1, n2 = len(F1["keys"]), len(F2["keys"])
    vn1 = np.full(n1, -1, np.int32); vn1[mp1["idx"]] = b1
    vn2 = np.full(n2, -1, np.int32); vn2[mp2["idx"]] = b2
    m12 = np.array([vn1[i] if vn1[i] >= 0 and vn2[vn1[i]] == i else -1 for i in range(n1)], np.int32)
    assert int((m12 >= 0).sum()) == case["n_sim3"] and np.array_equal(m12, fx["sim3"])
    assert case["n_fuse"] > 50 and case["n_sim3"] > 50


def test_logf_restatement(oracle):
    """glibc logf restated (MapPoint::PredictScale calls the float overload of log): sampled against the host libm here,
    exhaustively by tools/sweep_logf.c."""
    import ctypes, math
    L = oracle.lib(); L.orbo_logf.restype = ctypes.c_float; L.orbo_logf.argtypes = [ctypes.c_float]
    libm = ctypes.CDLL("libm.so.6"); libm.logf.restype = ctypes.c_float; libm.logf.argtypes = [ctypes.c_float]
    rng = np.random.default_rng(3)
    xs = np.concatenate([rng.uniform(0.01, 100, 20000), np.exp(rng.uniform(-80, 80, 20000)), [1.0, 1.2, 0.5, 2.0]]).astype(np.float32)
    for x in xs:
        assert np.float32(L.orbo_logf(float(x))).tobytes() == np.float32(libm.logf(float(x))).tobytes(), x


@pytest.mark.parametrize("case", GOLD_SLAM["kf"], ids=lambda c: f"kf{c['seed']}-L{c['L']}")
def test_oracle_remaining_matcher_members_match_reference(oracle, case):
    """SearchByProjection(pKF, Scw, ...) :290, SearchByProjection(CurrentFrame, pKF, sAlreadyFound, ...) :1472 and
    Fuse(pKF, Scw, ...) :977 of the reference on the same real KeyFrames / MapPoints."""
    c = _kc.build(oracle, case)
    F1, F2, fx = c["F1"], c["F2"], c["fx"]
    cam, bf, T2 = c["cam"], c["cfg"]["bf"], c["T2"]
    w, h = c["cfg"]["w"], c["cfg"]["h"]
    mp1 = fx["mps1"]; n2 = len(F2["keys"])
    fb = (0.0, float(w), 0.0, float(h))
    kfv = (cam[0], cam[1], cam[2], cam[3], bf, float(fx["log_scale"]), *fx["kf_bounds"])
    mpd = dict(valid=np.ones(len(mp1), np.uint8), xw=mp1["xw"], normal=mp1["normal"], dmin=mp1["dmin"], dmax=mp1["dmax"], desc=mp1["desc"],
               angle=F1["keys"]["angle"][mp1["idx"]])
    x = fx["scw_xf"]; Rcw, tcw, Ow = x[:9].reshape(3, 3), x[9:12], x[12:15]
    occ = np.array([c["has2"][i] and oracle.kf_rule(i, 7, 30) for i in range(n2)], np.uint8)
    a, n = oracle.project_match_ordered(F2["keys"], F2["desc"], kfv, fb, c["scale"], False, Rcw, tcw, Ow, mpd, 10.0, 50, False, occ)
    ref = fx["proj_scw"]
    assert n == case["n_proj_scw"] and np.array_equal(a, np.where(ref >= 0, ref, -1)) and np.array_equal(ref == -2, occ.astype(bool))
    occ = np.array([c["has2"][i] and oracle.kf_rule(i, 11, 20) for i in range(n2)], np.uint8)
    mpd2 = dict(mpd); mpd2["valid"] = np.array([not oracle.kf_rule(k, 9, 25) for k in range(len(mp1))], np.uint8)
    kfv2 = (cam[0], cam[1], cam[2], cam[3], bf, float(fx["log_scale"]), *fb)
    a, n = oracle.project_match_ordered(F2["keys"], F2["desc"], kfv2, fb, c["scale"], True, T2[:3, :3], T2[:3, 3], fx["reloc_Ow"], mpd2, 10.0, 100,
                                        case["check_ori"], occ)
    ref = fx["proj_kf"]
    assert n == case["n_proj_kf"] and np.array_equal(a, np.where(ref >= 0, ref, -1))
    best = oracle.project_search(F2["keys"], F2["desc"], F2["u_right"], kfv, c["scale"], fx["inv_sigma2"], dict(R1=Rcw, t1=tcw, Ow=Ow),
                                 dict(valid=mpd["valid"], xw=mp1["xw"], normal=mp1["normal"], dmin=mp1["dmin"], dmax=mp1["dmax"], desc=mp1["desc"]), 4.0, 2 | 4 | 16, 50)
    assert int((best >= 0).sum()) == case["n_fuse_scw"] and np.array_equal(best, fx["fuse_scw"])
    assert min(case["n_proj_scw"], case["n_proj_kf"], case["n_fuse_scw"]) > 50


# ---- distorted cameras: Frame::mnMinX / mnMinY != 0 (tests/golden/ref_dist.json, written by the reference compiled in place) ----
@pytest.mark.parametrize("case", _kc.GOLD_DIST["init"], ids=lambda c: f"init{c['seed']}")
def test_oracle_search_init_distorted_camera(oracle, case):
    """SearchForInitialization on TUM1 WITH its distortion: the grid of frame 2 is laid over the undistorted image bounds."""
    a, b, _ = synth.mono_pair(case["seed"], 640, 480)
    E = oracle.Extractor(1000)
    (k1, d1), (k2, d2) = E.extract(a), E.extract(b)
    ku1, bnd, _, _ = oracle.undistort_grid(case["cam"], k1, 640, 480)
    ku2 = oracle.undistort_grid(case["cam"], k2, 640, 480)[0]
    assert bnd[0] != 0.0 and bnd[2] != 0.0 and _sha(ku2) == case["keys_un2_sha256"]
    prev = np.stack([ku1["x"], ku1["y"]], 1).astype(np.float32)
    m12, n, prev = oracle.search_init(ku1, d1, ku2, d2, 640, 480, prev, case["ratio"], case["window"], case["check_ori"], bounds=bnd)
    assert n == case["nmatches"] and _sha(m12) == case["matches12_sha256"] and _sha(prev) == case["prev_sha256"]
    assert n > 50


@pytest.mark.parametrize("case", _kc.GOLD_DIST["kf"], ids=lambda c: f"kfd{c['seed']}")
def test_oracle_projection_searches_distorted_camera(oracle, case):
    """SearchByProjection(Frame, MapPoints) :45 and SearchByProjection(Current, Last) :1328 of the reference on a distorted
    stereo camera (mnMinX = -64, mnMinY = -22: 35 keypoints have negative undistorted x)."""
    cfg = synth.KITTI

    def extract_stereo(L, R):
        EL, ER = oracle.Extractor(cfg["nfeatures"]), oracle.Extractor(cfg["nfeatures"])
        kl, dl = EL.extract(L); kr, dr = ER.extract(R)
        ur, dp, _, _ = oracle.stereo(EL, ER, kl, dl, kr, dr, cfg["bf"], cfg["fx"])
        return dict(keys=kl, desc=dl, u_right=ur, depth=dp)

    c = _kc.build_dist(extract_stereo, lambda cam, k, w, h: oracle.undistort_grid(cam, k, w, h)[:2], case)
    F1, F2, fx, T2 = c["F1"], c["F2"], c["fx"], c["T2"]
    w, h = cfg["w"], cfg["h"]
    assert _sha(F1["keys_un"]) == case["keys_un1_sha256"] and _sha(F2["keys_un"]) == case["keys_un2_sha256"] and _sha(F2["u_right"]) == case["u_right2_sha256"]
    b = F2["bounds"]
    assert [float(int(v)) for v in b[:4]] == fx["kf_bounds"].tolist() and int((F2["keys_un"]["x"] < 0).sum()) == case["negative_x"] > 0
    scale = oracle.Extractor(cfg["nfeatures"]).tables()["scale"]
    P = _kc.gs.KF_PARAMS
    a, n, _ = oracle.search_projection(F2["keys_un"], F2["desc"], F2["u_right"], w, h, scale, c["mpd"], P["th_proj"], 0.8, bounds=b)
    assert n == case["n_proj"] and np.array_equal(a, fx["proj"])
    a, n, _ = oracle.search_projection_frame(F2["keys_un"], F2["desc"], F2["u_right"], w, h, scale, T2[:3, :3], T2[:3, 3], c["cam"] + (cfg["bf"],), c["last"],
                                             P["th_last"], 0, case["check_ori"], bounds=b)
    assert n == case["n_proj_last"] and np.array_equal(a, fx["proj_last"])
