"""GPU parity tests: the CUDA path, called through the C ABI (orb_slam2_b200.api -> liborbx.so),
against the CPU oracle on the same seeded synthetic inputs.  Integer, byte and index results must
be bit-exact; angles are bit-exact (fmad-off build + restated libm); stereo sub-pixel uRight is
checked both bit-for-bit and against the 1e-3 px tolerance BASELINE.json states.
Sizes are the three BASELINE shapes (KITTI 1241x376/2000, EuRoC 752x480/1200, TUM 640x480/1000)."""
import hashlib
import json
import os

import numpy as np
import pytest

from orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu

CONFIGS = [("kitti", 1241, 376, 2000), ("euroc", 752, 480, 1200), ("tum", 640, 480, 1000)]
UR_TOL = 1e-3   # px, BASELINE.json north_star


@pytest.fixture(scope="module")
def gx():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import orb_slam2_b200 as pkg
    return pkg


def _ext(gx, w, h, nf, batch=1):
    return gx.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=batch)


@pytest.mark.parametrize("name,w,h,nf", CONFIGS)
def test_pyramid_bit_exact(gx, oracle, name, w, h, nf):
    img = synth.mono_frame(0, w, h)
    E = oracle.Extractor(nf); E.extract(img)
    G = _ext(gx, w, h, nf); G(img)
    for l in range(8):
        assert np.array_equal(G.level(0, l), E.level(l)), f"{name} level {l}"


@pytest.mark.parametrize("name,w,h,nf", CONFIGS)
def test_fast_candidates_bit_exact_and_ordered(gx, oracle, name, w, h, nf):
    img = synth.mono_frame(1, w, h)
    E = oracle.Extractor(nf); E.extract(img)
    G = _ext(gx, w, h, nf); G(img)
    fallback = 0
    for l in range(8):
        c = E.candidates(l)
        ref = np.stack([c["x"], c["y"], c["response"]], 1).astype(np.int32)
        got = G.debug_candidates(0, l)
        assert got.shape == ref.shape and np.array_equal(got, ref), f"{name} level {l}: {len(got)} vs {len(ref)}"
        fallback += int((ref[:, 2] < 20).sum())
    assert fallback > 0


@pytest.mark.parametrize("name,w,h,nf", CONFIGS)
def test_quadtree_keys_bit_exact(gx, oracle, name, w, h, nf):
    img = synth.mono_frame(2, w, h)
    E = oracle.Extractor(nf); E.extract(img)
    G = _ext(gx, w, h, nf); G(img)
    for l in range(8):
        k = E.level_keys(l)
        ref = np.stack([k["x"], k["y"], k["response"]], 1).astype(np.int32)
        got = G.debug_level_keys(0, l)
        assert got.shape == ref.shape and np.array_equal(got, ref), f"{name} level {l}: {len(got)} vs {len(ref)}"


@pytest.mark.parametrize("name,w,h,nf", CONFIGS + [("tum_ini2x", 640, 480, 2000)])
def test_extract_keypoints_and_descriptors_bit_exact(gx, oracle, name, w, h, nf):
    G = _ext(gx, w, h, nf)
    E = oracle.Extractor(nf)
    for f in range(3):
        img = synth.mono_frame(10 + f, w, h)
        rk, rd = E.extract(img)
        gk, gd = G(img)
        assert len(gk) == len(rk)
        for field in ("octave", "x", "y", "size", "response", "class_id"):
            assert np.array_equal(gk[field], rk[field]), (name, f, field)
        assert np.array_equal(gk["angle"].view(np.uint32), rk["angle"].view(np.uint32)), (name, f, "angle bits")
        assert gk.tobytes() == rk.tobytes()
        assert np.array_equal(gd, rd), (name, f, "descriptors", int((gd != rd).any(1).sum()))


def test_reference_golden_digests_through_gpu(gx):
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "ref_extract.json")))
    cache = {}
    for case in gold["cases"]:
        key = (case["w"], case["h"], case["nfeatures"])
        if key not in cache:
            cache[key] = _ext(gx, *key)
        img = synth.mono_frame(case["frame"], case["w"], case["h"])
        k, d = cache[key](img)
        assert len(k) == case["count"]
        assert hashlib.sha256(k.tobytes()).hexdigest() == case["kps_sha256"], case["name"]
        assert hashlib.sha256(d.tobytes()).hexdigest() == case["desc_sha256"], case["name"]


def test_batch_equals_single_and_is_deterministic(gx, oracle):
    w, h, nf, n = 752, 480, 1200, 6
    imgs = np.stack([synth.mono_frame(20 + i, w, h) for i in range(n)])
    G = _ext(gx, w, h, nf, batch=n)
    G.extract_batch(imgs)
    first = G.fetch(0, n)
    G.extract_batch(imgs)
    again = G.fetch(0, n)
    E = oracle.Extractor(nf)
    for i in range(n):
        rk, rd = E.extract(imgs[i])
        assert first[i][0].tobytes() == rk.tobytes() and np.array_equal(first[i][1], rd), i
        assert again[i][0].tobytes() == rk.tobytes() and np.array_equal(again[i][1], rd), i
    # a device-resident batch gives the same bytes as the host-staged one
    import torch
    dev = torch.from_numpy(imgs).cuda()
    G.extract_batch(dev)
    third = G.fetch(0, n)
    for i in range(n):
        assert third[i][0].tobytes() == first[i][0].tobytes() and np.array_equal(third[i][1], first[i][1])


def test_strided_input_and_getters(gx, oracle):
    w, h, nf = 640, 480, 1000
    big = np.zeros((h, w + 37), np.uint8)
    big[:, :w] = synth.mono_frame(30, w, h)
    view = big[:, :w]                                   # row stride != width
    G = _ext(gx, w, h, nf)
    k, d = G(view)
    rk, rd = oracle.Extractor(nf).extract(np.ascontiguousarray(view))
    assert k.tobytes() == rk.tobytes() and np.array_equal(d, rd)
    t = oracle.Extractor(nf).tables()
    assert np.array_equal(G.GetScaleFactors(), t["scale"]) and np.array_equal(G.GetInverseScaleFactors(), t["inv"])
    assert np.array_equal(G.GetScaleSigmaSquares(), t["sigma2"]) and np.array_equal(G.GetInverseScaleSigmaSquares(), t["invsigma2"])
    assert np.array_equal(G.features_per_level(), t["quota"]) and G.GetLevels() == 8
    pyr = G.mvImagePyramid
    assert pyr[0].shape == (h, w) and np.array_equal(pyr[0], view)


@pytest.mark.parametrize("w,h,nf", [(366, 270, 600), (400, 300, 800), (1024, 768, 1500), (331, 250, 500)])
def test_odd_shapes_exercise_every_fast_variant(gx, oracle, w, h, nf):
    """Cell widths of 33-36 and 37+ px take the wide FAST kernels (366x270 and 400x300 have such
    cells on their top levels); 1024x768 is a 4:3 shape with one quad-tree root."""
    img = synth.mono_frame(33, w, h)
    G = _ext(gx, w, h, nf)
    E = oracle.Extractor(nf)
    rk, rd = E.extract(img)
    gk, gd = G(img)
    for l in range(8):
        c = E.candidates(l)
        ref = np.stack([c["x"], c["y"], c["response"]], 1).astype(np.int32)
        assert np.array_equal(G.debug_candidates(0, l), ref), f"level {l}"
    assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)


@pytest.mark.parametrize("nf,sf,nl,ini,mn", [(500, 1.2, 4, 20, 7), (3000, 1.2, 8, 30, 10), (800, 1.5, 5, 20, 7), (1000, 1.1, 8, 15, 5), (1, 1.2, 8, 20, 7)])
def test_other_extractor_parameters(gx, oracle, nf, sf, nl, ini, mn):
    """Other constructor arguments than the three dataset configs: fewer levels, a scale factor
    whose resize footprint does not fit the tiled kernel (1.5 -> generic resize kernel), other FAST
    thresholds, a degenerate one-feature budget."""
    w, h = 752, 480
    img = synth.mono_frame(35, w, h)
    G = gx.ORBextractor(nf, sf, nl, ini, mn, max_w=w, max_h=h)
    E = oracle.Extractor(nf, sf, nl, ini, mn)
    rk, rd = E.extract(img)
    gk, gd = G(img)
    for l in range(nl):
        assert np.array_equal(G.level(0, l), E.level(l)), f"pyramid level {l}"
    assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
    assert np.array_equal(G.features_per_level(), E.tables()["quota"][:nl])


def test_exact_2x_scale_is_rejected(gx):
    with pytest.raises(gx.OrbxError) as ei:              # cv::resize would switch to INTER_AREA
        gx.ORBextractor(1000, 2.0, 3, 20, 7, max_w=640, max_h=480)
    assert ei.value.status == 2 and "INTER_AREA" in str(ei.value)


def test_edge_cases(gx):
    G = _ext(gx, 640, 480, 1000)
    k, d = G(np.zeros((0, 0), np.uint8))                 # empty image: silent return
    assert len(k) == 0 and d.shape == (0, 32)
    k, d = G(np.full((480, 640), 77, np.uint8))          # no corner anywhere
    assert len(k) == 0
    with pytest.raises(gx.OrbxError) as ei:              # level 7 would be < 62 px wide
        gx.ORBextractor(500, 1.2, 8, 20, 7, max_w=200, max_h=160)
    assert ei.value.status == 2
    with pytest.raises(gx.OrbxError) as ei:              # larger than reserved
        G(np.zeros((600, 800), np.uint8))
    assert ei.value.status == 3
    # a smaller image on the same handle re-derives the level geometry
    img = synth.mono_frame(31, 400, 300)
    k, d = G(img)
    from oracle import pyoracle
    rk, rd = pyoracle.Extractor(1000).extract(img)
    assert k.tobytes() == rk.tobytes() and np.array_equal(d, rd)


@pytest.mark.parametrize("name,cfg", [("kitti", synth.KITTI), ("euroc", synth.EUROC)])
def test_stereo_matches(gx, oracle, name, cfg):
    w, h, nf, bf, fx = cfg["w"], cfg["h"], cfg["nfeatures"], cfg["bf"], cfg["fx"]
    n = 3
    pairs = [synth.stereo_pair(40 + i, w, h) for i in range(n)]
    G = _ext(gx, w, h, nf, batch=2 * n)
    G.extract_batch(np.stack([p[0] for p in pairs] + [p[1] for p in pairs]))
    ur, dp, nm = gx.ComputeStereoMatches(G, G, bf, fx, n=n, slotL0=0, slotR0=n)
    res = G.fetch(0, 2 * n)
    for i in range(n):
        EL, ER = oracle.Extractor(nf), oracle.Extractor(nf)
        kl, dl = EL.extract(pairs[i][0]); kr, dr = ER.extract(pairs[i][1])
        assert res[i][0].tobytes() == kl.tobytes() and res[n + i][0].tobytes() == kr.tobytes()
        rur, rdp, rsad, rm = oracle.stereo(EL, ER, kl, dl, kr, dr, bf, fx)
        cnt = len(kl)
        assert int(nm[i]) == rm, (name, i)
        assert np.array_equal(ur[i, :cnt] >= 0, rur >= 0)                       # same match set
        assert np.max(np.abs(ur[i, :cnt] - rur)) <= UR_TOL                      # stated tolerance
        assert np.array_equal(ur[i, :cnt].view(np.uint32), rur.view(np.uint32))  # and in fact bit-exact
        assert np.array_equal(dp[i, :cnt].view(np.uint32), rdp.view(np.uint32))
        assert rm > 100


def test_stereo_two_handles_like_the_reference(gx, oracle):
    cfg = synth.KITTI
    w, h, nf = cfg["w"], cfg["h"], cfg["nfeatures"]
    L, R, _ = synth.stereo_pair(50, w, h)
    GL, GR = _ext(gx, w, h, nf), _ext(gx, w, h, nf)       # mpORBextractorLeft / Right
    kl, dl = GL(L); kr, dr = GR(R)
    ur, dp, nm = gx.ComputeStereoMatches(GL, GR, cfg["bf"], cfg["fx"])
    EL, ER = oracle.Extractor(nf), oracle.Extractor(nf)
    okl, odl = EL.extract(L); okr, odr = ER.extract(R)
    rur, rdp, _, rm = oracle.stereo(EL, ER, okl, odl, okr, odr, cfg["bf"], cfg["fx"])
    assert int(nm[0]) == rm and np.array_equal(ur[0, :len(kl)].view(np.uint32), rur.view(np.uint32))
    flat = np.full((h, w), 90, np.uint8)                  # nothing to match: all -1, zero matches
    GL(flat); GR(flat)
    ur, dp, nm = gx.ComputeStereoMatches(GL, GR, cfg["bf"], cfg["fx"])
    assert int(nm[0]) == 0


def test_stereo_handles_with_different_feature_budgets(gx, oracle):
    """Left and right extractor with different nFeatures (different keypoint capacity per slot): the sorted
    right-keypoint records of K5 live in the right handle's scratch and are indexed by its capacity."""
    cfg = synth.EUROC
    w, h = cfg["w"], cfg["h"]
    L, R, _ = synth.stereo_pair(52, w, h)
    for nl, nr in [(1200, 700), (600, 1500)]:
        GL, GR = _ext(gx, w, h, nl), _ext(gx, w, h, nr)
        kl, dl = GL(L); kr, dr = GR(R)
        ur, dp, nm = gx.ComputeStereoMatches(GL, GR, cfg["bf"], cfg["fx"])
        EL, ER = oracle.Extractor(nl), oracle.Extractor(nr)
        okl, odl = EL.extract(L); okr, odr = ER.extract(R)
        assert kl.tobytes() == okl.tobytes() and kr.tobytes() == okr.tobytes()
        rur, rdp, _, rm = oracle.stereo(EL, ER, okl, odl, okr, odr, cfg["bf"], cfg["fx"])
        assert int(nm[0]) == rm and rm > 50
        assert np.array_equal(ur[0, :len(kl)].view(np.uint32), rur.view(np.uint32))
        assert np.array_equal(dp[0, :len(kl)].view(np.uint32), rdp.view(np.uint32))


def test_stereo_saturated_and_offset_right_images(gx, oracle):
    """The SAD of K5 is centre-subtracted, so right pixel + (cl - cr) spans 9 bits.  A right image that is 40 grey
    levels darker (or 35 brighter) with 1 % salt (255) and pepper (0) pixels drives those sums to their extremes;
    uRight / depth must still equal the oracle bit for bit."""
    cfg = synth.EUROC
    w, h, nf, bf, fx = cfg["w"], cfg["h"], cfg["nfeatures"], cfg["bf"], cfg["fx"]
    rng = np.random.default_rng(5)
    pairs = []
    for seed, off in [(41, 40), (42, -35)]:
        L, R, _ = synth.stereo_pair(seed, w, h)
        R2 = np.clip(R.astype(np.int32) - off, 0, 255).astype(np.uint8)
        m = rng.random(R2.shape)
        R2[m < 0.005] = 255
        R2[m > 0.995] = 0
        pairs.append((L, R2))
    n = len(pairs)
    G = _ext(gx, w, h, nf, batch=2 * n)
    G.extract_batch(np.stack([p[0] for p in pairs] + [p[1] for p in pairs]))
    ur, dp, nm = gx.ComputeStereoMatches(G, G, bf, fx, n=n, slotL0=0, slotR0=n)
    for i in range(n):
        EL, ER = oracle.Extractor(nf), oracle.Extractor(nf)
        kl, dl = EL.extract(pairs[i][0]); kr, dr = ER.extract(pairs[i][1])
        rur, rdp, _, rm = oracle.stereo(EL, ER, kl, dl, kr, dr, bf, fx)
        cnt = len(kl)
        assert int(nm[i]) == rm and rm > 100
        assert np.array_equal(ur[i, :cnt].view(np.uint32), rur.view(np.uint32))
        assert np.array_equal(dp[i, :cnt].view(np.uint32), rdp.view(np.uint32))


FORMS = pytest.mark.parametrize("form", ["latency", "throughput"])


def _pick_form(monkeypatch, form):
    """The order-dependent matchers have two device forms: the latency form (a 1024-thread CTA per problem, a few problems per
    call or many) and the small form (a warp or a 128-thread CTA per problem, taken
    when a problem's arrays do not fit the 1024-thread form's shared memory).  ORBX_PROJ_SERIAL / ORBX_BOW_SERIAL select it."""
    if form == "throughput":
        monkeypatch.setenv("ORBX_PROJ_SERIAL", "1"); monkeypatch.setenv("ORBX_BOW_SERIAL", "1")
    else:
        monkeypatch.delenv("ORBX_PROJ_SERIAL", raising=False); monkeypatch.delenv("ORBX_BOW_SERIAL", raising=False)


@FORMS
@pytest.mark.parametrize("nf", [1000, 2000])
def test_search_for_initialization(gx, oracle, nf, monkeypatch, form):
    _pick_form(monkeypatch, form)
    w, h = 640, 480
    G = _ext(gx, w, h, nf)
    F1s, F2s, prevs, refs = [], [], [], []
    for p in range(4):
        a, b, _ = synth.mono_pair(60 + p, w, h)
        k1, d1 = G(a); k2, d2 = G(b)
        prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
        refs.append(oracle.search_init(k1, d1, k2, d2, w, h, prev, 0.9, 100, True))
        F1s.append(gx.Frame(k1, d1, w, h)); F2s.append(gx.Frame(k2, d2, w, h)); prevs.append(prev.copy())
    nm, m12 = gx.ORBmatcher(0.9, True).SearchForInitializationBatch(F1s, F2s, prevs, 100)
    for p in range(4):
        rm12, rn, rprev = refs[p]
        assert nm[p] == rn and np.array_equal(m12[p], rm12), p
        assert np.array_equal(prevs[p], rprev), p
        assert rn > 20
    # no orientation check, tight window: different branch mix
    k1, d1 = F1s[0].mvKeysUn, F1s[0].mDescriptors
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    rm12, rn, _ = oracle.search_init(k1, d1, F2s[0].mvKeysUn, F2s[0].mDescriptors, w, h, prev, 0.9, 40, False)
    n1, g12 = gx.ORBmatcher(0.9, False).SearchForInitialization(F1s[0], F2s[0], prev.copy(), 40)
    assert n1 == rn and np.array_equal(g12, rm12)


def test_knn_ratio_bruteforce(gx, oracle):
    w, h, nf = 1241, 376, 2000
    G = _ext(gx, w, h, nf)
    As, Bs, refs = [], [], []
    for p in range(3):
        a, b, _ = synth.mono_pair(70 + p, w, h)
        k1, d1 = G(a); k2, d2 = G(b)
        As.append((d1, k1["angle"])); Bs.append((d2, k2["angle"]))
        refs.append(oracle.knn_ratio(d1, k1["angle"], d2, k2["angle"], 50, 0.75, True, True))
    nm, m12 = gx.ORBmatcher(0.75, True).SearchByBoWBatch(As, Bs, strict=True)
    for p in range(3):
        assert nm[p] == refs[p][1] and np.array_equal(m12[p], refs[p][0]), p
        assert refs[p][1] > 100
    r = oracle.knn_ratio(As[0][0], As[0][1], Bs[0][0], Bs[0][1], 50, 0.7, False, False)
    n1, g = gx.ORBmatcher(0.7, False).SearchByBoW(As[0][0], As[0][1], Bs[0][0], Bs[0][1], strict=False)
    assert n1 == r[1] and np.array_equal(g, r[0])
    # ragged: empty B side, single-feature A side
    n1, g = gx.ORBmatcher(0.75, True).SearchByBoW(As[0][0][:1], As[0][1][:1], Bs[0][0][:0], Bs[0][1][:0])
    assert n1 == 0 and g.tolist() == [-1]


@FORMS
def test_search_for_initialization_contention(gx, oracle, monkeypatch, form):
    """K6 under pressure: every frame-1 keypoint present five times with a few flipped descriptor bits, so rivals keep taking
    frame-2 features from each other (vnMatches21 steals, vMatchedDistance skips) inside and across the 32-query passes;
    windows of 100, 30 and 8 px; frames of 0, 1 and 33 keypoints."""
    _pick_form(monkeypatch, form)
    w, h = 640, 480
    G = _ext(gx, w, h, 1000)
    a, b, _ = synth.mono_pair(64, w, h)
    k1, d1 = G(a); k2, d2 = G(b)
    rng = np.random.default_rng(77)
    total = 0
    for case, (n1, dup, n2, win) in enumerate([(len(k1), 5, len(k2), 100), (300, 5, len(k2), 30), (len(k1), 3, 400, 8), (33, 2, 1, 100), (1, 1, 33, 100), (0, 1, 50, 100), (50, 1, 0, 100)]):
        src = np.repeat(np.arange(n1), dup)
        ka, da = k1[src].copy(), d1[src].copy()
        for i in range(len(src)):
            for bit in rng.integers(0, 256, rng.integers(0, 4)):
                da[i, bit >> 3] ^= np.uint8(1 << (bit & 7))
        kb, db = k2[:n2], d2[:n2]
        prev = np.stack([ka["x"], ka["y"]], 1).astype(np.float32)
        rm12, rn, rprev = oracle.search_init(ka, da, kb, db, w, h, prev.copy(), 0.9, win, True)
        p2 = prev.copy()
        gn, g12 = gx.ORBmatcher(0.9, True).SearchForInitialization(gx.Frame(ka, da, w, h), gx.Frame(kb, db, w, h), p2, win)
        assert gn == rn and np.array_equal(g12, rm12) and np.array_equal(p2, rprev), (case, gn, rn)
        total += rn
    assert total > 50


@pytest.mark.parametrize("th,ratio,strict", [(50, 0.75, True), (50, 0.95, False), (100, 0.9, False), (255, 0.999, False), (30, 0.2, True), (255, 1.5, False)])
def test_knn_ratio_crowded_neighbourhoods(gx, oracle, th, ratio, strict):
    """K7's ordered resolve under pressure: groups of identical queries facing ladders of near-copies in B
    (each accepted match consumes the nearest free copy, so a query's eight nearest B features are all
    taken and the warp has to rescan), exact duplicates (second == best, ratio test fails), and
    thresholds / ratios that move or disable the far-neighbour cut-off."""
    rng = np.random.default_rng(1234 + th)
    def flip(d, bits):
        o = d.copy()
        for b in bits: o[b >> 3] ^= np.uint8(1 << (b & 7))
        return o
    A, B = [], []
    for g in range(6):
        x = rng.integers(0, 256, 32, dtype=np.uint8)
        nq = 14 + g
        for _ in range(nq): A.append(x.copy())
        perm = rng.permutation(256)
        for k in range(nq + 4): B.append(flip(x, perm[:k * (1 + (g & 1))]))   # k (or 2k) bits away from x
    for _ in range(40):                                                       # unrelated descriptors and exact duplicates
        y = rng.integers(0, 256, 32, dtype=np.uint8)
        A.append(y.copy()); B.append(y.copy()); B.append(y.copy())
    for _ in range(300):
        A.append(rng.integers(0, 256, 32, dtype=np.uint8)); B.append(rng.integers(0, 256, 32, dtype=np.uint8))
    A = np.stack(A); B = np.stack(B)
    ia, ib = rng.permutation(len(A)), rng.permutation(len(B))
    A, B = np.ascontiguousarray(A[ia]), np.ascontiguousarray(B[ib])
    angA = rng.uniform(0, 360, len(A)).astype(np.float32); angB = rng.uniform(0, 360, len(B)).astype(np.float32)
    for ori in (True, False):
        r = oracle.knn_ratio(A, angA, B, angB, th, ratio, ori, strict)
        n1, g = gx.ORBmatcher(ratio, ori).SearchByBoW(A, angA, B, angB, th=th, strict=strict)
        assert n1 == r[1] and np.array_equal(g, r[0]), (ori, n1, r[1])
    if ratio >= 0.9:
        assert r[1] > 30                                                       # ladders deeper than the eight-entry lists
    # fewer B features than the list length, and a single B feature
    for nb in (1, 3, 7, 8, 9):
        r = oracle.knn_ratio(A[:50], angA[:50], B[:nb], angB[:nb], th, ratio, False, strict)
        n1, g = gx.ORBmatcher(ratio, False).SearchByBoW(A[:50], angA[:50], B[:nb], angB[:nb], th=th, strict=strict)
        assert n1 == r[1] and np.array_equal(g, r[0]), nb


@FORMS
@pytest.mark.parametrize("th", [1.0, 3.0])
def test_search_by_projection_local_map(gx, oracle, th, monkeypatch, form):
    """SearchByProjection(Frame, MapPoints): map points come from a second, shifted view of the scene
    (their descriptors, octaves and noisy predicted positions), some already-occupied keypoints, some
    points out of view, stereo uRight checks on."""
    _pick_form(monkeypatch, form)
    cfg = synth.KITTI
    w, h, nf = cfg["w"], cfg["h"], cfg["nfeatures"]
    G = _ext(gx, w, h, nf, batch=2)
    L, R, _ = synth.stereo_pair(90, w, h)
    a, b, (dx, dy) = synth.mono_pair(91, w, h)
    G.extract_batch(np.stack([L, R]))
    ur, dp, nm = gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=1, slotL0=0, slotR0=1)
    (kf, df), _ = G.fetch(0, 2)
    k2, d2 = G(L[::-1, ::-1].copy())                       # a different view as the "map": rotate by 180 degrees
    rng = np.random.default_rng(int(th * 10))
    m = len(k2)
    mp = dict(inview=(rng.random(m) < 0.9).astype(np.uint8),
              projx=(w - 1 - k2["x"] + rng.normal(0, 2.0, m)).astype(np.float32),
              projy=(h - 1 - k2["y"] + rng.normal(0, 2.0, m)).astype(np.float32),
              level=np.clip(k2["octave"] + rng.integers(0, 2, m), 0, 7).astype(np.int32),
              viewcos=rng.choice([0.9, 0.9985, 0.999], m).astype(np.float32),
              desc=d2.copy(), obs=rng.integers(0, 4, m).astype(np.int32))
    mp["projxr"] = (mp["projx"] - rng.uniform(0, 40, m)).astype(np.float32)
    # descriptors of a 180-degree view do not match; plant the frame's own descriptors with a few bit flips
    near = np.array([np.argmin((kf["x"] - px) ** 2 + (kf["y"] - py) ** 2) for px, py in zip(mp["projx"], mp["projy"])])
    mp["desc"] = df[near].copy()
    flips = rng.integers(0, 256, (m, 12))
    for i in range(m):
        for bit in flips[i][:rng.integers(0, 12)]:
            mp["desc"][i, bit >> 3] ^= np.uint8(1 << (bit & 7))
    n = len(kf)
    frame_obs = np.where(rng.random(n) < 0.1, rng.integers(0, 3, n), -1).astype(np.int32)
    scale = G.GetScaleFactors()
    ras, rn, rfo = oracle.search_projection(kf, df, ur[0, :n], w, h, scale, mp, th, 0.8, frame_obs)
    gn, gas, gfo = gx.ORBmatcher(0.8, True).SearchByProjection(gx.Frame(kf, df, w, h), ur[0, :n], scale, mp, th, frame_obs)
    assert gn == rn and np.array_equal(gas, ras) and np.array_equal(gfo, rfo)
    assert rn > 300
    if th == 1.0:       # the many-frames-per-launch instantiation (descriptors stay in global memory): 160 copies in one call
        ns, ass, fos = gx.ORBmatcher(0.8, True).SearchByProjection(gx.Frame(kf, df, w, h), ur[0, :n], scale, mp, th, frame_obs, replicas=160)
        assert all(v == rn for v in ns) and all(np.array_equal(x, ras) for x in ass) and all(np.array_equal(x, rfo) for x in fos)


@FORMS
def test_search_by_projection_contention_and_odd_sizes(gx, oracle, monkeypatch, form):
    """K8's ordered resolution under pressure and at odd sizes: every map point repeated up to seven times (more rivals for a
    keypoint than a candidate list has entries: lists run out and are re-scanned, rounds need several fixpoint iterations),
    Observations() mixed (claiming and non-claiming matches on the same keypoint), and frame / map sizes of 0, 1, 31, 33, ..."""
    _pick_form(monkeypatch, form)
    cfg = synth.KITTI
    w, h, nf = cfg["w"], cfg["h"], cfg["nfeatures"]
    G = _ext(gx, w, h, nf, batch=2)
    L, R, _ = synth.stereo_pair(93, w, h)
    G.extract_batch(np.stack([L, R]))
    ur, dp, nm = gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=1, slotL0=0, slotR0=1)
    (kf, df), _ = G.fetch(0, 2)
    scale = G.GetScaleFactors()
    rng = np.random.default_rng(4711)
    total = 0
    for case, (n_sub, m_sub, dup, th) in enumerate([(len(kf), 300, 7, 3.0), (len(kf), 600, 3, 1.0), (33, 64, 5, 3.0), (31, 1, 1, 3.0), (1, 33, 2, 3.0),
                                                     (0, 10, 1, 3.0), (500, 0, 1, 3.0), (len(kf), 97, 6, 7.0), (777, 777, 4, 3.0)]):
        k, d, u = kf[:n_sub], df[:n_sub], ur[0, :n_sub]
        src = rng.choice(max(n_sub, 1), m_sub) if n_sub else np.zeros(m_sub, np.int64)
        idx = np.repeat(src, dup)[:max(m_sub * dup, 0)]
        m = len(idx)
        base = kf[idx] if n_sub else np.zeros(m, kf.dtype)
        mp = dict(inview=(rng.random(m) < 0.95).astype(np.uint8),
                  projx=(base["x"] + rng.normal(0, 1.0, m)).astype(np.float32), projy=(base["y"] + rng.normal(0, 1.0, m)).astype(np.float32),
                  level=np.clip(base["octave"] + rng.integers(0, 2, m), 0, 7).astype(np.int32),
                  viewcos=rng.choice([0.9, 0.9985, 0.999], m).astype(np.float32),
                  desc=(df[idx].copy() if n_sub else np.zeros((m, 32), np.uint8)), obs=rng.integers(0, 3, m).astype(np.int32))
        mp["projxr"] = (mp["projx"] - rng.uniform(0, 40, m)).astype(np.float32)
        for i in range(m):                                   # rivals differ by a few bits: equal and near-equal distances
            for bit in rng.integers(0, 256, rng.integers(0, 6)):
                mp["desc"][i, bit >> 3] ^= np.uint8(1 << (bit & 7))
        frame_obs = np.where(rng.random(n_sub) < 0.1, rng.integers(0, 3, n_sub), -1).astype(np.int32)
        ras, rn, rfo = oracle.search_projection(k, d, u, w, h, scale, mp, th, 0.8, frame_obs)
        gn, gas, gfo = gx.ORBmatcher(0.8, True).SearchByProjection(gx.Frame(k, d, w, h), u, scale, mp, th, frame_obs)
        assert gn == rn and np.array_equal(gas, ras) and np.array_equal(gfo, rfo), (case, gn, rn)
        total += rn
    assert total > 500


@FORMS
@pytest.mark.parametrize("mode,th,ori", [(0, 15.0, True), (1, 7.0, True), (2, 15.0, False)])
def test_search_by_projection_motion_model(gx, oracle, mode, th, ori, monkeypatch, form):
    """SearchByProjection(Current, Last): last-frame stereo points are back-projected to 3-D, moved by a
    small pose change and searched for in a noisy copy of the same view."""
    _pick_form(monkeypatch, form)
    cfg = synth.KITTI
    w, h, nf, bf, fx = cfg["w"], cfg["h"], cfg["nfeatures"], cfg["bf"], cfg["fx"]
    fy, cx, cy = fx, 607.19, 185.22                              # Examples/Stereo/KITTI00-02.yaml
    L, R, _ = synth.stereo_pair(95, w, h)
    G = _ext(gx, w, h, nf, batch=2)
    G.extract_batch(np.stack([L, R]))
    ur, dp, nm = gx.ComputeStereoMatches(G, G, bf, fx, n=1, slotL0=0, slotR0=1)
    (kl, dl), _ = G.fetch(0, 2)
    nl = len(kl)
    z = dp[0, :nl]
    rng = np.random.default_rng(mode)
    last = dict(valid=((z > 0) & (rng.random(nl) < 0.9)).astype(np.uint8),
                xw=np.stack([(kl["x"] - cx) * z / fx, (kl["y"] - cy) * z / fy, z], 1).astype(np.float32),
                octave=kl["octave"].astype(np.int32), angle=kl["angle"].astype(np.float32), desc=dl.copy(),
                obs=rng.integers(0, 4, nl).astype(np.int32))
    cur_img = synth._add_noise(L, 4242, amp=3)
    cur_r = synth._add_noise(R, 4243, amp=3)
    G.extract_batch(np.stack([cur_img, cur_r]))
    ur2, dp2, _ = gx.ComputeStereoMatches(G, G, bf, fx, n=1, slotL0=0, slotR0=1)
    (kc, dc), _ = G.fetch(0, 2)
    n = len(kc)
    ang = 0.002
    Rcw = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]], np.float32)
    tcw = np.array([0.05, 0.01, 0.3 if mode == 2 else -0.3], np.float32)
    frame_obs = np.where(rng.random(n) < 0.08, rng.integers(0, 3, n), -1).astype(np.int32)
    scale = G.GetScaleFactors()
    cam = (fx, fy, cx, cy, bf)
    ras, rn, rfo = oracle.search_projection_frame(kc, dc, ur2[0, :n], w, h, scale, Rcw, tcw, cam, last, th, mode, ori, frame_obs)
    gn, gas, gfo = gx.ORBmatcher(0.9, ori).SearchByProjectionFrame(gx.Frame(kc, dc, w, h), ur2[0, :n], scale, Rcw, tcw, cam, last, th, mode, frame_obs)
    assert gn == rn and np.array_equal(gas, ras) and np.array_equal(gfo, rfo)
    assert rn > 100


@FORMS
def test_search_by_projection_frame_contention(gx, oracle, monkeypatch, form):
    """K9 under pressure: every last-frame point present six times over (nearly equal descriptors), so six map points want
    each current keypoint, eight-pixel to thirty-pixel windows, claiming and non-claiming Observations() mixed; and tiny frames."""
    _pick_form(monkeypatch, form)
    cfg = synth.KITTI
    w, h, nf, bf, fx = cfg["w"], cfg["h"], cfg["nfeatures"], cfg["bf"], cfg["fx"]
    fy, cx, cy = fx, 607.19, 185.22
    L, R, _ = synth.stereo_pair(97, w, h)
    G = _ext(gx, w, h, nf, batch=2)
    G.extract_batch(np.stack([L, R]))
    ur, dp, nm = gx.ComputeStereoMatches(G, G, bf, fx, n=1, slotL0=0, slotR0=1)
    (kl, dl), _ = G.fetch(0, 2)
    z = dp[0, :len(kl)]
    good = np.nonzero(z > 0)[0]
    scale = G.GetScaleFactors()
    cam = (fx, fy, cx, cy, bf)
    Rcw = np.eye(3, dtype=np.float32); tcw = np.array([0.02, 0.0, -0.1], np.float32)
    total = 0
    for case, (n_cur, n_src, dup, th, mode) in enumerate([(len(kl), 330, 6, 15.0, 0), (len(kl), 200, 6, 7.0, 1), (40, 30, 5, 15.0, 2), (1, 8, 2, 15.0, 0), (700, 1, 1, 15.0, 0)]):
        rng = np.random.default_rng(100 + case)
        src = np.repeat(rng.choice(good, n_src), dup)
        nl = len(src)
        desc = dl[src].copy()
        for i in range(nl):
            for bit in rng.integers(0, 256, rng.integers(0, 5)):
                desc[i, bit >> 3] ^= np.uint8(1 << (bit & 7))
        last = dict(valid=(rng.random(nl) < 0.9).astype(np.uint8),
                    xw=np.stack([(kl["x"][src] - cx) * z[src] / fx, (kl["y"][src] - cy) * z[src] / fy, z[src]], 1).astype(np.float32),
                    octave=kl["octave"][src].astype(np.int32), angle=kl["angle"][src].astype(np.float32), desc=desc,
                    obs=rng.integers(0, 3, nl).astype(np.int32))
        kc, dc, uc = kl[:n_cur], dl[:n_cur], ur[0, :n_cur]
        frame_obs = np.where(rng.random(n_cur) < 0.05, rng.integers(0, 3, n_cur), -1).astype(np.int32)
        ras, rn, rfo = oracle.search_projection_frame(kc, dc, uc, w, h, scale, Rcw, tcw, cam, last, th, mode, True, frame_obs)
        gn, gas, gfo = gx.ORBmatcher(0.9, True).SearchByProjectionFrame(gx.Frame(kc, dc, w, h), uc, scale, Rcw, tcw, cam, last, th, mode, frame_obs)
        assert gn == rn and np.array_equal(gas, ras) and np.array_equal(gfo, rfo), (case, gn, rn)
        total += rn
    assert total > 300


def test_euroc_shard_of_512_frames(gx, oracle):
    """BASELINE config 3: one GPU's contiguous 512-frame shard of the 4096-frame EuRoC-shaped sequence,
    processed in 4 batches of 128 stereo pairs.  Distinct frames are checked against the oracle; the
    frames that repeat (the synthetic sequence cycles through 32 pairs) must reproduce them byte for byte."""
    from orb_slam2_b200 import sharding
    cfg = synth.EUROC
    w, h, nf, B = cfg["w"], cfg["h"], cfg["nfeatures"], 128
    lo, hi = sharding.shard_range(4096, 8, 3)
    assert hi - lo == 512
    pairs = [synth.stereo_pair(lo + i, w, h)[:2] for i in range(32)]
    G = _ext(gx, w, h, nf, batch=2 * B)
    ref = {}
    for i in (0, 7, 31):
        EL, ER = oracle.Extractor(nf), oracle.Extractor(nf)
        kl, dl = EL.extract(pairs[i][0]); kr, dr = ER.extract(pairs[i][1])
        ref[i] = (kl, dl, kr, dr, oracle.stereo(EL, ER, kl, dl, kr, dr, cfg["bf"], cfg["fx"]))
    first = None
    for b in range(4):
        idx = [(b * B + j) % 32 for j in range(B)]
        G.extract_batch(np.stack([pairs[i][0] for i in idx] + [pairs[i][1] for i in idx]))
        ur, dp, nm = gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=B, slotL0=0, slotR0=B)
        res = G.fetch(0, 2 * B)
        for j, i in enumerate(idx):
            if i in ref and j < 32:
                kl, dl, kr, dr, (rur, rdp, _, rm) = ref[i]
                assert res[j][0].tobytes() == kl.tobytes() and np.array_equal(res[j][1], dl)
                assert res[B + j][0].tobytes() == kr.tobytes() and np.array_equal(res[B + j][1], dr)
                assert int(nm[j]) == rm and np.array_equal(ur[j, :len(kl)].view(np.uint32), rur.view(np.uint32))
            if j >= 32:                                    # repeats of frame j % 32 inside the same batch
                assert res[j][0].tobytes() == res[j % 32][0].tobytes() and np.array_equal(res[j][1], res[j % 32][1])
                assert np.array_equal(ur[j].view(np.uint32), ur[j % 32].view(np.uint32)) and nm[j] == nm[j % 32]
        sig = (res[0][0].tobytes(), ur[:32].tobytes())
        if first is None:
            first = sig
        else:
            assert sig == first                            # every batch of the shard starts on the same frame


def test_full_size_batch_properties(gx):
    """BASELINE-sized batch (KITTI stereo): properties that need no oracle — identical frames give
    identical bytes in every slot, output is level-major, per-level counts within [quota, quota+3],
    descriptors of a frame matched against themselves give distance 0 on the diagonal."""
    cfg = synth.KITTI
    w, h, nf, n = cfg["w"], cfg["h"], cfg["nfeatures"], 32
    L, R, _ = synth.stereo_pair(80, w, h)
    G = _ext(gx, w, h, nf, batch=2 * n)
    G.extract_batch(np.stack([L] * n + [R] * n))
    ur, dp, nm = gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=n, slotL0=0, slotR0=n)
    res = G.fetch(0, 2 * n)
    quota = G.features_per_level()
    for i in range(1, n):
        assert res[i][0].tobytes() == res[0][0].tobytes() and np.array_equal(res[i][1], res[0][1])
        assert res[n + i][0].tobytes() == res[n][0].tobytes()
        assert np.array_equal(ur[i].view(np.uint32), ur[0].view(np.uint32)) and nm[i] == nm[0]
    k = res[0][0]
    assert np.all(np.diff(k["octave"]) >= 0)
    per = [int((k["octave"] == l).sum()) for l in range(8)]
    assert all(quota[l] <= per[l] <= quota[l] + 3 for l in range(8)), per
    d = res[0][1]
    assert all(gx.ORBmatcher.DescriptorDistance(d[i], d[i]) == 0 for i in range(0, len(d), 97))
    assert nm[0] > 100


TUM1_CAM = (517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314)


@pytest.mark.parametrize("cam,w,h", [(TUM1_CAM, 640, 480), ((718.856, 718.856, 607.1928, 185.2157, 0, 0, 0, 0, 0), 1241, 376),
                                     ((458.654, 457.296, 367.215, 248.375, -0.28340811, 0.07395907, 0.00019359, 1.76187114e-05), 752, 480)])
def test_undistort_grid_and_features_in_area(gx, oracle, cam, w, h):
    """K10 / K11: mvKeysUn bit-exact (double-precision undistortPoints restated on the device),
    image bounds, the 64x48 grid in push_back order, GetFeaturesInArea in the reference's enumeration order.
    Distorted (TUM1, EuRoC raw) and rectified (KITTI: plain copy) cameras; frames of different sizes in one batch."""
    G = _ext(gx, w, h, 1000)
    frames = [synth.mono_frame(300 + i, w, h) for i in range(3)]
    kls = [G(f)[0] for f in frames]
    kls[1] = kls[1][:417]                                     # ragged batch
    kls.append(kls[0][:0])                                    # an empty frame
    FG = gx.FrameGrids(kls, cam, w, h)
    rng = np.random.default_rng(7)
    for f, k in enumerate(kls):
        kun, b, cs, items = oracle.undistort_grid(cam, k, w, h)
        got = FG.mvKeysUn[f]
        assert np.array_equal(got.view(np.uint8), kun.view(np.uint8)), f
        assert np.array_equal(FG.bounds.view(np.uint32), b.view(np.uint32))
        assert np.array_equal(FG.cell_start[f], cs), f
        assert np.array_equal(FG.cell_items[f, :cs[-1]], items), f
    if cam[4] != 0:
        assert not np.array_equal(FG.mvKeysUn[0]["x"], kls[0]["x"])
        assert FG.bounds[0] != 0.0
    else:
        assert np.array_equal(FG.mvKeysUn[0]["x"], kls[0]["x"]) and FG.bounds[0] == 0.0 and FG.bounds[1] == w
    queries = []
    for _ in range(300):
        f = int(rng.integers(0, len(kls)))
        lo, hi = (-1, -1) if rng.random() < 0.4 else (int(rng.integers(0, 3)), int(rng.integers(-1, 8)))
        queries.append((f, float(rng.uniform(-30, w + 30)), float(rng.uniform(-30, h + 30)), float(rng.uniform(1, 150)), lo, hi))
    queries.append((0, -500.0, 100.0, 10.0, -1, -1))           # window left of the grid
    queries.append((0, w / 2, h / 2, 5000.0, -1, -1))          # window covering everything
    got, cnt = FG.GetFeaturesInArea(queries)
    for (f, x, y, r, lo, hi), g, c in zip(queries, got, cnt):
        want = oracle.features_in_area(FG.mvKeysUn[f], FG.bounds, x, y, r, lo, hi)
        assert c == len(want) and np.array_equal(g, want), (f, x, y, r, lo, hi)
    assert cnt[-1] == len(kls[0])
    # truncation: count is still the full count, the first qcap indices are written
    got2, cnt2 = FG.GetFeaturesInArea(queries[-1:], qcap=10)
    assert cnt2[0] == len(kls[0]) and np.array_equal(got2[0], got[-1][:10])


def _vocab_nodes(desc, words):
    """A stand-in vocabulary: node = nearest of `words` in Hamming distance (ties -> lowest id), ids spread out."""
    x = np.unpackbits(desc, axis=1).astype(np.int16); wv = np.unpackbits(words, axis=1).astype(np.int16)
    d = (x[:, None, :] != wv[None, :, :]).sum(-1)
    return (d.argmin(1) * 37 + 5).astype(np.int32)


@FORMS
@pytest.mark.parametrize("strict,ratio,ori", [(True, 0.75, True), (False, 0.7, True), (False, 0.9, False)])
def test_search_by_bow_with_vocabulary_buckets(gx, oracle, strict, ratio, ori, monkeypatch, form):
    """K12: SearchByBoW(KF,KF) / (KF,Frame) over per-node buckets, in the reference's node / index
    order, with invalid map points on both sides, nodes present on one side only, ragged pairs and an empty side."""
    _pick_form(monkeypatch, form)
    w, h, nf = 1241, 376, 2000
    G = _ext(gx, w, h, nf)
    rng = np.random.default_rng(99)
    words = rng.integers(0, 256, (96, 32), dtype=np.uint8)
    sidesA, sidesB, refs = [], [], []
    for p in range(3):
        a, b, _ = synth.mono_pair(170 + p, w, h)
        k1, d1 = G(a); k2, d2 = G(b)
        if p == 1: k2, d2 = k2[:700], d2[:700]
        if p == 2: words[:, :] = rng.integers(0, 256, (96, 32), dtype=np.uint8); words = words[:7]   # few, crowded nodes
        n1, n2 = _vocab_nodes(d1, words), _vocab_nodes(d2, words)
        n2[n2 == n2.max()] = 10 ** 6                        # a node the A side does not have
        v1 = (rng.random(len(d1)) < 0.8).astype(np.uint8)
        v2 = np.ones(len(d2), np.uint8) if not strict else (rng.random(len(d2)) < 0.85).astype(np.uint8)
        sidesA.append((d1, k1["angle"], v1, n1)); sidesB.append((d2, k2["angle"], v2, n2))
        refs.append(oracle.search_by_bow(d1, k1["angle"], v1, n1, d2, k2["angle"], v2, n2, 50, ratio, ori, strict))
    sidesA.append((sidesA[0][0][:5], sidesA[0][1][:5], sidesA[0][2][:5], sidesA[0][3][:5]))
    sidesB.append((sidesB[0][0][:0], sidesB[0][1][:0], sidesB[0][2][:0], sidesB[0][3][:0]))
    refs.append((np.full(5, -1, np.int32), 0))
    nm, m12 = gx.ORBmatcher(ratio, ori).SearchByBoWNodes(sidesA, sidesB, strict=strict)
    for p in range(4):
        assert nm[p] == refs[p][1] and np.array_equal(m12[p], refs[p][0]), (p, nm[p], refs[p][1])
    assert refs[0][1] > 50 and refs[2][1] > 50
    # the same pairs 17 times over in one call (more CTAs than one wave of the 1024-thread form)
    nm2, m2 = gx.ORBmatcher(ratio, ori).SearchByBoWNodes(sidesA * 17, sidesB * 17, strict=strict)
    for q in range(68):
        assert nm2[q] == refs[q % 4][1] and np.array_equal(m2[q], refs[q % 4][0]), (q, nm2[q], refs[q % 4][1])
    # matches never cross nodes and never use an invalid feature
    m = m12[0]; ok = m >= 0
    assert np.array_equal(sidesA[0][3][ok], sidesB[0][3][m[ok]]) and sidesA[0][2][ok].all() and sidesB[0][2][m[ok]].all()


def test_distinctive_descriptors(gx, oracle):
    """K13: MapPoint::ComputeDistinctiveDescriptors, median by histogram vs the reference's sort.
    Clusters of noisy copies of a descriptor with 1..200 observations, exact duplicates (ties -> first), empty lists."""
    rng = np.random.default_rng(21)
    lists = []
    for n in [1, 2, 3, 4, 5, 8, 31, 32, 33, 64, 100, 200] + list(rng.integers(2, 60, 40)):
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        noise = np.packbits(rng.random((n, 256)) < rng.uniform(0.02, 0.25), axis=1)
        lists.append(base[None, :] ^ noise)
    lists.append(np.repeat(rng.integers(0, 256, (1, 32), dtype=np.uint8), 7, 0))     # all equal: index 0
    lists.append(np.zeros((0, 32), np.uint8))                                          # no observation
    lists.append(rng.integers(0, 256, (50, 32), dtype=np.uint8))                       # unrelated descriptors
    best, out = gx.ComputeDistinctiveDescriptors(lists)
    for m, d in enumerate(lists):
        want = oracle.distinctive_descriptor(d)
        assert best[m] == want, (m, len(d), best[m], want)
        if want >= 0:
            assert np.array_equal(out[m], d[want])
    assert best[-2] == -1 and best[-3] == 0


@FORMS
@pytest.mark.parametrize("only_stereo,ori", [(False, True), (True, True), (False, False)])
def test_search_for_triangulation(gx, oracle, only_stereo, ori, monkeypatch, form):
    """K14: SearchForTriangulation + CheckDistEpipolarLine.  The second keyframe is a translated
    view, so F12 = [t]x of the image translation makes true matches lie on their epipolar lines; the epipole is put
    inside the image so that the mono-mono exclusion zone fires; stereo / mono features and existing map points mixed."""
    _pick_form(monkeypatch, form)
    w, h, nf = 1241, 376, 2000
    G = _ext(gx, w, h, nf)
    scale, sigma2 = np.asarray(G.GetScaleFactors(), np.float32), np.asarray(G.GetScaleSigmaSquares(), np.float32)
    rng = np.random.default_rng(314)
    words = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    kf1, kf2, F12, poses, Cws, cams, refs = [], [], [], [], [], [], []
    for p in range(3):
        a, b, (dx, dy) = synth.mono_pair(230 + p, w, h)
        k1, d1 = G(a); k2, d2 = G(b)
        if p == 2: k1, d1 = k1[:900], d1[:900]
        mk = lambda k, d: dict(kps=k, desc=d, has_mp=(rng.random(len(k)) < 0.3).astype(np.uint8),
                               u_right=np.where(rng.random(len(k)) < 0.5, k["x"] - 3.0, -1.0).astype(np.float32), node=_vocab_nodes(d, words))
        A, B = mk(k1, d1), mk(k2, d2)
        t = np.array([dx, dy, 0.0]) / max(np.hypot(dx, dy), 1.0)
        F = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]], np.float32) + rng.normal(0, 1e-4, (3, 3)).astype(np.float32)
        R = np.eye(3, dtype=np.float32); tw = np.array([0.02 * p, -0.01, 1.0], np.float32); Cw = np.array([0.1, 0.05, 0.03 * p], np.float32)
        cam = np.array([718.856, 718.856, 607.1928, 185.2157], np.float32)
        kf1.append(A); kf2.append(B); F12.append(F); poses.append((R, tw)); Cws.append(Cw); cams.append(cam)
        refs.append(oracle.search_for_triangulation(k1, d1, A["has_mp"], A["u_right"], A["node"], k2, d2, B["has_mp"], B["u_right"], B["node"],
                                                    F, np.concatenate([R.reshape(9), tw]), Cw, cam, scale, sigma2, only_stereo, ori))
    nm, m12 = gx.ORBmatcher(0.6, ori).SearchForTriangulation(kf1, kf2, F12, poses, Cws, cams, scale, sigma2, bOnlyStereo=only_stereo)
    for p in range(3):
        assert nm[p] == refs[p][1] and np.array_equal(m12[p], refs[p][0]), (p, nm[p], refs[p][1])
    assert refs[0][1] > (5 if only_stereo else 30)
    # 66 pairs in one call
    nm2, m2 = gx.ORBmatcher(0.6, ori).SearchForTriangulation(kf1 * 22, kf2 * 22, F12 * 22, poses * 22, Cws * 22, cams * 22, scale, sigma2, bOnlyStereo=only_stereo)
    for q in range(66):
        assert nm2[q] == refs[q % 3][1] and np.array_equal(m2[q], refs[q % 3][0]), (q, nm2[q], refs[q % 3][1])


@pytest.mark.parametrize("k,L,levelsup,weighting,norm", [(10, 4, 2, 0, 1), (10, 3, 4, 0, 1), (6, 5, 3, 1, 2), (10, 4, 2, 2, 1), (8, 3, 1, 3, 0), (10, 4, 2, 1, 0)])
def test_bow_transform(gx, oracle, k, L, levelsup, weighting, norm):
    """K15: DBoW2 transform on a synthetic vocabulary tree (the reference ships no ORBvoc.txt):
    word and node per feature bit-exact, BowVector words in order and values equal to the last bit of the double
    (weights summed per feature in feature order, normalised by the sum in word order), stopped words dropped;
    all weighting / normalisation modes, levelsup beyond the tree (node = root), ragged and empty frames."""
    w, h = 1241, 376
    G = _ext(gx, w, h, 2000)
    lists = [G(synth.mono_frame(400 + i, w, h))[1] for i in range(2)]
    V = synth.vocabulary(17 + L, k, L, stop_fraction=0.05, train=lists[0])
    lists.append(lists[1][:333]); lists.append(lists[0][:0]); lists.append(np.repeat(lists[0][:1], 40, 0))   # one word, 40 times
    voc = gx.ORBVocabulary(V, weighting=weighting, norm=norm)
    got = voc.transform(lists, levelsup)
    for f, d in enumerate(lists):
        wo, no, bw, bv = oracle.voc_transform(V, d, levelsup, weighting, norm)
        g = got[f]
        assert np.array_equal(g[0], wo) and np.array_equal(g[1], no), f
        assert np.array_equal(g[2], bw), f
        assert np.array_equal(g[3].view(np.uint64), bv.view(np.uint64)), (f, np.abs(g[3] - bv).max() if len(bv) else 0)
    assert len(got[0][2]) > 100 and (got[0][0] < 0).any()
    voc.close()


def test_cuda_path_reproduces_reference_goldens(gx):
    """The CUDA path against tests/golden/ref_slam.json directly — digests written by the reference's own Frame.cc /
    ORBmatcher.cc compiled in place (tools/gen_golden_slam.py), no oracle in between: stereo uRight / depth bits of the
    stereo Frame constructor, SearchForInitialization matches, mvKeysUn / bounds / grid / GetFeaturesInArea."""
    import hashlib, importlib.util, json, os
    here = os.path.dirname(__file__)
    gold = json.load(open(os.path.join(here, "golden", "ref_slam.json")))
    spec = importlib.util.spec_from_file_location("gen_golden_slam", os.path.join(here, "..", "tools", "gen_golden_slam.py"))
    gs = importlib.util.module_from_spec(spec); spec.loader.exec_module(gs)
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    for case in gold["stereo"]:
        cfg = dict(kitti=synth.KITTI, euroc=synth.EUROC)[case["name"]]
        L, R, _ = synth.stereo_pair(case["seed"], cfg["w"], cfg["h"])
        G = _ext(gx, cfg["w"], cfg["h"], cfg["nfeatures"], batch=2)
        G.extract_batch(np.stack([L, R]))
        ur, dp, nm = gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=1, slotL0=0, slotR0=1)
        (kl, _), (kr, _) = G.fetch(0, 2)
        n = len(kl)
        assert n == case["n"] and len(kr) == case["n_right"] and sha(kl) == case["keys_sha256"]
        assert int(nm[0]) == case["matches"] and sha(ur[0, :n]) == case["u_right_sha256"] and sha(dp[0, :n]) == case["depth_sha256"]
    for case in gold["init"]:
        a, b, _ = synth.mono_pair(case["seed"], 640, 480)
        G = _ext(gx, 640, 480, 1000)
        k1, d1 = G(a); k2, d2 = G(b)
        prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
        n, m12 = gx.ORBmatcher(case["ratio"], case["check_ori"]).SearchForInitialization(gx.Frame(k1, d1, 640, 480), gx.Frame(k2, d2, 640, 480), prev, case["window"])
        assert n == case["nmatches"] and sha(m12.astype(np.int32)) == case["matches12_sha256"] and sha(prev) == case["prev_sha256"]
    for case in gold["grid"]:
        w, h = case["w"], case["h"]
        k, _ = _ext(gx, w, h, case["nfeatures"])(synth.mono_frame(case["frame"], w, h))
        FG = gx.FrameGrids([k], case["cam"], w, h)
        assert sha(FG.mvKeysUn[0]) == case["keys_un_sha256"] and [float(v).hex() for v in FG.bounds] == case["bounds_hex"]
        cells = [FG.cell(0, c // 48, c % 48) for c in range(64 * 48)]
        assert sha(np.concatenate([np.array([len(c)], np.int32) for c in cells] + cells)) == case["cells_sha256"]
        hits, cnt = FG.GetFeaturesInArea([(0,) + q for q in gs.grid_queries(case["frame"], w, h)])
        assert int(cnt.sum()) == case["total_hits"]
        assert sha(np.concatenate([np.array([len(x)], np.int32) for x in hits] + hits)) == case["hits_sha256"]


@FORMS
def test_cuda_keyframe_searches_reproduce_reference(gx, monkeypatch, form):
    """K7 / K8 / K9 / K12 / K13 / K14 / K15 against what the REFERENCE's ORBmatcher.cc / MapPoint.cc produced on real KeyFrames
    and MapPoints (tests/golden/ref_kf_*.npz, ref_slam.json written by oracle/_ref/orbref_slam): frames, stereo depth and
    feature-vector nodes come from the CUDA path itself, only the MapPoint / isInFrustum fields come from the fixture."""
    _pick_form(monkeypatch, form)
    import hashlib, importlib.util, os
    spec = importlib.util.spec_from_file_location("kf_cases", os.path.join(os.path.dirname(__file__), "kf_cases.py"))
    kc = importlib.util.module_from_spec(spec); spec.loader.exec_module(kc)
    from oracle.pyoracle import kf_rule                        # index rule only, no oracle arithmetic
    gs, P = kc.gs, kc.gs.KF_PARAMS
    cfg = synth.KITTI
    w, h, nf = cfg["w"], cfg["h"], cfg["nfeatures"]
    for case in kc.GOLD["kf"]:
        ims, V, T1, T2, F12 = gs.kf_case_inputs(case["seed"], case["L"])
        fx = np.load(os.path.join(kc.HERE, "golden", case["fixture"]))
        ori, only_stereo = case["check_ori"], case["only_stereo"]
        G = _ext(gx, w, h, nf, batch=4)
        G.extract_batch(np.stack([ims[0], ims[2], ims[1], ims[3]]))
        ur, dp, _ = gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=2, slotL0=0, slotR0=2)
        (k1, d1), (k2, d2) = G.fetch(0, 2)
        ur1, dp1, ur2, dp2 = ur[0, :len(k1)], dp[0, :len(k1)], ur[1, :len(k2)], dp[1, :len(k2)]
        scale, sigma2 = np.asarray(G.GetScaleFactors(), np.float32), np.asarray(G.GetScaleSigmaSquares(), np.float32)
        voc = gx.ORBVocabulary(V)
        (_, node1, _, _), (_, node2, _, _) = voc.transform([d1, d2], 4)
        voc.close()
        assert np.array_equal(node1, fx["node1"]) and np.array_equal(node2, fx["node2"])        # K15 vs the reference's mFeatVec
        has1 = np.array([dp1[i] > 0 and kf_rule(i, 0, P["pct1"]) for i in range(len(k1))], np.uint8)
        has2 = np.array([dp2[i] > 0 and kf_rule(i, 1, P["pct2"]) for i in range(len(k2))], np.uint8)
        assert np.array_equal(has1, fx["has1"]) and np.array_equal(has2, fx["has2"])
        # K12: SearchByBoW(KF, KF) and (KF, Frame)
        M = gx.ORBmatcher(P["ratio_kk"], ori)
        nm, m12 = M.SearchByBoWNodes([(d1, k1["angle"], has1, node1)], [(d2, k2["angle"], has2, node2)], strict=True)
        assert nm[0] == case["n_bow_kk"] and np.array_equal(m12[0], fx["bow_kk"])
        if case["L"] <= 4:                                       # K7: one bucket = brute force over the valid features
            va, vb = has1.astype(bool), has2.astype(bool)
            nk, mk = M.SearchByBoW(d1[va], k1["angle"][va], d2[vb], k2["angle"][vb], strict=True)
            full = np.full(len(va), -1, np.int32); full[np.nonzero(va)[0][mk >= 0]] = np.nonzero(vb)[0][mk[mk >= 0]]
            assert nk == case["n_bow_kk"] and np.array_equal(full, fx["bow_kk"])
        nm, m12 = gx.ORBmatcher(P["ratio_kf"], ori).SearchByBoWNodes([(d1, k1["angle"], has1, node1)], [(d2, k2["angle"], np.ones(len(d2), np.uint8), node2)], strict=False)
        inv = np.full(len(d2), -1, np.int32); ok = m12[0] >= 0; inv[m12[0][ok]] = np.nonzero(ok)[0]
        assert nm[0] == case["n_bow_kf"] and np.array_equal(inv, fx["bow_kf"])
        # K14: SearchForTriangulation
        nm, m12 = gx.ORBmatcher(0.6, ori).SearchForTriangulation([dict(kps=k1, desc=d1, has_mp=has1, u_right=ur1, node=node1)],
                                                                 [dict(kps=k2, desc=d2, has_mp=has2, u_right=ur2, node=node2)], [F12], [(T2[:3, :3], T2[:3, 3])],
                                                                 [fx["Ow1"]], [np.array(gs.KITTI_CAM, np.float32)], scale, sigma2, bOnlyStereo=only_stereo)
        assert nm[0] == case["n_tri"] and np.array_equal(np.stack([np.nonzero(m12[0] >= 0)[0], m12[0][m12[0] >= 0]], 1).astype(np.int32), fx["tri"])
        # K8: SearchByProjection(Frame, MapPoints) with the reference's isInFrustum fields
        mp, tr = fx["mps1"], fx["track"]
        mpd = dict(inview=tr["inview"].astype(np.uint8), projx=tr["projx"], projy=tr["projy"], projxr=tr["projxr"], level=tr["level"], viewcos=tr["viewcos"],
                   desc=mp["desc"], obs=mp["obs"])
        Fc = gx.Frame(k2, d2, w, h)
        n, a, _ = gx.ORBmatcher(0.8, True).SearchByProjection(Fc, ur2, scale, mpd, P["th_proj"])
        assert n == case["n_proj"] and np.array_equal(a, fx["proj"])
        # K9: SearchByProjection(CurrentFrame, LastFrame)
        n1 = len(k1)
        last = dict(valid=has1, xw=np.zeros((n1, 3), np.float32), octave=k1["octave"], angle=k1["angle"], desc=np.zeros((n1, 32), np.uint8), obs=np.zeros(n1, np.int32))
        last["xw"][mp["idx"]] = mp["xw"]; last["desc"][mp["idx"]] = mp["desc"]; last["obs"][mp["idx"]] = mp["obs"]
        n, a, _ = gx.ORBmatcher(0.9, ori).SearchByProjectionFrame(Fc, ur2, scale, T2[:3, :3], T2[:3, 3], gs.KITTI_CAM + (cfg["bf"],), last, P["th_last"], 0)
        assert n == case["n_proj_last"] and np.array_equal(a, fx["proj_last"])
    # K13 against MapPoint::ComputeDistinctiveDescriptors of the reference
    lists = gs.distinct_lists(kc.GOLD["distinct"]["seed"])
    best, out = gx.ComputeDistinctiveDescriptors(lists)
    assert hashlib.sha256(np.ascontiguousarray(out).tobytes()).hexdigest() == kc.GOLD["distinct"]["chosen_sha256"]


@pytest.mark.parametrize("grid", ["given", "built"])
@FORMS
def test_cuda_fuse_and_sim3_reproduce_reference(gx, oracle, monkeypatch, form, grid):
    """K16 against ORBmatcher::Fuse (search) and ORBmatcher::SearchBySim3 of the REFERENCE on real KeyFrames and MapPoints
    (tests/golden/ref_kf_*.npz): frames, stereo, the feature grid (K10) come from the CUDA path, MapPoint fields from the fixture."""
    _pick_form(monkeypatch, form)
    # grid = built: cell_start / cell_items are NULL and the library assigns keypoints to cells itself (in the kernel for the
    # 1024-thread form, on the host for the small form) - what the drop-in does
    monkeypatch.setattr(gx.ORBmatcher, "PASS_GRID", grid == "given")
    import importlib.util, os
    spec = importlib.util.spec_from_file_location("kf_cases", os.path.join(os.path.dirname(__file__), "kf_cases.py"))
    kc = importlib.util.module_from_spec(spec); spec.loader.exec_module(kc)
    gs = kc.gs
    cfg = synth.KITTI
    w, h, nf = cfg["w"], cfg["h"], cfg["nfeatures"]
    for case in kc.GOLD["kf"]:
        ims, V, T1, T2, F12 = gs.kf_case_inputs(case["seed"], case["L"])
        fx = np.load(os.path.join(kc.HERE, "golden", case["fixture"]))
        G = _ext(gx, w, h, nf, batch=4)
        G.extract_batch(np.stack([ims[0], ims[2], ims[1], ims[3]]))
        ur, dp, _ = gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=2, slotL0=0, slotR0=2)
        (k1, d1), (k2, d2) = G.fetch(0, 2)
        scale = np.asarray(G.GetScaleFactors(), np.float32)
        FG = gx.FrameGrids([k1, k2], gs.KITTI_CAM, w, h)
        mp1, mp2 = fx["mps1"], fx["mps2"]
        R2, t2 = T2[:3, :3], T2[:3, 3]
        cam = gs.KITTI_CAM + (cfg["bf"],)
        mpd = dict(valid=np.ones(len(mp1), np.uint8), xw=mp1["xw"], normal=mp1["normal"], dmin=mp1["dmin"], dmax=mp1["dmax"], desc=mp1["desc"])
        best = gx.ORBmatcher.FuseSearch(FG, 1, ur[1, :len(k2)], d2, fx["kf_bounds"], R2, t2, (-R2.T @ t2).astype(np.float32), cam, scale, fx["inv_sigma2"],
                                        float(fx["log_scale"]), mpd, 3.0)
        assert int((best >= 0).sum()) == case["n_fuse"] and np.array_equal(best, fx["fuse"])
        side = lambda mp: dict(index=mp["idx"], xw=mp["xw"], dmin=mp["dmin"], dmax=mp["dmax"], desc=mp["desc"])
        n, m12 = gx.ORBmatcher.SearchBySim3(FG, d1, d2, fx["kf_bounds"], (T1[:3, :3], T1[:3, 3]), (R2, t2), gs.KITTI_CAM, scale, float(fx["log_scale"]), fx["sim3_xf"],
                                            side(mp1), side(mp2), 7.5)
        assert n == case["n_sim3"] and np.array_equal(m12, fx["sim3"])
        # K16 with the Sim3 variant of Fuse (:977) and K17: SearchByProjection(pKF, Scw, ...) :290 / SearchByProjection(Frame, pKF, ...) :1472
        from oracle.pyoracle import kf_rule                   # index rule only
        x = fx["scw_xf"]; Rcw, tcw, Ow = x[:9].reshape(3, 3), x[9:12], x[12:15]
        best = gx.ORBmatcher.FuseSearch(FG, 1, ur[1, :len(k2)], d2, fx["kf_bounds"], Rcw, tcw, Ow, cam, scale, fx["inv_sigma2"], float(fx["log_scale"]), mpd, 4.0,
                                        sim3_variant=True)
        assert int((best >= 0).sum()) == case["n_fuse_scw"] and np.array_equal(best, fx["fuse_scw"])
        n2 = len(k2); has2 = fx["has2"]
        mpo = dict(mpd, angle=k1["angle"][mp1["idx"]])
        occ = np.array([has2[i] and kf_rule(i, 7, 30) for i in range(n2)], np.uint8)
        n, a = gx.ORBmatcher(0.75, True).SearchByProjectionOrdered(FG, 1, d2, fx["kf_bounds"], Rcw, tcw, Ow, gs.KITTI_CAM, scale, float(fx["log_scale"]), occ, mpo, 10.0, 50, False)
        assert n == case["n_proj_scw"] and np.array_equal(a, np.where(fx["proj_scw"] >= 0, fx["proj_scw"], -1))
        occ = np.array([has2[i] and kf_rule(i, 11, 20) for i in range(n2)], np.uint8)
        mpr = dict(mpo, valid=np.array([not kf_rule(k, 9, 25) for k in range(len(mp1))], np.uint8))
        n, a = gx.ORBmatcher(0.9, case["check_ori"]).SearchByProjectionOrdered(FG, 1, d2, (0.0, float(w), 0.0, float(h)), R2, t2, fx["reloc_Ow"], gs.KITTI_CAM, scale,
                                                                                 float(fx["log_scale"]), occ, mpr, 10.0, 100, True)
        assert n == case["n_proj_kf"] and np.array_equal(a, np.where(fx["proj_kf"] >= 0, fx["proj_kf"], -1))
        # K17 under contention: every map point six times over with a few flipped bits, so six rivals want each keypoint - more than
        # a candidate list holds (re-scans) and several per 32-point round (fixpoint iterations); against the CPU oracle
        rng = np.random.default_rng(5)
        dup = {key: np.repeat(np.asarray(v), 6, axis=0) for key, v in mpr.items()}
        dup["desc"] = dup["desc"].copy()
        for i in range(len(dup["desc"])):
            for bit in rng.integers(0, 256, rng.integers(0, 5)):
                dup["desc"][i, bit >> 3] ^= np.uint8(1 << (bit & 7))
        fb = (0.0, float(w), 0.0, float(h))
        for reloc, Rm, tm, Om, kb, thd in [(False, Rcw, tcw, Ow, tuple(fx["kf_bounds"]), 50), (True, R2, t2, fx["reloc_Ow"], fb, 100)]:
            kfv = (*gs.KITTI_CAM, cfg["bf"], float(fx["log_scale"]), *kb)
            ra, rn = oracle.project_match_ordered(FG.mvKeysUn[1], d2, kfv, fb, scale, reloc, Rm, tm, Om, dup, 10.0, thd, case["check_ori"], occ)
            n, a = gx.ORBmatcher(0.9, case["check_ori"]).SearchByProjectionOrdered(FG, 1, d2, kb, Rm, tm, Om, gs.KITTI_CAM, scale, float(fx["log_scale"]), occ, dup, 10.0, thd, reloc)
            assert n == rn and np.array_equal(a, ra), (reloc, n, rn)
            assert rn > 100


def test_stereo_direct_device_outputs_wider_than_kp_cap(gx, oracle):
    """Two handles (n == max_batch) writing uRight / depth straight into DEVICE buffers whose row pitch is larger than
    orbx_max_keypoints() (a caller rounding 2024 up to 2048): the SAD scratch of the left handle keeps its own pitch."""
    import torch
    cfg = synth.EUROC
    w, h, nf, n = cfg["w"], cfg["h"], cfg["nfeatures"], 3
    pairs = [synth.stereo_pair(70 + i, w, h) for i in range(n)]
    GL, GR = _ext(gx, w, h, nf, batch=n), _ext(gx, w, h, nf, batch=n)
    GL.extract_batch(np.stack([p[0] for p in pairs])); GR.extract_batch(np.stack([p[1] for p in pairs]))
    cap = GL.kp_cap + 24
    guard = 4096
    buf = torch.full((3, n * cap + guard), -7.0, dtype=torch.float32, device="cuda")      # guard floats behind each output
    ur, dp = buf[0, :n * cap].view(n, cap), buf[1, :n * cap].view(n, cap)
    nm = torch.zeros(n, dtype=torch.int32, device="cuda")
    gx.ComputeStereoMatches(GL, GR, cfg["bf"], cfg["fx"], n=n, out=(ur, dp, nm))
    torch.cuda.synchronize()
    assert bool((buf[:, n * cap:] == -7.0).all()), "wrote past the caller's buffers"
    ur, dp, nm = ur.cpu().numpy(), dp.cpu().numpy(), nm.cpu().numpy()
    resL = GL.fetch(0, n)
    for i in range(n):
        EL, ER = oracle.Extractor(nf), oracle.Extractor(nf)
        kl, dl = EL.extract(pairs[i][0]); kr, dr = ER.extract(pairs[i][1])
        assert resL[i][0].tobytes() == kl.tobytes()
        rur, rdp, _, rm = oracle.stereo(EL, ER, kl, dl, kr, dr, cfg["bf"], cfg["fx"])
        assert int(nm[i]) == rm and rm > 50
        assert np.array_equal(ur[i, :len(kl)].view(np.uint32), rur.view(np.uint32)) and np.array_equal(dp[i, :len(kl)].view(np.uint32), rdp.view(np.uint32))


def test_cuda_searches_on_distorted_cameras_reproduce_reference(gx):
    """K6 / K8 / K9 on cameras WITH distortion (Frame::mnMinX / mnMinY != 0) against tests/golden/ref_dist.json and
    ref_kfd_*.npz, which the reference's own Frame.cc / ORBmatcher.cc wrote (tools/gen_golden_dist.py): undistorted keypoints
    (K10), SearchForInitialization, SearchByProjection(Frame, MapPoints), SearchByProjection(Current, Last)."""
    import importlib.util
    here = os.path.dirname(__file__)
    spec = importlib.util.spec_from_file_location("kf_cases", os.path.join(here, "kf_cases.py"))
    kc = importlib.util.module_from_spec(spec); spec.loader.exec_module(kc)
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    for case in kc.GOLD_DIST["init"]:
        a, b, _ = synth.mono_pair(case["seed"], 640, 480)
        G = _ext(gx, 640, 480, 1000)
        k1, d1 = G(a); k2, d2 = G(b)
        FG = gx.FrameGrids([k1, k2], case["cam"], 640, 480)
        ku1, ku2 = FG.mvKeysUn
        assert sha(ku2) == case["keys_un2_sha256"] and FG.bounds[0] != 0.0
        prev = np.stack([ku1["x"], ku1["y"]], 1).astype(np.float32)
        n, m12 = gx.ORBmatcher(case["ratio"], case["check_ori"]).SearchForInitialization(gx.Frame(ku1, d1, 640, 480, FG.bounds), gx.Frame(ku2, d2, 640, 480, FG.bounds),
                                                                                         prev, case["window"])
        assert n == case["nmatches"] and sha(m12.astype(np.int32)) == case["matches12_sha256"] and sha(prev) == case["prev_sha256"]
    cfg = synth.KITTI
    w, h, nf = cfg["w"], cfg["h"], cfg["nfeatures"]
    for case in kc.GOLD_DIST["kf"]:
        def extract_stereo(L, R):
            G = _ext(gx, w, h, nf, batch=2)
            G.extract_batch(np.stack([L, R]))
            ur, dp, _ = gx.ComputeStereoMatches(G, G, cfg["bf"], cfg["fx"], n=1, slotL0=0, slotR0=1)
            (kl, dl), _ = G.fetch(0, 2)
            return dict(keys=kl, desc=dl, u_right=ur[0, :len(kl)].copy(), depth=dp[0, :len(kl)].copy())

        def undistort(cam, k, w_, h_):
            FG = gx.FrameGrids([k], cam, w_, h_)
            return FG.mvKeysUn[0], FG.bounds.copy()

        c = kc.build_dist(extract_stereo, undistort, case)
        F1, F2, fx, T2 = c["F1"], c["F2"], c["fx"], c["T2"]
        assert sha(F1["keys_un"]) == case["keys_un1_sha256"] and sha(F2["keys_un"]) == case["keys_un2_sha256"] and sha(F2["u_right"]) == case["u_right2_sha256"]
        scale = np.asarray(_ext(gx, w, h, nf).GetScaleFactors(), np.float32)
        Fc = gx.Frame(F2["keys_un"], F2["desc"], w, h, F2["bounds"])
        P = kc.gs.KF_PARAMS
        n, a, _ = gx.ORBmatcher(0.8, True).SearchByProjection(Fc, F2["u_right"], scale, c["mpd"], P["th_proj"])
        assert n == case["n_proj"] and np.array_equal(a, fx["proj"])
        n, a, _ = gx.ORBmatcher(0.9, case["check_ori"]).SearchByProjectionFrame(Fc, F2["u_right"], scale, T2[:3, :3], T2[:3, 3], c["cam"] + (cfg["bf"],), c["last"],
                                                                               P["th_last"], 0)
        assert n == case["n_proj_last"] and np.array_equal(a, fx["proj_last"])


def test_first_extract_on_another_host_thread(gx, oracle):
    """The reference constructs its extractors on the tracking thread and calls them on two fresh std::threads
    (src/Tracking.cc:120-126, src/Frame.cc:78-81).  A handle created here and first used on threads that never touched CUDA
    must work (the geometry set-up reaches the driver API, which needs a context bound to the calling thread)."""
    import ctypes as C
    import threading
    cfg = synth.EUROC
    w, h, nf = cfg["w"], cfg["h"], cfg["nfeatures"]
    L, R, _ = synth.stereo_pair(81, w, h)
    GL, GR = _ext(gx, w, h, nf), _ext(gx, w, h, nf)
    lib = gx.lib()
    sL, sR = C.c_void_p(), C.c_void_p()
    assert lib.orbx_stream_create(0, C.byref(sL)) == 0 and lib.orbx_stream_create(0, C.byref(sR)) == 0
    out, err = {}, []
    def run(name, G, im, st):
        try:
            out[name] = G(im, stream=st)
        except Exception as exc:
            err.append(exc)
    ts = [threading.Thread(target=run, args=("L", GL, L, sL)), threading.Thread(target=run, args=("R", GR, R, sR))]
    [t.start() for t in ts]; [t.join() for t in ts]
    assert not err, err
    ur, dp, nm = gx.ComputeStereoMatches(GL, GR, cfg["bf"], cfg["fx"])
    EL, ER = oracle.Extractor(nf), oracle.Extractor(nf)
    okl, odl = EL.extract(L); okr, odr = ER.extract(R)
    assert out["L"][0].tobytes() == okl.tobytes() and out["R"][0].tobytes() == okr.tobytes() and np.array_equal(out["R"][1], odr)
    rur, rdp, _, rm = oracle.stereo(EL, ER, okl, odl, okr, odr, cfg["bf"], cfg["fx"])
    assert int(nm[0]) == rm and np.array_equal(ur[0, :len(okl)].view(np.uint32), rur.view(np.uint32))
    lib.orbx_stream_destroy(sL); lib.orbx_stream_destroy(sR)


def test_matchers_read_extractor_slots_in_place(gx, oracle):
    """orbx_slot_results: the frames stay where the extractor left them on the device; UndistortKeyPoints + grid (K10) and
    SearchForInitialization (K6) read the slots in place (on_device = 1) - no fetch, no re-upload - and give what the host route
    and the oracle give."""
    import ctypes as C
    import torch
    w, h = 640, 480
    a, b, _ = synth.mono_pair(66, w, h)
    G = _ext(gx, w, h, 1000, batch=2)
    G.extract_batch(np.stack([a, b]))
    kd, dd, cd, cap = G.slot_results(0)
    lib = gx.lib()
    st = torch.cuda.current_stream().cuda_stream
    # K10 on the slots (undistorted camera: a copy, bounds = the image) -> device outputs
    kun = torch.empty((2, cap, 28), dtype=torch.uint8, device="cuda"); bounds = torch.empty(6, dtype=torch.float32, device="cuda")
    cs = torch.empty((2, 64 * 48 + 1), dtype=torch.int32, device="cuda"); ci = torch.empty((2, cap), dtype=torch.int32, device="cuda")
    cam = (C.c_float * 9)(517.3, 516.5, 318.6, 255.3, 0, 0, 0, 0, 0)
    assert lib.orbx_undistort_grid_batch(kd, cd, 2, cap, cam, w, h, kun.data_ptr(), bounds.data_ptr(), cs.data_ptr(), ci.data_ptr(), 1, st) == 0
    # K6: frame 1 = slot 0, frame 2 = slot 1, all device pointers
    prev = kun.view(torch.float32).reshape(2, cap, 7)[0, :, :2].contiguous()
    m12 = torch.empty(cap, dtype=torch.int32, device="cuda"); nm = torch.empty(1, dtype=torch.int32, device="cuda")
    rc = lib.orbx_search_init_batch(kd, dd, cd, kd + cap * 28, dd + cap * 32, cd + 4, 1, cap, bounds.data_ptr(), prev.data_ptr(), 0.9, 100, 1,
                                    m12.data_ptr(), nm.data_ptr(), 1, st)
    assert rc == 0, lib.orbx_last_error_string()
    torch.cuda.synchronize()
    (k1, d1), (k2, d2) = G.fetch(0, 2)
    p0 = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    rm12, rn, rprev = oracle.search_init(k1, d1, k2, d2, w, h, p0, 0.9, 100, True)
    assert int(nm.cpu()[0]) == rn and np.array_equal(m12.cpu().numpy()[:len(k1)], rm12) and rn > 50
    assert np.array_equal(prev.cpu().numpy()[:len(k1)], rprev)
    assert bounds.cpu().tolist()[:4] == [0.0, float(w), 0.0, float(h)]


def test_packed_fetch_and_host_unpack_equal_the_full_records(gx):
    """orbx_fetch_results_packed + orbx_unpack_keypoints (8 bytes per keypoint over the bus, cv::KeyPoint rebuilt on the host)
    against orbx_fetch_results, byte for byte, on frames of all three shapes (different level geometries and counts)."""
    for name, w, h, nf in CONFIGS:
        imgs = np.stack([synth.mono_frame(400 + i, w, h) for i in range(3)] + [np.full((h, w), 77, np.uint8)])   # the last one: no keypoints
        G = _ext(gx, w, h, nf, batch=4)
        G.extract_batch(imgs)
        full = G.fetch(0, 4)
        cap = G.kp_cap
        packed = np.zeros((4, cap), np.uint64); desc = np.zeros((4, cap, 32), np.uint8); lc = np.zeros((4, 8), np.int32)
        G.fetch_packed_into(0, 4, packed, desc, lc)
        import torch
        torch.cuda.synchronize()
        kps, cnt = G.unpack_keypoints(packed, lc)
        for i in range(4):
            assert cnt[i] == len(full[i][0]) == lc[i].sum(), (name, i)
            assert kps[i, :cnt[i]].tobytes() == full[i][0].tobytes(), (name, i)
            assert np.array_equal(desc[i, :cnt[i]], full[i][1])
        assert cnt[3] == 0 and cnt[0] >= nf


def test_fused_describe_kernel_kept_for_ab_gives_the_same_bytes(gx, oracle):
    """ORBX_DESCRIBE=fused (read at create) selects round 1's blur-on-demand kernel instead of K4a + K4b: same keypoints,
    angles and descriptors, and both equal the oracle."""
    w, h, nf = 752, 480, 1200
    img = synth.mono_frame(500, w, h)
    k0, d0 = _ext(gx, w, h, nf)(img)
    os.environ["ORBX_DESCRIBE"] = "fused"
    try:
        k1, d1 = _ext(gx, w, h, nf)(img)
    finally:
        del os.environ["ORBX_DESCRIBE"]
    ok, od = oracle.Extractor(nf).extract(img)
    assert k0.tobytes() == k1.tobytes() == ok.tobytes() and np.array_equal(d0, d1) and np.array_equal(d0, od)
