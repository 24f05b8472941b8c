"""Parity soak: thousands of DISTINCT frames per shape through the batched CUDA path (C ABI), every byte of every result
compared with the CPU oracle run on all host threads (oracle.batch_stereo).

The hazards this front end is built around are rare events - a descriptor bit that flips when sinf/cosf are one ULP off
(SURVEY.md H2 puts that at one per 10^2..10^3 frames), quad-tree ties (H1), cvRound half-way cases in the rotated taps,
FAST cells whose minThFAST retry fires, sub-pixel parabola edge cases - so a handful of frames per shape proves little.
Here: 2048 stereo pairs of the KITTI and EuRoC shapes (keypoints, descriptors, uRight / depth bits, match counts) and 2048
monocular frames of the TUM shape with the 2 x nFeatures initialisation extractor (src/Tracking.cc:126).  Frames are 32
painted scenes per shape and their roll / flip variants (synth.variant): distinct images, cheap to make.
"""
import numpy as np
import pytest

from orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu

CHUNK = 256
DEFAULT = (1.2, 8, 20, 7)            # scaleFactor, nLevels, iniThFAST, minThFAST of every dataset YAML
SHAPES = [("kitti", synth.KITTI["w"], synth.KITTI["h"], synth.KITTI["nfeatures"], synth.KITTI["bf"], synth.KITTI["fx"], True, 2048, DEFAULT),
          ("euroc", synth.EUROC["w"], synth.EUROC["h"], synth.EUROC["nfeatures"], synth.EUROC["bf"], synth.EUROC["fx"], True, 2048, DEFAULT),
          ("tum-ini", 640, 480, 2000, 0.0, 0.0, False, 2048, DEFAULT),
          # other constructor arguments: different level geometry, cell grid, quotas and both FAST thresholds (stereo, 512 pairs)
          ("svga-1.3x6", 800, 600, 1500, 60.0, 500.0, True, 512, (1.3, 6, 12, 5))]


@pytest.mark.parametrize("name,w,h,nf,bf,fx,stereo,count,ctor", SHAPES, ids=[s[0] for s in SHAPES])
def test_soak_every_result_byte_equals_the_oracle(oracle, name, w, h, nf, bf, fx, stereo, count, ctor):
    import torch
    assert torch.cuda.is_available()
    import orb_slam2_b200 as gx
    from orb_slam2_b200.api import KP_DTYPE
    nbase = 32
    if stereo:
        base = [synth.stereo_pair(20000 + i, w, h)[:2] for i in range(nbase)]
    else:
        base = [(synth.mono_frame(20000 + i, w, h), None) for i in range(nbase)]
    eyes = 2 if stereo else 1
    G = gx.ORBextractor(nf, *ctor, max_w=w, max_h=h, max_batch=eyes * CHUNK)
    cap = G.kp_cap
    params = (nf,) + ctor
    kps = np.zeros((eyes * CHUNK, cap), KP_DTYPE); desc = np.zeros((eyes * CHUNK, cap, 32), np.uint8); cnt = np.zeros(eyes * CHUNK, np.int32)
    seen = set()
    total_kp = total_m = frames_bad = 0
    for c0 in range(0, count, CHUNK):
        idx = range(c0, c0 + CHUNK)
        L = np.stack([synth.variant(base[i % nbase][0], i // nbase) for i in idx])
        R = np.stack([synth.variant(base[i % nbase][1], i // nbase) for i in idx]) if stereo else None
        seen.update(hash(L[j, ::7, ::5].tobytes()) for j in range(CHUNK))
        G.extract_batch(np.concatenate([L, R]) if stereo else L)
        if stereo:
            ur, dp, nm = gx.ComputeStereoMatches(G, G, bf, fx, n=CHUNK, slotL0=0, slotR0=CHUNK)
        G.fetch_into(0, eyes * CHUNK, kps, desc, cnt)
        torch.cuda.synchronize()
        okps, odesc, ocnt, our, odp, onm = oracle.batch_stereo(L, R, params, bf if stereo else 1.0, fx if stereo else 1.0, cap)
        assert np.array_equal(cnt, ocnt), (name, c0, np.nonzero(cnt != ocnt)[0][:8])
        for s in range(eyes * CHUNK):
            c = int(ocnt[s])
            same = kps[s, :c].tobytes() == okps[s, :c].tobytes() and desc[s, :c].tobytes() == odesc[s, :c].tobytes()
            if stereo and s < CHUNK:
                same = same and int(nm[s]) == int(onm[s]) and ur[s, :c].tobytes() == our[s, :c].tobytes() and dp[s, :c].tobytes() == odp[s, :c].tobytes()
            if not same:
                frames_bad += 1
                bad_k = np.nonzero(kps[s, :c] != okps[s, :c])[0][:3]
                bad_d = np.nonzero((desc[s, :c] != odesc[s, :c]).any(1))[0][:3]
                print(f"{name}: frame {c0 + s % CHUNK} eye {s // CHUNK}: keypoints differ at {bad_k}, descriptors at {bad_d}")
        total_kp += int(ocnt.sum()); total_m += int(onm.sum())
    G.close()
    assert frames_bad == 0, f"{name}: {frames_bad} of {eyes * count} frames differ from the oracle"
    assert len(seen) == count, "the soak frames are not all distinct"
    assert total_kp >= eyes * count * nf and (not stereo or total_m > 100 * count)


def test_soak_matchers_on_many_distinct_pairs(oracle):
    """BASELINE.json configs 4 and 5 in breadth: SearchForInitialization (ratio 0.9, window 100) on 192 distinct TUM-shaped
    frame pairs and the one-bucket SearchByBoW (brute-force kNN + ratio 0.75, TH_LOW strict) on 96 distinct KITTI-shaped
    pairs of ~2000 x 2000 features, every match array equal to the oracle's (which is pinned to the reference's ORBmatcher.cc)."""
    import torch
    assert torch.cuda.is_available()
    import orb_slam2_b200 as gx
    # config 4
    w, h, npairs = 640, 480, 192
    base = [synth.mono_pair(30000 + i, w, h)[:2] for i in range(16)]
    A = np.stack([synth.variant(base[i % 16][0], i // 16) for i in range(npairs)])
    B = np.stack([synth.variant(base[i % 16][1], i // 16) for i in range(npairs)])
    G = gx.ORBextractor(1000, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2 * npairs)
    G.extract_batch(np.concatenate([A, B]))
    res = G.fetch(0, 2 * npairs)
    F1 = [gx.Frame(res[i][0], res[i][1], w, h) for i in range(npairs)]
    F2 = [gx.Frame(res[npairs + i][0], res[npairs + i][1], w, h) for i in range(npairs)]
    prevs = [np.stack([f.mvKeysUn["x"], f.mvKeysUn["y"]], 1).astype(np.float32) for f in F1]
    refs = [oracle.search_init(a.mvKeysUn, a.mDescriptors, b.mvKeysUn, b.mDescriptors, w, h, p, 0.9, 100, True) for a, b, p in zip(F1, F2, prevs)]
    nm, m12 = gx.ORBmatcher(0.9, True).SearchForInitializationBatch(F1, F2, prevs, 100)
    bad = [i for i in range(npairs) if nm[i] != refs[i][1] or not np.array_equal(m12[i], refs[i][0]) or not np.array_equal(prevs[i], refs[i][2])]
    assert not bad, f"SearchForInitialization differs on pairs {bad[:8]}"
    assert sum(nm) > 20 * npairs
    G.close()
    # config 5
    w, h, npairs = 1241, 376, 96
    base = [synth.mono_pair(31000 + i, w, h)[:2] for i in range(16)]
    A = np.stack([synth.variant(base[i % 16][0], i // 16) for i in range(npairs)])
    B = np.stack([synth.variant(base[i % 16][1], i // 16) for i in range(npairs)])
    G = gx.ORBextractor(2000, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2 * npairs)
    G.extract_batch(np.concatenate([A, B]))
    res = G.fetch(0, 2 * npairs)
    As = [(res[i][1], res[i][0]["angle"]) for i in range(npairs)]
    Bs = [(res[npairs + i][1], res[npairs + i][0]["angle"]) for i in range(npairs)]
    nm, m12 = gx.ORBmatcher(0.75, True).SearchByBoWBatch(As, Bs, strict=True)
    bad = []
    for i in range(npairs):
        rm, rn = oracle.knn_ratio(As[i][0], As[i][1], Bs[i][0], Bs[i][1], 50, 0.75, True, True)
        if rn != nm[i] or not np.array_equal(rm, m12[i]):
            bad.append(i)
    assert not bad, f"kNN + ratio differs on pairs {bad[:8]}"
    assert sum(nm) > 100 * npairs
    G.close()
