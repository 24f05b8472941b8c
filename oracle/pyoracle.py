"""ctypes binding of oracle/liborb_oracle.so — TEST INFRASTRUCTURE ONLY.

Imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, never by the
orb_slam2_b200 package."""
import ctypes as C
import os
import sys
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


def build():
    subprocess.run(["make", "-s", "-C", _HERE], check=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liborb_oracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.orbo_create.restype = C.c_void_p
        _LIB.orbo_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        _LIB.orbo_destroy.argtypes = [C.c_void_p]
        _LIB.orbo_atan2.restype = C.c_float
        _LIB.orbo_atan2.argtypes = [C.c_float, C.c_float]
        _LIB.orbo_bench_stereo.restype = C.c_double
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Extractor:
    def __init__(self, nfeatures, scale=1.2, nlevels=8, ini=20, mn=7):
        self.nlevels, self.nfeatures = nlevels, nfeatures
        self.args = (nfeatures, scale, nlevels, ini, mn)
        self.h = C.c_void_p(lib().orbo_create(nfeatures, C.c_float(scale), nlevels, ini, mn))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orbo_destroy(self.h)
            self.h = None

    def tables(self):
        n = self.nlevels
        f = [np.zeros(n, np.float32) for _ in range(4)]
        q = np.zeros(n, np.int32); u = np.zeros(16, np.int32)
        lib().orbo_tables(self.h, *[_p(a) for a in f], _p(q), _p(u))
        return dict(scale=f[0], inv=f[1], sigma2=f[2], invsigma2=f[3], quota=q, umax=u)

    def extract(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        cap = self.nfeatures + 3 * self.nlevels + 64
        kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
        n = lib().orbo_extract(self.h, _p(img), img.shape[1], img.shape[0], img.strides[0], _p(kps), _p(desc), cap)
        if n < 0:
            raise RuntimeError(f"oracle extract failed: {n}")
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l):
        w, h = C.c_int(), C.c_int()
        lib().orbo_level_size(self.h, l, C.byref(w), C.byref(h))
        buf = np.zeros((h.value + 38, w.value + 38), np.uint8)
        lib().orbo_level_copy(self.h, l, _p(buf))
        return buf

    def candidates(self, l, cap=200000):
        out = np.zeros(cap, KP_DTYPE)
        n = lib().orbo_level_candidates(self.h, l, _p(out), cap)
        return out[:n].copy()

    def level_keys(self, l, cap=20000):
        out = np.zeros(cap, KP_DTYPE)
        n = lib().orbo_level_keys(self.h, l, _p(out), cap)
        return out[:n].copy()


def stereo(eL, eR, kL, dL, kR, dR, bf, fx):
    n = len(kL)
    ur = np.zeros(n, np.float32); dp = np.zeros(n, np.float32); sad = np.zeros(n, np.int32)
    kL = np.ascontiguousarray(kL); kR = np.ascontiguousarray(kR)
    dL = np.ascontiguousarray(dL); dR = np.ascontiguousarray(dR)
    m = lib().orbo_stereo(eL.h, eR.h, _p(kL), _p(dL), n, _p(kR), _p(dR), len(kR), C.c_float(bf), C.c_float(fx),
                          _p(ur), _p(dp), _p(sad))
    return ur, dp, sad, m


def _bounds4(cols, rows, bounds):
    """Frame::mnMinX, mnMaxX, mnMinY, mnMaxY: [0, cols] x [0, rows] for an undistorted camera."""
    return np.array([0, cols, 0, rows] if bounds is None else list(bounds)[:4], np.float32)


def search_init(k1, d1, k2, d2, cols, rows, prev, ratio=0.9, window=100, check_ori=True, bounds=None):
    k1 = np.ascontiguousarray(k1); k2 = np.ascontiguousarray(k2)
    d1 = np.ascontiguousarray(d1); d2 = np.ascontiguousarray(d2)
    prev = np.ascontiguousarray(prev, np.float32).copy()
    m12 = np.zeros(len(k1), np.int32)
    b = _bounds4(cols, rows, bounds)
    n = lib().orbo_search_init(_p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), _p(b), _p(prev),
                               C.c_float(ratio), window, int(check_ori), _p(m12))
    return m12, n, prev


def knn_ratio(dA, angA, dB, angB, th=50, ratio=0.75, check_ori=True, strict=True):
    dA = np.ascontiguousarray(dA); dB = np.ascontiguousarray(dB)
    angA = np.ascontiguousarray(angA, np.float32); angB = np.ascontiguousarray(angB, np.float32)
    m12 = np.zeros(len(dA), np.int32)
    n = lib().orbo_knn_ratio(_p(dA), _p(angA), len(dA), _p(dB), _p(angB), len(dB), th, C.c_float(ratio),
                             int(check_ori), int(strict), _p(m12))
    return m12, n


def search_by_bow(dA, angA, validA, nodeA, dB, angB, validB, nodeB, th=50, ratio=0.75, check_ori=True, strict=True):
    """ORBmatcher::SearchByBoW (KF,KF: strict=True; KF,Frame: strict=False) with per-feature vocabulary node ids."""
    dA = np.ascontiguousarray(dA); dB = np.ascontiguousarray(dB)
    angA = np.ascontiguousarray(angA, np.float32); angB = np.ascontiguousarray(angB, np.float32)
    validA = np.ascontiguousarray(validA, np.uint8); validB = np.ascontiguousarray(validB, np.uint8)
    nodeA = np.ascontiguousarray(nodeA, np.int32); nodeB = np.ascontiguousarray(nodeB, np.int32)
    m12 = np.zeros(max(len(dA), 1), np.int32)
    n = lib().orbo_search_by_bow(_p(dA), _p(angA), _p(validA), _p(nodeA), len(dA), _p(dB), _p(angB), _p(validB), _p(nodeB), len(dB),
                                 th, C.c_float(ratio), int(check_ori), int(strict), _p(m12))
    return m12[:len(dA)], n


def search_for_triangulation(k1, d1, has1, ur1, node1, k2, d2, has2, ur2, node2, F12, pose2, Cw, cam2, scale2, sigma2_2,
                             only_stereo=False, check_ori=True):
    """ORBmatcher::SearchForTriangulation on flat arrays -> (vMatches12, nmatches)."""
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    u8 = lambda a: np.ascontiguousarray(a, np.uint8)
    i32 = lambda a: np.ascontiguousarray(a, np.int32)
    k1 = np.ascontiguousarray(k1); k2 = np.ascontiguousarray(k2)
    args = [k1, u8(d1), u8(has1), f32(ur1), i32(node1)], [k2, u8(d2), u8(has2), f32(ur2), i32(node2)]
    tail = [f32(F12).reshape(9), f32(pose2).reshape(12), f32(Cw).reshape(3), f32(cam2), f32(scale2), f32(sigma2_2)]
    m12 = np.zeros(max(len(k1), 1), np.int32)
    n = lib().orbo_search_for_triangulation(*[_p(a) for a in args[0]], len(k1), *[_p(a) for a in args[1]], len(k2),
                                            *[_p(a) for a in tail], int(only_stereo), int(check_ori), _p(m12))
    return m12[:len(k1)], n


def voc_transform(voc, desc, levelsup=4, weighting=0, norm=1):
    """DBoW2 TemplatedVocabulary::transform on a flat vocabulary (dict: desc, child_first, child_count, children, word_id,
    weight, L) -> (word per feature, node per feature, BowVector words, BowVector values)."""
    desc = np.ascontiguousarray(desc, np.uint8); n = len(desc)
    wo = np.zeros(max(n, 1), np.int32); no = np.zeros(max(n, 1), np.int32)
    bw = np.zeros(max(n, 1), np.int32); bv = np.zeros(max(n, 1), np.float64)
    m = lib().orbo_voc_transform(_p(voc["desc"]), _p(voc["child_first"]), _p(voc["child_count"]), _p(voc["children"]), _p(voc["word_id"]),
                                 _p(voc["weight"]), int(voc["L"]), _p(desc), n, levelsup, weighting, norm, _p(wo), _p(no), _p(bw), _p(bv))
    return wo[:n], no[:n], bw[:m].copy(), bv[:m].copy()


REF_SLAM_BIN = os.path.join(_HERE, "_ref", "orbref_slam")


def ref_slam_available():
    return os.path.exists(REF_SLAM_BIN)


def _calib_bytes(cam, bf, th_depth=35.0):
    """cam: (fx, fy, cx, cy[, k1, k2, p1, p2, k3])"""
    c = np.zeros(11, np.float32); c[:len(cam)] = np.asarray(cam, np.float32); c[9] = bf; c[10] = th_depth
    return c.tobytes()


def _run_ref_slam(cmd, payload):
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        with open(fin, "wb") as f:
            f.write(payload)
        subprocess.run([REF_SLAM_BIN, cmd, fin, fout], check=True, capture_output=True)
        return open(fout, "rb").read()


class _Rd:
    def __init__(self, raw):
        self.raw, self.off = raw, 0

    def arr(self, dtype, count, shape=None):
        dt = np.dtype(dtype)
        a = np.frombuffer(self.raw, dt, count, self.off).copy(); self.off += dt.itemsize * count
        return a.reshape(shape) if shape is not None else a

    def i32(self):
        return int(self.arr(np.int32, 1)[0])

    def frame(self):
        n = self.i32()
        return dict(keys=self.arr(KP_DTYPE, n), keys_un=self.arr(KP_DTYPE, n), desc=self.arr(np.uint8, 32 * n, (n, 32)),
                    u_right=self.arr(np.float32, n), depth=self.arr(np.float32, n))


def ref_slam_stereo(imL, imR, nfeatures, cam, bf):
    """The REFERENCE's stereo Frame constructor (src/Frame.cc:61-117): extraction of both eyes + ComputeStereoMatches."""
    imL = np.ascontiguousarray(imL, np.uint8); imR = np.ascontiguousarray(imR, np.uint8); h, w = imL.shape
    rd = _Rd(_run_ref_slam("stereo", np.array([w, h, nfeatures], np.int32).tobytes() + _calib_bytes(cam, bf) + imL.tobytes() + imR.tobytes()))
    F = rd.frame(); nr = rd.i32()
    F["keys_right"] = rd.arr(KP_DTYPE, nr); F["desc_right"] = rd.arr(np.uint8, 32 * nr, (nr, 32))
    return F


def ref_slam_init(im1, im2, nfeatures, cam, ratio=0.9, window=100, check_ori=True):
    """The REFERENCE's monocular Frame constructor twice + ORBmatcher::SearchForInitialization.
    Returns (F1, F2, nmatches, vnMatches12, vbPrevMatched after the call)."""
    im1 = np.ascontiguousarray(im1, np.uint8); im2 = np.ascontiguousarray(im2, np.uint8); h, w = im1.shape
    pay = (np.array([w, h, nfeatures], np.int32).tobytes() + _calib_bytes(cam, 0.0) + np.float32(ratio).tobytes() +
           np.array([window, int(check_ori)], np.int32).tobytes() + im1.tobytes() + im2.tobytes())
    rd = _Rd(_run_ref_slam("init", pay))
    F1, F2 = rd.frame(), rd.frame()
    n = rd.i32(); n1 = len(F1["keys"])
    return F1, F2, n, rd.arr(np.int32, n1), rd.arr(np.float32, 2 * n1, (n1, 2))


def ref_slam_grid(im, nfeatures, cam, queries):
    """The REFERENCE's monocular Frame constructor on a (possibly distorted) camera: mvKeysUn, image bounds, mGrid, and
    Frame::GetFeaturesInArea for queries [(x, y, r, minLevel, maxLevel), ...]."""
    im = np.ascontiguousarray(im, np.uint8); h, w = im.shape
    q = b"".join(np.array(t[:3], np.float32).tobytes() + np.array(t[3:5], np.int32).tobytes() for t in queries)
    rd = _Rd(_run_ref_slam("grid", np.array([w, h, nfeatures], np.int32).tobytes() + _calib_bytes(cam, 0.0) + np.int32(len(queries)).tobytes() + q + im.tobytes()))
    F = rd.frame()
    bounds = rd.arr(np.float32, 6)
    cells = []
    for _ in range(64 * 48):
        c = rd.i32(); cells.append(rd.arr(np.int32, c))
    hits = []
    for _ in queries:
        c = rd.i32(); hits.append(rd.arr(np.int32, c))
    return F, bounds, cells, hits


def kf_rule(i, salt, pct):
    """Which stereo features get a MapPoint in ref_slam_kf (mirrors kf_rule in oracle/ref_slam_cmds.inc)."""
    return int((((i + 1) * 2654435761 + salt * 40503) & 0xFFFFFFFF) >> 16) % 100 < pct


def ref_slam_kf(L1, R1, L2, R2, nfeatures, cam, bf, voc, T1, T2, F12, pct1=60, pct2=50, ratio_kk=0.75, ratio_kf=0.7,
                th_proj=3.0, th_last=15.0, check_ori=True, only_stereo=False, mono=False, fuse_scw=False):
    """Two real KeyFrames with real MapPoints and DBoW2 feature vectors, and the REFERENCE's SearchByBoW (both forms),
    SearchForTriangulation and SearchByProjection (both forms) on them.  voc = flat vocabulary dict (written out in the
    reference's text format, L1_NORM + TF_IDF).  Returns a dict of everything the restatements need and the results."""
    import tempfile
    ims = [np.ascontiguousarray(x, np.uint8) for x in (L1, R1, L2, R2)]
    h, w = ims[0].shape
    with tempfile.TemporaryDirectory() as td:
        vp = os.path.join(td, "voc.txt")
        write_vocabulary_text(voc, vp, 0, 0)
        pb = vp.encode()
        pay = (np.array([w, h, nfeatures], np.int32).tobytes() + _calib_bytes(cam, bf) + np.array([pct1, pct2], np.int32).tobytes() +
               np.array([ratio_kk, ratio_kf, th_proj, th_last], np.float32).tobytes() +
               np.array([int(check_ori), int(only_stereo), int(mono), int(fuse_scw)], np.int32).tobytes() +
               np.asarray(T1, np.float32).reshape(16).tobytes() + np.asarray(T2, np.float32).reshape(16).tobytes() +
               np.asarray(F12, np.float32).reshape(9).tobytes() + np.int32(len(pb)).tobytes() + pb + b"".join(x.tobytes() for x in ims))
        rd = _Rd(_run_ref_slam("kf", pay))
    o = dict(F1=rd.frame(), F2=rd.frame())
    n1, n2 = len(o["F1"]["keys"]), len(o["F2"]["keys"])
    o["node1"], o["node2"], o["nodec"] = rd.arr(np.int32, n1), rd.arr(np.int32, n2), rd.arr(np.int32, n2)
    o["has1"], o["has2"] = rd.arr(np.uint8, n1), rd.arr(np.uint8, n2)
    m = rd.i32()
    mp = np.zeros(m, np.dtype([("idx", "<i4"), ("xw", "<f4", 3), ("normal", "<f4", 3), ("desc", "u1", 32), ("obs", "<i4"), ("dmin", "<f4"), ("dmax", "<f4")]))
    for k in range(m):
        mp[k] = (rd.i32(), rd.arr(np.float32, 3), rd.arr(np.float32, 3), rd.arr(np.uint8, 32), rd.i32(), rd.arr(np.float32, 1)[0], rd.arr(np.float32, 1)[0])
    o["mps1"] = mp; o["Ow1"] = rd.arr(np.float32, 3)
    o["bow_kk"] = (rd.i32(), rd.arr(np.int32, n1))
    o["bow_kf"] = (rd.i32(), rd.arr(np.int32, n2))
    nt, npairs = rd.i32(), rd.i32()
    o["tri"] = (nt, rd.arr(np.int32, 2 * npairs, (npairs, 2)))
    trk = np.zeros(m, np.dtype([("inview", "<i4"), ("projx", "<f4"), ("projy", "<f4"), ("projxr", "<f4"), ("level", "<i4"), ("viewcos", "<f4")]))
    for k in range(m):
        trk[k] = (rd.i32(), *rd.arr(np.float32, 3), rd.i32(), rd.arr(np.float32, 1)[0])
    o["track"] = trk
    o["proj"] = (rd.i32(), rd.arr(np.int32, n2))
    o["proj_last"] = (rd.i32(), rd.arr(np.int32, n2))
    m2 = rd.i32()
    mp2 = np.zeros(m2, mp.dtype)
    for k in range(m2):
        mp2[k] = (rd.i32(), rd.arr(np.float32, 3), rd.arr(np.float32, 3), rd.arr(np.uint8, 32), rd.i32(), rd.arr(np.float32, 1)[0], rd.arr(np.float32, 1)[0])
    o["mps2"] = mp2
    o["log_scale"] = float(rd.arr(np.float32, 1)[0]); o["kf_bounds"] = rd.arr(np.float32, 4); o["inv_sigma2"] = rd.arr(np.float32, 8)
    o["sim3_xf"] = dict(sR12=rd.arr(np.float32, 9, (3, 3)), t12=rd.arr(np.float32, 3), sR21=rd.arr(np.float32, 9, (3, 3)), t21=rd.arr(np.float32, 3))
    o["sim3"] = (rd.i32(), rd.arr(np.int32, n1))
    o["scw_xf"] = dict(Rcw=rd.arr(np.float32, 9, (3, 3)), tcw=rd.arr(np.float32, 3), Ow=rd.arr(np.float32, 3))
    o["proj_scw"] = (rd.i32(), rd.arr(np.int32, n2))
    o["reloc_Ow"] = rd.arr(np.float32, 3)
    o["proj_kf"] = (rd.i32(), rd.arr(np.int32, n2))
    o["fuse"] = (rd.i32(), rd.arr(np.int32, m))
    return o


def ref_slam_kfbench(L1, R1, L2, R2, nfeatures, cam, bf, voc, T1, T2, F12, reps=30, pct1=60, pct2=50, ratio_kk=0.75, ratio_kf=0.7,
                     th_proj=3.0, th_last=15.0, check_ori=True, only_stereo=False, mono=False, binary=None, prefix=()):
    """Per-call latency of the ORBmatcher members on the kf case (oracle/ref_slam_cmds.inc:cmd_kfbench) in the given build of the
    driver (`prefix`: a command to run it under, e.g. an ncu launch list).  Returns ({member: (p50_ms, p99_ms, result)}, (n1, n2, map points 1, map points 2))."""
    import tempfile
    ims = [np.ascontiguousarray(x, np.uint8) for x in (L1, R1, L2, R2)]
    h, w = ims[0].shape
    with tempfile.TemporaryDirectory() as td:
        vp = os.path.join(td, "voc.txt")
        write_vocabulary_text(voc, vp, 0, 0)
        pb = vp.encode()
        pay = (np.array([w, h, nfeatures], np.int32).tobytes() + _calib_bytes(cam, bf) + np.array([pct1, pct2], np.int32).tobytes() +
               np.array([ratio_kk, ratio_kf, th_proj, th_last], np.float32).tobytes() +
               np.array([int(check_ori), int(only_stereo), int(mono), int(reps)], np.int32).tobytes() +
               np.asarray(T1, np.float32).reshape(16).tobytes() + np.asarray(T2, np.float32).reshape(16).tobytes() +
               np.asarray(F12, np.float32).reshape(9).tobytes() + np.int32(len(pb)).tobytes() + pb + b"".join(x.tobytes() for x in ims))
        fin, fout = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        with open(fin, "wb") as f:
            f.write(pay)
        res = subprocess.run(list(prefix) + [binary or REF_SLAM_BIN, "kfbench", fin, fout], check=True, capture_output=True, text=True)
    if os.environ.get("ORBX_KFBENCH_STDOUT"):                       # experiments: device-side printf of an instrumented build
        sys.stderr.write(res.stdout)
    out, sizes = {}, None
    for ln in res.stdout.strip().splitlines():
        t = ln.split()
        if t[0] == "sizes":
            sizes = tuple(int(v) for v in t[1:])
        elif len(t) == 4:
            out[t[0]] = (float(t[1]), float(t[2]), int(t[3]))
    return out, sizes


def ref_slam_distinct(descriptor_lists, image, nfeatures=2000, cam=(700.0, 700.0, 600.0, 180.0)):
    """The REFERENCE's MapPoint::ComputeDistinctiveDescriptors on real MapPoints / KeyFrames: descriptor_lists[m] = (N_m, 32)
    observed descriptors of map point m (N_m >= 1).  Returns the (M, 32) chosen descriptors."""
    image = np.ascontiguousarray(image, np.uint8); h, w = image.shape
    M = len(descriptor_lists); nmax = max(len(d) for d in descriptor_lists)
    desc = np.zeros((M, nmax, 32), np.uint8); counts = np.zeros(M, np.int32)
    for m, d in enumerate(descriptor_lists):
        counts[m] = len(d); desc[m, :len(d)] = d
    pay = (np.array([w, h, nfeatures], np.int32).tobytes() + _calib_bytes(cam, 40.0) + np.array([M, nmax], np.int32).tobytes() +
           counts.tobytes() + desc.tobytes() + image.tobytes())
    rd = _Rd(_run_ref_slam("distinct", pay))
    out = np.zeros((M, 32), np.uint8)
    for m in range(M):
        rd.i32(); out[m] = rd.arr(np.uint8, 32)
    return out


REF_BENCH_MODES = {"stereo": 0, "mono": 1, "init": 2, "bow": 3}


def ref_slam_bench(imgsL, imgsR, nfeatures, cam, bf, threads, reps=1, mode="stereo", ratio=0.9, window=100, check_ori=True, binary=None,
                   detail=False):
    """Times the REFERENCE's own code on `threads` std::threads, one unit per thread at a time (oracle/ref_slam_cmds.inc:cmd_bench):
    mode "stereo" = the stereo Frame constructor (two extractor threads + ComputeStereoMatches), "mono" = the monocular Frame
    constructor (imgsR None), "init" = ORBmatcher::SearchForInitialization on (imgsL[i], imgsR[i]) frame pairs, "bow" =
    ORBmatcher::SearchByBoW(KF,KF) with one vocabulary bucket on such pairs.  binary: another build of the driver (the drop-in
    ones).  Returns (seconds, keypoints, matches), with detail=True also (p50_ms, p99_ms) of one call."""
    import tempfile
    imgsL = np.ascontiguousarray(imgsL, np.uint8)
    n, h, w = imgsL.shape
    m = REF_BENCH_MODES[mode]
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        with open(fin, "wb") as f:
            f.write(np.array([w, h, nfeatures], np.int32).tobytes() + _calib_bytes(cam, bf) + np.array([n, threads, reps, m], np.int32).tobytes())
            if m >= 2:
                f.write(np.float32(ratio).tobytes() + np.array([window, int(check_ori)], np.int32).tobytes())
            f.write(imgsL.tobytes())
            if m != 1:
                f.write(np.ascontiguousarray(imgsR, np.uint8).tobytes())
        res = subprocess.run([binary or REF_SLAM_BIN, "bench", fin, fout], check=True, capture_output=True, text=True)
    tok = res.stdout.strip().splitlines()[-1].split()
    out = (float(tok[0]), int(tok[1]), int(tok[2]))
    return out + ((float(tok[3]), float(tok[4])),) if detail else out


def ref_dbow_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "orbref_dbow"))


def write_vocabulary_text(voc, path, scoring, weighting):
    """The reference's text vocabulary format (TemplatedVocabulary::loadFromTextFile, TemplatedVocabulary.h:1337-1420):
    'k L scoring weighting', then one line per node in id order: parent, isLeaf, 32 descriptor bytes, weight.
    No trailing newline (the loader would read it as one more node)."""
    n = len(voc["desc"]); parent = np.zeros(n, np.int64)
    for p in range(n):
        for c in range(voc["child_count"][p]):
            parent[voc["children"][voc["child_first"][p] + c]] = p
    lines = [f'{voc["k"]} {voc["L"]} {scoring} {weighting}']
    for i in range(1, n):
        lines.append(f'{parent[i]} {1 if voc["child_count"][i] == 0 else 0} ' + " ".join(str(int(b)) for b in voc["desc"][i]) +
                     f' {voc["weight"][i]:.17g}')
    with open(path, "w") as f:
        f.write("\n".join(lines))


def ref_dbow_transform(voc, desc, levelsup, scoring, weighting):
    """Runs the REFERENCE's DBoW2 (oracle/_ref/orbref_dbow) -> (node per feature, BowVector words, BowVector values)."""
    import tempfile
    desc = np.ascontiguousarray(desc, np.uint8)
    with tempfile.TemporaryDirectory() as td:
        vp, dp = os.path.join(td, "voc.txt"), os.path.join(td, "desc.bin")
        write_vocabulary_text(voc, vp, scoring, weighting); desc.tofile(dp)
        out = subprocess.run([os.path.join(_HERE, "_ref", "orbref_dbow"), vp, dp, str(levelsup)], capture_output=True, text=True, check=True)
    bw, bv = [], []
    node = np.full(len(desc), -1, np.int32)
    for line in out.stdout.splitlines():
        t = line.split()
        if t[0] == "B":
            bw.append(int(t[1])); bv.append(float.fromhex(t[2]))
        elif t[0] == "V":
            idx = [int(x) for x in t[2:]]
            assert idx == sorted(idx)
            node[idx] = int(t[1])
    return node, np.array(bw, np.int32), np.array(bv, np.float64)


FUSE_MODE = 2 | 4 | 8            # Fuse(pKF, vpMapPoints, th): distance from the camera centre, viewing angle, chi2 gate, 1/z in float
SIM3_MODE = 1 | 16               # SearchBySim3 passes: two transforms, distance in the target camera, 1.0/z in double


def project_search(k, d, uright, kf, scale, inv_sigma2, xf, mp, th, mode, th_dist):
    """Fuse / SearchBySim3 search part.  kf = (fx, fy, cx, cy, bf, logScale, minX, maxX, minY, maxY); xf = dict(R1, t1[, R2, t2][, Ow]);
    mp = dict(valid, xw (m,3), normal (m,3), dmin, dmax, desc (m,32)).  Returns the best keypoint per map point (-1 = none)."""
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    k = np.ascontiguousarray(k); d = np.ascontiguousarray(d, np.uint8); m = len(mp["dmin"])
    x = np.concatenate([f32(xf["R1"]).reshape(9), f32(xf["t1"]).reshape(3), f32(xf.get("R2", np.eye(3))).reshape(9),
                        f32(xf.get("t2", np.zeros(3))).reshape(3), f32(xf.get("Ow", np.zeros(3))).reshape(3)])
    best = np.zeros(max(m, 1), np.int32)
    L = lib(); L.orbo_project_search.restype = None
    L.orbo_project_search(_p(k), _p(d), _p(f32(uright)), len(k), _p(f32(kf)), _p(f32(scale)), _p(f32(inv_sigma2)), len(scale), _p(x),
                          _p(np.ascontiguousarray(mp["valid"], np.uint8)), _p(f32(mp["xw"])), _p(f32(mp["normal"])), _p(f32(mp["dmin"])),
                          _p(f32(mp["dmax"])), _p(np.ascontiguousarray(mp["desc"], np.uint8)), m, C.c_float(th), int(mode), int(th_dist), _p(best))
    return best[:m]


def project_match_ordered(k, d, kf, grid_bounds, scale, reloc, Rcw, tcw, Ow, mp, th, th_dist, check_ori, occupied):
    """SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (reloc=False) / SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th,
    ORBdist) (reloc=True).  mp = dict(valid, xw, normal, dmin, dmax, desc, angle).  Returns (assigned per keypoint, nmatches)."""
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    k = np.ascontiguousarray(k); d = np.ascontiguousarray(d, np.uint8); m = len(mp["dmin"])
    xf = np.concatenate([f32(Rcw).reshape(9), f32(tcw).reshape(3), f32(Ow).reshape(3)])
    occ = np.ascontiguousarray(occupied, np.uint8).copy(); assigned = np.zeros(max(len(k), 1), np.int32)
    n = lib().orbo_project_match_ordered(_p(k), _p(d), len(k), _p(f32(kf)), _p(f32(grid_bounds)), _p(f32(scale)), len(scale), int(reloc), _p(xf),
                                         _p(np.ascontiguousarray(mp["valid"], np.uint8)), _p(f32(mp["xw"])), _p(f32(mp["normal"])), _p(f32(mp["dmin"])),
                                         _p(f32(mp["dmax"])), _p(np.ascontiguousarray(mp["desc"], np.uint8)), _p(f32(mp["angle"])), m, C.c_float(th),
                                         int(th_dist), int(check_ori), _p(occ), _p(assigned))
    return assigned[:len(k)], n


def distinctive_descriptor(desc):
    """MapPoint::ComputeDistinctiveDescriptors: index of the descriptor with the least median distance."""
    desc = np.ascontiguousarray(desc, np.uint8)
    return int(lib().orbo_distinctive_descriptor(_p(desc), len(desc)))


def _cam(cam):
    """cam: (fx, fy, cx, cy, k1, k2, p1, p2, k3) -> float32[9]"""
    c = np.zeros(9, np.float32); c[:len(cam)] = np.asarray(cam, np.float32)
    return c


def undistort_points(cam, xy):
    """cv::undistortPoints(xy, K, distCoef, noArray(), K) restated (src/Frame.cc:423)."""
    xy = np.ascontiguousarray(xy, np.float32); out = np.empty_like(xy); c = _cam(cam)
    lib().orbo_undistort_points(_p(c), _p(xy), len(xy), _p(out))
    return out


def undistort_grid(cam, k, cols, rows):
    """Frame::UndistortKeyPoints + ComputeImageBounds + AssignFeaturesToGrid.
    Returns (mvKeysUn, bounds[6] = minX,maxX,minY,maxY,invW,invH, cell_start[3073], items)."""
    k = np.ascontiguousarray(k); n = len(k); c = _cam(cam)
    kun = np.zeros(n, KP_DTYPE); bounds = np.zeros(6, np.float32)
    cs = np.zeros(64 * 48 + 1, np.int32); items = np.zeros(max(n, 1), np.int32)
    lib().orbo_undistort_grid(_p(c), _p(k), n, cols, rows, _p(kun), _p(bounds), _p(cs), _p(items))
    return kun, bounds, cs, items[:cs[-1]]


def features_in_area(kun, bounds, x, y, r, min_level=-1, max_level=-1):
    """Frame::GetFeaturesInArea (src/Frame.cc:327-380) -> indices in the reference's order."""
    kun = np.ascontiguousarray(kun); out = np.zeros(max(len(kun), 1), np.int32)
    L = lib(); L.orbo_features_in_area.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_int]
    n = L.orbo_features_in_area(_p(kun), len(kun), _p(np.ascontiguousarray(bounds, np.float32)), x, y, r, min_level, max_level, _p(out), len(out))
    return out[:n].copy()


def search_projection(k, d, uright, cols, rows, scale, mp, th=1.0, nnratio=0.8, frame_obs=None, bounds=None):
    """mp: dict of arrays inview(u8), projx, projy, projxr (f32), level (i32), viewcos (f32), desc (m,32), obs (i32)."""
    k = np.ascontiguousarray(k); d = np.ascontiguousarray(d); n = len(k); m = len(mp["level"])
    uright = np.ascontiguousarray(uright, np.float32); scale = np.ascontiguousarray(scale, np.float32)
    fo = np.full(n, -1, np.int32) if frame_obs is None else np.ascontiguousarray(frame_obs, np.int32).copy()
    assigned = np.zeros(n, np.int32)
    a = {key: np.ascontiguousarray(mp[key], dt) for key, dt in [("inview", np.uint8), ("projx", np.float32), ("projy", np.float32),
         ("projxr", np.float32), ("level", np.int32), ("viewcos", np.float32), ("desc", np.uint8), ("obs", np.int32)]}
    b = _bounds4(cols, rows, bounds)
    nm = lib().orbo_search_projection(_p(k), _p(d), _p(uright), n, _p(b), _p(scale), _p(a["inview"]), _p(a["projx"]), _p(a["projy"]),
                                      _p(a["projxr"]), _p(a["level"]), _p(a["viewcos"]), _p(a["desc"]), _p(a["obs"]), m,
                                      C.c_float(th), C.c_float(nnratio), _p(fo), _p(assigned))
    return assigned, nm, fo


def search_projection_frame(k, d, uright, cols, rows, scale, Rcw, tcw, cam, last, th=15.0, mode=0, check_ori=True, frame_obs=None, bounds=None):
    """cam = (fx, fy, cx, cy, bf); last: dict valid(u8), xw (n,3 f32), octave (i32), angle (f32), desc (n,32), obs (i32)."""
    k = np.ascontiguousarray(k); d = np.ascontiguousarray(d); n = len(k); nl = len(last["octave"])
    uright = np.ascontiguousarray(uright, np.float32); scale = np.ascontiguousarray(scale, np.float32)
    Rcw = np.ascontiguousarray(Rcw, np.float32); tcw = np.ascontiguousarray(tcw, np.float32)
    fo = np.full(n, -1, np.int32) if frame_obs is None else np.ascontiguousarray(frame_obs, np.int32).copy()
    assigned = np.zeros(n, np.int32)
    a = {key: np.ascontiguousarray(last[key], dt) for key, dt in [("valid", np.uint8), ("xw", np.float32), ("octave", np.int32),
         ("angle", np.float32), ("desc", np.uint8), ("obs", np.int32)]}
    fx, fy, cx, cy, bf = (C.c_float(v) for v in cam)
    b = _bounds4(cols, rows, bounds)
    nm = lib().orbo_search_projection_frame(_p(k), _p(d), _p(uright), n, _p(b), _p(scale), _p(Rcw), _p(tcw), fx, fy, cx, cy, bf,
                                            _p(a["valid"]), _p(a["xw"]), _p(a["octave"]), _p(a["angle"]), _p(a["desc"]), _p(a["obs"]), nl,
                                            C.c_float(th), mode, int(check_ori), _p(fo), _p(assigned))
    return assigned, nm, fo


def project_points(Rcw, tcw, xw):
    """The oracle's camera-frame transform (for pinning against cv2.gemm)."""
    Rcw = np.ascontiguousarray(Rcw, np.float32); tcw = np.ascontiguousarray(tcw, np.float32); xw = np.ascontiguousarray(xw, np.float32)
    out = np.zeros_like(xw)
    lib().orbo_transform_points(_p(Rcw), _p(tcw), _p(xw), _p(out), len(xw))
    return out


def resize(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((dh, dw), np.uint8)
    lib().orbo_resize(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def border(src, b=19):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((src.shape[0] + 2 * b, src.shape[1] + 2 * b), np.uint8)
    lib().orbo_border(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dst.shape[1], b)
    return dst


def blur(src):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros_like(src)
    lib().orbo_blur(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dst.strides[0])
    return dst


def fast(src, t, cap=100000):
    src = np.ascontiguousarray(src, np.uint8)
    out = np.zeros((cap, 3), np.int32)
    n = lib().orbo_fast(_p(src), src.shape[1], src.shape[0], src.strides[0], t, _p(out), cap)
    return out[:n].copy()


def atan2_array(y, x):
    y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(y)
    lib().orbo_atan2_array(_p(y), _p(x), _p(out), len(y))
    return out


def sincos_array(a, libm=False):
    a = np.ascontiguousarray(a, np.float32)
    s = np.zeros_like(a); c = np.zeros_like(a)
    (lib().orbo_libm_sincos_array if libm else lib().orbo_sincos_array)(_p(a), _p(s), _p(c), len(a))
    return s, c


def quadtree(cand, minX, maxX, minY, maxY, N, cap=20000):
    cand = np.ascontiguousarray(cand)
    out = np.zeros(cap, KP_DTYPE)
    n = lib().orbo_quadtree(_p(cand), len(cand), minX, maxX, minY, maxY, N, _p(out), cap)
    if n < 0:
        raise RuntimeError("degenerate quadtree geometry")
    return out[:n].copy()


def hamming256(a, b):
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    return lib().orbo_hamming256(_p(a), _p(b))


def bench_stereo(imgsL, imgsR, params, bf, fx, threads, reps=1):
    """Times extraction (+ stereo matching when imgsR is given) on the host; returns
    (seconds, total keypoints, total stereo matches)."""
    imgsL = np.ascontiguousarray(imgsL, np.uint8)
    n, h, w = imgsL.shape
    if imgsR is not None:
        imgsR = np.ascontiguousarray(imgsR, np.uint8)
    k = C.c_longlong(); m = C.c_longlong()
    nf, sf, nl, ini, mn = params
    s = lib().orbo_bench_stereo(_p(imgsL), _p(imgsR), n, w, h, nf, C.c_float(sf), nl, ini, mn, C.c_float(bf),
                                C.c_float(fx), threads, reps, C.byref(k), C.byref(m))
    return s, k.value, m.value


def batch_stereo(imgsL, imgsR, params, bf, fx, cap, threads=0):
    """Extraction (+ stereo matching when imgsR is given) of a batch on all host threads, full results in the layout of the
    CUDA path's batched fetch: kps [(1 or 2) * n][cap] (left eyes, then right eyes), desc, counts; uRight / depth [n][cap]
    (-1 filled), n_matches [n]."""
    imgsL = np.ascontiguousarray(imgsL, np.uint8)
    n, h, w = imgsL.shape
    eyes = 1 if imgsR is None else 2
    if imgsR is not None:
        imgsR = np.ascontiguousarray(imgsR, np.uint8)
    kps = np.zeros((eyes * n, cap), KP_DTYPE); desc = np.zeros((eyes * n, cap, 32), np.uint8); counts = np.zeros(eyes * n, np.int32)
    ur = np.full((n, cap), -1, np.float32); dp = np.full((n, cap), -1, np.float32); nm = np.zeros(n, np.int32)
    nf, sf, nl, ini, mn = params
    rc = lib().orbo_batch_stereo(_p(imgsL), _p(imgsR), n, w, h, nf, C.c_float(sf), nl, ini, mn, C.c_float(bf), C.c_float(fx),
                                 threads or hardware_threads(), cap, _p(kps), _p(desc), _p(counts), _p(ur), _p(dp), _p(nm))
    if rc != 0:
        raise RuntimeError(f"orbo_batch_stereo failed: {rc}")
    return kps, desc, counts, ur, dp, nm


def hardware_threads():
    return lib().orbo_hardware_threads()


# --- oracle/_ref: the reference's own ORBextractor.cc (built by `make -C oracle ref`) -----------
REF_BIN = os.path.join(_HERE, "_ref", "orbref")


def ref_available():
    return os.path.exists(REF_BIN)


def ref_extract(imgs, nfeatures, scale=1.2, nlevels=8, ini=20, mn=7, reps=1, use_malloc=False):
    """Runs the reference's ORBextractor::operator() (separate process) on a stack of equally
    sized images; returns ([(kps, desc), ...], seconds inside operator())."""
    import tempfile
    imgs = np.ascontiguousarray(imgs, np.uint8)
    if imgs.ndim == 2:
        imgs = imgs[None]
    n, h, w = imgs.shape
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        with open(fin, "wb") as f:
            f.write(np.array([w, h, n], np.int32).tobytes())
            f.write(imgs.tobytes())
        cmd = [REF_BIN, "extract", fin, fout, str(nfeatures), repr(float(scale)), str(nlevels), str(ini), str(mn),
               str(reps)] + (["malloc"] if use_malloc else [])
        res = subprocess.run(cmd, check=True, capture_output=True, text=True)
        raw = open(fout, "rb").read()
    out, off = [], 0
    for _ in range(n):
        cnt = int(np.frombuffer(raw, np.int32, 1, off)[0]); off += 4
        kps = np.frombuffer(raw, KP_DTYPE, cnt, off).copy(); off += 28 * cnt
        desc = np.frombuffer(raw, np.uint8, 32 * cnt, off).reshape(cnt, 32).copy(); off += 32 * cnt
        out.append((kps, desc))
    return out, float(res.stdout.strip().splitlines()[-1])
