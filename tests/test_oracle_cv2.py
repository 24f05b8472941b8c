"""Pins the oracle's OpenCV-primitive restatements against the real OpenCV on the box
(cv2 4.13.0).  The reference itself has no tests or golden
vectors, so these comparisons are what anchors the oracle."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
from orb_slam2_b200 import synth

SHAPES = [(1241, 376), (752, 480), (640, 480), (331, 97)]


def _img(i, w, h):
    return synth.synth_image(1000 + i, w, h)


def test_cv2_version_recorded():
    # H5: the Gaussian fixed-point path is version-sensitive; the oracle is pinned to 4.x.
    assert int(cv2.__version__.split(".")[0]) >= 4


@pytest.mark.parametrize("w,h", SHAPES)
def test_resize_matches_cv2(oracle, w, h):
    img = _img(0, w, h)
    for s in (1.2, 1.2 ** 2, 1.7):
        dw, dh = int(round(w / s)), int(round(h / s))
        ref = cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(oracle.resize(img, dw, dh), ref)


@pytest.mark.parametrize("w,h", SHAPES)
def test_border_matches_cv2(oracle, w, h):
    img = _img(1, w, h)
    ref = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    assert np.array_equal(oracle.border(img, 19), ref)


@pytest.mark.parametrize("w,h", SHAPES)
def test_blur_matches_cv2(oracle, w, h):
    img = _img(2, w, h)
    ref = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    assert np.array_equal(oracle.blur(img), ref)
    rnd = np.random.default_rng(5).integers(0, 256, (h, w), dtype=np.uint8)
    ref = cv2.GaussianBlur(rnd, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    assert np.array_equal(oracle.blur(rnd), ref)


def _cv_fast(img, t):
    det = cv2.FastFeatureDetector_create(threshold=t, nonmaxSuppression=True,
                                         type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    kps = det.detect(img, None)
    return np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kps], np.int32).reshape(-1, 3)


@pytest.mark.parametrize("t", [20, 12, 7])
def test_fast_matches_cv2_on_cells(oracle, t):
    img = _img(3, 640, 480)
    rng = np.random.default_rng(t)
    total = 0
    for _ in range(150):
        ww, hh = int(rng.integers(5, 44)), int(rng.integers(5, 47))
        x0, y0 = int(rng.integers(0, 640 - ww)), int(rng.integers(0, 480 - hh))
        win = np.ascontiguousarray(img[y0:y0 + hh, x0:x0 + ww])
        ours, ref = oracle.fast(win, t), _cv_fast(win, t)
        assert ours.shape == ref.shape and np.array_equal(ours, ref), (ww, hh, x0, y0)
        total += len(ref)
    assert total > 100


def test_fast_matches_cv2_on_noise_and_full_frame(oracle):
    rnd = np.random.default_rng(11).integers(0, 256, (120, 160), dtype=np.uint8)
    for t in (7, 20, 40):
        assert np.array_equal(oracle.fast(rnd, t), _cv_fast(rnd, t))
    img = _img(4, 752, 480)
    assert np.array_equal(oracle.fast(img, 20), _cv_fast(img, 20))


def test_fast_atan2_matches_cv2_scalar(oracle):
    rng = np.random.default_rng(3)
    y = rng.integers(-2_800_000, 2_800_000, 20000).astype(np.float32)
    x = rng.integers(-2_800_000, 2_800_000, 20000).astype(np.float32)
    y[:50] = 0; x[50:100] = 0; y[100:110] = 0; x[100:110] = 0
    ours = oracle.atan2_array(y, x)
    ref = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    assert np.array_equal(ours.view(np.uint32), ref.view(np.uint32))
    assert oracle.lib().orbo_atan2(0.0, 0.0) == 0.0


def test_sincos_port_equals_host_libm(oracle):
    # Exhaustive proof: tools/sweep_sincos.c (1 087 050 388 floats, 0 mismatches).  Here: a
    # dense sample of [0, 2*pi] plus every float reachable as fastAtan2(int,int)*(pi/180).
    a = np.linspace(0, 6.2831855, 2_000_001, dtype=np.float32)
    s, c = oracle.sincos_array(a)
    s2, c2 = oracle.sincos_array(a, libm=True)
    assert np.array_equal(s.view(np.uint32), s2.view(np.uint32))
    assert np.array_equal(c.view(np.uint32), c2.view(np.uint32))
    rng = np.random.default_rng(9)
    deg = oracle.atan2_array(rng.integers(-50000, 50000, 500000).astype(np.float32),
                             rng.integers(-50000, 50000, 500000).astype(np.float32))
    rad = deg * np.float32(np.float32(np.pi) / np.float32(180.0))
    s, c = oracle.sincos_array(rad)
    s2, c2 = oracle.sincos_array(rad, libm=True)
    assert np.array_equal(s.view(np.uint32), s2.view(np.uint32))
    assert np.array_equal(c.view(np.uint32), c2.view(np.uint32))


def _cv_pyramid_and_candidates(img, nlevels, inv, ini=20, mn=7):
    """ComputePyramid + the per-cell FAST loop of ComputeKeyPointsOctTree written with real cv2
    calls (src/ORBextractor.cc:1107-1132, 765-830)."""
    h, w = img.shape
    levels, cands = [], []
    for l in range(nlevels):
        sw = int(np.rint(np.float32(w) * inv[l])); sh = int(np.rint(np.float32(h) * inv[l]))
        if l == 0:
            cur = img
        else:
            cur = cv2.resize(levels[-1][19:-19, 19:-19], (sw, sh), interpolation=cv2.INTER_LINEAR)
        levels.append(cv2.copyMakeBorder(cur, 19, 19, 19, 19, cv2.BORDER_REFLECT_101))
        minB, maxBX, maxBY = 16, sw - 16, sh - 16
        width, height = np.float32(maxBX - minB), np.float32(maxBY - minB)
        nC, nR = int(width / np.float32(30)), int(height / np.float32(30))
        wC, hC = int(np.ceil(width / np.float32(nC))), int(np.ceil(height / np.float32(nR)))
        out = []
        for i in range(nR):
            iy = minB + i * hC; my = iy + hC + 6
            if iy >= maxBY - 3:
                continue
            my = min(my, maxBY)
            for j in range(nC):
                ix = minB + j * wC; mx = ix + wC + 6
                if ix >= maxBX - 6:
                    continue
                mx = min(mx, maxBX)
                win = np.ascontiguousarray(cur[iy:my, ix:mx])
                k = _cv_fast(win, ini)
                if len(k) == 0:
                    k = _cv_fast(win, mn)
                for x, y, r in k:
                    out.append((x + j * wC, y + i * hC, r))
        cands.append(np.array(out, np.int32).reshape(-1, 3))
    return levels, cands


@pytest.mark.parametrize("w,h,nf", [(1241, 376, 2000), (752, 480, 1200), (640, 480, 1000)])
def test_pyramid_and_candidates_match_cv2_pipeline(oracle, w, h, nf):
    img = synth.synth_image(77, w, h)
    E = oracle.Extractor(nf)
    E.extract(img)
    inv = E.tables()["inv"]
    levels, cands = _cv_pyramid_and_candidates(img, 8, inv)
    fallback = 0
    for l in range(8):
        assert np.array_equal(E.level(l), levels[l]), f"pyramid level {l}"
        c = E.candidates(l)
        ours = np.stack([c["x"], c["y"], c["response"]], 1).astype(np.int32)
        assert np.array_equal(ours, cands[l]), f"candidates level {l}"
        assert len(c) > E.tables()["quota"][l]          # the careful quad-tree phase is exercised
        fallback += int((c["response"] < 20).sum())
    assert fallback > 0                                   # the minThFAST retry is exercised


def test_camera_transform_matches_cv2_gemm(oracle):
    """SearchByProjection computes x3Dc = Rcw*x3Dw + tcw with cv::Mat operators (src/ORBmatcher.cc:1362),
    i.e. cv::gemm on a 3x3 and a 3x1 CV_32F matrix.  Pin the oracle's arithmetic (float products and
    sums, left to right, then + t) against the real gemm."""
    rng = np.random.default_rng(4)
    for _ in range(40):
        R = rng.normal(0, 1, (3, 3)).astype(np.float32)
        t = rng.normal(0, 5, (3, 1)).astype(np.float32)
        X = rng.normal(0, 30, (500, 3)).astype(np.float32)
        ours = oracle.project_points(R, t.ravel(), X)
        ref = np.stack([cv2.gemm(R, X[i].reshape(3, 1), 1.0, t, 1.0).ravel() for i in range(len(X))])
        assert np.array_equal(ours.view(np.uint32), ref.view(np.uint32))


TUM1 = (517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
TUM2 = (520.908620, 521.007327, 325.141442, 249.701764, 0.231222, -0.784899, -0.003257, -0.000105, 0.917205)


@pytest.mark.parametrize("cam", [TUM1, TUM2, (458.654, 457.296, 367.215, 248.375, -0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0)])
def test_undistort_points_matches_cv2(oracle, cam):
    O = oracle
    """Frame::UndistortKeyPoints / ComputeImageBounds call cv::undistortPoints(mat, mat, mK, mDistCoef, Mat(), mK)
    (src/Frame.cc:423, :449): the restatement must give the same float32 bits, corners and far-out points included."""
    rng = np.random.default_rng(11)
    pts = np.stack([rng.uniform(-40, 800, 40000), rng.uniform(-40, 560, 40000)], 1).astype(np.float32)
    pts[:4] = [[0, 0], [640, 0], [0, 480], [640, 480]]
    K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
    D = np.array(cam[4:], np.float32)
    ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    got = O.undistort_points(cam, pts)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_image_bounds_and_grid_restatement(oracle):
    O = oracle
    """bounds from the four undistorted corners (src/Frame.cc:436-464); the grid puts every keypoint in cell
    round((x - minX) * invW), round((y - minY) * invH), inside a cell in index order (src/Frame.cc:230-245)."""
    rng = np.random.default_rng(5)
    n = 1500
    k = np.zeros(n, O.KP_DTYPE)
    k["x"] = rng.uniform(19, 621, n).astype(np.float32); k["y"] = rng.uniform(19, 461, n).astype(np.float32)
    k["octave"] = rng.integers(0, 8, n)
    kun, b, cs, items = O.undistort_grid(TUM1, k, 640, 480)
    K = np.array([[TUM1[0], 0, TUM1[2]], [0, TUM1[1], TUM1[3]], [0, 0, 1]], np.float32)
    corners = cv2.undistortPoints(np.array([[[0, 0]], [[640, 0]], [[0, 480]], [[640, 480]]], np.float32), K, np.array(TUM1[4:], np.float32), None, K).reshape(4, 2)
    assert b[0] == min(corners[0, 0], corners[2, 0]) and b[1] == max(corners[1, 0], corners[3, 0])
    assert b[2] == min(corners[0, 1], corners[1, 1]) and b[3] == max(corners[2, 1], corners[3, 1])
    assert b[4] == np.float32(64) / (b[1] - b[0]) and b[5] == np.float32(48) / (b[3] - b[2])
    gx = np.floor((kun["x"] - b[0]) * b[4] + np.float32(0.5)).astype(int); gy = np.floor((kun["y"] - b[2]) * b[5] + np.float32(0.5)).astype(int)
    ok = (gx >= 0) & (gx < 64) & (gy >= 0) & (gy < 48)
    assert cs[-1] == ok.sum() == len(items)
    for c in rng.integers(0, 64 * 48, 200):
        want = np.nonzero(ok & (gx * 48 + gy == c))[0]
        assert np.array_equal(items[cs[c]:cs[c + 1]], want)
    # GetFeaturesInArea: brute force over the same predicate, order = cell column, cell row, index
    for _ in range(50):
        x, y, r = rng.uniform(0, 640), rng.uniform(0, 480), rng.uniform(5, 120)
        lo, hi = (-1, -1) if rng.random() < 0.5 else (int(rng.integers(0, 4)), int(rng.integers(3, 8)))
        got = O.features_in_area(kun, b, x, y, r, lo, hi)
        x32, y32, r32 = np.float32(x), np.float32(y), np.float32(r)
        m = ok & (np.abs(kun["x"] - x32) < r32) & (np.abs(kun["y"] - y32) < r32)
        if lo > 0 or hi >= 0:
            m &= kun["octave"] >= lo
            if hi >= 0: m &= kun["octave"] <= hi
        cx0 = max(0, int(np.floor((x32 - b[0] - r32) * b[4]))); cx1 = min(63, int(np.ceil((x32 - b[0] + r32) * b[4])))
        cy0 = max(0, int(np.floor((y32 - b[2] - r32) * b[5]))); cy1 = min(47, int(np.ceil((y32 - b[2] + r32) * b[5])))
        m &= (gx >= cx0) & (gx <= cx1) & (gy >= cy0) & (gy <= cy1)
        idx = np.nonzero(m)[0]
        want = idx[np.lexsort((idx, gy[idx], gx[idx]))]
        assert np.array_equal(got, want)
