"""Seeded synthetic inputs for the ORB front end (host side, numpy only).

The reference ships no images; every workload in BASELINE.json is "X-shaped
synthetic frames".  This generator is shared by the parity tests, the CPU baseline and bench.py
so that all of them see byte-identical inputs for a given (seed, shape).

Recipe per frame: flat-shaded rectangles and discs painted over mid grey (corner sources for
FAST at several scales) mixed 3:1 with low-pass noise (texture that only survives the
minThFAST fallback).  The right eye is the left one shifted by an integer disparity per 47-row
band plus +-2 grey levels of noise, so stereo SAD refinement and the median cut do real work.
"""
import numpy as np

_GOLD = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
BASE_SEED = 0x0B5EED00

KITTI = dict(w=1241, h=376, nfeatures=2000, bf=386.1448, fx=718.856)   # Examples/Stereo/KITTI00-02.yaml
EUROC = dict(w=752, h=480, nfeatures=1200, bf=47.90639, fx=435.2047)   # Examples/Stereo/EuRoC.yaml
TUM = dict(w=640, h=480, nfeatures=1000)                               # Examples/Monocular/TUM1.yaml


def splitmix64(seed, n, offset=0):
    """n outputs of the splitmix64 sequence started at `seed`, skipping `offset` (counter form)."""
    with np.errstate(over="ignore"):
        idx = np.arange(offset + 1, offset + n + 1, dtype=np.uint64)
        z = np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + idx * _GOLD
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


class _Stream:
    def __init__(self, seed):
        self.seed, self.pos = seed, 0

    def take(self, n):
        out = splitmix64(self.seed, n, self.pos)
        self.pos += n
        return out

    def ints(self, n, lo, hi):
        """n integers uniform in [lo, hi] (modulo bias is irrelevant here)."""
        return (self.take(n) % np.uint64(hi - lo + 1)).astype(np.int64) + lo

    def bytes(self, shape):
        n = int(np.prod(shape))
        words = self.take((n + 7) // 8)
        return words.view(np.uint8)[:n].reshape(shape)


def _blur7(img):
    """7x7 sigma-2 fixed-point blur with reflect-101 borders (taps 18,34,48,56,48,34,18 / 256)."""
    k = np.array([18, 34, 48, 56, 48, 34, 18], dtype=np.uint32)
    p = np.pad(img.astype(np.uint32), ((0, 0), (3, 3)), mode="reflect")
    hp = sum(k[i] * p[:, i:i + img.shape[1]] for i in range(7))
    p = np.pad(hp, ((3, 3), (0, 0)), mode="reflect")
    v = sum(k[i] * p[i:i + img.shape[0], :] for i in range(7))
    return ((v + 32768) >> 16).astype(np.uint8)


def synth_image(seed, w, h, nrect=1200, ndisc=400):
    s = _Stream(seed)
    img = np.full((h, w), 128, dtype=np.int32)
    rx = s.ints(nrect, 0, w - 1); ry = s.ints(nrect, 0, h - 1)
    rw = s.ints(nrect, 4, 70); rh = s.ints(nrect, 4, 50); rg = s.ints(nrect, 16, 240)
    cx = s.ints(ndisc, 0, w - 1); cy = s.ints(ndisc, 0, h - 1)
    cr = s.ints(ndisc, 2, 20); cg = s.ints(ndisc, 16, 240)
    for i in range(max(nrect, ndisc)):
        if i < nrect:
            img[ry[i]:ry[i] + rh[i], rx[i]:rx[i] + rw[i]] = rg[i]
        if i < ndisc:
            r = int(cr[i]); x0 = max(0, cx[i] - r); x1 = min(w, cx[i] + r + 1)
            y0 = max(0, cy[i] - r); y1 = min(h, cy[i] + r + 1)
            yy, xx = np.ogrid[y0:y1, x0:x1]
            m = (xx - cx[i]) ** 2 + (yy - cy[i]) ** 2 <= r * r
            img[y0:y1, x0:x1][m] = cg[i]
    noise = _blur7(s.bytes((h, w))).astype(np.int32)
    return np.clip((3 * img + noise + 2) >> 2, 0, 255).astype(np.uint8)


def _add_noise(img, seed, amp=2):
    s = _Stream(seed)
    n = (s.bytes(img.shape).astype(np.int32) % (2 * amp + 1)) - amp
    return np.ascontiguousarray(np.clip(img.astype(np.int32) + n, 0, 255).astype(np.uint8))


def stereo_pair(frame_idx, w, h, band=47, dmin=2, dmax=64):
    """(left, right, per-row planted disparity) for frame `frame_idx`."""
    seed = BASE_SEED + 16 * frame_idx
    left = synth_image(seed, w, h)
    nb = (h + band - 1) // band
    d = _Stream(seed + 2).ints(nb, dmin, dmax)
    rowd = np.repeat(d, band)[:h]
    cols = np.minimum(np.arange(w)[None, :] + rowd[:, None], w - 1)
    right = np.take_along_axis(left, cols, axis=1)
    return left, _add_noise(right, seed + 1), rowd


def mono_pair(frame_idx, w, h, maxshift=40):
    """(first, second, (dx, dy)): second(x, y) = first(x - dx, y - dy) + noise, edges clamped."""
    seed = BASE_SEED + 16 * frame_idx + 8
    a = synth_image(seed, w, h)
    dx, dy = (int(v) for v in _Stream(seed + 2).ints(2, -maxshift, maxshift))
    ys = np.clip(np.arange(h) - dy, 0, h - 1); xs = np.clip(np.arange(w) - dx, 0, w - 1)
    b = np.ascontiguousarray(a[ys][:, xs])
    return a, _add_noise(b, seed + 1), (dx, dy)


def mono_frame(frame_idx, w, h):
    return synth_image(BASE_SEED + 16 * frame_idx + 12, w, h)
