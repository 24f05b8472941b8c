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


def variant(img, v):
    """Distinct frame number v derived from a seeded base frame at almost no cost (large soaks and batches need thousands
    of different images; painting each from scratch costs 50 ms): the columns are rolled by 37 * v (a new alignment with the
    FAST cell grid and the quad-tree boxes, and a vertical seam), every other variant is turned upside down.  Applied alike
    to both eyes of a stereo pair it keeps the planted horizontal disparities.  v = 0 is the base frame."""
    out = np.roll(img, (37 * v) % img.shape[1], axis=1) if v else img
    return np.ascontiguousarray(out[::-1] if (v & 1) else out)


def stereo_set(first_seed, n_base, count, w, h):
    """count distinct stereo pairs as two arrays (count, h, w): n_base painted pairs (seeds first_seed ..) and their variants."""
    base = [stereo_pair(first_seed + i, w, h)[:2] for i in range(n_base)]
    L = np.stack([variant(base[i % n_base][0], i // n_base) for i in range(count)])
    R = np.stack([variant(base[i % n_base][1], i // n_base) for i in range(count)])
    return L, R


def mono_frame(frame_idx, w, h):
    return synth_image(BASE_SEED + 16 * frame_idx + 12, w, h)


def vocabulary(seed, k=10, L=4, stop_fraction=0.02, train=None):
    """A synthetic DBoW2-shaped vocabulary tree (there is no ORBvoc.txt in the reference tree): k children per node,
    L levels, node ids in breadth-first creation order with the root = 0 (as TemplatedVocabulary::create numbers them),
    leaves are the words (ids in leaf order).  Node descriptors: a child = its parent with a random 1/8 of the bits
    re-drawn, so that descents are decided by real distances; weights = idf-like positive doubles, a few words
    stopped (weight 0).  `train` (optional (N,32) uint8) seeds the first level with real descriptors."""
    rng = np.random.default_rng(seed)
    desc = [np.zeros(32, np.uint8)]; child_first = [0]; child_count = [0]; children = []; level = [0]
    frontier = [0]
    for lv in range(1, L + 1):
        nxt = []
        for pnode in frontier:
            child_first[pnode] = len(children); child_count[pnode] = k
            for c in range(k):
                nid = len(desc)
                if lv == 1:
                    d = train[rng.integers(0, len(train))].copy() if train is not None and len(train) else rng.integers(0, 256, 32, dtype=np.uint8)
                else:
                    flip = np.packbits(rng.random(256) < 0.125)
                    d = desc[pnode] ^ (flip & rng.integers(0, 256, 32, dtype=np.uint8))
                desc.append(d); child_first.append(0); child_count.append(0); level.append(lv)
                children.append(nid); nxt.append(nid)
        frontier = nxt
    n = len(desc)
    word_id = np.full(n, -1, np.int32); weight = np.zeros(n, np.float64)
    leaves = [i for i in range(n) if child_count[i] == 0]
    word_id[leaves] = np.arange(len(leaves), dtype=np.int32)
    w = np.log(rng.uniform(1.5, 4000.0, len(leaves)))
    w[rng.random(len(leaves)) < stop_fraction] = 0.0
    weight[leaves] = w
    return dict(desc=np.ascontiguousarray(np.stack(desc)), child_first=np.array(child_first, np.int32),
                child_count=np.array(child_count, np.int32), children=np.array(children, np.int32),
                word_id=word_id, weight=weight, L=L, k=k)


def vocabulary_fast(seed, k=10, L=6):
    """Same tree shape as vocabulary(), generated level by level with array operations (ORBvoc.txt's k = 10, L = 6
    is 1.1 M nodes); for throughput measurements, not tied to the golden vectors."""
    rng = np.random.default_rng(seed)
    levels = [np.zeros((1, 32), np.uint8)]
    for lv in range(1, L + 1):
        par = np.repeat(levels[-1], k, axis=0)
        if lv == 1:
            cur = rng.integers(0, 256, par.shape, dtype=np.uint8)
        else:
            flip = np.packbits(rng.random((len(par), 256)) < 0.125, axis=1)
            cur = par ^ (flip & rng.integers(0, 256, par.shape, dtype=np.uint8))
        levels.append(cur)
    desc = np.ascontiguousarray(np.concatenate(levels))
    n = len(desc)
    first = np.cumsum([0] + [len(x) for x in levels])            # id of the first node of each level
    child_first = np.zeros(n, np.int32); child_count = np.zeros(n, np.int32)
    inner = first[L]                                              # nodes below this id have children
    child_count[:inner] = k
    child_first[:inner] = np.arange(inner, dtype=np.int64) * k    # children list = ids 1.. in order
    children = np.arange(1, n, dtype=np.int32)
    word_id = np.full(n, -1, np.int32); word_id[inner:] = np.arange(n - inner, dtype=np.int32)
    weight = np.zeros(n, np.float64); weight[inner:] = np.log(rng.uniform(1.5, 4000.0, n - inner))
    return dict(desc=desc, child_first=child_first, child_count=child_count, children=children, word_id=word_id, weight=weight, L=L, k=k)
