"""Host-side mirror of the reference's front-end classes over the orbx C ABI (include/orbx.h).

The product's host language is C++ (include/orbx_cxx.hpp re-creates ORB_SLAM2::ORBextractor /
ORBmatcher for a C++ caller); this module is the same surface for Python so that tests and
bench.py read like calls into the reference:

    ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)   include/ORBextractor.h:51
    extractor(image, mask) -> keypoints, descriptors                      src/ORBextractor.cc:1043
    extractor.GetScaleFactors() ... , extractor.mvImagePyramid            include/ORBextractor.h:63-85
    ComputeStereoMatches(extL, extR, bf, fx)                              src/Frame.cc:466
    ORBmatcher(nnratio, checkOri).SearchForInitialization / SearchByBoW   src/ORBmatcher.cc:405,522

Everything here goes through liborbx.so (CUDA, sm_100a).  There is no CPU path: if the library
is missing or no GPU is present the calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])

OK, ERR_BAD_ARG, ERR_GEOMETRY, ERR_CAPACITY, ERR_CUDA, ERR_STATE = range(6)


class OrbxError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"orbx status {status}: {msg}")
        self.status = status


class _Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int),
                ("ini_th_fast", C.c_int), ("min_th_fast", C.c_int)]


EXPORTS = ["orbx_extractor_create", "orbx_extractor_destroy", "orbx_get_tables", "orbx_max_keypoints",
           "orbx_extract", "orbx_extract_batch", "orbx_fetch_results", "orbx_pyramid_level",
           "orbx_download_level", "orbx_debug_candidates", "orbx_debug_level_keys",
           "orbx_stereo_match_batch", "orbx_hamming256", "orbx_search_init_batch", "orbx_knn_ratio_batch",
           "orbx_last_error_string", "orbx_kernel_launch_count", "orbx_version", "orbx_set_profiling",
           "orbx_get_profile", "orbx_stream_synchronize", "orbx_knn_ratio_noalloc",
           "orbx_search_projection_batch", "orbx_search_projection_frame_batch"]

STAGES = ["pyramid", "fast", "quadtree", "describe", "stereo_match", "stereo_median", "blur"]


def library_path():
    return os.path.join(_HERE, "liborbx.so")


def lib():
    """Loads liborbx.so (building it in-tree with nvcc if it is not there).  Raises if it cannot."""
    global _LIB
    if _LIB is None:
        path = library_path()
        if not os.path.exists(path):
            from . import build as _build
            _build.build()
        L = C.CDLL(path)
        L.orbx_last_error_string.restype = C.c_char_p
        L.orbx_version.restype = C.c_char_p
        L.orbx_kernel_launch_count.restype = C.c_ulonglong
        L.orbx_extractor_destroy.restype = None
        L.orbx_extractor_destroy.argtypes = [C.c_void_p]
        L.orbx_max_keypoints.argtypes = [C.c_void_p]
        L.orbx_extract_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t,
                                         C.c_int, C.c_int, C.c_void_p]
        L.orbx_fetch_results.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                         C.c_int, C.c_void_p]
        L.orbx_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                   C.c_void_p, C.c_void_p]
        L.orbx_stereo_match_batch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orbx_search_init_batch.argtypes = [C.c_void_p] * 6 + [C.c_int] * 2 + [C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int,
                                                                             C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_knn_ratio_batch.argtypes = [C.c_void_p] * 6 + [C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                                                              C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_download_level.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orbx_debug_candidates.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_debug_level_keys.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_pyramid_level.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orbx_get_tables.argtypes = [C.c_void_p] * 7
        L.orbx_hamming256.argtypes = [C.c_void_p, C.c_void_p]
        L.orbx_stream_synchronize.argtypes = [C.c_void_p]
        L.orbx_fetch_results_packed.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orbx_unpack_keypoints.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_host_alloc.argtypes = [C.c_size_t, C.c_int, C.c_void_p]
        L.orbx_host_free.argtypes = [C.c_void_p]
        L.orbx_host_free.restype = None
        L.orbx_slot_results.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orbx_stream_create.argtypes = [C.c_int, C.c_void_p]
        L.orbx_stream_destroy.argtypes = [C.c_void_p]
        L.orbx_stream_destroy.restype = None
        L.orbx_search_projection_frame_batch.argtypes = ([C.c_void_p] * 4 + [C.c_int] * 2 + [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p] +
                                                         [C.c_float] * 5 + [C.c_void_p] * 7 + [C.c_int, C.c_float, C.c_int] +
                                                         [C.c_void_p] * 3 + [C.c_int, C.c_void_p])
        L.orbx_search_projection_batch.argtypes = ([C.c_void_p] * 4 + [C.c_int] * 2 + [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 9 +
                                                   [C.c_int, C.c_float, C.c_float] + [C.c_void_p] * 3 + [C.c_int, C.c_void_p])
        L.orbx_knn_ratio_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                             C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orbx_search_by_bow_batch.argtypes = [C.c_void_p] * 10 + [C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_search_triangulation_batch.argtypes = [C.c_void_p] * 12 + [C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_vocabulary_create.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_void_p, C.c_void_p] + [C.c_int] * 5 + [C.c_void_p]
        L.orbx_vocabulary_destroy.argtypes = [C.c_void_p]
        L.orbx_vocabulary_destroy.restype = None
        L.orbx_bow_transform_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 5 + [C.c_int, C.c_void_p]
        L.orbx_fuse_search_batch.argtypes = ([C.c_void_p] * 4 + [C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_float] * 5 + [C.c_void_p, C.c_void_p, C.c_int, C.c_float] +
                                             [C.c_void_p] * 7 + [C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_int, C.c_void_p])
        L.orbx_search_by_sim3_batch.argtypes = ([C.c_void_p] * 14 + [C.c_int, C.c_int, C.c_void_p] + [C.c_float] * 4 + [C.c_void_p, C.c_int, C.c_float, C.c_void_p] +
                                                [C.c_void_p] * 12 + [C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p])
        L.orbx_search_projection_ordered_batch.argtypes = ([C.c_void_p] * 3 + [C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_float] * 4 + [C.c_void_p, C.c_int, C.c_float] +
                                                           [C.c_void_p] * 9 + [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p])
        L.orbx_distinctive_descriptors_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_undistort_grid_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_void_p]
        L.orbx_features_in_area_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 9 + [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_set_profiling.argtypes = [C.c_void_p, C.c_int]
        L.orbx_get_profile.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _LIB = L
    return _LIB


def _check(status):
    if status != OK:
        raise OrbxError(status, lib().orbx_last_error_string().decode())


def _ptr(a):
    """Raw address of a numpy array or a torch tensor (None -> NULL)."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()          # torch.Tensor


def _on_device(a):
    return 0 if isinstance(a, np.ndarray) else int(a.is_cuda)


class HostBuffer:
    """Page-locked host memory from orbx_host_alloc as a numpy uint8 array (`.array`); write_combined for upload-only buffers."""

    def __init__(self, nbytes, write_combined=False):
        self.ptr = C.c_void_p()
        _check(lib().orbx_host_alloc(int(nbytes), int(write_combined), C.byref(self.ptr)))
        self.array = np.ctypeslib.as_array((C.c_uint8 * int(nbytes)).from_address(self.ptr.value))

    def close(self):
        if self.ptr:
            self.array = None
            lib().orbx_host_free(self.ptr); self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def kernel_launch_count():
    return int(lib().orbx_kernel_launch_count())


class ORBextractor:
    """ORB_SLAM2::ORBextractor.  The device handle is created on first use (the reference's
    constructor does not know the image size); pass max_w/max_h/max_batch to size it up front."""

    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_w=0, max_h=0, max_batch=1, device=0):
        self.params = _Params(int(nfeatures), float(scaleFactor), int(nlevels), int(iniThFAST), int(minThFAST))
        self.max_w, self.max_h, self.max_batch, self.device = max_w, max_h, max_batch, device
        self._h = None
        self.shape = None
        if max_w and max_h:
            self._create(max_w, max_h)

    def _create(self, w, h):
        h_ = C.c_void_p()
        _check(lib().orbx_extractor_create(C.byref(self.params), int(w), int(h), int(self.max_batch), int(self.device), C.byref(h_)))
        self._h = h_
        self.max_w, self.max_h = w, h
        self.kp_cap = lib().orbx_max_keypoints(self._h)

    def close(self):
        if self._h is not None:
            lib().orbx_extractor_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- the reference's getters (include/ORBextractor.h:63-83) ---
    def _tables(self):
        if self._h is None:
            raise OrbxError(ERR_STATE, "extractor handle not created yet (pass max_w/max_h or extract once)")
        n = C.c_int()
        arr = [np.zeros(8, np.float32) for _ in range(4)]
        q = np.zeros(8, np.int32)
        _check(lib().orbx_get_tables(self._h, C.byref(n), *[a.ctypes.data for a in arr], q.ctypes.data))
        return n.value, arr, q

    def GetLevels(self):
        return self.params.nlevels

    def GetScaleFactor(self):
        return self.params.scale_factor

    def GetScaleFactors(self):
        n, a, _ = self._tables(); return a[0][:n].copy()

    def GetInverseScaleFactors(self):
        n, a, _ = self._tables(); return a[1][:n].copy()

    def GetScaleSigmaSquares(self):
        n, a, _ = self._tables(); return a[2][:n].copy()

    def GetInverseScaleSigmaSquares(self):
        n, a, _ = self._tables(); return a[3][:n].copy()

    def features_per_level(self):
        n, _, q = self._tables(); return q[:n].copy()

    # --- operator() ---
    def __call__(self, image, mask=None, stream=None):
        """(keypoints, descriptors) for one host image; mask is ignored as in the reference.  stream: a cudaStream_t
        (e.g. from orbx_stream_create) so that two extractors on two host threads overlap on the GPU."""
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        if image.dtype != np.uint8 or image.ndim != 2:
            raise OrbxError(ERR_BAD_ARG, "image must be a 2-D uint8 array (CV_8UC1)")   # assert at src/ORBextractor.cc:1050
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)          # rows must be dense; the row stride may exceed the width
        h, w = image.shape
        if self._h is None:
            self._create(w, h)
        cap = self.kp_cap
        kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8); n = C.c_int()
        _check(lib().orbx_extract(self._h, image.ctypes.data, w, h, image.strides[0], kps.ctypes.data, desc.ctypes.data,
                                  cap, C.byref(n), stream))
        self.shape = (h, w)
        return kps[:n.value].copy(), desc[:n.value].copy()

    # --- batched path (many frames per launch) ---
    def extract_batch(self, images, first_slot=0, stream=None):
        """images: (n, h, w) uint8, numpy (host) or torch (host pinned / cuda).  Asynchronous."""
        n, h, w = images.shape
        if self._h is None:
            self._create(w, h)
        if isinstance(images, np.ndarray):
            stride, istride = images.strides[1], images.strides[0]
        else:
            stride, istride = images.stride(1), images.stride(0)
        _check(lib().orbx_extract_batch(self._h, _ptr(images), n, w, h, stride, istride, _on_device(images), first_slot, stream))
        self.shape = (h, w)

    def fetch_into(self, first_slot, n, kps, desc, counts, stream=None):
        """Copies results into caller arrays/tensors laid out [n][cap]; cap is read from `counts`' siblings."""
        ref = kps if kps is not None else desc
        cap = ref.shape[1] if ref is not None else self.kp_cap
        dev = _on_device(ref if ref is not None else counts)
        _check(lib().orbx_fetch_results(self._h, first_slot, n, _ptr(kps), _ptr(desc), _ptr(counts), cap, dev, stream))

    def fetch_packed_into(self, first_slot, n, packed, desc, level_counts, stream=None):
        """The 8-byte keypoint form: packed [n][cap] uint64, desc [n][cap][32], level_counts [n][8] (arrays / tensors, host or device)."""
        ref = packed if packed is not None else desc
        cap = ref.shape[1] if ref is not None else self.kp_cap
        dev = _on_device(ref if ref is not None else level_counts)
        _check(lib().orbx_fetch_results_packed(self._h, first_slot, n, _ptr(packed), _ptr(desc), _ptr(level_counts), cap, dev, stream))

    def unpack_keypoints(self, packed, level_counts, out=None, counts=None):
        """Host arrays from fetch_packed_into -> (keypoints [n][cap] KP_DTYPE, counts [n]): the cv::KeyPoint records."""
        n, cap = packed.shape
        out = np.zeros((n, cap), KP_DTYPE) if out is None else out
        counts = np.zeros(n, np.int32) if counts is None else counts
        _check(lib().orbx_unpack_keypoints(C.byref(self.params), _ptr(packed), _ptr(level_counts), n, cap, _ptr(out), out.shape[1], _ptr(counts)))
        return out, counts

    def fetch(self, first_slot=0, n=1):
        """Synchronous convenience: list of (keypoints, descriptors) per slot."""
        cap = self.kp_cap
        kps = np.zeros((n, cap), KP_DTYPE); desc = np.zeros((n, cap, 32), np.uint8); cnt = np.zeros(n, np.int32)
        self.fetch_into(first_slot, n, kps, desc, cnt)
        _sync(self.device)
        return [(kps[i, :cnt[i]].copy(), desc[i, :cnt[i]].copy()) for i in range(n)]

    def slot_results(self, first_slot=0):
        """Device addresses (ints) of the keypoints, descriptors and counts of slots first_slot.., and the row capacity:
        the in-place route into the matcher entry points (on_device = 1), no copy."""
        k, d, c, cap = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_int()
        _check(lib().orbx_slot_results(self._h, first_slot, C.byref(k), C.byref(d), C.byref(c), C.byref(cap)))
        return k.value, d.value, c.value, cap.value

    def set_profiling(self, enable=True):
        _check(lib().orbx_set_profiling(self._h, int(enable)))

    def get_profile(self):
        """{stage: (total ms, recorded passes)} since the last call (synchronises the device)."""
        ms = np.zeros(len(STAGES), np.float32); calls = np.zeros(len(STAGES), np.int32)
        _check(lib().orbx_get_profile(self._h, ms.ctypes.data, calls.ctypes.data))
        return {s: (float(ms[i]), int(calls[i])) for i, s in enumerate(STAGES)}

    def level(self, slot, level):
        """Bordered pyramid plane (h+38, w+38) of one slot, as the reference's mvImagePyramid backing store."""
        ww, hh = C.c_int(), C.c_int()
        _check(lib().orbx_pyramid_level(self._h, slot, level, None, C.byref(ww), C.byref(hh), None))
        out = np.zeros((hh.value + 38, ww.value + 38), np.uint8)
        _check(lib().orbx_download_level(self._h, slot, level, out.ctypes.data))
        return out

    @property
    def mvImagePyramid(self):
        """Host copies of slot 0's level ROIs (the reference's public member, include/ORBextractor.h:85)."""
        return [self.level(0, l)[19:-19, 19:-19] for l in range(self.params.nlevels)]

    def debug_candidates(self, slot, level, cap=1 << 20):
        out = np.zeros(cap, np.uint32); n = C.c_int()
        _check(lib().orbx_debug_candidates(self._h, slot, level, out.ctypes.data, cap, C.byref(n)))
        p = out[:n.value]
        return np.stack([p & 0xFFF, (p >> 12) & 0xFFF, p >> 24], 1).astype(np.int32)

    def debug_level_keys(self, slot, level, cap=1 << 16):
        out = np.zeros(cap, np.uint32); n = C.c_int()
        _check(lib().orbx_debug_level_keys(self._h, slot, level, out.ctypes.data, cap, C.byref(n)))
        p = out[:n.value]
        return np.stack([p & 0xFFF, (p >> 12) & 0xFFF, p >> 24], 1).astype(np.int32)


def _sync(device=0):
    import torch
    torch.cuda.synchronize(device)


def ComputeStereoMatches(extL, extR, bf, fx, n=1, slotL0=0, slotR0=0, out=None, stream=None):
    """Frame::ComputeStereoMatches for n pairs already extracted into the given slots.
    Returns (mvuRight [n][cap], mvDepth [n][cap], n_matches [n]).  With `out` = (uRight, depth,
    n_matches) torch/numpy buffers the call is asynchronous and returns them."""
    cap = extL.kp_cap
    if out is None:
        ur = np.zeros((n, cap), np.float32); dp = np.zeros((n, cap), np.float32); nm = np.zeros(n, np.int32)
        _check(lib().orbx_stereo_match_batch(extL._h, slotL0, extR._h, slotR0, n, bf, fx, ur.ctypes.data, dp.ctypes.data,
                                             nm.ctypes.data, cap, 0, stream))
        _sync(extL.device)
        return ur, dp, nm
    ur, dp, nm = out
    _check(lib().orbx_stereo_match_batch(extL._h, slotL0, extR._h, slotR0, n, bf, fx, _ptr(ur), _ptr(dp), _ptr(nm),
                                         ur.shape[1], _on_device(ur), stream))
    return out


class Frame:
    """The slice of ORB_SLAM2::Frame the matchers read: undistorted keypoints, descriptors, image bounds."""

    def __init__(self, keypoints, descriptors, cols, rows, bounds=None):
        """bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY) of a distorted camera (Frame::ComputeImageBounds, src/Frame.cc:436-464;
        FrameGrids.bounds[:4]); None = undistorted camera, [0, cols] x [0, rows]."""
        self.mvKeysUn, self.mDescriptors, self.cols, self.rows = keypoints, descriptors, cols, rows
        b = np.array([0, cols, 0, rows] if bounds is None else list(bounds)[:4], np.float32)
        # mfGridElementWidthInv / HeightInv, src/Frame.cc:212-213 (float division)
        self.grid_bounds = np.array([b[0], b[1], b[2], b[3], np.float32(GRID_COLS) / (b[1] - b[0]), np.float32(GRID_ROWS) / (b[3] - b[2])], np.float32)


class ORBVocabulary:
    """A DBoW2 ORB vocabulary on the device (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h) and its transform(), the
    call behind Frame::ComputeBoW (src/Frame.cc:395-402).  voc = dict(desc, child_first, child_count, children, word_id,
    weight, L) as flat arrays; weighting 0 = TF_IDF, norm 1 = L1 are ORBvoc.txt's settings."""

    def __init__(self, voc, weighting=0, norm=1, device=0):
        self.h = C.c_void_p()
        a = lambda k, t: np.ascontiguousarray(voc[k], t)
        d, cf, cc, ch, wi, wt = a("desc", np.uint8), a("child_first", np.int32), a("child_count", np.int32), a("children", np.int32), \
            a("word_id", np.int32), a("weight", np.float64)
        _check(lib().orbx_vocabulary_create(d.ctypes.data, cf.ctypes.data, cc.ctypes.data, ch.ctypes.data, len(ch), wi.ctypes.data,
                                            wt.ctypes.data, len(cf), int(voc["L"]), weighting, norm, device, C.byref(self.h)))

    def close(self):
        if self.h:
            lib().orbx_vocabulary_destroy(self.h); self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def transform(self, descriptor_lists, levelsup=4):
        """Returns per frame (word per feature, node per feature, BowVector words, BowVector values)."""
        nf = len(descriptor_lists)
        cap = max(max((len(d) for d in descriptor_lists), default=1), 1)
        desc = np.zeros((nf, cap, 32), np.uint8); n = np.zeros(nf, np.int32)
        for f, d in enumerate(descriptor_lists):
            n[f] = len(d); desc[f, :len(d)] = d
        wo = np.zeros((nf, cap), np.int32); no = np.zeros((nf, cap), np.int32); bw = np.zeros((nf, cap), np.int32)
        bv = np.zeros((nf, cap), np.float64); bc = np.zeros(nf, np.int32)
        _check(lib().orbx_bow_transform_batch(self.h, desc.ctypes.data, n.ctypes.data, nf, cap, levelsup, wo.ctypes.data, no.ctypes.data,
                                              bw.ctypes.data, bv.ctypes.data, bc.ctypes.data, 0, None))
        return [(wo[f, :n[f]].copy(), no[f, :n[f]].copy(), bw[f, :bc[f]].copy(), bv[f, :bc[f]].copy()) for f in range(nf)]


def ComputeDistinctiveDescriptors(descriptor_lists):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-305) for a batch of map points: descriptor_lists[m] is
    the (N_m, 32) uint8 array of map point m's observed descriptors.  Returns (best index per map point, (M, 32) descriptors)."""
    m = len(descriptor_lists)
    kcap = max(max((len(d) for d in descriptor_lists), default=1), 1)
    desc = np.zeros((m, kcap, 32), np.uint8); counts = np.zeros(m, np.int32)
    for i, d in enumerate(descriptor_lists):
        counts[i] = len(d); desc[i, :len(d)] = d
    best = np.zeros(m, np.int32); out = np.zeros((m, 32), np.uint8)
    _check(lib().orbx_distinctive_descriptors_batch(desc.ctypes.data, counts.ctypes.data, m, kcap, best.ctypes.data, out.ctypes.data, 0, None))
    return best, out


class _Camera(C.Structure):
    _fields_ = [(f, C.c_float) for f in ("fx", "fy", "cx", "cy", "k1", "k2", "p1", "p2", "k3")]


GRID_COLS, GRID_ROWS = 64, 48


class FrameGrids:
    """Frame::UndistortKeyPoints + ComputeImageBounds + AssignFeaturesToGrid (src/Frame.cc:404-464, 230-245) for a
    batch of frames of one camera, and GetFeaturesInArea (src/Frame.cc:327-380) on the result.
    camera = (fx, fy, cx, cy, k1, k2, p1, p2[, k3])."""

    def __init__(self, keypoint_lists, camera, cols, rows):
        cam = list(camera) + [0.0] * (9 - len(camera))
        self.camera = _Camera(*[float(v) for v in cam])
        nf = len(keypoint_lists)
        self.cap = cap = max(max((len(k) for k in keypoint_lists), default=1), 1)
        self.n = np.array([len(k) for k in keypoint_lists], np.int32)
        kps = np.zeros((nf, cap), KP_DTYPE)
        for f, k in enumerate(keypoint_lists):
            kps[f, :len(k)] = k
        self.kps_un = np.zeros((nf, cap), KP_DTYPE)
        self.bounds = np.zeros(6, np.float32)
        self.cell_start = np.zeros((nf, GRID_COLS * GRID_ROWS + 1), np.int32)
        self.cell_items = np.zeros((nf, cap), np.int32)
        _check(lib().orbx_undistort_grid_batch(kps.ctypes.data, self.n.ctypes.data, nf, cap, C.byref(self.camera), cols, rows,
                                               self.kps_un.ctypes.data, self.bounds.ctypes.data, self.cell_start.ctypes.data,
                                               self.cell_items.ctypes.data, 0, None))

    @property
    def mvKeysUn(self):
        return [self.kps_un[f, :self.n[f]] for f in range(len(self.n))]

    def cell(self, f, ix, iy):
        """mGrid[ix][iy] of frame f."""
        c = ix * GRID_ROWS + iy
        return self.cell_items[f, self.cell_start[f, c]:self.cell_start[f, c + 1]]

    def GetFeaturesInArea(self, queries, qcap=None):
        """queries: iterable of (frame, x, y, r, minLevel, maxLevel) -> list of index arrays in the reference's order."""
        q = list(queries)
        nq = len(q)
        qcap = int(qcap or self.cap)
        qf = np.array([v[0] for v in q], np.int32); qx = np.array([v[1] for v in q], np.float32)
        qy = np.array([v[2] for v in q], np.float32); qr = np.array([v[3] for v in q], np.float32)
        lo = np.array([v[4] for v in q], np.int32); hi = np.array([v[5] for v in q], np.int32)
        out = np.zeros((nq, qcap), np.int32); cnt = np.zeros(nq, np.int32)
        _check(lib().orbx_features_in_area_batch(self.kps_un.ctypes.data, self.n.ctypes.data, len(self.n), self.cap,
                                                 self.bounds.ctypes.data, self.cell_start.ctypes.data, self.cell_items.ctypes.data,
                                                 qf.ctypes.data, qx.ctypes.data, qy.ctypes.data, qr.ctypes.data, lo.ctypes.data,
                                                 hi.ctypes.data, nq, qcap, out.ctypes.data, cnt.ctypes.data, 0, None))
        return [out[i, :min(cnt[i], qcap)].copy() for i in range(nq)], cnt


class ORBmatcher:
    """ORB_SLAM2::ORBmatcher (include/ORBmatcher.h:41-89): the two searches on the hot path."""
    PASS_GRID = True     # False: the projection searches get cell_start = cell_items = NULL and build the feature grid themselves

    TH_LOW, TH_HIGH, HISTO_LENGTH = 50, 100, 30

    def __init__(self, nnratio=0.6, checkOri=True):
        self.mfNNratio, self.mbCheckOrientation = float(nnratio), bool(checkOri)

    @staticmethod
    def DescriptorDistance(a, b):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return int(lib().orbx_hamming256(a.ctypes.data, b.ctypes.data))

    def SearchForInitialization(self, F1, F2, vbPrevMatched, windowSize=10):
        """Returns (nmatches, vnMatches12); vbPrevMatched (N1 x 2 float32) is updated in place."""
        r = self.SearchForInitializationBatch([F1], [F2], [vbPrevMatched], windowSize)
        return r[0][0], r[1][0]

    def SearchForInitializationBatch(self, F1s, F2s, prevs, windowSize=10):
        npairs = len(F1s)
        cap = max(max(len(f.mvKeysUn) for f in F1s), max(len(f.mvKeysUn) for f in F2s), 1)
        k1 = np.zeros((npairs, cap), KP_DTYPE); k2 = np.zeros((npairs, cap), KP_DTYPE)
        d1 = np.zeros((npairs, cap, 32), np.uint8); d2 = np.zeros((npairs, cap, 32), np.uint8)
        n1 = np.zeros(npairs, np.int32); n2 = np.zeros(npairs, np.int32)
        pv = np.zeros((npairs, cap, 2), np.float32)
        for p, (a, b, pr) in enumerate(zip(F1s, F2s, prevs)):
            n1[p], n2[p] = len(a.mvKeysUn), len(b.mvKeysUn)
            k1[p, :n1[p]] = a.mvKeysUn; d1[p, :n1[p]] = a.mDescriptors
            k2[p, :n2[p]] = b.mvKeysUn; d2[p, :n2[p]] = b.mDescriptors
            pv[p, :n1[p]] = pr
        m12 = np.zeros((npairs, cap), np.int32); nm = np.zeros(npairs, np.int32)
        _check(lib().orbx_search_init_batch(k1.ctypes.data, d1.ctypes.data, n1.ctypes.data, k2.ctypes.data, d2.ctypes.data,
                                            n2.ctypes.data, npairs, cap, F2s[0].grid_bounds.ctypes.data, pv.ctypes.data,
                                            self.mfNNratio, int(windowSize), int(self.mbCheckOrientation),
                                            m12.ctypes.data, nm.ctypes.data, 0, None))
        for p, pr in enumerate(prevs):
            pr[:] = pv[p, :n1[p]]
        return [int(v) for v in nm], [m12[p, :n1[p]].copy() for p in range(npairs)]

    def SearchByProjection(self, F, uRight, scale_factors, mp, th=1.0, frame_obs=None, replicas=1):
        """SearchByProjection(Frame&, const vector<MapPoint*>&, th): mp is a dict of per-map-point arrays
        inview, projx, projy, projxr, level, viewcos, desc (m x 32), obs.  Returns (nmatches, assigned, frame_obs).
        replicas > 1 submits the same problem that many times in one call (the many-frames-per-launch form) and returns
        the lists of all results."""
        if replicas > 1:
            return self._projection_replicas(F, uRight, scale_factors, mp, th, frame_obs, replicas)
        k = np.ascontiguousarray(F.mvKeysUn); d = np.ascontiguousarray(F.mDescriptors, np.uint8)
        n, m = len(k), len(mp["level"])
        cap, mcap = max(n, 1), max(m, 1)
        ur = np.ascontiguousarray(uRight, np.float32); sc = np.ascontiguousarray(scale_factors, np.float32)
        fo = np.full(cap, -1, np.int32) if frame_obs is None else np.ascontiguousarray(frame_obs, np.int32).copy()
        a = {key: np.ascontiguousarray(mp[key], dt) for key, dt in [("inview", np.uint8), ("projx", np.float32), ("projy", np.float32),
             ("projxr", np.float32), ("level", np.int32), ("viewcos", np.float32), ("desc", np.uint8), ("obs", np.int32)]}
        cn = np.array([n], np.int32); cm = np.array([m], np.int32)
        assigned = np.zeros(cap, np.int32); nm = np.zeros(1, np.int32)
        _check(lib().orbx_search_projection_batch(k.ctypes.data, d.ctypes.data, ur.ctypes.data, cn.ctypes.data, 1, cap, F.grid_bounds.ctypes.data,
                                                  sc.ctypes.data, len(sc), a["inview"].ctypes.data, a["projx"].ctypes.data,
                                                  a["projy"].ctypes.data, a["projxr"].ctypes.data, a["level"].ctypes.data,
                                                  a["viewcos"].ctypes.data, a["desc"].ctypes.data, a["obs"].ctypes.data, cm.ctypes.data,
                                                  mcap, th, self.mfNNratio, fo.ctypes.data, assigned.ctypes.data, nm.ctypes.data, 0, None))
        return int(nm[0]), assigned[:n].copy(), fo[:n].copy()

    def _projection_replicas(self, F, uRight, scale_factors, mp, th, frame_obs, P):
        k = np.ascontiguousarray(F.mvKeysUn); d = np.ascontiguousarray(F.mDescriptors, np.uint8)
        n, m = len(k), len(mp["level"])
        cap, mcap = max(n, 1), max(m, 1)
        rep = lambda a: np.ascontiguousarray(np.broadcast_to(a, (P,) + a.shape))
        ur = rep(np.ascontiguousarray(uRight, np.float32)); sc = np.ascontiguousarray(scale_factors, np.float32)
        fo = rep(np.full(cap, -1, np.int32) if frame_obs is None else np.ascontiguousarray(frame_obs, np.int32)).copy()
        a = {key: rep(np.ascontiguousarray(mp[key], dt)) for key, dt in [("inview", np.uint8), ("projx", np.float32), ("projy", np.float32),
             ("projxr", np.float32), ("level", np.int32), ("viewcos", np.float32), ("desc", np.uint8), ("obs", np.int32)]}
        kk, dd = rep(k), rep(d)
        cn = np.full(P, n, np.int32); cm = np.full(P, m, np.int32)
        assigned = np.zeros((P, cap), np.int32); nm = np.zeros(P, np.int32)
        _check(lib().orbx_search_projection_batch(kk.ctypes.data, dd.ctypes.data, ur.ctypes.data, cn.ctypes.data, P, cap, F.grid_bounds.ctypes.data,
                                                  sc.ctypes.data, len(sc), a["inview"].ctypes.data, a["projx"].ctypes.data,
                                                  a["projy"].ctypes.data, a["projxr"].ctypes.data, a["level"].ctypes.data,
                                                  a["viewcos"].ctypes.data, a["desc"].ctypes.data, a["obs"].ctypes.data, cm.ctypes.data,
                                                  mcap, th, self.mfNNratio, fo.ctypes.data, assigned.ctypes.data, nm.ctypes.data, 0, None))
        return [int(v) for v in nm], [assigned[p, :n].copy() for p in range(P)], [fo[p, :n].copy() for p in range(P)]

    def SearchByProjectionFrame(self, F, uRight, scale_factors, Rcw, tcw, cam, last, th=15.0, mode=0, frame_obs=None):
        """SearchByProjection(Frame& Current, const Frame& Last, th, bMono).  cam = (fx, fy, cx, cy, bf); last is a dict
        of per-last-keypoint arrays valid, xw (n,3), octave, angle, desc (n,32), obs.  Returns (nmatches, assigned, frame_obs)."""
        k = np.ascontiguousarray(F.mvKeysUn); d = np.ascontiguousarray(F.mDescriptors, np.uint8)
        n, nl = len(k), len(last["octave"])
        cap, lcap = max(n, 1), max(nl, 1)
        ur = np.ascontiguousarray(uRight, np.float32); sc = np.ascontiguousarray(scale_factors, np.float32)
        pose = np.concatenate([np.asarray(Rcw, np.float32).ravel(), np.asarray(tcw, np.float32).ravel()]).astype(np.float32)
        fo = np.full(cap, -1, np.int32) if frame_obs is None else np.ascontiguousarray(frame_obs, np.int32).copy()
        a = {key: np.ascontiguousarray(last[key], dt) for key, dt in [("valid", np.uint8), ("xw", np.float32), ("octave", np.int32),
             ("angle", np.float32), ("desc", np.uint8), ("obs", np.int32)]}
        cn = np.array([n], np.int32); cl = np.array([nl], np.int32); md = np.array([mode], np.int32)
        assigned = np.zeros(cap, np.int32); nm = np.zeros(1, np.int32)
        fx, fy, cx, cy, bf = cam
        _check(lib().orbx_search_projection_frame_batch(k.ctypes.data, d.ctypes.data, ur.ctypes.data, cn.ctypes.data, 1, cap, F.grid_bounds.ctypes.data,
                                                        sc.ctypes.data, len(sc), pose.ctypes.data, md.ctypes.data, fx, fy, cx, cy, bf,
                                                        a["valid"].ctypes.data, a["xw"].ctypes.data, a["octave"].ctypes.data,
                                                        a["angle"].ctypes.data, a["desc"].ctypes.data, a["obs"].ctypes.data, cl.ctypes.data,
                                                        lcap, th, int(self.mbCheckOrientation), fo.ctypes.data, assigned.ctypes.data,
                                                        nm.ctypes.data, 0, None))
        return int(nm[0]), assigned[:n].copy(), fo[:n].copy()

    def SearchByBoW(self, descA, angA, descB, angB, strict=True, th=None):
        """SearchByBoW(KF,KF) with one bucket holding all features (brute-force kNN + ratio).
        th defaults to TH_LOW (src/ORBmatcher.cc:37,596)."""
        r = self.SearchByBoWBatch([(descA, angA)], [(descB, angB)], strict, th)
        return r[0][0], r[1][0]

    def SearchForTriangulation(self, kf1, kf2, F12, pose2, Cw, cam2, scale_factors2, level_sigma2_2, bOnlyStereo=False):
        """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-823) for a batch of keyframe pairs.
        kf1[p] / kf2[p] = dict(kps=mvKeysUn, desc, has_mp, u_right, node); F12[p] 3x3, pose2[p] = (R2w 3x3, t2w 3), Cw[p] 3,
        cam2[p] = (fx, fy, cx, cy).  Returns (nmatches per pair, vMatches12 per pair)."""
        npairs = len(kf1)
        cap = max(max(len(k["kps"]) for k in kf1), max(len(k["kps"]) for k in kf2), 1)
        def pack(kfs):
            k = np.zeros((npairs, cap), KP_DTYPE); d = np.zeros((npairs, cap, 32), np.uint8); hm = np.zeros((npairs, cap), np.uint8)
            ur = np.full((npairs, cap), -1, np.float32); nd = np.zeros((npairs, cap), np.int32); n = np.zeros(npairs, np.int32)
            for p, kf in enumerate(kfs):
                n[p] = len(kf["kps"]); k[p, :n[p]] = kf["kps"]; d[p, :n[p]] = kf["desc"]; hm[p, :n[p]] = kf["has_mp"]
                ur[p, :n[p]] = kf["u_right"]; nd[p, :n[p]] = kf["node"]
            return k, d, hm, ur, nd, n
        A = pack(kf1); B = pack(kf2)
        F = np.ascontiguousarray(np.asarray(F12, np.float32).reshape(npairs, 9))
        T = np.ascontiguousarray(np.stack([np.concatenate([np.asarray(R, np.float32).reshape(9), np.asarray(t, np.float32).reshape(3)]) for R, t in pose2]))
        Cc = np.ascontiguousarray(np.asarray(Cw, np.float32).reshape(npairs, 3)); K = np.ascontiguousarray(np.asarray(cam2, np.float32).reshape(npairs, 4))
        sc = np.ascontiguousarray(scale_factors2, np.float32); sg = np.ascontiguousarray(level_sigma2_2, np.float32)
        m12 = np.zeros((npairs, cap), np.int32); nm = np.zeros(npairs, np.int32)
        _check(lib().orbx_search_triangulation_batch(*[x.ctypes.data for x in A], *[x.ctypes.data for x in B], npairs, cap,
                                                     F.ctypes.data, T.ctypes.data, Cc.ctypes.data, K.ctypes.data, sc.ctypes.data, sg.ctypes.data,
                                                     len(sc), int(bOnlyStereo), int(self.mbCheckOrientation), m12.ctypes.data, nm.ctypes.data, 0, None))
        return [int(v) for v in nm], [m12[p, :A[5][p]].copy() for p in range(npairs)]

    @staticmethod
    def FuseSearch(grids, f, u_right, desc, kf_bounds, Rcw, tcw, Ow, cam, scale_factors, inv_level_sigma2, log_scale_factor, mp, th=3.0, sim3_variant=False):
        """The search of ORBmatcher::Fuse(pKF, vpMapPoints, th) (or, with sim3_variant, of Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)) (src/ORBmatcher.cc:825-975) for keyframe f of `grids` (a FrameGrids):
        cam = (fx, fy, cx, cy, bf); mp = dict(valid, xw, normal, dmin, dmax, desc) with dmin / dmax = mfMinDistance / mfMaxDistance.
        Returns bestIdx per candidate (-1 = no keypoint within TH_LOW)."""
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        cap = grids.cap; n = int(grids.n[f]); m = len(mp["dmin"]); mcap = max(m, 1)
        k = np.ascontiguousarray(grids.kps_un[f]); d = np.zeros((cap, 32), np.uint8); d[:n] = desc
        ur = np.full(cap, -1, np.float32); ur[:n] = u_right
        pose = np.concatenate([f32(Rcw).reshape(9), f32(tcw).reshape(3)])
        pad = lambda a, shape, dt: (lambda o: (o.__setitem__(slice(0, m), np.asarray(a, dt).reshape((m,) + shape)), o)[1])(np.zeros((mcap,) + shape, dt))
        best = np.zeros(mcap, np.int32)
        nn = np.array([n], np.int32); mm = np.array([m], np.int32)
        cs = np.ascontiguousarray(grids.cell_start[f]); ci = np.ascontiguousarray(grids.cell_items[f])
        kb = f32(kf_bounds); sc = f32(scale_factors); isg = f32(inv_level_sigma2); ow = f32(Ow)
        arrs = [pad(mp["valid"], (), np.uint8), pad(mp["xw"], (3,), np.float32), pad(mp["normal"], (3,), np.float32), pad(mp["dmin"], (), np.float32),
                pad(mp["dmax"], (), np.float32), pad(mp["desc"], (32,), np.uint8)]
        _check(lib().orbx_fuse_search_batch(k.ctypes.data, d.ctypes.data, ur.ctypes.data, nn.ctypes.data, 1, cap, grids.bounds.ctypes.data, cs.ctypes.data if ORBmatcher.PASS_GRID else None,
                                            ci.ctypes.data if ORBmatcher.PASS_GRID else None, kb.ctypes.data, pose.ctypes.data, ow.ctypes.data, *[float(v) for v in cam], sc.ctypes.data,
                                            isg.ctypes.data, len(sc), float(log_scale_factor), *[a.ctypes.data for a in arrs], mm.ctypes.data, mcap,
                                            float(th), int(sim3_variant), best.ctypes.data, 0, None))
        return best[:m].copy()

    def SearchByProjectionOrdered(self, grids, f, desc, target_bounds, Rcw, tcw, Ow, cam, scale_factors, log_scale_factor, occupied, mp, th, max_dist,
                                  relocalisation):
        """SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (relocalisation=False, src/ORBmatcher.cc:290) or
        SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (relocalisation=True, :1472) on target f of `grids`.
        mp = dict(valid, xw, normal, dmin, dmax, desc, angle).  Returns (nmatches, assigned per keypoint of the target)."""
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        cap = grids.cap; n = int(grids.n[f]); m = len(mp["dmin"]); mcap = max(m, 1)
        d = np.zeros((cap, 32), np.uint8); d[:n] = desc
        occ = np.zeros(cap, np.uint8); occ[:n] = occupied
        pose = np.concatenate([f32(Rcw).reshape(9), f32(tcw).reshape(3)]); ow = f32(Ow); tb = f32(target_bounds); sc = f32(scale_factors)
        def pad(a, shape, dt):
            o = np.zeros((mcap,) + shape, dt); o[:m] = np.asarray(a, dt).reshape((m,) + shape); return o
        arrs = [pad(mp["valid"], (), np.uint8), pad(mp["xw"], (3,), np.float32), pad(mp["normal"], (3,), np.float32), pad(mp["dmin"], (), np.float32),
                pad(mp["dmax"], (), np.float32), pad(mp["desc"], (32,), np.uint8), pad(mp["angle"], (), np.float32)]
        k = np.ascontiguousarray(grids.kps_un[f]); cs = np.ascontiguousarray(grids.cell_start[f]); ci = np.ascontiguousarray(grids.cell_items[f])
        nn = np.array([n], np.int32); mm = np.array([m], np.int32)
        assigned = np.zeros(cap, np.int32); nm = np.zeros(1, np.int32)
        _check(lib().orbx_search_projection_ordered_batch(k.ctypes.data, d.ctypes.data, nn.ctypes.data, 1, cap, grids.bounds.ctypes.data, cs.ctypes.data if ORBmatcher.PASS_GRID else None,
                                                          ci.ctypes.data if ORBmatcher.PASS_GRID else None, tb.ctypes.data, pose.ctypes.data, ow.ctypes.data, *[float(v) for v in cam[:4]],
                                                          sc.ctypes.data, len(sc), float(log_scale_factor), occ.ctypes.data, *[a.ctypes.data for a in arrs],
                                                          mm.ctypes.data, mcap, float(th), int(max_dist), int(relocalisation),
                                                          int(self.mbCheckOrientation), assigned.ctypes.data, nm.ctypes.data, 0, None))
        return int(nm[0]), assigned[:n].copy()

    @staticmethod
    def SearchBySim3(grids, desc1, desc2, kf_bounds, pose1, pose2, cam, scale_factors, log_scale_factor, xf, mp1, mp2, th=7.5):
        """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1102-1326) for the keyframe pair (0, 1) of `grids`.  pose = (Rcw, tcw);
        cam = (fx, fy, cx, cy) of keyframe 1; xf = sR12 (9), t12 (3), sR21 (9), t21 (3) as :1119-1122 compute them; mp1 / mp2 =
        dict(index, xw, dmin, dmax, desc) of the taking-part map points of each keyframe.  Returns (nFound, matches12 over KF1's features)."""
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        cap = grids.cap; n1, n2 = int(grids.n[0]), int(grids.n[1])
        mcap = max(len(mp1["index"]), len(mp2["index"]), 1)
        def side(f, desc, pose):
            d = np.zeros((cap, 32), np.uint8); d[:int(grids.n[f])] = desc
            return [np.ascontiguousarray(grids.kps_un[f]), d, np.array([grids.n[f]], np.int32), np.ascontiguousarray(grids.cell_start[f]),
                    np.ascontiguousarray(grids.cell_items[f]), f32(kf_bounds), np.concatenate([f32(pose[0]).reshape(9), f32(pose[1]).reshape(3)])]
        def mps(mp):
            m = len(mp["index"])
            idx = np.full(mcap, -1, np.int32); idx[:m] = mp["index"]
            xw = np.zeros((mcap, 3), np.float32); xw[:m] = mp["xw"]
            lo = np.zeros(mcap, np.float32); lo[:m] = mp["dmin"]; hi = np.zeros(mcap, np.float32); hi[:m] = mp["dmax"]
            dd = np.zeros((mcap, 32), np.uint8); dd[:m] = mp["desc"]
            return [idx, xw, lo, hi, dd, np.array([m], np.int32)]
        A, B, M1, M2 = side(0, desc1, pose1), side(1, desc2, pose2), mps(mp1), mps(mp2)
        sc = f32(scale_factors); x = f32(xf).reshape(24)
        m12 = np.zeros(cap, np.int32); nf = np.zeros(1, np.int32)
        gp = lambda S: [(None if (q in (3, 4) and not ORBmatcher.PASS_GRID) else a.ctypes.data) for q, a in enumerate(S)]
        _check(lib().orbx_search_by_sim3_batch(*gp(A), *gp(B), 1, cap, grids.bounds.ctypes.data,
                                               *[float(v) for v in cam[:4]], sc.ctypes.data, len(sc), float(log_scale_factor), x.ctypes.data,
                                               *[a.ctypes.data for a in M1], *[a.ctypes.data for a in M2], mcap, float(th), m12.ctypes.data,
                                               nf.ctypes.data, 0, None))
        return int(nf[0]), m12[:n1].copy()

    def SearchByBoWNodes(self, sidesA, sidesB, strict=True, th=None):
        """SearchByBoW with vocabulary buckets, batched.  Each side = (desc, angles, valid, node) per pair; strict=True is
        the (KeyFrame, KeyFrame) form (src/ORBmatcher.cc:522), strict=False the (KeyFrame, Frame) form (:159).
        Returns (nmatches per pair, matches12 per pair)."""
        npairs = len(sidesA)
        cap = max(max(len(a[0]) for a in sidesA), max(len(b[0]) for b in sidesB), 1)
        def pack(sides):
            d = np.zeros((npairs, cap, 32), np.uint8); ang = np.zeros((npairs, cap), np.float32)
            val = np.zeros((npairs, cap), np.uint8); node = np.zeros((npairs, cap), np.int32); n = np.zeros(npairs, np.int32)
            for p, (dd, aa, vv, nn) in enumerate(sides):
                n[p] = len(dd); d[p, :n[p]] = dd; ang[p, :n[p]] = aa; val[p, :n[p]] = vv; node[p, :n[p]] = nn
            return d, ang, val, node, n
        dA, aA, vA, nodA, nA = pack(sidesA); dB, aB, vB, nodB, nB = pack(sidesB)
        m12 = np.zeros((npairs, cap), np.int32); nm = np.zeros(npairs, np.int32)
        _check(lib().orbx_search_by_bow_batch(dA.ctypes.data, aA.ctypes.data, vA.ctypes.data, nodA.ctypes.data, nA.ctypes.data,
                                              dB.ctypes.data, aB.ctypes.data, vB.ctypes.data, nodB.ctypes.data, nB.ctypes.data,
                                              npairs, cap, self.TH_LOW if th is None else int(th), self.mfNNratio,
                                              int(self.mbCheckOrientation), int(strict), m12.ctypes.data, nm.ctypes.data, 0, None))
        return [int(v) for v in nm], [m12[p, :nA[p]].copy() for p in range(npairs)]

    def SearchByBoWBatch(self, As, Bs, strict=True, th=None):
        npairs = len(As)
        cap = max(max(len(a[0]) for a in As), max(len(b[0]) for b in Bs), 1)
        dA = np.zeros((npairs, cap, 32), np.uint8); dB = np.zeros((npairs, cap, 32), np.uint8)
        aA = np.zeros((npairs, cap), np.float32); aB = np.zeros((npairs, cap), np.float32)
        nA = np.zeros(npairs, np.int32); nB = np.zeros(npairs, np.int32)
        for p, ((da, aa), (db, ab)) in enumerate(zip(As, Bs)):
            nA[p], nB[p] = len(da), len(db)
            dA[p, :nA[p]] = da; aA[p, :nA[p]] = aa; dB[p, :nB[p]] = db; aB[p, :nB[p]] = ab
        m12 = np.zeros((npairs, cap), np.int32); nm = np.zeros(npairs, np.int32)
        _check(lib().orbx_knn_ratio_batch(dA.ctypes.data, aA.ctypes.data, nA.ctypes.data, dB.ctypes.data, aB.ctypes.data,
                                          nB.ctypes.data, npairs, cap, self.TH_LOW if th is None else int(th), self.mfNNratio,
                                          int(self.mbCheckOrientation), int(strict), m12.ctypes.data, nm.ctypes.data, 0, None))
        return [int(v) for v in nm], [m12[p, :nA[p]].copy() for p in range(npairs)]
