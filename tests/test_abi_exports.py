"""CPU-side checks of the drop-in boundary: liborbx.so builds/loads and exports exactly the entry
points include/orbx.h declares; argument validation that needs no GPU; the oracle is not reachable
from the product package."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "orbx.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from orb_slam2_b200 import api, build
    build.build()
    lib = ctypes.CDLL(api.library_path())
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/orbx.h but not exported"
    assert sorted(api.EXPORTS) == [n for n in names if n in api.EXPORTS]
    out = subprocess.run(["nm", "-D", "--defined-only", api.library_path()], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (orbx_[a-z0-9_]+)", out))
    assert exported == set(names), exported ^ set(names)


def test_header_cites_reference_for_each_entry_point():
    txt = open(os.path.join(ROOT, "include", "orbx.h")).read()
    for sym in ["orbx_extractor_create", "orbx_extract", "orbx_stereo_match_batch", "orbx_search_init_batch",
                "orbx_knn_ratio_batch", "orbx_hamming256", "orbx_pyramid_level", "orbx_get_tables"]:
        head = txt[:txt.index(sym + "(")]
        comment = head[head.rindex("/*"):]
        assert re.search(r"src/\w+\.cc:\d+|include/\w+\.h:\d+", comment), sym


def test_host_helpers_without_gpu():
    from orb_slam2_b200 import ORBmatcher
    import numpy as np
    z = np.zeros(32, np.uint8); o = np.full(32, 255, np.uint8)
    assert ORBmatcher.DescriptorDistance(z, o) == 256
    assert ORBmatcher.DescriptorDistance(z, z) == 0
    assert (ORBmatcher.TH_LOW, ORBmatcher.TH_HIGH, ORBmatcher.HISTO_LENGTH) == (50, 100, 30)


def test_tables_without_a_device(oracle):
    """orbx_compute_tables: the constructor tables (src/ORBextractor.cc:414-446) from the parameters alone - what every Frame
    constructor reads before the first extraction (src/Frame.cc:69-75).  Checked against the oracle (itself pinned to the reference)."""
    import numpy as np
    from orb_slam2_b200 import api
    L = api.lib()
    for nf, sf, nl in [(2000, 1.2, 8), (1000, 1.2, 8), (1200, 1.2, 8), (500, 1.5, 5)]:
        p = api._Params(nf, sf, nl, 20, 7)
        n = ctypes.c_int(); t = [np.zeros(8, np.float32) for _ in range(4)]; q = np.zeros(8, np.int32)
        assert L.orbx_compute_tables(ctypes.byref(p), ctypes.byref(n), *[a.ctypes.data_as(ctypes.c_void_p) for a in t], q.ctypes.data_as(ctypes.c_void_p)) == 0
        ref = oracle.Extractor(nf, sf, nl).tables()
        assert n.value == nl
        for a, key in zip(t, ("scale", "inv", "sigma2", "invsigma2")):
            assert np.array_equal(a[:nl].view(np.uint32), ref[key].view(np.uint32)), key
        assert np.array_equal(q[:nl], ref["quota"])
    assert L.orbx_compute_tables(ctypes.byref(api._Params(1000, 1.2, 9, 20, 7)), None, None, None, None, None, None) == api.ERR_BAD_ARG


def test_bad_arguments_rejected_before_touching_the_gpu():
    from orb_slam2_b200 import api
    L = api.lib()
    h = ctypes.c_void_p()
    p = api._Params(1000, 1.2, 9, 20, 7)
    assert L.orbx_extractor_create(ctypes.byref(p), 640, 480, 1, 0, ctypes.byref(h)) == api.ERR_BAD_ARG
    p = api._Params(1000, 1.2, 8, 300, 7)
    assert L.orbx_extractor_create(ctypes.byref(p), 640, 480, 1, 0, ctypes.byref(h)) == api.ERR_BAD_ARG
    assert b"threshold" in L.orbx_last_error_string()
    assert L.orbx_extractor_create(None, 640, 480, 1, 0, ctypes.byref(h)) == api.ERR_BAD_ARG


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from orb_slam2_b200 import ORBextractor, OrbxError
    with pytest.raises(OrbxError) as ei:
        ORBextractor(1000, 1.2, 8, 20, 7, max_w=640, max_h=480)
    assert ei.value.status == 4    # ORBX_ERR_CUDA: fails loudly, nothing is computed on the CPU


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "orb_slam2_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dp, fn)).read()
                assert "pyoracle" not in txt and "orb_oracle" not in txt and "liborb_oracle" not in txt, fn


def test_cxx_wrapper_header_compiles(tmp_path):
    """include/orbx_cxx.hpp (the ORB_SLAM2::ORBextractor / ORBmatcher stand-ins) must compile as
    C++11 against OpenCV-shaped headers; the oracle's cv shim stands in for OpenCV here."""
    src = tmp_path / "t.cpp"
    src.write_text('#include "orbx_cxx.hpp"\n'
                   'int main(){ ORB_SLAM2::ORBextractor e(1000,1.2f,8,20,7); ORB_SLAM2::ORBmatcher m(0.9f,true);\n'
                   ' std::vector<cv::KeyPoint> k; cv::Mat img, d; e(img, cv::Mat(), k, d); return e.GetLevels()==8 ? 0 : 1; }\n')
    r = subprocess.run(["g++", "-std=gnu++14", "-fsyntax-only", "-I", os.path.join(ROOT, "include"),
                        "-I", os.path.join(ROOT, "oracle", "cvshim"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_unpack_keypoints_is_host_code_and_rebuilds_the_records(oracle):
    """orbx_unpack_keypoints needs no device: the oracle's per-level keypoints (level coordinates), packed into the 8-byte form
    the CUDA path downloads, come back as exactly the cv::KeyPoint records the oracle's extractor returns."""
    import numpy as np
    from orb_slam2_b200 import api, synth
    img = synth.mono_frame(3, 640, 480)
    E = oracle.Extractor(1000)
    k, _ = E.extract(img)
    lv = [E.level_keys(l) for l in range(8)]
    cap = 1100
    packed = np.zeros((1, cap), np.uint64); lc = np.zeros((1, 8), np.int32)
    j = 0
    for l, q in enumerate(lv):
        lc[0, l] = len(q)
        for r in q:
            key = int(r["x"]) | (int(r["y"]) << 12) | (int(r["response"]) << 24)
            packed[0, j] = key | (int(np.float32(r["angle"]).view(np.uint32)) << 32)
            j += 1
    out = np.zeros((1, cap), api.KP_DTYPE); cnt = np.zeros(1, np.int32)
    p = api._Params(1000, 1.2, 8, 20, 7)
    rc = api.lib().orbx_unpack_keypoints(ctypes.byref(p), packed.ctypes.data, lc.ctypes.data, 1, cap, out.ctypes.data, cap, cnt.ctypes.data)
    assert rc == 0 and cnt[0] == len(k) == j
    assert out[0, :j].tobytes() == k.tobytes()
    # a buffer that is too small is refused
    small = np.zeros((1, 10), api.KP_DTYPE)
    assert api.lib().orbx_unpack_keypoints(ctypes.byref(p), packed.ctypes.data, lc.ctypes.data, 1, cap, small.ctypes.data, 10, cnt.ctypes.data) == api.ERR_CAPACITY
