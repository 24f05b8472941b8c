"""The drop-in boundary, end to end: oracle/_ref/orbref_slam is the reference's own Frame / KeyFrame / MapPoint / Map / ORBmatcher /
ORBextractor sources behind a small driver; oracle/_ref/orbref_slam_gpu is the SAME driver and the SAME reference sources with only
src/ORBextractor.cc and src/ORBmatcher.cc swapped for dropin/ORBextractor.h + dropin/ORBmatcher.cc, i.e. for the orbx library.
Both run here on the same inputs (real stereo / monocular Frame constructors, real KeyFrames and MapPoints, every ORBmatcher
member) and must write the same bytes.  The binaries are built where /root/reference exists (make -C oracle ref) and travel with
the snapshot; nothing here reads the reference tree."""
import os

import numpy as np
import pytest

from conftest import missing_reference_binary

from orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu


def _same(a, b, path=""):
    if isinstance(a, dict):
        assert a.keys() == b.keys(), path
        for k in a:
            _same(a[k], b[k], f"{path}/{k}")
    elif isinstance(a, (tuple, list)):
        assert len(a) == len(b), path
        for i, (x, y) in enumerate(zip(a, b)):
            _same(x, y, f"{path}[{i}]")
    elif isinstance(a, np.ndarray):
        assert a.dtype == b.dtype and a.shape == b.shape and a.tobytes() == b.tobytes(), path
    else:
        assert a == b, path


def _runner(oracle, names):
    """Runs an oracle.ref_slam_* call once per binary of oracle/_ref (the all-reference one first)."""
    d = os.path.dirname(oracle.REF_SLAM_BIN)
    bins = [os.path.join(d, n) for n in names]
    if not all(os.path.exists(b) for b in bins):
        missing_reference_binary("oracle/_ref/{" + ",".join(names) + "}")
    ref_bin = oracle.REF_SLAM_BIN

    def run(fn, *args, **kw):
        out = []
        for b in bins:
            oracle.REF_SLAM_BIN = b
            try:
                out.append(fn(*args, **kw))
            finally:
                oracle.REF_SLAM_BIN = ref_bin
        return out
    return run


@pytest.fixture(scope="module")
def both(oracle):
    return _runner(oracle, ["orbref_slam", "orbref_slam_gpu"])


@pytest.fixture(scope="module")
def all_three(oracle):
    """reference | drop-in extractor + matcher with the reference's CPU ComputeStereoMatches on a downloaded pyramid |
    the same with dropin/FrameStereo.cc: Frame::ComputeStereoMatches on the GPU, nothing copied back but the results."""
    return _runner(oracle, ["orbref_slam", "orbref_slam_gpu", "orbref_slam_gpustereo"])


def test_stereo_frame_constructor(oracle, all_three):
    """The reference's own stereo Frame constructor (src/Frame.cc:61-117) in all three builds: mvKeys, mvKeysUn, descriptors,
    mvuRight and mvDepth must be the same bytes (KITTI and EuRoC shapes)."""
    for cfg, cam, seeds in ((synth.KITTI, (718.856, 718.856, 607.1928, 185.2157), (5, 43)),
                            (synth.EUROC, (435.2047, 435.2047, 367.4517, 252.2008), (6,))):
        for seed in seeds:
            L, R, _ = synth.stereo_pair(seed, cfg["w"], cfg["h"])
            ref, gpu, gpu_stereo = all_three(oracle.ref_slam_stereo, L, R, cfg["nfeatures"], cam, cfg["bf"])
            _same(ref, gpu)
            _same(ref, gpu_stereo, "gpustereo")
            assert (ref["u_right"] >= 0).sum() > 200


def test_keyframe_searches_with_gpu_stereo_frames(oracle, all_three):
    """One keyframe case through the third binary as well: every Frame it builds gets its depth from the GPU stereo body."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("kf_cases", os.path.join(os.path.dirname(__file__), "kf_cases.py"))
    kc = importlib.util.module_from_spec(spec); spec.loader.exec_module(kc)
    gs = kc.gs
    cfg = synth.KITTI
    seed, L, only_stereo, mono, ori = gs.KF_CASES[2]
    ims, V, T1, T2, F12 = gs.kf_case_inputs(seed, L)
    ref, gpu, gpu_stereo = all_three(oracle.ref_slam_kf, *ims, cfg["nfeatures"], gs.KITTI_CAM, cfg["bf"], V, T1, T2, F12, check_ori=ori,
                                     only_stereo=only_stereo, mono=mono, **gs.KF_PARAMS)
    _same(ref, gpu_stereo, "gpustereo")
    _same(ref, gpu)


@pytest.mark.parametrize("form", ["latency", "small"])
def test_monocular_frames_and_initialisation(oracle, both, form, monkeypatch):
    if form == "small":
        monkeypatch.setenv("ORBX_PROJ_SERIAL", "1")
    a, b, _ = synth.mono_pair(60, 640, 480)
    ref, gpu = both(oracle.ref_slam_init, a, b, 1000, (517.3, 516.5, 318.6, 255.3), 0.9, 100, True)
    _same(ref, gpu)
    assert ref[2] > 50


def test_distorted_camera_grid(oracle, both):
    tum1 = (517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
    rng = np.random.default_rng(2)
    qs = [(float(rng.uniform(0, 640)), float(rng.uniform(0, 480)), float(rng.uniform(5, 100)), -1, -1) for _ in range(50)]
    ref, gpu = both(oracle.ref_slam_grid, synth.mono_frame(300, 640, 480), 1000, tum1, qs)
    _same(ref, gpu)


@pytest.mark.parametrize("form", ["latency", "small"])
@pytest.mark.parametrize("fuse_scw", [False, True])
def test_keyframes_mappoints_and_every_matcher_member(oracle, both, fuse_scw, form, monkeypatch):
    """Every ORBmatcher member of the drop-in against the reference's, in both device forms of the matcher kernels: the 1024-thread
    form (which builds the feature grid itself: the drop-in passes none) and, with ORBX_PROJ_SERIAL / ORBX_BOW_SERIAL in the child's
    environment, the small form (for which the library builds the grid on the host)."""
    if form == "small":
        monkeypatch.setenv("ORBX_PROJ_SERIAL", "1"); monkeypatch.setenv("ORBX_BOW_SERIAL", "1")
    import importlib.util
    spec = importlib.util.spec_from_file_location("kf_cases", os.path.join(os.path.dirname(__file__), "kf_cases.py"))
    kc = importlib.util.module_from_spec(spec); spec.loader.exec_module(kc)
    gs = kc.gs
    cfg = synth.KITTI
    for seed, L, only_stereo, mono, ori in gs.KF_CASES[:2]:
        ims, V, T1, T2, F12 = gs.kf_case_inputs(seed, L)
        ref, gpu = both(oracle.ref_slam_kf, *ims, cfg["nfeatures"], gs.KITTI_CAM, cfg["bf"], V, T1, T2, F12, check_ori=ori, only_stereo=only_stereo,
                        mono=mono, fuse_scw=fuse_scw, **gs.KF_PARAMS)
        _same(ref, gpu)
        assert ref["bow_kk"][0] > 20 and ref["fuse"][0] > 50 and ref["sim3"][0] > 50
