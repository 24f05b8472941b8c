"""The allocation-free brute-force matcher (orbx_knn_ratio_noalloc) against the regular batched
entry point (orbx_knn_ratio_batch), both called through the C ABI, and against the CPU oracle's
restatement of the reference loop (oracle.knn_ratio), so that a fault the two kernels shared would
still show.

Some users of this library may not allocate device memory while matching: a tracking thread on a
frame deadline cannot risk the device-wide lock behind cudaMalloc/cudaFree, and a process that
gave all its device memory to a pool cannot ask the driver for more.  For them the scratch of the
matcher lives in a fixed array inside the library, shared by every call on a GPU and never
cleared between calls.  So the check is made on two calls in a row with different inputs (the
second smaller than the first): whatever the first call left in the array must not leak into the
second.  Match indices are integers, so the comparison is exact.

This file sorts last on purpose: if it fails under `pytest -x`, every other test has already run.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _make_pair(seed, nA, nB, planted):
    """Random 256-bit descriptors; `planted` A rows are noisy copies of distinct B rows."""
    rng = np.random.default_rng(seed)
    dB = rng.integers(0, 256, (nB, 32), dtype=np.uint8)
    dA = rng.integers(0, 256, (nA, 32), dtype=np.uint8)
    src = rng.permutation(nB)[:planted]
    dst = rng.permutation(nA)[:planted]
    for a, b in zip(dst, src):
        row = dB[b].copy()
        for bit in rng.integers(0, 256, int(rng.integers(0, 40))):      # up to ~40 flipped bits
            row[bit >> 3] ^= np.uint8(1 << (bit & 7))
        dA[a] = row
    angB = rng.uniform(0, 360, nB).astype(np.float32)
    angA = rng.uniform(0, 360, nA).astype(np.float32)
    angA[dst] = (angB[src] + rng.normal(0, 4, planted)).astype(np.float32) % np.float32(360)
    return dA, angA, dB, angB


def _regular(lib, dA, angA, dB, angB, th, ratio, ori, strict):
    nA, nB = len(dA), len(dB)
    cap = max(nA, nB, 1)
    pa = np.zeros((cap, 32), np.uint8); pa[:nA] = dA
    pb = np.zeros((cap, 32), np.uint8); pb[:nB] = dB
    aa = np.zeros(cap, np.float32); aa[:nA] = angA
    ab = np.zeros(cap, np.float32); ab[:nB] = angB
    cA = np.array([nA], np.int32); cB = np.array([nB], np.int32)
    m12 = np.zeros(cap, np.int32); nm = np.zeros(1, np.int32)
    st = lib.orbx_knn_ratio_batch(pa.ctypes.data, aa.ctypes.data, cA.ctypes.data, pb.ctypes.data, ab.ctypes.data,
                                  cB.ctypes.data, 1, cap, th, ratio, int(ori), int(strict), m12.ctypes.data, nm.ctypes.data, 0, None)
    assert st == 0, lib.orbx_last_error_string()
    return m12[:nA].copy(), int(nm[0])


def _noalloc(lib, torch, dA, angA, dB, angB, th, ratio, ori, strict):
    nA, nB = len(dA), len(dB)
    tA = torch.from_numpy(np.ascontiguousarray(dA)).cuda(); tB = torch.from_numpy(np.ascontiguousarray(dB)).cuda()
    ta = torch.from_numpy(angA).cuda(); tb = torch.from_numpy(angB).cuda()
    m12 = torch.full((max(nA, 1),), -7, dtype=torch.int32, device="cuda"); nm = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    st = lib.orbx_knn_ratio_noalloc(tA.data_ptr(), ta.data_ptr(), nA, tB.data_ptr(), tb.data_ptr(), nB, th, ratio,
                                    int(ori), int(strict), m12.data_ptr(), nm.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert st == 0, lib.orbx_last_error_string()
    torch.cuda.synchronize()
    return m12.cpu().numpy()[:nA], int(nm.cpu()[0])


def test_noalloc_matches_regular_entry_point_twice_in_a_row(oracle):
    import torch
    assert torch.cuda.is_available()
    from orb_slam2_b200 import api
    lib = api.lib()
    # call 1: large, fills the static scratch; call 2: smaller and different, must not see call 1's leftovers
    cases = [(_make_pair(1, 2000, 2000, 900), 50, 0.75, True, True),
             (_make_pair(2, 700, 1100, 300), 50, 0.9, True, False),
             (_make_pair(3, 1500, 600, 250), 60, 0.8, False, True)]
    for (dA, angA, dB, angB), th, ratio, ori, strict in cases:
        ref_m, ref_n = _regular(lib, dA, angA, dB, angB, th, ratio, ori, strict)
        got_m, got_n = _noalloc(lib, torch, dA, angA, dB, angB, th, ratio, ori, strict)
        assert got_n == ref_n
        assert np.array_equal(got_m, ref_m)
        cpu_m, cpu_n = oracle.knn_ratio(dA, angA, dB, angB, th, ratio, ori, strict)      # src/ORBmatcher.cc:522-655, one bucket
        assert got_n == cpu_n and np.array_equal(got_m, cpu_m)
        assert ref_n > 50                      # the comparison is not vacuous
    # same input again: identical output although the scratch now holds its own stale state
    (dA, angA, dB, angB), th, ratio, ori, strict = cases[0]
    a = _noalloc(lib, torch, dA, angA, dB, angB, th, ratio, ori, strict)
    b = _noalloc(lib, torch, dA, angA, dB, angB, th, ratio, ori, strict)
    assert a[1] == b[1] and np.array_equal(a[0], b[0])


def test_noalloc_rejects_oversize_and_handles_empty():
    import torch
    from orb_slam2_b200 import api
    lib = api.lib()
    big = torch.zeros((4097, 32), dtype=torch.uint8, device="cuda"); ang = torch.zeros(4097, dtype=torch.float32, device="cuda")
    out = torch.zeros(4097, dtype=torch.int32, device="cuda"); nm = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = lib.orbx_knn_ratio_noalloc(big.data_ptr(), ang.data_ptr(), 4097, big.data_ptr(), ang.data_ptr(), 10, 50, 0.75, 1, 1,
                                    out.data_ptr(), nm.data_ptr(), None)
    assert st == api.ERR_CAPACITY and b"4096" in lib.orbx_last_error_string()
    for bad_th in (-1, 256, 1000):              # 256 is the scan's "no candidate" distance: it must not pass as a threshold
        st = lib.orbx_knn_ratio_noalloc(big.data_ptr(), ang.data_ptr(), 5, big.data_ptr(), ang.data_ptr(), 0, bad_th, 1.5, 1, 0,
                                        out.data_ptr(), nm.data_ptr(), None)
        assert st == api.ERR_BAD_ARG
    st = lib.orbx_knn_ratio_noalloc(big.data_ptr(), ang.data_ptr(), 5, big.data_ptr(), ang.data_ptr(), 0, 50, 0.75, 1, 1,
                                    out.data_ptr(), nm.data_ptr(), None)
    assert st == 0
    torch.cuda.synchronize()
    assert int(nm.cpu()[0]) == 0 and out[:5].cpu().tolist() == [-1] * 5
