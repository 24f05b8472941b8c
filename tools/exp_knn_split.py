#!/usr/bin/env python3
"""Config 5 (2000 x 2000 kNN + ratio) and config 4 (SearchForInitialization) alone, a few calls: run under an ncu launch list to
see how the step splits over the kernels (k_knn_topk / k_knn_resolve, k_search_init)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import orb_slam2_b200 as gx
from orb_slam2_b200 import synth
L = gx.lib()
w, h, P = 1241, 376, 128
G = gx.ORBextractor(2000, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2 * P)
cap = G.kp_cap
mp = [synth.mono_pair(800000 + i, w, h)[:2] for i in range(8)]
batch = np.stack([mp[i % 8][0] for i in range(P)] + [mp[i % 8][1] for i in range(P)])
G.extract_batch(torch.from_numpy(batch).cuda(), 0, 0)
kps = torch.empty((2 * P, cap, 28), dtype=torch.uint8, device="cuda"); desc = torch.empty((2 * P, cap, 32), dtype=torch.uint8, device="cuda")
cnt = torch.empty(2 * P, dtype=torch.int32, device="cuda")
G.fetch_into(0, 2 * P, kps, desc, cnt, 0)
torch.cuda.synchronize()
ang = kps.view(torch.float32).reshape(2 * P, cap, 7)[:, :, 3].contiguous()
m12 = torch.empty((P, cap), dtype=torch.int32, device="cuda"); nmm = torch.empty(P, dtype=torch.int32, device="cuda")
def knn():
    rc = L.orbx_knn_ratio_batch(desc.data_ptr(), ang.data_ptr(), cnt.data_ptr(), desc[P:].data_ptr(), ang[P:].data_ptr(), cnt[P:].data_ptr(),
                                P, cap, 50, 0.75, 1, 1, m12.data_ptr(), nmm.data_ptr(), 1, 0)
    assert rc == 0
for _ in range(3): knn()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5): knn()
torch.cuda.synchronize(); print("knn + ratio: %.3f ms per %d pairs" % ((time.perf_counter() - t0) / 5 * 1e3, P))
gb = torch.tensor([0, w, 0, h, 64.0 / w, 48.0 / h], dtype=torch.float32, device="cuda")
prev0 = kps.view(torch.float32).reshape(2 * P, cap, 7)[:P, :, :2].contiguous()
def init():
    prev = prev0.clone()
    rc = L.orbx_search_init_batch(kps.data_ptr(), desc.data_ptr(), cnt.data_ptr(), kps[P:].data_ptr(), desc[P:].data_ptr(), cnt[P:].data_ptr(),
                                  P, cap, gb.data_ptr(), prev.data_ptr(), 0.9, 100, 1, m12.data_ptr(), nmm.data_ptr(), 1, 0)
    assert rc == 0
for _ in range(3): init()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5): init()
torch.cuda.synchronize(); print("SearchForInitialization (2000 features): %.3f ms per %d pairs" % ((time.perf_counter() - t0) / 5 * 1e3, P))
