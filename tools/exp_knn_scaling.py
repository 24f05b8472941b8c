"""Side experiment: K7 pairs/s against the number of pairs per call (random descriptors, 2000 x 2000)."""
import sys, numpy as np, torch
sys.path.insert(0, ".")
import orb_slam2_b200 as gx

L = gx.lib()
cap = 2000
rng = np.random.default_rng(0)
st = torch.cuda.current_stream().cuda_stream
for P in (32, 64, 128, 256, 512, 1024):
    base = rng.integers(0, 256, (P, cap, 32), dtype=np.uint8)
    noise = (rng.random((P, cap, 32, 8)) < 0.06)
    other = base ^ np.packbits(noise, axis=-1, bitorder="little")[..., 0]
    dA = torch.from_numpy(base).cuda(); dB = torch.from_numpy(np.ascontiguousarray(other[:, ::-1])).cuda()
    ang = torch.zeros((P, cap), dtype=torch.float32, device="cuda")
    cnt = torch.full((P,), cap, dtype=torch.int32, device="cuda")
    m12 = torch.empty((P, cap), dtype=torch.int32, device="cuda"); nm = torch.empty(P, dtype=torch.int32, device="cuda")
    def step():
        rc = L.orbx_knn_ratio_batch(dA.data_ptr(), ang.data_ptr(), cnt.data_ptr(), dB.data_ptr(), ang.data_ptr(), cnt.data_ptr(),
                                    P, cap, 50, 0.75, 1, 1, m12.data_ptr(), nm.data_ptr(), 1, st)
        assert rc == 0
    for _ in range(2): step()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); [step() for _ in range(5)]; b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 5
    print(f"P={P:5d}  {ms:8.3f} ms/call  {P / ms * 1e3:10.0f} pairs/s  matches/pair {float(nm.float().mean()):.0f}", flush=True)
