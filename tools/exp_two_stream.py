import sys, numpy as np, torch, time
sys.path.insert(0, '.')
import orb_slam2_b200 as gx
from orb_slam2_b200 import synth
cfg = synth.KITTI; w, h = cfg['w'], cfg['h']
pairs = [synth.stereo_pair(i, w, h)[:2] for i in range(32)]
def mk(B): 
    return torch.from_numpy(np.stack([pairs[i % 32][0] for i in range(B)] + [pairs[i % 32][1] for i in range(B)])).cuda()
def run(nstreams, B, steps=20):
    Bs = B // nstreams
    G = [gx.ORBextractor(2000, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=2 * Bs) for _ in range(nstreams)]
    st = [torch.cuda.Stream() for _ in range(nstreams)]
    dev = mk(Bs); cap = G[0].kp_cap
    outs = [(torch.empty((Bs, cap), dtype=torch.float32, device='cuda'), torch.empty((Bs, cap), dtype=torch.float32, device='cuda'), torch.empty(Bs, dtype=torch.int32, device='cuda')) for _ in range(nstreams)]
    def step():
        for k in range(nstreams):
            G[k].extract_batch(dev, 0, st[k].cuda_stream)
            gx.ComputeStereoMatches(G[k], G[k], cfg['bf'], cfg['fx'], n=Bs, slotL0=0, slotR0=Bs, out=outs[k], stream=st[k].cuda_stream)
    for _ in range(3): step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps): step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    for g in G: g.close()
    return B * steps / dt
for ns, B in [(1, 128), (2, 128), (4, 128), (2, 256), (4, 256), (1, 256)]:
    print(ns, 'streams, total batch', B, '->', round(run(ns, B)), 'stereo frames/s')
