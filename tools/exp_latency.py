"""Single-frame latency of the drop-in calls (host image in, host keypoints/descriptors out)."""
import sys, time, numpy as np
sys.path.insert(0, '.')
import orb_slam2_b200 as gx
from orb_slam2_b200 import synth
cfg = synth.KITTI; w, h = cfg['w'], cfg['h']
L, R, _ = synth.stereo_pair(1, w, h)
GL = gx.ORBextractor(2000, 1.2, 8, 20, 7, max_w=w, max_h=h); GR = gx.ORBextractor(2000, 1.2, 8, 20, 7, max_w=w, max_h=h)
for _ in range(5): GL(L); GR(R); gx.ComputeStereoMatches(GL, GR, cfg['bf'], cfg['fx'])
t = []
for _ in range(50):
    t0 = time.perf_counter(); GL(L); t1 = time.perf_counter(); GR(R); t2 = time.perf_counter()
    gx.ComputeStereoMatches(GL, GR, cfg['bf'], cfg['fx']); t3 = time.perf_counter()
    t.append((t1 - t0, t2 - t1, t3 - t2))
t = np.array(t) * 1e3
print('operator() one KITTI image: median %.3f ms (min %.3f); ComputeStereoMatches: median %.3f ms; stereo frame total: %.3f ms' %
      (np.median(t[:, 0]), t[:, 0].min(), np.median(t[:, 2]), np.median(t.sum(1))))
