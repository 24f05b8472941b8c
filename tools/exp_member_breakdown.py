#!/usr/bin/env python3
"""Where does a drop-in ORBmatcher member call spend its time?  (GPU box)

Runs the kf-case latency driver (oracle/ref_slam_cmds.inc:cmd_kfbench) of the drop-in binary twice: plain (p50 per member)
and under an ncu launch list (`gpu__time_duration.sum`), then prints per kernel the number of launches and the summed /
mean duration.  Every member is called reps + 2 times, so (kernel time of a member's kernels) / (reps + 2) beside the
member's p50 says how much of a call is the device and how much is flattening, staging, launching and synchronising.
Output goes to stdout; keep it under profiles/ if it is quoted.
"""
import csv
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    from oracle import pyoracle as O
    import gen_golden_slam as gs
    from orb_slam2_b200 import synth
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    cfg = synth.KITTI
    import bench
    ims, V, T1, T2, F12 = gs.kf_case_inputs(5, 5)
    kw = dict(reps=reps, check_ori=True, **gs.KF_PARAMS)
    drop_bin = os.path.join(os.path.dirname(O.REF_SLAM_BIN), "orbref_slam_gpu")
    plain, sizes = O.ref_slam_kfbench(*ims, cfg["nfeatures"], bench.KITTI_CAM, cfg["bf"], V, T1, T2, F12, binary=drop_bin, **kw)
    print("sizes (n1, n2, map points 1, map points 2):", sizes)
    for k, v in plain.items():
        print("%-38s p50 %.3f ms  p99 %.3f ms  result %d" % (k, v[0], v[1], v[2]))
    with tempfile.TemporaryDirectory() as td:
        log = os.path.join(td, "launches.csv")
        prefix = ["ncu", "--metrics", "gpu__time_duration.sum", "--clock-control", "none", "--csv", "--log-file", log]
        O.ref_slam_kfbench(*ims, cfg["nfeatures"], bench.KITTI_CAM, cfg["bf"], V, T1, T2, F12, binary=drop_bin, prefix=prefix, **kw)
        rows = [r for r in csv.reader(l for l in open(log) if l.startswith('"'))]
    hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
    mul = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}
    order, agg = [], {}
    for r in rows[1:]:
        name = r[ki].split("(")[0].replace("void ", "").replace("orbx::", "")
        us = float(r[vi].replace(",", "")) * mul[r[ui]]
        if name not in agg:
            agg[name] = [0, 0.0]; order.append(name)
        agg[name][0] += 1; agg[name][1] += us
    print("\nkernel (in order of first launch), launches, total us, mean us; each member is called %d times" % (reps + 2))
    for n in order:
        print("%-44s %5d %10.1f %8.1f" % (n, agg[n][0], agg[n][1], agg[n][1] / agg[n][0]))


if __name__ == "__main__":
    main()
