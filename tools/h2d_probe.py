#!/usr/bin/env python3
"""Host-feed ceiling of this box: bare pinned-memory copies, no kernels.

    python tools/h2d_probe.py                                                         # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/h2d_probe.py

Per rank: pinned buffers of bench.py's sizes (239 MB up, 46 MB down per step of 256 KITTI-shaped stereo pairs), plain
cudaMemcpyAsync on 1 and on 3 streams, H2D alone / D2H alone / both at once, all ranks at the same time
(orb_slam2_b200/hostfeed.py).  Rank 0 prints one JSON line with per-GPU and aggregate GB/s and the GPU <-> PCIe / NUMA map
(`nvidia-smi topo -m`).  Output of this script is kept under profiles/.
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from orb_slam2_b200 import hostfeed
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if "--spread" in sys.argv and torch.cuda.device_count() > world and torch.cuda.device_count() % world == 0:
        local = local * (torch.cuda.device_count() // world)          # bench.py's placement: ranks spread over the visible GPUs
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B, w, h, cap = 256, 1241, 376, 2024
    h2d = 2 * B * w * h
    d2h = 2 * B * cap * 40 + 2 * B * 32 + 2 * B * cap * 4 + B * 4       # bench.py: 8-byte keypoints + descriptors, uRight + depth, counts
    wc = "--wc" in sys.argv
    res = hostfeed.probe(torch, h2d, d2h, streams=(1, 3), reps=8, dist=dist if world > 1 else None, write_combined=wc)
    flat = [res["h2d_gbs"][1], res["h2d_gbs"][3], res["d2h_gbs"][1], res["d2h_gbs"][3], *res["duplex_gbs"][1], *res["duplex_gbs"][3]]
    allv = hostfeed.gather(torch, dist if world > 1 else None, flat)
    if rank == 0:
        names = ["h2d_1stream", "h2d_3streams", "d2h_1stream", "d2h_3streams", "duplex1_h2d", "duplex1_d2h", "duplex3_h2d", "duplex3_d2h"]
        per_gpu = {n: [round(v[i], 2) for v in allv] for i, n in enumerate(names)}
        line = {"probe": "bare cudaMemcpyAsync, pinned host memory%s, all ranks concurrently" % (" (upload buffer write-combined)" if wc else ""), "n_gpus": world, "spread_over_visible_gpus": "--spread" in sys.argv, "h2d_bytes": h2d, "d2h_bytes": d2h,
                "per_gpu_gbs": per_gpu, "aggregate_gbs": {n: round(sum(v), 1) for n, v in per_gpu.items()},
                "min_per_gpu_gbs": {n: min(v) for n, v in per_gpu.items()}}
        try:
            line["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout.splitlines()
        except Exception as exc:
            line["topo"] = repr(exc)
        try:
            line["lscpu"] = [l for l in subprocess.run(["lscpu"], capture_output=True, text=True, timeout=20).stdout.splitlines()
                             if any(k in l for k in ("Model name", "Socket", "NUMA", "CPU(s):", "Thread"))]
        except Exception as exc:
            line["lscpu"] = repr(exc)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
