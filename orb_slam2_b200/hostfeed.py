"""Host-feed ceiling probe: what bare cudaMemcpyAsync reaches between pinned host memory and this rank's GPU.

The end-to-end leg of bench.py uploads every image of a step (239 MB for 256 KITTI-shaped stereo pairs) and downloads
66 MB of results; with N ranks on one box those copies share the host's memory system and PCIe root complexes.  This
module measures that ceiling in the same job, with no kernel running: H2D alone, D2H alone, and both directions at
once, on 1 or several streams, all ranks between the same two barriers.  bench.py reports the end-to-end leg's copy
rate as a fraction of it; tools/h2d_probe.py runs it stand-alone (under torch.distributed.run) and adds the topology.
"""
import time


def _barrier(dist):
    if dist is not None and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def probe(torch, h2d_bytes, d2h_bytes, streams=(1, 3), reps=6, dist=None, write_combined=False):
    """Returns {"h2d_gbs": {nstreams: GB/s}, "d2h_gbs": {...}, "duplex_gbs": {nstreams: (h2d GB/s, d2h GB/s)}} for THIS rank,
    every figure taken while all ranks run the same copy.  Buffers are split evenly over the streams."""
    dev = torch.device("cuda", torch.cuda.current_device())
    wc = None
    if write_combined:                                  # upload buffer in write-combined pinned memory (orbx_host_alloc)
        from . import api
        wc = api.HostBuffer(h2d_bytes, write_combined=True)
        wc.array[:] = 7
        up_h = torch.from_numpy(wc.array)
    else:
        up_h = torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory()
        up_h.fill_(7)
    up_d = torch.empty(h2d_bytes, dtype=torch.uint8, device=dev)
    dn_d = torch.zeros(d2h_bytes, dtype=torch.uint8, device=dev)
    dn_h = torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory()
    out = {"h2d_gbs": {}, "d2h_gbs": {}, "duplex_gbs": {}}

    def run(ns, up, down):
        sts = [torch.cuda.Stream() for _ in range(ns)]
        def issue():
            for i, s in enumerate(sts):
                with torch.cuda.stream(s):
                    if up:
                        a, b = i * h2d_bytes // ns, (i + 1) * h2d_bytes // ns
                        up_d[a:b].copy_(up_h[a:b], non_blocking=True)
                    if down:
                        a, b = i * d2h_bytes // ns, (i + 1) * d2h_bytes // ns
                        dn_h[a:b].copy_(dn_d[a:b], non_blocking=True)
        issue(); torch.cuda.synchronize()                       # warm-up
        _barrier(dist)
        t0 = time.perf_counter()
        for _ in range(reps):
            issue()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        _barrier(dist)
        return dt

    for ns in streams:
        out["h2d_gbs"][ns] = h2d_bytes * reps / run(ns, True, False) / 1e9
        out["d2h_gbs"][ns] = d2h_bytes * reps / run(ns, False, True) / 1e9
        dt = run(ns, True, True)
        out["duplex_gbs"][ns] = (h2d_bytes * reps / dt / 1e9, d2h_bytes * reps / dt / 1e9)
    del up_h
    if wc is not None:
        wc.close()
    return out


def gather(torch, dist, values):
    """values: list of floats of this rank -> [world][len] nested list (all ranks)."""
    t = torch.tensor(values, dtype=torch.float64, device="cuda")
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return [t.cpu().tolist()]
    outs = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(outs, t)
    return [o.cpu().tolist() for o in outs]
