"""N>1 host logic on CPU: two gloo ranks shard a frame range, each "processes" its frames (a hash of
the synthetic image stands in for the GPU work), and the gathered records must cover every frame
exactly once and agree with a single-process run."""
import hashlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from orb_slam2_b200 import sharding, synth

TOTAL = 11


def _frame_digest(i):
    img = synth.mono_frame(i, 96, 64)
    return int.from_bytes(hashlib.sha256(img.tobytes()).digest()[:8], "little") & 0x7FFFFFFFFFFFFFFF, int(img.sum() % 1000)


def _work(lo, hi):
    frames, kps, matches, chk = 0, 0, 0, 0
    for i in range(lo, hi):
        d, k = _frame_digest(i)
        frames += 1; kps += k; matches += k // 3; chk ^= d
    return sharding.make_record(frames, kps, matches, chk)


def _rank_main(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_range(TOTAL, world, rank)
    rec = _work(lo, hi)
    allrec = sharding.gather_records(rec)
    tmax = sharding.max_over_ranks([float(rank + 1), 5.0 - rank])
    q.put((rank, (lo, hi), allrec.tolist(), tmax))
    dist.destroy_process_group()


def test_shard_ranges_partition_the_frames():
    for total in (0, 1, 7, 64, 4096):
        for world in (1, 2, 3, 4, 8):
            spans = [sharding.shard_range(total, world, r) for r in range(world)]
            covered = [i for lo, hi in spans for i in range(lo, hi)]
            assert covered == list(range(total))
            assert max(hi - lo for lo, hi in spans) == (-(-total // world) if total else 0)
    assert sharding.shard_range(4096, 8, 3) == (1536, 2048)          # BASELINE config 3: 512 frames per GPU


def test_two_gloo_ranks_cover_every_frame_once():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    single = sharding.combine(_work(0, TOTAL)[None])
    for rank, span, allrec, tmax in out:
        assert sharding.combine(allrec) == single                     # every rank sees the same whole-job totals
        assert tmax == [2.0, 5.0]                                     # slowest rank wins
    assert out[0][1] == (0, 6) and out[1][1] == (6, 11)
    assert single["frames"] == TOTAL


def test_single_process_paths_need_no_process_group():
    rec = sharding.make_record(3, 10, 4, 0xFFFFFFFFFFFFFFFF)
    assert sharding.gather_records(rec).shape == (1, 4)
    assert sharding.max_over_ranks([1.5, 2.5]) == [1.5, 2.5]
    assert sharding.combine([rec])["checksum"] == 0x7FFFFFFFFFFFFFFF
