"""Host-side sharding of independent frames over ranks (one process per GPU).

The path has no exchange step: every frame goes through extraction and stereo
matching on exactly one GPU.  The only collectives are, after a batch, one all_gather of a small
per-rank record {frames, keypoints, matches, checksum} and a MAX over the ranks' elapsed times.
Works with any torch.distributed backend (nccl on the GPUs, gloo in the CPU tests).
"""
import numpy as np


def shard_range(total_frames, world, rank):
    """Contiguous block of ceil(F/G) frames per rank (the last ranks may get fewer or none)."""
    per = -(-total_frames // world)
    lo = min(rank * per, total_frames)
    return lo, min(lo + per, total_frames)


def make_record(frames, keypoints, matches, checksum):
    return np.array([frames, keypoints, matches, checksum & 0x7FFFFFFFFFFFFFFF], dtype=np.int64)


def gather_records(record, device="cpu"):
    """all_gather of the per-rank record -> array [world][4]; single process returns [1][4]."""
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(np.asarray(record, dtype=np.int64)).to(device)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return t.cpu().numpy()[None]
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return torch.stack(out).cpu().numpy()


def max_over_ranks(values, device="cpu"):
    """Element-wise MAX of a list of floats over all ranks (timing is the slowest rank's)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t.cpu()]


def combine(records):
    """Whole-job totals: sums of the counters, xor of the checksums."""
    r = np.asarray(records, dtype=np.int64)
    return dict(frames=int(r[:, 0].sum()), keypoints=int(r[:, 1].sum()), matches=int(r[:, 2].sum()),
                checksum=int(np.bitwise_xor.reduce(r[:, 3])))
