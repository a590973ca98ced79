"""Multi-GPU: streams are independent, so they shard across GPUs with no data-path collective
(SURVEY.md section 8e).  A stream is pinned to one rank for its lifetime by a stable hash of its
stream_id; torch.distributed (NCCL on GPUs, gloo in the CPU tests) is only used to start
together and to combine per-rank timings and counters."""
import zlib

import torch
import torch.distributed as dist


def owner_rank(stream_id, world_size):
    """Stable across processes and runs (Python's hash() is salted per process)."""
    return zlib.crc32(str(stream_id).encode()) % max(int(world_size), 1)


def shard_streams(stream_ids, world_size, rank):
    return [s for s in stream_ids if owner_rank(s, world_size) == rank]


def balanced_stream_ids(count, world_size, prefix="s"):
    """``count`` stream ids per rank whose owners are exactly balanced (for weak-scaling runs)."""
    per_rank = {r: [] for r in range(world_size)}
    i = 0
    while any(len(v) < count for v in per_rank.values()):
        sid = f"{prefix}{i:04d}"
        r = owner_rank(sid, world_size)
        if len(per_rank[r]) < count:
            per_rank[r].append(sid)
        i += 1
    return per_rank


def combine(values_max=(), values_sum=(), device="cpu"):
    """MAX-reduce ``values_max`` (timings) and SUM-reduce ``values_sum`` (frame counts) over ranks."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return list(values_max), list(values_sum)
    out_max, out_sum = list(values_max), list(values_sum)
    if values_max:
        t = torch.tensor(list(values_max), dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out_max = t.tolist()
    if values_sum:
        t = torch.tensor(list(values_sum), dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        out_sum = t.tolist()
    return out_max, out_sum
