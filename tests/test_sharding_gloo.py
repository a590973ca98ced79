"""CPU, world_size 2 over gloo: the N>1 host logic -- streams partition across ranks without
overlap, every rank agrees on the owner of a stream, and timings / counters combine as
max / sum over ranks (what bench.py does on N GPUs over NCCL)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from aiko_services_b200 import sharding


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ids = [f"s{k:02d}" for k in range(64)]
        mine = sharding.shard_streams(ids, world, rank)
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        (t_max,), (frames,) = sharding.combine([1.0 + rank], [len(mine) * 32])
        out.put((rank, gathered, t_max, frames))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_ranks_partition_streams_and_combine():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=100) for _ in procs]
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    for rank, gathered, t_max, frames in results:
        all_ids = sorted(gathered[0] + gathered[1])
        assert all_ids == [f"s{k:02d}" for k in range(64)]            # complete
        assert not set(gathered[0]) & set(gathered[1])                # disjoint
        assert t_max == 2.0 and frames == 64 * 32                     # max of timings, sum of counts
    assert results[0][1] == results[1][1]                             # both ranks computed the same partition


def test_owner_is_stable_and_balanced_ids():
    assert sharding.owner_rank("s00", 8) == sharding.owner_rank("s00", 8)
    assert sharding.owner_rank("anything", 1) == 0
    per_rank = sharding.balanced_stream_ids(8, 4)
    assert all(len(v) == 8 for v in per_rank.values())
    for r, ids in per_rank.items():
        assert all(sharding.owner_rank(s, 4) == r for s in ids)
    assert sharding.combine([3.0], [5]) == ([3.0], [5])               # no process group: identity
