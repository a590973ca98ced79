"""The box's host-to-device ceiling: every rank copies a 199 MB batch of 1080p frames from page-locked
memory to its GPU with plain cudaMemcpyAsync (one call per buffer, a ring of 3 buffers, one copy stream),
all ranks at once.  Run alone for one GPU or under torchrun for N; rank 0 prints one JSON line with the
aggregate and per-GPU GB/s -- the bound of the end-to-end numbers of bench.py.

    python benchmarks/h2d_ceiling.py
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 benchmarks/h2d_ceiling.py
"""
import json
import os

import torch
import torch.distributed as dist


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nbytes = 32 * 1080 * 1920 * 3
    host = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(3)]
    for h in host:
        h.fill_(7)
    gpu = [torch.empty(nbytes, dtype=torch.uint8, device=dev) for _ in range(3)]
    stream = torch.cuda.Stream(dev)
    steps = 60

    def run(n):
        with torch.cuda.stream(stream):
            for i in range(n):
                gpu[i % 3].copy_(host[i % 3], non_blocking=True)
    run(6)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    run(steps)
    e1.record(stream)
    torch.cuda.synchronize()
    seconds = torch.tensor([e0.elapsed_time(e1) * 1e-3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(seconds, op=dist.ReduceOp.MAX)
    if int(os.environ.get("RANK", "0")) == 0:
        total = nbytes * steps * world
        print(json.dumps({"n_gpus": world, "aggregate_GBps": total / seconds.item() / 1e9,
                          "per_gpu_GBps": total / seconds.item() / 1e9 / world,
                          "frames_per_s_bound_1080p": total / seconds.item() / (1080 * 1920 * 3),
                          "bytes_per_copy": nbytes, "copies_per_rank": steps,
                          "affinity": sorted(os.sched_getaffinity(0))[:4] + ["..."], "cpus": os.cpu_count()}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
