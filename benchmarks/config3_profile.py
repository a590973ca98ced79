"""cProfile of the config-3 step (64 streams, one 1080p frame each, Pipeline.process_frames): where the host time goes."""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from aiko_services_b200.pipeline import Pipeline  # noqa: E402


def main():
    ids = [f"stream_{k:03d}" for k in range(64)]
    p = Pipeline(bench.E2E_DEFINITION)
    frames = {}
    for k, sid in enumerate(ids):
        assert p.create_stream(sid)
        t = torch.empty((bench.SRC_H, bench.SRC_W, 3), dtype=torch.uint8).pin_memory()
        t.numpy()[...] = bench.synthetic_batch(30_000 + k, 1)[0]
        frames[sid] = t.numpy()

    def step(i):
        work = [({"stream_id": sid, "frame_id": i}, {"images": [frames[sid]], "overlay": bench.OVERLAY}) for sid in ids]
        results = p.process_frames(work)
        batches = {id(swag["images"][0].batch): swag["images"][0].batch for _, swag in results}
        return batches.values()

    def collect(batches):
        return sum(int(b.host_view()[0, 0, 0, 0]) for b in batches)

    for i in range(3):
        collect(step(i))
    t0 = time.perf_counter()
    host = 0.0
    pending = step(0)
    for i in range(1, 20):
        h0 = time.perf_counter()
        nxt = step(i)
        host += time.perf_counter() - h0
        collect(pending)
        pending = nxt
    collect(pending)
    dt = time.perf_counter() - t0
    print(f"config3: {64 * 20 / dt:.0f} frames/s, {dt / 20 * 1e3:.2f} ms per step, host part of a step {host / 19 * 1e3:.2f} ms")
    prof = cProfile.Profile()
    prof.enable()
    for i in range(10):
        collect(step(i))
    prof.disable()
    st = pstats.Stats(prof).sort_stats("tottime")
    st.print_stats(12)
    st.print_callers("copy_")
    st.print_callers("synchronize")


if __name__ == "__main__":
    main()
