#!/usr/bin/env python
"""bench.py -- BASELINE.json's headline metric on B200.

Workload (N=1): BASELINE config 2 -- ImageResize(1920x1080 -> 640x480, cv2 bilinear) ->
ImageOverlay (4 boxes + 4 anti-aliased labels) -> RGB->BGR, one stream, 32-frame batches of
synthetic 1080p RGB frames.  A "step" is one pass of that chain over one 32-frame batch.

  value     frames/s, device-resident inputs (CUDA events around K steps, max over ranks)
  e2e       frames/s through the public host API: pinned host frames -> H2D -> chain -> D2H
  roofline  HBM: algorithmic bytes per launch / measured kernel time, vs MEASURED_PEAKS.json
  cpu_baseline / --impl reference: the reference's CPU path (cv2 calls of its elements)

N>1 (torchrun, one rank per GPU): BASELINE config 3 -- independent streams sharded by
stream_id across ranks, no data-path collective; weak scaling (32 frames per rank per step).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

SRC_H, SRC_W, DST_H, DST_W = 1080, 1920, 480, 640
BATCH = 32
# 960 of 1080 source rows carry weight (5,529,600 B) + 921,600 B written once
ALGO_BYTES_PER_FRAME = 960 * SRC_W * 3 + DST_H * DST_W * 3
FALLBACK_HBM_GBS = 6650.0
OVERLAY_KERNEL = "overlay_cells_kernel"
USE_GRAPHS = os.environ.get("AIKO_BENCH_GRAPHS", "1") != "0"   # replay the timed steps as CUDA graphs
LANES = max(1, int(os.environ.get("AIKO_BENCH_LANES", "3")))   # CUDA streams the device-resident loop alternates between
N_ROTATING_BATCHES = 4     # 4 x 199 MB of distinct inputs: each step reads a batch the 126 MB L2 cannot hold

OVERLAY = {
    "objects": [{"name": "person", "confidence": 0.87}, {"name": "bicycle", "confidence": 0.5},
                {"name": "dog", "confidence": 0.33}, {"name": "traffic light", "confidence": 0.99}],
    "rectangles": [{"x": 100, "y": 120, "w": 200, "h": 150}, {"x": 300, "y": 10, "w": 120, "h": 90},
                   {"x": 560, "y": 300, "w": 120, "h": 100}, {"x": 20, "y": 400, "w": 260, "h": 120}],
}
WORKLOAD = ("config2: ImageResize(1920x1080->640x480 cv2 bilinear) -> ImageOverlay(4 boxes + 4 AA labels) "
            "-> RGB2BGR, 1 stream, 32-frame batches")


_IO = "aiko_services.elements.media.image_io"    # the module name existing pipeline definitions use
E2E_DEFINITION = {
    "version": 0, "name": "p_bench", "runtime": "python",
    "graph": ["(ImageResize ImageOverlay ImageConvert ImageOutput)"],
    "parameters": {"resolution": f"{DST_W}x{DST_H}"},
    "elements": [
        {"name": "ImageResize", "input": [{"name": "images", "type": "[image]"}],
         "output": [{"name": "images", "type": "[image]"}], "deploy": {"local": {"module": _IO}}},
        {"name": "ImageOverlay", "input": [{"name": "images", "type": "[image]"}, {"name": "overlay", "type": "dict"}],
         "output": [{"name": "images", "type": "[image]"}], "deploy": {"local": {"module": _IO}}},
        {"name": "ImageConvert", "parameters": {"conversion": "RGB2BGR"},
         "input": [{"name": "images", "type": "[image]"}], "output": [{"name": "images", "type": "[image]"}],
         "deploy": {"local": {"module": _IO}}},
        {"name": "ImageOutput", "input": [{"name": "images", "type": "[image]"}],
         "output": [{"name": "images", "type": "[image]"}], "deploy": {"local": {"module": _IO}}},
    ],
}


def synthetic_batch(seed, n=BATCH):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, (n, SRC_H, SRC_W, 3), dtype=np.uint8)


def natural_frame(seed, h=SRC_H, w=SRC_W):
    """Smooth random field + mild noise: a synthetic frame that compresses like camera content
    (uniform noise, the resize inputs above, would make every JPEG as large as the raw frame)."""
    from PIL import Image
    rng = np.random.default_rng(seed)
    low = (rng.random((h // 40 + 2, w // 40 + 2, 3)) * 255).astype(np.uint8)
    up = np.asarray(Image.fromarray(low, "RGB").resize((w, h), Image.BICUBIC)).astype(np.float32)
    return np.clip(up + rng.normal(0, 4, (h, w, 3)), 0, 255).astype(np.uint8)


def jpeg_blob(seed, restart_rows=0):
    """One 1080p baseline JPEG (quality 90, 4:2:0) as Pillow writes it by default: no restart markers
    (``restart_rows`` = 1: one after every MCU row, libjpeg's restart_in_rows)."""
    import io
    from PIL import Image
    buf = io.BytesIO()
    extra = {"restart_marker_rows": restart_rows} if restart_rows else {}
    Image.fromarray(natural_frame(seed), "RGB").save(buf, "JPEG", quality=90, subsampling=2, **extra)
    return buf.getvalue()


JPEG_WORKLOAD = ("config1: ImageReadFile -> ImageResize -> ImageWriteFile on JPEG files held in host memory: 1080p, quality "
                 "90, 4:2:0, Pillow's default encoder settings (no restart markers) -> decode -> ImageResize PIL branch "
                 "(antialiased bicubic 1920x1080->640x480) -> image.save as JPEG (Pillow's defaults: quality 75, 4:2:0) "
                 "-> file bytes in host memory, 32-frame batches")
JPEG_LANES = int(os.environ.get("AIKO_BENCH_JPEG_LANES", "8"))


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [s for t, s in self.samples if t0 - 0.05 <= t <= t1 + 0.15] or [s for _, s in self.samples]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, f[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------- CPU reference
CPU_ROTATE = 4        # distinct input sets per worker, rotated like the GPU arm's batches
CPU_ROUNDS = 3        # timed rounds of K steps each; the median round is reported
JPEG_SEED = 640_000


def _ref_worker(idx, count, warmup, steps, rounds, barrier, results):
    """One host core of the reference arm.  Its share of every 32-frame batch (``count`` frames,
    CPU_ROTATE distinct sets) is generated BEFORE the barrier; the timed region is the reference
    chain only, on this worker's own clock."""
    import cv2
    from oracle import reference_port
    cv2.setNumThreads(1)
    sets = [list(synthetic_batch(100_000 + 1000 * idx + r, count)) for r in range(CPU_ROTATE)]
    resolution = f"{DST_W}x{DST_H}"
    for i in range(max(warmup, 1)):
        reference_port.chain_resize_overlay_bgr(sets[i % CPU_ROTATE], resolution, OVERLAY)
    times = []
    for _ in range(rounds):
        barrier.wait()
        t0 = time.perf_counter()
        for i in range(steps):
            reference_port.chain_resize_overlay_bgr(sets[i % CPU_ROTATE], resolution, OVERLAY)
        times.append(time.perf_counter() - t0)
    results.put((idx, times))


def cpu_reference_rate(steps, warmup, frames_per_step=BATCH):
    """frames/s of the reference's CPU path on all host cores: one process per core (cv2 threads
    = 1), each running the reference elements' cv2 calls (oracle/reference_port.py) on its shard
    of every batch.  Inputs are resident before the clock starts; a round's time is the SLOWEST
    worker's own perf_counter span over its K steps (process start-up, imports and frame
    generation are outside it).  -> (frames/s of the median round, workers, s per step, rounds)"""
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    workers = max(1, min(cores, frames_per_step))
    per = [frames_per_step // workers + (1 if i < frames_per_step % workers else 0) for i in range(workers)]
    ctx = mp.get_context("spawn")
    barrier, results = ctx.Barrier(workers), ctx.Queue()
    procs = [ctx.Process(target=_ref_worker, args=(i, per[i], warmup, steps, CPU_ROUNDS, barrier, results))
             for i in range(workers)]
    for p in procs:
        p.start()
    times = {}
    deadline = time.time() + 1800
    while len(times) < workers:                      # never wait on a worker that died
        try:
            idx, t = results.get(timeout=2.0)
            times[idx] = t
        except Exception:
            if any(p.exitcode not in (None, 0) for p in procs) or time.time() > deadline:
                for p in procs:
                    p.kill()
                raise RuntimeError("bench.py: a CPU reference worker exited without a result")
    for p in procs:
        p.join()
    round_s = sorted(max(times[i][r] for i in range(workers)) for r in range(CPU_ROUNDS))
    dt = round_s[len(round_s) // 2]
    return frames_per_step * steps / dt, workers, dt / steps, [frames_per_step * steps / t for t in round_s]


def _ref_worker_jpeg(idx, seeds, warmup, steps, rounds, barrier, results):
    """One host core of the JPEG leg: Image.open + resize + save (R/elements/media/image_io.py:184,218,250;
    the decode runs lazily inside resize, as in the reference) on its share of every 32-blob batch."""
    import io
    from PIL import Image
    blobs = [jpeg_blob(s) for s in seeds]
    def step():
        out = None
        for b in blobs:
            out = io.BytesIO()
            Image.open(io.BytesIO(b)).resize((DST_W, DST_H)).save(out, "JPEG")
        return out
    for _ in range(max(warmup, 1)):
        step()
    times = []
    for _ in range(rounds):
        barrier.wait()
        t0 = time.perf_counter()
        for _ in range(steps):
            step()
        times.append(time.perf_counter() - t0)
    results.put((idx, times))


def cpu_reference_rate_jpeg(steps, warmup, frames_per_step=BATCH):
    """frames/s of the reference's CPU path from JPEG bytes (Pillow decode + bicubic resize), one process
    per host core, timed like cpu_reference_rate.  -> (frames/s, workers, rounds' frames/s)"""
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    workers = max(1, min(cores, frames_per_step))
    shards = [[JPEG_SEED + k for k in range(frames_per_step) if k % workers == i] for i in range(workers)]
    ctx = mp.get_context("spawn")
    barrier, results = ctx.Barrier(workers), ctx.Queue()
    procs = [ctx.Process(target=_ref_worker_jpeg, args=(i, shards[i], warmup, steps, CPU_ROUNDS, barrier, results))
             for i in range(workers)]
    for p in procs:
        p.start()
    times = {}
    deadline = time.time() + 1800
    while len(times) < workers:
        try:
            idx, t = results.get(timeout=2.0)
            times[idx] = t
        except Exception:
            if any(p.exitcode not in (None, 0) for p in procs) or time.time() > deadline:
                for p in procs:
                    p.kill()
                raise RuntimeError("bench.py: a CPU JPEG worker exited without a result")
    for p in procs:
        p.join()
    round_s = sorted(max(times[i][r] for i in range(workers)) for r in range(CPU_ROUNDS))
    dt = round_s[len(round_s) // 2]
    return frames_per_step * steps / dt, workers, [frames_per_step * steps / t for t in round_s]


def jpeg_steps(args):
    return max(2, min(args.steps, 10))


def cpu_jpeg_object(args):
    import PIL
    rate, workers, rounds = cpu_reference_rate_jpeg(jpeg_steps(args), min(args.warmup, 2))
    return {"value": rate, "unit": "frames/s", "cores": workers, "kind": "reference library calls",
            "workload": JPEG_WORKLOAD, "rounds_frames_per_s": [round(r, 1) for r in rounds],
            "sample": f"{CPU_ROUNDS} rounds x {jpeg_steps(args)} steps x {BATCH} frames: PIL.Image.open(bytes).resize(...).save(BytesIO, JPEG) "
                      f"(Pillow {PIL.__version__}), one process per core"}


def as_shipped_rate(batches=6):
    """The reference AS SHIPPED: its own ImageResize -> ImageOverlay inside its own PipelineImpl
    (R/main/pipeline.py:618-654,1018-1113), one Python process, cv2's default threading, driven by
    oracle/ref_harness.py from the unmodified copy in baseline/_ref (or /root/reference).  Runs in a
    child process; -> dict or None when the reference is not on this machine."""
    code = ("import json,sys,time;sys.path.insert(0,%r);import numpy as np;from oracle import ref_harness as H\n"
            "import bench,cv2\n"
            "if not H.available():\n    print(json.dumps(None));sys.exit(0)\n"
            "IO='aiko_services.elements.media.image_io'\n"
            "frames=[list(bench.synthetic_batch(9000+i)) for i in range(2)]\n"
            "b=[{'images':frames[i%%2],'overlay':bench.OVERLAY} for i in range(%d)]\n"
            "stamps=[]\n"
            "H.run_reference_pipeline([('ImageResize',IO,['images'],['images'],{'resolution':'%dx%d'}),"
            "('ImageOverlay',IO,['images','overlay'],['images'],None),('ImageOutput',IO,['images'],['images'],None)],"
            "b,timestamps=stamps)\n"
            "n=len(stamps)-1\n"
            "print(json.dumps({'value':bench.BATCH*n/(stamps[-1]-stamps[0]),'batches':n,"
            "'cv2_threads':cv2.getNumThreads()}))\n") % (ROOT, batches + 1, DST_W, DST_H)
    try:
        out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=ROOT)
        last = [l for l in out.stdout.splitlines() if l.startswith("{") or l == "null"]
        return json.loads(last[-1]) if last else None
    except Exception:
        return None


def cpu_baseline_objects(steps, warmup):
    import cv2
    rate, workers, step_s, rounds = cpu_reference_rate(steps, warmup)
    port = {"value": rate, "unit": "frames/s", "cores": workers, "kind": "port",
            "rounds_frames_per_s": [round(r, 1) for r in rounds],
            "sample": f"{CPU_ROUNDS} rounds x {steps} steps x {BATCH} frames (median round; a round is timed by its "
                      f"slowest worker's own clock, inputs resident): the reference elements' cv2 {cv2.__version__} "
                      f"calls (oracle/reference_port.py), one process per core, cv2 threads=1"}
    shipped = as_shipped_rate()
    if shipped:
        shipped = {"value": shipped["value"], "unit": "frames/s", "cores": shipped["cv2_threads"], "kind": "_ref",
                   "sample": f"{shipped['batches']} batches x {BATCH} frames through the unmodified reference's own "
                             f"PipelineImpl and elements (baseline/_ref), one process, cv2 default threads"}
    return port, shipped, step_s


def sharding_ids(world, total=64):
    """Config 3's streams: ids whose owners (crc32(stream_id) % world, sharding.owner_rank) split the
    64 streams evenly over the ranks; -> {rank: [stream ids]}"""
    from aiko_services_b200 import sharding
    return sharding.balanced_stream_ids(total // world, world)


def config_dict(world):
    return {"workload": WORKLOAD, "frames_per_step_per_gpu": BATCH,
            "cuda_streams": f"consecutive batches alternate between {LANES} CUDA streams with their own output buffers",
            "launch": "CUDA graphs of lcm(streams, rotating batches) steps captured from the C-ABI calls, replayed; "
                      "the remainder eagerly" if USE_GRAPHS else "eager C-ABI calls",
            "l2": f"{N_ROTATING_BATCHES} rotating input batches of 199 MB each (> 126 MB L2), no flush",
            "sharding": "independent streams per rank, no collective" if world > 1 else "single GPU"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    port, shipped, step_s = cpu_baseline_objects(args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": "1080p frames/sec through resize+convert+overlay", "value": port["value"],
        "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic", "config": config_dict(world),
        "cpu_baseline": port,
        "e2e": {"value": port["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if shipped:
        line["cpu_baseline_as_shipped"] = shipped
    line["e2e_jpeg"] = cpu_jpeg_object(args)
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- GPU arm
def run_gpu_arm(args):
    if int(os.environ.get("WORLD_SIZE", "1")) == 1 and "LOCAL_RANK" not in os.environ:
        # --gpus 1 on a multi-GPU box: every aiko stream of the e2e leg must land on the one GPU
        # under test (elements pin streams by crc32(stream_id) % visible devices)
        first = os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0] or "0"
        os.environ["CUDA_VISIBLE_DEVICES"] = first
    import torch
    import torch.distributed as dist
    from aiko_services_b200 import ops
    from aiko_services_b200.overlay_plan import build_ops
    from aiko_services_b200.text_atlas import TextAtlas

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the GPU arm has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")      # stdout carries the one JSON line only
        dist.init_process_group("nccl", device_id=device)

    atlas = TextAtlas.shared()
    frame_ops = build_ops(OVERLAY, DST_W, DST_H, "bgr", atlas)
    exp = os.environ.get("AIKO_BENCH_EXPERIMENT_OPS")          # experiments only: "rects" | "glyphs"
    if exp:
        if exp == "faraway":                # four rectangles that touch no row of the frame
            from aiko_services_b200._native import OverlayOp
            frame_ops = [OverlayOp(0, 100 + 50 * k, 10000, 300 + 50 * k, 10100, 0xffff, 0, 0) for k in range(4)]
        else:
            frame_ops = [] if exp == "none" else [o for o in frame_ops if (o.kind == 0) == (exp == "rects")]
    ov = ops.DeviceOverlay.from_host([frame_ops] * BATCH, device, atlas.device(device), width=DST_W, height=DST_H)

    # device-resident inputs: N_ROTATING_BATCHES distinct batches (streams sharded by rank for N>1)
    batches = [torch.from_numpy(synthetic_batch(1000 * rank + i)).to(device) for i in range(N_ROTATING_BATCHES)]
    out = torch.empty((BATCH, DST_H, DST_W, 3), dtype=torch.uint8, device=device)
    stream = torch.cuda.current_stream()
    # Consecutive batches alternate between LANES CUDA streams, each with its own output buffer: the
    # latency-bound overlay pass of batch i then runs beside the HBM-bound resize of batch i + 1, and
    # the launch ramp / tail of one resize hides behind the next (frames are independent, results stay
    # in order per lane).  LANES = 1 is the plain back-to-back sequence (value_single_stream).
    lane_streams = [torch.cuda.Stream(device) for _ in range(LANES)]
    lane_outs = [torch.empty((BATCH, DST_H, DST_W, 3), dtype=torch.uint8, device=device) for _ in range(LANES)]

    exp_ov = None if os.environ.get("AIKO_BENCH_EXPERIMENT_NO_OVERLAY") else ov   # experiments only

    def step(i, lanes=LANES):
        if lanes == 1:
            ops.resize_linear(batches[i % N_ROTATING_BATCHES], DST_W, DST_H, swap_rb=True, overlay=exp_ov, out=out)
        else:
            ops.resize_linear(batches[i % N_ROTATING_BATCHES], DST_W, DST_H, swap_rb=True, overlay=exp_ov,
                              out=lane_outs[i % lanes], stream=lane_streams[i % lanes])

    # The timed loop is issued as CUDA graphs of GRAPH_STEPS consecutive steps (captured once from the very
    # same C-ABI calls, fork / join over the lane streams included) so that the 35 us step is not paced by
    # the ~30 us of host time one call costs through Python; the remainder K % GRAPH_STEPS runs eagerly.
    graph_steps = math.lcm(LANES, N_ROTATING_BATCHES)
    graphs = {}

    def issue(first, count, lanes):
        """Fork from the current stream to the lanes, enqueue steps [first, first + count), join back."""
        cur = torch.cuda.current_stream()
        if lanes > 1:
            fork = torch.cuda.Event()
            fork.record(cur)
            for s in lane_streams[:lanes]:
                s.wait_event(fork)
        for i in range(first, first + count):
            step(i, lanes)
        if lanes > 1:
            for s in lane_streams[:lanes]:
                done = torch.cuda.Event()
                done.record(s)
                cur.wait_event(done)

    def graph_for(lanes, count=None):
        """A CUDA graph of ``count`` steps (default: the full rotation) starting at step index 0."""
        count = graph_steps if count is None else count
        key = (lanes, count)
        if key not in graphs:
            graphs[key] = None
            if USE_GRAPHS and count > 0:
                try:
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=torch.cuda.Stream(device), capture_error_mode="relaxed"):
                        issue(0, count, lanes)
                    graphs[key] = g
                except Exception as error:          # stay measurable: eager launches
                    print(f"bench.py: CUDA graph capture failed ({error!r}); launching eagerly", file=sys.stderr)
                    torch.cuda.synchronize()
        return graphs[key]

    def timed_steps(lanes):
        """-> CUDA events on the main stream around exactly args.steps steps: whole rotations as replays
        of one graph, the remaining K % rotation steps as one replay of a second, shorter graph."""
        full, rest = graph_for(lanes), graph_for(lanes, args.steps % graph_steps)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            ev0.record(stream)
            done = 0
            if full is not None:
                for _ in range(args.steps // graph_steps):
                    full.replay()
                done = args.steps // graph_steps * graph_steps
                if rest is not None:
                    rest.replay()
                    done = args.steps
            if done < args.steps:
                issue(done, args.steps - done, lanes)
            ev1.record(stream)
        return ev0, ev1

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        step(i)
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    for lanes in (LANES, 1):                  # capture (and upload) the graphs outside the timed region
        for g in (graph_for(lanes), graph_for(lanes, args.steps % graph_steps)):
            if g is not None:
                g.replay()
    barrier()
    wall0 = time.time()
    ev0, ev1 = timed_steps(LANES)
    barrier()
    wall1 = time.time()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop(wall0, wall1)
    for i in range(3):
        step(i, 1)
    barrier()
    s0, s1 = timed_steps(1)
    barrier()
    single_ms = s0.elapsed_time(s1)
    # A step launches two kernels: resize3_staged (resize + swap: reads every weighted source
    # sector and writes every output byte -- the algorithmic bytes) and the overlay kernel (the draw
    # list, in place on the L2-resident output).  The dominant kernel is timed INSIDE the real step
    # sequence: the same two launches issued as two C-ABI calls (resize + swap, then the draw list
    # in place -- what the fused call does internally) with an event pair around the first one.
    barrier()
    # The launches are issued eagerly: 10 untimed steps first, so that the stream's queue is full and an event pair
    # brackets the kernel alone (on an idle stream it would bracket the launch latency too), then at least 100 pairs
    # however few steps were asked for.
    kernel_lead, kernel_steps = 10, max(args.steps, 100)
    pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(kernel_steps)]
    for i in range(-kernel_lead, kernel_steps):
        if i >= 0:
            pairs[i][0].record(stream)
        ops.resize_linear(batches[i % N_ROTATING_BATCHES], DST_W, DST_H, swap_rb=True, out=out)
        if i >= 0:
            pairs[i][1].record(stream)
        if exp_ov is not None:
            ops.overlay_draw(out, exp_ov, out=out)
    barrier()
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in pairs]))
    step_ms = ms / args.steps
    single_step_ms = single_ms / args.steps

    # ---- the same step with a NEW overlay every step (labels carry the confidence, so detections
    # change the text every frame, R/elements/media/image_io.py:139): draw-list planning on the host
    # and its upload are inside the timed region.
    def changing_overlay(i):
        o = {"rectangles": [dict(r, x=r["x"] + (i * 7) % 23, y=r["y"] + (i * 5) % 17) for r in OVERLAY["rectangles"]],
             "objects": [dict(ob, confidence=0.01 + ((i * 13 + 29 * k) % 99) / 100.0)
                         for k, ob in enumerate(OVERLAY["objects"])]}
        return o

    rings = [ops.OverlayRing(device) for _ in range(LANES)]      # one ring of draw-list buffers per CUDA stream
    changing_steps = max(3, min(args.steps, 200))
    dicts = [changing_overlay(i) for i in range(changing_steps)]          # the inputs: one new dict per step

    def changing_step(i, overlay):
        # plan the new draw list + resize + swap + draw: ONE call of the library (aiko_b200_resize_linear_objects_u8)
        lane = i % LANES
        rings[lane].resize_draw(batches[i % N_ROTATING_BATCHES], overlay, DST_W, DST_H, swap_rb=True,
                                out=lane_outs[lane] if LANES > 1 else out,
                                stream=lane_streams[lane] if LANES > 1 else stream)

    # Warm-up: a label cut by the image border needs glyph variants rasterised against that border geometry
    # (~50 us per glyph, once, then cached by the library); a long-running pipeline is in that steady state, so
    # the cache is warmed with OTHER overlays (different boxes and confidences) before the timed steps.
    for i in range(max(args.warmup, 400)):
        changing_step(i, changing_overlay(100_000 + i))
    barrier()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record(stream)
    if LANES > 1:
        for s in lane_streams[:LANES]:
            s.wait_event(c0)
    stamps = [] if os.environ.get("AIKO_BENCH_DEBUG") else None
    tc0 = time.perf_counter()
    for i in range(changing_steps):
        changing_step(i, dicts[i])
        if stamps is not None:
            stamps.append(time.perf_counter())
    changing_host_s = time.perf_counter() - tc0          # host time to plan and enqueue
    if stamps:
        d = np.diff(np.array([tc0] + stamps)) * 1e6
        print(f"bench.py debug: changing-overlay host us per step: median {np.median(d):.1f} mean {d.mean():.1f} "
              f"p90 {np.percentile(d, 90):.1f} max {d.max():.1f}; slow steps {np.nonzero(d > 100)[0][:20].tolist()}",
              file=sys.stderr)
    if LANES > 1:
        for s in lane_streams[:LANES]:
            done = torch.cuda.Event()
            done.record(s)
            stream.wait_event(done)
    c1.record(stream)
    torch.cuda.synchronize()
    changing_s = time.perf_counter() - tc0               # until the last result is there
    changing_device_ms = c0.elapsed_time(c1)
    barrier()

    # ---- end to end through the public API: the reference-facing elements on this package's
    # pipeline runtime.  Frames sit in pinned host memory (FrameStager.allocate, where a decoder
    # would write them); every step does cudaMemcpyAsync H2D -> fused kernel -> D2H of the result.
    # Two aiko streams alternate, so the upload of step i overlaps the download of step i-1.
    from aiko_services_b200.elements.media.image_io import stager_for
    from aiko_services_b200.pipeline import Pipeline
    e2e_steps = max(3, min(args.steps, 20))
    pipeline = Pipeline(E2E_DEFINITION)
    e2e_frames = []
    for k in range(2):
        sid = f"r{rank}s{k}"
        assert pipeline.create_stream(sid)
        views = stager_for(pipeline.streams[sid]).allocate(BATCH, (SRC_H, SRC_W, 3))
        src = synthetic_batch(7000 + 10 * rank + k)
        for view, frame in zip(views, src):
            view[...] = frame
        e2e_frames.append((sid, views))
    checksum = 0

    def e2e_submit(i):
        sid, views = e2e_frames[i % 2]
        okay, swag = pipeline.process_frame({"stream_id": sid, "frame_id": i}, {"images": views, "overlay": OVERLAY})
        assert okay
        return swag["images"]

    def e2e_collect(images):
        result = images[0].batch.host_view()          # kernel launch + D2H into pinned memory + sync
        return int(result[0, 0, 0, 0])

    for i in range(4):                                    # warm-up, both streams: plan caches, atlas upload,
        checksum += e2e_collect(e2e_submit(i))            # first launches, pinned result buffers
    barrier()
    t0 = time.perf_counter()
    pending = e2e_submit(0)
    trace = [] if os.environ.get("AIKO_BENCH_E2E_TRACE") else None
    for i in range(1, e2e_steps):
        ta = time.perf_counter()
        nxt = e2e_submit(i)
        tb = time.perf_counter()
        checksum += e2e_collect(pending)
        if trace is not None:
            trace.append((round((tb - ta) * 1e3, 3), round((time.perf_counter() - tb) * 1e3, 3)))
        pending = nxt
    if trace:
        print("e2e per step (submit ms, collect ms):", trace, file=sys.stderr)
    checksum += e2e_collect(pending)
    barrier()
    e2e_s = time.perf_counter() - t0
    # bytes that cross PCIe per step: the stager ships only the source rows that carry weight in the resize
    # (8 of every 9 for 1080 -> 480), unless AIKO_B200_COPY_NEEDED_ROWS=0
    from aiko_services_b200.staging import needed_row_runs
    runs = needed_row_runs(SRC_H, DST_H) if os.environ.get("AIKO_B200_COPY_NEEDED_ROWS", "1") == "1" else None
    rows_copied = sum(runs[0][1::2]) if runs is not None else SRC_H
    h2d = BATCH * rows_copied * SRC_W * 3
    d2h = BATCH * DST_H * DST_W * 3

    # ---- BASELINE config 3: 64 concurrent 1080p streams sharded by stream_id across the ranks
    # (sharding.owner_rank, no collective), one frame per stream per step, all of a rank's streams
    # batched into one launch by Pipeline.process_frames; frames sit in page-locked host memory.
    config3 = None
    try:
        c3_ids = sharding_ids(world)[rank]
        c3_pipeline = Pipeline(E2E_DEFINITION)
        c3_frames = {}
        for k, sid in enumerate(c3_ids):
            assert c3_pipeline.create_stream(sid)
            t = torch.empty((SRC_H, SRC_W, 3), dtype=torch.uint8).pin_memory()
            t.numpy()[...] = synthetic_batch(30_000 + 64 * rank + k, 1)[0]
            c3_frames[sid] = t.numpy()

        def c3_step(i):
            work = [({"stream_id": sid, "frame_id": i}, {"images": [c3_frames[sid]], "overlay": OVERLAY})
                    for sid in c3_ids]
            results = c3_pipeline.process_frames(work)
            assert all(okay for okay, _ in results)
            batches = {id(swag["images"][0].batch): swag["images"][0].batch for _, swag in results}
            return batches.values()

        def c3_collect(batches):
            return sum(int(b.host_view()[0, 0, 0, 0]) for b in batches)

        c3_steps = max(3, min(args.steps, 20))
        for i in range(3):
            c3_collect(c3_step(i))
        barrier()
        t3 = time.perf_counter()
        pending = c3_step(0)
        for i in range(1, c3_steps):
            nxt = c3_step(i)
            c3_collect(pending)
            pending = nxt
        c3_collect(pending)
        barrier()
        config3 = {"seconds": time.perf_counter() - t3, "frames": len(c3_ids) * c3_steps, "steps": c3_steps,
                   "streams": len(c3_ids)}
    except Exception as error:
        config3 = {"error": repr(error)}

    # ---- end to end from JPEG bytes (BASELINE config 1): the frames are born on the device.  32 JPEG
    # blobs per step in page-locked host memory -> this library's decoder (csrc/jpeg_decode.cu: entropy decode
    # one thread per subsequence of the scan, IDCT, upsampling, colour conversion; pixels bit-identical to
    # Pillow's) -> antialiased bicubic resize (the reference's PIL branch) -> D2H of the 640x480 frames.  The
    # entropy kernel runs one CTA per image (32 of the 148 SMs per batch), so JPEG_LANES batches are in flight on
    # separate CUDA streams.
    jpeg_result = None
    try:
        from aiko_services_b200 import jpeg

        def pin(b):
            t = torch.empty(len(b), dtype=torch.uint8).pin_memory()
            t.numpy()[...] = np.frombuffer(b, np.uint8)
            return t

        JPEG_OUT_SLOT = 96 * 1024       # bytes copied back per file (a 640x480 file of this content is ~32 KB)

        def jpeg_leg(pinned_blobs, n_lanes, n_steps, jpeg_out):
            lanes = []
            for k in range(n_lanes):
                st = torch.cuda.Stream(device)
                lane = {"stream": st, "decoder": jpeg.GpuJpegDecoder(device), "busy": False, "oversize": 0,
                        "frames": torch.empty((BATCH, SRC_H, SRC_W, 3), dtype=torch.uint8, device=device),
                        "small": torch.empty((BATCH, DST_H, DST_W, 3), dtype=torch.uint8, device=device)}
                if jpeg_out:
                    lane["encoder"] = jpeg.GpuJpegEncoder(device)
                    lane["files"] = torch.empty((BATCH, JPEG_OUT_SLOT), dtype=torch.uint8).pin_memory()
                    lane["sizes"] = torch.empty(BATCH, dtype=torch.int32).pin_memory()
                else:
                    lane["host"] = torch.empty((BATCH, DST_H, DST_W, 3), dtype=torch.uint8).pin_memory()
                lanes.append(lane)

            def collect(lane):
                lane["stream"].synchronize()
                if jpeg_out and lane["busy"]:
                    sizes = lane["sizes"].numpy()
                    if sizes.min() < 0 or sizes.max() > JPEG_OUT_SLOT:     # (not with this content) a second, larger copy
                        lane["oversize"] += 1

            def submit(i):
                lane = lanes[i % n_lanes]
                if lane["busy"]:
                    collect(lane)
                with torch.cuda.stream(lane["stream"]):
                    lane["decoder"].decode(pinned_blobs, out=lane["frames"], stream=lane["stream"])
                    ops.resize_bicubic_aa(lane["frames"], DST_W, DST_H, out=lane["small"], stream=lane["stream"])
                    if jpeg_out:
                        files, sizes = lane["encoder"].encode_async(lane["small"], stream=lane["stream"])
                        lane["sizes"].copy_(sizes, non_blocking=True)
                        lane["files"].copy_(files[:, :JPEG_OUT_SLOT], non_blocking=True)
                    else:
                        lane["host"].copy_(lane["small"], non_blocking=True)
                lane["busy"] = True

            for i in range(2 * n_lanes):
                submit(i)
            for lane in lanes:
                collect(lane)
            barrier()
            t0 = time.perf_counter()
            for i in range(n_steps):
                submit(i)
            for lane in lanes:
                collect(lane)
            seconds = time.perf_counter() - t0
            barrier()
            for lane in lanes:
                lane["decoder"].check()                  # corrupt entropy data would raise here
            return seconds, lanes

        blobs = [jpeg_blob(JPEG_SEED + 100 * rank + k) for k in range(BATCH)]
        pinned_blobs = [pin(b) for b in blobs]
        n_jpeg = max(2 * JPEG_LANES, min(4 * args.steps, 160))
        jpeg_s, lanes = jpeg_leg(pinned_blobs, JPEG_LANES, n_jpeg, jpeg_out=True)
        sizes = lanes[0]["sizes"].numpy()
        first_file = lanes[0]["files"][0, :int(sizes[0])].numpy().tobytes()
        jpeg_result = {"seconds": jpeg_s, "frames": BATCH * n_jpeg, "steps": n_jpeg, "lanes": JPEG_LANES,
                       "h2d": sum(len(b) for b in blobs), "d2h": BATCH * JPEG_OUT_SLOT + 4 * BATCH,
                       "file_bytes": float(sizes.mean()), "oversize": sum(lane["oversize"] for lane in lanes),
                       "checksum": int(np.frombuffer(first_file, np.uint8).sum())}
        if rank == 0:                                    # the first file against the reference chain's (not timed)
            import io
            from PIL import Image
            ref = io.BytesIO()
            Image.open(io.BytesIO(blobs[0])).resize((DST_W, DST_H)).save(ref, "JPEG")
            jpeg_result["identical_to_reference_file"] = ref.getvalue() == first_file
        del lanes
        rgb_s, lanes = jpeg_leg(pinned_blobs, JPEG_LANES, n_jpeg, jpeg_out=False)
        jpeg_result.update({"rgb_seconds": rgb_s})
        del lanes
    except Exception as error:      # the headline does not depend on the JPEG leg
        jpeg_result = {"error": repr(error)}

    # max over ranks of the device / wall timings, sum over ranks of the frames processed
    from aiko_services_b200 import sharding
    (ms, e2e_s, changing_s), (frames_timed, frames_e2e, frames_changing) = sharding.combine(
        [ms, e2e_s, changing_s], [BATCH * args.steps, BATCH * e2e_steps, BATCH * changing_steps], device=device)
    c3_ok = sharding.combine([0.0 if (config3 and "error" not in config3) else 1.0], [], device=device)[0][0] == 0.0
    c3_totals = sharding.combine([config3["seconds"]], [config3["frames"], config3["streams"]], device=device) if c3_ok else None
    ok_everywhere = sharding.combine([0.0 if (jpeg_result and "error" not in jpeg_result) else 1.0], [], device=device)[0][0] == 0.0
    if ok_everywhere:
        jpeg_totals = sharding.combine([jpeg_result["seconds"], jpeg_result["rgb_seconds"]], [jpeg_result["frames"]],
                                       device=device)
    elif jpeg_result is not None and "error" not in jpeg_result:
        jpeg_result = {"error": "JPEG leg failed on another rank"}

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"
        achieved = ALGO_BYTES_PER_FRAME * BATCH / (kernel_ms * 1e-3) / 1e9
        # DRAM traffic of the dominant kernel: from the committed ncu --set full summary of the CURRENT
        # kernel (profiles/roofline_traffic.json names the capture it was read from), never typed in here
        traffic, traffic_src = None, "no ncu capture on record (profiles/roofline_traffic.json missing)"
        tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tpath):
            t = json.load(open(tpath))
            traffic = int(t["dram_bytes_read"]) + int(t["dram_bytes_write"])
            traffic_src = t["source"]
        line = {
            "metric": "1080p frames/sec through resize+convert+overlay", "value": frames_timed / (ms * 1e-3),
            "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic", "config": config_dict(world),
            "overlay": "static: one draw list built before the timed loop and reused by every step (the best case; "
                       "value_changing_overlay plans and uploads a new list every step)",
            "value_single_stream": {"value": BATCH * args.steps / (single_ms * 1e-3) if world == 1 else None,
                                    "unit": "frames/s", "ms_per_step": single_step_ms,
                                    "note": "the same steps back to back on ONE CUDA stream (rank 0): resize and "
                                            "overlay kernels strictly serialised"},
            "value_changing_overlay": {"value": frames_changing / changing_s, "unit": "frames/s",
                                       "steps": changing_steps,
                                       "host_us_per_step": changing_host_s / changing_steps * 1e6,
                                       "cuda_streams": LANES,
                                       "device_us_per_step": changing_device_ms / changing_steps * 1e3,
                                       "note": "same step, a different overlay dict every step (boxes move, label "
                                               "text changes); host planning + upload inside the timed region (plan + "
                                               "resize + draw = one library call, OverlayRing.resize_draw), wall clock "
                                               "with a final synchronize; warmed up on 400 other overlays (glyph "
                                               "variants of border-clipped labels are rasterised once and cached)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src, "kernel": "resize3_staged",
                         "algorithmic_bytes_per_launch": ALGO_BYTES_PER_FRAME * BATCH, "kernel_ms": kernel_ms,
                         "kernel_ms_how": "mean over max(steps, 100) steps (after 10 untimed ones) of a CUDA-event pair around the resize launch, the "
                                          "overlay launch running between consecutive pairs (real step sequence)",
                         "step_ms": step_ms, "step_frac": ALGO_BYTES_PER_FRAME * BATCH / (step_ms * 1e-3) / 1e9 / peak,
                         "single_stream_step_ms": single_step_ms,
                         "single_stream_step_frac": ALGO_BYTES_PER_FRAME * BATCH / (single_step_ms * 1e-3) / 1e9 / peak,
                         "kernels_per_step": ["resize3_staged", OVERLAY_KERNEL]},
            "e2e": {"value": frames_e2e / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                    "h2d_rows_per_frame": rows_copied,
                    "note": "Pipeline.process_frame over ImageResize -> ImageOverlay -> ImageConvert(RGB2BGR) -> "
                            "ImageOutput; pinned host frames -> cudaMemcpyAsync (only the source rows the resize reads: "
                            "two strided copies per batch) -> fused kernel -> D2H, 2 aiko streams alternating"},
            "gpu_launches": 2 * args.steps * world,
            "clocks": clocks,
        }
        if c3_totals is not None:
            (c3_s,), (c3_frames_total, c3_streams) = c3_totals
            line["config3"] = {"value": c3_frames_total / c3_s, "unit": "frames/s", "streams": int(c3_streams),
                               "steps": config3["steps"],
                               "workload": "config3: 64 concurrent 1080p streams sharded by stream_id over the ranks "
                                           "(no collective), one frame per stream per step, Pipeline.process_frames "
                                           "batches a rank's streams into one launch; end to end from page-locked host "
                                           "frames (one cudaMemcpyAsync per frame) to results in host memory"}
        elif config3 is not None:
            line["config3"] = {"unavailable": config3.get("error", "failed on another rank")}
        if jpeg_result is not None and "error" not in jpeg_result:
            (jpeg_s, rgb_s), (jpeg_frames,) = jpeg_totals
            line["e2e_jpeg"] = {"value": jpeg_frames / jpeg_s, "unit": "frames/s", "workload": JPEG_WORKLOAD,
                                "steps": jpeg_result["steps"], "batches_in_flight": jpeg_result["lanes"],
                                "h2d_bytes_per_step": jpeg_result["h2d"], "d2h_bytes_per_step": jpeg_result["d2h"],
                                "written_file_bytes_mean": jpeg_result["file_bytes"],
                                "files_larger_than_the_copied_slot": jpeg_result["oversize"],
                                "first_file_identical_to_the_reference_chain": jpeg_result.get("identical_to_reference_file"),
                                "value_rgb_out": jpeg_frames / rgb_s,
                                "kernels": "csrc/jpeg_decode.cu: jpeg_unstuff_kernel, jpeg_sync_huffman_kernel (one thread "
                                           "per 448-byte subsequence of the scan, boundary states by iteration), "
                                           "jpeg_idct_kernel, jpeg_colour_kernel; bicubic_mma_kernel; csrc/jpeg_encode.cu: "
                                           "jpeg_planes_kernel, jpeg_fdct_kernel, jpeg_entropy_kernel, jpeg_finish_kernel",
                                "note": "decoded pixels bit-identical to Pillow's decoder, written files byte-identical to "
                                        "Pillow's (tests/test_gpu_jpeg.py); value_rgb_out: the same without the encoder, "
                                        "640x480 RGB frames copied to host memory instead (29.5 MB per step)"}
        elif jpeg_result is not None:
            line["e2e_jpeg"] = {"unavailable": jpeg_result["error"]}
        if world == 1 and not args.no_cpu_baseline:
            port, shipped, _ = cpu_baseline_objects(args.steps, args.warmup)
            line["cpu_baseline"] = port
            if shipped:
                line["cpu_baseline_as_shipped"] = shipped
            if "e2e_jpeg" in line and "value" in line["e2e_jpeg"]:
                line["e2e_jpeg"]["cpu_baseline"] = cpu_jpeg_object(args)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
