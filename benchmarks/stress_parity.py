"""One-off parity sweep over random shapes (more than the test suite runs): every result must equal
the CPU oracle bit for bit.  python benchmarks/stress_parity.py [seed] [cases]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from aiko_services_b200 import ops  # noqa: E402
from oracle import image as O  # noqa: E402


def main():
    seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    cases = int(sys.argv[2]) if len(sys.argv) > 2 else 300
    rng = np.random.default_rng(seed)
    dev = torch.device("cuda:0")
    bad = refused = 0
    t0 = time.time()
    for i in range(cases):
        c = int(rng.choice([1, 3, 3, 4]))
        h, w = int(rng.integers(1, 400)), int(rng.integers(1, 700))
        H, W = int(rng.integers(1, 300)), int(rng.integers(1, 500))
        if rng.random() < 0.6:
            W = max(4, W & ~3)                       # register-resident row path of the bilinear kernel
        if rng.random() < 0.3:
            w = max(16, w & ~15)
        n = int(rng.integers(1, 4))
        frames = rng.integers(0, 256, (n, h, w, c), dtype=np.uint8)
        if c == 1:
            frames = frames[..., 0]
        off = int(rng.integers(0, 9))
        buf = torch.empty(frames.size + off, dtype=torch.uint8, device=dev)
        view = buf[off:].view(frames.shape)
        view.copy_(torch.from_numpy(frames).to(dev))
        kind = "linear" if i % 2 == 0 else "bicubic"
        swap = bool(c == 3 and rng.random() < 0.3 and kind == "linear")
        try:
            got = (ops.resize_linear(view, W, H, swap_rb=swap) if kind == "linear"
                   else ops.resize_bicubic_aa(view, W, H)).cpu().numpy()
        except Exception as error:                  # a documented refusal (shared-memory limits), not a mismatch
            refused += 1
            if "shared memory" not in str(error):
                raise
            continue
        for k in range(n):
            want = O.resize_linear(frames[k], W, H) if kind == "linear" else O.resize_bicubic_aa(frames[k], W, H)
            if swap:
                want = want[..., ::-1]
            if not np.array_equal(got[k], want):
                bad += 1
                print("MISMATCH", kind, frames.shape, "->", (H, W), "offset", off, "swap", swap,
                      "differing bytes", int((got[k] != want).sum()))
                break
    print(f"{cases} cases, {bad} mismatches, {refused} refused, {time.time() - t0:.1f} s")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
