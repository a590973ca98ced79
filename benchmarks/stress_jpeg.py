"""Random sweep of the JPEG codec against Pillow: sizes 1..400, every subsampling, qualities 1..100, gray, restart
intervals of every kind (decode), natural and noise content.  Encoder: file bytes; decoder: pixels.  Prints the
number of mismatches (expected: 0)."""
import io
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aiko_services_b200 import jpeg  # noqa: E402


def main():
    from PIL import Image
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 400
    dev = torch.device("cuda:0")
    enc, dec = jpeg.GpuJpegEncoder(dev), jpeg.GpuJpegDecoder(dev)
    rng = np.random.default_rng(2026)
    bad_enc = bad_dec = 0
    for t in range(cases):
        h, w = int(rng.integers(1, 400)), int(rng.integers(1, 400))
        if t % 3 == 0:
            frame = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        else:
            low = (rng.random((h // 16 + 2, w // 16 + 2, 3)) * 255).astype(np.uint8)
            frame = np.asarray(Image.fromarray(low, "RGB").resize((w, h), Image.BICUBIC))
            frame = np.clip(frame.astype(np.float32) + rng.normal(0, 5, frame.shape), 0, 255).astype(np.uint8)
        gray = t % 7 == 0
        sub = int(rng.integers(0, 3))
        quality = int(rng.integers(1, 101))
        image = Image.fromarray(frame, "RGB").convert("L") if gray else Image.fromarray(frame, "RGB")
        array = np.ascontiguousarray(np.asarray(image))
        # encoder: the file Pillow writes
        buf = io.BytesIO()
        image.save(buf, "JPEG", quality=quality, **({} if gray else {"subsampling": sub}))
        mine = enc.encode(torch.from_numpy(array[None]).to(dev), quality=quality, subsampling=sub)[0]
        bad_enc += mine != buf.getvalue()
        # decoder: the pixels Pillow reads, for a file with a random restart layout
        kw = [{}, {"restart_marker_rows": int(rng.integers(1, 4))}, {"restart_marker_blocks": int(rng.integers(1, 40))}][t % 3]
        buf = io.BytesIO()
        image.save(buf, "JPEG", quality=quality, **({} if gray else {"subsampling": sub}), **kw)
        blob = buf.getvalue()
        got = dec.decode([blob], check=True).cpu().numpy()[0]
        want = np.asarray(Image.open(io.BytesIO(blob)).convert("RGB"))
        if not np.array_equal(got, want):
            bad_dec += 1
            ys, xs = np.nonzero((got != want).any(axis=2))
            print(f"decoder mismatch: case {t}, {w}x{h}, gray {gray}, subsampling {sub}, quality {quality}, markers {kw}, "
                  f"{len(blob)} bytes, {len(ys)} pixels differ, first at y={ys[0]} x={xs[0]}, max |diff| "
                  f"{np.abs(got.astype(int) - want.astype(int)).max()}")
    print(f"stress_jpeg: {cases} cases, encoder mismatches {bad_enc}, decoder mismatches {bad_dec}")
    # batches: several files of one geometry but different content, quality, tables and restart layout in ONE call
    bad_batch = 0
    for t in range(cases // 10):
        h, w, n = int(rng.integers(1, 300)), int(rng.integers(1, 300)), int(rng.integers(2, 7))
        sub = int(rng.integers(0, 3))
        blobs, frames = [], []
        for k in range(n):
            frame = rng.integers(0, 256, (h, w, 3), dtype=np.uint8) if rng.random() < 0.3 else \
                np.clip(np.asarray(Image.fromarray((rng.random((h // 16 + 2, w // 16 + 2, 3)) * 255).astype(np.uint8), "RGB")
                                   .resize((w, h), Image.BICUBIC)).astype(np.float32) + rng.normal(0, 5, (h, w, 3)), 0, 255).astype(np.uint8)
            kw = [{}, {"restart_marker_rows": 1}, {"restart_marker_blocks": int(rng.integers(1, 10))}, {"optimize": True}][int(rng.integers(0, 4))]
            buf = io.BytesIO()
            Image.fromarray(frame, "RGB").save(buf, "JPEG", quality=int(rng.integers(5, 100)), subsampling=sub, **kw)
            blobs.append(buf.getvalue())
            frames.append(frame)
        got = dec.decode(blobs, check=True).cpu().numpy()
        for k in range(n):
            if not np.array_equal(got[k], np.asarray(Image.open(io.BytesIO(blobs[k])).convert("RGB"))):
                bad_batch += 1
                print(f"batch mismatch: batch {t}, image {k}, {w}x{h}, subsampling {sub}")
        files = enc.encode(torch.from_numpy(np.stack(frames)).to(dev), quality=77, subsampling=sub)
        for k in range(n):
            buf = io.BytesIO()
            Image.fromarray(frames[k], "RGB").save(buf, "JPEG", quality=77, subsampling=sub)
            bad_batch += files[k] != buf.getvalue()
    print(f"stress_jpeg: {cases // 10} batches, mismatches {bad_batch}")


if __name__ == "__main__":
    main()
