import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aiko_services_b200 import audio_ops
dev = torch.device("cuda:0")
t = np.arange(16000) / 16000.0
freqs = np.geomspace(50, 7000, 1024)
rng = np.random.default_rng(5000)
win = (3000 * np.sin(2 * np.pi * freqs[:, None] * t[None, :]) + rng.normal(0, 200, (1024, 16000))).astype(np.int16)
aud = [torch.from_numpy(np.roll(win, k, axis=0)).to(dev) for k in range(4)]
def timed(fn, steps=50):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps * 1e3
for bands, top in [(False, False), (True, False), (False, True), (True, True)]:
    print(bands, top, round(timed(lambda i: audio_ops.audio_chain(aud[i % 4], 16000, bands=bands, top=top)), 1), "us")
r = audio_ops.audio_chain(aud[0], 16000)
print("top counts:", r.top_count.float().mean().item() if hasattr(r, "top_count") else type(r))
