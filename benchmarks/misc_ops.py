import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aiko_services_b200 import ops
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
hd = [torch.randint(0, 256, (32, 1080, 1920, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(4)]
def timed(fn, steps=50):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps * 1e3
out = torch.empty_like(hd[0])
print("flip vertical", round(timed(lambda i: ops.flip(hd[i % 4], 0, out=out)), 1))
print("flip both", round(timed(lambda i: ops.flip(hd[i % 4], -1, out=out)), 1))
print("crop 1280x720", round(timed(lambda i: ops.crop(hd[i % 4], 320, 180, 1280, 720)), 1))
gray = torch.empty((32, 1080, 1920), dtype=torch.uint8, device=dev)
ops.bgr2gray(hd[0], out=gray)
print("gray2bgr", round(timed(lambda i: ops.gray2bgr(gray, out=out)), 1))
up = torch.empty((8, 1080, 1920, 3), dtype=torch.uint8, device=dev)
small = hd[0][:8, :720, :1280].contiguous()
print("720p->1080p x8", round(timed(lambda i: ops.resize_linear(small, 1920, 1080, out=up)), 1))
o720 = torch.empty((32, 720, 1280, 3), dtype=torch.uint8, device=dev)
print("1080p->720p x32", round(timed(lambda i: ops.resize_linear(hd[i % 4], 1280, 720, out=o720)), 1))
