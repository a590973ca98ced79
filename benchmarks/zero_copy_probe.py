"""Probe: resize3_staged reading its source rows straight from pinned host memory (UVA) versus
cudaMemcpyAsync H2D followed by the kernel on the device copy."""
import ctypes, sys, os, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # repo root
from aiko_services_b200 import _native, ops

lib = _native.lib()
dev = torch.device("cuda:0")
N, H, W = 32, 1080, 1920
host = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8).pin_memory()
dsrc = torch.empty((N, H, W, 3), dtype=torch.uint8, device=dev)
out_a = torch.empty((N, 480, 640, 3), dtype=torch.uint8, device=dev)
out_b = torch.empty_like(out_a)
stream = torch.cuda.current_stream().cuda_stream

def call(src_ptr, dst):
    rc = lib.aiko_b200_resize_linear_u8(ctypes.c_void_p(src_ptr), N, H, W, 3, W * 3, H * W * 3, ctypes.c_void_p(dst.data_ptr()),
                                        480, 640, 640 * 3, 480 * 640 * 3, 1, None, ctypes.c_void_p(stream))
    assert rc == 0, lib.aiko_b200_last_error()

def timed(fn, steps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps

def staged():
    dsrc.copy_(host, non_blocking=True)
    call(dsrc.data_ptr(), out_a)

def mapped():
    call(host.data_ptr(), out_b)

ms_copy = timed(lambda: dsrc.copy_(host, non_blocking=True))
ms_staged = timed(staged)
ms_mapped = timed(mapped)
torch.cuda.synchronize()
print(f"H2D copy alone {ms_copy:.3f} ms ({host.numel() / ms_copy / 1e6:.1f} GB/s); copy + kernel {ms_staged:.3f} ms; "
      f"kernel on mapped host memory {ms_mapped:.3f} ms; equal outputs: {torch.equal(out_a, out_b)}")
