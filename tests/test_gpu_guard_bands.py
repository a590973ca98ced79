"""GPU: out-of-bounds WRITES of the JPEG codec kernels would land in guard bands (compute-sanitizer is closed on this
pool): every device buffer the C ABI is given -- destination frames, workspace, file slots, sizes -- sits between two
bands of a known pattern inside one larger allocation, and the bands must come back untouched, for ordinary files,
for files of awkward sizes and for corrupt ones."""
import ctypes
import io

import numpy as np
import pytest
import torch

from aiko_services_b200 import _native

pytestmark = pytest.mark.gpu
GUARD = 1 << 16
PATTERN = 0xC3


@pytest.fixture(scope="module")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda:0")


class Guarded:
    """``nbytes`` of device memory (256-byte aligned) between two guard bands."""

    def __init__(self, nbytes, device):
        self.nbytes = int(nbytes)
        self.whole = torch.full((GUARD + self.nbytes + 255 + GUARD,), PATTERN, dtype=torch.uint8, device=device)
        base = self.whole.data_ptr() + GUARD
        self.offset = GUARD + (-base) % 256
        self.ptr = self.whole.data_ptr() + self.offset

    def view(self):
        return self.whole[self.offset:self.offset + self.nbytes]

    def intact(self):
        before = self.whole[:self.offset]
        after = self.whole[self.offset + self.nbytes:]
        return bool((before == PATTERN).all()) and bool((after == PATTERN).all())


def _files(rng):
    from PIL import Image
    out = []
    for (h, w, sub, kw) in ((270, 480, 2, {}), (67, 101, 1, {"restart_marker_rows": 1}), (9, 250, 0, {"restart_marker_blocks": 1}),
                            (1, 1, 2, {}), (600, 3, 2, {}), (64, 64, 2, {"optimize": True})):
        frame = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        buf = io.BytesIO()
        Image.fromarray(frame, "RGB").save(buf, "JPEG", quality=int(rng.integers(20, 100)), subsampling=sub, **kw)
        out.append((buf.getvalue(), frame))
    return out


def test_decoder_writes_stay_inside_its_buffers(cuda):
    lib = _native.lib()
    rng = np.random.default_rng(12)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for blob, _ in _files(rng):
        damaged = bytearray(blob)
        for _ in range(12):
            damaged[int(rng.integers(len(blob) * 2 // 3, len(blob)))] = int(rng.integers(256))
        for data in (blob, bytes(damaged), blob[:len(blob) * 3 // 4]):
            vals = [ctypes.c_int() for _ in range(6)]
            if lib.aiko_b200_jpeg_gpu_info_host(data, len(data), *[ctypes.byref(v) for v in vals]) != 0:
                continue
            w, h, comps, hmax, vmax, restart = (v.value for v in vals)
            mcus = -(-w // (8 * hmax)) * -(-h // (8 * vmax))
            segments = -(-mcus // restart) if restart > 0 else 1
            n = 2
            capacity = 2 * len(data) + 64
            frames = Guarded(n * h * w * 3, cuda)
            workspace = Guarded(lib.aiko_b200_jpeg_gpu_workspace_bytes(n, w, h, comps, hmax, vmax, capacity, segments), cuda)
            staging = torch.empty(lib.aiko_b200_jpeg_gpu_staging_bytes(n), dtype=torch.uint8).pin_memory()
            ptrs = (ctypes.c_void_p * n)(*[ctypes.cast(ctypes.c_char_p(data), ctypes.c_void_p).value] * n)
            lens = (ctypes.c_size_t * n)(*[len(data)] * n)
            rc = lib.aiko_b200_jpeg_gpu_decode_rgb_u8(ptrs, lens, n, frames.ptr, h, w, h * w * 3, staging.data_ptr(), staging.numel(),
                                                      workspace.ptr, workspace.nbytes, capacity, segments, stream)
            torch.cuda.synchronize()
            assert rc in (0, _native.ERR_INVALID_ARGUMENT, _native.ERR_UNSUPPORTED)
            assert frames.intact() and workspace.intact(), f"{w}x{h}: a write outside the decoder's buffers"


def test_encoder_writes_stay_inside_its_buffers(cuda):
    lib = _native.lib()
    rng = np.random.default_rng(13)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for (h, w, channels, hs, vs, quality, room) in ((270, 480, 3, 2, 2, 75, 1.0), (67, 101, 3, 2, 1, 100, 2.0), (9, 250, 3, 1, 1, 100, 2.0),
                                                    (1, 1, 3, 2, 2, 50, 1.0), (64, 64, 1, 1, 1, 90, 1.0), (64, 64, 3, 1, 1, 100, 0.05)):
        n = 3
        src = torch.from_numpy(rng.integers(0, 256, (n, h, w, channels), dtype=np.uint8)).to(cuda)
        scan_capacity = max(int(h * w * channels * room), 64)         # (the last case: far too small -- sizes must say -1)
        out_stride = scan_capacity + 2048
        workspace = Guarded(lib.aiko_b200_jpeg_gpu_encode_workspace_bytes(n, w, h, channels, hs, vs, scan_capacity), cuda)
        files = Guarded(n * out_stride, cuda)
        sizes = Guarded(4 * n, cuda)
        rc = lib.aiko_b200_jpeg_gpu_encode_u8(src.data_ptr(), n, h, w, channels, h * w * channels, quality, hs, vs, workspace.ptr,
                                              workspace.nbytes, scan_capacity, files.ptr, out_stride, sizes.ptr, stream)
        torch.cuda.synchronize()
        assert rc == 0
        assert workspace.intact() and files.intact() and sizes.intact(), f"{w}x{h}: a write outside the encoder's buffers"
        lengths = sizes.view().view(torch.int32).cpu().tolist()
        assert all(v == -1 or 0 < v <= out_stride for v in lengths), lengths
        if room < 0.1:
            assert lengths == [-1] * n


def test_overlay_boxes_and_alpha_paste_stay_inside_the_frames(cuda):
    """The device-side overlay planner (op records, offsets, status) and the alpha paste with boxes far outside
    the frame, huge boxes and pastes clipped by every border."""
    from aiko_services_b200 import ops
    from aiko_services_b200.text_atlas import TextAtlas
    lib = _native.lib()
    rng = np.random.default_rng(14)
    stream = torch.cuda.current_stream().cuda_stream
    atlas = TextAtlas.shared().device(cuda)
    keys, firsts, vboxes, n_variants = TextAtlas.shared().variants(cuda)
    names = ["person", "a much longer class name, 39 chars.."]
    det_names, stride = ops.DeviceDetections(torch.zeros((1, 1, 4), dtype=torch.int32, device=cuda),
                                             torch.zeros((1, 1), dtype=torch.int32, device=cuda),
                                             torch.zeros((1, 1), dtype=torch.float32, device=cuda), names).name_table(cuda)
    for trial in range(12):
        n, h, w = int(rng.integers(1, 4)), int(rng.integers(20, 300)), int(rng.integers(20, 400))
        lists, slots = (1 if trial % 2 else n), int(rng.integers(1, 40))
        boxes = torch.from_numpy(rng.integers(-5000, 5000, (lists, slots, 4)).astype(np.int32)).to(cuda)
        if trial % 3 == 0:
            boxes[..., 2:] = 1 << 30                                            # absurd sizes
        cls = torch.from_numpy(rng.integers(-3, 5, (lists, slots)).astype(np.int32)).to(cuda)
        conf = torch.from_numpy((rng.random((lists, slots)) * 3 - 1).astype(np.float32)).to(cuda)
        frames = Guarded(n * h * w * 3, cuda)
        workspace = Guarded(lib.aiko_b200_overlay_boxes_workspace_bytes(lists, slots, n), cuda)
        status = Guarded(4, cuda)
        rc = lib.aiko_b200_overlay_boxes_u8(frames.ptr, n, h, w, h * w * 3, trial & 1, boxes.data_ptr(), cls.data_ptr(),
                                            conf.data_ptr(), lists, slots, det_names.data_ptr(), len(names), stride,
                                            atlas.patch_ids.data_ptr(), atlas.luts.data_ptr(), atlas.n_patches, keys.data_ptr(),
                                            firsts.data_ptr(), vboxes.data_ptr(), n_variants, workspace.ptr, workspace.nbytes,
                                            status.ptr, ctypes.c_void_p(stream))
        torch.cuda.synchronize()
        assert rc == 0
        assert frames.intact() and workspace.intact() and status.intact(), f"trial {trial}: overlay_boxes wrote outside its buffers"
        # alpha paste at offsets that clip the image by every border (and miss the frame altogether)
        sh, sw = int(rng.integers(1, 200)), int(rng.integers(1, 300))
        rgba = torch.from_numpy(rng.integers(0, 256, (sh, sw, 4), dtype=np.uint8)).to(cuda)
        for (x, y) in ((-sw + 1, -sh + 1), (w - 1, h - 1), (-sw, 0), (0, h), (int(rng.integers(-sw, w)), int(rng.integers(-sh, h)))):
            rc = lib.aiko_b200_alpha_paste_u8(frames.ptr, n, h, w, h * w * 3, rgba.data_ptr(), sh, sw, x, y, ctypes.c_void_p(stream))
            torch.cuda.synchronize()
            assert rc == 0 and frames.intact(), f"trial {trial}: alpha_paste at ({x}, {y}) wrote outside the frames"
