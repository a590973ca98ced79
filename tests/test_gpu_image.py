"""GPU parity: the CUDA path through the C ABI against the CPU oracle (bit-exact), the golden
vectors of the reference, and size-independent properties at BASELINE's full sizes."""
import numpy as np
import pytest
import torch

from aiko_services_b200 import ops
from aiko_services_b200.overlay_plan import build_ops
from aiko_services_b200.text_atlas import TextAtlas
from golden import cases
from oracle import image as O
from util import check_golden, overlay_cpu, random_overlay

pytestmark = pytest.mark.gpu


def dev(arr, cuda):
    return torch.from_numpy(np.ascontiguousarray(arr)).to(cuda)


def host(t):
    return t.cpu().numpy()


@pytest.fixture(scope="module")
def atlas():
    return TextAtlas()


# ---------------------------------------------------------------- resize (a-1)
@pytest.mark.parametrize("case", cases.RESIZE_NDARRAY, ids=lambda c: c[0])
def test_resize_linear_golden(golden, cuda, case):
    name, seed, sh, sw, c, dw, dh = case
    img = cases.frame(seed, sh, sw, c)
    out = host(ops.resize_linear(dev(img[None], cuda), dw, dh))[0]
    check_golden(golden, "resize/" + name, out)


def test_resize_linear_random_shapes_vs_oracle(cuda):
    rng = np.random.default_rng(101)
    for i in range(60):
        h, w, H, W = (int(v) for v in rng.integers(1, 260, 4))
        c = int(rng.choice([1, 3, 4]))
        if i % 3 == 0:
            w = 16 * int(rng.integers(1, 20))      # 16-byte aligned rows take the TMA-staged kernel
        n = int(rng.integers(1, 4))
        frames = np.stack([cases.frame(1000 * i + k, h, w, c) for k in range(n)])
        out = host(ops.resize_linear(dev(frames, cuda), W, H))
        for k in range(n):
            np.testing.assert_array_equal(out[k], O.resize_linear(frames[k], W, H), err_msg=f"{(h, w, c)}->{(H, W)}")


def test_resize_linear_unaligned_base_pointer(cuda):
    frames = np.stack([cases.frame(77, 48, 64, 3)])
    buf = torch.empty(frames.size + 1, dtype=torch.uint8, device=cuda)
    view = buf[1:].view(frames.shape)
    view.copy_(dev(frames, cuda))
    assert view.data_ptr() % 16 != 0
    # a view is contiguous in the sense the ABI needs (tight strides)
    out = host(ops.resize_linear(view, 20, 30))
    np.testing.assert_array_equal(out[0], O.resize_linear(frames[0], 20, 30))


@pytest.mark.parametrize("shape", [(3, 270, 479, 160, 120), (2, 1080, 1918, 640, 480), (2, 97, 333, 84, 33),
                                   (1, 64, 1366, 1024, 31), (2, 50, 131, 400, 20),
                                   (1, 40, 3000, 1500, 20), (2, 32, 3840, 1920, 16), (1, 24, 5000, 2800, 12)])   # 2 / 4 groups per thread
def test_resize_linear_rows_off_16_byte_boundaries_take_the_row_pipeline(cuda, shape):
    """RGB rows that are not 16-byte aligned (odd widths, odd base pointers) are staged from the
    boundary below each row; the result is the oracle's, with and without the channel swap."""
    n, h, w, W, H = shape
    frames = np.stack([cases.frame(8100 + k, h, w, 3) for k in range(n)])
    want = np.stack([O.resize_linear(f, W, H) for f in frames])
    np.testing.assert_array_equal(host(ops.resize_linear(dev(frames, cuda), W, H)), want)
    buf = torch.empty(frames.size + 5, dtype=torch.uint8, device=cuda)
    view = buf[5:].view(frames.shape)
    view.copy_(dev(frames, cuda))
    np.testing.assert_array_equal(host(ops.resize_linear(view, W, H, swap_rb=True)), want[..., ::-1])


@pytest.mark.parametrize("shape", [(3, 270, 480, 160, 120), (2, 1080, 1920, 640, 480), (2, 97, 333, 84, 33),
                                   (1, 40, 50, 200, 90), (2, 64, 1366, 1024, 31)])
def test_resize_linear_gray_frames_take_the_row_pipeline(cuda, shape):
    """Gray frames (aligned or not) whose output width is a multiple of 4 run on the TMA ring."""
    n, h, w, W, H = shape
    frames = np.stack([cases.frame(8300 + k, h, w, 1).reshape(h, w) for k in range(n)])
    want = np.stack([O.resize_linear(f, W, H) for f in frames])
    np.testing.assert_array_equal(host(ops.resize_linear(dev(frames, cuda), W, H)), want)
    buf = torch.empty(frames.size + 7, dtype=torch.uint8, device=cuda)
    view = buf[7:].view(frames.shape)
    view.copy_(dev(frames, cuda))
    out = torch.empty((n, H, W), dtype=torch.uint8, device=cuda)
    np.testing.assert_array_equal(host(ops.resize_linear(view, W, H, out=out)), want)


def test_resize_linear_edge_values(cuda):
    for v in (0, 255):
        img = np.full((2, 90, 160, 3), v, np.uint8)
        assert (host(ops.resize_linear(dev(img, cuda), 37, 41)) == v).all()
    grad = np.tile(np.arange(256, dtype=np.uint8)[None, None, :, None], (1, 64, 1, 3))
    np.testing.assert_array_equal(host(ops.resize_linear(dev(grad, cuda), 100, 30))[0], O.resize_linear(grad[0], 100, 30))


def test_resize_linear_empty_batch_and_bad_args(cuda):
    empty = torch.empty((0, 8, 8, 3), dtype=torch.uint8, device=cuda)
    assert ops.resize_linear(empty, 4, 4).shape == (0, 4, 4, 3)
    with pytest.raises(RuntimeError):
        ops.resize_linear(torch.zeros((1, 8, 8, 3), dtype=torch.uint8), 4, 4)       # host tensor: no CPU path
    with pytest.raises(Exception):
        ops.resize_linear(torch.zeros((1, 8, 8, 2), dtype=torch.uint8, device=cuda), 4, 4)


def test_resize_cfg2_batch_properties(cuda):
    """BASELINE config 2 size (32 x 1080p -> 640x480): every frame equals the oracle; identical
    inputs give identical outputs wherever they sit in the batch; swap_rb commutes."""
    frames = np.stack([cases.frame(2000 + (i % 4), 1080, 1920, 3) for i in range(32)])
    d = dev(frames, cuda)
    out = host(ops.resize_linear(d, 640, 480))
    refs = [O.resize_linear(frames[i], 640, 480) for i in range(4)]
    for i in range(32):
        np.testing.assert_array_equal(out[i], refs[i % 4])
    swapped = host(ops.resize_linear(d, 640, 480, swap_rb=True))
    np.testing.assert_array_equal(swapped, out[..., ::-1])


# ---------------------------------------------------------------- PIL bicubic resize (a-2)
@pytest.mark.parametrize("case", cases.RESIZE_PIL, ids=lambda c: c[0])
def test_resize_bicubic_golden(golden, cuda, case):
    name, seed, sh, sw, dw, dh = case
    img = cases.frame(seed, sh, sw, 3)
    out = host(ops.resize_bicubic_aa(dev(img[None], cuda), dw, dh))[0]
    check_golden(golden, "resize_pil/" + name, out)


def test_resize_bicubic_random_shapes_vs_oracle(cuda):
    rng = np.random.default_rng(111)
    for i in range(40):
        h, w, H, W = (int(v) for v in rng.integers(1, 200, 4))
        c = int(rng.choice([1, 3, 4]))
        n = int(rng.integers(1, 4))
        frames = np.stack([cases.frame(5000 + 10 * i + k, h, w, c) for k in range(n)])
        out = host(ops.resize_bicubic_aa(dev(frames, cuda), W, H))
        for k in range(n):
            np.testing.assert_array_equal(out[k], O.resize_bicubic_aa(frames[k], W, H), err_msg=f"{(h, w, c)}->{(H, W)}")


def test_resize_bicubic_cfg1_batch(cuda):
    """BASELINE config 1 size: every frame of a 1080p batch equals the oracle; edge values saturate."""
    frames = np.stack([cases.frame(1000 + (i % 2), 1080, 1920, 3) for i in range(8)])
    out = host(ops.resize_bicubic_aa(dev(frames, cuda), 640, 480))
    refs = [O.resize_bicubic_aa(frames[i], 640, 480) for i in range(2)]
    for i in range(8):
        np.testing.assert_array_equal(out[i], refs[i % 2])
    for v in (0, 255):
        img = np.full((1, 200, 300, 3), v, np.uint8)
        assert (host(ops.resize_bicubic_aa(dev(img, cuda), 77, 51)) == v).all()


TENSOR_CORE_SHAPES = [   # (h, w, H, W): RGB frames take the tensor-core kernel (scale up to ~5)
    (270, 480, 120, 160),    # the config-1 ratio (3 x 2.25): two K steps along x, one along y
    (120, 160, 270, 480),    # upscale
    (200, 320, 200, 100),    # height unchanged (identity vertical table)
    (64, 64, 64, 64),        # both unchanged
    (301, 208, 77, 51),      # odd targets: partial last tile and last row group
    (480, 640, 100, 131),    # scale ~4.9 x 4.8: four K steps along x, two along y
    (90, 1600, 45, 1000),    # wide: more tiles than one CTA owns -> several column chunks
    (33, 48, 5, 7), (16, 16, 1, 1), (8, 16, 40, 300),
    (270, 479, 120, 160), (97, 333, 41, 50), (60, 1366, 45, 700)]   # rows off 16-byte boundaries


@pytest.mark.parametrize("shape", TENSOR_CORE_SHAPES, ids=lambda s: "%dx%d_to_%dx%d" % (s[1], s[0], s[3], s[2]))
def test_resize_bicubic_tensor_core_path_vs_oracle(cuda, shape, monkeypatch):
    h, w, H, W = shape
    frames = np.stack([cases.frame(7000 + k, h, w, 3) for k in range(3)])
    want = np.stack([O.resize_bicubic_aa(f, W, H) for f in frames])
    d = dev(frames, cuda)
    np.testing.assert_array_equal(host(ops.resize_bicubic_aa(d, W, H)), want)
    # band and chunk edges: one row group per CTA, three column chunks
    monkeypatch.setenv("AIKO_B200_BICUBIC_BAND_GROUPS", "1")
    monkeypatch.setenv("AIKO_B200_BICUBIC_CHUNKS", "3")
    np.testing.assert_array_equal(host(ops.resize_bicubic_aa(d, W, H)), want)
    monkeypatch.setenv("AIKO_B200_BICUBIC_BAND_GROUPS", "3")
    monkeypatch.setenv("AIKO_B200_BICUBIC_CHUNKS", "2")
    np.testing.assert_array_equal(host(ops.resize_bicubic_aa(d, W, H)), want)
    buf = torch.empty(frames.size + 3, dtype=torch.uint8, device=cuda)     # odd base pointer
    view = buf[3:].view(frames.shape)
    view.copy_(d)
    np.testing.assert_array_equal(host(ops.resize_bicubic_aa(view, W, H)), want)


@pytest.mark.parametrize("channels", [1, 4])
@pytest.mark.parametrize("shape", [(270, 480, 120, 160), (120, 161, 270, 480), (301, 207, 77, 51), (480, 640, 100, 131)])
def test_resize_bicubic_tensor_core_gray_and_rgba(cuda, shape, channels):
    """PIL 'L' and 'RGBA' frames take the same tensor-core kernel (one / four channel planes)."""
    h, w, H, W = shape
    frames = np.stack([cases.frame(7400 + k, h, w, channels) for k in range(2)])
    want = np.stack([O.resize_bicubic_aa(f, W, H) for f in frames])
    np.testing.assert_array_equal(host(ops.resize_bicubic_aa(dev(frames, cuda), W, H)), want)


def test_resize_bicubic_tensor_core_equals_scalar_kernel(cuda, monkeypatch):
    """Full-size batch: the two kernels behind the entry point agree byte for byte."""
    g = torch.Generator(device=cuda).manual_seed(3)
    src = torch.randint(0, 256, (8, 1080, 1920, 3), dtype=torch.uint8, device=cuda, generator=g)
    src[0, :, :960] = 255
    src[1, :540] = 0
    fast = ops.resize_bicubic_aa(src, 640, 480)
    monkeypatch.setenv("AIKO_B200_BICUBIC_SCALAR", "1")
    slow = ops.resize_bicubic_aa(src, 640, 480)
    torch.cuda.synchronize()
    assert torch.equal(fast, slow)
    monkeypatch.delenv("AIKO_B200_BICUBIC_SCALAR")
    up = ops.resize_bicubic_aa(src[:2, :480, :640].contiguous(), 1920, 1080)
    monkeypatch.setenv("AIKO_B200_BICUBIC_SCALAR", "1")
    assert torch.equal(up, ops.resize_bicubic_aa(src[:2, :480, :640].contiguous(), 1920, 1080))


def test_resize_bicubic_workspace_errors(cuda):
    from aiko_services_b200 import _native
    lib = _native.lib()
    assert lib.aiko_b200_resize_bicubic_workspace_bytes(0, 10, 10, 10) == 0
    src = torch.zeros((1, 8, 8, 3), dtype=torch.uint8, device=cuda)
    dst = torch.zeros((1, 4, 4, 3), dtype=torch.uint8, device=cuda)
    rc = lib.aiko_b200_resize_bicubic_aa_u8(src.data_ptr(), 1, 8, 8, 3, dst.data_ptr(), 4, 4, None, 0, None)
    assert rc == _native.ERR_WORKSPACE


# ---------------------------------------------------------------- config 4: resize + normalise -> fp32 NCHW
IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)


def test_resize_norm_f32_vs_oracle(cuda):
    rng = np.random.default_rng(121)
    for i in range(8):
        h, w, H, W = (int(v) for v in rng.integers(8, 300, 4))
        frames = np.stack([cases.frame(6000 + 10 * i + k, h, w, 3) for k in range(2)])
        out = host(ops.resize_norm_f32(dev(frames, cuda), W, H, IMAGENET_MEAN, IMAGENET_STD))
        for k in range(2):
            np.testing.assert_array_equal(out[k], O.resize_norm_f32(frames[k], W, H, IMAGENET_MEAN, IMAGENET_STD))


def test_resize_norm_f32_cfg4_size(cuda):
    """4K -> 224x224: bit-identical to cv2-exact resize followed by the numpy float32 transform."""
    frames = np.stack([cases.frame(4000 + k, 2160, 3840, 3) for k in range(2)])
    out = host(ops.resize_norm_f32(dev(frames, cuda), 224, 224, IMAGENET_MEAN, IMAGENET_STD))
    assert out.shape == (2, 3, 224, 224) and out.dtype == np.float32
    for k in range(2):
        u8 = O.resize_linear(frames[k], 224, 224)
        ref = ((u8.astype(np.float32) / np.float32(255) - np.asarray(IMAGENET_MEAN, np.float32)) /
               np.asarray(IMAGENET_STD, np.float32)).transpose(2, 0, 1)
        np.testing.assert_array_equal(out[k], ref)      # tolerance 0: every step is a correctly rounded fp32 op


# ---------------------------------------------------------------- convert / flip / crop (a-4)
@pytest.mark.parametrize("n_pixels", [0, 1, 15, 16, 17, 4097, 640 * 480, 1920 * 1080 + 5])
def test_convert_ops_vs_oracle(cuda, n_pixels):
    rng = np.random.default_rng(n_pixels)
    rgb = rng.integers(0, 256, (1, 1, n_pixels, 3), dtype=np.uint8)
    gray = rng.integers(0, 256, (1, 1, n_pixels), dtype=np.uint8)
    d = dev(rgb, cuda)
    if n_pixels == 0:
        assert ops.swap_rb(d).numel() == 0 and ops.bgr2gray(d).numel() == 0
        return
    np.testing.assert_array_equal(host(ops.swap_rb(d)), rgb[..., ::-1])
    np.testing.assert_array_equal(host(ops.bgr2gray(d))[0], O.bgr2gray(rgb[0]))
    np.testing.assert_array_equal(host(ops.bgr2gray(d, rgb_order=True))[0], O.bgr2gray(rgb[0], True))
    np.testing.assert_array_equal(host(ops.gray2bgr(dev(gray, cuda)))[0], O.gray2bgr(gray[0]))


def test_swap_is_involution_full_size(cuda):
    frames = dev(np.stack([cases.frame(9, 1080, 1920, 3)] * 2), cuda)
    assert torch.equal(ops.swap_rb(ops.swap_rb(frames)), frames)


@pytest.mark.parametrize("shape", [(2, 37, 53, 3), (1, 48, 64, 3), (3, 7, 32, 3), (3, 20, 33, 1), (1, 9, 16, 4), (1, 1080, 1920, 3)])
@pytest.mark.parametrize("code", [0, 1, -1])
def test_flip_vs_oracle(cuda, shape, code):
    n, h, w, c = shape
    frames = np.stack([cases.frame(50 + k, h, w, c) for k in range(n)])
    out = host(ops.flip(dev(frames, cuda), code))
    for k in range(n):
        np.testing.assert_array_equal(out[k], O.flip(frames[k], code))


def test_crop_vs_numpy(cuda):
    frames = np.stack([cases.frame(60 + k, 70, 90, 3) for k in range(3)])
    d = dev(frames, cuda)
    for (x0, y0, w, h) in [(0, 0, 90, 70), (16, 8, 32, 40), (5, 3, 17, 11), (89, 69, 1, 1), (10, 10, 0, 5)]:
        np.testing.assert_array_equal(host(ops.crop(d, x0, y0, w, h)), frames[:, y0:y0 + h, x0:x0 + w])
    with pytest.raises(Exception):
        ops.crop(d, 80, 0, 20, 10)


# ---------------------------------------------------------------- overlay (a-3)
def test_overlay_in_place(cuda, atlas):
    """src == dst: the frames are drawn where they are (no copy), RGB and gray."""
    rng = np.random.default_rng(77)
    for c in (3, 1):
        frames = np.stack([cases.frame(300 + k, 120, 200, c) for k in range(3)])
        overlay = random_overlay(rng, 200, 120)
        ov = device_overlay([overlay] * 3, 200, 120, "rgb" if c == 3 else "bgr", atlas, cuda)
        d = dev(frames, cuda)
        assert ops.overlay_draw(d, ov, out=d).data_ptr() == d.data_ptr()
        for k in range(3):
            np.testing.assert_array_equal(host(d)[k], overlay_cpu(frames[k], overlay, atlas))


def device_overlay(overlays, w, h, layout, atlas, cuda):
    per_frame = [build_ops(o, w, h, layout, atlas) for o in overlays]
    return ops.DeviceOverlay.from_host(per_frame, cuda, atlas.device(cuda), width=w, height=h)


@pytest.mark.parametrize("case", cases.OVERLAY, ids=lambda c: c[0])
def test_overlay_golden(golden, cuda, atlas, case):
    name, seed, h, w, c, overlay = case
    img = cases.frame(seed, h, w, c)
    ov = device_overlay([overlay], w, h, "rgb" if c == 3 else "bgr", atlas, cuda)
    out = host(ops.overlay_draw(dev(img[None], cuda), ov))[0]
    check_golden(golden, "overlay/" + name, out)


def test_overlay_random_vs_oracle_replay(cuda, atlas):
    rng = np.random.default_rng(202)
    for i in range(12):
        h, w = int(rng.integers(40, 200)), int(rng.integers(60, 320))
        c = 1 if i % 4 == 0 else 3
        n = 3
        frames = np.stack([cases.frame(300 + 10 * i + k, h, w, c) for k in range(n)])
        overlays = [random_overlay(rng, w, h, n=int(rng.integers(0, 9))) for _ in range(n)]
        ov = device_overlay(overlays, w, h, "rgb" if c == 3 else "bgr", atlas, cuda)
        out = host(ops.overlay_draw(dev(frames, cuda), ov))
        for k in range(n):
            np.testing.assert_array_equal(out[k], overlay_cpu(frames[k], overlays[k], atlas))


def test_overlay_many_ops_per_frame(cuda, atlas):
    rng = np.random.default_rng(203)
    h, w = 240, 320
    overlay = random_overlay(rng, w, h, n=70)     # > 256 ops: several cull passes per row
    img = cases.frame(204, h, w, 3)
    ov = device_overlay([overlay], w, h, "rgb", atlas, cuda)
    assert ov.n_ops > 256
    np.testing.assert_array_equal(host(ops.overlay_draw(dev(img[None], cuda), ov))[0], overlay_cpu(img, overlay, atlas))


# ---------------------------------------------------------------- fused chain (config 2)
def test_fused_chain_golden(golden, cuda, atlas):
    frames = np.stack([cases.frame(2000 + i, 1080, 1920, 3) for i in range(2)])
    ov = device_overlay([cases.OVERLAY_4] * 2, 640, 480, "rgb", atlas, cuda)
    out = host(ops.resize_linear(dev(frames, cuda), 640, 480, overlay=ov))
    for i in range(2):
        check_golden(golden, f"chain/cfg2_{i}", out[i])
    # ... and with the encoder-side RGB->BGR folded in (video_io.py:299)
    ov_bgr = device_overlay([cases.OVERLAY_4] * 2, 640, 480, "bgr", atlas, cuda)
    out_bgr = host(ops.resize_linear(dev(frames, cuda), 640, 480, swap_rb=True, overlay=ov_bgr))
    np.testing.assert_array_equal(out_bgr, out[..., ::-1])


def test_fused_equals_unfused_random(cuda, atlas):
    rng = np.random.default_rng(205)
    for i in range(6):
        h, w = int(rng.integers(60, 300)), 16 * int(rng.integers(4, 30))
        H, W = int(rng.integers(40, 200)), int(rng.integers(60, 320))
        frames = np.stack([cases.frame(400 + 10 * i + k, h, w, 3) for k in range(3)])
        overlays = [random_overlay(rng, W, H) for _ in range(3)]
        ov = device_overlay(overlays, W, H, "rgb", atlas, cuda)
        d = dev(frames, cuda)
        fused = ops.resize_linear(d, W, H, overlay=ov)
        unfused = ops.overlay_draw(ops.resize_linear(d, W, H), ov)
        assert torch.equal(fused, unfused)
        for k in range(3):
            np.testing.assert_array_equal(host(fused)[k], overlay_cpu(O.resize_linear(frames[k], W, H), overlays[k], atlas))


def test_fused_resize_with_long_draw_list_uses_device_side_culling(cuda, atlas):
    """More ops than the cell kernel takes per list (256): the tile kernel culls by itself."""
    rng = np.random.default_rng(206)
    H, W = 120, 160
    frames = np.stack([cases.frame(500 + k, 270, 480, 3) for k in range(2)])
    overlays = [random_overlay(rng, W, H, n=60) for _ in range(2)]
    ov = device_overlay(overlays, W, H, "rgb", atlas, cuda)
    assert ov.struct.max_ops_per_list > 256
    out = host(ops.resize_linear(dev(frames, cuda), W, H, overlay=ov))
    for k in range(2):
        np.testing.assert_array_equal(out[k], overlay_cpu(O.resize_linear(frames[k], W, H), overlays[k], atlas))


def test_overlay_variants_agree(cuda, atlas, monkeypatch):
    """Cell kernel (default), tile kernel (AIKO_B200_OVERLAY_TILES=1) and the single-kernel variant
    (AIKO_B200_OVERLAY_FUSED=1) are bit-identical."""
    frames = dev(np.stack([cases.frame(2000 + i, 1080, 1920, 3) for i in range(3)]), cuda)
    two_pass = ops.resize_linear(frames, 640, 480, swap_rb=True,
                                 overlay=device_overlay([cases.OVERLAY_4] * 3, 640, 480, "bgr", atlas, cuda))
    monkeypatch.setenv("AIKO_B200_OVERLAY_FUSED", "1")
    fused = ops.resize_linear(frames, 640, 480, swap_rb=True,
                              overlay=device_overlay([cases.OVERLAY_4] * 3, 640, 480, "bgr", atlas, cuda))
    assert torch.equal(two_pass, fused)
    monkeypatch.delenv("AIKO_B200_OVERLAY_FUSED")
    monkeypatch.setenv("AIKO_B200_OVERLAY_TILES", "1")
    tiles = ops.resize_linear(frames, 640, 480, swap_rb=True,
                              overlay=device_overlay([cases.OVERLAY_4] * 3, 640, 480, "bgr", atlas, cuda))
    assert torch.equal(two_pass, tiles)


def test_planned_overlay_per_frame_lists_and_shared_list(cuda, atlas):
    """overlay_plan.plan -> DeviceOverlay.from_plan: one dict shared by all frames, and distinct dicts
    per frame (cross-stream batch), both against the CPU replay; many marks (8 mask words)."""
    from aiko_services_b200.overlay_plan import plan
    rng = np.random.default_rng(77)
    W, H = 320, 200
    frames = np.stack([cases.frame(900 + k, H, W, 3) for k in range(5)])
    for n_obj in (1, 4, 12, 40):
        overlays = [random_overlay(rng, W, H, n=n_obj) for _ in range(5)]
        overlays[3] = overlays[1]                              # two frames share a dict
        ov = ops.DeviceOverlay.from_plan(plan(overlays, W, H, "rgb"), cuda)
        out = host(ops.overlay_draw(dev(frames, cuda), ov))
        for k in range(5):
            np.testing.assert_array_equal(out[k], overlay_cpu(frames[k], overlays[k], atlas), err_msg=f"{n_obj} {k}")
        ov = ops.DeviceOverlay.from_plan(plan(overlays[0], W, H, "rgb", n_frames=5), cuda)
        out = host(ops.overlay_draw(dev(frames, cuda), ov))
        for k in range(5):
            np.testing.assert_array_equal(out[k], overlay_cpu(frames[k], overlays[0], atlas))


def test_overlay_refuses_a_list_made_for_fewer_frames(cuda, atlas):
    from aiko_services_b200._native import NativeError
    ov = device_overlay([cases.OVERLAY_4] * 2, 640, 480, "rgb", atlas, cuda)
    assert ov.n_ops > 0
    with pytest.raises(NativeError):
        ops.overlay_draw(dev(np.zeros((3, 480, 640, 3), np.uint8), cuda), ov)


def test_resize_rgba_swap_and_widths_not_multiple_of_four(cuda, atlas):
    rgba = np.stack([cases.frame(70, 45, 64, 4)])
    out = host(ops.resize_linear(dev(rgba, cuda), 31, 17, swap_rb=True))[0]
    ref = O.resize_linear(rgba[0], 31, 17)
    np.testing.assert_array_equal(out, ref[..., [2, 1, 0, 3]])
    rng = np.random.default_rng(207)
    for W in (61, 62, 63, 130):
        frames = np.stack([cases.frame(71 + W, 96, 128, 3)])
        overlay = random_overlay(rng, W, 50)
        ov = device_overlay([overlay], W, 50, "rgb", atlas, cuda)
        out = host(ops.resize_linear(dev(frames, cuda), W, 50, overlay=ov))[0]
        np.testing.assert_array_equal(out, overlay_cpu(O.resize_linear(frames[0], W, 50), overlay, atlas))


def test_sizes_beyond_on_chip_limits_are_refused_not_mangled(cuda):
    """Rows that cannot be staged in shared memory give AIKO_B200_ERR_UNSUPPORTED with a message."""
    from aiko_services_b200 import _native
    tiny = torch.zeros((1, 2, 4, 3), dtype=torch.uint8, device=cuda)
    with pytest.raises(_native.NativeError) as e:
        ops.resize_linear(tiny, 40000, 2)                 # 40000-pixel output rows: tables exceed shared memory
    assert e.value.code == _native.ERR_UNSUPPORTED and "shared memory" in str(e.value)
    # a 10x reduction is outside the tensor-core kernel's tables: the scalar kernel, which splits rows that do not
    # fit in shared memory into column chunks; so does a wide row under a large vertical reduction
    for (h, w, H, W) in ((2, 60000, 2, 6000), (229, 696, 6, 430), (300, 2000, 7, 1999)):
        big = np.stack([cases.frame(83, h, w, 3)])
        np.testing.assert_array_equal(host(ops.resize_bicubic_aa(dev(big, cuda), W, H))[0], O.resize_bicubic_aa(big[0], W, H))
    wide = np.stack([cases.frame(82, 2, 4, 3)])                        # RGB: column chunks of the tensor-core kernel
    np.testing.assert_array_equal(host(ops.resize_bicubic_aa(dev(wide, cuda), 60000, 2))[0],
                                  O.resize_bicubic_aa(wide[0], 60000, 2))
    with pytest.raises(_native.NativeError) as e:
        ops.flip(torch.zeros((1, 1, 100001, 3), dtype=torch.uint8, device=cuda), 1)
    assert e.value.code == _native.ERR_INVALID_ARGUMENT
    wide = torch.arange(100000 * 3, device=cuda).to(torch.uint8).view(1, 1, 100000, 3)   # multiple of 16 pixels:
    assert torch.equal(ops.flip(wide, 1), wide.flip(2))                                   # register path, no row limit
    # the widest rows the staged kernel takes still resize exactly (8K frames)
    frames = np.stack([cases.frame(81, 24, 7680, 3)])
    out = host(ops.resize_linear(dev(frames, cuda), 3840, 12))
    np.testing.assert_array_equal(out[0], O.resize_linear(frames[0], 3840, 12))


def test_overlay_ring_changing_lists(cuda, atlas):
    """A new draw list per call through OverlayRing (planned into pinned memory, one async copy):
    results match the CPU replay while the ring wraps several times."""
    rng = np.random.default_rng(88)
    W, H = 160, 120
    frames = np.stack([cases.frame(700 + k, 270, 480, 3) for k in range(2)])
    d = dev(frames, cuda)
    ring = ops.OverlayRing(cuda, capacity=1 << 14, slots=3)
    for i in range(10):
        overlay = random_overlay(rng, W, H, n=3)
        ov = ring.plan(overlay, W, H, "rgb", n_frames=2)
        out = host(ops.resize_linear(d, W, H, overlay=ov))
        for k in range(2):
            np.testing.assert_array_equal(out[k], overlay_cpu(O.resize_linear(frames[k], W, H), overlay, atlas))


def test_overlay_ring_resize_draw_one_call(cuda, atlas):
    """OverlayRing.resize_draw = plan + resize + draw in ONE call of the library
    (aiko_b200_resize_linear_objects_u8): a new overlay per call, shared or one per frame, with and
    without the fused RGB->BGR; labels near the border make the planner grow the atlas in mid-run (the
    ATLAS_GREW path).  Bit-exact with the oracle; the ring wraps several times."""
    from aiko_services_b200.overlay_plan import build_ops
    from util import replay_ops_cpu
    rng = np.random.default_rng(89)
    W, H = 160, 120
    frames = np.stack([cases.frame(720 + k, 270, 480, 3) for k in range(2)])
    d = dev(frames, cuda)
    ring = ops.OverlayRing(cuda, capacity=1 << 14, slots=3)
    for i in range(12):
        swap = bool(i & 1)
        per_frame = i % 3 == 2
        overlays = [random_overlay(rng, W, H, n=3) for _ in range(2)] if per_frame else random_overlay(rng, W, H, n=3)
        out = host(ring.resize_draw(d, overlays, W, H, swap_rb=swap))
        for k in range(2):
            small = O.resize_linear(frames[k], W, H)
            o = overlays[k] if per_frame else overlays
            if swap:
                want = replay_ops_cpu(O.swap_rb(small), build_ops(o, W, H, "bgr", atlas), atlas)
            else:
                want = overlay_cpu(small, o, atlas)
            np.testing.assert_array_equal(out[k], want)
    # an empty overlay draws nothing; a list too long for a ring slot takes one-off buffers
    np.testing.assert_array_equal(host(ring.resize_draw(d, {}, W, H))[0], O.resize_linear(frames[0], W, H))
    big = random_overlay(rng, W, H, n=120)
    small_ring = ops.OverlayRing(cuda, capacity=1 << 10, slots=2)
    out = host(small_ring.resize_draw(d, big, W, H))
    np.testing.assert_array_equal(out[1], overlay_cpu(O.resize_linear(frames[1], W, H), big, atlas))


def _pil_array(im, mode):
    a = np.asarray(im)
    return a.astype(np.uint8) * 255 if mode == "1" else a


@pytest.mark.parametrize("case", cases.RESIZE_PIL_MODES, ids=lambda c: c[0])
def test_pil_mode_resize_golden(cuda, case):
    """ops.resize_pil against the reference's ImageResize on RGBA / LA / L / CMYK / P / 1 images
    (tests/golden/golden_pil_modes.npz): premultiplied alpha around the filter, nearest for P and 1."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_pil_modes.npz"))
    name, seed, mode, sh, sw, dw, dh = case
    src = _pil_array(cases.pil_image(seed, mode, sh, sw), mode)
    got = host(ops.resize_pil(dev(src[None], cuda), mode, dw, dh))[0]
    np.testing.assert_array_equal(got, g[name])


def test_pil_mode_resize_random_vs_oracle(cuda):
    rng = np.random.default_rng(19)
    for trial in range(24):
        mode = ("RGBA", "LA", "P", "1")[trial % 4]
        sh, sw = int(rng.integers(2, 200)), int(rng.integers(2, 300))
        dh, dw = int(rng.integers(1, 160)), int(rng.integers(1, 200))
        frames = np.stack([_pil_array(cases.pil_image(int(rng.integers(1 << 30)), mode, sh, sw), mode) for _ in range(3)])
        got = host(ops.resize_pil(dev(frames, cuda), mode, dw, dh))
        for k in range(3):
            np.testing.assert_array_equal(got[k], O.pil_resize(frames[k], mode, dw, dh), err_msg=f"{mode} {sw}x{sh}->{dw}x{dh}")


def test_yuv_conversions_vs_oracle(cuda):
    """RGB / BGR <-> YUV (cv2.cvtColor arithmetic, oracle pinned on all 2^24 colours): vector path,
    tails that are not a multiple of 16 pixels, unaligned buffers, and the ImageConvert recipe step."""
    rng = np.random.default_rng(31)
    for shape in ((3, 64, 96, 3), (1, 7, 13, 3), (2, 33, 50, 3)):
        frames = rng.integers(0, 256, shape, dtype=np.uint8)
        d = dev(frames, cuda)
        for bgr in (False, True):
            for inverse in (False, True):
                got = host(ops.yuv(d, bgr_order=bgr, inverse=inverse))
                np.testing.assert_array_equal(got, O.yuv(frames, bgr, inverse))
    flat = torch.from_numpy(rng.integers(0, 256, 3 * 1000 + 1, dtype=np.uint8)).to(cuda)
    src = flat[1:].view(1, 10, 100, 3)                      # deliberately off the 16-byte grid
    np.testing.assert_array_equal(host(ops.yuv(src.contiguous())), O.yuv(host(src)))
    out = torch.empty(3 * 1000 + 1, dtype=torch.uint8, device=cuda)[1:].view(1, 10, 100, 3)
    torch.ops.aiko_b200.yuv_u8(src, out, False, False, torch.cuda.current_stream().cuda_stream)
    np.testing.assert_array_equal(host(out), O.yuv(host(src)))


def test_alpha_paste_vs_oracle(cuda):
    """ops.alpha_paste (PIL's paste of an RGBA image by its own alpha; extension named by north_star) against the
    oracle pinned on Pillow: random sizes and offsets, clipped by every border, frame-sized images, batches; in
    place and into ``out``; also straight through torch.ops.aiko_b200."""
    rng = np.random.default_rng(93)
    for t in range(40):
        n, h, w = int(rng.integers(1, 4)), int(rng.integers(1, 90)), int(rng.integers(1, 130))
        sh, sw = int(rng.integers(1, 100)), int(rng.integers(1, 140))
        if t % 8 == 0:
            sh, sw = h, w
        frames = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
        rgba = rng.integers(0, 256, (sh, sw, 4), dtype=np.uint8)
        x, y = (0, 0) if t % 8 == 0 else (int(rng.integers(-sw - 2, w + 2)), int(rng.integers(-sh - 2, h + 2)))
        d, s = dev(frames, cuda), dev(rgba, cuda)
        out = ops.alpha_paste(d, s, x, y, out=torch.empty_like(d))
        np.testing.assert_array_equal(host(d), frames)                      # the source is untouched
        for k in range(n):
            np.testing.assert_array_equal(host(out)[k], O.pil_paste(frames[k], rgba, x, y), err_msg=f"case {t} frame {k}")
        torch.ops.aiko_b200.alpha_paste_u8(d, s, x, y, torch.cuda.current_stream().cuda_stream)       # in place
        np.testing.assert_array_equal(host(d), host(out))
    # rows that are 16-byte aligned in both images take the 16-pixel kernel
    for (h, w, sh, sw, x, y) in ((48, 64, 48, 64, 0, 0), (50, 96, 20, 32, 16, 5), (480, 640, 480, 640, 0, 0)):
        frames = rng.integers(0, 256, (3, h, w, 3), dtype=np.uint8)
        rgba = rng.integers(0, 256, (sh, sw, 4), dtype=np.uint8)
        got = host(ops.alpha_paste(dev(frames, cuda), dev(rgba, cuda), x, y))
        for k in range(3):
            np.testing.assert_array_equal(got[k], O.pil_paste(frames[k], rgba, x, y))


def _random_detections(rng, lists, slots, w, h, names, interior):
    boxes = np.zeros((lists, slots, 4), np.int32)
    cls = rng.integers(-1, len(names), (lists, slots)).astype(np.int32)
    conf = rng.random((lists, slots)).astype(np.float32)
    special = np.array([0.125, 0.875, 0.005, 0.995, 0.999, 1.0, 0.0, -0.3, 12.345, 0.015, 0.025, 0.5], np.float32)
    pick = rng.random((lists, slots)) < 0.3
    conf[pick] = rng.choice(special, size=int(pick.sum()))
    for li in range(lists):
        for k in range(slots):
            if interior:          # label and box well inside the frame: the device planner has every glyph
                bw, bh = int(rng.integers(10, w // 3)), int(rng.integers(10, h // 3))
                boxes[li, k] = (int(rng.integers(20, w - bw - 200)), int(rng.integers(60, h - bh - 60)), bw, bh)
            else:
                boxes[li, k] = (int(rng.integers(-30, w)), int(rng.integers(-30, h)), int(rng.integers(0, w)), int(rng.integers(0, h)))
    return boxes, cls, conf


def test_overlay_from_device_resident_detections(cuda, atlas):
    """ops.overlay_boxes: the draw lists are planned ON THE GPU from a detector's output tensors (boxes, class
    ids, float32 confidences) -- label text incl. the correctly rounded confidence, size, placement, glyph ops --
    and must equal the host planner's result (itself pinned on cv2) pixel for pixel: shared and per-frame lists,
    skipped slots, unnamed classes, ties in the rounding.  Labels cut by the frame border take the host path."""
    from aiko_services_b200.overlay_plan import plan
    names = ["person", "bicycle", "car", "", "traffic light", "Zq{|}~", "x"]
    rng = np.random.default_rng(2031)
    W, H = 640, 480
    for trial in range(12):
        n = int(rng.integers(1, 4))
        lists = 1 if trial % 2 == 0 else n
        slots = int(rng.integers(1, 9))
        interior = trial % 3 != 2
        boxes, cls, conf = _random_detections(rng, lists, slots, W, H, names, interior)
        frames = np.stack([cases.frame(9300 + 7 * trial + k, H, W, 3) for k in range(n)])
        det = ops.DeviceDetections(torch.from_numpy(boxes).to(cuda), torch.from_numpy(cls).to(cuda),
                                   torch.from_numpy(conf).to(cuda), names)
        dicts = det.to_overlay_dicts()
        for layout in ("rgb", "bgr"):
            d = dev(frames, cuda)
            got, status = ops.overlay_boxes(d, det, layout=layout, out=torch.empty_like(d), check=False)
            code = int(status.cpu()[0])
            planned = plan(dicts[0] if lists == 1 else dicts, W, H, layout, n_frames=n)
            want = host(ops.overlay_draw(d, ops.DeviceOverlay.from_plan(planned, cuda, atlas.device(cuda))))
            if interior:
                assert code == 0, f"trial {trial}: status {code}"
            if code == 0:
                np.testing.assert_array_equal(host(got), want, err_msg=f"trial {trial} {layout}")
            else:
                np.testing.assert_array_equal(host(got), frames)                       # nothing drawn, status says why
            checked = ops.overlay_boxes(d, det, layout=layout, out=torch.empty_like(d))   # falls back when needed
            np.testing.assert_array_equal(host(checked), want, err_msg=f"trial {trial} {layout} (checked)")
            # the host path rasterised the border variants this call needs; they are mirrored on the device now, so
            # the same detections (the next frame of a video) are planned on the GPU
            again, status = ops.overlay_boxes(d, det, layout=layout, out=torch.empty_like(d), check=False)
            assert int(status.cpu()[0]) == 0, f"trial {trial} {layout}: still status {int(status.cpu()[0])}"
            np.testing.assert_array_equal(host(again), want, err_msg=f"trial {trial} {layout} (second call)")
