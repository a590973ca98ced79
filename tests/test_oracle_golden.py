"""CPU: the C oracle against the golden vectors produced by the reference's own elements,
and directly against cv2 / Pillow where they are installed."""
import numpy as np
import pytest

from golden import cases
from oracle import image as O
from util import check_golden


@pytest.mark.parametrize("case", cases.RESIZE_NDARRAY, ids=lambda c: c[0])
def test_resize_linear_golden(golden, case):
    name, seed, sh, sw, c, dw, dh = case
    check_golden(golden, "resize/" + name, O.resize_linear(cases.frame(seed, sh, sw, c), dw, dh))


@pytest.mark.parametrize("case", cases.RESIZE_PIL, ids=lambda c: c[0])
def test_resize_bicubic_golden(golden, case):
    name, seed, sh, sw, dw, dh = case
    check_golden(golden, "resize_pil/" + name, O.resize_bicubic_aa(cases.frame(seed, sh, sw, 3), dw, dh))


def test_resize_linear_vs_cv2_random_shapes():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(7)
    for _ in range(120):
        h, w, H, W = (int(v) for v in rng.integers(1, 200, 4))
        c = int(rng.choice([1, 3, 4]))
        img = cases.frame(int(rng.integers(1 << 30)), h, w, c)
        np.testing.assert_array_equal(O.resize_linear(img, W, H), cv2.resize(img, (W, H)))


def test_resize_bicubic_vs_pillow_random_shapes():
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(8)
    for _ in range(40):
        h, w, H, W = (int(v) for v in rng.integers(1, 150, 4))
        img = cases.frame(int(rng.integers(1 << 30)), h, w, 3)
        ref = np.asarray(Image.fromarray(img, "RGB").resize((W, H)))
        np.testing.assert_array_equal(O.resize_bicubic_aa(img, W, H), ref)


def test_saturation_edges():
    cv2 = pytest.importorskip("cv2")
    for v in (0, 255):
        img = np.full((40, 50, 3), v, np.uint8)
        assert (O.resize_linear(img, 23, 17) == v).all()
        assert (O.resize_bicubic_aa(img, 23, 17) == v).all()
    grad = np.tile(np.arange(256, dtype=np.uint8)[None, :, None], (64, 1, 3))
    np.testing.assert_array_equal(O.resize_linear(grad, 100, 30), cv2.resize(grad, (100, 30)))


def test_convert_flip_vs_cv2():
    cv2 = pytest.importorskip("cv2")
    img = cases.frame(5, 67, 93, 3)
    gray = cases.frame(6, 67, 93, 1)
    np.testing.assert_array_equal(O.swap_rb(img), cv2.cvtColor(img, cv2.COLOR_RGB2BGR))
    np.testing.assert_array_equal(O.swap_rb(img), img[..., ::-1])
    np.testing.assert_array_equal(O.bgr2gray(img), cv2.cvtColor(img, cv2.COLOR_BGR2GRAY))
    np.testing.assert_array_equal(O.bgr2gray(img, True), cv2.cvtColor(img, cv2.COLOR_RGB2GRAY))
    np.testing.assert_array_equal(O.gray2bgr(gray), cv2.cvtColor(gray, cv2.COLOR_GRAY2BGR))
    for code in (0, 1, -1):
        np.testing.assert_array_equal(O.flip(img, code), cv2.flip(img, code))
        np.testing.assert_array_equal(O.flip(gray, code), cv2.flip(gray, code))


def test_rectangle_vs_cv2_random():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(9)
    for _ in range(1500):
        h, w = (int(v) for v in rng.integers(4, 80, 2))
        a = cases.frame(int(rng.integers(1 << 30)), h, w, 3)
        b = a.copy()
        x0, x1 = (int(v) for v in rng.integers(-20, w + 20, 2))
        y0, y1 = (int(v) for v in rng.integers(-20, h + 20, 2))
        cv2.rectangle(a, (x0, y0), (x1, y1), (0, 255, 255), 2)
        O.rectangle(b, x0, y0, x1, y1, (0, 255, 255))
        np.testing.assert_array_equal(a, b)


def test_resize_norm_f32_matches_numpy():
    img = cases.frame(3, 120, 160, 3)
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    u8 = O.resize_linear(img, 56, 56)
    ref = ((u8.astype(np.float32) / np.float32(255) - np.asarray(mean, np.float32)) /
           np.asarray(std, np.float32)).transpose(2, 0, 1)
    np.testing.assert_array_equal(O.resize_norm_f32(img, 56, 56, mean, std), ref)


# ---- audio chain: the restatement against vectors produced by the reference's own statements --------
def _audio_golden():
    import os
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_audio.npz"))


def test_audio_oracle_matches_the_reference_statements():
    """tests/golden/golden_audio.npz holds what R/elements/media/audio_io.py:216-227,253-277,307-313
    compute (exec'd from the reference file) on the config-5 windows; numpy's FFT may differ in the
    last bits between builds, hence 1e-12 instead of equality on the magnitudes."""
    from golden.make_golden_audio import windows_cfg5
    from oracle import audio as A
    g = _audio_golden()
    fs = int(g["sample_rate"])
    w = windows_cfg5(24)
    for k in range(24):
        amp, freq = A.pe_fft(w[k], fs)
        np.testing.assert_array_equal(freq, g["frequencies"])
        np.testing.assert_allclose(amp[::97], g[f"w{k}/amplitude_probe"], rtol=1e-12, atol=1e-15)
        if f"w{k}/amplitudes" in g:
            np.testing.assert_allclose(amp, g[f"w{k}/amplitudes"], rtol=1e-12, atol=1e-15)
        bf, ba = A.pe_audio_resampler(amp, freq)
        np.testing.assert_array_equal(bf, g["band_frequencies"])
        np.testing.assert_allclose(ba, g[f"w{k}/band_amplitudes"], rtol=1e-12)
        tf, ta = A.pe_audio_filter(amp, freq)
        np.testing.assert_array_equal(tf, g[f"w{k}/top_frequencies"])
        np.testing.assert_allclose(ta, g[f"w{k}/top_amplitudes"], rtol=1e-12)
    rng = np.random.default_rng(77)
    x = (rng.standard_normal(9600) * 4000).astype(np.float32)
    x += 30000 * np.sin(2 * np.pi * 440 * np.arange(9600) / 48000).astype(np.float32)
    amp, freq = A.pe_fft(x, 48000)
    np.testing.assert_allclose(amp, g["f32_9600/amplitudes"], rtol=1e-12, atol=1e-15)
    tf, ta = A.pe_audio_filter(amp, freq)
    np.testing.assert_array_equal(tf, g["f32_9600/top_frequencies"])


def _pil_array(im, mode):
    a = np.asarray(im)
    return a.astype(np.uint8) * 255 if mode == "1" else a


@pytest.mark.parametrize("case", cases.RESIZE_PIL_MODES, ids=lambda c: c[0])
def test_pil_mode_resize_oracle_vs_reference_golden(case):
    """oracle.pil_resize (premultiplied RGBA / LA, nearest P / 1, plain CMYK / L) against the reference's
    ImageResize run on PIL images of those modes (tests/golden/make_golden_pil_modes.py)."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_pil_modes.npz"))
    name, seed, mode, sh, sw, dw, dh = case
    src = _pil_array(cases.pil_image(seed, mode, sh, sw), mode)
    np.testing.assert_array_equal(O.pil_resize(src, mode, dw, dh), g[name])


def test_pil_mode_resize_oracle_vs_pillow_random():
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(9)
    for trial in range(40):
        mode = ("RGBA", "LA", "P", "1")[trial % 4]
        sh, sw = int(rng.integers(1, 90)), int(rng.integers(1, 120))
        dh, dw = int(rng.integers(1, 120)), int(rng.integers(1, 150))
        im = cases.pil_image(int(rng.integers(1 << 30)), mode, sh, sw)
        want = _pil_array(im.resize((dw, dh)), mode)
        np.testing.assert_array_equal(O.pil_resize(_pil_array(im, mode), mode, dw, dh), want, err_msg=f"{mode} {sw}x{sh}->{dw}x{dh}")


def test_yuv_oracle_equals_cv2_on_every_colour():
    """Extension (north_star names YUV; the reference never converts to it): the oracle of record is
    cv2.cvtColor on 8-bit images -- all 2^24 colours, both channel orders, both directions."""
    cv2 = pytest.importorskip("cv2")
    codes = {(False, False): cv2.COLOR_RGB2YUV, (True, False): cv2.COLOR_BGR2YUV,
             (False, True): cv2.COLOR_YUV2RGB, (True, True): cv2.COLOR_YUV2BGR}
    for c0 in range(0, 256, 64):
        cube = np.stack(np.meshgrid(np.arange(c0, c0 + 64), np.arange(256), np.arange(256), indexing="ij"), -1)
        cube = cube.astype(np.uint8).reshape(-1, 1, 3)
        for (bgr, inverse), code in codes.items():
            np.testing.assert_array_equal(O.yuv(cube, bgr, inverse), cv2.cvtColor(cube, code))


def _natural(rng, h, w):
    from PIL import Image
    low = (rng.random((h // 20 + 2, w // 20 + 2, 3)) * 255).astype(np.uint8)
    up = np.asarray(Image.fromarray(low, "RGB").resize((w, h), Image.BICUBIC)).astype(np.float32)
    return np.clip(up + rng.normal(0, 6, (h, w, 3)), 0, 255).astype(np.uint8)


def test_jpeg_oracle_is_bit_exact_with_pillow():
    """oracle/jpeg_oracle.c (Huffman, islow IDCT, fancy upsampling, YCbCr tables as libjpeg-turbo's default
    decoder) against Pillow's decoder -- the decode that runs inside the reference's image.resize
    (R/elements/media/image_io.py:184,218) -- on random images: every sampling layout, gray, odd and tiny
    sizes, any quality, with and without restart markers."""
    import io
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(5)
    for t in range(80):
        h, w = int(rng.integers(1, 200)), int(rng.integers(1, 300))
        if t % 10 == 9:
            w = int(rng.integers(1, 5))              # up to 4 pixels wide: chroma is replicated, not filtered
        img = _natural(rng, max(h, 2), max(w, 2))[:h, :w] if t % 3 else rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        kw = {}
        if t % 4 == 0:
            kw["restart_marker_rows"] = 1
        if t % 4 == 1:
            kw["restart_marker_blocks"] = int(rng.integers(1, 9))
        im = Image.fromarray(img, "RGB")
        if t % 7 == 0:
            im = im.convert("L")
        buf = io.BytesIO()
        im.save(buf, "JPEG", quality=int(rng.integers(30, 100)), subsampling=t % 3 if im.mode == "RGB" else -1, **kw)
        ref = np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("RGB"))
        np.testing.assert_array_equal(O.jpeg_decode(buf.getvalue()), ref, err_msg=f"case {t}: {w}x{h}")
    with pytest.raises(ValueError):
        O.jpeg_decode(b"\xff\xd8\xff\xe0 nonsense")
    buf = io.BytesIO()
    Image.fromarray(_natural(rng, 40, 40), "RGB").save(buf, "JPEG", progressive=True)
    with pytest.raises(ValueError):
        O.jpeg_decode(buf.getvalue())                 # progressive: outside the subset, refused


def test_jpeg_encode_oracle_writes_pillows_bytes():
    """oracle/jpeg_encode_oracle.c (restated libjpeg-turbo encoder: colour conversion, downsampling with edge
    replication and dummy blocks, islow forward DCT, quantisation, Huffman coding, headers) against the bytes
    Pillow writes -- what ImageWriteFile's image.save produces (R/elements/media/image_io.py:250): every
    subsampling, qualities 1..100, gray, sizes that are not multiples of the MCU."""
    import io
    from PIL import Image
    rng = np.random.default_rng(21)
    sizes = [(16, 16), (8, 8), (24, 40), (1, 1), (17, 33), (7, 100), (100, 7), (67, 101), (270, 480), (15, 15), (9, 24),
             (8, 24), (24, 8), (40, 24)]
    for h, w in sizes:
        for kind in ("noise", "natural"):
            if kind == "noise":
                a = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
            else:
                a = _natural(rng, max(h, 2), max(w, 2))[:h, :w]
            for quality in (75, 90, 30, 100, 1):
                for sub, factors in ((2, (2, 2)), (1, (2, 1)), (0, (1, 1))):
                    buf = io.BytesIO()
                    Image.fromarray(a, "RGB").save(buf, "JPEG", quality=quality, subsampling=sub)
                    assert O.jpeg_encode(a, quality, factors) == buf.getvalue(), (h, w, kind, quality, sub)
            gray = np.ascontiguousarray(a[..., 1])
            buf = io.BytesIO()
            Image.fromarray(gray, "L").save(buf, "JPEG")
            assert O.jpeg_encode(gray) == buf.getvalue(), (h, w, kind, "gray")
    # the default call of the reference: image.save(path) = quality 75, 4:2:0
    a = _natural(rng, 480, 640)
    buf = io.BytesIO()
    Image.fromarray(a, "RGB").save(buf, "JPEG")
    assert O.jpeg_encode(a) == buf.getvalue()


def test_alpha_paste_oracle_is_pillows_paste():
    """oracle.pil_paste (extension: alpha blending is named by north_star, the reference has no call site) against
    PIL.Image.paste(image, (x, y), image) for RGBA images at random offsets, clipped by every border."""
    from PIL import Image
    rng = np.random.default_rng(31)
    for t in range(200):
        h, w = int(rng.integers(1, 90)), int(rng.integers(1, 120))
        sh, sw = int(rng.integers(1, 100)), int(rng.integers(1, 130))
        frame = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        rgba = rng.integers(0, 256, (sh, sw, 4), dtype=np.uint8)
        if t % 5 == 0:
            rgba[..., 3] = rng.choice([0, 255, 128], size=(sh, sw))
        x, y = int(rng.integers(-sw - 2, w + 2)), int(rng.integers(-sh - 2, h + 2))
        ref = Image.fromarray(frame, "RGB")
        sticker = Image.fromarray(rgba, "RGBA")
        ref.paste(sticker, (x, y), sticker)
        np.testing.assert_array_equal(O.pil_paste(frame, rgba, x, y), np.asarray(ref), err_msg=f"case {t}")
