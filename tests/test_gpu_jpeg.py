"""GPU: JPEG frames decoded on the device against Pillow's decoder.
This library's own decoder (csrc/jpeg_decode.cu) restates libjpeg-turbo's arithmetic and is compared BIT-EXACT.
The nvJPEG path (libaiko_b200_jpeg.so; progressive files) is another decoder -- two JPEG decoders never agree
bit for bit (IDCT rounding; chroma upsampling for 4:2:0) -- so there the decoded pixels are compared within a
stated tolerance, and everything AFTER the decode bit-exact on the same decoded pixels."""
import io

import numpy as np
import pytest
import torch

from golden import cases
from oracle import image as O

pytestmark = pytest.mark.gpu


def natural_frame(seed, h, w):
    """Smooth random field + mild noise: compresses like camera content."""
    rng = np.random.default_rng(seed)
    low = rng.random((h // 40 + 2, w // 40 + 2, 3))
    from PIL import Image
    up = np.asarray(Image.fromarray((low * 255).astype(np.uint8), "RGB").resize((w, h), Image.BICUBIC)).astype(np.float32)
    return np.clip(up + rng.normal(0, 4, (h, w, 3)), 0, 255).astype(np.uint8)


def encode(frame, quality=90, subsampling=0):
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(frame, "RGB").save(buf, "JPEG", quality=quality, subsampling=subsampling)
    return buf.getvalue()


def test_library_loads_and_exports_its_symbols():
    from aiko_services_b200 import jpeg
    assert all(hasattr(jpeg.lib(), name) for name in jpeg.SIGNATURES)


@pytest.mark.parametrize("subsampling,mean_tol,max_tol", [(0, 0.6, 6), (2, 1.0, 16)], ids=["444", "420"])
def test_decode_vs_pillow(cuda, subsampling, mean_tol, max_tol):
    from PIL import Image
    from aiko_services_b200 import jpeg
    dec = jpeg.decoder_for(cuda, 8)
    frames = [natural_frame(50 + k, 270, 480) for k in range(5)]
    blobs = [encode(f, 90, subsampling) for f in frames]
    assert dec.info(blobs[0])[:3] == (480, 270, 3)
    got = dec.decode(blobs).cpu().numpy()
    for k in range(5):
        want = np.asarray(Image.open(io.BytesIO(blobs[k])).convert("RGB"))
        diff = np.abs(got[k].astype(np.int32) - want.astype(np.int32))
        assert diff.mean() <= mean_tol and diff.max() <= max_tol, (k, diff.mean(), diff.max())


def test_resize_of_the_decoded_pixels_is_bit_exact(cuda):
    """decode on the GPU -> antialiased bicubic resize (config 1) == oracle resize of the SAME pixels."""
    from aiko_services_b200 import jpeg, ops
    dec = jpeg.decoder_for(cuda, 8)
    blobs = [encode(natural_frame(80 + k, 540, 960), 85, 2) for k in range(3)]
    decoded = dec.decode(blobs)
    resized = ops.resize_bicubic_aa(decoded, 320, 240).cpu().numpy()
    pixels = decoded.cpu().numpy()
    for k in range(3):
        np.testing.assert_array_equal(resized[k], O.resize_bicubic_aa(pixels[k], 320, 240))


def test_bad_input_is_an_error_not_a_crash(cuda):
    from aiko_services_b200 import jpeg
    from aiko_services_b200._native import NativeError
    dec = jpeg.decoder_for(cuda, 8)
    with pytest.raises(NativeError):
        dec.decode([b"\xff\xd8\xff not a jpeg at all"])
    good = encode(natural_frame(1, 64, 96))
    other = encode(natural_frame(2, 48, 96))
    with pytest.raises(NativeError):
        dec.decode([good, other])            # sizes differ within one batch
    assert dec.decode([good]).shape == (1, 64, 96, 3)


def test_image_read_file_decodes_on_the_gpu(cuda, tmp_path):
    """ImageReadFile(decode=gpu) -> ImageResize -> ImageWriteFile: frames never exist as raw pixels on
    the host before the resize; the written files hold exactly the pixels of Pillow's decode + resize."""
    from PIL import Image
    from aiko_services_b200.pipeline import Pipeline
    mod = "aiko_services.elements.media.image_io"
    frames = [natural_frame(300 + k, 270, 480) for k in range(4)]
    for k, f in enumerate(frames[:3]):
        (tmp_path / f"in_{k:02d}.jpeg").write_bytes(encode(f, 92, 2))
    d = {"version": 0, "name": "p_jpeg", "runtime": "python", "graph": ["(ImageReadFile ImageResize ImageWriteFile)"],
         "parameters": {"resolution": "160x90"},
         "elements": [
             {"name": "ImageReadFile", "parameters": {"data_sources": f"(file://{tmp_path}/in_{{}}.jpeg)",
                                                      "data_batch_size": 2, "decode": "gpu"},
              "input": [{"name": "paths", "type": "[Path]"}], "output": [{"name": "images", "type": "[image]"}],
              "deploy": {"local": {"module": mod}}},
             {"name": "ImageResize", "input": [{"name": "images", "type": "[image]"}],
              "output": [{"name": "images", "type": "[image]"}], "deploy": {"local": {"module": mod}}},
             {"name": "ImageWriteFile", "parameters": {"data_targets": f"file://{tmp_path}/out_{{:02d}}.png"},
              "input": [{"name": "images", "type": "[image]"}], "output": [], "deploy": {"local": {"module": mod}}}]}
    p = Pipeline(d)
    assert p.create_stream("1")
    p.run()
    for k in range(3):
        got = np.asarray(Image.open(tmp_path / f"out_{k:02d}.png")).astype(np.int32)
        want = np.asarray(Image.open(tmp_path / f"in_{k:02d}.jpeg").resize((160, 90))).astype(np.int32)
        assert got.shape == (90, 160, 3)
        np.testing.assert_array_equal(got, want)


def test_jpeg_in_jpeg_out_files_are_the_references_bytes(cuda, tmp_path):
    """ImageReadFile(decode=gpu) -> ImageResize -> ImageWriteFile(encode=gpu) with JPEG files on both sides: only
    compressed bytes cross PCIe, and every written FILE is byte-identical to what the reference's chain writes
    (Image.open(path) -> image.resize((w, h)) -> image.save(path), R/elements/media/image_io.py:184,218,250)."""
    import io as _io
    from PIL import Image
    from aiko_services_b200.pipeline import Pipeline
    mod = "aiko_services.elements.media.image_io"
    for k in range(5):
        (tmp_path / f"in_{k:02d}.jpg").write_bytes(encode(natural_frame(400 + k, 360, 640), 88, 2))
    (tmp_path / "in_05.jpg").write_bytes(encode_rst(natural_frame(405, 360, 640), 80, mode="L", plain=True))   # a gray file
    d = {"version": 0, "name": "p_jpeg2", "runtime": "python", "graph": ["(ImageReadFile ImageResize ImageWriteFile)"],
         "parameters": {"resolution": "320x180"},
         "elements": [
             {"name": "ImageReadFile", "parameters": {"data_sources": f"(file://{tmp_path}/in_{{}}.jpg)",
                                                      "data_batch_size": 3, "decode": "gpu"},
              "input": [{"name": "paths", "type": "[Path]"}], "output": [{"name": "images", "type": "[image]"}],
              "deploy": {"local": {"module": mod}}},
             {"name": "ImageResize", "input": [{"name": "images", "type": "[image]"}],
              "output": [{"name": "images", "type": "[image]"}], "deploy": {"local": {"module": mod}}},
             {"name": "ImageWriteFile", "parameters": {"data_targets": f"file://{tmp_path}/out_{{:02d}}.jpg", "encode": "gpu"},
              "input": [{"name": "images", "type": "[image]"}], "output": [], "deploy": {"local": {"module": mod}}}]}
    p = Pipeline(d)
    assert p.create_stream("1")
    p.run()
    for k in range(6):
        buf = _io.BytesIO()
        Image.open(tmp_path / f"in_{k:02d}.jpg").resize((320, 180)).save(buf, "JPEG")
        assert (tmp_path / f"out_{k:02d}.jpg").read_bytes() == buf.getvalue(), k
    assert Image.open(tmp_path / "out_05.jpg").mode == "L"


# ---- this library's own decoder (csrc/jpeg_decode.cu): bit-exact with Pillow ------------------------------
def encode_rst(frame, quality=90, subsampling=2, mode="RGB", plain=False, **kw):
    """Baseline JPEG; restart markers per MCU row unless ``kw`` says otherwise or ``plain`` (none at all)."""
    from PIL import Image
    buf = io.BytesIO()
    im = Image.fromarray(frame, "RGB")
    if mode == "L":
        im = im.convert("L")
    extra = {} if plain else (kw or {"restart_marker_rows": 1})
    im.save(buf, "JPEG", quality=quality, subsampling=subsampling if mode == "RGB" else -1, **extra)
    return buf.getvalue()


def pillow_rgb(blob):
    from PIL import Image
    return np.asarray(Image.open(io.BytesIO(blob)).convert("RGB"))


@pytest.mark.parametrize("plain", [False, True], ids=["restart_rows", "no_markers"])
@pytest.mark.parametrize("subsampling", [0, 1, 2], ids=["444", "422", "420"])
def test_own_decoder_is_bit_exact_with_pillow(cuda, subsampling, plain):
    from aiko_services_b200 import jpeg
    dec = jpeg.GpuJpegDecoder(cuda)
    rng = np.random.default_rng(40 + subsampling)
    # (images up to 4 pixels wide: libjpeg replicates their chroma instead of filtering it)
    for h, w in ((270, 480), (64, 96), (67, 101), (8, 8), (1, 1), (17, 250), (600, 800), (68, 1), (175, 4), (40, 3), (9, 2)):
        frames = [natural_frame(int(rng.integers(1 << 30)), max(h, 2), max(w, 2))[:h, :w] for _ in range(3)]
        frames[1] = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)               # noise: long codes, big coefficients
        blobs = [encode_rst(f, int(rng.integers(40, 100)), subsampling, plain=plain) for f in frames]
        got = dec.decode(blobs, check=True).cpu().numpy()
        for k in range(3):
            np.testing.assert_array_equal(got[k], pillow_rgb(blobs[k]), err_msg=f"{w}x{h} image {k}")
            np.testing.assert_array_equal(got[k], O.jpeg_decode(blobs[k]))


def test_own_decoder_gray_restart_blocks_and_1080p(cuda):
    from aiko_services_b200 import jpeg
    dec = jpeg.GpuJpegDecoder(cuda)
    frame = natural_frame(7, 120, 200)
    for kw in ({"restart_marker_blocks": 1}, {"restart_marker_blocks": 7}, {"restart_marker_rows": 3}):
        blob = encode_rst(frame, 85, 2, **kw)
        np.testing.assert_array_equal(dec.decode([blob], check=True).cpu().numpy()[0], pillow_rgb(blob))
    gray = encode_rst(frame, 80, 0, mode="L")
    np.testing.assert_array_equal(dec.decode([gray, gray], check=True).cpu().numpy()[1], pillow_rgb(gray))
    gray = encode_rst(frame, 80, 0, mode="L", plain=True)
    np.testing.assert_array_equal(dec.decode([gray, gray], check=True).cpu().numpy()[0], pillow_rgb(gray))
    # BASELINE config 1 size; one batch mixes every kind of scan: a marker per MCU row and no markers at all (one
    # thread per 512-byte subsequence, states found by iteration), a marker every 8 MCUs (1020 segments: one thread
    # per segment), and white noise at quality 100 (3 MB of scan, long codes)
    big = [encode_rst(natural_frame(90, 1080, 1920), 90, 2), encode_rst(natural_frame(91, 1080, 1920), 90, 2, plain=True),
           encode_rst(natural_frame(92, 1080, 1920), 75, 2, restart_marker_blocks=8),
           encode_rst(np.random.default_rng(5).integers(0, 256, (1080, 1920, 3), dtype=np.uint8), 100, 2, plain=True)]
    got = dec.decode(big, check=True).cpu().numpy()
    for k in range(len(big)):
        np.testing.assert_array_equal(got[k], pillow_rgb(big[k]), err_msg=f"image {k}")


def test_own_decoder_refuses_what_it_cannot_do(cuda):
    from PIL import Image
    from aiko_services_b200 import jpeg
    from aiko_services_b200._native import ERR_UNSUPPORTED, NativeError
    dec = jpeg.GpuJpegDecoder(cuda)
    frame = natural_frame(3, 64, 96)
    buf = io.BytesIO()
    Image.fromarray(frame, "RGB").save(buf, "JPEG", progressive=True)
    with pytest.raises(NativeError) as e:
        dec.decode([buf.getvalue()])
    assert e.value.code == ERR_UNSUPPORTED
    good = encode_rst(frame)
    with pytest.raises(NativeError):
        dec.decode([good, encode_rst(natural_frame(4, 48, 96))])      # sizes differ within a batch
    with pytest.raises(RuntimeError):
        dec.decode([good], out=torch.empty((1, 32, 96, 3), dtype=torch.uint8, device=cuda))     # a destination of the wrong shape
    cut = good[:len(good) // 2]                                        # truncated scan: reported, no crash
    with pytest.raises(NativeError):
        dec.decode([cut], check=True)
    for cut in (good[:len(good) // 2], encode(frame)[:len(encode(frame)) // 2]):
        with pytest.raises(NativeError):
            dec.decode([cut], check=True)
    extras = io.BytesIO()                                              # comment / EXIF / ICC segments: skipped
    Image.fromarray(frame, "RGB").save(extras, "JPEG", comment=b"hello", icc_profile=b"x" * 70000,
                                       exif=b"Exif\x00\x00II*\x00\x08\x00\x00\x00\x00\x00")
    np.testing.assert_array_equal(dec.decode([extras.getvalue()], check=True).cpu().numpy()[0], pillow_rgb(extras.getvalue()))
    pinned = torch.frombuffer(bytearray(extras.getvalue()), dtype=torch.uint8).pin_memory()      # header beyond the first 64 KB
    np.testing.assert_array_equal(dec.decode([pinned], check=True).cpu().numpy()[0], pillow_rgb(extras.getvalue()))
    frames, which = jpeg.decode_on_gpu([buf.getvalue()], cuda)         # progressive: falls back to nvJPEG
    assert which == "nvjpeg" and frames.shape == (1, 64, 96, 3)
    for blob in (good, encode(frame)):                                 # with and without restart markers
        frames, which = jpeg.decode_on_gpu([blob], cuda)
        assert which == "aiko"
        np.testing.assert_array_equal(frames.cpu().numpy()[0], pillow_rgb(blob))


def test_own_decoder_survives_corrupt_files(cuda):
    """Random damage to valid files (flipped bytes in tables and scan, truncation, inserted markers): every call
    returns pixels or a NativeError -- no hang, no crash -- and a valid file decodes bit-exactly afterwards."""
    from aiko_services_b200 import jpeg
    from aiko_services_b200._native import NativeError
    dec = jpeg.GpuJpegDecoder(cuda)
    rng = np.random.default_rng(77)
    frame = natural_frame(11, 120, 200)
    sources = [encode_rst(frame, 85, 2, plain=True), encode_rst(frame, 85, 2), encode_rst(frame, 70, 0, restart_marker_blocks=1),
               encode_rst(frame, 95, 1, plain=True)]
    outcomes = {"pixels": 0, "error": 0}
    for trial in range(160):
        blob = bytearray(sources[trial % len(sources)])
        kind = trial % 4
        if kind == 0:                                   # damage anywhere after SOI (headers, tables, scan)
            for _ in range(int(rng.integers(1, 6))):
                blob[int(rng.integers(2, len(blob)))] = int(rng.integers(256))
        elif kind == 1:                                 # damage inside the scan only
            for _ in range(int(rng.integers(1, 20))):
                blob[int(rng.integers(len(blob) // 2, len(blob)))] = int(rng.integers(256))
        elif kind == 2:                                 # truncation
            blob = blob[:int(rng.integers(4, len(blob)))]
        else:                                           # stray markers in the scan
            at = int(rng.integers(len(blob) // 2, len(blob) - 2))
            blob[at:at + 2] = bytes([0xff, int(rng.choice([0xd0, 0xd3, 0xd9, 0xc4, 0xff, 0x01]))])
        try:
            out = dec.decode([bytes(blob)], check=True)
            assert out.shape[0] == 1
            outcomes["pixels"] += 1
        except NativeError:
            outcomes["error"] += 1
    assert outcomes["error"] > 20 and outcomes["pixels"] + outcomes["error"] == 160
    for blob in sources:
        np.testing.assert_array_equal(dec.decode([blob], check=True).cpu().numpy()[0], pillow_rgb(blob))


# ---- this library's encoder (csrc/jpeg_encode.cu): the bytes Pillow writes ---------------------------------
def pillow_bytes(frame, **kw):
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(frame, "RGB" if frame.ndim == 3 else "L").save(buf, "JPEG", **kw)
    return buf.getvalue()


@pytest.mark.parametrize("subsampling", [2, 1, 0], ids=["420", "422", "444"])
def test_encoder_writes_pillows_bytes(cuda, subsampling):
    """GpuJpegEncoder against PIL.Image.save (what ImageWriteFile calls, R/elements/media/image_io.py:250) and
    against the oracle: identical files for natural frames and noise, any quality, sizes that are not multiples
    of the MCU (replicated edges, dummy blocks), batches."""
    from aiko_services_b200 import jpeg
    enc = jpeg.GpuJpegEncoder(cuda)
    rng = np.random.default_rng(60 + subsampling)
    factors = {2: (2, 2), 1: (2, 1), 0: (1, 1)}[subsampling]
    for h, w in ((480, 640), (16, 16), (8, 8), (1, 1), (17, 33), (7, 100), (100, 7), (67, 101), (9, 24), (24, 8), (270, 480)):
        frames = np.stack([natural_frame(int(rng.integers(1 << 30)), max(h, 2), max(w, 2))[:h, :w] for _ in range(3)])
        frames[1] = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        for quality in (75, int(rng.integers(1, 101))):
            files = enc.encode(torch.from_numpy(frames).to(cuda), quality=quality, subsampling=subsampling)
            for k in range(3):
                want = pillow_bytes(frames[k], quality=quality, subsampling=subsampling)
                assert files[k] == want, f"{w}x{h} image {k} quality {quality}: {len(files[k])} vs {len(want)} bytes"
                assert files[k] == O.jpeg_encode(frames[k], quality, factors)


def test_encoder_defaults_gray_and_round_trip(cuda):
    from aiko_services_b200 import jpeg
    enc, dec = jpeg.GpuJpegEncoder(cuda), jpeg.GpuJpegDecoder(cuda)
    frames = np.stack([natural_frame(70 + k, 1080, 1920) for k in range(2)])        # BASELINE size, Pillow's defaults
    files = enc.encode(torch.from_numpy(frames).to(cuda))
    for k in range(2):
        assert files[k] == pillow_bytes(frames[k])
    back = dec.decode(files, check=True).cpu().numpy()                               # and back through the decoder
    for k in range(2):
        np.testing.assert_array_equal(back[k], pillow_rgb(files[k]))
    gray = np.ascontiguousarray(frames[:, :203, :301, 1])
    for k, f in enumerate(enc.encode(torch.from_numpy(gray).to(cuda), quality=90)):
        assert f == pillow_bytes(gray[k], quality=90)
    noise = np.random.default_rng(3).integers(0, 256, (1, 256, 256, 3), dtype=np.uint8)     # scan larger than the first buffers
    assert enc.encode(torch.from_numpy(noise).to(cuda), quality=100, subsampling=0)[0] == pillow_bytes(noise[0], quality=100, subsampling=0)
    assert enc.encode(torch.empty((0, 8, 8, 3), dtype=torch.uint8, device=cuda)) == []


def test_codec_on_large_and_extreme_shapes(cuda):
    """4K frames (BASELINE config 4's source size), a very wide strip and a very tall one: encoder bytes and decoder
    pixels against Pillow."""
    from aiko_services_b200 import jpeg
    enc, dec = jpeg.GpuJpegEncoder(cuda), jpeg.GpuJpegDecoder(cuda)
    for h, w in ((2160, 3840), (9, 4100), (4100, 9)):
        frame = natural_frame(h + w, max(h, 41), max(w, 41))[:h, :w]
        files = enc.encode(torch.from_numpy(np.ascontiguousarray(frame[None])).to(cuda), quality=85)
        assert files[0] == pillow_bytes(frame, quality=85), (h, w)
        np.testing.assert_array_equal(dec.decode(files, check=True).cpu().numpy()[0], pillow_rgb(files[0]))
