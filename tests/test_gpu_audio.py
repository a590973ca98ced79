"""GPU: the audio chain (PE_FFT -> PE_AudioResampler -> PE_AudioFilter) against the numpy float64
restatement of the reference's (dead) code, oracle/audio.py.  Tolerances (SURVEY.md 8c, fp32 FFT):
magnitudes abs(gpu - ref) <= 1e-4 * max|ref| per window, and per bin rtol 1e-4 with
atol 1e-4 * median|ref|; band sums rtol 1e-4; selected bins compared as values."""
import numpy as np
import pytest
import torch

from aiko_services_b200 import audio_ops
from oracle import audio as A

pytestmark = pytest.mark.gpu


def windows_cfg5(count, n=16000, fs=16000):
    t = np.arange(n) / fs
    freqs = np.geomspace(50, 7000, count)
    out = np.empty((count, n), np.int16)
    for k in range(count):
        rng = np.random.default_rng(5000 + k)
        out[k] = np.clip(3000 * np.sin(2 * np.pi * freqs[k] * t) + rng.normal(0, 200, n), -32768, 32767).astype(np.int16)
    return out


def check_window(res, k, window, fs):
    amp, freq = A.pe_fft(window, fs)
    got = res["amplitudes"][k].cpu().numpy().astype(np.float64)
    assert np.abs(got - amp).max() <= 1e-4 * np.abs(amp).max()
    np.testing.assert_allclose(got, amp, rtol=1e-4, atol=1e-4 * np.median(np.abs(amp)))
    np.testing.assert_array_equal(res["frequencies"], freq)
    bf, ba = A.pe_audio_resampler(amp, freq)
    np.testing.assert_allclose(res["band_frequencies"], bf, rtol=1e-12)
    np.testing.assert_allclose(res["band_amplitudes"][k].cpu().numpy(), ba, rtol=1e-4, atol=1e-4 * np.abs(ba).max())
    tf, ta = A.pe_audio_filter(amp, freq)
    count = int(res["top_count"][k])
    gf = res["top_frequencies"][k].cpu().numpy()[:count]
    ga = res["top_amplitudes"][k].cpu().numpy()[:count]
    # bins within 1e-4 of a gate edge or of the 100th value may legitimately differ in fp32
    lo_edge = np.sum((np.abs(amp - 0.1) < 2e-5) | (np.abs(amp - 12) < 2e-3))
    if lo_edge == 0 and (len(ta) < 100 or ta[99] - np.sort(amp[(amp <= 12) & (amp >= 0.1)])[::-1][100:101].sum() > 1e-4):
        assert count == len(ta)
        np.testing.assert_allclose(np.sort(ga)[::-1], ta, rtol=1e-4)
        np.testing.assert_allclose(np.sort(gf), np.sort(tf), rtol=1e-6, atol=1e-3)
    assert (np.diff(ga) <= 0).all()                      # sorted by amplitude, descending
    assert ((ga >= 0.1 - 1e-5) & (ga <= 12 + 1e-3)).all() and ((gf >= 10) & (gf <= 9000)).all()


def test_audio_chain_cfg5_windows(cuda):
    fs = 16000
    w = windows_cfg5(24)
    res = audio_ops.audio_chain(torch.from_numpy(w).to(cuda), fs)
    for k in range(24):
        check_window(res, k, w[k], fs)


def test_audio_chain_float_input_and_other_lengths(cuda):
    rng = np.random.default_rng(7)
    for n, fs in ((16000, 16000), (9600, 48000), (8192, 8000), (4000, 16000), (1000, 8000)):
        x = (rng.standard_normal((3, n)) * 4000).astype(np.float32)
        x[1] += 30000 * np.sin(2 * np.pi * 440 * np.arange(n) / fs).astype(np.float32)
        res = audio_ops.audio_chain(torch.from_numpy(x).to(cuda), fs)
        for k in range(3):
            amp, _ = A.pe_fft(x[k], fs)
            got = res["amplitudes"][k].cpu().numpy().astype(np.float64)
            assert np.abs(got - amp).max() <= 1e-4 * np.abs(amp).max(), (n, k)


def test_audio_filter_when_thousands_of_bins_pass_the_gate(cuda):
    """Noise-like windows: far more than 100 bins pass the gates, so the pruning path runs."""
    rng = np.random.default_rng(11)
    fs, n = 16000, 16000
    x = (rng.standard_normal((4, n)) * 9000).astype(np.float32)
    res = audio_ops.audio_chain(torch.from_numpy(x).to(cuda), fs)
    for k in range(4):
        amp, freq = A.pe_fft(x[k], fs)
        tf, ta = A.pe_audio_filter(amp, freq)
        assert ((amp >= 0.1) & (amp <= 12)).sum() > 2000 and len(ta) == 100
        assert int(res["top_count"][k]) == 100
        ga = res["top_amplitudes"][k].cpu().numpy()
        np.testing.assert_allclose(ga, ta, rtol=1e-4)             # same 100 values, same (descending) order
        gf = res["top_frequencies"][k].cpu().numpy()
        np.testing.assert_allclose(np.sort(gf), np.sort(tf), rtol=1e-6, atol=1e-3)


def test_audio_chain_properties_full_batch(cuda):
    """1024 windows (BASELINE config 5): linearity of the spectrum, Parseval, symmetry."""
    fs, n = 16000, 16000
    w = windows_cfg5(64)
    batch = np.tile(w, (16, 1))                                       # 1024 windows
    res = audio_ops.audio_chain(torch.from_numpy(batch).to(cuda), fs)
    amp = res["amplitudes"]
    assert torch.equal(amp[:64], amp[64:128])                         # same window, same result, anywhere in the batch
    a = amp[:64].double().cpu().numpy()
    np.testing.assert_allclose(a[:, 1:n // 2], a[:, :n // 2:-1], rtol=0, atol=0)      # |X[k]| == |X[n-k]| exactly
    energy_time = (w.astype(np.float64) ** 2).sum(1)
    energy_freq = ((a * 1e6) ** 2).sum(1) / n
    np.testing.assert_allclose(energy_freq, energy_time, rtol=1e-4)


def test_audio_chain_empty_and_errors(cuda):
    empty = torch.empty((0, 16000), dtype=torch.int16, device=cuda)
    assert audio_ops.audio_chain(empty, 16000)["amplitudes"].shape == (0, 16000)
    with pytest.raises(Exception):
        audio_ops.audio_chain(torch.zeros((1, 16001), dtype=torch.int16, device=cuda), 16000)      # odd length
    with pytest.raises(Exception):
        audio_ops.audio_chain(torch.zeros((1, 2 * 7 * 100), dtype=torch.int16, device=cuda), 16000)  # factor 7
    with pytest.raises(Exception):
        audio_ops.audio_chain(torch.zeros((1, 32000), dtype=torch.int16, device=cuda), 16000)      # too long


def test_audio_elements_in_pipeline(cuda, tmp_path):
    """The reference's graph (PE_FFT PE_AudioResampler PE_AudioFilter), 6 concurrent streams batched."""
    from aiko_services_b200.pipeline import Pipeline
    mod = "aiko_services.elements.media.audio_io"
    io3 = [{"name": "amplitudes", "type": "any"}, {"name": "frequencies", "type": "any"}, {"name": "spectrum", "type": "any"}]
    d = {"version": 0, "name": "p_audio", "runtime": "python", "graph": ["(PE_FFT PE_AudioFilter)"],
         "parameters": {"sample_rate": 16000},
         "elements": [{"name": "PE_FFT", "input": [{"name": "audio", "type": "any"}], "output": io3,
                       "deploy": {"local": {"module": mod}}},
                      {"name": "PE_AudioFilter", "input": io3, "output": io3[:2], "deploy": {"local": {"module": mod}}}]}
    p = Pipeline(d)
    w = windows_cfg5(6)
    work = []
    for k in range(6):
        assert p.create_stream(f"a{k}")
        work.append(({"stream_id": f"a{k}", "frame_id": 0}, {"audio": w[k]}))
    results = p.process_frames(work)
    for k, (okay, swag) in enumerate(results):
        assert okay
        amp, freq = A.pe_fft(w[k], 16000)
        tf, ta = A.pe_audio_filter(amp, freq)
        assert len(swag["amplitudes"]) == len(ta)
        np.testing.assert_allclose(np.sort(swag["amplitudes"])[::-1], ta, rtol=1e-4)


# ---- against the vectors computed by the reference's own statements (tests/golden/golden_audio.npz) ----
def _golden_audio():
    import os
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_audio.npz"))


def test_audio_chain_vs_reference_statement_golden(cuda):
    """The CUDA chain against what R/elements/media/audio_io.py:216-227,253-277,307-313 compute
    (exec'd from the reference file by tests/golden/make_golden_audio.py), fp32 tolerance 1e-4."""
    g = _golden_audio()
    fs = int(g["sample_rate"])
    w = windows_cfg5(24)
    res = audio_ops.audio_chain(torch.from_numpy(w).to(cuda), fs)
    np.testing.assert_array_equal(res["frequencies"], g["frequencies"])
    np.testing.assert_allclose(res["band_frequencies"], g["band_frequencies"], rtol=1e-12)
    amps = res["amplitudes"].cpu().numpy().astype(np.float64)
    for k in range(24):
        if f"w{k}/amplitudes" in g:
            ref = g[f"w{k}/amplitudes"]
            assert np.abs(amps[k] - ref).max() <= 1e-4 * np.abs(ref).max()
            np.testing.assert_allclose(amps[k], ref, rtol=1e-4, atol=1e-4 * np.median(np.abs(ref)))
        probe = g[f"w{k}/amplitude_probe"]
        assert np.abs(amps[k][::97] - probe).max() <= 1e-4 * np.abs(probe).max() + 1e-4 * np.abs(amps[k]).max()
        ba = g[f"w{k}/band_amplitudes"]
        np.testing.assert_allclose(res["band_amplitudes"][k].cpu().numpy(), ba, rtol=1e-4, atol=1e-4 * np.abs(ba).max())
        ta, tf = g[f"w{k}/top_amplitudes"], g[f"w{k}/top_frequencies"]
        count = int(res["top_count"][k])
        ga = res["top_amplitudes"][k].cpu().numpy()[:count]
        gf = res["top_frequencies"][k].cpu().numpy()[:count]
        # a bin within 1e-4 of a gate edge may fall on either side in fp32: compare the clear ones
        edge = np.sum((np.abs(amps[k] - 0.1) < 2e-5) | (np.abs(amps[k] - 12) < 2e-3))
        if edge == 0 and len(ta) < 100:
            assert count == len(ta)
            np.testing.assert_allclose(ga, ta, rtol=1e-4)
            np.testing.assert_allclose(np.sort(gf), np.sort(tf), rtol=1e-6, atol=1e-3)


def test_audio_filter_and_bands_on_plain_arrays(cuda):
    """float64 arrays from anywhere (not PE_FFT's fused results): same selection, order and ties as
    the reference's Python (stable sort), sums to 1e-12."""
    g = _golden_audio()
    fs = int(g["sample_rate"])
    w = windows_cfg5(24)
    amps, freqs = zip(*[A.pe_fft(w[k], fs) for k in range(24)])
    amps = np.stack(amps)
    tf, ta, tc = audio_ops.audio_filter(amps, freqs[0], cuda)
    bf, ba = audio_ops.audio_bands(amps, np.stack(freqs), cuda)
    tf, ta, tc, bf, ba = (t.cpu().numpy() for t in (tf, ta, tc, bf, ba))
    for k in range(24):
        n = len(g[f"w{k}/top_amplitudes"])
        assert tc[k] == n
        np.testing.assert_allclose(ta[k][:n], g[f"w{k}/top_amplitudes"], rtol=1e-12)
        np.testing.assert_array_equal(tf[k][:n], g[f"w{k}/top_frequencies"])
        np.testing.assert_allclose(ba[k], g[f"w{k}/band_amplitudes"], rtol=1e-12)
        np.testing.assert_allclose(bf[k], g["band_frequencies"], rtol=1e-15)
    # more than 100 pass the gate, with many exact ties (quantised amplitudes): stable order decides
    rng = np.random.default_rng(3)
    a = np.round(rng.random((5, 4000)) * 13, 1)
    f = rng.random((5, 4000)) * 20000 - 10000
    f[3] = np.fft.fftfreq(4000, 1 / 16000)
    tf, ta, tc = (t.cpu().numpy() for t in audio_ops.audio_filter(a, f, cuda))
    for k in range(5):
        rf, ra = A.pe_audio_filter(a[k], f[k])
        assert tc[k] == len(ra) == 100
        np.testing.assert_array_equal(ta[k], ra)
        np.testing.assert_array_equal(tf[k], rf)
    # nothing passes / a single window as a 1-D array
    tf, ta, tc = audio_ops.audio_filter(np.zeros(64), np.arange(64.0), cuda)
    assert int(tc) == 0
    rf, ra = A.pe_audio_resampler(a[3], f[3])
    bf, ba = audio_ops.audio_bands(a[3], f[3], cuda)
    np.testing.assert_allclose(ba.cpu().numpy(), ra, rtol=1e-12)
    np.testing.assert_allclose(bf.cpu().numpy(), rf, rtol=1e-15)


def test_reference_audio_graph_definition_unchanged(cuda):
    """R/examples/pipeline/pipeline_mic_fft_graph.json: the PE_FFT / PE_AudioResampler / PE_AudioFilter
    element definitions verbatim (inputs: amplitudes, frequencies only), without the microphone and
    graph-window elements; module name as existing definitions would spell it for the media package."""
    from aiko_services_b200.pipeline import Pipeline
    mod = "aiko_services.elements.media.audio_io"
    two = [{"name": "amplitudes", "type": "bytes"}, {"name": "frequencies", "type": "bytes"}]
    d = {"version": 0, "name": "p_mic_fft_graph", "runtime": "python",
         "graph": ["(PE_FFT PE_AudioResampler PE_AudioFilter)"], "parameters": {},
         "elements": [
             {"name": "PE_FFT", "input": [{"name": "audio", "type": "bytes"}], "output": two,
              "deploy": {"local": {"module": mod}}},
             {"name": "PE_AudioResampler", "input": two, "output": [], "deploy": {"local": {"module": mod}}},
             {"name": "PE_AudioFilter", "input": two, "output": two, "deploy": {"local": {"module": mod}}}]}
    p = Pipeline(d)
    g = _golden_audio()
    w = windows_cfg5(24)
    assert p.create_stream("mic")
    for k in (0, 5, 17):
        okay, swag = p.process_frame({"stream_id": "mic", "frame_id": k}, {"audio": w[k]})
        assert okay
        ta = g[f"w{k}/top_amplitudes"]
        assert len(swag["amplitudes"]) == len(ta) == len(swag["frequencies"])
        np.testing.assert_allclose(swag["amplitudes"], ta, rtol=1e-4)
        np.testing.assert_allclose(np.sort(swag["frequencies"]), np.sort(g[f"w{k}/top_frequencies"]), rtol=1e-6, atol=1e-3)
        bf, ba = p.streams["mic"].variables["audio_bands"]
        np.testing.assert_allclose(ba, g[f"w{k}/band_amplitudes"], rtol=1e-4, atol=1e-4 * np.abs(ba).max())
    # arrays from another element (plain float64 ndarrays) go through the float64 GPU path
    flt = p.elements["PE_AudioFilter"]
    amp, freq = A.pe_fft(w[3], 16000)
    p._enable_thread_local(p.streams["mic"], 99)
    event, out = flt.process_frame(p.streams["mic"], amp, freq)
    p._disable_thread_local()
    np.testing.assert_allclose(out["amplitudes"], g["w3/top_amplitudes"], rtol=1e-12)
    np.testing.assert_array_equal(out["frequencies"], g["w3/top_frequencies"])
