"""Generate tests/golden/golden_audio.npz by EXECUTING THE REFERENCE'S OWN STATEMENTS.

The reference's audio chain (PE_FFT, PE_AudioResampler, PE_AudioFilter) sits inside a module-level
string literal (R/elements/media/audio_io.py:162-642) and cannot be imported.  Its process_frame
bodies are nevertheless plain numpy / Python statements.  This script reads the reference file at
run time, cuts out the statement lines of the three bodies and the constants they use (located by
their first and last statement, so nothing is copied into this repository), and exec()s them on the
seeded windows of BASELINE config 5.  What is left out is only what cannot execute or is a side
effect: the broken ``return`` lines (:229-230, :319-320), the logger line (:315-317) and the MQTT /
LED block (:279-290).  ``PA_AUDIO_CHUNK_SIZE`` / ``PA_AUDIO_SAMPLE_RATE`` (module constants used by
np.fft.fftfreq at :312-313) are bound to the window's own length and rate.

Runs only where /root/reference exists (the build container):

    python tests/golden/make_golden_audio.py
"""
import os
import sys
import textwrap

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = "/root/reference/src/aiko_services/elements/media/audio_io.py"
OUT = os.path.join(HERE, "golden_audio.npz")
N_FULL = 3          # windows whose full amplitude arrays are stored (float64, 128 KB each)


def windows_cfg5(count, n=16000, fs=16000):
    """int16 windows 3000 sin(2 pi f_k t) + N(0, 200), f_k log-spaced 50..7000 Hz."""
    t = np.arange(n) / fs
    freqs = np.geomspace(50, 7000, count)
    out = np.empty((count, n), np.int16)
    for k in range(count):
        rng = np.random.default_rng(5000 + k)
        out[k] = np.clip(3000 * np.sin(2 * np.pi * freqs[k] * t) + rng.normal(0, 200, n), -32768, 32767).astype(np.int16)
    return out


def _cut(lines, first, last, after=0):
    """Statement lines from the one starting with ``first`` to the one starting with ``last`` (both
    inclusive), searching from line ``after``; -> (dedented source, (first line no, last line no))."""
    a = next(i for i in range(after, len(lines)) if lines[i].strip().startswith(first))
    b = next(i for i in range(a, len(lines)) if lines[i].strip().startswith(last))
    return textwrap.dedent("".join(lines[a:b + 1])), (a + 1, b + 1)


def reference_bodies():
    lines = open(REFERENCE).readlines()
    at = lambda text: next(i for i, l in enumerate(lines) if l.startswith(text))
    consts_filter, c1 = _cut(lines, "AF_AMPLITUDE_MINIMUM", "AF_SAMPLES_MAXIMUM")
    consts_bands, c2 = _cut(lines, "AR_BAND_COUNT", "AR_BAND_COUNT")
    consts_fft, c3 = _cut(lines, "FFT_AMPLITUDE_SCALER", "FFT_AMPLITUDE_SCALER")
    body_filter, b1 = _cut(lines, "data = list(zip(", "frequencies, amplitudes = [], []", at("class PE_AudioFilter"))
    body_bands, b2 = _cut(lines, "amplitudes = amplitudes[0:len(amplitudes) // 2]", "amplitudes = np.array(band_amplitudes)",
                          at("class PE_AudioResampler"))
    body_fft, b3 = _cut(lines, "fft_output = np.fft.fft(audio)", "PA_AUDIO_CHUNK_SIZE, 1 / PA_AUDIO_SAMPLE_RATE)",
                        at("class PE_FFT"))
    cited = {"filter_constants": c1, "band_constants": c2, "fft_constants": c3, "filter_body": b1, "band_body": b2,
             "fft_body": b3}
    return consts_filter + consts_bands + consts_fft, body_fft, body_bands, body_filter, cited


def run_reference(audio, fs, bodies=None):
    """-> dict of the three elements' outputs for one window, computed by the reference's statements."""
    consts, body_fft, body_bands, body_filter, _ = bodies or reference_bodies()
    env = {"np": np}
    exec(consts, env)
    ns = dict(env, audio=audio, PA_AUDIO_CHUNK_SIZE=len(audio), PA_AUDIO_SAMPLE_RATE=fs)
    exec(body_fft, ns)
    amplitudes, frequencies = ns["amplitudes"], ns["frequencies"]
    nb = dict(env, amplitudes=amplitudes, frequencies=frequencies)
    exec(body_bands, nb)
    nf = dict(env, amplitudes=amplitudes, frequencies=frequencies)
    exec(body_filter, nf)
    return {"amplitudes": amplitudes, "frequencies": frequencies,
            "band_frequencies": nb["frequencies"], "band_amplitudes": nb["amplitudes"],
            "top_frequencies": np.asarray(nf["frequencies"], np.float64),
            "top_amplitudes": np.asarray(nf["amplitudes"], np.float64)}


def main():
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    from oracle import audio as A
    bodies = reference_bodies()
    fs = 16000
    w = windows_cfg5(24)
    store = {"versions": np.array(f"numpy {np.__version__}"), "sample_rate": np.int64(fs),
             "cited_lines": np.array(repr(bodies[4]))}
    for k in range(24):
        ref = run_reference(w[k], fs, bodies)
        # the restatement must agree with the reference's own statements exactly (same numpy calls)
        amp, freq = A.pe_fft(w[k], fs)
        assert np.array_equal(amp, ref["amplitudes"]) and np.array_equal(freq, ref["frequencies"])
        bf, ba = A.pe_audio_resampler(amp, freq)
        assert np.array_equal(bf, ref["band_frequencies"]) and np.array_equal(ba, ref["band_amplitudes"])
        tf, ta = A.pe_audio_filter(amp, freq)
        assert np.array_equal(tf, ref["top_frequencies"]) and np.array_equal(ta, ref["top_amplitudes"])
        if k < N_FULL:
            store[f"w{k}/amplitudes"] = ref["amplitudes"]
        store[f"w{k}/band_amplitudes"] = ref["band_amplitudes"]
        store[f"w{k}/top_frequencies"] = ref["top_frequencies"]
        store[f"w{k}/top_amplitudes"] = ref["top_amplitudes"]
        store[f"w{k}/amplitude_probe"] = ref["amplitudes"][::97].copy()     # every 97th bin of every window
    store["frequencies"] = ref["frequencies"]
    store["band_frequencies"] = ref["band_frequencies"]
    # a float32 window of another length / rate (sounddevice path, audio_io.py:506-513)
    rng = np.random.default_rng(77)
    x = (rng.standard_normal(9600) * 4000).astype(np.float32)
    x += 30000 * np.sin(2 * np.pi * 440 * np.arange(9600) / 48000).astype(np.float32)
    ref = run_reference(x, 48000, bodies)
    for key, value in ref.items():
        store[f"f32_9600/{key}"] = value
    np.savez_compressed(OUT, **store)
    print(OUT, os.path.getsize(OUT), "bytes; reference lines used:", bodies[4])


if __name__ == "__main__":
    main()
