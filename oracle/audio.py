"""numpy float64 restatement of the reference's audio chain -- TEST INFRASTRUCTURE.

The reference code sits inside a module-level string literal and cannot be imported
(R/elements/media/audio_io.py:162-642).  PINNED all the same: tests/golden/make_golden_audio.py cuts
the statement lines of the three process_frame bodies out of the reference file at run time,
exec()s them on the seeded config-5 windows and stores the results (tests/golden/golden_audio.npz);
tests/test_oracle_golden.py checks this restatement against those vectors, and the generator
asserts exact equality when it runs.  Each function follows the cited lines; only what cannot
execute is changed
(``StreamEvent.AIKO`` does not exist; ``fftfreq`` uses the window's own length and rate instead
of the microphone constants of :312-313).  The MQTT / LED side effect of :279-290 is left out.
"""
import numpy as np

FFT_AMPLITUDE_SCALER = 1_000_000     # audio_io.py:299
AF_AMPLITUDE_MINIMUM = 0.1           # audio_io.py:202-206
AF_AMPLITUDE_MAXIMUM = 12
AF_FREQUENCY_MINIMUM = 10
AF_FREQUENCY_MAXIMUM = 9000
AF_SAMPLES_MAXIMUM = 100
AR_BAND_COUNT = 8                    # audio_io.py:237


def pe_fft(audio, sample_rate):
    """PE_FFT.process_frame, audio_io.py:306-320 -> (amplitudes, frequencies)"""
    audio = np.asarray(audio).reshape(-1)
    fft_output = np.fft.fft(audio)
    amplitudes = np.abs(fft_output / FFT_AMPLITUDE_SCALER)
    frequencies = np.fft.fftfreq(len(audio), 1 / sample_rate)
    return amplitudes, frequencies


def pe_audio_filter(amplitudes, frequencies):
    """PE_AudioFilter.process_frame, audio_io.py:213-230 -> (frequencies, amplitudes), <= 100 each"""
    data = list(zip(np.abs(frequencies), amplitudes))
    data = [(x, y) for x, y in data if y >= AF_AMPLITUDE_MINIMUM]
    data = [(x, y) for x, y in data if x >= AF_FREQUENCY_MINIMUM]
    data = [(x, y) for x, y in data if y <= AF_AMPLITUDE_MAXIMUM]
    data = [(x, y) for x, y in data if x <= AF_FREQUENCY_MAXIMUM]
    data = sorted(data, key=lambda x: x[1], reverse=True)
    data = data[:AF_SAMPLES_MAXIMUM]
    if len(data):
        f, a = zip(*data)
    else:
        f, a = [], []
    return np.asarray(f, np.float64), np.asarray(a, np.float64)


def band_edges(n, sample_rate):
    """Index ranges [lo, hi) of the 8 bands of PE_AudioResampler for an n-sample window
    (audio_io.py:250-268, evaluated in float64 exactly as the reference's masks)."""
    frequencies = np.fft.fftfreq(n, 1 / sample_rate)[0:n // 2]
    frequency_range = frequencies[-1] - frequencies[0]
    band_width = frequency_range / AR_BAND_COUNT / 10
    edges = []
    for band in range(AR_BAND_COUNT):
        start = band * band_width
        end = start + band_width
        idx = np.nonzero((frequencies >= start) & (frequencies < end))[0]
        edges.append((int(idx[0]), int(idx[-1]) + 1) if len(idx) else (0, 0))
    centres = [(b * band_width + (b * band_width + band_width)) / 2 for b in range(AR_BAND_COUNT)]
    return edges, np.asarray(centres)


def pe_audio_resampler(amplitudes, frequencies):
    """PE_AudioResampler.process_frame, audio_io.py:250-277 -> (band frequencies[8], band amplitudes[8])"""
    amplitudes = amplitudes[0:len(amplitudes) // 2]
    frequencies = frequencies[0:len(frequencies) // 2]
    frequency_range = frequencies[-1] - frequencies[0]
    band_width = frequency_range / AR_BAND_COUNT / 10
    band_frequencies, band_amplitudes = [], []
    for band in range(AR_BAND_COUNT):
        band_start = band * band_width
        band_end = band_start + band_width
        mask = (frequencies >= band_start) & (frequencies < band_end)
        band_frequencies.append((band_start + band_end) / 2)
        band_amplitudes.append(np.sum(amplitudes[mask]))
    return np.array(band_frequencies), np.array(band_amplitudes)
