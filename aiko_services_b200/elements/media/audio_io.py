"""GPU audio elements with the names of the reference's (dead-code) audio chain,
R/elements/media/audio_io.py:208-320: PE_FFT, PE_AudioResampler, PE_AudioFilter.

The three reference elements pass float64 arrays between them; here PE_FFT runs the whole fused
kernel once per frame (or once for all concurrent streams) and hands the later elements its
device-resident results, so PE_AudioResampler and PE_AudioFilter only pick their outputs.
``audio`` may be one window (int16 / float32 ndarray [n] or [n, 1]) or a batch [windows, n].
Parameter ``sample_rate`` (default 16000, PA_AUDIO_SAMPLE_RATE at audio_io.py:379).
"""
import numpy as np

from ...pipeline import PipelineElement
from ...stream import StreamEvent

__all__ = ["PE_AudioFilter", "PE_AudioResampler", "PE_FFT"]


class _Spectrum:
    """Result of the fused chain for one frame; arrays come to the host on first use."""

    def __init__(self, result, rows, single):
        self.result, self.rows, self.single = result, rows, single
        self._host = {}

    def get(self, key):
        if key not in self._host:
            value = self.result[key]
            if hasattr(value, "cpu"):
                value = value[self.rows[0]:self.rows[1]].cpu().numpy()
                value = value[0] if self.single else value
            self._host[key] = value
        return self._host[key]


def _as_windows(audio):
    a = np.asarray(audio)
    single = a.ndim == 1 or (a.ndim == 2 and a.shape[1] == 1)
    if single:
        a = a.reshape(1, -1)
    if a.dtype not in (np.int16, np.float32):
        a = a.astype(np.float32)
    return np.ascontiguousarray(a), single


def _run_chain(element, calls):
    import torch
    from ... import audio_ops
    from .image_io import device_for
    results = [None] * len(calls)
    groups = {}
    for ci, (stream, audio) in enumerate(calls):
        windows, single = _as_windows(audio)
        rate, _ = element.get_parameter("sample_rate", default=16000)
        key = (device_for(stream.stream_id), windows.dtype.str, windows.shape[1], float(rate))
        groups.setdefault(key, []).append((ci, windows, single))
    for (device_index, _, n, rate), members in groups.items():
        batch = np.concatenate([w for _, w, _ in members])
        device = torch.device("cuda", device_index)
        out = audio_ops.audio_chain(torch.from_numpy(batch).pin_memory().to(device, non_blocking=True), rate)
        row = 0
        for ci, windows, single in members:
            spectrum = _Spectrum(out, (row, row + len(windows)), single)
            row += len(windows)
            results[ci] = (StreamEvent.OKAY, {"amplitudes": _LazyArray(spectrum, "amplitudes"),
                                              "frequencies": out["frequencies"], "spectrum": spectrum})
    return results


class _LazyArray:
    """float32 amplitudes that stay on the GPU until a consumer needs them as an ndarray."""

    def __init__(self, spectrum, key):
        self.spectrum, self.key = spectrum, key

    def __array__(self, dtype=None, copy=None):
        a = self.spectrum.get(self.key)
        return a.astype(dtype) if dtype is not None else a

    def __len__(self):
        return len(np.asarray(self))

    def __getitem__(self, item):
        return np.asarray(self)[item]


class PE_FFT(PipelineElement):
    """audio_io.py:301-320.  Output ``spectrum`` carries the fused results for the elements below."""
    batch_across_streams = True

    def process_frame(self, stream, audio):
        return _run_chain(self, [(stream, audio)])[0]

    def process_frame_batch(self, calls, pipeline):
        prepared = []
        for stream, inputs in calls:
            pipeline._enable_thread_local(stream, stream.frame_id)
            prepared.append((stream, inputs["audio"]))
        try:
            return _run_chain(self, prepared)
        finally:
            pipeline._disable_thread_local()


class PE_AudioResampler(PipelineElement):
    """audio_io.py:239-292 without the LED / MQTT side effect: the 8 band sums."""

    def process_frame(self, stream, amplitudes, frequencies, spectrum=None):
        if spectrum is None:
            return StreamEvent.ERROR, {"diagnostic": "PE_AudioResampler needs PE_FFT's spectrum output"}
        return StreamEvent.OKAY, {"frequencies": spectrum.get("band_frequencies"),
                                  "amplitudes": spectrum.get("band_amplitudes")}


class PE_AudioFilter(PipelineElement):
    """audio_io.py:208-230: spectral gate and the 100 loudest bins, as (amplitudes, frequencies)."""

    def process_frame(self, stream, amplitudes, frequencies, spectrum=None):
        if spectrum is None:
            return StreamEvent.ERROR, {"diagnostic": "PE_AudioFilter needs PE_FFT's spectrum output"}
        count = spectrum.get("top_count")
        f, a = spectrum.get("top_frequencies"), spectrum.get("top_amplitudes")
        if np.ndim(count) == 0:
            f, a = tuple(f[:int(count)]), tuple(a[:int(count)])
        else:
            f = [tuple(row[:int(c)]) for row, c in zip(f, count)]
            a = [tuple(row[:int(c)]) for row, c in zip(a, count)]
        return StreamEvent.OKAY, {"amplitudes": a, "frequencies": f}
