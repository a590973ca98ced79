"""GPU audio elements with the names of the reference's (dead-code) audio chain,
R/elements/media/audio_io.py:208-320: PE_FFT, PE_AudioResampler, PE_AudioFilter.

The three reference elements pass float64 arrays between them and their pipeline definition
(R/examples/pipeline/pipeline_mic_fft_graph.json:19-44) gives the two later elements exactly
``amplitudes`` and ``frequencies``.  Here PE_FFT runs the whole fused kernel once per frame (or once
for all concurrent streams); its ``amplitudes`` output is a lazy array that remembers the fused
results, so PE_AudioResampler and PE_AudioFilter -- behind the unchanged definition -- only pick
theirs.  Arrays that come from anywhere else (any element may produce ``amplitudes`` /
``frequencies``) are filtered / banded on the GPU in float64 (csrc/audio_arrays.cu).
``audio`` may be one window (int16 / float32 ndarray [n] or [n, 1]) or a batch [windows, n].
Parameter ``sample_rate`` (default 16000, PA_AUDIO_SAMPLE_RATE at audio_io.py:379).
"""
import numpy as np

from ...pipeline import PipelineElement
from ...stream import StreamEvent

__all__ = ["PE_AudioFilter", "PE_AudioResampler", "PE_FFT"]


class _PinnedRing:
    """Page-locked staging for audio windows, allocated once per GPU and reused: the event-loop
    thread never calls cudaHostAlloc per frame."""

    def __init__(self, slots=4):
        self.buffers = [None] * slots
        self.events = [None] * slots
        self.next = 0

    def stage(self, batch, device, stream):
        import torch
        i = self.next
        self.next = (i + 1) % len(self.buffers)
        if self.events[i] is not None:
            self.events[i].synchronize()          # the copy out of this buffer has run
        nbytes = batch.nbytes
        if self.buffers[i] is None or self.buffers[i].numel() < nbytes:
            self.buffers[i] = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8).pin_memory()
        host = self.buffers[i][:nbytes].view(torch.int16 if batch.dtype == np.int16 else torch.float32).view(batch.shape)
        host.numpy()[...] = batch
        dev = host.to(device, non_blocking=True)
        if self.events[i] is None:
            self.events[i] = torch.cuda.Event()
        self.events[i].record(stream)
        return dev


_rings = {}


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
        ring = _rings.setdefault(device_index, _PinnedRing())
        with torch.cuda.device(device):
            out = audio_ops.audio_chain(ring.stage(batch, device, torch.cuda.current_stream(device)), rate)
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


def _fused(amplitudes, frequencies):
    """The fused results behind a PE_FFT ``amplitudes`` output, if that is what these two arrays are."""
    spectrum = getattr(amplitudes, "spectrum", None)
    if spectrum is not None and frequencies is spectrum.result["frequencies"]:
        return spectrum
    return None


def _device_of(stream):
    import torch
    from .image_io import device_for
    return torch.device("cuda", device_for(stream.stream_id))


class PE_AudioResampler(PipelineElement):
    """audio_io.py:239-292: the 8 band sums.  Like the reference it returns no outputs (its bands
    only fed the LED / MQTT side effect of :279-290, which is not reproduced); with the parameter
    ``emit_bands`` set it returns them as ``band_frequencies`` / ``band_amplitudes``.  The latest
    bands are also kept in ``stream.variables["audio_bands"]``."""

    def process_frame(self, stream, amplitudes, frequencies, spectrum=None):
        from ... import audio_ops
        fused = spectrum if spectrum is not None else _fused(amplitudes, frequencies)
        if fused is not None:
            bf, ba = fused.get("band_frequencies"), fused.get("band_amplitudes")
        else:
            bf, ba = audio_ops.audio_bands(np.asarray(amplitudes), np.asarray(frequencies), _device_of(stream))
            bf, ba = bf.cpu().numpy(), ba.cpu().numpy()
        stream.variables["audio_bands"] = (bf, ba)
        emit, _ = self.get_parameter("emit_bands", default=False)
        if str(emit).lower() in ("1", "true", "yes"):
            return StreamEvent.OKAY, {"band_frequencies": bf, "band_amplitudes": ba}
        return StreamEvent.OKAY, {}


class PE_AudioFilter(PipelineElement):
    """audio_io.py:208-230: spectral gate and the 100 loudest bins, as (amplitudes, frequencies)."""

    def process_frame(self, stream, amplitudes, frequencies, spectrum=None):
        from ... import audio_ops
        fused = spectrum if spectrum is not None else _fused(amplitudes, frequencies)
        if fused is not None:
            count = fused.get("top_count")
            f, a = fused.get("top_frequencies"), fused.get("top_amplitudes")
        else:
            tf, ta, tc = audio_ops.audio_filter(np.asarray(amplitudes), np.asarray(frequencies), _device_of(stream))
            f, a, count = tf.cpu().numpy(), ta.cpu().numpy(), tc.cpu().numpy()
        if np.ndim(count) == 0:
            f, a = tuple(f[:int(count)]), tuple(a[:int(count)])
        else:
            f = [tuple(row[:int(c)]) for row, c in zip(f, count)]
            a = [tuple(row[:int(c)]) for row, c in zip(a, count)]
        return StreamEvent.OKAY, {"amplitudes": a, "frequencies": f}
