"""Python-facing audio chain over the C ABI (PE_FFT -> PE_AudioResampler -> PE_AudioFilter,
R/elements/media/audio_io.py:202-320), batched over windows resident on the GPU."""
import ctypes

import numpy as np
import torch

from . import _native
from ._native import check
from .ops import _T, _require_cuda, _stream_int

AF_FREQUENCY_MINIMUM, AF_FREQUENCY_MAXIMUM = 10, 9000     # audio_io.py:204-205
AR_BAND_COUNT = 8                                         # audio_io.py:237

_plans = {}


class AudioPlan:
    """Per (device, window length, sample rate): twiddle workspace and the float64-evaluated
    constants of the chain (fftfreq, band edges, frequency gate)."""

    def __init__(self, device, n, sample_rate):
        self.n, self.sample_rate = n, float(sample_rate)
        lib = _native.lib()
        nbytes = lib.aiko_b200_audio_workspace_bytes(n)
        if nbytes == 0:
            raise ValueError(f"audio window length {n} must be even and >= 4")
        self.workspace = torch.empty(nbytes, dtype=torch.uint8, device=device)
        with torch.cuda.device(device):
            _T().audio_prepare(self.workspace, n, _stream_int())
        torch.cuda.current_stream(device).synchronize()
        self.frequencies = np.fft.fftfreq(n, 1 / self.sample_rate)             # PE_FFT, :312-313
        half = self.frequencies[0:n // 2]                                      # PE_AudioResampler, :253-257
        band_width = (half[-1] - half[0]) / AR_BAND_COUNT / 10
        ranges = []
        for band in range(AR_BAND_COUNT):
            start, end = band * band_width, band * band_width + band_width
            idx = np.nonzero((half >= start) & (half < end))[0]
            ranges += [int(idx[0]), int(idx[-1]) + 1] if len(idx) else [0, 0]
        self.band_ranges = [int(v) for v in ranges]
        self.band_frequencies = np.array([(b * band_width + b * band_width + band_width) / 2
                                          for b in range(AR_BAND_COUNT)])
        folded = np.abs(self.frequencies[:n // 2 + 1])                         # |f| of bins k = 0 .. n/2
        ok = np.nonzero((folded >= AF_FREQUENCY_MINIMUM) & (folded <= AF_FREQUENCY_MAXIMUM))[0]
        self.gate = (int(ok[0]), int(ok[-1])) if len(ok) else (1, 0)


def plan(device, n, sample_rate):
    key = (str(device), n, float(sample_rate))
    if key not in _plans:
        _plans[key] = AudioPlan(device, n, sample_rate)
    return _plans[key]


def audio_chain(audio, sample_rate, *, bands=True, top=True, stream=None):
    """audio: int16 or float32 CUDA tensor [windows, n].  Returns a dict with ``amplitudes``
    [windows, n] float32, ``frequencies`` (numpy float64 [n], the same for every window),
    and, when asked, ``band_frequencies`` [8], ``band_amplitudes`` [windows, 8],
    ``top_frequencies`` / ``top_amplitudes`` [windows, 100], ``top_count`` [windows]."""
    if audio.dtype not in (torch.int16, torch.float32):
        raise TypeError(f"audio_chain: int16 or float32 expected, got {audio.dtype}")
    _require_cuda(audio, "audio_chain audio", audio.dtype)
    if audio.dim() != 2:
        raise ValueError("audio_chain: expected [windows, samples]")
    b, n = audio.shape
    pl = plan(audio.device, n, sample_rate)
    dev = audio.device
    out = {"frequencies": pl.frequencies,
           "amplitudes": torch.empty((b, n), dtype=torch.float32, device=dev)}
    band = torch.empty((b, 8), dtype=torch.float32, device=dev) if bands else None
    tf = torch.empty((b, 100), dtype=torch.float32, device=dev) if top else None
    ta = torch.empty((b, 100), dtype=torch.float32, device=dev) if top else None
    tc = torch.empty((b,), dtype=torch.int32, device=dev) if top else None
    if b == 0:
        pass
    else:
        with torch.cuda.device(dev):
            _T().audio_chain(audio, pl.sample_rate, out["amplitudes"], band, pl.band_ranges, tf, ta, tc, pl.gate[0],
                             pl.gate[1], pl.workspace, _stream_int(stream))
    if bands:
        out["band_frequencies"], out["band_amplitudes"] = pl.band_frequencies, band
    if top:
        out["top_frequencies"], out["top_amplitudes"], out["top_count"] = tf, ta, tc
    return out


def _f64_windows(x, device, name):
    """ndarray / sequence / tensor -> contiguous float64 CUDA tensor [windows, n]; -> (tensor, single)"""
    if not isinstance(x, torch.Tensor):
        x = torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float64)))
    single = x.dim() == 1
    if single:
        x = x[None]
    if x.dim() != 2:
        raise ValueError(f"{name}: expected [n] or [windows, n]")
    return x.to(device=device, dtype=torch.float64).contiguous(), single


def audio_filter(amplitudes, frequencies, device=None, stream=None):
    """PE_AudioFilter (audio_io.py:216-227) on arrays from anywhere: float64 in and out, on the GPU.
    -> (frequencies [windows, 100], amplitudes [windows, 100], count [windows]) CUDA tensors
    (leading dimension dropped for a single window)."""
    device = torch.device(device if device is not None else
                          (amplitudes.device if isinstance(amplitudes, torch.Tensor) and amplitudes.is_cuda else "cuda"))
    a, single = _f64_windows(amplitudes, device, "audio_filter amplitudes")
    f, f_single = _f64_windows(frequencies, device, "audio_filter frequencies")
    b, n = a.shape
    if f.shape[1] != n or f.shape[0] not in (1, b):
        raise ValueError("audio_filter: amplitudes and frequencies differ in shape")
    tf = torch.empty((b, 100), dtype=torch.float64, device=device)
    ta = torch.empty((b, 100), dtype=torch.float64, device=device)
    tc = torch.empty((b,), dtype=torch.int32, device=device)
    with torch.cuda.device(device):
        _T().audio_filter_f64(a, f, tf, ta, tc, _stream_int(stream))
    return (tf[0], ta[0], tc[0]) if single else (tf, ta, tc)


def audio_bands(amplitudes, frequencies, device=None, stream=None):
    """PE_AudioResampler (audio_io.py:253-277) on arrays from anywhere.
    -> (band frequencies [windows, 8], band amplitudes [windows, 8]) float64 CUDA tensors."""
    device = torch.device(device if device is not None else
                          (amplitudes.device if isinstance(amplitudes, torch.Tensor) and amplitudes.is_cuda else "cuda"))
    a, single = _f64_windows(amplitudes, device, "audio_bands amplitudes")
    f, _ = _f64_windows(frequencies, device, "audio_bands frequencies")
    b, n = a.shape
    if f.shape[1] != n or f.shape[0] not in (1, b):
        raise ValueError("audio_bands: amplitudes and frequencies differ in shape")
    bf = torch.empty((b, 8), dtype=torch.float64, device=device)
    ba = torch.empty((b, 8), dtype=torch.float64, device=device)
    with torch.cuda.device(device):
        _T().audio_bands_f64(a, f, bf, ba, _stream_int(stream))
    return (bf[0], ba[0]) if single else (bf, ba)
