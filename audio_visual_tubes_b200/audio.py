"""The spectrogram front-end of the reference's datasets (datasets/dataloader.py:86-96), batched on
the GPU: what ``__getitem__`` does per clip on a DataLoader worker -- tile to 10 s, clip, 
``scipy.signal.spectrogram(nperseg=512, noverlap=1)``, ``log(. + 1e-7)``, ``Normalize(0, 12)`` --
for a whole batch of waveforms in one ``avt_log_spectrogram`` launch."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib

NPERSEG = 512
LOG_EPS = 1e-7          # dataloader.py:95
STD = 12.0              # dataloader.py:72

_window_cache = {}


def tukey_window_periodic(m: int = NPERSEG, alpha: float = 0.25) -> np.ndarray:
    """``scipy.signal.get_window(('tukey', alpha), m)`` (the default window of
    ``scipy.signal.spectrogram``): the symmetric Tukey window of m + 1 points without its last."""
    n = np.arange(0, m + 1)
    width = int(np.floor(alpha * m / 2.0))
    n1, n2, n3 = n[0:width + 1], n[width + 1:m - width], n[m - width:]
    w1 = 0.5 * (1 + np.cos(np.pi * (-1 + 2.0 * n1 / alpha / m)))
    w2 = np.ones(n2.shape)
    w3 = 0.5 * (1 + np.cos(np.pi * (-2.0 / alpha + 1 + 2.0 * n3 / alpha / m)))
    return np.concatenate((w1, w2, w3))[:-1]


def _window(device):
    key = str(device)
    if key not in _window_cache:
        w = tukey_window_periodic()
        _window_cache[key] = (torch.tensor(w, dtype=torch.float64, device=device), float((w * w).sum()))
    return _window_cache[key]


def log_spectrogram(samples, samplerate: int, lengths=None):
    """samples [B, >= min(length, 10 s)] mono waveforms on the GPU -- fp32 in [-1, 1], or int16 PCM
    as stored in the WAV file (taken as sample / 32768, what ``sf.read`` returns) -- ``samplerate`` as
    ``sf.read`` reports it, ``lengths`` [B] valid samples per clip (shorter clips are tiled as
    dataloader.py:88-91 does; None = every row already holds 10 s).  Returns [B, 1, 257, T] fp32,
    the ``spectrogram`` tensor of ``__getitem__`` (dataloader.py:97) stacked by the DataLoader."""
    _lib.require_cuda(samples)
    lib = _lib.load()
    if samples.dim() != 2 or samples.dtype not in (torch.float32, torch.int16):
        raise ValueError("log_spectrogram: samples must be a [B, n] float32 or int16 tensor")
    samples = samples.contiguous()
    b = samples.shape[0]
    n_keep = int(samplerate) * 10
    dev = samples.device
    if lengths is None:
        if samples.shape[1] < n_keep:
            raise ValueError(f"log_spectrogram: rows hold {samples.shape[1]} samples, 10 s needs {n_keep} "
                             "(give `lengths` to tile shorter clips)")
        lens = None
    else:
        lens = torch.as_tensor(lengths, dtype=torch.int32).to(dev).contiguous()
        if lens.numel() != b or int(lens.min()) < 1 or int(torch.minimum(lens, torch.tensor(n_keep, device=dev)).max()) > samples.shape[1]:
            raise ValueError("log_spectrogram: lengths must be [B], >= 1 and within the rows of `samples`")
    t = lib.avt_spectrogram_frames(n_keep)
    win, sumsq = _window(dev)
    out = torch.empty(b, 1, 257, t, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        fn = lib.avt_log_spectrogram if samples.dtype == torch.float32 else lib.avt_log_spectrogram_pcm16
        rc = fn(_lib.ptr(samples), _lib.ptr(lens), b, samples.stride(0), n_keep, _lib.ptr(win),
                1.0 / (samplerate * sumsq), LOG_EPS, STD, _lib.ptr(out), t, _lib.stream_ptr())
    _lib.check(rc, "avt_log_spectrogram")
    return out
