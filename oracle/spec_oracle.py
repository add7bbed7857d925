"""CPU oracle of the spectrogram front-end -- TEST INFRASTRUCTURE, never imported by the product.

Restates datasets/dataloader.py:86-96 of the reference:
    samples tiled to 10 s (:88-90), cut (:91), clipped to [-1, 1] (:92-93),
    scipy.signal.spectrogram(x, fs, nperseg=512, noverlap=1) (:94), log(S + 1e-7) (:95),
    ToTensor + Normalize(mean 0, std 12) (:72, :96).
The spectrogram itself lives in a third-party dependency (scipy, unpinned by the reference; 1.18
in this image); its published algorithm (scipy.signal.spectrogram -> _spectral_helper: hop
nperseg - noverlap segments, detrend='constant', window ('tukey', .25) periodic, rfft,
conj(X)*X * 1/(fs * sum(w^2)), interior bins doubled) is written out in numpy below and pinned
against scipy.signal.spectrogram itself in tests/test_spectrogram_cpu.py.
"""
from __future__ import annotations

import numpy as np

NPERSEG, NOVERLAP = 512, 1
LOG_EPS, STD = 1e-7, 12.0


def tukey_periodic(m=NPERSEG, alpha=0.25):
    """scipy.signal.windows.tukey(m + 1, alpha, sym=True)[:-1] == get_window(('tukey', alpha), m)."""
    M = m + 1
    n = np.arange(0, M)
    width = int(np.floor(alpha * (M - 1) / 2.0))
    n1, n2, n3 = n[0:width + 1], n[width + 1:M - width - 1], n[M - width - 1:]
    w1 = 0.5 * (1 + np.cos(np.pi * (-1 + 2.0 * n1 / alpha / (M - 1))))
    w2 = np.ones(n2.shape)
    w3 = 0.5 * (1 + np.cos(np.pi * (-2.0 / alpha + 1 + 2.0 * n3 / alpha / (M - 1))))
    return np.concatenate((w1, w2, w3))[:-1]


def prepare_samples(samples, samplerate):
    """dataloader.py:88-93 (float64, as sf.read returns)."""
    samples = np.asarray(samples, dtype=np.float64)
    if samples.shape[0] < samplerate * 10:
        n = int(samplerate * 10 / samples.shape[0]) + 1
        samples = np.tile(samples, n)
    re = samples[:samplerate * 10].copy()
    re[re > 1.] = 1.
    re[re < -1.] = -1.
    return re


def psd_spectrogram(x, fs):
    """scipy.signal.spectrogram(x, fs, nperseg=512, noverlap=1)[2] restated: [257, T] float64."""
    x = np.asarray(x, dtype=np.float64)
    step = NPERSEG - NOVERLAP
    T = (x.shape[0] - NOVERLAP) // step
    win = tukey_periodic()
    scale = 1.0 / (fs * (win * win).sum())
    segs = np.stack([x[t * step:t * step + NPERSEG] for t in range(T)])          # [T, 512]
    segs = segs - segs.mean(axis=-1, keepdims=True)                               # detrend='constant'
    X = np.fft.rfft(segs * win, n=NPERSEG, axis=-1)
    P = (np.conjugate(X) * X).real * scale
    P[:, 1:-1] *= 2                                                               # one-sided, nfft even
    return P.T


def log_spectrogram(samples, samplerate):
    """One clip through dataloader.py:86-96 -> [1, 257, T] float32 (after the loops' .float())."""
    x = prepare_samples(samples, samplerate)
    S = np.log(psd_spectrogram(x, samplerate) + LOG_EPS)
    return ((S - 0.0) / STD).astype(np.float32)[None]


def synthetic_waveforms(b, samplerate, seed=0, seconds=10.0):
    """Seeded test audio quantised like 16-bit PCM: a few chirps + noise, some samples beyond
    +-1 so the clip at dataloader.py:92-93 acts, silence in places (log floor)."""
    rng = np.random.default_rng(seed)
    n = int(samplerate * seconds)
    t = np.arange(n) / samplerate
    out = np.empty((b, n), np.float32)
    for i in range(b):
        f0, f1 = rng.uniform(80, 800), rng.uniform(1000, samplerate / 2.2)
        x = 0.5 * np.sin(2 * np.pi * (f0 + (f1 - f0) * t / t[-1] / 2) * t)
        x += rng.uniform(0.01, 0.4) * rng.standard_normal(n)
        x *= rng.uniform(0.3, 1.6)
        s0 = rng.integers(0, n - n // 8)
        x[s0:s0 + n // 16] = 0.0
        out[i] = (np.round(x * 32768.0) / 32768.0).astype(np.float32)
    return out
