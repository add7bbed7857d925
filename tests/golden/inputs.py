"""Seeded synthetic inputs shared by the golden-vector generator, the parity tests,
smoke() and bench.py.  Pure numpy (PCG64 streams are platform independent), so the
GPU box regenerates bit-identical inputs and only the *outputs* of the reference
need to be committed as fixtures."""
from __future__ import annotations

import numpy as np

C = 512
P = 196


def head_features(n_v, n_a, seed, mode="relu", c=C, p=P):
    """Raw (un-normalised) visual maps v[n_v,c,p] and pooled audio a[n_a,c], fp32.

    mode 'relu'  : |randn| for both -- post-ReLU-like, S ~ 0.75..0.82 (SURVEY 8d cfg 4)
    mode 'mix'   : each position is a random blend of its positive audio vector, one
                   random other audio vector and noise, so S spans roughly 0.2..0.98 and
                   both sigmoids (eps = .65, eps2 = .4) leave saturation
    mode 'signed': plain randn -- S ~ 0, all masks tiny (relative-accuracy stress)
    """
    rng = np.random.default_rng(seed)
    a = np.abs(rng.standard_normal((n_a, c))).astype(np.float32)
    if mode == "signed":
        v = rng.standard_normal((n_v, c, p)).astype(np.float32)
        a = rng.standard_normal((n_a, c)).astype(np.float32)
        return v, a
    noise = np.abs(rng.standard_normal((n_v, c, p))).astype(np.float32)
    if mode == "relu":
        return noise, a
    if mode == "mix":
        al = rng.uniform(0.0, 0.9, (n_v, 1, p)).astype(np.float32)
        be = rng.uniform(0.0, 0.5, (n_v, 1, p)).astype(np.float32)
        other = rng.integers(0, n_a, n_v)
        own = a[np.arange(n_v) % n_a][:, :, None]
        oth = a[other][:, :, None]
        sparse = (rng.uniform(size=(n_v, c, p)) < 0.5).astype(np.float32)
        v = al * own + be * oth + (1 - al) * noise * sparse * 1.5
        scale = rng.uniform(0.2, 3.0, (n_v, 1, p)).astype(np.float32)   # norms vary per position
        return (v * scale).astype(np.float32), a
    raise ValueError(mode)


def audio_maps_from_pooled(a, seed, hw=(17, 19)):
    """An un-pooled audio feature map [n_a,c,17,19] whose spatial max is exactly ``a``
    (the reference max-pools a 17x19 map, model.py:121)."""
    rng = np.random.default_rng(seed + 7919)
    n_a, c = a.shape
    m = a[:, :, None, None] - np.abs(rng.standard_normal((n_a, c) + hw)).astype(np.float32) - 1e-3
    iy = rng.integers(0, hw[0], (n_a, c)); ix = rng.integers(0, hw[1], (n_a, c))
    m[np.arange(n_a)[:, None], np.arange(c)[None, :], iy, ix] = a
    return m.astype(np.float32)
