"""Host logic of the de-duplicated tube mode (no GPU): the compact -> reference logits expansion
and its adjoint, checked against the CPU oracle run on T copies of every audio row
(train_3D.py:129-130, model.py:47-59)."""
import numpy as np
import torch

from inputs import head_features
from oracle import head_oracle as ho
from audio_visual_tubes_b200.head import (HeadConfig, expand_tube_logits, reduce_tube_dlogits,
                                          tube_positions)


def _compact(v, a, T):
    n = v.shape[0]
    pos, copy = tube_positions(n, T, "cpu")
    o = ho.head_forward(v, a, pos=pos.numpy())
    out = dict(logits=torch.tensor(o["logits"]), sim=torch.tensor(o["sim"]), pos_idx=pos, col_mult=T)
    return out, copy


def test_tube_positions_follow_the_b_t_flattening():
    pos, copy = tube_positions(12, 4, "cpu")
    assert pos.tolist() == [0] * 4 + [1] * 4 + [2] * 4
    assert copy.tolist() == [0, 1, 2, 3] * 3
    pos, copy = tube_positions(4, 4, "cpu", row_offset=8)         # a row shard of the same batch
    assert pos.tolist() == [2] * 4 and copy.tolist() == [0, 1, 2, 3]


def test_expanded_logits_equal_the_head_on_repeated_audio():
    b, T = 3, 4
    v, a = head_features(b * T, b, 11, "mix")
    cfg = HeadConfig()
    out, copy = _compact(v, a, T)
    full = expand_tube_logits(out, cfg, copy).numpy()
    ref = ho.head_forward(v, np.repeat(a, T, axis=0))["logits"]  # identity positives, (bT)+2 columns
    assert full.shape == ref.shape == (b * T, b * T + 2)
    assert np.abs(full - ref).max() <= 1e-9 * np.abs(ref).max()
    # exactly one masked (negative, -99x) entry per row, on the diagonal
    n = b * T
    assert np.all(full[np.arange(n), 1 + np.arange(n)] < 0)


def test_reduce_is_the_adjoint_of_expand():
    """<g_full, d full / d (sim, first, last)> == <reduced g, ...>: perturb the compact quantities,
    expand, and compare the directional derivative with the reduced gradient."""
    b, T = 2, 3
    v, a = head_features(b * T, b, 5, "mix")
    cfg = HeadConfig()
    out, copy = _compact(v, a, T)
    rng = np.random.default_rng(0)
    g_full = torch.tensor(rng.standard_normal((b * T, b * T + 2)))
    g_c = reduce_tube_dlogits(g_full, out, cfg, copy)            # wrt [first, sim / temp (x mask), last]
    d_sim = torch.tensor(rng.standard_normal(out["sim"].shape))
    eps = 1e-6
    out2 = dict(out)
    out2["sim"] = out["sim"] + eps * d_sim
    lg2 = out["logits"].clone()
    pos = out["pos_idx"].long()
    rows = torch.arange(b * T)
    mask = torch.ones_like(out["sim"]); mask[rows, pos] = 1.0 - cfg.mask_scale
    lg2[:, 1:1 + b] = out2["sim"] * mask / cfg.temp
    out2["logits"] = lg2
    d_full = (expand_tube_logits(out2, cfg, copy) - expand_tube_logits(out, cfg, copy)) / eps
    lhs = float((g_full * d_full).sum())
    rhs = float((g_c[:, 1:1 + b] * d_sim / cfg.temp).sum())     # kernels multiply by 1/temp only
    assert abs(lhs - rhs) <= 1e-5 * max(1.0, abs(lhs))
