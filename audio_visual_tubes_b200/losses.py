"""The auxiliary losses of the reference's 16-frame training (losses.py:7-23), same class names
and call signatures.  They act on the head's ``weighted_A`` maps -- [b, T, 14, 14], a few hundred
kilobytes -- so they stay stock torch ops; their autograd hands ``d weighted_A`` to
``avt_head_backward`` (``dWeighted``), where the work on the [N, C, P] features happens.
"""
from __future__ import annotations

import torch


class NPRatio(torch.nn.Module):
    """losses.py:7-14: mean |diff over frames of the spatial sum|."""

    def forward(self, heatmap):
        return torch.abs(torch.diff(torch.sum(heatmap, dim=(2, 3)), dim=1)).mean(dim=1).mean(dim=0)


class PropagationLoss(torch.nn.Module):
    """losses.py:16-23: mean |frame-to-frame difference| of the attention maps [b, T, h, w]."""

    def forward(self, heatmap):
        return torch.abs(torch.diff(heatmap, dim=1)).mean(dim=(2, 3)).mean(dim=1).mean(dim=0)


def hardway16_objective(head_out, head_out_aug, batch_size: int, frame_density: int,
                        loss_weight: float = 0.1):
    """train_hardway.py:134-142 on two head results ``(A, logits, Pos, Neg, weighted_A, loss_rows)``
    (frames and augmented frames, ``hardway_head(..., return_weighted=True)``): returns
    ``(combined, hardway, aug, l2, consistency)``.  The cross-entropy terms use the head's fused
    ``loss_rows`` (== CrossEntropyLoss(logits, 0) per row)."""
    w1, w2 = head_out[4], head_out_aug[4]
    hardway = head_out[5].mean() * loss_weight
    aug = head_out_aug[5].mean() * loss_weight
    l2 = torch.nn.functional.mse_loss(w1, w2) * (100 - loss_weight)
    side = int(round(w1.shape[-1] ** 0.5))
    crit = PropagationLoss()
    cons = crit(w1.reshape(batch_size, frame_density, side, side)) + \
        crit(w2.reshape(batch_size, frame_density, side, side))
    combined = (hardway + aug) / 2 + l2 + cons
    return combined, hardway, aug, l2, cons
