"""Golden fixture of the 16-frame training objective (train_hardway.py:132-142) from the
UNMODIFIED reference: model.AVENet (backbones swapped for nn.Identity(), as make_golden.py does)
and losses.PropagationLoss, on the seeded inputs of tests/golden/inputs.py.

    python tests/golden/make_golden_aux.py          # build container only (needs /root/reference)

Shims: ``Tensor.cuda`` -> identity (model.py:115), and empty stand-ins for the modules
losses.py imports but never uses on this path and that are absent here (``turtle``,
``pytorch_metric_learning``).  Only outputs are stored.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
REF = "/root/reference"


def main():
    import torch
    import torch.nn as nn
    torch.Tensor.cuda = lambda self, *a, **k: self
    for name, attrs in (("turtle", {"pos": None}), ("pytorch_metric_learning", {"losses": None})):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                m = types.ModuleType(name)
                for k, v in attrs.items():
                    setattr(m, k, v)
                sys.modules[name] = m
    sys.path.insert(0, REF)
    import model as ref_model
    import losses as ref_losses
    from inputs import head_features, audio_maps_from_pooled

    b, t, seed, loss_weight = 2, 4, 21, 0.1              # train_hardway.py:65 default loss_weight
    n = b * t
    v1, a = head_features(n, b, seed, "mix")
    v2, _ = head_features(n, b, seed + 1, "mix")
    amap = audio_maps_from_pooled(a, seed)                # [b, 512, h, w] whose max-pool is a
    args = types.SimpleNamespace(epsilon=0.65, epsilon2=0.4, tri_map=True, Neg=True)
    net = ref_model.AVENet(args, False)
    net.imgnet = nn.Identity()
    net.audnet = nn.Identity()
    f1 = torch.tensor(v1.reshape(n, 512, 14, 14), requires_grad=True)
    f2 = torch.tensor(v2.reshape(n, 512, 14, 14), requires_grad=True)
    spec = torch.tensor(amap, requires_grad=True)
    # train_hardway.py:128-129: every spectrogram repeated T times, (b t) flattening
    spec_rep = spec.unsqueeze(1).repeat(1, t, 1, 1, 1).flatten(0, 1)
    criterion, criterion2, criterion3 = nn.CrossEntropyLoss(), ref_losses.PropagationLoss(), nn.MSELoss()
    heatmap, out, weighted, _, _ = net(f1, spec_rep)
    heatmap2, out2, weighted2, _, _ = net(f2, spec_rep)
    target = torch.zeros(out.shape[0]).long()
    hardway_loss = criterion(out, target) * loss_weight
    aug_loss = criterion(out2, target) * loss_weight
    l2_loss = criterion3(weighted, weighted2) * (100 - loss_weight)
    attention = weighted.reshape(b, t, 14, 14)
    attention2 = weighted2.reshape(b, t, 14, 14)
    consistency_loss = torch.add(criterion2(attention), criterion2(attention2))
    combined = torch.add(torch.add(torch.div(torch.add(hardway_loss, aug_loss), 2), l2_loss), consistency_loss)
    combined.backward()
    da = spec.grad.numpy().reshape(b, 512, -1).sum(-1)
    out = {
        "meta": np.array([b, t, seed]), "loss_weight": np.float64(loss_weight),
        "combined": np.float64(combined.item()), "hardway": np.float64(hardway_loss.item()),
        "aug": np.float64(aug_loss.item()), "l2": np.float64(l2_loss.item()),
        "consistency": np.float64(consistency_loss.item()),
        "weighted": weighted.detach().numpy().reshape(n, 196),
        "weighted2": weighted2.detach().numpy().reshape(n, 196),
        "dv1_sub": f1.grad.numpy().reshape(n, 512, 196)[:, ::16, ::2].copy(),
        "dv2_sub": f2.grad.numpy().reshape(n, 512, 196)[:, ::16, ::2].copy(),
        "dv1_norm": np.sqrt((f1.grad.numpy().astype(np.float64) ** 2).sum((1, 2, 3))),
        "dv2_norm": np.sqrt((f2.grad.numpy().astype(np.float64) ** 2).sum((1, 2, 3))),
        "da": da,
    }
    path = os.path.join(HERE, "head_aux_golden.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", {k: float(v) for k, v in out.items() if np.ndim(v) == 0})


if __name__ == "__main__":
    main()
