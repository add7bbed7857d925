"""World-size-2 gloo test of the row-sharded head (audio_visual_tubes_b200/dist.py) on CPU.

The collective plumbing (all_gather of the audio rows, col_offset, loss all_reduce,
reduce_scatter of d_ahat, normalise Jacobian) is exercised with an injected CPU compute
callable (plain torch, test-only); the result must equal the single-process oracle at the
global batch -- i.e. GLOBAL negatives, not DataParallel's per-replica ones."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from inputs import head_features
from oracle import head_oracle as ho


def _cpu_compute():
    """Stand-in for the libavtubes entry points with the same contracts (test only)."""
    F = torch.nn.functional

    def normalize(a):
        nrm = a.norm(dim=1, keepdim=True).clamp_min(1e-12)
        return a / nrm, (1.0 / nrm).squeeze(1)

    def head(v, a_hat, cfg, col_offset):
        n_v, n_a = v.shape[0], a_hat.shape[0]
        vh = F.normalize(v, dim=1)
        pos = col_offset + torch.arange(n_v)
        S = torch.einsum('ncp,kc->nkp', vh, a_hat)
        A = torch.einsum('ncp,nc->np', vh, a_hat[pos])
        Pos = torch.sigmoid((A - cfg.eps) / cfg.tau)
        Neg = 1 - torch.sigmoid((A - cfg.eps2) / cfg.tau)
        Pall = torch.sigmoid((S - cfg.eps) / cfg.tau)
        sim = (Pall * S).sum(-1) / Pall.sum(-1)
        mask = torch.ones(n_v, n_a, dtype=v.dtype)
        mask[torch.arange(n_v), pos] = 1.0 - cfg.mask_scale
        sim1 = (Pos * A).sum(-1) / Pos.sum(-1)
        sim2 = (Neg * A).sum(-1) / Neg.sum(-1)
        logits = torch.cat([sim1[:, None], sim * mask, sim2[:, None]], 1) / cfg.temp
        return (torch.logsumexp(logits, 1) - logits[:, 0]).sum()

    def forward(v, a_hat, cfg, col_offset, need_grad):
        with torch.no_grad():
            return {"loss_sum": head(v, a_hat, cfg, col_offset)[None]}

    def backward(v, a_hat, cfg, saved, scale, col_offset):
        with torch.enable_grad():
            v2 = v.detach().clone().requires_grad_(True)
            a2 = a_hat.detach().clone().requires_grad_(True)
            (scale * head(v2, a2, cfg, col_offset)).backward()
        return v2.grad, a2.grad

    def normalize_backward(a_hat, rn, d):
        return rn[:, None] * (d - a_hat * (a_hat * d).sum(1, keepdim=True))

    return normalize, forward, backward, normalize_backward


def _worker(rank, world, port, b_local, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from audio_visual_tubes_b200 import HeadConfig
        from audio_visual_tubes_b200.dist import sharded_eval_counts, sharded_hardway_loss
        b = b_local * world
        v, a = head_features(b, b, 77, "mix", c=64, p=36)
        sl = slice(rank * b_local, (rank + 1) * b_local)
        vt = torch.tensor(v[sl], dtype=torch.float32, requires_grad=True)
        at = torch.tensor(a[sl], dtype=torch.float32, requires_grad=True)
        loss = sharded_hardway_loss(vt, at, HeadConfig(), compute=_cpu_compute())
        loss.backward()
        ciou = torch.tensor(np.linspace(0.0, 1.0, 11)[rank::world])
        half, auc, n_tot = sharded_eval_counts(ciou)
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), loss=float(loss.detach()), dv=vt.grad.numpy(),
                 da=at.grad.numpy(), half=half, auc=auc, n_tot=n_tot)
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.timeout(300)
def test_sharded_loss_equals_single_process_oracle(tmp_path):
    world, b_local = 2, 3
    mp.spawn(_worker, args=(world, _free_port(), b_local, str(tmp_path)), nprocs=world, join=True)
    b = world * b_local
    v, a = head_features(b, b, 77, "mix", c=64, p=36)
    o, dv, da = ho.head_forward_backward_torch(v, a)          # single process, global negatives
    for r in range(world):
        z = np.load(os.path.join(str(tmp_path), f"r{r}.npz"))
        sl = slice(r * b_local, (r + 1) * b_local)
        assert abs(float(z["loss"]) - o["loss"]) < 1e-5 * max(1.0, abs(o["loss"]))   # fp32 plumbing
        assert np.abs(z["dv"] - dv[sl]).max() <= 2e-4 * np.abs(dv).max()
        assert np.abs(z["da"] - da[sl]).max() <= 2e-4 * np.abs(da).max()
        # eval reduction: 11 cIoUs 0, .1, ..., 1 spread over the ranks
        from oracle import eval_oracle as eo
        c = np.linspace(0.0, 1.0, 11)
        _, res = eo.success_curve(c)
        assert int(z["n_tot"]) == 11
        assert abs(float(z["half"]) - eo.ciou_at_half(c)) < 1e-12
        assert abs(float(z["auc"]) - eo.auc_trapezoid(res)) < 1e-12


def test_per_replica_negatives_differ_from_global():
    """Documents the reference's DataParallel behaviour (model.py:114-115 inside each replica):
    its loss is NOT the global-batch loss -- the sharded head reproduces the latter."""
    v, a = head_features(6, 6, 77, "mix", c=64, p=36)
    g = ho.head_forward(v, a)["loss"]
    per = np.mean([ho.head_forward(v[s], a[s])["loss"] for s in (slice(0, 3), slice(3, 6))])
    assert abs(g - per) > 1e-3
