"""Batch-sharded hard-way loss: one process per GPU, global negatives.

The reference's only multi-GPU mode is ``nn.DataParallel`` (``train_hardway_1frame.py:93``),
whose replicas each build a ``B_local x B_local`` mask (``model.py:114-115``) and therefore
only see intra-replica negatives.  Here every rank owns ``B / R`` visual maps and scores them
against ALL ``B`` audio embeddings -- the single-device reference at the global batch
(SURVEY.md section 8e):

    forward   a_hat_local [B/R, C]  --all_gather-->  a_hat [B, C]          (2 MB at B = 1024)
              rows of this rank: positives sit at columns rank*B/R + n      (col_offset)
              loss = all_reduce(sum of the rank's row losses) / B
    backward  dv is local; d_ahat [B, C] is a per-rank partial
              --reduce_scatter-->  d_ahat_local, then the F.normalize Jacobian

No visual map, similarity or logit ever crosses the link.  The compute callables are
injectable so the collective plumbing is testable with gloo on CPU; the defaults are the
libavtubes entry points (no CPU fallback).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .head import (HeadConfig, as_head_input, audio_normalize, audio_normalize_backward,
                   head_backward_raw, head_forward_raw)


def _lib_compute():
    def normalize(a):
        return audio_normalize(a)

    def forward(v, a_hat, cfg, col_offset, need_grad):
        return head_forward_raw(v, a_hat, cfg, None, col_offset, False, for_backward=need_grad)

    def backward(v, a_hat, cfg, saved, scale, col_offset):
        return head_backward_raw(v, a_hat, cfg, saved, scale, None, None, col_offset)

    def normalize_backward(a_hat, rn, d_ahat):
        return audio_normalize_backward(a_hat, rn, d_ahat)

    return normalize, forward, backward, normalize_backward


class _ShardedHardWayLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v, a, cfg, group, compute):
        normalize, fwd, _, _ = compute
        world = dist.get_world_size(group)
        rank = dist.get_rank(group)
        b_local = a.shape[0]
        if v.shape[0] != b_local:
            raise RuntimeError("sharded head: every rank needs as many visual maps as audio rows")
        need_grad = v.requires_grad or a.requires_grad
        ctx.v_dtype = v.dtype
        v = as_head_input(v, cfg)
        a_hat_local, rn_local = normalize(a.contiguous().float())
        a_hat = torch.empty(world * b_local, a.shape[1], device=a.device, dtype=torch.float32)
        dist.all_gather_into_tensor(a_hat, a_hat_local.contiguous(), group=group)
        col_offset = rank * b_local
        out = fwd(v, a_hat, cfg, col_offset, need_grad)
        total = out["loss_sum"].clone()
        dist.all_reduce(total, group=group)
        ctx.cfg, ctx.group, ctx.compute = cfg, group, compute
        ctx.col_offset, ctx.world, ctx.b_local = col_offset, world, b_local
        ctx.saved = out
        ctx.save_for_backward(v, a_hat, a_hat_local, rn_local)
        return total[0] / float(world * b_local)

    @staticmethod
    def backward(ctx, gloss):
        _, _, bwd, normalize_backward = ctx.compute
        v, a_hat, a_hat_local, rn_local = ctx.saved_tensors
        scale = float(gloss) / float(ctx.world * ctx.b_local)
        dv, d_ahat = bwd(v, a_hat, ctx.cfg, ctx.saved, scale, ctx.col_offset)
        d_local = torch.empty_like(a_hat_local)
        dist.reduce_scatter_tensor(d_local, d_ahat.contiguous(), group=ctx.group)
        da = normalize_backward(a_hat_local, rn_local, d_local)
        return dv.to(ctx.v_dtype), da, None, None, None


def sharded_hardway_loss(v, a, cfg: HeadConfig = HeadConfig(), group=None, compute=None):
    """``CrossEntropyLoss()(logits, zeros)`` over the GLOBAL batch.  ``v [B/R, C, P]`` (or
    ``[B/R, C, H, W]``) and ``a [B/R, C]`` are this rank's clips; the return value is the same
    scalar on every rank and its gradient flows to the local ``v`` and ``a`` only.

    Gradient convention: the returned loss is the global mean, so (unlike DDP's averaged
    gradients) ``dv`` / ``da`` are the exact derivatives of that mean -- the sum over ranks of
    parameter gradients equals the single-device gradient at batch ``B``."""
    if v.dim() == 4:
        v = v.flatten(2)
    if not dist.is_initialized():
        raise RuntimeError("sharded_hardway_loss needs torch.distributed to be initialised")
    return _ShardedHardWayLoss.apply(v, a, cfg, group, compute or _lib_compute())


def sharded_forward(v, a, cfg: HeadConfig = HeadConfig(), group=None):
    """Forward only: this rank's rows of (A, logits [B/R, B+1(+1)], Pos, Neg, loss_rows)."""
    if v.dim() == 4:
        v = v.flatten(2)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    a_hat_local, _ = audio_normalize(a.contiguous().float())
    a_hat = torch.empty(world * a.shape[0], a.shape[1], device=a.device, dtype=torch.float32)
    dist.all_gather_into_tensor(a_hat, a_hat_local, group=group)
    out = head_forward_raw(v.contiguous().float(), a_hat, cfg, None, rank * a.shape[0], False, False)
    return out["A"], out["logits"], out["Pos"], out["Neg"], out["loss_rows"]


def sharded_eval_counts(ciou_local: torch.Tensor, group=None):
    """Evaluation shards trivially (maps are independent): every rank scores its clips, then
    the 21 success-curve counts and the cIoU@0.5 count are summed (integers) and the AUC is
    the trapezoid over the global curve (train_hardway_1frame.py:208-215).
    Returns (ciou_at_half, auc, n_total)."""
    c = ciou_local.double()
    thr = torch.tensor([0.05 * i for i in range(21)], dtype=torch.float64, device=c.device)
    counts = (c[None, :] >= thr[:, None]).sum(1)
    half = (c >= 0.5).sum()[None]
    n = torch.tensor([c.numel()], device=c.device, dtype=torch.int64)
    pack = torch.cat([counts.to(torch.int64), half.to(torch.int64), n])
    dist.all_reduce(pack, group=group)
    n_tot = int(pack[22])
    res = pack[:21].double() / n_tot
    dx = thr[1:] - thr[:-1]
    auc = float((dx * (res[1:] + res[:-1]) / 2.0).sum())
    return float(pack[21]) / n_tot, auc, n_tot


def sharded_value_and_grad(v, a, cfg: HeadConfig = HeadConfig(), group=None, want_grad=True):
    """One sharded training step of the head without autograd bookkeeping:
    ``(global mean loss [1], dv_local, da_local)``; see ``sharded_hardway_loss``."""
    if v.dim() == 4:
        v = v.flatten(2)
    v = as_head_input(v, cfg)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    b_local = a.shape[0]
    a_hat_local, rn_local = audio_normalize(a.contiguous().float())
    a_hat = torch.empty(world * b_local, a.shape[1], device=a.device, dtype=torch.float32)
    dist.all_gather_into_tensor(a_hat, a_hat_local, group=group)
    scale = 1.0 / float(world * b_local)
    out = head_forward_raw(v, a_hat, cfg, None, rank * b_local, False, for_backward=want_grad,
                           fuse_grad_scale=scale if want_grad else 0.0)
    total = out["loss_sum"].clone()
    dist.all_reduce(total, group=group)
    loss = total / float(world * b_local)
    if not want_grad:
        return loss, None, None
    dv, d_ahat = head_backward_raw(v, a_hat, cfg, out, 0.0, None, None, rank * b_local, rowgrad_done=True)
    d_local = torch.empty_like(a_hat_local)
    dist.reduce_scatter_tensor(d_local, d_ahat, group=group)
    da = audio_normalize_backward(a_hat_local, rn_local, d_local)
    return loss, dv, da
