"""Host side of the similarity + hard-way loss head.

Replaces the body of ``AVENet.forward`` after the backbones (reference ``model.py:114-154``),
``HardWayAttention.forward`` (``model.py:46-60``) and the ``nn.CrossEntropyLoss`` on the
logits (``train_hardway_1frame.py:113,130-131``) with calls into libavtubes (sm_100a CUDA).
The ``N_v x N_a x 14 x 14`` similarity tensor of ``model.py:125`` is never materialised.

Layouts: ``v [N_v, C, P]`` raw visual maps (NCHW with H*W flattened), ``a [N_a, C]`` pooled
audio embeddings, both fp32 CUDA tensors.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib

PRECISION = {"fp32": _lib.PRECISION_FP32, "f16": _lib.PRECISION_F16}
V_DTYPE = {torch.float32: _lib.DTYPE_F32, torch.float16: _lib.DTYPE_F16, torch.bfloat16: _lib.DTYPE_BF16}


def as_head_input(v: torch.Tensor, cfg) -> torch.Tensor:
    """The visual maps as the kernels take them: contiguous, fp32 -- or fp16 / bf16 left as they are
    on the tensor-core path (a backbone under autocast emits them; the prep pass reads half the
    bytes and a host->device copy moves half of them)."""
    if v.dtype in (torch.float16, torch.bfloat16) and cfg.precision == "f16":
        return v.contiguous()
    return v.contiguous().float()


@dataclass(frozen=True)
class HeadConfig:
    """Scalars of the head.  Reference defaults: ``--epsilon 0.65 --epsilon2 0.4``
    (``train_hardway_1frame.py:55-56``), ``tau = 0.03`` (``model.py:100``), logit temperature
    0.07 (``model.py:144``), diagonal mask ``1 - 100*eye`` (``model.py:115``)."""
    eps: float = 0.65
    eps2: float = 0.4
    tau: float = 0.03
    temp: float = 0.07
    mask_scale: float = 100.0
    trimap: bool = True
    use_neg: bool = True
    precision: str = "f16"


def audio_normalize(a: torch.Tensor):
    """``F.normalize(aud, dim=1)`` (``model.py:122``): returns (a_hat, rnorm_a)."""
    _lib.require_cuda(a)
    a = a.contiguous().float()
    n_a, c = a.shape
    a_hat = torch.empty_like(a)
    rn = torch.empty(n_a, device=a.device, dtype=torch.float32)
    lib = _lib.load()
    _lib.check(lib.avt_audio_normalize(_lib.ptr(a), n_a, c, _lib.ptr(a_hat), _lib.ptr(rn),
                                       _lib.stream_ptr()), "avt_audio_normalize")
    return a_hat, rn


def audio_normalize_backward(a_hat, rnorm_a, d_ahat):
    lib = _lib.load()
    n_a, c = a_hat.shape
    da = torch.empty_like(a_hat)
    _lib.check(lib.avt_audio_normalize_backward(_lib.ptr(a_hat), _lib.ptr(rnorm_a),
                                                _lib.ptr(d_ahat.contiguous()), n_a, c, _lib.ptr(da),
                                                _lib.stream_ptr()), "avt_audio_normalize_backward")
    return da


def head_forward_raw(v, a_hat, cfg: HeadConfig, pos_idx=None, col_offset=0, want_weighted=False,
                     for_backward=False, col_mult=1, fuse_grad_scale=0.0, stages=0, out=None, n_a_total=None):
    """One ``avt_head_forward`` call.  ``a_hat`` must already be L2-normalised rows.
    Returns a dict of output tensors (+ the workspace to hand to the backward).
    ``col_mult = T > 1``: tube mode, ``a_hat`` holds the DISTINCT audio rows, each standing for T
    identical logit columns (see include/avtubes.h); ``logits`` is then the compact row.
    ``fuse_grad_scale != 0``: fused training step, the row gradients of ``fuse_grad_scale * sum(loss_rows)``
    are produced by the forward; call ``head_backward_raw(..., rowgrad_done=True)``.
    ``stages`` / ``out`` / ``n_a_total``: split forward for row-sharded callers -- run
    ``stages=STAGE_PREP`` with the rank's own audio rows (``n_a_total`` = the global count, which
    sizes the outputs and the workspace), then ``stages=STAGE_SIM`` with the gathered rows and
    ``out=`` the dict the first call returned."""
    _lib.require_cuda(v, a_hat)
    lib = _lib.load()
    assert v.dtype in V_DTYPE and a_hat.dtype == torch.float32
    assert v.is_contiguous() and a_hat.is_contiguous()
    n_v, c, p = v.shape
    n_a = a_hat.shape[0]
    n_all = int(n_a_total or n_a)                 # sizes outputs / workspace (split forward: global count)
    dev = v.device
    prec = PRECISION[cfg.precision]
    if out is None:
        L = n_all + 1 + (1 if cfg.use_neg else 0)
        f32 = dict(device=dev, dtype=torch.float32)
        out = dict(
            A=torch.empty(n_v, p, **f32), Pos=torch.empty(n_v, p, **f32), Neg=torch.empty(n_v, p, **f32),
            logits=torch.empty(n_v, L, **f32), loss_rows=torch.empty(n_v, **f32),
            loss_sum=torch.empty(1, **f32), rnorm_v=torch.empty(n_v, p, **f32),
            sim=torch.empty(n_v, n_all, **f32), den=torch.empty(n_v, n_all, **f32),
            weighted_A=torch.empty(n_v, p, **f32) if want_weighted else None)
        nbytes = lib.avt_head_workspace_bytes(n_v, n_all, c, p, prec,
                                              1 if (for_backward or fuse_grad_scale) else 0)
        out["workspace"] = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    ws = out["workspace"]
    nbytes = ws.numel()
    if pos_idx is not None:
        pos_idx = pos_idx.to(device=dev, dtype=torch.int32).contiguous()
    with torch.cuda.device(dev):
        rc = lib.avt_head_forward(
            _lib.ptr(v), V_DTYPE[v.dtype], _lib.ptr(a_hat), n_v, n_a, c, p, _lib.ptr(pos_idx), int(col_offset),
            cfg.eps, cfg.eps2, cfg.tau, cfg.temp, cfg.mask_scale, int(cfg.trimap), int(cfg.use_neg),
            prec, int(col_mult), _lib.ptr(out["A"]), _lib.ptr(out["Pos"]), _lib.ptr(out["Neg"]),
            _lib.ptr(out["logits"]), _lib.ptr(out["loss_rows"]), _lib.ptr(out["loss_sum"]),
            _lib.ptr(out["weighted_A"]), _lib.ptr(out["rnorm_v"]), _lib.ptr(out["sim"]),
            _lib.ptr(out["den"]), int(stages), float(fuse_grad_scale), _lib.ptr(ws), C.c_size_t(nbytes),
            _lib.stream_ptr())
    _lib.check(rc, "avt_head_forward")
    out["pos_idx"] = pos_idx
    out["col_mult"] = int(col_mult)
    return out


def head_backward_raw(v, a_hat, cfg: HeadConfig, saved, dloss_scale=0.0, dlogits=None, dA=None,
                      col_offset=0, dW=None, dloss_rows=None, rowgrad_done=False, dv_stream=None):
    """One ``avt_head_backward`` call.  Returns (dv [N_v,C,P], d_ahat [N_a,C]).
    ``rowgrad_done``: the forward ran with ``fuse_grad_scale``.  ``dv_stream`` (torch.cuda.Stream):
    the dv contraction runs there and ``d_ahat`` is complete on the current stream without waiting
    for it; the caller orders ``dv_stream`` before consumers of ``dv``."""
    lib = _lib.load()
    n_v, c, p = v.shape
    n_a = a_hat.shape[0]
    prec = PRECISION[cfg.precision]
    ws = saved["workspace"]
    need = lib.avt_head_workspace_bytes(n_v, n_a, c, p, prec, 1)
    if ws.numel() < need:
        raise RuntimeError("head_backward_raw: the forward was not run with for_backward=True")
    dv = torch.empty(v.shape, device=v.device, dtype=torch.float32)      # always fp32, whatever v is
    d_ahat = torch.empty_like(a_hat)
    if dlogits is not None:
        dlogits = dlogits.contiguous().float()
    if dA is not None:
        dA = dA.contiguous().float().view(n_v, p)
    if dW is not None:
        dW = dW.contiguous().float().view(n_v, p)
    if dloss_rows is not None:
        dloss_rows = dloss_rows.contiguous().float().view(n_v)
    with torch.cuda.device(v.device):
        rc = lib.avt_head_backward(
            _lib.ptr(v), V_DTYPE[v.dtype], _lib.ptr(a_hat), n_v, n_a, c, p, _lib.ptr(saved["pos_idx"]), int(col_offset),
            cfg.eps, cfg.eps2, cfg.tau, cfg.temp, cfg.mask_scale, int(cfg.trimap), int(cfg.use_neg),
            prec, int(saved.get("col_mult", 1)), _lib.ptr(saved["A"]), _lib.ptr(saved["logits"]),
            _lib.ptr(saved["loss_rows"]),
            _lib.ptr(saved["rnorm_v"]), _lib.ptr(saved["sim"]), _lib.ptr(saved["den"]),
            float(dloss_scale), _lib.ptr(dloss_rows), _lib.ptr(dlogits), _lib.ptr(dA), _lib.ptr(dW), _lib.ptr(dv),
            _lib.ptr(d_ahat), _lib.BWD_ROWGRAD_DONE if rowgrad_done else 0,
            None if dv_stream is None else C.c_void_p(dv_stream.cuda_stream),
            _lib.ptr(ws), C.c_size_t(ws.numel()), _lib.stream_ptr())
    _lib.check(rc, "avt_head_backward")
    return dv, d_ahat


def tube_positions(n_v: int, col_mult: int, device, row_offset: int = 0):
    """(audio row, copy) of every frame under the reference's ``(b t)`` flattening
    (model.py:49) with each spectrogram repeated T times (train_3D.py:129-130): frame n belongs
    to clip n // T and is masked against copy n % T of that clip's audio."""
    rows = torch.arange(row_offset, row_offset + n_v, device=device)
    return (rows // col_mult).int(), (rows % col_mult).long()


def expand_tube_logits(out, cfg: HeadConfig, copy_idx):
    """Compact logits of tube mode -> the reference's ``[N_v, N_a*T + 1(+1)]`` row
    (model.py:56-59): every audio column T times, the frame's own copy carrying the mask."""
    T = out["col_mult"]
    lg, sim, pos = out["logits"], out["sim"], out["pos_idx"].long()
    n_v, n_a = sim.shape
    mid = (sim / cfg.temp).repeat_interleave(T, dim=1)
    rows = torch.arange(n_v, device=lg.device)
    ok = (pos >= 0) & (pos < n_a)
    mid[rows[ok], pos[ok] * T + copy_idx[ok]] = lg[rows[ok], 1 + pos[ok]]
    parts = [lg[:, :1], mid] + ([lg[:, -1:]] if cfg.use_neg else [])
    return torch.cat(parts, dim=1)


def reduce_tube_dlogits(g_full, out, cfg: HeadConfig, copy_idx):
    """Adjoint of ``expand_tube_logits`` for the kernels: per audio row the sum over its T copies
    of d/dlogit * d logit/d sim-scaled, i.e. with the mask factor of the masked copy applied."""
    T = out["col_mult"]
    n_v, n_a = out["sim"].shape
    pos = out["pos_idx"].long()
    mid = g_full[:, 1:1 + n_a * T].reshape(n_v, n_a, T).clone()
    rows = torch.arange(n_v, device=g_full.device)
    ok = (pos >= 0) & (pos < n_a)
    mid[rows[ok], pos[ok], copy_idx[ok]] *= (1.0 - cfg.mask_scale)
    parts = [g_full[:, :1], mid.sum(dim=2)] + ([g_full[:, -1:]] if cfg.use_neg else [])
    return torch.cat(parts, dim=1).contiguous()


class _HardWayHead(torch.autograd.Function):
    """autograd wrapper: (v, a) -> (A, logits, Pos, Neg, weighted_A, loss_rows).

    Gradients flow through ``logits``, ``A``, ``weighted_A`` (train_hardway.py:138-142 sends an
    MSE and the PropagationLoss back through it) and ``loss_rows`` (the fused
    CrossEntropyLoss(target 0) per row).  ``Pos`` / ``Neg`` are returned for inspection exactly
    like the reference returns them and are non-differentiable here (no reference training
    script back-propagates through them)."""

    @staticmethod
    def forward(ctx, v, a, cfg, pos_idx, col_offset, want_weighted, col_mult=1, copy_idx=None):
        need_grad = v.requires_grad or a.requires_grad
        ctx.set_materialize_grads(False)      # unused outputs arrive as None, not as zero tensors
        ctx.v_dtype = v.dtype
        v = as_head_input(v, cfg)
        a_hat, rn_a = audio_normalize(a)
        out = head_forward_raw(v, a_hat, cfg, pos_idx, col_offset, want_weighted, for_backward=need_grad,
                               col_mult=col_mult)
        ctx.cfg, ctx.col_offset = cfg, col_offset
        ctx.saved = out
        ctx.copy_idx = copy_idx
        ctx.save_for_backward(v, a_hat, rn_a)
        w = out["weighted_A"] if want_weighted else out["A"].new_empty(0)
        ctx.want_weighted = want_weighted
        if want_weighted:
            ctx.mark_non_differentiable(out["Pos"], out["Neg"])
        else:
            ctx.mark_non_differentiable(out["Pos"], out["Neg"], w)
        logits = out["logits"] if col_mult == 1 else expand_tube_logits(out, cfg, copy_idx)
        return out["A"], logits, out["Pos"], out["Neg"], w, out["loss_rows"]

    @staticmethod
    def backward(ctx, gA, glogits, gPos, gNeg, gW, gloss_rows):
        v, a_hat, rn_a = ctx.saved_tensors
        dlogits = glogits
        # d loss_rows[n] / d logits[n, :] = softmax - onehot(0) is formed inside row_grad_kernel from
        # the per-row upstream gradient (dloss_rows): no eager softmax on the host side
        if dlogits is not None and ctx.saved["col_mult"] > 1:
            dlogits = reduce_tube_dlogits(dlogits.contiguous().float(), ctx.saved, ctx.cfg, ctx.copy_idx)
        dv, d_ahat = head_backward_raw(v, a_hat, ctx.cfg, ctx.saved, 0.0, dlogits, gA,
                                       ctx.col_offset, gW if ctx.want_weighted else None,
                                       dloss_rows=gloss_rows)
        da = audio_normalize_backward(a_hat, rn_a, d_ahat)
        return dv.to(ctx.v_dtype), da, None, None, None, None, None, None


class _HardWayLoss(torch.autograd.Function):
    """(v, a) -> mean_n CrossEntropy(logits[n], 0), fully fused: the only upstream gradient
    is one scalar, so the backward needs no dlogits tensor at all."""

    @staticmethod
    def forward(ctx, v, a, cfg, pos_idx, col_offset, divisor, col_mult=1):
        need_grad = v.requires_grad or a.requires_grad
        ctx.v_dtype = v.dtype
        v = as_head_input(v, cfg)
        a_hat, rn_a = audio_normalize(a)
        out = head_forward_raw(v, a_hat, cfg, pos_idx, col_offset, False, for_backward=need_grad,
                               col_mult=col_mult)
        ctx.cfg, ctx.col_offset, ctx.divisor = cfg, col_offset, divisor
        ctx.saved = out
        ctx.save_for_backward(v, a_hat, rn_a)
        return out["loss_sum"][0] / divisor

    @staticmethod
    def backward(ctx, gloss):
        v, a_hat, rn_a = ctx.saved_tensors
        # one host read of the scalar upstream gradient (1.0 for loss.backward())
        scale = float(gloss) / ctx.divisor
        dv, d_ahat = head_backward_raw(v, a_hat, ctx.cfg, ctx.saved, scale, None, None, ctx.col_offset)
        da = audio_normalize_backward(a_hat, rn_a, d_ahat)
        return dv.to(ctx.v_dtype), da, None, None, None, None, None


def hardway_head(v, a, cfg: HeadConfig = HeadConfig(), pos_idx=None, col_offset=0,
                 return_weighted=False, col_mult=1):
    """v [N_v,C,P] (or [N_v,C,H,W]), a [N_a,C] -> (A [N_v,P], logits [N_v,N_a+1(+1)], Pos, Neg,
    weighted_A or None, loss_rows [N_v]).

    ``col_mult = T > 1`` (tube mode): ``a`` holds the N_v / T DISTINCT audio rows of clips whose T
    frames are consecutive rows of ``v``; the result is what the reference computes after
    repeating every spectrogram T times (train_3D.py:129-130) -- logits ``[N_v, N_a*T + 1(+1)]``
    included -- at 1/T of the similarity work, and ``a.grad`` is the gradient wrt the distinct
    rows (the sum over the copies)."""
    if v.dim() == 4:
        v = v.flatten(2)
    copy_idx = None
    if col_mult > 1:
        auto_pos, copy_idx = tube_positions(v.shape[0], col_mult, v.device, col_offset)
        pos_idx = auto_pos if pos_idx is None else pos_idx
    A, logits, Pos, Neg, w, loss_rows = _HardWayHead.apply(v, a, cfg, pos_idx, col_offset,
                                                           return_weighted, col_mult, copy_idx)
    return A, logits, Pos, Neg, (w if return_weighted else None), loss_rows


def hardway_loss(v, a, cfg: HeadConfig = HeadConfig(), pos_idx=None, col_offset=0, divisor=None,
                 col_mult=1):
    """Fused ``CrossEntropyLoss()(logits, zeros)`` of the head (``train_hardway_1frame.py:130-131``):
    mean over the N_v rows (or sum / divisor for row-sharded ranks).  ``col_mult``: tube mode, see
    ``hardway_head``."""
    if v.dim() == 4:
        v = v.flatten(2)
    if col_mult > 1 and pos_idx is None:
        pos_idx, _ = tube_positions(v.shape[0], col_mult, v.device, col_offset)
    return _HardWayLoss.apply(v, a, cfg, pos_idx, col_offset, float(divisor or v.shape[0]), col_mult)


def hardway_value_and_grad(v, a, cfg: HeadConfig = HeadConfig(), pos_idx=None, col_offset=0,
                           divisor=None, want_grad=True, col_mult=1):
    """One training step of the head without autograd bookkeeping: returns
    ``(loss [1] device tensor, dv [N_v,C,P], da [N_a,C])`` for
    ``loss = sum_n CE(logits[n], 0) / divisor`` (divisor defaults to N_v: the reference's
    ``CrossEntropyLoss`` mean, train_hardway_1frame.py:130-133).  Everything is enqueued on the
    current stream; nothing synchronises.  ``col_mult``: tube mode, see ``hardway_head``."""
    _lib.require_cuda(v, a)
    if v.dim() == 4:
        v = v.flatten(2)
    v = as_head_input(v, cfg)
    divisor = float(divisor or v.shape[0])
    if col_mult > 1 and pos_idx is None:
        pos_idx, _ = tube_positions(v.shape[0], col_mult, v.device, col_offset)
    a_hat, rn_a = audio_normalize(a)
    out = head_forward_raw(v, a_hat, cfg, pos_idx, col_offset, False, for_backward=want_grad,
                           col_mult=col_mult, fuse_grad_scale=(1.0 / divisor) if want_grad else 0.0)
    loss = out["loss_sum"] / divisor
    if not want_grad:
        return loss, None, None
    dv, d_ahat = head_backward_raw(v, a_hat, cfg, out, 0.0, None, None, col_offset, rowgrad_done=True)
    da = audio_normalize_backward(a_hat, rn_a, d_ahat)
    return loss, dv, da


def hardway_value_and_grad_host(v_host, a_host, cfg: HeadConfig = HeadConfig(), device=None,
                                want_grad=True):
    """The same step for HOST buffers (pinned memory recommended): copies the features to
    the device, runs the step, and returns ``(loss as a Python float, dv, da)`` with the
    gradients left on the device for the backbone's backward.  This is the end-to-end call
    bench.py times (H2D of v and a, D2H of the loss)."""
    if not torch.cuda.is_available():
        raise RuntimeError("libavtubes runs on the GPU only (there is no CPU fallback)")
    dev = torch.device(device or "cuda")
    v = v_host.to(dev, non_blocking=True)
    a = a_host.to(dev, non_blocking=True)
    loss, dv, da = hardway_value_and_grad(v, a, cfg, want_grad=want_grad)
    return float(loss.cpu()[0]), dv, da
