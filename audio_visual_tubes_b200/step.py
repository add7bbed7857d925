"""CUDA-graph-captured training step of the head (forward + backward), single GPU or row-sharded.

The head's kernels are short (tens to hundreds of microseconds; at 8 GPUs a rank's whole step is
~0.15 ms), so a step launched kernel by kernel from Python is bound by launch overhead and, when
sharded, by the latency of three small NCCL collectives.  ``HeadStep`` pre-allocates every
buffer, captures the whole step once and replays it:

    main stream :  normalise local audio -> all_gather(a_hat) -> prep + similarity + pooling + CE
                   -> backward (cell gradients, dv, d_ahat) -> reduce_scatter(d_ahat)
                   -> audio normalise Jacobian
    side stream :  all_reduce of the loss sum (nothing on the critical path consumes it), forked
                   after the forward and joined at the end, i.e. hidden under the backward

Inputs are read from the tensors given at construction (fixed addresses): copy new features into
``step.v`` / ``step.a`` (or hand in the backbone's output buffers) and call ``step()``.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import _lib
from .head import (HeadConfig, audio_normalize, audio_normalize_backward, head_backward_raw,
                   head_forward_raw)


class HeadStep:
    """``step() -> (loss [1], dv [N_v,C,P], da [N_a_local,C])`` for the fixed input tensors ``v``, ``a``.

    ``group`` (optional): a torch.distributed NCCL group; every rank owns ``v.shape[0]`` visual
    maps and as many audio rows, negatives are global (see ``dist.py``).  The loss is the global
    mean of ``CrossEntropyLoss(logits, 0)`` (train_hardway_1frame.py:130-131) and the gradients
    are its exact derivatives."""

    def __init__(self, v: torch.Tensor, a: torch.Tensor, cfg: HeadConfig = HeadConfig(), group=None,
                 sharded: bool | None = None, use_graph: bool = True):
        _lib.require_cuda(v, a)
        if v.dim() == 4:
            v = v.flatten(2)
        assert v.is_contiguous() and a.is_contiguous() and v.dtype == torch.float32 and a.dtype == torch.float32
        self.v, self.a, self.cfg, self.group = v, a, cfg, group
        self.sharded = dist.is_initialized() and dist.get_world_size(group) > 1 if sharded is None else sharded
        self.world = dist.get_world_size(group) if self.sharded else 1
        self.rank = dist.get_rank(group) if self.sharded else 0
        self.b_local = a.shape[0]
        self.divisor = float(self.world * v.shape[0])
        self.graph = None
        self.side = torch.cuda.Stream(device=v.device)
        self._out = None
        if use_graph:
            self._capture()

    # one eager step; also the body that gets captured
    def _body(self):
        v, a, cfg = self.v, self.a, self.cfg
        a_hat_local, rn_local = audio_normalize(a)
        if self.sharded:
            a_hat = torch.empty(self.world * self.b_local, a.shape[1], device=a.device, dtype=torch.float32)
            dist.all_gather_into_tensor(a_hat, a_hat_local, group=self.group)
        else:
            a_hat = a_hat_local
        col_offset = self.rank * self.b_local
        out = head_forward_raw(v, a_hat, cfg, None, col_offset, False, for_backward=True)
        main = torch.cuda.current_stream()
        total = out["loss_sum"]
        if self.sharded:                      # nobody on the critical path needs the loss value
            total = out["loss_sum"].clone()
            self.side.wait_stream(main)
            with torch.cuda.stream(self.side):
                dist.all_reduce(total, group=self.group)
        dv, d_ahat = head_backward_raw(v, a_hat, cfg, out, 1.0 / self.divisor, None, None, col_offset)
        if self.sharded:
            d_local = torch.empty_like(a_hat_local)
            dist.reduce_scatter_tensor(d_local, d_ahat, group=self.group)
            main.wait_stream(self.side)
        else:
            d_local = d_ahat
        da = audio_normalize_backward(a_hat_local, rn_local, d_local)
        loss = total / self.divisor
        return loss, dv, da, out

    def _capture(self):
        dev = self.v.device
        s = torch.cuda.Stream(device=dev)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):            # warm-up outside capture (NCCL, lazy attributes)
            for _ in range(2):
                self._body()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._out = self._body()
        self.graph = g

    def __call__(self):
        if self.graph is None:
            self._out = self._body()
        else:
            self.graph.replay()
        loss, dv, da, _ = self._out
        return loss, dv, da

    @property
    def outputs(self):
        """dict of the forward's output tensors of the last step (A, logits, Pos, Neg, ...)."""
        return self._out[3]
