"""CUDA-graph-captured training step of the head (forward + backward), single GPU or row-sharded.

The head's kernels are short (tens to hundreds of microseconds; at 8 GPUs a rank's whole step is
~0.15 ms), so a step launched kernel by kernel from Python is bound by launch overhead and, when
sharded, by the latency of three small NCCL collectives.  ``HeadStep`` pre-allocates every
buffer, captures the whole step once and replays it:

    main stream :  normalise local audio -> visual prep pass (needs only the rank's OWN audio rows)
                   -> [join all_gather] similarity + pooling + CE + row gradients (one row kernel)
                   -> cell gradients -> d_ahat -> reduce_scatter(d_ahat) -> audio normalise Jacobian
    side stream :  all_gather(a_hat), issued first and hidden under the prep pass; later the
                   all_reduce of the loss sum (nothing on the critical path consumes it)
    dv stream   :  the dv contraction, forked after the cell gradients, so the reduce_scatter of
                   d_ahat (which needs only the d_ahat kernel) runs under it; joined at the end

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
                 sharded: bool | None = None, use_graph: bool = True, overlap: bool = True):
        _lib.require_cuda(v, a)
        if v.dim() == 4:
            v = v.flatten(2)
        assert v.is_contiguous() and a.is_contiguous() and a.dtype == torch.float32
        assert v.dtype == torch.float32 or (v.dtype in (torch.float16, torch.bfloat16) and cfg.precision == "f16")
        self.overlap = overlap
        self.v, self.a, self.cfg, self.group = v, a, cfg, group
        self.sharded = dist.is_initialized() and dist.get_world_size(group) > 1 if sharded is None else sharded
        self.world = dist.get_world_size(group) if self.sharded else 1
        self.rank = dist.get_rank(group) if self.sharded else 0
        self.b_local = a.shape[0]
        self.divisor = float(self.world * v.shape[0])
        self.graph = None
        self.side = torch.cuda.Stream(device=v.device)
        self.dv_stream = torch.cuda.Stream(device=v.device)
        self._out = None
        if use_graph:
            self._capture()

    # one eager step; also the body that gets captured
    def _body(self):
        v, a, cfg = self.v, self.a, self.cfg
        main = torch.cuda.current_stream()
        scale = 1.0 / self.divisor
        a_hat_local, rn_local = audio_normalize(a)
        col_offset = self.rank * self.b_local
        if self.sharded and self.overlap:
            # all_gather first, on the side stream; the one pass over the visual maps needs only this
            # rank's own audio rows (its positives) and runs under it
            n_all = self.world * self.b_local
            a_hat = torch.empty(n_all, a.shape[1], device=a.device, dtype=torch.float32)
            self.side.wait_stream(main)
            with torch.cuda.stream(self.side):
                dist.all_gather_into_tensor(a_hat, a_hat_local, group=self.group)
            out = head_forward_raw(v, a_hat_local, cfg, None, 0, False, for_backward=True,
                                   stages=_lib.STAGE_PREP, n_a_total=n_all)
            main.wait_stream(self.side)
            out = head_forward_raw(v, a_hat, cfg, None, col_offset, False, stages=_lib.STAGE_SIM, out=out,
                                   fuse_grad_scale=scale)
        else:
            if self.sharded:
                a_hat = torch.empty(self.world * self.b_local, a.shape[1], device=a.device, dtype=torch.float32)
                dist.all_gather_into_tensor(a_hat, a_hat_local, group=self.group)
            else:
                a_hat = a_hat_local
            out = head_forward_raw(v, a_hat, cfg, None, col_offset, False, for_backward=True,
                                   fuse_grad_scale=scale)
        total = out["loss_sum"]
        if self.sharded:                      # nobody on the critical path needs the loss value
            total = out["loss_sum"].clone()
            self.side.wait_stream(main)
            with torch.cuda.stream(self.side):
                dist.all_reduce(total, group=self.group)
        # only the tensor-core backward has a separate dv contraction to move; forking the stream here keeps
        # it inside a capture even if the library leaves it unused (joining an un-forked stream under
        # capture is cudaErrorStreamCaptureIsolation)
        ext = self.dv_stream if (self.sharded and self.overlap and cfg.precision == "f16") else None
        if ext is not None:
            ext.wait_stream(main)
        dv, d_ahat = head_backward_raw(v, a_hat, cfg, out, 0.0, None, None, col_offset, rowgrad_done=True,
                                       dv_stream=ext)
        if self.sharded:
            d_local = torch.empty_like(a_hat_local)
            dist.reduce_scatter_tensor(d_local, d_ahat, group=self.group)     # under the dv kernel
        else:
            d_local = d_ahat
        da = audio_normalize_backward(a_hat_local, rn_local, d_local)
        if self.sharded:
            main.wait_stream(self.side)
            if ext is not None:
                main.wait_stream(ext)
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


class HostHeadPipeline:
    """The training step of the head fed from HOST buffers, double-buffered: the host->device copy
    of step i+1 (413 MB of fp32 features at B=1024: ~7.5 ms over PCIe) runs on a copy stream while
    the GPU computes step i (~1 ms), so a stream of steps costs max(copy, compute) instead of their
    sum.  Every step still copies its own inputs and reads its own loss back.

        pipe = HostHeadPipeline(v_shape, a_shape, cfg, group)
        h = pipe.submit(v_host, a_host)        # pinned fp32 tensors; returns immediately
        ...                                    # submit the next step before reading this one
        loss = h.loss()                        # float; waits for that step only
        dv, da = h.grads()                     # device tensors of that slot (valid until its reuse)
    """

    class _Handle:
        def __init__(self, slot):
            self._slot = slot

        def loss(self) -> float:
            self._slot["read"].synchronize()
            return float(self._slot["loss_host"][0])

        def grads(self):
            self._slot["done"].synchronize()
            _, dv, da = self._slot["step"]._out[:3]
            return dv, da

    def __init__(self, v_shape, a_shape, cfg: HeadConfig = HeadConfig(), group=None, device=None,
                 use_graph: bool = True, v_dtype: torch.dtype = torch.float32):
        dev = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.slots = []
        for _ in range(2):
            v = torch.zeros(*v_shape, device=dev, dtype=v_dtype)    # fp16 / bf16 features: half the H2D bytes
            a = torch.zeros(*a_shape, device=dev, dtype=torch.float32)
            if v.dim() == 4:
                v = v.flatten(2)
            a.normal_()                                         # warm-up inside HeadStep needs non-zero rows
            v.normal_()
            self.slots.append(dict(
                v=v, a=a, step=HeadStep(v, a, cfg, group=group, use_graph=use_graph),
                copied=torch.cuda.Event(), done=torch.cuda.Event(), read=torch.cuda.Event(),
                loss_host=torch.zeros(1, dtype=torch.float32).pin_memory()))
        self._i = 0

    def submit(self, v_host: torch.Tensor, a_host: torch.Tensor):
        s = self.slots[self._i & 1]
        self._i += 1
        main = torch.cuda.current_stream()
        # the slot's buffers are free once its previous step (and loss read) has been enqueued-and-run
        self.copy_stream.wait_event(s["read"])
        with torch.cuda.stream(self.copy_stream):
            s["v"].copy_(v_host.view_as(s["v"]), non_blocking=True)
            s["a"].copy_(a_host, non_blocking=True)
            s["copied"].record(self.copy_stream)
        main.wait_event(s["copied"])
        loss, _, _ = s["step"]()
        s["done"].record(main)
        s["loss_host"].copy_(loss, non_blocking=True)          # D2H of this step's loss
        s["read"].record(main)
        return HostHeadPipeline._Handle(s)
