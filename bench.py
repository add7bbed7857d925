#!/usr/bin/env python
"""bench.py -- the head's training step (similarity + hard-way loss, forward + backward) on
N B200s, plus the evaluation kernel, the end-to-end (host buffers) leg, the roofline of the
dominant kernel and the reference's CPU path timed on the same box.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg5|cfg2] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A *step* is one pass of the hot path over one synthetic batch: L2-normalise, all-pairs
similarity, tri-map pooling, logits, cross-entropy and the full backward (dv, da) -- the
head's share of ``train_hardway_1frame.py:129-133``.  ``value`` is similarity cells per second
(cells = N_v * N_a * 196, ``model.py:125``) summed over all ranks, inputs resident in HBM.

Workloads:
  cfg5  B = 1024 single-frame (== config 4's 64 clips x 16 frames cell count), rows sharded
        1024/N per rank, audio embeddings all-gathered with NCCL; strong scaling.  Default:
        it is the configuration the metric's tensor-core target is quoted on and fits one GPU.
  cfg2  B = 128 per rank (weak scaling), the reference's 1-GPU training batch.

One JSON line is printed by rank 0 (see the task contract for the keys).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

METRIC = "av_sim_hardway_loss_cells_per_s"
UNIT = "cells/s"
C, P = 512, 196
FLOP_FWD_PER_CELL = 1024.0          # 2*C   
FLOP_BWD_PER_CELL = 2048.0          # the two gradient contractions; recompute earns no credit
EVAL_BYTES_PER_MAP = 784 + 50176 + 200704 + 50176 + 24     # materialising mode


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg5", choices=["cfg5", "cfg2"])
    ap.add_argument("--precision", default="f16", choices=["f16", "fp32"])
    ap.add_argument("--no-extras", action="store_true", help="skip eval / cfg2 / train-step extras")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch the step kernel by kernel instead of replaying a CUDA graph")
    ap.add_argument("--no-overlap", action="store_true", help="sharded step without the all-gather / reduce-scatter overlap (A/B)")
    return ap.parse_args()


def workload_shape(name, world):
    if name == "cfg5":
        b_global = 1024
        assert b_global % world == 0
        return b_global // world, b_global, "strong"
    return 128, 128 * world, "weak"


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=d["hbm_gbs"], tc_burst=d["bf16_tflops"], tc_sustained=d["bf16_tflops_sustained"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tc_burst=1590.0, tc_sustained=1400.0, source="fallback (B200_PROFILING.md)")


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons while the timed region runs (NVML)."""

    def __init__(self, index=0, period=0.05):
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(self.period)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ----------------------------------------------------------------------------- CPU baseline
def workload_string(name, b_local, b_global):
    return (f"{name}: head fwd+bwd, global B={b_global} ({b_local} visual rows/rank x {b_global} audio), "
            f"C=512, P=196, features |randn| fp32 NCHW")


def cpu_head_baseline(b_sample, steps, warmup, seed=1234):
    """The reference's own head on the host cores: the UNMODIFIED ``AVENet.forward`` (model.py:112-154,
    from oracle/_ref, backbones swapped for nn.Identity so that features are the inputs) +
    ``nn.CrossEntropyLoss`` + ``backward()`` exactly as train_hardway_1frame.py:129-133 chains them,
    eager fp32 torch on all host cores.  The reference can only score SQUARE batches (its mask is
    B x B, model.py:115) and materialises ~10 A0-sized fp32 temporaries (8 GB at B=1024), so the
    timed sample is a square batch of ``b_sample`` clips; cells/s is what is compared.  Falls back
    to the oracle port (same torch ops, restated) when oracle/_ref did not travel."""
    import numpy as np
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    g = torch.Generator().manual_seed(seed)
    v = torch.randn(b_sample, C, 14, 14, generator=g).abs_()
    a = torch.randn(b_sample, C, 1, 1, generator=g).abs_()
    kind = "port"
    try:
        from oracle import make_ref
        if make_ref.available():
            net = make_ref.reference_head()
            crit = torch.nn.CrossEntropyLoss()
            tgt = torch.zeros(b_sample).long()

            def one():
                vt = v.clone().requires_grad_(True)
                at = a.clone().requires_grad_(True)
                out = net(vt, at)
                loss = crit(out[1], tgt)
                loss.backward()
                return float(loss.detach())
            kind = "reference"
    except Exception as e:                                   # noqa: BLE001
        sys.stderr.write(f"[bench] oracle/_ref unusable ({e!r}); timing the oracle port instead\n")
    if kind == "port":
        from oracle import head_oracle as ho

        def one():
            return ho.reference_style_forward_torch(v.flatten(2), a.view(b_sample, C), backward=True)
    from oracle import make_ref as _mr
    with _mr.cpu_shim():                                     # host tensors even where a GPU is visible
        for _ in range(warmup):
            one()
        ts = []
        for _ in range(steps):
            t0 = time.perf_counter()
            one()
            ts.append(time.perf_counter() - t0)
    t = float(np.median(ts))
    cells = b_sample * b_sample * P
    what = ("the UNMODIFIED reference AVENet.forward (oracle/_ref, Identity backbones) + CrossEntropyLoss + backward"
            if kind == "reference" else "eager fp32 torch port of model.py:117-146 (oracle/head_oracle.py) + CE + backward")
    return dict(value=cells / t, unit=UNIT, cores=cores, kind=kind,
                sample=f"square batch of {b_sample} clips ({cells} cells) per step, fwd+bwd, median of {steps} "
                       f"after {warmup} warm-up; {what}; eager fp32 on {cores} host threads",
                seconds_per_sample=t), t


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on all host cores,
    same metric / unit / config string / steps / warm-up as the GPU arm; each step is a bounded
    sample (a square batch of 256 clips) of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = args.gpus
    b_local, b_global, scaling = workload_shape(args.workload, world)
    b_sample = min(256, b_global)
    base, t = cpu_head_baseline(b_sample, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT,
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3,
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_string(args.workload, b_local, b_global),
                   "sample": f"each step = a square batch of {b_sample} clips on the host cores "
                             f"(the reference's B x B mask admits no row block; full B=1024 needs ~8 GB of temporaries)"},
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ----------------------------------------------------------------------------- GPU arm
def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    import audio_visual_tubes_b200 as avt
    from audio_visual_tubes_b200 import _lib, evaluation as ev
    from audio_visual_tubes_b200.dist import sharded_value_and_grad
    lib = _lib.load()
    assert lib.avt_device_ok() == 1, "libavtubes needs an sm_100 device"
    pk = peaks()
    cfg = avt.HeadConfig(precision=args.precision)

    # ONE seeded GLOBAL batch, sliced by rank: the loss at N ranks is the loss at 1 (checked below)
    b_local, b_global, scaling = workload_shape(args.workload, world)
    g = torch.Generator(device=dev).manual_seed(1234)
    v_glob = torch.randn(b_global, C, P, device=dev, generator=g).abs_()      # post-ReLU-like maps
    a_glob = torch.randn(b_global, C, device=dev, generator=g).abs_()
    sl = slice(rank * b_local, (rank + 1) * b_local)
    loss_global_1gpu = float(avt.hardway_value_and_grad(v_glob, a_glob, cfg, want_grad=False)[0][0])
    v = v_glob[sl].clone()
    a = a_glob[sl].clone()
    del v_glob, a_glob
    torch.cuda.empty_cache()
    cells_per_step = b_global * b_global * P
    in_bytes = v.numel() * 4
    flush = None
    if in_bytes * 3 < 200e6:          # inputs + outputs fit L2: flush it between timed steps
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    # the product's fast path: the whole step (incl. the NCCL collectives) captured in a CUDA graph
    from audio_visual_tubes_b200.step import HeadStep
    hs = HeadStep(v, a, cfg, use_graph=not args.no_graph, overlap=not args.no_overlap)

    def step():
        return hs()

    def eager_step():
        if world > 1:
            return sharded_value_and_grad(v, a, cfg)
        return avt.hardway_value_and_grad(v, a, cfg)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()

    # ---- stage times and launch count: one profiled eager pass (events cannot live in a graph; with
    # profiling on the library runs every stage on ONE stream, so each time is that kernel alone)
    _lib.profile_enable(True)
    n0 = _lib.launch_count()
    for _ in range(3):
        eager_step()
    barrier()
    launches_per_step = (_lib.launch_count() - n0) // 3
    prof = _lib.profile_read()
    _lib.profile_enable(False)

    # ---- timed region: K steps, device time, max over ranks
    with ClockSampler(local_rank) as clk:
        if flush is None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record()
            for _ in range(args.steps):
                loss, dv, da = step()
            e1.record()
            barrier()
            ms_total = e0.elapsed_time(e1)
        else:
            evs = []
            barrier()
            for _ in range(args.steps):
                flush.fill_(1)
                s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s0.record()
                loss, dv, da = step()
                s1.record()
                evs.append((s0, s1))
            barrier()
            ms_total = sum(x.elapsed_time(y) for x, y in evs)
    launches = launches_per_step * args.steps
    t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t[0])
    ms_step = ms_total / args.steps
    value = cells_per_step / (ms_step * 1e-3)
    loss_val = float(loss.cpu()[0])
    # the same global batch on 1 GPU and on N: same loss (sum order differs only)
    assert abs(loss_val - loss_global_1gpu) <= 2e-6 * max(1.0, abs(loss_global_1gpu)), (loss_val, loss_global_1gpu)

    # ---- rooflines.  Event-timed kernels inside a millisecond-long step run at burst clocks, so the
    # burst bf16 / fp16 tensor peak is the denominator; per kernel and for the whole step.
    stage_ms = {k: (ms / n if n else None) for k, (ms, n) in prof.items()}
    local_cells = b_local * b_global * P
    gflop = local_cells * FLOP_FWD_PER_CELL / 1e9            # one contraction = 2*C flop per cell
    tensor_bound = b_global >= 512 and args.precision == "f16"
    kern = {}
    for key, name, credit in (("sim", "sim_tc_kernel<false> (forward contraction + pooling)", 1.0),
                              ("bwd_g", "sim_tc_kernel<true> (recompute + cell gradients)", 0.0),
                              ("bwd_dv", "bwd_dv_pair_kernel (dv contraction + normalise Jacobian)", 1.0),
                              ("bwd_da", "bwd_da_pair_kernel (d_ahat contraction)", 1.0)):
        ms = stage_ms.get(key)
        if ms:
            tf = gflop / ms                                   # GFLOP / ms = TFLOP/s of tensor work EXECUTED
            kern[key] = {"kernel": name, "ms": ms, "executed_tflops": tf, "frac_of_burst_peak": tf / pk["tc_burst"],
                         "algorithmic_credit": credit}
    step_tflops = cells_per_step * (FLOP_FWD_PER_CELL + FLOP_BWD_PER_CELL) / (ms_step * 1e-3) / 1e12 / world
    if tensor_bound and kern:
        top = max(kern, key=lambda k: kern[k]["ms"])
        roofline = {"kernel": kern[top]["kernel"], "bound": "tensor", "achieved": kern[top]["executed_tflops"],
                    "peak": pk["tc_burst"], "unit": "TFLOP/s", "frac": kern[top]["frac_of_burst_peak"],
                    "traffic": None, "ms_per_launch": kern[top]["ms"],
                    "peak_source": pk["source"] + ", BURST fp16/bf16 dense (kernel event-timed inside a ~1 ms step)",
                    "note": "the longest kernel of the step; 'achieved' counts the MMA flops it executes "
                            "(the recompute of S in the cell-gradient kernel earns no algorithmic credit: see step_frac)",
                    "step_frac": step_tflops / pk["tc_burst"], "step_tflops_per_gpu": step_tflops,
                    "step_frac_of_sustained": step_tflops / pk["tc_sustained"], "kernels": kern}
        tr = NCU_TRAFFIC.get(top) if (world == 1 and args.workload == "cfg5") else None
        if tr:
            roofline["traffic"], roofline["traffic_source"] = tr
    else:
        bytes_alg = b_local * C * P * 4 + b_global * C * 4 + b_local * (b_global + 2) * 4 + 3 * b_local * P * 4
        tot = (stage_ms["prep"] or 0) + (stage_ms["sim"] or 0) + (stage_ms["finish"] or 0)
        ach = bytes_alg / (tot * 1e-3) / 1e9
        roofline = {"kernel": "head forward (prep + sim + finish)", "bound": "hbm", "achieved": ach,
                    "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"], "traffic": None,
                    "peak_source": pk["source"], "ms_per_launch": tot, "kernels": kern}

    # ---- e2e: the same step from HOST buffers (pinned), H2D inside the timed region, loss D2H.
    e2e = e2e_leg(args, avt, cfg, hs, v, a, dev, world, barrier, loss_val, cells_per_step)

    # ---- extras.  Collective ones (train clips/s with DDP backbones, sharded eval) run on every rank;
    # the single-GPU ones on rank 0 only.
    also = {}
    if not args.no_extras:
        try:
            also.update(extras_collective(avt, ev, _lib, cfg, dev, world, rank, barrier))
        except Exception as e:                               # noqa: BLE001  keep the headline line alive
            also["collective_error"] = repr(e)[:300]
        if rank == 0 and world == 1:
            also.update(extras(avt, ev, _lib, cfg, dev, pk))

    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu_base, _ = cpu_head_baseline(min(256, b_global), 3, 1)
            cpu_base = {k: cpu_base[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as e:                               # noqa: BLE001  never lose the headline line
            cpu_base = {"error": repr(e)[:300]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": {"workload": workload_string(args.workload, b_local, b_global),
                       "parallelism": f"rows sharded over {world} rank(s), NCCL all-gather of audio"
                                      if world > 1 else "single GPU",
                       "launch": "eager" if args.no_graph else "CUDA graph replay (step incl. NCCL captured once)",
                       "arithmetic": "fp16 operands (tcgen05 kind::f16), fp32 accumulate" if args.precision == "f16"
                                     else "fp32 CUDA cores",
                       "batch": "one global batch, seed 1234, sliced by rank",
                       "l2": "inputs (%.0f MB/rank) exceed L2" % (in_bytes / 1e6) if flush is None
                             else "256 MiB L2 flush between timed steps (per-step events)"},
            "loss": loss_val, "loss_global_batch_on_1_gpu": loss_global_1gpu,
            "roofline": roofline,
            "stage_ms": stage_ms,
            "cpu_baseline": cpu_base,
            "e2e": e2e,
            "gpu_launches": launches,
            "clocks": clk.summary(),
            "also": also,
        }
        emit(line)
    if world > 1:
        # a captured graph holds NCCL work: tearing the communicator down under it can block, so
        # drain, meet the other ranks once more and leave without running destructors
        sys.stdout.flush()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        os._exit(0)


# DRAM bytes per launch of the step's kernels from the committed `ncu --set full` capture of this same
# command (dram__bytes_read.sum + dram__bytes_write.sum); filled from profiles/ by hand after each capture
NCU_TRAFFIC = {
    # profiles/r2_ncu_head.txt (B = 1024): dram__bytes_read.sum + dram__bytes_write.sum per launch
    "sim": (206696192 + 12365312, "profiles/r2_ncu_head.txt"),
    "bwd_g": (220016384 + 361715968, "profiles/r2_ncu_head.txt"),
    "bwd_dv": (620157952 + 381518080, "profiles/r2_ncu_head.txt"),
    "bwd_da": (618686464 + 4047104, "profiles/r2_ncu_head.txt"),
}


def e2e_leg(args, avt, cfg, hs, v, a, dev, world, barrier, loss_val, cells_per_step):
    """The step through the host-facing call, HostHeadPipeline.submit (step.py): every step copies its
    own features host->device from pinned memory and reads its own loss back; the copy of step i+1
    runs on a copy stream under the compute of step i.  Timed with fp16 host features (what a
    backbone under autocast hands over: half the PCIe bytes -- the headline) and with fp32 ones;
    the serial copy -> compute -> read variant beside them."""
    import torch
    import torch.distributed as dist
    from audio_visual_tubes_b200.step import HostHeadPipeline
    e2e_steps = max(4, min(args.steps, 10))

    def timed(fn_loop):
        barrier()
        t0 = time.perf_counter()
        fn_loop()
        barrier()
        te = torch.tensor([(time.perf_counter() - t0) / e2e_steps], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        return float(te[0])

    ah = a.cpu().pin_memory()
    out = {}
    for tag, dt in (("f16_features", torch.float16), ("f32_features", torch.float32)):
        if dt != torch.float32 and args.precision != "f16":
            continue
        vh = v.to(dt).cpu().pin_memory()
        pipe = HostHeadPipeline(tuple(v.shape), tuple(a.shape), cfg, use_graph=not args.no_graph, v_dtype=dt)
        losses = []

        def pipelined(n):
            pend = None
            for _ in range(n):
                h = pipe.submit(vh, ah)               # H2D of this step + compute, asynchronous
                if pend is not None:
                    losses.append(pend.loss())        # D2H read of the previous step's loss
                pend = h
            losses.append(pend.loss())

        pipelined(3)
        losses.clear()
        t_pipe = timed(lambda: pipelined(e2e_steps))
        tol = 1e-5 if dt == torch.float32 else 2e-4           # fp16 maps: one more 2^-11 rounding of v
        assert len(losses) == e2e_steps and all(abs(x - loss_val) <= tol * max(1.0, abs(loss_val)) for x in losses), \
            (tag, losses, loss_val)
        nbytes = vh.numel() * vh.element_size() + ah.numel() * 4
        out[tag] = {"value": cells_per_step / t_pipe, "ms_per_step": t_pipe * 1e3, "h2d_bytes_per_step": nbytes,
                    "h2d_GBps": nbytes / t_pipe / 1e9, "loss": losses[-1]}
        del pipe, vh
    # serial: copy, compute, read, one after another on one stream (fp32 features)
    vh = v.cpu().pin_memory()

    def serial_step():
        v.copy_(vh, non_blocking=True)
        a.copy_(ah, non_blocking=True)
        l, _, _ = hs()
        return float(l.cpu()[0])
    for _ in range(2):
        serial_step()
    t_serial = timed(lambda: [serial_step() for _ in range(e2e_steps)])
    head = out.get("f16_features") or out["f32_features"]
    e2e = {"value": head["value"], "unit": UNIT, "h2d_bytes_per_step": head["h2d_bytes_per_step"],
           "d2h_bytes_per_step": 4, "ms_per_step": head["ms_per_step"], "steps": e2e_steps,
           "h2d_GBps": head["h2d_GBps"],
           "path": "HostHeadPipeline.submit (double-buffered: H2D of step i+1 under compute of step i), "
                   + ("fp16 host features" if "f16_features" in out else "fp32 host features")
                   + "; dv stays on the device for the backbone's backward, the loss (4 B) is read back every step",
           "variants": out, "serial_fp32_ms_per_step": t_serial * 1e3, "pcie": pcie_link(dev)}
    return e2e


def pcie_link(dev):
    """Generation and width the GPU's PCIe link is running at: the e2e leg is the H2D copy, so a box
    whose link trained at half rate (seen in this pool: 28.6 instead of 54 GB/s) halves it."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(dev.index or 0)
        return {"gen": pynvml.nvmlDeviceGetCurrPcieLinkGeneration(h), "width": pynvml.nvmlDeviceGetCurrPcieLinkWidth(h),
                "gen_max": pynvml.nvmlDeviceGetMaxPcieLinkGeneration(h), "width_max": pynvml.nvmlDeviceGetMaxPcieLinkWidth(h)}
    except Exception as e:                                   # noqa: BLE001
        return {"error": repr(e)[:120]}


# ----------------------------------------------------------------------------- CPU eval baseline
_EV = {}


def _ref_eval_init(n_maps):
    """Worker initialiser: the seeded config-3 maps, the real VGG-SS boxes and the reference's utils
    (oracle/_ref when it travelled -> kind 'reference'; else the numpy oracle -> kind 'port')."""
    import numpy as np
    sys.path.insert(0, ROOT)
    from oracle import eval_oracle as eo
    g = np.load(os.path.join(ROOT, "tests", "golden", "vggss_golden.npz"))
    _EV["maps"] = eo.synthetic_maps(n_maps, 0)
    _EV["boxes"] = [g["boxes"][i, :g["count"][i]].tolist() for i in range(n_maps)]
    _EV["eo"] = eo
    _EV["kind"] = "port"
    try:
        import cv2                                            # noqa: F401
        from oracle import make_ref
        if make_ref.available():
            _EV["utils"] = make_ref.load()[1]
            _EV["kind"] = "reference"
    except Exception:                                         # noqa: BLE001
        pass
    return _EV["kind"]


def _ref_eval_range(lo_hi):
    """train_hardway_1frame.py:198-207 for maps lo..hi-1, one Python iteration per clip as the
    reference runs it: real cv2.resize + the reference's normalize_img / testset_gt / Evaluator."""
    import types
    import numpy as np
    lo, hi = lo_hi
    out = []
    if _EV["kind"] == "reference":
        import cv2
        U = _EV["utils"]
        args = types.SimpleNamespace(testset="vggss", gt_all={str(i): _EV["boxes"][i] for i in range(lo, hi)})
        for i in range(lo, hi):
            heatmap_now = cv2.resize(_EV["maps"][i], dsize=(224, 224), interpolation=cv2.INTER_LINEAR)
            heatmap_now = U.normalize_img(-heatmap_now)
            gt_map = U.testset_gt(args, str(i) + ".mp4")
            pred = 1 - heatmap_now
            threshold = np.sort(pred.flatten())[int(pred.shape[0] * pred.shape[1] / 2)]
            pred[pred > threshold] = 1
            pred[pred < 1] = 0
            with np.errstate(divide="ignore", invalid="ignore"):
                ciou, _, _ = U.Evaluator().cal_CIOU(pred, gt_map, 0.5)
            out.append(ciou)
    else:
        eo = _EV["eo"]
        for i in range(lo, hi):
            out.append(eo.evaluate_map(_EV["maps"][i], eo.rasterise_gt(_EV["boxes"][i], "vggss")[0])["ciou"])
    return out


def eval_worker(spec):
    lo, hi, start_at = spec.split(":")
    lo, hi, start_at = int(lo), int(hi), float(start_at)
    kind = _ref_eval_init(hi)
    _ref_eval_range((lo, lo + 4))
    late = time.time() > start_at
    while time.time() < start_at:
        time.sleep(0.005)
    t0 = time.time()
    _ref_eval_range((lo, hi))
    print(json.dumps({"t0": t0, "t1": time.time(), "late": late, "kind": kind}))


def cpu_eval_baseline(n_single=300, per_worker=96):
    """Reference eval glue on the host cores: single process (as the reference runs it) and one
    worker process per core."""
    cores = os.cpu_count() or 1
    kind = _ref_eval_init(max(n_single, 1))
    _ref_eval_range((0, 8))
    t0 = time.perf_counter()
    _ref_eval_range((0, n_single))
    t1 = time.perf_counter() - t0
    res = {"single_process": {"maps_per_s": n_single / t1, "cores": 1, "kind": kind,
                              "sample": f"{n_single} of the 5158 config-3 maps with their real VGG-SS boxes, "
                                        "train_hardway_1frame.py:198-207 per clip (cv2.resize, "
                                        + ("the reference's utils.normalize_img / testset_gt / Evaluator from oracle/_ref)"
                                           if kind == "reference" else "oracle/eval_oracle.py port)")}}
    # all cores: one worker PROCESS per core (the multiprocessing.Pool variant of SURVEY 8d, written with
    # plain subprocesses + a common start time so that no fork happens under a live CUDA context and a
    # stuck worker is killed by a timeout instead of hanging the bench)
    try:
        import subprocess
        nw = min(cores, 32)
        start_at = time.time() + 40.0                         # workers import cv2 / torch / sklearn first
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--eval-worker",
                                   f"{i * per_worker}:{(i + 1) * per_worker}:{start_at}"],
                                  stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
                 for i in range(nw)]
        durs = []
        for p in procs:
            try:
                o, _ = p.communicate(timeout=150)
                durs.append(json.loads(o.strip().splitlines()[-1]))
            except Exception:                                 # noqa: BLE001
                p.kill()
        if len(durs) == nw and all(d["late"] is False for d in durs):
            t_end = max(d["t1"] for d in durs); t_beg = min(d["t0"] for d in durs)
            res["pool_all_cores"] = {"maps_per_s": nw * per_worker / (t_end - t_beg), "cores": nw, "kind": durs[0]["kind"],
                                     "sample": f"{nw} worker processes x {per_worker} maps each, started together"}
        else:
            res["pool_all_cores"] = {"error": f"{len(durs)}/{nw} workers reported in time"}
    except Exception as e:                                    # noqa: BLE001
        res["pool_all_cores"] = {"error": repr(e)[:200]}
    return res


def extras(avt, ev, _lib, cfg, dev, pk):
    import numpy as np
    import torch
    from oracle import eval_oracle as eo
    out = {}

    def timed(fn, n, warm, flush=None):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(n):
            if flush is not None:
                flush.fill_(1)
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record(); fn(); s1.record()
            torch.cuda.synchronize()
            tot += s0.elapsed_time(s1)
        return tot / n

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    # config 3: 5158 maps against the REAL VGG-SS box lists (tests/golden/vggss_golden.npz: data copied
    # from metadata/vggss.json), materialising mode (dense GT in, up224 + mask out) -- HBM roofline
    n = 5158
    maps = torch.tensor(eo.synthetic_maps(n, 0), device=dev)
    gold = np.load(os.path.join(ROOT, "tests", "golden", "vggss_golden.npz"))
    real_boxes = [gold["boxes"][i, :gold["count"][i]].tolist() for i in range(n)]
    gt = avt.rasterise_gt(real_boxes, "vggss").contiguous()            # utils.py:264-273 on the device
    up = torch.empty(n, 224, 224, device=dev)
    mk = torch.empty(n, 224, 224, dtype=torch.uint8, device=dev)
    iw = torch.empty(n, dtype=torch.int32, device=dev); uw = torch.empty_like(iw)
    ci = torch.empty(n, dtype=torch.float64, device=dev)
    tkt = torch.empty(64, dtype=torch.uint8, device=dev)
    lib = _lib.load()

    def run_eval(materialise=True):
        _lib.check(lib.avt_eval_maps(_lib.ptr(maps), n, _lib.ptr(gt), None, None, 0, 0,
                                     _lib.ptr(up) if materialise else None,
                                     _lib.ptr(mk) if materialise else None, _lib.ptr(iw), _lib.ptr(uw),
                                     _lib.ptr(ci), _lib.ptr(tkt), _lib.stream_ptr()), "avt_eval_maps")
    ms = timed(run_eval, 10, 3)            # 1.56 GB of traffic per launch: larger than L2
    gbs = n * EVAL_BYTES_PER_MAP / (ms * 1e-3) / 1e9
    # end-to-end numbers of the same maps: must equal what the unmodified reference scored (1e-3)
    res3 = avt.evaluate_maps(maps, boxes=real_boxes, gt_mode="vggss")
    out["eval_cfg3_result"] = {"auc": res3["auc"], "ciou_at_half": res3["ciou_at_half"],
                               "reference_auc": float(gold["auc"]), "reference_ciou_at_half": float(gold["ciou_at_half"]),
                               "boxes": "real VGG-SS (5158 clips, 12 with negative coordinates, 2 with empty GT)"}
    assert abs(res3["auc"] - float(gold["auc"])) < 1e-3 and abs(res3["ciou_at_half"] - float(gold["ciou_at_half"])) < 1e-3
    out["eval_cfg3"] = {"maps_per_s": n / (ms * 1e-3), "ms": ms, "mode": "materialising (dense GT, up224 + mask out)",
                        "roofline": {"kernel": "eval_maps_kernel<true>", "bound": "hbm", "achieved": gbs,
                                     "peak": pk["hbm"], "unit": "GB/s", "frac": gbs / pk["hbm"],
                                     "traffic": 293586432 + 1246250000, "traffic_source": "profiles/r2_ncu_eval.txt"}}
    ms2 = timed(lambda: run_eval(False), 10, 3)
    out["eval_cfg3_metrics_only"] = {"maps_per_s": n / (ms2 * 1e-3), "ms": ms2}
    # the pipeline mode of the reference's test loop: boxes in (utils.py:264-273 rasterised inside
    # the kernel as bit planes), scalars out -- nothing but 784 B + a few boxes read per clip
    bx_np, bc_np = ev.pack_boxes(real_boxes, "vggss")
    bxs = torch.as_tensor(bx_np, device=dev).contiguous()
    bcs = torch.as_tensor(bc_np, device=dev).contiguous()
    max_b = int(bxs.shape[1])

    def run_eval_boxes():
        _lib.check(lib.avt_eval_maps(_lib.ptr(maps), n, None, _lib.ptr(bxs), _lib.ptr(bcs), max_b, 0, None, None,
                                     _lib.ptr(iw), _lib.ptr(uw), _lib.ptr(ci), _lib.ptr(tkt), _lib.stream_ptr()), "avt_eval_maps")
    ms3 = timed(run_eval_boxes, 10, 3)
    out["eval_cfg3_boxes_in_scalars_out"] = {"maps_per_s": n / (ms3 * 1e-3), "ms": ms3}
    # the reference's CPU eval glue on the host cores: single process and one process per core
    try:
        out["eval_cpu_baseline"] = cpu_eval_baseline()
    except Exception as e:                                   # noqa: BLE001
        out["eval_cpu_baseline"] = {"error": repr(e)[:200]}
    del gt, up, mk

    # config 2: B = 128 head step on one GPU, L2 flushed between steps
    g = torch.Generator(device=dev).manual_seed(7)
    v = torch.randn(128, C, P, device=dev, generator=g).abs_()
    a = torch.randn(128, C, device=dev, generator=g).abs_()
    _lib.profile_enable(True)
    ms = timed(lambda: avt.hardway_value_and_grad(v, a, cfg), 10, 3, flush)
    prof = _lib.profile_read(); _lib.profile_enable(False)
    cells = 128 * 128 * P
    fwd_ms = sum(prof[k][0] / max(prof[k][1], 1) for k in ("prep", "sim", "finish"))
    bytes_alg = 128 * C * P * 4 + 128 * C * 4 + 128 * 130 * 4 + 3 * 128 * P * 4
    out["head_cfg2_b128"] = {"cells_per_s": cells / (ms * 1e-3), "ms_fwd_bwd": ms, "ms_fwd": fwd_ms,
                             "fwd_hbm_gbs": bytes_alg / (fwd_ms * 1e-3) / 1e9,
                             "fwd_hbm_frac": bytes_alg / (fwd_ms * 1e-3) / 1e9 / pk["hbm"]}

    # the kernel-to-beat on the same box: the reference's eager PyTorch head on this B200
    try:
        out["reference_eager_b200"] = reference_eager_gpu(avt, cfg, dev)
    except Exception as e:                                   # noqa: BLE001
        out["reference_eager_b200"] = {"error": repr(e)[:300]}
    # config 1 (BASELINE.json configs[0]): the reference's own AVENet, B = 8, on the host cores
    try:
        out["config1_reference_cpu"] = config1_reference_cpu(avt, cfg, dev)
    except Exception as e:                                   # noqa: BLE001
        out["config1_reference_cpu"] = {"error": repr(e)[:300]}

    # config 4: tube model, b = 64 clips x T = 16 frames = 1024 frames.  Reference-faithful mode scores
    # every frame against the 1024 REPEATED audio rows (train_3D.py:129-130); tube mode scores it
    # against the 64 distinct ones with column multiplicity 16 -- same logits, loss and gradients.
    try:
        bt, T = 64, 16
        g4 = torch.Generator(device=dev).manual_seed(4)
        v4 = torch.randn(bt * T, C, P, device=dev, generator=g4).abs_()
        a4 = torch.randn(bt, C, device=dev, generator=g4).abs_()
        a4_rep = a4.repeat_interleave(T, dim=0).contiguous()
        ms_rep = timed(lambda: avt.hardway_value_and_grad(v4, a4_rep, cfg), 10, 3)
        ms_ded = timed(lambda: avt.hardway_value_and_grad(v4, a4, cfg, col_mult=T), 10, 3)
        l_rep = float(avt.hardway_value_and_grad(v4, a4_rep, cfg, want_grad=False)[0][0])
        l_ded = float(avt.hardway_value_and_grad(v4, a4, cfg, col_mult=T, want_grad=False)[0][0])
        cells4 = (bt * T) * (bt * T) * P
        out["tube_cfg4"] = {"frames": bt * T, "T": T, "reference_cells": cells4, "unique_cells": cells4 // T,
                            "repeated_audio": {"ms_fwd_bwd": ms_rep, "cells_per_s": cells4 / (ms_rep * 1e-3), "loss": l_rep},
                            "tube_mode": {"ms_fwd_bwd": ms_ded, "reference_cells_per_s": cells4 / (ms_ded * 1e-3),
                                          "unique_cells_per_s": cells4 / T / (ms_ded * 1e-3), "loss": l_ded},
                            "speedup": ms_rep / ms_ded}
        del v4, a4, a4_rep
    except Exception as e:
        out["tube_cfg4"] = {"error": repr(e)[:200]}

    # next-4: the dataloader's spectrogram (datasets/dataloader.py:86-96) for a config-2 batch of
    # 128 ten-second clips at 16 kHz: 4 B read per sample + 4 B written per bin, L2 flushed
    try:
        from oracle import spec_oracle as so
        from audio_visual_tubes_b200.audio import log_spectrogram
        fs, bs = 16000, 128
        wav = torch.from_numpy(so.synthetic_waveforms(4, fs, 9)).to(dev).repeat(bs // 4, 1).contiguous()
        ms = timed(lambda: log_spectrogram(wav, fs), 10, 3, flush)
        tfr = (fs * 10 - 1) // 511
        bytes_alg = bs * (fs * 10 * 4 + 257 * tfr * 4)
        gbs = bytes_alg / (ms * 1e-3) / 1e9
        # the same call from pinned host waveforms, spectrograms read back to the host
        wav_h = wav.cpu().pin_memory()
        spec_h = torch.empty(bs, 1, 257, tfr, dtype=torch.float32).pin_memory()

        wav_stage = torch.empty_like(wav)                  # device staging buffers, as a loader would keep

        def spec_e2e():
            wav_stage.copy_(wav_h, non_blocking=True)
            spec_h.copy_(log_spectrogram(wav_stage, fs), non_blocking=True)
        ms_e2e = timed(spec_e2e, 10, 3)
        pcm_h = (wav.cpu() * 32768.0).round().clamp_(-32768, 32767).to(torch.int16).pin_memory()
        pcm_d = pcm_h.to(dev)
        ms_pcm = timed(lambda: log_spectrogram(pcm_d, fs), 10, 3, flush)

        def spec_e2e_pcm():
            pcm_d.copy_(pcm_h, non_blocking=True)
            spec_h.copy_(log_spectrogram(pcm_d, fs), non_blocking=True)
        ms_e2e_pcm = timed(spec_e2e_pcm, 10, 3)
        t0 = time.perf_counter()
        w_np = wav[:4].cpu().numpy()
        reps = 0
        while time.perf_counter() - t0 < 2.0:
            so.log_spectrogram(w_np[reps % 4], fs); reps += 1
        cpu_cps = reps / (time.perf_counter() - t0)
        out["spectrogram_b128_16k"] = {"clips_per_s": bs / (ms * 1e-3), "ms": ms, "dtype": "f64 arithmetic, f32 out",
                                       "e2e": {"clips_per_s": bs / (ms_e2e * 1e-3), "ms": ms_e2e, "h2d_bytes": bs * fs * 10 * 4,
                                               "d2h_bytes": bs * 257 * tfr * 4},
                                       "pcm16_in": {"ms": ms_pcm, "clips_per_s": bs / (ms_pcm * 1e-3),
                                                    "e2e_ms": ms_e2e_pcm, "e2e_clips_per_s": bs / (ms_e2e_pcm * 1e-3),
                                                    "h2d_bytes": bs * fs * 10 * 2},
                                       "roofline": {"kernel": "spectrogram_kernel", "bound": "hbm", "achieved": gbs,
                                                    "peak": pk["hbm"], "unit": "GB/s", "frac": gbs / pk["hbm"],
                                                    "traffic": 81938944 + 11092224, "traffic_source": "profiles/r1_ncu_spectrogram.txt (the 41 MB output stays in L2)",
                                                    "note": "fp64 arithmetic: bound by the L1/shared data path (84 %) and the fp64 pipe (52 %), not HBM"},
                                       "cpu_baseline": {"clips_per_s": cpu_cps, "cores": 1, "kind": "port",
                                                        "sample": f"{reps} clips, oracle/spec_oracle.py log_spectrogram (numpy rfft)"}}
    except Exception as e:
        out["spectrogram_b128_16k"] = {"error": repr(e)[:200]}

    return out


def reference_eager_gpu(avt, cfg, dev):
    """'the bar to beat on the same box is the reference's own eager PyTorch
    execution'.  The UNMODIFIED reference head (oracle/_ref AVENet, Identity backbones; model.py:112-154)
    + nn.CrossEntropyLoss + backward, fp32 eager on this GPU (allow_tf32 off, as the reference never
    enables it), at B = 128 (config 2) and B = 1024 (configs 4/5: A0 is 822 MB, ~10 such temporaries
    with autograd).  Same features as the product arm; ours is timed beside it on the same tensors."""
    import torch
    from oracle import make_ref
    torch.backends.cuda.matmul.allow_tf32 = False
    net = make_ref.reference_head().to(dev)
    crit = torch.nn.CrossEntropyLoss()
    res = {"kind": "reference (oracle/_ref, unmodified model.py) on cuda, fp32 eager"}
    for b in (128, 1024):
        g = torch.Generator(device=dev).manual_seed(1234)
        v = torch.randn(b, C, P, device=dev, generator=g).abs_()
        a = torch.randn(b, C, device=dev, generator=g).abs_()
        tgt = torch.zeros(b, dtype=torch.long, device=dev)

        def ref_step():
            vt = v.view(b, C, 14, 14).clone().requires_grad_(True)
            at = a.view(b, C, 1, 1).clone().requires_grad_(True)
            o = net(vt, at)
            loss = crit(o[1], tgt)
            loss.backward()
            return loss, vt.grad

        def ours_step():
            return avt.hardway_value_and_grad(v, a, cfg)

        def timed(fn, n, warm):
            for _ in range(warm):
                r = fn()
            torch.cuda.synchronize()
            ts = []
            for _ in range(n):
                s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s0.record(); r = fn(); s1.record()
                torch.cuda.synchronize()
                ts.append(s0.elapsed_time(s1))
            ts.sort()
            return ts[len(ts) // 2], r
        torch.cuda.reset_peak_memory_stats(dev)
        m0 = torch.cuda.memory_allocated(dev)
        ms_ref, (l_ref, dv_ref) = timed(ref_step, 5, 2)
        peak = torch.cuda.max_memory_allocated(dev) - m0
        ms_ours, (l_ours, dv_ours, _) = timed(ours_step, 10, 3)
        cells = b * b * P
        err = float((dv_ours - dv_ref.view(b, C, P)).norm() / dv_ref.norm())
        res[f"b{b}"] = {"reference_ms_fwd_bwd": ms_ref, "reference_cells_per_s": cells / (ms_ref * 1e-3),
                        "reference_peak_extra_bytes": int(peak),
                        "ours_ms_fwd_bwd": ms_ours, "ours_cells_per_s": cells / (ms_ours * 1e-3),
                        "speedup": ms_ref / ms_ours, "launch": "eager (both)",
                        "loss_reference": float(l_ref), "loss_ours": float(l_ours[0]),
                        "dv_rel_err_vs_reference_fp32": err}
        del v, a, dv_ref, dv_ours
        torch.cuda.empty_cache()
    return res


def config1_reference_cpu(avt, cfg, dev):
    """BASELINE.json configs[0] / SURVEY 8d config 1: the reference ITSELF -- AVENet(args, False) with its
    ResNet-18 backbones (oracle/_ref), random init, image randn(8,3,224,224), audio randn(8,1,257,300),
    fp32 on all host cores: (a) forward + CE, (b) + backward, (c) the similarity stage alone on the
    pre-computed features; beside it this repo's AVENet on the GPU with the same inputs."""
    import types
    import numpy as np
    import torch
    from oracle import make_ref
    ref_model, _ = make_ref.load()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    args = types.SimpleNamespace(epsilon=0.65, epsilon2=0.4, tri_map=True, Neg=True)
    shim = make_ref.cpu_shim().__enter__()                   # keep model.py:115's mask on the CPU for this leg
    try:
        net = ref_model.AVENet(args, False)
        image = torch.randn(8, 3, 224, 224)
        audio = torch.randn(8, 1, 257, 300)
        crit = torch.nn.CrossEntropyLoss()
        tgt = torch.zeros(8).long()

        def med(fn, n=3, warm=1):
            for _ in range(warm):
                fn()
            ts = []
            for _ in range(n):
                t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
            return float(np.median(ts))

        def fwd():
            with torch.no_grad():
                return float(crit(net(image, audio)[1], tgt))

        def fwd_bwd():
            net.zero_grad(set_to_none=True)
            loss = crit(net(image, audio)[1], tgt)
            loss.backward()
        t_a = med(fwd)
        t_b = med(fwd_bwd, 2, 1)
        with torch.no_grad():
            img = net.imgnet(image)
            aud = net.audnet(audio)
        head = make_ref.reference_head()
        aud_p = torch.nn.functional.adaptive_max_pool2d(aud, 1)

        def stage():
            vt = img.clone().requires_grad_(True); at = aud_p.clone().requires_grad_(True)
            crit(head(vt, at)[1], tgt).backward()
        t_c = med(stage, 10, 2)
        loss_ref = fwd()
    finally:
        shim.__exit__()
    res = {"cores": cores, "kind": "reference (oracle/_ref: unmodified model.py + models/base_models.py)",
           "cells": 8 * 8 * P,
           "a_forward_plus_ce_s": t_a, "b_forward_backward_s": t_b, "c_similarity_stage_fwd_bwd_s": t_c,
           "clips_per_s_forward": 8 / t_a, "clips_per_s_train": 8 / t_b, "stage_cells_per_s": 8 * 8 * P / t_c,
           "loss": loss_ref}
    # this repo on the GPU, same inputs and the reference's weights (state-dict compatible names)
    try:
        ours = avt.AVENet(args, False, precision=cfg.precision).to(dev)
        missing = ours.load_state_dict(net.state_dict(), strict=False)
        ours.eval(); net.eval()
        im_d, au_d = image.to(dev), audio.to(dev)

        def g_fwd():
            with torch.no_grad():
                return ours.loss(im_d, au_d)
        for _ in range(3):
            l = g_fwd()
        torch.cuda.synchronize()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for _ in range(10):
            l = g_fwd()
        s1.record(); torch.cuda.synchronize()
        with make_ref.cpu_shim(), torch.no_grad():           # the reference in eval mode, same weights
            res["loss_eval_mode"] = float(crit(net(image, audio)[1], tgt))
        res["ours_gpu"] = {"forward_plus_ce_ms": s0.elapsed_time(s1) / 10, "loss_eval_mode": float(l),
                           "state_dict_missing": list(missing.missing_keys)[:4],
                           "state_dict_unexpected": list(missing.unexpected_keys)[:4]}
    except Exception as e:                                   # noqa: BLE001
        res["ours_gpu"] = {"error": repr(e)[:300]}
    return res


def extras_collective(avt, ev, _lib, cfg, dev, world, rank, barrier):
    """Legs every rank takes part in: (1) train clips/s of the whole step -- stock-PyTorch ResNet-18
    backbones under DDP (NCCL all-reduce of their gradients) + the row-sharded head with GLOBAL
    negatives + Adam, 128 clips per rank (config 2, weak scaling; train_hardway_1frame.py:125-134);
    (2) evaluation sharded over the ranks (maps are independent; integer success-curve counts
    all-reduced, train_hardway_1frame.py:208-215)."""
    import types
    import numpy as np
    import torch
    import torch.distributed as dist
    out = {}

    def timed(fn, n, warm):
        for _ in range(warm):
            fn()
        barrier()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for _ in range(n):
            fn()
        s1.record()
        barrier()
        t = torch.tensor([s0.elapsed_time(s1) / n], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    # ---- (1) train clips/s
    args = types.SimpleNamespace(epsilon=0.65, epsilon2=0.4, tri_map=True, Neg=True)
    torch.manual_seed(0)
    net = avt.AVENet(args, False, precision=cfg.precision).to(dev).train()
    model = net
    # the stock ResNets keep their `fc` layers for checkpoint compatibility; nothing on this path uses
    # them (in the reference they simply never get a gradient) and DDP must not wait for theirs
    # (nor the stems of the other modality: conv1_a / conv1_flow / audnet.conv1).  One small probe backward
    # finds them: whatever gets no gradient from the features is frozen.
    f_i, f_a = net.features(torch.randn(2, 3, 224, 224, device=dev), torch.randn(2, 1, 257, 300, device=dev))
    (f_i.float().sum() + f_a.float().sum()).backward()
    for _, prm in net.named_parameters():
        if prm.grad is None:
            prm.requires_grad_(False)
    net.zero_grad(set_to_none=True)
    del f_i, f_a
    if world > 1:
        from torch.nn.parallel import DistributedDataParallel as DDP

        class _Features(torch.nn.Module):                     # DDP wraps the part that has parameters
            def __init__(self, n):
                super().__init__(); self.n = n

            def forward(self, image, audio):
                return self.n.features(image, audio)
        model = DDP(_Features(net), device_ids=[dev.index])
    opt = torch.optim.Adam([q for q in net.parameters() if q.requires_grad], lr=1e-6, weight_decay=1e-4)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    img = torch.randn(128, 3, 224, 224, device=dev, generator=g)
    spec = torch.randn(128, 1, 257, 300, device=dev, generator=g)
    for tag, amp in (("fp32_backbones", False), ("bf16_autocast_backbones", True)):
        if amp:
            # stock PyTorch's fast path for convolutions under autocast: NHWC weights and inputs
            try:
                net = net.to(memory_format=torch.channels_last)
                img = img.contiguous(memory_format=torch.channels_last)
                spec = spec.contiguous(memory_format=torch.channels_last)
                if world > 1:
                    model = DDP(_Features(net), device_ids=[dev.index])
            except Exception as e:                           # noqa: BLE001
                out[f"train_cfg2_b128_per_gpu_{tag}"] = {"error": repr(e)[:300]}
                continue

        def train_step():
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                if world > 1:
                    f_img, f_aud = model(img, spec)
                else:
                    f_img, f_aud = net.features(img, spec)
            if world > 1:
                loss = avt.dist.sharded_hardway_loss(f_img, f_aud.float(), net.head_config)
            else:
                loss = avt.hardway_loss(f_img, f_aud.float(), net.head_config)
            loss.backward()
            opt.step()
        try:
            ms = timed(train_step, 5, 2)
            out[f"train_cfg2_b128_per_gpu_{tag}"] = {
                "clips_per_s": world * 128 / (ms * 1e-3), "ms_per_step": ms, "n_gpus": world,
                "global_batch": world * 128,
                "note": "stock PyTorch ResNet-18 backbones" + (" under bf16 autocast + channels_last (features reach the head as bf16)" if amp else " in fp32 (as the reference runs them)")
                        + (", DDP gradient all-reduce; head row-sharded with global negatives" if world > 1 else "; head = libavtubes")
                        + "; Adam(lr 1e-6, wd 1e-4)"}
        except Exception as e:                               # noqa: BLE001
            out[f"train_cfg2_b128_per_gpu_{tag}"] = {"error": repr(e)[:300]}
    del net, model, opt, img, spec
    torch.cuda.empty_cache()

    # ---- (2) sharded evaluation: config 3's 5158 maps, real boxes, split over the ranks
    from oracle import eval_oracle as eo
    gold = np.load(os.path.join(ROOT, "tests", "golden", "vggss_golden.npz"))
    n = 5158
    idx = np.arange(rank, n, world)
    maps = torch.tensor(eo.synthetic_maps(n, 0)[idx], device=dev)
    boxes = [gold["boxes"][i, :gold["count"][i]].tolist() for i in idx]
    bx, bc = ev.pack_boxes(boxes, "vggss", max_boxes=7)
    bx, bc = bx.to(dev), bc.to(dev)
    state = {}

    def eval_step():
        r = ev.eval_maps_raw(maps, boxes=bx, box_count=bc, gt_mode="vggss")
        if world > 1:
            state["res"] = avt.dist.sharded_eval_counts(r["ciou"])
        else:
            state["res"] = ev.success_auc(r["ciou"])
    ms = timed(eval_step, 5, 2)
    e = {"maps_per_s": n / (ms * 1e-3), "ms": ms, "n_gpus": world,
         "mode": "boxes in, scalars out; integer counts all-reduced" if world > 1 else "boxes in, scalars out"}
    if world > 1:
        half, auc, n_tot = state["res"]
        e.update(auc=auc, ciou_at_half=half, n_total=n_tot, reference_auc=float(gold["auc"]))
        assert n_tot == n and abs(auc - float(gold["auc"])) < 1e-3
    out["eval_cfg3_sharded"] = e
    return out


def emit(line: dict):
    """The ONE JSON line goes to the real stdout; everything else (NCCL banners, library chatter)
    was re-routed to stderr at start-up."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1
if __name__ == "__main__":
    if len(sys.argv) == 3 and sys.argv[1] == "--eval-worker":
        eval_worker(sys.argv[2])
        sys.exit(0)
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)                      # fd 1 -> stderr for C libraries and stray prints
    main()
