"""The row-sharded head on real GPUs under NCCL (one process per GPU), against the single-GPU
result at the global batch -- the check round 1 lacked: a wrong col_offset, a mis-scattered
d_ahat or a stale graph buffer under NCCL would show up here.

ONE seeded global batch is built on every rank and sliced by rank.  Each rank runs
  * ``HeadStep(group=...)`` (the product's CUDA-graph-captured step, replayed twice) and
  * ``dist.sharded_value_and_grad`` (the eager step),
then computes the global-batch step by itself on its own GPU with ``hardway_value_and_grad`` and
compares its slice: loss, dv rows, da rows.  The global problem is also compared with the fp64
oracle on rank 0.  Skipped unless >= 2 GPUs are visible (world 2; world 8 when 8 are).
"""
import os
import socket

import numpy as np
import pytest
import torch

from inputs import head_features

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _relnorm(x, y):
    return float((x.double() - y.double()).norm() / y.double().norm().clamp_min(1e-30))


def _worker(rank, world, port, b_global, precision, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    os.environ.setdefault("TORCH_NCCL_HEARTBEAT_TIMEOUT_SEC", "120")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    def mark(what):                          # progress trail on stderr: a hang names its step
        torch.cuda.synchronize()
        print(f"[test_dist_gpu rank {rank}] {what}", file=__import__("sys").stderr, flush=True)
    try:
        from audio_visual_tubes_b200 import HeadConfig, HeadStep, hardway_value_and_grad
        from audio_visual_tubes_b200.dist import sharded_value_and_grad
        cfg = HeadConfig(precision=precision)
        v, a = head_features(b_global, b_global, 2024, "mix")
        b_local = b_global // world
        sl = slice(rank * b_local, (rank + 1) * b_local)
        v_all = torch.tensor(v, device=dev); a_all = torch.tensor(a, device=dev)
        v_loc = v_all[sl].contiguous(); a_loc = a_all[sl].contiguous()
        # single-GPU result at the global batch, on this rank's own GPU
        loss1, dv1, da1 = hardway_value_and_grad(v_all, a_all, cfg)
        mark("single-GPU step done")
        # eager sharded step
        loss_e, dv_e, da_e = sharded_value_and_grad(v_loc, a_loc, cfg)
        mark("eager sharded step done")
        # graph-captured step (collectives inside the graph), replayed twice
        hs = HeadStep(v_loc, a_loc, cfg)
        assert hs.sharded and hs.world == world
        mark("graph captured")
        hs()
        loss_g, dv_g, da_g = hs()
        mark("graph replayed twice")
        res = dict(
            loss1=float(loss1[0]), loss_e=float(loss_e[0]), loss_g=float(loss_g[0]),
            dv_e=_relnorm(dv_e, dv1[sl]), da_e=_relnorm(da_e, da1[sl]),
            dv_g=_relnorm(dv_g, dv1[sl]), da_g=_relnorm(da_g, da1[sl]),
            dv_bits=bool(torch.equal(dv_g, dv_e)))
        # forward outputs of the rank's rows == rows of the single-GPU forward
        from audio_visual_tubes_b200.dist import sharded_forward
        from audio_visual_tubes_b200.head import audio_normalize, head_forward_raw
        A_s, lg_s, _, _, rows_s = sharded_forward(v_loc, a_loc, cfg)
        full = head_forward_raw(v_all, audio_normalize(a_all)[0], cfg)
        res["logits_bits"] = bool(torch.equal(lg_s, full["logits"][sl]))
        res["A_bits"] = bool(torch.equal(A_s, full["A"][sl]))
        if rank == 0:
            from oracle import head_oracle as ho
            o, dv_o, da_o = ho.head_forward_backward_blocks(v, a, row_block=64)
            res["oracle_loss"] = float(o["loss"])
            res["oracle_dv"] = float(np.linalg.norm(dv1.cpu().numpy() - dv_o) / np.linalg.norm(dv_o))
            res["oracle_da"] = float(np.linalg.norm(da1.cpu().numpy() - da_o) / np.linalg.norm(da_o))
            # the sharded result held to the oracle by itself, on this rank's rows
            res["oracle_da_sharded"] = float(np.linalg.norm(da_g.cpu().numpy() - da_o[sl]) / np.linalg.norm(da_o[sl]))
            res["oracle_dv_sharded"] = float(np.linalg.norm(dv_g.cpu().numpy() - dv_o[sl]) / np.linalg.norm(dv_o[sl]))
        np.save(os.path.join(out_dir, f"r{rank}.npy"), res, allow_pickle=True)
        mark("results saved")
        dist.barrier()
        mark("final barrier passed")
    except BaseException:
        import traceback
        traceback.print_exc()
        os._exit(1)
    # A captured graph holds NCCL work: tearing the communicator down under it (destroy_process_group,
    # or interpreter shutdown running destructors in some order) can block for the watchdog's whole
    # timeout.  Results are on disk and both ranks have met: leave without running destructors, as
    # bench.py does.
    os._exit(0)


def _run(world, b_global, precision, tmp_path, limit_s=300):
    import time
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, b_global, precision, str(tmp_path))) for r in range(world)]
    [p.start() for p in procs]
    t0 = time.time()
    while any(p.is_alive() for p in procs) and time.time() - t0 < limit_s:
        time.sleep(0.5)
    stuck = [p for p in procs if p.is_alive()]
    for p in stuck:                      # never leave a rank behind spinning on a GPU
        p.kill()
    [p.join() for p in procs]
    assert not stuck, f"{len(stuck)} of {world} ranks still running after {limit_s} s (killed)"
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    out = [np.load(os.path.join(str(tmp_path), f"r{r}.npy"), allow_pickle=True).item() for r in range(world)]
    for r, z in enumerate(out):
        print(f"world {world} rank {r} [{precision}]: {z}")
        ref = z["loss1"]
        # the loss is a sum in a different order (per-rank partial sums, then all_reduce)
        assert abs(z["loss_e"] - ref) <= 2e-6 * max(1.0, abs(ref)), (r, z)
        assert abs(z["loss_g"] - ref) <= 2e-6 * max(1.0, abs(ref)), (r, z)
        # rows are computed by the same kernels on the same operands; only the power-of-two scale of
        # the fp16 cell gradients (max over the rank's rows vs over all rows) and the order of the
        # d_ahat atomics / reduce_scatter differ
        assert z["dv_e"] < 1e-5 and z["dv_g"] < 1e-5, (r, z)
        # d_ahat sums over ALL visual rows in fp32 inside tensor memory.  One GPU runs the K = N_v*196 rows as
        # a few long accumulation chains (7 K splits beside the dv kernel), a rank of `world` runs 1/world of
        # them and the partial sums meet in the reduce_scatter: same terms, different association.  At
        # B = 256 / world 2 the chains are short and the two agree to 1e-5; at B = 1024 / world 8 the single-GPU
        # chains are 8x longer (1792 accumulations each) and the two differ by ~1e-4 while each sits 1.5e-4
        # from the fp64 oracle (measured: profiles/r2_pytest_dist_gpu.txt) -- the 2e-3 contract is checked
        # against the oracle below for BOTH.  A mis-scattered or dropped block would be >= 3e-4 coherent and
        # would also break the loss / dv equalities above.
        da_tol = 1e-5 if b_global <= 256 else 3e-4
        assert z["da_e"] < da_tol and z["da_g"] < da_tol, (r, z)
        if precision == "f16":     # (the fp32 CUDA-core path accumulates dv with atomics)
            assert z["dv_bits"], "graph replay and eager sharded step disagree on dv"
        assert z["logits_bits"] and z["A_bits"], (r, z)
    z = out[0]
    gtol = 2e-3 if precision == "f16" else 2e-4
    assert abs(z["loss1"] - z["oracle_loss"]) < 2e-3 * max(1.0, abs(z["oracle_loss"]))
    assert z["oracle_dv"] < gtol and z["oracle_da"] < gtol, z
    assert z["oracle_dv_sharded"] < gtol and z["oracle_da_sharded"] < gtol, z


@pytest.mark.timeout(600)
@pytest.mark.parametrize("precision", ["f16", "fp32"])
def test_sharded_step_equals_single_gpu_world2(precision, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    _run(2, 256, precision, tmp_path)


@pytest.mark.timeout(600)
def test_sharded_step_equals_single_gpu_world8(tmp_path):
    if torch.cuda.device_count() < 8:
        pytest.skip("needs 8 GPUs")
    _run(8, 1024, "f16", tmp_path)
