"""Developer probe: quick device timings of the individual entry points (not the bench)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden"))
import numpy as np
import torch
import audio_visual_tubes_b200 as avt
from audio_visual_tubes_b200 import evaluation as ev
from oracle import eval_oracle as eo


def timeit(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(n):
        fn()
    t1.record(); torch.cuda.synchronize()
    return t0.elapsed_time(t1) / n


def main():
    what = sys.argv[1:] or ["eval", "head"]
    if "eval" in what:
        n = 5158
        maps = torch.tensor(eo.synthetic_maps(n, 0), device="cuda")
        gt = torch.zeros(n, 224, 224, dtype=torch.uint8, device="cuda"); gt[:, 50:150, 60:170] = 1
        ms = timeit(lambda: ev.eval_maps_raw(maps, gt_dense=gt, want_up224=True, want_mask=True))
        print(f"eval materialising: {ms:.3f} ms  {n/ms*1e3:.0f} maps/s  {n*301840/ms/1e6:.1f} GB/s")
        ms = timeit(lambda: ev.eval_maps_raw(maps, gt_dense=gt))
        print(f"eval dense-gt, scalars only: {ms:.3f} ms  {n/ms*1e3:.0f} maps/s")
    if "head" in what:
        for prec in ("fp32", "f16"):
            for B in (128, 1024):
                if prec == "fp32" and B == 1024:
                    continue
                try:
                    v = torch.randn(B, 512, 196, device="cuda").abs_()
                    a = torch.randn(B, 512, device="cuda").abs_()
                    cfg = avt.HeadConfig(precision=prec)
                    a_hat, _ = avt.audio_normalize(a)
                    from audio_visual_tubes_b200.head import head_forward_raw, head_backward_raw
                    ms = timeit(lambda: head_forward_raw(v, a_hat, cfg))
                    cells = B * B * 196
                    print(f"head fwd {prec} B={B}: {ms*1e3:.1f} us  {cells/ms*1e3:.3e} cells/s  {cells*1024/ms/1e9:.1f} TFLOP/s")
                    s = head_forward_raw(v, a_hat, cfg, for_backward=True)
                    ms = timeit(lambda: head_backward_raw(v, a_hat, cfg, s, 1.0 / B))
                    print(f"head bwd {prec} B={B}: {ms*1e3:.1f} us")
                except Exception as e:
                    print(f"head {prec} B={B}: FAILED {e}")


if __name__ == "__main__":
    main()
