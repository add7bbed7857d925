"""Developer probe: per-stage device times of one rank's share of the B=1024 step at N = 1, 2, 4, 8
(rows/rank = 1024/N against all 1024 audio columns), on a single GPU."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import audio_visual_tubes_b200 as avt
from audio_visual_tubes_b200 import _lib

g = torch.Generator(device="cuda").manual_seed(1)
cfg = avt.HeadConfig()
for n in (1, 2, 4, 8):
    b = 1024 // n
    v = torch.randn(b, 512, 196, device="cuda", generator=g).abs_()
    a = torch.randn(1024, 512, device="cuda", generator=g).abs_()
    pos = torch.arange(b, device="cuda").int()
    for _ in range(3):
        avt.hardway_value_and_grad(v, a, cfg, pos_idx=pos)
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    K = 10
    for _ in range(K):
        avt.hardway_value_and_grad(v, a, cfg, pos_idx=pos)
    e1.record(); torch.cuda.synchronize()
    st = {k: ms / c * 1e3 for k, (ms, c) in _lib.profile_read().items() if c}
    _lib.profile_enable(False)
    tot = sum(st.values())
    print(f"N={n} rows/rank={b}: eager step {e0.elapsed_time(e1)/K*1e3:7.1f} us, kernels {tot:7.1f} us  " +
          " ".join(f"{k}={x:.1f}" for k, x in st.items()))
