"""Developer probe: per-stage device times of the head step at a given batch (library timers)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import audio_visual_tubes_b200 as avt
from audio_visual_tubes_b200 import _lib

b = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
n_a = int(sys.argv[2]) if len(sys.argv) > 2 else b
g = torch.Generator(device="cuda").manual_seed(1)
v = torch.randn(b, 512, 196, device="cuda", generator=g).abs_()
a = torch.randn(n_a, 512, device="cuda", generator=g).abs_()
pos = (torch.arange(b, device="cuda") % n_a).int()
cfg = avt.HeadConfig()
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
print(f"B={b} N_a={n_a}: step {e0.elapsed_time(e1)/K:.4f} ms")
for k, (ms, n) in _lib.profile_read().items():
    if n:
        print(f"  {k:10s} {ms/n*1e3:9.1f} us")
