"""Dev: device time of the un-profiled head step (forward + backward) and of the backward alone at a
given shape, as the env switches of this process set it up (one process per variant)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import audio_visual_tubes_b200 as avt
from audio_visual_tubes_b200.head import audio_normalize, head_forward_raw, head_backward_raw

b = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
n_a = int(sys.argv[2]) if len(sys.argv) > 2 else b
g = torch.Generator(device="cuda").manual_seed(1234)
v = torch.randn(b, 512, 196, device="cuda", generator=g).abs_()
a = torch.randn(n_a, 512, device="cuda", generator=g).abs_()
pos = (torch.arange(b, device="cuda") % n_a).int()
cfg = avt.HeadConfig(precision="f16")
ah = audio_normalize(a)[0]


def fwd():
    return head_forward_raw(v, ah, cfg, pos, 0, False, for_backward=True, fuse_grad_scale=1.0 / b)


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


out = fwd()
t_b = timed(lambda: head_backward_raw(v, ah, cfg, out, 0.0, rowgrad_done=True))
t_f = timed(fwd)
dv, da = head_backward_raw(v, ah, cfg, out, 0.0, rowgrad_done=True)
torch.cuda.synchronize()
sw = " ".join(f"{k}={os.environ[k]}" for k in sorted(os.environ) if k.startswith("AVT_"))
print(f"[{sw}] B={b} N_a={n_a}: fwd {t_f*1e3:7.1f} us  bwd {t_b*1e3:7.1f} us  |dv| {float(dv.double().norm()):.6e} |da| {float(da.double().norm()):.6e}")
