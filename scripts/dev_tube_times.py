"""Developer probe: config 4 (b=64 clips x T=16 frames) -- per-stage device times of the step with
repeated audio (reference-faithful) and in tube mode (64 distinct rows, col_mult = 16)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import audio_visual_tubes_b200 as avt
from audio_visual_tubes_b200 import _lib

b, T = 64, 16
g = torch.Generator(device="cuda").manual_seed(4)
v = torch.randn(b * T, 512, 196, device="cuda", generator=g).abs_()
a = torch.randn(b, 512, device="cuda", generator=g).abs_()
a_rep = a.repeat_interleave(T, dim=0).contiguous()
cfg = avt.HeadConfig()
for name, fn in (("repeated", lambda: avt.hardway_value_and_grad(v, a_rep, cfg)),
                 ("tube", lambda: avt.hardway_value_and_grad(v, a, cfg, col_mult=T))):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record(); torch.cuda.synchronize()
    st = {k: ms / c * 1e3 for k, (ms, c) in _lib.profile_read().items() if c}
    _lib.profile_enable(False)
    print(f"{name:9s}: eager step {e0.elapsed_time(e1)*100:7.1f} us, kernels {sum(st.values()):7.1f} us  " +
          " ".join(f"{k}={x:.1f}" for k, x in st.items()))
