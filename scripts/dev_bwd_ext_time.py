"""Dev: the backward as a row-sharded rank runs it (dv on a caller stream, d_ahat on the main one): when
is d_ahat complete (the reduce-scatter can start) and when has dv joined?  One process per env variant."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import audio_visual_tubes_b200 as avt
from audio_visual_tubes_b200.head import audio_normalize, head_forward_raw, head_backward_raw

b = int(sys.argv[1]); n_a = int(sys.argv[2])
g = torch.Generator(device="cuda").manual_seed(1234)
v = torch.randn(b, 512, 196, device="cuda", generator=g).abs_()
a = torch.randn(n_a, 512, device="cuda", generator=g).abs_()
pos = (torch.arange(b, device="cuda") % n_a).int()
cfg = avt.HeadConfig(precision="f16")
ah = audio_normalize(a)[0]
out = head_forward_raw(v, ah, cfg, pos, 0, False, for_backward=True, fuse_grad_scale=1.0 / b)
ext = torch.cuda.Stream()
cur = torch.cuda.current_stream()
t_da = t_all = 0.0
N = 20
for i in range(N + 3):
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    torch.cuda.synchronize()
    e0.record(cur)
    ext.wait_stream(cur)
    dv, da = head_backward_raw(v, ah, cfg, out, 0.0, rowgrad_done=True, dv_stream=ext)
    e1.record(cur)                       # d_ahat complete: the reduce-scatter would start here
    cur.wait_stream(ext)
    e2.record(cur)                       # dv joined
    torch.cuda.synchronize()
    if i >= 3:
        t_da += e0.elapsed_time(e1); t_all += e0.elapsed_time(e2)
sw = " ".join(f"{k}={os.environ[k]}" for k in sorted(os.environ) if k.startswith("AVT_"))
tail = 25.0                              # reduce-scatter + audio normalise Jacobian behind d_ahat (us)
print(f"[{sw}] {b} x {n_a}: d_ahat done {t_da/N*1e3:6.1f} us  dv joined {t_all/N*1e3:6.1f} us  "
      f"critical path with a {tail:.0f} us tail behind d_ahat: {max(t_all/N*1e3, t_da/N*1e3 + tail):6.1f} us")
