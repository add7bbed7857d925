"""One B=1024 head step (forward + backward), a few times: the command ncu profiles."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import audio_visual_tubes_b200 as avt
b = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
g = torch.Generator(device="cuda").manual_seed(1234)
v = torch.randn(b, 512, 196, device="cuda", generator=g).abs_()
a = torch.randn(b, 512, device="cuda", generator=g).abs_()
cfg = avt.HeadConfig(precision="f16")
for _ in range(reps):
    loss, dv, da = avt.hardway_value_and_grad(v, a, cfg)
torch.cuda.synchronize()
print("loss", float(loss[0]))
