"""Dev: cycle breakdown of sim_tc_kernel (needs a library built with AVT_NVCC_EXTRA=-DAVT_TC_TRACE)."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import audio_visual_tubes_b200 as avt
from audio_visual_tubes_b200 import _lib
from audio_visual_tubes_b200.head import audio_normalize, head_forward_raw, head_backward_raw
lib = _lib.load()
b = 1024
g = torch.Generator(device="cuda").manual_seed(1234)
v = torch.randn(b, 512, 196, device="cuda", generator=g).abs_()
a = torch.randn(b, 512, device="cuda", generator=g).abs_()
cfg = avt.HeadConfig(precision="f16")
ah = audio_normalize(a)[0]
def dump(tag):
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * (8 * 148))()
    assert lib.avt_debug_tc_trace(buf, 148) == 0
    t = np.array(buf, dtype=np.float64).reshape(148, 8)
    lead = t[t[:, 2] > 0]
    print(tag, "blocks with an issuer:", len(lead), "tiles/block", lead[:, 6].mean())
    print("  issuer: total %.0f  wait operands %.0f (%.1f%%)  wait accumulator %.0f (%.1f%%)  per tile %.0f clk" % (
        lead[:, 2].mean(), lead[:, 0].mean(), 100 * lead[:, 0].mean() / lead[:, 2].mean(), lead[:, 1].mean(),
        100 * lead[:, 1].mean() / lead[:, 2].mean(), lead[:, 2].mean() / lead[:, 6].mean()))
    print("  epilogue warp 2: total %.0f  wait accumulator %.0f (%.1f%%)  sweep %.0f (%.1f%%)" % (
        t[:, 5].mean(), t[:, 3].mean(), 100 * t[:, 3].mean() / t[:, 5].mean(), t[:, 4].mean(), 100 * t[:, 4].mean() / t[:, 5].mean()))
for _ in range(2):
    out = head_forward_raw(v, ah, cfg, for_backward=True, fuse_grad_scale=1.0 / b)
dump("forward ")
for _ in range(2):
    dv, da = head_backward_raw(v, ah, cfg, out, 0.0, rowgrad_done=True)
dump("backward")


def dump_dv():
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * (8 * 148))()
    assert lib.avt_debug_dv_trace(buf, 148) == 0
    t = np.array(buf, dtype=np.float64).reshape(148, 8)
    lead = t[t[:, 2] > 0]
    tiles = lead[:, 6].mean()
    print("dv: blocks with an issuer:", len(lead), "tile pairs/block", tiles)
    print("  issuer: total %.0f  wait operands %.0f (%.1f%%)  wait accumulator %.0f (%.1f%%)  per tile %.0f clk" % (
        lead[:, 2].mean(), lead[:, 0].mean(), 100 * lead[:, 0].mean() / lead[:, 2].mean(), lead[:, 1].mean(),
        100 * lead[:, 1].mean() / lead[:, 2].mean(), lead[:, 2].mean() / tiles))
    e = t[t[:, 5] > 0]
    print("  epilogue warp 0: total %.0f  wait accumulators %.0f (%.1f%%)  pass 1 %.0f (%.1f%%)  pass 2 %.0f (%.1f%%)  per tile: wait %.0f p1 %.0f p2 %.0f" % (
        e[:, 5].mean(), e[:, 3].mean(), 100 * e[:, 3].mean() / e[:, 5].mean(), e[:, 4].mean(), 100 * e[:, 4].mean() / e[:, 5].mean(),
        e[:, 7].mean(), 100 * e[:, 7].mean() / e[:, 5].mean(), e[:, 3].mean() / tiles, e[:, 4].mean() / tiles, e[:, 7].mean() / tiles))


if hasattr(lib, "avt_debug_dv_trace"):
    dump_dv()

