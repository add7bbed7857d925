"""Developer probe: evaluation kernel at config 3 (5158 maps) -- parity on a sample, then timings
of the materialising (dense GT in, up224 + mask out), dense metrics-only and boxes modes."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from audio_visual_tubes_b200 import _lib, evaluate_maps
if os.environ.get('AVT_LIB_PATH'):
    _lib.LIB_PATH = os.environ['AVT_LIB_PATH']
from oracle import eval_oracle as eo

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5158
dev = "cuda"
maps_np = eo.synthetic_maps(n, 0)
maps = torch.tensor(maps_np, device=dev)
gt = torch.zeros(n, 224, 224, dtype=torch.uint8, device=dev); gt[:, 40:160, 50:180] = 1
up = torch.empty(n, 224, 224, device=dev); mk = torch.empty(n, 224, 224, dtype=torch.uint8, device=dev)
iw = torch.empty(n, dtype=torch.int32, device=dev); uw = torch.empty_like(iw)
ci = torch.empty(n, dtype=torch.float64, device=dev)
tk = torch.empty(64, dtype=torch.uint8, device=dev)
bx = torch.tensor([[[50, 40, 180, 160]]], dtype=torch.int16, device=dev).repeat(n, 1, 1).contiguous()
bc = torch.ones(n, dtype=torch.uint8, device=dev)
lib = _lib.load()

def run(mode):
    if mode == "boxes":
        a = (_lib.ptr(maps), n, None, _lib.ptr(bx), _lib.ptr(bc), 1, 0, None, None)
    else:
        a = (_lib.ptr(maps), n, _lib.ptr(gt), None, None, 0, 0,
             _lib.ptr(up) if mode == "mat" else None, _lib.ptr(mk) if mode == "mat" else None)
    _lib.check(lib.avt_eval_maps(*a, _lib.ptr(iw), _lib.ptr(uw), _lib.ptr(ci), _lib.ptr(tk), _lib.stream_ptr()), "avt_eval_maps")

import ctypes
st = (ctypes.c_ulonglong * 2)()
lib.avt_eval_stats(st, 1)
run("mat"); torch.cuda.synchronize()
lib.avt_eval_stats(st, 1)
print(f"generic-path maps: {st[0]} of {st[1]}")
g_np = gt[0].cpu().numpy()
bad = 0
for i in range(0, n, max(1, n // 40)):
    o = eo.evaluate_map(maps_np[i], g_np)
    ok = np.array_equal(mk[i].cpu().numpy(), o["mask"]) and int(iw[i]) == o["inter_w"] and int(uw[i]) == o["union_w"] \
        and np.array_equal(up[i].cpu().numpy(), o["up224"])
    bad += (not ok)
print("parity sample: mismatches =", bad)
ref = (iw.clone(), uw.clone())
for mode in ("mat", "dense", "boxes"):
    for _ in range(3): run(mode)
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(mode); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    same = torch.equal(iw, ref[0]) and torch.equal(uw, ref[1])
    gb = n * 301840 / 1e9 if mode == "mat" else 0
    print(f"{mode:6s} {ms:8.4f} ms  {n/ms*1e3/1e6:7.2f} Mmaps/s  {gb/ms*1e3:8.1f} GB/s  counts_same={same}")
