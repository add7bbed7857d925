"""Developer probe: bf16 tensor-core backward against the fp32 CUDA-core backward on the GPU."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch
import audio_visual_tubes_b200 as avt
from inputs import head_features


def run(n_v, n_a, prec, mode="mix", seed=3):
    v, a = head_features(n_v, n_a, seed, mode)
    pos = torch.tensor(np.arange(n_v) % n_a, device="cuda", dtype=torch.int32)
    vt = torch.tensor(v, device="cuda"); at = torch.tensor(a, device="cuda")
    loss, dv, da = avt.hardway_value_and_grad(vt, at, avt.HeadConfig(precision=prec), pos_idx=pos)
    torch.cuda.synchronize()
    return float(loss), dv.double().cpu().numpy(), da.double().cpu().numpy()


def rn(a, b):
    return np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30)


for n_v, n_a in [(8, 8), (3, 5), (33, 17), (7, 130), (128, 128), (64, 1024)]:
    for mode in ("mix", "relu"):
        l0, dv0, da0 = run(n_v, n_a, "fp32", mode)
        l1, dv1, da1 = run(n_v, n_a, "f16", mode)
        print(f"N_v={n_v} N_a={n_a} {mode}: loss {l0:.6f}/{l1:.6f}  dv rel {rn(dv1, dv0):.2e}  da rel {rn(da1, da0):.2e}"
              f"  |dv| {np.linalg.norm(dv0):.3e} |da| {np.linalg.norm(da0):.3e}", flush=True)
