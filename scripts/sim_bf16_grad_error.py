"""CPU emulation of the bf16 tensor-core head (forward + backward) with each rounding site
switchable, to attribute the gradient error against the fp64 autograd oracle.
Dev tool (imports the oracle: not product code)."""
import sys, os, itertools
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden"))
from inputs import head_features
from oracle import head_oracle as ho

torch.set_num_threads(8)
def bf(x): return x.to(torch.bfloat16).to(torch.float32)
def tf32(x):  # round-to-nearest tf32 (10-bit mantissa)
    i = x.contiguous().view(torch.int32)
    i = (i + 0x1000) & ~0x1FFF
    return i.view(torch.float32)
def split2(x):
    hi = bf(x); lo = bf(x - hi); return hi, lo

def emulate(v, a, opt, eps=0.65, eps2=0.4, tau=0.03, temp=0.07, ms=100.0):
    v = torch.tensor(v); a = torch.tensor(a)
    N, C, P = v.shape
    rn = 1.0 / v.pow(2).sum(1).sqrt().clamp_min(1e-12)            # [N,P]
    vh = v * rn[:, None, :]
    ah = torch.nn.functional.normalize(a, dim=1)
    rv = {"bf16": bf, "tf32": tf32, "fp32": lambda x: x}
    vS = rv[opt["S_v"]](vh); aS = rv[opt["S_a"]](ah)
    if opt.get("S_split"):        # 3-term split bf16
        vhi, vlo = split2(vh); ahi, alo = split2(ah)
        S = torch.einsum('ncp,kc->nkp', vhi, ahi) + torch.einsum('ncp,kc->nkp', vhi, alo) + torch.einsum('ncp,kc->nkp', vlo, ahi)
    else:
        S = torch.einsum('ncp,kc->nkp', vS, aS)
    A = torch.einsum('ncp,nc->np', vh, ah)                         # fp32 prep
    W = torch.sigmoid((S - eps) / tau)
    den = W.sum(-1); sim = (W * S).sum(-1) / den
    Pos = torch.sigmoid((A - eps) / tau); Neg = 1 - torch.sigmoid((A - eps2) / tau)
    sim1 = (Pos * A).sum(-1) / Pos.sum(-1); sim2 = (Neg * A).sum(-1) / Neg.sum(-1)
    mask = torch.ones(N, N); mask[torch.arange(N), torch.arange(N)] = 1 - ms
    logits = torch.cat([sim1[:, None], sim * mask, sim2[:, None]], 1) / temp
    sm = torch.softmax(logits, 1); sm[:, 0] -= 1
    g = sm / N                                                     # dloss/dlogits
    gsim = g[:, 1:1 + N] * mask / temp
    # GA: grad through A of sim1, sim2
    def pooled_grad(Wt, dWdA, pooled, gout):
        d = Wt.sum(-1, keepdim=True)
        return gout[:, None] / temp * (Wt + (A - pooled[:, None]) * dWdA) / d
    GA = pooled_grad(Pos, Pos * (1 - Pos) / tau, sim1, g[:, 0]) + \
         pooled_grad(Neg, -(1 - Neg) * Neg / tau, sim2, g[:, -1])
    coef = gsim / den
    G = coef[:, :, None] * W * (1 + (S - sim[:, :, None]) * (1 - W) / tau)
    rG = rv[opt["G"]]
    if opt.get("GA_double", True) and opt["G"] == "bf16":
        G = rG(G)
        G[torch.arange(N), torch.arange(N)] = rG(G[torch.arange(N), torch.arange(N)] + GA)
    else:
        G[torch.arange(N), torch.arange(N)] += GA
        G = rG(G)
    a_dv = rv[opt["dv_a"]](ah)
    dvh = torch.einsum('nkp,kc->ncp', G, a_dv)
    vj = rv[opt["jac_v"]](vh)
    t = (vj * dvh).sum(1, keepdim=True)
    dv = rn[:, None, :] * (dvh - vj * t)
    v_da = rv[opt["da_v"]](vh)
    dah = torch.einsum('nkp,ncp->kc', G, v_da)
    na = 1.0 / a.pow(2).sum(1).sqrt().clamp_min(1e-12)
    da = na[:, None] * (dah - ah * (ah * dah).sum(1, keepdim=True))
    return logits.numpy(), dv.numpy(), da.numpy()

def relnorm(x, y): return float(np.linalg.norm((x - y).ravel()) / np.linalg.norm(y.ravel()))

base = dict(S_v="bf16", S_a="bf16", G="bf16", dv_a="bf16", jac_v="bf16", da_v="bf16", GA_double=True)
variants = {
  "current(all bf16)": base,
  "jac fp32 + GA single": dict(base, jac_v="fp32", GA_double=False),
  "S fp32 only": dict(base, S_v="fp32", S_a="fp32"),
  "S tf32": dict(base, S_v="tf32", S_a="tf32"),
  "S split3": dict(base, S_split=True),
  "S tf32 + jac fp32": dict(base, S_v="tf32", S_a="tf32", jac_v="fp32", GA_double=False),
  "S fp32 + jac fp32": dict(base, S_v="fp32", S_a="fp32", jac_v="fp32", GA_double=False),
  "S fp32+jac fp32+G fp32": dict(base, S_v="fp32", S_a="fp32", jac_v="fp32", G="fp32"),
  "only G bf16": dict(S_v="fp32", S_a="fp32", G="bf16", dv_a="fp32", jac_v="fp32", da_v="fp32", GA_double=False),
  "only GEMM operands bf16": dict(S_v="fp32", S_a="fp32", G="fp32", dv_a="bf16", jac_v="fp32", da_v="bf16"),
  "S split3 + jac fp32": dict(base, S_split=True, jac_v="fp32", GA_double=False),
}
for n, mode, seed in [(64, "relu", 3), (64, "mix", 31), (128, "mix", 31)]:
    v, a = head_features(n, n, seed, mode)
    o, dv, da = ho.head_forward_backward_torch(v, a, dtype="float64", threads=8)
    print(f"--- n={n} {mode}")
    for name, opt in variants.items():
        lg, dv_e, da_e = emulate(v, a, opt)
        print(f"{name:28s} logits {np.abs(lg-o['logits']).max()/np.abs(o['logits']).max():.2e}  dv {relnorm(dv_e, dv):.2e}  da {relnorm(da_e, da):.2e}")

# ---- fp16 operands (kind::f16 with F16 formats): same tensor-core rate, 11-bit mantissa
print("=== fp16 emulation (G scaled by a power of two)")
def f16(x): return x.to(torch.float16).to(torch.float32)
def emulate16(v, a, scaleG=True, eps=0.65, eps2=0.4, tau=0.03, temp=0.07, ms=100.0, upstream=1.0):
    v = torch.tensor(v); a = torch.tensor(a)
    N, C, P = v.shape
    rn = 1.0 / v.pow(2).sum(1).sqrt().clamp_min(1e-12)
    vh = v * rn[:, None, :]; ah = torch.nn.functional.normalize(a, dim=1)
    v16, a16 = f16(vh), f16(ah)
    S = torch.einsum('ncp,kc->nkp', v16, a16)
    A = torch.einsum('ncp,nc->np', vh, ah)
    W = torch.sigmoid((S - eps) / tau)
    den = W.sum(-1); sim = (W * S).sum(-1) / den
    Pos = torch.sigmoid((A - eps) / tau); Neg = 1 - torch.sigmoid((A - eps2) / tau)
    sim1 = (Pos * A).sum(-1) / Pos.sum(-1); sim2 = (Neg * A).sum(-1) / Neg.sum(-1)
    mask = torch.ones(N, N); mask[torch.arange(N), torch.arange(N)] = 1 - ms
    logits = torch.cat([sim1[:, None], sim * mask, sim2[:, None]], 1) / temp
    sm = torch.softmax(logits, 1); sm[:, 0] -= 1
    g = sm / N * upstream
    gsim = g[:, 1:1 + N] * mask / temp
    def pooled_grad(Wt, dWdA, pooled, gout):
        d = Wt.sum(-1, keepdim=True)
        return gout[:, None] / temp * (Wt + (A - pooled[:, None]) * dWdA) / d
    GA = pooled_grad(Pos, Pos * (1 - Pos) / tau, sim1, g[:, 0]) + pooled_grad(Neg, -(1 - Neg) * Neg / tau, sim2, g[:, -1])
    coef = gsim / den
    G = coef[:, :, None] * W * (1 + (S - sim[:, :, None]) * (1 - W) / tau)
    G[torch.arange(N), torch.arange(N)] += GA
    bound = 17.7 * coef.abs().max() + GA.abs().max()
    s = 2.0 ** torch.floor(torch.log2(32768.0 / bound)) if scaleG else torch.tensor(1.0)
    G16 = f16(G * s)
    assert torch.isfinite(G16).all()
    dvh = torch.einsum('nkp,kc->ncp', G16, a16) / s
    t = (v16 * dvh).sum(1, keepdim=True)
    dv = rn[:, None, :] * (dvh - v16 * t)
    dah = torch.einsum('nkp,ncp->kc', G16, v16) / s
    na = 1.0 / a.pow(2).sum(1).sqrt().clamp_min(1e-12)
    da = na[:, None] * (dah - ah * (ah * dah).sum(1, keepdim=True))
    return logits.numpy(), dv.numpy(), da.numpy(), float(s), float((G16 == 0).float().mean()), float((G == 0).float().mean())

for n, mode, seed in [(64, "relu", 3), (64, "mix", 31), (128, "mix", 31), (128, "relu", 5), (32, "signed", 9)]:
    v, a = head_features(n, n, seed, mode)
    o, dv, da = ho.head_forward_backward_torch(v, a, dtype="float64", threads=8)
    for sc in (True, False):
        lg, dv_e, da_e, s, z16, z = emulate16(v, a, sc)
        print(f"n={n} {mode:6s} scale={s:9.0f} logits {np.abs(lg-o['logits']).max()/np.abs(o['logits']).max():.2e}  dv {relnorm(dv_e, dv):.2e}  da {relnorm(da_e, da):.2e}  zeroG16 {z16:.3f} zeroG {z:.3f}")
