"""GPU parity of the similarity + hard-way loss head: libavtubes (through the C ABI) against
the golden vectors of the unmodified reference and against the CPU oracle.

Tolerances (BASELINE.json north_star): logits / loss / maps within 2e-3 relative and
gradients within 2e-3 norm-wise for the tensor-core path (fp16 operands, fp32 accumulate); the
fp32 CUDA-core path is held to fp32 round-off."""
import os

import numpy as np
import pytest
import torch

from inputs import head_features
from oracle import head_oracle as ho

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
HEAD = np.load(os.path.join(GOLD, "head_golden.npz"))
HEAD_CASES = [("cfg1_relu", "relu"), ("cfg1_mix", "mix"), ("b32_mix", "mix"),
              ("b16_notrimap", "mix"), ("b16_noneg", "mix"), ("b8_signed", "signed")]
# (max-norm relative tolerance of logits / maps, norm-wise gradient tolerance).  SURVEY 8c states
# 2e-3 for both on the tensor-core path.  Round 1 ran bf16 operands and could only hold gradients
# to 1e-2: a row's gradient is dominated by ONE term (the -99-weighted diagonal column), so the
# 2^-9 rounding of each bf16 factor (a_hat, G, v_hat) shows up un-averaged (measured 3e-3..6e-3;
# scripts/sim_bf16_grad_error.py attributes it).  fp16 operands (2^-11, same tensor-core rate) with a
# power-of-two scale on the cell gradients measure 4e-4..8e-4 and meet the stated 2e-3.
TOL = {"fp32": (2e-5, 2e-4), "f16": (2e-3, 2e-3)}
PRECISIONS = ["fp32", "f16"]


def rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def relnorm(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30)


def run_head(v, a, precision, trimap=True, use_neg=True, pos=None, col_offset=0, grad=False,
             dlogits=None, dA=None, loss_scale=1.0, weighted=True):
    from audio_visual_tubes_b200 import HeadConfig, hardway_head
    cfg = HeadConfig(trimap=trimap, use_neg=use_neg, precision=precision)
    vt = torch.tensor(v, device="cuda", requires_grad=grad)
    at = torch.tensor(a, device="cuda", requires_grad=grad)
    pt = None if pos is None else torch.tensor(pos, device="cuda", dtype=torch.int32)
    A, logits, Pos, Neg, wA, loss_rows = hardway_head(vt, at, cfg, pt, col_offset, return_weighted=weighted)
    out = dict(A=A, logits=logits, Pos=Pos, Neg=Neg, weighted_A=wA, loss_rows=loss_rows)
    res = {k: (None if t is None else t.detach().cpu().numpy()) for k, t in out.items()}
    if grad:
        total = loss_scale * loss_rows.mean()
        if dlogits is not None:
            total = total + (torch.tensor(dlogits, device="cuda") * logits).sum()
        if dA is not None:
            total = total + (torch.tensor(dA, device="cuda") * A).sum()
        total.backward()
        res["dv"] = vt.grad.cpu().numpy()
        res["da"] = at.grad.cpu().numpy()
    torch.cuda.synchronize()
    return res


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name,mode", HEAD_CASES)
def test_forward_matches_reference_golden(name, mode, precision):
    n, seed, trimap, neg = HEAD[f"{name}/meta"]
    v, a = head_features(n, n, seed, mode)
    r = run_head(v, a, precision, bool(trimap), bool(neg))
    tol = TOL[precision][0]
    assert r["logits"].shape == (n, n + 1 + int(neg))
    assert rel(r["A"], HEAD[f"{name}/A"].reshape(n, -1)) < tol
    assert rel(r["logits"], HEAD[f"{name}/logits"]) < tol
    assert rel(r["weighted_A"], HEAD[f"{name}/weighted_A"].reshape(n, -1)) < max(tol, 5e-5)
    # Pos / Neg are sigmoids of (A - eps)/tau: an error dA becomes dA/tau/4 absolute
    ptol = 2e-5 if precision == "fp32" else 2e-3
    assert np.abs(r["Pos"] - HEAD[f"{name}/Pos"].reshape(n, -1)).max() < ptol
    assert np.abs(r["Neg"] - HEAD[f"{name}/Neg"].reshape(n, -1)).max() < ptol
    loss = float(r["loss_rows"].mean())
    assert abs(loss - float(HEAD[f"{name}/loss"])) < tol * max(1.0, abs(loss))


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name,mode", HEAD_CASES)
def test_backward_matches_reference_autograd(name, mode, precision):
    n, seed, trimap, neg = HEAD[f"{name}/meta"]
    v, a = head_features(n, n, seed, mode)
    r = run_head(v, a, precision, bool(trimap), bool(neg), grad=True)
    tol = TOL[precision][1]
    assert relnorm(r["dv"][:, ::16, ::7], HEAD[f"{name}/dv_sub"]) < tol
    assert rel(np.sqrt((r["dv"].astype(np.float64) ** 2).sum((1, 2))), HEAD[f"{name}/dv_norm"]) < tol
    assert relnorm(r["da"], HEAD[f"{name}/da"]) < tol


@pytest.mark.parametrize("precision", PRECISIONS)
def test_backward_with_external_grads_vs_oracle(precision):
    """dlogits and dA upstream gradients (not only the fused CE) against fp64 autograd."""
    n = 24
    v, a = head_features(n, n, 21, "mix")
    rng = np.random.default_rng(5)
    dlog = (rng.standard_normal((n, n + 2)) * 0.05).astype(np.float32)
    dA = (rng.standard_normal((n, 196)) * 0.01).astype(np.float32)
    r = run_head(v, a, precision, grad=True, dlogits=dlog, dA=dA, loss_scale=0.5)
    _, dv, da = ho.head_forward_backward_torch(v, a, dlogits=dlog, dA=dA, loss_scale=0.5)
    tol = TOL[precision][1]
    assert relnorm(r["dv"], dv) < tol
    assert relnorm(r["da"], da) < tol


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("mode", ["relu", "mix"])
def test_config2_b128_vs_oracle(mode, precision):
    """BASELINE config 2 shape (B=128): forward + backward against the oracle."""
    n = 128
    v, a = head_features(n, n, 31, mode)
    r = run_head(v, a, precision, grad=True)
    o, dv, da = ho.head_forward_backward_torch(v, a, dtype="float64", threads=8)
    tol, gtol = TOL[precision]
    assert rel(r["logits"], o["logits"]) < tol
    assert rel(r["A"], o["A"]) < tol
    assert abs(r["loss_rows"].mean() - o["loss"]) < tol * max(1.0, abs(o["loss"]))
    assert relnorm(r["dv"], dv) < gtol
    assert relnorm(r["da"], da) < gtol


@pytest.mark.parametrize("precision", PRECISIONS)
def test_row_shard_of_b1024(precision):
    """Config 5 as one rank sees it: 64 visual rows against all 1024 audio columns, the
    positives starting at column 320 (col_offset).  Must equal rows 320..383 of the full
    problem, which the oracle evaluates with pos = 320 + arange."""
    n_a, n_v, off = 1024, 64, 320
    v_all, a = head_features(n_v, n_a, 41, "mix")
    pos = off + np.arange(n_v)
    # make v's own positives the shifted columns (head_features built them for pos = arange)
    o = ho.head_forward(v_all, a, pos=pos, dtype=np.float64)
    r = run_head(v_all, a, precision, col_offset=off)
    tol = TOL[precision][0]
    assert rel(r["logits"], o["logits"]) < tol
    assert rel(r["A"], o["A"]) < tol
    r2 = run_head(v_all, a, precision, pos=pos)              # explicit pos_idx, same thing
    assert np.array_equal(r["logits"], r2["logits"])


@pytest.mark.parametrize("precision", PRECISIONS)
def test_tube_mode_and_broadcast(precision):
    from audio_visual_tubes_b200 import HardWayAttention
    b, t, seed = HEAD["tube_b2t4/meta"]
    n = b * t
    v, a = head_features(n, b, seed, "mix")
    vid = torch.tensor(v.reshape(b, t, 512, 14, 14).transpose(0, 2, 1, 3, 4).copy(), device="cuda")
    aud = torch.tensor(np.repeat(a, t, axis=0), device="cuda")           # train_3D.py:129-130
    att = HardWayAttention(precision)
    A, logits = att(aud, vid)
    tol = TOL[precision][0]
    assert A.shape == (n, 1, 14, 14) and logits.shape == (n, n + 2)
    assert rel(A.cpu().numpy().reshape(n, -1), HEAD["tube_b2t4/A"].reshape(n, -1)) < tol
    assert rel(logits.cpu().numpy(), HEAD["tube_b2t4/logits"]) < tol
    # de-duplicated tube mode: the b DISTINCT embeddings instead of T copies of each -- same
    # outputs as the unmodified reference's golden vectors, 1/T of the similarity work
    A2, l2 = att(torch.tensor(a, device="cuda"), vid)
    assert A2.shape == (n, 1, 14, 14) and l2.shape == (n, n + 2)
    assert rel(A2.cpu().numpy().reshape(n, -1), HEAD["tube_b2t4/A"].reshape(n, -1)) < tol
    assert rel(l2.cpu().numpy(), HEAD["tube_b2t4/logits"]) < tol
    # test time: one audio row against T frames (train_3D.py:153-154)
    A1, l1 = att(torch.tensor(a[:1], device="cuda"), vid[:1])
    assert rel(l1.cpu().numpy(), HEAD["tube_b2t4/logits_bcast"]) < tol
    assert rel(A1.cpu().numpy().reshape(t, -1), HEAD["tube_b2t4/A_bcast"].reshape(t, -1)) < tol


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("b,t", [(3, 4), (5, 16), (9, 2)])
def test_tube_dedup_forward_backward(b, t, precision):
    """col_mult = T: distinct audio rows against (b T) frames == the head on T copies of every row
    (train_3D.py:129-130, model.py:47-59): logits, per-row loss, dv, and the audio gradient summed
    over the copies -- through logits, through A and through the fused per-row loss."""
    from audio_visual_tubes_b200 import HeadConfig, hardway_head, hardway_loss
    n = b * t
    v, a = head_features(n, b, 40 + b, "mix")
    rng = np.random.default_rng(b * 100 + t)
    dlog = (rng.standard_normal((n, n + 2)) * 1e-2).astype(np.float32)
    dA = (rng.standard_normal((n, 196)) * 1e-3).astype(np.float32)
    a_full = np.repeat(a, t, axis=0)
    o, dv_o, da_o = ho.head_forward_backward_torch(v, a_full, dlogits=dlog, dA=dA, loss_scale=0.7)
    da_o = da_o.reshape(b, t, -1).sum(1)
    tol, gtol = TOL[precision]
    cfg = HeadConfig(precision=precision)
    vt = torch.tensor(v, device="cuda", requires_grad=True)
    at = torch.tensor(a, device="cuda", requires_grad=True)
    A, logits, Pos, Neg, _, loss_rows = hardway_head(vt, at, cfg, col_mult=t)
    assert logits.shape == (n, n + 2)
    assert rel(logits.detach().cpu().numpy(), o["logits"]) < tol
    assert rel(A.detach().cpu().numpy(), o["A"]) < tol
    assert rel(loss_rows.detach().cpu().numpy(), o["loss_rows"]) < 5 * tol
    total = 0.7 * loss_rows.mean() + (torch.tensor(dlog, device="cuda") * logits).sum() + \
        (torch.tensor(dA, device="cuda") * A).sum()
    total.backward()
    assert relnorm(vt.grad.cpu().numpy(), dv_o) < gtol
    assert relnorm(at.grad.cpu().numpy(), da_o) < gtol
    # fused loss path: the kernels see only dloss_scale
    o2, dv2, da2 = ho.head_forward_backward_torch(v, a_full)
    v2 = torch.tensor(v, device="cuda", requires_grad=True)
    a2 = torch.tensor(a, device="cuda", requires_grad=True)
    loss = hardway_loss(v2, a2, cfg, col_mult=t)
    loss.backward()
    assert abs(float(loss.detach()) - o2["loss"]) < 5 * tol * max(1.0, abs(o2["loss"]))
    assert relnorm(v2.grad.cpu().numpy(), dv2) < gtol
    assert relnorm(a2.grad.cpu().numpy(), da2.reshape(b, t, -1).sum(1)) < gtol


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("n_v,n_a", [(1, 1), (3, 5), (7, 130), (33, 17), (2, 257)])
def test_ragged_shapes(n_v, n_a, precision):
    v, a = head_features(n_v, n_a, 51, "mix")
    pos = np.arange(n_v) % n_a
    o = ho.head_forward(v, a, pos=pos, dtype=np.float64)
    r = run_head(v, a, precision, pos=pos)
    assert rel(r["logits"], o["logits"]) < TOL[precision][0]


@pytest.mark.parametrize("precision", PRECISIONS)
def test_anticorrelated_features_take_the_exact_path(precision):
    """Every similarity negative (v = -|.|, a = +|.|): all tri-map weights fall below 2^-30,
    the regime where the shared-reciprocal epilogue must hand over to the exact one
    (weights of a ratio need RELATIVE accuracy)."""
    n = 16
    v, a = head_features(n, n, 71, "relu")
    v = -v
    v[: n // 2] = np.abs(v[: n // 2]) * np.where(np.arange(196) % 2 == 0, 1.0, -1.0).astype(np.float32)
    o, dv, da = ho.head_forward_backward_torch(v, a)
    r = run_head(v, a, precision, grad=True)
    tol, gtol = TOL[precision]
    assert np.isfinite(r["logits"]).all()
    assert rel(r["logits"], o["logits"]) < max(tol, 5e-5)
    assert abs(r["loss_rows"].mean() - o["loss"]) < max(tol, 5e-5) * max(1.0, abs(o["loss"]))
    assert relnorm(r["dv"], dv) < gtol
    assert relnorm(r["da"], da) < gtol


def test_fused_loss_matches_head_and_is_linear_in_scale():
    from audio_visual_tubes_b200 import HeadConfig, hardway_loss
    v, a = head_features(16, 16, 61, "mix")
    cfg = HeadConfig(precision="fp32")
    vt = torch.tensor(v, device="cuda", requires_grad=True)
    at = torch.tensor(a, device="cuda", requires_grad=True)
    loss = hardway_loss(vt, at, cfg)
    (3.0 * loss).backward()
    o, dv, da = ho.head_forward_backward_torch(v, a, loss_scale=3.0)
    assert abs(float(loss) - o["loss"]) < 1e-5 * max(1, abs(o["loss"]))
    assert relnorm(vt.grad.cpu().numpy(), dv) < 2e-4
    assert relnorm(at.grad.cpu().numpy(), da) < 2e-4


def test_graph_captured_step_equals_eager():
    """HeadStep replays one captured CUDA graph; its outputs must equal the eager calls bit for
    bit (same kernels, same order) and track new inputs written into the captured buffers."""
    from audio_visual_tubes_b200 import HeadConfig, HeadStep, hardway_value_and_grad
    v, a = head_features(24, 24, 81, "mix")
    vt, at = torch.tensor(v, device="cuda"), torch.tensor(a, device="cuda")
    cfg = HeadConfig()
    hs = HeadStep(vt, at, cfg)
    for seed in (81, 82):
        v2, a2 = head_features(24, 24, seed, "mix")
        vt.copy_(torch.tensor(v2)); at.copy_(torch.tensor(a2))
        l0, dv0, da0 = hardway_value_and_grad(vt, at, cfg)
        l1, dv1, da1 = hs()
        torch.cuda.synchronize()
        assert torch.equal(l0, l1) and torch.equal(dv0, dv1)
        # d_ahat is reduced with floating-point atomics: equal up to summation order
        assert relnorm(da1.cpu().numpy(), da0.cpu().numpy()) < 1e-5


def test_host_pipeline_tracks_each_steps_inputs():
    """HostHeadPipeline overlaps the H2D of step i+1 with the compute of step i on two slots:
    every handle must report the loss and gradients of ITS OWN host inputs, in any read order."""
    from audio_visual_tubes_b200 import HeadConfig, HostHeadPipeline, hardway_value_and_grad
    cfg = HeadConfig()
    feats = [head_features(24, 24, s, "mix") for s in (91, 92, 93, 94, 95)]
    host = [(torch.tensor(v).pin_memory(), torch.tensor(a).pin_memory()) for v, a in feats]
    want = []
    for v, a in feats:
        l, dv, da = hardway_value_and_grad(torch.tensor(v, device="cuda"), torch.tensor(a, device="cuda"), cfg)
        want.append((float(l.cpu()[0]), dv.cpu(), da.cpu()))
    pipe = HostHeadPipeline(feats[0][0].shape, feats[0][1].shape, cfg)
    pend, got = None, []
    for vh, ah in host:
        h = pipe.submit(vh, ah)
        if pend is not None:
            dv, da = pend.grads()
            got.append((pend.loss(), dv.cpu(), da.cpu()))
        pend = h
    dv, da = pend.grads()
    got.append((pend.loss(), dv.cpu(), da.cpu()))
    assert len({w[0] for w in want}) == len(want)           # the five steps really differ
    for (l0, dv0, da0), (l1, dv1, da1) in zip(want, got):
        assert l0 == l1 and torch.equal(dv0, dv1)
        assert relnorm(da1.numpy(), da0.numpy()) < 1e-5


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("n,mode", [(12, "mix"), (8, "relu"), (5, "signed")])
def test_weighted_map_gradient_vs_oracle(n, mode, precision):
    """avt_head_backward's dWeighted (model.py:148-152 back-propagated as train_hardway.py:138-142
    does): <dW, weighted_A> alone, and together with the loss, dlogits and dA."""
    from audio_visual_tubes_b200 import HeadConfig, hardway_head
    v, a = head_features(n, n, 70 + n, mode)
    rng = np.random.default_rng(n)
    dW = rng.standard_normal((n, 196)).astype(np.float32)
    dA = (0.01 * rng.standard_normal((n, 196))).astype(np.float32)
    cfg = HeadConfig(precision=precision)
    for scale, use_dA in ((0.0, False), (1.0, True)):
        o, dv, da = ho.head_forward_backward_torch(v, a, loss_scale=scale, dW=dW, dA=dA if use_dA else None)
        vt = torch.tensor(v, device="cuda", requires_grad=True)
        at = torch.tensor(a, device="cuda", requires_grad=True)
        A, logits, Pos, Neg, wA, loss_rows = hardway_head(vt, at, cfg, return_weighted=True)
        total = (torch.tensor(dW, device="cuda") * wA).sum()
        if scale:
            total = total + scale * loss_rows.mean() + (torch.tensor(dA, device="cuda") * A).sum()
        total.backward()
        assert rel(wA.detach().cpu().numpy(), o["weighted_A"]) < TOL[precision][0]
        # the channel-uniform part of dv is exact fp32 in both precisions; the part through
        # Pos -> A rides the fp16 contraction like every other map gradient
        assert relnorm(vt.grad.cpu().numpy(), dv) < TOL[precision][1], (scale, relnorm(vt.grad.cpu().numpy(), dv))
        assert relnorm(at.grad.cpu().numpy(), da) < TOL[precision][1], (scale, relnorm(at.grad.cpu().numpy(), da))


@pytest.mark.parametrize("precision", PRECISIONS)
def test_hardway16_objective_matches_reference_golden(precision):
    """The 16-frame training objective (train_hardway.py:132-142) through the product: two head
    passes with return_weighted, the losses.py mirror, one backward -- against the UNMODIFIED
    reference's values and gradients (tests/golden/make_golden_aux.py)."""
    from audio_visual_tubes_b200 import HeadConfig, hardway_head, hardway16_objective
    G = np.load(os.path.join(GOLD, "head_aux_golden.npz"))
    b, t, seed = (int(x) for x in G["meta"])
    lw = float(G["loss_weight"])
    n = b * t
    v1, a = head_features(n, b, seed, "mix")
    v2, _ = head_features(n, b, seed + 1, "mix")
    cfg = HeadConfig(precision=precision)
    f1 = torch.tensor(v1, device="cuda", requires_grad=True)
    f2 = torch.tensor(v2, device="cuda", requires_grad=True)
    spec = torch.tensor(a, device="cuda", requires_grad=True)
    spec_rep = spec.repeat_interleave(t, dim=0)                  # train_hardway.py:128-129
    out1 = hardway_head(f1, spec_rep, cfg, return_weighted=True)
    out2 = hardway_head(f2, spec_rep, cfg, return_weighted=True)
    combined, hardway, aug, l2, cons = hardway16_objective(out1, out2, b, t, lw)
    combined.backward()
    tol_f, tol_g = TOL[precision]
    assert rel(out1[4].detach().cpu().numpy(), G["weighted"]) < max(tol_f, 2e-5)
    assert abs(float(hardway) - G["hardway"]) < tol_f and abs(float(aug) - G["aug"]) < tol_f
    assert abs(float(cons) - G["consistency"]) < tol_f * G["consistency"] + 1e-7
    assert abs(float(l2) - G["l2"]) < 2 * tol_f * G["l2"] + 1e-8
    assert abs(float(combined) - G["combined"]) < tol_f
    dv1 = f1.grad.cpu().numpy()
    assert relnorm(dv1[:, ::16, ::2], G["dv1_sub"]) < tol_g
    assert relnorm(f2.grad.cpu().numpy()[:, ::16, ::2], G["dv2_sub"]) < tol_g
    assert relnorm(np.sqrt((dv1.astype(np.float64) ** 2).sum((1, 2))), G["dv1_norm"]) < tol_g
    assert relnorm(spec.grad.cpu().numpy(), G["da"]) < tol_g


def test_errors_are_loud():
    from audio_visual_tubes_b200 import HeadConfig, hardway_head
    v, a = head_features(2, 2, 1, "relu")
    with pytest.raises(RuntimeError):                       # CPU tensors: no fallback
        hardway_head(torch.tensor(v), torch.tensor(a), HeadConfig())
    with pytest.raises(RuntimeError):                       # C not a multiple of 64
        hardway_head(torch.zeros(2, 100, 196, device="cuda"), torch.zeros(2, 100, device="cuda"),
                     HeadConfig(precision="fp32"))


def test_avenet_contract():
    """AVENet(args, pretrained).forward(image, audio) -> (A, logits, Pos, Neg), model.py:112-154."""
    import types
    from audio_visual_tubes_b200 import AVENet
    args = types.SimpleNamespace(epsilon=0.65, epsilon2=0.4, tri_map=True, Neg=True)
    torch.manual_seed(0)
    net = AVENet(args, False).cuda().eval()
    img = torch.randn(4, 3, 224, 224, device="cuda")
    aud = torch.randn(4, 1, 257, 300, device="cuda")
    with torch.no_grad():
        A, logits, Pos, Neg = net(img, aud)
        fi, fa = net.features(img, aud)
    assert A.shape == (4, 1, 14, 14) and logits.shape == (4, 6)
    assert Pos.shape == (4, 1, 14, 14) and Neg.shape == (4, 1, 14, 14)
    o = ho.head_forward(fi.flatten(2).cpu().numpy(), fa.cpu().numpy(), dtype=np.float64)
    assert rel(logits.cpu().numpy(), o["logits"]) < 2e-3
    assert rel(A.cpu().numpy().reshape(4, -1), o["A"]) < 2e-3
    net.train()
    loss = net.loss(img, aud)
    loss.backward()
    assert torch.isfinite(loss) and net.imgnet.conv1.weight.grad is not None


# ---------------------------------------------------------------------------------------------
# The shapes bench.py measures ("the benchmarked shape has no backward parity").
def _run_value_and_grad(v, a, precision, **kw):
    from audio_visual_tubes_b200 import HeadConfig, hardway_value_and_grad
    vt = torch.tensor(v, device="cuda"); at = torch.tensor(a, device="cuda")
    loss, dv, da = hardway_value_and_grad(vt, at, HeadConfig(precision=precision), **kw)
    torch.cuda.synchronize()
    return float(loss[0]), dv.cpu().numpy(), da.cpu().numpy()


@pytest.fixture(scope="module")
def cfg5_case():
    """Config 5 at full size, N_v = N_a = 1024, bench.py's |randn| features; fp64 autograd oracle in
    row blocks (rows are independent given all audio columns; da accumulates across blocks)."""
    rng = np.random.default_rng(1234)
    v = np.abs(rng.standard_normal((1024, 512, 196), dtype=np.float32))
    a = np.abs(rng.standard_normal((1024, 512), dtype=np.float32))
    o, dv, da = ho.head_forward_backward_blocks(v, a, row_block=64)
    return v, a, o, dv, da


@pytest.mark.parametrize("precision", PRECISIONS)
def test_cfg5_b1024_forward_backward(cfg5_case, precision):
    """n_a_pad = 1024: eight audio tiles (a_full / a_empty reloads), cta_group::2 d_ahat kernel with
    18 K splits, dv contraction with K = 1024 -- the kernels of the headline bench, against the
    fp64 oracle: logits / loss / A at 2e-3, gradients at 2e-3 norm-wise."""
    v, a, o, dv, da = cfg5_case
    from audio_visual_tubes_b200 import HeadConfig
    from audio_visual_tubes_b200.head import audio_normalize, head_forward_raw
    cfg = HeadConfig(precision=precision)
    tol, gtol = TOL[precision]
    vt = torch.tensor(v, device="cuda"); at = torch.tensor(a, device="cuda")
    out = head_forward_raw(vt, audio_normalize(at)[0], cfg)
    assert rel(out["logits"].cpu().numpy(), o["logits"]) < tol
    assert rel(out["A"].cpu().numpy(), o["A"]) < tol
    assert rel(out["loss_rows"].cpu().numpy(), o["loss_rows"]) < 5 * tol
    loss, dv_g, da_g = _run_value_and_grad(v, a, precision)
    assert abs(loss - o["loss"]) < tol * max(1.0, abs(o["loss"]))
    e_dv, e_da = relnorm(dv_g, dv), relnorm(da_g, da)
    print(f"cfg5 B=1024 [{precision}] loss {loss:.6f} vs {o['loss']:.6f}; norm-wise grad err dv {e_dv:.2e} da {e_da:.2e}")
    assert e_dv < gtol and e_da < gtol
    # per-row gradient norms too (a wrong tile / split shows up in SOME rows, not in the norm)
    rn = np.sqrt((dv_g.astype(np.float64) ** 2).sum((1, 2))); rn_o = np.sqrt((dv ** 2).sum((1, 2)))
    assert np.abs(rn - rn_o).max() / rn_o.max() < gtol


@pytest.mark.parametrize("precision", PRECISIONS)
def test_cfg4_tube_64x16_both_modes(precision):
    """Config 4: b = 64 clips x T = 16 frames = 1024 frames.  Reference-faithful mode (every
    spectrogram repeated 16 times, train_3D.py:129-130: 1024 audio rows, 15 unmasked duplicate
    columns per frame) and the de-duplicated tube mode (64 distinct rows, col_mult = 16) against the
    same fp64 oracle of the repeated problem, forward and backward."""
    b, t = 64, 16
    v, a = head_features(b * t, b, 77, "mix")
    a_rep = np.repeat(a, t, axis=0)
    o, dv, da_rep = ho.head_forward_backward_blocks(v, a_rep, row_block=64)
    da = da_rep.reshape(b, t, -1).sum(1)
    tol, gtol = TOL[precision]
    l1, dv1, da1 = _run_value_and_grad(v, a_rep, precision)
    l2, dv2, da2 = _run_value_and_grad(v, a, precision, col_mult=t)
    print(f"cfg4 [{precision}] repeated: dv {relnorm(dv1, dv):.2e} da {relnorm(da1, da_rep):.2e}; "
          f"tube: dv {relnorm(dv2, dv):.2e} da {relnorm(da2, da):.2e}")
    for l in (l1, l2):
        assert abs(l - o["loss"]) < tol * max(1.0, abs(o["loss"]))
    assert relnorm(dv1, dv) < gtol and relnorm(da1, da_rep) < gtol
    assert relnorm(dv2, dv) < gtol and relnorm(da2, da) < gtol


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("n_v,n_a,off", [(64, 1024, 320), (7, 130, 60), (2, 257, 255), (33, 300, 0)])
def test_backward_ragged_shapes_with_col_offset(n_v, n_a, off, precision):
    """Backward with col_offset != 0 (mask column and the positive-column fix-up), ragged audio
    counts (pad rows of the last 128 / 256-row tile) and multi-tile n_a_pad: what a rank of the
    sharded step runs."""
    v, a = head_features(n_v, n_a, 100 + n_v, "mix")
    pos = off + np.arange(n_v)
    o, dv, da = ho.head_forward_backward_torch(v, a, pos=pos)
    tol, gtol = TOL[precision]
    r = run_head(v, a, precision, col_offset=off, grad=True)
    assert rel(r["logits"], o["logits"]) < tol
    assert relnorm(r["dv"], dv) < gtol, relnorm(r["dv"], dv)
    assert relnorm(r["da"], da) < gtol, relnorm(r["da"], da)


@pytest.mark.parametrize("precision", PRECISIONS)
def test_loss_rows_upstream_gradient_in_kernel(precision):
    """A non-uniform gradient into loss_rows (<w, loss_rows>) is applied inside row_grad_kernel
    (dloss_rows of the C ABI), not by an eager softmax on the host side."""
    n = 40
    v, a = head_features(n, n, 8, "mix")
    w = np.random.default_rng(2).uniform(-0.5, 1.5, n).astype(np.float32)
    _, dv, da = ho.head_forward_backward_torch(v, a, loss_scale=0.0, row_weights=w)
    from audio_visual_tubes_b200 import HeadConfig, hardway_head
    vt = torch.tensor(v, device="cuda", requires_grad=True)
    at = torch.tensor(a, device="cuda", requires_grad=True)
    out = hardway_head(vt, at, HeadConfig(precision=precision))
    (torch.tensor(w, device="cuda") * out[5]).sum().backward()
    gtol = TOL[precision][1]
    assert relnorm(vt.grad.cpu().numpy(), dv) < gtol
    assert relnorm(at.grad.cpu().numpy(), da) < gtol


def test_two_host_threads_one_device():
    """The nn.DataParallel caller of SURVEY 8b: two host threads, each on its own stream of the
    same device, run forward + backward concurrently (the side stream / events of the backward are
    per caller stream) and both get the single-threaded result bit for bit."""
    import threading
    from audio_visual_tubes_b200 import HeadConfig, hardway_value_and_grad
    cfg = HeadConfig(precision="f16")
    data = []
    for seed in (3, 4):
        v, a = head_features(48, 300, seed, "mix")
        vt = torch.tensor(v, device="cuda"); at = torch.tensor(a, device="cuda")
        pos = torch.arange(48, device="cuda", dtype=torch.int32) + 17
        ref = [hardway_value_and_grad(vt, at, cfg, pos_idx=pos)[1].clone() for _ in range(1)][0]
        data.append((vt, at, pos, ref))
    torch.cuda.synchronize()
    bad = []

    def work(i):
        vt, at, pos, ref = data[i]
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for _ in range(40):
                dv = hardway_value_and_grad(vt, at, cfg, pos_idx=pos)[1]
                if not torch.equal(dv, ref):
                    bad.append(i)
        s.synchronize()
    th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    [t.start() for t in th]; [t.join() for t in th]
    assert not bad


def test_fused_row_gradients_equal_separate_kernel():
    """The fused training step (row gradients out of finish_rows_kernel, AVT_BWD_ROWGRAD_DONE) gives
    the bits of the two-kernel path (row_grad_kernel after the forward), tube mode included."""
    from audio_visual_tubes_b200 import HeadConfig, hardway_loss, hardway_value_and_grad
    for prec in PRECISIONS:
        for (nv, na, t) in [(40, 40, 1), (48, 6, 8), (130, 130, 1)]:
            v, a = head_features(nv, na, 5 + nv, "mix")
            cfg = HeadConfig(precision=prec)
            vt = torch.tensor(v, device="cuda", requires_grad=True)
            at = torch.tensor(a, device="cuda", requires_grad=True)
            loss = hardway_loss(vt, at, cfg, col_mult=t)                  # autograd: separate row_grad_kernel
            loss.backward()
            l2, dv2, da2 = hardway_value_and_grad(vt.detach(), at.detach(), cfg, col_mult=t)   # fused
            assert float(loss) == float(l2[0])
            if prec == "f16":
                assert torch.equal(vt.grad, dv2)
            else:
                assert relnorm(dv2.cpu().numpy(), vt.grad.cpu().numpy()) < 1e-6
            assert relnorm(da2.cpu().numpy(), at.grad.cpu().numpy()) < 1e-5   # d_ahat: atomics order


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_half_precision_feature_maps(dtype):
    """fp16 / bf16 visual maps (a backbone under autocast; half the H2D bytes of the host pipeline)
    give exactly what the fp32 entry gives on the same rounded values, and gradients flow back in
    the maps' dtype."""
    from audio_visual_tubes_b200 import HeadConfig, hardway_head, hardway_value_and_grad
    v, a = head_features(24, 24, 19, "mix")
    cfg = HeadConfig(precision="f16")
    vh = torch.tensor(v, device="cuda").to(dtype)
    at = torch.tensor(a, device="cuda")
    l_h, dv_h, da_h = hardway_value_and_grad(vh, at, cfg)
    l_f, dv_f, da_f = hardway_value_and_grad(vh.float(), at, cfg)
    assert float(l_h[0]) == float(l_f[0]) and torch.equal(dv_h, dv_f) and dv_h.dtype == torch.float32
    vg = vh.clone().requires_grad_(True)
    out = hardway_head(vg, at, cfg)
    out[5].mean().backward()
    assert vg.grad.dtype == dtype
    # the CUDA-core fp32 path takes them too, by an explicit up-cast on the host side
    lg32 = hardway_head(vh, at, HeadConfig(precision="fp32"))[1]
    assert rel(out[1].detach().cpu().numpy(), lg32.cpu().numpy()) < 2e-3


def test_head_step_graph_equals_eager():
    from audio_visual_tubes_b200 import HeadConfig, HeadStep, hardway_value_and_grad
    v, a = head_features(96, 96, 23, "mix")
    vt = torch.tensor(v, device="cuda"); at = torch.tensor(a, device="cuda")
    cfg = HeadConfig(precision="f16")
    l0, dv0, da0 = hardway_value_and_grad(vt, at, cfg)
    hs = HeadStep(vt, at, cfg)
    hs()
    l1, dv1, da1 = hs()
    torch.cuda.synchronize()
    assert float(l0[0]) == float(l1[0]) and torch.equal(dv0, dv1)
    assert relnorm(da1.cpu().numpy(), da0.cpu().numpy()) < 1e-5
