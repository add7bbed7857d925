"""Pin the CPU oracle (oracle/) to the golden vectors produced by the unmodified
reference (tests/golden/make_golden.py).  CPU only."""
import json
import os

import numpy as np
import pytest

from inputs import head_features
from oracle import eval_oracle as eo
from oracle import head_oracle as ho

GOLD = os.path.join(os.path.dirname(__file__), "golden")
HEAD = np.load(os.path.join(GOLD, "head_golden.npz"))
EVAL = np.load(os.path.join(GOLD, "eval_golden.npz"))

HEAD_CASES = [("cfg1_relu", "relu"), ("cfg1_mix", "mix"), ("b32_mix", "mix"),
              ("b16_notrimap", "mix"), ("b16_noneg", "mix"), ("b8_signed", "signed")]


def rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


@pytest.mark.parametrize("name,mode", HEAD_CASES)
def test_head_forward_matches_reference(name, mode):
    n, seed, trimap, neg = HEAD[f"{name}/meta"]
    v, a = head_features(n, n, seed, mode)
    o = ho.head_forward(v, a, trimap=bool(trimap), use_neg=bool(neg), dtype=np.float64)
    # the reference ran in fp32; the fp64 oracle must agree to fp32 round-off
    assert rel(o["A"], HEAD[f"{name}/A"]) < 2e-6
    assert rel(o["logits"], HEAD[f"{name}/logits"]) < 2e-5
    assert rel(o["weighted_A"], HEAD[f"{name}/weighted_A"]) < 2e-5
    assert np.abs(o["Pos"] - HEAD[f"{name}/Pos"]).max() < 2e-5
    assert np.abs(o["Neg"] - HEAD[f"{name}/Neg"]).max() < 2e-5
    assert abs(o["loss"] - HEAD[f"{name}/loss"]) < 1e-4 * max(1.0, abs(o["loss"]))
    assert o["logits"].shape == (n, n + 1 + int(neg))


@pytest.mark.parametrize("name,mode", HEAD_CASES)
def test_head_backward_matches_reference_autograd(name, mode):
    n, seed, trimap, neg = HEAD[f"{name}/meta"]
    v, a = head_features(n, n, seed, mode)
    outs, dv, da = ho.head_forward_backward_torch(v, a, trimap=bool(trimap), use_neg=bool(neg))
    g_dv = HEAD[f"{name}/dv_sub"]
    assert rel(dv[:, ::16, ::7], g_dv) < 5e-4
    assert rel(np.sqrt((dv ** 2).sum((1, 2))), HEAD[f"{name}/dv_norm"]) < 5e-4
    assert rel(da, HEAD[f"{name}/da"]) < 5e-4
    assert rel(outs["logits"], HEAD[f"{name}/logits"]) < 2e-5


def test_numpy_and_torch_oracles_agree():
    v, a = head_features(12, 12, 3, "mix")
    o = ho.head_forward(v, a)
    t, _, _ = ho.head_forward_backward_torch(v, a)
    for k in ("A", "Pos", "Neg", "logits", "loss_rows"):
        assert rel(o[k], t[k]) < 1e-10


def test_tube_mode_matches_reference():
    b, t, seed = HEAD["tube_b2t4/meta"]
    n = b * t
    v, a = head_features(n, b, seed, "mix")
    aud = np.repeat(a, t, axis=0)                    # train_3D.py:129-130
    o = ho.head_forward(v, aud, dtype=np.float64)
    assert rel(o["A"], HEAD["tube_b2t4/A"]) < 2e-6
    assert rel(o["logits"], HEAD["tube_b2t4/logits"]) < 2e-5
    # same-clip duplicate columns are unmasked and equal the positive
    lg = o["logits"]
    assert np.allclose(lg[0, 2:2 + t - 1], lg[0, 0], rtol=1e-12)
    assert lg[0, 1] < -100
    # test-time broadcast: 1 audio row vs T frames == T copies of that row
    ob = ho.head_forward(v[:t], np.repeat(a[:1], t, axis=0), dtype=np.float64)
    assert rel(ob["logits"], HEAD["tube_b2t4/logits_bcast"]) < 2e-5
    assert rel(ob["A"], HEAD["tube_b2t4/A_bcast"]) < 2e-6


# ------------------------------------------------------------------ eval
def _eval_inputs():
    n, seed = EVAL["eval/meta"]
    maps = eo.synthetic_maps(int(n), int(seed))
    maps[5] = 0.25; maps[6] = -0.25; maps[7] = np.round(maps[7], 1)
    boxes = json.loads(str(EVAL["eval/boxes_json"]))
    names = [str(x) for x in EVAL["eval/names"]]
    return maps, [boxes[k] for k in names]


def test_upsample_within_one_ulp_of_cv2():
    maps, _ = _eval_inputs()
    up = eo.upsample_14_to_224(maps[:4])
    g = EVAL["eval/up224_cv2_first4"]
    assert np.abs(up - g).max() <= 1.2e-7
    assert np.abs(eo.upsample_14_to_224(maps[7]) - EVAL["eval/up224_cv2_ties"]).max() <= 1.2e-7


def test_gt_rasterisation_bit_exact():
    _, boxes = _eval_inputs()
    n = len(boxes)
    gts = np.stack([eo.rasterise_gt(b, "vggss")[0] for b in boxes])
    g = np.unpackbits(EVAL["eval/gt_packed"])[: n * 224 * 224].reshape(n, 224, 224)
    assert np.array_equal(gts, g)
    fb = EVAL["gt/flickr_boxes"]
    w, den = eo.rasterise_gt(fb, "flickr")
    assert den == 2 and np.array_equal(w / 2.0, EVAL["gt/flickr"])
    w, den = eo.rasterise_gt(fb, "flickr_frame")
    assert den == 1 and np.array_equal(w.astype(np.float64), EVAL["gt/flickr_frame"])


def test_eval_integer_stage_bit_exact_given_identical_maps():
    """Feed the reference's own cv2 maps: masks and counts must be bit-exact."""
    maps, boxes = _eval_inputs()
    n = len(boxes)
    gmask = np.unpackbits(EVAL["eval/mask_packed"])[: n * 224 * 224].reshape(n, 224, 224)
    for i, up in enumerate(EVAL["eval/up224_cv2_first4"]):
        infer, _, _ = eo.binarise_from_up224(up)
        assert np.array_equal(infer, gmask[i])
        gw, den = eo.rasterise_gt(boxes[i], "vggss")
        iw, uw = eo.ciou_counts(infer, gw, den)
        assert iw == EVAL["eval/inter"][i] and uw == EVAL["eval/union"][i]
    infer, _, _ = eo.binarise_from_up224(EVAL["eval/up224_cv2_ties"])
    assert np.array_equal(infer, gmask[7])


def test_eval_end_to_end_vs_reference():
    maps, boxes = _eval_inputs()
    gts = [eo.rasterise_gt(b, "vggss")[0] for b in boxes]
    r = eo.evaluate_maps(maps, gts, 1)
    g = EVAL["eval/ciou"]
    both_nan = np.isnan(r["ciou"]) & np.isnan(g)
    assert np.all(both_nan | (np.abs(r["ciou"] - g) <= 1e-3))
    assert abs(r["ciou_at_half"] - float(EVAL["eval/ciou_at_half"])) <= 1e-3 + 1.0 / len(g)
    assert abs(r["auc"] - float(EVAL["eval/auc"])) <= 1e-3 + 1.0 / len(g)
    # <= 2 mask pixels may flip per map because cv2's SIMD order differs by <= 1 ulp
    n = len(boxes)
    gmask = np.unpackbits(EVAL["eval/mask_packed"])[: n * 224 * 224].reshape(n, 224, 224)
    for i in range(n):
        m = eo.evaluate_map(maps[i], gts[i])["mask"]
        assert (m != gmask[i]).sum() <= 4
    # degenerate maps: constant >= 0 -> everything selected; constant < 0 -> nothing
    assert eo.evaluate_map(maps[5], gts[5])["mask"].all()
    assert not eo.evaluate_map(maps[6], gts[6])["mask"].any()


def test_auc_formula_and_float_ciou():
    g = EVAL["eval/ciou"]
    counts, results = eo.success_curve(g)
    assert np.array_equal(results, EVAL["eval/results"])
    assert eo.auc_trapezoid(results) == float(EVAL["eval/auc"])
    assert eo.ciou_at_half(g) == float(EVAL["eval/ciou_at_half"])
    n = len(g)
    gmask = np.unpackbits(EVAL["eval/mask_packed"])[: n * 224 * 224].reshape(n, 224, 224)
    ggt = np.unpackbits(EVAL["eval/gt_packed"])[: n * 224 * 224].reshape(n, 224, 224)
    for i in (0, 1, 2, n - 1, n - 2):
        c, inter, union = eo.cal_ciou_float(gmask[i].astype(np.float32), ggt[i].astype(np.float64), 0.5)
        iw, uw = eo.ciou_counts(gmask[i], ggt[i], 1)
        assert (inter, union) == (iw, uw)
        assert (np.isnan(c) and np.isnan(g[i])) or c == g[i]


def test_mtc_matches_reference():
    n = int(EVAL["eval/meta"][0])
    gmask = np.unpackbits(EVAL["eval/mask_packed"])[: n * 224 * 224].reshape(n, 224, 224)
    assert abs(eo.mtc(list(gmask[16:24])) - float(EVAL["eval/mtc_16_24"])) < 1e-6


def test_hardway16_objective_matches_reference():
    """train_hardway.py:132-142 (two AVENet passes, CE + MSE(weighted) + PropagationLoss) from the
    UNMODIFIED reference (tests/golden/make_golden_aux.py) pins the oracle's weighted_A and its
    gradient: one oracle pass per frame set, the other set's weighted map held constant."""
    G = np.load(os.path.join(GOLD, "head_aux_golden.npz"))
    b, t, seed = (int(x) for x in G["meta"])
    lw = float(G["loss_weight"])
    n = b * t
    v1, a = head_features(n, b, seed, "mix")
    v2, _ = head_features(n, b, seed + 1, "mix")
    a_rep = np.repeat(a, t, axis=0)                              # train_hardway.py:128-129
    o1, dv1, da1 = ho.head_forward_backward_torch(v1, a_rep, loss_scale=lw / 2,
                                                  aux=(b, t, 1.0, 100 - lw, G["weighted2"]))
    o2, dv2, da2 = ho.head_forward_backward_torch(v2, a_rep, loss_scale=lw / 2,
                                                  aux=(b, t, 1.0, 100 - lw, G["weighted"]))
    assert rel(o1["weighted_A"], G["weighted"]) < 2e-5 and rel(o2["weighted_A"], G["weighted2"]) < 2e-5
    assert abs(o1["loss"] * lw - G["hardway"]) < 1e-5 and abs(o2["loss"] * lw - G["aug"]) < 1e-5
    assert abs(o1["consistency"] + o2["consistency"] - G["consistency"]) < 1e-6 * max(1, G["consistency"])
    assert abs(o1["l2"] * (100 - lw) - G["l2"]) < 2e-4 * G["l2"]
    assert rel(dv1[:, ::16, ::2], G["dv1_sub"]) < 5e-4 and rel(dv2[:, ::16, ::2], G["dv2_sub"]) < 5e-4
    assert rel(np.sqrt((dv1 ** 2).sum((1, 2))), G["dv1_norm"]) < 5e-4
    da = (da1 + da2).reshape(b, t, -1).sum(1)
    assert rel(da, G["da"]) < 5e-4


def test_per_video_loop_is_the_single_frame_glue_per_video():
    """oracle evaluate_videos (train_hardway.py:158-203) == evaluate_maps / mtc applied per video."""
    maps = eo.synthetic_maps(6, 4).reshape(2, 3, 14, 14)
    gt = [[eo.rasterise_gt([[20 + 10 * f, 30, 150, 200]], "flickr_frame")[0] for f in range(3)] for _ in range(2)]
    o = eo.evaluate_videos(maps, gt, 1)
    for v in range(2):
        e = eo.evaluate_maps(maps[v], gt[v], 1)
        assert np.array_equal(o["ciou"][v], e["ciou"])
        assert o["per_video"][v, 0] == e["ciou_at_half"] and o["per_video"][v, 1] == e["auc"]
        masks = [eo.evaluate_map(m, g, 1)["mask"] for m, g in zip(maps[v], gt[v])]
        assert o["per_video"][v, 2] == eo.mtc(masks)
    assert o["auc_mean"] == o["per_video"][:, 1].sum() / 2


def test_eval_oracle_on_all_real_vggss_boxes():
    """Oracle vs the UNMODIFIED reference on config 3's real data: the GT rasterisation of all 5158
    VGG-SS box lists (pixel sums), and cIoU / AUC of the seeded maps within the 1e-3 that cv2's
    <= 1-ulp bilinear rounding allows (tests/golden/make_golden_vggss.py made the fixture)."""
    g = np.load(os.path.join(GOLD, "vggss_golden.npz"))
    n = len(g["count"])
    boxes = [g["boxes"][i, :g["count"][i]].tolist() for i in range(n)]
    gts = [eo.rasterise_gt(b, "vggss")[0] for b in boxes]
    assert np.array_equal(np.array([int(x.sum()) for x in gts]), g["gt_sum"])
    idx = np.r_[0:n:7, np.nonzero(g["gt_sum"] == 0)[0], np.nonzero(g["count"] >= 4)[0]]
    maps = eo.synthetic_maps(n, 0)
    o = eo.evaluate_maps(maps[idx], [gts[i] for i in idx], 1)
    ok = ~np.isnan(g["ciou"][idx])
    assert np.abs(o["ciou"][ok] - g["ciou"][idx][ok]).max() < 1e-3
    assert np.abs(o["inter_w"] - g["inter"][idx]).max() <= 8
