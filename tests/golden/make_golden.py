"""Generate the golden fixtures by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

It imports /root/reference/model.py and /root/reference/utils.py as they are (with
the one CPU shim SURVEY.md section 8c describes: ``Tensor.cuda`` -> identity, because
model.py:115 hard-codes ``.cuda()``), swaps the backbones for ``nn.Identity()`` so
raw features are the inputs, and stores the reference's outputs on the seeded inputs
of ``tests/golden/inputs.py``.  Only outputs are stored; inputs are regenerated from
the seed.  The eval fixtures use the real cv2.resize / sklearn.metrics.auc and the
real VGG-SS boxes from /root/reference/metadata/vggss.json (a 96-clip subset of the
box lists is copied into the fixture as *data*).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
REF = "/root/reference"


def load_reference():
    import torch
    torch.Tensor.cuda = lambda self, *a, **k: self          # CPU shim (model.py:115,48-51)
    sys.path.insert(0, REF)
    import model as ref_model                                # noqa: E402
    import utils as ref_utils                                # noqa: E402
    return torch, ref_model, ref_utils


def head_case(torch, ref_model, name, n, seed, mode, trimap=True, neg=True):
    import torch.nn as nn
    from inputs import head_features, audio_maps_from_pooled
    v, a = head_features(n, n, seed, mode)
    amap = audio_maps_from_pooled(a, seed)
    args = types.SimpleNamespace(epsilon=0.65, epsilon2=0.4, tri_map=trimap, Neg=neg)
    net = ref_model.AVENet(args, False)
    net.imgnet = nn.Identity()
    net.audnet = nn.Identity()
    vt = torch.tensor(v.reshape(n, 512, 14, 14), requires_grad=True)
    at = torch.tensor(amap, requires_grad=True)
    A, logits, wA, Pos, Neg = net(vt, at)
    loss = nn.CrossEntropyLoss()(logits, torch.zeros(n).long())
    loss.backward()
    dv = vt.grad.numpy().reshape(n, 512, 196)
    da = at.grad.numpy().reshape(n, 512, -1).sum(-1)          # max-pool routes grad to one cell
    return {
        f"{name}/A": A.detach().numpy().reshape(n, 196),
        f"{name}/logits": logits.detach().numpy(),
        f"{name}/weighted_A": wA.detach().numpy().reshape(n, 196),
        f"{name}/Pos": Pos.detach().numpy().reshape(n, 196),
        f"{name}/Neg": Neg.detach().numpy().reshape(n, 196),
        f"{name}/loss": np.float32(loss.item()),
        f"{name}/dv_sub": dv[:, ::16, ::7].copy(),
        f"{name}/dv_norm": np.sqrt((dv.astype(np.float64) ** 2).sum((1, 2))),
        f"{name}/da": da,
        f"{name}/meta": np.array([n, seed, int(trimap), int(neg)]),
    }


def tube_case(torch, ref_model, name, b, t, seed, mode):
    from inputs import head_features
    n = b * t
    v, a = head_features(n, b, seed, mode)
    # reference tube layout: video_features [b,512,T,14,14]; audio repeated T times
    vid = v.reshape(b, t, 512, 14, 14).transpose(0, 2, 1, 3, 4).copy()
    aud = np.repeat(a, t, axis=0)
    # FullModel L2-normalises both before the attention (model.py:31,35)
    vt = torch.nn.functional.normalize(torch.tensor(vid), dim=1)
    at = torch.nn.functional.normalize(torch.tensor(aud), dim=1)
    att = ref_model.HardWayAttention()
    A, logits = att(at, vt)
    # test-time broadcast: ONE audio row against T frames (train_3D.py:153-154)
    A1, logits1 = att(at[:1], vt[:1])
    return {
        f"{name}/A": A.numpy().reshape(n, 196),
        f"{name}/logits": logits.numpy(),
        f"{name}/A_bcast": A1.numpy().reshape(t, 196),
        f"{name}/logits_bcast": logits1.numpy(),
        f"{name}/meta": np.array([b, t, seed]),
    }


def eval_case(ref_utils, n_maps=96, seed=0):
    import cv2
    from sklearn.metrics import auc as sk_auc
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    from oracle.eval_oracle import synthetic_maps
    maps = synthetic_maps(n_maps, seed)
    maps[5] = 0.25                      # constant map, h >= 0: normalisation skipped, pred >= 1
    maps[6] = -0.25                     # constant map, h < 0: empty mask
    maps[7] = np.round(maps[7], 1)      # heavy ties around the median
    meta = json.load(open(os.path.join(REF, "metadata", "vggss.json")))
    # pick a subset that contains the interesting entries: multi-box, negative coords, empty GT
    idx = list(range(n_maps - 12))
    multi = [i for i, m in enumerate(meta) if len(m["bbox"]) >= 4][:4]
    negc = [i for i, m in enumerate(meta) if any(x < 0 for b in m["bbox"] for x in b)][:4]
    idx += multi + negc
    args = types.SimpleNamespace(testset="vggss", gt_all={})
    names = []
    for j, i in enumerate(idx):
        names.append(meta[i]["file"])
        args.gt_all[meta[i]["file"]] = meta[i]["bbox"]
    # find clips whose GT rasterises to empty and append up to 4 of them
    empties = []
    for i, m in enumerate(meta):
        a2 = types.SimpleNamespace(testset="vggss", gt_all={m["file"]: m["bbox"]})
        if ref_utils.testset_gt(a2, m["file"] + ".mp4").sum() == 0:
            empties.append(i)
    for i in empties[:4]:
        names.append(meta[i]["file"])
        args.gt_all[meta[i]["file"]] = meta[i]["bbox"]
    j = n_maps
    while len(names) < n_maps:          # top up with further ordinary clips
        if meta[j]["file"] not in args.gt_all:
            names.append(meta[j]["file"])
            args.gt_all[meta[j]["file"]] = meta[j]["bbox"]
        j += 1
    names = names[:n_maps]
    cious, inters, unions, pops, ups, gts, masks = [], [], [], [], [], [], []
    for i in range(n_maps):
        # ---- the reference glue, train_hardway_1frame.py:198-207, line for line ----
        heatmap_now = cv2.resize(maps[i], dsize=(224, 224), interpolation=cv2.INTER_LINEAR)
        ups.append(heatmap_now.copy())
        heatmap_now = ref_utils.normalize_img(-heatmap_now)
        gt_map = ref_utils.testset_gt(args, names[i] + ".mp4")
        pred = 1 - heatmap_now
        threshold = np.sort(pred.flatten())[int(pred.shape[0] * pred.shape[1] / 2)]
        pred[pred > threshold] = 1
        pred[pred < 1] = 0
        evaluator = ref_utils.Evaluator()
        with np.errstate(divide="ignore", invalid="ignore"):
            ciou, inter, union = evaluator.cal_CIOU(pred, gt_map, 0.5)
        cious.append(ciou); inters.append(inter); unions.append(union)
        pops.append(int((pred >= 0.5).sum()))
        gts.append(gt_map.astype(np.uint8)); masks.append((pred >= 0.5).astype(np.uint8))
    results = []
    for i in range(21):
        result = np.sum(np.array(cious) >= 0.05 * i)
        results.append(result / len(cious))
    x = [0.05 * i for i in range(21)]
    auc_ = sk_auc(x, results)
    ciou_half = np.sum(np.array(cious) >= 0.5) / len(cious)
    import torch
    mtc = float(ref_utils.mTC([m.astype(np.float64) for m in masks[16:24]], [None] * 8))
    return {
        "eval/names": np.array(names),
        "eval/boxes_json": np.array(json.dumps({n: args.gt_all[n] for n in names})),
        "eval/ciou": np.array(cious, dtype=np.float64),
        "eval/inter": np.array(inters, dtype=np.float64),
        "eval/union": np.array(unions, dtype=np.float64),
        "eval/popcount": np.array(pops),
        "eval/results": np.array(results),
        "eval/auc": np.float64(auc_),
        "eval/ciou_at_half": np.float64(ciou_half),
        "eval/up224_cv2_first4": np.stack(ups[:4]),
        "eval/up224_cv2_ties": ups[7],
        "eval/gt_packed": np.packbits(np.stack(gts), axis=None),
        "eval/mask_packed": np.packbits(np.stack(masks), axis=None),
        "eval/mtc_16_24": np.float64(mtc),
        "eval/meta": np.array([n_maps, seed]),
    }


def flickr_gt_case(ref_utils):
    """testset_gt / testset_gt_frame in 'flickr' mode need XML files; write tiny ones."""
    import tempfile
    out = {}
    d = tempfile.mkdtemp()
    boxes = [[20, 30, 200, 180], [40, 60, 220, 256], [0, 0, 90, 100]]
    xml = "<annotation>" + "".join(
        f"<object><bbox><i>{j}</i><a>{b[0]}</a><b>{b[1]}</b><c>{b[2]}</c><d>{b[3]}</d></bbox></object>"
        for j, b in enumerate(boxes)) + "</annotation>"
    open(os.path.join(d, "clip.xml"), "w").write(xml)
    open(os.path.join(d, "clip_3.xml"), "w").write(xml)
    args = types.SimpleNamespace(testset="flickr", og_gt_path=d + "/")
    out["gt/flickr_boxes"] = np.array(boxes)
    out["gt/flickr"] = ref_utils.testset_gt(args, "clip.jpg")
    out["gt/flickr_frame"] = ref_utils.testset_gt_frame(args, "clip.jpg", 3)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=HERE)
    a = ap.parse_args()
    torch, ref_model, ref_utils = load_reference()
    torch.manual_seed(0)
    head = {}
    head.update(head_case(torch, ref_model, "cfg1_relu", 8, 11, "relu"))
    head.update(head_case(torch, ref_model, "cfg1_mix", 8, 12, "mix"))
    head.update(head_case(torch, ref_model, "b32_mix", 32, 13, "mix"))
    head.update(head_case(torch, ref_model, "b16_notrimap", 16, 14, "mix", trimap=False))
    head.update(head_case(torch, ref_model, "b16_noneg", 16, 15, "mix", neg=False))
    head.update(head_case(torch, ref_model, "b8_signed", 8, 16, "signed"))
    head.update(tube_case(torch, ref_model, "tube_b2t4", 2, 4, 17, "mix"))
    np.savez_compressed(os.path.join(a.out, "head_golden.npz"), **head)
    ev = eval_case(ref_utils)
    ev.update(flickr_gt_case(ref_utils))
    np.savez_compressed(os.path.join(a.out, "eval_golden.npz"), **ev)
    for f in ("head_golden.npz", "eval_golden.npz"):
        print(f, os.path.getsize(os.path.join(a.out, f)), "bytes")


if __name__ == "__main__":
    main()
