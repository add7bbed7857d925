"""Config 3 on the REAL box lists: all 5158 VGG-SS clips of /root/reference/metadata/vggss.json
through the UNMODIFIED reference glue (train_hardway_1frame.py:198-216 line for line, with the
real cv2.resize / utils.normalize_img / utils.testset_gt / utils.Evaluator / sklearn auc).

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_vggss.py

Writes tests/golden/vggss_golden.npz:
  boxes [5158,7,4] float64 + count [5158] uint8   the json's box lists, copied as DATA (the 12
                                                  negative-coordinate entries and the 2 clips whose
                                                  GT rasterises to nothing included)
  gt_sum [5158] int32        utils.testset_gt(...).sum() per clip (utils.py:264-273)
  ciou [5158] float64, inter / union [5158] int32, auc, ciou_at_half
                             of the seeded config-3 maps (oracle.eval_oracle.synthetic_maps(5158, 0))
Only data and reference OUTPUTS are stored; the maps are regenerated from the seed.
"""
from __future__ import annotations

import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, "..", ".."))
REF = "/root/reference"


def main():
    import cv2
    from sklearn.metrics import auc as sk_auc
    from make_golden import load_reference
    from oracle.eval_oracle import synthetic_maps
    _, _, ref_utils = load_reference()
    meta = json.load(open(os.path.join(REF, "metadata", "vggss.json")))
    n = len(meta)
    maps = synthetic_maps(n, 0)
    args = types.SimpleNamespace(testset="vggss", gt_all={m["file"]: m["bbox"] for m in meta})
    boxes = np.zeros((n, 7, 4), np.float64)
    count = np.zeros(n, np.uint8)
    gt_sum = np.zeros(n, np.int32)
    cious = []
    inter_a = np.zeros(n, np.int32); union_a = np.zeros(n, np.int32)
    for i, m in enumerate(meta):
        bb = m["bbox"]
        count[i] = len(bb)
        boxes[i, :len(bb)] = np.asarray(bb, np.float64)
        # ---- the reference glue, train_hardway_1frame.py:198-207 ----
        heatmap_now = cv2.resize(maps[i], dsize=(224, 224), interpolation=cv2.INTER_LINEAR)
        heatmap_now = ref_utils.normalize_img(-heatmap_now)
        gt_map = ref_utils.testset_gt(args, m["file"] + ".mp4")
        gt_sum[i] = int(gt_map.sum())
        pred = 1 - heatmap_now
        threshold = np.sort(pred.flatten())[int(pred.shape[0] * pred.shape[1] / 2)]
        pred[pred > threshold] = 1
        pred[pred < 1] = 0
        evaluator = ref_utils.Evaluator()
        with np.errstate(divide="ignore", invalid="ignore"):
            ciou, inter, union = evaluator.cal_CIOU(pred, gt_map, 0.5)
        cious.append(ciou); inter_a[i] = int(inter); union_a[i] = int(union)
    cious = np.array(cious, np.float64)
    results = [np.sum(cious >= 0.05 * i) / len(cious) for i in range(21)]    # :208-212
    auc_ = sk_auc([0.05 * i for i in range(21)], results)                    # :213-214
    half = np.sum(cious >= 0.5) / len(cious)                                 # :215
    out = os.path.join(HERE, "vggss_golden.npz")
    np.savez_compressed(out, boxes=boxes, count=count, gt_sum=gt_sum, ciou=cious, inter=inter_a,
                        union=union_a, auc=np.float64(auc_), ciou_at_half=np.float64(half),
                        results=np.array(results))
    print(out, os.path.getsize(out), "bytes; empty GT clips:", int((gt_sum == 0).sum()),
          "negative-coordinate clips:", int(sum(any(x < 0 for b in m["bbox"] for x in b) for m in meta)),
          "auc", auc_, "cIoU@0.5", half)


if __name__ == "__main__":
    main()
