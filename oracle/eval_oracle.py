"""CPU oracle for the localisation-map evaluation (14->224 upsample, median
threshold, cIoU, AUC, mTC, ground-truth box rasterisation).

TEST INFRASTRUCTURE ONLY.  Imported by ``tests/``, ``__graft_entry__.smoke()``
and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``; nothing on
the product path touches it.

Parity status: PINNED against the reference itself.  The reference ships no
tests; the pin is ``tests/golden/eval_*.npz``, produced
by ``tests/golden/make_golden.py`` by running the *unmodified* reference
``utils.normalize_img`` / ``utils.Evaluator`` / ``utils.testset_gt`` /
``utils.mTC`` plus the reference's inline glue with the real ``cv2.resize`` and
``sklearn.metrics.auc`` on seeded maps and the real VGG-SS boxes.

What is restated (reference file:line):
  * ``cv2.resize(h, (224,224), INTER_LINEAR)``      train_hardway_1frame.py:198, test.py:97
      (OpenCV 4.x, third party, not vendored by the reference: half-pixel centres,
       edge clamp; restated here in a fixed fp32 operation order -- horizontal lerp
       then vertical lerp, every product and sum rounded to fp32.  OpenCV's own SIMD
       paths differ from any fixed order by <= 1 ulp, see tests.)
  * ``normalize_img(-h)``                           utils.py:234-239
  * ``pred = 1 - h``; ``np.sort(pred.flatten())[25088]``; binarise
                                                    train_hardway_1frame.py:201-204
  * ``Evaluator.cal_CIOU(pred, gt, 0.5)``           utils.py:209-214
  * 21-point success curve, ``sklearn.metrics.auc``, cIoU@0.5
                                                    train_hardway_1frame.py:208-215, utils.py:216-229
  * ``testset_gt`` / ``testset_gt_frame``           utils.py:241-274 / 276-309
  * ``mTC``                                         utils.py:311-318
"""
from __future__ import annotations

import numpy as np

SRC = 14
DST = 224
NPIX = DST * DST
K_RANK = NPIX // 2          # int(224*224/2) = 25088, 0-based rank of the threshold


def _taps():
    """Source index and fractional weight for each of the 224 destination
    coordinates (scale 1/16, half-pixel centres, clamped)."""
    x = np.arange(DST, dtype=np.float64)
    f = (x + 0.5) * (SRC / DST) - 0.5
    s = np.floor(f).astype(np.int64)
    w = (f - s).astype(np.float32)            # multiples of 1/32 -> exact in fp32
    lo = s < 0
    s[lo] = 0
    w[lo] = 0.0
    hi = s >= SRC - 1
    s[hi] = SRC - 1
    w[hi] = 0.0
    s1 = np.minimum(s + 1, SRC - 1)
    return s, s1, w


_S0, _S1, _W = _taps()


def upsample_14_to_224(maps14):
    """[N,14,14] (or [14,14]) fp32 -> [N,224,224] fp32, fixed op order:
    t = m[:, sx]*(1-fx) + m[:, sx+1]*fx  (horizontal), then the same vertically."""
    m = np.asarray(maps14, dtype=np.float32)
    single = m.ndim == 2
    if single:
        m = m[None]
    w1 = _W
    w0 = (np.float32(1.0) - _W).astype(np.float32)
    rows = m[:, :, _S0] * w0[None, None, :] + m[:, :, _S1] * w1[None, None, :]   # [N,14,224]
    out = rows[:, _S0, :] * w0[None, :, None] + rows[:, _S1, :] * w1[None, :, None]
    out = out.astype(np.float32)
    return out[0] if single else out


def normalize_img(value, vmax=None, vmin=None):
    """utils.py:234-239 -- min/max normalisation, skipped for a constant map."""
    vmin = value.min() if vmin is None else vmin
    vmax = value.max() if vmax is None else vmax
    if not (vmax - vmin) == 0:
        value = (value - vmin) / (vmax - vmin)
    return value


def binarise_from_up224(up224):
    """train_hardway_1frame.py:199-204 on one already-upsampled fp32 map.
    Returns (infer uint8 [224,224], pred_before_threshold fp32, threshold)."""
    h = normalize_img(-np.asarray(up224, dtype=np.float32))
    pred = np.float32(1.0) - h
    thr = np.sort(pred.flatten())[K_RANK]
    final = pred.copy()
    final[final > thr] = 1
    final[final < 1] = 0
    infer = (final >= 0.5)                      # utils.py:211 with thres = 0.5
    return infer.astype(np.uint8), pred, thr


def ciou_counts(infer, gt_w, gt_den=1):
    """utils.py:209-214 in integer arithmetic.  gt = gt_w / gt_den elementwise.
    Returns (inter_w, union_w) with ciou = inter_w / union_w exactly equal to the
    reference's float64 ``sum(infer*gt) / (sum(gt) + sum(infer*(gt==0)))``."""
    infer = np.asarray(infer).astype(np.int64)
    gt_w = np.asarray(gt_w).astype(np.int64)
    inter_w = int((infer * gt_w).sum())
    union_w = int(gt_w.sum() + gt_den * (infer * (gt_w == 0)).sum())
    return inter_w, union_w


def ciou_value(inter_w, union_w):
    with np.errstate(divide='ignore', invalid='ignore'):
        return np.float64(inter_w) / np.float64(union_w)


def cal_ciou_float(infer, gtmap, thres=0.01):
    """Literal float64 restatement of Evaluator.cal_CIOU (utils.py:209-214)."""
    infer_map = np.zeros((DST, DST))
    infer_map[infer >= thres] = 1
    inter = np.sum(infer_map * gtmap)
    union = np.sum(gtmap) + np.sum(infer_map * (gtmap == 0))
    with np.errstate(divide='ignore', invalid='ignore'):
        return inter / union, inter, union


def success_curve(ciou):
    """results[i] = mean(ciou >= 0.05*i), i = 0..20; float64 thresholds exactly as
    Python computes ``0.05*i`` (train_hardway_1frame.py:208-212)."""
    c = np.asarray(ciou, dtype=np.float64)
    counts = np.array([int(np.sum(c >= 0.05 * i)) for i in range(21)], dtype=np.int64)
    return counts, counts / len(c)


def auc_trapezoid(results):
    """sklearn.metrics.auc(x, y) for increasing x == trapezoid rule."""
    x = np.array([0.05 * i for i in range(21)], dtype=np.float64)
    y = np.asarray(results, dtype=np.float64)
    dx = np.diff(x)
    return float((dx * (y[1:] + y[:-1]) / 2.0).sum())


def ciou_at_half(ciou):
    c = np.asarray(ciou, dtype=np.float64)
    return float(np.sum(c >= 0.5) / len(c))     # train_hardway_1frame.py:215


# --------------------------------------------------------------------------- GT
def boxes_to_pixel_rects(boxes, mode):
    """Integer pixel boxes (xmin,ymin,xmax,ymax) as the reference derives them.
    mode 'vggss': normalised floats, int(224*max(x,0))        utils.py:267
    mode 'flickr' / 'flickr_frame': 256-grid ints, int(224*int(t)/256)   utils.py:254"""
    out = []
    for b in boxes:
        if mode == 'vggss':
            out.append([int(224 * max(float(x), 0)) for x in b])
        else:
            out.append([int(224 * int(x) / 256) for x in b])
    return out


def rasterise_gt(boxes, mode):
    """utils.py:241-274 (testset_gt) and :276-309 (testset_gt_frame).
    Returns (weights uint8 [224,224], denominator) with gt = weights / denominator:
      vggss         union of boxes                     -> {0,1}, den 1
      flickr        sum of annotators / 2, clipped 1   -> {0,1,2}, den 2
      flickr_frame  raw sum of boxes                   -> counts, den 1"""
    acc = np.zeros((DST, DST), dtype=np.int64)
    for xmin, ymin, xmax, ymax in boxes_to_pixel_rects(boxes, mode):
        tmp = np.zeros((DST, DST), dtype=np.int64)
        tmp[ymin:ymax, xmin:xmax] = 1            # numpy slice semantics, as the reference
        acc += tmp
    if mode == 'vggss':
        return (acc > 0).astype(np.uint8), 1
    if mode == 'flickr':
        return np.minimum(acc, 2).astype(np.uint8), 2
    if mode == 'flickr_frame':
        return acc.astype(np.uint8), 1
    raise ValueError(mode)


# ------------------------------------------------------------------- pipelines
def evaluate_map(map14, gt_w, gt_den=1):
    """One clip through the reference glue (train_hardway_1frame.py:198-207)."""
    up = upsample_14_to_224(map14)
    infer, _, _ = binarise_from_up224(up)
    iw, uw = ciou_counts(infer, gt_w, gt_den)
    return dict(up224=up, mask=infer, inter_w=iw, union_w=uw, ciou=ciou_value(iw, uw))


def evaluate_maps(maps14, gts, gt_den=1):
    """Whole test set: per-clip cIoU, then cIoU@0.5 and AUC
    (train_hardway_1frame.py:196-216)."""
    cious, inter, union = [], [], []
    for m, g in zip(maps14, gts):
        r = evaluate_map(m, g, gt_den)
        cious.append(r['ciou']); inter.append(r['inter_w']); union.append(r['union_w'])
    cious = np.array(cious, dtype=np.float64)
    counts, results = success_curve(cious)
    return dict(ciou=cious, inter_w=np.array(inter), union_w=np.array(union),
                counts=counts, ciou_at_half=ciou_at_half(cious), auc=auc_trapezoid(results))


def mtc(masks):
    """utils.py:311-318 -- mean cIoU between consecutive binary predictions."""
    vals = []
    for i in range(len(masks) - 1):
        iw, uw = ciou_counts(masks[i], masks[i + 1], 1)
        vals.append(np.float32(ciou_value(iw, uw)))      # stored into a float32 torch tensor
    return float(np.sum(np.array(vals, dtype=np.float32), dtype=np.float32) / np.float32(len(masks) - 1))


def synthetic_maps(n, seed=0):
    """0.5 + 0.02*randn plus one Gaussian bump."""
    rng = np.random.default_rng(seed)
    base = 0.5 + 0.02 * rng.standard_normal((n, SRC, SRC))
    amp = rng.uniform(0.0, 0.3, n)
    cx = rng.uniform(0, SRC, n); cy = rng.uniform(0, SRC, n)
    sg = rng.uniform(1.0, 4.0, n)
    yy, xx = np.mgrid[0:SRC, 0:SRC]
    bump = amp[:, None, None] * np.exp(-((xx[None] - cx[:, None, None]) ** 2 +
                                         (yy[None] - cy[:, None, None]) ** 2) /
                                       (2 * sg[:, None, None] ** 2))
    return (base + bump).astype(np.float32)


def evaluate_videos(maps14, gts, gt_den=1):
    """The PerFrameLabels test loop, train_hardway.py:158-203: per video, every sampled frame goes
    through the single-frame glue (:168-178), then the video's cIoU@0.5 fraction (:200), the AUC
    of its own 21-point success curve (:193-199) and mTC of its predictions (:185); the dataset
    numbers are plain means over videos (:201-203).  maps14 [V][F,14,14], gts [V][F] weights."""
    per_video, cious = [], []
    for vm, vg in zip(maps14, gts):
        iou, preds = [], []
        for m, g in zip(vm, vg):
            r = evaluate_map(m, g, gt_den)
            iou.append(r['ciou']); preds.append(r['mask'])
        iou = np.array(iou, dtype=np.float64)
        _, results = success_curve(iou)
        m_tc = mtc(preds) if len(preds) > 1 else float('nan')
        per_video.append((float(np.sum(iou >= 0.5) / len(iou)), auc_trapezoid(results), m_tc))
        cious.append(iou)
    pv = np.array(per_video, dtype=np.float64)
    return dict(ciou=np.array(cious), per_video=pv, ciou_mean=float(np.sum(pv[:, 0]) / len(pv)),
                auc_mean=float(np.sum(pv[:, 1]) / len(pv)), mtc_mean=float(np.sum(pv[:, 2]) / len(pv)))
