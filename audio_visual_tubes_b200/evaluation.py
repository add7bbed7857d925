"""cIoU / AUC evaluation with the reference's API (``/root/reference/utils.py:203-318`` and
the glue every script repeats, ``train_hardway_1frame.py:187-219`` == ``test.py:87-172``),
running in libavtubes on the GPU.

    Evaluator().cal_CIOU(infer, gtmap, thres) -> (ciou, inter, union)      utils.py:209-214
    Evaluator().cal_AUC() / .final() / .clear()                             utils.py:216-232
    normalize_img(value, vmax=None, vmin=None)                              utils.py:234-239
    testset_gt(args, name) / testset_gt_frame(args, name, frame)            utils.py:241-309
    mTC(predictions, gt_maps)                                               utils.py:311-318
    evaluate_maps(maps14, boxes=.. | gt_dense=..) -> per-clip cIoU, cIoU@0.5, AUC
                                                  train_hardway_1frame.py:196-216 in one launch

Host arrays are accepted wherever the reference takes them (they are copied to the device);
CUDA tensors are used in place.  There is no CPU implementation here.
"""
from __future__ import annotations

import xml.etree.ElementTree as ET

import numpy as np
import torch

from . import _lib

GT_MODES = {"vggss": _lib.GT_VGGSS, "flickr": _lib.GT_FLICKR, "flickr_frame": _lib.GT_FLICKR_FRAME}
GT_DEN = {"vggss": 1, "flickr": 2, "flickr_frame": 1}
NPIX = 224 * 224


def _dev(device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("libavtubes evaluation needs a CUDA device (no CPU fallback)")
    return torch.device(device or "cuda")


def _to_dev(x, dtype, device):
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=dtype).contiguous()
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype).to(device)


# ------------------------------------------------------------------ boxes
def boxes_to_pixel_rects(boxes, mode):
    """(xmin, ymin, xmax, ymax) integer pixels as the reference derives them:
    VGG-SS normalised floats ``int(224*max(x,0))`` (utils.py:267); Flickr 256-grid
    ``int(224*int(t)/256)`` (utils.py:254, 290)."""
    if mode == "vggss":
        return [[int(224 * max(float(x), 0)) for x in b] for b in boxes]
    return [[int(224 * int(x) / 256) for x in b] for b in boxes]


def pack_boxes(box_lists, mode, max_boxes=None):
    """List (per clip) of box lists -> (int16 [N,max_boxes,4], uint8 [N]) host tensors."""
    rects = [boxes_to_pixel_rects(b, mode) for b in box_lists]
    mb = max(1, max((len(r) for r in rects), default=1))
    mb = max_boxes or mb
    if mb > 32:
        raise ValueError("at most 32 boxes per clip")
    arr = np.zeros((len(rects), mb, 4), dtype=np.int16)
    cnt = np.zeros(len(rects), dtype=np.uint8)
    for i, r in enumerate(rects):
        if len(r) > mb:
            raise ValueError(f"clip {i} has {len(r)} boxes > max_boxes={mb}")
        cnt[i] = len(r)
        if r:
            arr[i, :len(r)] = np.clip(np.asarray(r, dtype=np.int64), 0, 32767)
    return torch.from_numpy(arr), torch.from_numpy(cnt)


def rasterise_gt(box_lists, mode="vggss", device=None):
    """utils.py:264-273 / 252-262 / 291-297 on the device: uint8 weights [N,224,224]
    (gt = weights / GT_DEN[mode])."""
    dev = _dev(device)
    lib = _lib.load()
    boxes, cnt = pack_boxes(box_lists, mode)
    boxes, cnt = boxes.to(dev), cnt.to(dev)
    n = boxes.shape[0]
    out = torch.empty(n, 224, 224, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.avt_rasterise_gt(_lib.ptr(boxes), _lib.ptr(cnt), n, boxes.shape[1],
                                        GT_MODES[mode], _lib.ptr(out), _lib.stream_ptr()),
                   "avt_rasterise_gt")
    return out


def _flickr_boxes(xml_path):
    """utils.py:243-251: every <bbox> of the annotation file, children 1..4."""
    root = ET.parse(xml_path).getroot()
    out = []
    for child in root:
        for childs in child:
            if childs.tag == "bbox":
                out.append([int(childs[i].text) for i in range(1, 5)])
    return out


def testset_gt(args, name):
    """utils.py:241-274 -- float64 [224,224] ground-truth map of one clip."""
    if args.testset == "flickr":
        boxes = _flickr_boxes(args.og_gt_path + "%s.xml" % name[:-4])
        w = rasterise_gt([boxes], "flickr")
        return (w[0].cpu().numpy().astype(np.float64)) / 2.0
    if args.testset == "vggss":
        w = rasterise_gt([args.gt_all[name[:-4]]], "vggss")
        return w[0].cpu().numpy().astype(np.float64)
    raise ValueError(args.testset)


def testset_gt_frame(args, name, frame):
    """utils.py:276-309 -- per-frame variant (Flickr: raw box counts, no /2, no clip)."""
    if args.testset == "flickr":
        boxes = _flickr_boxes(args.og_gt_path + name[:-4] + "_" + str(frame) + ".xml")
        w = rasterise_gt([boxes], "flickr_frame")
        return w[0].cpu().numpy().astype(np.float64)
    return testset_gt(args, name)


def normalize_img(value, vmax=None, vmin=None):
    """utils.py:234-239 (works on numpy arrays and torch tensors alike)."""
    vmin = value.min() if vmin is None else vmin
    vmax = value.max() if vmax is None else vmax
    if not (vmax - vmin) == 0:
        value = (value - vmin) / (vmax - vmin)
    return value


# ------------------------------------------------------------------ kernels
def ciou_counts(infer, gt_w, thres, gt_den=1, device=None):
    """cal_CIOU on ready-made maps.  infer [N,224,224] float, gt_w [N,224,224] uint8 weights.
    Returns (ciou float64 [N], inter_w int32 [N], union_w int32 [N]) device tensors."""
    dev = _dev(device if device is not None else (infer.device if isinstance(infer, torch.Tensor) and infer.is_cuda else None))
    lib = _lib.load()
    infer = _to_dev(infer, torch.float32, dev).view(-1, 224, 224)
    gt_w = _to_dev(gt_w, torch.uint8, dev).view(-1, 224, 224)
    n = infer.shape[0]
    inter = torch.empty(n, dtype=torch.int32, device=dev)
    union = torch.empty(n, dtype=torch.int32, device=dev)
    ciou = torch.empty(n, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.avt_ciou_counts(_lib.ptr(infer), float(thres), _lib.ptr(gt_w), int(gt_den), n,
                                       _lib.ptr(inter), _lib.ptr(union), _lib.ptr(ciou),
                                       _lib.stream_ptr()), "avt_ciou_counts")
    return ciou, inter, union


def success_auc(ciou):
    """train_hardway_1frame.py:208-215: (counts int32 [21], cIoU@0.5, AUC) from float64 cIoUs."""
    lib = _lib.load()
    dev = ciou.device
    counts = torch.empty(21, dtype=torch.int32, device=dev)
    out = torch.empty(2, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.avt_auc(_lib.ptr(ciou), ciou.numel(), _lib.ptr(counts), _lib.ptr(out),
                               _lib.stream_ptr()), "avt_auc")
    return counts, out


def eval_maps_raw(maps14, gt_dense=None, boxes=None, box_count=None, gt_mode="vggss",
                  want_up224=False, want_mask=False):
    """One ``avt_eval_maps`` launch on device tensors; returns a dict of device tensors."""
    lib = _lib.load()
    _lib.require_cuda(maps14, gt_dense, boxes, box_count)
    dev = maps14.device
    maps14 = maps14.contiguous().float().view(-1, 14, 14)
    n = maps14.shape[0]
    up = torch.empty(n, 224, 224, dtype=torch.float32, device=dev) if want_up224 else None
    mk = torch.empty(n, 224, 224, dtype=torch.uint8, device=dev) if want_mask else None
    inter = torch.empty(n, dtype=torch.int32, device=dev)
    union = torch.empty(n, dtype=torch.int32, device=dev)
    ciou = torch.empty(n, dtype=torch.float64, device=dev)
    mb = 0 if boxes is None else boxes.shape[1]
    ws = torch.empty(_lib.EVAL_WORKSPACE_BYTES, dtype=torch.uint8, device=dev)   # this call's ticket word
    with torch.cuda.device(dev):
        _lib.check(lib.avt_eval_maps(_lib.ptr(maps14), n, _lib.ptr(gt_dense), _lib.ptr(boxes),
                                     _lib.ptr(box_count), mb, GT_MODES[gt_mode], _lib.ptr(up),
                                     _lib.ptr(mk), _lib.ptr(inter), _lib.ptr(union), _lib.ptr(ciou),
                                     _lib.ptr(ws), _lib.stream_ptr()), "avt_eval_maps")
    return dict(up224=up, mask=mk, inter_w=inter, union_w=union, ciou=ciou)


def evaluate_maps(maps14, boxes=None, gt_dense=None, gt_mode="vggss", want_up224=False,
                  want_mask=False, device=None):
    """The reference's whole test loop (train_hardway_1frame.py:196-216) for N clips.

    maps14: [N,14,14] (or [N,1,14,14]) localisation maps ``A``; give either ``boxes`` (per-clip
    lists in the dataset's own units, rasterised on the fly) or ``gt_dense`` uint8 weights.
    Returns dict(ciou [N] float64, inter_w, union_w, counts [21], ciou_at_half, auc [, up224, mask])
    with device tensors; ``ciou_at_half`` / ``auc`` are Python floats (one D2H of 2 doubles)."""
    dev = _dev(device if device is not None else (maps14.device if isinstance(maps14, torch.Tensor) and maps14.is_cuda else None))
    m = _to_dev(maps14, torch.float32, dev)
    bx = bc = None
    if gt_dense is not None:
        gt_dense = _to_dev(gt_dense, torch.uint8, dev)
    else:
        bx, bc = pack_boxes(boxes, gt_mode)
        bx, bc = bx.to(dev), bc.to(dev)
    r = eval_maps_raw(m, gt_dense, bx, bc, gt_mode, want_up224, want_mask)
    counts, out = success_auc(r["ciou"])
    o = out.cpu()
    r.update(counts=counts, ciou_at_half=float(o[0]), auc=float(o[1]))
    return r


def evaluate_videos(maps14, boxes=None, gt_dense=None, gt_mode="flickr_frame", device=None):
    """The PerFrameLabels test loop of the 16-frame model (train_hardway.py:158-203) for V videos
    of F sampled frames each, without the reference's per-frame ``.cpu().numpy()``.

    maps14: [V,F,14,14] localisation maps (``heatmap[0,0]`` of every sampled frame); ``boxes``:
    V lists of F per-frame box lists (``testset_gt_frame`` units, utils.py:276-309) or ``gt_dense``
    [V,F,224,224] uint8 weights.  Returns dict(ciou [V,F] float64, per_video [V,3] float64 =
    (cIoU@0.5 fraction, AUC, mTC) per video, ciou_mean, auc_mean, mtc_mean = the three numbers the
    reference prints (:201-203)); device tensors except the three means (one D2H of V*3 doubles)."""
    dev = _dev(device if device is not None else (maps14.device if isinstance(maps14, torch.Tensor) and maps14.is_cuda else None))
    m = _to_dev(maps14, torch.float32, dev)
    if m.dim() != 4 or m.shape[2:] != (14, 14):
        raise ValueError(f"evaluate_videos: maps14 must be [V,F,14,14], got {tuple(m.shape)}")
    nv, nf = int(m.shape[0]), int(m.shape[1])
    bx = bc = None
    if gt_dense is not None:
        gt_dense = _to_dev(gt_dense, torch.uint8, dev).reshape(nv * nf, 224, 224)
    else:
        if len(boxes) != nv or any(len(b) != nf for b in boxes):
            raise ValueError("evaluate_videos: boxes must be V lists of F per-frame box lists")
        bx, bc = pack_boxes([fb for vb in boxes for fb in vb], gt_mode)
        bx, bc = bx.to(dev), bc.to(dev)
    r = eval_maps_raw(m.reshape(nv * nf, 14, 14), gt_dense, bx, bc, gt_mode, False, True)
    lib = _lib.load()
    pi = torch.empty(max(nv * nf - 1, 1), dtype=torch.int32, device=dev)
    pu = torch.empty_like(pi)
    per_video = torch.empty(nv, 3, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.avt_video_metrics(_lib.ptr(r["ciou"]), _lib.ptr(r["mask"]), nv, nf, _lib.ptr(pi),
                                         _lib.ptr(pu), _lib.ptr(per_video), _lib.stream_ptr()),
                   "avt_video_metrics")
    pv = per_video.cpu().numpy()
    return dict(ciou=r["ciou"].view(nv, nf), inter_w=r["inter_w"].view(nv, nf),
                union_w=r["union_w"].view(nv, nf), mask=r["mask"].view(nv, nf, 224, 224),
                per_video=per_video,
                ciou_mean=float(np.sum(pv[:, 0]) / nv), auc_mean=float(np.sum(pv[:, 1]) / nv),
                mtc_mean=float(np.sum(pv[:, 2]) / nv))


def mTC(predictions, gt_maps=None, device=None):
    """utils.py:311-318 -- mean cIoU between consecutive binary predictions (gt_maps unused,
    as in the reference)."""
    dev = _dev(device)
    lib = _lib.load()
    if isinstance(predictions, torch.Tensor):
        masks = (predictions.to(dev) >= 0.5).to(torch.uint8).contiguous()
    else:
        masks = torch.as_tensor(np.stack([np.asarray(p) >= 0.5 for p in predictions]).astype(np.uint8)).to(dev)
    t = masks.shape[0]
    pi = torch.empty(t - 1, dtype=torch.int32, device=dev)
    pu = torch.empty(t - 1, dtype=torch.int32, device=dev)
    out = torch.empty(1, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.avt_mtc(_lib.ptr(masks), t, _lib.ptr(pi), _lib.ptr(pu), _lib.ptr(out),
                               _lib.stream_ptr()), "avt_mtc")
    return out[0]


class Evaluator:
    """utils.py:203-232, same methods and return values."""

    def __init__(self):
        self.ciou = []

    def cal_CIOU(self, infer, gtmap, thres=0.01):
        gt = np.asarray(gtmap.cpu() if isinstance(gtmap, torch.Tensor) else gtmap, dtype=np.float64)
        # the reference's GT maps hold multiples of 1/2 (Flickr) or integers: carry them as
        # integer weights so the counts are exact
        den = 1 if np.all(gt == np.round(gt)) else 2
        w = np.round(gt * den)
        if np.any(np.abs(w - gt * den) > 0) or w.max(initial=0) > 255:
            raise ValueError("cal_CIOU: gtmap values must be multiples of 1/2 in [0, 127.5]")
        ciou, inter, union = ciou_counts(np.asarray(infer.cpu() if isinstance(infer, torch.Tensor) else infer,
                                                    dtype=np.float32)[None], w.astype(np.uint8)[None],
                                         thres, den)
        c = float(ciou[0].item())
        self.ciou.append(c)
        return c, float(inter[0].item()) / den, float(union[0].item()) / den

    def cal_AUC(self):
        """utils.py:216-225 (21 thresholds 0.05*i, trapezoid)."""
        c = torch.as_tensor(np.asarray(self.ciou, dtype=np.float64)).to(_dev())
        _, out = success_auc(c)
        return float(out[1].item())

    def final(self):
        return float(np.mean(np.array(self.ciou) >= 0.5))

    def clear(self):
        self.ciou = []
