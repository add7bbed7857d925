"""Recipe for ``oracle/_ref/``: the UNMODIFIED reference modules the hot path lives in, staged so
they travel to the GPU box like a built artefact (``oracle/_ref/`` is git-ignored, NOT
gpurun-ignored; nothing under it is ever committed).

TEST / MEASUREMENT INFRASTRUCTURE ONLY (see oracle/__init__.py): used by ``bench.py --impl
reference``, bench.py's ``cpu_baseline`` / ``also.reference_*`` legs and by tests as a checker.

    python oracle/make_ref.py            # needs /root/reference (the build container)

The reference is pure Python -- there is nothing to compile.  The modules are packed byte for byte
into ONE archive, ``oracle/_ref/reference_modules.zip`` (a built artefact like a .so; Python imports
straight from it), with the sha256 of every source recorded in ``oracle/_ref/MANIFEST.json``:
    model.py                 AVENet / HardWayAttention / FullModel          (the head, model.py:17-154)
    utils.py                 Evaluator / normalize_img / testset_gt / mTC   (the metric, utils.py:203-318)
    models/base_models.py    2-D ResNet-18 (config 1 runs the whole AVENet)
    models/resnet3D.py       3-D ResNet-18 (imported by model.py:6)
``load()`` imports them under private module names with the one CPU shim SURVEY.md 8c describes
(``Tensor.cuda`` -> identity when no GPU is present, because model.py:115 hard-codes ``.cuda()``).
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import zipfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
DST = os.path.join(HERE, "_ref")
ZIP = os.path.join(DST, "reference_modules.zip")
FILES = ["model.py", "utils.py", "models/base_models.py", "models/resnet3D.py"]


def _sha(p):
    return hashlib.sha256(open(p, "rb").read()).hexdigest()


def stage(force: bool = False) -> bool:
    """Pack the reference modules into oracle/_ref/reference_modules.zip.  Returns False when
    /root/reference is absent (the GPU box: it then uses what travelled with the snapshot)."""
    if not os.path.isdir(REF):
        return False
    os.makedirs(DST, exist_ok=True)
    man = {f: _sha(os.path.join(REF, f)) for f in FILES}
    mpath = os.path.join(DST, "MANIFEST.json")
    if not force and os.path.exists(ZIP) and os.path.exists(mpath):
        try:
            if json.load(open(mpath)).get("sha256") == man:
                return True
        except Exception:
            pass
    with zipfile.ZipFile(ZIP, "w", zipfile.ZIP_DEFLATED) as z:
        for f in FILES:
            z.write(os.path.join(REF, f), f)
        z.writestr("models/__init__.py", "")                  # the reference has a namespace package
    with zipfile.ZipFile(ZIP) as z:                           # byte-for-byte check of what was packed
        for f in FILES:
            assert hashlib.sha256(z.read(f)).hexdigest() == man[f], f
    json.dump({"reference": REF, "sha256": man}, open(mpath, "w"), indent=1)
    return True


def available() -> bool:
    return os.path.exists(ZIP)


_loaded = None


class cpu_shim:
    """``with cpu_shim():`` -- run the reference on HOST tensors where a GPU is present: model.py:115
    (and :48-51) call ``.cuda()`` on the mask they build, which would put it on the GPU while the
    features are on the CPU.  Inside the block ``Tensor.cuda`` is the identity (SURVEY.md 8c's shim)."""

    def __enter__(self):
        import torch
        self._orig = torch.Tensor.cuda
        torch.Tensor.cuda = lambda t, *a, **k: t
        return self

    def __exit__(self, *exc):
        import torch
        torch.Tensor.cuda = self._orig
        return False


def load():
    """(ref_model, ref_utils): the staged reference modules, imported unmodified."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("oracle/_ref is not staged: run `python oracle/make_ref.py` where /root/reference exists")
    import torch
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self                  # CPU shim (model.py:115,48-51)
    saved = {k: sys.modules.get(k) for k in ("model", "utils", "models", "models.base_models", "models.resnet3D")}
    sys.path.insert(0, ZIP)                # zipimport: the modules are imported from the archive
    try:
        for k in saved:
            sys.modules.pop(k, None)
        import model as ref_model          # noqa: E402
        import utils as ref_utils          # noqa: E402
    finally:
        sys.path.remove(ZIP)
        for k, m in saved.items():         # leave no 'model' / 'utils' / 'models' names behind
            sys.modules.pop(k, None)
            if m is not None:
                sys.modules[k] = m
    _loaded = (ref_model, ref_utils)
    return _loaded


def reference_head(trimap=True, neg=True, eps=0.65, eps2=0.4):
    """The reference's ``AVENet`` with both backbones replaced by ``nn.Identity()`` so that raw
    features are the inputs: ``net(v [B,512,14,14], a [B,512,1,1])`` runs model.py:112-154 verbatim
    (the audio max-pool of a 1x1 map is the identity)."""
    import types
    import torch.nn as nn
    ref_model, _ = load()
    args = types.SimpleNamespace(epsilon=eps, epsilon2=eps2, tri_map=trimap, Neg=neg)
    net = ref_model.AVENet(args, False)
    net.imgnet = nn.Identity()
    net.audnet = nn.Identity()
    return net


if __name__ == "__main__":
    ok = stage(force="--force" in sys.argv)
    print("staged" if ok else "no /root/reference here", DST, "available:", available())
