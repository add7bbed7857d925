"""CPU-only checks of the boundary: the C-ABI library builds for sm_100a, loads, and exports
every symbol include/avtubes.h declares (no compute without a GPU); host-side box logic
matches the oracle."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from audio_visual_tubes_b200 import build, _lib
    build.build()
    return _lib.load()


def _declared():
    src = open(os.path.join(ROOT, "include", "avtubes.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(avt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(lib):
    from audio_visual_tubes_b200 import _lib
    names = _declared()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/avtubes.h but not exported"
    assert sorted(_lib.SIGNATURES) == names, "ctypes prototypes out of sync with the header"


def test_header_constants_match_the_binding():
    from audio_visual_tubes_b200 import _lib
    src = open(os.path.join(ROOT, "include", "avtubes.h")).read()
    consts = {k: int(v) for k, v in re.findall(r"#define\s+(AVT_[A-Z0-9_]+)\s+(-?\d+)\b", src)}
    assert consts["AVT_MTC_STATIC_MAX_FRAMES"] == _lib.MTC_STATIC_MAX_FRAMES
    assert consts["AVT_EVAL_WORKSPACE_BYTES"] == _lib.EVAL_WORKSPACE_BYTES
    assert (consts["AVT_PRECISION_FP32"], consts["AVT_PRECISION_F16"]) == (_lib.PRECISION_FP32, _lib.PRECISION_F16)
    assert (consts["AVT_DTYPE_F32"], consts["AVT_DTYPE_F16"], consts["AVT_DTYPE_BF16"]) == \
        (_lib.DTYPE_F32, _lib.DTYPE_F16, _lib.DTYPE_BF16)
    assert (consts["AVT_STAGE_PREP"], consts["AVT_STAGE_SIM"], consts["AVT_BWD_ROWGRAD_DONE"]) == \
        (_lib.STAGE_PREP, _lib.STAGE_SIM, _lib.BWD_ROWGRAD_DONE)
    assert (consts["AVT_GT_VGGSS"], consts["AVT_GT_FLICKR"], consts["AVT_GT_FLICKR_FRAME"]) == \
        (_lib.GT_VGGSS, _lib.GT_FLICKR, _lib.GT_FLICKR_FRAME)


def test_version_and_error_paths_without_gpu(lib):
    assert lib.avt_version() >= 100
    # argument validation happens before any CUDA call: bad shapes report AVT_ERR_INVALID
    rc = lib.avt_head_forward(None, 0, None, 0, 0, 512, 196, None, 0, 0.65, 0.4, 0.03, 0.07, 100.0,
                              1, 1, 0, 1, None, None, None, None, None, None, None, None, None, None,
                              0, 0.0, None, ctypes.c_size_t(0), None)
    assert rc == -1 and b"N_v" in lib.avt_last_error()
    assert lib.avt_head_workspace_bytes(128, 128, 512, 196, 1, 1) > 128 * 196 * 512 * 2
    assert lib.avt_head_workspace_bytes(128, 128, 512, 196, 0, 0) < 1 << 20
    assert lib.avt_eval_maps(None, 0, None, None, None, 0, 0, None, None, None, None, None, None, None) == 0
    assert lib.avt_eval_maps(None, 4, None, None, None, 0, 0, None, None, None, None, None, None, None) == -1


def test_product_path_refuses_cpu_tensors():
    import torch
    from audio_visual_tubes_b200 import HeadConfig, hardway_head, evaluate_maps
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        hardway_head(torch.zeros(2, 512, 196), torch.zeros(2, 512), HeadConfig())
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            evaluate_maps(np.zeros((1, 14, 14), np.float32), boxes=[[[0, 0, 1, 1]]])


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "audio_visual_tubes_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("eval_oracle", "oracle") or "import oracle" not in txt
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f


def test_box_packing_matches_oracle():
    from audio_visual_tubes_b200.evaluation import boxes_to_pixel_rects, pack_boxes
    from oracle import eval_oracle as eo
    vg = [[0.1, 0.2, 0.55, 0.9], [-0.3, 0.0, 1.2, 0.5]]
    fl = [[20, 30, 200, 180], [0, 0, 256, 256]]
    assert boxes_to_pixel_rects(vg, "vggss") == eo.boxes_to_pixel_rects(vg, "vggss")
    assert boxes_to_pixel_rects(fl, "flickr") == eo.boxes_to_pixel_rects(fl, "flickr")
    b, c = pack_boxes([vg, [], vg[:1]], "vggss")
    assert b.shape == (3, 2, 4) and c.tolist() == [2, 0, 1]
    assert b[0, 1].tolist() == [0, 0, 268, 112]
