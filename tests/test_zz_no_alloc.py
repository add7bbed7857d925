"""avt_mtc_static: the allocation-free variant of avt_mtc (scratch in a fixed device array inside the
library; no counterpart in the reference).  Collected
last on purpose (file name).

Checked through the C ABI against the existing entry point avt_mtc on the same masks, bit for bit
(both run the same integer counts and the same float32 accumulation), on two calls in a row with
different frame counts so that what the first call leaves in the shared array can not be what the
second one reads.  Passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest


def _lib():
    from audio_visual_tubes_b200 import _lib as L
    return L, L.load()


@pytest.mark.gpu
def test_mtc_static_matches_mtc_on_two_calls_in_a_row():
    import torch
    L, lib = _lib()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(77)
    out_s = torch.full((1,), -1.0, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        for t, p_on in ((16, 0.5), (5, 0.2), (16, 0.7)):          # long, short (stale tail behind it), long again
            masks = (torch.rand(t, 224, 224, device=dev, generator=g) < p_on).to(torch.uint8).contiguous()
            pi = torch.empty(t - 1, dtype=torch.int32, device=dev)
            pu = torch.empty(t - 1, dtype=torch.int32, device=dev)
            out_r = torch.empty(1, dtype=torch.float32, device=dev)
            L.check(lib.avt_mtc(L.ptr(masks), t, L.ptr(pi), L.ptr(pu), L.ptr(out_r), L.stream_ptr()), "avt_mtc")
            out_s.fill_(-1.0)
            L.check(lib.avt_mtc_static(L.ptr(masks), t, L.ptr(out_s), L.stream_ptr()), "avt_mtc_static")
            torch.cuda.synchronize(dev)
            a, b = float(out_r[0]), float(out_s[0])
            assert 0.0 < a < 1.0, a
            assert a == b, (t, a, b)                               # bit-exact, as the eval tests require
            # and against plain torch arithmetic of utils.py:311-318 on the same masks (1e-6, float32 sums)
            m = masks.view(t, -1).bool()
            inter = (m[:-1] & m[1:]).sum(1).double()
            union = m[1:].sum(1).double() + (m[:-1] & ~m[1:]).sum(1).double()
            ref = float((inter / union).float().sum() / (t - 1))
            assert abs(b - ref) < 1e-6, (b, ref)


@pytest.mark.gpu
def test_mtc_static_rejects_sizes_beyond_its_scratch():
    import torch
    L, lib = _lib()
    dev = torch.device("cuda", 0)
    masks = torch.zeros(2, 224, 224, dtype=torch.uint8, device=dev)
    out = torch.zeros(1, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        # the size check precedes any launch, so the (too short) masks buffer is never touched
        rc = lib.avt_mtc_static(L.ptr(masks), L.MTC_STATIC_MAX_FRAMES + 1, L.ptr(out), L.stream_ptr())
        assert rc == -1 and "AVT_MTC_STATIC_MAX_FRAMES" in L.last_error()
        assert lib.avt_mtc_static(L.ptr(masks), 1, L.ptr(out), L.stream_ptr()) == -1
        assert lib.avt_mtc_static(None, 4, L.ptr(out), L.stream_ptr()) == -1
    torch.cuda.synchronize(dev)


def test_mtc_static_is_declared_and_exported():
    L, lib = _lib()
    assert hasattr(lib, "avt_mtc_static") and L.SIGNATURES["avt_mtc_static"][1] == [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    # argument checks need no GPU
    assert lib.avt_mtc_static(None, 4, None, None) == -1
    assert lib.avt_mtc_static(C.c_void_p(256), L.MTC_STATIC_MAX_FRAMES + 1, C.c_void_p(256), None) == -1
    assert "AVT_MTC_STATIC_MAX_FRAMES" in L.last_error()
