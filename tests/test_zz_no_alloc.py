"""pltb_film_develop_noalloc (include/pltb.h): the develop step without any device allocation, its scratch in a static __device__
array.  No counterpart in the reference.  Checked through the C ABI against the
existing entry point -- pltb_render with PLTB_RENDER_DEVELOP -- on calls in a row that share the static array (different film
sizes and channel counts, so stale contents of an earlier call would show).  Collected last on
purpose."""
import ctypes as C

import numpy as np
import pytest


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_noalloc_rejects_bad_sizes_before_touching_the_device():
    """the limit and argument checks come before any CUDA call: they hold without a GPU"""
    from mitsuba3_plt_b200 import capi
    L = capi.lib()
    x = np.zeros(16, np.float32)
    side = 1025   # 1025 * 1025 * 4 floats > PLTB_NOALLOC_MAX_FLOATS
    assert side * side * 4 > capi.NOALLOC_MAX_FLOATS
    assert L.pltb_film_develop_noalloc(x.ctypes.data, side, side, 4, 0, x.ctypes.data, None) == -5
    msg = L.pltb_last_error().decode()
    assert "PLTB_NOALLOC_MAX_FLOATS" in msg and str(capi.NOALLOC_MAX_FLOATS) in msg
    assert L.pltb_film_develop_noalloc(x.ctypes.data, 2, 2, 5, 0, x.ctypes.data, None) == -1
    assert L.pltb_film_develop_noalloc(None, 2, 2, 4, 0, x.ctypes.data, None) == -1
    assert L.pltb_film_develop_noalloc(x.ctypes.data, 2, 2, 4, 8, x.ctypes.data, None) == -1
    assert L.pltb_film_develop_noalloc(x.ctypes.data, 0, 7, 4, 0, x.ctypes.data, None) == 0   # empty film: nothing to do


@pytest.mark.gpu
def test_noalloc_develop_equals_pltb_render_develop(mi):
    import torch
    from mitsuba3_plt_b200 import capi, scenes
    L = capi.lib()
    ctx = mi.context()

    def render(sc, W, H, flags, nch_out, spp=8):
        prm = capi.RenderParams(seed=3, spp=spp, spp_begin=0, spp_count=spp, max_depth=7, rr_depth=50, flags=flags)
        out = np.empty((H, W, nch_out), np.float32)
        capi.check(L.pltb_render(ctx._h, sc.device(ctx), C.byref(prm), out.ctypes.data, None, None))
        return out

    def device_accum(n_expected):
        p, n = C.c_void_p(), C.c_uint64()
        capi.check(L.pltb_film_device_ptr(ctx._h, C.byref(p), C.byref(n)))
        assert n.value == n_expected
        return p

    # call 1: 96 x 64 RGBW, input = the DEVICE accumulator pltb_render just developed -> bit-identical image
    W, H = 96, 64
    sc = scenes.load("cbox-grating", resx=W, resy=H)
    want = render(sc, W, H, capi.RENDER_DEVELOP, 3)
    got = np.full((H, W, 3), np.nan, np.float32)
    capi.check(L.pltb_film_develop_noalloc(device_accum(W * H * 4), W, H, 4, 0, got.ctypes.data, None))
    assert np.isfinite(want).mean() > 0.9 and np.array_equal(_bits(got), _bits(want))

    # call 2, right after: a smaller film from HOST memory (a raw render).  Exact against the develop arithmetic, and equal to
    # pltb_render's developed image of the same samples up to the order of the film's float atomics (the tolerance of
    # test_gpu_parity.py::test_strip_splat_equals_per_tap_atomics)
    W2, H2 = 40, 24
    sc2 = scenes.load("cbox", resx=W2, resy=H2)
    raw = render(sc2, W2, H2, 0, 4)
    want2 = render(sc2, W2, H2, capi.RENDER_DEVELOP, 3)
    got2 = np.full((H2, W2, 3), np.nan, np.float32)
    capi.check(L.pltb_film_develop_noalloc(raw.ctypes.data, W2, H2, 4, 0, got2.ctypes.data, None))
    w = raw[..., 3:4]
    with np.errstate(all="ignore"):
        exact = raw[..., :3] / np.where(w == 0, np.float32(1), w)
    assert np.array_equal(_bits(got2), _bits(exact))
    fin = np.isfinite(got2).all(-1) & np.isfinite(want2).all(-1)
    assert fin.mean() > 0.9
    np.testing.assert_allclose(got2[fin], want2[fin], rtol=2e-4, atol=1e-4)

    # call 3: the 16-channel Stokes film (15 developed channels), asynchronous, device input and DEVICE output
    flags = capi.RENDER_POLARIZED | capi.RENDER_STOKES | capi.RENDER_DEVELOP
    want3 = render(sc2, W2, H2, flags, 15, spp=4)
    out_dev = torch.full((H2, W2, 15), float("nan"), device="cuda:0")
    torch.cuda.synchronize()
    capi.check(L.pltb_film_develop_noalloc(device_accum(W2 * H2 * 16), W2, H2, 16, capi.DEVELOP_ASYNC, C.c_void_p(out_dev.data_ptr()), None))
    torch.cuda.synchronize()
    assert np.array_equal(_bits(out_dev.cpu().numpy()), _bits(want3))

    # call 4: call 1's input again -- the array held other films in between
    want4 = render(sc, W, H, capi.RENDER_DEVELOP, 3)
    got4 = np.full((H, W, 3), np.nan, np.float32)
    capi.check(L.pltb_film_develop_noalloc(device_accum(W * H * 4), W, H, 4, 0, got4.ctypes.data, None))
    assert np.array_equal(_bits(got4), _bits(want4))
    # the largest film the array holds
    side = 1024
    big = np.random.default_rng(0).random((side, side, 4), dtype=np.float32)
    big[::7, ::5, 3] = 0
    got5 = np.empty((side, side, 3), np.float32)
    capi.check(L.pltb_film_develop_noalloc(big.ctypes.data, side, side, 4, 0, got5.ctypes.data, None))
    assert np.array_equal(got5, big[..., :3] / np.where(big[..., 3:4] == 0, np.float32(1), big[..., 3:4]))
