"""ctypes binding of oracle/libplt_oracle.so -- the CPU oracle (test infrastructure only; see oracle/plt_oracle.cpp).
Imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs, never by the product package."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
SO = os.path.join(ORACLE_DIR, "libplt_oracle.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "libplt_oracle.so"])


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(ORACLE_DIR, "plt_oracle.cpp")
        if not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(src):
            build()
        from mitsuba3_plt_b200 import capi
        L = C.CDLL(SO)
        L.plt_oracle_render.restype = C.c_double
        L.plt_oracle_render.argtypes = [C.POINTER(capi.SceneDesc), C.POINTER(capi.RenderParams), C.c_void_p,
                                        C.POINTER(capi.SampleDump), C.POINTER(capi.Stats), C.c_int]
        L.plt_oracle_develop.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        L.plt_oracle_develop_n.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
        f, i, u = C.c_float, C.c_int, C.c_uint32
        PF = C.POINTER(C.c_float)
        for name, res, args in [
            ("plt_oracle_bessel_j", f, [f, i]), ("plt_oracle_erf", f, [f]), ("plt_oracle_erfinv", f, [f]),
            ("plt_oracle_fresnel_conductor", f, [f, f, f]), ("plt_oracle_fresnel", None, [f, f, PF]),
            ("plt_oracle_mf_eval", f, [u, f, f, PF]), ("plt_oracle_mf_g1", f, [u, f, f, PF, PF]),
            ("plt_oracle_mf_sample", None, [u, f, f, i, PF, PF, PF]),
            ("plt_oracle_grating_intensity", f, [PF, u, u, i, i, PF, f]),
            ("plt_oracle_grating_diffract", i, [PF, u, u, i, i, PF, f, PF]),
            ("plt_oracle_grating_lobe_pdf", f, [PF, u, u, i, i, PF, f]),
            ("plt_oracle_grating_sample_lobe", None, [PF, u, u, PF, PF, f, C.POINTER(C.c_int), PF]),
            ("plt_oracle_grating_alpha", f, [PF, u, u, PF, f]),
            ("plt_oracle_sample_wavelengths", None, [f, PF, PF]),
            ("plt_oracle_spectrum_to_srgb", None, [PF, PF, PF, PF]),
            ("plt_oracle_srgb_model_eval", f, [PF, f]),
            ("plt_oracle_tea32", None, [u, u, i, C.POINTER(u)]),
            ("plt_oracle_pcg32_floats", None, [C.c_uint64, C.c_uint64, i, PF, C.POINTER(u)]),
            ("plt_oracle_sampler_floats", None, [u, u, i, PF]),
        ]:
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def farr(*v):
    return (C.c_float * len(v))(*v)


def render(flat, spp, seed=0, max_depth=7, rr_depth=50, spp_begin=0, spp_count=None, crop=None, dump=False,
           film=True, n_threads=0, polarized=False, stokes=False):
    """Run the oracle on a FlatScene.  Returns dict(film_rgbw, image, wavelengths, radiance, n_rays, stats, seconds).
    polarized: the `*_spectral_polarized` variant (Mueller-matrix Spectrum); stokes: the `stokes` wrapper integrator around it
    (film_rgbw then has 16 channels R G B W S0.RGB .. S3.RGB, image 15, and the dump gains 'stokes' (n, 4, 4))."""
    from mitsuba3_plt_b200 import capi
    L = lib()
    d = flat.desc
    W, H = d.film.width * d.n_sensors, d.film.height
    prm = capi.RenderParams()
    prm.seed, prm.spp, prm.spp_begin = (seed + flat.base_seed) & 0xFFFFFFFF, spp, spp_begin
    prm.spp_count = spp if spp_count is None else spp_count
    prm.max_depth, prm.rr_depth = max_depth, rr_depth
    cx, cy, cw, ch = crop if crop else (0, 0, W, H)
    prm.crop_x, prm.crop_y, prm.crop_w, prm.crop_h = cx, cy, cw, ch
    if polarized:
        prm.flags |= capi.RENDER_POLARIZED | (capi.RENDER_STOKES if stokes else 0)
    if getattr(flat, "rgb", False):   # scene flattened for an RGB variant
        prm.flags |= capi.RENDER_RGB
    nch = capi.film_channels(prm.flags)
    out = {}
    dmp = None
    if dump:
        n = cw * ch * prm.spp_count
        out["wavelengths"] = np.zeros((n, 4), np.float32); out["radiance"] = np.zeros((n, 4), np.float32)
        out["n_rays"] = np.zeros((n, 2), np.uint32)
        dmp = capi.SampleDump(capi.fptr(out["wavelengths"]), capi.fptr(out["radiance"]), capi.uptr(out["n_rays"]), n)
        if polarized:
            out["stokes"] = np.zeros((n, 4, 4), np.float32)
            dmp.stokes = capi.fptr(out["stokes"])
    fbuf = np.zeros((H, W, nch), np.float32) if film else None
    if not film:
        prm.flags |= capi.RENDER_NO_SPLAT
    st = capi.Stats()
    secs = L.plt_oracle_render(C.byref(d), C.byref(prm), fbuf.ctypes.data if film else None,
                               C.byref(dmp) if dmp else None, C.byref(st), n_threads)
    out["seconds"], out["stats"] = secs, st.as_dict()
    if film:
        img = np.zeros((H, W, nch - 1), np.float32)
        L.plt_oracle_develop_n(fbuf.ctypes.data, W * H, nch, img.ctypes.data)
        out["film_rgbw"], out["image"] = fbuf, img
    return out
