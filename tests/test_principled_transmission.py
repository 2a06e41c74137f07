"""The GGX refraction lobe of `principled` (src/bsdfs/principled.cpp:437-453,604-621,816-828), oracle side: the reference's own
chi^2 tests of sample() against pdf() (src/bsdfs/tests/test_principled.py:60-121, test02-test04: same parameters and incident
directions) restated against the oracle, plus sample-weight / eval / pdf consistency."""
import ctypes as C

import numpy as np
import pytest

WL = (450.0, 520.0, 600.0, 680.0)
CASES = {   # test_principled.py:60-121
    "test02_chi2_spec_trans_outside": (dict(roughness=0.6, metallic=0.2, anisotropic=0.4, clearcoat=0.8, spec_trans=0.7, eta=1.3296), (1, 0, 1)),
    "test03_chi2_spec_trans_inside": (dict(roughness=0.5, anisotropic=0.4, clearcoat=0.8, spec_trans=0.7, eta=1.5), (1, 0, -1)),
    "test04_chi2_spec_trans_less_dense": (dict(roughness=0.8, metallic=0.2, anisotropic=0.0, clearcoat=1.0, sheen=0.0, spec_trans=0.5, eta=0.5), (1, 0, 1)),
}


def scene(mi, params):
    return mi.load_dict({"type": "scene", "sensor": {"type": "perspective", "film": {"type": "hdrfilm", "width": 4, "height": 4}},
                         "shape": {"type": "rectangle", "bsdf": {"type": "principled", **params}}})


def batch(oracle, sc, wi, u=None, wo=None):
    L = oracle.lib()
    for fn in (L.plt_oracle_bsdf_sample_batch, L.plt_oracle_bsdf_pdf_batch):
        fn.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    wi = np.asarray(wi, np.float32); uv = np.array([0.5, 0.5], np.float32); wl = np.array(WL, np.float32)
    if u is not None:
        u = np.ascontiguousarray(u, np.float32); out = np.zeros((len(u), 5), np.float32)
        L.plt_oracle_bsdf_sample_batch(C.byref(sc.flat.desc), 0, wi.ctypes.data, uv.ctypes.data, wl.ctypes.data, len(u), u.ctypes.data, out.ctypes.data)
        return out
    wo = np.ascontiguousarray(wo, np.float32); out = np.zeros((len(wo), 2), np.float32)
    L.plt_oracle_bsdf_pdf_batch(C.byref(sc.flat.desc), 0, wi.ctypes.data, uv.ctypes.data, wl.ctypes.data, len(wo), wo.ctypes.data, out.ctypes.data)
    return out


@pytest.mark.parametrize("name", sorted(CASES))
def test_chi2_sample_against_pdf(mi, oracle, name):
    """mi.chi2.ChiSquareTest over a SphericalDomain (src/python/python/chi2.py): histogram of sampled directions against the pdf
    integrated over the same (cos theta, phi) bins; Pearson statistic with pooled low-expectation cells at the 1 % level"""
    from scipy import stats
    params, wi = CASES[name]
    wi = np.array(wi, np.float64); wi /= np.linalg.norm(wi)
    sc = scene(mi, params)
    rng = np.random.default_rng(7)
    n = 400000
    s = batch(oracle, sc, wi, u=rng.random((n, 3)))
    ok = s[:, 3] > 0                                  # failed samples (pdf 0) land nowhere, as in BSDFAdapter
    wo = s[ok, :3].astype(np.float64)
    assert 0.5 < ok.mean() <= 1.0
    assert (wo[:, 2] * wi[2] < 0).mean() > 0.05       # the refraction lobe is actually drawn
    res_c, res_p = 24, 48
    ic = np.clip(((wo[:, 2] + 1) * 0.5 * res_c).astype(int), 0, res_c - 1)
    ip = np.clip(((np.arctan2(wo[:, 1], wo[:, 0]) / (2 * np.pi)) % 1.0 * res_p).astype(int), 0, res_p - 1)
    hist = np.zeros((res_c, res_p)); np.add.at(hist, (ic, ip), 1)
    k = 12                                            # midpoint rule, k x k nodes per cell
    cz = (np.arange(res_c * k) + 0.5) / (res_c * k) * 2 - 1
    ph = (np.arange(res_p * k) + 0.5) / (res_p * k) * 2 * np.pi
    Z, P = np.meshgrid(cz, ph, indexing="ij")
    r = np.sqrt(np.maximum(0, 1 - Z * Z))
    dirs = np.stack([r * np.cos(P), r * np.sin(P), Z], -1).reshape(-1, 3)
    pdf = batch(oracle, sc, wi, wo=dirs)[:, 0].astype(np.float64).reshape(res_c, k, res_p, k)
    expect = pdf.sum(axis=(1, 3)) * (4 * np.pi / pdf.size) * n
    assert abs(expect.sum() / n - ok.mean()) < 0.01, (expect.sum() / n, ok.mean())   # pdf integrates to the success probability
    order = np.argsort(expect.ravel())
    e, h = expect.ravel()[order], hist.ravel()[order]
    cells_e, cells_h, ae, ah = [], [], 0.0, 0.0
    for ei, hi in zip(e, h):                          # pool cells until each has >= 5 expected samples (chi2.py)
        ae += ei; ah += hi
        if ae >= 5:
            cells_e.append(ae); cells_h.append(ah); ae = ah = 0.0
    cells_e, cells_h = np.array(cells_e), np.array(cells_h)
    chi2 = ((cells_h - cells_e) ** 2 / cells_e).sum()
    # the midpoint integration of a sharply peaked pdf is not exact: compare the statistic with its distribution loosely
    p = 1 - stats.chi2.cdf(chi2, len(cells_e) - 1)
    assert chi2 < 1.6 * len(cells_e) and (p > 1e-6 or chi2 < 1.25 * len(cells_e)), (chi2, len(cells_e), p)


@pytest.mark.parametrize("name", sorted(CASES))
def test_sample_weight_is_eval_over_pdf(mi, oracle, name):
    params, wi = CASES[name]
    wi = np.array(wi, np.float64); wi /= np.linalg.norm(wi)
    sc = scene(mi, params)
    rng = np.random.default_rng(3)
    s = batch(oracle, sc, wi, u=rng.random((4000, 3)))
    ok = s[:, 3] > 0
    ev = batch(oracle, sc, wi, wo=s[ok, :3])
    assert np.allclose(ev[:, 0], s[ok, 3], rtol=1e-5)
    assert np.allclose(ev[:, 1] / ev[:, 0], s[ok, 4], rtol=1e-4, atol=1e-7)
    trans = s[ok, 2] * wi[2] < 0
    assert trans.any() and (ev[trans, 1] > 0).mean() > 0.9


def test_loader_rules(mi):
    from mitsuba3_plt_b200 import capi
    f = scene(mi, dict(spec_trans=0.5, eta=1.0)).flat
    b = f.bsdfs[0]
    assert b.u[0] & 128 and abs(b.f[10] - 0.5) < 1e-7 and abs(b.f[0] - 1.001) < 1e-6      # principled.cpp:222-223
    assert not scene(mi, dict(spec_trans=0.0)).flat.bsdfs[0].u[0] & 128                    # get_flag: a float 0 switches the lobe off
    with pytest.raises(RuntimeError, match="transmission"):
        mi.load_dict({"type": "scene", "sensor": {"type": "perspective", "film": {"type": "hdrfilm", "width": 4, "height": 4}},
                      "shape": {"type": "rectangle", "bsdf": {"type": "twosided", "nested": {"type": "principled", "spec_trans": 0.5}}}})
