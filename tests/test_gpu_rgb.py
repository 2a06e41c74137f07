"""GPU parity of the RGB variants (`cuda_ad_rgb` / `cuda_ad_rgb_polarized`, the second mode of
scripts/rendering/main-headless.py:131-133) against the oracle's RGB mode: per-sample radiance (lanes R, G, B; lane 3 repeats B),
the path's live `sampling_wavelengths` (plt.py:64-70), ray counts and the image, with roughgrating on its RGB fallback
(roughgrating.cpp:499-506,716-839,1027-1031)."""
import numpy as np
import pytest

from parity import MAX_RMSE, MIN_FRACTION, max_finiteness_mismatch, nonfinite_counts, relative_mse, sample_agreement

pytestmark = pytest.mark.gpu

CASES = [("cbox", 64, 8, 0), ("cbox-grating", 64, 8, 0), ("cbox-grating", 128, 16, 3), ("gratingcap", 96, 8, 0),
         ("gratings", 96, 16, 1),                 # the reference's scenes/gratings/gratings.xml: `directionalspot` emitter
         ("veach-mis", 64, 8, 1), ("cbox-grating:envmap", 48, 8, 5), ("disk", 64, 8, 2)]


@pytest.fixture()
def rgb(mi):
    mi.set_variant("cuda_ad_rgb")
    yield mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
    mi.set_variant("cuda_spectral")


@pytest.fixture()
def rgb_polarized(mi):
    mi.set_variant("cuda_ad_rgb_polarized")
    yield mi.load_dict({"type": "stokes", "nested": {"type": "plt", "max_depth": 7, "rr_depth": 50}})
    mi.set_variant("cuda_spectral")


@pytest.mark.parametrize("name,res,spp,seed", CASES)
def test_rgb_per_sample(mi, oracle, rgb, name, res, spp, seed):
    from mitsuba3_plt_b200 import scenes
    from test_gpu_parity import ray_count_agreement
    sc = scenes.load(name, resx=res, resy=res)
    assert sc.flat.rgb
    dump = {}
    img = rgb.render(sc, spp=spp, seed=seed, dump=dump)
    ref = oracle.render(sc.flat, spp, seed=seed, dump=True)
    assert (dump["wavelengths"] == ref["wavelengths"]).all()          # sampling wavelengths: bit-equal draws * 680
    assert (dump["wavelengths"][:, 2] == dump["wavelengths"][:, 3]).all() and dump["wavelengths"].max() < 680.0
    assert (dump["radiance"][:, 2] == dump["radiance"][:, 3]).all() or np.isnan(dump["radiance"]).any()
    frac = sample_agreement(dump["radiance"], ref["radiance"])
    assert frac >= MIN_FRACTION, f"{name}: only {frac * 100:.4f}% of samples within 1e-4"
    nf_a, nf_b, nf_both = nonfinite_counts(dump["radiance"], ref["radiance"])
    assert abs(nf_a - nf_b) <= max(2, 0.02 * nf_b) and nf_both >= min(nf_a, nf_b) - max(2, 0.02 * nf_b), (nf_a, nf_b, nf_both)
    ray_count_agreement(name, dump["n_rays"], ref["n_rays"])
    for k in ("closest_rays", "shadow_rays", "vertices", "grating_vertices"):
        a, b = rgb.last_stats[k], ref["stats"][k]
        assert abs(a - b) <= max(4, 1e-3 * b), (k, a, b)
    rmse, mism = relative_mse(img, ref["image"])
    assert rmse <= MAX_RMSE and mism <= max_finiteness_mismatch(img.shape[0] * img.shape[1]), (rmse, mism)


@pytest.mark.parametrize("name,res,spp,seed", [("cbox-grating", 64, 8, 0), ("gratings", 96, 16, 1), ("gratingcap", 64, 8, 2)])
def test_rgb_polarized_stokes(mi, oracle, rgb_polarized, name, res, spp, seed):
    """`cuda_ad_rgb_polarized` + stokes{plt}: what main-headless.py runs without --spectral"""
    from mitsuba3_plt_b200 import scenes
    from test_gpu_polarized import stokes_agreement
    sc = scenes.load(name, resx=res, resy=res)
    dump = {}
    img = rgb_polarized.render(sc, spp=spp, seed=seed, dump=dump)
    ref = oracle.render(sc.flat, spp, seed=seed, dump=True, polarized=True, stokes=True)
    assert img.shape == ref["image"].shape and img.shape[-1] == 15
    assert (dump["wavelengths"] == ref["wavelengths"]).all()
    assert sample_agreement(dump["radiance"], ref["radiance"]) >= MIN_FRACTION
    assert stokes_agreement(dump["stokes"], ref["stokes"]) >= MIN_FRACTION
    rmse, mism = relative_mse(img[..., :3], ref["image"][..., :3])
    assert rmse <= MAX_RMSE and mism <= 0.01, (rmse, mism)
    s0 = np.abs(ref["image"][..., 3:6]).mean()
    for c in range(3, 15):
        d = np.nanmean(np.abs(img[..., c] - ref["image"][..., c]))
        assert d <= 0.01 * max(s0, 1e-6), (c, d, s0)
    # in an RGB variant the S0 AOV is the radiance itself (stokes.cpp:114-115)
    fin = np.isfinite(img).all(-1)
    assert np.allclose(img[fin][:, 3:6], img[fin][:, :3], rtol=1e-5, atol=1e-7)


def test_colour_mode_is_checked(mi, rgb):
    """a scene flattened for a spectral variant cannot be rendered by an RGB one (and vice versa), at the Python layer and at the ABI"""
    import ctypes as C
    from mitsuba3_plt_b200 import capi, scenes
    sc = scenes.load("cbox", resx=16, resy=16)           # loaded under cuda_ad_rgb
    mi.set_variant("cuda_spectral")
    with pytest.raises(RuntimeError, match="variant"):
        mi.load_dict({"type": "plt"}).render(sc, spp=1)
    prm = capi.RenderParams()
    prm.spp = prm.spp_count = 1
    prm.max_depth, prm.rr_depth = 4, 4
    out = np.zeros((16, 16, 4), np.float32)
    rc = capi.lib().pltb_render(mi.context()._h, sc.device(), C.byref(prm), out.ctypes.data, None, None)
    assert rc == -1 and b"RGB" in capi.lib().pltb_last_error()
