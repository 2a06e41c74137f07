"""RGB variants, CPU side: the load-time colour conversions of an RGB variant (src/core/xml.cpp:1146-1260,
src/spectra/{srgb,d65}.cpp, src/core/spectrum.cpp:85-130), the oracle's RGB mode and the RGB fallback of roughgrating
(src/bsdfs/roughgrating.cpp:499-506,716-839,1027-1031)."""
import ctypes as C

import numpy as np
import pytest


@pytest.fixture()
def rgb_variant(mi):
    mi.set_variant("cuda_ad_rgb")
    yield mi
    mi.set_variant("cuda_spectral")


def test_spectrum_list_to_srgb_whitepoint():
    """xml.cpp:1196-1199: "Values should be scaled so that integrating the spectrum against the CIE curves and converting to sRGB
    yields (1, 1, 1) for D65 in non-spectral modes" -- a unit reflectance under the D65 product, and the D65 curve itself as an
    emitter, both come out white"""
    from mitsuba3_plt_b200 import spectra
    wl = np.array([360.0, 830.0], np.float32)
    refl = spectra.spectrum_list_to_srgb(wl, np.array([1.0, 1.0], np.float32) * np.float32(spectra.MI_CIE_Y_NORMALIZATION), bounded=True, d65=True)
    assert np.allclose(refl, 1.0, atol=2e-3), refl
    tabs = spectra.cie_tables()
    lam = np.linspace(360, 830, 95).astype(np.float32)
    d65 = (tabs["d65"] * np.float32(spectra.MI_CIE_D65_NORMALIZATION) * np.float32(spectra.MI_CIE_Y_NORMALIZATION)).astype(np.float32)
    em = spectra.spectrum_list_to_srgb(lam, d65, bounded=False, d65=False)
    assert np.allclose(em, 1.0, atol=2e-3), em
    # a uniform emitter spectrum is the purple-ish xyz_to_srgb(1, 1, 1) (xml.cpp:1176-1188)
    white = spectra.xyz_to_srgb(np.ones(3, np.float32))
    assert np.allclose(white, [1.204794, 0.948291, 0.908916], atol=1e-5)


def test_rgb_flatten_of_the_reference_scenes(rgb_variant):
    from mitsuba3_plt_b200 import capi, scenes
    sc = scenes.load("cbox", resx=16, resy=16)
    f = sc.flat
    assert f.rgb and all(t.type in (capi.TEX_UNIFORM, capi.TEX_RGB) for t in f.textures)
    light = f.textures[f.emitters[0].radiance]
    assert light.type == capi.TEX_RGB and list(light.p)[:3] == [10.0, 5.0, 5.0]      # cbox.xml:71-78, d65 -> srgb(color * scale)
    g = scenes.load("cbox-grating", resx=16, resy=16).flat
    bm = [t for t in g.textures if t.type == capi.TEX_BITMAP]
    assert bm and all(t.flags & capi.TEXFLAG_RGB for t in bm)
    data = np.asarray(g.a_tex_data[bm[0].offset:bm[0].offset + bm[0].count])
    assert data.min() >= 0.0 and data.max() <= 1.0                                      # linear RGB texels, not rgb2spec coefficients
    s = scenes.load("gratings", resx=16, resy=16).flat                                   # scenes/gratings/gratings.xml
    e = s.emitters[0]
    assert e.type == capi.EMITTER_DIRECTIONALSPOT and abs(e.env_pad[0] - np.sin(np.float32(0.003))) < 1e-9
    assert np.allclose(list(e.direction), [0.0, -0.707107, 0.707107], atol=1e-6) and e.bsphere_radius > 0
    with pytest.raises(RuntimeError, match="spectral"):
        rgb_variant.load_dict({"type": "scene", "sensor": {"type": "perspective", "film": {"type": "hdrfilm", "width": 4, "height": 4}},
                               "s": {"type": "sphere", "bsdf": {"type": "diffuse", "reflectance": {"type": "blackbody", "temperature": 3000.0}}}})


def test_lobe_colour_table(oracle):
    """xyz_to_srgb(cie1931_xyz(wl)) of the oracle against a float64 evaluation of the same tables (spectrum.h:236-264,406-412)"""
    from mitsuba3_plt_b200 import spectra
    from mitsuba3_plt_b200.scene import FlatScene
    tabs = spectra.cie_tables()
    cie = np.ascontiguousarray(np.concatenate([tabs["srgb"], tabs["xyz"][1:2], tabs["xyz"][0:1], tabs["xyz"][2:3]]), np.float32)
    M = np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875991, 0.041556], [0.055648, -0.204043, 1.057311]])
    L = oracle.lib()
    L.plt_oracle_wavelength_to_srgb.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
    out = np.zeros(3, np.float32)
    for wl in (360.0, 400.0, 455.5, 550.0, 612.3, 700.0, 830.0):
        L.plt_oracle_wavelength_to_srgb(cie.ctypes.data, wl, out.ctypes.data)
        t = (wl - 360.0) * 94 / 470.0
        i0 = min(int(t), 93)
        xyz = (1 - (t - i0)) * tabs["xyz"][:, i0].astype(np.float64) + (t - i0) * tabs["xyz"][:, i0 + 1]
        assert np.allclose(out, M @ xyz, rtol=2e-5, atol=2e-6), (wl, out, M @ xyz)
    for wl in (100.0, 359.9, 830.1):     # outside [360, 830]: the masked lookup returns 0 (plt.py draws u * 680 nm, so this is common)
        L.plt_oracle_wavelength_to_srgb(cie.ctypes.data, wl, out.ctypes.data)
        assert (out == 0).all()


def test_rgb_fallback_of_roughgrating(oracle, rgb_variant):
    """RoughGrating::wbsdf_eval in an RGB variant (roughgrating.cpp:716-839) on the reference's gratings.xml: only the mirror
    direction survives the angular coherence factor (coherence 6e5 => exp(-0.5 theta^2 1.44e6 / wl_um^2 ...)), every lobe is
    coloured by its wavelength, a wavelength outside the CIE range contributes nothing, and wbsdf_pdf takes k from the FIRST
    sampling wavelength (:1027-1031)."""
    from mitsuba3_plt_b200 import capi, scenes
    sc = scenes.load("gratings", resx=16, resy=16)
    f = sc.flat
    shape = next(i for i, s in enumerate(f.shapes) if f.bsdfs[s.bsdf].type == capi.BSDF_ROUGHGRATING)
    L = oracle.lib()
    L.plt_oracle_bsdf_probe_rgb.argtypes = [C.c_void_p, C.c_int32] + [C.c_void_p] * 5
    arr = lambda *v: np.array(v, np.float32)
    def probe(wi, wo, swl):
        out = np.zeros(7, np.float32)
        a = [arr(*wi), arr(0.3, 0.6), arr(*swl), arr(*wo)]
        L.plt_oracle_bsdf_probe_rgb(C.byref(f.desc), shape, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, a[3].ctypes.data, out.ctypes.data)
        return out
    wi = np.array([0.3, 0.1, 0.0]); wi[2] = np.sqrt(1 - wi[0] ** 2 - wi[1] ** 2)
    mirror = [-wi[0], -wi[1], wi[2]]
    swl = (450.0, 550.0, 650.0)
    m = probe(wi, mirror, swl)
    # at the mirror direction only lobe (0, 0) is inside its window: intensity = multiplier (10) at each of the 3 wavelengths, coherence 1
    from mitsuba3_plt_b200 import spectra
    tabs = spectra.cie_tables()
    def colour(wl):
        t = (wl - 360.0) * 94 / 470.0; i0 = min(int(t), 93)
        xyz = (1 - (t - i0)) * tabs["xyz"][:, i0].astype(np.float64) + (t - i0) * tabs["xyz"][:, i0 + 1]
        return np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875991, 0.041556], [0.055648, -0.204043, 1.057311]]) @ xyz
    expect = 10.0 * sum(colour(w) for w in swl)
    assert np.allclose(m[4:7], expect, rtol=1e-4), (m[4:7], expect)
    assert np.allclose(m[:3], m[4:7], rtol=1e-5)          # eta = 0, k = 1: the conductor's Fresnel factor is 1
    off = np.array(mirror) + [0.02, 0.0, 0.0]; off /= np.linalg.norm(off)
    assert (probe(wi, off, swl)[4:7] == 0).all()           # 20 mrad off the mirror direction: the coherence factor underflows
    out_of_range = probe(wi, mirror, (100.0, 200.0, 300.0))
    assert (out_of_range[4:7] == 0).all()                  # colours are 0 below 360 nm
    k = 2 * np.pi / (450.0 * 1e-3)
    assert np.isclose(m[3], np.exp(-(wi[2] * 0.04 * k) ** 2), rtol=1e-4)
    assert not np.isclose(probe(wi, mirror, (650.0, 550.0, 450.0))[3], m[3])


def test_rgb_and_spectral_variants_agree_in_expectation(oracle, mi):
    """"consistency between RGB and spectral variants of Mitsuba" (xml.cpp:1178-1183): the cbox (no grating) rendered by the oracle in
    both colour modes gives the same image up to the spectral-upsampling difference; same geometry, same ray statistics"""
    from mitsuba3_plt_b200 import scenes
    res = {}
    try:
        for v in ("cuda_ad_spectral", "cuda_ad_rgb"):
            mi.set_variant(v)
            sc = scenes.load("cbox", resx=48, resy=48)
            res[v] = oracle.render(sc.flat, 32, seed=3)
    finally:
        mi.set_variant("cuda_spectral")
    a, b = res["cuda_ad_spectral"], res["cuda_ad_rgb"]
    ma, mb = np.nanmean(a["image"], axis=(0, 1)), np.nanmean(b["image"], axis=(0, 1))
    assert np.allclose(ma, mb, rtol=0.12), (ma, mb)
    for k in ("closest_rays", "shadow_rays", "vertices"):
        assert abs(a["stats"][k] - b["stats"][k]) <= 0.03 * a["stats"][k], k
