"""The reference's own plugin tests, run against the oracle.

Each test restates the inputs of one reference test and checks the oracle against what that test asserts: the literal
Mitsuba-0.6 tables of src/render/tests/test_microfacet.py (extracted by oracle/gen_golden_tables.py into
tests/golden/ref_test_tables.json), and the closed-form expectations of the diffuse / twosided / area / perspective / regular /
irregular tests.  The reference tests run in scalar_rgb or *_spectral variants; the oracle is 4-wavelength spectral, so RGB
expectations are checked on grey inputs, where the spectral upsampling is exact."""
import ctypes as C
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TAB = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_test_tables.json")))
MF = "src/render/tests/test_microfacet.py::"
BECKMANN, GGX = 0, 1
PF = C.POINTER(C.c_float)
WL = (450.0, 520.0, 590.0, 660.0)


def table(test, k):
    return np.array(TAB[MF + test][k]["values"], np.float64)


@pytest.fixture(scope="module")
def L(oracle):
    lib = oracle.lib()
    lib.plt_oracle_bsdf_probe.argtypes = [C.c_void_p, C.c_int32, PF, PF, PF, PF, PF, PF]
    lib.plt_oracle_area_emitter_probe.argtypes = [C.c_void_p, C.c_uint32, PF, PF, PF, PF]
    lib.plt_oracle_sensor_ray.argtypes = [C.c_void_p, C.c_uint32, PF, PF]
    lib.plt_oracle_tex_probe.argtypes = [C.c_void_p, C.c_int32, PF, PF, PF]
    return lib


def sph(theta, phi):
    return np.stack([np.cos(phi) * np.sin(theta), np.sin(phi) * np.sin(theta), np.cos(theta)], -1).astype(np.float32)


def mf_eval(L, oracle, typ, au, av, v):
    return np.array([L.plt_oracle_mf_eval(typ, au, av, oracle.farr(*x)) for x in v])


def mf_g1(L, oracle, typ, au, av, v):
    return np.array([L.plt_oracle_mf_g1(typ, au, av, oracle.farr(*x), oracle.farr(0, 0, 1)) for x in v])


def test_microfacet_beckmann_eval_pdf_tables(L, oracle):
    """test_microfacet.py:15-88 (test02_eval_pdf_beckmann): pdf(wi, m) = eval(m) cos_theta(m) without visible-normal sampling"""
    steps = 20
    v = sph(np.linspace(0, np.pi, steps, dtype=np.float32), np.full(steps, np.pi / 2, np.float32))
    cos = v[:, 2].astype(np.float64)
    for k, (au, av) in ((0, (0.1, 0.3)), (2, (0.1, 0.1))):
        e = mf_eval(L, oracle, BECKMANN, au, av, v)
        np.testing.assert_allclose(e, table("test02_eval_pdf_beckmann", k), rtol=1e-5, atol=1e-8)
        np.testing.assert_allclose(e * cos, table("test02_eval_pdf_beckmann", k + 1), rtol=1e-5, atol=1e-8)
    v = sph(np.full(steps, 0.1, np.float32), np.linspace(0, 2 * np.pi, steps, dtype=np.float32))
    e = mf_eval(L, oracle, BECKMANN, 0.1, 0.3, v)
    np.testing.assert_allclose(e, table("test02_eval_pdf_beckmann", 4), rtol=2e-5)
    np.testing.assert_allclose(e * v[:, 2], table("test02_eval_pdf_beckmann", 5) * np.cos(0.1), rtol=2e-5)   # :81-85 scales the table
    np.testing.assert_allclose(mf_eval(L, oracle, BECKMANN, 0.1, 0.1, v), 11.86709118, rtol=2e-5)   # :87


@pytest.mark.parametrize("typ,test,iso", [(BECKMANN, "test03_smith_g1_beckmann", 0.67333597), (GGX, "test03_smith_g1_ggx", 0.46130955)])
def test_microfacet_smith_g1_tables(L, oracle, typ, test, iso):
    """test_microfacet.py:91-129 and :191-227"""
    steps = 20
    v = sph(np.linspace(np.pi / 3, np.pi / 2, steps, dtype=np.float32), np.full(steps, np.pi / 2, np.float32))
    np.testing.assert_allclose(mf_g1(L, oracle, typ, 0.1, 0.3, v), table(test, 0), rtol=1e-4, atol=5e-6)   # last entry: theta = pi / 2 in float32
    np.testing.assert_allclose(mf_g1(L, oracle, typ, 0.1, 0.1, v), table(test, 1), rtol=1e-4, atol=5e-6)
    v = sph(np.full(steps, np.pi / 2 * 0.98, np.float32), np.linspace(0, 2 * np.pi, steps, dtype=np.float32))
    np.testing.assert_allclose(mf_g1(L, oracle, typ, 0.1, 0.3, v), table(test, 2), rtol=1e-4)
    np.testing.assert_allclose(mf_g1(L, oracle, typ, 0.1, 0.1, v), iso, rtol=1e-4)


@pytest.mark.parametrize("typ,test", [(BECKMANN, "test04_sample_beckmann"), (GGX, "test05_sample_ggx")])
def test_microfacet_sample_tables(L, oracle, typ, test):
    """test_microfacet.py:132-188 and :230-285: a 6 x 6 grid of samples (u1 fastest), wi = +z, all normals sampled"""
    u = np.arange(6, dtype=np.float32) / np.float32(6)
    out = (C.c_float * 4)()
    got = []
    for u2 in u:
        for u1 in u:
            L.plt_oracle_mf_sample(typ, 0.1, 0.3, 0, oracle.farr(0, 0, 1), oracle.farr(u1, u2), out)
            got.append(list(out))
    got = np.array(got)
    np.testing.assert_allclose(got[:, :3], table(test, 0), atol=5e-4)     # the reference's own tolerances (:187-188)
    np.testing.assert_allclose(got[:, 3], table(test, 1), rtol=1e-4, atol=1e-4)


def one_shape_scene(mi, bsdf, emitter=None, shape="rectangle", to_world=None):
    d = {"type": "scene", "sensor": {"type": "perspective", "film": {"type": "hdrfilm", "width": 4, "height": 4}},
         "shape": {"type": shape, "bsdf": bsdf}}
    if emitter:
        d["shape"]["emitter"] = emitter
    if to_world is not None:
        d["shape"]["to_world"] = to_world
    return mi.load_dict(d)


def bsdf_probe(L, oracle, sc, wi, wo, u=(0.3, 0.4, 0.6, 0.5, 0.5)):
    out = (C.c_float * 18)()
    L.plt_oracle_bsdf_probe(C.byref(sc.flat.desc), 0, oracle.farr(*wi), oracle.farr(0.5, 0.5), oracle.farr(*WL), oracle.farr(*wo),
                            oracle.farr(*u), out)
    o = np.array(list(out))
    return {"eval": o[0:4], "pdf": o[4], "weight": o[5:9], "wo": o[9:12], "s_pdf": o[12], "s_weight": o[13:17], "eta": o[17]}


def test_diffuse_eval_pdf(mi, L, oracle):
    """src/bsdfs/tests/test_diffuse.py:13-35: pdf = cos / pi, eval = 0.5 cos / pi for the default reflectance 0.5"""
    sc = one_shape_scene(mi, {"type": "diffuse"})
    for i in range(20):
        theta = i / 19.0 * (np.pi / 2)
        wo = (np.sin(theta), 0.0, np.cos(theta))
        r = bsdf_probe(L, oracle, sc, (0, 0, 1), wo)
        assert np.isclose(r["pdf"], wo[2] / np.pi, rtol=1e-5, atol=1e-8)
        np.testing.assert_allclose(r["eval"], 0.5 * wo[2] / np.pi, rtol=1e-5, atol=1e-8)


def test_twosided_pdf_and_sample_consistency(mi, L, oracle):
    """src/bsdfs/tests/test_twosided.py:29-45 (test02_pdf) and :48-93 (test03: sample weight * pdf == eval, both sides)"""
    sc = one_shape_scene(mi, {"type": "twosided", "bsdf": {"type": "diffuse"}})
    assert np.isclose(bsdf_probe(L, oracle, sc, (0, 0, 1), (0, 0, 1))["pdf"], 1 / np.pi, rtol=1e-6)
    assert bsdf_probe(L, oracle, sc, (0, 0, 1), (0, 0, -1))["pdf"] == 0.0
    for bsdf in ({"type": "twosided", "bsdf": {"type": "diffuse", "reflectance": 0.2}},
                 {"type": "twosided", "bsdf": {"type": "roughconductor", "alpha": 0.3, "distribution": "ggx"}}):
        sc = one_shape_scene(mi, bsdf)
        for wi in ((0.0, 0.6, 0.8), (0.0, 0.6, -0.8)):
            for u in ((0.3, 0.2, 0.7, 0.5, 0.5), (0.9, 0.65, 0.15, 0.5, 0.5)):
                s = bsdf_probe(L, oracle, sc, wi, (0, 0, 1), u)
                assert s["s_pdf"] > 0 and np.sign(s["wo"][2]) == np.sign(wi[2])
                e = bsdf_probe(L, oracle, sc, wi, s["wo"], u)
                assert np.isclose(e["pdf"], s["s_pdf"], rtol=1e-4)
                np.testing.assert_allclose(s["s_weight"] * s["s_pdf"], e["eval"], rtol=2e-3, atol=1e-6)   # atol 1e-2 in the reference


def test_regular_and_irregular_spectrum_eval(mi, L, oracle):
    """src/spectra/tests/test_regular.py:19-26, test_irregular.py:18-25: linear interpolation, 0 outside the range"""
    for spec, want in (({"type": "regular", "wavelength_min": 500, "wavelength_max": 600, "values": "1, 2"}, [0, 1, 1.5, 2, 0]),
                       ({"type": "irregular", "wavelengths": "500, 600, 650", "values": "1, 2, .5"}, [0, 1, 1.5, 2, .5, 0])):
        sc = one_shape_scene(mi, {"type": "diffuse", "reflectance": spec})
        tex = sc.flat.desc.bsdfs[0].tex[0]
        out = (C.c_float * 4)()
        got = []
        for i in range(len(want)):
            L.plt_oracle_tex_probe(C.byref(sc.flat.desc), tex, oracle.farr(*([450.0 + 50 * i] * 4)), oracle.farr(0.5, 0.5), out)
            got.append(out[0])
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("origin", [[1.0, 0.0, 1.5], [1.0, 4.0, 1.5]])
@pytest.mark.parametrize("direction", [[0.0, 0.0, 1.0], [1.0, 0.0, 0.0]])
def test_perspective_sample_ray(mi, L, oracle, origin, direction):
    """src/sensors/tests/test_perspective.py:57-86 (test02_sample_ray): rays start on the near plane along their own direction,
    the film centre looks down the camera axis"""
    near = 1.0
    T = mi.Transform().look_at(origin=origin, target=[o + d for o, d in zip(origin, direction)], up=[0, 1, 0])
    sc = mi.load_dict({"type": "scene", "sensor": {"type": "perspective", "near_clip": near, "far_clip": 35.0, "fov": 34.0, "to_world": T,
                                                   "film": {"type": "hdrfilm", "width": 512, "height": 256}},
                       "shape": {"type": "rectangle"}})
    out = (C.c_float * 6)()
    inv = np.linalg.inv(np.asarray(T.matrix, np.float64))
    for pos in ((0.2, 0.6), (0.1, 0.9), (0.2, 0.2)):
        L.plt_oracle_sensor_ray(C.byref(sc.flat.desc), 0, oracle.farr(*pos), out)
        o, d = np.array(out[0:3]), np.array(out[3:6])
        inv_z = 1.0 / (inv[:3, :3] @ d)[2]
        np.testing.assert_allclose(o, np.array(origin) + near * inv_z * d, atol=1e-4)
        assert np.isclose(np.linalg.norm(d), 1.0, atol=1e-6)
    L.plt_oracle_sensor_ray(C.byref(sc.flat.desc), 0, oracle.farr(0.5, 0.5), out)
    np.testing.assert_allclose(out[3:6], direction, atol=1e-6)


def test_area_emitter_eval_and_sample_direction(mi, L, oracle):
    """src/emitters/tests/test_area.py:51-127 on its sphere emitter translated by (5, 0, 0): eval returns the radiance spectrum and 0
    for directions that point into the shape; sample_direction agrees with the shape's, pdf_direction(ds) == ds.pdf,
    eval_direction == radiance (value / pdf is what the sampler returns)"""
    sc = one_shape_scene(mi, {"type": "diffuse"}, emitter={"type": "area", "radiance": {"type": "uniform", "value": 3.0}},
                         shape="sphere", to_world=mi.Transform().translate([5, 0, 0]))
    out = (C.c_float * 24)()
    for p in ((0.2, 0.1, 0.2), (0.6, -0.9, 0.2), (0.4, 0.9, -0.2)):
        for u in ((0.4, 0.1), (0.5, 0.4), (0.3, 0.9)):
            L.plt_oracle_area_emitter_probe(C.byref(sc.flat.desc), 0, oracle.farr(*p), oracle.farr(*u), oracle.farr(*WL), out)
            o = np.array(list(out))
            d, pdf, dist, q, n = o[0:3], o[3], o[4], o[5:8], o[8:11]
            assert pdf > 0 and np.isclose(o[11], pdf, rtol=1e-4)                       # pdf_direction(it, ds) == ds.pdf
            np.testing.assert_allclose(q - np.array(p), d * dist, atol=2e-5)            # ds.d, ds.dist consistent with ds.p
            assert np.isclose(np.linalg.norm(q - np.array([5, 0, 0])), 1.0, atol=1e-5)  # on the sphere
            np.testing.assert_allclose(n, q - np.array([5, 0, 0]), atol=1e-5)
            assert np.dot(d, n) < 0                                                     # the sampled point faces the reference point
            np.testing.assert_allclose(o[12:16], 3.0, rtol=1e-6)                        # eval_direction == spectrum.eval
            np.testing.assert_allclose(o[16:20], 3.0, rtol=1e-6)                        # eval, front side
            np.testing.assert_allclose(o[20:24], 0.0)                                   # eval, direction into the shape


@pytest.mark.parametrize("name,radius", [("gaussian", 2.0), ("tent", 1.0), ("box", 0.5)])
def test_imageblock_put_against_brute_force(L, oracle, name, radius):
    """src/render/tests/test_imageblock.py:37-120 (test02_put; border=False, offset 0, normalize=False, coalesced): the block after
    one put equals the separable filter evaluated at every pixel centre; one of the 25 positions lies exactly on a pixel boundary"""
    from mitsuba3_plt_b200 import capi
    typ = {"gaussian": capi.RFILTER_GAUSSIAN, "tent": capi.RFILTER_TENT, "box": capi.RFILTER_BOX}[name]
    L.plt_oracle_splat.argtypes = [C.c_uint32, C.c_float, C.c_float, C.c_uint32, C.c_uint32, PF, PF]
    sigma = 0.5

    def f(x):
        x = np.asarray(x, np.float64)
        if name == "gaussian":
            return np.maximum(0, np.exp(-x * x / (2 * sigma ** 2)) - np.exp(-radius ** 2 / (2 * sigma ** 2)))
        if name == "tent":
            return np.maximum(0, 1 - np.abs(x))
        return ((x >= -0.5) & (x < 0.5)).astype(np.float64)
    for j in range(5):
        for i in range(5):
            pos = (np.float32(3.3 + 0.25 * i), np.float32(3 + 0.25 * j))
            film = np.zeros((6, 6), np.float32)
            L.plt_oracle_splat(typ, radius, sigma, 6, 6, oracle.farr(*pos), film.ctypes.data_as(PF))
            c = np.arange(6) + 0.5
            ref = np.outer(f(c - float(pos[1])), f(c - float(pos[0])))
            if name == "box":   # the box filter is a point deposit into the pixel containing the sample
                ref = np.zeros((6, 6)); ref[int(np.floor(pos[1])), int(np.floor(pos[0]))] = 1
            np.testing.assert_allclose(film, ref, atol=1e-5)
