"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on the same scenes and seeds."""
import numpy as np
import pytest

from parity import MAX_RMSE, MIN_FRACTION, max_finiteness_mismatch, nonfinite_counts, relative_mse, sample_agreement

pytestmark = pytest.mark.gpu

CASES = [  # (scene, res, spp, seed); "cbox", "cbox-grating", "gratingcap", "veach-mis" are the reference's own scene files
    ("cbox", 256, 16, 0),                    # C1 at its configured size (BASELINE.json configs[0])
    ("cbox", 64, 8, 0),
    ("cbox-grating", 64, 8, 0),
    ("cbox-grating", 48, 16, 7),
    ("cbox-grating", 256, 16, 1),            # C2's scene, full frame, 1.05e6 samples: the margin of the 99.9 % bar at scale
    ("cbox-grating:procedural", 64, 8, 0),   # the procedural stand-ins (other tessellation / UVs) stay covered
    ("cbox:procedural", 64, 8, 0),
    ("veach-mis", 64, 8, 1),                 # C4: rough conductors, sphere lights, tent filter
    ("spectrograph:bbody_3000", 1, 256, 0),  # C3: batch of 150 orthographic sensors with SRFs, directional blackbody light
    ("spectrograph:d65", 1, 64, 3),
    ("gratingcap", 96, 8, 0),                # C5: principled screen + non-radial grating strip, directional light
    ("cbox-grating:sky", 64, 8, 2),          # `constant` environment emitter: escaped rays, uniform-sphere NEE (8f f3, first step)
    ("cbox:sky", 48, 8, 4),
    ("cbox-grating:envmap", 64, 8, 5),       # `envmap` emitter (EXR file, rotated): Hierarchical2D sampling, bilinear spectral lookup
    ("cbox:envmap", 48, 8, 6),
    ("disk", 96, 8, 0),                      # scenes/disk equivalent: radial grating disc, textured metallic principled, envmap only
]


def ray_count_agreement(name, a, b):
    """Closest-hit rays per sample must agree like the radiance does.  Shadow rays are counted when `ds.pdf != 0`
    (scene.cpp:330-340), and for a vertex that lies ON a flat area light (paths continue through emitters, N8) the sampled
    direction is in the light's plane: |dot(d, n)| is 0 or one ulp, so pdf is 0 or huge -- with zero radiance either way
    (`dot(d, n) < 0` fails, area.cpp:198-208).  Device sinf / cosf differ from glibc by an ulp often enough to flip that test on
    ~1.5 % of the samples of the closed procedural box: those counts may differ by one, everything else must be equal."""
    a, b = a.astype(np.int64), b.astype(np.int64)
    same_closest = (a[:, 0] == b[:, 0]).mean()
    assert same_closest >= MIN_FRACTION, f"{name}: closest-hit ray counts agree on {same_closest * 100:.4f}% of samples"
    d = np.abs(a[:, 1] - b[:, 1])
    assert (d <= 1).mean() >= MIN_FRACTION and (d == 0).mean() >= 0.97, (name, (d == 0).mean(), (d <= 1).mean())
    assert abs(int(a[:, 1].sum()) - int(b[:, 1].sum())) <= max(4, 2e-3 * b[:, 1].sum()), name


@pytest.fixture(scope="module")
def integ(mi):
    return mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})


@pytest.mark.parametrize("name,res,spp,seed", CASES)
def test_per_sample_radiance(mi, oracle, integ, name, res, spp, seed):
    from mitsuba3_plt_b200 import scenes
    sc = scenes.load(name, resx=res, resy=res)
    dump = {}
    img = integ.render(sc, spp=spp, seed=seed, dump=dump)
    ref = oracle.render(sc.flat, spp, seed=seed, dump=True)
    assert integ.last_stats["kernel_launches"] >= 4
    wl_frac = sample_agreement(dump["wavelengths"], ref["wavelengths"])
    assert wl_frac == 1.0, f"sampled wavelengths differ: {wl_frac}"
    frac = sample_agreement(dump["radiance"], ref["radiance"])
    assert frac >= MIN_FRACTION, f"{name}: only {frac * 100:.4f}% of samples within 1e-4"
    # non-finite samples (invalid-lobe quirk) are counted, not waved through: same samples up to branch flips
    nf_a, nf_b, nf_both = nonfinite_counts(dump["radiance"], ref["radiance"])
    assert abs(nf_a - nf_b) <= max(2, 0.02 * nf_b) and nf_both >= min(nf_a, nf_b) - max(2, 0.02 * nf_b), (nf_a, nf_b, nf_both)
    # per-sample ray counts (pltb_sample_dump.n_rays): closest-hit and shadow rays of every sample
    ray_count_agreement(name, dump["n_rays"], ref["n_rays"])
    assert int(dump["n_rays"][:, 0].sum()) == integ.last_stats["closest_rays"]
    assert int(dump["n_rays"][:, 1].sum()) == integ.last_stats["shadow_rays"]
    # ray counters are part of the metric (Mrays/s): must match the oracle's up to branch flips
    for k in ("closest_rays", "shadow_rays", "vertices", "grating_vertices"):
        a, b = integ.last_stats[k], ref["stats"][k]
        assert abs(a - b) <= max(4, 1e-3 * b), (k, a, b)
    rmse, mism = relative_mse(img, ref["image"])
    assert rmse <= MAX_RMSE and mism <= max_finiteness_mismatch(img.shape[0] * img.shape[1]), (rmse, mism)


def test_environment_furnace(mi, oracle, integ):
    """a large diffuse plane (albedo 0.5) under a uniform `constant` sky of radiance 1 reflects 0.5: emission of escaped rays
    and uniform-sphere NEE with their MIS weights (constant.cpp:92-157) sum to the analytic value, and match the oracle"""
    T = mi.Transform
    sc = mi.load_dict({"type": "scene",
                       "sensor": {"type": "perspective", "fov": 20, "to_world": T().look_at([0, 0, 5], [0, 0, 0], [0, 1, 0]),
                                  "film": {"type": "hdrfilm", "width": 32, "height": 32}},
                       "plane": {"type": "rectangle", "to_world": T().scale([50, 50, 1]),
                                 "bsdf": {"type": "diffuse", "reflectance": {"type": "uniform", "value": 0.5}}},
                       "sky": {"type": "constant", "radiance": {"type": "uniform", "value": 1.0}}})
    dump = {}
    integ.render(sc, spp=64, seed=1, dump=dump)
    assert abs(dump["radiance"].mean() - 0.5) < 5e-3
    ref = oracle.render(sc.flat, 64, seed=1, dump=True, film=False)
    assert sample_agreement(dump["radiance"], ref["radiance"]) >= MIN_FRACTION


def test_many_small_passes(mi, oracle, integ, monkeypatch):
    """the wavefront pass loop with passes that straddle sample indices (2304 pixels, 4096-sample passes): per-sample results
    and the film do not depend on the pass size"""
    from mitsuba3_plt_b200 import scenes
    sc = scenes.load("cbox-grating", resx=48, resy=48)
    monkeypatch.setenv("PLTB_CHUNK_LOG2", "12")
    ctx = mi.Context()          # fresh context: the pass size is fixed when the wavefront state is first allocated
    try:
        dump = {}
        img = integ.render(sc, spp=16, seed=9, dump=dump, ctx=ctx)
        assert integ.last_stats["passes"] == 9
    finally:
        sc.release()
        ctx.close()
    monkeypatch.delenv("PLTB_CHUNK_LOG2")
    ref = oracle.render(sc.flat, 16, seed=9, dump=True)
    assert sample_agreement(dump["radiance"], ref["radiance"]) >= MIN_FRACTION
    rmse, mism = relative_mse(img, ref["image"])
    assert rmse <= MAX_RMSE and mism <= max_finiteness_mismatch(48 * 48)
    one = {}
    integ.render(sc, spp=16, seed=9, dump=one)       # default context, one pass
    np.testing.assert_array_equal(one["radiance"].view(np.uint32), dump["radiance"].view(np.uint32))


def test_rr_and_depth_variants(mi, oracle):
    """plugin defaults max_depth/rr_depth (plt.py:37-38): Russian roulette active from depth 4"""
    from mitsuba3_plt_b200 import scenes
    sc = scenes.load("cbox", resx=48, resy=48)
    for md, rr in ((16, 4), (3, 2), (1, 1)):
        it = mi.load_dict({"type": "plt", "max_depth": md, "rr_depth": rr})
        dump = {}
        it.render(sc, spp=4, seed=3, dump=dump)
        ref = oracle.render(sc.flat, 4, seed=3, max_depth=md, rr_depth=rr, dump=True, film=False)
        frac = sample_agreement(dump["radiance"], ref["radiance"])
        assert frac >= MIN_FRACTION, (md, rr, frac)


def test_spp_window_union_is_bit_identical(mi, integ):
    """rank shards [spp_begin, spp_begin+count) reproduce the single-GPU per-sample radiance bit for bit."""
    from mitsuba3_plt_b200 import capi, scenes
    import ctypes as C
    sc = scenes.load("cbox-grating", resx=32, resy=32)
    full = {}
    integ.render(sc, spp=8, seed=1, dump=full)
    ctx = mi.context()
    parts = []
    for lo in (0, 4):
        prm = capi.RenderParams(seed=1, spp=8, spp_begin=lo, spp_count=4, max_depth=7, rr_depth=50, flags=capi.RENDER_NO_SPLAT)
        n = 32 * 32 * 4
        L = np.zeros((n, 4), np.float32)
        dmp = capi.SampleDump(None, capi.fptr(L), None, n)
        capi.check(capi.lib().pltb_render(ctx._h, sc.device(ctx), C.byref(prm), None, C.byref(dmp), None))
        parts.append(L.reshape(32 * 32, 4, 4))
    union = np.concatenate(parts, axis=1).reshape(-1, 4)
    np.testing.assert_array_equal(union.view(np.uint32), full["radiance"].view(np.uint32))


def test_animation_frames(mi, oracle, integ):
    """C5: camera animation of disk_animation.py:108-134 -- sensor to_world updated in place on the device copy
    (pltb_scene_update_sensor), seed = frame index; each frame must match the oracle run on the updated host scene."""
    from mitsuba3_plt_b200 import animation, scenes
    sc = scenes.load("gratingcap", resx=64, resy=48)
    sc.device()   # create the device copy first so that the in-place update path is what gets exercised
    params = mi.traverse(sc)
    key = next(iter(params))
    initial = params[key]
    for i in (0, 7, 23):
        params[key] = animation.camera_to_world(initial, i, 24)
        params.update()
        dump = {}
        img = integ.render(sc, spp=4, seed=i, dump=dump)
        ref = oracle.render(sc.flat, 4, seed=i, dump=True)
        assert sample_agreement(dump["wavelengths"], ref["wavelengths"]) == 1.0
        frac = sample_agreement(dump["radiance"], ref["radiance"])
        assert frac >= MIN_FRACTION, (i, frac)
    params[key] = initial
    params.update()
    frames = animation.render_frames(sc, integ, frames=24, spp=8, indices=[2, 12])
    assert sorted(frames) == [2, 12] and frames[2].shape == (48, 64, 3)
    assert np.nanmax(frames[2]) > 0 and np.nanmax(frames[12]) > 0 and not np.array_equal(frames[2], frames[12])
    # the loop restores the initial camera; frame 12 rendered by hand is the same image (up to the order of the film's atomics)
    assert np.array_equal(mi.traverse(sc)[key].matrix, initial.matrix)
    params[key] = animation.camera_to_world(initial, 12, 24)
    params.update()
    again = integ.render(sc, spp=8, seed=12)
    np.testing.assert_allclose(again, frames[12], rtol=1e-4, atol=1e-6)
    params[key] = initial
    params.update()
    # several frames in flight (one context and device scene copy per lane): the same images as the serial loop, frame by frame
    idx = list(range(0, 24, 2))
    serial = animation.render_frames(sc, integ, frames=24, spp=8, indices=idx)
    for lanes in (3, 5):
        seen = []
        flying = animation.render_frames(sc, integ, frames=24, spp=8, indices=idx, inflight=lanes)
        assert sorted(flying) == idx
        for i in idx:
            np.testing.assert_allclose(flying[i], serial[i], rtol=1e-4, atol=1e-6)
        animation.render_frames(sc, integ, frames=24, spp=8, indices=idx, inflight=lanes, on_frame=lambda i, img: seen.append((i, float(np.nansum(img)))))
        assert sorted(i for i, _ in seen) == idx
        for i, total in seen:
            assert abs(total - float(np.nansum(serial[i]))) <= 1e-3 * abs(float(np.nansum(serial[i])))
    assert np.array_equal(mi.traverse(sc)[key].matrix, initial.matrix)


def test_accumulate_and_raw_film(mi, oracle, integ):
    """develop=False returns the raw RGBW accumulator (hdrfilm.cpp:304-311); weights sum to ~spp per interior pixel"""
    from mitsuba3_plt_b200 import scenes
    sc = scenes.load("cbox", resx=32, resy=32)
    raw = integ.render(sc, spp=16, seed=0, develop=False)
    assert raw.shape == (32, 32, 4)
    ref = oracle.render(sc.flat, 16, seed=0)
    np.testing.assert_allclose(raw[..., 3], ref["film_rgbw"][..., 3], rtol=1e-4, atol=1e-4)
    rmse, _ = relative_mse(raw[..., :3], ref["film_rgbw"][..., :3])
    assert rmse <= MAX_RMSE


def test_errors(mi, integ):
    from mitsuba3_plt_b200 import capi, scenes
    import ctypes as C
    sc = scenes.load("cbox", resx=16, resy=16)
    ctx = mi.context()
    prm = capi.RenderParams(seed=0, spp=4, spp_begin=2, spp_count=4, max_depth=7, rr_depth=50)
    assert capi.lib().pltb_render(ctx._h, sc.device(ctx), C.byref(prm), None, None, None) == -1
    assert b"spp" in capi.lib().pltb_last_error()
    prm = capi.RenderParams(seed=0, spp=4, spp_begin=0, spp_count=4, max_depth=0, rr_depth=50)
    assert capi.lib().pltb_render(ctx._h, sc.device(ctx), C.byref(prm), None, None, None) == -1


def test_strip_splat_equals_per_tap_atomics(mi, integ, monkeypatch):
    """k_splat_strip merges 32 register footprints into a shared-memory strip with plain read-modify-writes ordered by
    __syncwarp; a lost update would show as missing film weight.  Against the per-tap vector-atomic path (PLTB_SPLAT_DIRECT):
    the W channel (a sum of filter weights, no cancellation) must agree to float32 summation-order accuracy, RGB likewise.
    (compute-sanitizer's racecheck is closed on the GPU pool; this and the -DPLTB_DEBUG_CHECKS run stand in for it.)"""
    from mitsuba3_plt_b200 import scenes
    for name, res, spp in (("cbox-grating", 96, 64), ("veach-mis", 80, 32)):   # 5x5 Gaussian and 3x3 tent footprints
        sc = scenes.load(name, resx=res, resy=res)
        a = integ.render(sc, spp=spp, seed=2, develop=False)
        monkeypatch.setenv("PLTB_SPLAT_DIRECT", "1")
        b = integ.render(sc, spp=spp, seed=2, develop=False)
        monkeypatch.delenv("PLTB_SPLAT_DIRECT")
        np.testing.assert_allclose(a[..., 3], b[..., 3], rtol=2e-5, atol=1e-5)
        fin = np.isfinite(a).all(-1) & np.isfinite(b).all(-1)
        assert (np.isfinite(a).all(-1) == np.isfinite(b).all(-1)).all()
        np.testing.assert_allclose(a[fin], b[fin], rtol=2e-4, atol=1e-4)


def glass_scene(mi, res=64):
    """a rough refractive `principled` sphere and slab (spec_trans lobe, principled.cpp:437-453,604-621,816-828; outside and inside
    incidence, eta > 1) over a diffuse floor under an area light"""
    T = mi.Transform
    glass = {"type": "principled", "base_color": {"type": "rgb", "value": [0.8, 0.9, 0.7]}, "roughness": 0.35, "spec_trans": 0.8,
             "eta": 1.45, "clearcoat": 0.3, "metallic": 0.1, "anisotropic": 0.2}
    return mi.load_dict({
        "type": "scene",
        "sensor": {"type": "perspective", "fov": 40, "to_world": T().look_at([0, 1.2, 4.5], [0, 0.6, 0], [0, 1, 0]),
                   "film": {"type": "hdrfilm", "width": res, "height": res}},
        "floor": {"type": "rectangle", "to_world": T().rotate([1, 0, 0], -90).scale([4, 4, 1]),
                  "bsdf": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.6, 0.5, 0.4]}}},
        "ball": {"type": "sphere", "center": [-0.7, 0.7, 0], "radius": 0.7, "bsdf": glass},
        # (the slab floats: a face coplanar with the floor would make closest hits depend on traversal order)
        "slab": {"type": "cube", "to_world": T().translate([0.9, 0.75, 0.2]).rotate([0, 1, 0], 20).scale([0.5, 0.6, 0.08]), "bsdf": glass},
        "light": {"type": "rectangle", "to_world": T().translate([0, 3.0, 0]).rotate([1, 0, 0], 90).scale([1.0, 1.0, 1]),
                  "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [12, 12, 12]}}}})


def test_principled_transmission(mi, oracle, integ):
    sc = glass_scene(mi)
    dump = {}
    img = integ.render(sc, spp=16, seed=5, dump=dump)
    ref = oracle.render(sc.flat, 16, seed=5, dump=True)
    frac = sample_agreement(dump["radiance"], ref["radiance"])
    assert frac >= MIN_FRACTION, f"principled glass: only {frac * 100:.4f}% of samples within 1e-4"
    ray_count_agreement("glass", dump["n_rays"], ref["n_rays"])
    assert ref["stats"]["vertices"] > 1.5 * ref["stats"]["samples"] * 0.3      # paths do go through the glass
    rmse, mism = relative_mse(img, ref["image"])
    assert rmse <= MAX_RMSE and mism <= max_finiteness_mismatch(img.shape[0] * img.shape[1]), (rmse, mism)
    try:   # polarized variant: principled stays a depolarizer (principled.cpp:708)
        mi.set_variant("cuda_ad_spectral_polarized")
        dump = {}
        integ.render(sc, spp=8, seed=6, dump=dump)
        refp = oracle.render(sc.flat, 8, seed=6, dump=True, polarized=True)
        assert sample_agreement(dump["radiance"], refp["radiance"]) >= MIN_FRACTION
    finally:
        mi.set_variant("cuda_spectral")
