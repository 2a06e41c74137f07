"""The reference's headless render script on this backend:

    python -m mitsuba3_plt_b200.headless scene.xml -s -i plt --spp 64 -o out/ [-r name=value ...]

mirrors scripts/rendering/main-headless.py:23-138 -- same arguments, same output files (result.exr/.png, result_s{0..3}.exr/.png,
params.json with bitmap_size / samples / time / time_per_sample) -- with `render_scene` as in
scripts/rendering/utils/render.py:18-44 (a `stokes` integrator around the chosen one, max_depth 7, rr_depth 50,
samples_per_pass 512), `stokes_to_bitmaps` as in scripts/utils/polarization.py:6-33 and `format_time` as in
scripts/utils/misc.py:6-30.  --plot (matplotlib windows) and --denoise (OptiX) have no counterpart here and are rejected.
"""
import argparse
import json
import os
import time

import numpy as np


def format_time(time_ns):
    """nanoseconds -> "<value> <unit>", stepping ns/us/ms by 1e3, s/m by 60, h by 24 (scripts/utils/misc.py:6-30)"""
    value = time_ns
    for unit, step in (("ns", 1e3), ("us", 1e3), ("ms", 1e3), ("s", 60), ("m", 60), ("h", 24), ("d", None)):
        if step is None or value < step:
            return f"{value} {unit}"
        value /= step


def stokes_to_bitmaps(result):
    """(H, W, 15) stokes result -> (L, [s0, s1, s2, s3]) as (H, W, 3) arrays (scripts/utils/polarization.py:6-33)"""
    val = np.asarray(result)
    if val.ndim != 3 or val.shape[2] < 15:
        raise RuntimeError(f"stokes_to_bitmaps(): expected (H, W, 15) channels, got {val.shape}")
    return val[:, :, 0:3], [val[:, :, 3 * i + 3:3 * i + 6] for i in range(4)]


def render_scene(mi, scene, samples, integrator="plt", denoise=False):
    """scripts/rendering/utils/render.py:18-44"""
    if denoise:
        raise RuntimeError("--denoise uses the OptiX denoiser: unavailable on this backend")
    it = mi.load_dict({"type": "stokes",
                       "nested": {"type": integrator, "max_depth": 7, "rr_depth": 50, "samples_per_pass": 512}})
    result = it.render(scene, spp=samples)
    L, sp = stokes_to_bitmaps(result)
    return L, sp, result


def write_outputs(mi, folder, L, sp, spp, elapsed_ns):
    """main-headless.py:57-87: the image files and params.json"""
    if folder:
        os.makedirs(folder, exist_ok=True)
    for ext in ("exr", "png"):
        mi.util.write_bitmap(os.path.join(folder, f"result.{ext}"), L)
        for i, s in enumerate(sp):
            mi.util.write_bitmap(os.path.join(folder, f"result_s{i}.{ext}"), s)
    data = {"bitmap_size": {"width": int(L.shape[1]), "height": int(L.shape[0])}, "samples": spp,
            "time": format_time(elapsed_ns), "time_per_sample": format_time(elapsed_ns / spp)}
    with open(os.path.join(folder, "params.json"), "w") as f:
        json.dump(data, f, indent=2)
    return data


def main(argv=None):
    ap = argparse.ArgumentParser(description="Wave-based scripted renderer.")
    ap.add_argument("scene", help="Scene description file")
    ap.add_argument("--spectral", "-s", action="store_true", help="Whether to perform spectral rendering.")
    ap.add_argument("--verbose", "-v", action="store_true")
    ap.add_argument("--plot", "-p", action="store_true")
    ap.add_argument("--integrator", "-i", type=str, default="plt")
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--outdir", "-o", type=str, default="")
    ap.add_argument("--denoise", "-d", action="store_true")
    ap.add_argument("--override", "-r", type=str, nargs="*", default=[])
    a = ap.parse_args(argv)
    if a.plot:
        raise SystemExit("--plot opens matplotlib windows: unavailable on this (headless) backend")
    import mitsuba3_plt_b200 as mi
    if a.spectral:   # main-headless.py:128-133
        print("Loading spectral variant...")
        mi.set_variant("cuda_ad_spectral_polarized")
    else:
        print("Loading RGB variant...")
        mi.set_variant("cuda_ad_rgb_polarized")
    overrides = {k: float(v) for k, v in (o.split("=") for o in a.override)}
    print("Loading scene...")
    start = time.perf_counter_ns()
    scene = mi.load_file(a.scene)
    print(f"...done. ({format_time(time.perf_counter_ns() - start)})")
    params = mi.traverse(scene)
    if a.verbose:
        print(params)
    for k, v in overrides.items():
        print("overriding", k, "=", v)
        params[k] = v
    params.update()
    print("Rendering...")
    L, sp, _ = render_scene(mi, scene, a.spp, a.integrator, a.denoise)
    el = time.perf_counter_ns() - start     # the reference's "time" includes the scene load (main-headless.py:36, 54)
    print(f"...done. ({format_time(el)})")
    write_outputs(mi, a.outdir, L, sp, a.spp, el)


if __name__ == "__main__":
    main()
