#!/usr/bin/env python
"""Times every BASELINE.json config at its full size on one GPU (device time from pltb_stats, scene load / BVH build excluded)
next to the CPU oracle on a bounded sample of the same workload (all host cores).  One JSON line per config.
usage: python tools/config_bench.py [--cpu-seconds 6] [--only C2,C4]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import mitsuba3_plt_b200 as mi
from mitsuba3_plt_b200 import animation, scenes
import oracle_lib

ap = argparse.ArgumentParser()
ap.add_argument("--cpu-seconds", type=float, default=6.0)
ap.add_argument("--only", default="")
ap.add_argument("--profile", action="store_true", help="add per-kernel device times (a second, event-bracketed render)")
args = ap.parse_args()

CONFIGS = [  # id, builtin scene, resx, resy, spp, note    (ids ending in "pol": polarized spectral variant + `stokes` wrapper, 8f f1)
    ("C1", "cbox", 256, 256, 16, "scenes/cbox, 256x256, 16 spp"),
    ("C1L", "cbox", 1024, 1024, 64, "scenes/cbox, 1024x1024, 64 spp (C1's scene at a size that fills the GPU)"),
    ("C1pol", "cbox", 500, 500, 1024, "scenes/cbox 500x500, 1024 spp, stokes{plt}, polarized: the reference's published results/cbox-plt "
                                    "(params.json: 1.654 min on the authors' GPU, JIT included)"),
    ("C2pol", "cbox-grating", 1024, 1024, 128, "scenes/cbox-grating 1024x1024, 128 spp, stokes{plt}, polarized"),
    ("C2", "cbox-grating", 1024, 1024, 1024, "scenes/cbox-grating 1024x1024, 1024 spp"),
    ("C2rgb", "cbox-grating", 1024, 1024, 256, "scenes/cbox-grating 1024x1024, 256 spp, cuda_ad_rgb (RGB fallback of roughgrating, 8f f2)"),
    ("Grgbpol", "gratings", 800, 600, 8192, "scenes/gratings 800x600, 8192 spp, stokes{plt}, cuda_ad_rgb_polarized: the reference's published "
                                         "results/gratings/rgb/period=0.6 (params.json: 12.93 min on the authors' GPU, JIT included)"),
    ("C3", "spectrograph:bbody_3000", 1, 1, 1 << 20, "spectrograph, 150 sensors x 2^20 spp (largest row of synthetic-times.csv)"),
    ("C3p", "spectrograph:bbody_3000", 1, 1, 10_000_000, "spectrograph prior pass, 150 sensors x 1e7 spp"),
    ("C4", "veach-mis", 1920, 1080, 4096, "scenes/veach-mis 1920x1080, 4096 spp (2 sub-renders)"),
]
it = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
only = set(filter(None, args.only.split(",")))


def oracle_rate(sc, spp, W, H, pol=False):
    kw = {"polarized": pol, "stokes": pol}
    probe = oracle_lib.render(sc.flat, spp, spp_begin=0, spp_count=1, film=False, **kw) if W * H > 4096 else \
        oracle_lib.render(sc.flat, spp, spp_begin=0, spp_count=min(spp, 256), film=False, **kw)
    rate = probe["stats"]["samples"] / max(probe["seconds"], 1e-9)
    n = int(max(1, min(spp, args.cpu_seconds * rate / (W * H))))
    r = oracle_lib.render(sc.flat, spp, spp_begin=0, spp_count=n, **kw)
    st = r["stats"]
    return st["samples"] / r["seconds"] / 1e6, (st["closest_rays"] + st["shadow_rays"]) / r["seconds"] / 1e6, n


for cid, name, rx, ry, spp, note in CONFIGS:
    if only and cid not in only:
        continue
    pol, rgb = cid.endswith("pol"), "rgb" in cid
    mi.set_variant(("cuda_ad_rgb" if rgb else "cuda_ad_spectral") + ("_polarized" if pol else ""))
    it = mi.load_dict({"type": "stokes", "nested": {"type": "plt", "max_depth": 7, "rr_depth": 50}} if pol else
                      {"type": "plt", "max_depth": 7, "rr_depth": 50})
    sc = scenes.load(name, resx=rx, resy=ry)
    W, H = sc.film_size()
    # warm-up at the pass size of the timed render (2^25 samples): a smaller one makes the timed render re-allocate the wavefront
    # state, and the first render on freshly allocated device memory was an intermittent 1.5-2.5x outlier (profiles/r02_configs_7a.jsonl)
    warm = min(spp, max(1, -(-(1 << 25) // (W * H))))
    it.render(sc, spp=warm, seed=0, download=False)
    t0 = time.perf_counter()
    img = it.render(sc, spp=spp, seed=0)
    wall = time.perf_counter() - t0
    st = it.last_stats
    ms = st["ms_total"]
    cpu_ms, cpu_mr, n = oracle_rate(sc, spp, W, H, pol) if not args.profile else (0.0, 0.0, 0)
    kern = None
    if args.profile:
        it.render(sc, spp=spp, seed=0, profile=True, download=False)
        kern = {k: (round(v["ms"], 2), v["launches"]) for k, v in it.last_stats["kernels"].items() if v["launches"]}
    print(json.dumps({"config": cid, "kernels_ms_launches": kern, "workload": note, "film": [W, H], "spp": spp, "samples": st["samples"],
                      "gpu_ms": round(ms, 3), "gpu_wall_ms": round(wall * 1e3, 3),
                      "gpu_msamples_per_s": round(st["samples"] / ms / 1e3, 2),
                      "gpu_mrays_per_s": round((st["closest_rays"] + st["shadow_rays"]) / ms / 1e3, 2),
                      "rays_per_sample": round((st["closest_rays"] + st["shadow_rays"]) / st["samples"], 3),
                      "grating_vertices_per_sample": round(st["grating_vertices"] / st["samples"], 4),
                      "passes": st["passes"], "launches": st["kernel_launches"],
                      "image_mean": float(np.nanmean(img)), "nan_pixels": int(np.isnan(img).any(-1).sum()),
                      "cpu_oracle_msamples_per_s": round(cpu_ms, 4), "cpu_oracle_mrays_per_s": round(cpu_mr, 4),
                      "cpu_cores": os.cpu_count(), "cpu_sample": f"spp indices [0,{n}) of {spp}, full frame"}), flush=True)

mi.set_variant("cuda_spectral")
it = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
for c5_id, c5_scene in (("C5", "gratingcap"), ("C5d", "disk")):
    if only and c5_id not in only:
        continue
    sc = scenes.load(c5_scene, resx=800, resy=600)
    animation.render_frames(sc, it, frames=240, spp=1, indices=list(range(8)))
    dev = []
    t0 = time.perf_counter()
    animation.render_frames(sc, it, frames=240, spp=1, on_frame=lambda i, img: dev.append(it.last_stats["ms_total"]))
    wall = time.perf_counter() - t0
    animation.render_frames(sc, it, frames=240, spp=1, indices=list(range(32)), inflight=4)
    t0 = time.perf_counter()
    animation.render_frames(sc, it, frames=240, spp=1, on_frame=lambda i, img: None, inflight=4)
    wall4 = time.perf_counter() - t0   # four frames in flight (animation.render_frames(inflight=4)): same frames, same images
    t1 = time.perf_counter()
    for i in range(4):
        oracle_lib.render(sc.flat, 1, seed=i)
    cpu = (time.perf_counter() - t1) / 4
    print(json.dumps({"config": c5_id, "workload": f"scenes/{c5_scene} 800x600, 1 spp per frame x 240 frames, camera path of "
                      "disk_animation.py", "frames_per_s": round(240 / wall, 1), "ms_per_frame_wall": round(1e3 * wall / 240, 3),
                      "ms_per_frame_device": round(float(np.mean(dev)), 3), "gpu_msamples_per_s": round(480000 * 240 / wall / 1e6, 2),
                      "frames_per_s_inflight4": round(240 / wall4, 1), "ms_per_frame_wall_inflight4": round(1e3 * wall4 / 240, 3),
                      "cpu_oracle_ms_per_frame": round(cpu * 1e3, 2), "cpu_cores": os.cpu_count()}), flush=True)
