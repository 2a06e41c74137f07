#!/usr/bin/env python
"""Integrator-level cross-check against the reference's OWN published render (container only: needs /root/reference).

results/cbox-plt/result.exr is the reference's render of scenes/cbox/cbox.xml with `stokes{plt}` (polarized CUDA variant, 500x500,
1024 spp, `main-headless.py`).  This renders the same scene file with the CPU oracle and reports
image-level agreement, now with the oracle's polarized + stokes restatement (the variant the published files were made with), incl.
the structure of the Stokes components against result_s{0..3}.exr.  The two differ by Monte-Carlo noise (the oracle uses fewer spp)
and by the authors' post-processing of the published files, so this is a sanity bound on integrator-level behaviour (geometry,
light transport, polarization structure, film), not a bit-level pin.
usage: python tools/compare_reference_image.py [spp] [ref=/root/reference]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import mitsuba3_plt_b200 as mi
from mitsuba3_plt_b200 import imageio
import oracle_lib

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
ref = sys.argv[2] if len(sys.argv) > 2 else "/root/reference"
want, _ = imageio.read_exr(f"{ref}/results/cbox-plt/result.exr")
params = json.load(open(f"{ref}/results/cbox-plt/params.json"))
sc = mi.load_file(f"{ref}/scenes/cbox/cbox.xml", resx=want.shape[1], resy=want.shape[0])
# the variant the published render used: polarized spectral + the `stokes` wrapper (15 channels: R G B, S0.RGB .. S3.RGB)
r = oracle_lib.render(sc.flat, spp, seed=0, max_depth=7, rr_depth=50, polarized=True, stokes=True)
img = r["image"][..., :3]


def box(a, k=8):   # k x k box filter: compares the signal, not the per-pixel noise of the lower-spp image
    H, W = a.shape[0] // k * k, a.shape[1] // k * k
    return a[:H, :W].reshape(H // k, k, W // k, k, a.shape[2]).mean(axis=(1, 3))


a, b = box(img), box(want)
rel_mse = float(np.mean((a - b) ** 2 / (b ** 2 + 1e-2)))
out = {"reference_image": "results/cbox-plt/result.exr", "reference_params": params, "oracle_spp": spp, "oracle_variant": "polarized + stokes",
       "oracle_seconds": r["seconds"], "mean_rgb_oracle": [float(x) for x in img.mean(axis=(0, 1))],
       "mean_rgb_reference": [float(x) for x in want.mean(axis=(0, 1))],
       "mean_ratio": [float(x) for x in img.mean(axis=(0, 1)) / want.mean(axis=(0, 1))],
       "relative_mse_8x8_blocks": rel_mse,
       "per_pixel_relative_mse": float(np.mean((img - want) ** 2 / (want ** 2 + 1e-2))),
       "nonfinite_pixels_oracle": int((~np.isfinite(r["image"])).any(-1).sum())}
# Stokes components against results/cbox-plt/result_s{0..3}.exr: structure (correlation of 16 x 16 block means, sign agreement
# where the reference's |S_i| is above its noise) -- the published files are tone-mapped by the authors' script, so levels differ
for i in range(4):
    path = f"{ref}/results/cbox-plt/result_s{i}.exr"
    if not os.path.exists(path):
        continue
    si_ref, _ = imageio.read_exr(path)
    x, y = box(r["image"][..., 3 + 3 * i:6 + 3 * i], 16)[..., 0].ravel(), box(si_ref[..., :3], 16)[..., 0].ravel()
    fin = np.isfinite(x) & np.isfinite(y)
    x, y = x[fin], y[fin]
    strong = np.abs(y) > 0.25 * np.abs(y).max()
    out[f"S{i}"] = {"corr_16x16_blocks": float(np.corrcoef(x, y)[0, 1]), "sign_agreement_where_strong": float((np.sign(x[strong]) == np.sign(y[strong])).mean()),
                    "n_strong_blocks": int(strong.sum()), "mean_oracle": float(x.mean()), "mean_reference": float(y.mean()),
                    "nonfinite_pixels_reference": int((~np.isfinite(si_ref)).any(-1).sum())}
print(json.dumps(out))
if len(sys.argv) > 3:
    imageio.write_bitmap(sys.argv[3], np.concatenate([img, want], axis=1))
