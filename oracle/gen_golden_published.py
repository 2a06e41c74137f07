#!/usr/bin/env python
"""Fixtures from the reference's OWN published render (container only: needs /root/reference/results).

results/cbox-plt/result_s{0..3}.exr are the Stokes components of scenes/cbox/cbox.xml rendered by the reference with
`stokes{plt}` in the polarized spectral CUDA variant (500x500, 1024 spp; scripts/rendering/main-headless.py, params.json next to
them) -- linear float images written by `mi.Bitmap(s_i)` (scripts/utils/polarization.py:24-36), the only integrator-level outputs
of the plt path the reference holds.  This script stores
  tests/golden/published_cbox_plt_s0.exr        S0 at full resolution (float32, ZIP), for per-pixel comparison
  tests/golden/published_cbox_plt_s{1,2,3}.exr  S1..S3 at full resolution (half floats, ZIP)
  tests/golden/published_cbox_plt_stokes.json    20 x 20 block means of S0..S3 (first channel), for structure checks
so that tests can compare the oracle and the CUDA path with what the reference itself produced.
    python oracle/gen_golden_published.py [/root/reference]"""
import json, os, sys
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
import numpy as np
from mitsuba3_plt_b200 import imageio

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
gold = os.path.join(HERE, "..", "tests", "golden")
out = {"_source": "results/cbox-plt/result_s{0..3}.exr of ZenithGD/mitsuba3-plt @ bc454f3d (scenes/cbox/cbox.xml, stokes{plt}, polarized "
                  "spectral CUDA variant, 500x500, 1024 spp)", "params": json.load(open(f"{ref}/results/cbox-plt/params.json")),
       "block": 20}
for i in range(4):
    img, _ = imageio.read_exr(f"{ref}/results/cbox-plt/result_s{i}.exr")
    img = img[..., :3].astype(np.float32)
    assert img.shape == (500, 500, 3) and np.isfinite(img).all()
    out[f"S{i}"] = img.reshape(25, 20, 25, 20, 3).mean(axis=(1, 3))[..., 0].round(7).tolist()
    imageio.write_exr(os.path.join(gold, f"published_cbox_plt_s{i}.exr"), img, half=i > 0, compression="zip")
json.dump(out, open(os.path.join(gold, "published_cbox_plt_stokes.json"), "w"), separators=(",", ":"))
print({k: (np.array(v).shape if isinstance(v, list) else v) for k, v in out.items()}, os.path.getsize(os.path.join(gold, "published_cbox_plt_s0.exr")))
