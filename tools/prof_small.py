#!/usr/bin/env python
"""Per-kernel device times (CUDA events after every launch, PLTB_RENDER_PROFILE) of one render of a builtin scene.
usage: python tools/prof_small.py [scene] [res] [spp] [polarized]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mitsuba3_plt_b200 as mi
from mitsuba3_plt_b200 import scenes

name = sys.argv[1] if len(sys.argv) > 1 else "cbox-grating"
res, spp = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((2, 1024), (3, 8)))
pol = len(sys.argv) > 4 and sys.argv[4] == "polarized"
if pol:
    mi.set_variant("cuda_ad_spectral_polarized")
it = mi.load_dict({"type": "stokes", "nested": {"type": "plt", "max_depth": 7, "rr_depth": 50}} if pol else {"type": "plt", "max_depth": 7, "rr_depth": 50})
sc = scenes.load(name, resx=res, resy=res)
it.render(sc, spp=spp, seed=0, download=False)
it.render(sc, spp=spp, seed=0, download=False, profile=True)
st = it.last_stats
tot = sum(v["ms"] for v in st["kernels"].values())
print(f"{name} {res}x{res} x {spp} spp{' polarized' if pol else ''}: {st['samples'] / st['ms_total'] / 1e3:.1f} Msamples/s, {st['ms_total']:.2f} ms, launches {st['kernel_launches']}")
for k, v in st["kernels"].items():
    if v["launches"]:
        print(f"  {k:22s} {v['ms']:9.3f} ms {100 * v['ms'] / tot:5.1f} %  {v['launches']:4d} launches")
