#!/usr/bin/env python
"""One render of a builtin scene, sized to be captured under ncu (one pass unless spp is raised).
usage: python tools/ncu_target.py [scene] [res] [spp] [repeats]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mitsuba3_plt_b200 as mi
from mitsuba3_plt_b200 import scenes

name = sys.argv[1] if len(sys.argv) > 1 else "cbox-grating"
res, spp, rep = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((2, 1024), (3, 8), (4, 1)))
sc = scenes.load(name, resx=res, resy=res, spp=spp)
it = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
for _ in range(rep):
    it.render(sc, spp=spp, seed=0, download=False, develop=False)
print(it.last_stats)
