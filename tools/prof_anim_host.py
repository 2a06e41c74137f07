#!/usr/bin/env python
"""cProfile of the C5 frame loop (host side): where the wall time per frame goes beyond the device time."""
import cProfile, pstats, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mitsuba3_plt_b200 as mi
from mitsuba3_plt_b200 import animation, scenes
sc = scenes.load("gratingcap", resx=800, resy=600)
it = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
animation.render_frames(sc, it, frames=240, spp=1, indices=list(range(8)))
pr = cProfile.Profile()
pr.enable()
animation.render_frames(sc, it, frames=240, spp=1, on_frame=lambda i, img: None)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(12)
