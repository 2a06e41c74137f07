#!/usr/bin/env python
"""C5 (BASELINE.json configs[4]): gratingcap, 800x600, 1 spp per frame x 240 frames with the camera path of
disk_animation.py:112-123 -- the launch/overhead-bound path.  Prints frames/s and ms/frame through the public API (sensor
update H2D + developed image D2H inside each frame), plus the device-only time per frame.
usage: python tools/anim_bench.py [frames] [spp] [resx] [resy]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import mitsuba3_plt_b200 as mi
from mitsuba3_plt_b200 import animation, scenes

frames, spp, rx, ry = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((1, 240), (2, 1), (3, 800), (4, 600)))
sc = scenes.load("gratingcap", resx=rx, resy=ry)
it = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
animation.render_frames(sc, it, frames=frames, spp=spp, indices=list(range(8)))   # warm-up
dev_ms, launches = [], []
def keep(i, img):
    dev_ms.append(it.last_stats["ms_total"]); launches.append(it.last_stats["kernel_launches"])
t0 = time.perf_counter()
animation.render_frames(sc, it, frames=frames, spp=spp, on_frame=keep)
dt = time.perf_counter() - t0
print(json.dumps({"workload": f"gratingcap {rx}x{ry}, {spp} spp/frame x {frames} frames", "frames_per_s": frames / dt,
                  "ms_per_frame_wall": 1e3 * dt / frames, "ms_per_frame_device": float(np.mean(dev_ms)),
                  "kernel_launches_per_frame": float(np.mean(launches)),
                  "Msamples_per_s_wall": rx * ry * spp * frames / dt / 1e6}))
