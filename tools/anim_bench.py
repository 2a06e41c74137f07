#!/usr/bin/env python
"""C5 (BASELINE.json configs[4]): gratingcap, 800x600, 1 spp per frame x 240 frames with the camera path of
disk_animation.py:112-123 -- the launch/overhead-bound path.  Prints frames/s and ms/frame through the public API (sensor
update H2D + developed image D2H inside each frame), plus the device-only time per frame.
usage: python tools/anim_bench.py [frames] [spp] [resx] [resy] [--inflight=K] [--dist]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import mitsuba3_plt_b200 as mi
from mitsuba3_plt_b200 import animation, scenes

args = [a for a in sys.argv[1:] if not a.startswith("--")]
frames, spp, rx, ry = (int(args[i]) if len(args) > i else d for i, d in ((0, 240), (1, 1), (2, 800), (3, 600)))
inflight = next((int(a.split("=")[1]) for a in sys.argv if a.startswith("--inflight=")), 1)   # frames in flight per GPU
dist_mode = "--dist" in sys.argv    # under torchrun: frames dealt round-robin to the ranks, no collective on the data path
ctx = mi.init_distributed() if dist_mode else mi.context()
sc = scenes.load("gratingcap", resx=rx, resy=ry)
it = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
animation.render_frames(sc, it, frames=frames, spp=spp, indices=list(range(8 * inflight)), ctx=ctx, inflight=inflight)   # warm-up
dev_ms, launches = [], []
def keep(i, img):
    dev_ms.append(it.last_stats["ms_total"]); launches.append(it.last_stats["kernel_launches"])
if dist_mode:
    import torch, torch.distributed as dist
    dist.barrier(); torch.cuda.synchronize()
t0 = time.perf_counter()
animation.render_frames(sc, it, frames=frames, spp=spp, on_frame=keep, ctx=ctx, inflight=inflight)
dt = time.perf_counter() - t0
if dist_mode:
    t = torch.tensor([dt, float(len(dev_ms))], dtype=torch.float64, device="cuda")
    tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    tsum = t.clone(); dist.all_reduce(tsum)
    dt = tmax[0].item()
    assert int(tsum[1].item()) == frames, "every frame rendered exactly once across the ranks"
    if ctx.rank != 0:
        sys.exit(0)
print(json.dumps({"workload": f"gratingcap {rx}x{ry}, {spp} spp/frame x {frames} frames", "ranks": ctx.world, "inflight": inflight, "frames_per_s": frames / dt,
                  "ms_per_frame_wall": 1e3 * dt / frames, "ms_per_frame_device": float(np.mean(dev_ms)),
                  "kernel_launches_per_frame": float(np.mean(launches)),
                  "Msamples_per_s_wall": rx * ry * spp * frames / dt / 1e6}))
