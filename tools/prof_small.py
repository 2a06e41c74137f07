import sys; sys.path.insert(0,'.')
import mitsuba3_plt_b200 as mi, json
from mitsuba3_plt_b200 import scenes
for name in ("gratingcap","disk","cbox"):
    sc = scenes.load(name, resx=800 if name!="cbox" else 256, resy=600 if name!="cbox" else 256)
    it = mi.load_dict({"type":"plt","max_depth":7,"rr_depth":50})
    spp = 16 if name=="cbox" else 1
    for _ in range(3): it.render(sc, spp=spp, seed=0)
    it.render(sc, spp=spp, seed=0, profile=True)
    st = it.last_stats
    print(name, round(st["ms_total"],3), {k:(round(v["ms"]*1e3),v["launches"]) for k,v in st["kernels"].items()}, "develop/copy us", round(st["ms_develop"]*1e3), round(st["ms_copy"]*1e3))
