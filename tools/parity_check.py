#!/usr/bin/env python
"""Large-sample parity check: CUDA path vs CPU oracle on a builtin scene (per-sample radiance + image rMSE).
usage: python tools/parity_check.py [scene] [res] [spp] [seed]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import mitsuba3_plt_b200 as mi
from mitsuba3_plt_b200 import scenes
import oracle_lib
from parity import relative_mse, sample_agreement

name = sys.argv[1] if len(sys.argv) > 1 else "cbox-grating"
res, spp, seed = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((2, 256), (3, 16), (4, 0)))
sc = scenes.load(name, resx=res, resy=res)
it = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
dump = {}
img = it.render(sc, spp=spp, seed=seed, dump=dump)
t = time.time()
ref = oracle_lib.render(sc.flat, spp, seed=seed, dump=True)
a, b = dump["radiance"], ref["radiance"]
frac = sample_agreement(a, b)
rmse, mism = relative_mse(img, ref["image"])
nan_gpu, nan_ref = int(np.isnan(a).any(1).sum()), int(np.isnan(b).any(1).sum())
out = {"scene": name, "res": res, "spp": spp, "seed": seed, "samples": len(a), "fraction_within_1e-4": frac, "image_rel_mse": rmse,
       "nan_mask_mismatch": mism, "nan_samples_gpu": nan_gpu, "nan_samples_oracle": nan_ref,
       "stats_gpu": {k: it.last_stats[k] for k in ("closest_rays", "shadow_rays", "vertices", "grating_vertices")},
       "stats_oracle": {k: ref["stats"][k] for k in ("closest_rays", "shadow_rays", "vertices", "grating_vertices")},
       "oracle_seconds": ref["seconds"]}
print(json.dumps(out))
sys.exit(0 if frac >= 0.999 and rmse <= 0.01 else 1)
